"""Tokenizer loading.  The reference uses ``T5TokenizerFast.from_pretrained('t5-small')`` with ``pad = eos``
(reference task.py:58-59).  The t5-small files are not cached on isolated hosts, so when ``tokenizer_path`` cannot be
loaded a deterministic stand-in with the same vocabulary size / special ids is used (synthetic-data runs only need
ids of the right range)."""
from __future__ import annotations

import hashlib
import logging
import os
import re
from typing import List

logger = logging.getLogger(__name__)


class HashingTokenizer:
    """Whitespace/punctuation pieces hashed into [2, vocab_size); id 0 = <pad>, 1 = </s> like T5."""

    vocab_size = 32100
    eos_token, eos_token_id = "</s>", 1
    pad_token, pad_token_id = "</s>", 1

    def _encode(self, text: str) -> List[int]:
        pieces = re.findall(r"\w+|[^\w\s]", text.lower())
        return [2 + int.from_bytes(hashlib.blake2s(p.encode(), digest_size=4).digest(), "little") % (self.vocab_size - 2)
                for p in pieces]

    def __call__(self, texts, add_special_tokens=False, max_length=None, truncation=False, **kw):
        single = isinstance(texts, str)
        items = [texts] if single else list(texts)
        ids = [self._encode(t) for t in items]
        if truncation and max_length is not None:
            ids = [i[:max_length] for i in ids]
        out = {"input_ids": ids, "attention_mask": [[1] * len(i) for i in ids]}
        return {k: v[0] for k, v in out.items()} if single else out


def load_tokenizer(path: str):
    if path and (os.path.isdir(path) or os.path.isfile(path)):
        try:
            from transformers import T5TokenizerFast

            tok = T5TokenizerFast.from_pretrained(path)
            tok.pad_token = tok.eos_token
            return tok
        except Exception as e:
            logger.warning(f"could not load tokenizer from {path}: {e!r}")
    try:
        os.environ.setdefault("HF_HUB_OFFLINE", "1")
        from transformers import T5TokenizerFast

        tok = T5TokenizerFast.from_pretrained(path, local_files_only=True)
        tok.pad_token = tok.eos_token
        return tok
    except Exception:
        logger.warning(f"tokenizer {path!r} is not available offline; using the hashing stand-in "
                       f"(same vocabulary size and special ids)")
        return HashingTokenizer()
