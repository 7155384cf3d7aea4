"""Training data (reference data.py:11-47): streaming ``laion/laion_100m_vqgan_f8`` records
``{caption, NSFW, original_width, original_height, code}`` -> ``{input_ids, attention_mask, image}``.

Filters and decoding follow the reference: caption >= 3 chars, NSFW == 'UNLIKELY', aspect ratio <= 2, tokenise
without special tokens truncated to ``max_sequence_length``, VQGAN codes = 1024 little-endian int16 -> int64.
Without network access (``--synthetic_data``, or when ``datasets`` cannot reach the hub) an infinite synthetic
stream of the same shapes and dtypes is produced.
"""
from __future__ import annotations

import itertools
import logging
from typing import Iterator, Optional

import numpy as np
import torch

logger = logging.getLogger(__name__)


def keep_example(caption, nsfw, width, height) -> bool:
    return (caption is not None and len(caption) >= 3 and nsfw == "UNLIKELY" and width > 0 and height > 0
            and max(height / width, width / height) <= 2)


def preprocess_batch(batch, tokenizer, max_sequence_length: int):
    keep = [keep_example(c, n, w, h) for c, n, w, h in
            zip(batch["caption"], batch["NSFW"], batch["original_width"], batch["original_height"])]
    logger.debug(f"{100 * np.mean(keep):.1f}% of examples left after filtering")
    captions = list(itertools.compress(batch["caption"], keep))
    if captions:
        out = dict(tokenizer(captions, add_special_tokens=False, max_length=max_sequence_length, truncation=True))
    else:
        out = {"input_ids": [], "attention_mask": []}  # tokenizer([]) raises
    out["image"] = [np.frombuffer(code, np.int16).astype(np.int64) for code in itertools.compress(batch["code"], keep)]
    return out


class SyntheticDalleDataset(torch.utils.data.IterableDataset):
    """Random text ids (ragged, to exercise padding) and VQGAN codes of the LAION record shape."""

    def __init__(self, vocab_size: int, max_sequence_length: int, image_tokens: int = 1024, image_vocab: int = 8192,
                 seed: int = 0):
        self.vocab, self.seq, self.img, self.img_vocab, self.seed = vocab_size, max_sequence_length, image_tokens, image_vocab, seed

    def __iter__(self) -> Iterator[dict]:
        rng = np.random.default_rng(self.seed)
        while True:
            n = int(rng.integers(3, self.seq + 1))
            ids = rng.integers(2, self.vocab, size=n, dtype=np.int64)
            yield {"input_ids": torch.from_numpy(ids), "attention_mask": torch.ones(n, dtype=torch.int64),
                   "image": torch.from_numpy(rng.integers(0, self.img_vocab, size=self.img, dtype=np.int64))}


def make_dataset(tokenizer, *, shuffle_buffer_size: int = 8192, shuffle_seed: Optional[int],
                 preprocessing_batch_size: int = 256, max_sequence_length: int, synthetic: bool = False,
                 image_tokens: int = 1024, image_vocab: int = 8192):
    if not synthetic:
        try:
            from datasets import load_dataset

            ds = load_dataset("laion/laion_100m_vqgan_f8", split="train", streaming=True)
            ds = ds.shuffle(shuffle_buffer_size, seed=shuffle_seed)
            ds = ds.map(lambda b: preprocess_batch(b, tokenizer, max_sequence_length), batched=True,
                        batch_size=preprocessing_batch_size)
            return ds.with_format("torch")
        except Exception as e:
            logger.warning(f"cannot stream laion/laion_100m_vqgan_f8 ({type(e).__name__}); using synthetic data")
    vocab = getattr(tokenizer, "vocab_size", 32100)
    return SyntheticDalleDataset(vocab, max_sequence_length, image_tokens, image_vocab, seed=shuffle_seed or 0)


class PadCollator:
    """``DataCollatorWithPadding(padding='max_length')`` with pad id = eos id (reference task.py:59, 178-181)."""

    def __init__(self, pad_token_id: int, max_length: int):
        self.pad, self.max_length = pad_token_id, max_length

    def __call__(self, features):
        n = len(features)
        ids = torch.full((n, self.max_length), self.pad, dtype=torch.int64)
        mask = torch.zeros((n, self.max_length), dtype=torch.int64)
        for i, f in enumerate(features):
            t = torch.as_tensor(f["input_ids"], dtype=torch.int64)[: self.max_length]
            ids[i, : t.numel()] = t
            mask[i, : t.numel()] = 1
        image = torch.stack([torch.as_tensor(f["image"], dtype=torch.int64) for f in features])
        return {"input_ids": ids, "attention_mask": mask, "image": image}
