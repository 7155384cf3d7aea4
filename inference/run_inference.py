#!/usr/bin/env python3
"""Text -> images with a trained DALL-E checkpoint (same flags as reference inference/run_inference.py:95-146).

Per query: tokenize, pad to 256 with id 1, repeat to a batch, run ``--n-iters`` rounds of KV-cached sampling of 1024
image tokens, decode with the VQGAN, score with CLIP ViT-B/32 and pickle
``{'query', 'temperature', 'images', 'clip_scores'}`` to ``<output-dir>/<query>.pickle``.
Without ``--vqgan`` the sampled codes are stored instead of images; without ``--clip`` scores are omitted.
"""
import argparse
import os
import pickle
import sys
from datetime import datetime

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dalle_b200.config import get_config  # noqa: E402
from dalle_b200.models.dalle import DALLE, ModelWrapper  # noqa: E402
from dalle_b200.training.tokenizer import load_tokenizer  # noqa: E402


def make_model(preset: str = "ttt", tokenizer_path: str = "t5-small"):
    tokenizer = load_tokenizer(tokenizer_path)
    cfg = get_config(preset, num_text_tokens=tokenizer.vocab_size) if preset == "ttt" else get_config(preset)
    return tokenizer, ModelWrapper(DALLE(cfg))


def encode_query(query: str, tokenizer, text_seq_len: int, batch_size: int, device) -> torch.Tensor:
    ids = tokenizer(query, add_special_tokens=False, max_length=text_seq_len, truncation=True)["input_ids"]
    ids = torch.tensor(ids, dtype=torch.int64)
    ids = torch.nn.functional.pad(ids, (0, text_seq_len - len(ids)), value=1)   # pad id = </s> = 1, as in training
    return ids.repeat(batch_size, 1).to(device)


def generate(query, *, tokenizer, model, batch_size, n_iters, temperature, top_k, top_p, device):
    text = encode_query(query, tokenizer, model.model.cfg.text_seq_len, batch_size, device)
    out = []
    # On the GPU several iterations go into ONE call: generate_images cuts a batch above 32 into slices that decode
    # concurrently (one engine and stream each), which is ~1.5x the tokens/s of running the iterations back to back
    # (the reference loops over n_iters, inference/run_inference.py:87-91).  At most 128 prompts per call.
    per_call = max(1, 128 // max(1, batch_size)) if text.is_cuda else 1
    done = 0
    while done < n_iters:
        k = min(per_call, n_iters - done)
        res = model.model.generate_images(text.repeat(k, 1) if k > 1 else text, temperature=temperature, top_k=top_k,
                                          top_p=top_p, use_cache=True)
        if res.dim() == 4:
            res = res.permute(0, 2, 3, 1)                      # b c h w -> b h w c
        out.extend(res.float().cpu().numpy())
        done += k
    return out


def main(argv=None):
    # (the reference script disables autograd at import time; doing it here keeps importing this module from
    # tests or other tools free of that process-wide side effect)
    with torch.no_grad():
        return _main(argv)


def _main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--queries", type=str, help="List of queries (*.txt, newline-separated)")
    ap.add_argument("--temperature", type=float, default=1.0, help="Sampling temperature")
    ap.add_argument("--top-k", type=int, default=0)
    ap.add_argument("--top-p", type=float, default=1.0)
    ap.add_argument("--model", type=str, help="DALL-E checkpoint (*.pt)")
    ap.add_argument("--vqgan", type=str, help="VQGAN checkpoint (*.ckpt)")
    ap.add_argument("--vqgan-config", type=str, help="VQGAN config (*.yaml)")
    ap.add_argument("--output-dir", type=str, help="Output directory")
    # additions (optional)
    ap.add_argument("--clip", type=str, default=None, help="CLIP ViT-B/32 weights; enables reranking scores")
    ap.add_argument("--clip-bpe", type=str, default=None, help="CLIP BPE vocabulary (bpe_simple_vocab_16e6.txt.gz)")
    ap.add_argument("--batch-size", type=int, default=16)
    ap.add_argument("--n-iters", type=int, default=8)
    ap.add_argument("--model-preset", type=str, default="ttt")
    ap.add_argument("--tokenizer", type=str, default="t5-small")
    args = ap.parse_args(argv)

    with open(args.queries) as f:
        queries = [q for q in (line.rstrip() for line in f) if q]
    print(f"[*] Loaded {len(queries)} queries")
    device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    tokenizer, model = make_model(args.model_preset, args.tokenizer)
    if args.model:
        print(f"[*] Model modification time: {datetime.fromtimestamp(os.stat(args.model).st_mtime)}")
        ok = model.load_state_dict(torch.load(args.model, map_location="cpu", weights_only=False))  # either key spelling
        print(f"[*] Loaded model: {ok}")
    model = model.to(device)
    if args.vqgan:
        from dalle_b200.models.vqgan import VQGanVAE

        model.model.vae = VQGanVAE(args.vqgan, args.vqgan_config).to(device).eval()
    clip_model = clip_tok = None
    if args.clip and args.clip_bpe:
        from dalle_b200.models import clip as clip_mod

        clip_model, clip_tok = clip_mod.load(args.clip, device), clip_mod.SimpleTokenizer(args.clip_bpe)
    os.makedirs(args.output_dir, exist_ok=True)
    print(f"[*] Saving results to `{args.output_dir}`")
    for query in queries:
        images = generate(query, tokenizer=tokenizer, model=model, batch_size=args.batch_size, n_iters=args.n_iters,
                          temperature=args.temperature, top_k=args.top_k, top_p=args.top_p, device=device)
        scores = None
        if clip_model is not None and images and np.asarray(images[0]).ndim == 3:
            from dalle_b200.models import clip as clip_mod

            batch = clip_mod.preprocess(torch.from_numpy(np.stack(images)).to(device))
            text = clip_mod.tokenize(clip_tok, [query]).to(device)
            _, logits_per_text = clip_model(batch, text)
            scores = logits_per_text[0].softmax(dim=-1).cpu().numpy()
        key = "images" if np.asarray(images[0]).ndim == 3 else "codes"
        with open(os.path.join(args.output_dir, f"{query}.pickle"), "wb") as f:
            pickle.dump({"query": query, "temperature": args.temperature, key: images, "clip_scores": scores}, f)
    return 0


if __name__ == "__main__":
    sys.exit(main())
