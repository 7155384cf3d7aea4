"""Decode benchmark (BASELINE config 4; workload of reference inference/run_inference.py:79-92, 132): KV-cached sampling
of 1024 image tokens for a batch of 32 prompts per GPU on the TTT model, top-p 0.95 at temperature 1.

    python tools/bench_decode.py [--batch 32] [--reps 2] [--eager] [--out profiles/decode_r2_n1.json]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_decode.py --out ...

Under torchrun every rank samples its own 32 prompts (data parallel over prompts, as the reference's inference workers
do); the time is the device time of the slowest rank, the rate the whole job's.  Random-init weights, random prompts.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from dalle_b200.config import ttt_config
from dalle_b200.models import generate as G
from dalle_b200.models.dalle import DALLE
from dalle_b200.models.decode_engine import DecodeEngine


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32, help="prompts per engine (<= 64)")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--eager", action="store_true", help="also time the PyTorch cached decode (40 positions)")
    ap.add_argument("--engines", type=int, default=1, help="concurrent batches per GPU (one engine + stream each)")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", 1), ("RANK", 0), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = args.batch
    cfg = ttt_config()
    model = DALLE(cfg, seed=0).cuda()
    g = torch.Generator(device="cuda").manual_seed(1000 + rank)
    text = torch.randint(2, cfg.num_text_tokens, (B, cfg.text_seq_len), device="cuda", generator=g)
    from dalle_b200.models.decode_engine import generate_many

    engs = [DecodeEngine(model, B) for _ in range(args.engines)]
    texts = [text] + [torch.randint(2, cfg.num_text_tokens, (B, cfg.text_seq_len), device="cuda", generator=g)
                      for _ in range(args.engines - 1)]
    eng = engs[0]

    def run():
        if args.engines == 1:
            return eng.generate(text, temperature=1.0, top_k=0, top_p=0.95)
        return generate_many(engs, texts, temperature=1.0, top_k=0, top_p=0.95)[0]

    run()   # builds the graphs, warms everything up
    torch.cuda.synchronize()
    times = []
    for _ in range(args.reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(True), torch.cuda.Event(True)
        s.record()
        codes = run()
        e.record()
        torch.cuda.synchronize()
        ms = torch.tensor([s.elapsed_time(e)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        times.append(ms.item())
    assert int(codes.min()) >= 0 and int(codes.max()) < cfg.num_image_tokens
    ms = min(times)
    B_all = B * args.engines
    out = {"metric": "decode_tokens_per_s", "value": world * B_all * cfg.image_seq_len / ms * 1e3, "unit": "image tokens/s",
           "n_gpus": world, "batch_per_gpu": B_all, "concurrent_batches_per_gpu": args.engines,
           "images_per_s": world * B_all / ms * 1e3,
           "ms_per_1024_tokens": ms, "ms_per_position": ms / cfg.seq_len, "ms_all_reps": times,
           "engine": type(eng).__name__, "kv_cache_gib": getattr(eng, "cache_bytes", lambda: 0)() / 2 ** 30,
           "dtype": "bf16", "data": "synthetic (random-init TTT weights, random prompts)",
           "sampling": "top_p 0.95, temperature 1", "timing": "CUDA events, max over ranks, best of reps"}
    if args.eager and rank == 0:
        from dalle_b200.models import reference as R

        cache = G.DecodeCache(cfg, B, text.device, torch.bfloat16)
        tokens, _ = R.embed(model, text, None, torch.bfloat16)
        with torch.no_grad():
            for t in range(24):
                h = G.decode_step_reference(model, tokens[:, t], cache)
            torch.cuda.synchronize()
            t0 = time.time()
            for t in range(24, 64):
                h = G.decode_step_reference(model, tokens[:, t], cache)
                lg = G.logits_from_hidden(model, h, True)
                G.sample_from_logits(lg, 1.0, 0, 0.95)
            torch.cuda.synchronize()
        ms_e = (time.time() - t0) / 40 * 1e3
        out.update({"eager_ms_per_position": ms_e, "speedup_vs_eager": ms_e / out["ms_per_position"]})
    if rank == 0:
        line = json.dumps(out)
        print(line)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            with open(args.out, "w") as f:
                f.write(line + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
