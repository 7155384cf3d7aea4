#!/usr/bin/env python3
"""Headline benchmark: DALL-E (TTT config) training samples/s on N B200s of one box.

    python bench.py --gpus N --steps K --warmup W            # N > 1: launched by torchrun, one rank per GPU
    python bench.py --impl reference ...                     # the unmodified reference, if it could be installed

One "step" = forward + backward of the full 64-layer model on every rank's micro-batch (bf16 compute, fp32 master
weights), gradient averaging across the ranks, ONE LAMB-8bit optimizer update (global-norm clip, 8-bit moments,
trust ratio), gradients zeroed.  Weak scaling: the per-GPU micro-batch is fixed.  Synthetic VQGAN codes / text ids,
random-init weights (no network on the box).

Prints ONE JSON line (rank 0).  ``value`` is device-timed (CUDA events, max over ranks) on device-resident inputs;
``e2e`` is the same metric through the public API ``GroupTrainer.step_from_host`` with the per-step host->device
copy from pinned memory and a device->host read of the loss inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--checkpoint", action="store_true", help="eager arm: checkpoint every layer")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "eager", "naive"],
                    help="b200: this framework; reference: the unmodified reference (not installable offline); eager: baseline "
                         "B1 = the same model on cuBLAS / SDPA / DDP-NCCL / foreach-LAMB (dalle_b200/baseline_ddp.py); naive: "
                         "the fp32-oracle model run in bf16 with per-tensor all-reduce (the slow stand-in of round 1)")
    ap.add_argument("--batch", type=int, default=32, help="micro-batch per GPU")
    ap.add_argument("--preset", default="ttt")
    ap.add_argument("--averaging", default="auto", choices=["auto", "fused", "nccl"])
    ap.add_argument("--payload", default="fp32", choices=["fp32", "bf16", "u8", "fp16", "u8h", "adaptive"],
                    help="gradient payload of the fused intra-box averaging kernel; adaptive = hivemind's size-adaptive "
                         "fp16 / uniform-8-bit codecs (BASELINE config 3)")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i",
                 str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for n, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def reference_arm(args):
    """The reference cannot be installed on this image: it has no setup.py/pyproject (pip refuses the directory) and
    its pinned dependencies hivemind / dalle-pytorch / bitsandbytes-cuda111 are GitHub archives and a CUDA-11.1
    wheel that are neither in /opt/wheelhouse nor importable ("Reference install")."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    why = None
    if not os.path.isdir(ref_dir) or not os.listdir(ref_dir):
        why = ("reference not installable offline: /root/reference has no setup.py/pyproject and needs hivemind, "
               "dalle-pytorch, bitsandbytes-cuda111 which are absent from /opt/wheelhouse")
    else:
        sys.path.insert(0, ref_dir)
        try:
            import hivemind  # noqa: F401
            import dalle_pytorch  # noqa: F401
        except Exception as e:  # pragma: no cover
            why = f"reference dependencies not importable: {type(e).__name__}: {e}"
    print(json.dumps({"impl": "reference", "unavailable": why or "reference present but no runnable entry point"}))
    return 0


def activation_policy(trainer, B, cfg):
    """How many layers kept their activations for the backward (the rest are recomputed, reversible-style)."""
    try:
        eng = trainer.model.engine
        bufs = eng._bufs.get((B, str(trainer.device), True))
        n = len(bufs.saved) if bufs is not None else 0
        return f"stored for {n}/{cfg.depth} layers, reversible recompute for the rest (same gradients)"
    except Exception:  # reference arm / eager baseline
        return "n/a"


def main():
    args = parse()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist

    from dalle_b200.config import get_config
    from dalle_b200.ops import lib
    from dalle_b200.trainer import GroupTrainer, init_distributed_from_env

    rank, local, world = init_distributed_from_env()
    assert world == args.gpus or world == 1 and args.gpus == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    cfg = get_config(args.preset)
    B, K, W = args.batch, args.steps, max(args.warmup, 3)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    baseline_note = None
    if args.impl == "naive":
        from dalle_b200.baseline_eager import EagerTrainer

        trainer = EagerTrainer(cfg, dev)
    elif args.impl == "eager":
        from dalle_b200.baseline_ddp import DDPTrainer

        # store every activation if the batch fits (what the product does); otherwise checkpoint per layer
        def build(ckpt):
            return DDPTrainer(cfg, dev, use_checkpoint=ckpt)

        trainer = build(args.checkpoint)
        if not args.checkpoint:
            g0 = torch.Generator().manual_seed(7)
            tt = torch.randint(1, cfg.num_text_tokens, (B, cfg.text_seq_len), generator=g0).to(dev)
            ti = torch.randint(0, cfg.num_image_tokens, (B, cfg.image_seq_len), generator=g0).to(dev)
            fits = torch.ones(1, device=dev)
            try:
                trainer.step(tt, ti)
            except torch.cuda.OutOfMemoryError:
                fits.zero_()
            if world > 1:
                dist.all_reduce(fits, op=dist.ReduceOp.MIN)
            if fits.item() == 0:
                del trainer
                import gc

                gc.collect()
                torch.cuda.empty_cache()
                trainer = build(True)
        baseline_note = ("activations stored by autograd" if not trainer.net.use_checkpoint
                         else "per-layer activation checkpointing (the stored-activation graph does not fit at this batch)")
    else:
        trainer = GroupTrainer(cfg, dev, averaging=args.averaging, seed=0, payload=args.payload)

    gen = torch.Generator().manual_seed(1234 + rank)
    n_batches = 4
    host_text = [torch.randint(1, cfg.num_text_tokens, (B, cfg.text_seq_len), generator=gen).pin_memory()
                 for _ in range(n_batches)]
    host_img = [torch.randint(0, cfg.num_image_tokens, (B, cfg.image_seq_len), generator=gen).pin_memory()
                for _ in range(n_batches)]
    dev_text = [t.to(dev) for t in host_text]
    dev_img = [t.to(dev) for t in host_img]

    # ---- device-timed headline ------------------------------------------------------------------
    for i in range(W):
        trainer.step(dev_text[i % n_batches], dev_img[i % n_batches])
    sync_all()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = lib.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    # exposed averaging + optimizer time = the part of a step between the end of backward and the next forward,
    # bracketed by events INSIDE the timed loop (so it stays "exposed" time if the update is ever overlapped)
    marks = []
    for i in range(K):
        if args.impl == "b200":
            loss = trainer.accumulate(dev_text[i % n_batches], dev_img[i % n_batches])
            ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ea.record()
            trainer.optimizer_step()
            eb.record()
            marks.append((ea, eb))
        else:
            loss = trainer.step(dev_text[i % n_batches], dev_img[i % n_batches])
    ev1.record()
    sync_all()
    launches = lib.launch_count - n0
    exposed_ms = None
    if marks:
        t = torch.tensor([sum(a.elapsed_time(b) for a, b in marks) / len(marks)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        exposed_ms = t.item()
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = ms.item()
    value = world * B * K / (ms_total / 1e3)
    final_loss = float(loss.item())

    # ---- end-to-end through the public API ---------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        for i in range(2):
            trainer.step_from_host(host_text[i % n_batches], host_img[i % n_batches])
        sync_all()
        t0 = time.perf_counter()
        for i in range(K):
            trainer.step_from_host(host_text[i % n_batches], host_img[i % n_batches])
        sync_all()
        dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": world * B * K / dt.item(), "unit": "samples/s",
               "h2d_bytes_per_step": (host_text[0].numel() + host_img[0].numel()) * 8,
               "d2h_bytes_per_step": 4}

    if rank == 0:
        out = {
            "metric": "DALL-E samples/sec at 1/2/4/8 B200 (device-timed); exposed averaging ms/step",
            "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic (random text ids + VQGAN codes, random-init weights)",
            "impl": args.impl,
            "config": {"model": "TTT DALL-E dim1024 depth64 heads16 shared-attn/ff axial_row/col+conv_like, "
                                "256 text + 1024 image tokens, vocab 8192" if args.preset == "ttt" else args.preset,
                       "global_batch": B * world, "micro_batch_per_gpu": B, "seq_len": cfg.seq_len,
                       "parallelism": f"dp{world}",
                       "optimizer": "LAMB (fp32 moments, torch._foreach), one update per step" if args.impl == "eager"
                       else "LAMB-8bit, one update per step",
                       "averaging": getattr(trainer, "averaging", "n/a"),
                       "averaging_payload": getattr(trainer, "payload", "n/a"),
                       "activations": baseline_note or activation_policy(trainer, B, cfg),
                       "l2": "per-step working set (>5 GB activations/weights) exceeds the 126 MB L2; no flush"},
            "exposed_averaging_ms_per_step": exposed_ms,
            "exposed_averaging_share": (exposed_ms / (ms_total / K)) if exposed_ms else None,
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "final_loss": final_loss,
        }
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
