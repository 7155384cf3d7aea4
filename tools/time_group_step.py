"""Device time of the fused averaging + LAMB step (parallel/group.py) on the TTT parameter layout for every payload
codec.  world = 1 (a self-peer: shows encode / LAMB / barrier costs without NVLink) or under torchrun."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "multigpu"))
import torch
import torch.distributed as dist

from check_group import init_single_rank, make_ttt
from dalle_b200.parallel.group import FusedGroupAverager
from dalle_b200.trainer import init_distributed_from_env

if int(os.environ.get("WORLD_SIZE", "1")) > 1:
    rank, local, world = init_distributed_from_env()
else:
    rank, local, world = init_single_rank()
dev = torch.device("cuda", local)
for payload in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["fp32", "bf16", "u8", "fp16", "u8h", "adaptive"]):
    ps, opt = make_ttt(dev)
    group = FusedGroupAverager(opt, payload=payload)
    L = opt._fused.layout
    g = torch.Generator(device=dev).manual_seed(rank)
    times = []
    for it in range(6):
        L.flat_grad.copy_(torch.randn(L.flat_grad.shape, generator=g, device=dev) * 0.01)
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        group.step(grad_scale=1.0 / world)
        b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    t = torch.tensor([min(times[2:])], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"world {world} payload {payload:9s}: fused step {t.item():.2f} ms", flush=True)
    del group, opt, ps
    torch.cuda.empty_cache()
dist.destroy_process_group()
