"""Run under torchrun (N ranks, one GPU each): fused NVLink averaging + sharded LAMB vs NCCL all-reduce + the
single-GPU LAMB kernels (T4).  Prints GROUP_OK on rank 0 when everything matches."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import torch.distributed as dist

from dalle_b200.optim.lamb8bit import LAMB8Bit
from dalle_b200.parallel.group import FusedGroupAverager
from dalle_b200.trainer import init_distributed_from_env


def make(dev, seed=0):
    g = torch.Generator().manual_seed(seed)
    shapes = [(300, 256), (70001,), (1024,), (1, 1, 256), (4096 * 21 + 3,), (513, 1024)]
    ps = [torch.nn.Parameter((torch.randn(*s, generator=g) * 0.1).to(dev)) for s in shapes]
    opt = LAMB8Bit([{"params": ps[:4], "weight_decay": 0.045}, {"params": ps[4:], "weight_decay": 0.0}], lr=1e-2,
                   betas=(0.9, 0.96), eps=1e-6, weight_decay=0.045, max_grad_norm=0.5, clamp_value=10000.0)
    opt._build_fused()
    return ps, opt


def make_ttt(dev, seed=0):
    """The real parameter layout of the TTT model: 423 tensors, 125,894,244 elements, tensors straddling shards."""
    from dalle_b200.config import ttt_config
    from dalle_b200.models.dalle import DALLE
    from dalle_b200.trainer import param_groups

    model = DALLE(ttt_config(), seed=seed).to(dev)
    opt = LAMB8Bit(param_groups(model, 0.045), lr=1e-2, betas=(0.9, 0.96), eps=1e-6, weight_decay=0.045,
                   max_grad_norm=0.5, clamp_value=10000.0)
    opt._build_fused()
    return opt._fused.layout.params, opt


def init_single_rank():
    """world = 1 (a self-peer): lets a one-GPU box execute the fused kernel."""
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(29650 + os.getpid() % 200))
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    return 0, 0, 1


TOL = {"fp32": 1e-3, "bf16": 3e-3, "u8": 2e-2, "fp16": 3e-3, "u8h": 3e-2, "adaptive": 3e-2}
GN_TOL = {"fp32": 1e-4, "bf16": 1e-2, "u8": 2e-2, "fp16": 1e-2, "u8h": 4e-2, "adaptive": 4e-2}


def main():
    payload = sys.argv[1] if len(sys.argv) > 1 else "fp32"
    layout = sys.argv[2] if len(sys.argv) > 2 else "small"
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        rank, local, world = init_distributed_from_env()
    else:
        rank, local, world = init_single_rank()
    dev = torch.device("cuda", local)
    mk = make_ttt if layout == "ttt" else make
    ps_a, opt_a = mk(dev)
    ps_b, opt_b = mk(dev)
    group = FusedGroupAverager(opt_a, payload=payload)
    for step in range(3 if layout == "small" else 2):
        g = torch.Generator(device=dev).manual_seed(1000 * step + rank)
        grads = [torch.randn(p.shape, generator=g, device=dev) * 0.05 for p in ps_a]
        for p, q, gr in zip(ps_a, ps_b, grads):
            p.grad.copy_(gr)
            q.grad.copy_(gr)
        group.step(grad_scale=1.0 / world)
        flat_b = opt_b._fused.layout.flat_grad
        dist.all_reduce(flat_b)
        opt_b.step(grad_scale=1.0 / world, zero_grad=True)
        torch.cuda.synchronize()
        assert not group.step_was_skipped()
        La, Lb = opt_a._fused.layout, opt_b._fused.layout
        tol = TOL[payload]
        err = (La.flat - Lb.flat).abs().max().item()
        assert err < tol, f"step {step}: fused vs nccl path differ by {err}"
        assert La.flat_grad.abs().max().item() == 0.0
        assert torch.equal(La.flat_bf16.float(), La.flat.to(torch.bfloat16).float())
        # every rank must hold bit-identical parameters after the all-gather
        mine = La.flat.clone()
        ref = mine.clone()
        dist.broadcast(ref, src=0)
        assert torch.equal(mine, ref), "ranks diverged"
        gn_a, gn_b = group.grad_norm(), opt_b.grad_norm() / world  # the single-GPU entry reports the unscaled norm
        assert abs(gn_a - gn_b) / gn_b < GN_TOL[payload], (gn_a, gn_b)
    # non-finite gradient on ONE rank: every rank skips the step
    before = opt_a._fused.layout.flat.clone()
    for p in ps_a:
        p.grad.normal_()
    if rank == world - 1:
        ps_a[0].grad[0, 0] = float("inf")
    group.step(grad_scale=1.0 / world)
    torch.cuda.synchronize()
    assert group.step_was_skipped() and torch.equal(before, opt_a._fused.layout.flat)
    dist.barrier()
    if rank == 0:
        print(f"GROUP_OK world={world} payload={payload} layout={layout}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
