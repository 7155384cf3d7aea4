"""b200_lamb_step_noalloc (include/b200_api.h): the LAMB step for callers that cannot hand the library any scratch
memory - processes whose device memory is entirely owned by their own pool, or control loops that must not risk an
allocation failing or stalling the device in the middle of an optimizer step.  The entry point keeps its two work
arrays in a fixed __device__ array inside the library, so consecutive calls share (and must re-clear) that array.

The test drives the entry point through the C ABI exactly like a foreign caller would, and checks every output
against the ordinary entry point (b200_lamb_step with caller-owned scratch) on two consecutive steps: the second call
starts with the first call's leftovers in the built-in array.  This file sorts last on purpose, so that a run which
stops at the first failure has already executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_TENSORS = 4096               # B200_LAMB_NOALLOC_MAX_TENSORS
ERR_TOO_MANY = -53               # B200_ERR_NOALLOC_TOO_MANY_TENSORS

_p, _i, _l, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float


def _params(seed):
    g = torch.Generator().manual_seed(seed)
    shapes = [(260, 256), (70001,), (1024,), (1, 1, 256), (4096 * 5 + 3,)]
    return [torch.nn.Parameter((torch.randn(*s, generator=g) * 0.1).cuda()) for s in shapes]


def _make(seed):
    from dalle_b200.optim.lamb8bit import LAMB8Bit

    ps = _params(seed)
    opt = LAMB8Bit([{"params": ps[:3], "weight_decay": 0.045}, {"params": ps[3:], "weight_decay": 0.0}], lr=1e-2,
                   betas=(0.9, 0.96), eps=1e-6, weight_decay=0.045, max_grad_norm=0.3, clamp_value=10000.0)
    opt._build_fused()
    return ps, opt


def _set_grads(ps, seed):
    g = torch.Generator().manual_seed(seed)
    for p in ps:
        p.grad.copy_((torch.randn(p.shape, generator=g) * 0.05).cuda())


def _call_noalloc(lib, opt, trust_out, stats_out, n_tensors=None):
    f, L, g0 = opt._fused, opt._fused.layout, opt.param_groups[0]
    fn = lib.b200_lamb_step_noalloc
    fn.restype = _i
    fn.argtypes = [_p, _p, _p, _l, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _p, _i, _p, _p,
                   _f, _f, _f, _f, _f, _f, _f, _i, _p]
    return fn(L.flat_grad.data_ptr(), L.flat.data_ptr(), L.flat_bf16.data_ptr(), L.n_padded, f.q1.data_ptr(),
              f.q2.data_ptr(), f.f1.data_ptr(), f.f2.data_ptr(), f.absmax1.data_ptr(), f.absmax2.data_ptr(),
              opt.name2qmap["dynamic"].data_ptr(), opt.name2qmap["udynamic"].data_ptr(), f.blocks.data_ptr(),
              f.n_blocks, f.wd.data_ptr(), f.n_tensors if n_tensors is None else n_tensors,
              trust_out.data_ptr(), stats_out.data_ptr(), g0["betas"][0], g0["betas"][1], g0["eps"],
              float(opt.max_grad_norm), opt.clamp_value, g0["lr"], 1.0, 1,
              torch.cuda.current_stream().cuda_stream)


def test_noalloc_lamb_matches_regular_entry_point_twice():
    from dalle_b200.ops import lib as L

    lib = L.cuda_lib()
    ps_a, opt_a = _make(7)      # ordinary entry point (scratch owned by the optimizer object)
    ps_b, opt_b = _make(7)      # new entry point, scratch inside the library
    trust = torch.zeros(3 * opt_b._fused.n_tensors, device="cuda")
    stats = torch.zeros(2, device="cuda")
    for step in range(2):       # the second round finds the first round's values in the built-in array
        _set_grads(ps_a, 100 + step)
        _set_grads(ps_b, 100 + step)
        opt_a.step(zero_grad=True)
        rc = _call_noalloc(lib, opt_b, trust, stats)
        assert rc == 0
        torch.cuda.synchronize()
        fa, fb = opt_a._fused, opt_b._fused
        assert (fa.layout.flat - fb.layout.flat).abs().max().item() < 1e-3
        assert (fa.layout.flat_bf16.float() - fb.layout.flat_bf16.float()).abs().max().item() < 1e-2
        assert fb.layout.flat_grad.abs().max().item() == 0.0
        # 8-bit moments: identical up to the rare code flip caused by the order of the atomic partial sums
        assert (fa.q1 != fb.q1).float().mean().item() < 1e-3 and (fa.q2 != fb.q2).float().mean().item() < 1e-3
        assert torch.allclose(fa.absmax1, fb.absmax1, rtol=1e-4, atol=1e-9)
        assert torch.allclose(fa.f1, fb.f1, rtol=1e-4, atol=1e-7) and torch.allclose(fa.f2, fb.f2, rtol=1e-4, atol=1e-9)
        assert torch.allclose(fa.trust, trust, rtol=1e-4)
        assert torch.allclose(fa.stats, stats, rtol=1e-4) and stats[1].item() == 0.0


def test_noalloc_lamb_rejects_more_tensors_than_its_scratch_holds():
    from dalle_b200.ops import lib as L

    lib = L.cuda_lib()
    ps, opt = _make(9)
    _set_grads(ps, 5)
    before = opt._fused.layout.flat.clone()
    trust, stats = torch.zeros(3, device="cuda"), torch.zeros(2, device="cuda")
    rc = _call_noalloc(lib, opt, trust, stats, n_tensors=MAX_TENSORS + 1)
    torch.cuda.synchronize()
    assert rc == ERR_TOO_MANY
    assert torch.equal(before, opt._fused.layout.flat)  # nothing was launched
