"""8-bit LAMB: quantisation maps, block-wise quantiser, optimizer state layout (T0) and fused GPU step (T3)."""
import copy

import pytest
import torch

from dalle_b200.optim.lamb8bit import (BLOCK, LAMB8Bit, create_dynamic_map, dequantize_blockwise_ref,
                                       get_linear_schedule_with_warmup, prepared_code, quantize_blockwise_ref)


def test_dynamic_maps_pinned():
    s, u = create_dynamic_map(True), create_dynamic_map(False)
    assert s.numel() == u.numel() == 256
    assert torch.all(s[1:] > s[:-1]) and torch.all(u[1:] > u[:-1])
    assert s[0].item() == pytest.approx(-0.9929688, abs=1e-6) and s[-1].item() == 1.0 and u[0].item() == 0.0
    assert (s == 0).sum() == 1 and s[127].item() == 0.0
    # smallest magnitudes: level i=0 is one midpoint (0.55) x 1e-6 (signed) / two midpoints (0.325, 0.775) x 1e-6
    assert s[128].item() == pytest.approx(0.55e-6, rel=1e-5)
    assert u[1].item() == pytest.approx(0.325e-6, rel=1e-5) and u[2].item() == pytest.approx(0.775e-6, rel=1e-5)
    assert prepared_code(False)[0].item() == -1.0


def test_blockwise_roundtrip_error():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(3 * BLOCK + 100, generator=g) * 0.01
    code = prepared_code(True)
    q, absmax = quantize_blockwise_ref(x, code)
    assert q.dtype == torch.uint8 and absmax.numel() == 4
    y = dequantize_blockwise_ref(q, absmax, code)
    rel = (x - y).abs() / absmax.repeat_interleave(BLOCK)[: x.numel()]
    assert rel.max().item() < 0.04  # worst-case half-gap of the dynamic code near 1.0
    assert (x - y).abs().mean().item() < 0.02 * x.abs().mean().item() + 1e-6


def _make_params(device, seed=0):
    g = torch.Generator().manual_seed(seed)
    shapes = [(300, 256), (70000,), (1024,), (1, 1, 256), (BLOCK * 16 + 5,)]
    params = [torch.nn.Parameter(torch.randn(*s, generator=g).to(device) * 0.1) for s in shapes]
    return params


def _grads(params, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(p.shape, generator=g) * 0.05 for p in params]


def test_state_layout_and_schedule():
    params = _make_params("cpu")
    opt = LAMB8Bit([{"params": params[:3], "weight_decay": 0.045}, {"params": params[3:], "weight_decay": 0.0}],
                   lr=2.5e-3, betas=(0.9, 0.96), eps=1e-6, weight_decay=0.045, max_grad_norm=4.0, clamp_value=10000.0,
                   reuse_grad_buffers=True)
    sched = get_linear_schedule_with_warmup(opt, 10, 100)
    for p, g in zip(params, _grads(params, 1)):
        p.grad = g.clone()
    opt.step()
    sched.step()
    sd = opt.state_dict()
    assert len(sd["param_groups"]) == 2 and sd["param_groups"][0]["weight_decay"] == 0.045
    st_big, st_small = sd["state"][0], sd["state"][2]
    assert st_big["state1"].dtype == torch.uint8 and st_big["state1"].shape == (300, 256)
    assert st_big["qmap1"].numel() == 256 and st_big["absmax1"].numel() == (300 * 256 + BLOCK - 1) // BLOCK
    assert st_small["state1"].dtype == torch.float32 and "qmap1" not in st_small
    assert {"weight_norm", "step_norm", "trust_ratio", "step"} <= set(st_big)
    assert opt.param_groups[0]["lr"] == pytest.approx(2.5e-3 * 1 / 10)
    opt2 = LAMB8Bit([{"params": _make_params("cpu")[:3]}, {"params": _make_params("cpu")[3:]}], lr=1.0)
    opt2.load_state_dict(copy.deepcopy(sd))
    assert opt2.state_dict()["state"][0]["step"] == 1


def test_quantized_step_tracks_fp32_lamb():
    """8-bit statistics must stay close to an exact-state LAMB over a few steps."""
    pa, pb = _make_params("cpu"), _make_params("cpu")
    kw = dict(lr=1e-2, betas=(0.9, 0.96), eps=1e-6, weight_decay=0.045, max_grad_norm=4.0, clamp_value=10000.0)
    a = LAMB8Bit(pa, **kw)
    b = LAMB8Bit(pb, min_8bit_size=10 ** 12, **kw)
    for it in range(4):
        for p, q, g in zip(pa, pb, _grads(pa, 10 + it)):
            p.grad, q.grad = g.clone(), g.clone()
        a.step(); b.step()
    for p, q in zip(pa, pb):
        assert (p - q).abs().max().item() < 2e-3


@pytest.mark.gpu
def test_fused_gpu_step_matches_cpu_reference():
    pc = _make_params("cpu")
    pg = [torch.nn.Parameter(p.detach().clone().cuda()) for p in pc]
    kw = dict(lr=1e-2, betas=(0.9, 0.96), eps=1e-6, weight_decay=0.045, max_grad_norm=0.3, clamp_value=10000.0)
    groups = lambda ps: [{"params": ps[:3], "weight_decay": 0.045}, {"params": ps[3:], "weight_decay": 0.0}]
    oc, og = LAMB8Bit(groups(pc), **kw), LAMB8Bit(groups(pg), **kw)
    for it in range(4):
        for p, q, g in zip(pc, pg, _grads(pc, 20 + it)):
            p.grad = g.clone()
            if q.grad is None:
                q.grad = g.clone().cuda()
            else:
                q.grad.copy_(g)
        oc.step()
        og.step(zero_grad=True)
        assert not og.step_was_skipped()
    for p, q in zip(pc, pg):
        assert q.grad.abs().max().item() == 0.0
        assert (p - q.detach().cpu()).abs().max().item() < 1e-3, (p - q.detach().cpu()).abs().max().item()
        assert torch.equal(q._bf16_shadow.float(), q.detach().to(torch.bfloat16).float())
    sc, sg = oc.state_dict(), og.state_dict()
    for i in range(len(pc)):
        assert sc["state"][i]["state1"].dtype == sg["state"][i]["state1"].dtype
        tr_c, tr_g = float(sc["state"][i]["trust_ratio"]), float(sg["state"][i]["trust_ratio"])
        assert abs(tr_c - tr_g) / tr_c < 2e-2
    # non-finite gradients: step is skipped, flag raised, parameters untouched
    before = [q.detach().clone() for q in pg]
    pg[0].grad[0, 0] = float("nan")
    og.step()
    assert og.step_was_skipped()
    for q, b in zip(pg, before):
        assert torch.equal(q.detach(), b)
