"""Fused executor on the GPU kernels vs the fp32 oracle (T3)."""
import pytest
import torch

from dalle_b200.config import small_config, tiny_config
from dalle_b200.models.dalle import DALLE

pytestmark = pytest.mark.gpu


def _batch(cfg, b=2, seed=0):
    g = torch.Generator().manual_seed(seed)
    text = torch.randint(1, cfg.num_text_tokens, (b, cfg.text_seq_len), generator=g)
    text[:, -3:] = 0
    image = torch.randint(0, cfg.num_image_tokens, (b, cfg.image_seq_len), generator=g)
    return text, image


def _boost(model, v=0.25):
    with torch.no_grad():
        for p in list(model.scale_f) + list(model.scale_g):
            p.fill_(v)
        for p in list(model.norm_f_w) + list(model.norm_g_w):
            p.uniform_(0.7, 1.3)


@pytest.mark.parametrize("store_layers", [None, 0, 1])   # stored activations / reversible recompute / mixed
@pytest.mark.parametrize("cfg_fn", [tiny_config, small_config])
def test_gpu_executor_matches_fp32_oracle(cfg_fn, store_layers):
    cfg = cfg_fn()
    ref = DALLE(cfg, seed=0)
    _boost(ref)
    model = DALLE(cfg, seed=0)
    model.engine.store_layers = store_layers
    model.load_state_dict(ref.state_dict())
    model.cuda()
    text, image = _batch(cfg)
    ref_loss = ref.forward_reference(text, image, return_loss=True)
    ref_grads = torch.autograd.grad(ref_loss, list(ref.parameters()))
    loss = model(text.cuda(), image.cuda(), return_loss=True)
    assert abs(loss.item() - ref_loss.item()) < 3e-2, (loss.item(), ref_loss.item())
    loss.backward()
    torch.cuda.synchronize()
    bad = []
    for (name, p), g in zip(model.named_parameters(), ref_grads):
        assert p.grad is not None and torch.isfinite(p.grad).all(), name
        err = (p.grad.cpu() - g).abs().max().item()
        den = g.abs().max().item() + 1e-8
        if err / den > 6e-2:
            bad.append((name, err / den))
    assert not bad, bad


def test_gpu_elementwise_kernels_vs_cpu():
    from dalle_b200.ops import elementwise as E

    cfg = small_config()
    B, D = 2, cfg.dim
    M = B * cfg.seq_len
    x = torch.randn(M, D)
    w, b = torch.rand(D) + 0.5, torch.randn(D) * 0.1
    outs = {}
    for dev in ("cpu", "cuda"):
        dt = torch.float32 if dev == "cpu" else torch.bfloat16
        out = torch.zeros(M, D, device=dev, dtype=dt)
        mean, rstd = torch.zeros(M, device=dev), torch.zeros(M, device=dev)
        E.ln_shift_fwd(x.to(dev), None, w.to(dev), b.to(dev), out, mean, rstd, cfg, True)
        dh = torch.randn(M, D, generator=torch.Generator().manual_seed(1)).to(dev).to(dt)
        dx = torch.ones(M, D, device=dev)
        dw, db = torch.zeros(D, device=dev), torch.zeros(D, device=dev)
        E.ln_shift_bwd(dh, x.to(dev), None, mean, rstd, w.to(dev), dx, None, dw, db, cfg, True, 0)
        outs[dev] = [t.float().cpu() for t in (out, mean, rstd, dx, dw, db)]
    for a, c, tol in zip(outs["cpu"], outs["cuda"], (3e-2, 1e-4, 1e-3, 3e-2, 3e-2, 3e-2)):
        assert (a - c).abs().max().item() <= tol * (a.abs().max().item() + 1.0)


def test_gpu_cross_entropy_segment():
    from dalle_b200.ops import elementwise as E

    rows, n, ld = 64, 1000, 1024
    z = torch.randn(rows, ld) * 3
    labels = torch.randint(0, n, (rows,))
    res = {}
    for dev in ("cpu", "cuda"):
        lg = z.to(dev).to(torch.bfloat16)
        if dev == "cpu":
            lg = lg.float()
        loss = torch.zeros(1, device=dev)
        gs = torch.full((1,), 2.0, device=dev)
        E.ce_fwd_bwd(lg, labels.to(dev), n, 0.01, gs, loss, True)
        res[dev] = (loss.cpu(), lg[:, :n].float().cpu())
    assert abs(res["cpu"][0].item() - res["cuda"][0].item()) < 1e-3
    assert (res["cpu"][1] - res["cuda"][1]).abs().max().item() < 2e-4


def test_ttt_dimension_stack_matches_fp32_oracle():
    """Depth-2 stack at the TTT dimensions (dim 1024, 16 heads, 257 + 1024 tokens, vocab 40548), batch 1, against the
    fp32 oracle: loss and EVERY gradient, by maximum error and by per-tensor cosine."""
    from dalle_b200.config import ttt_config

    cfg = ttt_config(depth=2, attn_types=("axial_row", "axial_col"), shared_attn_ids=(0, 1), shared_ff_ids=(0, 1))
    ref = DALLE(cfg, seed=0)
    _boost(ref)
    model = DALLE(cfg, seed=0)
    model.load_state_dict(ref.state_dict())
    model.cuda()
    text, image = _batch(cfg, b=1)
    ref.cuda()   # the oracle runs in fp32 on the GPU (cuBLAS / ATen): 40548-wide logits on the CPU would take minutes
    torch.backends.cuda.matmul.allow_tf32 = False
    ref_loss = ref.forward_reference(text.cuda(), image.cuda(), return_loss=True)
    ref_grads = torch.autograd.grad(ref_loss, list(ref.parameters()))
    loss = model(text.cuda(), image.cuda(), return_loss=True)
    assert abs(loss.item() - ref_loss.item()) < 3e-2, (loss.item(), ref_loss.item())
    loss.backward()
    torch.cuda.synchronize()
    bad = []
    for (name, p), g in zip(model.named_parameters(), ref_grads):
        assert p.grad is not None and torch.isfinite(p.grad).all(), name
        err = (p.grad - g).abs().max().item()
        den = g.abs().max().item() + 1e-12
        cos = torch.nn.functional.cosine_similarity(p.grad.flatten().double(), g.flatten().double(), dim=0).item()
        if err / den > 6e-2 or cos < 0.999:
            bad.append((name, round(err / den, 4), round(cos, 5)))
    assert not bad, bad


def test_head_gradient_in_the_forward_pass_gives_the_same_gradients():
    """engine.expected_grad_out: the loss gradient of the head is formed during the forward pass (one logits GEMM and
    one cross-entropy pass fewer) - same loss, same gradients as the two-pass form, also for a scaled backward."""
    cfg = small_config()
    text, image = _batch(cfg)
    grads = {}
    base = DALLE(cfg, seed=0)
    _boost(base)
    for mode, scale in (("two-pass", 0.5), ("fused", 0.5)):
        model = DALLE(cfg, seed=0)
        model.load_state_dict(base.state_dict())
        model.cuda()
        if mode == "fused":
            model.engine.expected_grad_out = scale
        loss = model(text.cuda(), image.cuda(), return_loss=True)
        (loss * scale).backward()
        torch.cuda.synchronize()
        grads[mode] = (loss.item(), {n: p.grad.clone() for n, p in model.named_parameters()})
    assert abs(grads["two-pass"][0] - grads["fused"][0]) < 1e-5
    # Two identical steps are not bit-identical: the convolution-like layer's backward sums shared-key gradients with
    # fp32 red.add before the bf16 store, and the order of the atomics flips single bf16 ulps (2^-8) that travel down
    # the stack (tools/exp/flake_attn_bwd.py).  Hence a bound of a few ulps on the maximum and a cosine on the whole.
    for n, g in grads["two-pass"][1].items():
        f = grads["fused"][1][n]
        d = (g - f).abs().max().item()
        assert d <= 2e-2 * (g.abs().max().item() + 1e-8), (n, d)
        if g.numel() > 8 and g.abs().max().item() > 0:
            cos = torch.nn.functional.cosine_similarity(g.flatten().double(), f.flatten().double(), dim=0).item()
            assert cos > 0.99999, (n, cos)
