"""The fused executor's hand-written forward/backward (run on the CPU oracle ops) vs plain autograd through the
naive reference model (SURVEY §4.3 T1 ii)."""
import torch
import pytest

from dalle_b200.config import small_config, tiny_config
from dalle_b200.models.dalle import DALLE


def _batch(cfg, b=2, seed=0):
    g = torch.Generator().manual_seed(seed)
    text = torch.randint(1, cfg.num_text_tokens, (b, cfg.text_seq_len), generator=g)
    text[:, -3:] = 0  # exercise the unique-pad remap
    image = torch.randint(0, cfg.num_image_tokens, (b, cfg.image_seq_len), generator=g)
    return text, image


def _boost(model, v=0.25):
    with torch.no_grad():
        for p in list(model.scale_f) + list(model.scale_g):
            p.fill_(v)
        for p in list(model.norm_f_w) + list(model.norm_g_w):
            p.uniform_(0.7, 1.3)
        for p in list(model.norm_f_b) + list(model.norm_g_b):
            p.uniform_(-0.2, 0.2)


# store_layers: how many layers keep their activations instead of being recomputed by the reversible backward
# (None = all that fit; 0 = the reference's pure reversible scheme; 1 = one stored layer in front of reversible ones)
@pytest.mark.parametrize("store_layers", [None, 0, 1])
@pytest.mark.parametrize("cfg_fn", [tiny_config, small_config])
def test_executor_matches_autograd(cfg_fn, store_layers):
    cfg = cfg_fn()
    model = DALLE(cfg, seed=0)
    model.engine.store_layers = store_layers
    _boost(model)
    text, image = _batch(cfg)
    ref_loss = model.forward_reference(text, image, return_loss=True)
    ref_grads = torch.autograd.grad(ref_loss, list(model.parameters()))
    loss = model(text, image, return_loss=True)
    assert abs(loss.item() - ref_loss.item()) < 1e-4
    loss.backward()
    for (name, p), g in zip(model.named_parameters(), ref_grads):
        assert p.grad is not None, name
        err = (p.grad - g).abs().max().item()
        den = g.abs().max().item() + 1e-8
        assert err / den < 2e-3, f"{name}: {err} / {den}"


def test_executor_accumulates_and_scales_gradients():
    cfg = tiny_config()
    model = DALLE(cfg, seed=1)
    _boost(model)
    text, image = _batch(cfg, seed=3)
    model(text, image, return_loss=True).backward()
    g1 = [p.grad.clone() for p in model.parameters()]
    (0.5 * model(text, image, return_loss=True)).backward()
    for p, g in zip(model.parameters(), g1):
        assert torch.allclose(p.grad, 1.5 * g, rtol=1e-4, atol=1e-7)


def test_executor_eval_and_logits():
    cfg = tiny_config()
    model = DALLE(cfg, seed=2)
    _boost(model)
    text, image = _batch(cfg)
    with torch.no_grad():
        a = model(text, image, return_loss=True)
        b = model.forward_reference(text, image, return_loss=True)
        la = model(text, image[:, :-1])
        lb = model.forward_reference(text, image[:, :-1])
    assert abs(a.item() - b.item()) < 1e-4
    from dalle_b200.models.reference import logits_mask
    m = ~logits_mask(cfg, cfg.seq_len, "cpu")
    assert torch.allclose(la[:, m], lb[:, m], atol=1e-3, rtol=1e-3)
