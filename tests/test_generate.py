"""T1 (iii, iv): cached decode logits == full re-forward logits at every position; sampler filters."""
import torch

from dalle_b200.config import DalleConfig
from dalle_b200.models.dalle import DALLE
from dalle_b200.models import generate as G


def _cfg():
    return DalleConfig(num_text_tokens=200, text_seq_len=8, dim=128, depth=3, heads=2, dim_head=64, image_fmap_size=4,
                       num_image_tokens=64, attn_types=("axial_row", "axial_col", "conv_like"),
                       shared_attn_ids=(0, 1, "w_conv"), shared_ff_ids=(0, 1, "w_conv"), conv_kernel_size=3)


def test_cached_decode_matches_full_forward():
    cfg = _cfg()
    model = DALLE(cfg, seed=0)
    with torch.no_grad():
        for p in list(model.scale_f) + list(model.scale_g):
            p.fill_(0.4)
    text = torch.randint(1, cfg.num_text_tokens, (2, cfg.text_seq_len))
    g1, g2 = torch.Generator().manual_seed(5), torch.Generator().manual_seed(5)
    codes_c, logits_c = G.generate_codes_reference(model, text, use_cache=True, generator=g1, return_logits=True,
                                                   dtype=torch.float32)
    codes_f, logits_f = G.generate_codes_reference(model, text, use_cache=False, generator=g2, return_logits=True,
                                                   dtype=torch.float32)
    assert torch.equal(codes_c, codes_f)
    for a, b in zip(logits_c, logits_f):
        assert torch.allclose(a, b, atol=2e-4, rtol=1e-4)
    assert codes_c.shape == (2, cfg.image_seq_len) and codes_c.max() < cfg.num_image_tokens


def test_top_k_top_p_filters():
    logits = torch.tensor([[2.0, 1.0, 0.5, -1.0, -3.0]])
    k2 = G.filter_logits(logits, top_k=2)
    assert torch.isfinite(k2).sum() == 2 and torch.isfinite(k2[0, :2]).all()
    p = G.filter_logits(logits, top_p=0.6)
    probs = logits.softmax(-1)
    assert torch.isfinite(p[0, 0]) and (probs[0, 0] > 0.6) == (torch.isfinite(p).sum() == 1)
    none = G.filter_logits(logits, top_k=0, top_p=1.0)
    assert torch.equal(none, logits)
    s = G.sample_from_logits(torch.tensor([[0.0, 50.0, 0.0]]), temperature=0.5)
    assert s.item() == 1
