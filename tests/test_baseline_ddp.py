"""The library-kernel baseline arm (bench.py --impl eager) states the same function as the oracle model."""
import torch

from dalle_b200.baseline_ddp import EagerDalle, ForeachLamb
from dalle_b200.config import tiny_config


def test_eager_baseline_matches_oracle_loss_and_lamb_moves_weights():
    cfg = tiny_config()
    net = EagerDalle(cfg, seed=3)
    g = torch.Generator().manual_seed(0)
    text = torch.randint(1, cfg.num_text_tokens, (2, cfg.text_seq_len), generator=g)
    image = torch.randint(0, cfg.num_image_tokens, (2, cfg.image_seq_len), generator=g)
    loss = net(text, image)
    ref = net.p.forward_reference(text, image, return_loss=True, build="dense")
    assert abs(loss.item() - ref.item()) < 1e-4, (loss.item(), ref.item())
    loss.backward()
    named = list(net.p.reference_named_parameters())
    opt = ForeachLamb([p for n, p in named if "bias" not in n], [p for n, p in named if "bias" in n], lr=1e-2,
                      betas=(0.9, 0.96), eps=1e-6, weight_decay=0.045, max_grad_norm=4.0, clamp_value=10000.0)
    before = net.p.emb_weight.detach().clone()
    opt.step()
    assert not torch.equal(before, net.p.emb_weight)
    assert net(text, image).item() < loss.item()
