"""Sanity run at full TTT scale: 30 optimizer steps on ONE fixed synthetic batch must drive the loss down
(end-to-end check of every forward/backward kernel + LAMB-8bit at the benchmark shapes)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import get_config
from dalle_b200.trainer import GroupTrainer

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
cfg = get_config("ttt")
tr = GroupTrainer(cfg, torch.device("cuda", 0), averaging="nccl", lr=2e-3, warmup_steps=2, total_steps=1000)
g = torch.Generator().manual_seed(0)
text = torch.randint(1, cfg.num_text_tokens, (B, cfg.text_seq_len), generator=g).cuda()
image = torch.randint(0, cfg.num_image_tokens, (B, cfg.image_seq_len), generator=g).cuda()
losses = []
for step in range(30):
    losses.append(float(tr.step(text, image)))
print("losses", " ".join(f"{x:.3f}" for x in losses[::3]), "->", f"{losses[-1]:.3f}")
assert all(map(lambda v: v == v, losses)), "NaN"
assert losses[-1] < losses[0] - 0.5, "loss did not go down"
print("OVERFIT_OK")
