"""Tiny training step (forward + backward + optimizer) for compute-sanitizer runs:
   compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_step.py [store_layers]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import small_config
from dalle_b200.trainer import GroupTrainer

store = None if len(sys.argv) < 2 else int(sys.argv[1])
cfg = small_config()
tr = GroupTrainer(cfg, torch.device("cuda", 0), averaging="nccl")
tr.model.engine.store_layers = store
g = torch.Generator().manual_seed(0)
text = torch.randint(1, cfg.num_text_tokens, (2, cfg.text_seq_len), generator=g).cuda()
image = torch.randint(0, cfg.num_image_tokens, (2, cfg.image_seq_len), generator=g).cuda()
for _ in range(2):
    loss = tr.step(text, image)
torch.cuda.synchronize()
print("loss", float(loss))
