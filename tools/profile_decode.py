"""Short program for an ncu launch list: three eager (non-graph) decode steps of the TTT decode engine at position 600."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import ttt_config
from dalle_b200.models.dalle import DALLE
from dalle_b200.models.decode_engine import DecodeEngine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
cfg = ttt_config()
model = DALLE(cfg, seed=0).cuda()
eng = DecodeEngine(model, B)
with torch.no_grad():
    for _ in range(3):
        eng.pos.fill_(600)
        eng._step(True, 1.0, 0, 0.95)
torch.cuda.synchronize()
print("ok")
