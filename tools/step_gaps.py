"""How much of a training step is NOT kernel time: CUPTI trace of two steps of the benchmark configuration (TTT, batch 32,
1 GPU); prints the span of the trace, the summed kernel durations and the idle share (launch bubbles + host stalls),
and the per-kernel totals of one step (the step breakdown at the real depth)."""
import collections
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

from dalle_b200.config import get_config
from dalle_b200.trainer import GroupTrainer

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
cfg = get_config("ttt")
tr = GroupTrainer(cfg, torch.device("cuda", 0), averaging="nccl")
text = torch.randint(1, cfg.num_text_tokens, (B, cfg.text_seq_len), device="cuda")
image = torch.randint(0, cfg.num_image_tokens, (B, cfg.image_seq_len), device="cuda")
for _ in range(3):
    tr.step(text, image)
torch.cuda.synchronize()
N = 2
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(N):
        tr.step(text, image)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.device_time > 0]
ev.sort(key=lambda e: e.time_range.start)
span = ev[-1].time_range.end - ev[0].time_range.start
busy = sum(e.device_time for e in ev)
gaps = [max(0.0, b.time_range.start - a.time_range.end) for a, b in zip(ev, ev[1:])]
print(f"{N} steps: span {span / N / 1e3:.2f} ms per step, kernels {busy / N / 1e3:.2f} ms per step ({len(ev) // N} launches), "
      f"idle {100 * (span - busy) / span:.2f} %; gaps: median {sorted(gaps)[len(gaps) // 2]:.2f} us, "
      f"sum of gaps under 20 us {sum(g for g in gaps if g < 20) / N / 1e3:.2f} ms per step")
agg = collections.defaultdict(lambda: [0, 0.0])
for e in ev:
    k = re.sub(r"\(.*", "", e.name).replace("void ", "").replace("b200::", "")[:70]
    agg[k][0] += 1
    agg[k][1] += e.device_time
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:32]:
    print(f"{t / N / 1e3:9.2f} ms {100 * t / busy:5.1f}%  n={n // N:5d}  {k}")
