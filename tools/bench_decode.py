"""Decode benchmark (BASELINE config 4): KV-cached sampling of 1024 image tokens, batch 32, TTT model, 1 GPU.
Prints tokens/s and images/s for the CUDA-graph decode engine and (short run) the eager PyTorch cached decode."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import ttt_config
from dalle_b200.models.dalle import DALLE
from dalle_b200.models.decode_engine import DecodeEngine
from dalle_b200.models import generate as G

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
cfg = ttt_config()
model = DALLE(cfg, seed=0).cuda()
text = torch.randint(2, cfg.num_text_tokens, (B, cfg.text_seq_len), device="cuda")
eng = DecodeEngine(model, B)
eng.generate(text, temperature=1.0, top_k=0, top_p=0.95)   # builds graphs + warm-up
torch.cuda.synchronize()
s, e = torch.cuda.Event(True), torch.cuda.Event(True)
s.record()
codes = eng.generate(text, temperature=1.0, top_k=0, top_p=0.95)
e.record(); torch.cuda.synchronize()
ms = s.elapsed_time(e)
out = {"batch": B, "engine_ms_per_1024_tokens": ms, "engine_tokens_per_s": B * cfg.image_seq_len / ms * 1e3,
       "engine_images_per_s": B / ms * 1e3, "ms_per_position": ms / cfg.seq_len}
# eager cached decode: time 40 image positions after the prompt
cache = G.DecodeCache(cfg, B, text.device, torch.bfloat16)
from dalle_b200.models import reference as R
tokens, _ = R.embed(model, text, None, torch.bfloat16)
with torch.no_grad():
    for t in range(24):
        h = G.decode_step_reference(model, tokens[:, t], cache)
    torch.cuda.synchronize(); t0 = time.time()
    for t in range(24, 64):
        h = G.decode_step_reference(model, tokens[:, t], cache)
        lg = G.logits_from_hidden(model, h, True)
        nxt = G.sample_from_logits(lg, 1.0, 0, 0.95)
    torch.cuda.synchronize()
ms_e = (time.time() - t0) / 40 * 1e3
out.update({"eager_ms_per_position": ms_e, "eager_tokens_per_s": B / ms_e * 1e3, "speedup_vs_eager": ms_e / (ms / cfg.seq_len)})
print(json.dumps(out))
