"""Timeline of CTA 0 of the pipelined attention forward (b200_attn3_set_trace): per role, events with SM-clock deltas.
usage: attn3_trace.py [B] [kind]"""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import ttt_config
from dalle_b200.ops import attention as A, lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
kind = sys.argv[2] if len(sys.argv) > 2 else "axial_row"
cfg = ttt_config()
M, I = B * cfg.seq_len, cfg.inner_dim
qkv = torch.zeros(M + 128, 3 * I, device="cuda", dtype=torch.bfloat16); qkv[:M] = (torch.randn(M, 3 * I, device="cuda") * 0.5).bfloat16()
out = torch.zeros(M, I, device="cuda", dtype=torch.bfloat16); lse = torch.zeros(B, cfg.heads, cfg.seq_len, device="cuda")
for _ in range(3): A.attn_fwd(qkv, out, lse, cfg, kind)
torch.cuda.synchronize()
buf = torch.zeros(4 * 2048 * 2, device="cuda", dtype=torch.int64)
fn = lib.cuda_lib().b200_attn3_set_trace
fn.argtypes = [ctypes.c_void_p]; fn.restype = None
fn(buf.data_ptr())
A.attn_fwd(qkv, out, lse, cfg, kind)
torch.cuda.synchronize()
fn(None)
t = buf.cpu().view(4, 2048, 2)
names = {9: "p_ready", 10: "pv_go", 12: "s_go", 13: "s_issued", 1: "S", 2: "PV", 3: "wait_s", 4: "got_s", 5: "p_done", 6: "got_o", 7: "Q", 8: "L"}
ev = []
for role in range(4):
    for tag, clk in t[role].tolist():
        if tag == 0 and clk == 0: break
        ev.append((clk, role, tag))
ev.sort()
t0 = ev[0][0]
limit = int(sys.argv[3]) if len(sys.argv) > 3 else 400
last = {}
for clk, role, tag in ev[:limit]:
    kindc = tag >> 24
    if role == 0: desc = f"{names[kindc]} stage {(tag >> 20) & 15} c{(tag >> 16) & 15} g{tag & 0xFFFF}"
    elif role == 3: desc = f"{names[kindc]} g{tag & 0xFFFF}"
    else: desc = f"{names[kindc]} c{(tag >> 16) & 255} g{tag & 0xFFFF}"
    d = clk - last.get(role, clk); last[role] = clk
    print(f"{clk - t0:9d} (+{d:6d})  {'mma wg0 wg1 load'.split()[role]:5s} {desc}")
print("total events", len(ev), "span", ev[-1][0] - t0)
