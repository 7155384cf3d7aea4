"""GPU diagnostic: fwd / fwd+bwd time of the TTT model through the fused executor."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import get_config
from dalle_b200.models.dalle import DALLE
from dalle_b200.ops import lib

def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    depth = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    cfg = get_config("ttt", depth=depth) if depth != 64 else get_config("ttt")
    model = DALLE(cfg, seed=0).cuda()
    text = torch.randint(1, cfg.num_text_tokens, (B, cfg.text_seq_len), device="cuda")
    image = torch.randint(0, cfg.num_image_tokens, (B, cfg.image_seq_len), device="cuda")
    for it in range(3):
        torch.cuda.synchronize(); t0 = time.time(); n0 = lib.launch_count
        s, e1, e2 = torch.cuda.Event(True), torch.cuda.Event(True), torch.cuda.Event(True)
        s.record()
        loss = model(text, image, return_loss=True)
        e1.record()
        loss.backward()
        e2.record(); t_enq = time.time() - t0; torch.cuda.synchronize()
        print(f"B={B} depth={depth} it={it} loss={loss.item():.4f} fwd {s.elapsed_time(e1):.1f} ms bwd {e1.elapsed_time(e2):.1f} ms "
              f"total {s.elapsed_time(e2):.1f} ms -> {B / s.elapsed_time(e2) * 1e3:.1f} samples/s; wall {time.time()-t0:.3f}s (host enqueue {t_enq:.3f}s); launches {lib.launch_count-n0}; "
              f"mem {torch.cuda.max_memory_allocated()/2**30:.1f} GiB", flush=True)
    g = model.emb_weight.grad
    print("grad finite:", all(torch.isfinite(p.grad).all().item() for p in model.parameters()), "emb grad norm", g.norm().item())

if __name__ == "__main__":
    main()
