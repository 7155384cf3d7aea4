"""Short program for ncu: two training steps of a depth-4 TTT-shaped model (row, col, row, conv) at micro-batch B."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import get_config
from dalle_b200.trainer import GroupTrainer

def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    cfg = get_config("ttt", depth=4, attn_types=("axial_row", "axial_col", "axial_row", "conv_like"),
                     shared_attn_ids=(0, 1, 0, "w_conv"), shared_ff_ids=(0, 1, 0, "w_conv"))
    tr = GroupTrainer(cfg, torch.device("cuda", 0), averaging="nccl")
    text = torch.randint(1, cfg.num_text_tokens, (B, cfg.text_seq_len), device="cuda")
    image = torch.randint(0, cfg.num_image_tokens, (B, cfg.image_seq_len), device="cuda")
    for _ in range(steps):
        loss = tr.step(text, image)
    torch.cuda.synchronize()
    print("loss", loss.item())

if __name__ == "__main__":
    main()
