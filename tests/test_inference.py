"""Inference entry point, VQGAN decoder and CLIP scorer (restated architectures) on tiny shapes."""
import os
import pickle
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "inference"))


def test_vqgan_decoder_shapes_and_keys():
    from dalle_b200.models.vqgan import VQGanVAE

    vae = VQGanVAE(n_embed=64, embed_dim=32, ddconfig=dict(ch=32, out_ch=3, ch_mult=(1, 2), num_res_blocks=1,
                                                            attn_resolutions=(4,), resolution=8, z_channels=32))
    keys = set(vae.state_dict())
    assert {"quantize.embed.weight", "post_quant_conv.weight", "decoder.conv_in.weight",
            "decoder.mid.attn_1.q.weight", "decoder.up.1.upsample.conv.weight", "decoder.norm_out.weight"} <= keys
    img = vae.decode(torch.randint(0, 64, (2, 16)))
    assert img.shape == (2, 3, 8, 8) and 0 <= img.min() and img.max() <= 1
    assert vae.num_layers == 1 and vae.num_tokens == 64


def test_clip_restated_forward():
    from dalle_b200.models import clip

    m = clip.CLIP(embed_dim=32, image_resolution=64, vision_layers=1, vision_width=64, vision_patch_size=32,
                  context_length=8, vocab_size=100, transformer_width=32, transformer_heads=2, transformer_layers=1)
    imgs = clip.preprocess(torch.rand(3, 40, 50, 3), 64)
    text = torch.randint(1, 99, (1, 8))
    li, lt = m(imgs, text)
    assert li.shape == (3, 1) and lt.shape == (1, 3)
    assert {"visual.conv1.weight", "visual.transformer.resblocks.0.attn.in_proj_weight", "text_projection",
            "logit_scale"} <= set(m.state_dict())


def test_run_inference_writes_reference_pickle(tmp_path):
    import run_inference
    from dalle_b200.config import get_config
    from dalle_b200.models.dalle import DALLE, ModelWrapper

    q = tmp_path / "queries.txt"
    q.write_text("a cat\n\na dog on a mat\n")
    ckpt = tmp_path / "model.pt"
    torch.save(ModelWrapper(DALLE(get_config("tiny"), seed=0)).state_dict(), ckpt)
    out = tmp_path / "out"
    run_inference.main(["--queries", str(q), "--temperature", "1.0", "--top-k", "8", "--model", str(ckpt),
                        "--output-dir", str(out), "--model-preset", "tiny", "--batch-size", "2", "--n-iters", "1"])
    rec = pickle.load(open(out / "a cat.pickle", "rb"))
    assert rec["query"] == "a cat" and rec["temperature"] == 1.0 and len(rec["codes"]) == 2
    assert rec["codes"][0].shape == (256,) and (out / "a dog on a mat.pickle").exists()
