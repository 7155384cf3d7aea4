"""GPU decode engine (CUDA-graph replay of sm_100a kernels) vs the PyTorch cached-decode oracle."""
import pytest
import torch

from dalle_b200.config import DalleConfig
from dalle_b200.models.dalle import DALLE
from dalle_b200.models import generate as G

pytestmark = pytest.mark.gpu


def _cfg():
    return DalleConfig(num_text_tokens=500, text_seq_len=31, dim=256, depth=4, heads=4, dim_head=64,
                       image_fmap_size=16, num_image_tokens=512,
                       attn_types=("axial_row", "axial_col", "axial_row", "conv_like"),
                       shared_attn_ids=(0, 1, 0, "w_conv"), shared_ff_ids=(0, 1, 0, "w_conv"))


def test_greedy_decode_matches_oracle():
    from dalle_b200.models.decode_engine import DecodeEngine

    cfg = _cfg()
    torch.manual_seed(0)   # the prompt is random: an unseeded draw occasionally lands on a near-tie argmax
    model = DALLE(cfg, seed=0)
    with torch.no_grad():
        for p in list(model.scale_f) + list(model.scale_g):
            p.fill_(0.3)
        model.emb_weight.mul_(4.0)  # sharper logits -> robust argmax under bf16
    model.cuda()
    text = torch.randint(1, cfg.num_text_tokens, (3, cfg.text_seq_len), device="cuda")
    eng = DecodeEngine(model, 3)
    codes = eng.generate(text, temperature=1.0, top_k=1)                   # greedy
    ref, ref_logits = G.generate_codes_reference(model, text, top_k=1, use_cache=True, dtype=torch.float32,
                                                 return_logits=True)
    agree = (codes == ref).float().mean().item()
    # the first divergence changes everything after it; demand the first 64 tokens and most of the rest
    assert torch.equal(codes[:, :32], ref[:, :32]), (codes[:, :32], ref[:, :32])
    assert agree > 0.5, agree
    codes2 = eng.generate(text, temperature=1.0, top_k=1)
    assert torch.equal(codes, codes2)                                        # replay is deterministic


def test_sampler_distribution():
    from dalle_b200.ops import lib
    import ctypes

    n, B, trials = 512, 1, 4000
    logits = torch.full((trials, 512), -30.0, device="cuda", dtype=torch.bfloat16)
    logits[:, 7], logits[:, 100], logits[:, 300] = 2.0, 1.0, 0.0
    ids = torch.zeros(trials, 8, dtype=torch.int64, device="cuda")
    codes = torch.zeros(trials, 8, dtype=torch.int64, device="cuda")
    pos = torch.zeros(1, dtype=torch.int32, device="cuda")
    L = lib.cuda_lib()
    seed = torch.tensor([1234], dtype=torch.int64, device="cuda")
    rc = L.b200_decode_sample(logits.data_ptr(), 512, n, 1.0, 0, 1.0, seed.data_ptr(), ids.data_ptr(), 8, 1000, codes.data_ptr(),
                              8, pos.data_ptr(), 1, trials, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    tok = codes[:, 0]
    assert set(tok.unique().tolist()) <= {7, 100, 300}
    p = torch.tensor([2.0, 1.0, 0.0]).softmax(0)
    freq = torch.tensor([(tok == 7).float().mean(), (tok == 100).float().mean(), (tok == 300).float().mean()])
    assert (freq - p).abs().max().item() < 0.04
    assert torch.equal(ids[:, 1], tok + 1000)
    # top-k = 1 is greedy; top-p small keeps only the mode
    rc = L.b200_decode_sample(logits.data_ptr(), 512, n, 1.0, 0, 0.3, seed.data_ptr(), ids.data_ptr(), 8, 0, codes.data_ptr(), 8,
                              pos.data_ptr(), 1, trials, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert (codes[:, 0] == 7).all()


@pytest.mark.parametrize("temperature", [0.5, 1.0, 2.0])
@pytest.mark.parametrize("top_p", [0.5, 0.9, 1.0])
def test_sampler_filters_then_applies_temperature(temperature, top_p):
    """The kept set comes from the raw logits (temperature 1), the draw from the kept set at `temperature`: the order
    of models/generate.py:sample_from_logits and of dalle-pytorch's generate_images."""
    from dalle_b200.ops import lib

    n, trials = 512, 6000
    g = torch.Generator().manual_seed(3)
    # 512 DISTINCT values exactly representable in bf16 (ties at the nucleus edge would make the kept set depend on the
    # sort's tie-breaking, which differs between torch.sort and the kernel's index order)
    row = ((torch.randperm(n, generator=g) - n // 2).float() / 64.0).to(torch.bfloat16)
    logits = row.cuda().repeat(trials, 1).contiguous()
    want = (G.filter_logits(row.float()[None], 0, top_p) / temperature).softmax(-1)[0]
    ids = torch.zeros(trials, 8, dtype=torch.int64, device="cuda")
    codes = torch.zeros(trials, 8, dtype=torch.int64, device="cuda")
    pos = torch.zeros(1, dtype=torch.int32, device="cuda")
    seed = torch.tensor([99], dtype=torch.int64, device="cuda")
    rc = lib.cuda_lib().b200_decode_sample(logits.data_ptr(), n, n, temperature, 0, top_p, seed.data_ptr(), ids.data_ptr(), 8, 0,
                                           codes.data_ptr(), 8, pos.data_ptr(), 1, trials,
                                           torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    freq = torch.bincount(codes[:, 0].cpu(), minlength=n).float() / trials
    assert freq[want == 0].sum().item() == 0.0, "drew a token outside the nucleus of the raw logits"
    assert (freq - want).abs().max().item() < 0.03


# ---- second-generation decode path (ops/csrc/decode2.cu) -------------------------------------------------------------
def _ttt_like():
    return DalleConfig(num_text_tokens=500, text_seq_len=31, dim=1024, depth=1, heads=16, dim_head=64,
                       image_fmap_size=16, num_image_tokens=512, attn_types=("axial_row",), shared_attn_ids=(0,),
                       shared_ff_ids=(0,))


def _skinny(kind, W, N, K, **kw):
    """One b200_decode2_gemm call with the keyword arguments of DecodeEngine._gemm, on bare tensors."""
    from dalle_b200.ops import lib

    L = lib.cuda_lib()
    ptr = lambda t: None if t is None else t.data_ptr()
    g = lambda k: kw.get(k)
    rc = L.b200_decode2_gemm(kind, W.data_ptr(), N, K, ptr(g("xin")), ptr(g("x")), ptr(g("stats")), ptr(g("ln_w")),
                             ptr(g("ln_b")), ptr(g("ring")), g("pos").data_ptr(), g("T") or 32, g("F") or 16, 1e-5,
                             ptr(g("bias")), ptr(g("scale")), ptr(g("xout")), ptr(g("stats_out")), ptr(g("out")),
                             g("ldo") or 0, ptr(g("rot")), g("q_cols") or 0, g("q_scale") or 1.0, g("nb") or 32,
                             torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
    torch.cuda.synchronize()


@pytest.mark.parametrize("nb", [32, 64])
@pytest.mark.parametrize("N,K", [(1024, 1024), (1024, 4096), (256, 256), (8192, 1024)])
def test_skinny_gemm_residual_and_logits(N, K, nb):
    """TMA-operand forms at the production shapes (out-proj K=1024 / 8 splits of 128, FF2 K=4096 / 8 splits of 512, the
    head N=8192) against fp32 matmul; the residual form also leaves the row statistics of its result behind."""
    g = torch.Generator(device="cuda").manual_seed(N + K)
    W = (torch.randn(N, K, device="cuda", generator=g) * K ** -0.5).to(torch.bfloat16)
    xin = torch.zeros(nb, K, device="cuda", dtype=torch.bfloat16)
    xin[:nb - 3] = torch.randn(nb - 3, K, device="cuda", generator=g).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda", generator=g)
    scale = torch.rand(N, device="cuda", generator=g) + 0.5
    pos = torch.zeros(1, dtype=torch.int32, device="cuda")
    ref = xin.float() @ W.float().t() + bias
    x0 = torch.randn(nb, N, device="cuda", generator=g)
    xout = x0.clone()
    stats = torch.zeros(nb, 2, device="cuda")
    _skinny(1, W, N, K, xin=xin, bias=bias, scale=scale, xout=xout, stats_out=stats, pos=pos, nb=nb)
    want = x0 + scale * ref
    assert (xout - want).abs().max().item() < 2e-2
    assert torch.allclose(stats[:, 0], xout.sum(1), rtol=1e-4, atol=1e-2)
    assert torch.allclose(stats[:, 1], (xout * xout).sum(1), rtol=1e-4, atol=1e-2)
    out = torch.zeros(nb, N, device="cuda", dtype=torch.bfloat16)
    _skinny(3, W, N, K, xin=xin, bias=bias, out=out, ldo=N, pos=pos, nb=nb)
    assert (out.float() - ref).abs().max().item() < 3e-2 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("nb", [32, 64])
@pytest.mark.parametrize("D", [1024, 256])
@pytest.mark.parametrize("pos_kind", ["text", "image_first_col", "image_inner"])
def test_skinny_gemm_layernorm_prologue(D, pos_kind, nb):
    """QKV (rotary epilogue) and GEGLU forms: layer norm from the (sum, sum of squares) left by the producer, token
    shift from the 64-slot ring, against the same statement in fp32 PyTorch."""
    from dalle_b200.models.rotary import rotary_pair_table

    cfg = DalleConfig(num_text_tokens=500, text_seq_len=31, dim=D, depth=1, heads=D // 64, dim_head=64,
                      image_fmap_size=16, num_image_tokens=512, attn_types=("axial_row",), shared_attn_ids=(0,),
                      shared_ff_ids=(0,))
    T, F = cfg.text_len, cfg.image_fmap_size
    p = {"text": 9, "image_first_col": T + 3 * F, "image_inner": T + 3 * F + 5}[pos_kind]
    g = torch.Generator(device="cuda").manual_seed(D + p)
    x = torch.zeros(nb, D, device="cuda")
    x[:nb - 2] = torch.randn(nb - 2, D, device="cuda", generator=g) * 2 + 0.3
    stats = torch.stack((x.sum(1), (x * x).sum(1)), dim=1).contiguous()
    ln_w, ln_b = torch.rand(D, device="cuda", generator=g) + 0.5, torch.randn(D, device="cuda", generator=g) * 0.1
    ring = (torch.randn(nb, 64, D // 2, device="cuda", generator=g)).to(torch.bfloat16)
    ring0 = ring.clone()
    pos = torch.full((1,), p, dtype=torch.int32, device="cuda")
    # oracle of the prologue
    hn = torch.nn.functional.layer_norm(x, (D,), ln_w, ln_b, 1e-5)
    h = hn.clone()
    if p < T:
        src_a = src_b = p - 1
    else:
        r, c = divmod(p - T, F)
        src_a = p - F if r >= 1 else -1
        src_b = p - 1 if c >= 1 else -1
    h[:, : D // 4] = ring0[:, src_a % 64, : D // 4].float() if src_a >= 0 else 0.0
    h[:, D // 4: D // 2] = ring0[:, src_b % 64, D // 4: D // 2].float() if src_b >= 0 else 0.0
    hb = h.to(torch.bfloat16).float()
    # --- QKV + rotary
    I = D
    Wq = (torch.randn(3 * I, D, device="cuda", generator=g) * D ** -0.5).to(torch.bfloat16)
    rot = rotary_pair_table(cfg).cuda()
    out = torch.zeros(nb, 3 * I, device="cuda", dtype=torch.bfloat16)
    _skinny(0, Wq, 3 * I, D, x=x, stats=stats, ln_w=ln_w, ln_b=ln_b, ring=ring, pos=pos, T=T, F=F, out=out, ldo=3 * I,
            rot=rot, q_cols=I, q_scale=0.125, nb=nb)
    raw = (hb @ Wq.float().t()).view(nb, 3 * I // 2, 2)
    cs = rot[p].repeat(3 * I // 64, 1)                       # [pairs of the row, 2]
    ref = torch.stack((raw[..., 0] * cs[:, 0] - raw[..., 1] * cs[:, 1], raw[..., 1] * cs[:, 0] + raw[..., 0] * cs[:, 1]),
                      dim=-1).view(nb, 3 * I)
    ref[:, :I] *= 0.125
    assert (out.float() - ref).abs().max().item() < 3e-2 * max(1.0, ref.abs().max().item())
    # the token's own normalised first half went into the ring, other slots untouched
    assert (ring[:, p % 64].float() - hn[:, : D // 2]).abs().max().item() < 3e-2
    keep = [s for s in range(64) if s != p % 64]
    assert torch.equal(ring[:, keep], ring0[:, keep])
    # --- GEGLU
    hid = 4 * D
    W1 = (torch.randn(2 * hid, D, device="cuda", generator=g) * D ** -0.5).to(torch.bfloat16)
    b1 = torch.randn(2 * hid, device="cuda", generator=g) * 0.1
    u = torch.zeros(nb, hid, device="cuda", dtype=torch.bfloat16)
    ring.copy_(ring0)
    _skinny(2, W1, hid, D, x=x, stats=stats, ln_w=ln_w, ln_b=ln_b, ring=ring, pos=pos, T=T, F=F, bias=b1, out=u, ldo=hid, nb=nb)
    pre = hb @ W1.float().t() + b1
    ref_u = pre[:, :hid] * torch.nn.functional.gelu(pre[:, hid:])
    assert (u.float() - ref_u).abs().max().item() < 3e-2 * max(1.0, ref_u.abs().max().item())


@pytest.mark.parametrize("use_graph", [False, True])
def test_decode2_prefill_and_stepwise_prompt_agree_with_the_oracle(use_graph):
    from dalle_b200.models.decode_engine import DecodeEngine, DecodeEngineV1

    cfg = _cfg()
    torch.manual_seed(0)
    model = DALLE(cfg, seed=0)
    with torch.no_grad():
        for p in list(model.scale_f) + list(model.scale_g):
            p.fill_(0.3)
        model.emb_weight.mul_(4.0)
    model.cuda()
    text = torch.randint(1, cfg.num_text_tokens, (3, cfg.text_seq_len), device="cuda")
    eng = DecodeEngine(model, 3)
    assert not isinstance(eng, DecodeEngineV1)
    ref = G.generate_codes_reference(model, text, top_k=1, use_cache=True, dtype=torch.float32)
    a = eng.generate(text, temperature=1.0, top_k=1, use_graph=use_graph, prefill=True)
    b = eng.generate(text, temperature=1.0, top_k=1, use_graph=False, prefill=False)
    assert torch.equal(a[:, :32], ref[:, :32]), (a[:, :32], ref[:, :32])
    assert torch.equal(b[:, :32], ref[:, :32]), (b[:, :32], ref[:, :32])
    assert (a == ref).float().mean().item() > 0.5
    assert eng.cache_bytes() < 0.5 * 2 * cfg.depth * 3 * cfg.heads * cfg.seq_len * 64 * 2   # under half of dense


def test_first_generation_engine_still_matches():
    from dalle_b200.models.decode_engine import DecodeEngineV1

    cfg = _cfg()
    torch.manual_seed(0)
    model = DALLE(cfg, seed=0)
    with torch.no_grad():
        for p in list(model.scale_f) + list(model.scale_g):
            p.fill_(0.3)
        model.emb_weight.mul_(4.0)
    model.cuda()
    text = torch.randint(1, cfg.num_text_tokens, (3, cfg.text_seq_len), device="cuda")
    ref = G.generate_codes_reference(model, text, top_k=1, use_cache=True, dtype=torch.float32)
    codes = DecodeEngineV1(model, 3).generate(text, temperature=1.0, top_k=1)
    assert torch.equal(codes[:, :32], ref[:, :32])


def test_run_inference_on_the_gpu_goes_through_the_decode_engine(tmp_path):
    """The public entry point (reference inference/run_inference.py:79-92) on CUDA: greedy codes written to the pickle
    equal the PyTorch cached-decode oracle on the same checkpoint."""
    import os
    import pickle
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "inference"))
    import run_inference
    from dalle_b200.config import get_config
    from dalle_b200.models.dalle import ModelWrapper
    from dalle_b200.models import decode_engine

    cfg = get_config("tiny")
    model = DALLE(cfg, seed=0)
    with torch.no_grad():
        for p in list(model.scale_f) + list(model.scale_g):
            p.fill_(0.3)
        model.emb_weight.mul_(4.0)
    ckpt = tmp_path / "model.pt"
    torch.save(ModelWrapper(model).state_dict(), ckpt)
    q = tmp_path / "queries.txt"
    q.write_text("a cat\n")
    out = tmp_path / "out"
    made = []
    orig = decode_engine.DecodeEngine.__init__
    decode_engine.DecodeEngine.__init__ = lambda self, *a, **k: (made.append(1), orig(self, *a, **k))[1]
    try:
        run_inference.main(["--queries", str(q), "--temperature", "1.0", "--top-k", "1", "--model", str(ckpt),
                            "--output-dir", str(out), "--model-preset", "tiny", "--batch-size", "2", "--n-iters", "1"])
    finally:
        decode_engine.DecodeEngine.__init__ = orig
    assert made, "the CUDA path did not build a decode engine"
    rec = pickle.load(open(out / "a cat.pickle", "rb"))
    codes = torch.as_tensor(rec["codes"][0])
    assert codes.shape == (cfg.image_seq_len,) and int(codes.min()) >= 0 and int(codes.max()) < cfg.num_image_tokens
    # the oracle on the same prompt and weights
    from dalle_b200.training.tokenizer import load_tokenizer

    text = run_inference.encode_query("a cat", load_tokenizer("t5-small"), cfg.text_seq_len, 2, "cuda")
    ref = G.generate_codes_reference(model.cuda(), text, top_k=1, use_cache=True, dtype=torch.float32)
    assert torch.equal(codes[:32].cuda(), ref[0, :32]), (codes[:32], ref[0, :32])


def test_batches_above_64_run_as_concurrent_engines():
    """generate_images with 70 prompts: slices of 64 + 6 on two engines / streams, interleaved position by position;
    greedy codes of every prompt equal the oracle's (the slices must not disturb each other's caches or position)."""
    cfg = _cfg()
    torch.manual_seed(0)
    model = DALLE(cfg, seed=0)
    with torch.no_grad():
        for p in list(model.scale_f) + list(model.scale_g):
            p.fill_(0.3)
        model.emb_weight.mul_(4.0)
    model.cuda()
    text = torch.randint(1, cfg.num_text_tokens, (70, cfg.text_seq_len), device="cuda")
    codes = G.generate_images(model, text, top_k=1, return_codes=True)
    assert codes.shape == (70, cfg.image_seq_len) and len(model._decode_engines) == 2
    ref = G.generate_codes_reference(model, text, top_k=1, use_cache=True, dtype=torch.float32)
    assert torch.equal(codes[:, :24], ref[:, :24]), (codes[:, :24] != ref[:, :24]).nonzero()[:5]
    assert (codes == ref).float().mean().item() > 0.5


def test_run_inference_iterations_share_one_concurrent_call(tmp_path):
    """--batch-size 20 --n-iters 2 on CUDA: the 40 prompts go through one generate_images call (two concurrent
    engines); greedy sampling of the same query must give the same codes in every row of both iterations."""
    import os
    import pickle
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "inference"))
    import run_inference
    from dalle_b200.config import get_config
    from dalle_b200.models.dalle import ModelWrapper

    cfg = get_config("tiny")
    model = DALLE(cfg, seed=0)
    with torch.no_grad():
        for p in list(model.scale_f) + list(model.scale_g):
            p.fill_(0.3)
        model.emb_weight.mul_(4.0)
    ckpt = tmp_path / "model.pt"
    torch.save(ModelWrapper(model).state_dict(), ckpt)
    q = tmp_path / "queries.txt"
    q.write_text("a dog\n")
    out = tmp_path / "out"
    run_inference.main(["--queries", str(q), "--temperature", "1.0", "--top-k", "1", "--model", str(ckpt),
                        "--output-dir", str(out), "--model-preset", "tiny", "--batch-size", "20", "--n-iters", "2"])
    rec = pickle.load(open(out / "a dog.pickle", "rb"))
    codes = torch.as_tensor(__import__("numpy").stack(rec["codes"]))
    assert codes.shape == (40, cfg.image_seq_len)
    assert (codes[:, :32] == codes[0, :32]).all(), "rows of one greedy query differ"
