"""GPU executor of KV-cached sampling (workload of reference inference/run_inference.py:79-92).

``DecodeEngine`` (second generation, ops/csrc/decode2.cu):

* the 257 prompt positions run ONCE through the training forward kernels (``Engine.run_stack``), which leaves every
  layer's rotated keys and values of the text positions in the cache - no per-position replays for the prompt;
* one decode step = embed -> depth x (QKV GEMM with LN + shift prologue and rotary epilogue, attention over the
  sparsity-aware cache, out-proj GEMM + residual, GEGLU GEMM with LN + shift prologue, FF2 GEMM + residual) -> final
  norm -> image-segment logits GEMM -> top-k / top-p sampler: 5 launches per layer, all with programmatic dependent
  launch, position read from device memory, captured once as a CUDA graph and replayed for the 1023 image positions;
* K/V cache per layer: text keys + one image row (row layers), + conv_half + 1 rows (convolution-like), + the
  transposed image (column layers) instead of a dense [B, H, seq, 64] block; token-shift history in a 64-slot ring.

``DecodeEngineV1`` is the first generation (decode.cu + the training GEMM at M = 32, dense cache, prompt decoded
position by position); it serves batches above 32 and dense-causal layers, and ``DALLE_B200_DECODE=1`` selects it."""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

from ..config import DalleConfig
from ..ops import gemm as G
from ..ops import lib
from .masks import ATTN_TYPE_CODES


class DecodeEngineV1:
    def __init__(self, model, batch: int):
        cfg: DalleConfig = model.cfg
        assert cfg.dim_head == 64 and cfg.conv_dilation == 1
        self.model, self.cfg, self.batch = model, cfg, batch
        dev = model.emb_weight.device
        B, D, I, n = batch, cfg.dim, cfg.inner_dim, cfg.seq_len
        hid = D * cfg.ff_mult
        bf = torch.bfloat16
        z = lambda *s, dtype=bf: torch.zeros(*s, device=dev, dtype=dtype)
        self.x1, self.x2 = z(B, D, dtype=torch.float32), z(B, D, dtype=torch.float32)
        self.h, self.qkv, self.ao, self.u, self.hf = z(B, D), z(B, 3 * I), z(B, I), z(B, hid), z(B, D)
        self.kc = [z(B, cfg.heads, n, 64) for _ in range(cfg.depth)]
        self.vc = [z(B, cfg.heads, n, 64) for _ in range(cfg.depth)]
        self.shift = [[z(B, n, D // 2) for _ in range(2)] for _ in range(cfg.depth)]
        self.ld = (cfg.num_image_tokens + 63) // 64 * 64
        self.logits = z(B, self.ld)
        self.ids = z(B, n + 1, dtype=torch.int64)
        self.codes = z(B, cfg.image_seq_len, dtype=torch.int64)
        self.pos = z(1, dtype=torch.int32)
        self.seed = z(1, dtype=torch.int64)
        self.dummy_mean, self.dummy = z(B, dtype=torch.float32), None
        self._graphs = {}
        self._sampling = None
        from .executor import Engine

        self._w = model.engine.w if model._engine is not None else Engine(model).w

    # ---- one decode step (enqueued on the current stream) ------------------------------------------
    def _step(self, with_head: bool, temperature: float, top_k: int, top_p: float):
        m, cfg, L = self.model, self.cfg, lib.cuda_lib()
        st = torch.cuda.current_stream().cuda_stream
        B, D, I = self.batch, cfg.dim, cfg.inner_dim
        pp = self.pos.data_ptr()
        ck = lambda rc, what: lib.check(rc, what)
        ck(L.b200_decode_embed(self.ids.data_ptr(), self.ids.shape[1], m.emb_weight.data_ptr(), self.x1.data_ptr(),
                               self.x2.data_ptr(), pp, B, D, st), "decode_embed")
        for i in range(cfg.depth):
            a, f = m.attn_index[i], m.ff_index[i]
            ck(L.b200_decode_ln_shift(self.x2.data_ptr(), None, m.norm_f_w[i].data_ptr(), m.norm_f_b[i].data_ptr(),
                                      self.shift[i][0].data_ptr(), self.h.data_ptr(), pp, B, D, cfg.seq_len,
                                      cfg.text_len, cfg.image_fmap_size, cfg.ln_eps, 1, st), "decode_ln_shift")
            G.gemm(G.NT, G.EPI_ROTARY, self.h, self._w(m.attn_qkv_w[a]), self.qkv, rot=m.rotary_pairs,
                   seq_len=cfg.seq_len, q_cols=I, q_scale=cfg.dim_head ** -0.5, pos_ptr=self.pos)
            ck(L.b200_decode_attn(self.qkv.data_ptr(), self.kc[i].data_ptr(), self.vc[i].data_ptr(),
                                  self.ao.data_ptr(), pp, B, cfg.heads, cfg.seq_len, cfg.text_len,
                                  cfg.image_fmap_size, ATTN_TYPE_CODES[cfg.attn_types[i]] + 1, cfg.conv_kernel_size,
                                  st), "decode_attn")
            G.gemm(G.NT, G.EPI_RESID, self.ao, self._w(m.attn_out_w[a]), self.x1, bias=m.attn_out_b[a],
                   scale=m.scale_f[i].view(-1), res=self.x1, alpha=1.0)
            ck(L.b200_decode_ln_shift(self.x1.data_ptr(), None, m.norm_g_w[i].data_ptr(), m.norm_g_b[i].data_ptr(),
                                      self.shift[i][1].data_ptr(), self.h.data_ptr(), pp, B, D, cfg.seq_len,
                                      cfg.text_len, cfg.image_fmap_size, cfg.ln_eps, 1, st), "decode_ln_shift")
            G.gemm(G.NT, G.EPI_GEGLU, self.h, self._w(m.ff_w1[f]), self.u, bias=m.ff_b1[f], n_half=D * cfg.ff_mult)
            G.gemm(G.NT, G.EPI_RESID, self.u, self._w(m.ff_w2[f]), self.x2, bias=m.ff_b2[f],
                   scale=m.scale_g[i].view(-1), res=self.x2, alpha=1.0)
        if with_head:
            ck(L.b200_decode_ln_shift(self.x1.data_ptr(), self.x2.data_ptr(), m.final_norm_w.data_ptr(),
                                      m.final_norm_b.data_ptr(), None, self.hf.data_ptr(), pp, B, D, cfg.seq_len,
                                      cfg.text_len, cfg.image_fmap_size, cfg.ln_eps, 0, st), "decode_final_norm")
            v0, nv = cfg.total_text_tokens, cfg.num_image_tokens
            G.gemm(G.NT, G.EPI_BF16, self.hf, self._w(m.emb_weight)[v0:v0 + nv], self.logits[:, :nv],
                   bias=m.emb_bias.detach()[v0:v0 + nv])
            ck(L.b200_decode_sample(self.logits.data_ptr(), self.ld, nv, temperature, top_k, top_p, self.seed.data_ptr(),
                                    self.ids.data_ptr(), self.ids.shape[1], v0, self.codes.data_ptr(),
                                    self.codes.shape[1], pp, cfg.text_len, B, st), "decode_sample")
        ck(L.b200_decode_advance(pp, st), "decode_advance")

    def _graph(self, with_head: bool, temperature, top_k, top_p):
        key = (with_head, float(temperature), int(top_k), float(top_p))
        g = self._graphs.get(key)
        if g is None:
            # warm-up outside capture (function attributes, lazy module loading), then capture.  Both run one
            # real decode step, so the position must be a valid one.
            self.pos.zero_()
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                self._step(*key[:1], *key[1:])
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            self.pos.zero_()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._step(*key[:1], *key[1:])
            self._graphs[key] = g
        return g

    @torch.no_grad()
    def generate(self, text: torch.Tensor, *, temperature=1.0, top_k=0, top_p=1.0, generator=None,
                 use_graph: bool = True) -> torch.Tensor:
        cfg = self.cfg
        assert text.shape[0] == self.batch
        text = text[:, : cfg.text_seq_len]
        pad_ids = torch.arange(cfg.text_seq_len, device=text.device) + cfg.num_text_tokens
        text = torch.where(text == 0, pad_ids, text)
        seed = int(generator.initial_seed()) if generator is not None else int(torch.seed() & 0x7FFFFFFFFFFFFFFF)
        T = cfg.text_len
        if use_graph:
            g_body = self._graph(False, 1.0, 0, 1.0)
            g_full = self._graph(True, temperature, top_k, top_p)
        # (graph construction ran warm-up steps: reset the state afterwards)
        self.ids.zero_()
        self.ids[:, 1:T] = text
        self.pos.zero_()
        self.codes.zero_()
        self.seed.fill_(seed)
        for pos in range(cfg.seq_len):
            head = pos >= T - 1
            if use_graph:
                (g_full if head else g_body).replay()
            else:
                self._step(head, temperature, top_k, top_p)
        return self.codes.clone()



class DecodeEngine:
    """See the module docstring.  ``DecodeEngine(model, batch)`` returns a first-generation engine when the
    configuration is outside what decode2.cu tiles (batch > 64, dense-causal layers, long local windows)."""

    MAX_BATCH = 64

    def __new__(cls, model, batch: int):
        cfg = model.cfg
        local = max(cfg.image_fmap_size, (cfg.conv_kernel_size // 2 + 1) * (2 * (cfg.conv_kernel_size // 2) + 1))
        ok = (batch <= cls.MAX_BATCH and "full" not in cfg.attn_types and cfg.text_len + local + 1 <= 304
              and cfg.dim % 128 == 0 and cfg.dim <= 4096 and os.environ.get("DALLE_B200_DECODE", "2") != "1")
        return super().__new__(cls) if ok else DecodeEngineV1(model, batch)

    def __init__(self, model, batch: int):
        cfg: DalleConfig = model.cfg
        assert cfg.dim_head == 64 and cfg.conv_dilation == 1
        self.model, self.cfg, self.batch = model, cfg, batch
        dev = model.emb_weight.device
        R = self.rows = 32 if batch <= 32 else 64     # batch tile of the skinny GEMMs: every activation buffer has R rows
        D, I, n = cfg.dim, cfg.inner_dim, cfg.seq_len
        hid = D * cfg.ff_mult
        bf = torch.bfloat16
        z = lambda *s, dtype=bf: torch.zeros(*s, device=dev, dtype=dtype)
        self.x1, self.x2 = z(R, D, dtype=torch.float32), z(R, D, dtype=torch.float32)
        self.qkv, self.ao, self.u, self.hf = z(R, 3 * I), z(R, I), z(R, hid), z(R, D)
        self.n_sites = 2 * cfg.depth + 1
        self.stats = z(self.n_sites, R, 2, dtype=torch.float32)
        self.ring = [[z(R, 64, D // 2) for _ in range(2)] for _ in range(cfg.depth)]
        T, F, half = cfg.text_len, cfg.image_fmap_size, cfg.conv_kernel_size // 2
        local = {"axial_row": F, "axial_col": F * F, "conv_like": (half + 1) * F}
        self.slots = [T + local[t] for t in cfg.attn_types]
        self.kc = [z(batch, cfg.heads, s, 64) for s in self.slots]
        self.vc = [z(batch, cfg.heads, s, 64) for s in self.slots]
        self.ld = (cfg.num_image_tokens + 63) // 64 * 64
        self.logits = z(R, self.ld)
        self.ids = z(batch, n + 1, dtype=torch.int64)
        self.codes = z(batch, cfg.image_seq_len, dtype=torch.int64)
        self.pos = z(1, dtype=torch.int32)
        self.seed = z(1, dtype=torch.int64)
        self._graphs = {}
        self._w = model.engine.w
        L = lib.cuda_lib()
        L.b200_decode2_set_pdl(0 if os.environ.get("DALLE_B200_DECODE_PDL", "1") == "0" else 1)

    def cache_bytes(self) -> int:
        """K/V cache only (the token-shift ring adds 32 x 64 x dim/2 bf16 per layer norm)."""
        return sum(t.numel() * 2 for t in self.kc + self.vc)

    # ---- one decode step (enqueued on the current stream) ------------------------------------------
    def _gemm(self, kind, W, N, K, *, xin=None, x=None, site=None, ln=None, ring=None, bias=None, scale=None,
              xout=None, site_out=None, out=None, ldo=0, rot=None, q_cols=0, q_scale=1.0):
        cfg, L = self.cfg, lib.cuda_lib()
        ptr = lambda t: None if t is None else t.data_ptr()
        lib.check(L.b200_decode2_gemm(
            kind, W.data_ptr(), N, K, ptr(xin), ptr(x), None if site is None else self.stats[site].data_ptr(),
            None if ln is None else ln[0].data_ptr(), None if ln is None else ln[1].data_ptr(), ptr(ring),
            self.pos.data_ptr(), cfg.text_len, cfg.image_fmap_size, cfg.ln_eps, ptr(bias), ptr(scale), ptr(xout),
            None if site_out is None else self.stats[site_out].data_ptr(), ptr(out), ldo, ptr(rot), q_cols, q_scale,
            self.rows, torch.cuda.current_stream().cuda_stream), "decode2_gemm")

    def _layers(self):
        m, cfg, L = self.model, self.cfg, lib.cuda_lib()
        st = torch.cuda.current_stream().cuda_stream
        D, I, hid = cfg.dim, cfg.inner_dim, cfg.dim * cfg.ff_mult
        for i in range(cfg.depth):
            a, f = m.attn_index[i], m.ff_index[i]
            self._gemm(0, self._w(m.attn_qkv_w[a]), 3 * I, D, x=self.x2, site=2 * i, ln=(m.norm_f_w[i], m.norm_f_b[i]),
                       ring=self.ring[i][0], out=self.qkv, ldo=3 * I, rot=m.rotary_pairs, q_cols=I,
                       q_scale=cfg.dim_head ** -0.5)
            lib.check(L.b200_decode2_attn(self.qkv.data_ptr(), self.kc[i].data_ptr(), self.vc[i].data_ptr(),
                                          self.ao.data_ptr(), self.pos.data_ptr(), self.batch, cfg.heads, cfg.text_len,
                                          cfg.image_fmap_size, ATTN_TYPE_CODES[cfg.attn_types[i]] + 1,
                                          cfg.conv_kernel_size, self.slots[i], st), "decode2_attn")
            self._gemm(1, self._w(m.attn_out_w[a]), D, I, xin=self.ao, bias=m.attn_out_b[a], scale=m.scale_f[i].view(-1),
                       xout=self.x1, site_out=2 * i + 1)
            self._gemm(2, self._w(m.ff_w1[f]), hid, D, x=self.x1, site=2 * i + 1, ln=(m.norm_g_w[i], m.norm_g_b[i]),
                       ring=self.ring[i][1], bias=m.ff_b1[f], out=self.u, ldo=hid)
            self._gemm(1, self._w(m.ff_w2[f]), D, hid, xin=self.u, bias=m.ff_b2[f], scale=m.scale_g[i].view(-1),
                       xout=self.x2, site_out=2 * i + 2)

    def _head(self, temperature: float, top_k: int, top_p: float):
        m, cfg, L = self.model, self.cfg, lib.cuda_lib()
        st = torch.cuda.current_stream().cuda_stream
        pp = self.pos.data_ptr()
        lib.check(L.b200_decode_ln_shift(self.x1.data_ptr(), self.x2.data_ptr(), m.final_norm_w.data_ptr(),
                                         m.final_norm_b.data_ptr(), None, self.hf.data_ptr(), pp, self.rows, cfg.dim,
                                         cfg.seq_len, cfg.text_len, cfg.image_fmap_size, cfg.ln_eps, 0, st),
                  "decode_final_norm")
        v0, nv = cfg.total_text_tokens, cfg.num_image_tokens
        self._gemm(3, self._w(m.emb_weight)[v0:v0 + nv], nv, cfg.dim, xin=self.hf, bias=m.emb_bias.detach()[v0:v0 + nv],
                   out=self.logits, ldo=self.ld)
        lib.check(L.b200_decode_sample(self.logits.data_ptr(), self.ld, nv, temperature, top_k, top_p,
                                       self.seed.data_ptr(), self.ids.data_ptr(), self.ids.shape[1], v0,
                                       self.codes.data_ptr(), self.codes.shape[1], pp, cfg.text_len, self.batch, st),
                  "decode_sample")
        lib.check(L.b200_decode_advance(pp, st), "decode_advance")

    def _step(self, with_head: bool, temperature: float, top_k: int, top_p: float):
        m, cfg, L = self.model, self.cfg, lib.cuda_lib()
        st = torch.cuda.current_stream().cuda_stream
        lib.check(L.b200_decode2_embed(self.ids.data_ptr(), self.ids.shape[1], m.emb_weight.data_ptr(),
                                       self.x1.data_ptr(), self.x2.data_ptr(), self.stats.data_ptr(), self.n_sites,
                                       self.pos.data_ptr(), self.batch, cfg.dim, self.rows, st), "decode2_embed")
        self._layers()
        if with_head:
            self._head(temperature, top_k, top_p)
        else:
            lib.check(L.b200_decode_advance(self.pos.data_ptr(), st), "decode_advance")

    def _graph(self, kind: str, temperature, top_k, top_p):
        """kind: "step" (one whole position with the sampler) or "head" (final norm + logits + sampler only)."""
        key = (kind, float(temperature), int(top_k), float(top_p))
        g = self._graphs.get(key)
        if g is None:
            run = (lambda: self._step(True, *key[1:])) if kind == "step" else (lambda: self._head(*key[1:]))
            self.pos.fill_(self.cfg.text_len)       # warm-up + capture run one real step: a valid image position
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                run()
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            self.pos.fill_(self.cfg.text_len)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                run()
            self._graphs[key] = g
        return g

    # ---- prompt: one pass of the training forward kernels --------------------------------------------
    def _prefill(self, text: torch.Tensor):
        """Runs positions 0..T-1 (<bos> + text) through ``Engine.run_stack`` (the image positions of that pass are
        padding: attention is causal, nothing flows from them to the text), copies every layer's text keys / values
        into the cache and leaves the two residual streams of position T-1 in x1 / x2."""
        cfg, eng = self.cfg, self.model.engine
        B, T, n = self.batch, cfg.text_len, cfg.seq_len
        ids = torch.full((B, n), cfg.total_text_tokens, device=text.device, dtype=torch.int64)
        ids[:, 0] = 0
        ids[:, 1:T] = text
        b = eng.buffers(B, text.device, False)

        def keep_text_kv(i, qkv):
            rows = qkv[: B * n].view(B, n, 3, cfg.heads, 64)[:, :T]
            self.kc[i][:, :, :T].copy_(rows[:, :, 1].transpose(1, 2))
            self.vc[i][:, :, :T].copy_(rows[:, :, 2].transpose(1, 2))

        eng.run_stack(ids, b, after_attention=keep_text_kv)
        self.x1[:B].copy_(b.x1.view(B, n, cfg.dim)[:, T - 1])
        self.x2[:B].copy_(b.x2.view(B, n, cfg.dim)[:, T - 1])

    @torch.no_grad()
    def begin(self, text: torch.Tensor, *, temperature=1.0, top_k=0, top_p=1.0, generator=None, use_graph: bool = True,
              prefill: bool = True) -> int:
        """First half of ``generate``: prompt pass and first image code; returns the first position still to decode.
        ``step_once`` then advances one position per call and ``result`` returns the codes - split so that several
        engines (streams) can be interleaved by one host loop (``generate_many``)."""
        cfg = self.cfg
        assert text.shape[0] == self.batch
        text = text[:, : cfg.text_seq_len]
        pad_ids = torch.arange(cfg.text_seq_len, device=text.device) + cfg.num_text_tokens
        text = torch.where(text == 0, pad_ids, text)
        seed = int(generator.initial_seed()) if generator is not None else int(torch.seed() & 0x7FFFFFFFFFFFFFFF)
        T = cfg.text_len
        self._sampling = (temperature, top_k, top_p)
        self._g_step = self._g_head = None
        if use_graph:
            self._g_step = self._graph("step", temperature, top_k, top_p)
            self._g_head = self._graph("head", temperature, top_k, top_p)
        # (graph construction ran warm-up steps: reset the state afterwards)
        self.ids.zero_()
        self.ids[:, 1:T] = text
        self.codes.zero_()
        self.seed.fill_(seed)
        if prefill:
            self._prefill(text)
            self.pos.fill_(T - 1)
            if use_graph:
                self._g_head.replay()        # first image code from the prompt's last position
            else:
                self._head(temperature, top_k, top_p)
            return T
        # position-by-position prompt (kept for tests: must give the same codes as the batched prefill)
        self.pos.zero_()
        for pos in range(T - 1):
            self._step(False, temperature, top_k, top_p)
        return T - 1

    def step_once(self):
        if self._g_step is not None:
            self._g_step.replay()
        else:
            self._step(True, *self._sampling)

    def result(self) -> torch.Tensor:
        return self.codes.clone()

    @torch.no_grad()
    def generate(self, text: torch.Tensor, *, temperature=1.0, top_k=0, top_p=1.0, generator=None,
                 use_graph: bool = True, prefill: bool = True) -> torch.Tensor:
        first = self.begin(text, temperature=temperature, top_k=top_k, top_p=top_p, generator=generator,
                           use_graph=use_graph, prefill=prefill)
        for pos in range(first, self.cfg.seq_len):
            self.step_once()
        return self.result()


@torch.no_grad()
def generate_many(engines, texts, *, temperature=1.0, top_k=0, top_p=1.0, generator=None):
    """Several batches decoded CONCURRENTLY on one GPU: one engine and one CUDA stream per batch, graph replays
    interleaved position by position.  A batch-32 decode step is a latency chain that leaves most SMs idle most of the
    time; a second (third) chain fills them, so the GPU's tokens/s grows while a request's latency barely moves."""
    streams = [torch.cuda.Stream() for _ in engines]
    cur = torch.cuda.current_stream()
    firsts = []
    for eng, text in zip(engines, texts):        # prompt passes use the shared training-forward buffers: one at a time
        if not hasattr(eng, "begin"):   # a first-generation engine (configuration outside decode2.cu): no interleaving
            return [e.generate(t, temperature=temperature, top_k=top_k, top_p=top_p, generator=generator)
                    for e, t in zip(engines, texts)]
        firsts.append(eng.begin(text, temperature=temperature, top_k=top_k, top_p=top_p, generator=generator))
    assert len(set(firsts)) == 1
    for s in streams:
        s.wait_stream(cur)
    n = engines[0].cfg.seq_len
    for pos in range(firsts[0], n):
        for eng, s in zip(engines, streams):
            with torch.cuda.stream(s):
                eng.step_once()
    for s in streams:
        cur.wait_stream(s)
    return [eng.result() for eng in engines]
