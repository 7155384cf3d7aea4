"""GPU executor of KV-cached sampling: one decode step = embed -> 64 x (LN+shift, QKV GEMM+rotary, cache append +
sparse attention, out-proj GEMM+residual, LN+shift, GEGLU GEMM, FF2 GEMM+residual) -> final norm -> image-segment
logits GEMM -> top-k/top-p sampler, all sm_100a kernels reading the position from device memory, captured ONCE as a
CUDA graph and replayed for the 257 prompt positions and the 1024 image positions."""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from ..config import DalleConfig
from ..ops import gemm as G
from ..ops import lib
from .masks import ATTN_TYPE_CODES


class DecodeEngine:
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
