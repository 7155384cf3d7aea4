// Work decomposition shared by the third-generation (software-pipelined, persistent) sparse-attention kernels:
// attn3_fwd_sm100.cu (forward), attn3_bwd_sm100.cu (backward).  Masks: SURVEY Appendix D.3 (dalle-pytorch
// SparseAxialCausalAttention as configured at reference task.py:63-64).  conv_like keeps the second-generation
// kernels (one layer in 64).
//
// A persistent CTA walks ITEMS; an item is one (batch, head) together with either a run of image query blocks or the
// text query blocks, so that the text keys/values of that (batch, head) are fetched into shared memory once per item
// and stay there for all of its query blocks.  Text keys are cut into PIECES: full 128-key tiles and one padded tail.
#pragma once
#include "attn_common.cuh"

namespace b200 {

struct Attn3Plan {
  AttnGeom g;
  int nfull;        // full 128-key text tiles
  int tail_pad;     // rows of the tail piece (T - 128 * nfull rounded up to 16; 0 = no tail)
  int merge_tail;   // the tail rides in the same score chunk as the 128 local keys (tail_pad <= 16)
  int nkc;          // text pieces = nfull + (tail_pad > 0)
  int ib;           // image query blocks per item
  int parts_img;    // image items per (batch, head)
  int n_img_items;  // B * H * parts_img; items [n_img_items, n_items) are the text items
  int n_items;
};

// returns 0 when the geometry fits these kernels
inline int attn3_make_plan(Attn3Plan* p, int B, int H, int seq, int T, int F) {
  AttnGeom g;
  g.B = B; g.H = H; g.seq = seq; g.T = T; g.T_pad = (T + 15) / 16 * 16; g.F = F; g.P = F * F;
  g.n_text_blocks = (T + 127) / 128; g.n_img_blocks = g.P / 128; g.conv_half = 0;
  if (g.P % 128 || 32 % F || seq != T + g.P - 1) return -1;
  p->g = g;
  p->nfull = T / 128;
  p->tail_pad = g.T_pad - 128 * p->nfull;
  if (p->tail_pad > 128 || 128 * p->nfull + p->tail_pad > 272) return -1;   // text keys of one head stay in 34 KB
  p->merge_tail = (p->tail_pad > 0 && p->tail_pad <= 16) ? 1 : 0;
  p->nkc = p->nfull + (p->tail_pad > 0 ? 1 : 0);
  if (p->nkc > 3 || p->nkc < 1) return -1;
  p->ib = g.n_img_blocks >= 4 ? 4 : g.n_img_blocks;
  if (g.n_img_blocks % p->ib) return -1;
  p->parts_img = g.n_img_blocks / p->ib;
  p->n_img_items = B * H * p->parts_img;
  p->n_items = p->n_img_items + B * H;
  return 0;
}

struct A3Chunk {
  int piece;     // text piece of the chunk's text columns, -1 = none
  int ntext;     // text columns (multiple of 16)
  int row0;      // first text key of the piece
  int has_local; // 128 local-key columns follow the text columns
};

__device__ __forceinline__ int a3_piece_rows(const Attn3Plan& p, int piece) {
  return piece < p.nfull ? 128 : p.tail_pad;
}

// score chunks of an image query block
__device__ __forceinline__ int a3_img_chunks(const Attn3Plan& p) {
  return p.nfull + ((p.tail_pad > 0 && !p.merge_tail) ? 1 : 0) + 1;
}
__device__ __forceinline__ A3Chunk a3_img_chunk(const Attn3Plan& p, int c) {
  A3Chunk d;
  d.piece = -1; d.ntext = 0; d.row0 = 0; d.has_local = 0;
  if (c < p.nfull) { d.piece = c; d.ntext = 128; d.row0 = 128 * c; return d; }
  if (p.tail_pad > 0 && (p.merge_tail || c == p.nfull)) {
    d.piece = p.nfull; d.ntext = p.tail_pad; d.row0 = 128 * p.nfull;
    d.has_local = p.merge_tail;
    return d;
  }
  d.has_local = 1;
  return d;
}
// score chunks of text query block t (causal: pieces 0..t)
__device__ __forceinline__ int a3_txt_chunks(const Attn3Plan& p, int t) { return min(t + 1, p.nkc); }
__device__ __forceinline__ A3Chunk a3_txt_chunk(const Attn3Plan& p, int c) {
  A3Chunk d;
  d.piece = c; d.row0 = 128 * c; d.has_local = 0;
  d.ntext = c < p.nfull ? 128 : p.tail_pad;
  return d;
}

// Walks the query blocks of a persistent CTA in processing order.  Every role (loaders, MMA issuer, softmax groups)
// runs its own copy, so all of them agree on block numbering without communicating.
struct A3Iter {
  int it;       // item id
  int k;        // how many items this CTA has started before this one
  int j;        // block within the item
  int nblk;     // blocks of the item
  int g;        // running block number of this CTA
  int stride;
  bool valid;
  __device__ __forceinline__ void load_item(const Attn3Plan& p) {
    valid = it < p.n_items;
    nblk = valid ? (it < p.n_img_items ? p.ib : p.g.n_text_blocks) : 0;
  }
  __device__ __forceinline__ void init(const Attn3Plan& p, int first, int stride_) {
    it = first; k = 0; j = 0; g = 0; stride = stride_;
    load_item(p);
  }
  __device__ __forceinline__ void next(const Attn3Plan& p) {
    ++g;
    if (++j == nblk) { it += stride; ++k; j = 0; load_item(p); }
  }
  __device__ __forceinline__ bool is_text(const Attn3Plan& p) const { return it >= p.n_img_items; }
  // (batch, head, query block index inside its family)
  __device__ __forceinline__ void where(const Attn3Plan& p, int& b, int& h, int& blk) const {
    int bh;
    if (it < p.n_img_items) { bh = it / p.parts_img; blk = (it - bh * p.parts_img) * p.ib + j; }
    else { bh = it - p.n_img_items; blk = j; }
    b = bh / p.g.H;
    h = bh - b * p.g.H;
  }
  __device__ __forceinline__ int nchunks(const Attn3Plan& p) const {
    return is_text(p) ? a3_txt_chunks(p, j) : a3_img_chunks(p);
  }
  __device__ __forceinline__ A3Chunk chunk(const Attn3Plan& p, int c) const {
    return is_text(p) ? a3_txt_chunk(p, c) : a3_img_chunk(p, c);
  }
  // how many MMAs of this item read text piece `piece` as keys (and as many read it as values)
  __device__ __forceinline__ int piece_uses(const Attn3Plan& p, int piece) const {
    return is_text(p) ? (p.g.n_text_blocks - piece) : nblk;
  }
};

}  // namespace b200
