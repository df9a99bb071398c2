// Shared by the flash-attention forward and backward kernels: per-sequence geometry (dense batches and packed
// variable-length batches go through the same code), element-type helpers (bf16 / fp16), and the dropout mask.
//
// Geometry.  Q/K/V/O are addressed as 2-D [rows][heads * D] tensors: rows = b * s for the dense layout
// [b, s, h, d], rows = total tokens for the packed layout [total, h, d] with cu_seqlens (reference:
// paddle/phi/kernels/gpu/flash_attn_kernel.cu:36-170).  A 128-row TMA box that starts inside a sequence may run past
// its end: the extra rows are the next sequence's tokens or, past the end of the tensor, zeros.  Nothing is done about
// that on the load side; every kernel masks by the sequence lengths instead - scores of keys >= k_len become -inf
// (probability 0), statistics of queries >= q_len are forced to (+inf, 0) so their probabilities are 0, and stores are
// limited to valid rows.
#pragma once
#include "common.cuh"

namespace pdb {
namespace fac {

struct SeqArgs {
  const int* cu_q;          // [b + 1] prefix sums of query lengths, or nullptr for the dense layout
  const int* cu_k;
  int sq, sk;               // dense: the sequence lengths; packed: the maximum lengths (grid sizing only)
  int64_t stat_bs, stat_hs; // LSE / Delta strides: dense [b, hq, sq] -> (hq*sq, sq); packed [hq, total_q] -> (0, total_q)
};

struct Seq {
  int q_start, q_len, k_start, k_len;
  int64_t stat0;            // index of (this sequence, head 0, query 0) in LSE / Delta
};

__device__ __forceinline__ Seq seq_of(const SeqArgs& a, int b) {
  Seq s;
  if (a.cu_q) {
    s.q_start = a.cu_q[b];
    s.q_len = a.cu_q[b + 1] - s.q_start;
    s.k_start = a.cu_k[b];
    s.k_len = a.cu_k[b + 1] - s.k_start;
    s.stat0 = s.q_start;
  } else {
    s.q_start = b * a.sq;
    s.q_len = a.sq;
    s.k_start = b * a.sk;
    s.k_len = a.sk;
    s.stat0 = (int64_t)b * a.stat_bs;
  }
  return s;
}

// Two floats -> one packed 32-bit word of the kernel's element type.
template <bool kHalf> __device__ __forceinline__ uint32_t pack2(float a, float b) {
  if constexpr (kHalf) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  } else {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
}
template <bool kHalf> __device__ __forceinline__ float2 unpack2(uint32_t w) {
  if constexpr (kHalf) {
    return __half22float2(*reinterpret_cast<__half2*>(&w));
  } else {
    return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&w));
  }
}

// Dropout.  The keep/drop decision of score (bh, q, k) is a pure function of (seed, offset, bh, q, k), so the backward
// kernels regenerate exactly the forward's mask.  One 32-bit hash serves a 2 x 2 patch (two queries x two keys), one
// byte each, compared with an 8-bit threshold: both the row-per-thread forward / dQ kernels and the key-per-thread
// dK/dV kernel use half of every hash.  keep <=> byte < keep_threshold, keep_threshold = round((1 - p) * 256).
struct Dropout {
  uint32_t seed_lo, seed_hi;    // seed ^ offset folded by the host
  uint32_t threshold;           // 0 => dropout disabled
  float inv_keep;               // 1 / (1 - p)
};

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}
// hash of the 2x2 patch (q >> 1, k >> 1) of head-batch bh; byte ((q & 1) * 2 + (k & 1)) belongs to (q, k)
__device__ __forceinline__ uint32_t dropout_patch(const Dropout& d, uint32_t bh, uint32_t q, uint32_t k) {
  const uint32_t a = mix32((q >> 1) * 0x9e3779b9u + d.seed_lo);
  return mix32(a ^ ((k >> 1) * 0x85ebca6bu + bh * 0xc2b2ae35u + d.seed_hi));
}
__device__ __forceinline__ bool dropout_keep(const Dropout& d, uint32_t bh, uint32_t q, uint32_t k) {
  const uint32_t w = dropout_patch(d, bh, q, k);
  return ((w >> (((q & 1) * 2 + (k & 1)) * 8)) & 0xffu) < d.threshold;
}

}  // namespace fac
namespace fa {
// host: keep probability -> 8-bit threshold + output scale; (seed, offset) folded into two 32-bit words (flash_attn.cu)
fac::Dropout make_dropout(float p_drop, uint64_t seed, uint64_t offset);
}  // namespace fa
}  // namespace pdb
