// K17: flash-attention backward on tcgen05 tensor cores for sm_100a (bf16 / fp16, head_dim 64/128, causal or full, GQA,
// dense batches of any length and packed variable-length batches, dropout; geometry and masking rules in
// flash_common.cuh).
//
// Replaces in the reference: paddle/phi/kernels/gpu/flash_attn_grad_kernel.cu -> vendored FlashAttention-2 backward
// (mma.sync, one kernel, fp32 atomic adds for dQ).  Here the work is split so that nothing needs atomics or an fp32
// dQ buffer and every result is deterministic:
//   delta_kernel          Delta[b,h,q] = sum_d dO * O                                  (memory-bound pre-pass)
//   flash_bwd_dq_kernel   one CTA per 128-row query tile, loops over key tiles:   S = Q K^T, dP = dO V^T on the tensor
//                         core, dS = P o (dP - Delta) * scale in registers, dQ += dS K accumulated in TMEM
//   flash_bwd_dkv_kernel  one CTA per 128-key tile, loops over the group's query heads x 128-row query tiles, with the
//                         TRANSPOSED products S^T = K Q^T and dP^T = V dO^T so that P^T and dS^T come out of TMEM with
//                         keys on the lanes, ready to be the A operand of dV += P^T dO and dK += dS^T Q (both in TMEM).
//                         GQA needs no reduction across heads: the group's heads are just more loop iterations.
// The probabilities are recomputed from the forward's LSE (P = exp2(S*scale*log2e - LSE*log2e)); no running max, no
// rescale, so the score columns of a tile can be split between two softmax warpgroups without any exchange.
// Warp roles in both kernels: 0 and 3 TMA producers, 1 MMA issuer, 2 TMEM allocator, 4-11 the element-wise stage
// (lane quarter = warp % 4, column half = (warp - 4) / 4).  Scores are pulled into registers with one tcgen05.ld pass
// and the TMEM buffers are handed back immediately, so the next tile's MMAs overlap this tile's exponentials.
#include "common.cuh"
#include "flash_common.cuh"
#include "tc.cuh"

namespace pdb {
namespace fab {
using namespace pdb::tc;
using namespace pdb::fac;

constexpr int kThreads = 384;
constexpr float kLog2e = 1.4426950408889634f;

struct BwdParams {
  SeqArgs seq;
  int hq, hk;
  float scale_log2, scale;
  int causal;
  const float* lse;      // [b, hq, sq] or [hq, total_q]
  const float* delta;    // same layout
  uint16_t* dq;          // [rows_q, hq, D]   bf16 or fp16
  uint16_t* dk;          // [rows_k, hk, D]
  uint16_t* dv;
  Dropout drop;
};

// ---------------------------------------------------------------------------------------------- Delta
// One warp per (token, head) row.  `sq` is the dense sequence length, or 0 for the packed layout, where the output is
// [hq, rows / hq] (token index = r / hq).
template <int D, bool kHalf>
__global__ void __launch_bounds__(256) delta_kernel(const uint16_t* __restrict__ o, const uint16_t* __restrict__ d_o,
                                                   float* __restrict__ delta, int sq, int hq, int64_t rows) {
  constexpr int E = D / 32;                               // elements per lane
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);     // row of the [tokens*hq, D] view
  if (r >= rows) return;
  const uint16_t* po = o + r * D + lane * E;
  const uint16_t* pd = d_o + r * D + lane * E;
  float acc = 0.f;
  if constexpr (E == 4) {
    const uint2 a = *reinterpret_cast<const uint2*>(po), c = *reinterpret_cast<const uint2*>(pd);
    const float2 x0 = unpack2<kHalf>(a.x), x1 = unpack2<kHalf>(a.y), y0 = unpack2<kHalf>(c.x), y1 = unpack2<kHalf>(c.y);
    acc = x0.x * y0.x + x0.y * y0.y + x1.x * y1.x + x1.y * y1.y;
  } else {
    const float2 x = unpack2<kHalf>(*reinterpret_cast<const uint32_t*>(po));
    const float2 y = unpack2<kHalf>(*reinterpret_cast<const uint32_t*>(pd));
    acc = x.x * y.x + x.y * y.y;
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
  if (lane == 0) {
    const int64_t tok = r / hq;                           // b * sq + s, or the packed token index
    const int h = (int)(r - tok * hq);
    if (sq > 0) {
      const int64_t bb = tok / sq;
      const int s = (int)(tok - bb * sq);
      delta[(bb * hq + h) * sq + s] = acc;
    } else {
      delta[(int64_t)h * (rows / hq) + tok] = acc;
    }
  }
}

// ---------------------------------------------------------------------------------------------- dQ
template <int D> struct DqSmem {
  static constexpr int kTile = 128 * D * 2;               // Q, dO, one K or V stage
  static constexpr int kKStages = 3;                        // a K tile is held from S = Q K^T until dQ += dS K retires
  static constexpr int kTotal = (4 + kKStages) * kTile + 1024 + 256;
};

template <int D, bool kHalf, bool kDrop>
__global__ void __launch_bounds__(kThreads, 1)
flash_bwd_dq_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_do,
                    const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                    const BwdParams p) {
  using S = DqSmem<D>;
  constexpr int KB = D / 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sDO = sQ + S::kTile;
  uint8_t* sK = sDO + S::kTile;              // [3]
  uint8_t* sV = sK + S::kKStages * S::kTile; // [2]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 2 * S::kTile);
  uint64_t* qdo_full = bars;
  uint64_t* k_full = bars + 1;               // [3]
  uint64_t* k_empty = bars + 4;              // [3]
  uint64_t* v_full = bars + 7;               // [2]
  uint64_t* v_empty = bars + 9;              // [2]
  uint64_t* sdp_full = bars + 11;
  uint64_t* sdp_empty = bars + 12;
  uint64_t* ds_ready = bars + 13;
  uint64_t* dq_done = bars + 14;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_qt = (int)gridDim.x;
  const int qt = p.causal ? (n_qt - 1 - (int)blockIdx.x) : (int)blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int hk = h / (p.hq / p.hk);
  const Seq sq_ = seq_of(p.seq, b);
  if (qt * 128 >= sq_.q_len) return;
  const int q_len = sq_.q_len, k_len = sq_.k_len;
  const int q_row0 = sq_.q_start + qt * 128;
  const int kv_row0 = sq_.k_start;
  const int diag = k_len - q_len;
  const int n_kv_all = (k_len + 127) / 128;
  const int n_kv = p.causal ? min((qt * 128 + 128 + diag + 127) / 128, n_kv_all) : n_kv_all;
  if (n_kv <= 0) {                                        // no visible key for any row of the tile: dQ = 0
    for (int i = threadIdx.x; i < 128 * (D / 8); i += kThreads) {
      const int row = i / (D / 8), c8 = i % (D / 8);
      if (qt * 128 + row < q_len)
        *reinterpret_cast<uint4*>(p.dq + ((int64_t)(q_row0 + row) * p.hq + h) * D + c8 * 8) = make_uint4(0, 0, 0, 0);
    }
    return;
  }

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_q);
    prefetch_tmap(&tm_do);
    prefetch_tmap(&tm_k);
    prefetch_tmap(&tm_v);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(qdo_full, 1);
    for (int i = 0; i < S::kKStages; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    mbar_init(sdp_full, 1);
    mbar_init(sdp_empty, 8);
    mbar_init(ds_ready, 8);
    mbar_init(dq_done, 1);
    fence_barrier_init();
  }
  constexpr int kTmemCols = 512;             // S | dP | dQ | dS (bf16 pairs, 64 columns)
  if (warp == 2) tmem_alloc(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_s = *tmem_ptr_smem;
  const uint32_t tmem_dp = tmem_s + 128;
  const uint32_t tmem_dq = tmem_s + 256;
  const uint32_t tmem_ds = tmem_s + 384;

  if (warp < 4) {
  reg_dealloc<56>();
  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(qdo_full, 2 * S::kTile);
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) {
        tma_load_2d(sQ + kb * (128 * 128), &tm_q, qdo_full, h * D + kb * 64, q_row0);
        tma_load_2d(sDO + kb * (128 * 128), &tm_do, qdo_full, h * D + kb * 64, q_row0);
      }
      for (int j = 0; j < n_kv; ++j) {
        const int st = j % S::kKStages;
        mbar_wait(&k_empty[st], ((j / S::kKStages) & 1) ^ 1);
        mbar_expect_tx(&k_full[st], S::kTile);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb)
          tma_load_2d(sK + st * S::kTile + kb * (128 * 128), &tm_k, &k_full[st], hk * D + kb * 64, kv_row0 + j * 128);
      }
    }
  } else if (warp == 3) {
    if (elect_one()) {
      for (int j = 0; j < n_kv; ++j) {
        const int st = j & 1;
        mbar_wait(&v_empty[st], ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(&v_full[st], S::kTile);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb)
          tma_load_2d(sV + st * S::kTile + kb * (128 * 128), &tm_v, &v_full[st], hk * D + kb * 64, kv_row0 + j * 128);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {   // elect.sync, not a lane test: ptxas then knows one lane runs this and issues UTCHMMA / UTCBAR from uniform registers
      constexpr uint32_t idesc_s = make_idesc_16(kHalf, 128, 128, 0, 0);    // [q][d] x [keys][d]^T
      constexpr uint32_t idesc_dq = make_idesc_16(kHalf, 128, D, 0, 1);     // dS [q][keys] x K [keys][d] (MN-major)
      auto issue_sdp = [&](int j) {
        const int st = j & 1, ph = (j >> 1) & 1, kst = j % S::kKStages;
        mbar_wait(&k_full[kst], (j / S::kKStages) & 1);
        mbar_wait(&v_full[st], ph);
        mbar_wait(sdp_empty, (j & 1) ^ 1);
        tc_fence_after();
        const uint32_t aq = smem_u32(sQ), ad = smem_u32(sDO);
        const uint32_t bk = smem_u32(sK + kst * S::kTile), bv = smem_u32(sV + st * S::kTile);
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint32_t off = (k >> 2) * (128 * 128) + (k & 3) * 32;
          umma_f16(tmem_s, make_smem_desc(aq + off, 16, 1024), make_smem_desc(bk + off, 16, 1024), idesc_s, k != 0);
        }
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint32_t off = (k >> 2) * (128 * 128) + (k & 3) * 32;
          umma_f16(tmem_dp, make_smem_desc(ad + off, 16, 1024), make_smem_desc(bv + off, 16, 1024), idesc_s, k != 0);
        }
        umma_commit(sdp_full);
        umma_commit(&v_empty[st]);
      };
      auto issue_dq = [&](int j) {
        const int st = j % S::kKStages;
        mbar_wait(ds_ready, j & 1);
        tc_fence_after();
        const uint32_t bk = smem_u32(sK + st * S::kTile);
#pragma unroll
        for (int k = 0; k < 8; ++k)       // A = dS straight from TMEM: 16 keys per step = 8 packed columns
          umma_f16_ts(tmem_dq, tmem_ds + k * 8, make_smem_desc(bk + k * 2048, 128 * 128, 1024), idesc_dq,
                      (j > 0 || k > 0) ? 1u : 0u);
        umma_commit(dq_done);
        umma_commit(&k_empty[st]);
      };
      mbar_wait(qdo_full, 0);
      issue_sdp(0);
      for (int j = 0; j < n_kv; ++j) {
        if (j + 1 < n_kv) issue_sdp(j + 1);
        issue_dq(j);
      }
    }
  }
  } else {
    reg_alloc<216>();
    const int q = warp & 3, half = (warp - 4) >> 2;
    const int row = q * 32 + lane;
    const int q_idx = qt * 128 + row;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const bool row_ok = q_idx < q_len;
    const int64_t stat = sq_.stat0 + (int64_t)h * p.seq.stat_hs + q_idx;
    // rows past the end of the sequence (and rows that saw no key: LSE = -inf) get probability 0 everywhere
    float lse2 = row_ok ? p.lse[stat] * kLog2e : INFINITY;
    if (lse2 == -INFINITY) lse2 = INFINITY;
    const float dls = row_ok ? -p.delta[stat] * p.scale : 0.f;
    const float dp_scale = kDrop ? p.scale * p.drop.inv_keep : p.scale;
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(sdp_full, j & 1);
      tc_fence_after();
      uint32_t rs[64], rd[64];
      tmem_ld_32x32(tmem_s + lane_off + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&rs[0]));
      tmem_ld_32x32(tmem_s + lane_off + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&rs[32]));
      tmem_ld_32x32(tmem_dp + lane_off + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&rd[0]));
      tmem_ld_32x32(tmem_dp + lane_off + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&rd[32]));
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(sdp_empty);
      const bool on_diag = p.causal && (j * 128 + 127 > qt * 128 + diag);
      if (on_diag || j * 128 + 128 > k_len) {                      // diagonal tile or end of the keys: probability 0
        int lim = k_len - 1 - j * 128 - half * 64;                 // local key index > lim is masked
        if (on_diag) lim = min(lim, q_idx + diag - j * 128 - half * 64);
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i > lim) rs[i] = 0xff800000u;
      }
      if constexpr (kDrop) {
        // dP of a dropped score is 0 (the kept ones carry 1 / keep, folded into dp_scale)
        const uint32_t bh = (uint32_t)(b * p.hq + h);
        const int sh = (q_idx & 1) * 16;
#pragma unroll
        for (int i = 0; i < 64; i += 2) {
          const uint32_t w = dropout_patch(p.drop, bh, (uint32_t)q_idx, (uint32_t)(j * 128 + half * 64 + i)) >> sh;
          if ((w & 0xffu) >= p.drop.threshold) rd[i] = 0u;
          if (((w >> 8) & 0xffu) >= p.drop.threshold) rd[i + 1] = 0u;
        }
      }
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 64; i += 2) {
        float x0, x1, t0, t1;
        ffma2(x0, x1, __uint_as_float(rs[i]), __uint_as_float(rs[i + 1]), p.scale_log2, -lse2);
        ffma2(t0, t1, __uint_as_float(rd[i]), __uint_as_float(rd[i + 1]), dp_scale, dls);    // (dP - Delta) * scale
        x0 = ex2_approx(x0);
        x1 = ex2_approx(x1);
        fmul2(t0, t1, t0, t1, x0, x1);
        pk[i >> 1] = pack2<kHalf>(t0, t1);
      }
      if (j >= 1) {
        mbar_wait(dq_done, (j - 1) & 1);                          // previous dS has been consumed
      }
      tmem_st_32x32(tmem_ds + lane_off + half * 32, pk);           // 64 keys of this row = 32 packed columns
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(ds_ready);
    }
    mbar_wait(dq_done, (n_kv - 1) & 1);
    tc_fence_after();
    uint16_t* out = p.dq + ((int64_t)(q_row0 + row) * p.hq + h) * D + half * (D / 2);
#pragma unroll
    for (int c = 0; c < D / 64; ++c) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_dq + lane_off + half * (D / 2) + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        uint4 v;
        v.x = pack2<kHalf>(__uint_as_float(r[i]), __uint_as_float(r[i + 1]));
        v.y = pack2<kHalf>(__uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
        v.z = pack2<kHalf>(__uint_as_float(r[i + 4]), __uint_as_float(r[i + 5]));
        v.w = pack2<kHalf>(__uint_as_float(r[i + 6]), __uint_as_float(r[i + 7]));
        if (row_ok) *reinterpret_cast<uint4*>(out + c * 32 + i) = v;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_s, kTmemCols);
}

}  // namespace fab
}  // namespace pdb

#include "flash_attn_bwd_dkv.cuh"

namespace pdb {
namespace fab {

template <int D, bool kHalf, bool kDrop>
static int launch_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                      void* dq, void* dk, void* dv, float* delta, const SeqArgs& seq, int64_t rows_q, int64_t rows_k, int b,
                      int hq, int hk, float scale, int causal, const Dropout& drop, cudaStream_t st) {
  const int64_t rows = rows_q * hq;
  delta_kernel<D, kHalf><<<(unsigned)((rows + 7) / 8), 256, 0, st>>>((const uint16_t*)o, (const uint16_t*)d_o, delta,
                                                                     seq.cu_q ? 0 : seq.sq, hq, rows);
  if (int e = check_launch("flash_attn_bwd delta")) return e;
  CUtensorMap tq, tdo, tk, tv;
  const uint64_t qc = (uint64_t)hq * D, kc = (uint64_t)hk * D;
  if (get_tmap_2d(&tq, q, (uint64_t)rows_q, qc, qc * 2, 128, 64, 2, true, "flash_attn q")) return -1;
  if (get_tmap_2d(&tdo, d_o, (uint64_t)rows_q, qc, qc * 2, 128, 64, 2, true, "flash_attn d_o")) return -1;
  if (get_tmap_2d(&tk, k, (uint64_t)rows_k, kc, kc * 2, 128, 64, 2, true, "flash_attn k")) return -1;
  if (get_tmap_2d(&tv, v, (uint64_t)rows_k, kc, kc * 2, 128, 64, 2, true, "flash_attn v")) return -1;
  BwdParams p{seq, hq, hk, scale * kLog2e, scale, causal, lse, delta, (uint16_t*)dq, (uint16_t*)dk, (uint16_t*)dv, drop};
  auto kq = flash_bwd_dq_kernel<D, kHalf, kDrop>;
  auto kkv = flash_bwd_dkv_kernel<D, kHalf, kDrop>;
  PDB_ENSURE_SMEM(kq, DqSmem<D>::kTotal);
  PDB_ENSURE_SMEM(kkv, DkvSmem<D>::kTotal);
  kkv<<<dim3((seq.sk + 127) / 128, hk, b), kThreads, DkvSmem<D>::kTotal, st>>>(tq, tdo, tk, tv, p);
  if (int e = check_launch("flash_attn_bwd dkv")) return e;
  kq<<<dim3((seq.sq + 127) / 128, hq, b), kThreads, DqSmem<D>::kTotal, st>>>(tq, tdo, tk, tv, p);
  return check_launch("flash_attn_bwd dq");
}

}  // namespace fab
}  // namespace pdb

using namespace pdb;

#define PDB_FA_BWD_ARGS q, k, v, o, d_o, lse, dq, dk, dv, (float*)workspace, seq, rows_q, rows_k, (int)b, (int)hq, (int)hk, scale, causal, drop, st

static int fa_bwd_common(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                         void* dq, void* dk, void* dv, void* workspace, int64_t workspace_bytes, const fac::SeqArgs& seq,
                         int64_t rows_q, int64_t rows_k, int64_t b, int64_t hq, int64_t hk, int64_t d, float scale,
                         int causal, float p_drop, uint64_t seed, uint64_t offset, int dtype, void* stream) {
  PDB_REQUIRE(dtype == PDB_DTYPE_BF16 || dtype == PDB_DTYPE_F16, "flash_attn_bwd: bfloat16 or float16 only");
  PDB_REQUIRE(d == 64 || d == 128, "flash_attn_bwd: head_dim must be 64 or 128 (got %lld)", (long long)d);
  PDB_REQUIRE(b > 0 && seq.sq > 0 && seq.sk > 0 && rows_q > 0 && rows_k > 0, "flash_attn_bwd: empty problem");
  PDB_REQUIRE(hq % hk == 0, "flash_attn_bwd: q heads must be a multiple of kv heads");
  PDB_REQUIRE(p_drop >= 0.f && p_drop < 1.f, "flash_attn_bwd: dropout must be in [0, 1)");
  PDB_REQUIRE(workspace_bytes >= rows_q * hq * (int64_t)sizeof(float) && ((uintptr_t)workspace & 15) == 0,
              "flash_attn_bwd: workspace too small or misaligned");
  const fac::Dropout drop = fa::make_dropout(p_drop, seed, offset);
  cudaStream_t st = (cudaStream_t)stream;
  const bool half = dtype == PDB_DTYPE_F16, dr = drop.threshold != 0;
  if (d == 128) {
    if (half) return dr ? fab::launch_bwd<128, true, true>(PDB_FA_BWD_ARGS) : fab::launch_bwd<128, true, false>(PDB_FA_BWD_ARGS);
    return dr ? fab::launch_bwd<128, false, true>(PDB_FA_BWD_ARGS) : fab::launch_bwd<128, false, false>(PDB_FA_BWD_ARGS);
  }
  if (half) return dr ? fab::launch_bwd<64, true, true>(PDB_FA_BWD_ARGS) : fab::launch_bwd<64, true, false>(PDB_FA_BWD_ARGS);
  return dr ? fab::launch_bwd<64, false, true>(PDB_FA_BWD_ARGS) : fab::launch_bwd<64, false, false>(PDB_FA_BWD_ARGS);
}

PDB_EXPORT int64_t pdb200_flash_attn_bwd_workspace_bytes(int64_t b, int64_t sq, int64_t sk, int64_t hq, int64_t hk, int64_t d) {
  (void)sk; (void)hk; (void)d;
  return b * hq * sq * (int64_t)sizeof(float);            // Delta (packed layout: pass b = 1, sq = total_q)
}

PDB_EXPORT int pdb200_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                                     const float* lse, void* dq, void* dk, void* dv, void* workspace,
                                     int64_t workspace_bytes, int64_t b, int64_t sq, int64_t sk, int64_t hq, int64_t hk,
                                     int64_t d, float scale, int causal, int dtype, void* stream) {
  fac::SeqArgs seq{nullptr, nullptr, (int)sq, (int)sk, hq * sq, sq};
  return fa_bwd_common(q, k, v, o, d_o, lse, dq, dk, dv, workspace, workspace_bytes, seq, b * sq, b * sk, b, hq, hk, d, scale,
                       causal, 0.f, 0, 0, dtype, stream);
}

PDB_EXPORT int pdb200_flash_attn_bwd_dropout(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                                             const float* lse, void* dq, void* dk, void* dv, void* workspace,
                                             int64_t workspace_bytes, int64_t b, int64_t sq, int64_t sk, int64_t hq,
                                             int64_t hk, int64_t d, float scale, int causal, float p_drop, uint64_t seed,
                                             uint64_t offset, int dtype, void* stream) {
  fac::SeqArgs seq{nullptr, nullptr, (int)sq, (int)sk, hq * sq, sq};
  return fa_bwd_common(q, k, v, o, d_o, lse, dq, dk, dv, workspace, workspace_bytes, seq, b * sq, b * sk, b, hq, hk, d, scale,
                       causal, p_drop, seed, offset, dtype, stream);
}

PDB_EXPORT int pdb200_flash_attn_varlen_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                                            const float* lse, void* dq, void* dk, void* dv, void* workspace,
                                            int64_t workspace_bytes, const int* cu_seqlens_q, const int* cu_seqlens_k,
                                            int64_t b, int64_t total_q, int64_t total_k, int64_t max_seqlen_q,
                                            int64_t max_seqlen_k, int64_t hq, int64_t hk, int64_t d, float scale, int causal,
                                            float p_drop, uint64_t seed, uint64_t offset, int dtype, void* stream) {
  PDB_REQUIRE(cu_seqlens_q && cu_seqlens_k, "flash_attn_varlen_bwd: cu_seqlens are required");
  fac::SeqArgs seq{cu_seqlens_q, cu_seqlens_k, (int)max_seqlen_q, (int)max_seqlen_k, 0, total_q};
  return fa_bwd_common(q, k, v, o, d_o, lse, dq, dk, dv, workspace, workspace_bytes, seq, total_q, total_k, b, hq, hk, d, scale,
                       causal, p_drop, seed, offset, dtype, stream);
}
