// K17: flash-attention backward on tcgen05 tensor cores for sm_100a (bf16, head_dim 64/128, causal or full, GQA).
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
#include "tc.cuh"

namespace pdb {
namespace fab {
using namespace pdb::tc;

constexpr int kThreads = 384;
constexpr float kLog2e = 1.4426950408889634f;

struct BwdParams {
  int sq, sk, hq, hk;
  float scale_log2, scale;
  int causal;
  const float* lse;      // [b, hq, sq]
  const float* delta;    // [b, hq, sq]
  __nv_bfloat16* dq;     // [b, sq, hq, D]
  __nv_bfloat16* dk;     // [b, sk, hk, D]
  __nv_bfloat16* dv;
};

// ---------------------------------------------------------------------------------------------- Delta
template <int D>
__global__ void __launch_bounds__(256) delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                                                   float* __restrict__ delta, int sq, int hq, int64_t rows) {
  constexpr int E = D / 32;                               // elements per lane
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);     // row of the [b*sq*hq, D] view
  if (r >= rows) return;
  const __nv_bfloat16* po = o + r * D + lane * E;
  const __nv_bfloat16* pd = d_o + r * D + lane * E;
  float acc = 0.f;
  if constexpr (E == 4) {
    const uint2 a = *reinterpret_cast<const uint2*>(po), c = *reinterpret_cast<const uint2*>(pd);
    const __nv_bfloat162* a2 = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* c2 = reinterpret_cast<const __nv_bfloat162*>(&c);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const float2 x = __bfloat1622float2(a2[i]), y = __bfloat1622float2(c2[i]);
      acc += x.x * y.x + x.y * y.y;
    }
  } else {
    const float2 x = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(po));
    const float2 y = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(pd));
    acc = x.x * y.x + x.y * y.y;
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
  if (lane == 0) {
    const int64_t bs = r / hq;                            // b * sq + s
    const int h = (int)(r - bs * hq);
    const int64_t bb = bs / sq;
    const int s = (int)(bs - bb * sq);
    delta[(bb * hq + h) * sq + s] = acc;
  }
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h2 = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h2);
}

// ---------------------------------------------------------------------------------------------- dQ
template <int D> struct DqSmem {
  static constexpr int kTile = 128 * D * 2;               // Q, dO, one K or V stage
  static constexpr int kKStages = 3;                        // a K tile is held from S = Q K^T until dQ += dS K retires
  static constexpr int kTotal = (4 + kKStages) * kTile + 1024 + 256;
};

template <int D>
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
  const int n_qt = p.sq / 128;
  const int qt = p.causal ? (n_qt - 1 - (int)blockIdx.x) : (int)blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int hk = h / (p.hq / p.hk);
  const int q_row0 = b * p.sq + qt * 128;
  const int kv_row0 = b * p.sk;
  const int diag = p.sk - p.sq;
  const int n_kv = p.causal ? min((qt * 128 + 128 + diag + 127) / 128, p.sk / 128) : p.sk / 128;

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
    if (lane == 0) {
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
    if (lane == 0) {
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
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, 0, 0);    // [q][d] x [keys][d]^T
      constexpr uint32_t idesc_dq = make_idesc_bf16(128, D, 0, 1);     // dS [q][keys] x K [keys][d] (MN-major)
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
    const int64_t stat = ((int64_t)b * p.hq + h) * p.sq + q_idx;
    const float lse2 = p.lse[stat] * kLog2e;
    const float dls = -p.delta[stat] * p.scale;
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
      if (p.causal && (j * 128 + 127 > qt * 128 + diag)) {         // diagonal tile: masked keys get probability 0
        const int lim = q_idx + diag - j * 128 - half * 64;        // local key index > lim is masked
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i > lim) rs[i] = 0xff800000u;
      }
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 64; i += 2) {
        float x0, x1, t0, t1;
        ffma2(x0, x1, __uint_as_float(rs[i]), __uint_as_float(rs[i + 1]), p.scale_log2, -lse2);
        ffma2(t0, t1, __uint_as_float(rd[i]), __uint_as_float(rd[i + 1]), p.scale, dls);     // (dP - Delta) * scale
        x0 = ex2_approx(x0);
        x1 = ex2_approx(x1);
        fmul2(t0, t1, t0, t1, x0, x1);
        pk[i >> 1] = pack_bf16(t0, t1);
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
    __nv_bfloat16* out = p.dq + ((int64_t)(b * p.sq + q_idx) * p.hq + h) * D + half * (D / 2);
#pragma unroll
    for (int c = 0; c < D / 64; ++c) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_dq + lane_off + half * (D / 2) + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        uint4 v;
        v.x = pack_bf16(__uint_as_float(r[i]), __uint_as_float(r[i + 1]));
        v.y = pack_bf16(__uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
        v.z = pack_bf16(__uint_as_float(r[i + 4]), __uint_as_float(r[i + 5]));
        v.w = pack_bf16(__uint_as_float(r[i + 6]), __uint_as_float(r[i + 7]));
        *reinterpret_cast<uint4*>(out + c * 32 + i) = v;
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

template <int D>
static int launch_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                      void* dq, void* dk, void* dv, float* delta, int b, int sq, int sk, int hq, int hk, float scale,
                      int causal, cudaStream_t st) {
  const int64_t rows = (int64_t)b * sq * hq;
  delta_kernel<D><<<(unsigned)((rows + 7) / 8), 256, 0, st>>>((const __nv_bfloat16*)o, (const __nv_bfloat16*)d_o, delta, sq, hq, rows);
  if (int e = check_launch("flash_attn_bwd delta")) return e;
  CUtensorMap tq, tdo, tk, tv;
  const uint64_t qc = (uint64_t)hq * D, kc = (uint64_t)hk * D;
  if (get_tmap_2d(&tq, q, (uint64_t)b * sq, qc, qc * 2, 128, 64, 2, true, "flash_attn q")) return -1;
  if (get_tmap_2d(&tdo, d_o, (uint64_t)b * sq, qc, qc * 2, 128, 64, 2, true, "flash_attn d_o")) return -1;
  if (get_tmap_2d(&tk, k, (uint64_t)b * sk, kc, kc * 2, 128, 64, 2, true, "flash_attn k")) return -1;
  if (get_tmap_2d(&tv, v, (uint64_t)b * sk, kc, kc * 2, 128, 64, 2, true, "flash_attn v")) return -1;
  BwdParams p{sq, sk, hq, hk, scale * kLog2e, scale, causal, lse, delta, (__nv_bfloat16*)dq, (__nv_bfloat16*)dk,
              (__nv_bfloat16*)dv};
  auto kq = flash_bwd_dq_kernel<D>;
  auto kkv = flash_bwd_dkv_kernel<D>;
  PDB_ENSURE_SMEM(kq, DqSmem<D>::kTotal);
  PDB_ENSURE_SMEM(kkv, DkvSmem<D>::kTotal);
  kkv<<<dim3(sk / 128, hk, b), kThreads, DkvSmem<D>::kTotal, st>>>(tq, tdo, tk, tv, p);
  if (int e = check_launch("flash_attn_bwd dkv")) return e;
  kq<<<dim3(sq / 128, hq, b), kThreads, DqSmem<D>::kTotal, st>>>(tq, tdo, tk, tv, p);
  return check_launch("flash_attn_bwd dq");
}

}  // namespace fab
}  // namespace pdb

using namespace pdb;

PDB_EXPORT int64_t pdb200_flash_attn_bwd_workspace_bytes(int64_t b, int64_t sq, int64_t sk, int64_t hq, int64_t hk, int64_t d) {
  (void)sk; (void)hk; (void)d;
  return b * hq * sq * (int64_t)sizeof(float);            // Delta
}

PDB_EXPORT int pdb200_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                                     const float* lse, void* dq, void* dk, void* dv, void* workspace,
                                     int64_t workspace_bytes, int64_t b, int64_t sq, int64_t sk, int64_t hq, int64_t hk,
                                     int64_t d, float scale, int causal, int dtype, void* stream) {
  PDB_REQUIRE(dtype == PDB_DTYPE_BF16, "flash_attn_bwd: bfloat16 only");
  PDB_REQUIRE(d == 64 || d == 128, "flash_attn_bwd: head_dim must be 64 or 128 (got %lld)", (long long)d);
  PDB_REQUIRE(sq % 128 == 0 && sk % 128 == 0 && sq > 0 && sk > 0, "flash_attn_bwd: sequence lengths must be multiples of 128");
  PDB_REQUIRE(hq % hk == 0, "flash_attn_bwd: q heads must be a multiple of kv heads");
  PDB_REQUIRE(!causal || sk >= sq, "flash_attn_bwd: causal needs sk >= sq");
  PDB_REQUIRE(workspace_bytes >= pdb200_flash_attn_bwd_workspace_bytes(b, sq, sk, hq, hk, d) &&
                  ((uintptr_t)workspace & 15) == 0,
              "flash_attn_bwd: workspace too small or misaligned");
  cudaStream_t st = (cudaStream_t)stream;
  if (d == 128)
    return fab::launch_bwd<128>(q, k, v, o, d_o, lse, dq, dk, dv, (float*)workspace, (int)b, (int)sq, (int)sk, (int)hq,
                                (int)hk, scale, causal, st);
  return fab::launch_bwd<64>(q, k, v, o, d_o, lse, dq, dk, dv, (float*)workspace, (int)b, (int)sq, (int)sk, (int)hq,
                             (int)hk, scale, causal, st);
}
