// K16: flash-attention forward on tcgen05 tensor cores for sm_100a (bf16, head_dim 64/128, causal or full, GQA).
//
// Replaces in the reference: paddle/phi/kernels/gpu/flash_attn_kernel.cu:305-447 -> vendored FlashAttention-2
// (mma.sync m16n8k16 + cp.async, third_party/flashattn/csrc/flash_attn/src/flash_fwd_kernel.h).
//
// One CTA per (batch, q-head, 128-row query tile); K/V tiles of 128 keys stream through a 2-stage TMA ring.
//   warp 0    TMA producer   Q once, then K_j / V_j (128-B swizzled boxes)
//   warp 1    MMA issuer     S_j = Q K_j^T into TMEM (double-buffered, so S_{j+1} is computed while the softmax of
//                            tile j runs), then O_j = P_j V_j into TMEM once P_j is in shared memory
//   warp 2    TMEM allocator (2 x 128 columns for S, D columns for O_j)
//   warps 4-7 softmax        one query row per thread: tcgen05.ld the score row, online max / exp2 / row sum in fp32,
//                            write P_j (bf16) into shared memory in the UMMA K-major swizzled layout, fold the previous
//                            tile's O_{j-1} from TMEM into the fp32 register accumulator with the running rescale
// K_j and V_j share one TMA box geometry ([keys][64 d-columns]); the descriptor reads the same bytes K-major for
// Q K^T and MN-major for P V, so V needs no transpose.
// Outputs: O [b, sq, hq, d] bf16 and LSE [b, hq, sq] fp32 (natural log), the contract of the backward kernels.
#include "common.cuh"
#include "tc.cuh"

namespace pdb {
namespace fa {
using namespace pdb::tc;

constexpr int BQ = 128;     // query rows per CTA
constexpr int BKV = 128;    // keys per tile
constexpr int kThreads = 256;

template <int D> struct FwdSmem {
  static constexpr int kQBytes = BQ * D * 2;
  static constexpr int kKVBytes = BKV * D * 2;            // one of K or V
  static constexpr int kPBytes = BQ * BKV * 2;
  static constexpr int kStages = 2;
  static constexpr int kTotal = kQBytes + kStages * 2 * kKVBytes + kPBytes + 1024 + 256;
};

struct FwdParams {
  int sq, sk, hq, hk;
  float scale_log2;     // softmax_scale * log2(e)
  float scale;          // softmax_scale
  int causal;
  __nv_bfloat16* o;
  float* lse;
};

template <int D>
__global__ void __launch_bounds__(kThreads, 1)
flash_fwd_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                 const __grid_constant__ CUtensorMap tm_v, const FwdParams p) {
  using S = FwdSmem<D>;
  constexpr int KB = D / 64;                 // 64-column k-blocks of the head dimension
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + S::kQBytes;            // stage s: K at sKV + s*2*kKVBytes, V right after it
  uint8_t* sP = sKV + S::kStages * 2 * S::kKVBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + S::kPBytes);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;              // [2]
  uint64_t* kv_empty = bars + 3;             // [2]
  uint64_t* s_full = bars + 5;               // [2]
  uint64_t* s_empty = bars + 7;              // [2]
  uint64_t* p_ready = bars + 9;
  uint64_t* o_full = bars + 10;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 11);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heavy (late) causal tiles first so the tail of the grid is made of short CTAs
  const int n_qt = p.sq / BQ;
  const int qt = p.causal ? (n_qt - 1 - (int)blockIdx.x) : (int)blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int hk = h / (p.hq / p.hk);
  const int q_row0 = b * p.sq + qt * BQ;                 // row in the [b*sq, hq*D] view
  const int kv_row0 = b * p.sk;
  const int diag = p.sk - p.sq;                          // key index aligned with query index 0
  const int n_kv = p.causal ? min((qt * BQ + BQ + diag + BKV - 1) / BKV, p.sk / BKV) : p.sk / BKV;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_q);
    prefetch_tmap(&tm_k);
    prefetch_tmap(&tm_v);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&s_empty[i], 4);
    }
    mbar_init(p_ready, 4);
    mbar_init(o_full, 1);
    fence_barrier_init();
  }
  constexpr int kTmemCols = 512;             // S0 | S1 | O  (2*128 + D, rounded to a power of two)
  if (warp == 2) tmem_alloc(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tmem_o = tmem_base + 2 * BKV;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, S::kQBytes);
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) tma_load_2d(sQ + kb * (BQ * 128), &tm_q, q_full, h * D + kb * 64, q_row0);
      for (int j = 0; j < n_kv; ++j) {
        const int st = j & 1;
        mbar_wait(&kv_empty[st], ((j >> 1) & 1) ^ 1);
        uint8_t* sK = sKV + st * 2 * S::kKVBytes;
        uint8_t* sV = sK + S::kKVBytes;
        mbar_expect_tx(&kv_full[st], 2 * S::kKVBytes);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb) {
          tma_load_2d(sK + kb * (BKV * 128), &tm_k, &kv_full[st], hk * D + kb * 64, kv_row0 + j * BKV);
          tma_load_2d(sV + kb * (BKV * 128), &tm_v, &kv_full[st], hk * D + kb * 64, kv_row0 + j * BKV);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(BQ, BKV, 0, 0);     // Q (K-major) x K (K-major)
      constexpr uint32_t idesc_o = make_idesc_bf16(BQ, D, 0, 1);       // P (K-major) x V (MN-major)
      auto issue_pv = [&](int j) {
        const uint32_t sV = smem_u32(sKV + (j & 1) * 2 * S::kKVBytes + S::kKVBytes);
        const uint32_t sp = smem_u32(sP);
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {
          const uint64_t da = make_smem_desc(sp + (k >> 2) * (BQ * 128) + (k & 3) * 32, 16, 1024);
          const uint64_t db = make_smem_desc(sV + k * 2048, BKV * 128, 1024);
          umma_f16(tmem_o, da, db, idesc_o, k != 0 ? 1u : 0u);
        }
        umma_commit(o_full);
        umma_commit(&kv_empty[j & 1]);
      };
      mbar_wait(q_full, 0);
      for (int j = 0; j < n_kv; ++j) {
        const int st = j & 1;
        mbar_wait(&kv_full[st], (j >> 1) & 1);
        mbar_wait(&s_empty[st], ((j >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t sq_ = smem_u32(sQ);
        const uint32_t sK = smem_u32(sKV + st * 2 * S::kKVBytes);
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint64_t da = make_smem_desc(sq_ + (k >> 2) * (BQ * 128) + (k & 3) * 32, 16, 1024);
          const uint64_t db = make_smem_desc(sK + (k >> 2) * (BKV * 128) + (k & 3) * 32, 16, 1024);
          umma_f16(tmem_base + st * BKV, da, db, idesc_s, k != 0 ? 1u : 0u);
        }
        umma_commit(&s_full[st]);
        if (j >= 1) {
          mbar_wait(p_ready, (j - 1) & 1);
          tc_fence_after();
          issue_pv(j - 1);
        }
      }
      mbar_wait(p_ready, (n_kv - 1) & 1);
      tc_fence_after();
      issue_pv(n_kv - 1);
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int row = q * 32 + lane;                          // row inside the tile
    const int q_idx = qt * BQ + row;                        // query position in the sequence
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    float o_acc[D];
#pragma unroll
    for (int i = 0; i < D; ++i) o_acc[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f, alpha_prev = 1.f;
    auto fold_o = [&](int jprev, float alpha) {
      // O_acc = O_acc * alpha + O_jprev   (alpha rescales everything accumulated BEFORE tile jprev)
      mbar_wait(o_full, jprev & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_o + lane_off + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] = o_acc[c * 32 + i] * alpha + __uint_as_float(r[i]);
      }
    };
    for (int j = 0; j < n_kv; ++j) {
      const int st = j & 1;
      const uint32_t ts = tmem_base + st * BKV + lane_off;
      mbar_wait(&s_full[st], (j >> 1) & 1);
      tc_fence_after();
      const bool need_mask = p.causal && (j * BKV + BKV - 1 > qt * BQ + diag);   // tile crosses the diagonal for some row
      // ---- pass 1: row max
      float m_tile = -INFINITY;
#pragma unroll
      for (int c = 0; c < BKV / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(ts + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float s = __uint_as_float(r[i]);
          if (need_mask && (j * BKV + c * 32 + i > q_idx + diag)) s = -INFINITY;
          m_tile = fmaxf(m_tile, s);
        }
      }
      const float m_new = fmaxf(m_run, m_tile);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = exp2f((m_run - m_use) * p.scale_log2);      // m_run = -inf -> 0
      // ---- previous tile's P V product joins the accumulator (P buffer becomes free)
      if (j >= 1) fold_o(j - 1, alpha_prev);
      // ---- pass 2: probabilities -> shared memory (UMMA K-major, 128-byte swizzle), row sum
      float l_tile = 0.f;
      const float mb = m_use * p.scale_log2;
#pragma unroll
      for (int c = 0; c < BKV / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(ts + c * 32, r);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float s0 = __uint_as_float(r[i]), s1 = __uint_as_float(r[i + 1]);
          float p0 = exp2f(s0 * p.scale_log2 - mb), p1 = exp2f(s1 * p.scale_log2 - mb);
          if (need_mask) {
            if (j * BKV + c * 32 + i > q_idx + diag) p0 = 0.f;
            if (j * BKV + c * 32 + i + 1 > q_idx + diag) p1 = 0.f;
          }
          l_tile += p0 + p1;
          __nv_bfloat162 h2 = __floats2bfloat162_rn(p0, p1);
          pk[i >> 1] = *reinterpret_cast<uint32_t*>(&h2);
        }
        // 32 keys = 4 chunks of 16 B inside k-block (c / 2), chunk index ((c & 1) * 4 + t), XOR-swizzled by row % 8
        uint8_t* base = sP + (c >> 1) * (BQ * 128) + row * 128;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int chunk = ((c & 1) * 4 + t) ^ (row & 7);
          *reinterpret_cast<uint4*>(base + chunk * 16) = make_uint4(pk[t * 4], pk[t * 4 + 1], pk[t * 4 + 2], pk[t * 4 + 3]);
        }
      }
      l_run = l_run * alpha + l_tile;
      m_run = m_new;
      alpha_prev = alpha;
      // hand the score buffer back and publish P to the tensor core (async proxy)
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&s_empty[st]);
        mbar_arrive(p_ready);
      }
      // alpha rescales what was accumulated before tile j; it is applied when O_j is folded (next iteration)
    }
    // last tile
    fold_o(n_kv - 1, alpha_prev);
    const float inv_l = l_run > 0.f ? 1.f / l_run : 0.f;
    __nv_bfloat16* orow = p.o + ((int64_t)(b * p.sq + q_idx) * p.hq + h) * D;
#pragma unroll
    for (int i = 0; i < D; i += 8) {
      uint4 v;
      __nv_bfloat162 a0 = __floats2bfloat162_rn(o_acc[i] * inv_l, o_acc[i + 1] * inv_l);
      __nv_bfloat162 a1 = __floats2bfloat162_rn(o_acc[i + 2] * inv_l, o_acc[i + 3] * inv_l);
      __nv_bfloat162 a2 = __floats2bfloat162_rn(o_acc[i + 4] * inv_l, o_acc[i + 5] * inv_l);
      __nv_bfloat162 a3 = __floats2bfloat162_rn(o_acc[i + 6] * inv_l, o_acc[i + 7] * inv_l);
      v.x = *reinterpret_cast<uint32_t*>(&a0); v.y = *reinterpret_cast<uint32_t*>(&a1);
      v.z = *reinterpret_cast<uint32_t*>(&a2); v.w = *reinterpret_cast<uint32_t*>(&a3);
      *reinterpret_cast<uint4*>(orow + i) = v;
    }
    if (p.lse) {
      const float m_fin = (m_run == -INFINITY) ? 0.f : m_run;
      p.lse[((int64_t)b * p.hq + h) * p.sq + q_idx] = m_fin * p.scale + __logf(fmaxf(l_run, 1e-37f));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

template <int D>
static int launch_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int b, int sq, int sk, int hq,
                      int hk, float scale, int causal, cudaStream_t st) {
  using S = FwdSmem<D>;
  CUtensorMap tm_q, tm_k, tm_v;
  if (encode_tmap_2d(&tm_q, q, (uint64_t)b * sq, (uint64_t)hq * D, (uint64_t)hq * D * 2, BQ, 64, 2, true)) return -1;
  if (encode_tmap_2d(&tm_k, k, (uint64_t)b * sk, (uint64_t)hk * D, (uint64_t)hk * D * 2, BKV, 64, 2, true)) return -1;
  if (encode_tmap_2d(&tm_v, v, (uint64_t)b * sk, (uint64_t)hk * D, (uint64_t)hk * D * 2, BKV, 64, 2, true)) return -1;
  auto kern = flash_fwd_kernel<D>;
  static bool attr_set = false;
  if (!attr_set) {
    PDB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
    attr_set = true;
  }
  FwdParams p{sq, sk, hq, hk, scale * 1.4426950408889634f, scale, causal, (__nv_bfloat16*)o, lse};
  dim3 grid(sq / BQ, hq, b);
  kern<<<grid, kThreads, S::kTotal, st>>>(tm_q, tm_k, tm_v, p);
  return check_launch("flash_attn_fwd");
}

}  // namespace fa
}  // namespace pdb

using namespace pdb;

PDB_EXPORT int pdb200_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int64_t b,
                                     int64_t sq, int64_t sk, int64_t hq, int64_t hk, int64_t d, float scale, int causal,
                                     int dtype, void* stream) {
  PDB_REQUIRE(dtype == PDB_DTYPE_BF16, "flash_attn_fwd: bfloat16 only");
  PDB_REQUIRE(d == 64 || d == 128, "flash_attn_fwd: head_dim must be 64 or 128 (got %lld)", (long long)d);
  PDB_REQUIRE(sq % 128 == 0 && sk % 128 == 0 && sq > 0 && sk > 0, "flash_attn_fwd: sequence lengths must be multiples of 128");
  PDB_REQUIRE(hq % hk == 0, "flash_attn_fwd: q heads must be a multiple of kv heads");
  PDB_REQUIRE(!causal || sk >= sq, "flash_attn_fwd: causal needs sk >= sq");
  cudaStream_t st = (cudaStream_t)stream;
  if (d == 128) return fa::launch_fwd<128>(q, k, v, o, lse, (int)b, (int)sq, (int)sk, (int)hq, (int)hk, scale, causal, st);
  return fa::launch_fwd<64>(q, k, v, o, lse, (int)b, (int)sq, (int)sk, (int)hq, (int)hk, scale, causal, st);
}
