// K23 / K13, CTA-pair variant: persistent bf16 GEMM where two SMs of one TPC work on one 256 x 256 output tile with
// tcgen05.mma.cta_group::2 (the shape the library's own kernels use on this chip: *_256x256_64x4_*_2cta_*).
//
// Why a pair: with one CTA per tile, every 128 x 256 x 16 MMA reads 12 KB of operands from that SM's shared memory in
// 128 cycles (96 B/clk), which is what held the single-CTA kernel at ~78 % of the library.  In the pair, each CTA
// stages its own 128 rows of A and only HALF of B (128 of the 256 columns); each SM's tensor core reads the other
// half of B from its partner's shared memory, so the per-SM feed drops to 64 B/clk and a k-slab costs 32 KB of shared
// memory instead of 48 KB (6 stages deep instead of 4).
//
// Cluster = (2,1,1).  Roles per CTA (256 threads): warp 0 TMA producer (own A rows, own half of B; both CTAs' loads
// complete on the LEADER's "full" mbarrier), warp 1 MMA issuer (leader CTA only; tcgen05.commit arrives on both CTAs' barriers, so both
// CTAs see "slab free" / "accumulator ready"), warp 2 TMEM allocator (cta_group::2), warps 4-7 epilogue of the CTA's
// own 128 rows (accumulator double-buffered in TMEM: 2 x 256 columns; the partner's epilogue warps arrive remotely on
// the leader's "accumulator free" barrier).  Operand majors as in gemm_tc.cu (both, via descriptor bits).
#include <cooperative_groups.h>

#include "common.cuh"
#include "tc.cuh"

namespace pdb {
namespace tc2 {
using namespace pdb::tc;

constexpr int BM = 128;          // rows per CTA (256 per pair)
constexpr int BN = 256;          // columns per pair
constexpr int BNH = 128;         // columns of B staged by each CTA
constexpr int BK = 64;
constexpr int kThreads = 256;
constexpr int kStages = 6;
constexpr int kABytes = BM * BK * 2;
constexpr int kBBytes = BNH * BK * 2;
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kTmemCols = 2 * BN;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
static int g_group_m = 0;        // pair-tiles (256 rows) per rasterisation group; 0 = pick per shape (see launch)

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t saddr, uint32_t rank) {     // same offset in CTA `rank`
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose completion bytes land on an mbarrier that may live in the partner CTA (cluster address).
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint32_t bar_cluster_addr, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {      // arrives on `bar` in BOTH CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}

__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int group_m, int& m_blk, int& n_blk) {
  const int per_group = group_m * num_n;
  const int group = tile / per_group;
  const int first_m = group * group_m;
  const int gsz = min(num_m - first_m, group_m);
  const int in_group = tile - group * per_group;
  m_blk = first_m + in_group % gsz;
  n_blk = in_group / gsz;
}

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == 1) return 0.5f * v * (1.f + erff(v * 0.70710678118654752f));
  if (act == 2) return fmaxf(v, 0.f);
  return v;
}

template <bool A_MN, bool B_MN, bool OUT_F32>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, void* __restrict__ Cptr,
                const __nv_bfloat16* __restrict__ bias, int M, int N, int K, int64_t ldc, float beta, int act, int group_m) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  uint64_t* full_bar = bars;                       // used on the leader: both CTAs' TMA bytes land here
  uint64_t* empty_bar = bars + kStages;            // each CTA has its own, fed by the two-CTA tcgen05.commit
  uint64_t* tmem_full = bars + 2 * kStages;        // [2] each CTA its own (two-CTA tcgen05.commit)
  uint64_t* tmem_empty = tmem_full + 2;            // [2] used on the leader: 8 epilogue warps of the pair arrive
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int num_m = (M + 2 * BM - 1) / (2 * BM), num_n = (N + BN - 1) / BN;      // in pair-tiles
  const int num_tiles = num_m * num_n;
  const int num_k = (K + BK - 1) / BK;
  const int first_tile = blockIdx.x >> 1, tile_step = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_a);
    prefetch_tmap(&tm_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 8);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();                              // barriers of BOTH CTAs are initialised; TMEM is allocated
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ================================================================= TMA producer (both CTAs)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = first_tile; tile < num_tiles; tile += tile_step) {
        int m_blk, n_blk;
        tile_coords(tile, num_m, num_n, group_m, m_blk, n_blk);
        const int m0 = m_blk * (2 * BM) + (int)rank * BM;        // this CTA's rows
        const int n0 = n_blk * BN + (int)rank * BNH;             // this CTA's half of B
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb = sa + kABytes;
          const uint32_t full_leader = mapa_rank(smem_u32(&full_bar[stage]), 0);
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * kStageBytes);      // both CTAs' bytes
          const int k0 = kb * BK;
          if (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) tma_load_2d_pair(sa + j * (BK * 128), &tm_a, full_leader, m0 + j * 64, k0);
          } else {
            tma_load_2d_pair(sa, &tm_a, full_leader, k0, m0);
          }
          if (B_MN) {
#pragma unroll
            for (int j = 0; j < BNH / 64; ++j) tma_load_2d_pair(sb + j * (BK * 128), &tm_b, full_leader, n0 + j * 64, k0);
          } else {
            tma_load_2d_pair(sb, &tm_b, full_leader, k0, n0);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer (leader CTA only)
    if (rank == 0 && elect_one()) {   // elect.sync, not a lane test: ptxas then knows one lane runs this and issues UTCHMMA / UTCBAR from uniform registers
      constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = first_tile; tile < num_tiles; tile += tile_step) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * kStageBytes);
          const uint32_t sb = sa + kABytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = A_MN ? make_smem_desc(sa + k * 2048, BK * 128, 1024) : make_smem_desc(sa + k * 32, 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc(sb + k * 2048, BK * 128, 1024) : make_smem_desc(sb + k * 32, 16, 1024);
            umma_f16_pair(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[stage]);      // slab reusable in BOTH CTAs once these MMAs retire
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_pair(&tmem_full[acc]);          // accumulator complete, both CTAs' epilogues may read
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // ================================================================= epilogue (own 128 rows)
    const int q = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = first_tile; tile < num_tiles; tile += tile_step) {
      int m_blk, n_blk;
      tile_coords(tile, num_m, num_n, group_m, m_blk, n_blk);
      const int row = m_blk * (2 * BM) + (int)rank * BM + q * 32 + lane;
      const int n0 = n_blk * BN;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c * 32, r);
        tmem_ld_wait();
        if (c == BN / 32 - 1) {
          // all TMEM reads of this accumulator are done: tell the leader's MMA warp (remote arrive from the partner)
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_rank(smem_u32(&tmem_empty[acc]), 0));
        }
        const int col0 = n0 + c * 32;
        if (row < M && col0 < N) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          if (bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < N) v[j] += __bfloat162float(bias[col0 + j]);
          }
          if (act != 0) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = apply_act(v[j], act);
          }
          const bool full = (col0 + 32 <= N);
          if (OUT_F32) {
            float* dst = reinterpret_cast<float*>(Cptr) + (int64_t)row * ldc + col0;
            if (full && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                float4 o = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                if (beta != 0.f) {
                  const float4 pz = *reinterpret_cast<const float4*>(dst + j);
                  o.x += beta * pz.x; o.y += beta * pz.y; o.z += beta * pz.z; o.w += beta * pz.w;
                }
                *reinterpret_cast<float4*>(dst + j) = o;
              }
            } else {
              for (int j = 0; j < 32 && col0 + j < N; ++j) dst[j] = v[j] + (beta != 0.f ? beta * dst[j] : 0.f);
            }
          } else {
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(Cptr) + (int64_t)row * ldc + col0;
            if (full && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                uint4 o;
                __nv_bfloat162 p0 = __floats2bfloat162_rn(v[j], v[j + 1]), p1 = __floats2bfloat162_rn(v[j + 2], v[j + 3]);
                __nv_bfloat162 p2 = __floats2bfloat162_rn(v[j + 4], v[j + 5]), p3 = __floats2bfloat162_rn(v[j + 6], v[j + 7]);
                o.x = *reinterpret_cast<uint32_t*>(&p0); o.y = *reinterpret_cast<uint32_t*>(&p1);
                o.z = *reinterpret_cast<uint32_t*>(&p2); o.w = *reinterpret_cast<uint32_t*>(&p3);
                *reinterpret_cast<uint4*>(dst + j) = o;
              }
            } else {
              for (int j = 0; j < 32 && col0 + j < N; ++j) dst[j] = __float2bfloat16_rn(v[j]);
            }
          }
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();                              // nobody leaves while the partner can still touch its smem / TMEM
  if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
}

template <bool A_MN, bool B_MN, bool OUT_F32>
static int launch(const void* a, const void* b, void* c, const void* bias, int M, int N, int K, float beta, int act,
                  cudaStream_t st) {
  CUtensorMap tm_a, tm_b;
  // Both operands and the output must be this GPU's own memory (get_tmap_2d / require_local refuse anything else).
  if (A_MN) {
    if (get_tmap_2d(&tm_a, a, K, M, (uint64_t)M * 2, BK, 64, 2, true, "gemm_pair A")) return -1;
  } else {
    if (get_tmap_2d(&tm_a, a, M, K, (uint64_t)K * 2, BM, BK, 2, true, "gemm_pair A")) return -1;
  }
  if (B_MN) {
    if (get_tmap_2d(&tm_b, b, K, N, (uint64_t)N * 2, BK, 64, 2, true, "gemm_pair B")) return -1;
  } else {
    if (get_tmap_2d(&tm_b, b, N, K, (uint64_t)K * 2, BNH, BK, 2, true, "gemm_pair B")) return -1;
  }
  if (require_local(c, (size_t)M * N * (OUT_F32 ? 4 : 2), "gemm_pair C")) return -1;
  auto kern = gemm_tc2_kernel<A_MN, B_MN, OUT_F32>;
  PDB_ENSURE_SMEM(kern, kSmemBytes);
  const int num_tiles = ((M + 2 * BM - 1) / (2 * BM)) * ((N + BN - 1) / BN);
  const int pairs = sm_count() / 2;
  const int grid = 2 * (num_tiles < pairs ? num_tiles : pairs);
  // Raster: tiles walk down a band of group_m pair-rows, then step one tile column.  A band's A rows stay in L2 while B
  // streams, so B is re-read from HBM once per band: the taller the band, the less HBM traffic (and power), as long
  // as one wave (pairs tiles) still spans enough columns to keep its B tiles shared.  Default: about sqrt(pairs).
  int group_m = g_group_m > 0 ? g_group_m : 8;
  const int num_m = (M + 2 * BM - 1) / (2 * BM);
  if (group_m > num_m) group_m = num_m;
  kern<<<grid, kThreads, kSmemBytes, st>>>(tm_a, tm_b, c, (const __nv_bfloat16*)bias, M, N, K, (int64_t)N, beta, act, group_m);
  return check_launch("gemm_tc2");
}

}  // namespace tc2
}  // namespace pdb

using namespace pdb;

PDB_EXPORT int pdb200_gemm_set_group_m(int n) {     // tuning knob: raster band height of the CTA-pair kernel (0 = default)
  pdb::tc2::g_group_m = n;
  return 0;
}

// Same contract as pdb200_gemm_bf16 (gemm_tc.cu); 256 x 256 tiles on CTA pairs.
PDB_EXPORT int pdb200_gemm_bf16_pair(const void* a, const void* b, void* c, const void* bias, int64_t M, int64_t N, int64_t K,
                                     int trans_a, int trans_b, int out_dtype, float beta, int act, void* stream) {
  PDB_REQUIRE(M > 0 && N > 0 && K > 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "gemm: bad shape");
  PDB_REQUIRE(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0, "gemm: A and B must be 16-byte aligned");
  PDB_REQUIRE((trans_a ? M : K) % 8 == 0 && (trans_b ? K : N) % 8 == 0,
              "gemm: contiguous dims must be multiples of 8 (TMA 16-byte stride rule)");
  PDB_REQUIRE(out_dtype == 0 || (out_dtype == 1 && beta == 0.f), "gemm: beta != 0 needs fp32 output");
  cudaStream_t st = (cudaStream_t)stream;
  const bool a_mn = trans_a != 0, b_mn = trans_b == 0;
  const int m = (int)M, n = (int)N, k = (int)K;
#define PDB_TC2(AM, BMN)                                                                        \
  return out_dtype == 0 ? tc2::launch<AM, BMN, true>(a, b, c, bias, m, n, k, beta, act, st)     \
                        : tc2::launch<AM, BMN, false>(a, b, c, bias, m, n, k, beta, act, st)
  if (a_mn && b_mn) { PDB_TC2(true, true); }
  if (a_mn && !b_mn) { PDB_TC2(true, false); }
  if (!a_mn && b_mn) { PDB_TC2(false, true); }
  PDB_TC2(false, false);
#undef PDB_TC2
}
