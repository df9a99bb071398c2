// K9 / K10: grouped bf16 GEMMs for the local experts of an MoE layer, scheduled on the device.
//
//   forward / dgrad   C[r, :] = A[r, :] . op(B[e])          for the rows r of expert e   (pdb200_gemm_grouped)
//   wgrad             dW[e]  (=|+=) A_e^T . G_e             reduction over the rows of expert e (pdb200_gemm_grouped_wgrad)
//
// Rows arrive expert-major in one slab (what the dispatch kernel of moe_nvl.cu writes); how many rows each expert got is
// only known on the device (`counts`, int64[E]).  Every CTA rebuilds the same tile list from `counts` in shared memory,
// so one launch covers all local experts and the host never reads the counts.  Replaces the reference's per-expert
// Python loop after a blocking count read (incubate/distributed/models/moe/moe_layer.py:441-470) and the three D2H
// syncs around global_scatter (fluid/operators/collective/global_scatter_op.cu.cc:55-75).
//
// Built on the single-CTA tcgen05 kernel of gemm_tc.cu: warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM
// allocator, warp 3 builds the schedule, warps 4-7 epilogue (row-guarded stores: a 128-row tile may run past the
// expert's last row into the next expert's rows, which are computed and discarded).  The stacked weights [E][K][N] are
// addressed as one 2-D tensor [E*K][N], so no per-expert tensor maps are needed.  All tensor maps cover local memory.
#include "common.cuh"
#include "tc.cuh"

namespace pdb {
namespace grp {
using namespace pdb::tc;

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;
constexpr int kThreads = 256;
constexpr int kStages = 4;
constexpr int kABytes = BM * BK * 2;
constexpr int kBBytes = BN * BK * 2;
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kTmemCols = 2 * BN;
constexpr int kMaxExperts = 64;
constexpr int kMaxSegs = 512;                    // (expert, band) segments of the forward schedule
constexpr int kBand = 8;                         // 128-row tiles per raster band (A band stays in L2 while B streams)
constexpr int kSchedBytes = (kMaxExperts + 1) * 8 + kMaxSegs * 16 + 16;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256 + kSchedBytes;

struct Seg {            // a band of `h` row tiles of one expert; its tiles are numbered n_blk * h + i
  int expert, row0, h, tile0;
};
struct Sched {
  long long off[kMaxExperts + 1];   // first slab row of every expert (prefix sum of counts)
  Seg seg[kMaxSegs];
  int nseg, ntiles;
};

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == 1) return 0.5f * v * (1.f + erff(v * 0.70710678118654752f));
  if (act == 2) return fmaxf(v, 0.f);
  return v;
}

// ------------------------------------------------------------------------------------------------ forward / dgrad
template <bool B_MN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_grouped_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                    __nv_bfloat16* __restrict__ C, const long long* __restrict__ counts, int E, int N, int K,
                    long long rows_cap, int act) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tmem_full = bars + 2 * kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  Sched* S = reinterpret_cast<Sched*>(smem + kStages * kStageBytes + 256);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_n = (N + BN - 1) / BN, num_k = K / BK;

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
      mbar_init(&tmem_empty[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_ptr_smem, kTmemCols);
  if (warp == 3 && lane == 0) {
    // the schedule: every CTA derives the same list from the device-side counts
    long long o = 0;
    int nseg = 0, tiles = 0;
    for (int e = 0; e < E; ++e) {
      S->off[e] = o;
      long long c = counts[e];
      if (c < 0) c = 0;
      if (o + c > rows_cap) c = rows_cap - o > 0 ? rows_cap - o : 0;      // never walk past the slab
      const int mt = (int)((c + BM - 1) / BM);
      for (int b = 0; b < mt && nseg < kMaxSegs; b += kBand) {
        const int h = min(kBand, mt - b);
        S->seg[nseg] = Seg{e, (int)(o + (long long)b * BM), h, tiles};
        tiles += h * num_n;
        ++nseg;
      }
      o += c;
    }
    S->off[E] = o;
    S->nseg = nseg;
    S->ntiles = tiles;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const int ntiles = S->ntiles;

  // tile number -> (segment, row tile inside the band, column block); segments are few, a linear walk is enough
  auto locate = [&](int tile, int& seg_i, int& m_i, int& n_blk) {
    int s = seg_i;
    while (s + 1 < S->nseg && S->seg[s + 1].tile0 <= tile) ++s;
    seg_i = s;
    const int in_seg = tile - S->seg[s].tile0;
    n_blk = in_seg / S->seg[s].h;
    m_i = in_seg - n_blk * S->seg[s].h;
  };

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0, seg_i = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        int m_i, n_blk;
        locate(tile, seg_i, m_i, n_blk);
        const Seg sg = S->seg[seg_i];
        const int m0 = sg.row0 + m_i * BM, n0 = n_blk * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb = sa + kABytes;
          mbar_expect_tx(&full_bar[stage], kStageBytes);
          const int k0 = kb * BK;
          tma_load_2d(sa, &tm_a, &full_bar[stage], k0, m0);
          if (B_MN) {                              // B stored [E][K][N]: rows e*K + k, 64 x 64 sub-tiles
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (BK * 128), &tm_b, &full_bar[stage], n0 + j * 64, sg.expert * K + k0);
          } else {                                 // B stored [E][N][K]: rows e*N + n
            tma_load_2d(sb, &tm_b, &full_bar[stage], k0, sg.expert * N + n0);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {   // elect.sync, not a lane test: ptxas then issues UTCHMMA / UTCBAR from uniform registers
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, 0, B_MN ? 1 : 0);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
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
            const uint64_t da = make_smem_desc(sa + k * 32, 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc(sb + k * 2048, BK * 128, 1024) : make_smem_desc(sb + k * 32, 16, 1024);
            umma_f16(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[acc]);
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    int acc = 0, seg_i = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      int m_i, n_blk;
      locate(tile, seg_i, m_i, n_blk);
      const Seg sg = S->seg[seg_i];
      const long long row_end = S->off[sg.expert + 1];           // rows of the next expert are computed, not stored
      const long long row = (long long)sg.row0 + m_i * BM + q * 32 + lane;
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
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        }
        const int col0 = n0 + c * 32;
        if (row < row_end && col0 < N) {
          __nv_bfloat16* dst = C + row * (long long)N + col0;
          if (col0 + 32 <= N && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              uint32_t p[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                __nv_bfloat162 t = __floats2bfloat162_rn(apply_act(__uint_as_float(r[j + 2 * e]), act),
                                                         apply_act(__uint_as_float(r[j + 2 * e + 1]), act));
                p[e] = *reinterpret_cast<uint32_t*>(&t);
              }
              *reinterpret_cast<uint4*>(dst + j) = make_uint4(p[0], p[1], p[2], p[3]);
            }
          } else {
            for (int j = 0; j < 32 && col0 + j < N; ++j) dst[j] = __float2bfloat16_rn(apply_act(__uint_as_float(r[j]), act));
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
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------------------------------------ wgrad
// dW[e][m, n] (=|+=) sum over the rows r of expert e of A[r, m] * G[r, n].  A: [rows][Kin] slab (the layer input),
// G: [rows][N] slab (the output gradient); both expert-major with the same counts.  Tile = (expert, 128 of Kin,
// BN of N); its reduction runs over ceil(count_e / 64) row slabs that start at the expert's first row.  The last slab
// of an expert reaches into the next expert's rows: the MMA warp zeroes those k-rows of the A tile in shared memory
// before issuing, so they contribute nothing.
template <bool OUT_F32>
__global__ void __launch_bounds__(kThreads, 1)
gemm_grouped_wgrad_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_g, void* __restrict__ Wout,
                          const long long* __restrict__ counts, int E, int Kin, int N, long long rows_cap, float beta) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tmem_full = bars + 2 * kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  Sched* S = reinterpret_cast<Sched*>(smem + kStages * kStageBytes + 256);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_m = (Kin + BM - 1) / BM, num_n = (N + BN - 1) / BN;
  const int per_expert = num_m * num_n, ntiles = E * per_expert;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_a);
    prefetch_tmap(&tm_g);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_ptr_smem, kTmemCols);
  if (warp == 3 && lane == 0) {
    long long o = 0;
    for (int e = 0; e < E; ++e) {
      S->off[e] = o;
      long long c = counts[e];
      if (c < 0) c = 0;
      if (o + c > rows_cap) c = rows_cap - o > 0 ? rows_cap - o : 0;
      o += c;
    }
    S->off[E] = o;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // tile -> (expert, m block of Kin, n block); n fastest so that CTAs running together share the A (input) columns
  auto coords = [&](int tile, int& e, int& m_blk, int& n_blk) {
    e = tile / per_expert;
    const int in_e = tile - e * per_expert;
    m_blk = in_e / num_n;
    n_blk = in_e - m_blk * num_n;
  };

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        int e, m_blk, n_blk;
        coords(tile, e, m_blk, n_blk);
        const long long r0 = S->off[e], cnt = S->off[e + 1] - r0;
        const int nk = (int)((cnt + BK - 1) / BK);
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb = sa + kABytes;
          mbar_expect_tx(&full_bar[stage], kStageBytes);
          const int row = (int)(r0 + (long long)kb * BK);
#pragma unroll
          for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * (BK * 128), &tm_a, &full_bar[stage], m_blk * BM + j * 64, row);
#pragma unroll
          for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (BK * 128), &tm_g, &full_bar[stage], n_blk * BN + j * 64, row);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // the whole warp walks the tiles: all lanes help zeroing the tail rows, lane 0 issues the MMAs
    constexpr uint32_t idesc = make_idesc_bf16(BM, BN, 1, 1);
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      int e, m_blk, n_blk;
      coords(tile, e, m_blk, n_blk);
      const long long cnt = S->off[e + 1] - S->off[e];
      const int nk = (int)((cnt + BK - 1) / BK);
      if (nk == 0) continue;                       // empty expert: the epilogue writes zeros / leaves C alone
      if (lane == 0) mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
      __syncwarp();
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int kb = 0; kb < nk; ++kb) {
        if (lane == 0) mbar_wait(&full_bar[stage], phase);
        __syncwarp();
        tc_fence_after();
        uint8_t* sa_ptr = smem + stage * kStageBytes;
        const int valid = (int)min((long long)BK, cnt - (long long)kb * BK);     // k-rows of this slab that belong to e
        if (valid < BK) {
          // both tiles are sub-blocks of [64 k-rows][128 B] (A: BM/64, G: BN/64, back to back); zero k-rows valid..63 of
          // each (the 128-B swizzle stays inside a row).  Both operands: a stale row may hold NaN, and 0 * NaN != 0.
          const int nzero = (BK - valid) * 8;                                    // 16-byte pieces per sub-block
          for (int j = 0; j < (BM + BN) / 64; ++j) {
            const uint32_t base = smem_u32(sa_ptr + j * (BK * 128) + valid * 128);
            for (int i = lane; i < nzero; i += 32) sts128(base + i * 16, 0u, 0u, 0u, 0u);
          }
          fence_proxy_async();                     // generic-proxy zeros visible to the tensor core's reads
          __syncwarp();
        }
        if (elect_one()) {
          const uint32_t sa = smem_u32(sa_ptr);
          const uint32_t sb = sa + kABytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = make_smem_desc(sa + k * 2048, BK * 128, 1024);
            const uint64_t db = make_smem_desc(sb + k * 2048, BK * 128, 1024);
            umma_f16(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
        }
        __syncwarp();
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (lane == 0) umma_commit(&tmem_full[acc]);
      __syncwarp();
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      int e, m_blk, n_blk;
      coords(tile, e, m_blk, n_blk);
      const long long cnt = S->off[e + 1] - S->off[e];
      const int m = m_blk * BM + q * 32 + lane, n0 = n_blk * BN;
      const long long out_row = ((long long)e * Kin + m) * (long long)N;
      if (cnt == 0) {
        if (beta == 0.f && m < Kin) {              // nothing routed to this expert: its gradient is zero
          for (int j = 0; j < BN && n0 + j < N; ++j) {
            if (OUT_F32) reinterpret_cast<float*>(Wout)[out_row + n0 + j] = 0.f;
            else reinterpret_cast<__nv_bfloat16*>(Wout)[out_row + n0 + j] = __float2bfloat16_rn(0.f);
          }
        }
        continue;
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c * 32, r);
        tmem_ld_wait();
        if (c == BN / 32 - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        }
        const int col0 = n0 + c * 32;
        if (m < Kin && col0 < N) {
          if (OUT_F32) {
            float* dst = reinterpret_cast<float*>(Wout) + out_row + col0;
            if (col0 + 32 <= N && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                float4 o = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
                if (beta != 0.f) {
                  const float4 p = *reinterpret_cast<const float4*>(dst + j);
                  o.x += beta * p.x; o.y += beta * p.y; o.z += beta * p.z; o.w += beta * p.w;
                }
                *reinterpret_cast<float4*>(dst + j) = o;
              }
            } else {
              for (int j = 0; j < 32 && col0 + j < N; ++j) dst[j] = __uint_as_float(r[j]) + (beta != 0.f ? beta * dst[j] : 0.f);
            }
          } else {
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(Wout) + out_row + col0;
            if (col0 + 32 <= N && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                uint32_t p[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                  __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(r[j + 2 * t]), __uint_as_float(r[j + 2 * t + 1]));
                  p[t] = *reinterpret_cast<uint32_t*>(&h);
                }
                *reinterpret_cast<uint4*>(dst + j) = make_uint4(p[0], p[1], p[2], p[3]);
              }
            } else {
              for (int j = 0; j < 32 && col0 + j < N; ++j) dst[j] = __float2bfloat16_rn(__uint_as_float(r[j]));
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
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace grp
}  // namespace pdb

using namespace pdb;

// C[rows_cap, N] (bf16): rows of expert e (expert-major, counts[e] of them, device int64) = A[rows, K] . B[e].
// B: stacked [E][K][N], or [E][N][K] when trans_b.  Rows past sum(counts) are left untouched.  act: 0 none, 1 gelu, 2 relu.
PDB_EXPORT int pdb200_gemm_grouped(const void* a, const void* b, void* c, const int64_t* counts, int E, int64_t rows_cap,
                                   int64_t N, int64_t K, int trans_b, int act, void* stream) {
  using namespace pdb::grp;
  PDB_REQUIRE(E >= 1 && E <= kMaxExperts, "gemm_grouped: 1..%d experts per launch (got %d)", kMaxExperts, E);
  PDB_REQUIRE(rows_cap > 0 && N > 0 && K > 0 && rows_cap < (1ll << 31) && E * K < (1ll << 31) && E * N < (1ll << 31), "gemm_grouped: bad shape");
  PDB_REQUIRE(K % BK == 0 && N % 8 == 0, "gemm_grouped: K must be a multiple of %d and N of 8 (got K=%lld N=%lld)", BK, (long long)K, (long long)N);
  PDB_REQUIRE(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0 && ((uintptr_t)c & 15) == 0, "gemm_grouped: 16-byte alignment");
  CUtensorMap tm_a, tm_b;
  if (tc::get_tmap_2d(&tm_a, a, rows_cap, K, (uint64_t)K * 2, BM, BK, 2, true, "gemm_grouped A")) return -1;
  if (trans_b) {
    if (tc::get_tmap_2d(&tm_b, b, (uint64_t)E * N, K, (uint64_t)K * 2, BN, BK, 2, true, "gemm_grouped B")) return -1;
  } else {
    if (tc::get_tmap_2d(&tm_b, b, (uint64_t)E * K, N, (uint64_t)N * 2, BK, 64, 2, true, "gemm_grouped B")) return -1;
  }
  if (require_local(c, (size_t)rows_cap * N * 2, "gemm_grouped C")) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = sm_count();
  if (trans_b) {
    auto kern = gemm_grouped_kernel<false>;
    PDB_ENSURE_SMEM(kern, kSmemBytes);
    kern<<<grid, kThreads, kSmemBytes, st>>>(tm_a, tm_b, (__nv_bfloat16*)c, (const long long*)counts, E, (int)N, (int)K, rows_cap, act);
  } else {
    auto kern = gemm_grouped_kernel<true>;
    PDB_ENSURE_SMEM(kern, kSmemBytes);
    kern<<<grid, kThreads, kSmemBytes, st>>>(tm_a, tm_b, (__nv_bfloat16*)c, (const long long*)counts, E, (int)N, (int)K, rows_cap, act);
  }
  return check_launch("gemm_grouped");
}

// dW[E][Kin][N] (out_dtype 0 = fp32, beta 1 accumulates; 1 = bf16) = per-expert A_e^T . G_e, reduction over the rows of
// expert e.  A: [rows_cap][Kin], G: [rows_cap][N], both bf16 expert-major slabs.
PDB_EXPORT int pdb200_gemm_grouped_wgrad(const void* a, const void* g, void* dw, const int64_t* counts, int E, int64_t rows_cap,
                                         int64_t Kin, int64_t N, int out_dtype, float beta, void* stream) {
  using namespace pdb::grp;
  PDB_REQUIRE(E >= 1 && E <= kMaxExperts, "gemm_grouped_wgrad: 1..%d experts per launch (got %d)", kMaxExperts, E);
  PDB_REQUIRE(rows_cap > 0 && N > 0 && Kin > 0 && rows_cap < (1ll << 31), "gemm_grouped_wgrad: bad shape");
  PDB_REQUIRE(Kin % 8 == 0 && N % 8 == 0, "gemm_grouped_wgrad: Kin and N must be multiples of 8");
  PDB_REQUIRE(out_dtype == 0 || (out_dtype == 1 && beta == 0.f), "gemm_grouped_wgrad: beta != 0 needs fp32 output");
  PDB_REQUIRE(((uintptr_t)a & 15) == 0 && ((uintptr_t)g & 15) == 0 && ((uintptr_t)dw & 15) == 0, "gemm_grouped_wgrad: 16-byte alignment");
  CUtensorMap tm_a, tm_g;
  if (tc::get_tmap_2d(&tm_a, a, rows_cap, Kin, (uint64_t)Kin * 2, BK, 64, 2, true, "gemm_grouped_wgrad A")) return -1;
  if (tc::get_tmap_2d(&tm_g, g, rows_cap, N, (uint64_t)N * 2, BK, 64, 2, true, "gemm_grouped_wgrad G")) return -1;
  if (require_local(dw, (size_t)E * Kin * N * (out_dtype == 0 ? 4 : 2), "gemm_grouped_wgrad dW")) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = sm_count();
  if (out_dtype == 0) {
    auto kern = gemm_grouped_wgrad_kernel<true>;
    PDB_ENSURE_SMEM(kern, kSmemBytes);
    kern<<<grid, kThreads, kSmemBytes, st>>>(tm_a, tm_g, dw, (const long long*)counts, E, (int)Kin, (int)N, rows_cap, beta);
  } else {
    auto kern = gemm_grouped_wgrad_kernel<false>;
    PDB_ENSURE_SMEM(kern, kSmemBytes);
    kern<<<grid, kThreads, kSmemBytes, st>>>(tm_a, tm_g, dw, (const long long*)counts, E, (int)Kin, (int)N, rows_cap, beta);
  }
  return check_launch("gemm_grouped_wgrad");
}
