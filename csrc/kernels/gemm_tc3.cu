// K23 / K13, macro-tile variant of the CTA-pair GEMM: one pair of SMs computes a 256 x 512 output block as two
// 256 x 256 tcgen05.mma.cta_group::2 accumulators that share the A operand, and writes C with TMA stores.
//
// Why (measured against the 256 x 256 kernel of gemm_tc2.cu and the library kernel on the same shapes, B200 at its
// 1 kW power cap, where throughput is set by energy per flop):
//   * A is staged once for two accumulators: shared-memory fill traffic (L2 -> SM) drops from 64 KB to 48 KB per
//     256 x 256 x 64 block of work (-25 %).
//   * the epilogue goes TMEM -> registers -> swizzled shared memory -> cp.async.bulk.tensor store: whole 128-byte
//     lines leave the SM instead of 16-byte pieces with a row stride between lanes (4x fewer L2 write transactions,
//     10x fewer LSU wavefronts); fp32 accumulate-into-C (K13, beta = 1) is a TMA reduce-add, C is never read.
//   * waits are single-lane with back-off instead of 128 threads spinning on try_wait.
// The accumulators take all 512 TMEM columns, so the epilogue of block i is no longer fully hidden behind block i+1;
// each 256-column accumulator has its own "drained" barrier so the next block's first MMAs start as soon as the first
// half is out.
//
// Work list per pair (static, identical in every role): full rounds of macro blocks in band-raster order, then the
// leftover blocks split into 256-column halves so the last round costs at most half a block.
//
// Roles per CTA (256 threads): warp 0 TMA producer (own 128 rows of A, own 128-column half of each B tile; bytes of
// both CTAs complete on the LEADER's "full" barrier), warp 1 MMA issuer (leader only; tcgen05.commit arrives on both
// CTAs' barriers), warp 2 TMEM allocator, warps 4-7 epilogue of the CTA's own 128 rows.
// Every tensor map (A, B and C) covers this GPU's own memory only (get_tmap_2d refuses anything else).
#include <cooperative_groups.h>
#include <stdio.h>

#include "common.cuh"
#include "tc.cuh"

namespace pdb {
namespace tc3 {
using namespace pdb::tc;

constexpr int BM = 128;          // rows per CTA (256 per pair)
constexpr int BN = 256;          // columns per accumulator
constexpr int BNH = 128;         // columns of each B tile staged by one CTA
constexpr int BK = 64;
constexpr int kThreads = 256;
constexpr int kStages = 4;
constexpr int kABytes = BM * BK * 2;                 // 16 KB
constexpr int kBBytes = BNH * BK * 2;                // 16 KB per accumulator
constexpr int kStageBytes = kABytes + 2 * kBBytes;   // 48 KB
constexpr int kTmemCols = 2 * BN;
constexpr int kStoreBufBytes = 4096;                 // 32 rows x 128 B per epilogue warp, double-buffered
constexpr int kStoreBytes = 4 * 2 * kStoreBufBytes;  // 32 KB
constexpr int kSmemBytes = kStages * kStageBytes + kStoreBytes + 1024 + 256;

static int g_band = 0;           // pair-rows (256 rows) per raster band; 0 = default

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t saddr, uint32_t rank) {
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
// Wait with back-off: try_wait already suspends for a hardware time slice; a short sleep between polls keeps the
// waiting lane out of the issue slots.
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (done) break;
    __nanosleep(64);
  }
}
// TMA load (local memory) whose completion bytes land on an mbarrier that may live in the partner CTA.
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint32_t bar_cluster_addr, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
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

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == 1) return 0.5f * v * (1.f + erff(v * 0.70710678118654752f));
  if (act == 2) return fmaxf(v, 0.f);
  return v;
}

// bias + activation on 32 accumulator columns (in place, fp32 bit patterns); kept out of line so the common
// no-epilogue drain loop stays small
__device__ __noinline__ void epilogue_math(uint32_t (&r)[32], const __nv_bfloat16* __restrict__ bias, int act, int col, int N) {
#pragma unroll 4
  for (int j = 0; j < 32; ++j) {
    float v = __uint_as_float(r[j]);
    if (bias != nullptr && col + j < N) v += __bfloat162float(bias[col + j]);
    r[j] = __float_as_uint(apply_act(v, act));
  }
}

// K5 gate: spin (bounded) until the chunk holding pair-row `row0` has landed in local memory, then order the TMA
// (async proxy) reads after the flag's acquire.
struct Shape;
__device__ __forceinline__ void gate_wait_impl(const uint32_t* flags, int gate_rows, uint32_t epoch, int row0) {
  const uint32_t* f = flags + row0 / gate_rows;
  unsigned long long spins = 0;
  while (true) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
    if ((int32_t)(v - epoch) >= 0) break;
    if (++spins > 400000000ull) {
      printf("[pdb200 gemm gate] chunk %d never arrived (flag %u, want %u)\n", row0 / gate_rows, v, epoch);
      __trap();
    }
    __nanosleep(100);
  }
  asm volatile("fence.proxy.async.global;" ::: "memory");
}

// ---- work list -------------------------------------------------------------------------------------------------
struct Shape {
  int M, N, K;
  int num_m;        // pair-rows (256 rows)
  int num_c;        // macro columns (512 wide)
  int band;         // pair-rows per raster band
  int full;         // macro blocks handled in full rounds: (num_m * num_c / pairs) * pairs
  int total;        // num_m * num_c
  // Optional gate on the A operand (K5 all-gather -> GEMM): rows [i * gate_rows, (i + 1) * gate_rows) of A may only be
  // loaded once gate_flags[i] has reached gate_epoch.  The flags live in THIS GPU's memory; peers set them after
  // their copy of that row chunk into this GPU's gathered buffer has landed.  m_rot rotates the pair-row order so
  // that the rows this rank contributed itself come first.
  const uint32_t* gate_flags;
  int gate_rows;
  uint32_t gate_epoch;
  int m_rot;
  // Optional completion counters (K6 GEMM -> reduce-scatter): after the stores of a work item have completed, every
  // epilogue warp adds the item's number of 256-column tiles to done_counters[row / done_rows] (device scope).  A row
  // block of done_rows rows is complete when its counter has grown by (done_rows / 256) * ceil(N / 256) * 8.
  unsigned int* done_counters;
  int done_rows;
};
struct Item {
  int m_blk, n_base, nslots;
};
// Block index -> (pair-row, macro column): walk down a band of `band` pair-rows, then step one macro column.
__device__ __forceinline__ void block_coords(const Shape& s, int t, int& m_blk, int& c_blk) {
  const int per_band = s.band * s.num_c;
  const int b = t / per_band;
  const int first_m = b * s.band;
  const int h = min(s.num_m - first_m, s.band);
  const int in_band = t - b * per_band;
  m_blk = first_m + in_band % h;
  c_blk = in_band / h;
}
__device__ __forceinline__ void gate_wait(const Shape& s, int row0) { gate_wait_impl(s.gate_flags, s.gate_rows, s.gate_epoch, row0); }

// Work item number w of this pair's list (w = pair, pair + pairs, ...); false when the list is exhausted.
__device__ __forceinline__ bool get_item(const Shape& s, int w, Item& it) {
  int t, half = -1;
  if (w < s.full) {
    t = w;
  } else {
    const int u = w - s.full;
    t = s.full + (u >> 1);
    half = u & 1;
    if (t >= s.total) return false;
  }
  int c_blk;
  block_coords(s, t, it.m_blk, c_blk);
  if (s.m_rot) it.m_blk = (it.m_blk + s.m_rot) % s.num_m;
  it.n_base = c_blk * (2 * BN);
  it.nslots = 2;
  if (half >= 0) {
    it.n_base += half * BN;
    it.nslots = 1;
  }
  if (it.n_base + BN >= s.N && it.nslots == 2) it.nslots = 1;      // right half entirely past the last column
  if (it.n_base >= s.N) it.nslots = 0;                             // nothing to do (kept so every role counts alike)
  return true;
}

template <bool A_MN, bool B_MN, bool OUT_F32>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_tc3_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                const __grid_constant__ CUtensorMap tm_c, const __nv_bfloat16* __restrict__ bias, const Shape shape,
                int accumulate_c, int act) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* store_smem = smem + kStages * kStageBytes;                     // 1024-byte aligned (48 KB stages)
  uint64_t* bars = reinterpret_cast<uint64_t*>(store_smem + kStoreBytes);
  uint64_t* full_bar = bars;                       // used on the leader: both CTAs' TMA bytes land here
  uint64_t* empty_bar = bars + kStages;            // each CTA its own, fed by the two-CTA tcgen05.commit
  uint64_t* tmem_full = bars + 2 * kStages;        // [2] per accumulator, each CTA its own
  uint64_t* tmem_empty = tmem_full + 2;            // [2] per accumulator, used on the leader: 8 epilogue warps arrive
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int num_k = (shape.K + BK - 1) / BK;
  const int pair = blockIdx.x >> 1, pairs = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_a);
    prefetch_tmap(&tm_b);
    prefetch_tmap(&tm_c);
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
      Item it;
      for (int w = pair; get_item(shape, w, it); w += pairs) {
        if (it.nslots == 0) continue;
        const int m0 = it.m_blk * (2 * BM) + (int)rank * BM;        // this CTA's rows
        if (shape.gate_flags != nullptr) gate_wait(shape, it.m_blk * (2 * BM));
        const uint32_t tx = 2u * (uint32_t)(kABytes + it.nslots * kBBytes);
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          const uint32_t full_leader = mapa_rank(smem_u32(&full_bar[stage]), 0);
          if (rank == 0) mbar_expect_tx(&full_bar[stage], tx);      // both CTAs' bytes
          const int k0 = kb * BK;
          if (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) tma_load_2d_pair(sa + j * (BK * 128), &tm_a, full_leader, m0 + j * 64, k0);
          } else {
            tma_load_2d_pair(sa, &tm_a, full_leader, k0, m0);
          }
          for (int s = 0; s < it.nslots; ++s) {
            uint8_t* sb = sa + kABytes + s * kBBytes;
            const int n0 = it.n_base + s * BN + (int)rank * BNH;    // this CTA's half of accumulator s's B tile
            if (B_MN) {
#pragma unroll
              for (int j = 0; j < BNH / 64; ++j) tma_load_2d_pair(sb + j * (BK * 128), &tm_b, full_leader, n0 + j * 64, k0);
            } else {
              tma_load_2d_pair(sb, &tm_b, full_leader, k0, n0);
            }
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
      uint32_t uses0 = 0, uses1 = 0;               // how often each accumulator has been filled so far
      int singles = 0;
      Item it;
      // One 64-deep slab of one accumulator: 4 MMAs (K = 16 each) reading A and this accumulator's B tile from `st`.
      auto issue = [&](int st, int slot, int bsel, bool first_slab) {
        const uint32_t sa = smem_u32(smem + st * kStageBytes);
        const uint32_t sb = sa + kABytes + bsel * kBBytes;
        const uint32_t tmem_d = tmem_base + slot * BN;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t da = A_MN ? make_smem_desc(sa + k * 2048, BK * 128, 1024) : make_smem_desc(sa + k * 32, 16, 1024);
          const uint64_t db = B_MN ? make_smem_desc(sb + k * 2048, BK * 128, 1024) : make_smem_desc(sb + k * 32, 16, 1024);
          umma_f16_pair(tmem_d, da, db, idesc, (first_slab && k == 0) ? 0u : 1u);
        }
      };
      for (int w = pair; get_item(shape, w, it); w += pairs) {
        if (it.nslots == 0) continue;
        if (it.nslots == 1) {
          const int slot = singles++ & 1;
          mbar_wait(&tmem_empty[slot], ((slot ? uses1 : uses0) & 1) ^ 1);      // the epilogue has drained it
          tc_fence_after();
          for (int kb = 0; kb < num_k; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            issue(stage, slot, 0, kb == 0);
            umma_commit_pair(&empty_bar[stage]);   // slab reusable in BOTH CTAs once these MMAs retire
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
          umma_commit_pair(&tmem_full[slot]);
          if (slot) ++uses1; else ++uses0;
          continue;
        }
        // Two accumulators, skewed by one slab: step j issues accumulator 0 on slab j and accumulator 1 on slab j-1.
        // Accumulator 0 therefore completes one slab early and is drained while accumulator 1 finishes, and at the
        // start of the next block accumulator 0 runs its first slab while accumulator 1 is still being drained: of
        // the two drains (TMEM reads at 64 B/clk, ~2 K cycles each) only part of one is left exposed.
        int prev_stage = 0;
        for (int j = 0; j <= num_k; ++j) {
          if (j < num_k) {
            if (j == 0) {
              mbar_wait(&tmem_empty[0], (uses0 & 1) ^ 1);
              tc_fence_after();
            }
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            issue(stage, 0, 0, j == 0);
            if (j == num_k - 1) {
              umma_commit_pair(&tmem_full[0]);     // accumulator 0 complete
              ++uses0;
            }
          }
          if (j >= 1) {
            if (j == 1) {
              mbar_wait(&tmem_empty[1], (uses1 & 1) ^ 1);
              tc_fence_after();
            }
            issue(prev_stage, 1, 1, j == 1);
            umma_commit_pair(&empty_bar[prev_stage]);   // both accumulators are done with that slab
          }
          if (j < num_k) {
            prev_stage = stage;
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        umma_commit_pair(&tmem_full[1]);           // accumulator 1 complete
        ++uses1;
      }
    }
  } else if (warp >= 4) {
    // ================================================================= epilogue (own 128 rows)
    const int q = warp & 3;                         // TMEM lane quarter this warp may touch
    uint8_t* my_store = store_smem + q * (2 * kStoreBufBytes);
    uint32_t uses0 = 0, uses1 = 0;
    int singles = 0;
    uint32_t nstore = 0;                            // TMA stores issued by this warp so far (buffer = nstore & 1)
    constexpr int CW = OUT_F32 ? 32 : 64;           // columns per store box (128 bytes per row)
    Item it;
    for (int w = pair; get_item(shape, w, it); w += pairs) {
      if (it.nslots == 0) continue;
      const int slot0 = it.nslots == 2 ? 0 : (singles++ & 1);
      const int row0 = it.m_blk * (2 * BM) + (int)rank * BM + q * 32;          // first row of this warp
      for (int s = 0; s < it.nslots; ++s) {
        const int slot = slot0 + s;
        if (lane == 0) mbar_wait_relaxed(&tmem_full[slot], (slot ? uses1 : uses0) & 1);
        __syncwarp();
        if (slot) ++uses1; else ++uses0;
        tc_fence_after();
        const uint32_t taddr = tmem_base + slot * BN + ((uint32_t)(q * 32) << 16);
        // Software pipeline over 32-column steps: the tcgen05.ld of step i+1 is in flight while step i is converted
        // and staged, so the accumulator is drained at TMEM speed rather than at one load latency per step.  The loop
        // is rolled (two steps per trip, one per register set): unrolled it was ~8 K instructions per accumulator and
        // the instruction cache, not the TMEM, set the drain time.
        constexpr int SPB = CW / 32;                 // steps per store box
        constexpr int NSTEP = BN / 32;
        const bool has_epi = bias != nullptr || act != 0;
        uint32_t r[2][32];
        tmem_ld_32x32(taddr, r[0]);
#pragma unroll 1
        for (int i2 = 0; i2 < NSTEP; i2 += 2) {
#pragma unroll
          for (int par = 0; par < 2; ++par) {
            const int i = i2 + par;
            const int c = i / SPB, h = i % SPB;
            const int col0 = it.n_base + s * BN + c * CW;
            const bool inside = row0 < shape.M && col0 < shape.N;     // warp-uniform: some of the box lies in C
            uint8_t* buf = my_store + (nstore & 1) * kStoreBufBytes;
            const uint32_t row_addr = smem_u32(buf) + lane * 128;
            const uint32_t sw = lane & 7;
            tmem_ld_wait();                                          // step i is in r[par]
            if (i + 1 < NSTEP) {
              tmem_ld_32x32(taddr + (i + 1) * 32, r[par ^ 1]);
            } else {
              // every TMEM read of this accumulator is done: hand it back to the MMA warp (remote arrive from the partner)
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive_cluster(mapa_rank(smem_u32(&tmem_empty[slot]), 0));
            }
            if (!inside) continue;
            if (h == 0) {
              // the buffer we are about to fill was the source of the store issued two boxes ago
              if (lane == 0) tma_store_wait_read<1>();
              __syncwarp();
            }
            if (has_epi) {                              // slow path: through a local array, the fast path keeps r in registers
              uint32_t t[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) t[j] = r[par][j];
              epilogue_math(t, bias, act, col0 + h * 32, shape.N);
#pragma unroll
              for (int j = 0; j < 32; ++j) r[par][j] = t[j];
            }
            if (OUT_F32) {
#pragma unroll
              for (int ch = 0; ch < 8; ++ch)
                sts128(row_addr + ((ch ^ sw) << 4), r[par][ch * 4], r[par][ch * 4 + 1], r[par][ch * 4 + 2], r[par][ch * 4 + 3]);
            } else {
#pragma unroll
              for (int ch = 0; ch < 4; ++ch) {
                uint32_t p[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  __nv_bfloat162 t = __floats2bfloat162_rn(__uint_as_float(r[par][ch * 8 + 2 * e]), __uint_as_float(r[par][ch * 8 + 2 * e + 1]));
                  p[e] = *reinterpret_cast<uint32_t*>(&t);
                }
                sts128(row_addr + (((h * 4 + ch) ^ sw) << 4), p[0], p[1], p[2], p[3]);
              }
            }
            if (h == SPB - 1) {
              fence_proxy_async();                   // generic-proxy writes visible to the TMA (async proxy)
              __syncwarp();
              if (lane == 0) {
                if (OUT_F32 && accumulate_c) tma_reduce_add_2d(&tm_c, buf, col0, row0);
                else tma_store_2d(&tm_c, buf, col0, row0);
                tma_store_commit();
              }
              ++nstore;
            }
          }
        }
      }
      if (shape.done_counters != nullptr) {
        // this warp's part of the item is in global memory once its bulk stores have completed: count it
        if (lane == 0) {
          tma_store_wait<0>();
          asm volatile("fence.proxy.async.global;" ::: "memory");
          __threadfence();
          atomicAdd(shape.done_counters + (it.m_blk * (2 * BM)) / shape.done_rows, (unsigned int)it.nslots);
        }
        __syncwarp();
      }
    }
    if (lane == 0) tma_store_wait<0>();               // all of this warp's stores have completed
    __syncwarp();
  }

  tc_fence_before();
  cluster_sync_all();                              // nobody leaves while the partner can still touch its smem / TMEM
  if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
}

struct Gate {
  const uint32_t* flags = nullptr;
  int rows = 0;
  uint32_t epoch = 0;
  int m_rot = 0;
  unsigned int* done = nullptr;     // completion counters, one per done_rows rows
  int done_rows = 0;
  int band = 0;                     // raster band override (pair-rows), 0 = default
};

template <bool A_MN, bool B_MN, bool OUT_F32>
static int launch(const void* a, const void* b, void* c, const void* bias, int M, int N, int K, float beta, int act,
                  cudaStream_t st, const Gate& gate = Gate()) {
  CUtensorMap tm_a, tm_b, tm_c;
  // All three tensor maps are built through get_tmap_2d: local memory of the current device or an error.
  if (A_MN) {
    if (get_tmap_2d(&tm_a, a, K, M, (uint64_t)M * 2, BK, 64, 2, true, "gemm_macro A")) return -1;
  } else {
    if (get_tmap_2d(&tm_a, a, M, K, (uint64_t)K * 2, BM, BK, 2, true, "gemm_macro A")) return -1;
  }
  if (B_MN) {
    if (get_tmap_2d(&tm_b, b, K, N, (uint64_t)N * 2, BK, 64, 2, true, "gemm_macro B")) return -1;
  } else {
    if (get_tmap_2d(&tm_b, b, N, K, (uint64_t)K * 2, BNH, BK, 2, true, "gemm_macro B")) return -1;
  }
  if (OUT_F32) {
    if (get_tmap_2d(&tm_c, c, M, N, (uint64_t)N * 4, 32, 32, 4, true, "gemm_macro C")) return -1;
  } else {
    if (get_tmap_2d(&tm_c, c, M, N, (uint64_t)N * 2, 32, 64, 2, true, "gemm_macro C")) return -1;
  }
  auto kern = gemm_tc3_kernel<A_MN, B_MN, OUT_F32>;
  PDB_ENSURE_SMEM(kern, kSmemBytes);
  Shape s{};
  s.M = M; s.N = N; s.K = K;
  s.num_m = (M + 2 * BM - 1) / (2 * BM);
  s.num_c = (N + 2 * BN - 1) / (2 * BN);
  s.total = s.num_m * s.num_c;
  const int pairs_hw = sm_count() / 2;
  // the leftover round works on half blocks, so it pays to use every pair as soon as there are half as many blocks
  const int pairs = 2 * s.total < pairs_hw ? (2 * s.total > 0 ? 2 * s.total : 1) : pairs_hw;
  s.full = (s.total / pairs) * pairs;
  s.band = g_band > 0 ? g_band : 16;
  if (s.band > s.num_m) s.band = s.num_m;
  if (gate.band > 0) s.band = gate.band < s.num_m ? gate.band : s.num_m;
  s.done_counters = gate.done;
  s.done_rows = gate.done_rows;
  if (gate.done && require_local(gate.done, sizeof(unsigned int) * (size_t)((M + gate.done_rows - 1) / gate.done_rows), "gemm done counters")) return -1;
  s.gate_flags = gate.flags;
  s.gate_rows = gate.rows;
  s.gate_epoch = gate.epoch;
  s.m_rot = gate.m_rot;
  if (gate.flags && require_local(gate.flags, sizeof(uint32_t) * (size_t)((M + gate.rows - 1) / gate.rows), "gemm gate flags")) return -1;
  kern<<<2 * pairs, kThreads, kSmemBytes, st>>>(tm_a, tm_b, tm_c, (const __nv_bfloat16*)bias, s, beta != 0.f ? 1 : 0, act);
  return check_launch("gemm_tc3");
}

}  // namespace tc3
}  // namespace pdb

using namespace pdb;

PDB_EXPORT int pdb200_gemm_set_band(int n) {        // tuning knob: raster band height of the macro-tile kernel (0 = default)
  pdb::tc3::g_band = n;
  return 0;
}

// Same contract as pdb200_gemm_bf16 (gemm_tc.cu) except beta must be 0 or 1; 256 x 512 blocks on CTA pairs, TMA-store epilogue.
PDB_EXPORT int pdb200_gemm_bf16_macro(const void* a, const void* b, void* c, const void* bias, int64_t M, int64_t N, int64_t K,
                                      int trans_a, int trans_b, int out_dtype, float beta, int act, void* stream) {
  PDB_REQUIRE(M > 0 && N > 0 && K > 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "gemm: bad shape");
  PDB_REQUIRE(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0 && ((uintptr_t)c & 15) == 0, "gemm: A, B and C must be 16-byte aligned");
  PDB_REQUIRE((trans_a ? M : K) % 8 == 0 && (trans_b ? K : N) % 8 == 0 && N % 8 == 0,
              "gemm: contiguous dims must be multiples of 8 (TMA 16-byte stride rule)");
  PDB_REQUIRE(out_dtype == 0 || (out_dtype == 1 && beta == 0.f), "gemm: beta != 0 needs fp32 output");
  PDB_REQUIRE(beta == 0.f || beta == 1.f, "gemm_macro: beta must be 0 or 1 (C += is a TMA reduce-add)");
  cudaStream_t st = (cudaStream_t)stream;
  const bool a_mn = trans_a != 0, b_mn = trans_b == 0;
  const int m = (int)M, n = (int)N, k = (int)K;
#define PDB_TC3(AM, BMN)                                                                        \
  return out_dtype == 0 ? tc3::launch<AM, BMN, true>(a, b, c, bias, m, n, k, beta, act, st)     \
                        : tc3::launch<AM, BMN, false>(a, b, c, bias, m, n, k, beta, act, st)
  if (a_mn && b_mn) { PDB_TC3(true, true); }
  if (a_mn && !b_mn) { PDB_TC3(true, false); }
  if (!a_mn && b_mn) { PDB_TC3(false, true); }
  PDB_TC3(false, false);
#undef PDB_TC3
}

// K5 all-gather -> GEMM, compute half: C = A . op(B) where A [M, K] is the gathered activation in LOCAL memory that
// peers fill chunk by chunk; rows [i * gate_rows, (i+1) * gate_rows) are loaded only after gate_flags[i] (local uint32
// array) has reached gate_epoch, and the row blocks are visited starting at pair-row m_rot (the caller passes the first
// block of this rank's own shard, which is there from the start).  gate_rows must be a multiple of 256.  Reference
// sequence: ncclAllGather, then GEMM (fleet/utils/sequence_parallel_utils.py:255-285).
PDB_EXPORT int pdb200_gemm_bf16_gated(const void* a, const void* b, void* c, const void* bias, int64_t M, int64_t N, int64_t K,
                                      int trans_b, int act, const void* gate_flags, int gate_rows, int64_t gate_epoch,
                                      int m_rot, void* stream) {
  PDB_REQUIRE(M > 0 && N > 0 && K > 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "gemm_gated: bad shape");
  PDB_REQUIRE(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0 && ((uintptr_t)c & 15) == 0, "gemm_gated: 16-byte alignment");
  PDB_REQUIRE(K % 8 == 0 && N % 8 == 0, "gemm_gated: K and N must be multiples of 8");
  PDB_REQUIRE(gate_flags != nullptr && gate_rows > 0 && gate_rows % 256 == 0, "gemm_gated: gate_rows must be a positive multiple of 256");
  tc3::Gate g;
  g.flags = (const uint32_t*)gate_flags;
  g.rows = gate_rows;
  g.epoch = (uint32_t)gate_epoch;
  g.m_rot = m_rot;
  cudaStream_t st = (cudaStream_t)stream;
  if (trans_b) return tc3::launch<false, false, false>(a, b, c, bias, (int)M, (int)N, (int)K, 0.f, act, st, g);
  return tc3::launch<false, true, false>(a, b, c, bias, (int)M, (int)N, (int)K, 0.f, act, st, g);
}

// K6 GEMM -> reduce-scatter, compute half: C = A . op(B) (bf16, local memory) visiting the row blocks of block_rows rows
// in the order m_rot, m_rot+1, ... (the caller starts one block after its own, so its own block comes last) and
// counting finished work into done_counters[row / block_rows] (see Shape::done_counters).  block_rows % 256 == 0.
PDB_EXPORT int pdb200_gemm_bf16_counted(const void* a, const void* b, void* c, int64_t M, int64_t N, int64_t K, int trans_b,
                                        void* done_counters, int block_rows, int m_rot, void* stream) {
  PDB_REQUIRE(M > 0 && N > 0 && K > 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "gemm_counted: bad shape");
  PDB_REQUIRE(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0 && ((uintptr_t)c & 15) == 0, "gemm_counted: 16-byte alignment");
  PDB_REQUIRE(K % 8 == 0 && N % 8 == 0, "gemm_counted: K and N must be multiples of 8");
  PDB_REQUIRE(done_counters != nullptr && block_rows > 0 && block_rows % 256 == 0 && M % block_rows == 0,
              "gemm_counted: block_rows must be a positive multiple of 256 that divides M");
  tc3::Gate g;
  g.done = (unsigned int*)done_counters;
  g.done_rows = block_rows;
  g.m_rot = m_rot;
  g.band = block_rows / 256;                 // finish one owner's block before starting the next
  cudaStream_t st = (cudaStream_t)stream;
  if (trans_b) return tc3::launch<false, false, false>(a, b, c, nullptr, (int)M, (int)N, (int)K, 0.f, 0, st, g);
  return tc3::launch<false, true, false>(a, b, c, nullptr, (int)M, (int)N, (int)K, 0.f, 0, st, g);
}
