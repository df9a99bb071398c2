// K16: flash-attention forward on tcgen05 tensor cores for sm_100a (bf16 / fp16, head_dim 64/128, causal or full, GQA,
// dense [b, s, h, d] batches of any length and packed variable-length batches with cu_seqlens, dropout).
//
// Replaces in the reference: paddle/phi/kernels/gpu/flash_attn_kernel.cu:36-170 (unpadded), :305-447 (dense + dropout)
// -> vendored FlashAttention-2 (mma.sync m16n8k16 + cp.async, third_party/flashattn/csrc/flash_attn/src/flash_fwd_kernel.h).
//
// One CTA per (batch, q-head, 128-row query tile); K tiles of 128 keys stream through a 3-stage TMA ring, V tiles
// through a 2-stage ring.  What shapes the kernel (clock-stamp timelines in profiles/flash_attn_timeline.md):
//  * the softmax of a tile costs more than its two MMAs (one exp2 per score on 16 SFU lanes/clk/SM alone is as long),
//    so the softmax warps must never wait for scores.  There are THREE score buffers in TMEM: Q K^T runs up to three
//    key tiles ahead of the softmax, and all eight softmax warps work on the same tile, half a row per thread, then
//    move straight to the next one.
//  * shared-memory bandwidth.  A 128x128x16 tcgen05.mma with both operands in shared memory reads 8 KB per 64 clocks,
//    which is the whole 128 B/clk of the SM.  P therefore never goes through shared memory: each thread writes its
//    half row of P with tcgen05.st over the score columns it has just read, and P V takes its A operand from TMEM.
//   warp 0      TMA producer for Q and the K ring            warp 3   TMA producer for the V ring
//   warp 1      MMA issuer: S_j = Q K_j^T into score buffer j % 3; O += P_j V_j once P_j is in TMEM
//   warp 2      TMEM allocator (S_0 | S_1 | S_2 | O = 512 columns)
//   warps 4-11  softmax: lane quarter = warp % 4, column half = (warp - 4) / 4; a thread owns 64 keys of one query row
// The two threads of a row must use the same reference maximum.  They do not exchange it per tile: the probabilities
// are computed against a stale shared maximum, which is exact as long as no score exceeds it by more than 2^8, and one
// barrier-with-OR over the 256 softmax threads per tile tells whether any row outgrew that bound.  Only then (and on the
// first tile) are the half-row maxima exchanged through shared memory, the accumulator rescaled in TMEM and the row
// sums corrected.  The half-row sums are combined once, at the end.
// The tensor pipe executes in issue order, so S_(j+3) (issued behind P_j V_j) cannot overwrite P_j before it is read.
// K_j and V_j share one TMA box geometry ([keys][64 d-columns]); the descriptor reads the same bytes K-major for
// Q K^T and MN-major for P V, so V needs no transpose.
// Outputs: O [rows, hq, d] and LSE fp32 (natural log), the contract of the backward kernels.
#include "common.cuh"
#include "flash_common.cuh"
#include "tc.cuh"

#ifndef PDB_FA_POLY
#define PDB_FA_POLY 2      // share of the exponentials computed on the FMA pipe: 0 none, 1: 1/8, 2: 1/4, 3: 1/2
#endif
#ifndef PDB_FA_DIAG
#define PDB_FA_DIAG 0      // 1: timing experiment only, no exponentials (wrong results)
#endif

namespace pdb {
namespace fa {
using namespace pdb::tc;
using namespace pdb::fac;

// -DPDB_FA_TRACE: CTA 0 stamps clock64() at the hand-off points of its MMA thread (role 0) and of lane 0 of the first
// warp of each softmax group (roles 1, 2) into a caller-provided buffer [3][4096] (pdb200_flash_attn_set_trace);
// bench/attn_trace.py turns the stamps into a per-tile timeline.  Not compiled into the product library.
#ifdef PDB_FA_TRACE
__device__ unsigned long long* g_trace = nullptr;
#define PDB_TR(role, tag)                                                                              \
  do {                                                                                                 \
    if (blockIdx.x == 1 && blockIdx.y == 0 && blockIdx.z == 0 && g_trace && tr_n < 4096) g_trace[(role) * 4096 + tr_n++] = ((unsigned long long)(tag) << 56) | (clock64() & 0xffffffffffffffull); \
  } while (0)
#else
#define PDB_TR(role, tag) do { } while (0)
#endif

constexpr int BQ = 128;     // query rows per CTA
constexpr int BKV = 128;    // keys per tile
constexpr int kThreads = 384;
constexpr float kRescaleThreshold = 8.f;   // log2 units

template <int D> struct FwdSmem {
  static constexpr int kQBytes = BQ * D * 2;
  static constexpr int kKVBytes = BKV * D * 2;            // one of K or V
  static constexpr int kKStages = 3;                     // K_j is needed two tiles before V_j
  static constexpr int kVStages = 2;
  static constexpr int kStatsBytes = 2 * BQ * 4;          // [column half][row]: half-row maxima / sums in flight
  static constexpr int kTotal = kQBytes + (kKStages + kVStages) * kKVBytes + kStatsBytes + 1024 + 256;
};

struct FwdParams {
  SeqArgs seq;
  int hq, hk;
  float scale_log2;     // softmax_scale * log2(e)
  float scale;          // softmax_scale
  int causal;
  uint16_t* o;          // bf16 or fp16
  float* lse;
  Dropout drop;
};

template <int D, bool kHalf, bool kDrop>
__global__ void __launch_bounds__(kThreads, 1)
flash_fwd_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                 const __grid_constant__ CUtensorMap tm_v, const FwdParams p) {
  using S = FwdSmem<D>;
  constexpr int KB = D / 64;                 // 64-column k-blocks of the head dimension
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + S::kQBytes;             // [3] stages
  uint8_t* sV = sK + S::kKStages * S::kKVBytes;   // [2] stages
  uint8_t* sStats = sV + S::kVStages * S::kKVBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sStats + S::kStatsBytes);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;               // [3]
  uint64_t* k_empty = bars + 4;              // [3]
  uint64_t* v_full = bars + 7;               // [2]
  uint64_t* v_empty = bars + 9;              // [2]
  uint64_t* s_full = bars + 11;              // [3]
  uint64_t* p_ready = bars + 14;             // [3]
  uint64_t* o_full = bars + 17;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 18);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heavy (late) causal tiles first so the tail of the grid is made of short CTAs
  const int n_qt = (int)gridDim.x;
  const int qt = p.causal ? (n_qt - 1 - (int)blockIdx.x) : (int)blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int hk = h / (p.hq / p.hk);
  const Seq sq_ = seq_of(p.seq, b);
  if (qt * BQ >= sq_.q_len) return;                      // packed batch: this sequence is shorter than the longest
  const int q_len = sq_.q_len, k_len = sq_.k_len;
  const int q_row0 = sq_.q_start + qt * BQ;              // row in the [rows, hq*D] view
  const int kv_row0 = sq_.k_start;
  const int diag = k_len - q_len;                        // key index aligned with query index 0
  const int n_kv_all = (k_len + BKV - 1) / BKV;
  const int n_kv = p.causal ? min((qt * BQ + BQ + diag + BKV - 1) / BKV, n_kv_all) : n_kv_all;
  const int64_t stat_row0 = sq_.stat0 + (int64_t)h * p.seq.stat_hs + qt * BQ;
  if (n_kv <= 0) {
    // no key is visible to any row of this tile (causal with fewer keys than queries, or an empty key sequence)
    for (int i = threadIdx.x; i < BQ * (D / 8); i += kThreads) {
      const int row = i / (D / 8), c8 = i % (D / 8);
      if (qt * BQ + row < q_len)
        *reinterpret_cast<uint4*>(p.o + ((int64_t)(q_row0 + row) * p.hq + h) * D + c8 * 8) = make_uint4(0, 0, 0, 0);
    }
    if (p.lse)
      for (int row = threadIdx.x; row < BQ; row += kThreads)
        if (qt * BQ + row < q_len) p.lse[stat_row0 + row] = -INFINITY;
    return;
  }

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_q);
    prefetch_tmap(&tm_k);
    prefetch_tmap(&tm_v);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < S::kKStages; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    for (int i = 0; i < 3; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_ready[i], 8);
    }
    mbar_init(o_full, 1);
    fence_barrier_init();
  }
  constexpr int kTmemCols = 512;             // S_0 | S_1 | S_2 | O
  if (warp == 2) tmem_alloc(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tmem_o = tmem_base + 3 * BKV;

  if (warp < 4) {
  reg_dealloc<56>();
  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(q_full, S::kQBytes);
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) tma_load_2d(sQ + kb * (BQ * 128), &tm_q, q_full, h * D + kb * 64, q_row0);
      for (int j = 0; j < n_kv; ++j) {
        const int st = j % S::kKStages;
        mbar_wait(&k_empty[st], ((j / S::kKStages) & 1) ^ 1);
        mbar_expect_tx(&k_full[st], S::kKVBytes);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb)
          tma_load_2d(sK + st * S::kKVBytes + kb * (BKV * 128), &tm_k, &k_full[st], hk * D + kb * 64, kv_row0 + j * BKV);
      }
    }
  } else if (warp == 3) {
    if (elect_one()) {
      for (int j = 0; j < n_kv; ++j) {
        const int st = j & 1;
        mbar_wait(&v_empty[st], ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(&v_full[st], S::kKVBytes);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb)
          tma_load_2d(sV + st * S::kKVBytes + kb * (BKV * 128), &tm_v, &v_full[st], hk * D + kb * 64, kv_row0 + j * BKV);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {   // elect.sync, not a lane test: ptxas then knows one lane runs this and issues UTCHMMA / UTCBAR from uniform registers
      [[maybe_unused]] int tr_n = 0;
      constexpr uint32_t idesc_s = make_idesc_16(kHalf, BQ, BKV, 0, 0);     // Q (K-major) x K (K-major)
      constexpr uint32_t idesc_o = make_idesc_16(kHalf, BQ, D, 0, 1);       // P (K-major) x V (MN-major)
      auto issue_pv = [&](int j) {
        const int sb = j % 3, vst = j & 1;
        PDB_TR(0, 8);
        mbar_wait(&v_full[vst], (j >> 1) & 1);
        PDB_TR(0, 9);
        mbar_wait(&p_ready[sb], (j / 3) & 1);
        tc_fence_after();
        PDB_TR(0, 10);
        const uint32_t sv = smem_u32(sV + vst * S::kKVBytes);
        // P_j: 16 keys = 8 packed columns; keys 0-63 sit at columns 0-31 of S_sb, keys 64-127 at columns 64-95 (each
        // softmax thread writes over the start of the 64 score columns it has read itself)
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {
          const uint64_t db = make_smem_desc(sv + k * 2048, BKV * 128, 1024);
          umma_f16_ts(tmem_o, tmem_base + sb * BKV + (k >> 2) * 64 + (k & 3) * 8, db, idesc_o, (k != 0 || j > 0) ? 1u : 0u);
        }
        umma_commit(o_full);
        umma_commit(&v_empty[vst]);
        PDB_TR(0, 11);
      };
      auto issue_qk = [&](int j) {
        // S_(j%3) is free: every thread read S_(j-3) before it produced P_(j-3), and P_(j-3) V was issued ahead of this
        const int sb = j % 3, kst = j % S::kKStages;
        mbar_wait(&k_full[kst], (j / S::kKStages) & 1);
        tc_fence_after();
        PDB_TR(0, 12);
        const uint32_t sq_a = smem_u32(sQ);
        const uint32_t sk_ = smem_u32(sK + kst * S::kKVBytes);
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint64_t da = make_smem_desc(sq_a + (k >> 2) * (BQ * 128) + (k & 3) * 32, 16, 1024);
          const uint64_t db = make_smem_desc(sk_ + (k >> 2) * (BKV * 128) + (k & 3) * 32, 16, 1024);
          umma_f16(tmem_base + sb * BKV, da, db, idesc_s, k != 0 ? 1u : 0u);
        }
        umma_commit(&s_full[sb]);
        umma_commit(&k_empty[kst]);
        PDB_TR(0, 13);
      };
      mbar_wait(q_full, 0);
      issue_qk(0);
      if (n_kv > 1) issue_qk(1);
      if (n_kv > 2) issue_qk(2);
      for (int j = 0; j < n_kv; ++j) {
        issue_pv(j);
        if (j + 3 < n_kv) issue_qk(j + 3);
      }
    }
  }
  } else {
    reg_alloc<216>();
    const int half = (warp - 4) >> 2;                       // column half of the tile this thread works on
    const int q = warp & 3;                                 // TMEM lane quarter this warp may touch
    const int row = q * 32 + lane;                          // row inside the tile
    const int q_idx = qt * BQ + row;                        // query position in the sequence
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const uint32_t to = tmem_o + lane_off + half * (D / 2); // this thread's half of the accumulator row
    const uint32_t xch = smem_u32(sStats);                  // [half][row] floats
    float m_used = -INFINITY, l_part = 0.f;
    [[maybe_unused]] int tr_n = 0;
    [[maybe_unused]] const bool tr_on = (q == 0 && lane == 0);
    for (int j = 0; j < n_kv; ++j) {
      const int sb = j % 3;
      const uint32_t ts = tmem_base + sb * BKV + half * 64 + lane_off;
      if (tr_on) PDB_TR(1 + half, 0);
      mbar_wait(&s_full[sb], (j / 3) & 1);
      tc_fence_after();
      if (tr_on) PDB_TR(1 + half, 1);
      uint32_t r[64];
      tmem_ld_32x32(ts, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
      tmem_ld_32x32(ts + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
      tmem_ld_wait();
      if (tr_on) PDB_TR(1 + half, 2);
      const int key0 = j * BKV + half * 64;                     // first key of this thread's 64
      const bool on_diag = p.causal && (j * BKV + BKV - 1 > qt * BQ + diag);   // tile crosses the diagonal for some row
      if (on_diag || j * BKV + BKV > k_len) {                   // ... or holds the end of the key sequence
        int lim = k_len - 1 - key0;                             // keys with local index > lim are masked
        if (on_diag) lim = min(lim, q_idx + diag - key0);
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i > lim) r[i] = 0xff800000u;                      // -inf
      }
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < 64; i += 4) {
#pragma unroll
        for (int u = 0; u < 4; ++u) mx[u] = fmaxf(mx[u], __uint_as_float(r[i + u]));
      }
      const float m_half = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      // does any row of the tile exceed the shared stale maximum by more than the bound?  (-inf - -inf = NaN: false)
      const bool grow = (m_half - m_used) * p.scale_log2 > kRescaleThreshold;
      const bool resync = (j == 0) || bar_red_or(1, 256, grow);
      if (resync) {
        // rare after the first tile: agree on the row maximum, rescale what has been accumulated under the old one
        sts32f(xch + (half * BQ + row) * 4, m_half);
        named_bar_sync(2, 256);
        const float m_new = fmaxf(m_used, fmaxf(lds32f(xch + row * 4), lds32f(xch + (BQ + row) * 4)));
        named_bar_sync(2, 256);                                 // the exchange words may be written again
        if (j > 0) {
          const float alpha = (m_new == -INFINITY) ? 1.f : ex2_approx((m_used - m_new) * p.scale_log2);
          l_part *= alpha;
          mbar_wait(o_full, (j - 1) & 1);                       // P_(j-1) V has been accumulated
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < D / 64; ++c) {
            uint32_t o[32];
            tmem_ld_32x32(to + c * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32(to + c * 32, o);
          }
          tmem_st_wait();
        }
        m_used = m_new;
      }
      // probabilities (packed, kept in registers) + partial row sum
      const float mb = (m_used == -INFINITY) ? 0.f : m_used * p.scale_log2;   // all scores are -inf then: exp2 -> 0
      float ls[4] = {0.f, 0.f, 0.f, 0.f};
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 64; i += 4) {
        float x0, x1, x2, x3;
        ffma2(x0, x1, __uint_as_float(r[i]), __uint_as_float(r[i + 1]), p.scale_log2, -mb);
        ffma2(x2, x3, __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]), p.scale_log2, -mb);
#if PDB_FA_DIAG == 0
        x0 = ex2_approx(x0);
        x1 = ex2_approx(x1);
        // PDB_FA_POLY: share of the exponentials computed on the FMA pipe (1: 1/8, 2: 1/4, 3: 1/2)
        if ((PDB_FA_POLY == 3) || (PDB_FA_POLY == 2 && (i & 4)) || (PDB_FA_POLY == 1 && (i & 12) == 12)) {
          ex2_poly2(x2, x3, x2, x3);
        } else {
          x2 = ex2_approx(x2);
          x3 = ex2_approx(x3);
        }
#endif
        fadd2(ls[0], ls[1], x0, x1);
        fadd2(ls[2], ls[3], x2, x3);
        pk[i >> 1] = pack2<kHalf>(x0, x1);
        pk[(i >> 1) + 1] = pack2<kHalf>(x2, x3);
      }
      if constexpr (kDrop) {
        // the row sum above is of the undropped probabilities; the P that multiplies V loses its dropped entries
        // (the 1 / keep factor is applied to the output once, at the end)
        const uint32_t bh = (uint32_t)(b * p.hq + h);
        const int sh = (q_idx & 1) * 16;
#pragma unroll
        for (int i = 0; i < 64; i += 2) {
          const uint32_t w = dropout_patch(p.drop, bh, (uint32_t)q_idx, (uint32_t)(key0 + i)) >> sh;
          uint32_t m = 0;
          if ((w & 0xffu) < p.drop.threshold) m |= 0x0000ffffu;
          if (((w >> 8) & 0xffu) < p.drop.threshold) m |= 0xffff0000u;
          pk[i >> 1] &= m;
        }
      }
      l_part += (ls[0] + ls[1]) + (ls[2] + ls[3]);
      if (tr_on) PDB_TR(1 + half, 3);
      // P -> TMEM over the first 32 of the 64 score columns this thread has read (its own lanes and columns only)
      tmem_st_32x32(ts, pk);
      tmem_st_wait();
      if (j > 0) {
        // keeps the parity of o_full unambiguous (one completion per tile, observed every tile); P_(j-1) V was issued a
        // whole softmax ago, so this does not wait in practice
        mbar_wait(o_full, (j - 1) & 1);
      }
      // publish P and any rescale of O to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_ready[sb]);
      if (tr_on) PDB_TR(1 + half, 5);
    }
    // ---- combine the half-row sums, normalise, store --------------------------------------------------------
    mbar_wait(o_full, (n_kv - 1) & 1);
    tc_fence_after();
    sts32f(xch + (half * BQ + row) * 4, l_part);
    named_bar_sync(2, 256);
    const float l_fin = lds32f(xch + row * 4) + lds32f(xch + (BQ + row) * 4);
    // a row with a visible key has a sum of at least 2^-8 (its largest probability); a row without one only collects
    // the ~1e-38 that the polynomial exp2 returns for -inf
    const bool seen = l_fin > 1e-20f;
    const float inv_l = (seen ? 1.f / l_fin : 0.f) * (kDrop ? p.drop.inv_keep : 1.f);
    const bool row_ok = q_idx < q_len;
    // this thread stores columns [half * D/2, (half+1) * D/2) of its row
    uint16_t* orow = p.o + ((int64_t)(q_row0 + row) * p.hq + h) * D + half * (D / 2);
#pragma unroll
    for (int c = 0; c < D / 64; ++c) {
      uint32_t oa[32];
      tmem_ld_32x32(to + c * 32, oa);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        uint4 pk;
        pk.x = pack2<kHalf>(__uint_as_float(oa[i]) * inv_l, __uint_as_float(oa[i + 1]) * inv_l);
        pk.y = pack2<kHalf>(__uint_as_float(oa[i + 2]) * inv_l, __uint_as_float(oa[i + 3]) * inv_l);
        pk.z = pack2<kHalf>(__uint_as_float(oa[i + 4]) * inv_l, __uint_as_float(oa[i + 5]) * inv_l);
        pk.w = pack2<kHalf>(__uint_as_float(oa[i + 6]) * inv_l, __uint_as_float(oa[i + 7]) * inv_l);
        if (row_ok) *reinterpret_cast<uint4*>(orow + c * 32 + i) = pk;
      }
    }
    if (tr_on) PDB_TR(1 + half, 7);
    if (half == 0 && p.lse && row_ok)
      p.lse[stat_row0 + row] = seen ? m_used * p.scale + __logf(l_fin) : -INFINITY;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

template <int D, bool kHalf>
static int launch_fwd(const void* q, const void* k, const void* v, void* o, float* lse, const SeqArgs& seq, int64_t rows_q,
                      int64_t rows_k, int b, int hq, int hk, float scale, int causal, const Dropout& drop, cudaStream_t st) {
  using S = FwdSmem<D>;
  CUtensorMap tm_q, tm_k, tm_v;
  if (get_tmap_2d(&tm_q, q, (uint64_t)rows_q, (uint64_t)hq * D, (uint64_t)hq * D * 2, BQ, 64, 2, true, "flash_attn q")) return -1;
  if (get_tmap_2d(&tm_k, k, (uint64_t)rows_k, (uint64_t)hk * D, (uint64_t)hk * D * 2, BKV, 64, 2, true, "flash_attn k")) return -1;
  if (get_tmap_2d(&tm_v, v, (uint64_t)rows_k, (uint64_t)hk * D, (uint64_t)hk * D * 2, BKV, 64, 2, true, "flash_attn v")) return -1;
  FwdParams p{seq, hq, hk, scale * 1.4426950408889634f, scale, causal, (uint16_t*)o, lse, drop};
  dim3 grid((seq.sq + BQ - 1) / BQ, hq, b);
  if (drop.threshold) {
    auto kern = flash_fwd_kernel<D, kHalf, true>;
    PDB_ENSURE_SMEM(kern, S::kTotal);
    kern<<<grid, kThreads, S::kTotal, st>>>(tm_q, tm_k, tm_v, p);
  } else {
    auto kern = flash_fwd_kernel<D, kHalf, false>;
    PDB_ENSURE_SMEM(kern, S::kTotal);
    kern<<<grid, kThreads, S::kTotal, st>>>(tm_q, tm_k, tm_v, p);
  }
  return check_launch("flash_attn_fwd");
}

// keep probability -> 8-bit threshold + scale; the (seed, offset) pair of the caller's generator is folded into two words
Dropout make_dropout(float p_drop, uint64_t seed, uint64_t offset) {
  Dropout d{0, 0, 0, 1.f};
  if (p_drop > 0.f) {
    int thr = (int)((1.0 - (double)p_drop) * 256.0 + 0.5);
    thr = thr < 1 ? 1 : (thr > 255 ? 255 : thr);
    const uint64_t z = seed * 0x9e3779b97f4a7c15ull + offset * 0xbf58476d1ce4e5b9ull + 0x94d049bb133111ebull;
    d.seed_lo = (uint32_t)z;
    d.seed_hi = (uint32_t)(z >> 32);
    d.threshold = (uint32_t)thr;
    d.inv_keep = 256.f / (float)thr;
  }
  return d;
}

template <int D>
static int launch_fwd_dt(bool half, const void* q, const void* k, const void* v, void* o, float* lse, const SeqArgs& seq,
                         int64_t rows_q, int64_t rows_k, int b, int hq, int hk, float scale, int causal, const Dropout& drop,
                         cudaStream_t st) {
  return half ? launch_fwd<D, true>(q, k, v, o, lse, seq, rows_q, rows_k, b, hq, hk, scale, causal, drop, st)
              : launch_fwd<D, false>(q, k, v, o, lse, seq, rows_q, rows_k, b, hq, hk, scale, causal, drop, st);
}

}  // namespace fa
}  // namespace pdb

using namespace pdb;

#ifdef PDB_FA_TRACE
PDB_EXPORT int pdb200_flash_attn_set_trace(void* buf) {
  unsigned long long* p = (unsigned long long*)buf;
  return cudaMemcpyToSymbol(fa::g_trace, &p, sizeof(p)) == cudaSuccess ? 0 : -1;
}
#endif

static int fa_fwd_common(const void* q, const void* k, const void* v, void* o, float* lse, const fac::SeqArgs& seq,
                         int64_t rows_q, int64_t rows_k, int64_t b, int64_t hq, int64_t hk, int64_t d, float scale, int causal,
                         float p_drop, uint64_t seed, uint64_t offset, int dtype, void* stream) {
  PDB_REQUIRE(dtype == PDB_DTYPE_BF16 || dtype == PDB_DTYPE_F16, "flash_attn_fwd: bfloat16 or float16 only");
  PDB_REQUIRE(d == 64 || d == 128, "flash_attn_fwd: head_dim must be 64 or 128 (got %lld)", (long long)d);
  PDB_REQUIRE(b > 0 && seq.sq > 0 && seq.sk > 0 && rows_q > 0 && rows_k > 0, "flash_attn_fwd: empty problem");
  PDB_REQUIRE(hq % hk == 0, "flash_attn_fwd: q heads must be a multiple of kv heads");
  PDB_REQUIRE(p_drop >= 0.f && p_drop < 1.f, "flash_attn_fwd: dropout must be in [0, 1)");
  PDB_REQUIRE(rows_q < (1ll << 31) && rows_k < (1ll << 31), "flash_attn_fwd: more than 2^31 tokens");
  const fac::Dropout drop = fa::make_dropout(p_drop, seed, offset);
  cudaStream_t st = (cudaStream_t)stream;
  const bool half = dtype == PDB_DTYPE_F16;
  if (d == 128) return fa::launch_fwd_dt<128>(half, q, k, v, o, lse, seq, rows_q, rows_k, (int)b, (int)hq, (int)hk, scale, causal, drop, st);
  return fa::launch_fwd_dt<64>(half, q, k, v, o, lse, seq, rows_q, rows_k, (int)b, (int)hq, (int)hk, scale, causal, drop, st);
}

// Dense [b, s, h, d]; any sq / sk >= 1.  LSE [b, hq, sq].  Causal masks are aligned to the last key (query i sees keys
// <= i + sk - sq); rows that see no key get output 0 and LSE -inf.
PDB_EXPORT int pdb200_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int64_t b,
                                     int64_t sq, int64_t sk, int64_t hq, int64_t hk, int64_t d, float scale, int causal,
                                     int dtype, void* stream) {
  fac::SeqArgs seq{nullptr, nullptr, (int)sq, (int)sk, hq * sq, sq};
  return fa_fwd_common(q, k, v, o, lse, seq, b * sq, b * sk, b, hq, hk, d, scale, causal, 0.f, 0, 0, dtype, stream);
}

PDB_EXPORT int pdb200_flash_attn_fwd_dropout(const void* q, const void* k, const void* v, void* o, float* lse, int64_t b,
                                             int64_t sq, int64_t sk, int64_t hq, int64_t hk, int64_t d, float scale,
                                             int causal, float p_drop, uint64_t seed, uint64_t offset, int dtype,
                                             void* stream) {
  fac::SeqArgs seq{nullptr, nullptr, (int)sq, (int)sk, hq * sq, sq};
  return fa_fwd_common(q, k, v, o, lse, seq, b * sq, b * sk, b, hq, hk, d, scale, causal, p_drop, seed, offset, dtype, stream);
}

// Packed [total, h, d] with device-side int32 prefix sums cu_seqlens_q / cu_seqlens_k of b + 1 entries each; the grid
// is sized from max_seqlen_q, nothing is read back to the host.  LSE [hq, total_q].
PDB_EXPORT int pdb200_flash_attn_varlen_fwd(const void* q, const void* k, const void* v, void* o, float* lse,
                                            const int* cu_seqlens_q, const int* cu_seqlens_k, int64_t b, int64_t total_q,
                                            int64_t total_k, int64_t max_seqlen_q, int64_t max_seqlen_k, int64_t hq,
                                            int64_t hk, int64_t d, float scale, int causal, float p_drop, uint64_t seed,
                                            uint64_t offset, int dtype, void* stream) {
  PDB_REQUIRE(cu_seqlens_q && cu_seqlens_k, "flash_attn_varlen_fwd: cu_seqlens are required");
  fac::SeqArgs seq{cu_seqlens_q, cu_seqlens_k, (int)max_seqlen_q, (int)max_seqlen_k, 0, total_q};
  return fa_fwd_common(q, k, v, o, lse, seq, total_q, total_k, b, hq, hk, d, scale, causal, p_drop, seed, offset, dtype, stream);
}
