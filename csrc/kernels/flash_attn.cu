// K16: flash-attention forward on tcgen05 tensor cores for sm_100a (bf16 / fp16, head_dim 64/128, causal or full, GQA,
// dense [b, s, h, d] batches of any length and packed variable-length batches with cu_seqlens, dropout).
//
// Replaces in the reference: paddle/phi/kernels/gpu/flash_attn_kernel.cu:36-170 (unpadded), :305-447 (dense + dropout)
// -> vendored FlashAttention-2 (mma.sync m16n8k16 + cp.async, third_party/flashattn/csrc/flash_attn/src/flash_fwd_kernel.h).
//
// One CTA per (batch, q-head, 128-row query tile); K and V tiles of 128 keys stream through separate 2-stage TMA rings.
// The softmax is the bottleneck on this machine (one exp2 per score on 16 SFU lanes/clk/SM costs as many cycles as the
// two 128x128x128 MMAs of a tile), so the CTA runs TWO independent softmax pipelines:
//   warp 0      TMA producer for Q and the K ring            warp 3   TMA producer for the V ring
//   warp 1      MMA issuer: S_j = Q K_j^T into TMEM buffer (j & 1); O_(j&1) += P_j V_j once P_j is in shared memory
//   warp 2      TMEM allocator (S_A | S_B | O_A | O_B = 512 columns)
//   warps 4-7   softmax group A: even key tiles            warps 8-11   softmax group B: odd key tiles
// Each group keeps its own running (max, sum) and its own output accumulator IN TMEM (the MMA accumulates across the
// group's tiles), so the groups never wait for each other; the two partial results are merged once at the end
// (the split-KV combination, inside the CTA).  A group thread owns one query row: a single tcgen05.ld pass brings the
// 128 scores into registers (the S buffer is handed back to the MMA warp immediately), then max / exp2 / row sum /
// bf16 pack, and P is written to shared memory in the UMMA K-major swizzled layout.  The accumulator is rescaled
// lazily: only when a row's maximum grew by more than 2^8 since the last rescale does the warp read-modify-write its
// O rows in TMEM (exact: P and the row sum use the same stale maximum until then).
// K_j and V_j share one TMA box geometry ([keys][64 d-columns]); the descriptor reads the same bytes K-major for
// Q K^T and MN-major for P V, so V needs no transpose.
// Outputs: O [b, sq, hq, d] bf16 and LSE [b, hq, sq] fp32 (natural log), the contract of the backward kernels.
#include "common.cuh"
#include "flash_common.cuh"
#include "tc.cuh"

namespace pdb {
namespace fa {
using namespace pdb::tc;
using namespace pdb::fac;

constexpr int BQ = 128;     // query rows per CTA
constexpr int BKV = 128;    // keys per tile
constexpr int kThreads = 384;
constexpr float kRescaleThreshold = 8.f;   // log2 units

template <int D> struct FwdSmem {
  static constexpr int kQBytes = BQ * D * 2;
  static constexpr int kKVBytes = BKV * D * 2;            // one of K or V
  static constexpr int kPBytes = BQ * BKV * 2;
  static constexpr int kTotal = kQBytes + 4 * kKVBytes + 2 * kPBytes + 1024 + 256;
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
  uint8_t* sK = sQ + S::kQBytes;             // [2] stages
  uint8_t* sV = sK + 2 * S::kKVBytes;        // [2] stages
  uint8_t* sP = sV + 2 * S::kKVBytes;        // [2] one per softmax group
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * S::kPBytes);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;               // [2]
  uint64_t* k_empty = bars + 3;              // [2]
  uint64_t* v_full = bars + 5;               // [2]
  uint64_t* v_empty = bars + 7;              // [2]
  uint64_t* s_full = bars + 9;               // [2]
  uint64_t* s_empty = bars + 11;             // [2]
  uint64_t* p_ready = bars + 13;             // [2]
  uint64_t* o_full = bars + 15;              // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 17);

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
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&s_empty[i], 4);
      mbar_init(&p_ready[i], 4);
      mbar_init(&o_full[i], 1);
    }
    fence_barrier_init();
  }
  constexpr int kTmemCols = 512;             // S_A | S_B | O_A | O_B
  if (warp == 2) tmem_alloc(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tmem_o = tmem_base + 2 * BKV;           // group g: tmem_o + g * D

  if (warp < 4) {
  reg_dealloc<56>();
  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, S::kQBytes);
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) tma_load_2d(sQ + kb * (BQ * 128), &tm_q, q_full, h * D + kb * 64, q_row0);
      for (int j = 0; j < n_kv; ++j) {
        const int st = j & 1;
        mbar_wait(&k_empty[st], ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(&k_full[st], S::kKVBytes);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb)
          tma_load_2d(sK + st * S::kKVBytes + kb * (BKV * 128), &tm_k, &k_full[st], hk * D + kb * 64, kv_row0 + j * BKV);
      }
    }
  } else if (warp == 3) {
    if (lane == 0) {
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
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_16(kHalf, BQ, BKV, 0, 0);     // Q (K-major) x K (K-major)
      constexpr uint32_t idesc_o = make_idesc_16(kHalf, BQ, D, 0, 1);       // P (K-major) x V (MN-major)
      auto issue_pv = [&](int t) {
        const int g = t & 1, ph = (t >> 1) & 1;
        mbar_wait(&v_full[g], ph);
        mbar_wait(&p_ready[g], ph);
        tc_fence_after();
        const uint32_t sv = smem_u32(sV + g * S::kKVBytes);
        const uint32_t sp = smem_u32(sP + g * S::kPBytes);
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {
          const uint64_t da = make_smem_desc(sp + (k >> 2) * (BQ * 128) + (k & 3) * 32, 16, 1024);
          const uint64_t db = make_smem_desc(sv + k * 2048, BKV * 128, 1024);
          umma_f16(tmem_o + g * D, da, db, idesc_o, (k != 0 || t >= 2) ? 1u : 0u);
        }
        umma_commit(&o_full[g]);
        umma_commit(&v_empty[g]);
      };
      auto issue_qk = [&](int j) {
        const int g = j & 1, ph = (j >> 1) & 1;
        mbar_wait(&k_full[g], ph);
        mbar_wait(&s_empty[g], ph ^ 1);
        tc_fence_after();
        const uint32_t sq_ = smem_u32(sQ);
        const uint32_t sk_ = smem_u32(sK + g * S::kKVBytes);
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint64_t da = make_smem_desc(sq_ + (k >> 2) * (BQ * 128) + (k & 3) * 32, 16, 1024);
          const uint64_t db = make_smem_desc(sk_ + (k >> 2) * (BKV * 128) + (k & 3) * 32, 16, 1024);
          umma_f16(tmem_base + g * BKV, da, db, idesc_s, k != 0 ? 1u : 0u);
        }
        umma_commit(&s_full[g]);
        umma_commit(&k_empty[g]);
      };
      mbar_wait(q_full, 0);
      // scores run two tiles ahead of the P V products: a group finds its next S ready when it finishes a tile
      issue_qk(0);
      if (n_kv > 1) issue_qk(1);
      for (int j = 2; j < n_kv; ++j) {
        issue_qk(j);
        issue_pv(j - 2);
      }
      if (n_kv > 1) issue_pv(n_kv - 2);
      issue_pv(n_kv - 1);
    }
  }
  } else {
    reg_alloc<216>();
    const int g = (warp - 4) >> 2;                          // softmax group: 0 = even tiles, 1 = odd tiles
    const int q = warp & 3;                                 // TMEM lane quarter this warp may touch
    const int row = q * 32 + lane;                          // row inside the tile
    const int q_idx = qt * BQ + row;                        // query position in the sequence
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const uint32_t ts = tmem_base + g * BKV + lane_off;
    const uint32_t to = tmem_o + g * D + lane_off;
    const uint32_t sPg_u32 = smem_u32(sP + g * S::kPBytes);
    float m_used = -INFINITY, l_run = 0.f;
    int it = 0;
    for (int j = g; j < n_kv; j += 2, ++it) {
      mbar_wait(&s_full[g], it & 1);
      tc_fence_after();
      uint32_t r[BKV];
#pragma unroll
      for (int c = 0; c < BKV / 32; ++c) tmem_ld_32x32(ts + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&r[c * 32]));
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_empty[g]);             // scores are in registers: the MMA warp may refill the buffer
      const bool on_diag = p.causal && (j * BKV + BKV - 1 > qt * BQ + diag);   // tile crosses the diagonal for some row
      if (on_diag || j * BKV + BKV > k_len) {                   // ... or holds the end of the key sequence
        int lim = k_len - 1 - j * BKV;                          // keys with local index > lim are masked
        if (on_diag) lim = min(lim, q_idx + diag - j * BKV);
#pragma unroll
        for (int i = 0; i < BKV; ++i)
          if (i > lim) r[i] = 0xff800000u;                      // -inf
      }
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < BKV; i += 4) {
#pragma unroll
        for (int u = 0; u < 4; ++u) mx[u] = fmaxf(mx[u], __uint_as_float(r[i + u]));
      }
      const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      float alpha = 1.f;
      bool rescale = false;
      if (it == 0) {
        m_used = m_tile;                                        // may be -inf: no visible key for this row so far
      } else {
        const bool grow = (m_tile - m_used) * p.scale_log2 > kRescaleThreshold;
        rescale = __any_sync(0xffffffffu, grow);
        if (rescale) {
          const float m_new = fmaxf(m_used, m_tile);
          alpha = (m_new == -INFINITY) ? 1.f : ex2_approx((m_used - m_new) * p.scale_log2);
          l_run *= alpha;
          m_used = m_new;
        }
      }
      // probabilities (packed bf16, kept in registers) + row sum; all of this overlaps the previous P V of the group
      const float mb = (m_used == -INFINITY) ? 0.f : m_used * p.scale_log2;   // all scores are -inf then: exp2 -> 0
      float ls[4] = {0.f, 0.f, 0.f, 0.f};
      uint32_t pk[BKV / 2];
#pragma unroll
      for (int i = 0; i < BKV; i += 4) {
        float x0, x1, x2, x3;
        ffma2(x0, x1, __uint_as_float(r[i]), __uint_as_float(r[i + 1]), p.scale_log2, -mb);
        ffma2(x2, x3, __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]), p.scale_log2, -mb);
        x0 = ex2_approx(x0);
        x1 = ex2_approx(x1);
        x2 = ex2_approx(x2);
        x3 = ex2_approx(x3);
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
        for (int i = 0; i < BKV; i += 2) {
          const uint32_t w = dropout_patch(p.drop, bh, (uint32_t)q_idx, (uint32_t)(j * BKV + i)) >> sh;
          uint32_t m = 0;
          if ((w & 0xffu) < p.drop.threshold) m |= 0x0000ffffu;
          if (((w >> 8) & 0xffu) < p.drop.threshold) m |= 0xffff0000u;
          pk[i >> 1] &= m;
        }
      }
      if (it > 0) {
        // the previous P V of this group must be done before P is overwritten or O is touched
        mbar_wait(&o_full[g], (it - 1) & 1);
        tc_fence_after();
        if (rescale) {
#pragma unroll
          for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            tmem_ld_32x32(to + c * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32(to + c * 32, o);
          }
          tmem_st_wait();
        }
      }
      // P -> shared memory (UMMA K-major, 128-byte swizzle): 32 keys = 4 chunks of 16 B inside k-block (c / 2),
      // chunk index ((c & 1) * 4 + t), XOR-swizzled by row % 8
#pragma unroll
      for (int c = 0; c < BKV / 32; ++c) {
        const uint32_t base = sPg_u32 + (c >> 1) * (BQ * 128) + row * 128;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int chunk = ((c & 1) * 4 + t) ^ (row & 7);
          sts128(base + chunk * 16, pk[c * 16 + t * 4], pk[c * 16 + t * 4 + 1], pk[c * 16 + t * 4 + 2], pk[c * 16 + t * 4 + 3]);
        }
      }
      l_run += (ls[0] + ls[1]) + (ls[2] + ls[3]);
      // publish P (generic-proxy writes -> async proxy) and any TMEM rescale to the MMA warp
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_ready[g]);
    }
    // ---- merge the two groups' partial results ------------------------------------------------------------
    if (it > 0) {
      mbar_wait(&o_full[g], (it - 1) & 1);
      tc_fence_after();
    }
    const uint32_t stats = smem_u32(sP);                   // P buffers are free now: [group][row][2] floats
    // group B may still be reading... no: each group only touches its own P buffer; the stats live in group A's
    // buffer, which group A's last P V (awaited above by group A) has finished reading before A writes here, and B
    // writes only after the named barrier below orders it behind A's wait.
    named_bar_sync(1, 256);
    sts32f(stats + ((g * BQ + row) * 2) * 4, (it > 0) ? m_used : -INFINITY);
    sts32f(stats + ((g * BQ + row) * 2 + 1) * 4, l_run);
    tc_fence_before();
    named_bar_sync(1, 256);
    tc_fence_after();
    const float m_a = lds32f(stats + (row * 2) * 4), l_a = lds32f(stats + (row * 2 + 1) * 4);
    const float m_b = lds32f(stats + ((BQ + row) * 2) * 4), l_b = lds32f(stats + ((BQ + row) * 2 + 1) * 4);
    const bool has_b = n_kv > 1;
    const float m_fin = fmaxf(m_a, m_b);                    // -inf only for a row that sees no key at all
    const float f_a = (m_a == -INFINITY) ? 0.f : ex2_approx((m_a - m_fin) * p.scale_log2);
    const float f_b = (m_b == -INFINITY) ? 0.f : ex2_approx((m_b - m_fin) * p.scale_log2);
    const float l_fin = l_a * f_a + l_b * f_b;
    const float inv_l = (l_fin > 0.f ? 1.f / l_fin : 0.f) * (kDrop ? p.drop.inv_keep : 1.f);
    const float w_a = f_a * inv_l, w_b = f_b * inv_l;
    const bool row_ok = q_idx < q_len;
    // group g writes columns [g * D/2, (g+1) * D/2) of the row
    uint16_t* orow = p.o + ((int64_t)(q_row0 + row) * p.hq + h) * D + g * (D / 2);
#pragma unroll
    for (int c = 0; c < D / 64; ++c) {
      uint32_t oa[32], ob[32];
      tmem_ld_32x32(tmem_o + lane_off + g * (D / 2) + c * 32, oa);
      if (has_b) tmem_ld_32x32(tmem_o + D + lane_off + g * (D / 2) + c * 32, ob);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          v[u] = __uint_as_float(oa[i + u]) * w_a;
          if (has_b) v[u] = fmaf(__uint_as_float(ob[i + u]), w_b, v[u]);
        }
        uint4 pk;
        pk.x = pack2<kHalf>(v[0], v[1]);
        pk.y = pack2<kHalf>(v[2], v[3]);
        pk.z = pack2<kHalf>(v[4], v[5]);
        pk.w = pack2<kHalf>(v[6], v[7]);
        if (row_ok) *reinterpret_cast<uint4*>(orow + c * 32 + i) = pk;
      }
    }
    if (g == 0 && p.lse && row_ok)
      p.lse[stat_row0 + row] = l_fin > 0.f ? m_fin * p.scale + __logf(l_fin) : -INFINITY;
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
