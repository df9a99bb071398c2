// K17, dK/dV half: included by flash_attn_bwd.cu (shares BwdParams / pack_bf16 / kThreads with the dQ kernel).
//
// One CTA per (128-key tile, kv head, batch).  It loops over the group's query heads x 128-row query tiles and keeps
// dV and dK in TMEM for the whole loop.  Per iteration, on the tensor core:
//     S^T  = K_j Q_i^T      (region A, 128 fp32 columns)        dP^T = V_j dO_i^T    (region B, 128 fp32 columns)
//     dV  += P^T dO_i                                             dK  += dS^T Q_i
// P^T and dS^T never touch shared memory: each element-wise warp reads its 64 columns of S^T and dP^T into registers,
// then stores P^T and dS^T as packed bf16 over the dP^T columns it has just read (32 + 32 columns), and the two
// accumulating MMAs take their A operand from there (tcgen05.mma with A in TMEM).  Region A is free again as soon as
// the scores are in registers, so S^T of the next tile is computed while this tile's exponentials run; dP^T of the
// next tile is issued behind the accumulating MMAs (in-order tensor pipe, so it cannot overtake their reads).
#pragma once

namespace pdb {
namespace fab {

template <int D> struct DkvSmem {
  static constexpr int kTile = 128 * D * 2;               // K_j, V_j, one Q or dO stage
  static constexpr int kStats = 2 * 2 * 128 * 4;          // [stage][-lse*log2e | -delta*scale][128]
  static constexpr int kTotal = 6 * kTile + kStats + 1024 + 256;
};

template <int D, bool kHalf, bool kDrop>
__global__ void __launch_bounds__(kThreads, 1)
flash_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_do,
                     const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                     const BwdParams p) {
  using S = DkvSmem<D>;
  constexpr int KB = D / 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + S::kTile;
  uint8_t* sQ = sV + S::kTile;               // [2]
  uint8_t* sDO = sQ + 2 * S::kTile;          // [2]
  uint8_t* stats = sDO + 2 * S::kTile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stats + S::kStats);
  uint64_t* kv_full = bars;
  uint64_t* q_full = bars + 1;               // [2]
  uint64_t* q_empty = bars + 3;              // [2]
  uint64_t* do_full = bars + 5;              // [2]
  uint64_t* do_empty = bars + 7;             // [2]
  uint64_t* s_full = bars + 9;
  uint64_t* dp_full = bars + 10;
  uint64_t* regs_loaded = bars + 11;
  uint64_t* pds_ready = bars + 12;
  uint64_t* all_done = bars + 13;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int jt = blockIdx.x, hk = blockIdx.y, b = blockIdx.z;
  const int G = p.hq / p.hk;
  const Seq sq_ = seq_of(p.seq, b);
  if (jt * 128 >= sq_.k_len) return;                                   // packed batch: shorter key sequence
  const int q_len = sq_.q_len, k_len = sq_.k_len;
  const int diag = k_len - q_len;
  const int i0 = p.causal ? max(0, jt * 128 - diag) / 128 : 0;        // first query tile that sees this key tile
  const int n_qi = (q_len + 127) / 128 - i0;
  const int T = G * n_qi;
  const int kv_row0 = sq_.k_start + jt * 128;
  const int q_base = sq_.q_start;
  if (n_qi <= 0) {                                                     // no query sees these keys: dK = dV = 0
    for (int i = threadIdx.x; i < 128 * (D / 8); i += kThreads) {
      const int row = i / (D / 8), c8 = i % (D / 8);
      if (jt * 128 + row < k_len) {
        const int64_t at = ((int64_t)(kv_row0 + row) * p.hk + hk) * D + c8 * 8;
        *reinterpret_cast<uint4*>(p.dk + at) = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(p.dv + at) = make_uint4(0, 0, 0, 0);
      }
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
    mbar_init(kv_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
      mbar_init(&do_full[i], 1);
      mbar_init(&do_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(dp_full, 1);
    mbar_init(regs_loaded, 8);
    mbar_init(pds_ready, 8);
    mbar_init(all_done, 1);
    fence_barrier_init();
  }
  constexpr int kTmemCols = 512;             // A: S^T | B: dP^T -> P^T, dS^T | dV | dK
  if (warp == 2) tmem_alloc(tmem_ptr_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_a = *tmem_ptr_smem;
  const uint32_t tmem_b = tmem_a + 128;
  const uint32_t tmem_dv = tmem_a + 256;
  const uint32_t tmem_dk = tmem_a + 256 + D;

  if (warp < 4) {
    reg_dealloc<56>();
    if (warp == 0) {
      if (elect_one()) {
        mbar_expect_tx(kv_full, 2 * S::kTile);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb) {
          tma_load_2d(sK + kb * (128 * 128), &tm_k, kv_full, hk * D + kb * 64, kv_row0);
          tma_load_2d(sV + kb * (128 * 128), &tm_v, kv_full, hk * D + kb * 64, kv_row0);
        }
        for (int t = 0; t < T; ++t) {
          const int st = t & 1;
          const int h = hk * G + t / n_qi, i = i0 + t % n_qi;
          mbar_wait(&q_empty[st], ((t >> 1) & 1) ^ 1);
          mbar_expect_tx(&q_full[st], S::kTile);
#pragma unroll
          for (int kb = 0; kb < KB; ++kb)
            tma_load_2d(sQ + st * S::kTile + kb * (128 * 128), &tm_q, &q_full[st], h * D + kb * 64, q_base + i * 128);
        }
      }
    } else if (warp == 3) {
      if (elect_one()) {
        for (int t = 0; t < T; ++t) {
          const int st = t & 1;
          const int h = hk * G + t / n_qi, i = i0 + t % n_qi;
          mbar_wait(&do_empty[st], ((t >> 1) & 1) ^ 1);
          mbar_expect_tx(&do_full[st], S::kTile);
#pragma unroll
          for (int kb = 0; kb < KB; ++kb)
            tma_load_2d(sDO + st * S::kTile + kb * (128 * 128), &tm_do, &do_full[st], h * D + kb * 64, q_base + i * 128);
        }
      }
    } else if (warp == 1) {
      if (elect_one()) {   // elect.sync, not a lane test: ptxas then knows one lane runs this and issues UTCHMMA / UTCBAR from uniform registers
        constexpr uint32_t idesc_t = make_idesc_16(kHalf, 128, 128, 0, 0);    // [keys][d] x [q][d]^T
        constexpr uint32_t idesc_acc = make_idesc_16(kHalf, 128, D, 0, 1);    // [keys][q] (TMEM) x [q][d] (MN-major smem)
        auto issue_s = [&](int t) {
          const int st = t & 1;
          mbar_wait(&q_full[st], (t >> 1) & 1);
          mbar_wait(regs_loaded, (t & 1) ^ 1);                           // region A of tile t-1 is in registers
          tc_fence_after();
          const uint32_t ak = smem_u32(sK), bq = smem_u32(sQ + st * S::kTile);
#pragma unroll
          for (int k = 0; k < D / 16; ++k) {
            const uint32_t off = (k >> 2) * (128 * 128) + (k & 3) * 32;
            umma_f16(tmem_a, make_smem_desc(ak + off, 16, 1024), make_smem_desc(bq + off, 16, 1024), idesc_t, k != 0);
          }
          umma_commit(s_full);
        };
        auto issue_dp = [&](int t) {
          const int st = t & 1;
          mbar_wait(&do_full[st], (t >> 1) & 1);
          tc_fence_after();
          const uint32_t av = smem_u32(sV), bd = smem_u32(sDO + st * S::kTile);
#pragma unroll
          for (int k = 0; k < D / 16; ++k) {
            const uint32_t off = (k >> 2) * (128 * 128) + (k & 3) * 32;
            umma_f16(tmem_b, make_smem_desc(av + off, 16, 1024), make_smem_desc(bd + off, 16, 1024), idesc_t, k != 0);
          }
          umma_commit(dp_full);
        };
        auto issue_acc = [&](int t) {
          const int st = t & 1;
          mbar_wait(pds_ready, t & 1);
          tc_fence_after();
          const uint32_t bd = smem_u32(sDO + st * S::kTile), bq = smem_u32(sQ + st * S::kTile);
          // 16 queries per step = 8 packed columns; half h of the queries sits at B + 64*h (+32 for dS^T)
#pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_f16_ts(tmem_dv, tmem_b + (k >> 2) * 64 + (k & 3) * 8, make_smem_desc(bd + k * 2048, 128 * 128, 1024),
                        idesc_acc, (t > 0 || k > 0) ? 1u : 0u);
#pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_f16_ts(tmem_dk, tmem_b + (k >> 2) * 64 + 32 + (k & 3) * 8, make_smem_desc(bq + k * 2048, 128 * 128, 1024),
                        idesc_acc, (t > 0 || k > 0) ? 1u : 0u);
          umma_commit(&q_empty[st]);
          umma_commit(&do_empty[st]);
        };
        mbar_wait(kv_full, 0);
        issue_s(0);
        issue_dp(0);
        for (int t = 0; t < T; ++t) {
          if (t + 1 < T) issue_s(t + 1);
          issue_acc(t);
          if (t + 1 < T) issue_dp(t + 1);       // behind the accumulating MMAs: they read region B first
        }
        umma_commit(all_done);
      }
    }
  } else {
    reg_alloc<216>();
    const int q = warp & 3, half = (warp - 4) >> 2;
    const int row = q * 32 + lane;                       // key row inside the tile
    const int sidx = (warp - 4) * 32 + lane;             // 0..255
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const uint32_t stats_u32 = smem_u32(stats);
    // per-query statistics of the NEXT iteration are fetched while the current one computes (global latency hidden)
    float nxt = 0.f;
    auto fetch_stats = [&](int t) {
      if (t < T) {
        const int qi = (i0 + t % n_qi) * 128 + (sidx & 127);
        const int64_t at = sq_.stat0 + (int64_t)(hk * G + t / n_qi) * p.seq.stat_hs + qi;
        if (qi < q_len) {
          nxt = (sidx < 128) ? -p.lse[at] * kLog2e : -p.delta[at] * p.scale;
          if (nxt == INFINITY) nxt = -INFINITY;           // LSE = -inf (row saw no key): probability 0, not inf
        } else {
          nxt = (sidx < 128) ? -INFINITY : 0.f;           // queries past the end of the sequence: probability 0
        }
      }
    };
    fetch_stats(0);
    for (int t = 0; t < T; ++t) {
      const int st = t & 1;
      const int i = i0 + t % n_qi;
      const int h = hk * G + t / n_qi;
      (void)h;
      sts32f(stats_u32 + (st * 256 + sidx) * 4, nxt);     // threads 0-127: -lse*log2e, threads 128-255: -delta*scale
      named_bar_sync(1, 256);
      fetch_stats(t + 1);
      mbar_wait(s_full, t & 1);
      mbar_wait(dp_full, t & 1);
      tc_fence_after();
      uint32_t rs[64], rd[64];
      tmem_ld_32x32(tmem_a + lane_off + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&rs[0]));
      tmem_ld_32x32(tmem_a + lane_off + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&rs[32]));
      tmem_ld_32x32(tmem_b + lane_off + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&rd[0]));
      tmem_ld_32x32(tmem_b + lane_off + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&rd[32]));
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(regs_loaded);
      if (p.causal && (jt * 128 + 127 > i * 128 + diag)) {
        // key (jt*128 + row) is visible to query (i*128 + c) iff key <= query + diag  <=>  c >= row + jt*128 - diag - i*128
        const int cmin = jt * 128 + row - diag - i * 128 - half * 64;      // local columns below cmin are masked
#pragma unroll
        for (int c = 0; c < 64; ++c)
          if (c < cmin) rs[c] = 0xff800000u;
      }
      uint32_t keep[2] = {0xffffffffu, 0xffffffffu};     // bit c: score (key row, query column c) survived dropout
      if constexpr (kDrop) {
        // dP^T of a dropped score is 0 here; its P^T entry (the dV operand) is zeroed after the exponentials
        const uint32_t bh = (uint32_t)(b * p.hq + h);
        const uint32_t key = (uint32_t)(jt * 128 + row);
        const int sh = (int)(key & 1) * 8;
        keep[0] = keep[1] = 0u;
#pragma unroll
        for (int c = 0; c < 64; c += 2) {
          // patch (q >> 1, k >> 1): byte (q & 1) * 2 + (k & 1); q = even column c -> byte (k & 1), c + 1 -> byte 2 + (k & 1)
          const uint32_t w = dropout_patch(p.drop, bh, (uint32_t)(i * 128 + half * 64 + c), key) >> sh;
          if ((w & 0xffu) < p.drop.threshold) keep[c >> 5] |= 1u << (c & 31);
          else rd[c] = 0u;
          if (((w >> 16) & 0xffu) < p.drop.threshold) keep[c >> 5] |= 1u << ((c + 1) & 31);
          else rd[c + 1] = 0u;
        }
      }
      const float dp_scale = kDrop ? p.scale * p.drop.inv_keep : p.scale;
      const uint32_t lse2 = stats_u32 + (st * 256 + half * 64) * 4;
      const uint32_t dl = stats_u32 + (st * 256 + 128 + half * 64) * 4;
      uint32_t pp[32], pd[32];
#pragma unroll
      for (int c = 0; c < 64; c += 4) {
        const float4 l4 = lds128f(lse2 + c * 4);
        const float4 d4 = lds128f(dl + c * 4);
        float x0, x1, x2, x3, t0, t1, t2, t3;
        ffma2v(x0, x1, __uint_as_float(rs[c]), __uint_as_float(rs[c + 1]), p.scale_log2, p.scale_log2, l4.x, l4.y);
        ffma2v(x2, x3, __uint_as_float(rs[c + 2]), __uint_as_float(rs[c + 3]), p.scale_log2, p.scale_log2, l4.z, l4.w);
        ffma2v(t0, t1, __uint_as_float(rd[c]), __uint_as_float(rd[c + 1]), dp_scale, dp_scale, d4.x, d4.y);
        ffma2v(t2, t3, __uint_as_float(rd[c + 2]), __uint_as_float(rd[c + 3]), dp_scale, dp_scale, d4.z, d4.w);
        x0 = ex2_approx(x0);
        x1 = ex2_approx(x1);
        x2 = ex2_approx(x2);
        x3 = ex2_approx(x3);
        fmul2(t0, t1, t0, t1, x0, x1);
        fmul2(t2, t3, t2, t3, x2, x3);
        if constexpr (kDrop) {
          const uint32_t kb = keep[c >> 5] >> (c & 31);
          x0 = (kb & 1u) ? x0 : 0.f;
          x1 = (kb & 2u) ? x1 : 0.f;
          x2 = (kb & 4u) ? x2 : 0.f;
          x3 = (kb & 8u) ? x3 : 0.f;
        }
        pp[c >> 1] = pack2<kHalf>(x0, x1);
        pp[(c >> 1) + 1] = pack2<kHalf>(x2, x3);
        pd[c >> 1] = pack2<kHalf>(t0, t1);
        pd[(c >> 1) + 1] = pack2<kHalf>(t2, t3);
      }
      // over the dP^T columns this warp has just read (no other warp touches them): P^T then dS^T
      tmem_st_32x32(tmem_b + lane_off + half * 64, pp);
      tmem_st_32x32(tmem_b + lane_off + half * 64 + 32, pd);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_ready);
    }
    mbar_wait(all_done, 0);
    tc_fence_after();
    const int64_t orow = ((int64_t)(kv_row0 + row) * p.hk + hk) * D + half * (D / 2);
    const bool row_ok = jt * 128 + row < k_len;
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      uint16_t* out = (which == 0 ? p.dv : p.dk) + orow;
      const float osc = (kDrop && which == 0) ? p.drop.inv_keep : 1.f;       // dV = (mask o P / keep)^T dO
      const uint32_t src = (which == 0 ? tmem_dv : tmem_dk) + lane_off + half * (D / 2);
#pragma unroll
      for (int c = 0; c < D / 64; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(src + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 v;
          v.x = pack2<kHalf>(__uint_as_float(r[i]) * osc, __uint_as_float(r[i + 1]) * osc);
          v.y = pack2<kHalf>(__uint_as_float(r[i + 2]) * osc, __uint_as_float(r[i + 3]) * osc);
          v.z = pack2<kHalf>(__uint_as_float(r[i + 4]) * osc, __uint_as_float(r[i + 5]) * osc);
          v.w = pack2<kHalf>(__uint_as_float(r[i + 6]) * osc, __uint_as_float(r[i + 7]) * osc);
          if (row_ok) *reinterpret_cast<uint4*>(out + c * 32 + i) = v;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_a, kTmemCols);
}

}  // namespace fab
}  // namespace pdb
