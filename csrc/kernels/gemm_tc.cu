// K23 / K13: persistent warp-specialised bf16 GEMM on tcgen05 tensor cores for sm_100a.
//
//   C[M,N] (=|+=) act(op(A) op(B) + bias),  fp32 accumulation in TMEM.
//
// Replaces in the reference: cuBLASLt calls behind `linear` / fused_gemm_epilogue
// (paddle/phi/kernels/funcs/blas/blaslt_impl.cu.h, fusion/gpu/fused_gemm_epilogue_kernel.cu:63-125) and the
// accumulate-into-fp32 weight-gradient GEMM of fused_linear_param_grad_add_kernel.cu:35-289 (beta = 1 path).
//
// Structure (one CTA per SM, persistent over output tiles of 128 x BN):
//   warp 0   TMA producer: cp.async.bulk.tensor loads of A/B k-slabs (128-B swizzle) into a STAGES-deep smem ring
//   warp 1   MMA issuer:   one elected thread issues tcgen05.mma (M=128, N=BN, K=16) x4 per slab, commits the slab's
//                          "empty" mbarrier, and after the last slab the accumulator's "full" mbarrier
//   warp 2   TMEM allocator (2 x BN fp32 columns: the accumulator is double-buffered so the epilogue of tile i
//                          overlaps the MMAs of tile i+1)
//   warps 4-7 epilogue:    tcgen05.ld 32 lanes x 32 columns -> registers -> bias/activation/convert -> global
// Both operand majors are supported natively (descriptor major bits + matching TMA boxes), so forward (X W),
// dgrad (dY W^T) and wgrad (X^T dY) all run without a transpose copy:
//   A: [M,K] row-major (K-major)  or  [K,M] row-major (M-major, trans_a)
//   B: [K,N] row-major (N-major)  or  [N,K] row-major (K-major, trans_b)
#include "common.cuh"
#include "tc.cuh"

namespace pdb {
namespace tc {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

int encode_tmap_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t row_stride_bytes,
                   uint32_t box_rows, uint32_t box_cols, int elem_bytes, bool swizzle128) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return -1;
  }
  // Driver API calls need a current context on THIS thread; autograd's backward threads only ever used the runtime
  // API lazily, so bind the primary context once per thread (cuTensorMapEncodeTiled returned INVALID_CONTEXT there).
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    cudaFree(nullptr);
    ctx_bound = true;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = fn(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): rows=%llu cols=%llu stride=%llu box=%ux%u", (int)r,
              (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)row_stride_bytes, box_rows, box_cols);
    return -1;
  }
  return 0;
}

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kThreads = 256;
constexpr int kGroupM = 8;

template <int BN> struct Cfg {
  static constexpr int kStages = BN == 256 ? 4 : 6;
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTmemCols = 2 * BN;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
};

__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int& m_blk, int& n_blk) {
  const int per_group = kGroupM * num_n;
  const int group = tile / per_group;
  const int first_m = group * kGroupM;
  const int gsz = min(num_m - first_m, kGroupM);
  const int in_group = tile - group * per_group;
  m_blk = first_m + in_group % gsz;
  n_blk = in_group / gsz;
}

// Communication fused into the GEMM (K5 / K6).  mode 1 = all-gather -> GEMM: the M dimension is the concatenation of
// every rank's [rows_per_rank, K] activation shard living in symmetric memory; the TMA producer reads each A tile
// straight from the owning peer through that peer's tensor map (no gathered copy, local tiles first).  mode 2 = GEMM
// -> reduce-scatter: the epilogue stores each partial tile into the OWNER rank's staging slab [src rank][rows][N]
// over NVLink as soon as it leaves TMEM; a small follow-up kernel on the owner sums the slabs.
struct CommArgs {
  int mode;
  int nranks;
  int rank;
  int rows_per_rank;
  int m_rot;                    // tile-row rotation so every rank starts on a different chunk
  uint64_t out_ptr[8];          // mode 2: staging base address on each owner (as mapped in this process)
  CUtensorMap tm_a_peer[8];     // mode 1: A shard of each peer
};

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == 1) return 0.5f * v * (1.f + erff(v * 0.70710678118654752f));
  if (act == 2) return fmaxf(v, 0.f);
  return v;
}

template <bool A_MN, bool B_MN, int BN, bool OUT_F32>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, void* __restrict__ Cptr,
               const __nv_bfloat16* __restrict__ bias, int M, int N, int K, int64_t ldc, float beta, int act,
               const __grid_constant__ CommArgs comm) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + C::kStages;
  uint64_t* tmem_full = bars + 2 * C::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_m = (M + BM - 1) / BM, num_n = (N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_k = (K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_a);
    prefetch_tmap(&tm_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < C::kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);   // one arrive per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_ptr_smem, C::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ================================================================= TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        tile_coords(tile, num_m, num_n, m_blk, n_blk);
        if (comm.m_rot) m_blk = (m_blk + comm.m_rot) % num_m;
        const int m0 = m_blk * BM, n0 = n_blk * BN;
        const int a_owner = comm.mode == 1 ? m0 / comm.rows_per_rank : 0;
        const int a_row = comm.mode == 1 ? m0 - a_owner * comm.rows_per_rank : m0;
        const CUtensorMap* tma_a = comm.mode == 1 ? &comm.tm_a_peer[a_owner] : &tm_a;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + C::kABytes;
          mbar_expect_tx(&full_bar[stage], C::kStageBytes);
          const int k0 = kb * BK;
          if (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * (BK * 128), &tm_a, &full_bar[stage], m0 + j * 64, k0);
          } else {
            tma_load_2d(sa, tma_a, &full_bar[stage], k0, a_row);
          }
          if (B_MN) {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (BK * 128), &tm_b, &full_bar[stage], n0 + j * 64, k0);
          } else {
            tma_load_2d(sb, &tm_b, &full_bar[stage], k0, n0);
          }
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
          const uint32_t sb = sa + C::kABytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = A_MN ? make_smem_desc(sa + k * 2048, BK * 128, 1024) : make_smem_desc(sa + k * 32, 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc(sb + k * 2048, BK * 128, 1024) : make_smem_desc(sb + k * 32, 16, 1024);
            umma_f16(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);          // smem slab reusable once these MMAs retire
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[acc]);              // accumulator complete
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // ================================================================= epilogue
    const int q = warp & 3;                        // TMEM lane quarter this warp may touch
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      tile_coords(tile, num_m, num_n, m_blk, n_blk);
      if (comm.m_rot) m_blk = (m_blk + comm.m_rot) % num_m;
      const int row = m_blk * BM + q * 32 + lane;
      const int n0 = n_blk * BN;
      // mode 2: this row lands in the owner's staging slab, slot = my rank
      void* c_base = Cptr;
      int64_t c_row = row;
      if (comm.mode == 2) {
        const int owner = (m_blk * BM) / comm.rows_per_rank;
        c_base = reinterpret_cast<void*>(comm.out_ptr[owner]);
        c_row = (int64_t)comm.rank * comm.rows_per_rank + (row - owner * comm.rows_per_rank);
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
          // all TMEM reads of this accumulator are done: hand it back before the global stores
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty[acc]);
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
            float* dst = reinterpret_cast<float*>(c_base) + c_row * ldc + col0;
            if (full && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                float4 o = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                if (beta != 0.f) {
                  const float4 p = *reinterpret_cast<const float4*>(dst + j);
                  o.x += beta * p.x; o.y += beta * p.y; o.z += beta * p.z; o.w += beta * p.w;
                }
                *reinterpret_cast<float4*>(dst + j) = o;
              }
            } else {
              for (int j = 0; j < 32 && col0 + j < N; ++j) dst[j] = v[j] + (beta != 0.f ? beta * dst[j] : 0.f);
            }
          } else {
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(c_base) + c_row * ldc + col0;
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
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, C::kTmemCols);
}

template <bool A_MN, bool B_MN, int BN, bool OUT_F32>
static int launch(const void* a, const void* b, void* c, const void* bias, int M, int N, int K, float beta, int act,
                  cudaStream_t st, const CommArgs* comm_in = nullptr) {
  using C = Cfg<BN>;
  CUtensorMap tm_a, tm_b;
  static const CommArgs kNoComm{};
  const CommArgs& comm = comm_in ? *comm_in : kNoComm;
  // A: K-major [M][K] -> box 128 rows x 64 k ; M-major stored [K][M] -> box 64 k-rows x 64 m
  if (comm.mode == 1) {
    tm_a = comm.tm_a_peer[comm.rank];
  } else if (A_MN) {
    if (encode_tmap_2d(&tm_a, a, K, M, (uint64_t)M * 2, BK, 64, 2, true)) return -1;
  } else {
    if (encode_tmap_2d(&tm_a, a, M, K, (uint64_t)K * 2, BM, BK, 2, true)) return -1;
  }
  if (B_MN) {
    if (encode_tmap_2d(&tm_b, b, K, N, (uint64_t)N * 2, BK, 64, 2, true)) return -1;
  } else {
    if (encode_tmap_2d(&tm_b, b, N, K, (uint64_t)K * 2, BN, BK, 2, true)) return -1;
  }
  auto kern = gemm_tc_kernel<A_MN, B_MN, BN, OUT_F32>;
  static bool attr_set = false;
  if (!attr_set) {
    PDB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    attr_set = true;
  }
  const int num_tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int grid = num_tiles < sm_count() ? num_tiles : sm_count();
  kern<<<grid, kThreads, C::kSmemBytes, st>>>(tm_a, tm_b, c, (const __nv_bfloat16*)bias, M, N, K, (int64_t)N, beta, act, comm);
  return check_launch("gemm_tc");
}

template <bool A_MN, bool B_MN>
static int dispatch_bn(const void* a, const void* b, void* c, const void* bias, int M, int N, int K, int out_f32,
                       float beta, int act, cudaStream_t st) {
  const bool big = (N % 256 == 0) || N >= 1024;
  if (big) {
    return out_f32 ? launch<A_MN, B_MN, 256, true>(a, b, c, bias, M, N, K, beta, act, st)
                   : launch<A_MN, B_MN, 256, false>(a, b, c, bias, M, N, K, beta, act, st);
  }
  return out_f32 ? launch<A_MN, B_MN, 128, true>(a, b, c, bias, M, N, K, beta, act, st)
                 : launch<A_MN, B_MN, 128, false>(a, b, c, bias, M, N, K, beta, act, st);
}

}  // namespace tc
}  // namespace pdb

using namespace pdb;

// out_dtype: 0 = fp32 (beta may be 1 to accumulate into C), 1 = bf16.  act: 0 none, 1 gelu, 2 relu.  bias: bf16[N].
PDB_EXPORT int pdb200_gemm_bf16(const void* a, const void* b, void* c, const void* bias, int64_t M, int64_t N, int64_t K,
                                int trans_a, int trans_b, int out_dtype, float beta, int act, void* stream) {
  PDB_REQUIRE(M > 0 && N > 0 && K > 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "gemm: bad shape");
  PDB_REQUIRE(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0, "gemm: A and B must be 16-byte aligned");
  PDB_REQUIRE((trans_a ? M : K) % 8 == 0 && (trans_b ? K : N) % 8 == 0,
              "gemm: contiguous dims must be multiples of 8 (TMA 16-byte stride rule)");
  PDB_REQUIRE(out_dtype == 0 || (out_dtype == 1 && beta == 0.f), "gemm: beta != 0 needs fp32 output");
  cudaStream_t st = (cudaStream_t)stream;
  const bool a_mn = trans_a != 0;      // A stored [K][M]
  const bool b_mn = trans_b == 0;      // B stored [K][N]
  const int of = out_dtype == 0;
  if (a_mn && b_mn) return tc::dispatch_bn<true, true>(a, b, c, bias, (int)M, (int)N, (int)K, of, beta, act, st);
  if (a_mn && !b_mn) return tc::dispatch_bn<true, false>(a, b, c, bias, (int)M, (int)N, (int)K, of, beta, act, st);
  if (!a_mn && b_mn) return tc::dispatch_bn<false, true>(a, b, c, bias, (int)M, (int)N, (int)K, of, beta, act, st);
  return tc::dispatch_bn<false, false>(a, b, c, bias, (int)M, (int)N, (int)K, of, beta, act, st);
}


// ------------------------------------------------------------------------------------------------ K5 / K6 entry points
#include "nvl.cuh"
namespace pdb {
namespace nvl {
// same body as nvl_collectives.cu's barrier (kernels cannot be shared across translation units without -rdc)
static __global__ void gemm_barrier_kernel(Peers P, uint64_t pad_off, uint32_t epoch, unsigned long long spin_limit) {
  barrier_all(P, pad_off, epoch, spin_limit);
}
}
namespace tc {
// out[rows, N] = sum over src of staging[src][rows][N]   (bf16 in, fp32 sum, bf16 out)
__global__ void __launch_bounds__(256) rs_fold_kernel(const __nv_bfloat16* __restrict__ staging, __nv_bfloat16* __restrict__ out,
                                                     int64_t per_rank_elems, int nranks) {
  const int64_t nvec = per_rank_elems / 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int r = 0; r < nranks; ++r) {
      Vec<__nv_bfloat16> x;
      x.load_nc(staging + r * per_rank_elems + i * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += x.f(j);
    }
    Vec<__nv_bfloat16> o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.set(j, acc[j]);
    o.store(out + i * 8);
  }
}
// K4 tail: same fold, but the summed rows are stored into every rank's full output (all-gather half of the all-reduce).
struct OutPtrs { uint64_t p[8]; };
__global__ void __launch_bounds__(256) ar_fold_kernel(const __nv_bfloat16* __restrict__ staging, OutPtrs outs,
                                                     int64_t per_rank_elems, int nranks, int rank) {
  const int64_t nvec = per_rank_elems / 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int r = 0; r < nranks; ++r) {
      Vec<__nv_bfloat16> x;
      x.load_nc(staging + r * per_rank_elems + i * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += x.f(j);
    }
    Vec<__nv_bfloat16> o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.set(j, acc[j]);
    for (int r = 0; r < nranks; ++r) {
      const int dst = (rank + r) % nranks;   // start with the local copy, then walk the peers
      o.store((__nv_bfloat16*)outs.p[dst] + (int64_t)rank * per_rank_elems + i * 8);
    }
  }
}
}  // namespace tc
}  // namespace pdb

static pdb::nvl::Peers make_peers_(const uint64_t* ptrs, int nranks, int rank) {
  pdb::nvl::Peers P{};
  for (int i = 0; i < nranks && i < pdb::nvl::kMaxRanks; ++i) P.ptr[i] = ptrs[i];
  P.nranks = nranks;
  P.rank = rank;
  return P;
}
static const unsigned long long kGemmSpinLimit = 20000000ull;

// K5 all-gather -> GEMM.  Every rank's activation shard [rows_per_rank, K] (bf16, row-major) lives at byte offset a_off
// of its symmetric segment; C[nranks * rows_per_rank, N] = gather(A) @ B is written locally.  B: [K,N] or [N,K] (trans_b).
PDB_EXPORT int pdb200_gemm_bf16_ag(const uint64_t* peer_ptrs, int nranks, int rank, int64_t a_off, int64_t rows_per_rank,
                                   const void* b, void* c, const void* bias, int64_t N, int64_t K, int trans_b, int act,
                                   int64_t pad_off, int64_t epoch, void* stream) {
  using namespace pdb::tc;
  PDB_REQUIRE(nranks >= 1 && nranks <= 8, "gemm_ag: nranks %d out of range", nranks);
  PDB_REQUIRE(rows_per_rank % BM == 0, "gemm_ag: rows per rank (%lld) must be a multiple of %d", (long long)rows_per_rank, BM);
  PDB_REQUIRE(K % 8 == 0 && N % 8 == 0 && a_off % 16 == 0, "gemm_ag: K, N multiples of 8 and 16-byte aligned shard required");
  cudaStream_t st = (cudaStream_t)stream;
  CommArgs comm{};
  comm.mode = 1;
  comm.nranks = nranks;
  comm.rank = rank;
  comm.rows_per_rank = (int)rows_per_rank;
  comm.m_rot = (int)(rank * (rows_per_rank / BM));          // start on the local shard: remote latency hides behind it
  for (int p = 0; p < nranks; ++p)
    if (encode_tmap_2d(&comm.tm_a_peer[p], (const void*)(peer_ptrs[p] + a_off), rows_per_rank, K, (uint64_t)K * 2, BM, BK, 2, true)) return -1;
  const int M = (int)(rows_per_rank * nranks);
  pdb::nvl::Peers P = make_peers_(peer_ptrs, nranks, rank);
  pdb::nvl::gemm_barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)epoch, kGemmSpinLimit);   // shards are written
  int rc;
  const bool big = (N % 256 == 0) || N >= 1024;
  if (trans_b)
    rc = big ? launch<false, false, 256, false>(nullptr, b, c, bias, M, (int)N, (int)K, 0.f, act, st, &comm)
             : launch<false, false, 128, false>(nullptr, b, c, bias, M, (int)N, (int)K, 0.f, act, st, &comm);
  else
    rc = big ? launch<false, true, 256, false>(nullptr, b, c, bias, M, (int)N, (int)K, 0.f, act, st, &comm)
             : launch<false, true, 128, false>(nullptr, b, c, bias, M, (int)N, (int)K, 0.f, act, st, &comm);
  if (rc) return rc;
  pdb::nvl::gemm_barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kGemmSpinLimit);  // shards reusable
  return check_launch("gemm_ag");
}

// K6 GEMM -> reduce-scatter.  A: local [M, K] (M = nranks * rows_per_rank), B as above.  Partial tiles are pushed into
// the owner's staging slab at stage_off ([nranks][rows_per_rank][N] bf16 in every segment); out[rows_per_rank, N] (local,
// bf16) receives the sum.
PDB_EXPORT int pdb200_gemm_bf16_rs(const uint64_t* peer_ptrs, int nranks, int rank, const void* a, const void* b,
                                   int64_t stage_off, void* out, int64_t rows_per_rank, int64_t N, int64_t K, int trans_b,
                                   int64_t pad_off, int64_t epoch, void* stream) {
  using namespace pdb::tc;
  PDB_REQUIRE(nranks >= 1 && nranks <= 8, "gemm_rs: nranks %d out of range", nranks);
  PDB_REQUIRE(rows_per_rank % BM == 0, "gemm_rs: rows per rank (%lld) must be a multiple of %d", (long long)rows_per_rank, BM);
  PDB_REQUIRE(K % 8 == 0 && N % 8 == 0 && stage_off % 16 == 0 && ((uintptr_t)a & 15) == 0, "gemm_rs: alignment");
  cudaStream_t st = (cudaStream_t)stream;
  CommArgs comm{};
  comm.mode = 2;
  comm.nranks = nranks;
  comm.rank = rank;
  comm.rows_per_rank = (int)rows_per_rank;
  comm.m_rot = (int)(((rank + 1) % nranks) * (rows_per_rank / BM));   // spread the owners: each rank starts one chunk ahead
  for (int p = 0; p < nranks; ++p) comm.out_ptr[p] = peer_ptrs[p] + (uint64_t)stage_off;
  const int M = (int)(rows_per_rank * nranks);
  pdb::nvl::Peers P = make_peers_(peer_ptrs, nranks, rank);
  pdb::nvl::gemm_barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)epoch, kGemmSpinLimit);   // slabs free again
  int rc;
  const bool big = (N % 256 == 0) || N >= 1024;
  if (trans_b)
    rc = big ? launch<false, false, 256, false>(a, b, nullptr, nullptr, M, (int)N, (int)K, 0.f, 0, st, &comm)
             : launch<false, false, 128, false>(a, b, nullptr, nullptr, M, (int)N, (int)K, 0.f, 0, st, &comm);
  else
    rc = big ? launch<false, true, 256, false>(a, b, nullptr, nullptr, M, (int)N, (int)K, 0.f, 0, st, &comm)
             : launch<false, true, 128, false>(a, b, nullptr, nullptr, M, (int)N, (int)K, 0.f, 0, st, &comm);
  if (rc) return rc;
  pdb::nvl::gemm_barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kGemmSpinLimit);  // all partials landed
  const int64_t per = rows_per_rank * N;
  int64_t blocks = (per / 8 + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  rs_fold_kernel<<<(int)(blocks < 1 ? 1 : blocks), 256, 0, st>>>((const __nv_bfloat16*)(peer_ptrs[rank] + stage_off),
                                                             (__nv_bfloat16*)out, per, nranks);
  return check_launch("gemm_rs");
}

// K4 GEMM -> all-reduce (RowParallelLinear forward, ColumnParallelLinear dX).  Same tile-by-tile push of partial tiles to
// the owning rank as K6; the owner then folds its rows and stores the sum into EVERY rank's output, which lives at byte
// offset out_off ([M, N] bf16) of each symmetric segment.  Uses epochs epoch, epoch+1, epoch+2.
PDB_EXPORT int pdb200_gemm_bf16_ar(const uint64_t* peer_ptrs, int nranks, int rank, const void* a, const void* b,
                                   int64_t stage_off, int64_t out_off, int64_t rows_per_rank, int64_t N, int64_t K,
                                   int trans_b, int64_t pad_off, int64_t epoch, void* stream) {
  using namespace pdb::tc;
  PDB_REQUIRE(nranks >= 1 && nranks <= 8, "gemm_ar: nranks %d out of range", nranks);
  PDB_REQUIRE(rows_per_rank % BM == 0, "gemm_ar: rows per rank (%lld) must be a multiple of %d", (long long)rows_per_rank, BM);
  PDB_REQUIRE(K % 8 == 0 && N % 8 == 0 && stage_off % 16 == 0 && out_off % 16 == 0 && ((uintptr_t)a & 15) == 0,
              "gemm_ar: alignment");
  cudaStream_t st = (cudaStream_t)stream;
  CommArgs comm{};
  comm.mode = 2;
  comm.nranks = nranks;
  comm.rank = rank;
  comm.rows_per_rank = (int)rows_per_rank;
  comm.m_rot = (int)(((rank + 1) % nranks) * (rows_per_rank / BM));
  for (int p = 0; p < nranks; ++p) comm.out_ptr[p] = peer_ptrs[p] + (uint64_t)stage_off;
  const int M = (int)(rows_per_rank * nranks);
  pdb::nvl::Peers P = make_peers_(peer_ptrs, nranks, rank);
  pdb::nvl::gemm_barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)epoch, kGemmSpinLimit);   // slabs + outputs free
  int rc;
  const bool big = (N % 256 == 0) || N >= 1024;
  if (trans_b)
    rc = big ? launch<false, false, 256, false>(a, b, nullptr, nullptr, M, (int)N, (int)K, 0.f, 0, st, &comm)
             : launch<false, false, 128, false>(a, b, nullptr, nullptr, M, (int)N, (int)K, 0.f, 0, st, &comm);
  else
    rc = big ? launch<false, true, 256, false>(a, b, nullptr, nullptr, M, (int)N, (int)K, 0.f, 0, st, &comm)
             : launch<false, true, 128, false>(a, b, nullptr, nullptr, M, (int)N, (int)K, 0.f, 0, st, &comm);
  if (rc) return rc;
  pdb::nvl::gemm_barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kGemmSpinLimit);  // partials landed
  const int64_t per = rows_per_rank * N;
  int64_t blocks = (per / 8 + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  OutPtrs outs{};
  for (int p = 0; p < nranks; ++p) outs.p[p] = peer_ptrs[p] + (uint64_t)out_off;
  ar_fold_kernel<<<(int)(blocks < 1 ? 1 : blocks), 256, 0, st>>>((const __nv_bfloat16*)(peer_ptrs[rank] + stage_off), outs, per,
                                                             nranks, rank);
  if (int e = check_launch("gemm_ar fold")) return e;
  pdb::nvl::gemm_barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)(epoch + 2), kGemmSpinLimit);  // every row landed
  return check_launch("gemm_ar");
}
