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
#include <mutex>
#include <unordered_map>

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

// Encoded tensor maps are pure functions of (address, shape, box): cache them, a training step re-issues the same
// ~800 GEMM shapes on recurring allocator addresses.  A miss also runs the slow ownership check; every call runs the
// cheap foreign-range check, so a map is only ever built over, or used on, this GPU's own memory.
namespace {
struct TmapKey {
  const void* base;
  uint64_t rows, cols, stride;
  uint32_t box_rows, box_cols;
  int esz, swz, dev;
  bool operator==(const TmapKey& o) const {
    return base == o.base && rows == o.rows && cols == o.cols && stride == o.stride && box_rows == o.box_rows &&
           box_cols == o.box_cols && esz == o.esz && swz == o.swz && dev == o.dev;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    uint64_t h = (uint64_t)(uintptr_t)k.base * 0x9E3779B97F4A7C15ull;
    h ^= (k.rows * 0xC2B2AE3D27D4EB4Full) ^ (k.cols << 17) ^ (k.stride << 3) ^ ((uint64_t)k.box_rows << 40) ^
         ((uint64_t)k.box_cols << 50) ^ ((uint64_t)k.esz << 60) ^ ((uint64_t)k.swz << 59) ^ ((uint64_t)k.dev << 54);
    return (size_t)(h ^ (h >> 29));
  }
};
std::mutex g_tmap_mu;
std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_tmap_cache;
}  // namespace

int get_tmap_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t row_stride_bytes,
                uint32_t box_rows, uint32_t box_cols, int elem_bytes, bool swizzle128, const char* what) {
  if (require_local(base, (size_t)(rows * row_stride_bytes), what)) return -1;
  int dev = 0;
  cudaGetDevice(&dev);
  const TmapKey key{base, rows, cols, row_stride_bytes, box_rows, box_cols, elem_bytes, swizzle128 ? 1 : 0, dev};
  {
    std::lock_guard<std::mutex> g(g_tmap_mu);
    auto it = g_tmap_cache.find(key);
    if (it != g_tmap_cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  if (require_local_device(base, what)) return -1;
  if (encode_tmap_2d(out, base, rows, cols, row_stride_bytes, box_rows, box_cols, elem_bytes, swizzle128)) return -1;
  std::lock_guard<std::mutex> g(g_tmap_mu);
  if (g_tmap_cache.size() > 8192) g_tmap_cache.clear();
  g_tmap_cache.emplace(key, *out);
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

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == 1) return 0.5f * v * (1.f + erff(v * 0.70710678118654752f));
  if (act == 2) return fmaxf(v, 0.f);
  return v;
}

template <bool A_MN, bool B_MN, int BN, bool OUT_F32>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, void* __restrict__ Cptr,
               const __nv_bfloat16* __restrict__ bias, int M, int N, int K, int64_t ldc, float beta, int act) {
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
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        tile_coords(tile, num_m, num_n, m_blk, n_blk);
        const int m0 = m_blk * BM, n0 = n_blk * BN;
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
            tma_load_2d(sa, &tm_a, &full_bar[stage], k0, m0);
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
    if (elect_one()) {   // elect.sync, not a lane test: ptxas then knows one lane runs this and issues UTCHMMA / UTCBAR from uniform registers

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
      const int row = m_blk * BM + q * 32 + lane;
      const int n0 = n_blk * BN;
      void* c_base = Cptr;
      const int64_t c_row = row;
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
                  cudaStream_t st) {
  using C = Cfg<BN>;
  CUtensorMap tm_a, tm_b;
  // A: K-major [M][K] -> box 128 rows x 64 k ; M-major stored [K][M] -> box 64 k-rows x 64 m
  if (A_MN) {
    if (get_tmap_2d(&tm_a, a, K, M, (uint64_t)M * 2, BK, 64, 2, true, "gemm A")) return -1;
  } else {
    if (get_tmap_2d(&tm_a, a, M, K, (uint64_t)K * 2, BM, BK, 2, true, "gemm A")) return -1;
  }
  if (B_MN) {
    if (get_tmap_2d(&tm_b, b, K, N, (uint64_t)N * 2, BK, 64, 2, true, "gemm B")) return -1;
  } else {
    if (get_tmap_2d(&tm_b, b, N, K, (uint64_t)K * 2, BN, BK, 2, true, "gemm B")) return -1;
  }
  if (require_local(c, (size_t)M * N * (OUT_F32 ? 4 : 2), "gemm C")) return -1;
  auto kern = gemm_tc_kernel<A_MN, B_MN, BN, OUT_F32>;
  PDB_ENSURE_SMEM(kern, C::kSmemBytes);
  const int num_tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int grid = num_tiles < sm_count() ? num_tiles : sm_count();
  kern<<<grid, kThreads, C::kSmemBytes, st>>>(tm_a, tm_b, c, (const __nv_bfloat16*)bias, M, N, K, (int64_t)N, beta, act);
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
