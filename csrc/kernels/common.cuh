// Shared device helpers for the sm_100a kernel library (libpdb200_kernels.so).
// Everything here is SIMT-side plumbing: 16-byte vector access, warp/block reductions, dtype conversion,
// error plumbing for the C ABI.  Tensor-core / TMA / multimem helpers live in tc.cuh and nvl.cuh.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#define PDB_DTYPE_F32 0
#define PDB_DTYPE_BF16 1
#define PDB_DTYPE_F16 2

#define PDB_EXPORT extern "C" __attribute__((visibility("default")))

namespace pdb {

// ---- C-ABI error plumbing -------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}
#define PDB_CUDA_OK(expr)                                                  \
  do {                                                                     \
    cudaError_t _e = (expr);                                               \
    if (_e != cudaSuccess) {                                               \
      pdb::set_error("%s failed: %s", #expr, cudaGetErrorString(_e));      \
      return (int)_e;                                                      \
    }                                                                      \
  } while (0)
#define PDB_REQUIRE(cond, ...)            \
  do {                                    \
    if (!(cond)) {                        \
      pdb::set_error(__VA_ARGS__);        \
      return -1;                          \
    }                                     \
  } while (0)

int sm_count();  // cached multiProcessorCount of the current device
extern uint32_t* g_nvl_progress;   // mapped host word for the barrier-progress trace (runtime_glue.cu), nullptr = off

// Local-memory guard (runtime_glue.cu): 0 when [p, p+bytes) touches no registered mapping of another GPU's memory.
int require_local(const void* p, size_t bytes, const char* what);
// Slow check (cudaPointerGetAttributes): p is device memory owned by the current GPU.  Run once per new tensor map.
int require_local_device(const void* p, const char* what);

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (call site, device) - the attribute is per device.
#define PDB_ENSURE_SMEM(kern, bytes)                                                                       \
  do {                                                                                                     \
    static std::atomic<uint64_t> _pdb_done{0};                                                             \
    int _dev = 0;                                                                                          \
    PDB_CUDA_OK(cudaGetDevice(&_dev));                                                                     \
    const uint64_t _bit = 1ull << (_dev & 63);                                                             \
    if (!(_pdb_done.load(std::memory_order_acquire) & _bit)) {                                             \
      PDB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
      _pdb_done.fetch_or(_bit, std::memory_order_release);                                                 \
    }                                                                                                      \
  } while (0)

// ---- conversion -----------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// ---- 16-byte vectors ------------------------------------------------------------------------------------
// Vec<T> holds 16 bytes worth of T: 4 floats or 8 half/bf16.  Loads/stores are single LDG.128/STG.128.
template <typename T> struct Vec {
  static constexpr int N = 16 / sizeof(T);
  union {
    uint4 raw;
    T v[N];
  };
  __device__ __forceinline__ void load(const T* p) { raw = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void load_nc(const T* p) {  // streaming read, do not pollute L1
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(raw.x), "=r"(raw.y), "=r"(raw.z), "=r"(raw.w) : "l"(p));
  }
  __device__ __forceinline__ void store(T* p) const { *reinterpret_cast<uint4*>(p) = raw; }
  __device__ __forceinline__ float f(int i) const { return to_f32<T>(v[i]); }
  __device__ __forceinline__ void set(int i, float x) { v[i] = from_f32<T>(x); }
};

// ---- reductions -----------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// Block-wide sum broadcast to every thread.  `red` is >= 32 floats of shared memory; safe to call repeatedly.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from the previous call's readers
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.f;
  return warp_sum(r);
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : -INFINITY;
  return warp_max(r);
}

// Dispatch a functor template over the three float dtypes.
#define PDB_DISPATCH_FLOAT(dtype, T, ...)                                 \
  switch (dtype) {                                                        \
    case PDB_DTYPE_F32: { using T = float; __VA_ARGS__; break; }          \
    case PDB_DTYPE_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; } \
    case PDB_DTYPE_F16: { using T = __half; __VA_ARGS__; break; }         \
    default: pdb::set_error("unsupported dtype code %d", (int)(dtype)); return -1; \
  }

}  // namespace pdb
