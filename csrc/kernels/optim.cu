// K12/K15 + AMP helpers: multi-tensor AdamW, multi-tensor squared-L2-norm, check_finite_and_unscale for sm_100a.
//
// Replaces in the reference: paddle/phi/kernels/gpu/adamw_kernel.cu:35-372 (one launch per parameter + a 1-thread
// beta-pow kernel), fused_adam_kernel.cu + funcs/multi_tensor_apply.h (48-tensor / 320-block launches),
// squared_l2_norm_kernel.cu (one launch per parameter), amp_kernel.cu (check_finite_and_unscale).
// Design: ONE launch per optimizer step regardless of parameter count.  The host uploads a descriptor table
// (pointers, sizes, per-tensor lr ratio / decay switch, dtypes) and a chunk map; each CTA owns one 16K-element
// chunk of one tensor.  Gradient scaling (1/loss_scale x clip coefficient) and the found_inf skip are read from
// device memory, so the step never synchronises with the host.  Bias-correction powers are host scalars.
#include "common.cuh"

#include <atomic>

namespace pdb {

struct MTDesc {          // 64 bytes, mirrored by paddle_b200/ops/multi_tensor.py (numpy structured dtype)
  void* p;               // parameter (bf16/f16/f32)           | l2norm/unscale: tensor
  void* g;               // gradient (bf16/f16/f32)
  float* m;              // moment1 fp32
  float* v;              // moment2 fp32
  float* master;         // fp32 master weight or null
  int64_t numel;
  float lr_ratio;        // per-tensor learning-rate multiplier
  float decay;           // 1 = apply decoupled weight decay, 0 = skip
  int32_t p_dtype;
  int32_t g_dtype;
};
static_assert(sizeof(MTDesc) == 64, "MTDesc layout is part of the ABI");

constexpr int kChunk = 16384;   // elements per CTA
constexpr int kThreads = 256;

__device__ __forceinline__ float load_any(const void* p, int64_t i, int dt) {
  if (dt == PDB_DTYPE_F32) return ((const float*)p)[i];
  if (dt == PDB_DTYPE_BF16) return __bfloat162float(((const __nv_bfloat16*)p)[i]);
  return __half2float(((const __half*)p)[i]);
}
__device__ __forceinline__ void store_any(void* p, int64_t i, int dt, float v) {
  if (dt == PDB_DTYPE_F32) ((float*)p)[i] = v;
  else if (dt == PDB_DTYPE_BF16) ((__nv_bfloat16*)p)[i] = __float2bfloat16_rn(v);
  else ((__half*)p)[i] = __float2half_rn(v);
}
// 4 consecutive elements (i % 4 == 0, pointer 16B/8B aligned for the dtype)
__device__ __forceinline__ void load4(const void* p, int64_t i, int dt, float (&o)[4]) {
  if (dt == PDB_DTYPE_F32) {
    const float4 t = *reinterpret_cast<const float4*>((const float*)p + i);
    o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
  } else {
    const uint2 t = *reinterpret_cast<const uint2*>((const uint16_t*)p + i);
    if (dt == PDB_DTYPE_BF16) {
      const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&t.x), b = *reinterpret_cast<const __nv_bfloat162*>(&t.y);
      o[0] = __low2float(a); o[1] = __high2float(a); o[2] = __low2float(b); o[3] = __high2float(b);
    } else {
      const __half2 a = *reinterpret_cast<const __half2*>(&t.x), b = *reinterpret_cast<const __half2*>(&t.y);
      o[0] = __low2float(a); o[1] = __high2float(a); o[2] = __low2float(b); o[3] = __high2float(b);
    }
  }
}
__device__ __forceinline__ void store4(void* p, int64_t i, int dt, const float (&o)[4]) {
  if (dt == PDB_DTYPE_F32) {
    *reinterpret_cast<float4*>((float*)p + i) = make_float4(o[0], o[1], o[2], o[3]);
  } else {
    uint2 t;
    if (dt == PDB_DTYPE_BF16) {
      __nv_bfloat162 a = __floats2bfloat162_rn(o[0], o[1]), b = __floats2bfloat162_rn(o[2], o[3]);
      t.x = *reinterpret_cast<uint32_t*>(&a); t.y = *reinterpret_cast<uint32_t*>(&b);
    } else {
      __half2 a = __floats2half2_rn(o[0], o[1]), b = __floats2half2_rn(o[2], o[3]);
      t.x = *reinterpret_cast<uint32_t*>(&a); t.y = *reinterpret_cast<uint32_t*>(&b);
    }
    *reinterpret_cast<uint2*>((uint16_t*)p + i) = t;
  }
}
__device__ __forceinline__ bool aligned_for(const void* p, int dt) {
  return ((uintptr_t)p % (dt == PDB_DTYPE_F32 ? 16 : 8)) == 0;
}

struct AdamHyper {
  float lr, beta1, beta2, eps, weight_decay, bc1 /* 1-beta1^t */, bc2_sqrt /* sqrt(1-beta2^t) */;
  const float* beta_pows;  // optional device {beta1^t, beta2^t}; overrides bc1/bc2_sqrt (keeps the step count on device)
};

__device__ __forceinline__ void adamw_one(float& p, float g, float& m, float& v, const AdamHyper& h, float lr,
                                          float decay_mul) {
  p *= decay_mul;
  m = h.beta1 * m + (1.f - h.beta1) * g;
  v = h.beta2 * v + (1.f - h.beta2) * g * g;
  const float denom = sqrtf(v) / h.bc2_sqrt + h.eps;
  p -= (lr / h.bc1) * (m / denom);
}

__global__ void __launch_bounds__(kThreads) adamw_multi_kernel(const MTDesc* __restrict__ table,
                                                               const int2* __restrict__ chunk_map, AdamHyper h,
                                                               const float* __restrict__ grad_scale,
                                                               const float* __restrict__ found_inf) {
  if (found_inf && *found_inf != 0.f) return;   // SkipUpdate (adamw_kernel.cu found_inf semantics)
  if (h.beta_pows) {
    h.bc1 = 1.f - h.beta_pows[0];
    h.bc2_sqrt = sqrtf(1.f - h.beta_pows[1]);
  }
  const int2 cm = chunk_map[blockIdx.x];
  const MTDesc d = table[cm.x];
  const int64_t begin = (int64_t)cm.y * kChunk;
  const int64_t end = begin + kChunk < d.numel ? begin + kChunk : d.numel;
  const float gs = grad_scale ? *grad_scale : 1.f;
  const float lr = h.lr * d.lr_ratio;
  const float decay_mul = 1.f - lr * h.weight_decay * d.decay;
  const bool vec = aligned_for(d.p, d.p_dtype) && aligned_for(d.g, d.g_dtype) && ((uintptr_t)d.m % 16 == 0) &&
                   ((uintptr_t)d.v % 16 == 0) && (!d.master || (uintptr_t)d.master % 16 == 0);
  if (vec) {
    const int64_t end4 = begin + ((end - begin) & ~3ll);
    for (int64_t i = begin + threadIdx.x * 4; i < end4; i += kThreads * 4) {
      float p[4], g[4], m[4], v[4];
      if (d.master) load4(d.master, i, PDB_DTYPE_F32, p); else load4(d.p, i, d.p_dtype, p);
      load4(d.g, i, d.g_dtype, g);
      load4(d.m, i, PDB_DTYPE_F32, m);
      load4(d.v, i, PDB_DTYPE_F32, v);
#pragma unroll
      for (int j = 0; j < 4; ++j) adamw_one(p[j], g[j] * gs, m[j], v[j], h, lr, decay_mul);
      store4(d.m, i, PDB_DTYPE_F32, m);
      store4(d.v, i, PDB_DTYPE_F32, v);
      if (d.master) store4(d.master, i, PDB_DTYPE_F32, p);
      store4(d.p, i, d.p_dtype, p);
    }
    for (int64_t i = end4 + threadIdx.x; i < end; i += kThreads) {
      float p = d.master ? d.master[i] : load_any(d.p, i, d.p_dtype);
      float m = d.m[i], v = d.v[i];
      adamw_one(p, load_any(d.g, i, d.g_dtype) * gs, m, v, h, lr, decay_mul);
      d.m[i] = m; d.v[i] = v;
      if (d.master) d.master[i] = p;
      store_any(d.p, i, d.p_dtype, p);
    }
  } else {
    for (int64_t i = begin + threadIdx.x; i < end; i += kThreads) {
      float p = d.master ? d.master[i] : load_any(d.p, i, d.p_dtype);
      float m = d.m[i], v = d.v[i];
      adamw_one(p, load_any(d.g, i, d.g_dtype) * gs, m, v, h, lr, decay_mul);
      d.m[i] = m; d.v[i] = v;
      if (d.master) d.master[i] = p;
      store_any(d.p, i, d.p_dtype, p);
    }
  }
}

// Momentum / SGD share the table: m = velocity.
__global__ void __launch_bounds__(kThreads) momentum_multi_kernel(const MTDesc* __restrict__ table,
                                                                  const int2* __restrict__ chunk_map, float lr,
                                                                  float mu, float l2_decay, int nesterov,
                                                                  const float* __restrict__ grad_scale,
                                                                  const float* __restrict__ found_inf) {
  if (found_inf && *found_inf != 0.f) return;
  const int2 cm = chunk_map[blockIdx.x];
  const MTDesc d = table[cm.x];
  const int64_t begin = (int64_t)cm.y * kChunk;
  const int64_t end = begin + kChunk < d.numel ? begin + kChunk : d.numel;
  const float gs = grad_scale ? *grad_scale : 1.f;
  const float lrr = lr * d.lr_ratio;
  for (int64_t i = begin + threadIdx.x; i < end; i += kThreads) {
    float p = d.master ? d.master[i] : load_any(d.p, i, d.p_dtype);
    float g = load_any(d.g, i, d.g_dtype) * gs + l2_decay * d.decay * p;
    if (d.m) {
      float vel = mu * d.m[i] + g;
      d.m[i] = vel;
      p -= nesterov ? (g + mu * vel) * lrr : lrr * vel;
    } else {
      p -= lrr * g;
    }
    if (d.master) d.master[i] = p;
    store_any(d.p, i, d.p_dtype, p);
  }
}

// Sum of squares of every table entry's `g` tensor: per-chunk partial -> scratch, then one CTA folds.
__global__ void __launch_bounds__(kThreads) l2norm_partial_kernel(const MTDesc* __restrict__ table,
                                                                  const int2* __restrict__ chunk_map,
                                                                  float* __restrict__ partials) {
  __shared__ float red[32];
  const int2 cm = chunk_map[blockIdx.x];
  const MTDesc d = table[cm.x];
  const int64_t begin = (int64_t)cm.y * kChunk;
  const int64_t end = begin + kChunk < d.numel ? begin + kChunk : d.numel;
  float acc = 0.f;
  if (aligned_for(d.g, d.g_dtype)) {
    const int64_t end4 = begin + ((end - begin) & ~3ll);
    for (int64_t i = begin + threadIdx.x * 4; i < end4; i += kThreads * 4) {
      float g[4];
      load4(d.g, i, d.g_dtype, g);
      acc += g[0] * g[0] + g[1] * g[1] + g[2] * g[2] + g[3] * g[3];
    }
    for (int64_t i = end4 + threadIdx.x; i < end; i += kThreads) {
      const float g = load_any(d.g, i, d.g_dtype);
      acc += g * g;
    }
  } else {
    for (int64_t i = begin + threadIdx.x; i < end; i += kThreads) {
      const float g = load_any(d.g, i, d.g_dtype);
      acc += g * g;
    }
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
}

__global__ void __launch_bounds__(1024) fold_sum_kernel(const float* __restrict__ partials, int n,
                                                       float* __restrict__ out, int accumulate) {
  __shared__ float red[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += partials[i];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) out[0] = accumulate ? out[0] + acc : acc;
}

// g *= inv_scale (read from device), found_inf |= any non-finite.  Table `g` tensors, in place.
__global__ void __launch_bounds__(kThreads) unscale_check_kernel(const MTDesc* __restrict__ table,
                                                                 const int2* __restrict__ chunk_map,
                                                                 const float* __restrict__ inv_scale,
                                                                 float* __restrict__ found_inf) {
  const int2 cm = chunk_map[blockIdx.x];
  const MTDesc d = table[cm.x];
  const int64_t begin = (int64_t)cm.y * kChunk;
  const int64_t end = begin + kChunk < d.numel ? begin + kChunk : d.numel;
  const float s = *inv_scale;
  bool bad = false;
  for (int64_t i = begin + threadIdx.x; i < end; i += kThreads) {
    const float g = load_any(d.g, i, d.g_dtype);
    bad |= !isfinite(g);
    store_any(d.g, i, d.g_dtype, g * s);
  }
  if (__syncthreads_or(bad) && threadIdx.x == 0) *found_inf = 1.f;
}

}  // namespace pdb

using namespace pdb;

PDB_EXPORT int pdb200_mt_chunk_elems() { return kChunk; }

PDB_EXPORT int pdb200_adamw_multi(const void* table, const void* chunk_map, int n_chunks, float lr, float beta1,
                                  float beta2, float eps, float weight_decay, float beta1_pow, float beta2_pow,
                                  const float* beta_pows_dev, const float* grad_scale, const float* found_inf,
                                  void* stream) {
  if (n_chunks <= 0) return 0;
  AdamHyper h{lr, beta1, beta2, eps, weight_decay, 1.f - beta1_pow, sqrtf(1.f - beta2_pow), beta_pows_dev};
  adamw_multi_kernel<<<n_chunks, kThreads, 0, (cudaStream_t)stream>>>((const MTDesc*)table, (const int2*)chunk_map, h,
                                                                     grad_scale, found_inf);
  return check_launch("adamw_multi");
}

PDB_EXPORT int pdb200_momentum_multi(const void* table, const void* chunk_map, int n_chunks, float lr, float mu,
                                     float l2_decay, int nesterov, const float* grad_scale, const float* found_inf,
                                     void* stream) {
  if (n_chunks <= 0) return 0;
  momentum_multi_kernel<<<n_chunks, kThreads, 0, (cudaStream_t)stream>>>((const MTDesc*)table, (const int2*)chunk_map,
                                                                        lr, mu, l2_decay, nesterov, grad_scale, found_inf);
  return check_launch("momentum_multi");
}

// scratch: >= n_chunks floats.  out[0] = sum of squares (+= when accumulate).
PDB_EXPORT int pdb200_l2norm_multi(const void* table, const void* chunk_map, int n_chunks, float* scratch,
                                   int64_t scratch_floats, float* out, int accumulate, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  PDB_REQUIRE(n_chunks >= 0 && scratch_floats >= n_chunks, "l2norm_multi: scratch holds %lld floats, need %d",
              (long long)scratch_floats, n_chunks);
  if (n_chunks > 0) {
    l2norm_partial_kernel<<<n_chunks, kThreads, 0, st>>>((const MTDesc*)table, (const int2*)chunk_map, scratch);
    if (int e = check_launch("l2norm_partial")) return e;
  }
  fold_sum_kernel<<<1, 1024, 0, st>>>(scratch, n_chunks, out, accumulate);
  return check_launch("l2norm_fold");
}

PDB_EXPORT int pdb200_unscale_check_multi(const void* table, const void* chunk_map, int n_chunks,
                                          const float* inv_scale, float* found_inf, void* stream) {
  if (n_chunks <= 0) return 0;
  unscale_check_kernel<<<n_chunks, kThreads, 0, (cudaStream_t)stream>>>((const MTDesc*)table, (const int2*)chunk_map,
                                                                       inv_scale, found_inf);
  return check_launch("unscale_check_multi");
}

// ---------------------------------------------------------------------------------------------------------------
// pdb200_l2norm_multi_static: the same sum of squares as pdb200_l2norm_multi, for callers that cannot hand over a
// scratch buffer because they cannot allocate device memory at the point of the call: a plain C host program that
// links this library without a tensor framework (so there is no caching allocator to ask), or a step loop that has
// already handed all free memory to its own pools and must not let cudaMalloc synchronise the device in the middle
// of a step.  The per-chunk partial sums live in a fixed array inside this library instead.
//
//  * Limit: kStaticScratchChunks chunks (1 Mi chunks x 16384 elements = 17.2 G elements, 4 MiB of device memory
//    per device that loads the library).  A larger n_chunks is refused with an error; nothing is launched.
//  * Set-up: the first call on each device resolves the array's device address (cudaGetSymbolAddress, with that
//    device current) and caches it.  That lookup may initialise the module on the device, so make the first call
//    on every device before any stream capture begins.
//  * Nothing has to start at zero: the first kernel overwrites partials[0, n_chunks) before the second kernel reads
//    exactly that range, so whatever the previous call left behind is never observed.
//  * The array is shared by every caller on the device.  Two calls whose kernels can overlap (different streams
//    without an ordering between them, or two host threads on different streams) race on it - order them on one
//    stream or with an event.  Calls on the same stream are safe back to back.
//  * This entry point performs no allocation of any kind (no cudaMalloc/cudaFree, no library allocator).
// ---------------------------------------------------------------------------------------------------------------
namespace {
constexpr int kStaticScratchChunks = 1 << 20;
constexpr int kMaxDevices = 64;
static __device__ __align__(256) float g_l2norm_static_partials[kStaticScratchChunks];
std::atomic<float*> g_static_partials_addr[kMaxDevices];
}  // namespace

PDB_EXPORT int pdb200_l2norm_multi_static_max_chunks() { return kStaticScratchChunks; }

PDB_EXPORT int pdb200_l2norm_multi_static(const void* table, const void* chunk_map, int n_chunks, float* out,
                                          int accumulate, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  PDB_REQUIRE(n_chunks >= 0 && n_chunks <= kStaticScratchChunks,
              "l2norm_multi_static: %d chunks requested, the built-in scratch holds at most %d (use "
              "pdb200_l2norm_multi with caller-provided scratch for larger inputs)",
              n_chunks, kStaticScratchChunks);
  int dev = -1;
  PDB_CUDA_OK(cudaGetDevice(&dev));
  PDB_REQUIRE(dev >= 0 && dev < kMaxDevices, "l2norm_multi_static: device ordinal %d out of range", dev);
  float* partials = g_static_partials_addr[dev].load(std::memory_order_acquire);
  if (partials == nullptr) {
    // once per device (a concurrent first call resolves the same address, so the race is benign)
    void* addr = nullptr;
    PDB_CUDA_OK(cudaGetSymbolAddress(&addr, g_l2norm_static_partials));
    PDB_REQUIRE(addr != nullptr, "l2norm_multi_static: cudaGetSymbolAddress returned a null address");
    partials = (float*)addr;
    g_static_partials_addr[dev].store(partials, std::memory_order_release);
  }
  if (n_chunks > 0) {
    l2norm_partial_kernel<<<n_chunks, kThreads, 0, st>>>((const MTDesc*)table, (const int2*)chunk_map, partials);
    if (int e = check_launch("l2norm_partial(static)")) return e;
  }
  fold_sum_kernel<<<1, 1024, 0, st>>>(partials, n_chunks, out, accumulate);
  return check_launch("l2norm_fold(static)");
}
