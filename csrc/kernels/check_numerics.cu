// A2: numerical check of one tensor in a single pass - counts of NaN / Inf / zero elements and max / min / mean of the
// finite ones, all left on the device (the host decides whether and when to read them).  Behind FLAGS_check_nan_inf and
// paddle.amp.debugging.check_numerics (reference: paddle/phi/kernels/gpu/check_numerics_kernel.cu, which also reduces
// per block and then prints from the device; here the Python layer formats the log line).
#include "common.cuh"

namespace pdb {

// stats: int64[3] = {nan, inf, zero};  values: float[3] = {max, min, sum} over finite elements (mean = sum / finite)
template <typename T>
__global__ void __launch_bounds__(256) check_numerics_kernel(const T* __restrict__ x, int64_t n, unsigned long long* __restrict__ stats,
                                                            float* __restrict__ values) {
  __shared__ float red[32];
  __shared__ unsigned long long cnt[3];
  if (threadIdx.x < 3) cnt[threadIdx.x] = 0;
  __syncthreads();
  unsigned int nan = 0, inf = 0, zero = 0;
  float mx = -INFINITY, mn = INFINITY, sum = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = to_f32<T>(x[i]);
    if (isnan(v)) ++nan;
    else if (isinf(v)) ++inf;
    else {
      if (v == 0.f) ++zero;
      mx = fmaxf(mx, v);
      mn = fminf(mn, v);
      sum += v;
    }
  }
  if (nan) atomicAdd(&cnt[0], (unsigned long long)nan);
  if (inf) atomicAdd(&cnt[1], (unsigned long long)inf);
  if (zero) atomicAdd(&cnt[2], (unsigned long long)zero);
  const float bmx = block_max(mx, red);
  const float bmn = -block_max(-mn, red);
  const float bsum = block_sum(sum, red);
  __syncthreads();
  if (threadIdx.x < 3 && cnt[threadIdx.x]) atomicAdd(&stats[threadIdx.x], cnt[threadIdx.x]);
  if (threadIdx.x == 0) {
    // float atomic max / min through the ordered-int trick (values start at -inf / +inf)
    if (bmx > -INFINITY) {
      if (bmx >= 0.f) atomicMax(reinterpret_cast<int*>(&values[0]), __float_as_int(bmx));
      else atomicMin(reinterpret_cast<unsigned int*>(&values[0]), __float_as_uint(bmx));
    }
    if (bmn < INFINITY) {
      if (bmn >= 0.f) atomicMin(reinterpret_cast<int*>(&values[1]), __float_as_int(bmn));
      else atomicMax(reinterpret_cast<unsigned int*>(&values[1]), __float_as_uint(bmn));
    }
    atomicAdd(&values[2], bsum);
  }
}

__global__ void check_numerics_init_kernel(unsigned long long* stats, float* values) {
  if (threadIdx.x < 3) stats[threadIdx.x] = 0;
  if (threadIdx.x == 0) {
    values[0] = -INFINITY;
    values[1] = INFINITY;
    values[2] = 0.f;
  }
}

}  // namespace pdb

using namespace pdb;

// stats: device int64[3] (nan, inf, zero counts), values: device float[3] (max, min, SUM of the finite elements).
PDB_EXPORT int pdb200_check_numerics(const void* x, int64_t numel, int dtype, int64_t* stats, float* values, void* stream) {
  PDB_REQUIRE(stats != nullptr && values != nullptr, "check_numerics: output buffers required");
  cudaStream_t st = (cudaStream_t)stream;
  check_numerics_init_kernel<<<1, 32, 0, st>>>(reinterpret_cast<unsigned long long*>(stats), values);
  if (numel > 0) {
    int64_t blocks = (numel + 256 * 8 - 1) / (256 * 8);
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    PDB_DISPATCH_FLOAT(dtype, T, (check_numerics_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(
                                     (const T*)x, numel, reinterpret_cast<unsigned long long*>(stats), values)));
  }
  return check_launch("check_numerics");
}
