// K8 (local part): fused softmax-cross-entropy forward/backward, usable both as the plain single-GPU loss and as
// the per-rank half of the vocab-parallel loss (ParallelCrossEntropy).
//
// Replaces in the reference: paddle/phi/kernels/gpu/cross_entropy_kernel.cu:82-1320 (hard-label path) and
// paddle/fluid/operators/collective/c_softmax_with_cross_entropy_op.cu:47-114,133-340 (>=7 launches over N x V/n
// plus three N x 1 all-reduces).  Here one pass over the logits per row produces the online-softmax statistics
// (max, sum-exp, picked target logit); the cross-rank step only ever moves 3 floats per row (see nvl_ce.cu for the
// in-kernel exchange), and the backward writes (softmax - onehot) * dloss in one pass.
#include "common.cuh"

namespace pdb {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// stats[row] = {local max, local sum exp(x - max), target logit (0 if the label is not in this shard)}
template <typename T>
__global__ void __launch_bounds__(256) ce_stats_kernel(const T* __restrict__ logits, const int64_t* __restrict__ labels,
                                                      float* __restrict__ stats, float* __restrict__ loss,
                                                      float* __restrict__ lse, int64_t rows, int V, int64_t ld,
                                                      int64_t vocab_start, int64_t ignore_index) {
  constexpr int N = Vec<T>::N;
  __shared__ float red[32];
  const bool vec_ok = (V % N == 0) && (ld % N == 0) && ((uintptr_t)logits % 16 == 0);
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* xr = logits + row * ld;
    float m = -INFINITY, s = 0.f;
    if (vec_ok) {
      const int nv = V / N;
      for (int vi = threadIdx.x; vi < nv; vi += blockDim.x) {
        Vec<T> x;
        x.load_nc(xr + vi * N);
        float vm = x.f(0);
#pragma unroll
        for (int j = 1; j < N; ++j) vm = fmaxf(vm, x.f(j));
        const float mn = fmaxf(m, vm);
        float add = 0.f;
#pragma unroll
        for (int j = 0; j < N; ++j) add += exp2f((x.f(j) - mn) * kLog2e);
        s = s * exp2f((m - mn) * kLog2e) + add;
        m = mn;
      }
    } else {
      for (int c = threadIdx.x; c < V; c += blockDim.x) {
        const float xv = to_f32(xr[c]);
        const float mn = fmaxf(m, xv);
        s = s * exp2f((m - mn) * kLog2e) + exp2f((xv - mn) * kLog2e);
        m = mn;
      }
    }
    const float gm = block_max(m, red);
    const float contrib = (m == -INFINITY) ? 0.f : s * exp2f((m - gm) * kLog2e);
    const float gs = block_sum(contrib, red);
    if (threadIdx.x == 0) {
      const int64_t lab = labels[row];
      const int64_t local = lab - vocab_start;
      const bool in_range = lab != ignore_index && local >= 0 && local < V;
      const float tgt = in_range ? to_f32(xr[local]) : 0.f;
      if (stats) {
        stats[row * 3 + 0] = gm;
        stats[row * 3 + 1] = gs;
        stats[row * 3 + 2] = tgt;
      }
      const float l = gm + __logf(gs);
      if (lse) lse[row] = l;
      if (loss) loss[row] = (lab == ignore_index) ? 0.f : l - tgt;
    }
  }
}

// Combine per-rank stats gathered as [nranks, rows, 3] -> loss[rows], lse[rows]
__global__ void ce_combine_kernel(const float* __restrict__ all_stats, const int64_t* __restrict__ labels,
                                  float* __restrict__ loss, float* __restrict__ lse, int64_t rows, int nranks,
                                  int64_t ignore_index) {
  const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= rows) return;
  float gm = -INFINITY;
  for (int r = 0; r < nranks; ++r) gm = fmaxf(gm, all_stats[((int64_t)r * rows + row) * 3]);
  float gs = 0.f, tgt = 0.f;
  for (int r = 0; r < nranks; ++r) {
    const float* st = all_stats + ((int64_t)r * rows + row) * 3;
    gs += st[1] * exp2f((st[0] - gm) * kLog2e);
    tgt += st[2];
  }
  const float l = gm + __logf(gs);
  lse[row] = l;
  loss[row] = (labels[row] == ignore_index) ? 0.f : l - tgt;
}

// dlogits = (exp(x - lse) - onehot) * dloss ; rows with ignore_index get zeros.  May run in place (dlogits == logits).
template <typename T>
__global__ void __launch_bounds__(256) ce_bwd_kernel(const T* logits, const int64_t* __restrict__ labels,
                                                    const float* __restrict__ lse, const float* __restrict__ dloss,
                                                    T* dlogits, int64_t rows, int V, int64_t ld, int64_t vocab_start,
                                                    int64_t ignore_index, float dloss_scalar) {
  constexpr int N = Vec<T>::N;
  const bool vec_ok = (V % N == 0) && (ld % N == 0) && ((uintptr_t)logits % 16 == 0) && ((uintptr_t)dlogits % 16 == 0);
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const int64_t lab = labels[row];
    const float g = (lab == ignore_index) ? 0.f : (dloss ? dloss[row] : dloss_scalar);
    const float l = lse[row] * kLog2e;
    const int64_t local = lab - vocab_start;
    const T* xr = logits + row * ld;
    T* dr = dlogits + row * ld;
    if (vec_ok) {
      const int nv = V / N;
      for (int vi = threadIdx.x; vi < nv; vi += blockDim.x) {
        Vec<T> x, o;
        x.load_nc(xr + vi * N);
#pragma unroll
        for (int j = 0; j < N; ++j) {
          float p = exp2f(x.f(j) * kLog2e - l);
          if ((int64_t)vi * N + j == local) p -= 1.f;
          o.set(j, p * g);
        }
        o.store(dr + vi * N);
      }
    } else {
      for (int c = threadIdx.x; c < V; c += blockDim.x) {
        float p = exp2f(to_f32(xr[c]) * kLog2e - l);
        if (c == local) p -= 1.f;
        dr[c] = from_f32<T>(p * g);
      }
    }
  }
}

}  // namespace pdb

using namespace pdb;

static int ce_grid(int64_t rows) {
  int64_t g = (int64_t)sm_count() * 8;
  return (int)(rows < g ? rows : g);
}

PDB_EXPORT int pdb200_ce_fwd(const void* logits, const int64_t* labels, float* stats, float* loss, float* lse,
                             int64_t rows, int64_t V, int64_t ld, int64_t vocab_start, int64_t ignore_index, int dtype,
                             void* stream) {
  if (rows == 0) return 0;
  PDB_REQUIRE(V > 0 && V < (1ll << 31), "ce_fwd: bad vocab size %lld", (long long)V);
  cudaStream_t st = (cudaStream_t)stream;
  PDB_DISPATCH_FLOAT(dtype, T, {
    ce_stats_kernel<T><<<ce_grid(rows), 256, 0, st>>>((const T*)logits, labels, stats, loss, lse, rows, (int)V, ld,
                                                      vocab_start, ignore_index);
  });
  return check_launch("ce_fwd");
}

PDB_EXPORT int pdb200_ce_combine(const float* all_stats, const int64_t* labels, float* loss, float* lse, int64_t rows,
                                 int nranks, int64_t ignore_index, void* stream) {
  if (rows == 0) return 0;
  ce_combine_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(all_stats, labels, loss, lse, rows,
                                                                                     nranks, ignore_index);
  return check_launch("ce_combine");
}

PDB_EXPORT int pdb200_ce_bwd(const void* logits, const int64_t* labels, const float* lse, const float* dloss,
                             float dloss_scalar, void* dlogits, int64_t rows, int64_t V, int64_t ld,
                             int64_t vocab_start, int64_t ignore_index, int dtype, void* stream) {
  if (rows == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  PDB_DISPATCH_FLOAT(dtype, T, {
    ce_bwd_kernel<T><<<ce_grid(rows), 256, 0, st>>>((const T*)logits, labels, lse, dloss, (T*)dlogits, rows, (int)V, ld,
                                                    vocab_start, ignore_index, dloss_scalar);
  });
  return check_launch("ce_bwd");
}
