// K18/K19: fused RMSNorm / LayerNorm forward and backward for sm_100a.
//
// Replaces in the reference: paddle/phi/kernels/gpu/rms_norm_kernel.cu:420-520 (RmsNormBlockSMemImpl),
// rms_norm_grad_kernel.cu + rms_norm_funcs.h:507-942 (two-stage gamma/beta reduction), layer_norm_kernel.cu.
// Design for B200 (HBM-bound op): persistent CTAs (grid = k x 148) stride over rows, the row stays in registers
// between the statistics pass and the normalise pass (one HBM read, one write), all global traffic is LDG/STG.128,
// gamma/beta are read once per CTA and kept in registers, optional residual-add is fused in front
// (fused_rms_norm's `residual` argument).  Backward keeps per-thread dgamma/dbeta partials in registers across all
// rows of the CTA and spills them once to a [grid, cols] fp32 scratch that a second tiny kernel folds.
#include "common.cuh"

namespace pdb {

template <typename T, int VPT, bool LN>
__global__ void __launch_bounds__(256) norm_fwd_kernel(
    const T* __restrict__ x, const T* __restrict__ residual, const void* __restrict__ w, const void* __restrict__ b,
    T* __restrict__ y, T* __restrict__ residual_out, float* __restrict__ mean_out, float* __restrict__ rstd_out,
    int64_t rows, int cols, float eps, int w_is_f32) {
  constexpr int N = Vec<T>::N;
  __shared__ float red[32];
  const int nvec = cols / N;
  const int tid = threadIdx.x;
  float wr[VPT][N], br[VPT][N];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int vi = tid + i * blockDim.x;
#pragma unroll
    for (int j = 0; j < N; ++j) {
      const int c = vi * N + j;
      float wv = 1.f, bv = 0.f;
      if (vi < nvec) {
        if (w) wv = w_is_f32 ? ((const float*)w)[c] : to_f32(((const T*)w)[c]);
        if (b) bv = w_is_f32 ? ((const float*)b)[c] : to_f32(((const T*)b)[c]);
      }
      wr[i][j] = wv;
      br[i][j] = bv;
    }
  }
  const float inv_cols = 1.f / (float)cols;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* xr = x + row * cols;
    Vec<T> xv[VPT];
    float s = 0.f, ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = tid + i * blockDim.x;
      if (vi < nvec) {
        xv[i].load_nc(xr + vi * N);
        if (residual) {
          Vec<T> rv;
          rv.load_nc(residual + row * cols + vi * N);
#pragma unroll
          for (int j = 0; j < N; ++j) xv[i].set(j, xv[i].f(j) + rv.f(j));
          xv[i].store(residual_out + row * cols + vi * N);
        }
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const float f = xv[i].f(j);
          s += f;
          ss += f * f;
        }
      }
    }
    float mean = 0.f, rstd;
    if (LN) {
      mean = block_sum(s, red) * inv_cols;
      // two-pass variance on the register copy: no cancellation, no extra memory traffic
      float vs = 0.f;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int vi = tid + i * blockDim.x;
        if (vi < nvec) {
#pragma unroll
          for (int j = 0; j < N; ++j) {
            const float d = xv[i].f(j) - mean;
            vs += d * d;
          }
        }
      }
      rstd = rsqrtf(block_sum(vs, red) * inv_cols + eps);
    } else {
      rstd = rsqrtf(block_sum(ss, red) * inv_cols + eps);
    }
    if (tid == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    T* yr = y + row * cols;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = tid + i * blockDim.x;
      if (vi < nvec) {
        Vec<T> o;
#pragma unroll
        for (int j = 0; j < N; ++j) o.set(j, (xv[i].f(j) - mean) * rstd * wr[i][j] + br[i][j]);
        o.store(yr + vi * N);
      }
    }
  }
}

// Generic fallback (any cols, scalar access, re-reads the row): correctness path for odd shapes.
template <typename T, bool LN>
__global__ void norm_fwd_generic(const T* __restrict__ x, const T* __restrict__ residual, const void* w, const void* b,
                                 T* y, T* residual_out, float* mean_out, float* rstd_out, int64_t rows, int cols,
                                 float eps, int w_is_f32) {
  __shared__ float red[32];
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* xr = x + row * cols;
    float s = 0.f;
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
      float f = to_f32(xr[c]);
      if (residual) {
        f = to_f32(from_f32<T>(f + to_f32(residual[row * cols + c])));
        residual_out[row * cols + c] = from_f32<T>(f);
      }
      s += f;
    }
    const T* src = residual ? residual_out + row * cols : xr;
    float mean = 0.f;
    if (LN) mean = block_sum(s, red) / cols;
    float vs = 0.f;
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
      const float d = to_f32(src[c]) - mean;
      vs += d * d;
    }
    const float rstd = rsqrtf(block_sum(vs, red) / cols + eps);
    if (threadIdx.x == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
      const float wv = w ? (w_is_f32 ? ((const float*)w)[c] : to_f32(((const T*)w)[c])) : 1.f;
      const float bv = b ? (w_is_f32 ? ((const float*)b)[c] : to_f32(((const T*)b)[c])) : 0.f;
      y[row * cols + c] = from_f32<T>((to_f32(src[c]) - mean) * rstd * wv + bv);
    }
  }
}

// ---------------------------------------------------------------------------------------------- backward
template <typename T, int VPT, bool LN>
__global__ void __launch_bounds__(256) norm_bwd_kernel(
    const T* __restrict__ dy, const T* __restrict__ x, const void* __restrict__ w, const float* __restrict__ mean_in,
    const float* __restrict__ rstd_in, T* __restrict__ dx, float* __restrict__ part_dw, float* __restrict__ part_db,
    int64_t rows, int cols, int w_is_f32) {
  constexpr int N = Vec<T>::N;
  __shared__ float red[32];
  const int nvec = cols / N;
  const int tid = threadIdx.x;
  float wr[VPT][N], dwa[VPT][N], dba[VPT][N];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int vi = tid + i * blockDim.x;
#pragma unroll
    for (int j = 0; j < N; ++j) {
      const int c = vi * N + j;
      wr[i][j] = (vi < nvec && w) ? (w_is_f32 ? ((const float*)w)[c] : to_f32(((const T*)w)[c])) : 1.f;
      dwa[i][j] = 0.f;
      dba[i][j] = 0.f;
    }
  }
  const float inv_cols = 1.f / (float)cols;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const float mean = LN ? mean_in[row] : 0.f;
    const float rstd = rstd_in[row];
    Vec<T> xv[VPT], gv[VPT];
    float sg = 0.f, sgx = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = tid + i * blockDim.x;
      if (vi < nvec) {
        xv[i].load_nc(x + row * cols + vi * N);
        gv[i].load_nc(dy + row * cols + vi * N);
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const float xh = (xv[i].f(j) - mean) * rstd;
          const float g = gv[i].f(j);
          dwa[i][j] += g * xh;
          if (LN) dba[i][j] += g;
          const float gw = g * wr[i][j];
          sg += gw;
          sgx += gw * xh;
        }
      }
    }
    float mg = 0.f;
    if (LN) mg = block_sum(sg, red) * inv_cols;
    const float mgx = block_sum(sgx, red) * inv_cols;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = tid + i * blockDim.x;
      if (vi < nvec) {
        Vec<T> o;
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const float xh = (xv[i].f(j) - mean) * rstd;
          o.set(j, rstd * (gv[i].f(j) * wr[i][j] - mg - xh * mgx));
        }
        o.store(dx + row * cols + vi * N);
      }
    }
  }
  if (part_dw) {
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = tid + i * blockDim.x;
      if (vi < nvec) {
        float* pw = part_dw + (int64_t)blockIdx.x * cols + vi * N;
#pragma unroll
        for (int j = 0; j < N; j += 4) *reinterpret_cast<float4*>(pw + j) = make_float4(dwa[i][j], dwa[i][j + 1], dwa[i][j + 2], dwa[i][j + 3]);
        if (LN && part_db) {
          float* pb = part_db + (int64_t)blockIdx.x * cols + vi * N;
#pragma unroll
          for (int j = 0; j < N; j += 4) *reinterpret_cast<float4*>(pb + j) = make_float4(dba[i][j], dba[i][j + 1], dba[i][j + 2], dba[i][j + 3]);
        }
      }
    }
  }
}

template <typename T, bool LN>
__global__ void norm_bwd_generic(const T* dy, const T* x, const void* w, const float* mean_in, const float* rstd_in,
                                 T* dx, float* part_dw, float* part_db, int64_t rows, int cols, int w_is_f32) {
  __shared__ float red[32];
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    if (part_dw) part_dw[(int64_t)blockIdx.x * cols + c] = 0.f;
    if (LN && part_db) part_db[(int64_t)blockIdx.x * cols + c] = 0.f;
  }
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const float mean = LN ? mean_in[row] : 0.f, rstd = rstd_in[row];
    float sg = 0.f, sgx = 0.f;
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
      const float wv = w ? (w_is_f32 ? ((const float*)w)[c] : to_f32(((const T*)w)[c])) : 1.f;
      const float xh = (to_f32(x[row * cols + c]) - mean) * rstd, g = to_f32(dy[row * cols + c]);
      if (part_dw) part_dw[(int64_t)blockIdx.x * cols + c] += g * xh;
      if (LN && part_db) part_db[(int64_t)blockIdx.x * cols + c] += g;
      sg += g * wv;
      sgx += g * wv * xh;
    }
    float mg = 0.f;
    if (LN) mg = block_sum(sg, red) / cols;
    const float mgx = block_sum(sgx, red) / cols;
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
      const float wv = w ? (w_is_f32 ? ((const float*)w)[c] : to_f32(((const T*)w)[c])) : 1.f;
      const float xh = (to_f32(x[row * cols + c]) - mean) * rstd, g = to_f32(dy[row * cols + c]);
      dx[row * cols + c] = from_f32<T>(rstd * (g * wv - mg - xh * mgx));
    }
  }
}

// Fold [parts, cols] fp32 partials into gamma/beta grads.  One warp-row of 32 columns x 8 part-lanes per CTA.
template <typename WT>
__global__ void norm_bwd_fold(const float* __restrict__ part_dw, const float* __restrict__ part_db, WT* dw, WT* db,
                              int parts, int cols) {
  __shared__ float sw[8][33], sb[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  float aw = 0.f, ab = 0.f;
  if (c < cols) {
    for (int p = threadIdx.y; p < parts; p += 8) {
      aw += part_dw[(int64_t)p * cols + c];
      if (part_db) ab += part_db[(int64_t)p * cols + c];
    }
  }
  sw[threadIdx.y][threadIdx.x] = aw;
  sb[threadIdx.y][threadIdx.x] = ab;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
#pragma unroll
    for (int k = 1; k < 8; ++k) {
      aw += sw[k][threadIdx.x];
      ab += sb[k][threadIdx.x];
    }
    if (dw) dw[c] = from_f32<WT>(aw);
    if (db && part_db) db[c] = from_f32<WT>(ab);
  }
}

static int pick_grid(int64_t rows, int ctas_per_sm) {
  int64_t g = (int64_t)sm_count() * ctas_per_sm;
  return (int)(rows < g ? rows : g);
}

struct NormCfg {
  int threads, vpt;
  bool fast;
};
static NormCfg norm_cfg(int cols, int elt, int max_vpt) {
  // <=256 threads always (register budget: gamma/beta/partials live in registers), 1..max_vpt vectors per thread.
  const int N = 16 / elt;
  NormCfg c{256, 1, false};
  if (cols % N != 0) return c;
  const int nvec = cols / N;
  if (nvec <= 256) {
    c.threads = ((nvec + 31) / 32) * 32;
    c.vpt = 1;
  } else {
    c.threads = 256;
    int vpt = (nvec + 255) / 256, v = 1;
    while (v < vpt) v *= 2;
    if (v > max_vpt) return c;
    c.vpt = v;
  }
  c.fast = true;
  return c;
}

template <typename T, bool LN>
static int launch_fwd(const void* x, const void* residual, const void* w, const void* b, void* y, void* residual_out,
                      float* mean, float* rstd, int64_t rows, int cols, float eps, int w_is_f32, cudaStream_t st) {
  NormCfg c = norm_cfg(cols, sizeof(T), 8);
  const bool aligned = ((uintptr_t)x % 16 == 0) && ((uintptr_t)y % 16 == 0) &&
                       (!residual || ((uintptr_t)residual % 16 == 0 && (uintptr_t)residual_out % 16 == 0));
  const int ctas = 4;
  const int grid = pick_grid(rows, ctas);
#define LAUNCH(V)                                                                                              \
  norm_fwd_kernel<T, V, LN><<<grid, c.threads, 0, st>>>((const T*)x, (const T*)residual, w, b, (T*)y,          \
                                                        (T*)residual_out, mean, rstd, rows, cols, eps, w_is_f32)
  if (c.fast && aligned) {
    switch (c.vpt) {
      case 1: LAUNCH(1); break;
      case 2: LAUNCH(2); break;
      case 4: LAUNCH(4); break;
      default: LAUNCH(8); break;
    }
  } else {
    norm_fwd_generic<T, LN><<<grid, 256, 0, st>>>((const T*)x, (const T*)residual, w, b, (T*)y, (T*)residual_out,
                                                  mean, rstd, rows, cols, eps, w_is_f32);
  }
#undef LAUNCH
  return check_launch("norm_fwd");
}

static int bwd_grid(int64_t rows) { return pick_grid(rows, 2); }

template <typename T, bool LN>
static int launch_bwd(const void* dy, const void* x, const void* w, const float* mean, const float* rstd, void* dx,
                      void* dw, void* db, float* scratch, int64_t rows, int cols, int w_is_f32, cudaStream_t st) {
  NormCfg c = norm_cfg(cols, sizeof(T), 4);
  const bool aligned = ((uintptr_t)x % 16 == 0) && ((uintptr_t)dy % 16 == 0) && ((uintptr_t)dx % 16 == 0);
  const int grid = bwd_grid(rows);
  float* part_dw = (dw || db) ? scratch : nullptr;
  float* part_db = (LN && part_dw) ? scratch + (int64_t)grid * cols : nullptr;
#define LAUNCH(V)                                                                                                   \
  norm_bwd_kernel<T, V, LN><<<grid, c.threads, 0, st>>>((const T*)dy, (const T*)x, w, mean, rstd, (T*)dx, part_dw, \
                                                        part_db, rows, cols, w_is_f32)
  if (c.fast && aligned) {
    switch (c.vpt) {
      case 1: LAUNCH(1); break;
      case 2: LAUNCH(2); break;
      default: LAUNCH(4); break;
    }
  } else {
    norm_bwd_generic<T, LN><<<grid, 256, 0, st>>>((const T*)dy, (const T*)x, w, mean, rstd, (T*)dx, part_dw, part_db,
                                                  rows, cols, w_is_f32);
  }
#undef LAUNCH
  if (int e = check_launch("norm_bwd")) return e;
  if (part_dw) {
    dim3 blk(32, 8);
    const int g = (cols + 31) / 32;
    if (w_is_f32)
      norm_bwd_fold<float><<<g, blk, 0, st>>>(part_dw, part_db, (float*)dw, (float*)db, grid, cols);
    else
      norm_bwd_fold<T><<<g, blk, 0, st>>>(part_dw, part_db, (T*)dw, (T*)db, grid, cols);
    return check_launch("norm_bwd_fold");
  }
  return 0;
}

}  // namespace pdb

using namespace pdb;

PDB_EXPORT int pdb200_norm_fwd(const void* x, const void* residual, const void* w, const void* b, void* y,
                               void* residual_out, float* mean, float* rstd, int64_t rows, int64_t cols, float eps,
                               int x_dtype, int w_is_f32, int is_layernorm, void* stream) {
  PDB_REQUIRE(rows >= 0 && cols > 0 && cols < (1ll << 31), "norm_fwd: bad shape rows=%lld cols=%lld", (long long)rows, (long long)cols);
  if (rows == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  PDB_DISPATCH_FLOAT(x_dtype, T, {
    const int wf = (x_dtype == PDB_DTYPE_F32) ? 1 : w_is_f32;
    return is_layernorm ? launch_fwd<T, true>(x, residual, w, b, y, residual_out, mean, rstd, rows, (int)cols, eps, wf, st)
                        : launch_fwd<T, false>(x, residual, w, b, y, residual_out, mean, rstd, rows, (int)cols, eps, wf, st);
  });
  return 0;
}

PDB_EXPORT int64_t pdb200_norm_bwd_scratch_bytes(int64_t rows, int64_t cols, int is_layernorm) {
  if (rows <= 0) return 0;
  return (int64_t)bwd_grid(rows) * cols * sizeof(float) * (is_layernorm ? 2 : 1);
}

PDB_EXPORT int pdb200_norm_bwd(const void* dy, const void* x, const void* w, const float* mean, const float* rstd,
                               void* dx, void* dw, void* db, void* scratch, int64_t scratch_bytes, int64_t rows,
                               int64_t cols, int x_dtype, int w_is_f32, int is_layernorm, void* stream) {
  PDB_REQUIRE(rows >= 0 && cols > 0 && cols < (1ll << 31), "norm_bwd: bad shape");
  if (rows == 0) return 0;
  if (dw || db)
    PDB_REQUIRE(scratch && scratch_bytes >= pdb200_norm_bwd_scratch_bytes(rows, cols, is_layernorm),
                "norm_bwd: scratch too small (%lld < %lld)", (long long)scratch_bytes,
                (long long)pdb200_norm_bwd_scratch_bytes(rows, cols, is_layernorm));
  cudaStream_t st = (cudaStream_t)stream;
  PDB_DISPATCH_FLOAT(x_dtype, T, {
    const int wf = (x_dtype == PDB_DTYPE_F32) ? 1 : w_is_f32;
    return is_layernorm ? launch_bwd<T, true>(dy, x, w, mean, rstd, dx, dw, db, (float*)scratch, rows, (int)cols, wf, st)
                        : launch_bwd<T, false>(dy, x, w, mean, rstd, dx, dw, db, (float*)scratch, rows, (int)cols, wf, st);
  });
  return 0;
}
