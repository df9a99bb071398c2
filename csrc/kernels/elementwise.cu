// K20/K21/K22: fused rotary embedding, SwiGLU (fwd+bwd) and fused bias+activation for sm_100a.
//
// Replaces in the reference: paddle/phi/kernels/fusion/gpu/fused_rope_kernel.cu + fused_rope_utils.h:235-426,
// paddle/phi/kernels/gpu/swiglu_kernel.cu / swiglu_grad_kernel.cu, fusion/gpu/fused_bias_act_kernel.cu:30-545.
// All three are pure HBM streamers: every access is a 16-byte vector, loads use the no-allocate path, grids are
// sized to a multiple of the SM count and stride.  RoPE handles q and k (and optionally v) in ONE launch; its
// backward is the same kernel with the sine negated.
#include "common.cuh"

namespace pdb {

// ------------------------------------------------------------------------------------------------ SwiGLU
// out = silu(x) * y.  x,y are rows of width H with row strides (in elements) ldx/ldy so the "packed" form
// (x | y halves of one [rows, 2H] tensor) needs no split copy.
template <typename T>
__global__ void __launch_bounds__(256) swiglu_fwd_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                        T* __restrict__ out, int64_t rows, int H, int64_t ldx,
                                                        int64_t ldy) {
  constexpr int N = Vec<T>::N;
  const int hv = H / N;
  const int64_t total = rows * hv;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / hv;
    const int c = (int)(i - r * hv) * N;
    Vec<T> a, b, o;
    a.load_nc(x + r * ldx + c);
    b.load_nc(y + r * ldy + c);
#pragma unroll
    for (int j = 0; j < N; ++j) {
      const float xv = a.f(j);
      o.set(j, xv / (1.f + __expf(-xv)) * b.f(j));
    }
    o.store(out + r * (int64_t)H + c);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) swiglu_bwd_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                        const T* __restrict__ dout, T* __restrict__ dx,
                                                        T* __restrict__ dy, int64_t rows, int H, int64_t ldx,
                                                        int64_t ldy, int64_t lddx, int64_t lddy) {
  constexpr int N = Vec<T>::N;
  const int hv = H / N;
  const int64_t total = rows * hv;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / hv;
    const int c = (int)(i - r * hv) * N;
    Vec<T> a, b, g, oa, ob;
    a.load_nc(x + r * ldx + c);
    b.load_nc(y + r * ldy + c);
    g.load_nc(dout + r * (int64_t)H + c);
#pragma unroll
    for (int j = 0; j < N; ++j) {
      const float xv = a.f(j), yv = b.f(j), gv = g.f(j);
      const float sig = 1.f / (1.f + __expf(-xv));
      const float silu = xv * sig;
      oa.set(j, gv * yv * (sig + silu * (1.f - sig)));
      ob.set(j, gv * silu);
    }
    oa.store(dx + r * lddx + c);
    ob.store(dy + r * lddy + c);
  }
}

template <typename T>
__global__ void swiglu_fwd_scalar(const T* x, const T* y, T* out, int64_t rows, int H, int64_t ldx, int64_t ldy) {
  const int64_t total = rows * H;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / H;
    const int c = (int)(i - r * H);
    const float xv = to_f32(x[r * ldx + c]);
    out[i] = from_f32<T>(xv / (1.f + __expf(-xv)) * to_f32(y[r * ldy + c]));
  }
}
template <typename T>
__global__ void swiglu_bwd_scalar(const T* x, const T* y, const T* dout, T* dx, T* dy, int64_t rows, int H, int64_t ldx,
                                  int64_t ldy, int64_t lddx, int64_t lddy) {
  const int64_t total = rows * H;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / H;
    const int c = (int)(i - r * H);
    const float xv = to_f32(x[r * ldx + c]), yv = to_f32(y[r * ldy + c]), gv = to_f32(dout[i]);
    const float sig = 1.f / (1.f + __expf(-xv)), silu = xv * sig;
    dx[r * lddx + c] = from_f32<T>(gv * yv * (sig + silu * (1.f - sig)));
    dy[r * lddy + c] = from_f32<T>(gv * silu);
  }
}

static int stream_grid(int64_t work_items, int threads) {
  int64_t blocks = (work_items + threads - 1) / threads;
  int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// ------------------------------------------------------------------------------------------------ RoPE
// Layout [b, s, h, d] (or [s, b, h, d] when time_major).  sin/cos are fp32 tables [s_table, d] ("full-width":
// both halves / both elements of a pair hold the same angle, the reference's convention) or null, in which case
// the angle is computed from rotary_emb_base.  position_ids [b, s] (int64) optional.
// neox style:  out[i] = x[i] cos[i] - x[i + d/2] sin[i]           (i <  d/2)
//              out[i] = x[i] cos[i] + x[i - d/2] sin[i]           (i >= d/2)
// GPT-J style: pairs (2i, 2i+1) rotate together.
struct RopeArgs {
  const void* in[3];
  void* out[3];
  int heads[3];
  int64_t in_tok_stride[3];    // elements between consecutive tokens (heads*d when dense, packed width for qkv)
  int64_t out_tok_stride[3];
  int rotate[3];               // 0 = plain copy (v in the fused qkv split)
  int n_tensors;
};

template <typename T, bool NEOX /* true = rotate-half (HF/NeoX layout), false = rotate-every-two */>
__global__ void __launch_bounds__(256) rope_kernel(RopeArgs a, const float* __restrict__ sin_t,
                                                  const float* __restrict__ cos_t,
                                                  const int64_t* __restrict__ pos_ids, int64_t batch, int64_t seq, int d,
                                                  int time_major, float base, float sign) {
  constexpr int N = Vec<T>::N;
  const int half = d / 2;
  // work item = one (token, tensor, head, vector-of-the-first-half) for neox; one vector for GPT-J
  const int vec_per_head = NEOX ? half / N : d / N;
  int total_heads = 0;
  for (int t = 0; t < a.n_tensors; ++t) total_heads += a.heads[t];
  const int64_t per_token = (int64_t)total_heads * vec_per_head;
  const int64_t total = batch * seq * per_token;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t tok = i / per_token;
    int rem = (int)(i - tok * per_token);
    int hh = rem / vec_per_head;
    const int v = rem - hh * vec_per_head;
    int t = 0;
    while (hh >= a.heads[t]) {
      hh -= a.heads[t];
      ++t;
    }
    const int64_t bi = time_major ? tok % batch : tok / seq;
    const int64_t si = time_major ? tok / batch : tok % seq;
    const int64_t pos = pos_ids ? pos_ids[bi * seq + si] : si;
    const T* src = (const T*)a.in[t] + tok * a.in_tok_stride[t] + hh * (int64_t)d;
    T* dst = (T*)a.out[t] + tok * a.out_tok_stride[t] + hh * (int64_t)d;
    if (!a.rotate[t]) {
      if (NEOX) {
        Vec<T> lo, hi;
        lo.load_nc(src + v * N);
        hi.load_nc(src + half + v * N);
        lo.store(dst + v * N);
        hi.store(dst + half + v * N);
      } else {
        Vec<T> x;
        x.load_nc(src + v * N);
        x.store(dst + v * N);
      }
      continue;
    }
    if (NEOX) {
      Vec<T> lo, hi, olo, ohi;
      lo.load_nc(src + v * N);
      hi.load_nc(src + half + v * N);
#pragma unroll
      for (int j = 0; j < N; ++j) {
        const int idx = v * N + j;
        float sn, cs, sn1, cs1;
        if (sin_t) {
          sn = sin_t[pos * d + idx];
          cs = cos_t[pos * d + idx];
          sn1 = sin_t[pos * d + half + idx];
          cs1 = cos_t[pos * d + half + idx];
        } else {
          const float inv_freq = __powf(base, -2.f * (float)idx / (float)d);
          __sincosf((float)pos * inv_freq, &sn, &cs);
          sn1 = sn;
          cs1 = cs;
        }
        sn *= sign;
        sn1 *= sign;
        const float x0 = lo.f(j), x1 = hi.f(j);
        olo.set(j, x0 * cs - x1 * sn);
        ohi.set(j, x1 * cs1 + x0 * sn1);
      }
      olo.store(dst + v * N);
      ohi.store(dst + half + v * N);
    } else {
      Vec<T> x, o;
      x.load_nc(src + v * N);
#pragma unroll
      for (int j = 0; j < N; j += 2) {
        const int idx = v * N + j;
        float sn, cs, sn1, cs1;
        if (sin_t) {
          sn = sin_t[pos * d + idx];
          cs = cos_t[pos * d + idx];
          sn1 = sin_t[pos * d + idx + 1];
          cs1 = cos_t[pos * d + idx + 1];
        } else {
          const float inv_freq = __powf(base, -(float)idx / (float)d);
          __sincosf((float)pos * inv_freq, &sn, &cs);
          sn1 = sn;
          cs1 = cs;
        }
        sn *= sign;
        sn1 *= sign;
        const float x0 = x.f(j), x1 = x.f(j + 1);
        o.set(j, x0 * cs - x1 * sn);
        o.set(j + 1, x1 * cs1 + x0 * sn1);
      }
      o.store(dst + v * N);
    }
  }
}

// ------------------------------------------------------------------------------------------------ bias + act
enum { ACT_GELU = 0, ACT_RELU = 1, ACT_SWIGLU = 2, ACT_GEGLU = 3, ACT_SILU = 4, ACT_IDENTITY = 5 };

__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }

template <typename T, int ACT>
__global__ void __launch_bounds__(256) bias_act_kernel(const T* __restrict__ x, const T* __restrict__ bias,
                                                      T* __restrict__ out, int64_t rows, int cols_out) {
  constexpr int N = Vec<T>::N;
  constexpr bool GLU = (ACT == ACT_SWIGLU || ACT == ACT_GEGLU);
  const int cols_in = GLU ? cols_out * 2 : cols_out;
  const int cv = cols_out / N;
  const int64_t total = rows * cv;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cv;
    const int c = (int)(i - r * cv) * N;
    Vec<T> a, b, ba, bb, o;
    a.load_nc(x + r * cols_in + c);
    if (bias) ba.load(bias + c);
    if (GLU) {
      b.load_nc(x + r * cols_in + cols_out + c);
      if (bias) bb.load(bias + cols_out + c);
    }
#pragma unroll
    for (int j = 0; j < N; ++j) {
      float v = a.f(j) + (bias ? ba.f(j) : 0.f);
      float res;
      if (ACT == ACT_GELU) res = gelu_f(v);
      else if (ACT == ACT_RELU) res = fmaxf(v, 0.f);
      else if (ACT == ACT_SILU) res = v / (1.f + __expf(-v));
      else if (ACT == ACT_IDENTITY) res = v;
      else {
        const float g = b.f(j) + (bias ? bb.f(j) : 0.f);
        res = (ACT == ACT_SWIGLU ? v / (1.f + __expf(-v)) : gelu_f(v)) * g;
      }
      o.set(j, res);
    }
    o.store(out + r * (int64_t)cols_out + c);
  }
}

}  // namespace pdb

using namespace pdb;

PDB_EXPORT int pdb200_swiglu_fwd(const void* x, const void* y, void* out, int64_t rows, int64_t H, int64_t ldx,
                                 int64_t ldy, int dtype, void* stream) {
  if (rows * H == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  PDB_DISPATCH_FLOAT(dtype, T, {
    constexpr int N = 16 / sizeof(T);
    const bool vec = H % N == 0 && ldx % N == 0 && ldy % N == 0 && (uintptr_t)x % 16 == 0 && (uintptr_t)y % 16 == 0 &&
                     (uintptr_t)out % 16 == 0;
    if (vec)
      swiglu_fwd_kernel<T><<<stream_grid(rows * (H / N), 256), 256, 0, st>>>((const T*)x, (const T*)y, (T*)out, rows, (int)H, ldx, ldy);
    else
      swiglu_fwd_scalar<T><<<stream_grid(rows * H, 256), 256, 0, st>>>((const T*)x, (const T*)y, (T*)out, rows, (int)H, ldx, ldy);
  });
  return check_launch("swiglu_fwd");
}

PDB_EXPORT int pdb200_swiglu_bwd(const void* x, const void* y, const void* dout, void* dx, void* dy, int64_t rows,
                                 int64_t H, int64_t ldx, int64_t ldy, int64_t lddx, int64_t lddy, int dtype,
                                 void* stream) {
  if (rows * H == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  PDB_DISPATCH_FLOAT(dtype, T, {
    constexpr int N = 16 / sizeof(T);
    const bool vec = H % N == 0 && ldx % N == 0 && ldy % N == 0 && lddx % N == 0 && lddy % N == 0 &&
                     (uintptr_t)x % 16 == 0 && (uintptr_t)y % 16 == 0 && (uintptr_t)dout % 16 == 0 &&
                     (uintptr_t)dx % 16 == 0 && (uintptr_t)dy % 16 == 0;
    if (vec)
      swiglu_bwd_kernel<T><<<stream_grid(rows * (H / N), 256), 256, 0, st>>>((const T*)x, (const T*)y, (const T*)dout, (T*)dx, (T*)dy, rows, (int)H, ldx, ldy, lddx, lddy);
    else
      swiglu_bwd_scalar<T><<<stream_grid(rows * H, 256), 256, 0, st>>>((const T*)x, (const T*)y, (const T*)dout, (T*)dx, (T*)dy, rows, (int)H, ldx, ldy, lddx, lddy);
  });
  return check_launch("swiglu_bwd");
}

// q/k/v: up to three tensors rotated in one launch (null = absent).  sign=+1 forward, -1 backward.
PDB_EXPORT int pdb200_rope(const void* q, const void* k, const void* v, void* oq, void* ok, void* ov, int hq, int hk,
                           int hv, const float* sin_t, const float* cos_t, const int64_t* pos_ids, int64_t batch,
                           int64_t seq, int d, int rotate_half, int time_major, float base, float sign, int dtype,
                           void* stream) {
  const int neox = rotate_half;
  RopeArgs a{};
  const void* ins[3] = {q, k, v};
  void* outs[3] = {oq, ok, ov};
  const int hs[3] = {hq, hk, hv};
  for (int i = 0; i < 3; ++i)
    if (ins[i]) {
      a.in[a.n_tensors] = ins[i];
      a.out[a.n_tensors] = outs[i];
      a.heads[a.n_tensors] = hs[i];
      a.in_tok_stride[a.n_tensors] = (int64_t)hs[i] * d;
      a.out_tok_stride[a.n_tensors] = (int64_t)hs[i] * d;
      a.rotate[a.n_tensors] = 1;
      a.n_tensors++;
      PDB_REQUIRE((uintptr_t)ins[i] % 16 == 0 && (uintptr_t)outs[i] % 16 == 0, "rope: tensors must be 16-byte aligned");
    }
  if (a.n_tensors == 0 || batch * seq == 0) return 0;
  PDB_REQUIRE((sin_t == nullptr) == (cos_t == nullptr), "rope: sin and cos must be given together");
  cudaStream_t st = (cudaStream_t)stream;
  int total_heads = 0;
  for (int t = 0; t < a.n_tensors; ++t) total_heads += a.heads[t];
  PDB_DISPATCH_FLOAT(dtype, T, {
    constexpr int N = 16 / sizeof(T);
    PDB_REQUIRE(d % (2 * N) == 0, "rope: head_dim %d must be a multiple of %d", d, 2 * N);
    const int64_t items = batch * seq * total_heads * (neox ? d / 2 / N : d / N);
    if (neox)
      rope_kernel<T, true><<<stream_grid(items, 256), 256, 0, st>>>(a, sin_t, cos_t, pos_ids, batch, seq, d, time_major, base, sign);
    else
      rope_kernel<T, false><<<stream_grid(items, 256), 256, 0, st>>>(a, sin_t, cos_t, pos_ids, batch, seq, d, time_major, base, sign);
  });
  return check_launch("rope");
}

// Fused QKV split + rotary embedding (forward) / merge + inverse rotary (backward, sign = -1):
//   split = 1: packed [tokens, (hq + 2 hk) * d] -> q [tokens, hq, d], k, v contiguous, q/k rotated, v copied
//   split = 0: q, k, v contiguous -> packed, q/k rotated with `sign`, v copied
// One pass over the data instead of three strided copies + a rope launch (+ zero/copy/add kernels in backward).
PDB_EXPORT int pdb200_qkv_rope(const void* packed, void* q, void* k, void* v, int hq, int hk, const float* sin_t,
                               const float* cos_t, const int64_t* pos_ids, int64_t batch, int64_t seq, int d,
                               int rotate_half, float base, float sign, int split, int dtype, void* stream) {
  RopeArgs a{};
  const int64_t W = (int64_t)(hq + 2 * hk) * d;
  const int hs[3] = {hq, hk, hk};
  const int64_t offs[3] = {0, (int64_t)hq * d, (int64_t)(hq + hk) * d};
  void* sep[3] = {q, k, v};
  const int esz = dtype == PDB_DTYPE_F32 ? 4 : 2;
  for (int i = 0; i < 3; ++i) {
    char* pk = (char*)const_cast<void*>(packed) + offs[i] * esz;
    a.in[i] = split ? (const void*)pk : (const void*)sep[i];
    a.out[i] = split ? sep[i] : (void*)pk;
    a.heads[i] = hs[i];
    a.in_tok_stride[i] = split ? W : (int64_t)hs[i] * d;
    a.out_tok_stride[i] = split ? (int64_t)hs[i] * d : W;
    a.rotate[i] = i < 2;
    PDB_REQUIRE((uintptr_t)a.in[i] % 16 == 0 && (uintptr_t)a.out[i] % 16 == 0, "qkv_rope: tensors must be 16-byte aligned");
  }
  a.n_tensors = 3;
  if (batch * seq == 0) return 0;
  PDB_REQUIRE((sin_t == nullptr) == (cos_t == nullptr), "qkv_rope: sin and cos must be given together");
  cudaStream_t st = (cudaStream_t)stream;
  const int total_heads = hq + 2 * hk;
  PDB_DISPATCH_FLOAT(dtype, T, {
    constexpr int N = 16 / sizeof(T);
    PDB_REQUIRE(d % (2 * N) == 0, "qkv_rope: head_dim %d must be a multiple of %d", d, 2 * N);
    const int64_t items = batch * seq * total_heads * (rotate_half ? d / 2 / N : d / N);
    if (rotate_half)
      rope_kernel<T, true><<<stream_grid(items, 256), 256, 0, st>>>(a, sin_t, cos_t, pos_ids, batch, seq, d, 0, base, sign);
    else
      rope_kernel<T, false><<<stream_grid(items, 256), 256, 0, st>>>(a, sin_t, cos_t, pos_ids, batch, seq, d, 0, base, sign);
  });
  return check_launch("qkv_rope");
}

template <typename T>
static int launch_bias_act(const void* x, const void* bias, void* out, int64_t rows, int64_t cols_out, int act,
                           cudaStream_t st) {
  constexpr int N = 16 / sizeof(T);
  PDB_REQUIRE(cols_out % N == 0, "bias_act: cols %lld must be a multiple of %d", (long long)cols_out, N);
  const int g = stream_grid(rows * (cols_out / N), 256);
#define PDB_BA(A) bias_act_kernel<T, A><<<g, 256, 0, st>>>((const T*)x, (const T*)bias, (T*)out, rows, (int)cols_out)
  switch (act) {
    case ACT_GELU: PDB_BA(ACT_GELU); break;
    case ACT_RELU: PDB_BA(ACT_RELU); break;
    case ACT_SWIGLU: PDB_BA(ACT_SWIGLU); break;
    case ACT_GEGLU: PDB_BA(ACT_GEGLU); break;
    case ACT_SILU: PDB_BA(ACT_SILU); break;
    case ACT_IDENTITY: PDB_BA(ACT_IDENTITY); break;
    default: set_error("bias_act: unknown act %d", act); return -1;
  }
#undef PDB_BA
  return check_launch("bias_act");
}

PDB_EXPORT int pdb200_bias_act(const void* x, const void* bias, void* out, int64_t rows, int64_t cols_out, int act,
                               int dtype, void* stream) {
  if (rows * cols_out == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  PDB_DISPATCH_FLOAT(dtype, T, { return launch_bias_act<T>(x, bias, out, rows, cols_out, act, st); });
  return 0;
}
