// Fused collectives over NVLink symmetric memory:
//   K1  dp_bucket_allreduce : scale + all-reduce of a gradient bucket in ONE kernel (NVLS two-shot: every rank pulls
//                             its 1/n slice through multimem.ld_reduce - the switch does the sum - scales it and
//                             multimem.st broadcasts it back).  Replaces concat -> scale_ -> ncclAllReduce -> split
//                             (paddle/fluid/distributed/collective/reducer.cc:1093-1127).
//   K2+K3 rs_adamw_ag       : reduce-scatter of the flat gradient buffer + partitioned AdamW on the local shard
//                             (fp32 master / m / v) + all-gather of the updated bf16 parameters, ONE kernel.
//                             Replaces ncclReduceScatter + per-parameter adamw launches + ncclAllGather
//                             (fleet/utils/tensor_fusion_helper.py:559-645, phi/kernels/gpu/adamw_kernel.cu).
//   K8  ce_combine_symm     : vocab-parallel cross-entropy statistics exchange: every rank reads the 3 floats per
//                             row of all peers straight from their symmetric scratch (replaces three N x 1
//                             all-reduces, c_softmax_with_cross_entropy_op.cu:211-314).
// Cross-GPU ordering uses system-scope flag barriers on a signal pad inside the segment (nvl.cuh).  When the fabric
// has no multicast support the same kernels fall back to peer loads/stores (P2P two-shot).
#include <stdlib.h>

#include "common.cuh"
#include "nvl.cuh"

namespace pdb {
namespace nvl {

__global__ void barrier_kernel(Peers P, uint64_t pad_off, uint32_t epoch, unsigned long long spin_limit) {
  barrier_all(P, pad_off, epoch, spin_limit);
}

// ---------------------------------------------------------------------------------------------- K1 all-reduce
template <typename T>
__global__ void __launch_bounds__(512) allreduce_scale_kernel(Peers P, uint64_t data_off, int64_t nvec, float scale) {
  // nvec 16-byte vectors in total; rank r owns vectors [r*per, min((r+1)*per, nvec))
  const int64_t per = (nvec + P.nranks - 1) / P.nranks;
  const int64_t lo = per * P.rank, hi = (lo + per < nvec) ? lo + per : nvec;
  constexpr int N = 16 / sizeof(T);
  for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t off = data_off + (uint64_t)i * 16;
    Vec<T> out;
    if (P.mc) {
      if (sizeof(T) == 4) {
        float v[4];
        mm_ld_reduce_f32x4(reinterpret_cast<const void*>(P.mc + off), v);
#pragma unroll
        for (int j = 0; j < 4; ++j) out.set(j, v[j] * scale);
      } else {
        uint32_t r[4];
        mm_ld_reduce_bf16x8(reinterpret_cast<const void*>(P.mc + off), r);
        out.raw = make_uint4(r[0], r[1], r[2], r[3]);
        if (scale != 1.f) {
#pragma unroll
          for (int j = 0; j < N; ++j) out.set(j, out.f(j) * scale);
        }
      }
      uint32_t w[4] = {out.raw.x, out.raw.y, out.raw.z, out.raw.w};
      mm_st_b32x4(reinterpret_cast<void*>(P.mc + off), w);
    } else {
      float acc[N];
#pragma unroll
      for (int j = 0; j < N; ++j) acc[j] = 0.f;
      for (int p = 0; p < P.nranks; ++p) {
        Vec<T> x;
        x.load(reinterpret_cast<const T*>(P.ptr[p] + off));
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] += x.f(j);
      }
#pragma unroll
      for (int j = 0; j < N; ++j) out.set(j, acc[j] * scale);
      for (int p = 0; p < P.nranks; ++p) out.store(reinterpret_cast<T*>(P.ptr[p] + off));
    }
  }
}

// ---------------------------------------------------------------------------------------------- K2 + K3
struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay;
  const float* beta_pows;    // device {beta1^t, beta2^t}
  const float* grad_scale;   // optional device scalar (clip coefficient / inverse loss scale)
  const float* found_inf;    // optional device flag: skip the whole update when != 0
};

// Gradients: symmetric flat buffer of G (bf16 or fp32) at grad_off; parameters: symmetric flat bf16 buffer at
// param_off; this rank's shard is elements [shard_start, shard_start + shard_numel) of both (shard_numel % 8 == 0).
// master / m / v: local fp32 arrays of shard_numel.
template <typename G>
__global__ void __launch_bounds__(256) rs_adamw_ag_kernel(Peers P, uint64_t grad_off, uint64_t param_off,
                                                         int64_t shard_start, int64_t shard_numel,
                                                         float* __restrict__ master, float* __restrict__ m,
                                                         float* __restrict__ v, AdamArgs a,
                                                         const float* __restrict__ local_grad) {
  if (a.found_inf && *a.found_inf != 0.f) return;
  // local_grad != null: the shard's gradient was already reduced (two-pass mode used with global-norm clipping)
  const float gs = (a.grad_scale ? *a.grad_scale : 1.f) / (local_grad ? 1.f : (float)P.nranks);    // AVG reduce
  const float bc1 = 1.f - a.beta_pows[0];
  const float bc2s = sqrtf(1.f - a.beta_pows[1]);
  const float decay_mul = 1.f - a.lr * a.weight_decay;
  const float step = a.lr / bc1;
  const int64_t n8 = shard_numel / 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = shard_start + i * 8;          // global element index
    float g[8];
    if (local_grad) {
      const float4 x0 = *reinterpret_cast<const float4*>(local_grad + i * 8);
      const float4 x1 = *reinterpret_cast<const float4*>(local_grad + i * 8 + 4);
      g[0] = x0.x; g[1] = x0.y; g[2] = x0.z; g[3] = x0.w; g[4] = x1.x; g[5] = x1.y; g[6] = x1.z; g[7] = x1.w;
    } else if (sizeof(G) == 4) {
      const uint64_t off = grad_off + (uint64_t)e * 4;
      if (P.mc) {
        float t0[4], t1[4];
        mm_ld_reduce_f32x4(reinterpret_cast<const void*>(P.mc + off), t0);
        mm_ld_reduce_f32x4(reinterpret_cast<const void*>(P.mc + off + 16), t1);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          g[j] = t0[j];
          g[4 + j] = t1[j];
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = 0.f;
        for (int p = 0; p < P.nranks; ++p) {
          const float4 x0 = *reinterpret_cast<const float4*>(P.ptr[p] + off);
          const float4 x1 = *reinterpret_cast<const float4*>(P.ptr[p] + off + 16);
          g[0] += x0.x; g[1] += x0.y; g[2] += x0.z; g[3] += x0.w;
          g[4] += x1.x; g[5] += x1.y; g[6] += x1.z; g[7] += x1.w;
        }
      }
    } else {
      const uint64_t off = grad_off + (uint64_t)e * 2;
      if (P.mc) {
        uint32_t r[4];
        mm_ld_reduce_bf16x8(reinterpret_cast<const void*>(P.mc + off), r);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&r[j]);
          g[2 * j] = __low2float(h);
          g[2 * j + 1] = __high2float(h);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = 0.f;
        for (int p = 0; p < P.nranks; ++p) {
          Vec<__nv_bfloat16> x;
          x.load(reinterpret_cast<const __nv_bfloat16*>(P.ptr[p] + off));
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] += x.f(j);
        }
      }
    }
    float pw[8], mm_[8], vv[8];
    {
      const float4 p0 = *reinterpret_cast<const float4*>(master + i * 8), p1 = *reinterpret_cast<const float4*>(master + i * 8 + 4);
      const float4 m0 = *reinterpret_cast<const float4*>(m + i * 8), m1 = *reinterpret_cast<const float4*>(m + i * 8 + 4);
      const float4 v0 = *reinterpret_cast<const float4*>(v + i * 8), v1 = *reinterpret_cast<const float4*>(v + i * 8 + 4);
      pw[0] = p0.x; pw[1] = p0.y; pw[2] = p0.z; pw[3] = p0.w; pw[4] = p1.x; pw[5] = p1.y; pw[6] = p1.z; pw[7] = p1.w;
      mm_[0] = m0.x; mm_[1] = m0.y; mm_[2] = m0.z; mm_[3] = m0.w; mm_[4] = m1.x; mm_[5] = m1.y; mm_[6] = m1.z; mm_[7] = m1.w;
      vv[0] = v0.x; vv[1] = v0.y; vv[2] = v0.z; vv[3] = v0.w; vv[4] = v1.x; vv[5] = v1.y; vv[6] = v1.z; vv[7] = v1.w;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = g[j] * gs;
      pw[j] *= decay_mul;
      mm_[j] = a.beta1 * mm_[j] + (1.f - a.beta1) * gj;
      vv[j] = a.beta2 * vv[j] + (1.f - a.beta2) * gj * gj;
      pw[j] -= step * (mm_[j] / (sqrtf(vv[j]) / bc2s + a.eps));
    }
    *reinterpret_cast<float4*>(master + i * 8) = make_float4(pw[0], pw[1], pw[2], pw[3]);
    *reinterpret_cast<float4*>(master + i * 8 + 4) = make_float4(pw[4], pw[5], pw[6], pw[7]);
    *reinterpret_cast<float4*>(m + i * 8) = make_float4(mm_[0], mm_[1], mm_[2], mm_[3]);
    *reinterpret_cast<float4*>(m + i * 8 + 4) = make_float4(mm_[4], mm_[5], mm_[6], mm_[7]);
    *reinterpret_cast<float4*>(v + i * 8) = make_float4(vv[0], vv[1], vv[2], vv[3]);
    *reinterpret_cast<float4*>(v + i * 8 + 4) = make_float4(vv[4], vv[5], vv[6], vv[7]);
    // all-gather: the updated bf16 parameters go to every rank's copy of the flat parameter buffer
    Vec<__nv_bfloat16> o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.set(j, pw[j]);
    const uint64_t poff = param_off + (uint64_t)e * 2;
    if (P.mc) {
      uint32_t w[4] = {o.raw.x, o.raw.y, o.raw.z, o.raw.w};
      mm_st_b32x4(reinterpret_cast<void*>(P.mc + poff), w);
    } else {
      for (int p = 0; p < P.nranks; ++p) o.store(reinterpret_cast<__nv_bfloat16*>(P.ptr[p] + poff));
    }
  }
}

// Reduce-scatter (AVG) of the flat gradient buffer into a local fp32 shard + per-CTA sum of squares (for the
// global-norm clip that must see the reduced gradient before the update).
template <typename G>
__global__ void __launch_bounds__(256) reduce_scatter_norm_kernel(Peers P, uint64_t grad_off, int64_t shard_start,
                                                                 int64_t shard_numel, float* __restrict__ out,
                                                                 float* __restrict__ partial_sq) {
  __shared__ float red[32];
  const float inv = 1.f / (float)P.nranks;
  const int64_t n8 = shard_numel / 8;
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = shard_start + i * 8;
    float g[8];
    if (sizeof(G) == 4) {
      const uint64_t off = grad_off + (uint64_t)e * 4;
      if (P.mc) {
        float t0[4], t1[4];
        mm_ld_reduce_f32x4(reinterpret_cast<const void*>(P.mc + off), t0);
        mm_ld_reduce_f32x4(reinterpret_cast<const void*>(P.mc + off + 16), t1);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          g[j] = t0[j];
          g[4 + j] = t1[j];
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = 0.f;
        for (int p = 0; p < P.nranks; ++p) {
          const float4 x0 = *reinterpret_cast<const float4*>(P.ptr[p] + off);
          const float4 x1 = *reinterpret_cast<const float4*>(P.ptr[p] + off + 16);
          g[0] += x0.x; g[1] += x0.y; g[2] += x0.z; g[3] += x0.w;
          g[4] += x1.x; g[5] += x1.y; g[6] += x1.z; g[7] += x1.w;
        }
      }
    } else {
      const uint64_t off = grad_off + (uint64_t)e * 2;
      if (P.mc) {
        uint32_t r[4];
        mm_ld_reduce_bf16x8(reinterpret_cast<const void*>(P.mc + off), r);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&r[j]);
          g[2 * j] = __low2float(h);
          g[2 * j + 1] = __high2float(h);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = 0.f;
        for (int p = 0; p < P.nranks; ++p) {
          Vec<__nv_bfloat16> x;
          x.load(reinterpret_cast<const __nv_bfloat16*>(P.ptr[p] + off));
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] += x.f(j);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      g[j] *= inv;
      acc += g[j] * g[j];
    }
    *reinterpret_cast<float4*>(out + i * 8) = make_float4(g[0], g[1], g[2], g[3]);
    *reinterpret_cast<float4*>(out + i * 8 + 4) = make_float4(g[4], g[5], g[6], g[7]);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0 && partial_sq) partial_sq[blockIdx.x] = acc;
}

// ---------------------------------------------------------------------------------------------- K8
constexpr float kLog2e = 1.4426950408889634f;
__global__ void ce_combine_symm_kernel(Peers P, uint64_t stats_off, const int64_t* __restrict__ labels,
                                       float* __restrict__ loss, float* __restrict__ lse, int64_t rows,
                                       int64_t ignore_index) {
  const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= rows) return;
  float mx[kMaxRanks], sm[kMaxRanks], tgt = 0.f, gm = -INFINITY;
  for (int p = 0; p < P.nranks; ++p) {
    const float* st = reinterpret_cast<const float*>(P.ptr[p] + stats_off) + row * 3;
    mx[p] = st[0];
    sm[p] = st[1];
    tgt += st[2];
    gm = fmaxf(gm, mx[p]);
  }
  float gs = 0.f;
  for (int p = 0; p < P.nranks; ++p) gs += sm[p] * exp2f((mx[p] - gm) * kLog2e);
  const float l = gm + __logf(gs);
  lse[row] = l;
  loss[row] = (labels[row] == ignore_index) ? 0.f : l - tgt;
}

// Sum (not average) reduce-scatter of a symmetric bf16 buffer into a local bf16 shard: the tail of GEMM ->
// reduce-scatter when the GEMM wrote its partial product into symmetric memory.  With NVLS the switch adds the n
// copies (fp32 accumulation) and each GPU only receives its own slice once.
__global__ void __launch_bounds__(256) rs_sum_bf16_kernel(Peers P, uint64_t src_off, int64_t shard_start, int64_t shard_numel,
                                                         __nv_bfloat16* __restrict__ out) {
  const int64_t n8 = shard_numel / 8;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const uint64_t off = src_off + (uint64_t)(shard_start + i * 8) * 2;
    uint4 o;
    if (P.mc) {
      uint32_t r[4];
      mm_ld_reduce_bf16x8(reinterpret_cast<const void*>(P.mc + off), r);
      o = make_uint4(r[0], r[1], r[2], r[3]);
    } else {
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
      for (int p = 0; p < P.nranks; ++p) {
        Vec<__nv_bfloat16> x;
        x.load(reinterpret_cast<const __nv_bfloat16*>(P.ptr[p] + off));
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += x.f(j);
      }
      Vec<__nv_bfloat16> v;
#pragma unroll
      for (int j = 0; j < 8; ++j) v.set(j, acc[j]);
      o = v.raw;
    }
    *reinterpret_cast<uint4*>(out + i * 8) = o;
  }
}

// ---------------------------------------------------------------------------------------------- K3 (stand-alone)
// Parameter all-gather as a push: every rank stores its shard into slot `rank` of the gathered buffer of ALL ranks -
// one multimem.st per 16 bytes when the segment has a multicast mapping (the switch replicates it), a loop of peer
// stores otherwise.  4 vectors per thread per trip keep enough bytes in flight to cover the NVLink round trip.
__global__ void __launch_bounds__(256) allgather_push_kernel(Peers P, const uint4* __restrict__ src, uint64_t dst_off, int64_t nvec) {
  const uint64_t dst0 = dst_off + (uint64_t)P.rank * (uint64_t)nvec * 16;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += 4 * stride) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) v[u] = src[i + u * stride];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t j = i + u * stride;
      if (j >= nvec) break;
      const uint64_t off = dst0 + (uint64_t)j * 16;
      if (P.mc) {
        uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
        mm_st_b32x4(reinterpret_cast<void*>(P.mc + off), w);
      } else {
        for (int p = 0; p < P.nranks; ++p) *reinterpret_cast<uint4*>(P.ptr[(P.rank + p) % P.nranks] + off) = v[u];
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------- K7
// Vocab-parallel embedding forward.  The reference looks rows up locally (zeros for ids another rank owns) and then
// all-reduces [tokens, hidden]; since exactly one rank owns each row, the sum is really a gather: every rank reads the
// row straight from the owner's table shard over NVLink.  One warp per token, 16-byte vectors.
template <typename T>
__global__ void __launch_bounds__(256) vocab_embed_gather_kernel(Peers P, uint64_t table_off, const int64_t* __restrict__ ids,
                                                                T* __restrict__ out, int64_t n_tokens, int hidden,
                                                                int64_t vocab_per_rank) {
  constexpr int V = 16 / sizeof(T);
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31;
  for (int64_t t = (int64_t)blockIdx.x * warps + (threadIdx.x >> 5); t < n_tokens; t += (int64_t)gridDim.x * warps) {
    const int64_t id = ids[t];
    const int64_t owner = id / vocab_per_rank;
    T* dst = out + t * hidden;
    if (id < 0 || owner >= P.nranks) {   // padding ids produce zeros, like the masked lookup they replace
      for (int c = lane * V; c < hidden; c += 32 * V) *reinterpret_cast<uint4*>(dst + c) = make_uint4(0, 0, 0, 0);
      continue;
    }
    const T* src = reinterpret_cast<const T*>(P.ptr[owner] + table_off) + (id - owner * vocab_per_rank) * hidden;
    for (int c = lane * V; c < hidden; c += 32 * V)
      *reinterpret_cast<uint4*>(dst + c) = *reinterpret_cast<const uint4*>(src + c);
  }
}

// ---------------------------------------------------------------------------------------------- stage-3 unit transport
// One launch moves every parameter (all-gather) or every gradient (reduce-scatter) of a ZeRO-3 unit: a table of up to
// kMaxSegs segments is passed by value, and the cross-GPU barriers run inside the kernel (grid_barrier_begin / _end),
// so a unit costs one launch instead of 3 per parameter.  Replaces the per-parameter all_gather / all_reduce + slice
// of the reference (group_sharded_stage3.py:741-803, 849-1118).
constexpr int kMaxSegs = 16;
struct AgSeg { uint64_t src_off, dst_off; int64_t nvec; };      // nvec 16-byte vectors per rank shard
struct AgSegs { AgSeg s[kMaxSegs]; int n; };
__global__ void __launch_bounds__(256, 2) allgather_multi_kernel(Peers P, AgSegs S, uint64_t pad_off, uint32_t epoch,
                                                             unsigned long long spin_limit) {
  grid_barrier_begin(P, pad_off, epoch, spin_limit);         // every rank may have its gathered buffer overwritten
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  constexpr int U = 8;                                       // 128 B per thread in flight
  for (int k = 0; k < S.n; ++k) {
    const uint4* src = reinterpret_cast<const uint4*>(P.ptr[P.rank] + S.s[k].src_off);
    const int64_t nvec = S.s[k].nvec;
    const uint64_t dst0 = S.s[k].dst_off + (uint64_t)P.rank * (uint64_t)nvec * 16;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += U * stride) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i + u * stride < nvec) v[u] = __ldcs(src + i + u * stride);
      if (P.mc) {
        // NVLS: one store per vector, the switch replicates it into every rank's buffer (this one included)
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t j = i + u * stride;
          if (j < nvec) {
            uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
            mm_st_b32x4(reinterpret_cast<void*>(P.mc + dst0 + (uint64_t)j * 16), w);
          }
        }
      } else {
        // P2P: the local copy stays on this GPU, one store per peer crosses NVLink (no loop back through the switch)
        for (int p = 0; p < P.nranks; ++p) {
          uint8_t* base = reinterpret_cast<uint8_t*>(P.ptr[(P.rank + p) % P.nranks] + dst0);
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int64_t j = i + u * stride;
            if (j < nvec) *reinterpret_cast<uint4*>(base + (uint64_t)j * 16) = v[u];
          }
        }
      }
    }
  }
  grid_barrier_end(P, pad_off, epoch + 1, spin_limit);       // all shards have landed everywhere
}

// Reduce-scatter of a unit: segment k is a padded gradient [nranks * shard_numel] (G = bf16 or fp32) staged at stage_off
// of every rank's segment; this rank's shard is summed over the ranks (in the switch with NVLS, by peer loads without),
// scaled and written - or, with accumulate, added - to its local fp32 accumulator `out` (the optimizer's main_grad
// slice).  U independent 16-byte loads per thread are in flight before the first is consumed: the loads cross NVLink
// (~2-3 us round trip), so bytes in flight, not CTAs, set the bandwidth.
struct RsSeg { uint64_t stage_off; int64_t shard_numel; float* out; };
struct RsSegs { RsSeg s[kMaxSegs]; int n; };

template <typename G> struct RsTraits;
template <> struct RsTraits<__nv_bfloat16> { static constexpr int kElems = 8; };   // per 16-byte vector
template <> struct RsTraits<float> { static constexpr int kElems = 4; };

template <typename G> __device__ __forceinline__ void rs_unpack(const uint4& r, float* g);
template <> __device__ __forceinline__ void rs_unpack<__nv_bfloat16>(const uint4& r, float* g) {
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[j]);
    g[2 * j] = __low2float(h);
    g[2 * j + 1] = __high2float(h);
  }
}
template <> __device__ __forceinline__ void rs_unpack<float>(const uint4& r, float* g) {
  g[0] = __uint_as_float(r.x); g[1] = __uint_as_float(r.y); g[2] = __uint_as_float(r.z); g[3] = __uint_as_float(r.w);
}
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {       // streaming 16-byte load (peer or local)
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

template <typename G>
__device__ __forceinline__ void rs_store(float* __restrict__ o, const float* g, float scale, int accumulate) {
  constexpr int E = RsTraits<G>::kElems;
#pragma unroll
  for (int e = 0; e < E; e += 4) {
    float4 v = make_float4(g[e] * scale, g[e + 1] * scale, g[e + 2] * scale, g[e + 3] * scale);
    if (accumulate) {
      const float4 a = *reinterpret_cast<const float4*>(o + e);
      v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
    }
    __stcs(reinterpret_cast<float4*>(o + e), v);
  }
}

template <typename G>
__global__ void __launch_bounds__(256, 2) reduce_scatter_multi_kernel(Peers P, RsSegs S, float scale, int accumulate,
                                                                     uint64_t pad_off, uint32_t epoch,
                                                                     unsigned long long spin_limit) {
  grid_barrier_begin(P, pad_off, epoch, spin_limit);         // every rank's staged gradients are complete
  constexpr int E = RsTraits<G>::kElems;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int k = 0; k < S.n; ++k) {
    const int64_t nvec = S.s[k].shard_numel / E;
    float* __restrict__ out = S.s[k].out;
    const uint64_t base = S.s[k].stage_off + (uint64_t)P.rank * (uint64_t)S.s[k].shard_numel * sizeof(G);
    if (P.mc) {
      // NVLS: the switch returns the sum; 8 requests per thread in flight
      constexpr int U = 8;
      for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += U * stride) {
        uint4 r[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t j = i + u * stride;
          if (j < nvec) {
            const void* a = reinterpret_cast<const void*>(P.mc + base + (uint64_t)j * 16);
            if (sizeof(G) == 4) {
              float t[4];
              mm_ld_reduce_f32x4(a, t);
              r[u] = make_uint4(__float_as_uint(t[0]), __float_as_uint(t[1]), __float_as_uint(t[2]), __float_as_uint(t[3]));
            } else {
              uint32_t t[4];
              mm_ld_reduce_bf16x8(a, t);
              r[u] = make_uint4(t[0], t[1], t[2], t[3]);
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t j = i + u * stride;
          if (j < nvec) {
            float g[E];
            rs_unpack<G>(r[u], g);
            rs_store<G>(out + j * E, g, scale, accumulate);
          }
        }
      }
    } else if (P.nranks <= 4) {
      // peer loads, up to 4 ranks: all ranks' copies of 4 vectors are requested before the first is consumed
      constexpr int U = 4;
      for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += U * stride) {
        uint4 r[4][U];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          if (p < P.nranks) {
            const uint8_t* src = reinterpret_cast<const uint8_t*>(P.ptr[(P.rank + p) % P.nranks] + base);
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const int64_t j = i + u * stride;
              if (j < nvec) r[p][u] = ld_peer_v4(src + (uint64_t)j * 16);
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t j = i + u * stride;
          if (j >= nvec) continue;
          float g[E];
          rs_unpack<G>(r[0][u], g);
#pragma unroll
          for (int p = 1; p < 4; ++p) {
            if (p < P.nranks) {
              float t[E];
              rs_unpack<G>(r[p][u], t);
#pragma unroll
              for (int e = 0; e < E; ++e) g[e] += t[e];
            }
          }
          rs_store<G>(out + j * E, g, scale, accumulate);
        }
      }
    } else {
      // peer loads, more than 4 ranks and no NVLS: one vector at a time per rank pair (fallback, not tuned)
      for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < nvec; j += stride) {
        float g[E];
#pragma unroll
        for (int e = 0; e < E; ++e) g[e] = 0.f;
        for (int p = 0; p < P.nranks; ++p) {
          float t[E];
          rs_unpack<G>(ld_peer_v4(reinterpret_cast<const uint8_t*>(P.ptr[(P.rank + p) % P.nranks] + base) + (uint64_t)j * 16), t);
#pragma unroll
          for (int e = 0; e < E; ++e) g[e] += t[e];
        }
        rs_store<G>(out + j * E, g, scale, accumulate);
      }
    }
  }
  grid_barrier_end(P, pad_off, epoch + 1, spin_limit);       // every rank has pulled: the stages may be rewritten
}

// Local half of the copy-engine reduce-scatter: slot r of `recv` holds rank r's contribution to this rank's shard
// (landed there by pdb200_symm_alltoall_ce); out (+)= scale * sum of the nranks slots.  Reads local memory only.
struct SlotSeg { const void* recv; int64_t shard_numel; float* out; };
struct SlotSegs { SlotSeg s[kMaxSegs]; int n; };
template <typename G>
__global__ void __launch_bounds__(256, 2) reduce_slots_multi_kernel(SlotSegs S, int nranks, float scale, int accumulate, int out_bf16) {
  constexpr int E = RsTraits<G>::kElems;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int k = 0; k < S.n; ++k) {
    const int64_t nvec = S.s[k].shard_numel / E;
    const uint8_t* base = reinterpret_cast<const uint8_t*>(S.s[k].recv);
    const int64_t slot_bytes = S.s[k].shard_numel * (int64_t)sizeof(G);
    float* __restrict__ out = S.s[k].out;
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < nvec; j += stride) {
      float g[E];
#pragma unroll
      for (int e = 0; e < E; ++e) g[e] = 0.f;
      for (int p0 = 0; p0 < nranks; p0 += 4) {             // 4 slots' loads in flight together
        uint4 r[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (p0 + u < nranks) r[u] = ld_peer_v4(base + (int64_t)(p0 + u) * slot_bytes + j * 16);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (p0 + u < nranks) {
            float t[E];
            rs_unpack<G>(r[u], t);
#pragma unroll
            for (int e = 0; e < E; ++e) g[e] += t[e];
          }
        }
      }
      if (out_bf16) {                                      // K6: the reduced activation goes back to bf16
        __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(out) + j * E;
#pragma unroll
        for (int e = 0; e < E; e += 2) *reinterpret_cast<__nv_bfloat162*>(ob + e) = __floats2bfloat162_rn(g[e] * scale, g[e + 1] * scale);
      } else {
        rs_store<G>(out + j * E, g, scale, accumulate);
      }
    }
  }
}

// NVLS or plain peer access?  An all-gather through the switch loops this rank's own shard back over NVLink and a
// switch reduction pulls every rank's full buffer out of its memory, so with few ranks peer loads / stores move
// fewer bytes; with many ranks the switch saves (n-1)x the store / load instructions and, for the reduction, the
// ingress.  PDB200_NVL_P2P_MAX_RANKS (default 4): groups up to this size use peer access.
static int p2p_max_ranks() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("PDB200_NVL_P2P_MAX_RANKS");
    v = e ? atoi(e) : 4;
  }
  return v;
}

static Peers make_peers(const uint64_t* ptrs, uint64_t mc, int nranks, int rank) {
  Peers P{};
  for (int i = 0; i < nranks && i < kMaxRanks; ++i) P.ptr[i] = ptrs[i];
  P.mc = mc;
  P.progress = g_nvl_progress;
  P.nranks = nranks;
  P.rank = rank;
  return P;
}

}  // namespace nvl
}  // namespace pdb

using namespace pdb;
using namespace pdb::nvl;

static const unsigned long long kSpinLimit = 20000000ull;   // ~ tens of seconds of system-scope polling, then trap

// Single-GPU emulation (tests): all "ranks" are segments of one GPU's memory and the entry points are called once per
// rank, one after the other, on one stream.  Nothing may wait for another rank then (the launches are serialised), so
// every cross-rank barrier is skipped; the data movement and arithmetic are the real kernels (peer-access paths).
static int g_emulate = 0;
PDB_EXPORT int pdb200_nvl_set_emulation(int on) {
  g_emulate = on;
  return 0;
}
#define PDB_BARRIER(...)                                    \
  do {                                                      \
    if (!g_emulate) barrier_kernel<<<1, 32, 0, st>>>(__VA_ARGS__); \
  } while (0)

// peer_ptrs: host array of nranks device addresses (this process's mappings).  Every entry point brackets its data
// kernel with signal-pad barriers using epochs `epoch` and `epoch + 1`; callers advance their counter by 2.
PDB_EXPORT int pdb200_nvl_barrier(const uint64_t* peer_ptrs, int nranks, int rank, int64_t pad_off, int64_t epoch,
                                  void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  Peers P = make_peers(peer_ptrs, 0, nranks, rank);
  if (!g_emulate) barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(P, (uint64_t)pad_off, (uint32_t)epoch, kSpinLimit);
  return check_launch("nvl_barrier");
}

PDB_EXPORT int pdb200_nvl_allreduce(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, int64_t data_off,
                                    int64_t numel, int dtype, float scale, int64_t pad_off, int64_t epoch, int max_ctas,
                                    void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  const int esz = dtype == PDB_DTYPE_F32 ? 4 : 2;
  PDB_REQUIRE((numel * esz) % 16 == 0 && data_off % 16 == 0, "nvl_allreduce: buffer must be a multiple of 16 bytes");
  PDB_REQUIRE(dtype == PDB_DTYPE_F32 || dtype == PDB_DTYPE_BF16, "nvl_allreduce: float32 / bfloat16 only");
  cudaStream_t st = (cudaStream_t)stream;
  Peers P = make_peers(peer_ptrs, mc_ptr, nranks, rank);
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)epoch, kSpinLimit);
  const int64_t nvec = numel * esz / 16;
  const int64_t per = (nvec + nranks - 1) / nranks;
  int grid = (int)((per + 511) / 512);
  const int cap = max_ctas > 0 ? max_ctas : sm_count();     // few CTAs when overlapping with backward compute
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  if (dtype == PDB_DTYPE_F32)
    allreduce_scale_kernel<float><<<grid, 512, 0, st>>>(P, (uint64_t)data_off, nvec, scale);
  else
    allreduce_scale_kernel<__nv_bfloat16><<<grid, 512, 0, st>>>(P, (uint64_t)data_off, nvec, scale);
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kSpinLimit);
  return check_launch("nvl_allreduce");
}

PDB_EXPORT int pdb200_nvl_rs_adamw_ag(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, int64_t grad_off,
                                      int grad_dtype, int64_t param_off, int64_t shard_start, int64_t shard_numel,
                                      float* master, float* m, float* v, float lr, float beta1, float beta2, float eps,
                                      float weight_decay, const float* beta_pows, const float* grad_scale,
                                      const float* found_inf, const float* local_grad, int64_t pad_off, int64_t epoch,
                                      void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  PDB_REQUIRE(shard_numel % 8 == 0 && shard_start % 8 == 0, "rs_adamw_ag: shard must be a multiple of 8 elements");
  PDB_REQUIRE(grad_dtype == PDB_DTYPE_F32 || grad_dtype == PDB_DTYPE_BF16, "rs_adamw_ag: grads must be fp32 or bf16");
  PDB_REQUIRE(beta_pows != nullptr, "rs_adamw_ag: beta_pows (device) is required");
  cudaStream_t st = (cudaStream_t)stream;
  Peers P = make_peers(peer_ptrs, mc_ptr, nranks, rank);
  AdamArgs a{lr, beta1, beta2, eps, weight_decay, beta_pows, grad_scale, found_inf};
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)epoch, kSpinLimit);
  int64_t blocks = (shard_numel / 8 + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (grad_dtype == PDB_DTYPE_F32)
    rs_adamw_ag_kernel<float><<<(int)blocks, 256, 0, st>>>(P, (uint64_t)grad_off, (uint64_t)param_off, shard_start, shard_numel, master, m, v, a, local_grad);
  else
    rs_adamw_ag_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, st>>>(P, (uint64_t)grad_off, (uint64_t)param_off, shard_start, shard_numel, master, m, v, a, local_grad);
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kSpinLimit);
  return check_launch("nvl_rs_adamw_ag");
}

// Reduce-scatter (AVG) into a local fp32 shard; partial_sq: >= pdb200_nvl_rs_grid(shard_numel) floats of scratch whose
// sum is the shard's squared L2 norm (fold with pdb200_l2norm / torch.sum).
static int g_rs_max_ctas = 0;     // 0 = 4 CTAs per SM; set lower when the kernel should leave the SMs to overlapping GEMMs
PDB_EXPORT int pdb200_nvl_set_rs_max_ctas(int n) {
  g_rs_max_ctas = n;
  return 0;
}
PDB_EXPORT int pdb200_nvl_rs_grid(int64_t shard_numel) {
  int64_t blocks = (shard_numel / 8 + 255) / 256;
  const int64_t cap = g_rs_max_ctas > 0 ? g_rs_max_ctas : (int64_t)sm_count() * 4;
  if (blocks > cap) blocks = cap;
  return (int)(blocks < 1 ? 1 : blocks);
}
PDB_EXPORT int pdb200_nvl_reduce_scatter(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, int64_t grad_off,
                                         int grad_dtype, int64_t shard_start, int64_t shard_numel, float* out,
                                         float* partial_sq, int64_t pad_off, int64_t epoch, void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  PDB_REQUIRE(shard_numel % 8 == 0 && shard_start % 8 == 0, "nvl_reduce_scatter: shard must be a multiple of 8 elements");
  cudaStream_t st = (cudaStream_t)stream;
  Peers P = make_peers(peer_ptrs, mc_ptr, nranks, rank);
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)epoch, kSpinLimit);
  const int grid = pdb200_nvl_rs_grid(shard_numel);
  if (grad_dtype == PDB_DTYPE_F32)
    reduce_scatter_norm_kernel<float><<<grid, 256, 0, st>>>(P, (uint64_t)grad_off, shard_start, shard_numel, out, partial_sq);
  else
    reduce_scatter_norm_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(P, (uint64_t)grad_off, shard_start, shard_numel, out, partial_sq);
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kSpinLimit);
  return check_launch("nvl_reduce_scatter");
}

PDB_EXPORT int pdb200_nvl_ce_combine(const uint64_t* peer_ptrs, int nranks, int rank, int64_t stats_off,
                                     const int64_t* labels, float* loss, float* lse, int64_t rows, int64_t ignore_index,
                                     int64_t pad_off, int64_t epoch, void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  cudaStream_t st = (cudaStream_t)stream;
  Peers P = make_peers(peer_ptrs, 0, nranks, rank);
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)epoch, kSpinLimit);
  if (rows > 0)
    ce_combine_symm_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(P, (uint64_t)stats_off, labels, loss, lse, rows, ignore_index);
  // peers may overwrite their stats only after everyone has read them
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kSpinLimit);
  return check_launch("nvl_ce_combine");
}

// K7: out[t, :] = table_of_owner(ids[t])[ids[t] - owner * vocab_per_rank, :].  Every rank's table shard
// [vocab_per_rank, hidden] sits at byte offset table_off of its symmetric segment.  The leading barrier orders the
// gather after every rank's last write to its shard (optimizer step); the trailing one keeps shards stable while read.
PDB_EXPORT int pdb200_nvl_vocab_embed(const uint64_t* peer_ptrs, int nranks, int rank, int64_t table_off,
                                      const int64_t* ids, void* out, int64_t n_tokens, int hidden,
                                      int64_t vocab_per_rank, int dtype, int64_t pad_off, int64_t epoch, void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  const int esz = dtype == PDB_DTYPE_F32 ? 4 : 2;
  PDB_REQUIRE((hidden * esz) % 16 == 0 && table_off % 16 == 0 && ((uintptr_t)out & 15) == 0,
              "vocab_embed: rows must be 16-byte multiples and 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  Peers P = make_peers(peer_ptrs, 0, nranks, rank);
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)epoch, kSpinLimit);
  if (n_tokens > 0) {
    int64_t blocks = (n_tokens + 7) / 8;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (dtype == PDB_DTYPE_F32)
      vocab_embed_gather_kernel<float><<<(unsigned)blocks, 256, 0, st>>>(P, (uint64_t)table_off, ids, (float*)out, n_tokens, hidden, vocab_per_rank);
    else
      vocab_embed_gather_kernel<uint16_t><<<(unsigned)blocks, 256, 0, st>>>(P, (uint64_t)table_off, ids, (uint16_t*)out, n_tokens, hidden, vocab_per_rank);
    if (int e = check_launch("vocab_embed_gather")) return e;
  }
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kSpinLimit);
  return check_launch("nvl_vocab_embed");
}

// K3 stand-alone: gathered[r * shard_bytes ...] on every rank = rank r's shard.  src_off: this rank's shard inside its
// own segment; dst_off: the gathered buffer (same offset in every segment).  The leading barrier certifies that every
// rank is ready to have that buffer overwritten, the trailing one that every shard has landed everywhere.
PDB_EXPORT int pdb200_nvl_allgather(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, int64_t src_off,
                                    int64_t dst_off, int64_t shard_bytes, int64_t pad_off, int64_t epoch, int max_ctas,
                                    void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  PDB_REQUIRE(shard_bytes % 16 == 0 && src_off % 16 == 0 && dst_off % 16 == 0, "nvl_allgather: 16-byte granularity");
  cudaStream_t st = (cudaStream_t)stream;
  Peers P = make_peers(peer_ptrs, mc_ptr, nranks, rank);
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)epoch, kSpinLimit);
  const int64_t nvec = shard_bytes / 16;
  if (nvec > 0) {
    int64_t blocks = (nvec + 1023) / 1024;
    const int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)sm_count() * 4;
    if (blocks > cap) blocks = cap;
    allgather_push_kernel<<<(unsigned)(blocks < 1 ? 1 : blocks), 256, 0, st>>>(
        P, reinterpret_cast<const uint4*>(peer_ptrs[rank] + (uint64_t)src_off), (uint64_t)dst_off, nvec);
    if (int e = check_launch("allgather_push")) return e;
  }
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kSpinLimit);
  return check_launch("nvl_allgather");
}

// Same push, but the shard is read from ordinary (non-symmetric) device memory: only the gathered buffer has to live
// in the symmetric segment.  Used by the sequence-parallel all-gather in front of a GEMM (K5).
PDB_EXPORT int pdb200_nvl_allgather_from(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, const void* src,
                                         int64_t dst_off, int64_t shard_bytes, int64_t pad_off, int64_t epoch,
                                         int max_ctas, void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  PDB_REQUIRE(shard_bytes % 16 == 0 && ((uintptr_t)src & 15) == 0 && dst_off % 16 == 0, "nvl_allgather_from: 16-byte granularity");
  cudaStream_t st = (cudaStream_t)stream;
  Peers P = make_peers(peer_ptrs, mc_ptr, nranks, rank);
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)epoch, kSpinLimit);
  const int64_t nvec = shard_bytes / 16;
  if (nvec > 0) {
    int64_t blocks = (nvec + 1023) / 1024;
    const int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)sm_count() * 4;
    if (blocks > cap) blocks = cap;
    allgather_push_kernel<<<(unsigned)(blocks < 1 ? 1 : blocks), 256, 0, st>>>(P, reinterpret_cast<const uint4*>(src), (uint64_t)dst_off, nvec);
    if (int e = check_launch("allgather_push")) return e;
  }
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kSpinLimit);
  return check_launch("nvl_allgather_from");
}

// out[shard_numel] (bf16, local) = sum over ranks of the symmetric bf16 buffer at src_off, elements
// [shard_start, shard_start + shard_numel).  Leading barrier: every rank's buffer is complete; trailing: all pulled.
PDB_EXPORT int pdb200_nvl_rs_sum_bf16(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank, int64_t src_off,
                                      int64_t shard_start, int64_t shard_numel, void* out, int64_t pad_off, int64_t epoch,
                                      void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  PDB_REQUIRE(shard_numel % 8 == 0 && shard_start % 8 == 0 && src_off % 16 == 0 && ((uintptr_t)out & 15) == 0,
              "nvl_rs_sum_bf16: 16-byte granularity");
  cudaStream_t st = (cudaStream_t)stream;
  Peers P = make_peers(peer_ptrs, mc_ptr, nranks, rank);
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)epoch, kSpinLimit);
  if (shard_numel > 0) {
    int64_t blocks = (shard_numel / 8 + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    rs_sum_bf16_kernel<<<(unsigned)(blocks < 1 ? 1 : blocks), 256, 0, st>>>(P, (uint64_t)src_off, shard_start, shard_numel,
                                                                          (__nv_bfloat16*)out);
    if (int e = check_launch("rs_sum_bf16")) return e;
  }
  PDB_BARRIER(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kSpinLimit);
  return check_launch("nvl_rs_sum_bf16");
}

// Stage-3 unit transport.  segs (HOST memory): nseg triples of int64.
//   all-gather:      {src_off, dst_off, shard_bytes}: this rank's shard at src_off of its own segment goes to
//                    [dst_off + r * shard_bytes) of every rank's segment (r = sender's rank).
//   reduce-scatter:  {stage_off, shard_numel, out_ptr}: see reduce_scatter_multi_kernel; out_ptr is a local fp32 pointer.
// Both use epochs `epoch` and `epoch + 1` of the pad and run their cross-GPU barriers inside the one kernel.
PDB_EXPORT int pdb200_nvl_allgather_multi(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank,
                                          const int64_t* segs, int nseg, int64_t pad_off, int64_t epoch, int max_ctas,
                                          void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  PDB_REQUIRE(nseg >= 1 && nseg <= kMaxSegs, "nvl_allgather_multi: 1..%d segments per launch (got %d)", kMaxSegs, nseg);
  AgSegs S{};
  S.n = nseg;
  int64_t most = 0;
  for (int k = 0; k < nseg; ++k) {
    const int64_t so = segs[3 * k], dof = segs[3 * k + 1], nb = segs[3 * k + 2];
    PDB_REQUIRE(nb >= 0 && nb % 16 == 0 && so % 16 == 0 && dof % 16 == 0, "nvl_allgather_multi: segment %d breaks the 16-byte granularity", k);
    S.s[k] = AgSeg{(uint64_t)so, (uint64_t)dof, nb / 16};
    if (nb / 16 > most) most = nb / 16;
  }
  Peers P = make_peers(peer_ptrs, nranks > p2p_max_ranks() ? mc_ptr : 0, nranks, rank);
  int64_t blocks = (most + 2047) / 2048;
  const int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)sm_count();
  if (blocks > cap) blocks = cap;
  allgather_multi_kernel<<<(unsigned)(blocks < 1 ? 1 : blocks), 256, 0, (cudaStream_t)stream>>>(P, S, (uint64_t)pad_off, (uint32_t)epoch, g_emulate ? kNoBarrier : kSpinLimit);
  return check_launch("nvl_allgather_multi");
}

PDB_EXPORT int pdb200_nvl_reduce_scatter_multi(const uint64_t* peer_ptrs, uint64_t mc_ptr, int nranks, int rank,
                                               const int64_t* segs, int nseg, int grad_dtype, float scale, int accumulate,
                                               int64_t pad_off, int64_t epoch, int max_ctas, void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl: nranks %d out of range (1..%d)", nranks, kMaxRanks);
  PDB_REQUIRE(nseg >= 1 && nseg <= kMaxSegs, "nvl_reduce_scatter_multi: 1..%d segments per launch (got %d)", kMaxSegs, nseg);
  PDB_REQUIRE(grad_dtype == PDB_DTYPE_F32 || grad_dtype == PDB_DTYPE_BF16, "nvl_reduce_scatter_multi: grads must be fp32 or bf16");
  RsSegs S{};
  S.n = nseg;
  int64_t most = 0;
  for (int k = 0; k < nseg; ++k) {
    const int64_t so = segs[3 * k], ne = segs[3 * k + 1];
    float* out = reinterpret_cast<float*>((uintptr_t)segs[3 * k + 2]);
    PDB_REQUIRE(ne >= 0 && ne % 8 == 0 && so % 16 == 0 && ((uintptr_t)out & 15) == 0 && out != nullptr,
                "nvl_reduce_scatter_multi: segment %d: shard must be a multiple of 8 elements, 16-byte aligned", k);
    S.s[k] = RsSeg{(uint64_t)so, ne, out};
    if (ne / (grad_dtype == PDB_DTYPE_F32 ? 4 : 8) > most) most = ne / (grad_dtype == PDB_DTYPE_F32 ? 4 : 8);
  }
  Peers P = make_peers(peer_ptrs, nranks > p2p_max_ranks() ? mc_ptr : 0, nranks, rank);
  int64_t blocks = (most + 2047) / 2048;
  const int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)sm_count();
  if (blocks > cap) blocks = cap;
  const unsigned grid = (unsigned)(blocks < 1 ? 1 : blocks);
  cudaStream_t st = (cudaStream_t)stream;
  if (grad_dtype == PDB_DTYPE_F32)
    reduce_scatter_multi_kernel<float><<<grid, 256, 0, st>>>(P, S, scale, accumulate, (uint64_t)pad_off, (uint32_t)epoch, g_emulate ? kNoBarrier : kSpinLimit);
  else
    reduce_scatter_multi_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(P, S, scale, accumulate, (uint64_t)pad_off, (uint32_t)epoch, g_emulate ? kNoBarrier : kSpinLimit);
  return check_launch("nvl_reduce_scatter_multi");
}

PDB_EXPORT int pdb200_reduce_slots_multi2(const int64_t* segs, int nseg, int nranks, int grad_dtype, float scale, int accumulate,
                                          int out_dtype, int max_ctas, void* stream);
// Local sum of receive slots: segs (HOST) = nseg triples {recv_ptr (local device pointer), shard_numel, out_ptr (fp32)}.
PDB_EXPORT int pdb200_reduce_slots_multi(const int64_t* segs, int nseg, int nranks, int grad_dtype, float scale, int accumulate,
                                         int max_ctas, void* stream) {
  return pdb200_reduce_slots_multi2(segs, nseg, nranks, grad_dtype, scale, accumulate, 0, max_ctas, stream);
}
// Same with the output precision chosen: out_dtype 0 = fp32 (accumulate allowed), 1 = bf16 (out pointers are bf16).
PDB_EXPORT int pdb200_reduce_slots_multi2(const int64_t* segs, int nseg, int nranks, int grad_dtype, float scale, int accumulate,
                                          int out_dtype, int max_ctas, void* stream) {
  PDB_REQUIRE(out_dtype == 0 || (out_dtype == 1 && !accumulate), "reduce_slots_multi: bf16 output cannot accumulate");
  const int out_bf16 = out_dtype == 1;
  PDB_REQUIRE(nseg >= 1 && nseg <= kMaxSegs, "reduce_slots_multi: 1..%d segments per launch (got %d)", kMaxSegs, nseg);
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "reduce_slots_multi: nranks %d out of range", nranks);
  PDB_REQUIRE(grad_dtype == PDB_DTYPE_F32 || grad_dtype == PDB_DTYPE_BF16, "reduce_slots_multi: grads must be fp32 or bf16");
  SlotSegs S{};
  S.n = nseg;
  int64_t most = 0;
  const int E = grad_dtype == PDB_DTYPE_F32 ? 4 : 8;
  for (int k = 0; k < nseg; ++k) {
    const void* recv = reinterpret_cast<const void*>((uintptr_t)segs[3 * k]);
    const int64_t ne = segs[3 * k + 1];
    float* out = reinterpret_cast<float*>((uintptr_t)segs[3 * k + 2]);
    PDB_REQUIRE(ne >= 0 && ne % 8 == 0 && ((uintptr_t)recv & 15) == 0 && ((uintptr_t)out & 15) == 0 && out && recv,
                "reduce_slots_multi: segment %d: shard must be a multiple of 8 elements, 16-byte aligned", k);
    S.s[k] = SlotSeg{recv, ne, out};
    if (ne / E > most) most = ne / E;
  }
  int64_t blocks = (most + 255) / 256;
  const int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)sm_count() * 2;
  if (blocks > cap) blocks = cap;
  const unsigned grid = (unsigned)(blocks < 1 ? 1 : blocks);
  cudaStream_t st = (cudaStream_t)stream;
  if (grad_dtype == PDB_DTYPE_F32) reduce_slots_multi_kernel<float><<<grid, 256, 0, st>>>(S, nranks, scale, accumulate, out_bf16);
  else reduce_slots_multi_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(S, nranks, scale, accumulate, out_bf16);
  return check_launch("reduce_slots_multi");
}

// One-thread-per-peer signal: flag word `flag_off + 4 * index` of EVERY rank's segment <- value, released at system
// scope.  Enqueued behind the copies of a chunk (stream order = the copies have completed), it tells the consumers -
// a gated GEMM polling its local flag, or a later kernel - that the chunk is in their memory.
namespace pdb { namespace nvl {
__global__ void signal_peers_kernel(Peers P, uint64_t flag_off, int index, uint32_t value, int only) {
  const int t = threadIdx.x;
  if (t < P.nranks && (only < 0 || t == only)) {
    fence_sys();
    st_release_sys(reinterpret_cast<uint32_t*>(P.ptr[t] + flag_off) + index, value);
  }
}
} }
PDB_EXPORT int pdb200_nvl_signal(const uint64_t* peer_ptrs, int nranks, int rank, int64_t flag_off, int index, int64_t value, void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "nvl_signal: nranks %d out of range", nranks);
  Peers P = make_peers(peer_ptrs, 0, nranks, rank);
  signal_peers_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(P, (uint64_t)flag_off, index, (uint32_t)value, -1);
  return check_launch("nvl_signal");
}
PDB_EXPORT int pdb200_nvl_signal_one(const uint64_t* peer_ptrs, int nranks, int rank, int target, int64_t flag_off, int index, int64_t value,
                                     void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks && target >= 0 && target < nranks, "nvl_signal_one: bad rank");
  Peers P = make_peers(peer_ptrs, 0, nranks, rank);
  signal_peers_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(P, (uint64_t)flag_off, index, (uint32_t)value, target);
  return check_launch("nvl_signal_one");
}

// Stream-ordered waits used to chain copy-engine transfers behind in-kernel progress:
//   wait_counter: spin until *counter (this GPU's memory, written with device-scope atomics by a running kernel) has
//                 reached `target`; what follows on the stream (copies of the finished block) starts only then.
//   wait_flags:   spin until every one of `count` local flag words has reached `epoch` (set by peers' signal kernels).
namespace pdb { namespace nvl {
__global__ void wait_counter_kernel(const unsigned int* counter, unsigned int target, unsigned long long spin_limit) {
  unsigned long long spins = 0;
  while (true) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
    if ((int)(v - target) >= 0) break;
    if (spin_limit && ++spins > spin_limit) {
      printf("[pdb200 nvl] wait_counter timed out (saw %u, want %u)\n", v, target);
      __trap();
    }
    __nanosleep(200);
  }
  __threadfence_system();
}
__global__ void wait_flags_kernel(const uint32_t* flags, int count, uint32_t epoch, int skip, unsigned long long spin_limit) {
  const int t = threadIdx.x;
  if (t < count && t != skip) {
    unsigned long long spins = 0;
    while ((int32_t)(ld_acquire_sys(flags + t) - epoch) < 0) {
      if (spin_limit && ++spins > spin_limit) {
        printf("[pdb200 nvl] wait_flags timed out on flag %d (saw %u, want %u)\n", t, ld_acquire_sys(flags + t), epoch);
        __trap();
      }
      __nanosleep(200);
    }
  }
  __syncthreads();
  __threadfence_system();
}
} }
PDB_EXPORT int pdb200_nvl_wait_counter(const void* counter, int64_t target, void* stream) {
  wait_counter_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((const unsigned int*)counter, (unsigned int)target, 200000000ull);
  return check_launch("nvl_wait_counter");
}
PDB_EXPORT int pdb200_nvl_wait_flags(const void* flags, int count, int64_t epoch, int skip_index, void* stream) {
  PDB_REQUIRE(count >= 1 && count <= 1024, "nvl_wait_flags: 1..1024 flags (got %d)", count);
  wait_flags_kernel<<<1, (count + 31) / 32 * 32, 0, (cudaStream_t)stream>>>((const uint32_t*)flags, count, (uint32_t)epoch, skip_index,
                                                                        200000000ull);
  return check_launch("nvl_wait_flags");
}
