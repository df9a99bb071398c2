// K11: MoE routing utilities for sm_100a.
//
// Replaces in the reference: paddle/phi/kernels/gpu/number_count_kernel.cu:37-80, assign_pos_kernel.cu:32-43,
// limit_by_capacity_kernel.cu:24-45, prune_gate_by_capacity_kernel.cu:33-50, random_routing_kernel.cu:34-60.
// Differences by design: (1) `moe_route` does count + exclusive scan + stable position assignment in ONE launch so
// the layer never copies counts to the host between them (the reference syncs D2H three times per layer, SURVEY
// §3.6); (2) assign_pos is stable (token order preserved inside an expert) so runs are bit-reproducible.
#include "common.cuh"

namespace pdb {

// histogram of ids in [0, n_bins); negative ids ignored.  Shared-memory bins then one global atomic per bin/CTA.
__global__ void number_count_kernel(const int64_t* __restrict__ ids, int64_t n, int64_t* __restrict__ out, int n_bins) {
  extern __shared__ int sbins[];
  for (int i = threadIdx.x; i < n_bins; i += blockDim.x) sbins[i] = 0;
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t v = ids[i];
    if (v >= 0 && v < n_bins) atomicAdd(&sbins[v], 1);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n_bins; i += blockDim.x)
    if (sbins[i]) atomicAdd((unsigned long long*)&out[i], (unsigned long long)sbins[i]);
}

// Stable assign_pos: out[offset(e) + rank_of_token_within_e] = token index.  One CTA per expert-bin walks the id
// array in order with a block-wide ballot scan, so the result is deterministic.  `cum` is the INCLUSIVE cumsum of
// counts (the reference's contract), out has cum[last] entries.
__global__ void __launch_bounds__(256) assign_pos_stable_kernel(const int64_t* __restrict__ ids, int64_t n,
                                                               const int64_t* __restrict__ cum,
                                                               int64_t* __restrict__ out) {
  __shared__ int warp_cnt[8];
  __shared__ int64_t base;
  const int e = blockIdx.x;
  if (threadIdx.x == 0) base = e == 0 ? 0 : cum[e - 1];
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int64_t start = 0; start < n; start += blockDim.x) {
    const int64_t i = start + threadIdx.x;
    const bool hit = i < n && ids[i] == e;
    const unsigned bal = __ballot_sync(0xffffffffu, hit);
    if (lane == 0) warp_cnt[wid] = __popc(bal);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      if (w < wid) before += warp_cnt[w];
      total += warp_cnt[w];
    }
    if (hit) out[base + before + __popc(bal & ((1u << lane) - 1))] = i;
    __syncthreads();
    if (threadIdx.x == 0) base += total;
    __syncthreads();
  }
}

// out[w, e] = min(count[w, e], remaining capacity of e after workers < w took theirs)  (deterministic by worker).
__global__ void limit_by_capacity_kernel(const int64_t* __restrict__ expc, const int64_t* __restrict__ cap,
                                         int64_t* __restrict__ out, int n_expert, int n_worker) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_expert) return;
  int64_t left = cap[e];
  for (int w = 0; w < n_worker; ++w) {
    const int64_t want = expc[(int64_t)w * n_expert + e];
    const int64_t take = left >= want ? want : (left > 0 ? left : 0);
    out[(int64_t)w * n_expert + e] = take;
    left -= want;
  }
}

// new_gate[i] = gate[i] if a slot of expert gate[i] is still free else -1.  `remaining` is consumed (atomics).
__global__ void prune_gate_kernel(const int64_t* __restrict__ gate, int64_t* __restrict__ new_gate,
                                  int* __restrict__ remaining, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t g = gate[i];
    if (g < 0) {
      new_gate[i] = -1;
      continue;
    }
    const int old = atomicSub(&remaining[g], 1);
    new_gate[i] = old > 0 ? g : -1;
  }
}

__global__ void i64_to_i32_kernel(const int64_t* in, int* out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (int)in[i];
}

template <typename T>
__global__ void random_routing_kernel(int64_t* __restrict__ topk_idx, const T* __restrict__ topk_val,
                                      const T* __restrict__ prob, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && 2.f * to_f32(topk_val[i * 2 + 1]) < to_f32(prob[i])) topk_idx[i * 2 + 1] = -1;
}

// Fused router for top-k gating: per token softmax over E experts (E <= 256), top-k selection, renormalised
// weights, and the expert histogram - one warp per token, no intermediate [tokens, E] probability tensor.
template <typename T>
__global__ void __launch_bounds__(256) topk_softmax_kernel(const T* __restrict__ logits, int64_t tokens, int E, int k,
                                                          int renorm, float* __restrict__ weights,
                                                          int64_t* __restrict__ idx, int64_t* __restrict__ counts) {
  const int lane = threadIdx.x & 31;
  const int64_t tok = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (tok >= tokens) return;
  float v[8];  // E <= 256
  float m = -INFINITY;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int e = lane + j * 32;
    v[j] = e < E ? to_f32(logits[tok * E + e]) : -INFINITY;
    m = fmaxf(m, v[j]);
  }
  m = warp_max(m);
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    v[j] = (lane + j * 32 < E) ? __expf(v[j] - m) : 0.f;
    s += v[j];
  }
  s = warp_sum(s);
  float wsum = 0.f;
  float picked_w[8];
  int picked_e[8];
  for (int r = 0; r < k; ++r) {
    float best = -1.f;
    int be = -1;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (v[j] > best) {
        best = v[j];
        be = lane + j * 32;
      }
    // warp arg-max, ties to the lower expert id
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oe = __shfl_xor_sync(0xffffffffu, be, o);
      if (ob > best || (ob == best && oe >= 0 && (be < 0 || oe < be))) {
        best = ob;
        be = oe;
      }
    }
    if (be >= 0 && (be & 31) == lane) v[be >> 5] = -1.f;
    picked_w[r] = best / s;
    picked_e[r] = be;
    wsum += best / s;
  }
  if (lane == 0) {
    for (int r = 0; r < k; ++r) {
      weights[tok * k + r] = renorm ? picked_w[r] / wsum : picked_w[r];
      idx[tok * k + r] = picked_e[r];
      if (counts && picked_e[r] >= 0) atomicAdd((unsigned long long*)&counts[picked_e[r]], 1ull);
    }
  }
}

}  // namespace pdb

using namespace pdb;

PDB_EXPORT int pdb200_number_count(const int64_t* ids, int64_t n, int64_t* out, int n_bins, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  PDB_REQUIRE(n_bins > 0 && n_bins <= 8192, "number_count: n_bins %d out of range (1..8192)", n_bins);
  PDB_CUDA_OK(cudaMemsetAsync(out, 0, sizeof(int64_t) * n_bins, st));
  if (n == 0) return 0;
  int g = (int)((n + 255) / 256);
  const int cap = sm_count() * 4;
  if (g > cap) g = cap;
  number_count_kernel<<<g, 256, n_bins * sizeof(int), st>>>(ids, n, out, n_bins);
  return check_launch("number_count");
}

PDB_EXPORT int pdb200_assign_pos(const int64_t* ids, int64_t n, const int64_t* cum_count, int n_bins, int64_t* out,
                                 void* stream) {
  if (n == 0 || n_bins == 0) return 0;
  assign_pos_stable_kernel<<<n_bins, 256, 0, (cudaStream_t)stream>>>(ids, n, cum_count, out);
  return check_launch("assign_pos");
}

PDB_EXPORT int pdb200_limit_by_capacity(const int64_t* expert_count, const int64_t* capacity, int64_t* out,
                                        int n_expert, int n_worker, void* stream) {
  if (n_expert == 0) return 0;
  limit_by_capacity_kernel<<<(n_expert + 127) / 128, 128, 0, (cudaStream_t)stream>>>(expert_count, capacity, out,
                                                                                    n_expert, n_worker);
  return check_launch("limit_by_capacity");
}

// remaining_i32: caller-provided int32[n_bins] workspace (overwritten).
PDB_EXPORT int pdb200_prune_gate_by_capacity(const int64_t* gate, int64_t* new_gate, const int64_t* expert_count,
                                             int* remaining_i32, int n_bins, int64_t n, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  i64_to_i32_kernel<<<(n_bins + 255) / 256, 256, 0, st>>>(expert_count, remaining_i32, n_bins);
  if (n > 0) {
    int g = (int)((n + 255) / 256);
    prune_gate_kernel<<<g, 256, 0, st>>>(gate, new_gate, remaining_i32, n);
  }
  return check_launch("prune_gate_by_capacity");
}

PDB_EXPORT int pdb200_random_routing(int64_t* topk_idx, const void* topk_val, const void* prob, int64_t n, int dtype,
                                     void* stream) {
  if (n == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  PDB_DISPATCH_FLOAT(dtype, T, {
    random_routing_kernel<T><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(topk_idx, (const T*)topk_val, (const T*)prob, n);
  });
  return check_launch("random_routing");
}

PDB_EXPORT int pdb200_topk_softmax(const void* logits, int64_t tokens, int E, int k, int renorm, float* weights,
                                   int64_t* idx, int64_t* counts, int dtype, void* stream) {
  PDB_REQUIRE(E > 0 && E <= 256 && k > 0 && k <= 8 && k <= E, "topk_softmax: need E<=256, k<=8 (E=%d k=%d)", E, k);
  cudaStream_t st = (cudaStream_t)stream;
  if (counts) PDB_CUDA_OK(cudaMemsetAsync(counts, 0, sizeof(int64_t) * E, st));
  if (tokens == 0) return 0;
  PDB_DISPATCH_FLOAT(dtype, T, {
    topk_softmax_kernel<T><<<(unsigned)((tokens + 7) / 8), 256, 0, st>>>((const T*)logits, tokens, E, k, renorm, weights, idx, counts);
  });
  return check_launch("topk_softmax");
}
