// K9 / K10: MoE token dispatch and combine over NVLink symmetric memory, with the routing done on the device.
//
// Replaces in the reference: global_scatter_op.cu.cc:33-200 / global_gather_op.cu.cc (the per-(expert, rank) counts are
// copied to the HOST, then E x W grouped ncclSend/ncclRecv move the rows) together with the count all-to-all and the
// local index_select / scatter around them (moe_layer.py:408-503).
//
// Here every rank publishes its [W x E] send counts in its symmetric segment; after one in-stream barrier each sender
// computes, on the device, where its rows for (destination rank r, local expert i) start inside r's expert-major
// receive slab (prefix sums over the published counts of ALL ranks) and stores the rows straight into that slab over
// NVLink, fusing the local gather by `pos`.  No count all-to-all, no host round trip, no reorder copies: the slab is
// already laid out expert by expert, sources in rank order, which is what the expert GEMMs consume.
// The way back is a pull: each token (or token x k entry) reads its expert-output row from the slab of the rank that
// computed it, using the (rank, row) route recorded during dispatch; `sum` mode adds the k rows (dispatch backward).
#include "common.cuh"
#include "nvl.cuh"

namespace pdb {
namespace moe_nvl {
using namespace pdb::nvl;

__global__ void publish_counts_kernel(Peers P, uint64_t counts_off, const int64_t* __restrict__ lec, int n) {
  int32_t* mine = reinterpret_cast<int32_t*>(P.ptr[P.rank] + counts_off);
  for (int i = threadIdx.x; i < n; i += blockDim.x) mine[i] = (int32_t)lec[i];
}

__global__ void barrier_kernel(Peers P, uint64_t pad_off, uint32_t epoch, unsigned long long spin_limit) {
  barrier_all(P, pad_off, epoch, spin_limit);
}

// recv_counts[i] = rows this rank receives for its local expert i (from all ranks); recv_counts[E] = their sum.
__global__ void recv_counts_kernel(Peers P, uint64_t counts_off, int E, int64_t* __restrict__ recv_counts) {
  __shared__ int64_t tot[64];
  const int i = threadIdx.x;
  if (i < E) {
    int64_t t = 0;
    for (int j = 0; j < P.nranks; ++j) t += reinterpret_cast<const int32_t*>(P.ptr[j] + counts_off)[i + P.rank * E];
    recv_counts[i] = t;
    tot[i] = t;
  }
  __syncthreads();
  if (i == 0) {
    int64_t s = 0;
    for (int k = 0; k < E; ++k) s += tot[k];
    recv_counts[E] = s;
  }
}

// grid = (x, W*E): block row b = r*E + i serves destination rank r, its local expert i.  One warp per row.
__global__ void __launch_bounds__(256) push_rows_kernel(Peers P, uint64_t counts_off, uint64_t slab_off,
                                                       const uint8_t* __restrict__ src, const int64_t* __restrict__ pos,
                                                       int div, int E, int row_bytes, int64_t slab_rows,
                                                       int32_t* __restrict__ route_rank, int32_t* __restrict__ route_row,
                                                       int32_t* __restrict__ overflow) {
  __shared__ int s_cnt, s_src, s_dst;
  const int b = blockIdx.y, r = b / E, i = b - r * E;
  if (threadIdx.x == 0) {
    const int32_t* mine = reinterpret_cast<const int32_t*>(P.ptr[P.rank] + counts_off);
    int src_start = 0;
    for (int k = 0; k < b; ++k) src_start += mine[k];
    int dst = 0;
    for (int j = 0; j < P.nranks; ++j) {
      const int32_t* cj = reinterpret_cast<const int32_t*>(P.ptr[j] + counts_off) + r * E;
      for (int k = 0; k < i; ++k) dst += cj[k];          // every source's rows for the experts before i
      if (j < P.rank) dst += cj[i];                       // lower-ranked sources of expert i come first
    }
    s_cnt = mine[b];
    s_src = src_start;
    s_dst = dst;
  }
  __syncthreads();
  const int cnt = s_cnt, src_start = s_src, dst_off = s_dst;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* slab = reinterpret_cast<uint8_t*>(P.ptr[r] + slab_off);
  const int nvec = row_bytes / 16;
  for (int k = blockIdx.x * 8 + warp; k < cnt; k += gridDim.x * 8) {
    const int s = src_start + k;
    const int64_t drow = (int64_t)dst_off + k;
    if (drow >= slab_rows) {
      if (lane == 0) *overflow = 1;
      continue;
    }
    const uint4* sp = reinterpret_cast<const uint4*>(src + (pos[s] / div) * (int64_t)row_bytes);
    uint4* dp = reinterpret_cast<uint4*>(slab + drow * (int64_t)row_bytes);
    for (int v = lane; v < nvec; v += 32) dp[v] = sp[v];
    if (lane == 0) {
      route_rank[s] = r;
      route_row[s] = (int32_t)drow;
    }
  }
}

// gather mode: out[e] = slab row routed for entry e (token*topk + k), zeros where the entry was dropped.
__global__ void __launch_bounds__(256) pull_rows_kernel(Peers P, uint64_t slab_off, const int32_t* __restrict__ route_rank,
                                                       const int32_t* __restrict__ route_row, const int64_t* __restrict__ inv,
                                                       int64_t n_entries, uint8_t* __restrict__ out, int row_bytes) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = row_bytes / 16;
  for (int64_t e = (int64_t)blockIdx.x * 8 + warp; e < n_entries; e += (int64_t)gridDim.x * 8) {
    const int64_t s = inv[e];
    uint4* dp = reinterpret_cast<uint4*>(out + e * (int64_t)row_bytes);
    if (s < 0) {
      for (int v = lane; v < nvec; v += 32) dp[v] = make_uint4(0, 0, 0, 0);
      continue;
    }
    const uint4* sp = reinterpret_cast<const uint4*>(P.ptr[route_rank[s]] + slab_off + (int64_t)route_row[s] * row_bytes);
    for (int v = lane; v < nvec; v += 32) dp[v] = sp[v];
  }
}

// sum mode: out[t] = sum_k slab row of entry (t, k); fp32 accumulation.
template <typename T>
__global__ void __launch_bounds__(256) pull_sum_kernel(Peers P, uint64_t slab_off, const int32_t* __restrict__ route_rank,
                                                      const int32_t* __restrict__ route_row, const int64_t* __restrict__ inv,
                                                      int64_t n_tokens, int topk, T* __restrict__ out, int d) {
  constexpr int V = 16 / sizeof(T);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int64_t t = (int64_t)blockIdx.x * 8 + warp; t < n_tokens; t += (int64_t)gridDim.x * 8) {
    for (int c = lane * V; c < d; c += 32 * V) {
      float acc[V];
#pragma unroll
      for (int u = 0; u < V; ++u) acc[u] = 0.f;
      for (int k = 0; k < topk; ++k) {
        const int64_t s = inv[t * topk + k];
        if (s < 0) continue;
        Vec<T> x;
        x.load(reinterpret_cast<const T*>(P.ptr[route_rank[s]] + slab_off) + (int64_t)route_row[s] * d + c);
#pragma unroll
        for (int u = 0; u < V; ++u) acc[u] += x.f(u);
      }
      Vec<T> o;
#pragma unroll
      for (int u = 0; u < V; ++u) o.set(u, acc[u]);
      o.store(out + t * d + c);
    }
  }
}

static Peers make_peers(const uint64_t* ptrs, int nranks, int rank) {
  Peers P{};
  for (int i = 0; i < nranks && i < kMaxRanks; ++i) P.ptr[i] = ptrs[i];
  P.progress = g_nvl_progress;
  P.nranks = nranks;
  P.rank = rank;
  return P;
}
static const unsigned long long kSpin = 20000000ull;

}  // namespace moe_nvl
}  // namespace pdb

using namespace pdb;
using namespace pdb::moe_nvl;

// K9 dispatch.  src: [*, row_bytes] rows (local), pos: [n_slots] int64 sorted slot -> flattened (token*topk + k)
// index, row taken = pos / div (div = topk when src is per token, 1 when src is per entry).  lec: device int64
// [nranks * n_expert_local] send counts (rank-major, expert-minor).  counts_off / slab_off: offsets inside every rank's
// segment (4 * nranks * n_expert_local bytes of counts; slab_rows * row_bytes of slab).  Outputs: route_rank/route_row
// [n_slots] (where each slot landed), recv_counts [n_expert_local + 1] (device), overflow flag (device int32, set when
// a destination slab was too small; those rows are dropped).  Uses epochs epoch, epoch + 1.
PDB_EXPORT int pdb200_moe_dispatch(const uint64_t* peer_ptrs, int nranks, int rank, int64_t counts_off, int64_t slab_off,
                                   int64_t slab_rows, const void* src, const int64_t* pos, int div, const int64_t* lec,
                                   int n_expert_local, int64_t n_slots, int row_bytes, int32_t* route_rank,
                                   int32_t* route_row, int64_t* recv_counts, int32_t* overflow, int64_t pad_off,
                                   int64_t epoch, void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "moe_dispatch: nranks %d out of range", nranks);
  PDB_REQUIRE(n_expert_local >= 1 && n_expert_local <= 64, "moe_dispatch: 1..64 local experts supported (got %d)", n_expert_local);
  PDB_REQUIRE(row_bytes % 16 == 0 && slab_off % 16 == 0 && ((uintptr_t)src & 15) == 0, "moe_dispatch: rows must be 16-byte multiples");
  PDB_REQUIRE(div >= 1, "moe_dispatch: div must be >= 1");
  cudaStream_t st = (cudaStream_t)stream;
  Peers P = make_peers(peer_ptrs, nranks, rank);
  const int nb = nranks * n_expert_local;
  publish_counts_kernel<<<1, 128, 0, st>>>(P, (uint64_t)counts_off, lec, nb);
  moe_nvl::barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)epoch, kSpin);      // counts visible, slabs free
  if (n_slots > 0) {
    int64_t x = (n_slots / nb + 7) / 8;          // warps needed if the rows were spread evenly; imbalance just loops
    if (x < 1) x = 1;
    if (x > 32) x = 32;
    push_rows_kernel<<<dim3((unsigned)x, (unsigned)nb), 256, 0, st>>>(P, (uint64_t)counts_off, (uint64_t)slab_off, (const uint8_t*)src, pos,
                                                                  div, n_expert_local, row_bytes, slab_rows, route_rank,
                                                                  route_row, overflow);
    if (int e = check_launch("moe push_rows")) return e;
  }
  recv_counts_kernel<<<1, 64, 0, st>>>(P, (uint64_t)counts_off, n_expert_local, recv_counts);
  moe_nvl::barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kSpin);  // every row has landed
  return check_launch("moe_dispatch");
}

// K10 combine (and dispatch backward).  Every rank's slab (at slab_off) holds its expert-major rows; entry e of this
// rank reads row route_*[inv[e]] from the rank that holds it.  sum_over_k = 0: out is [n_entries, row_bytes] (gather);
// sum_over_k = topk > 0: out is [n_entries / topk, d] with the k rows added (dtype: 0 fp32, 1 bf16).
PDB_EXPORT int pdb200_moe_pull(const uint64_t* peer_ptrs, int nranks, int rank, int64_t slab_off, const int32_t* route_rank,
                               const int32_t* route_row, const int64_t* inv, int64_t n_entries, void* out, int row_bytes,
                               int sum_over_k, int dtype, int64_t pad_off, int64_t epoch, void* stream) {
  PDB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, "moe_pull: nranks %d out of range", nranks);
  PDB_REQUIRE(row_bytes % 16 == 0 && slab_off % 16 == 0 && ((uintptr_t)out & 15) == 0, "moe_pull: rows must be 16-byte multiples");
  cudaStream_t st = (cudaStream_t)stream;
  Peers P = make_peers(peer_ptrs, nranks, rank);
  moe_nvl::barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)epoch, kSpin);      // every slab is filled
  if (n_entries > 0) {
    const int64_t rows = sum_over_k > 0 ? n_entries / sum_over_k : n_entries;
    int64_t blocks = (rows + 7) / 8;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (sum_over_k > 0) {
      if (dtype == PDB_DTYPE_F32)
        pull_sum_kernel<float><<<(unsigned)blocks, 256, 0, st>>>(P, (uint64_t)slab_off, route_rank, route_row, inv, rows, sum_over_k,
                                                                (float*)out, row_bytes / 4);
      else
        pull_sum_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(P, (uint64_t)slab_off, route_rank, route_row, inv, rows,
                                                                        sum_over_k, (__nv_bfloat16*)out, row_bytes / 2);
    } else {
      pull_rows_kernel<<<(unsigned)blocks, 256, 0, st>>>(P, (uint64_t)slab_off, route_rank, route_row, inv, n_entries, (uint8_t*)out, row_bytes);
    }
    if (int e = check_launch("moe pull")) return e;
  }
  moe_nvl::barrier_kernel<<<1, 32, 0, st>>>(P, (uint64_t)pad_off, (uint32_t)(epoch + 1), kSpin);  // slabs may be overwritten
  return check_launch("moe_pull");
}
