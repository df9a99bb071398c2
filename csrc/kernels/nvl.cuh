// NVLink 5 / NVSwitch device-side primitives: system-scope flags for cross-GPU barriers on symmetric signal pads,
// and NVLS multimem load-reduce / store / reduce on a multicast address.  Used by the fused compute+collective kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace pdb {
namespace nvl {

constexpr int kMaxRanks = 8;

struct Peers {              // passed by value to kernels
  uint64_t ptr[kMaxRanks];  // base address of every rank's symmetric segment as mapped in THIS process
  uint64_t mc;              // multicast mapping of the same segment (0 when NVLS is unavailable)
  uint32_t* progress;       // optional mapped host word: last barrier epoch this rank passed (watchdog trace)
  int nranks;
  int rank;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }

// All-ranks barrier on a pad of kMaxRanks uint32 flags living at `pad_off` bytes inside every rank's segment.
// Thread t < nranks publishes `epoch` into peer t's slot [my rank], then waits until its own slot [t] reaches it.
// Epochs only grow, so the pad never needs a reset.  Call from ONE block; `spin_limit` bounds the wait (0 = forever)
// and traps with a diagnostic so a lost peer becomes an error instead of a hang.
__device__ __forceinline__ void barrier_all(const Peers& P, uint64_t pad_off, uint32_t epoch, unsigned long long spin_limit) {
  const int t = threadIdx.x;
  __syncthreads();
  if (t < P.nranks) {
    fence_sys();
    uint32_t* remote = reinterpret_cast<uint32_t*>(P.ptr[t] + pad_off) + P.rank;
    st_release_sys(remote, epoch);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(P.ptr[P.rank] + pad_off) + t;
    unsigned long long spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (spin_limit && ++spins > spin_limit) {
        printf("[pdb200 nvl] rank %d: barrier epoch %u timed out waiting for rank %d (saw %u)\n", P.rank, epoch, t,
               ld_acquire_sys(mine));
        __trap();
      }
    }
    fence_sys();
  }
  __syncthreads();
  if (t == 0 && P.progress) *reinterpret_cast<volatile uint32_t*>(P.progress) = epoch;
}

// ---- barriers embedded in multi-CTA data kernels ---------------------------------------------------------------
// Same pad protocol as barrier_all, callable from EVERY CTA of a grid so that a collective needs one launch instead
// of barrier + data + barrier.  pad layout: [0, 32) the kMaxRanks epoch flags, [64, 68) a CTA-arrival counter that
// is zero between launches.  All launches that share a pad must be ordered on one stream.
//
// grid_barrier_begin: CTA 0 announces "this rank reached epoch" to every peer; every CTA then waits until all peers
// have announced it (each CTA polls its own GPU's pad, which is local memory).
constexpr unsigned long long kNoBarrier = ~0ull;     // spin_limit value that turns the grid barriers into no-ops
__device__ __forceinline__ void grid_barrier_begin(const Peers& P, uint64_t pad_off, uint32_t epoch, unsigned long long spin_limit) {
  if (spin_limit == kNoBarrier) return;              // single-GPU emulation: the launches of the ranks are serialised
  const int t = threadIdx.x;
  if (t < P.nranks) {
    if (blockIdx.x == 0) {
      fence_sys();
      st_release_sys(reinterpret_cast<uint32_t*>(P.ptr[t] + pad_off) + P.rank, epoch);
    }
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(P.ptr[P.rank] + pad_off) + t;
    unsigned long long spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (spin_limit && ++spins > spin_limit) {
        printf("[pdb200 nvl] rank %d cta %d: begin-barrier epoch %u timed out waiting for rank %d (saw %u)\n", P.rank,
               (int)blockIdx.x, epoch, t, ld_acquire_sys(mine));
        __trap();
      }
    }
    fence_sys();
  }
  __syncthreads();
}
// grid_barrier_end: every CTA calls it after its last store.  The last CTA of this rank to arrive announces "all my
// stores are issued and fenced" (epoch) to every peer and waits for the same announcement from all of them, so when
// the kernel completes, every peer's stores into this GPU have landed and every peer has finished reading from it.
__device__ __forceinline__ void grid_barrier_end(const Peers& P, uint64_t pad_off, uint32_t epoch, unsigned long long spin_limit) {
  if (spin_limit == kNoBarrier) return;
  __shared__ int s_last;
  __threadfence_system();                       // this thread's stores (peer / multimem) before the arrival below
  __syncthreads();
  unsigned int* counter = reinterpret_cast<unsigned int*>(P.ptr[P.rank] + pad_off + 64);
  if (threadIdx.x == 0) {
    __threadfence_system();                     // cumulative: covers the whole CTA's stores (ordered by the barrier)
    const unsigned int prev = atomicAdd(counter, 1u);
    s_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  const int t = threadIdx.x;
  if (t == 0) *counter = 0;                     // ready for the next launch on this pad (stream-ordered)
  if (t < P.nranks) {
    fence_sys();                                // acquire the other CTAs' arrivals, then release to the peers
    st_release_sys(reinterpret_cast<uint32_t*>(P.ptr[t] + pad_off) + P.rank, epoch);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(P.ptr[P.rank] + pad_off) + t;
    unsigned long long spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (spin_limit && ++spins > spin_limit) {
        printf("[pdb200 nvl] rank %d: end-barrier epoch %u timed out waiting for rank %d (saw %u)\n", P.rank, epoch, t,
               ld_acquire_sys(mine));
        __trap();
      }
    }
    fence_sys();
  }
  __syncthreads();
}

// ---- NVLS multimem ------------------------------------------------------------------------------------------
// 16-byte in-switch reductions over all ranks bound to the multicast object.
__device__ __forceinline__ void mm_ld_reduce_f32x4(const void* mc_addr, float (&v)[4]) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3])
               : "l"(mc_addr)
               : "memory");
}
__device__ __forceinline__ void mm_ld_reduce_bf16x8(const void* mc_addr, uint32_t (&v)[4]) {   // fp32 accumulation
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
               : "l"(mc_addr)
               : "memory");
}
__device__ __forceinline__ void mm_st_b32x4(void* mc_addr, const uint32_t (&v)[4]) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(__uint_as_float(v[0])),
               "f"(__uint_as_float(v[1])), "f"(__uint_as_float(v[2])), "f"(__uint_as_float(v[3]))
               : "memory");
}
__device__ __forceinline__ void mm_st_f32x4(void* mc_addr, const float (&v)[4]) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v[0]), "f"(v[1]),
               "f"(v[2]), "f"(v[3])
               : "memory");
}

}  // namespace nvl
}  // namespace pdb
