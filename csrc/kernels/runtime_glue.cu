// C-ABI glue shared by every kernel file: last-error string, SM count cache, version probe.
#include <stdarg.h>
#include <stdio.h>

#include <mutex>
#include <vector>

#include "common.cuh"

namespace pdb {
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// ---- local-memory guard ------------------------------------------------------------------------------------------
// Address ranges of this process that map ANOTHER GPU's memory (peer segments of a symmetric arena, NVLS multicast
// windows).  The symmetric-memory layer registers them; every launcher that builds a TMA tensor map or hands a buffer
// to a tensor-core kernel refuses operands that touch one of them.  Rule of this library: TMA tensor maps cover this
// GPU's own memory only; a peer's data is first landed in local memory and loaded from there.
namespace {
struct Range { uintptr_t lo, hi; };
std::mutex g_foreign_mu;
std::vector<Range> g_foreign;
}  // namespace
int require_local(const void* p, size_t bytes, const char* what) {
  if (p == nullptr) return 0;
  const uintptr_t lo = (uintptr_t)p, hi = lo + (bytes ? bytes : 1);
  std::lock_guard<std::mutex> g(g_foreign_mu);
  for (const Range& r : g_foreign)
    if (lo < r.hi && r.lo < hi) {
      set_error("%s: operand %p (+%zu B) lies in a mapping of another GPU's memory [%p, %p); tensor-core / TMA operands "
                "must be local - copy or gather the data into this GPU's memory first", what, p, bytes, (void*)r.lo, (void*)r.hi);
      return -1;
    }
  return 0;
}
int require_local_device(const void* p, const char* what) {
  if (p == nullptr) return 0;
  cudaPointerAttributes at{};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    set_error("%s: no current CUDA device", what);
    return -1;
  }
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    set_error("%s: operand %p is not a CUDA allocation", what, p);
    return -1;
  }
  if (at.type != cudaMemoryTypeDevice || at.device != dev) {
    set_error("%s: operand %p is not device memory of the current GPU %d (type %d, device %d)", what, p, dev, (int)at.type, at.device);
    return -1;
  }
  return 0;
}
}  // namespace pdb

PDB_EXPORT int pdb200_register_foreign_range(uint64_t base, uint64_t bytes) {
  PDB_REQUIRE(base != 0 && bytes != 0, "register_foreign_range: empty range");
  std::lock_guard<std::mutex> g(pdb::g_foreign_mu);
  pdb::g_foreign.push_back({(uintptr_t)base, (uintptr_t)(base + bytes)});
  return 0;
}
PDB_EXPORT int pdb200_unregister_foreign_range(uint64_t base) {
  std::lock_guard<std::mutex> g(pdb::g_foreign_mu);
  for (size_t i = 0; i < pdb::g_foreign.size(); ++i)
    if (pdb::g_foreign[i].lo == (uintptr_t)base) {
      pdb::g_foreign.erase(pdb::g_foreign.begin() + i);
      return 0;
    }
  return 0;
}
PDB_EXPORT int pdb200_check_local(const void* p, int64_t bytes) { return pdb::require_local(p, (size_t)bytes, "check_local"); }

// Mapped host word the fused kernels' barriers write their last epoch to (watchdog trace); 0 = off.
namespace pdb { uint32_t* g_nvl_progress = nullptr; }
PDB_EXPORT int pdb200_nvl_set_progress(uint64_t host_mapped_ptr) {
  pdb::g_nvl_progress = reinterpret_cast<uint32_t*>((uintptr_t)host_mapped_ptr);
  return 0;
}

PDB_EXPORT const char* pdb200_last_error() { return pdb::g_err; }
PDB_EXPORT int pdb200_abi_version() { return 2; }
PDB_EXPORT int pdb200_sm_count() { return pdb::sm_count(); }
