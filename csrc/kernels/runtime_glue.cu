// C-ABI glue shared by every kernel file: last-error string, SM count cache, version probe.
#include <stdarg.h>
#include <stdio.h>

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
}  // namespace pdb

PDB_EXPORT const char* pdb200_last_error() { return pdb::g_err; }
PDB_EXPORT int pdb200_abi_version() { return 1; }
PDB_EXPORT int pdb200_sm_count() { return pdb::sm_count(); }
