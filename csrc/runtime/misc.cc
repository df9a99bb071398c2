// Error plumbing, flag mirror and gradient-bucket assignment of the host runtime.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "rt_common.h"

namespace pdbrt {
static thread_local char g_err[2048] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

static std::mutex g_flag_mu;
static std::map<std::string, std::string> g_flags;
}  // namespace pdbrt

PDB_RT_EXPORT const char* pdb200_rt_last_error() { return pdbrt::last_error(); }

PDB_RT_EXPORT int pdb200_flag_set(const char* name, const char* value) {
  std::lock_guard<std::mutex> g(pdbrt::g_flag_mu);
  pdbrt::g_flags[name] = value;
  return 0;
}
PDB_RT_EXPORT int64_t pdb200_flag_get(const char* name, char* out, int64_t capacity) {
  std::lock_guard<std::mutex> g(pdbrt::g_flag_mu);
  auto it = pdbrt::g_flags.find(name);
  if (it == pdbrt::g_flags.end()) return -1;
  const int64_t n = (int64_t)it->second.size();
  if (n + 1 > capacity) return -(n + 1);
  memcpy(out, it->second.c_str(), n + 1);
  return n;
}

// Per-dtype greedy bucketing with a size limit that advances after the first bucket of each dtype
// ([last_comm_buffer_size, comm_buffer_size]): a bucket closes as soon as its byte size reaches the current limit;
// sparse gradients get a bucket of their own; leftovers close at the end; buckets are ordered by first tensor.
PDB_RT_EXPORT int pdb200_assign_group_by_size(const int64_t* sizes, const int* dtype_ids, const int* is_sparse, int n,
                                              const int64_t* limits, int n_limits, int* out_group) {
  if (n < 0 || n_limits <= 0) {
    pdbrt::set_error("assign_group_by_size: bad arguments n=%d n_limits=%d", n, n_limits);
    return -1;
  }
  struct Open {
    std::vector<int> idx;
    int64_t bytes = 0;
    int limit_index = 0;
  };
  std::map<int, Open> open;
  std::vector<std::vector<int>> res;
  for (int i = 0; i < n; ++i) {
    if (is_sparse && is_sparse[i]) {
      res.push_back({i});
      continue;
    }
    Open& o = open[dtype_ids[i]];
    o.idx.push_back(i);
    o.bytes += sizes[i];
    if (o.bytes >= limits[o.limit_index]) {
      res.push_back(std::move(o.idx));
      o.idx.clear();
      o.bytes = 0;
      o.limit_index = std::min(o.limit_index + 1, n_limits - 1);
    }
  }
  for (auto& kv : open)
    if (!kv.second.idx.empty()) res.push_back(std::move(kv.second.idx));
  std::sort(res.begin(), res.end(), [](const std::vector<int>& a, const std::vector<int>& b) { return a.front() < b.front(); });
  for (size_t g = 0; g < res.size(); ++g)
    for (int i : res[g]) out_group[i] = (int)g;
  return (int)res.size();
}
