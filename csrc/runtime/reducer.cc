// Bookkeeping core of the data-parallel gradient reducer (N6).
//
// Role of the reference's EagerReducer state (paddle/fluid/distributed/collective/reducer.cc: MarkVarReady :834-930,
// MarkGroupReady :932-990, next_group_ / pending counts, local_used_vars_ :1129-1180, FinalizeBackward :1068-1127) with
// the tensor work and the collectives left to the caller: parameters are marked ready from autograd hooks, a bucket
// whose last parameter arrives becomes ready, and buckets are handed out for reduction strictly in bucket order (every
// rank must issue its collectives in the same order, whatever order its gradients arrive in).  finalize() closes the
// iteration: buckets that never filled hold unused parameters - an error unless the reducer was built to tolerate them,
// in which case they are released in order behind the filled ones and the used-parameter bitmap is available for the
// cross-rank MAX reduction.
#include <stdint.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "rt_common.h"

using pdbrt::set_error;

namespace {

struct Bucket {
  int size = 0;        // parameters in the bucket
  int pending = 0;
  bool ready = false;
  bool sparse = false; // a single sparse-gradient parameter: reduced on its own as soon as it is marked
};

struct Reducer {
  std::vector<int> bucket_of;     // parameter -> bucket
  std::vector<Bucket> buckets;
  std::vector<uint8_t> marked;    // parameter received a gradient in this iteration
  int next_bucket = 0;
  bool armed = false;
  std::mutex mu;                  // hooks may fire from autograd worker threads

  int drain(int32_t* out, int cap) {
    int n = 0;
    while (next_bucket < (int)buckets.size() && buckets[next_bucket].ready) {
      if (n < cap) out[n] = next_bucket;
      ++n;
      ++next_bucket;
    }
    return n;
  }
};

}  // namespace

PDB_RT_EXPORT void* pdb200_reducer_create(const int32_t* bucket_of_param, int n_params, const uint8_t* bucket_sparse,
                                          int n_buckets) {
  if (n_params < 0 || n_buckets < 0) return nullptr;
  auto* r = new Reducer();
  r->bucket_of.assign(bucket_of_param, bucket_of_param + n_params);
  r->buckets.resize(n_buckets);
  for (int p = 0; p < n_params; ++p) {
    const int b = bucket_of_param[p];
    if (b < 0 || b >= n_buckets) {
      set_error("reducer: parameter %d names bucket %d of %d", p, b, n_buckets);
      delete r;
      return nullptr;
    }
    r->buckets[b].size++;
  }
  for (int b = 0; b < n_buckets; ++b) r->buckets[b].sparse = bucket_sparse && bucket_sparse[b] != 0;
  r->marked.assign(n_params, 0);
  return r;
}

PDB_RT_EXPORT void pdb200_reducer_destroy(void* h) { delete (Reducer*)h; }

// Start of an iteration (before backward): every bucket waits for all of its parameters again.
PDB_RT_EXPORT void pdb200_reducer_prepare(void* h) {
  auto* r = (Reducer*)h;
  std::lock_guard<std::mutex> g(r->mu);
  for (auto& b : r->buckets) {
    b.pending = b.size;
    b.ready = false;
  }
  std::fill(r->marked.begin(), r->marked.end(), 0);
  r->next_bucket = 0;
  r->armed = true;
}

// A parameter's gradient is final.  Returns -1 if the reducer is not armed or the parameter was already marked in
// this iteration, else the number of buckets that may be reduced NOW, in order, their ids in out_launch.
PDB_RT_EXPORT int pdb200_reducer_mark_ready(void* h, int param, int32_t* out_launch, int cap) {
  auto* r = (Reducer*)h;
  std::lock_guard<std::mutex> g(r->mu);
  if (!r->armed || param < 0 || param >= (int)r->marked.size() || r->marked[param]) return -1;
  r->marked[param] = 1;
  Bucket& b = r->buckets[r->bucket_of[param]];
  if (b.sparse) {
    b.pending = 0;
    b.ready = true;
  } else if (--b.pending == 0) {
    b.ready = true;
  }
  return r->drain(out_launch, cap);
}

// End of backward.  Parameters that were never marked are written to out_missing (up to cap_missing; the return value
// is their full count).  With tolerate_unused != 0 the buckets they block are released and every remaining bucket id
// is appended to out_launch (*n_launch).  Without it, a non-zero return leaves the state untouched (caller raises).
PDB_RT_EXPORT int pdb200_reducer_finalize(void* h, int tolerate_unused, int32_t* out_missing, int cap_missing,
                                          int32_t* out_launch, int cap_launch, int32_t* n_launch) {
  auto* r = (Reducer*)h;
  std::lock_guard<std::mutex> g(r->mu);
  *n_launch = 0;
  if (!r->armed) return 0;
  int missing = 0;
  for (int p = 0; p < (int)r->marked.size(); ++p)
    if (!r->marked[p]) {
      if (missing < cap_missing) out_missing[missing] = p;
      ++missing;
    }
  if (missing && !tolerate_unused) {
    r->armed = false;
    return missing;
  }
  for (auto& b : r->buckets) b.ready = true;
  *n_launch = r->drain(out_launch, cap_launch);
  r->armed = false;
  return missing;
}

// 1 for every parameter that received a gradient in the iteration just finalized (input of the cross-rank MAX).
PDB_RT_EXPORT void pdb200_reducer_used_bitmap(void* h, int32_t* out) {
  auto* r = (Reducer*)h;
  std::lock_guard<std::mutex> g(r->mu);
  for (size_t p = 0; p < r->marked.size(); ++p) out[p] = r->marked[p];
}
