// Offset heap for the symmetric-memory arena: best-fit allocation with immediate coalescing on free.
//
// The arena is one VA range per rank, mapped at the same size everywhere; what the ranks must agree on is the OFFSET
// of every buffer.  The heap is therefore purely a function of the call sequence (no addresses, no time, no thread
// ids): ranks that issue the same alloc/free calls in the same order get the same offsets.  Role of the reference's
// auto-growth best-fit allocator (paddle/phi/core/memory/allocation/auto_growth_best_fit_allocator.cc) for this one
// pool; growth is not possible here (the peer mappings are fixed at creation), so exhaustion is reported to the
// caller, which evicts cached staging buffers and retries.
#include <stdint.h>

#include <map>
#include <mutex>

#include "rt_common.h"

namespace {

class Heap {
 public:
  Heap(int64_t capacity, int64_t reserved) : capacity_(capacity) {
    if (capacity > reserved) free_[reserved] = capacity - reserved;
  }

  int64_t alloc(int64_t nbytes, int64_t align) {
    if (nbytes <= 0) nbytes = 1;
    if (align < 1) align = 1;
    std::lock_guard<std::mutex> g(mu_);
    // best fit: smallest free block that holds the aligned request; ties go to the lowest offset (map order)
    int64_t best_off = -1, best_sz = 0, best_start = 0;
    for (auto& kv : free_) {
      int64_t start = (kv.first + align - 1) / align * align;
      int64_t need = (start - kv.first) + nbytes;
      if (need <= kv.second && (best_off < 0 || kv.second < best_sz)) {
        best_off = kv.first;
        best_sz = kv.second;
        best_start = start;
      }
    }
    if (best_off < 0) return -1;
    free_.erase(best_off);
    if (best_start > best_off) free_[best_off] = best_start - best_off;               // alignment gap stays free
    int64_t end = best_start + nbytes;
    if (end < best_off + best_sz) free_[end] = best_off + best_sz - end;
    live_[best_start] = nbytes;
    in_use_ += nbytes;
    if (in_use_ > peak_) peak_ = in_use_;
    return best_start;
  }

  int release(int64_t off) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = live_.find(off);
    if (it == live_.end()) return -1;
    int64_t sz = it->second;
    live_.erase(it);
    in_use_ -= sz;
    auto nxt = free_.lower_bound(off);
    if (nxt != free_.end() && off + sz == nxt->first) {                               // merge with the block after
      sz += nxt->second;
      nxt = free_.erase(nxt);
    }
    if (nxt != free_.begin()) {                                                       // merge with the block before
      auto prv = std::prev(nxt);
      if (prv->first + prv->second == off) {
        prv->second += sz;
        return 0;
      }
    }
    free_[off] = sz;
    return 0;
  }

  int64_t stat(int what) {
    std::lock_guard<std::mutex> g(mu_);
    switch (what) {
      case 0: {                       // free bytes
        int64_t s = 0;
        for (auto& kv : free_) s += kv.second;
        return s;
      }
      case 1: {                       // largest free block
        int64_t s = 0;
        for (auto& kv : free_) s = kv.second > s ? kv.second : s;
        return s;
      }
      case 2: return (int64_t)live_.size();
      case 3: return in_use_;
      case 4: return peak_;
      case 5: return (int64_t)free_.size();
      case 6: return capacity_;
    }
    return -1;
  }

 private:
  std::mutex mu_;
  std::map<int64_t, int64_t> free_;   // offset -> size, no two neighbours adjacent
  std::map<int64_t, int64_t> live_;   // offset -> size
  int64_t capacity_, in_use_ = 0, peak_ = 0;
};

}  // namespace

PDB_RT_EXPORT void* pdb200_heap_create(int64_t capacity, int64_t reserved) { return new Heap(capacity, reserved); }
PDB_RT_EXPORT void pdb200_heap_destroy(void* h) { delete (Heap*)h; }
PDB_RT_EXPORT int64_t pdb200_heap_alloc(void* h, int64_t nbytes, int64_t align) { return ((Heap*)h)->alloc(nbytes, align); }
PDB_RT_EXPORT int pdb200_heap_free(void* h, int64_t offset) { return ((Heap*)h)->release(offset); }
PDB_RT_EXPORT int64_t pdb200_heap_stat(void* h, int what) { return ((Heap*)h)->stat(what); }
