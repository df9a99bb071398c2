// Native process group over NCCL (N1 ProcessGroup + Task, N2 ProcessGroupNCCL, N3 NCCLCommContext / manager; reference:
// paddle/fluid/distributed/collective/process_group_nccl.cc:160-1016, paddle/phi/core/distributed/nccl_comm_context.cc,
// comm_context_manager.cc:61-110).
//
// One CommContext per (group id) holds the ncclComm_t, the device and a dedicated non-blocking communication stream.
// The unique id is published by the group's first rank through the TCP store of this library (store.cc) and fetched
// by the others, exactly the reference's bootstrap.  Every collective follows the reference's stream contract:
//   use_calc_stream  -> the collective is enqueued on the caller's (compute) stream; nothing to wait for;
//   otherwise        -> the comm stream first waits for an event recorded on the compute stream (the inputs are
//                       ready), runs the collective, and records the task's end event; Task::wait makes the compute
//                       stream wait for that event on the DEVICE - the host never blocks; Task::is_completed is an
//                       event query.
// Each task is also registered with the watchdog of this library with CUDA events on the stream it runs on, so hangs
// are classified from the device, and the watchdog's sweep asks NCCL for asynchronous errors (ncclCommGetAsyncError) -
// a communicator in error is aborted (ncclCommAbort) so blocked kernels return instead of hanging the process.
// NCCL is loaded with dlopen (the copy torch already mapped when there is one): the library carries no link-time
// dependency on it and loads on machines without NCCL.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "rt_common.h"

extern "C" {
int pdb200_store_set(void* store, const char* key, const void* value, int64_t nbytes);
int64_t pdb200_store_get(void* store, const char* key, void* out, int64_t capacity, int timeout_ms);
int pdb200_wd_task_begin_ev(int group, int64_t seq, const char* op, int64_t numel, int sync_op, void* stream);
int pdb200_wd_task_end_ev(int group, int64_t seq, void* stream);
}

namespace pdbrt {

struct NcclApi {
  void* lib = nullptr;
#define PDB_NCCL(name) decltype(&::name) name = nullptr;
  PDB_NCCL(ncclGetUniqueId)
  PDB_NCCL(ncclCommInitRank)
  PDB_NCCL(ncclCommDestroy)
  PDB_NCCL(ncclCommAbort)
  PDB_NCCL(ncclCommGetAsyncError)
  PDB_NCCL(ncclGetErrorString)
  PDB_NCCL(ncclAllReduce)
  PDB_NCCL(ncclBroadcast)
  PDB_NCCL(ncclReduce)
  PDB_NCCL(ncclAllGather)
  PDB_NCCL(ncclReduceScatter)
  PDB_NCCL(ncclSend)
  PDB_NCCL(ncclRecv)
  PDB_NCCL(ncclGroupStart)
  PDB_NCCL(ncclGroupEnd)
  PDB_NCCL(ncclGetVersion)
#undef PDB_NCCL
};

static NcclApi* nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* n : names) {
      h = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);      // the copy torch has already mapped, if any
      if (h) break;
    }
    if (!h) {
      const char* env = getenv("PDB200_NCCL_LIB");
      if (env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
    }
    for (const char* n : names) {
      if (h) break;
      h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    }
    if (!h) return;
    api.lib = h;
#define PDB_NCCL(name) api.name = (decltype(api.name))dlsym(h, #name);
    PDB_NCCL(ncclGetUniqueId)
    PDB_NCCL(ncclCommInitRank)
    PDB_NCCL(ncclCommDestroy)
    PDB_NCCL(ncclCommAbort)
    PDB_NCCL(ncclCommGetAsyncError)
    PDB_NCCL(ncclGetErrorString)
    PDB_NCCL(ncclAllReduce)
    PDB_NCCL(ncclBroadcast)
    PDB_NCCL(ncclReduce)
    PDB_NCCL(ncclAllGather)
    PDB_NCCL(ncclReduceScatter)
    PDB_NCCL(ncclSend)
    PDB_NCCL(ncclRecv)
    PDB_NCCL(ncclGroupStart)
    PDB_NCCL(ncclGroupEnd)
    PDB_NCCL(ncclGetVersion)
#undef PDB_NCCL
  });
  return api.ncclAllReduce ? &api : nullptr;
}

#define PDB_NCCL_OK(expr)                                                                   \
  do {                                                                                      \
    ncclResult_t _r = (expr);                                                               \
    if (_r != ncclSuccess) {                                                                \
      set_error("%s failed: %s", #expr, nccl()->ncclGetErrorString(_r));                    \
      return -1;                                                                            \
    }                                                                                       \
  } while (0)
#define PDB_CU_OK(expr)                                                                     \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      set_error("%s failed: %s", #expr, cudaGetErrorString(_e));                            \
      return -1;                                                                            \
    }                                                                                       \
  } while (0)

// dtype codes of the C ABI: 0 f32, 1 bf16, 2 f16, 3 f64, 4 i32, 5 i64, 6 u8, 7 i8
static bool to_nccl_dtype(int code, ncclDataType_t* out) {
  static const ncclDataType_t t[] = {ncclFloat32, ncclBfloat16, ncclFloat16, ncclFloat64, ncclInt32, ncclInt64, ncclUint8, ncclInt8};
  if (code < 0 || code > 7) return false;
  *out = t[code];
  return true;
}
// reduce codes: 0 sum, 1 max, 2 min, 3 prod, 4 avg
static bool to_nccl_op(int code, ncclRedOp_t* out) {
  static const ncclRedOp_t t[] = {ncclSum, ncclMax, ncclMin, ncclProd, ncclAvg};
  if (code < 0 || code > 4) return false;
  *out = t[code];
  return true;
}

struct Task {                      // N1 ProcessGroup::Task
  cudaEvent_t done = nullptr;      // null: ran on the calc stream (nothing to wait for)
  int device = 0;
  int group = 0;
  int64_t seq = 0;
};

class CommContext {                // N3 NCCLCommContext
 public:
  int gid = 0, rank = 0, nranks = 1, device = 0;
  ncclComm_t comm = nullptr;
  cudaStream_t stream = nullptr;   // the group's communication stream
  cudaEvent_t ready = nullptr;     // "inputs ready on the calc stream", re-recorded per call
  std::atomic<int64_t> seq{0};
  std::atomic<bool> aborted{false};
  std::vector<cudaEvent_t> pool;
  std::mutex mu;

  cudaEvent_t take_event() {
    std::lock_guard<std::mutex> g(mu);
    if (!pool.empty()) {
      cudaEvent_t e = pool.back();
      pool.pop_back();
      return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    return e;
  }
  void give_event(cudaEvent_t e) {
    std::lock_guard<std::mutex> g(mu);
    pool.push_back(e);
  }
};

class CommContextManager {        // N3 CommContextManager: group id -> context
 public:
  static CommContextManager& get() {
    static CommContextManager m;
    return m;
  }
  CommContext* find(int gid) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = ctx_.find(gid);
    return it == ctx_.end() ? nullptr : it->second;
  }
  void put(int gid, CommContext* c) {
    std::lock_guard<std::mutex> g(mu_);
    ctx_[gid] = c;
  }
  void erase(int gid) {
    std::lock_guard<std::mutex> g(mu_);
    ctx_.erase(gid);
  }
  // ncclCommGetAsyncError over every live communicator; a communicator in error is aborted.  Returns how many.
  int check_async_errors() {
    NcclApi* n = nccl();
    if (!n || !n->ncclCommGetAsyncError) return 0;
    std::lock_guard<std::mutex> g(mu_);
    int bad = 0;
    for (auto& kv : ctx_) {
      CommContext* c = kv.second;
      if (!c->comm || c->aborted.load()) continue;
      ncclResult_t st = ncclSuccess;
      if (n->ncclCommGetAsyncError(c->comm, &st) == ncclSuccess && st != ncclSuccess && st != ncclInProgress) {
        fprintf(stderr, "[native pg] rank %d group %d: NCCL asynchronous error: %s - aborting the communicator\n", c->rank,
                c->gid, n->ncclGetErrorString(st));
        fflush(stderr);
        c->aborted.store(true);
        n->ncclCommAbort(c->comm);
        ++bad;
      }
    }
    return bad;
  }

 private:
  std::mutex mu_;
  std::map<int, CommContext*> ctx_;
};

}  // namespace pdbrt

using namespace pdbrt;

PDB_RT_EXPORT int pdb200_pg_available() { return nccl() ? 1 : 0; }
PDB_RT_EXPORT int pdb200_pg_nccl_version() {
  int v = 0;
  if (nccl() && nccl()->ncclGetVersion) nccl()->ncclGetVersion(&v);
  return v;
}

// Create the communicator of group `gid` (ranks 0..nranks-1 of the group; `rank` is this process's index in it) on the
// current device.  `store`: a pdb200_store handle shared by the ranks; the group's first rank publishes the NCCL unique
// id under "<prefix>/nccl_id/<gid>".  Collective: every rank of the group must call it.
// Comm streams live as long as the process.  A caching allocator that was told "this block is in use on the comm
// stream" records an event on that stream when the block is freed, which can be long after the group is gone; a
// destroyed stream handle there is a use-after-free.  So a closed group hands its stream back to a per-device pool
// and the next group on that device takes it from there.
static std::mutex g_stream_mu;
static std::vector<std::pair<int, cudaStream_t>> g_stream_pool;

static cudaStream_t acquire_comm_stream(int device) {
  {
    std::lock_guard<std::mutex> g(g_stream_mu);
    for (size_t i = 0; i < g_stream_pool.size(); ++i)
      if (g_stream_pool[i].first == device) {
        cudaStream_t s = g_stream_pool[i].second;
        g_stream_pool.erase(g_stream_pool.begin() + i);
        return s;
      }
  }
  int lo = 0, hi = 0;
  cudaStream_t s = nullptr;
  cudaDeviceGetStreamPriorityRange(&lo, &hi);
  cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, hi);     // highest priority: comm kernels go first
  return s;
}

static void release_comm_stream(int device, cudaStream_t s) {
  std::lock_guard<std::mutex> g(g_stream_mu);
  g_stream_pool.emplace_back(device, s);
}

PDB_RT_EXPORT void* pdb200_pg_create(void* store, const char* prefix, int gid, int rank, int nranks, int timeout_ms) {
  NcclApi* n = nccl();
  if (!n) {
    set_error("NCCL library not found (set PDB200_NCCL_LIB)");
    return nullptr;
  }
  std::string key = std::string(prefix ? prefix : "") + "/nccl_id/" + std::to_string(gid);
  ncclUniqueId id;
  if (rank == 0) {
    if (n->ncclGetUniqueId(&id) != ncclSuccess) {
      set_error("ncclGetUniqueId failed");
      return nullptr;
    }
    if (pdb200_store_set(store, key.c_str(), &id, sizeof(id)) != 0) return nullptr;
  } else {
    if (pdb200_store_get(store, key.c_str(), &id, sizeof(id), timeout_ms > 0 ? timeout_ms : 60000) != (int64_t)sizeof(id)) {
      set_error("native pg: could not fetch the NCCL unique id of group %d from the store", gid);
      return nullptr;
    }
  }
  auto* c = new CommContext();
  c->gid = gid;
  c->rank = rank;
  c->nranks = nranks;
  if (cudaGetDevice(&c->device) != cudaSuccess) {
    set_error("native pg: no current CUDA device");
    delete c;
    return nullptr;
  }
  ncclResult_t r = n->ncclCommInitRank(&c->comm, nranks, id, rank);
  if (r != ncclSuccess) {
    set_error("ncclCommInitRank failed: %s", n->ncclGetErrorString(r));
    delete c;
    return nullptr;
  }
  c->stream = acquire_comm_stream(c->device);
  cudaEventCreateWithFlags(&c->ready, cudaEventDisableTiming);
  CommContextManager::get().put(gid, c);
  return c;
}

PDB_RT_EXPORT int pdb200_pg_destroy(void* pg) {
  auto* c = (CommContext*)pg;
  if (!c) return 0;
  CommContextManager::get().erase(c->gid);
  NcclApi* n = nccl();
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->comm && n) {
    if (c->aborted.load()) {
      // already aborted
    } else {
      n->ncclCommDestroy(c->comm);
    }
  }
  for (cudaEvent_t e : c->pool) cudaEventDestroy(e);
  if (c->ready) cudaEventDestroy(c->ready);
  if (c->stream) release_comm_stream(c->device, c->stream);
  delete c;
  return 0;
}

PDB_RT_EXPORT uint64_t pdb200_pg_comm_stream(void* pg) { return (uint64_t)(uintptr_t)((CommContext*)pg)->stream; }
PDB_RT_EXPORT int pdb200_pg_abort(void* pg) {
  auto* c = (CommContext*)pg;
  if (!c || !c->comm || c->aborted.exchange(true)) return 0;
  return nccl()->ncclCommAbort(c->comm) == ncclSuccess ? 0 : -1;
}
PDB_RT_EXPORT int pdb200_pg_check_async_errors() { return CommContextManager::get().check_async_errors(); }

namespace {
// Pick the stream (calc or comm), order it after the calc stream when needed, register the task with the watchdog.
struct Launch {
  CommContext* c;
  cudaStream_t st;
  Task* task = nullptr;
  int64_t seq;
  bool calc;
  int begin(void* pg, void* calc_stream, int use_calc_stream, const char* op, int64_t numel, int sync_op) {
    c = (CommContext*)pg;
    if (c->aborted.load()) {
      set_error("native pg: communicator of group %d was aborted", c->gid);
      return -1;
    }
    calc = use_calc_stream != 0;
    st = calc ? (cudaStream_t)calc_stream : c->stream;
    seq = ++c->seq;
    if (!calc) {
      PDB_CU_OK(cudaEventRecord(c->ready, (cudaStream_t)calc_stream));
      PDB_CU_OK(cudaStreamWaitEvent(c->stream, c->ready, 0));
    }
    pdb200_wd_task_begin_ev(c->gid, seq, op, numel, sync_op, st);
    return 0;
  }
  void* end() {
    pdb200_wd_task_end_ev(c->gid, seq, st);
    auto* t = new Task();
    t->device = c->device;
    t->group = c->gid;
    t->seq = seq;
    if (!calc) {
      t->done = c->take_event();
      cudaEventRecord(t->done, c->stream);
    }
    return t;
  }
};
}  // namespace

#define PDB_PG_PROLOGUE(opname, numel)                                                        \
  NcclApi* n = nccl();                                                                        \
  if (!n) { set_error("NCCL unavailable"); return nullptr; }                                  \
  ncclDataType_t dt;                                                                          \
  if (!to_nccl_dtype(dtype, &dt)) { set_error("native pg: unsupported dtype code %d", dtype); return nullptr; } \
  Launch L;                                                                                   \
  if (L.begin(pg, calc_stream, use_calc_stream, opname, numel, sync_op)) return nullptr;
#define PDB_PG_CALL(expr)                                                                     \
  do {                                                                                        \
    ncclResult_t _r = (expr);                                                                 \
    if (_r != ncclSuccess) { set_error("%s failed: %s", #expr, n->ncclGetErrorString(_r)); return nullptr; } \
  } while (0)

// Every collective returns a task handle (pdb200_pg_task_*), or null on failure.
PDB_RT_EXPORT void* pdb200_pg_all_reduce(void* pg, const void* in, void* out, int64_t numel, int dtype, int op, void* calc_stream,
                                         int use_calc_stream, int sync_op) {
  PDB_PG_PROLOGUE("all_reduce", numel)
  ncclRedOp_t rop;
  if (!to_nccl_op(op, &rop)) { set_error("native pg: unsupported reduce op %d", op); return nullptr; }
  PDB_PG_CALL(n->ncclAllReduce(in, out, (size_t)numel, dt, rop, L.c->comm, L.st));
  return L.end();
}
PDB_RT_EXPORT void* pdb200_pg_broadcast(void* pg, const void* in, void* out, int64_t numel, int dtype, int root, void* calc_stream,
                                        int use_calc_stream, int sync_op) {
  PDB_PG_PROLOGUE("broadcast", numel)
  PDB_PG_CALL(n->ncclBroadcast(in, out, (size_t)numel, dt, root, L.c->comm, L.st));
  return L.end();
}
PDB_RT_EXPORT void* pdb200_pg_reduce(void* pg, const void* in, void* out, int64_t numel, int dtype, int op, int root, void* calc_stream,
                                     int use_calc_stream, int sync_op) {
  PDB_PG_PROLOGUE("reduce", numel)
  ncclRedOp_t rop;
  if (!to_nccl_op(op, &rop)) { set_error("native pg: unsupported reduce op %d", op); return nullptr; }
  PDB_PG_CALL(n->ncclReduce(in, out, (size_t)numel, dt, rop, root, L.c->comm, L.st));
  return L.end();
}
PDB_RT_EXPORT void* pdb200_pg_all_gather(void* pg, const void* in, void* out, int64_t numel_per_rank, int dtype, void* calc_stream,
                                         int use_calc_stream, int sync_op) {
  PDB_PG_PROLOGUE("all_gather", numel_per_rank)
  PDB_PG_CALL(n->ncclAllGather(in, out, (size_t)numel_per_rank, dt, L.c->comm, L.st));
  return L.end();
}
PDB_RT_EXPORT void* pdb200_pg_reduce_scatter(void* pg, const void* in, void* out, int64_t numel_per_rank, int dtype, int op,
                                             void* calc_stream, int use_calc_stream, int sync_op) {
  PDB_PG_PROLOGUE("reduce_scatter", numel_per_rank)
  ncclRedOp_t rop;
  if (!to_nccl_op(op, &rop)) { set_error("native pg: unsupported reduce op %d", op); return nullptr; }
  PDB_PG_CALL(n->ncclReduceScatter(in, out, (size_t)numel_per_rank, dt, rop, L.c->comm, L.st));
  return L.end();
}
// all-to-all of equal splits: numel_per_rank elements go to / come from every rank (grouped send/recv, the reference's
// AllToAll, process_group_nccl.cc:700-760)
PDB_RT_EXPORT void* pdb200_pg_all_to_all(void* pg, const void* in, void* out, int64_t numel_per_rank, int dtype, void* calc_stream,
                                         int use_calc_stream, int sync_op) {
  PDB_PG_PROLOGUE("all_to_all", numel_per_rank)
  const size_t esz = (dt == ncclFloat64 || dt == ncclInt64) ? 8 : (dt == ncclFloat32 || dt == ncclInt32) ? 4 : (dt == ncclUint8 || dt == ncclInt8) ? 1 : 2;
  PDB_PG_CALL(n->ncclGroupStart());
  for (int p = 0; p < L.c->nranks; ++p) {
    n->ncclSend((const char*)in + (size_t)p * numel_per_rank * esz, (size_t)numel_per_rank, dt, p, L.c->comm, L.st);
    n->ncclRecv((char*)out + (size_t)p * numel_per_rank * esz, (size_t)numel_per_rank, dt, p, L.c->comm, L.st);
  }
  PDB_PG_CALL(n->ncclGroupEnd());
  return L.end();
}
PDB_RT_EXPORT void* pdb200_pg_send(void* pg, const void* in, int64_t numel, int dtype, int peer, void* calc_stream, int use_calc_stream,
                                   int sync_op) {
  PDB_PG_PROLOGUE("send", numel)
  PDB_PG_CALL(n->ncclSend(in, (size_t)numel, dt, peer, L.c->comm, L.st));
  return L.end();
}
PDB_RT_EXPORT void* pdb200_pg_recv(void* pg, void* out, int64_t numel, int dtype, int peer, void* calc_stream, int use_calc_stream,
                                   int sync_op) {
  PDB_PG_PROLOGUE("recv", numel)
  PDB_PG_CALL(n->ncclRecv(out, (size_t)numel, dt, peer, L.c->comm, L.st));
  return L.end();
}

// ---- Task (N1): wait = the given stream waits for the collective on the device; the host never blocks
PDB_RT_EXPORT int pdb200_pg_task_wait(void* task, void* calc_stream) {
  auto* t = (Task*)task;
  if (t && t->done) PDB_CU_OK(cudaStreamWaitEvent((cudaStream_t)calc_stream, t->done, 0));
  return 0;
}
PDB_RT_EXPORT int pdb200_pg_task_is_completed(void* task) {
  auto* t = (Task*)task;
  if (!t || !t->done) return 1;
  const cudaError_t q = cudaEventQuery(t->done);
  if (q == cudaSuccess) return 1;
  cudaGetLastError();
  return 0;
}
PDB_RT_EXPORT int pdb200_pg_task_synchronize(void* task) {            // host-blocking variant (Task::Synchronize)
  auto* t = (Task*)task;
  if (t && t->done) PDB_CU_OK(cudaEventSynchronize(t->done));
  return 0;
}
PDB_RT_EXPORT int pdb200_pg_task_free(void* pg, void* task) {
  auto* t = (Task*)task;
  if (!t) return 0;
  if (t->done) {
    if (pg) ((CommContext*)pg)->give_event(t->done);
    else cudaEventDestroy(t->done);
  }
  delete t;
  return 0;
}
