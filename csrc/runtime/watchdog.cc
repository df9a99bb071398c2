// Communication watchdog: per-collective records with start time, a polling thread that flags tasks older than
// the timeout and prints one trace line per offender (group, seq, op, numel, sync flag, age) followed by the last
// finished task of that group - the same information the reference's CommTaskManager logs
// (paddle/phi/core/distributed/comm_task_manager.cc:135-232, nccl_comm_task.cc:238-268).  Optional abort ends the
// process so the launcher's restart policy takes over (hung in-kernel spins have no NCCL task to abort).
//
// Device truth: a task registered with a stream carries two CUDA events (nccl_comm_task.cc:90-140 uses the same
// pair).  The start event is recorded before the collective is enqueued, the end event after; the polling thread
// queries them, so "not started" (start event pending: the stream is stuck on earlier work), "started, not finished"
// (the collective itself hangs) and "finished" are read from the GPU, and a finished task closes itself even if
// nobody ever calls wait() on it.  The in-kernel barriers of the fused NVLink kernels publish the last epoch they
// passed in a pinned host word (pdb200_wd_progress_ptr) that the trace lines print.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <map>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>

#include <cuda_runtime.h>

#include <vector>

#include "rt_common.h"

extern "C" int pdb200_pg_check_async_errors();      // process_group.cc: ncclCommGetAsyncError over the live communicators

namespace pdbrt {

struct TaskRec {
  int group;
  int64_t seq;
  std::string op;
  int64_t numel;
  int sync_op;
  std::chrono::steady_clock::time_point start;
  bool reported = false;
  // device-side state (has_ev): events recorded around the collective on its stream
  bool has_ev = false, end_recorded = false;
  int device = 0;
  cudaEvent_t ev_start = nullptr, ev_end = nullptr;
};

class Watchdog {
 public:
  static Watchdog& get() {
    static Watchdog w;
    return w;
  }
  ~Watchdog() { stop(); }     // a process that never called stop() must not die in std::terminate at exit
  int start(int rank, int timeout_ms, int poll_ms) {
    std::lock_guard<std::mutex> g(mu_);
    if (running_) return 0;
    rank_ = rank;
    timeout_ms_ = timeout_ms;
    poll_ms_ = poll_ms > 0 ? poll_ms : 10000;
    running_ = true;
    thread_ = std::thread([this] { loop(); });
    return 0;
  }
  int stop() {
    {
      std::lock_guard<std::mutex> g(mu_);
      if (!running_) return 0;
      running_ = false;
    }
    cv_.notify_all();
    if (thread_.joinable()) thread_.join();
    std::lock_guard<std::mutex> g(mu_);
    inflight_.clear();
    return 0;
  }
  void begin(int group, int64_t seq, const char* op, int64_t numel, int sync_op) {
    std::lock_guard<std::mutex> g(mu_);
    inflight_[{group, seq}] = TaskRec{group, seq, op ? op : "", numel, sync_op, std::chrono::steady_clock::now()};
  }
  void end(int group, int64_t seq) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = inflight_.find({group, seq});
    if (it != inflight_.end()) {
      release_events(it->second);
      last_done_[group] = it->second;
      inflight_.erase(it);
    }
  }
  // Register a task whose progress is read from the device: records the start event on `stream` now.
  int begin_ev(int group, int64_t seq, const char* op, int64_t numel, int sync_op, void* stream) {
    std::lock_guard<std::mutex> g(mu_);
    TaskRec t{group, seq, op ? op : "", numel, sync_op, std::chrono::steady_clock::now()};
    if (cudaGetDevice(&t.device) == cudaSuccess && take_event(&t.ev_start) && take_event(&t.ev_end) &&
        cudaEventRecord(t.ev_start, (cudaStream_t)stream) == cudaSuccess) {
      t.has_ev = true;
    } else {
      cudaGetLastError();
      release_events(t);
    }
    if (inflight_.size() > 512) sweep_locked();
    inflight_[{group, seq}] = t;
    return t.has_ev ? 1 : 0;
  }
  // The collective has been enqueued: record the end event; the polling thread (or the next sweep) retires the task
  // once the GPU has passed it.
  int end_ev(int group, int64_t seq, void* stream) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = inflight_.find({group, seq});
    if (it == inflight_.end()) return 0;
    TaskRec& t = it->second;
    if (!t.has_ev || cudaEventRecord(t.ev_end, (cudaStream_t)stream) != cudaSuccess) {
      cudaGetLastError();
      release_events(t);
      last_done_[group] = t;
      inflight_.erase(it);
      return 0;
    }
    t.end_recorded = true;
    return 1;
  }
  int sweep() {
    std::lock_guard<std::mutex> g(mu_);
    return sweep_locked();
  }
  volatile unsigned int* progress_ptr() {
    std::lock_guard<std::mutex> g(mu_);
    if (!progress_) {
      void* p = nullptr;
      if (cudaHostAlloc(&p, 64, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) {
        memset(p, 0, 64);
        progress_ = (volatile unsigned int*)p;
      } else {
        cudaGetLastError();
      }
    }
    return progress_;
  }
  int inflight() {
    std::lock_guard<std::mutex> g(mu_);
    return (int)inflight_.size();
  }
  int timeouts() const { return timeouts_.load(); }
  void set_abort(bool a) { abort_ = a; }
  std::string dump() {
    std::lock_guard<std::mutex> g(mu_);
    std::ostringstream os;
    auto now = std::chrono::steady_clock::now();
    sweep_locked();
    for (auto& kv : inflight_) {
      const TaskRec& t = kv.second;
      os << "[in-flight] rank=" << rank_ << " group=" << t.group << " seq=" << t.seq << " op=" << t.op
         << " numel=" << t.numel << " sync=" << t.sync_op << " state=" << state_of(t) << " age_ms="
         << std::chrono::duration_cast<std::chrono::milliseconds>(now - t.start).count() << "\n";
    }
    if (progress_) os << "[fused-kernel barrier] rank=" << rank_ << " last epoch passed=" << progress_[0] << "\n";
    for (auto& kv : last_done_)
      os << "[last-done] rank=" << rank_ << " group=" << kv.first << " seq=" << kv.second.seq << " op=" << kv.second.op
         << " numel=" << kv.second.numel << "\n";
    return os.str();
  }

 private:
  void loop() {
    std::unique_lock<std::mutex> lk(mu_);
    while (running_) {
      cv_.wait_for(lk, std::chrono::milliseconds(poll_ms_));
      if (!running_) break;
      auto now = std::chrono::steady_clock::now();
      bool fatal = false;
      sweep_locked();
      lk.unlock();
      pdb200_pg_check_async_errors();               // a communicator in error is aborted so blocked kernels return
      lk.lock();
      for (auto& kv : inflight_) {
        TaskRec& t = kv.second;
        auto age = std::chrono::duration_cast<std::chrono::milliseconds>(now - t.start).count();
        if (age > timeout_ms_ && !t.reported) {
          t.reported = true;
          timeouts_++;
          fprintf(stderr, "[comm watchdog] rank %d: collective timed out (%s): group=%d seq=%lld op=%s "
                          "numel=%lld sync_op=%d age=%lldms\n", rank_, state_of(t), t.group, (long long)t.seq, t.op.c_str(),
                  (long long)t.numel, t.sync_op, (long long)age);
          if (progress_)
            fprintf(stderr, "[comm watchdog] rank %d: fused-kernel barriers: last epoch passed = %u\n", rank_, progress_[0]);
          auto ld = last_done_.find(t.group);
          if (ld != last_done_.end())
            fprintf(stderr, "[comm watchdog] rank %d: last finished task of group %d: seq=%lld op=%s numel=%lld\n", rank_,
                    t.group, (long long)ld->second.seq, ld->second.op.c_str(), (long long)ld->second.numel);
          fflush(stderr);
          fatal = true;
        }
      }
      if (fatal && abort_) {
        fprintf(stderr, "[comm watchdog] rank %d: aborting the process\n", rank_);
        fflush(stderr);
        _exit(113);
      }
    }
  }

  // ---- device-event helpers (mu_ held) ----
  bool take_event(cudaEvent_t* e) {
    if (!pool_.empty()) {
      *e = pool_.back();
      pool_.pop_back();
      return true;
    }
    return cudaEventCreateWithFlags(e, cudaEventDisableTiming) == cudaSuccess;
  }
  void release_events(TaskRec& t) {
    if (t.ev_start) pool_.push_back(t.ev_start);
    if (t.ev_end) pool_.push_back(t.ev_end);
    t.ev_start = t.ev_end = nullptr;
    t.has_ev = false;
  }
  const char* state_of(const TaskRec& t) {
    if (!t.has_ev) return "host clock only";
    int cur = -1;
    cudaGetDevice(&cur);
    if (cur != t.device) cudaSetDevice(t.device);
    const bool started = cudaEventQuery(t.ev_start) == cudaSuccess;
    cudaGetLastError();
    return started ? "started, not finished" : "not started";
  }
  // retire every task whose end event the GPU has passed
  int sweep_locked() {
    int n = 0;
    for (auto it = inflight_.begin(); it != inflight_.end();) {
      TaskRec& t = it->second;
      if (t.has_ev && t.end_recorded) {
        int cur = -1;
        cudaGetDevice(&cur);
        if (cur != t.device) cudaSetDevice(t.device);
        const cudaError_t q = cudaEventQuery(t.ev_end);
        if (q == cudaSuccess) {
          release_events(t);
          last_done_[t.group] = t;
          it = inflight_.erase(it);
          ++n;
          continue;
        }
        cudaGetLastError();
      }
      ++it;
    }
    return n;
  }

  std::mutex mu_;
  std::vector<cudaEvent_t> pool_;
  volatile unsigned int* progress_ = nullptr;
  std::condition_variable cv_;
  std::thread thread_;
  bool running_ = false;
  int rank_ = 0, timeout_ms_ = 1800000, poll_ms_ = 10000;
  std::atomic<int> timeouts_{0};
  std::atomic<bool> abort_{false};
  std::map<std::pair<int, int64_t>, TaskRec> inflight_;
  std::map<int, TaskRec> last_done_;
};

}  // namespace pdbrt

using pdbrt::Watchdog;

PDB_RT_EXPORT int pdb200_wd_start(int rank, int timeout_ms, int poll_ms) { return Watchdog::get().start(rank, timeout_ms, poll_ms); }
PDB_RT_EXPORT int pdb200_wd_stop() { return Watchdog::get().stop(); }
PDB_RT_EXPORT int pdb200_wd_task_begin(int group, int64_t seq, const char* op, int64_t numel, int sync_op) {
  Watchdog::get().begin(group, seq, op, numel, sync_op);
  return 0;
}
PDB_RT_EXPORT int pdb200_wd_task_end(int group, int64_t seq) {
  Watchdog::get().end(group, seq);
  return 0;
}
PDB_RT_EXPORT int pdb200_wd_task_begin_ev(int group, int64_t seq, const char* op, int64_t numel, int sync_op, void* stream) {
  return Watchdog::get().begin_ev(group, seq, op, numel, sync_op, stream);
}
PDB_RT_EXPORT int pdb200_wd_task_end_ev(int group, int64_t seq, void* stream) { return Watchdog::get().end_ev(group, seq, stream); }
PDB_RT_EXPORT int pdb200_wd_sweep() { return Watchdog::get().sweep(); }
PDB_RT_EXPORT uint64_t pdb200_wd_progress_ptr() { return (uint64_t)(uintptr_t)Watchdog::get().progress_ptr(); }
PDB_RT_EXPORT int pdb200_wd_inflight() { return Watchdog::get().inflight(); }
PDB_RT_EXPORT int pdb200_wd_timeouts() { return Watchdog::get().timeouts(); }
PDB_RT_EXPORT int pdb200_wd_set_abort(int a) {
  Watchdog::get().set_abort(a != 0);
  return 0;
}
PDB_RT_EXPORT int64_t pdb200_wd_dump(char* out, int64_t capacity) {
  std::string s = Watchdog::get().dump();
  if ((int64_t)s.size() + 1 > capacity) return -(int64_t)s.size() - 1;
  memcpy(out, s.c_str(), s.size() + 1);
  return (int64_t)s.size();
}
