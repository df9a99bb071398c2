// Communication watchdog: per-collective records with start time, a polling thread that flags tasks older than
// the timeout and prints one trace line per offender (group, seq, op, numel, sync flag, age) followed by the last
// finished task of that group - the same information the reference's CommTaskManager logs
// (paddle/phi/core/distributed/comm_task_manager.cc:135-232, nccl_comm_task.cc:238-268).  Optional abort ends the
// process so the launcher's restart policy takes over (hung in-kernel spins have no NCCL task to abort).
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

#include "rt_common.h"

namespace pdbrt {

struct TaskRec {
  int group;
  int64_t seq;
  std::string op;
  int64_t numel;
  int sync_op;
  std::chrono::steady_clock::time_point start;
  bool reported = false;
};

class Watchdog {
 public:
  static Watchdog& get() {
    static Watchdog w;
    return w;
  }
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
      last_done_[group] = it->second;
      inflight_.erase(it);
    }
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
    for (auto& kv : inflight_) {
      const TaskRec& t = kv.second;
      os << "[in-flight] rank=" << rank_ << " group=" << t.group << " seq=" << t.seq << " op=" << t.op
         << " numel=" << t.numel << " sync=" << t.sync_op << " age_ms="
         << std::chrono::duration_cast<std::chrono::milliseconds>(now - t.start).count() << "\n";
    }
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
      for (auto& kv : inflight_) {
        TaskRec& t = kv.second;
        auto age = std::chrono::duration_cast<std::chrono::milliseconds>(now - t.start).count();
        if (age > timeout_ms_ && !t.reported) {
          t.reported = true;
          timeouts_++;
          fprintf(stderr, "[comm watchdog] rank %d: collective timed out (started, not finished): group=%d seq=%lld op=%s "
                          "numel=%lld sync_op=%d age=%lldms\n", rank_, t.group, (long long)t.seq, t.op.c_str(),
                  (long long)t.numel, t.sync_op, (long long)age);
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

  std::mutex mu_;
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
