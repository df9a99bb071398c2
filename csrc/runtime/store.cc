// TCP key/value store: rank 0 hosts a poll()-driven daemon thread, every rank (rank 0 included) talks to it
// through a client socket.  Semantics follow the reference's TCPStore (paddle/phi/core/distributed/store/
// tcp_store.{h,cc}: commands ADD / GET / CHECK / SET / WAIT / STOP, blocked waiters are notified when the key
// appears, clients retry their connect until the timeout).  Wire format: u8 command, u32 key length, key bytes,
// then per command: SET u64 n + bytes; ADD i64 delta; replies: GET/WAIT u64 n + bytes, ADD i64, CHECK u8.
#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "rt_common.h"

namespace pdbrt {

enum Cmd : uint8_t { CMD_SET = 1, CMD_GET = 2, CMD_ADD = 3, CMD_CHECK = 4, CMD_WAIT = 5, CMD_DELETE = 6, CMD_STOP = 7 };

static bool send_all(int fd, const void* buf, size_t n) {
  const char* p = (const char*)buf;
  while (n > 0) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k <= 0) {
      if (k < 0 && (errno == EINTR)) continue;
      return false;
    }
    p += k;
    n -= (size_t)k;
  }
  return true;
}
static bool recv_all(int fd, void* buf, size_t n) {
  char* p = (char*)buf;
  while (n > 0) {
    ssize_t k = ::recv(fd, p, n, 0);
    if (k <= 0) {
      if (k < 0 && errno == EINTR) continue;
      return false;
    }
    p += k;
    n -= (size_t)k;
  }
  return true;
}

class MasterDaemon {
 public:
  explicit MasterDaemon(int listen_fd) : listen_fd_(listen_fd) {
    if (pipe(wake_) != 0) wake_[0] = wake_[1] = -1;
    thread_ = std::thread([this] { run(); });
  }
  ~MasterDaemon() {
    stop_ = true;
    if (wake_[1] >= 0) {
      char c = 1;
      (void)!write(wake_[1], &c, 1);
    }
    if (thread_.joinable()) thread_.join();
    for (int fd : clients_) ::close(fd);
    ::close(listen_fd_);
    if (wake_[0] >= 0) ::close(wake_[0]);
    if (wake_[1] >= 0) ::close(wake_[1]);
  }

 private:
  void run() {
    while (!stop_) {
      std::vector<pollfd> fds;
      fds.push_back({listen_fd_, POLLIN, 0});
      fds.push_back({wake_[0], POLLIN, 0});
      for (int fd : clients_) fds.push_back({fd, POLLIN, 0});
      int rc = ::poll(fds.data(), fds.size(), 1000);
      if (rc <= 0) continue;
      if (fds[0].revents & POLLIN) {
        int c = ::accept(listen_fd_, nullptr, nullptr);
        if (c >= 0) {
          int one = 1;
          setsockopt(c, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
          clients_.push_back(c);
        }
      }
      std::vector<int> dead;
      for (size_t i = 2; i < fds.size(); ++i) {
        if (fds[i].revents & (POLLIN | POLLHUP | POLLERR)) {
          if (!handle(fds[i].fd)) dead.push_back(fds[i].fd);
        }
      }
      for (int fd : dead) drop(fd);
    }
  }
  void drop(int fd) {
    ::close(fd);
    for (auto it = clients_.begin(); it != clients_.end(); ++it)
      if (*it == fd) {
        clients_.erase(it);
        break;
      }
    for (auto& kv : waiters_) {
      auto& v = kv.second;
      for (auto it = v.begin(); it != v.end();)
        it = (it->first == fd) ? v.erase(it) : it + 1;
    }
  }
  bool reply_value(int fd, const std::string& val) {
    uint64_t n = val.size();
    return send_all(fd, &n, 8) && (n == 0 || send_all(fd, val.data(), n));
  }
  void notify(const std::string& key) {
    auto it = waiters_.find(key);
    if (it == waiters_.end()) return;
    const std::string& val = kv_[key];
    for (auto& w : it->second) {
      if (w.second == CMD_GET) reply_value(w.first, val);
      else {
        uint8_t ok = 1;
        send_all(w.first, &ok, 1);
      }
    }
    waiters_.erase(it);
  }
  bool handle(int fd) {
    uint8_t cmd;
    uint32_t klen;
    if (!recv_all(fd, &cmd, 1) || !recv_all(fd, &klen, 4) || klen > (1u << 20)) return false;
    std::string key(klen, '\0');
    if (klen && !recv_all(fd, &key[0], klen)) return false;
    switch (cmd) {
      case CMD_SET: {
        uint64_t n;
        if (!recv_all(fd, &n, 8) || n > (1ull << 31)) return false;
        std::string val(n, '\0');
        if (n && !recv_all(fd, &val[0], n)) return false;
        kv_[key] = std::move(val);
        notify(key);
        return true;
      }
      case CMD_GET: {
        auto it = kv_.find(key);
        if (it != kv_.end()) return reply_value(fd, it->second);
        waiters_[key].push_back({fd, CMD_GET});     // answered when the key is set
        return true;
      }
      case CMD_WAIT: {
        if (kv_.count(key)) {
          uint8_t ok = 1;
          return send_all(fd, &ok, 1);
        }
        waiters_[key].push_back({fd, CMD_WAIT});
        return true;
      }
      case CMD_ADD: {
        int64_t delta;
        if (!recv_all(fd, &delta, 8)) return false;
        int64_t cur = 0;
        auto it = kv_.find(key);
        if (it != kv_.end() && it->second.size() == 8) memcpy(&cur, it->second.data(), 8);
        cur += delta;
        kv_[key] = std::string((const char*)&cur, 8);
        notify(key);
        return send_all(fd, &cur, 8);
      }
      case CMD_CHECK: {
        uint8_t ok = kv_.count(key) ? 1 : 0;
        return send_all(fd, &ok, 1);
      }
      case CMD_DELETE: {
        uint8_t ok = kv_.erase(key) ? 1 : 0;
        return send_all(fd, &ok, 1);
      }
      case CMD_STOP:
        stop_ = true;
        return true;
      default:
        return false;
    }
  }

  int listen_fd_;
  int wake_[2];
  std::atomic<bool> stop_{false};
  std::thread thread_;
  std::vector<int> clients_;
  std::map<std::string, std::string> kv_;
  std::map<std::string, std::vector<std::pair<int, uint8_t>>> waiters_;
};

class TCPStore {
 public:
  TCPStore(const std::string& host, int port, bool is_master, int world, int timeout_ms)
      : host_(host), port_(port), timeout_ms_(timeout_ms) {
    if (is_master) {
      int lfd = ::socket(AF_INET, SOCK_STREAM, 0);
      int one = 1;
      setsockopt(lfd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
      sockaddr_in addr{};
      addr.sin_family = AF_INET;
      addr.sin_addr.s_addr = INADDR_ANY;
      addr.sin_port = htons((uint16_t)port);
      if (::bind(lfd, (sockaddr*)&addr, sizeof(addr)) != 0 || ::listen(lfd, 1024) != 0) {
        ::close(lfd);
        throw std::runtime_error(std::string("TCPStore: cannot listen on port ") + std::to_string(port) + ": " + strerror(errno));
      }
      if (port == 0) {
        socklen_t len = sizeof(addr);
        getsockname(lfd, (sockaddr*)&addr, &len);
        port_ = ntohs(addr.sin_port);
      }
      daemon_.reset(new MasterDaemon(lfd));
    }
    connect_with_retry();
    (void)world;
  }
  ~TCPStore() {
    if (fd_ >= 0) ::close(fd_);
    daemon_.reset();
  }
  int port() const { return port_; }

  bool set(const std::string& key, const void* val, uint64_t n) {
    std::lock_guard<std::mutex> g(mu_);
    return send_header(CMD_SET, key) && send_all(fd_, &n, 8) && (n == 0 || send_all(fd_, val, n));
  }
  int64_t get(const std::string& key, void* out, int64_t cap, int timeout_ms) {
    std::lock_guard<std::mutex> g(mu_);
    if (!send_header(CMD_GET, key)) return -1;
    if (!wait_readable(timeout_ms)) return -2;
    uint64_t n;
    if (!recv_all(fd_, &n, 8)) return -1;
    std::string tmp(n, '\0');
    if (n && !recv_all(fd_, &tmp[0], n)) return -1;
    if ((int64_t)n > cap) return -(int64_t)n - 16;   // caller retries with a larger buffer via check/size protocol
    if (n) memcpy(out, tmp.data(), n);
    return (int64_t)n;
  }
  int64_t add(const std::string& key, int64_t delta) {
    std::lock_guard<std::mutex> g(mu_);
    if (!send_header(CMD_ADD, key) || !send_all(fd_, &delta, 8)) return INT64_MIN;
    int64_t v;
    return recv_all(fd_, &v, 8) ? v : INT64_MIN;
  }
  int check(const std::string& key) {
    std::lock_guard<std::mutex> g(mu_);
    uint8_t ok;
    if (!send_header(CMD_CHECK, key) || !recv_all(fd_, &ok, 1)) return -1;
    return ok;
  }
  int del(const std::string& key) {
    std::lock_guard<std::mutex> g(mu_);
    uint8_t ok;
    if (!send_header(CMD_DELETE, key) || !recv_all(fd_, &ok, 1)) return -1;
    return ok;
  }
  int wait(const std::string& key, int timeout_ms) {
    std::lock_guard<std::mutex> g(mu_);
    if (!send_header(CMD_WAIT, key)) return -1;
    if (!wait_readable(timeout_ms)) return -2;
    uint8_t ok;
    return recv_all(fd_, &ok, 1) ? 0 : -1;
  }

 private:
  bool wait_readable(int timeout_ms) {
    pollfd p{fd_, POLLIN, 0};
    int rc = ::poll(&p, 1, timeout_ms < 0 ? timeout_ms_ : timeout_ms);
    return rc > 0;
  }
  bool send_header(uint8_t cmd, const std::string& key) {
    uint32_t klen = (uint32_t)key.size();
    return send_all(fd_, &cmd, 1) && send_all(fd_, &klen, 4) && (klen == 0 || send_all(fd_, key.data(), klen));
  }
  void connect_with_retry() {
    auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms_);
    addrinfo hints{}, *res = nullptr;
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    std::string port = std::to_string(port_);
    while (true) {
      if (getaddrinfo(host_.c_str(), port.c_str(), &hints, &res) == 0 && res) {
        int fd = ::socket(res->ai_family, res->ai_socktype, res->ai_protocol);
        if (fd >= 0 && ::connect(fd, res->ai_addr, res->ai_addrlen) == 0) {
          int one = 1;
          setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
          fd_ = fd;
          freeaddrinfo(res);
          return;
        }
        if (fd >= 0) ::close(fd);
        freeaddrinfo(res);
        res = nullptr;
      }
      if (std::chrono::steady_clock::now() > deadline)
        throw std::runtime_error("TCPStore: connect to " + host_ + ":" + port + " timed out");
      std::this_thread::sleep_for(std::chrono::milliseconds(100));
    }
  }

  std::string host_;
  int port_;
  int timeout_ms_;
  int fd_ = -1;
  std::mutex mu_;
  std::unique_ptr<MasterDaemon> daemon_;
};

}  // namespace pdbrt

using pdbrt::TCPStore;

PDB_RT_EXPORT void* pdb200_store_create(const char* host, int port, int is_master, int world_size, int timeout_ms) {
  try {
    return new TCPStore(host ? host : "127.0.0.1", port, is_master != 0, world_size, timeout_ms > 0 ? timeout_ms : 900000);
  } catch (const std::exception& e) {
    pdbrt::set_error("%s", e.what());
    return nullptr;
  }
}
PDB_RT_EXPORT void pdb200_store_destroy(void* s) { delete (TCPStore*)s; }
PDB_RT_EXPORT int pdb200_store_set(void* s, const char* key, const void* value, int64_t nbytes) {
  return ((TCPStore*)s)->set(key, value, (uint64_t)nbytes) ? 0 : -1;
}
PDB_RT_EXPORT int64_t pdb200_store_get(void* s, const char* key, void* out, int64_t capacity, int timeout_ms) {
  return ((TCPStore*)s)->get(key, out, capacity, timeout_ms);
}
PDB_RT_EXPORT int64_t pdb200_store_add(void* s, const char* key, int64_t delta) { return ((TCPStore*)s)->add(key, delta); }
PDB_RT_EXPORT int pdb200_store_check(void* s, const char* key) { return ((TCPStore*)s)->check(key); }
PDB_RT_EXPORT int pdb200_store_wait(void* s, const char* key, int timeout_ms) { return ((TCPStore*)s)->wait(key, timeout_ms); }
PDB_RT_EXPORT int pdb200_store_delete(void* s, const char* key) { return ((TCPStore*)s)->del(key); }
PDB_RT_EXPORT int pdb200_store_port(void* s) { return ((TCPStore*)s)->port(); }
