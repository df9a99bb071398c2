// Symmetric-memory arena over NVLink / NVSwitch (the reference has no equivalent: every collective
// there is a host-launched NCCL call).  Each rank of a single-node group
//   1. allocates one physical segment with cuMemCreate (POSIX-FD shareable),
//   2. hands the FD to every peer over a Unix-domain socket (SCM_RIGHTS),
//   3. imports and maps all peers' segments, so device code gets `nranks` base pointers into the same layout,
//   4. optionally binds all segments to one NVLS multicast object (created by rank 0, FD passed the same way) and
//      maps it: stores to that address are replicated to every GPU by the switch, `multimem.ld_reduce` loads
//      return the in-switch reduction.
// The CUDA driver is reached through dlopen (no link-time libcuda), rendezvous uses files + sockets in a directory
// all ranks of the node can see.
#include <cuda.h>
#include <dlfcn.h>
#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/un.h>
#include <unistd.h>

#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

#include "rt_common.h"

namespace pdbrt {

struct Driver {
  void* lib = nullptr;
#define PDB_DRV(name) decltype(&::name) name = nullptr;
  PDB_DRV(cuInit)
  PDB_DRV(cuDeviceGet)
  PDB_DRV(cuDeviceGetAttribute)
  PDB_DRV(cuDevicePrimaryCtxRetain)
  PDB_DRV(cuCtxSetCurrent)
  PDB_DRV(cuMemGetAllocationGranularity)
  PDB_DRV(cuMemCreate)
  PDB_DRV(cuMemRelease)
  PDB_DRV(cuMemExportToShareableHandle)
  PDB_DRV(cuMemImportFromShareableHandle)
  PDB_DRV(cuMemAddressReserve)
  PDB_DRV(cuMemAddressFree)
  PDB_DRV(cuMemMap)
  PDB_DRV(cuMemUnmap)
  PDB_DRV(cuMemSetAccess)
  PDB_DRV(cuMulticastCreate)
  PDB_DRV(cuMulticastAddDevice)
  PDB_DRV(cuMulticastBindMem)
  PDB_DRV(cuMulticastGetGranularity)
  PDB_DRV(cuGetErrorString)
  PDB_DRV(cuMemsetD32Async)
  PDB_DRV(cuStreamWaitValue32)
#undef PDB_DRV
  bool ok = false;
};

static Driver& drv() {
  static Driver d;
  static bool tried = false;
  if (tried) return d;
  tried = true;
  d.lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
  if (!d.lib) return d;
  bool all = true;
#define PDB_LOAD(name, sym)                              \
  d.name = (decltype(d.name))dlsym(d.lib, sym);          \
  if (!d.name) all = false;
  PDB_LOAD(cuInit, "cuInit")
  PDB_LOAD(cuDeviceGet, "cuDeviceGet")
  PDB_LOAD(cuDeviceGetAttribute, "cuDeviceGetAttribute")
  PDB_LOAD(cuDevicePrimaryCtxRetain, "cuDevicePrimaryCtxRetain")
  PDB_LOAD(cuCtxSetCurrent, "cuCtxSetCurrent")
  PDB_LOAD(cuMemGetAllocationGranularity, "cuMemGetAllocationGranularity")
  PDB_LOAD(cuMemCreate, "cuMemCreate")
  PDB_LOAD(cuMemRelease, "cuMemRelease")
  PDB_LOAD(cuMemExportToShareableHandle, "cuMemExportToShareableHandle")
  PDB_LOAD(cuMemImportFromShareableHandle, "cuMemImportFromShareableHandle")
  PDB_LOAD(cuMemAddressReserve, "cuMemAddressReserve")
  PDB_LOAD(cuMemAddressFree, "cuMemAddressFree")
  PDB_LOAD(cuMemMap, "cuMemMap")
  PDB_LOAD(cuMemUnmap, "cuMemUnmap")
  PDB_LOAD(cuMemSetAccess, "cuMemSetAccess")
  PDB_LOAD(cuGetErrorString, "cuGetErrorString")
  d.ok = all;
  // multicast entry points are optional (older drivers / no NVSwitch)
  d.cuMulticastCreate = (decltype(d.cuMulticastCreate))dlsym(d.lib, "cuMulticastCreate");
  d.cuMulticastAddDevice = (decltype(d.cuMulticastAddDevice))dlsym(d.lib, "cuMulticastAddDevice");
  d.cuMulticastBindMem = (decltype(d.cuMulticastBindMem))dlsym(d.lib, "cuMulticastBindMem");
  d.cuMulticastGetGranularity = (decltype(d.cuMulticastGetGranularity))dlsym(d.lib, "cuMulticastGetGranularity");
  // stream-ordered flag writes / waits (optional: the overlapped GEMM+collective paths need them)
  d.cuMemsetD32Async = (decltype(d.cuMemsetD32Async))dlsym(d.lib, "cuMemsetD32Async");
  d.cuStreamWaitValue32 = (decltype(d.cuStreamWaitValue32))dlsym(d.lib, "cuStreamWaitValue32_v2");
  if (!d.cuStreamWaitValue32) d.cuStreamWaitValue32 = (decltype(d.cuStreamWaitValue32))dlsym(d.lib, "cuStreamWaitValue32");
#undef PDB_LOAD
  return d;
}

static std::string cu_err(CUresult r) {
  const char* s = nullptr;
  if (drv().cuGetErrorString) drv().cuGetErrorString(r, &s);
  return std::string(s ? s : "unknown") + " (" + std::to_string((int)r) + ")";
}
#define CU_CHECK(expr)                                                                   \
  do {                                                                                   \
    CUresult _r = (expr);                                                                \
    if (_r != CUDA_SUCCESS) throw std::runtime_error(std::string(#expr) + ": " + cu_err(_r)); \
  } while (0)

// ---- fd passing ---------------------------------------------------------------------------------------------
static bool send_fd(int sock, int fd) {
  char byte = 'F';
  iovec iov{&byte, 1};
  char ctrl[CMSG_SPACE(sizeof(int))] = {0};
  msghdr msg{};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  return sendmsg(sock, &msg, 0) == 1;
}
static int recv_fd(int sock) {
  char byte = 0;
  iovec iov{&byte, 1};
  char ctrl[CMSG_SPACE(sizeof(int))] = {0};
  msghdr msg{};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  if (recvmsg(sock, &msg, 0) != 1) return -1;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_type != SCM_RIGHTS) return -1;
  int fd = -1;
  memcpy(&fd, CMSG_DATA(c), sizeof(int));
  return fd;
}

class Arena {
 public:
  Arena(int rank, int nranks, int device, int64_t bytes, const std::string& dir, const std::string& tag, bool want_mc,
        int timeout_ms)
      : rank_(rank), n_(nranks), dev_(device), dir_(dir), tag_(tag), timeout_ms_(timeout_ms) {
    Driver& d = drv();
    if (!d.ok) throw std::runtime_error("CUDA driver library (libcuda.so.1) not available");
    CU_CHECK(d.cuInit(0));
    CUdevice cudev;
    CU_CHECK(d.cuDeviceGet(&cudev, device));
    CUcontext ctx;
    CU_CHECK(d.cuDevicePrimaryCtxRetain(&ctx, cudev));
    CU_CHECK(d.cuCtxSetCurrent(ctx));

    CUmemAllocationProp prop{};
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = device;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t gran = 0;
    CU_CHECK(d.cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    int mc_supported = 0;
    d.cuDeviceGetAttribute(&mc_supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev);
    use_mc_ = want_mc && mc_supported && d.cuMulticastCreate && nranks > 1;
    CUmulticastObjectProp mcprop{};
    if (use_mc_) {
      mcprop.numDevices = (unsigned)nranks;
      mcprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      mcprop.size = (size_t)bytes;
      size_t mcg = 0;
      if (d.cuMulticastGetGranularity(&mcg, &mcprop, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mcg > gran) gran = mcg;
    }
    size_ = ((size_t)bytes + gran - 1) / gran * gran;
    mcprop.size = size_;

    CU_CHECK(d.cuMemCreate(&local_, size_, &prop, 0));
    int my_fd = -1;
    CU_CHECK(d.cuMemExportToShareableHandle(&my_fd, local_, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));

    // rank 0 owns the multicast object
    int mc_fd = -1;
    if (use_mc_ && rank == 0) {
      CU_CHECK(d.cuMulticastCreate(&mc_, &mcprop));
      CU_CHECK(d.cuMemExportToShareableHandle(&mc_fd, mc_, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    }

    // ---- serve my FDs to peers, fetch theirs
    start_server(my_fd, mc_fd);
    ptrs_.assign(n_, 0);
    handles_.assign(n_, 0);
    for (int p = 0; p < n_; ++p) {
      CUmemGenericAllocationHandle h = local_;
      if (p != rank_) {
        int fd = fetch_fd(p, 'M');
        CU_CHECK(d.cuMemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
        ::close(fd);
      }
      handles_[p] = h;
      CUdeviceptr va = 0;
      CU_CHECK(d.cuMemAddressReserve(&va, size_, gran, 0, 0));
      CU_CHECK(d.cuMemMap(va, size_, 0, h, 0));
      CUmemAccessDesc acc{};
      acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
      acc.location.id = device;
      acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
      CU_CHECK(d.cuMemSetAccess(va, size_, &acc, 1));
      ptrs_[p] = va;
    }
    if (use_mc_) {
      if (rank != 0) {
        int fd = fetch_fd(0, 'C');
        CU_CHECK(d.cuMemImportFromShareableHandle(&mc_, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
        ::close(fd);
      }
      CU_CHECK(d.cuMulticastAddDevice(mc_, cudev));
      file_barrier("mcadd");                      // every device must be added before any bind
      CU_CHECK(d.cuMulticastBindMem(mc_, 0, local_, 0, size_, 0));
      CUdeviceptr va = 0;
      CU_CHECK(d.cuMemAddressReserve(&va, size_, gran, 0, 0));
      CU_CHECK(d.cuMemMap(va, size_, 0, mc_, 0));
      CUmemAccessDesc acc{};
      acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
      acc.location.id = device;
      acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
      CU_CHECK(d.cuMemSetAccess(va, size_, &acc, 1));
      mc_ptr_ = va;
    }
    file_barrier("done");                         // peers may still be fetching my FDs
    stop_server();
    ::close(my_fd);
    if (mc_fd >= 0) ::close(mc_fd);
  }

  ~Arena() {
    Driver& d = drv();
    stop_server();
    if (mc_ptr_) {
      d.cuMemUnmap(mc_ptr_, size_);
      d.cuMemAddressFree(mc_ptr_, size_);
    }
    for (int p = 0; p < (int)ptrs_.size(); ++p) {
      if (ptrs_[p]) {
        d.cuMemUnmap(ptrs_[p], size_);
        d.cuMemAddressFree(ptrs_[p], size_);
      }
      if (p != rank_ && handles_[p]) d.cuMemRelease(handles_[p]);
    }
    if (mc_) d.cuMemRelease(mc_);
    if (local_) d.cuMemRelease(local_);
  }

  int64_t bytes() const { return (int64_t)size_; }
  uint64_t peer(int p) const { return (p >= 0 && p < n_) ? (uint64_t)ptrs_[p] : 0; }
  int nranks() const { return n_; }
  int rank() const { return rank_; }
  uint64_t mc() const { return (uint64_t)mc_ptr_; }

 private:
  std::string sock_path(int r) const { return dir_ + "/" + tag_ + "." + std::to_string(r) + ".sock"; }

  void start_server(int mem_fd, int mc_fd) {
    std::string path = sock_path(rank_);
    ::unlink(path.c_str());
    listen_fd_ = ::socket(AF_UNIX, SOCK_STREAM, 0);
    sockaddr_un addr{};
    addr.sun_family = AF_UNIX;
    snprintf(addr.sun_path, sizeof(addr.sun_path), "%s", path.c_str());
    if (::bind(listen_fd_, (sockaddr*)&addr, sizeof(addr)) != 0 || ::listen(listen_fd_, 64) != 0)
      throw std::runtime_error("symm: cannot listen on " + path + ": " + strerror(errno));
    serving_ = true;
    server_ = std::thread([this, mem_fd, mc_fd] {
      while (serving_) {
        fd_set rf;
        FD_ZERO(&rf);
        FD_SET(listen_fd_, &rf);
        timeval tv{0, 100000};
        if (select(listen_fd_ + 1, &rf, nullptr, nullptr, &tv) <= 0) continue;
        int c = ::accept(listen_fd_, nullptr, nullptr);
        if (c < 0) continue;
        char what = 0;
        if (::recv(c, &what, 1, 0) == 1) send_fd(c, what == 'C' ? mc_fd : mem_fd);
        ::close(c);
      }
    });
  }
  void stop_server() {
    if (!serving_) return;
    serving_ = false;
    if (server_.joinable()) server_.join();
    ::close(listen_fd_);
    ::unlink(sock_path(rank_).c_str());
  }
  int fetch_fd(int peer, char what) {
    auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms_);
    std::string path = sock_path(peer);
    while (true) {
      int s = ::socket(AF_UNIX, SOCK_STREAM, 0);
      sockaddr_un addr{};
      addr.sun_family = AF_UNIX;
      snprintf(addr.sun_path, sizeof(addr.sun_path), "%s", path.c_str());
      if (::connect(s, (sockaddr*)&addr, sizeof(addr)) == 0) {
        if (::send(s, &what, 1, 0) == 1) {
          int fd = recv_fd(s);
          ::close(s);
          if (fd >= 0) return fd;
        } else {
          ::close(s);
        }
      } else {
        ::close(s);
      }
      if (std::chrono::steady_clock::now() > deadline)
        throw std::runtime_error("symm: timed out fetching handle from rank " + std::to_string(peer));
      std::this_thread::sleep_for(std::chrono::milliseconds(20));
    }
  }
  void file_barrier(const char* phase) {
    std::string mine = dir_ + "/" + tag_ + "." + phase + "." + std::to_string(rank_);
    int f = ::open(mine.c_str(), O_CREAT | O_WRONLY, 0600);
    if (f >= 0) ::close(f);
    auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms_);
    for (int r = 0; r < n_; ++r) {
      std::string p = dir_ + "/" + tag_ + "." + phase + "." + std::to_string(r);
      struct stat st;
      while (::stat(p.c_str(), &st) != 0) {
        if (std::chrono::steady_clock::now() > deadline) throw std::runtime_error(std::string("symm: barrier '") + phase + "' timed out");
        std::this_thread::sleep_for(std::chrono::milliseconds(5));
      }
    }
  }

  int rank_, n_, dev_;
  std::string dir_, tag_;
  int timeout_ms_;
  size_t size_ = 0;
  CUmemGenericAllocationHandle local_ = 0, mc_ = 0;
  std::vector<CUmemGenericAllocationHandle> handles_;
  std::vector<CUdeviceptr> ptrs_;
  CUdeviceptr mc_ptr_ = 0;
  bool use_mc_ = false;
  int listen_fd_ = -1;
  std::atomic<bool> serving_{false};
  std::thread server_;
};

}  // namespace pdbrt

using pdbrt::Arena;

PDB_RT_EXPORT void* pdb200_symm_create(int rank, int nranks, int device, int64_t bytes, const char* rendezvous_dir,
                                       const char* tag, int want_multicast, int timeout_ms) {
  try {
    return new Arena(rank, nranks, device, bytes, rendezvous_dir ? rendezvous_dir : "/tmp", tag ? tag : "pdb200symm",
                     want_multicast != 0, timeout_ms > 0 ? timeout_ms : 60000);
  } catch (const std::exception& e) {
    pdbrt::set_error("%s", e.what());
    return nullptr;
  }
}
PDB_RT_EXPORT void pdb200_symm_destroy(void* a) { delete (Arena*)a; }
PDB_RT_EXPORT int64_t pdb200_symm_bytes(void* a) { return ((Arena*)a)->bytes(); }
PDB_RT_EXPORT uint64_t pdb200_symm_peer_ptr(void* a, int peer) { return ((Arena*)a)->peer(peer); }
PDB_RT_EXPORT uint64_t pdb200_symm_multicast_ptr(void* a) { return ((Arena*)a)->mc(); }
PDB_RT_EXPORT int pdb200_symm_supported(int device) {
  pdbrt::Driver& d = pdbrt::drv();
  if (!d.ok) return 0;
  if (d.cuInit(0) != CUDA_SUCCESS) return 0;
  CUdevice dev;
  if (d.cuDeviceGet(&dev, device) != CUDA_SUCCESS) return 0;
  int vmm = 0, fd = 0, mc = 0;
  d.cuDeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
  d.cuDeviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
  d.cuDeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  return (vmm && fd) ? (mc ? 2 : 1) : 0;
}

// Copy-engine all-gather over the arena's peer mappings: for every segment {src_off, dst_off, shard_bytes} this rank's
// shard (at src_off of its own segment) is copied to [dst_off + rank * shard_bytes) of EVERY rank's segment with plain
// device-to-device copies on `stream` - the DMA engines move the bytes over NVLink, no SM runs a copy kernel, so the
// transfer does not take issue slots or L2 bandwidth from the GEMMs it overlaps.  The caller brackets the call with
// the cross-rank barriers (pdb200_nvl_barrier of the kernel library) like every other collective on the arena.
// (The reference moves parameters with ncclAllGather kernels on SMs: group_sharded_stage3.py:849-896.)
PDB_RT_EXPORT int pdb200_symm_allgather_ce(void* a, const int64_t* segs, int nseg, void* stream) {
  pdbrt::Arena* A = (pdbrt::Arena*)a;
  const int n = A->nranks(), r = A->rank();
  cudaStream_t st = (cudaStream_t)stream;
  for (int k = 0; k < nseg; ++k) {
    const int64_t so = segs[3 * k], dof = segs[3 * k + 1], nb = segs[3 * k + 2];
    if (nb <= 0) continue;
    const char* src = (const char*)(uintptr_t)A->peer(r) + so;
    for (int i = 0; i < n; ++i) {
      const int p = (r + i) % n;                       // local copy first, then walk the peers (spreads the ingress)
      char* dst = (char*)(uintptr_t)A->peer(p) + dof + (int64_t)r * nb;
      cudaError_t e = cudaMemcpyAsync(dst, src, (size_t)nb, cudaMemcpyDeviceToDevice, st);
      if (e != cudaSuccess) {
        pdbrt::set_error("symm_allgather_ce: cudaMemcpyAsync to rank %d failed: %s", p, cudaGetErrorString(e));
        return -1;
      }
    }
  }
  return 0;
}

// Copy-engine all-to-all, the transfer half of a reduce-scatter: for every segment {src_off, dst_off, shard_bytes} shard p
// of this rank's buffer (src_off + p * shard_bytes) is copied into slot `rank` of rank p's receive buffer
// (dst_off + rank * shard_bytes).  A local kernel then sums the nranks slots (pdb200_reduce_slots_multi of the kernel
// library), so the only SM work of the reduce-scatter reads local memory.
PDB_RT_EXPORT int pdb200_symm_alltoall_ce(void* a, const int64_t* segs, int nseg, void* stream) {
  pdbrt::Arena* A = (pdbrt::Arena*)a;
  const int n = A->nranks(), r = A->rank();
  cudaStream_t st = (cudaStream_t)stream;
  for (int k = 0; k < nseg; ++k) {
    const int64_t so = segs[3 * k], dof = segs[3 * k + 1], nb = segs[3 * k + 2];
    if (nb <= 0) continue;
    for (int i = 0; i < n; ++i) {
      const int p = (r + i) % n;
      const char* src = (const char*)(uintptr_t)A->peer(r) + so + (int64_t)p * nb;
      char* dst = (char*)(uintptr_t)A->peer(p) + dof + (int64_t)r * nb;
      cudaError_t e = cudaMemcpyAsync(dst, src, (size_t)nb, cudaMemcpyDeviceToDevice, st);
      if (e != cudaSuccess) {
        pdbrt::set_error("symm_alltoall_ce: cudaMemcpyAsync to rank %d failed: %s", p, cudaGetErrorString(e));
        return -1;
      }
    }
  }
  return 0;
}

// Copy-engine broadcast of raw byte ranges: {src_off, dst_off, nbytes}: [src_off, +nbytes) of this rank's segment goes to
// [dst_off, +nbytes) of every rank's segment (this one included).
PDB_RT_EXPORT int pdb200_symm_bcast_ce(void* a, const int64_t* segs, int nseg, void* stream) {
  pdbrt::Arena* A = (pdbrt::Arena*)a;
  const int n = A->nranks(), r = A->rank();
  cudaStream_t st = (cudaStream_t)stream;
  for (int k = 0; k < nseg; ++k) {
    const int64_t so = segs[3 * k], dof = segs[3 * k + 1], nb = segs[3 * k + 2];
    if (nb <= 0) continue;
    const char* src = (const char*)(uintptr_t)A->peer(r) + so;
    for (int i = 0; i < n; ++i) {
      const int p = (r + i) % n;
      cudaError_t e = cudaMemcpyAsync((char*)(uintptr_t)A->peer(p) + dof, src, (size_t)nb, cudaMemcpyDeviceToDevice, st);
      if (e != cudaSuccess) {
        pdbrt::set_error("symm_bcast_ce: cudaMemcpyAsync to rank %d failed: %s", p, cudaGetErrorString(e));
        return -1;
      }
    }
  }
  return 0;
}

// Copy-engine point-to-point: {src_off, dst_off, nbytes} from this rank's segment into rank `peer`'s segment.
PDB_RT_EXPORT int pdb200_symm_put_ce(void* a, int peer, int64_t src_off, int64_t dst_off, int64_t nbytes, void* stream) {
  pdbrt::Arena* A = (pdbrt::Arena*)a;
  if (peer < 0 || peer >= A->nranks() || nbytes < 0) {
    pdbrt::set_error("symm_put_ce: bad arguments");
    return -1;
  }
  if (nbytes == 0) return 0;
  cudaError_t e = cudaMemcpyAsync((char*)(uintptr_t)A->peer(peer) + dst_off, (const char*)(uintptr_t)A->peer(A->rank()) + src_off, (size_t)nbytes,
                                  cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    pdbrt::set_error("symm_put_ce: cudaMemcpyAsync to rank %d failed: %s", peer, cudaGetErrorString(e));
    return -1;
  }
  return 0;
}


// ---- overlapped GEMM + collective plumbing: everything below is enqueued from ONE native call per collective -------
// Flags are 32-bit words in the arena; a flag write is a stream-ordered 32-bit memset into the receiver's mapping, so it
// becomes visible after the copies enqueued before it on the same stream have completed.
PDB_RT_EXPORT int pdb200_symm_streamops_supported() {
  pdbrt::Driver& d = pdbrt::drv();
  return (d.cuMemsetD32Async && d.cuStreamWaitValue32) ? 1 : 0;
}

// K5 push: chunk c (c < nchunks) = [src_off + c*chunk_bytes, +chunk_bytes) of this rank's segment goes to
// [dst_off + c*chunk_bytes, +chunk_bytes) of EVERY rank's segment, then flag word (flag_off + 4*(flag_index0 + c)) of
// every rank <- value.
PDB_RT_EXPORT int pdb200_symm_push_signal(void* a, int64_t src_off, int64_t dst_off, int64_t chunk_bytes, int nchunks, int64_t flag_off,
                                          int flag_index0, int64_t value, void* stream) {
  pdbrt::Arena* A = (pdbrt::Arena*)a;
  pdbrt::Driver& d = pdbrt::drv();
  if (!d.cuMemsetD32Async) {
    pdbrt::set_error("symm_push_signal: cuMemsetD32Async unavailable");
    return -1;
  }
  const int n = A->nranks(), r = A->rank();
  cudaStream_t st = (cudaStream_t)stream;
  for (int c = 0; c < nchunks; ++c) {
    const char* src = (const char*)(uintptr_t)A->peer(r) + src_off + (int64_t)c * chunk_bytes;
    for (int i = 0; i < n; ++i) {
      const int p = (r + i) % n;
      if (cudaMemcpyAsync((char*)(uintptr_t)A->peer(p) + dst_off + (int64_t)c * chunk_bytes, src, (size_t)chunk_bytes,
                          cudaMemcpyDeviceToDevice, st) != cudaSuccess) {
        pdbrt::set_error("symm_push_signal: copy to rank %d failed: %s", p, cudaGetErrorString(cudaGetLastError()));
        return -1;
      }
    }
    for (int i = 0; i < n; ++i) {
      const int p = (r + i) % n;
      CUresult e = d.cuMemsetD32Async((CUdeviceptr)(A->peer(p) + flag_off + 4ull * (uint64_t)(flag_index0 + c)), (unsigned int)value, 1,
                                      (CUstream)st);
      if (e != CUDA_SUCCESS) {
        pdbrt::set_error("symm_push_signal: flag write to rank %d failed: %s", p, pdbrt::cu_err(e).c_str());
        return -1;
      }
    }
  }
  return 0;
}

// K6 ship: for the owners o = rank+1, rank+2, ... (n-1 of them): wait until the local 32-bit counter
// counters_off[o] has reached `target` (a running GEMM counts finished work there), copy block o
// ([part_off + o*blk_bytes, +blk_bytes) of this rank) into slot `rank` of owner o's receive buffer
// ([recv_off + rank*blk_bytes)), then owner o's flag word (flag_off + 4*rank) <- value.
PDB_RT_EXPORT int pdb200_symm_ship_blocks(void* a, int64_t counters_off, int64_t target, int64_t part_off, int64_t recv_off,
                                          int64_t blk_bytes, int64_t flag_off, int64_t value, void* stream) {
  pdbrt::Arena* A = (pdbrt::Arena*)a;
  pdbrt::Driver& d = pdbrt::drv();
  if (!d.cuMemsetD32Async || !d.cuStreamWaitValue32) {
    pdbrt::set_error("symm_ship_blocks: stream memory operations unavailable");
    return -1;
  }
  const int n = A->nranks(), r = A->rank();
  cudaStream_t st = (cudaStream_t)stream;
  for (int t = 0; t < n - 1; ++t) {
    const int o = (r + 1 + t) % n;
    CUresult e = d.cuStreamWaitValue32((CUstream)st, (CUdeviceptr)(A->peer(r) + counters_off + 4ull * (uint64_t)o), (cuuint32_t)target,
                                       CU_STREAM_WAIT_VALUE_GEQ);
    if (e != CUDA_SUCCESS) {
      pdbrt::set_error("symm_ship_blocks: cuStreamWaitValue32 failed: %s", pdbrt::cu_err(e).c_str());
      return -1;
    }
    if (cudaMemcpyAsync((char*)(uintptr_t)A->peer(o) + recv_off + (int64_t)r * blk_bytes,
                        (const char*)(uintptr_t)A->peer(r) + part_off + (int64_t)o * blk_bytes, (size_t)blk_bytes, cudaMemcpyDeviceToDevice,
                        st) != cudaSuccess) {
      pdbrt::set_error("symm_ship_blocks: copy to rank %d failed: %s", o, cudaGetErrorString(cudaGetLastError()));
      return -1;
    }
    e = d.cuMemsetD32Async((CUdeviceptr)(A->peer(o) + flag_off + 4ull * (uint64_t)r), (unsigned int)value, 1, (CUstream)st);
    if (e != CUDA_SUCCESS) {
      pdbrt::set_error("symm_ship_blocks: flag write to rank %d failed: %s", o, pdbrt::cu_err(e).c_str());
      return -1;
    }
  }
  return 0;
}

// Stream waits until every local flag word flag_off[i], i < nranks, i != skip, has reached `value`.
PDB_RT_EXPORT int pdb200_symm_wait_flags(void* a, int64_t flag_off, int64_t value, int skip, void* stream) {
  pdbrt::Arena* A = (pdbrt::Arena*)a;
  pdbrt::Driver& d = pdbrt::drv();
  if (!d.cuStreamWaitValue32) {
    pdbrt::set_error("symm_wait_flags: cuStreamWaitValue32 unavailable");
    return -1;
  }
  for (int i = 0; i < A->nranks(); ++i) {
    if (i == skip) continue;
    CUresult e = d.cuStreamWaitValue32((CUstream)stream, (CUdeviceptr)(A->peer(A->rank()) + flag_off + 4ull * (uint64_t)i), (cuuint32_t)value,
                                       CU_STREAM_WAIT_VALUE_GEQ);
    if (e != CUDA_SUCCESS) {
      pdbrt::set_error("symm_wait_flags: cuStreamWaitValue32 failed: %s", pdbrt::cu_err(e).c_str());
      return -1;
    }
  }
  return 0;
}
