// Symmetric heap: one equally sized device allocation per rank, mapped into every peer
// process of the NVSwitch domain, plus an NVLS multicast alias when the driver has one.
//
// This is what stands in for the reference's ActorRef peer map (AllreduceMessage.scala:8,
// AllreduceWorker.scala:12): instead of an address you can send messages to, a peer is a
// device pointer you can store into over NVLink.
//
// Allocation goes through the CUDA VMM driver API (cuMemCreate) so the memory can be exported
// as a POSIX file descriptor; descriptors travel between the rank processes over unix datagram
// sockets (SCM_RIGHTS) in a directory all ranks share.  If the platform cannot export VMM
// handles, the heap falls back to cudaMalloc + cudaIpc handles (no multicast in that mode).
// Driver entry points are resolved at run time through cudaGetDriverEntryPoint, so the library
// itself does not link against libcuda and loads on machines without a driver.
#include <cuda.h>
#include <cuda_runtime.h>
#include <errno.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <string>
#include <vector>

#include "aar_b200.h"

namespace aar {
extern thread_local std::string g_error;
int fail(int code, const std::string& what);

namespace {

struct Driver {
  decltype(&cuMemCreate) MemCreate = nullptr;
  decltype(&cuMemRelease) MemRelease = nullptr;
  decltype(&cuMemAddressReserve) MemAddressReserve = nullptr;
  decltype(&cuMemAddressFree) MemAddressFree = nullptr;
  decltype(&cuMemMap) MemMap = nullptr;
  decltype(&cuMemUnmap) MemUnmap = nullptr;
  decltype(&cuMemSetAccess) MemSetAccess = nullptr;
  decltype(&cuMemGetAllocationGranularity) MemGetAllocationGranularity = nullptr;
  decltype(&cuMemExportToShareableHandle) MemExportToShareableHandle = nullptr;
  decltype(&cuMemImportFromShareableHandle) MemImportFromShareableHandle = nullptr;
  decltype(&cuMulticastCreate) MulticastCreate = nullptr;
  decltype(&cuMulticastAddDevice) MulticastAddDevice = nullptr;
  decltype(&cuMulticastBindMem) MulticastBindMem = nullptr;
  decltype(&cuMulticastGetGranularity) MulticastGetGranularity = nullptr;
  decltype(&cuDeviceGet) DeviceGet = nullptr;
  decltype(&cuDeviceGetAttribute) DeviceGetAttribute = nullptr;
  decltype(&cuGetErrorString) GetErrorString = nullptr;
  bool ok = false;
};

template <class F>
bool resolve(const char* name, F& fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || p == nullptr) return false;
  fn = reinterpret_cast<F>(p);
  return true;
}

Driver resolve_driver() {
  Driver d;
  bool ok = true;
  ok &= resolve("cuMemCreate", d.MemCreate);
  ok &= resolve("cuMemRelease", d.MemRelease);
  ok &= resolve("cuMemAddressReserve", d.MemAddressReserve);
  ok &= resolve("cuMemAddressFree", d.MemAddressFree);
  ok &= resolve("cuMemMap", d.MemMap);
  ok &= resolve("cuMemUnmap", d.MemUnmap);
  ok &= resolve("cuMemSetAccess", d.MemSetAccess);
  ok &= resolve("cuMemGetAllocationGranularity", d.MemGetAllocationGranularity);
  ok &= resolve("cuMemExportToShareableHandle", d.MemExportToShareableHandle);
  ok &= resolve("cuMemImportFromShareableHandle", d.MemImportFromShareableHandle);
  ok &= resolve("cuDeviceGet", d.DeviceGet);
  ok &= resolve("cuDeviceGetAttribute", d.DeviceGetAttribute);
  ok &= resolve("cuGetErrorString", d.GetErrorString);
  // Multicast entry points are optional (driver >= 12.1).
  resolve("cuMulticastCreate", d.MulticastCreate);
  resolve("cuMulticastAddDevice", d.MulticastAddDevice);
  resolve("cuMulticastBindMem", d.MulticastBindMem);
  resolve("cuMulticastGetGranularity", d.MulticastGetGranularity);
  d.ok = ok;
  return d;
}

Driver& driver() {
  static Driver d = resolve_driver();  // initialised once, thread-safe
  return d;
}

std::string cu_err(CUresult r) {
  const char* s = nullptr;
  if (driver().GetErrorString) driver().GetErrorString(r, &s);
  return s ? s : ("CUresult " + std::to_string(static_cast<int>(r)));
}

// ---- rank-to-rank side channel: unix datagram sockets in a shared directory -------------
struct Packet {
  int32_t kind;   // 0 memory fd, 1 multicast fd, 2 ipc handle, 3 agreement value
  int32_t from;
  int32_t epoch;
  int32_t value;
  char blob[64];
};

class SideChannel {
 public:
  SideChannel(int rank, int world, std::string dir, int seq)
      : rank_(rank), world_(world), dir_(std::move(dir)), seq_(seq) {}
  ~SideChannel() {
    if (fd_ >= 0) {
      close(fd_);
      unlink(path(rank_).c_str());
    }
  }
  bool open_socket(std::string& err) {
    fd_ = socket(AF_UNIX, SOCK_DGRAM | SOCK_CLOEXEC, 0);
    if (fd_ < 0) {
      err = "socket(AF_UNIX) failed";
      return false;
    }
    sockaddr_un sa{};
    sa.sun_family = AF_UNIX;
    const std::string p = path(rank_);
    if (p.size() >= sizeof(sa.sun_path)) {
      err = "unix socket path too long: " + p;
      return false;
    }
    unlink(p.c_str());
    strncpy(sa.sun_path, p.c_str(), sizeof(sa.sun_path) - 1);
    if (bind(fd_, reinterpret_cast<sockaddr*>(&sa), sizeof(sa)) != 0) {
      err = "bind " + p + ": " + strerror(errno);
      return false;
    }
    timeval tv{120, 0};
    setsockopt(fd_, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
    return true;
  }
  bool send_to(int peer, const Packet& pk, int pass_fd, std::string& err) {
    sockaddr_un sa{};
    sa.sun_family = AF_UNIX;
    strncpy(sa.sun_path, path(peer).c_str(), sizeof(sa.sun_path) - 1);
    msghdr msg{};
    iovec iov{const_cast<Packet*>(&pk), sizeof(pk)};
    msg.msg_name = &sa;
    msg.msg_namelen = sizeof(sa);
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    alignas(cmsghdr) char ctl[CMSG_SPACE(sizeof(int))];
    if (pass_fd >= 0) {
      memset(ctl, 0, sizeof(ctl));
      msg.msg_control = ctl;
      msg.msg_controllen = sizeof(ctl);
      cmsghdr* c = CMSG_FIRSTHDR(&msg);
      c->cmsg_level = SOL_SOCKET;
      c->cmsg_type = SCM_RIGHTS;
      c->cmsg_len = CMSG_LEN(sizeof(int));
      memcpy(CMSG_DATA(c), &pass_fd, sizeof(int));
    }
    for (int attempt = 0; attempt < 2400; ++attempt) {  // the peer may not have bound yet: ~120 s
      if (sendmsg(fd_, &msg, 0) >= 0) return true;
      if (errno != ENOENT && errno != ECONNREFUSED && errno != EAGAIN && errno != ENOBUFS) break;
      timespec ts{0, 50 * 1000 * 1000};
      nanosleep(&ts, nullptr);
    }
    err = "sendmsg to rank " + std::to_string(peer) + ": " + strerror(errno);
    return false;
  }
  // Receives until a packet of `kind` (and epoch, for kind 3) shows up; others are stashed.
  bool recv_kind(int kind, int epoch, Packet& pk, int& got_fd, std::string& err) {
    for (size_t i = 0; i < stash_.size(); ++i) {
      if (stash_[i].first.kind == kind && (kind != 3 || stash_[i].first.epoch == epoch)) {
        pk = stash_[i].first;
        got_fd = stash_[i].second;
        stash_.erase(stash_.begin() + i);
        return true;
      }
    }
    for (;;) {
      Packet q;
      int f = -1;
      if (!recv_raw(q, f, err)) return false;
      if (q.kind == kind && (kind != 3 || q.epoch == epoch)) {
        pk = q;
        got_fd = f;
        return true;
      }
      stash_.emplace_back(q, f);
    }
  }
  // min over all ranks of `value`; doubles as a barrier.
  bool agree_min(int value, int& out, std::string& err) {
    const int epoch = ++epoch_;
    Packet pk{};
    pk.kind = 3;
    pk.from = rank_;
    pk.epoch = epoch;
    pk.value = value;
    for (int p = 0; p < world_; ++p)
      if (p != rank_ && !send_to(p, pk, -1, err)) return false;
    out = value;
    for (int i = 0; i < world_ - 1; ++i) {
      Packet q;
      int f;
      if (!recv_kind(3, epoch, q, f, err)) return false;
      if (q.value < out) out = q.value;
    }
    return true;
  }

 private:
  // One socket name per (heap sequence number, rank): ranks create their heaps in the same order.
  std::string path(int r) const { return dir_ + "/aar_heap_" + std::to_string(seq_) + "_" + std::to_string(r) + ".sock"; }
  bool recv_raw(Packet& pk, int& got_fd, std::string& err) {
    msghdr msg{};
    iovec iov{&pk, sizeof(pk)};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    alignas(cmsghdr) char ctl[CMSG_SPACE(sizeof(int))];
    msg.msg_control = ctl;
    msg.msg_controllen = sizeof(ctl);
    const ssize_t n = recvmsg(fd_, &msg, 0);
    if (n != static_cast<ssize_t>(sizeof(pk))) {
      err = std::string("recvmsg on the heap side channel: ") + (n < 0 ? strerror(errno) : "short packet");
      return false;
    }
    got_fd = -1;
    for (cmsghdr* c = CMSG_FIRSTHDR(&msg); c != nullptr; c = CMSG_NXTHDR(&msg, c))
      if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) memcpy(&got_fd, CMSG_DATA(c), sizeof(int));
    return true;
  }
  int rank_, world_;
  std::string dir_;
  int seq_;
  int fd_ = -1;
  int epoch_ = 0;
  std::vector<std::pair<Packet, int>> stash_;
};

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace
}  // namespace aar

using namespace aar;

struct AarHeap {
  int rank = 0, world = 1, device = 0;
  size_t bytes = 0;
  bool vmm = true;
  CUmemGenericAllocationHandle local = 0;
  std::vector<CUmemGenericAllocationHandle> imported;
  std::vector<CUdeviceptr> va;  // [world] mapping of every rank's allocation in this process
  void* ipc_local = nullptr;    // cudaMalloc'd memory in IPC mode
  CUmemGenericAllocationHandle mc = 0;
  CUdeviceptr mc_va = 0;
  bool has_mc = false;
};

extern "C" {

static int map_handle(AarHeap* h, CUmemGenericAllocationHandle handle, CUdeviceptr* out) {
  Driver& d = driver();
  CUdeviceptr va = 0;
  CUresult r = d.MemAddressReserve(&va, h->bytes, 0, 0, 0);
  if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMemAddressReserve: " + cu_err(r));
  r = d.MemMap(va, h->bytes, 0, handle, 0);
  if (r != CUDA_SUCCESS) {
    d.MemAddressFree(va, h->bytes);
    return fail(AAR_ECUDA, "cuMemMap: " + cu_err(r));
  }
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = h->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.MemSetAccess(va, h->bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    d.MemUnmap(va, h->bytes);
    d.MemAddressFree(va, h->bytes);
    return fail(AAR_ECUDA, "cuMemSetAccess: " + cu_err(r));
  }
  *out = va;
  return AAR_OK;
}

int aar_heap_create(AarHeap** out, int rank, int world, int device, int64_t bytes, const char* uds_dir,
                    int want_multicast) {
  if (out == nullptr || world < 1 || world > AAR_MAX_RANKS || rank < 0 || rank >= world || bytes <= 0)
    return fail(AAR_EINVAL, "aar_heap_create: bad arguments");
  if (cudaSetDevice(device) != cudaSuccess || cudaFree(nullptr) != cudaSuccess)
    return fail(AAR_ECUDA, "cannot make the CUDA device current");
  Driver& d = driver();
  AarHeap* h = new AarHeap;
  h->rank = rank;
  h->world = world;
  h->device = device;
  h->va.assign(world, 0);
  h->imported.assign(world, 0);
  // Every early return below goes through this: whatever was created, mapped or imported so far is undone.
  struct Cleanup {
    AarHeap* h;
    ~Cleanup() {
      if (h != nullptr) aar_heap_destroy(h);
    }
  } cleanup{h};

  std::string err;
  static std::atomic<int> heap_seq{0};
  SideChannel ch(rank, world, uds_dir ? uds_dir : "/tmp", heap_seq.fetch_add(1));
  if (world > 1 && !ch.open_socket(err)) return fail(AAR_ECUDA, err);

  // ---- can this platform export VMM handles as file descriptors? (agreed by all ranks)
  int vmm_ok = 0, mc_ok = 0;
  CUdevice cudev = 0;
  if (d.ok && d.DeviceGet(&cudev, device) == CUDA_SUCCESS) {
    int v = 0;
    if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, cudev) == CUDA_SUCCESS)
      vmm_ok = v;
    v = 0;
    if (want_multicast && world > 1 && d.MulticastCreate != nullptr &&
        d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev) == CUDA_SUCCESS)
      mc_ok = v;
  }
  if (getenv("AAR_HEAP_FORCE_IPC") != nullptr) vmm_ok = 0;
  if (world > 1) {
    int agreed = 0;
    if (!ch.agree_min(vmm_ok * 2 + (vmm_ok ? mc_ok : 0), agreed, err)) return fail(AAR_ECUDA, err);
    vmm_ok = agreed >= 2;
    mc_ok = agreed == 3;
  }
  h->vmm = vmm_ok != 0;

  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;

  if (h->vmm) {
    size_t gran = 0;
    CUresult r = d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
    if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMemGetAllocationGranularity: " + cu_err(r));
    CUmulticastObjectProp mprop{};
    if (mc_ok) {
      mprop.numDevices = static_cast<unsigned>(world);
      mprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      mprop.size = round_up(static_cast<size_t>(bytes), gran);
      size_t mgran = 0;
      if (d.MulticastGetGranularity(&mgran, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mgran > gran)
        gran = mgran;
    }
    h->bytes = round_up(static_cast<size_t>(bytes), gran);
    r = d.MemCreate(&h->local, h->bytes, &prop, 0);
    if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMemCreate(" + std::to_string(h->bytes) + "): " + cu_err(r));
    int rc = map_handle(h, h->local, &h->va[rank]);
    if (rc != AAR_OK) return rc;
    if (cudaMemset(reinterpret_cast<void*>(h->va[rank]), 0, h->bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess)
      return fail(AAR_ECUDA, "zero-fill of the symmetric region failed");

    if (world > 1) {
      int my_fd = -1;
      r = d.MemExportToShareableHandle(&my_fd, h->local, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
      if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMemExportToShareableHandle: " + cu_err(r));
      Packet pk{};
      pk.kind = 0;
      pk.from = rank;
      for (int p = 0; p < world; ++p)
        if (p != rank && !ch.send_to(p, pk, my_fd, err)) return fail(AAR_ECUDA, err);
      for (int i = 0; i < world - 1; ++i) {
        Packet q;
        int fd = -1;
        if (!ch.recv_kind(0, 0, q, fd, err)) return fail(AAR_ECUDA, err);
        if (fd < 0 || q.from < 0 || q.from >= world) return fail(AAR_ECUDA, "peer sent no file descriptor");
        r = d.MemImportFromShareableHandle(&h->imported[q.from], reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                           CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
        close(fd);
        if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMemImportFromShareableHandle: " + cu_err(r));
        rc = map_handle(h, h->imported[q.from], &h->va[q.from]);
        if (rc != AAR_OK) return rc;
      }
      close(my_fd);

      // ---- NVLS multicast alias (best effort, all-or-nothing across ranks)
      int step_ok = mc_ok;
      if (mc_ok) {
        mprop.size = h->bytes;
        if (rank == 0) {
          int mfd = -1;
          if (d.MulticastCreate(&h->mc, &mprop) != CUDA_SUCCESS ||
              d.MemExportToShareableHandle(&mfd, h->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS)
            step_ok = 0;
          Packet mk{};
          mk.kind = 1;
          mk.from = 0;
          mk.value = step_ok;
          for (int p = 1; p < world; ++p)
            if (!ch.send_to(p, mk, step_ok ? mfd : -1, err)) return fail(AAR_ECUDA, err);
          if (mfd >= 0) close(mfd);
        } else {
          Packet mk;
          int mfd = -1;
          if (!ch.recv_kind(1, 0, mk, mfd, err)) return fail(AAR_ECUDA, err);
          step_ok = mk.value && mfd >= 0;
          if (step_ok && d.MemImportFromShareableHandle(&h->mc, reinterpret_cast<void*>(static_cast<uintptr_t>(mfd)),
                                                        CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) != CUDA_SUCCESS)
            step_ok = 0;
          if (mfd >= 0) close(mfd);
        }
        if (step_ok && d.MulticastAddDevice(h->mc, cudev) != CUDA_SUCCESS) step_ok = 0;
        int all = 0;
        if (!ch.agree_min(step_ok, all, err)) return fail(AAR_ECUDA, err);  // every device added
        step_ok = all;
        if (step_ok && d.MulticastBindMem(h->mc, 0, h->local, 0, h->bytes, 0) != CUDA_SUCCESS) step_ok = 0;
        if (!ch.agree_min(step_ok, all, err)) return fail(AAR_ECUDA, err);  // every rank bound
        step_ok = all;
        if (step_ok) {
          CUdeviceptr mva = 0;
          if (map_handle(h, h->mc, &mva) == AAR_OK) {
            h->mc_va = mva;
            h->has_mc = true;
          } else {
            step_ok = 0;
          }
        }
        if (!ch.agree_min(step_ok, all, err)) return fail(AAR_ECUDA, err);
        if (!all && h->mc_va != 0) {  // some other rank could not map it: nobody uses it
          d.MemUnmap(h->mc_va, h->bytes);
          d.MemAddressFree(h->mc_va, h->bytes);
          h->has_mc = false;
          h->mc_va = 0;
        }
      }
      int done = 0;
      if (!ch.agree_min(1, done, err)) return fail(AAR_ECUDA, err);  // everyone mapped everyone
    }
  } else {
    // ---- legacy cudaIpc path
    h->bytes = round_up(static_cast<size_t>(bytes), 2u << 20);
    if (cudaMalloc(&h->ipc_local, h->bytes) != cudaSuccess) return fail(AAR_ECUDA, "cudaMalloc of the symmetric region failed");
    if (cudaMemset(h->ipc_local, 0, h->bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess)
      return fail(AAR_ECUDA, "zero-fill of the symmetric region failed");
    h->va[rank] = reinterpret_cast<CUdeviceptr>(h->ipc_local);
    if (world > 1) {
      cudaIpcMemHandle_t mine;
      if (cudaIpcGetMemHandle(&mine, h->ipc_local) != cudaSuccess) return fail(AAR_ECUDA, "cudaIpcGetMemHandle failed");
      Packet pk{};
      pk.kind = 2;
      pk.from = rank;
      static_assert(sizeof(cudaIpcMemHandle_t) <= sizeof(pk.blob), "ipc handle does not fit the packet");
      memcpy(pk.blob, &mine, sizeof(mine));
      for (int p = 0; p < world; ++p)
        if (p != rank && !ch.send_to(p, pk, -1, err)) return fail(AAR_ECUDA, err);
      for (int i = 0; i < world - 1; ++i) {
        Packet q;
        int fd;
        if (!ch.recv_kind(2, 0, q, fd, err)) return fail(AAR_ECUDA, err);
        cudaIpcMemHandle_t theirs;
        memcpy(&theirs, q.blob, sizeof(theirs));
        void* p = nullptr;
        if (cudaIpcOpenMemHandle(&p, theirs, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
          return fail(AAR_ECUDA, "cudaIpcOpenMemHandle failed for rank " + std::to_string(q.from));
        h->va[q.from] = reinterpret_cast<CUdeviceptr>(p);
      }
      int done = 0;
      if (!ch.agree_min(1, done, err)) return fail(AAR_ECUDA, err);
    }
  }
  cleanup.h = nullptr;
  *out = h;
  return AAR_OK;
}

int aar_heap_ptrs(AarHeap* h, uint8_t** peer_bases, uint8_t** mc_base) {
  if (h == nullptr || peer_bases == nullptr) return fail(AAR_EINVAL, "aar_heap_ptrs: null argument");
  for (int p = 0; p < h->world; ++p) peer_bases[p] = reinterpret_cast<uint8_t*>(h->va[p]);
  if (mc_base != nullptr) *mc_base = h->has_mc ? reinterpret_cast<uint8_t*>(h->mc_va) : nullptr;
  return AAR_OK;
}

int64_t aar_heap_bytes(AarHeap* h) { return h ? static_cast<int64_t>(h->bytes) : 0; }

int aar_heap_destroy(AarHeap* h) {
  if (h == nullptr) return AAR_OK;
  cudaDeviceSynchronize();
  Driver& d = driver();
  if (h->vmm) {
    if (h->has_mc && h->mc_va) {
      d.MemUnmap(h->mc_va, h->bytes);
      d.MemAddressFree(h->mc_va, h->bytes);
    }
    if (h->mc) d.MemRelease(h->mc);
    for (int p = 0; p < h->world; ++p) {
      if (h->va[p]) {
        d.MemUnmap(h->va[p], h->bytes);
        d.MemAddressFree(h->va[p], h->bytes);
      }
      if (p != h->rank && h->imported[p]) d.MemRelease(h->imported[p]);
    }
    if (h->local) d.MemRelease(h->local);
  } else {
    for (int p = 0; p < h->world; ++p)
      if (p != h->rank && h->va[p]) cudaIpcCloseMemHandle(reinterpret_cast<void*>(h->va[p]));
    if (h->ipc_local) cudaFree(h->ipc_local);
  }
  delete h;
  return AAR_OK;
}

}  // extern "C"
