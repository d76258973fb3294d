// Symmetric-memory runtime: device windows that every rank of the node maps into its own address space,
// plus a host watchdog for kernels that spin on peer flags.
//
// Two transports:
//   * CUDA IPC  (cudaIpcGetMemHandle / cudaIpcOpenMemHandle): 64-byte handles, exchanged by the Python side over
//     the process group's control plane (gloo / TCPStore).  Enough for P2P loads/stores over NVLink.
//   * VMM + POSIX file descriptors (cuMemCreate / cuMemExportToShareableHandle) with optional multicast
//     (cuMulticastCreate / BindMem) for the NVLS path; descriptors travel over a UNIX socket (SCM_RIGHTS).
//
// Native stand-in for ProcessGroupNCCL's communicator/buffer registration and its watchdog
// (reference resnet/main.py:74 init_process_group("nccl")).
#include "b200ddp.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>

#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

// ------------------------------------------------------------------------------------------------ CUDA IPC
extern "C" int b2_ipc_alloc(long long bytes, void** dptr, unsigned char* handle64) {
  cudaError_t e = cudaMalloc(dptr, (size_t)bytes);
  if (e != cudaSuccess) return e;
  e = cudaMemset(*dptr, 0, (size_t)bytes);
  if (e != cudaSuccess) return e;
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, *dptr);
  if (e != cudaSuccess) return e;
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  std::memcpy(handle64, &h, 64);
  return cudaDeviceSynchronize();
}
extern "C" int b2_ipc_open(const unsigned char* handle64, void** dptr) {
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle64, 64);
  return cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess);
}
extern "C" int b2_ipc_close(void* dptr) { return cudaIpcCloseMemHandle(dptr); }
extern "C" int b2_ipc_free(void* dptr) { return cudaFree(dptr); }

// ------------------------------------------------------------------------------------------------ VMM + multicast
namespace {
struct Drv {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                         unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  bool ok = false;
};

template <typename F>
bool load(F& fn, const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
    return false;
  fn = reinterpret_cast<F>(p);
  return true;
}

Drv& drv() {
  static Drv d;
  static std::once_flag once;
  std::call_once(once, [] {
    bool ok = true;
    ok &= load(d.MemCreate, "cuMemCreate");
    ok &= load(d.MemRelease, "cuMemRelease");
    ok &= load(d.MemAddressReserve, "cuMemAddressReserve");
    ok &= load(d.MemAddressFree, "cuMemAddressFree");
    ok &= load(d.MemMap, "cuMemMap");
    ok &= load(d.MemUnmap, "cuMemUnmap");
    ok &= load(d.MemSetAccess, "cuMemSetAccess");
    ok &= load(d.MemGetAllocationGranularity, "cuMemGetAllocationGranularity");
    ok &= load(d.MemExportToShareableHandle, "cuMemExportToShareableHandle");
    ok &= load(d.MemImportFromShareableHandle, "cuMemImportFromShareableHandle");
    ok &= load(d.MulticastCreate, "cuMulticastCreate");
    ok &= load(d.MulticastAddDevice, "cuMulticastAddDevice");
    ok &= load(d.MulticastBindMem, "cuMulticastBindMem");
    ok &= load(d.MulticastGetGranularity, "cuMulticastGetGranularity");
    ok &= load(d.DeviceGetAttribute, "cuDeviceGetAttribute");
    load(d.MulticastUnbind, "cuMulticastUnbind");   // optional: only needed by the teardown
    d.ok = ok;
  });
  return d;
}

CUmemAllocationProp vmm_prop(int device) {
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_handle(CUmemGenericAllocationHandle h, size_t bytes, int device, void** dptr) {
  Drv& d = drv();
  CUdeviceptr va = 0;
  CUresult r = d.MemAddressReserve(&va, bytes, 0, 0, 0);
  if (r != CUDA_SUCCESS) return r;
  r = d.MemMap(va, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) return r;
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.MemSetAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) return r;
  *dptr = reinterpret_cast<void*>(va);
  return 0;
}
}  // namespace

// Rounds `bytes` up to what both plain and multicast mappings need on this device.
extern "C" long long b2_vmm_round(long long bytes, int device, int world) {
  Drv& d = drv();
  if (!d.ok) return -1;
  CUmemAllocationProp prop = vmm_prop(device);
  size_t g = 0;
  if (d.MemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) return -1;
  CUmulticastObjectProp mp{};
  mp.numDevices = world;
  mp.size = (size_t)bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t mg = 0;
  if (world > 1 && d.MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g)
    g = mg;
  return (long long)(((size_t)bytes + g - 1) / g * g);
}
extern "C" int b2_vmm_multicast_supported(int device) {
  Drv& d = drv();
  if (!d.ok) return 0;
  int v = 0;
  if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device) != CUDA_SUCCESS) return 0;
  return v;
}
// Allocate a shareable window on `device`, map it locally, export a POSIX fd. `handle` is the generic handle
// (needed later to bind into a multicast object).
extern "C" int b2_vmm_alloc(long long bytes, int device, void** dptr, unsigned long long* handle, int* fd) {
  Drv& d = drv();
  if (!d.ok) return cudaErrorNotSupported;
  cudaFree(0);
  CUmemAllocationProp prop = vmm_prop(device);
  CUmemGenericAllocationHandle h;
  CUresult r = d.MemCreate(&h, (size_t)bytes, &prop, 0);
  if (r != CUDA_SUCCESS) return r;
  int e = map_handle(h, (size_t)bytes, device, dptr);
  if (e) return e;
  r = d.MemExportToShareableHandle(fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) return r;
  *handle = (unsigned long long)h;
  e = cudaMemset(*dptr, 0, (size_t)bytes);
  if (e) return e;
  return cudaDeviceSynchronize();
}
// Import a peer's window from a received fd and map it for `device`.
extern "C" int b2_vmm_import(int fd, long long bytes, int device, void** dptr) {
  Drv& d = drv();
  if (!d.ok) return cudaErrorNotSupported;
  CUmemGenericAllocationHandle h;
  CUresult r = d.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return r;
  return map_handle(h, (size_t)bytes, device, dptr);
}
// Multicast: rank 0 creates the object and exports an fd; every rank imports it, adds its device, and (after all
// ranks added their device — caller barriers) binds its own window and maps the multicast address.
extern "C" int b2_mc_create(long long bytes, int world, unsigned long long* mc_handle, int* fd) {
  Drv& d = drv();
  if (!d.ok) return cudaErrorNotSupported;
  CUmulticastObjectProp mp{};
  mp.numDevices = world;
  mp.size = (size_t)bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  CUresult r = d.MulticastCreate(&h, &mp);
  if (r != CUDA_SUCCESS) return r;
  r = d.MemExportToShareableHandle(fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) return r;
  *mc_handle = (unsigned long long)h;
  return 0;
}
extern "C" int b2_mc_import(int fd, unsigned long long* mc_handle) {
  Drv& d = drv();
  if (!d.ok) return cudaErrorNotSupported;
  CUmemGenericAllocationHandle h;
  CUresult r = d.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return r;
  *mc_handle = (unsigned long long)h;
  return 0;
}
extern "C" int b2_mc_add_device(unsigned long long mc_handle, int device) {
  return drv().MulticastAddDevice((CUmemGenericAllocationHandle)mc_handle, device);
}
extern "C" int b2_mc_bind_and_map(unsigned long long mc_handle, unsigned long long mem_handle, long long bytes,
                                  int device, void** mc_ptr) {
  Drv& d = drv();
  CUresult r = d.MulticastBindMem((CUmemGenericAllocationHandle)mc_handle, 0,
                                  (CUmemGenericAllocationHandle)mem_handle, 0, (size_t)bytes, 0);
  if (r != CUDA_SUCCESS) return r;
  return map_handle((CUmemGenericAllocationHandle)mc_handle, (size_t)bytes, device, mc_ptr);
}

// ---- teardown (everything the set-up calls above created, in reverse order) ------------------------------------
// Unmap a mapping made by b2_vmm_alloc / b2_vmm_import / b2_mc_bind_and_map and give its address range back.
extern "C" int b2_vmm_unmap(void* dptr, long long bytes) {
  Drv& d = drv();
  if (!d.ok) return cudaErrorNotSupported;
  if (!dptr) return 0;
  CUresult r = d.MemUnmap((CUdeviceptr)(uintptr_t)dptr, (size_t)bytes);
  if (r != CUDA_SUCCESS) return r;
  return d.MemAddressFree((CUdeviceptr)(uintptr_t)dptr, (size_t)bytes);
}
// Drop a generic allocation handle (the local window's, or a multicast object's). The physical memory / the
// multicast object goes away when the last mapping and the last handle (on any rank) are gone.
extern "C" int b2_vmm_release(unsigned long long handle) {
  Drv& d = drv();
  if (!d.ok) return cudaErrorNotSupported;
  return d.MemRelease((CUmemGenericAllocationHandle)handle);
}
// Imported peer windows: b2_vmm_import maps and forgets the handle; this variant also returns it so that the
// teardown can release it (an imported handle holds a reference on the peer's physical memory).
extern "C" int b2_vmm_import_h(int fd, long long bytes, int device, void** dptr, unsigned long long* handle) {
  Drv& d = drv();
  if (!d.ok) return cudaErrorNotSupported;
  CUmemGenericAllocationHandle h;
  CUresult r = d.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return r;
  int e = map_handle(h, (size_t)bytes, device, dptr);
  if (e) {
    d.MemRelease(h);
    return e;
  }
  *handle = (unsigned long long)h;
  return 0;
}
// Unbind this device's window from the multicast object (before unmapping / releasing either).
extern "C" int b2_mc_unbind(unsigned long long mc_handle, int device, long long bytes) {
  Drv& d = drv();
  if (!d.ok || !d.MulticastUnbind) return cudaErrorNotSupported;
  return d.MulticastUnbind((CUmemGenericAllocationHandle)mc_handle, device, 0, (size_t)bytes);
}

// ------------------------------------------------------------------------------------------------ fd passing
static int make_addr(const char* name, sockaddr_un* addr, socklen_t* len) {
  std::memset(addr, 0, sizeof(*addr));
  addr->sun_family = AF_UNIX;
  const size_t n = std::strlen(name);
  if (n + 1 >= sizeof(addr->sun_path)) return -1;
  addr->sun_path[0] = '\0';  // abstract namespace: no filesystem entry, vanishes with the process
  std::memcpy(addr->sun_path + 1, name, n);
  *len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
  return 0;
}
// Listening socket bound to the abstract name; returns the socket fd or -1.
extern "C" int b2_fd_listen(const char* name) {
  sockaddr_un addr;
  socklen_t len;
  if (make_addr(name, &addr, &len)) return -1;
  int s = socket(AF_UNIX, SOCK_STREAM, 0);
  if (s < 0) return -1;
  if (bind(s, (sockaddr*)&addr, len) || listen(s, 64)) {
    close(s);
    return -1;
  }
  return s;
}
// Serve `count` connections: each peer sends one int (index into fds[]) and receives that descriptor.
extern "C" int b2_fd_serve(int listen_sock, const int* fds, int nfds, int count) {
  for (int i = 0; i < count; ++i) {
    int c = accept(listen_sock, nullptr, nullptr);
    if (c < 0) return -1;
    int which = -1;
    if (read(c, &which, sizeof(which)) != (ssize_t)sizeof(which) || which < 0 || which >= nfds) {
      close(c);
      return -2;
    }
    msghdr msg{};
    char ctrl[CMSG_SPACE(sizeof(int))];
    std::memset(ctrl, 0, sizeof(ctrl));
    char byte = 'f';
    iovec io{&byte, 1};
    msg.msg_iov = &io;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    cmsghdr* cm = CMSG_FIRSTHDR(&msg);
    cm->cmsg_level = SOL_SOCKET;
    cm->cmsg_type = SCM_RIGHTS;
    cm->cmsg_len = CMSG_LEN(sizeof(int));
    std::memcpy(CMSG_DATA(cm), &fds[which], sizeof(int));
    if (sendmsg(c, &msg, 0) < 0) {
      close(c);
      return -3;
    }
    close(c);
  }
  return 0;
}
// Connect to a peer's abstract socket and fetch descriptor number `which`. Returns the received fd or < 0.
extern "C" int b2_fd_fetch(const char* name, int which, int timeout_ms) {
  sockaddr_un addr;
  socklen_t len;
  if (make_addr(name, &addr, &len)) return -1;
  int s = socket(AF_UNIX, SOCK_STREAM, 0);
  if (s < 0) return -1;
  int waited = 0;
  while (connect(s, (sockaddr*)&addr, len)) {
    if (waited >= timeout_ms) {
      close(s);
      return -2;
    }
    usleep(10000);
    waited += 10;
  }
  if (write(s, &which, sizeof(which)) != (ssize_t)sizeof(which)) {
    close(s);
    return -3;
  }
  msghdr msg{};
  char ctrl[CMSG_SPACE(sizeof(int))];
  char byte;
  iovec io{&byte, 1};
  msg.msg_iov = &io;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  if (recvmsg(s, &msg, 0) <= 0) {
    close(s);
    return -4;
  }
  close(s);
  cmsghdr* cm = CMSG_FIRSTHDR(&msg);
  if (!cm || cm->cmsg_type != SCM_RIGHTS) return -5;
  int fd;
  std::memcpy(&fd, CMSG_DATA(cm), sizeof(int));
  return fd;
}
extern "C" void b2_fd_close(int fd) { close(fd); }

// ------------------------------------------------------------------------------------------------ watchdog
// Kernels that wait for peers have a device-side spin bound; the host side of failure detection is this thread:
// it polls the most recently registered "step finished" event and, if that stays pending longer than the
// timeout (a peer died or stalled), reports and exits non-zero so the launcher (torchrun) can restart the job.
namespace {
struct Watchdog {
  std::thread th;
  std::atomic<bool> run{false};
  std::atomic<cudaEvent_t> ev{nullptr};
  std::atomic<long long> since_ms{0};
  std::atomic<int> fired{0};
  int timeout_s = 0;
  int device = 0;
  int abort_on_fire = 1;
};
Watchdog g_wd;
long long now_ms() {
  return std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now().time_since_epoch())
      .count();
}
}  // namespace

extern "C" int b2_watchdog_start(int timeout_s, int device, int abort_on_fire) {
  if (g_wd.run.load()) return 0;
  g_wd.timeout_s = timeout_s;
  g_wd.device = device;
  g_wd.abort_on_fire = abort_on_fire;
  g_wd.run.store(true);
  g_wd.th = std::thread([] {
    cudaSetDevice(g_wd.device);
    while (g_wd.run.load()) {
      std::this_thread::sleep_for(std::chrono::milliseconds(100));
      cudaEvent_t e = g_wd.ev.load();
      if (!e) continue;
      cudaError_t q = cudaEventQuery(e);
      if (q == cudaSuccess) {
        g_wd.ev.compare_exchange_strong(e, nullptr);
        continue;
      }
      const bool timed_out = q == cudaErrorNotReady && (now_ms() - g_wd.since_ms.load()) > 1000LL * g_wd.timeout_s;
      if (timed_out || (q != cudaErrorNotReady)) {
        g_wd.fired.store(1);
        std::fprintf(stderr,
                     "[b200ddp watchdog] device %d: %s (event pending %.1f s, limit %d s) — a peer rank is gone or "
                     "stalled; exiting so the launcher can restart from the last checkpoint\n",
                     g_wd.device, timed_out ? "collective timed out" : cudaGetErrorString(q),
                     (now_ms() - g_wd.since_ms.load()) / 1000.0, g_wd.timeout_s);
        std::fflush(stderr);
        if (g_wd.abort_on_fire) _exit(86);
        g_wd.ev.store(nullptr);
      }
    }
  });
  return 0;
}
extern "C" void b2_watchdog_watch(void* cuda_event) {
  g_wd.since_ms.store(now_ms());
  g_wd.ev.store((cudaEvent_t)cuda_event);
}
extern "C" int b2_watchdog_fired(void) { return g_wd.fired.load(); }
extern "C" void b2_watchdog_stop(void) {
  if (!g_wd.run.load()) return;
  g_wd.run.store(false);
  if (g_wd.th.joinable()) g_wd.th.join();
  g_wd.ev.store(nullptr);
}
