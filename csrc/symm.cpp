// Symmetric heap: one allocation per rank, mapped into every peer of the node, plus an
// optional NVSwitch multicast alias — the B200 replacement for "MPI.COMM_WORLD + device
// buffers" (/root/reference/src/mpi_extensions.jl:97-106; the reference never touches peer
// memory, it stages through the host).
//
// Three modes:
//   VMM    cuMemCreate (POSIX-fd shareable) -> fd passed over a Unix socket (csrc/fdx.cpp) ->
//          cuMemImportFromShareableHandle + cuMemMap in the peer. Needed for multicast:
//          cuMulticastCreate / AddDevice / BindMem give a second VA whose loads reduce and
//          whose stores replicate inside the switch (NVLS).
//   IPC    cudaMalloc + cudaIpcGetMemHandle / cudaIpcOpenMemHandle. No multicast; the robust
//          fallback when VMM export is not permitted in a container.
//   LOCAL  plain cudaMalloc (world of one, single-GPU emulation).
//
// The driver API is resolved through cudaGetDriverEntryPoint, so this object has no link-time
// dependency on libcuda and loads on a machine without a GPU (CPU build check).
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <mutex>
#include <unistd.h>

#include "../include/fmpi.h"

namespace fmpi_host {
int set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
}  // namespace fmpi_host
using fmpi_host::check_cuda;
using fmpi_host::set_error;

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
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*GetErrorString)(CUresult, const char**);
  bool ok = false;
  bool mc_ok = false;
};

Drv g_drv;
std::once_flag g_drv_once;

template <typename F>
bool resolve(const char* name, F& fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || !p ||
      q != cudaDriverEntryPointSuccess) {
    (void)cudaGetLastError();
    return false;
  }
  fn = reinterpret_cast<F>(p);
  return true;
}

void load_driver() {
  Drv& d = g_drv;
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
  d.ok = ok;
  bool mc = true;
  mc &= resolve("cuMulticastCreate", d.MulticastCreate);
  mc &= resolve("cuMulticastAddDevice", d.MulticastAddDevice);
  mc &= resolve("cuMulticastBindMem", d.MulticastBindMem);
  mc &= resolve("cuMulticastGetGranularity", d.MulticastGetGranularity);
  mc &= resolve("cuMulticastUnbind", d.MulticastUnbind);
  d.mc_ok = ok && mc;
}

Drv& drv() {
  std::call_once(g_drv_once, load_driver);
  return g_drv;
}

int check_cu(CUresult r, const char* what) {
  if (r == CUDA_SUCCESS) return 0;
  const char* s = nullptr;
  if (g_drv.GetErrorString) g_drv.GetErrorString(r, &s);
  return set_error("%s: %s (CUresult %d)", what, s ? s : "unknown", int(r));
}

size_t round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

CUmemAllocationProp alloc_prop(int device) {
  CUmemAllocationProp p;
  memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = device;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

int map_handle(CUmemGenericAllocationHandle h, size_t size, size_t gran, int device, void** out) {
  Drv& d = drv();
  CUdeviceptr va = 0;
  if (check_cu(d.MemAddressReserve(&va, size, gran, 0, 0), "cuMemAddressReserve")) return 1;
  if (check_cu(d.MemMap(va, size, 0, h, 0), "cuMemMap")) {
    d.MemAddressFree(va, size);
    return 1;
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  if (check_cu(d.MemSetAccess(va, size, &acc, 1), "cuMemSetAccess")) {
    d.MemUnmap(va, size);
    d.MemAddressFree(va, size);
    return 1;
  }
  *out = reinterpret_cast<void*>(va);
  return 0;
}

}  // namespace

struct fmpi_symm {
  int device = 0, mode = FMPI_SYMM_LOCAL, world = 1, rank = -1;
  size_t size = 0, gran = 0;
  void* local = nullptr;
  CUmemGenericAllocationHandle handle = 0;
  bool has_handle = false;
  struct Peer {
    void* ptr = nullptr;
    CUmemGenericAllocationHandle h = 0;
    bool vmm = false, ipc = false;
  } peers[FMPI_MAX_RANKS];
  CUmemGenericAllocationHandle mc_handle = 0;
  bool has_mc = false, mc_bound = false;
  void* mc_ptr = nullptr;
  size_t mc_gran = 0;
};

extern "C" {

int fmpi_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  return n;
}

int fmpi_symm_mc_supported(int device) {
  Drv& d = drv();
  if (!d.mc_ok) return 0;
  if (cudaSetDevice(device) != cudaSuccess) return 0;
  cudaFree(0);
  CUdevice dev;
  if (d.DeviceGet(&dev, device) != CUDA_SUCCESS) return 0;
  int v = 0;
  if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return 0;
  return v;
}

int fmpi_symm_create(int device, size_t bytes, int mode, int want_multicast, int world,
                     fmpi_symm** out) {
  if (!out || bytes == 0) return set_error("fmpi_symm_create: bad arguments");
  if (world < 1 || world > FMPI_MAX_RANKS) return set_error("world %d out of range", world);
  if (check_cuda(cudaSetDevice(device), "cudaSetDevice")) return 1;
  if (check_cuda(cudaFree(0), "context init")) return 1;
  auto* s = new fmpi_symm;
  s->device = device;
  s->mode = mode;
  s->world = world;
  if (mode == FMPI_SYMM_VMM) {
    Drv& d = drv();
    if (!d.ok) { delete s; return set_error("CUDA driver VMM entry points unavailable"); }
    CUmemAllocationProp prop = alloc_prop(device);
    size_t gran = 0;
    if (check_cu(d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED),
                 "cuMemGetAllocationGranularity")) { delete s; return 1; }
    if (want_multicast && d.mc_ok) {
      CUmulticastObjectProp mp;
      memset(&mp, 0, sizeof(mp));
      mp.numDevices = unsigned(world);
      mp.size = round_up(bytes, gran);
      mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      size_t mg = 0;
      if (d.MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg) {
        s->mc_gran = mg;
        if (mg > gran) gran = mg;
      }
    }
    s->gran = gran;
    s->size = round_up(bytes, gran);
    if (check_cu(d.MemCreate(&s->handle, s->size, &prop, 0), "cuMemCreate")) { delete s; return 1; }
    s->has_handle = true;
    if (map_handle(s->handle, s->size, gran, device, &s->local)) {
      d.MemRelease(s->handle);
      delete s;
      return 1;
    }
  } else {
    s->gran = size_t(2) << 20;
    s->size = round_up(bytes, s->gran);
    if (check_cuda(cudaMalloc(&s->local, s->size), "cudaMalloc(symmetric window)")) { delete s; return 1; }
  }
  if (check_cuda(cudaMemset(s->local, 0, s->size), "cudaMemset(symmetric window)")) {
    fmpi_symm_destroy(s);
    return 1;
  }
  if (check_cuda(cudaDeviceSynchronize(), "cudaDeviceSynchronize")) { fmpi_symm_destroy(s); return 1; }
  *out = s;
  return 0;
}

size_t fmpi_symm_size(const fmpi_symm* s) { return s ? s->size : 0; }
void* fmpi_symm_local_ptr(const fmpi_symm* s) { return s ? s->local : nullptr; }
int fmpi_symm_mode(const fmpi_symm* s) { return s ? s->mode : -1; }

int fmpi_symm_set_self(fmpi_symm* s, int rank) {
  if (!s || rank < 0 || rank >= s->world) return set_error("bad rank");
  s->rank = rank;
  s->peers[rank].ptr = s->local;
  return 0;
}

void* fmpi_symm_peer_ptr(const fmpi_symm* s, int r) {
  if (!s || r < 0 || r >= s->world) return nullptr;
  return s->peers[r].ptr;
}

int fmpi_symm_export_fd(fmpi_symm* s, int* fd_out) {
  if (!s || s->mode != FMPI_SYMM_VMM || !s->has_handle) return set_error("not a VMM allocation");
  int fd = -1;
  if (check_cu(drv().MemExportToShareableHandle(&fd, s->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0),
               "cuMemExportToShareableHandle"))
    return 1;
  *fd_out = fd;
  return 0;
}

int fmpi_symm_export_ipc(fmpi_symm* s, void* handle64) {
  if (!s || s->mode != FMPI_SYMM_IPC) return set_error("not an IPC allocation");
  cudaIpcMemHandle_t h;
  if (check_cuda(cudaIpcGetMemHandle(&h, s->local), "cudaIpcGetMemHandle")) return 1;
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return 0;
}

int fmpi_symm_import_fd(fmpi_symm* s, int r, int fd) {
  if (!s || r < 0 || r >= s->world || r == s->rank) return set_error("bad peer %d", r);
  if (check_cuda(cudaSetDevice(s->device), "cudaSetDevice")) return 1;
  Drv& d = drv();
  CUmemGenericAllocationHandle h;
  if (check_cu(d.MemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                              CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
               "cuMemImportFromShareableHandle"))
    return 1;
  close(fd);
  void* p = nullptr;
  if (map_handle(h, s->size, s->gran, s->device, &p)) {
    d.MemRelease(h);
    return 1;
  }
  s->peers[r].ptr = p;
  s->peers[r].h = h;
  s->peers[r].vmm = true;
  return 0;
}

int fmpi_symm_import_ipc(fmpi_symm* s, int r, const void* handle64) {
  if (!s || r < 0 || r >= s->world || r == s->rank) return set_error("bad peer %d", r);
  if (check_cuda(cudaSetDevice(s->device), "cudaSetDevice")) return 1;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  if (check_cuda(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle"))
    return 1;
  s->peers[r].ptr = p;
  s->peers[r].ipc = true;
  return 0;
}

int fmpi_symm_mc_create(fmpi_symm* s, int* fd_out) {
  if (!s || s->mode != FMPI_SYMM_VMM) return set_error("multicast needs a VMM allocation");
  Drv& d = drv();
  if (!d.mc_ok) return set_error("multicast driver entry points unavailable");
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = unsigned(s->world);
  mp.size = s->size;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  if (check_cu(d.MulticastCreate(&s->mc_handle, &mp), "cuMulticastCreate")) return 1;
  s->has_mc = true;
  int fd = -1;
  if (check_cu(d.MemExportToShareableHandle(&fd, s->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0),
               "cuMemExportToShareableHandle(multicast)"))
    return 1;
  *fd_out = fd;
  return 0;
}

int fmpi_symm_mc_import(fmpi_symm* s, int fd) {
  if (!s || s->mode != FMPI_SYMM_VMM) return set_error("multicast needs a VMM allocation");
  Drv& d = drv();
  if (!d.mc_ok) return set_error("multicast driver entry points unavailable");
  if (check_cu(d.MemImportFromShareableHandle(&s->mc_handle, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                              CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
               "cuMemImportFromShareableHandle(multicast)"))
    return 1;
  close(fd);
  s->has_mc = true;
  return 0;
}

int fmpi_symm_mc_add_device(fmpi_symm* s) {
  if (!s || !s->has_mc) return set_error("no multicast object");
  Drv& d = drv();
  CUdevice dev;
  if (check_cu(d.DeviceGet(&dev, s->device), "cuDeviceGet")) return 1;
  return check_cu(d.MulticastAddDevice(s->mc_handle, dev), "cuMulticastAddDevice");
}

int fmpi_symm_mc_bind_map(fmpi_symm* s) {
  if (!s || !s->has_mc) return set_error("no multicast object");
  Drv& d = drv();
  if (check_cu(d.MulticastBindMem(s->mc_handle, 0, s->handle, 0, s->size, 0), "cuMulticastBindMem")) return 1;
  s->mc_bound = true;
  return map_handle(s->mc_handle, s->size, s->gran, s->device, &s->mc_ptr);
}

void* fmpi_symm_mc_ptr(const fmpi_symm* s) { return s ? s->mc_ptr : nullptr; }

int fmpi_symm_destroy(fmpi_symm* s) {
  if (!s) return 0;
  cudaSetDevice(s->device);
  cudaDeviceSynchronize();
  Drv& d = g_drv;
  if (s->mc_ptr) {
    d.MemUnmap(reinterpret_cast<CUdeviceptr>(s->mc_ptr), s->size);
    d.MemAddressFree(reinterpret_cast<CUdeviceptr>(s->mc_ptr), s->size);
  }
  if (s->mc_bound) {
    CUdevice dev;
    if (d.DeviceGet(&dev, s->device) == CUDA_SUCCESS) d.MulticastUnbind(s->mc_handle, dev, 0, s->size);
  }
  if (s->has_mc) d.MemRelease(s->mc_handle);
  for (int r = 0; r < s->world; ++r) {
    auto& p = s->peers[r];
    if (r == s->rank || !p.ptr) continue;
    if (p.vmm) {
      d.MemUnmap(reinterpret_cast<CUdeviceptr>(p.ptr), s->size);
      d.MemAddressFree(reinterpret_cast<CUdeviceptr>(p.ptr), s->size);
      d.MemRelease(p.h);
    } else if (p.ipc) {
      cudaIpcCloseMemHandle(p.ptr);
    }
  }
  if (s->local) {
    if (s->mode == FMPI_SYMM_VMM) {
      d.MemUnmap(reinterpret_cast<CUdeviceptr>(s->local), s->size);
      d.MemAddressFree(reinterpret_cast<CUdeviceptr>(s->local), s->size);
      if (s->has_handle) d.MemRelease(s->handle);
    } else {
      cudaFree(s->local);
    }
  }
  (void)cudaGetLastError();
  delete s;
  return 0;
}

}  // extern "C"
