// Topology discovery for Init: devices, peer-access matrix, NVLink link state (NVML, loaded
// at run time), multicast capability. The reference picks a device blindly as
// (rank + 1) % device_count (/root/reference/src/common.jl:33-34) and knows nothing about the
// links; here Init records what the box looks like and the collectives' algorithm choice
// (P2P vs NVLS) starts from it.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>

#include "../include/fmpi.h"

namespace fmpi_host {
int set_error(const char* fmt, ...);
}

namespace {

struct Nvml {
  void* lib = nullptr;
  int (*Init)() = nullptr;
  int (*Shutdown)() = nullptr;
  int (*GetHandleByPciBusId)(const char*, void**) = nullptr;
  int (*GetNvLinkState)(void*, unsigned, unsigned*) = nullptr;
  int (*GetNvLinkVersion)(void*, unsigned, unsigned*) = nullptr;
  bool ok = false;
  Nvml() {
    lib = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!lib) return;
    Init = reinterpret_cast<int (*)()>(dlsym(lib, "nvmlInit_v2"));
    Shutdown = reinterpret_cast<int (*)()>(dlsym(lib, "nvmlShutdown"));
    GetHandleByPciBusId = reinterpret_cast<int (*)(const char*, void**)>(dlsym(lib, "nvmlDeviceGetHandleByPciBusId_v2"));
    GetNvLinkState = reinterpret_cast<int (*)(void*, unsigned, unsigned*)>(dlsym(lib, "nvmlDeviceGetNvLinkState"));
    GetNvLinkVersion = reinterpret_cast<int (*)(void*, unsigned, unsigned*)>(dlsym(lib, "nvmlDeviceGetNvLinkVersion"));
    ok = Init && Shutdown && GetHandleByPciBusId && GetNvLinkState && Init() == 0;
  }
  ~Nvml() {
    if (ok) Shutdown();
    if (lib) dlclose(lib);
  }
};

void append(std::string& s, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  s += buf;
}

}  // namespace

extern "C" int fmpi_symm_mc_supported(int device);

extern "C" int fmpi_topo_json(char* buf, size_t buflen) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    (void)cudaGetLastError();
    n = 0;
  }
  int cur = 0;
  if (n) cudaGetDevice(&cur);
  Nvml nvml;
  std::string s = "{";
  append(s, "\"device_count\": %d, \"nvml\": %s, \"devices\": [", n, nvml.ok ? "true" : "false");
  for (int d = 0; d < n; ++d) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, d) != cudaSuccess) {
      (void)cudaGetLastError();
      continue;
    }
    int links = -1, version = -1;
    if (nvml.ok) {
      char bus[32];
      if (cudaDeviceGetPCIBusId(bus, sizeof(bus), d) == cudaSuccess) {
        void* h = nullptr;
        if (nvml.GetHandleByPciBusId(bus, &h) == 0) {
          links = 0;
          for (unsigned l = 0; l < 18; ++l) {
            unsigned active = 0;
            if (nvml.GetNvLinkState(h, l, &active) == 0 && active) {
              ++links;
              unsigned v = 0;
              if (nvml.GetNvLinkVersion && nvml.GetNvLinkVersion(h, l, &v) == 0) version = int(v);
            }
          }
        }
      }
    }
    append(s, "%s{\"index\": %d, \"name\": \"%s\", \"sm_count\": %d, \"cc\": \"%d.%d\", "
              "\"mem_gb\": %.1f, \"l2_mb\": %.1f, \"nvlink_active\": %d, \"nvlink_version\": %d, "
              "\"multicast\": %d}",
           d ? ", " : "", d, p.name, p.multiProcessorCount, p.major, p.minor,
           double(p.totalGlobalMem) / 1e9, double(p.l2CacheSize) / 1048576.0, links, version,
           fmpi_symm_mc_supported(d));
  }
  s += "], \"p2p\": [";
  for (int a = 0; a < n; ++a) {
    s += a ? ", [" : "[";
    for (int b = 0; b < n; ++b) {
      int can = a == b ? 1 : 0, rankp = 0, atomics = 0;
      if (a != b) {
        cudaDeviceCanAccessPeer(&can, a, b);
        cudaDeviceGetP2PAttribute(&rankp, cudaDevP2PAttrPerformanceRank, a, b);
        cudaDeviceGetP2PAttribute(&atomics, cudaDevP2PAttrNativeAtomicSupported, a, b);
        (void)cudaGetLastError();
      }
      append(s, "%s{\"access\": %d, \"perf_rank\": %d, \"atomics\": %d}", b ? ", " : "", can, rankp, atomics);
    }
    s += "]";
  }
  s += "]}";
  if (n) cudaSetDevice(cur);
  if (s.size() + 1 > buflen) return fmpi_host::set_error("topology JSON needs %zu bytes", s.size() + 1);
  memcpy(buf, s.c_str(), s.size() + 1);
  return 0;
}
