#include "common.cuh"
#include "driver_api.h"
#include <dlfcn.h>
#include <mutex>

namespace kml {
static DriverApi g_api;
static bool g_ok = false;
static std::once_flag g_once;

static void load() {
  void* h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libcuda.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) { set_error("cannot load libcuda.so.1: %s", dlerror()); return; }
  bool ok = true;
#define SYM(name, sym)                                                            \
  g_api.name = reinterpret_cast<decltype(g_api.name)>(dlsym(h, sym));             \
  if (!g_api.name) { set_error("libcuda lacks %s", sym); ok = false; }
  SYM(cuMemCreate, "cuMemCreate")
  SYM(cuMemRelease, "cuMemRelease")
  SYM(cuMemAddressReserve, "cuMemAddressReserve")
  SYM(cuMemAddressFree, "cuMemAddressFree")
  SYM(cuMemMap, "cuMemMap")
  SYM(cuMemUnmap, "cuMemUnmap")
  SYM(cuMemSetAccess, "cuMemSetAccess")
  SYM(cuMemGetAllocationGranularity, "cuMemGetAllocationGranularity")
  SYM(cuMemExportToShareableHandle, "cuMemExportToShareableHandle")
  SYM(cuMemImportFromShareableHandle, "cuMemImportFromShareableHandle")
  SYM(cuMulticastCreate, "cuMulticastCreate")
  SYM(cuMulticastAddDevice, "cuMulticastAddDevice")
  SYM(cuMulticastBindMem, "cuMulticastBindMem")
  SYM(cuMulticastGetGranularity, "cuMulticastGetGranularity")
  SYM(cuDeviceGet, "cuDeviceGet")
  SYM(cuDeviceGetAttribute, "cuDeviceGetAttribute")
  SYM(cuGetErrorString, "cuGetErrorString")
  SYM(cuCtxGetCurrent, "cuCtxGetCurrent")
  SYM(cuTensorMapEncodeTiled, "cuTensorMapEncodeTiled")
#undef SYM
  g_ok = ok;
}

const DriverApi* driver() {
  std::call_once(g_once, load);
  return g_ok ? &g_api : nullptr;
}

const char* cu_err(CUresult r) {
  const char* s = nullptr;
  if (g_ok && g_api.cuGetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown CUDA driver error";
}
}  // namespace kml
