// Weight arena on the CUDA virtual-memory-management API + NVSwitch multicast objects.
//
// Why VMM and not cudaMalloc + cudaIpc: a VMM allocation can be (a) exported as a POSIX file
// descriptor and mapped by the other per-GPU worker processes at any address, (b) bound to a
// multicast object so that one `multimem.st` lands in every replica and one `multimem.ld_reduce`
// returns the in-switch sum.  Both are what replaces the reference's RedisAI hop
// (python/kubeml/kubeml/network.py:416-461; ml/pkg/model/model.go:135-161,249-269).
#include "common.cuh"
#include "driver_api.h"
#include "../include/kubeml_b200.h"
#include <unistd.h>
#include <cstdlib>
#include <cstdint>

struct kml_arena {
  CUmemGenericAllocationHandle handle = 0;
  CUdeviceptr va = 0;
  size_t size = 0;
  int device = -1;
  bool imported = false;
};

struct kml_mcast {
  CUmemGenericAllocationHandle handle = 0;
  size_t size = 0;
  int num_devices = 0;
  CUdeviceptr va = 0;
  bool mapped = false;
};

namespace {

CUmemAllocationProp alloc_prop(int device, bool exportable) {
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = exportable ? CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR : CU_MEM_HANDLE_TYPE_NONE;
  return prop;
}

int map_rw(const kml::DriverApi* d, kml_arena* a, int device) {
  KML_CU(d->cuMemAddressReserve(&a->va, a->size, 0, 0, 0));
  KML_CU(d->cuMemMap(a->va, a->size, 0, a->handle, 0));
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  KML_CU(d->cuMemSetAccess(a->va, a->size, &acc, 1));
  return 0;
}

}  // namespace

extern "C" {

int kml_arena_create(int device, size_t bytes, int exportable, kml_arena_t** out) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  KML_CHECK(out && bytes > 0, "kml_arena_create: bad arguments");
  KML_CUDA(cudaSetDevice(device));
  KML_CUDA(cudaFree(0));  // make sure the primary context exists
  CUmemAllocationProp prop = alloc_prop(device, exportable != 0);
  size_t gran = 0;
  KML_CU(d->cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  // Multicast binding wants the (usually larger) multicast granularity; round to the max of both
  // when the device supports multicast so the same arena can be bound later.
  int mc = 0;
  CUdevice cudev;
  KML_CU(d->cuDeviceGet(&cudev, device));
  d->cuDeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev);
  if (mc) {
    CUmulticastObjectProp mp = {};
    mp.numDevices = 2;
    mp.size = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d->cuMulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran)
      gran = mg;
  }
  kml_arena* a = new kml_arena();
  a->size = (bytes + gran - 1) / gran * gran;
  a->device = device;
  CUresult r = d->cuMemCreate(&a->handle, a->size, &prop, 0);
  if (r != CUDA_SUCCESS) {
    kml::set_error("cuMemCreate(%zu bytes) failed: %s", a->size, kml::cu_err(r));
    delete a;
    return 3;
  }
  int rc = map_rw(d, a, device);
  if (rc) { d->cuMemRelease(a->handle); delete a; return rc; }
  KML_CUDA(cudaMemset(reinterpret_cast<void*>(a->va), 0, a->size));
  *out = a;
  return 0;
}

int kml_arena_destroy(kml_arena_t* a) {
  if (!a) return 0;
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  if (a->va) {
    d->cuMemUnmap(a->va, a->size);
    d->cuMemAddressFree(a->va, a->size);
  }
  if (a->handle) d->cuMemRelease(a->handle);
  delete a;
  return 0;
}

void* kml_arena_ptr(const kml_arena_t* a) { return a ? reinterpret_cast<void*>(a->va) : nullptr; }
size_t kml_arena_size(const kml_arena_t* a) { return a ? a->size : 0; }

int kml_arena_export_fd(kml_arena_t* a, int* fd_out) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  KML_CHECK(a && fd_out, "kml_arena_export_fd: bad arguments");
  int fd = -1;
  KML_CU(d->cuMemExportToShareableHandle(&fd, a->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd_out = fd;
  return 0;
}

int kml_arena_import_fd(int device, int fd, size_t bytes, kml_arena_t** out) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  KML_CHECK(out && fd >= 0 && bytes > 0, "kml_arena_import_fd: bad arguments");
  KML_CUDA(cudaSetDevice(device));
  KML_CUDA(cudaFree(0));
  kml_arena* a = new kml_arena();
  a->size = bytes;
  a->device = device;
  a->imported = true;
  CUresult r = d->cuMemImportFromShareableHandle(&a->handle, reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                                 CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    kml::set_error("cuMemImportFromShareableHandle failed: %s", kml::cu_err(r));
    delete a;
    return 3;
  }
  int rc = map_rw(d, a, device);
  if (rc) { d->cuMemRelease(a->handle); delete a; return rc; }
  *out = a;
  return 0;
}

int kml_arena_grant_access(kml_arena_t* a, int device) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  KML_CHECK(a, "kml_arena_grant_access: null arena");
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  KML_CU(d->cuMemSetAccess(a->va, a->size, &acc, 1));
  return 0;
}

int kml_mcast_supported(int device, int* supported) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  CUdevice cudev;
  KML_CU(d->cuDeviceGet(&cudev, device));
  int mc = 0;
  KML_CU(d->cuDeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev));
  *supported = mc;
  return 0;
}

int kml_mcast_create(int num_devices, size_t bytes, kml_mcast_t** out) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  KML_CHECK(out && num_devices >= 2, "kml_mcast_create: bad arguments");
  CUmulticastObjectProp mp = {};
  mp.numDevices = static_cast<unsigned>(num_devices);
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  kml_mcast* m = new kml_mcast();
  m->size = bytes;
  m->num_devices = num_devices;
  CUresult r = d->cuMulticastCreate(&m->handle, &mp);
  if (r != CUDA_SUCCESS) {
    kml::set_error("cuMulticastCreate failed: %s", kml::cu_err(r));
    delete m;
    return 3;
  }
  *out = m;
  return 0;
}

int kml_mcast_export_fd(kml_mcast_t* m, int* fd_out) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  int fd = -1;
  KML_CU(d->cuMemExportToShareableHandle(&fd, m->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd_out = fd;
  return 0;
}

int kml_mcast_import_fd(int fd, int num_devices, size_t bytes, kml_mcast_t** out) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  kml_mcast* m = new kml_mcast();
  m->size = bytes;
  m->num_devices = num_devices;
  CUresult r = d->cuMemImportFromShareableHandle(&m->handle, reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                                 CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    kml::set_error("multicast import failed: %s", kml::cu_err(r));
    delete m;
    return 3;
  }
  *out = m;
  return 0;
}

int kml_mcast_add_device(kml_mcast_t* m, int device) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  CUdevice cudev;
  KML_CU(d->cuDeviceGet(&cudev, device));
  KML_CU(d->cuMulticastAddDevice(m->handle, cudev));
  return 0;
}

int kml_mcast_bind(kml_mcast_t* m, int device, kml_arena_t* arena, void** mc_ptr_out) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  KML_CHECK(m && arena && mc_ptr_out, "kml_mcast_bind: bad arguments");
  KML_CHECK(arena->size >= m->size, "arena (%zu) smaller than multicast object (%zu)", arena->size, m->size);
  KML_CU(d->cuMulticastBindMem(m->handle, 0, arena->handle, 0, m->size, 0));
  if (!m->mapped) {
    KML_CU(d->cuMemAddressReserve(&m->va, m->size, 0, 0, 0));
    KML_CU(d->cuMemMap(m->va, m->size, 0, m->handle, 0));
    CUmemAccessDesc acc = {};
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = device;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    KML_CU(d->cuMemSetAccess(m->va, m->size, &acc, 1));
    m->mapped = true;
  }
  *mc_ptr_out = reinterpret_cast<void*>(m->va);
  return 0;
}

int kml_mcast_destroy(kml_mcast_t* m) {
  if (!m) return 0;
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  if (m->mapped) {
    d->cuMemUnmap(m->va, m->size);
    d->cuMemAddressFree(m->va, m->size);
  }
  if (m->handle) d->cuMemRelease(m->handle);
  delete m;
  return 0;
}

}  // extern "C"
