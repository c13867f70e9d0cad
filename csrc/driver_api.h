// Lazily resolved CUDA driver entry points.  The library links only the static CUDA runtime, so it
// loads on a machine without a driver (CPU-side build check); driver symbols are looked up the
// first time a VMM / multicast / tensor-map call needs them.
#pragma once
#include <cuda.h>

namespace kml {
struct DriverApi {
  CUresult (*cuMemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*cuMemRelease)(CUmemGenericAllocationHandle);
  CUresult (*cuMemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*cuMemAddressFree)(CUdeviceptr, size_t);
  CUresult (*cuMemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*cuMemUnmap)(CUdeviceptr, size_t);
  CUresult (*cuMemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*cuMemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*cuMemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*cuMemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*cuMulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*cuMulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*cuMulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*cuMulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*cuDeviceGet)(CUdevice*, int);
  CUresult (*cuDeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*cuGetErrorString)(CUresult, const char**);
  CUresult (*cuCtxGetCurrent)(CUcontext*);
  CUresult (*cuTensorMapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                     CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
};
// Returns nullptr (and sets the error string) if libcuda cannot be loaded.
const DriverApi* driver();
const char* cu_err(CUresult r);
}  // namespace kml

#define KML_CU(expr)                                                                          \
  do {                                                                                        \
    CUresult _r = (expr);                                                                     \
    if (_r != CUDA_SUCCESS) {                                                                 \
      kml::set_error("%s failed: %s (%s:%d)", #expr, kml::cu_err(_r), __FILE__, __LINE__);    \
      return 3;                                                                               \
    }                                                                                         \
  } while (0)
