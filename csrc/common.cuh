// Shared helpers for the kubeml_b200 native runtime (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <atomic>
#include <cstdlib>

namespace kml {

void set_error(const char* fmt, ...);
extern std::atomic<unsigned long long> g_launches;

inline void count_launch(unsigned long long n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define KML_CUDA(expr)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      kml::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,   \
                     __LINE__);                                                          \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)

#define KML_CHECK(cond, ...)                  \
  do {                                        \
    if (!(cond)) {                            \
      kml::set_error(__VA_ARGS__);            \
      return 2;                               \
    }                                         \
  } while (0)

inline int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev;
}

// Function attributes (dynamic shared-memory limit, cluster size, carve-out) belong to the *current
// device's* copy of a kernel.  One process may drive several GPUs (the in-process Cluster runs one
// worker thread per GPU), so "configured" is one bit per device ordinal, not one flag per process.
// Setting an attribute twice is harmless, so a lost race between two threads only repeats the call.
struct PerDeviceOnce {
  std::atomic<unsigned long long> done{0};
  bool need(int dev) const { return dev < 0 || dev >= 64 || !((done.load(std::memory_order_acquire) >> dev) & 1ull); }
  void mark(int dev) { if (dev >= 0 && dev < 64) done.fetch_or(1ull << dev, std::memory_order_release); }
};

inline int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// ---- programmatic dependent launch (PDL) -------------------------------------------------------
// Kernels launched with launch_pdl() may start while the previous kernel in the stream is still
// running: everything before pdl_wait() (barrier init, TMEM allocation, descriptor prefetch, smem
// zeroing) overlaps the predecessor's tail; pdl_wait() returns once the predecessor has completed
// and its writes are visible, so no global memory may be touched before it.  Each kernel calls
// pdl_launch_dependents() at its start so that its own successor can be scheduled early.  A kernel
// launched normally after a PDL kernel simply waits for full completion, so mixing is safe.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

inline bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("KUBEML_PDL");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

// Every kernel of this library asks for the same shared-memory carve-out (maximum shared memory).
// Consecutive kernels with different carve-outs cannot overlap on an SM and force a drain between
// them; the training step alternates 100-130 KB GEMM CTAs with tiny element-wise CTAs ~200 times.
void note_kernel(const void* fn);

// `cluster`: thread-block cluster shape (all ones = no cluster attribute).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                      cudaStream_t stream, dim3 cluster, Args&&... args) {
  note_kernel(reinterpret_cast<const void*>(kernel));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.numAttrs = 1;
  if (cluster.x * cluster.y * cluster.z > 1) {
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = cluster.x;
    attr[1].val.clusterDim.y = cluster.y;
    attr[1].val.clusterDim.z = cluster.z;
    cfg.numAttrs = 2;
  }
  cfg.attrs = attr;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  return launch_pdl_cluster(kernel, grid, block, smem, stream, dim3(1, 1, 1), static_cast<Args&&>(args)...);
}

}  // namespace kml
