// Fused softmax-cross-entropy forward + backward + accuracy count, one launch.
// Replaces F.cross_entropy -> log_softmax + nll_loss (+ autograd's two backward kernels) and the
// argmax/eq/sum().item() chain of the reference's validate()
// (ml/experiments/kubeml/function_resnet34.py:69,82-93; function_lenet.py:86-127).
//
// One warp per row; the row's loss goes to a scratch array and the last block to finish (ticket
// counter in the same scratch) adds the rows up in index order, so the mean is deterministic.
#include "common.cuh"
#include "../include/kubeml_b200.h"
#include <mutex>

namespace {

constexpr int kWarpsPerBlock = 8;

struct CeScratch {          // layout of the caller-provided scratch
  unsigned int ticket;      // self-resetting
  int correct;              // self-resetting
  unsigned int pad[2];
  // followed by float row_loss[rows]
};

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

template <typename T>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
softmax_ce_kernel(const T* __restrict__ logits, const long long* __restrict__ labels, int rows, int classes,
                  float* __restrict__ loss_out, T* __restrict__ dlogits, int* __restrict__ correct_out,
                  CeScratch* __restrict__ scratch) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  float* row_loss = reinterpret_cast<float*>(scratch + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * kWarpsPerBlock + warp;
  __shared__ bool is_last;

  if (row < rows) {
    const T* x = logits + static_cast<size_t>(row) * classes;
    const int label = static_cast<int>(labels[row]);
    float mx = -INFINITY;
    int arg = 0;
    for (int c = lane; c < classes; c += 32) {
      float v = ldf(x + c);
      if (v > mx) { mx = v; arg = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {   // argmax with lowest-index tie-break, like torch
      float om = __shfl_xor_sync(0xffffffffu, mx, o);
      int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
    }
    float sum = 0.f;
    for (int c = lane; c < classes; c += 32) sum += __expf(ldf(x + c) - mx);
    sum = kml::warp_sum(sum);
    const float lse = mx + __logf(sum);
    if (dlogits) {
      T* dx = dlogits + static_cast<size_t>(row) * classes;
      const float inv_rows = 1.f / static_cast<float>(rows);
      const float inv_sum = 1.f / sum;
      for (int c = lane; c < classes; c += 32) {
        float pr = __expf(ldf(x + c) - mx) * inv_sum;
        stf(dx + c, (pr - (c == label ? 1.f : 0.f)) * inv_rows);
      }
    }
    if (lane == 0) {
      const bool valid = label >= 0 && label < classes;
      row_loss[row] = valid ? lse - ldf(x + label) : 0.f;
      if (correct_out && arg == label) atomicAdd(&scratch->correct, 1);
    }
  }

  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(&scratch->ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // deterministic final reduction by the last block: fixed assignment, fixed tree
  float acc = 0.f;
  for (int r = threadIdx.x; r < rows; r += blockDim.x) acc += __ldcg(row_loss + r);
  acc = kml::warp_sum(acc);
  __shared__ float part[kWarpsPerBlock];
  if (lane == 0) part[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kWarpsPerBlock; ++i) t += part[i];
    *loss_out = t / static_cast<float>(rows);
    if (correct_out) *correct_out = atomicExch(&scratch->correct, 0);
    scratch->ticket = 0;   // ready for the next launch on this stream
  }
}

}  // namespace

extern "C" {

size_t kml_softmax_ce_scratch_bytes(int rows) {
  return sizeof(CeScratch) + static_cast<size_t>(rows > 0 ? rows : 0) * sizeof(float);
}

int kml_softmax_ce(const void* logits, int logits_bf16, const long long* labels, int rows, int classes, float* loss,
                   void* dlogits, int* correct, void* scratch, void* stream) {
  KML_CHECK(logits && labels && loss && scratch, "kml_softmax_ce: null pointer");
  KML_CHECK(rows > 0 && classes > 0, "kml_softmax_ce: empty problem (%d x %d)", rows, classes);
  KML_CHECK((reinterpret_cast<uintptr_t>(scratch) & 15u) == 0, "kml_softmax_ce: scratch must be 16-byte aligned");
  const int grid = (rows + kWarpsPerBlock - 1) / kWarpsPerBlock;
  auto s = static_cast<cudaStream_t>(stream);
  if (logits_bf16)
    KML_CUDA(kml::launch_pdl(softmax_ce_kernel<__nv_bfloat16>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, s, static_cast<const __nv_bfloat16*>(logits), labels, rows, classes, loss, static_cast<__nv_bfloat16*>(dlogits), correct, static_cast<CeScratch*>(scratch)));
  else
    KML_CUDA(kml::launch_pdl(softmax_ce_kernel<float>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, s, static_cast<const float*>(logits), labels, rows, classes, loss, static_cast<float*>(dlogits), correct, static_cast<CeScratch*>(scratch)));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant (see include/kubeml_b200.h: kml_softmax_ce_noalloc).
// ------------------------------------------------------------------------------------------------
}  // extern "C"

namespace {

// The scratch of kml_softmax_ce (ticket + correct counter + one float per row), but owned by the
// library instead of the caller.  Every GPU that loads this library carries these bytes until the
// process exits, which is why the row limit is kept modest.
static __device__ __align__(256) unsigned char g_ce_noalloc_scratch[sizeof(CeScratch) +
                                                                    KML_SOFTMAX_CE_NOALLOC_MAX_ROWS * sizeof(float)];

constexpr int kMaxDevices = 64;
void* g_ce_noalloc_ptr[kMaxDevices] = {};   // per device: device address of the array above
std::mutex g_ce_noalloc_mu;

// In host code `g_ce_noalloc_scratch` is only a handle for the symbol, not an address a kernel can
// use; the device address has to be asked for, per device, with that device current.
int ce_noalloc_scratch_for_current_device(void** out) {
  int dev = -1;
  KML_CUDA(cudaGetDevice(&dev));
  KML_CHECK(dev >= 0 && dev < kMaxDevices, "kml_softmax_ce_noalloc: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_ce_noalloc_mu);
  if (g_ce_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    KML_CUDA(cudaGetSymbolAddress(&p, g_ce_noalloc_scratch));
    KML_CHECK(p != nullptr, "kml_softmax_ce_noalloc: cudaGetSymbolAddress returned a null address");
    g_ce_noalloc_ptr[dev] = p;
  }
  *out = g_ce_noalloc_ptr[dev];
  return 0;
}

}  // namespace

extern "C" {

int kml_softmax_ce_noalloc_max_rows(void) { return KML_SOFTMAX_CE_NOALLOC_MAX_ROWS; }

int kml_softmax_ce_noalloc(const void* logits, int logits_bf16, const long long* labels, int rows, int classes,
                           float* loss, void* dlogits, int* correct, void* stream) {
  KML_CHECK(logits && labels && loss, "kml_softmax_ce_noalloc: null pointer");
  KML_CHECK(rows > 0 && classes > 0, "kml_softmax_ce_noalloc: empty problem (%d x %d)", rows, classes);
  KML_CHECK(rows <= KML_SOFTMAX_CE_NOALLOC_MAX_ROWS,
            "kml_softmax_ce_noalloc: %d rows exceed the built-in scratch (max %d rows); use kml_softmax_ce with "
            "caller-provided scratch for larger batches", rows, KML_SOFTMAX_CE_NOALLOC_MAX_ROWS);
  void* scratch = nullptr;
  int rc = ce_noalloc_scratch_for_current_device(&scratch);
  if (rc) return rc;
  auto s = static_cast<cudaStream_t>(stream);
  // The array is shared by every call and keeps what the previous one left behind.  Row losses are
  // fully rewritten before they are read; the ticket and the hit counter must start from zero, so
  // they are cleared here in stream order (a call that was cut short cannot poison the next one).
  KML_CUDA(cudaMemsetAsync(scratch, 0, sizeof(CeScratch), s));
  const int grid = (rows + kWarpsPerBlock - 1) / kWarpsPerBlock;
  if (logits_bf16)
    KML_CUDA(kml::launch_pdl(softmax_ce_kernel<__nv_bfloat16>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, s, static_cast<const __nv_bfloat16*>(logits), labels, rows, classes, loss, static_cast<__nv_bfloat16*>(dlogits), correct, static_cast<CeScratch*>(scratch)));
  else
    KML_CUDA(kml::launch_pdl(softmax_ce_kernel<float>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, s, static_cast<const float*>(logits), labels, rows, classes, loss, static_cast<float*>(dlogits), correct, static_cast<CeScratch*>(scratch)));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

}  // extern "C"
