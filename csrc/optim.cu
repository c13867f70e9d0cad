// Optimizers over the flat weight arena.  One launch updates every parameter of the model —
// the reference runs torch 1.7's per-tensor Python loop (218 tensors for ResNet-34;
// ml/experiments/kubeml/function_resnet34.py:62, function_lenet.py:79-81).
// Semantics follow torch.optim.SGD / Adam (L2 weight decay folded into the gradient).
#include "common.cuh"
#include "../include/kubeml_b200.h"

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ uint2 pack_bf16x4(const float4& v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
  __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
  uint2 r;
  r.x = *reinterpret_cast<uint32_t*>(&a);
  r.y = *reinterpret_cast<uint32_t*>(&b);
  return r;
}

template <bool kMomentum, bool kNesterov, bool kBf16>
__global__ void __launch_bounds__(kThreads)
sgd_kernel(float4* __restrict__ w, const float4* __restrict__ g, float4* __restrict__ mbuf, uint2* __restrict__ wb,
           size_t n4, float lr, float wd, float mom) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 wv = w[i];
    float4 d = __ldcs(g + i);
    d.x = fmaf(wd, wv.x, d.x); d.y = fmaf(wd, wv.y, d.y); d.z = fmaf(wd, wv.z, d.z); d.w = fmaf(wd, wv.w, d.w);
    if constexpr (kMomentum) {
      float4 b = mbuf[i];
      // a zeroed buffer reproduces torch's "first step: buf = grad" (mom*0 + d)
      b.x = fmaf(mom, b.x, d.x); b.y = fmaf(mom, b.y, d.y); b.z = fmaf(mom, b.z, d.z); b.w = fmaf(mom, b.w, d.w);
      mbuf[i] = b;
      if constexpr (kNesterov) {
        d.x = fmaf(mom, b.x, d.x); d.y = fmaf(mom, b.y, d.y); d.z = fmaf(mom, b.z, d.z); d.w = fmaf(mom, b.w, d.w);
      } else {
        d = b;
      }
    }
    wv.x = fmaf(-lr, d.x, wv.x); wv.y = fmaf(-lr, d.y, wv.y); wv.z = fmaf(-lr, d.z, wv.z); wv.w = fmaf(-lr, d.w, wv.w);
    w[i] = wv;
    if constexpr (kBf16) wb[i] = pack_bf16x4(wv);
  }
}

// step_dev == nullptr: bias corrections come from the host (bc1, bc2_sqrt).  Otherwise the step count
// lives on the device ({count, ticket}): every block reads count at its start, the last block to
// finish publishes count+1 — so a CUDA-graph replay of the step advances the bias correction.
template <bool kBf16>
__global__ void __launch_bounds__(kThreads)
adam_kernel(float4* __restrict__ w, const float4* __restrict__ g, float4* __restrict__ m1, float4* __restrict__ m2,
            uint2* __restrict__ wb, size_t n4, float lr, float b1, float b2, float eps, float wd, float bc1,
            float bc2_sqrt, int* __restrict__ step_dev) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  int t = 0;
  if (step_dev) {
    t = *reinterpret_cast<volatile int*>(step_dev) + 1;
    bc1 = 1.f - powf(b1, static_cast<float>(t));
    bc2_sqrt = sqrtf(1.f - powf(b2, static_cast<float>(t)));
  }
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const float step_size = lr / bc1;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 wv = w[i], gv = __ldcs(g + i), a = m1[i], b = m2[i];
    float* wp = &wv.x; float* gp = &gv.x; float* ap = &a.x; float* bp = &b.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float gr = fmaf(wd, wp[k], gp[k]);
      ap[k] = fmaf(b1, ap[k], (1.f - b1) * gr);
      bp[k] = fmaf(b2, bp[k], (1.f - b2) * gr * gr);
      float denom = sqrtf(bp[k]) / bc2_sqrt + eps;
      wp[k] -= step_size * (ap[k] / denom);
    }
    w[i] = wv; m1[i] = a; m2[i] = b;
    if constexpr (kBf16) wb[i] = pack_bf16x4(wv);
  }
  if (step_dev) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(step_dev + 1, 1) == static_cast<int>(gridDim.x) - 1) {
        step_dev[1] = 0;
        step_dev[0] = t;
      }
    }
  }
}

__global__ void __launch_bounds__(kThreads)
cast_kernel(const float4* __restrict__ src, uint2* __restrict__ dst, size_t n4) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride)
    dst[i] = pack_bf16x4(src[i]);
}

__global__ void __launch_bounds__(kThreads)
fill_kernel(float4* __restrict__ dst, float v, size_t n4) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const float4 f = make_float4(v, v, v, v);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) dst[i] = f;
}

inline int grid_for(size_t n4) {
  size_t want = (n4 + kThreads - 1) / kThreads;
  size_t cap = static_cast<size_t>(kml::num_sms()) * 8;
  return static_cast<int>(want < cap ? (want ? want : 1) : cap);
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

extern "C" {

int kml_sgd_step(float* w, const float* g, float* mbuf, size_t n, float lr, float wd, float mom, int nesterov,
                 void* w_bf16, void* stream) {
  KML_CHECK(w && g, "kml_sgd_step: null pointer");
  KML_CHECK(n % 4 == 0 && aligned16(w) && aligned16(g), "kml_sgd_step: buffers must be 16-byte aligned, n %% 4 == 0");
  KML_CHECK(mom == 0.f || mbuf != nullptr, "kml_sgd_step: momentum without a buffer");
  if (n == 0) return 0;
  const size_t n4 = n / 4;
  auto s = static_cast<cudaStream_t>(stream);
  auto W = reinterpret_cast<float4*>(w);
  auto G = reinterpret_cast<const float4*>(g);
  auto M = reinterpret_cast<float4*>(mbuf);
  auto B = reinterpret_cast<uint2*>(w_bf16);
  const int grid = grid_for(n4);
#define LAUNCH(MOM, NES, BF) KML_CUDA(kml::launch_pdl(sgd_kernel<MOM, NES, BF>, dim3(grid), dim3(kThreads), 0, s, W, G, M, B, n4, lr, wd, mom))
  const bool m = mom != 0.f, nes = nesterov != 0, bf = w_bf16 != nullptr;
  if (m && nes && bf) LAUNCH(true, true, true);
  else if (m && nes) LAUNCH(true, true, false);
  else if (m && bf) LAUNCH(true, false, true);
  else if (m) LAUNCH(true, false, false);
  else if (bf) LAUNCH(false, false, true);
  else LAUNCH(false, false, false);
#undef LAUNCH
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

static int adam_launch(float* w, const float* g, float* m1, float* m2, size_t n, float lr, float b1, float b2,
                       float eps, float wd, float bc1, float bc2s, int* step_dev, void* w_bf16, void* stream) {
  if (n == 0) return 0;
  const size_t n4 = n / 4;
  auto s = static_cast<cudaStream_t>(stream);
  const int grid = grid_for(n4);
  if (w_bf16)
    KML_CUDA(kml::launch_pdl(adam_kernel<true>, dim3(grid), dim3(kThreads), 0, s, reinterpret_cast<float4*>(w),
                             reinterpret_cast<const float4*>(g), reinterpret_cast<float4*>(m1),
                             reinterpret_cast<float4*>(m2), reinterpret_cast<uint2*>(w_bf16), n4, lr, b1, b2, eps, wd,
                             bc1, bc2s, step_dev));
  else
    KML_CUDA(kml::launch_pdl(adam_kernel<false>, dim3(grid), dim3(kThreads), 0, s, reinterpret_cast<float4*>(w),
                             reinterpret_cast<const float4*>(g), reinterpret_cast<float4*>(m1),
                             reinterpret_cast<float4*>(m2), static_cast<uint2*>(nullptr), n4, lr, b1, b2, eps, wd, bc1,
                             bc2s, step_dev));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_adam_step(float* w, const float* g, float* m1, float* m2, size_t n, float lr, float b1, float b2, float eps,
                  float wd, int step, void* w_bf16, void* stream) {
  KML_CHECK(w && g && m1 && m2, "kml_adam_step: null pointer");
  KML_CHECK(n % 4 == 0 && aligned16(w) && aligned16(g) && aligned16(m1) && aligned16(m2),
            "kml_adam_step: buffers must be 16-byte aligned, n %% 4 == 0");
  KML_CHECK(step >= 1, "kml_adam_step: step starts at 1");
  const float bc1 = 1.f - powf(b1, static_cast<float>(step));
  const float bc2s = sqrtf(1.f - powf(b2, static_cast<float>(step)));
  return adam_launch(w, g, m1, m2, n, lr, b1, b2, eps, wd, bc1, bc2s, nullptr, w_bf16, stream);
}

int kml_adam_step_dev(float* w, const float* g, float* m1, float* m2, size_t n, float lr, float b1, float b2,
                      float eps, float wd, int* step_dev, void* w_bf16, void* stream) {
  KML_CHECK(w && g && m1 && m2 && step_dev, "kml_adam_step_dev: null pointer");
  KML_CHECK(n % 4 == 0 && aligned16(w) && aligned16(g) && aligned16(m1) && aligned16(m2),
            "kml_adam_step_dev: buffers must be 16-byte aligned, n %% 4 == 0");
  return adam_launch(w, g, m1, m2, n, lr, b1, b2, eps, wd, 1.f, 1.f, step_dev, w_bf16, stream);
}

int kml_cast_f32_to_bf16(const float* src, void* dst, size_t n, void* stream) {
  KML_CHECK(src && dst && n % 4 == 0 && aligned16(src) && (reinterpret_cast<uintptr_t>(dst) & 7u) == 0,
            "kml_cast_f32_to_bf16: bad arguments");
  if (n == 0) return 0;
  KML_CUDA(kml::launch_pdl(cast_kernel, dim3(grid_for(n / 4)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), reinterpret_cast<const float4*>(src), reinterpret_cast<uint2*>(dst), n / 4));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_fill_f32(float* dst, float v, size_t n, void* stream) {
  KML_CHECK(dst && n % 4 == 0 && aligned16(dst), "kml_fill_f32: bad arguments");
  if (n == 0) return 0;
  KML_CUDA(kml::launch_pdl(fill_kernel, dim3(grid_for(n / 4)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), reinterpret_cast<float4*>(dst), v, n / 4));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

}  // extern "C"
