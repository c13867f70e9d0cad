// BatchNorm2d for NHWC bf16 activations, training and inference, with ReLU and the residual add of
// a ResNet block fused in.  The reference gets these through cuDNN/ATen inside torch 1.7.1
// (torchvision resnet34 in ml/experiments/kubeml/function_resnet34.py:101; SURVEY 2.4 "BatchNorm2d
// fwd/bwd (train: batch stats + running-stat update + num_batches_tracked++)").
//
// Training forward, one cooperative launch:
//   phase 1  every CTA sums x and x^2 per channel over its rows -> partial[cta][2][C]
//   grid.sync
//   phase 2  every CTA folds the partials in the same fixed order (deterministic, no atomics, nothing
//            to zero), normalises its rows, adds the residual, applies ReLU, stores bf16;
//            CTA 0 also stores mean / inv-std for backward, updates the running statistics
//            (momentum, unbiased variance) and bumps num_batches_tracked.
// Backward has the same shape: phase 1 sums g and g*xhat (g = dy masked by the ReLU output),
// phase 2 writes dx (and g itself for the residual branch); CTA 0 writes dgamma / dbeta.
#include "common.cuh"
#include "../include/kubeml_b200.h"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace {

constexpr int kThreads = 256;
constexpr int kVec = 8;   // channels per thread (one 16-byte bf16 vector)

struct alignas(16) Bf16x8 { __nv_bfloat162 h[4]; };

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 v = __bfloat1622float2(h[i]);
    f[2 * i] = v.x;
    f[2 * i + 1] = v.y;
  }
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&f)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = u;
}

struct BnFwdArgs {
  const __nv_bfloat16* x;       // [M][C] convolution output
  const __nv_bfloat16* res;     // [M][C] residual added before ReLU, or null
  __nv_bfloat16* y;             // [M][C]
  const float* gamma;
  const float* beta;
  float* running_mean;
  float* running_var;
  long long* num_batches_tracked;   // may be null
  float* save_mean;             // [C]
  float* save_invstd;           // [C]
  float* partial;               // [grid][2][C] scratch
  int M, C, relu;
  float eps, momentum;
};

// thread layout: groups = C/8 channel vectors, lanes = kThreads/groups rows in flight per CTA
__global__ void __launch_bounds__(kThreads) bn_fwd_train_kernel(const BnFwdArgs a) {
  cg::grid_group grid = cg::this_grid();
  extern __shared__ float sm[];              // [2][C] block sums, then [2][C] scale/shift
  const int groups = a.C / kVec;
  const int lanes = kThreads / groups;
  const int g = threadIdx.x % groups;
  const int lane = threadIdx.x / groups;
  const int c0 = g * kVec;
  const int row_stride = gridDim.x * lanes;
  const bool active = lane < lanes;

  for (int i = threadIdx.x; i < 2 * a.C; i += kThreads) sm[i] = 0.f;
  __syncthreads();
  float s[kVec], q[kVec];
#pragma unroll
  for (int i = 0; i < kVec; ++i) s[i] = q[i] = 0.f;
  if (active) {
    for (int r = blockIdx.x * lanes + lane; r < a.M; r += row_stride) {
      float f[kVec];
      load8(a.x + static_cast<size_t>(r) * a.C + c0, f);
#pragma unroll
      for (int i = 0; i < kVec; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
    }
#pragma unroll
    for (int i = 0; i < kVec; ++i) { atomicAdd(&sm[c0 + i], s[i]); atomicAdd(&sm[a.C + c0 + i], q[i]); }
  }
  __syncthreads();
  float* mine = a.partial + static_cast<size_t>(blockIdx.x) * 2 * a.C;
  for (int i = threadIdx.x; i < 2 * a.C; i += kThreads) mine[i] = sm[i];
  __threadfence();
  grid.sync();

  // fold partials: thread i owns channel i (and i + kThreads, ...)
  const float inv_m = 1.f / static_cast<float>(a.M);
  for (int c = threadIdx.x; c < a.C; c += kThreads) {
    float ts = 0.f, tq = 0.f;
    for (int b = 0; b < gridDim.x; ++b) {
      ts += __ldcg(a.partial + static_cast<size_t>(b) * 2 * a.C + c);
      tq += __ldcg(a.partial + static_cast<size_t>(b) * 2 * a.C + a.C + c);
    }
    const float mean = ts * inv_m;
    const float var = fmaxf(tq * inv_m - mean * mean, 0.f);
    const float invstd = rsqrtf(var + a.eps);
    const float sc = a.gamma[c] * invstd;
    sm[c] = sc;
    sm[a.C + c] = a.beta[c] - mean * sc;
    if (blockIdx.x == 0) {
      a.save_mean[c] = mean;
      a.save_invstd[c] = invstd;
      if (a.running_mean) {
        const float unbiased = a.M > 1 ? var * static_cast<float>(a.M) / static_cast<float>(a.M - 1) : var;
        a.running_mean[c] = (1.f - a.momentum) * a.running_mean[c] + a.momentum * mean;
        a.running_var[c] = (1.f - a.momentum) * a.running_var[c] + a.momentum * unbiased;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.num_batches_tracked) *a.num_batches_tracked += 1;
  __syncthreads();

  if (active) {
    float sc[kVec], sh[kVec];
#pragma unroll
    for (int i = 0; i < kVec; ++i) { sc[i] = sm[c0 + i]; sh[i] = sm[a.C + c0 + i]; }
    for (int r = blockIdx.x * lanes + lane; r < a.M; r += row_stride) {
      const size_t off = static_cast<size_t>(r) * a.C + c0;
      float f[kVec];
      load8(a.x + off, f);
#pragma unroll
      for (int i = 0; i < kVec; ++i) f[i] = fmaf(f[i], sc[i], sh[i]);
      if (a.res) {
        float rr[kVec];
        load8(a.res + off, rr);
#pragma unroll
        for (int i = 0; i < kVec; ++i) f[i] += rr[i];
      }
      if (a.relu) {
#pragma unroll
        for (int i = 0; i < kVec; ++i) f[i] = fmaxf(f[i], 0.f);
      }
      store8(a.y + off, f);
    }
  }
}

// Inference: running statistics, no reduction.
__global__ void __launch_bounds__(kThreads)
bn_fwd_eval_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                   __nv_bfloat16* __restrict__ y, const float* __restrict__ gamma, const float* __restrict__ beta,
                   const float* __restrict__ rmean, const float* __restrict__ rvar, int M, int C, int relu, float eps) {
  const int groups = C / kVec;
  const size_t total = static_cast<size_t>(M) * groups;
  for (size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
       idx += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(idx % groups) * kVec;
    const size_t off = (idx / groups) * C + c0;
    float f[kVec];
    load8(x + off, f);
#pragma unroll
    for (int i = 0; i < kVec; ++i) {
      const float sc = gamma[c0 + i] * rsqrtf(rvar[c0 + i] + eps);
      f[i] = fmaf(f[i] - rmean[c0 + i], sc, beta[c0 + i]);
    }
    if (res) {
      float rr[kVec];
      load8(res + off, rr);
#pragma unroll
      for (int i = 0; i < kVec; ++i) f[i] += rr[i];
    }
    if (relu) {
#pragma unroll
      for (int i = 0; i < kVec; ++i) f[i] = fmaxf(f[i], 0.f);
    }
    store8(y + off, f);
  }
}

struct BnBwdArgs {
  const __nv_bfloat16* dy;      // [M][C] gradient w.r.t. the block output
  const __nv_bfloat16* y;       // [M][C] block output (ReLU mask), or null when there was no ReLU
  const __nv_bfloat16* x;       // [M][C] convolution output saved by forward
  const float* gamma;
  const float* save_mean;
  const float* save_invstd;
  __nv_bfloat16* dx;            // [M][C]
  __nv_bfloat16* dres;          // [M][C] masked gradient for the residual branch, or null
  float* dgamma;
  float* dbeta;
  float* partial;               // [grid][2][C]
  int M, C;
};

__global__ void __launch_bounds__(kThreads) bn_bwd_kernel(const BnBwdArgs a) {
  cg::grid_group grid = cg::this_grid();
  extern __shared__ float sm[];
  const int groups = a.C / kVec;
  const int lanes = kThreads / groups;
  const int g = threadIdx.x % groups;
  const int lane = threadIdx.x / groups;
  const int c0 = g * kVec;
  const int row_stride = gridDim.x * lanes;
  const bool active = lane < lanes;

  for (int i = threadIdx.x; i < 2 * a.C; i += kThreads) sm[i] = 0.f;
  __syncthreads();
  float mean[kVec], istd[kVec];
#pragma unroll
  for (int i = 0; i < kVec; ++i) { mean[i] = a.save_mean[c0 + i]; istd[i] = a.save_invstd[c0 + i]; }
  if (active) {
    float s1[kVec], s2[kVec];
#pragma unroll
    for (int i = 0; i < kVec; ++i) s1[i] = s2[i] = 0.f;
    for (int r = blockIdx.x * lanes + lane; r < a.M; r += row_stride) {
      const size_t off = static_cast<size_t>(r) * a.C + c0;
      float gy[kVec], xv[kVec];
      load8(a.dy + off, gy);
      load8(a.x + off, xv);
      if (a.y) {
        float yv[kVec];
        load8(a.y + off, yv);
#pragma unroll
        for (int i = 0; i < kVec; ++i) gy[i] = yv[i] > 0.f ? gy[i] : 0.f;
      }
#pragma unroll
      for (int i = 0; i < kVec; ++i) {
        s1[i] += gy[i];
        s2[i] = fmaf(gy[i], (xv[i] - mean[i]) * istd[i], s2[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < kVec; ++i) { atomicAdd(&sm[c0 + i], s1[i]); atomicAdd(&sm[a.C + c0 + i], s2[i]); }
  }
  __syncthreads();
  float* mine = a.partial + static_cast<size_t>(blockIdx.x) * 2 * a.C;
  for (int i = threadIdx.x; i < 2 * a.C; i += kThreads) mine[i] = sm[i];
  __threadfence();
  grid.sync();

  for (int c = threadIdx.x; c < a.C; c += kThreads) {
    float t1 = 0.f, t2 = 0.f;
    for (int b = 0; b < gridDim.x; ++b) {
      t1 += __ldcg(a.partial + static_cast<size_t>(b) * 2 * a.C + c);
      t2 += __ldcg(a.partial + static_cast<size_t>(b) * 2 * a.C + a.C + c);
    }
    sm[c] = t1;
    sm[a.C + c] = t2;
    if (blockIdx.x == 0) { a.dbeta[c] = t1; a.dgamma[c] = t2; }
  }
  __syncthreads();

  if (active) {
    const float inv_m = 1.f / static_cast<float>(a.M);
    float k1[kVec], k2[kVec], sc[kVec];
#pragma unroll
    for (int i = 0; i < kVec; ++i) {
      sc[i] = a.gamma[c0 + i] * istd[i];
      k1[i] = sm[c0 + i] * inv_m;
      k2[i] = sm[a.C + c0 + i] * inv_m;
    }
    for (int r = blockIdx.x * lanes + lane; r < a.M; r += row_stride) {
      const size_t off = static_cast<size_t>(r) * a.C + c0;
      float gy[kVec], xv[kVec];
      load8(a.dy + off, gy);
      load8(a.x + off, xv);
      if (a.y) {
        float yv[kVec];
        load8(a.y + off, yv);
#pragma unroll
        for (int i = 0; i < kVec; ++i) gy[i] = yv[i] > 0.f ? gy[i] : 0.f;
      }
      if (a.dres) store8(a.dres + off, gy);
      float d[kVec];
#pragma unroll
      for (int i = 0; i < kVec; ++i) d[i] = sc[i] * (gy[i] - k1[i] - (xv[i] - mean[i]) * istd[i] * k2[i]);
      store8(a.dx + off, d);
    }
  }
}

int bn_grid(int M, int C, const void* kernel, size_t smem) {
  const int groups = C / kVec;
  const int lanes = kThreads / groups;
  int want = (M + lanes * 4 - 1) / (lanes * 4);          // ~4 rows per thread
  if (want < 1) want = 1;
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kThreads, smem);
  if (occ < 1) occ = 1;
  const int cap = kml::num_sms() * (occ > 2 ? 2 : occ);
  return want < cap ? want : cap;
}

}  // namespace

extern "C" {

size_t kml_bn_scratch_bytes(int C) { return static_cast<size_t>(kml::num_sms()) * 2 * 2 * C * sizeof(float); }

int kml_bn_fwd_train(const void* x, const void* res, void* y, const float* gamma, const float* beta,
                     float* running_mean, float* running_var, long long* num_batches_tracked, float* save_mean,
                     float* save_invstd, float* scratch, int M, int C, int relu, float eps, float momentum,
                     void* stream) {
  KML_CHECK(x && y && gamma && beta && save_mean && save_invstd && scratch, "kml_bn_fwd_train: null pointer");
  KML_CHECK(C % kVec == 0 && C / kVec <= kThreads && M > 0, "kml_bn_fwd_train: C (%d) must be a multiple of 8 and <= %d", C,
            kThreads * kVec);
  BnFwdArgs a{static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(res),
              static_cast<__nv_bfloat16*>(y), gamma, beta, running_mean, running_var, num_batches_tracked,
              save_mean, save_invstd, scratch, M, C, relu, eps, momentum};
  const size_t smem = 2 * C * sizeof(float);
  const int grid = bn_grid(M, C, reinterpret_cast<const void*>(bn_fwd_train_kernel), smem);
  void* args[] = {&a};
  KML_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(bn_fwd_train_kernel), dim3(grid), dim3(kThreads), args,
                                       smem, static_cast<cudaStream_t>(stream)));
  kml::count_launch();
  return 0;
}

int kml_bn_fwd_eval(const void* x, const void* res, void* y, const float* gamma, const float* beta,
                    const float* running_mean, const float* running_var, int M, int C, int relu, float eps,
                    void* stream) {
  KML_CHECK(x && y && gamma && beta && running_mean && running_var, "kml_bn_fwd_eval: null pointer");
  KML_CHECK(C % kVec == 0 && M > 0, "kml_bn_fwd_eval: C must be a multiple of 8");
  const size_t total = static_cast<size_t>(M) * (C / kVec);
  size_t want = (total + kThreads - 1) / kThreads;
  const size_t cap = static_cast<size_t>(kml::num_sms()) * 8;
  const int grid = static_cast<int>(want < cap ? want : cap);
  bn_fwd_eval_kernel<<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(res), static_cast<__nv_bfloat16*>(y),
      gamma, beta, running_mean, running_var, M, C, relu, eps);
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_bn_bwd(const void* dy, const void* y, const void* x, const float* gamma, const float* save_mean,
               const float* save_invstd, void* dx, void* dres, float* dgamma, float* dbeta, float* scratch, int M,
               int C, void* stream) {
  KML_CHECK(dy && x && gamma && save_mean && save_invstd && dx && dgamma && dbeta && scratch,
            "kml_bn_bwd: null pointer");
  KML_CHECK(C % kVec == 0 && C / kVec <= kThreads && M > 0, "kml_bn_bwd: C (%d) must be a multiple of 8 and <= %d", C,
            kThreads * kVec);
  BnBwdArgs a{static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(y),
              static_cast<const __nv_bfloat16*>(x), gamma, save_mean, save_invstd, static_cast<__nv_bfloat16*>(dx),
              static_cast<__nv_bfloat16*>(dres), dgamma, dbeta, scratch, M, C};
  const size_t smem = 2 * C * sizeof(float);
  const int grid = bn_grid(M, C, reinterpret_cast<const void*>(bn_bwd_kernel), smem);
  void* args[] = {&a};
  KML_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(bn_bwd_kernel), dim3(grid), dim3(kThreads), args, smem,
                                       static_cast<cudaStream_t>(stream)));
  kml::count_launch();
  return 0;
}

}  // extern "C"
