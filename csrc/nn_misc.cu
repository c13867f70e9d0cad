// Small NHWC bf16 layers and data-movement kernels around the tcgen05 GEMMs:
//   max-pool (any k/stride/pad) forward with saved arg-max + backward as a gather,
//   global average pool forward/backward,
//   im2col / col2im for the convolutions the shifted-TMA path cannot express (stride 2, Cin < 64,
//   non power-of-two grids), column sums (bias gradients), padded fp32->bf16 packing, uint8 image
//   normalisation.  The reference reaches all of these through ATen/cuDNN.
#include <algorithm>
#include "common.cuh"
#include "../include/kubeml_b200.h"

namespace {

constexpr int kThreads = 256;

inline int grid_for(size_t work, int per_block = kThreads) {
  size_t want = (work + per_block - 1) / per_block;
  const size_t cap = static_cast<size_t>(kml::num_sms()) * 16;
  if (want < 1) want = 1;
  return static_cast<int>(want < cap ? want : cap);
}

// ---------------------------------------------------------------------------------- max pool
__global__ void __launch_bounds__(kThreads)
maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, unsigned char* __restrict__ idx,
                   int N, int H, int W, int C, int Ho, int Wo, int k, int stride, int pad) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t total = static_cast<size_t>(N) * Ho * Wo * C;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    size_t p = i / C;
    const int wo = static_cast<int>(p % Wo); p /= Wo;
    const int ho = static_cast<int>(p % Ho);
    const int n = static_cast<int>(p / Ho);
    float best = -INFINITY;
    int arg = 0;
    for (int kh = 0; kh < k; ++kh) {
      const int h = ho * stride - pad + kh;
      if (h < 0 || h >= H) continue;
      for (int kw = 0; kw < k; ++kw) {
        const int w = wo * stride - pad + kw;
        if (w < 0 || w >= W) continue;
        const float v = __bfloat162float(x[((static_cast<size_t>(n) * H + h) * W + w) * C + c]);
        if (v > best) { best = v; arg = kh * k + kw; }      // first maximum wins, like ATen
      }
    }
    y[i] = __float2bfloat16_rn(best);
    if (idx) idx[i] = static_cast<unsigned char>(arg);
  }
}

// dx[n,h,w,c] = sum of dy over the windows whose saved arg-max is (h,w): a gather, no atomics
__global__ void __launch_bounds__(kThreads)
maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const unsigned char* __restrict__ idx,
                   __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int Ho, int Wo, int k, int stride,
                   int pad) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t total = static_cast<size_t>(N) * H * W * C;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    size_t p = i / C;
    const int w = static_cast<int>(p % W); p /= W;
    const int h = static_cast<int>(p % H);
    const int n = static_cast<int>(p / H);
    float acc = 0.f;
    for (int kh = 0; kh < k; ++kh) {
      const int hn = h + pad - kh;
      if (hn < 0 || hn % stride) continue;
      const int ho = hn / stride;
      if (ho >= Ho) continue;
      for (int kw = 0; kw < k; ++kw) {
        const int wn = w + pad - kw;
        if (wn < 0 || wn % stride) continue;
        const int wo = wn / stride;
        if (wo >= Wo) continue;
        const size_t o = ((static_cast<size_t>(n) * Ho + ho) * Wo + wo) * C + c;
        if (idx[o] == kh * k + kw) acc += __bfloat162float(dy[o]);
      }
    }
    dx[i] = __float2bfloat16_rn(acc);
  }
}

// ---------------------------------------------------------------------------------- avg pool
__global__ void __launch_bounds__(kThreads)
avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int HW, int C) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t total = static_cast<size_t>(N) * C;
  const float inv = 1.f / static_cast<float>(HW);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    const size_t n = i / C;
    float acc = 0.f;
    for (int p = 0; p < HW; ++p) acc += __bfloat162float(x[(n * HW + p) * C + c]);
    y[i] = __float2bfloat16_rn(acc * inv);
  }
}
__global__ void __launch_bounds__(kThreads)
avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, int N, int HW, int C) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t total = static_cast<size_t>(N) * HW * C;
  const float inv = 1.f / static_cast<float>(HW);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    const size_t n = i / (static_cast<size_t>(HW) * C);
    dx[i] = __float2bfloat16_rn(__bfloat162float(dy[n * C + c]) * inv);
  }
}

// ---------------------------------------------------------------------------------- im2col / col2im
// cols[m][(kh*KW + kw)*C + c], m = (n, ho, wo); row length ldk >= KH*KW*C, tail zero-filled.
__global__ void __launch_bounds__(kThreads)
im2col_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ cols, int N, int H, int W, int C, int Ho,
              int Wo, int KH, int KW, int stride, int pad, int ldk) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t total = static_cast<size_t>(N) * Ho * Wo * ldk;
  const int K = KH * KW * C;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int kk = static_cast<int>(i % ldk);
    size_t m = i / ldk;
    __nv_bfloat16 v = __float2bfloat16_rn(0.f);
    if (kk < K) {
      const int c = kk % C;
      const int tap = kk / C;
      const int kw = tap % KW, kh = tap / KW;
      const int wo = static_cast<int>(m % Wo); m /= Wo;
      const int ho = static_cast<int>(m % Ho);
      const int n = static_cast<int>(m / Ho);
      const int h = ho * stride - pad + kh, w = wo * stride - pad + kw;
      if (h >= 0 && h < H && w >= 0 && w < W) v = x[((static_cast<size_t>(n) * H + h) * W + w) * C + c];
    }
    cols[i] = v;
  }
}

// dx[n,h,w,c] = sum over taps of dcols[m(n,ho,wo)][tap*C + c] (+ addend), a gather over <= KH*KW terms
__global__ void __launch_bounds__(kThreads)
col2im_kernel(const __nv_bfloat16* __restrict__ dcols, const __nv_bfloat16* __restrict__ addend,
              __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int Ho, int Wo, int KH, int KW, int stride,
              int pad, int ldk) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t total = static_cast<size_t>(N) * H * W * C;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    size_t p = i / C;
    const int w = static_cast<int>(p % W); p /= W;
    const int h = static_cast<int>(p % H);
    const int n = static_cast<int>(p / H);
    float acc = addend ? __bfloat162float(addend[i]) : 0.f;
    for (int kh = 0; kh < KH; ++kh) {
      const int hn = h + pad - kh;
      if (hn < 0 || hn % stride) continue;
      const int ho = hn / stride;
      if (ho >= Ho) continue;
      for (int kw = 0; kw < KW; ++kw) {
        const int wn = w + pad - kw;
        if (wn < 0 || wn % stride) continue;
        const int wo = wn / stride;
        if (wo >= Wo) continue;
        const size_t m = (static_cast<size_t>(n) * Ho + ho) * Wo + wo;
        acc += __bfloat162float(dcols[m * ldk + (kh * KW + kw) * C + c]);
      }
    }
    dx[i] = __float2bfloat16_rn(acc);
  }
}


// ---------------------------------------------------------------------------------- 8-channel vector variants
__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float (&f)[8]) {
  const uint4 u = __ldg(reinterpret_cast<const uint4*>(p));
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { const float2 v = __bfloat1622float2(h[i]); f[2 * i] = v.x; f[2 * i + 1] = v.y; }
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float (&f)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = u;
}

__global__ void __launch_bounds__(kThreads)
maxpool_fwd_vec_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, unsigned char* __restrict__ idx,
                       int N, int H, int W, int C, int Ho, int Wo, int k, int stride, int pad) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const int C8 = C / 8;
  const size_t total = static_cast<size_t>(N) * Ho * Wo * C8;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C8) * 8;
    size_t p = i / C8;
    const int wo = static_cast<int>(p % Wo); p /= Wo;
    const int ho = static_cast<int>(p % Ho);
    const int n = static_cast<int>(p / Ho);
    float best[8];
    int arg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; arg[j] = 0; }
    for (int kh = 0; kh < k; ++kh) {
      const int h = ho * stride - pad + kh;
      if (h < 0 || h >= H) continue;
      for (int kw = 0; kw < k; ++kw) {
        const int w = wo * stride - pad + kw;
        if (w < 0 || w >= W) continue;
        float v[8];
        ld8(x + ((static_cast<size_t>(n) * H + h) * W + w) * C + c, v);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (v[j] > best[j]) { best[j] = v[j]; arg[j] = kh * k + kw; }
      }
    }
    const size_t o = (i / C8) * C + c;
    st8(y + o, best);
    if (idx) {
      uint2 u;
      u.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
      u.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
      *reinterpret_cast<uint2*>(idx + o) = u;
    }
  }
}

__global__ void __launch_bounds__(kThreads)
maxpool_bwd_vec_kernel(const __nv_bfloat16* __restrict__ dy, const unsigned char* __restrict__ idx,
                       __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int Ho, int Wo, int k, int stride,
                       int pad) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const int C8 = C / 8;
  const size_t total = static_cast<size_t>(N) * H * W * C8;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C8) * 8;
    size_t p = i / C8;
    const int w = static_cast<int>(p % W); p /= W;
    const int h = static_cast<int>(p % H);
    const int n = static_cast<int>(p / H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int kh = 0; kh < k; ++kh) {
      const int hn = h + pad - kh;
      if (hn < 0 || hn % stride) continue;
      const int ho = hn / stride;
      if (ho >= Ho) continue;
      for (int kw = 0; kw < k; ++kw) {
        const int wn = w + pad - kw;
        if (wn < 0 || wn % stride) continue;
        const int wo = wn / stride;
        if (wo >= Wo) continue;
        const size_t o = ((static_cast<size_t>(n) * Ho + ho) * Wo + wo) * C + c;
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(idx + o));
        float g[8];
        ld8(dy + o, g);
        const int tap = kh * k + kw;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int a = ((j < 4 ? u.x : u.y) >> (8 * (j & 3))) & 0xff;
          if (a == tap) acc[j] += g[j];
        }
      }
    }
    st8(dx + (i / C8) * C + c, acc);
  }
}

__global__ void __launch_bounds__(kThreads)
im2col_vec_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ cols, int N, int H, int W, int C,
                  int Ho, int Wo, int KH, int KW, int stride, int pad, int ldk) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const int L8 = ldk / 8, C8 = C / 8, K8 = KH * KW * C8;
  const size_t total = static_cast<size_t>(N) * Ho * Wo * L8;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int k8 = static_cast<int>(i % L8);
    size_t m = i / L8;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (k8 < K8) {
      const int c = (k8 % C8) * 8;
      const int tap = k8 / C8;
      const int kw = tap % KW, kh = tap / KW;
      const int wo = static_cast<int>(m % Wo); m /= Wo;
      const int ho = static_cast<int>(m % Ho);
      const int n = static_cast<int>(m / Ho);
      const int h = ho * stride - pad + kh, w = wo * stride - pad + kw;
      if (h >= 0 && h < H && w >= 0 && w < W)
        v = __ldg(reinterpret_cast<const uint4*>(x + ((static_cast<size_t>(n) * H + h) * W + w) * C + c));
    }
    *reinterpret_cast<uint4*>(cols + i * 8) = v;
  }
}

// any C (the 3-channel stem): one thread gathers 8 consecutive columns and stores them as one vector
__global__ void __launch_bounds__(kThreads)
im2col_gather8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ cols, int N, int H, int W, int C,
                      int Ho, int Wo, int KH, int KW, int stride, int pad, int ldk) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const int L8 = ldk / 8, K = KH * KW * C;
  const size_t total = static_cast<size_t>(N) * Ho * Wo * L8;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int k0 = static_cast<int>(i % L8) * 8;
    size_t m = i / L8;
    const int wo = static_cast<int>(m % Wo); m /= Wo;
    const int ho = static_cast<int>(m % Ho);
    const int n = static_cast<int>(m / Ho);
    const __nv_bfloat16* img = x + static_cast<size_t>(n) * H * W * C;
    __align__(16) __nv_bfloat16 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int kk = k0 + j;
      __nv_bfloat16 e = __float2bfloat16_rn(0.f);
      if (kk < K) {
        const int c = kk % C, tap = kk / C;
        const int h = ho * stride - pad + tap / KW, w = wo * stride - pad + tap % KW;
        if (h >= 0 && h < H && w >= 0 && w < W) e = img[(static_cast<size_t>(h) * W + w) * C + c];
      }
      v[j] = e;
    }
    *reinterpret_cast<uint4*>(cols + i * 8) = *reinterpret_cast<const uint4*>(v);
  }
}

__global__ void __launch_bounds__(kThreads)
col2im_vec_kernel(const __nv_bfloat16* __restrict__ dcols, const __nv_bfloat16* __restrict__ addend,
                  __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int Ho, int Wo, int KH, int KW, int stride,
                  int pad, int ldk) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const int C8 = C / 8;
  const size_t total = static_cast<size_t>(N) * H * W * C8;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C8) * 8;
    size_t p = i / C8;
    const int w = static_cast<int>(p % W); p /= W;
    const int h = static_cast<int>(p % H);
    const int n = static_cast<int>(p / H);
    const size_t o = (i / C8) * C + c;
    float acc[8];
    if (addend) ld8(addend + o, acc);
    else {
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    }
    for (int kh = 0; kh < KH; ++kh) {
      const int hn = h + pad - kh;
      if (hn < 0 || hn % stride) continue;
      const int ho = hn / stride;
      if (ho >= Ho) continue;
      for (int kw = 0; kw < KW; ++kw) {
        const int wn = w + pad - kw;
        if (wn < 0 || wn % stride) continue;
        const int wo = wn / stride;
        if (wo >= Wo) continue;
        const size_t m = (static_cast<size_t>(n) * Ho + ho) * Wo + wo;
        float g[8];
        ld8(dcols + m * ldk + (kh * KW + kw) * C + c, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += g[j];
      }
    }
    st8(dx + o, acc);
  }
}

// ---------------------------------------------------------------------------------- misc
// Column pass over a row-major [M][C] matrix, 8 columns (16 B) per thread, rows split over the grid:
//   v = dy (bf16) or the sum of S fp32 split-K slabs;  a = (y ? (y > 0 ? v * scale : 0) : v);
//   g (optional) <- a;   colsum (optional, zeroed by the launcher) += sum over rows of a.
// One kernel serves the bias gradient (colsum only), the ReLU(+dropout) gradient, both at once, and
// folds the slabs of a split-K data-gradient GEMM on the way (no reduction pass, no atomics there).
template <bool kSlabs>
__global__ void __launch_bounds__(kThreads)
colpass_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ dys, int S, size_t slab_stride,
               const __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ g, float scale,
               float* __restrict__ colsum, int M, int C, long long ld, int rows_per_block) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  __shared__ float red[kThreads][9];
  const int cg = C / 8;
  const int cgw = cg < kThreads ? cg : kThreads;           // column groups handled by this block
  const int rpi = kThreads / cgw;                          // rows per iteration
  const int lane_row = threadIdx.x / cgw;
  const int cgi = blockIdx.y * kThreads + threadIdx.x % cgw;
  const bool active = lane_row < rpi && cgi < cg;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(M, r0 + rows_per_block);
  if (active) {
    int r = r0 + lane_row;
    if constexpr (!kSlabs) {
      // four rows in flight per thread: the pass is latency-bound otherwise (2 loads, 1 store per row)
      for (; r + 3 * rpi < r1; r += 4 * rpi) {
        float v[4][8], m[4][8];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const size_t off = static_cast<size_t>(r + u * rpi) * ld + static_cast<size_t>(cgi) * 8;
          ld8(dy + off, v[u]);
          if (y) ld8(y + off, m[u]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const size_t off = static_cast<size_t>(r + u * rpi) * ld + static_cast<size_t>(cgi) * 8;
          if (y) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[u][k] = m[u][k] > 0.f ? v[u][k] * scale : 0.f;
          }
          if (g) st8(g + off, v[u]);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += v[u][k];
        }
      }
    }
    for (; r < r1; r += rpi) {
      const size_t off = static_cast<size_t>(r) * ld + static_cast<size_t>(cgi) * 8;
      float v[8];
      if constexpr (kSlabs) {
        const float4* p = reinterpret_cast<const float4*>(dys + off);
        float4 a = p[0], b = p[1];
        for (int s2 = 1; s2 < S; ++s2) {
          const float4* q = reinterpret_cast<const float4*>(dys + s2 * slab_stride + off);
          const float4 c = q[0], d = q[1];
          a.x += c.x; a.y += c.y; a.z += c.z; a.w += c.w;
          b.x += d.x; b.y += d.y; b.z += d.z; b.w += d.w;
        }
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
      } else {
        ld8(dy + off, v);
      }
      if (y) {
        float m[8];
        ld8(y + off, m);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = m[k] > 0.f ? v[k] * scale : 0.f;
      }
      if (g) st8(g + off, v);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += v[k];
    }
  }
  if (colsum == nullptr) return;
#pragma unroll
  for (int k = 0; k < 8; ++k) red[threadIdx.x][k] = acc[k];
  __syncthreads();
  if (active && lane_row == 0) {
    for (int l = 1; l < rpi; ++l)
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += red[threadIdx.x + l * cgw][k];
#pragma unroll
    for (int k = 0; k < 8; ++k) atomicAdd(colsum + static_cast<size_t>(cgi) * 8 + k, acc[k]);
  }
}

__global__ void zero_f32_kernel(float* __restrict__ p, int n) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = 0.f;
}

// scalar fallback (columns not a multiple of 8): one block per 32 columns
__global__ void __launch_bounds__(kThreads)
colsum_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, int rows, int cols, long long ld) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  __shared__ float part[kThreads / 32][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float acc = 0.f;
  if (c < cols)
    for (int r = warp; r < rows; r += kThreads / 32) acc += __bfloat162float(x[static_cast<size_t>(r) * ld + c]);
  part[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) t += part[w][lane];
    out[c] = t;
  }
}

// y = relu?(sum of S fp32 split-K slabs + bias) as bf16: the epilogue a split GEMM cannot apply itself
__global__ void __launch_bounds__(kThreads)
bias_act_kernel(const float* __restrict__ xs, int S, size_t slab_stride, const float* __restrict__ bias,
                __nv_bfloat16* __restrict__ y, int relu, size_t M, int C, long long ld_x, long long ld_y) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const int cg = C / 8;
  const size_t total = M * cg;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t r = i / cg;
    const int c = static_cast<int>(i % cg) * 8;
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = bias ? bias[c + k] : 0.f;
    for (int s2 = 0; s2 < S; ++s2) {
      const float4* q = reinterpret_cast<const float4*>(xs + s2 * slab_stride + r * ld_x + c);
      const float4 a = q[0], b = q[1];
      v[0] += a.x; v[1] += a.y; v[2] += a.z; v[3] += a.w; v[4] += b.x; v[5] += b.y; v[6] += b.z; v[7] += b.w;
    }
    if (relu) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
    }
    st8(y + r * ld_y + c, v);
  }
}

// dst bf16 [rows][ld] <- src fp32 [rows][cols], zero padded
__global__ void __launch_bounds__(kThreads)
pack_rows_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int rows, int cols, int ld) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t total = static_cast<size_t>(rows) * ld;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % ld);
    const size_t r = i / ld;
    dst[i] = __float2bfloat16_rn(c < cols ? src[r * cols + c] : 0.f);
  }
}

// uint8 HWC image batch -> normalised bf16 NHWC, optional per-sample horizontal flip and crop offset
// (RandomHorizontalFlip + RandomCrop(size, padding) of the reference's CIFAR pipeline,
// ml/experiments/kubeml/function_resnet34.py:17-30, done on whole batches on the GPU)
__global__ void __launch_bounds__(kThreads)
image_prep_kernel(const unsigned char* __restrict__ x, __nv_bfloat16* __restrict__ y, const int* __restrict__ aug, int N,
                  int H, int W, int C, int pad, float m0, float m1, float m2, float s0, float s1, float s2) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t total = static_cast<size_t>(N) * H * W * C;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    size_t p = i / C;
    int w = static_cast<int>(p % W); p /= W;
    int h = static_cast<int>(p % H);
    const int n = static_cast<int>(p / H);
    float v = 0.f;
    bool inside = true;
    if (aug) {
      const int a = aug[n];                 // bit 0: flip, bits 8..15: dy, bits 16..23: dx (0..2*pad)
      h += ((a >> 8) & 0xff) - pad;
      w += ((a >> 16) & 0xff) - pad;
      inside = h >= 0 && h < H && w >= 0 && w < W;
      if (a & 1) w = W - 1 - w;
    }
    const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2);
    const float istd = c == 0 ? s0 : (c == 1 ? s1 : s2);
    if (inside) v = static_cast<float>(x[((static_cast<size_t>(n) * H + h) * W + w) * C + c]) * (1.f / 255.f);
    // padding pixels are zeros *before* normalisation in the reference pipeline (crop pads the PIL image)
    y[i] = __float2bfloat16_rn((v - mean) * istd);
  }
}


// The same with the flip / crop offsets drawn inside the kernel: sample n of call number `ctr` gets a hash of
// (seed, ctr, n).  state = {calls so far, finished-block ticket} lives on the device and is advanced by the last
// block to finish, so the kernel can sit in a replayed CUDA graph and no host-side random tensors (and the
// handful of tiny kernels that build them) are needed per step.
__global__ void __launch_bounds__(kThreads)
image_prep_rand_kernel(const unsigned char* __restrict__ x, __nv_bfloat16* __restrict__ y, unsigned long long* state,
                       unsigned seed, int N, int H, int W, int C, int pad, float m0, float m1, float m2, float s0, float s1,
                       float s2, const long long* __restrict__ lab_in, long long* __restrict__ lab_out) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  // the batch's labels ride along into their fixed buffer (the captured step reads both from fixed addresses):
  // no separate device-to-device copy between two training steps
  if (lab_out != nullptr && blockIdx.x == 0)
    for (int i = threadIdx.x; i < N; i += blockDim.x) lab_out[i] = lab_in[i];
  const unsigned ctr = static_cast<unsigned>(*reinterpret_cast<volatile unsigned long long*>(state));
  const unsigned span = 2u * pad + 1u;
  const size_t total = static_cast<size_t>(N) * H * W * C;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    size_t p = i / C;
    int w = static_cast<int>(p % W); p /= W;
    int h = static_cast<int>(p % H);
    const int n = static_cast<int>(p / H);
    unsigned r = (static_cast<unsigned>(n) * 0x9E3779B1u) ^ (ctr * 0x85EBCA77u) ^ seed;
    r ^= r >> 16; r *= 0x7feb352du; r ^= r >> 15; r *= 0x846ca68bu; r ^= r >> 16;
    h += static_cast<int>((r >> 8) % span) - pad;
    w += static_cast<int>((r >> 20) % span) - pad;
    const bool inside = h >= 0 && h < H && w >= 0 && w < W;
    if (r & 1u) w = W - 1 - w;
    const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2);
    const float istd = c == 0 ? s0 : (c == 1 ? s1 : s2);
    float v = 0.f;
    if (inside) v = static_cast<float>(x[((static_cast<size_t>(n) * H + h) * W + w) * C + c]) * (1.f / 255.f);
    y[i] = __float2bfloat16_rn((v - mean) * istd);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // every block read `ctr` at its start; the last one to finish moves the call counter on
    const unsigned ticket = atomicInc(reinterpret_cast<unsigned*>(state + 1), gridDim.x - 1);
    if (ticket == gridDim.x - 1) *state = static_cast<unsigned long long>(ctr) + 1ull;
  }
}

// ---------------------------------------------------------------------------------- VGG-style layers
// g = scale * dy * (y > 0): gradient through a ReLU that was fused into the producing GEMM's epilogue;
// with y the *dropped* output and scale = 1/(1-p) it is also the gradient through ReLU -> Dropout
// (a kept, positive activation is exactly a positive dropped output).
__global__ void __launch_bounds__(kThreads)
relu_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ g,
                float scale, size_t n8) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float a[8], b[8];
    ld8(dy + i * 8, a);
    ld8(y + i * 8, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = b[j] > 0.f ? a[j] * scale : 0.f;
    st8(g + i * 8, a);
  }
}

__device__ __forceinline__ uint32_t mix32(uint64_t z) {        // splitmix64 finaliser
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return static_cast<uint32_t>((z ^ (z >> 31)) >> 32);
}

// Inverted dropout.  The keep-mask is drawn in the kernel from (device-resident step counter, layer id,
// element index), stored as bytes for the backward pass; the counter is bumped on the device so a
// CUDA-graph replay draws a fresh mask every step.
__global__ void __launch_bounds__(kThreads)
dropout_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, unsigned char* __restrict__ mask,
                   unsigned long long* __restrict__ counter, unsigned int layer, float p, size_t n) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const unsigned long long step = *counter;
  const float scale = 1.f / (1.f - p);
  const uint32_t thresh = static_cast<uint32_t>(p * 4294967296.0);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const uint32_t r = mix32((step << 40) ^ (static_cast<unsigned long long>(layer) << 34) ^ i);
    const bool keep = r >= thresh;
    if (mask) mask[i] = keep ? 1 : 0;
    y[i] = __float2bfloat16_rn(keep ? __bfloat162float(x[i]) * scale : 0.f);
  }
}
__global__ void __launch_bounds__(kThreads)
dropout_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const unsigned char* __restrict__ mask,
                   __nv_bfloat16* __restrict__ dx, float p, size_t n) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const float scale = 1.f / (1.f - p);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    dx[i] = __float2bfloat16_rn(mask[i] ? __bfloat162float(dy[i]) * scale : 0.f);
}
__global__ void bump_counter_kernel(unsigned long long* counter) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  if (threadIdx.x == 0 && blockIdx.x == 0) *counter += 1;
}

// AdaptiveAvgPool2d((R,R)) on a 1x1 feature map is a broadcast: y[n][c*R2 + r] = x[n][c]
// (torchvision vgg on 32x32 inputs), and its gradient a sum over the R2 copies.
__global__ void __launch_bounds__(kThreads)
replicate_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int C, int R2) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t total = static_cast<size_t>(N) * C * R2;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    y[i] = x[i / R2];
}
__global__ void __launch_bounds__(kThreads)
replicate_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, int N, int C, int R2) {
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const size_t total = static_cast<size_t>(N) * C;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float acc = 0.f;
    for (int r = 0; r < R2; ++r) acc += __bfloat162float(dy[i * R2 + r]);
    dx[i] = __float2bfloat16_rn(acc);
  }
}

}  // namespace

extern "C" {

int kml_relu_bwd(const void* dy, const void* y, void* g, float scale, size_t n, void* stream) {
  KML_CHECK(dy && y && g && n % 8 == 0, "kml_relu_bwd: bad arguments (n must be a multiple of 8)");
  KML_CUDA(kml::launch_pdl(relu_bwd_kernel, dim3(grid_for(n / 8)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream),
                           static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(y),
                           static_cast<__nv_bfloat16*>(g), scale, n / 8));
  kml::count_launch();
  return 0;
}

int kml_dropout_fwd(const void* x, void* y, void* mask_u8, unsigned long long* counter, unsigned int layer, float p,
                    size_t n, void* stream) {
  KML_CHECK(x && y && counter && p >= 0.f && p < 1.f, "kml_dropout_fwd: bad arguments");
  KML_CUDA(kml::launch_pdl(dropout_fwd_kernel, dim3(grid_for(n)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream),
                           static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y),
                           static_cast<unsigned char*>(mask_u8), counter, layer, p, n));
  kml::count_launch();
  return 0;
}

int kml_dropout_bwd(const void* dy, const void* mask_u8, void* dx, float p, size_t n, void* stream) {
  KML_CHECK(dy && mask_u8 && dx, "kml_dropout_bwd: null pointer");
  KML_CUDA(kml::launch_pdl(dropout_bwd_kernel, dim3(grid_for(n)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream),
                           static_cast<const __nv_bfloat16*>(dy), static_cast<const unsigned char*>(mask_u8),
                           static_cast<__nv_bfloat16*>(dx), p, n));
  kml::count_launch();
  return 0;
}

int kml_bump_counter(unsigned long long* counter, void* stream) {
  KML_CHECK(counter, "kml_bump_counter: null pointer");
  KML_CUDA(kml::launch_pdl(bump_counter_kernel, dim3(1), dim3(32), 0, static_cast<cudaStream_t>(stream), counter));
  kml::count_launch();
  return 0;
}

int kml_replicate_fwd(const void* x, void* y, int N, int C, int R2, void* stream) {
  KML_CHECK(x && y && R2 > 0, "kml_replicate_fwd: bad arguments");
  KML_CUDA(kml::launch_pdl(replicate_fwd_kernel, dim3(grid_for(static_cast<size_t>(N) * C * R2)), dim3(kThreads), 0,
                           static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x),
                           static_cast<__nv_bfloat16*>(y), N, C, R2));
  kml::count_launch();
  return 0;
}

int kml_replicate_bwd(const void* dy, void* dx, int N, int C, int R2, void* stream) {
  KML_CHECK(dy && dx && R2 > 0, "kml_replicate_bwd: bad arguments");
  KML_CUDA(kml::launch_pdl(replicate_bwd_kernel, dim3(grid_for(static_cast<size_t>(N) * C)), dim3(kThreads), 0,
                           static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(dy),
                           static_cast<__nv_bfloat16*>(dx), N, C, R2));
  kml::count_launch();
  return 0;
}

int kml_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride, int pad,
                    void* stream) {
  KML_CHECK(x && y && k * k <= 255, "kml_maxpool_fwd: bad arguments");
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const size_t total = static_cast<size_t>(N) * Ho * Wo * C;
  if (C % 8 == 0)
    KML_CUDA(kml::launch_pdl(maxpool_fwd_vec_kernel, dim3(grid_for(total / 8)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<unsigned char*>(idx), N, H, W, C, Ho, Wo, k, stride, pad));
  else
  KML_CUDA(kml::launch_pdl(maxpool_fwd_kernel, dim3(grid_for(total)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<unsigned char*>(idx), N, H, W, C, Ho, Wo, k, stride, pad));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_maxpool_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int k, int stride, int pad,
                    void* stream) {
  KML_CHECK(dy && idx && dx, "kml_maxpool_bwd: null pointer");
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const size_t total = static_cast<size_t>(N) * H * W * C;
  if (C % 8 == 0)
    KML_CUDA(kml::launch_pdl(maxpool_bwd_vec_kernel, dim3(grid_for(total / 8)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(dy), static_cast<const unsigned char*>(idx), static_cast<__nv_bfloat16*>(dx), N, H, W, C, Ho, Wo, k, stride, pad));
  else
  KML_CUDA(kml::launch_pdl(maxpool_bwd_kernel, dim3(grid_for(total)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(dy), static_cast<const unsigned char*>(idx), static_cast<__nv_bfloat16*>(dx), N, H, W, C, Ho, Wo, k, stride, pad));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream) {
  KML_CHECK(x && y && HW > 0, "kml_avgpool_fwd: bad arguments");
  KML_CUDA(kml::launch_pdl(avgpool_fwd_kernel, dim3(grid_for(static_cast<size_t>(N) * C)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), N, HW, C));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream) {
  KML_CHECK(dy && dx && HW > 0, "kml_avgpool_bwd: bad arguments");
  KML_CUDA(kml::launch_pdl(avgpool_bwd_kernel, dim3(grid_for(static_cast<size_t>(N) * HW * C)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), N, HW, C));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_im2col(const void* x, void* cols, int N, int H, int W, int C, int KH, int KW, int stride, int pad, int ldk,
               void* stream) {
  KML_CHECK(x && cols && ldk >= KH * KW * C, "kml_im2col: bad arguments");
  const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
  const size_t total = static_cast<size_t>(N) * Ho * Wo * ldk;
  if (C % 8 == 0 && ldk % 8 == 0)
    KML_CUDA(kml::launch_pdl(im2col_vec_kernel, dim3(grid_for(total / 8)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(cols), N, H, W, C, Ho, Wo, KH, KW, stride, pad, ldk));
  else if (ldk % 8 == 0)
    KML_CUDA(kml::launch_pdl(im2col_gather8_kernel, dim3(grid_for(total / 8)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(cols), N, H, W, C, Ho, Wo, KH, KW, stride, pad, ldk));
  else
  KML_CUDA(kml::launch_pdl(im2col_kernel, dim3(grid_for(total)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(cols), N, H, W, C, Ho, Wo, KH, KW, stride, pad, ldk));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_col2im(const void* dcols, const void* addend, void* dx, int N, int H, int W, int C, int KH, int KW, int stride,
               int pad, int ldk, void* stream) {
  KML_CHECK(dcols && dx && ldk >= KH * KW * C, "kml_col2im: bad arguments");
  const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
  const size_t total = static_cast<size_t>(N) * H * W * C;
  if (C % 8 == 0 && ldk % 8 == 0)
    KML_CUDA(kml::launch_pdl(col2im_vec_kernel, dim3(grid_for(total / 8)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(dcols), static_cast<const __nv_bfloat16*>(addend), static_cast<__nv_bfloat16*>(dx), N, H, W, C, Ho, Wo, KH, KW, stride, pad, ldk));
  else
  KML_CUDA(kml::launch_pdl(col2im_kernel, dim3(grid_for(total)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(dcols), static_cast<const __nv_bfloat16*>(addend), static_cast<__nv_bfloat16*>(dx), N, H, W, C, Ho, Wo, KH, KW, stride, pad, ldk));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

static int colpass_launch(const void* dy, const float* dys, int S, size_t slab_stride, const void* y, void* g, float scale,
                          float* colsum, int M, int C, long long ld, cudaStream_t s) {
  const int cg = C / 8;
  const int cgw = cg < kThreads ? cg : kThreads;
  const int rpi = kThreads / cgw;
  const int gy = (cg + kThreads - 1) / kThreads;
  // enough blocks to fill the GPU, at least 4 iterations of rows each
  int want = std::max(1, (kml::num_sms() * 2) / gy);      // fewer, longer blocks: fewer atomics per column
  int rows_per_block = std::max(rpi * 4, (M + want - 1) / want);
  rows_per_block = (rows_per_block + rpi - 1) / rpi * rpi;
  const int gx = (M + rows_per_block - 1) / rows_per_block;
  if (colsum) {
    KML_CUDA(kml::launch_pdl(zero_f32_kernel, dim3((C + 255) / 256), dim3(256), 0, s, colsum, C));
    kml::count_launch();
  }
  if (dys)
    KML_CUDA(kml::launch_pdl(colpass_kernel<true>, dim3(gx, gy), dim3(kThreads), 0, s,
                             static_cast<const __nv_bfloat16*>(nullptr), dys, S, slab_stride,
                             static_cast<const __nv_bfloat16*>(y), static_cast<__nv_bfloat16*>(g), scale, colsum, M, C, ld,
                             rows_per_block));
  else
    KML_CUDA(kml::launch_pdl(colpass_kernel<false>, dim3(gx, gy), dim3(kThreads), 0, s,
                             static_cast<const __nv_bfloat16*>(dy), static_cast<const float*>(nullptr), 0, size_t{0},
                             static_cast<const __nv_bfloat16*>(y), static_cast<__nv_bfloat16*>(g), scale, colsum, M, C, ld,
                             rows_per_block));
  kml::count_launch();
  return 0;
}

int kml_colsum_bf16(const void* x, float* out, int rows, int cols, long long ld, void* stream) {
  KML_CHECK(x && out && rows > 0 && cols > 0, "kml_colsum_bf16: bad arguments");
  auto s = static_cast<cudaStream_t>(stream);
  if (cols % 8 == 0 && ld % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15u) == 0)
    return colpass_launch(x, nullptr, 0, 0, nullptr, nullptr, 1.f, out, rows, cols, ld, s);
  KML_CUDA(kml::launch_pdl(colsum_kernel, dim3((cols + 31) / 32), dim3(kThreads), 0, s,
                           static_cast<const __nv_bfloat16*>(x), out, rows, cols, ld));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_relu_bwd_cols(const void* dy, const float* dy_slabs, int nslabs, size_t slab_stride, const void* y, void* g,
                      float scale, float* dbias, int M, int C, void* stream) {
  KML_CHECK((dy != nullptr) != (dy_slabs != nullptr && nslabs > 0), "kml_relu_bwd_cols: give dy or dy_slabs");
  KML_CHECK(y && g && M > 0 && C > 0 && C % 8 == 0, "kml_relu_bwd_cols: bad arguments (C must be a multiple of 8)");
  return colpass_launch(dy, dy_slabs, nslabs, slab_stride, y, g, scale, dbias, M, C, C, static_cast<cudaStream_t>(stream));
}

int kml_bias_act(const float* x_slabs, int nslabs, size_t slab_stride, const float* bias, void* y, int relu, int M,
                 int C, long long ld_x, long long ld_y, void* stream) {
  KML_CHECK(x_slabs && y && nslabs > 0 && M > 0 && C > 0 && C % 8 == 0 && ld_x % 4 == 0 && ld_y % 8 == 0,
            "kml_bias_act: bad arguments");
  KML_CUDA(kml::launch_pdl(bias_act_kernel, dim3(grid_for(static_cast<size_t>(M) * (C / 8))), dim3(kThreads), 0,
                           static_cast<cudaStream_t>(stream), x_slabs, nslabs, slab_stride, bias,
                           static_cast<__nv_bfloat16*>(y), relu, static_cast<size_t>(M), C, ld_x, ld_y));
  kml::count_launch();
  return 0;
}

// Writes the device's global nanosecond timer into slot[idx]: a timeline of a captured training step without a
// profiler (bench/step_timeline.py).  A regular launch: it orders behind everything before it on the stream.
__global__ void stamp_kernel(unsigned long long* slots, int idx) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  slots[idx] = t;
}
int kml_stamp(unsigned long long* slots, int idx, void* stream) {
  KML_CHECK(slots != nullptr && idx >= 0, "kml_stamp: bad arguments");
  stamp_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(slots, idx);
  KML_CUDA(cudaGetLastError());
  return 0;
}

// A step's scalar result (its loss) goes to page-locked host memory with a plain store from a one-thread kernel
// instead of a copy-engine transfer: a 4-byte cudaMemcpyAsync in the middle of a compute stream costs two engine
// hand-overs (compute -> copy -> compute, several microseconds each) between consecutive training steps.
__global__ void post_f32_kernel(const float* __restrict__ src, float* dst_host, int n) {
  for (int i = threadIdx.x; i < n; i += blockDim.x) dst_host[i] = src[i];
  __threadfence_system();
}
int kml_post_f32(const float* src, float* dst_host, int n, void* stream) {
  KML_CHECK(src != nullptr && dst_host != nullptr && n > 0, "kml_post_f32: bad arguments");
  void* dev_view = nullptr;          // identical to dst_host under unified addressing; asked for, not assumed
  KML_CUDA(cudaHostGetDevicePointer(&dev_view, dst_host, 0));
  post_f32_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(src, static_cast<float*>(dev_view), n);
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_pack_rows_bf16(const float* src, void* dst, int rows, int cols, int ld, void* stream) {
  KML_CHECK(src && dst && ld >= cols, "kml_pack_rows_bf16: bad arguments");
  KML_CUDA(kml::launch_pdl(pack_rows_kernel, dim3(grid_for(static_cast<size_t>(rows) * ld)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), src, static_cast<__nv_bfloat16*>(dst), rows, cols, ld));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_image_prep(const void* x_u8, void* y_bf16, const int* aug, int N, int H, int W, int C, int pad,
                   const float* mean3, const float* std3, void* stream) {
  KML_CHECK(x_u8 && y_bf16 && mean3 && std3 && C <= 3, "kml_image_prep: bad arguments");
  const size_t total = static_cast<size_t>(N) * H * W * C;
  KML_CUDA(kml::launch_pdl(image_prep_kernel, dim3(grid_for(total)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const unsigned char*>(x_u8), static_cast<__nv_bfloat16*>(y_bf16), aug, N, H, W, C, pad, mean3[0], mean3[1], mean3[2], 1.f / std3[0], 1.f / std3[1], 1.f / std3[2]));
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_image_prep_rand(const void* x_u8, void* y_bf16, unsigned long long* state, unsigned seed, int N, int H, int W,
                        int C, int pad, const float* mean3, const float* std3, const long long* labels_in,
                        long long* labels_out, void* stream) {
  KML_CHECK(x_u8 && y_bf16 && state && mean3 && std3 && C <= 3 && pad >= 0, "kml_image_prep_rand: bad arguments");
  KML_CHECK((labels_in == nullptr) == (labels_out == nullptr), "kml_image_prep_rand: labels_in and labels_out go together");
  const size_t total = static_cast<size_t>(N) * H * W * C;
  KML_CUDA(kml::launch_pdl(image_prep_rand_kernel, dim3(grid_for(total)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream),
                           static_cast<const unsigned char*>(x_u8), static_cast<__nv_bfloat16*>(y_bf16), state, seed, N, H, W, C,
                           pad, mean3[0], mean3[1], mean3[2], 1.f / std3[0], 1.f / std3[1], 1.f / std3[2], labels_in,
                           labels_out));
  kml::count_launch();
  return 0;
}

}  // extern "C"
