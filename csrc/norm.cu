// BatchNorm2d for NHWC activations, training and inference, with ReLU and the residual add of a
// ResNet block fused in.  The reference gets these through cuDNN/ATen inside torch 1.7.1
// (torchvision resnet34 in ml/experiments/kubeml/function_resnet34.py:101; SURVEY 2.4 "BatchNorm2d
// fwd/bwd (train: batch stats + running-stat update + num_batches_tracked++)").
//
// The layers are tiny (64 K .. 2 M elements) and sit between latency-bound GEMMs, so the design goal
// is the fewest dependent memory round trips:
//   * one launch per direction with a software grid barrier (not a cooperative launch, see grid_barrier); each thread owns <= 4 rows x 8 channels which it keeps
//     in registers across the grid barrier, so the tensor is read exactly once;
//   * the input may be the split-K *slabs* of the producing GEMM (fp32 [S][M][C]): the partial sums
//     are folded here, on the fly, instead of by a reduction kernel or atomics in the GEMM;
//   * per-channel statistics are combined across CTAs with red.add into a double-buffered
//     accumulator: launch k uses buffer k&1 and clears buffer (k+1)&1 for its successor, the launch
//     counter lives in device memory, so nothing has to be zeroed by the host and the kernel is
//     safe to replay from a CUDA graph.
// Forward (training): y = relu?(gamma*(x-mean)*invstd + beta + res?), saves mean/invstd, x (bf16)
// when it came as slabs, updates running stats (momentum, unbiased variance) and num_batches_tracked.
// Backward: g = dy * (y > 0); dbeta = sum g; dgamma = sum g*xhat;
//           dx = gamma*invstd*(g - dbeta/M - xhat*dgamma/M); optionally stores g for the residual branch.
#include "common.cuh"
#include "bn_scratch.cuh"
#include "../include/kubeml_b200.h"

namespace {

constexpr int kThreads = 256;
constexpr int kVec = 8;        // channels per thread (one 16-byte bf16 vector)
constexpr int kRows = 4;       // rows a thread keeps in registers
constexpr int kMaxC = kml::kBnMaxC;

using kml::BnScratch;

// Grid-wide barrier for a *regular* launch (see csrc/bn_scratch.cuh for the co-residency argument and the
// bounded spin).  A cooperative launch costs ~7 us on this part (measured: 9.3 us vs 1.9 us for an 8-row
// problem), which is more than the whole layer.
__device__ __forceinline__ void grid_barrier(BnScratch* sc, unsigned base) {
  __syncthreads();
  if (threadIdx.x == 0) kml::grid_arrive_and_wait(sc, base, gridDim.x);
  __syncthreads();
}

__device__ __forceinline__ void cvt8(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 v = __bfloat1622float2(h[i]);
    f[2 * i] = v.x;
    f[2 * i + 1] = v.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
  cvt8(__ldcg(reinterpret_cast<const uint4*>(p)), f);
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&f)[8]) {
  *reinterpret_cast<uint4*>(p) = pack8(f);
}
// 8 channels of one row from either a bf16 tensor or the sum of S fp32 slabs
__device__ __forceinline__ void load_row(const __nv_bfloat16* xb, const float* xs, int nslabs, size_t slab_stride,
                                         size_t off, float (&f)[8]) {
  if (xs == nullptr) {
    load8(xb + off, f);
    return;
  }
  float4 a = __ldcg(reinterpret_cast<const float4*>(xs + off));
  float4 b = __ldcg(reinterpret_cast<const float4*>(xs + off + 4));
  for (int s = 1; s < nslabs; ++s) {
    const float4 c = __ldcg(reinterpret_cast<const float4*>(xs + s * slab_stride + off));
    const float4 d = __ldcg(reinterpret_cast<const float4*>(xs + s * slab_stride + off + 4));
    a.x += c.x; a.y += c.y; a.z += c.z; a.w += c.w;
    b.x += d.x; b.y += d.y; b.z += d.z; b.w += d.w;
  }
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

// Block-level then grid-level combination of two per-channel sums.  On return sm[0..2C) holds the
// grid totals.  `sm` needs 2C floats.
// `state` = {launches, bar_base} read by the caller at kernel entry, so the round trip overlaps the data
// loads (nobody can bump them before every CTA has passed the barrier below).
__device__ __forceinline__ void combine_stats(BnScratch* sc, uint2 state, int C, int c0, bool active,
                                              const float (&s1)[kVec], const float (&s2)[kVec], float* sm,
                                              bool wide_smem) {
  const unsigned par = state.x & 1u;
  const unsigned base = state.y;
  const int groups = C / kVec;
  // Shared-memory float atomics compile to compare-and-swap loops, so avoid them where the layout
  // allows: with a power-of-two group count <= 32 the lanes of a warp that share a channel group fold
  // with shuffles and every warp covers ALL channel groups, so each warp owns one row of a
  // [warps][2C] table (wide_smem) that is then summed; otherwise fall back to atomics.
  // (More than 32 groups: `wpg` consecutive warps cover the groups between them, so warps w and
  // w + wpg write the same entries and take different rows.)
  const bool tabled = wide_smem && groups <= kThreads && (groups & (groups - 1)) == 0;
  const int wpg = groups > 32 ? groups / 32 : 1;
  const int nrows = (kThreads / 32) / wpg;
  float r1[kVec], r2[kVec];
#pragma unroll
  for (int i = 0; i < kVec; ++i) { r1[i] = active ? s1[i] : 0.f; r2[i] = active ? s2[i] : 0.f; }
  bool writer = active;
  if (groups < 32 && (groups & (groups - 1)) == 0) {
    for (int off = groups; off < 32; off <<= 1) {
#pragma unroll
      for (int i = 0; i < kVec; ++i) {
        r1[i] += __shfl_xor_sync(0xffffffffu, r1[i], off);
        r2[i] += __shfl_xor_sync(0xffffffffu, r2[i], off);
      }
    }
    writer = (threadIdx.x & 31) < groups;
  }
  if (tabled) {
    float* row = sm + 2 * C + ((threadIdx.x >> 5) / wpg) * 2 * C;     // table starts after the totals
    if (writer) {
#pragma unroll
      for (int i = 0; i < kVec; ++i) { row[c0 + i] = r1[i]; row[C + c0 + i] = r2[i]; }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * C; i += kThreads) {
      float t = 0.f;
      for (int w = 0; w < nrows; ++w) t += sm[2 * C + w * 2 * C + i];
      sm[i] = t;
    }
  } else {
    for (int i = threadIdx.x; i < 2 * C; i += kThreads) sm[i] = 0.f;
    __syncthreads();
    if (writer) {
#pragma unroll
      for (int i = 0; i < kVec; ++i) { atomicAdd(&sm[c0 + i], r1[i]); atomicAdd(&sm[C + c0 + i], r2[i]); }
    }
  }
  __syncthreads();
  float* acc = sc->acc[par];
  for (int i = threadIdx.x; i < 2 * C; i += kThreads) atomicAdd(acc + i, sm[i]);
  if (blockIdx.x == gridDim.x - 1) {               // clear the buffer of the *next* launch: all of it,
    float4* nxt = reinterpret_cast<float4*>(sc->acc[par ^ 1u]);   // the next layer may be wider than this one
    for (int i = threadIdx.x; i < 2 * kMaxC / 4; i += kThreads) nxt[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  grid_barrier(sc, base);
  for (int i = threadIdx.x; i < 2 * C; i += kThreads) sm[i] = __ldcg(acc + i);
  if (blockIdx.x == 0 && threadIdx.x == 0) {       // everyone has arrived: publish the next launch's state
    sc->launches += 1;
    sc->bar_base = base + gridDim.x;
  }
  __syncthreads();
}

struct BnFwdArgs {
  const __nv_bfloat16* x;       // [M][C] convolution output (bf16), or null when xs is given
  const float* xs;              // [S][M][C] split-K slabs of the convolution output, or null
  int nslabs;
  long long slab_stride;
  __nv_bfloat16* x_out;         // [M][C] bf16 copy of the folded input (needed by backward) or null
  const __nv_bfloat16* res;     // [M][C] residual added before ReLU, or null
  __nv_bfloat16* y;             // [M][C]
  const float* gamma;
  const float* beta;
  float* running_mean;
  float* running_var;
  long long* num_batches_tracked;
  float* save_mean;
  float* save_invstd;
  BnScratch* scratch;
  int M, C, relu;
  float eps, momentum;
  int wide_smem;                // dynamic smem holds [2C] totals + [warps][2C] table
};

__global__ void __launch_bounds__(kThreads) bn_fwd_train_kernel(const BnFwdArgs a) {
  extern __shared__ float sm[];                    // [2C] totals, then reused as scale/shift
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const uint2 bn_state = __ldcg(reinterpret_cast<const uint2*>(&a.scratch->launches));
  const int groups = a.C / kVec;
  const int lanes = kThreads / groups;
  const int g = threadIdx.x % groups;
  const int lane = threadIdx.x / groups;
  const int c0 = g * kVec;
  const int row_stride = gridDim.x * lanes;
  const int row0 = blockIdx.x * lanes + lane;
  const bool active = lane < lanes;
  const size_t ss = static_cast<size_t>(a.slab_stride);

  float v[kRows][kVec];
  float s[kVec], q[kVec];
  // Shifted sums: (x - pivot) and (x - pivot)^2 with pivot = running mean, so that E[x^2] - E[x]^2 does not
  // cancel catastrophically when |mean| >> std and M is large (the sums are combined with fp32 atomics in
  // arbitrary order).  Read here, before the grid barrier; the running statistics are updated after it.
  float pv[kVec];
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    s[i] = q[i] = 0.f;
    pv[i] = (a.running_mean != nullptr && active) ? a.running_mean[c0 + i] : 0.f;
  }
  if (active) {
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      const int row = row0 + r * row_stride;
      if (row < a.M) load_row(a.x, a.xs, a.nslabs, ss, static_cast<size_t>(row) * a.C + c0, v[r]);
    }
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      const int row = row0 + r * row_stride;
      if (row < a.M) {
        if (a.x_out) {
          const uint4 u = pack8(v[r]);
          *reinterpret_cast<uint4*>(a.x_out + static_cast<size_t>(row) * a.C + c0) = u;
          cvt8(u, v[r]);                           // statistics of exactly what backward will read
        }
#pragma unroll
        for (int i = 0; i < kVec; ++i) { const float d = v[r][i] - pv[i]; s[i] += d; q[i] = fmaf(d, d, q[i]); }
      }
    }
    for (int row = row0 + kRows * row_stride; row < a.M; row += row_stride) {   // rare: very large M
      float f[kVec];
      load_row(a.x, a.xs, a.nslabs, ss, static_cast<size_t>(row) * a.C + c0, f);
      if (a.x_out) { const uint4 u = pack8(f); *reinterpret_cast<uint4*>(a.x_out + static_cast<size_t>(row) * a.C + c0) = u; cvt8(u, f); }
#pragma unroll
      for (int i = 0; i < kVec; ++i) { const float d = f[i] - pv[i]; s[i] += d; q[i] = fmaf(d, d, q[i]); }
    }
  }
  combine_stats(a.scratch, bn_state, a.C, c0, active, s, q, sm, a.wide_smem != 0);

  const float inv_m = 1.f / static_cast<float>(a.M);
  float sc[kVec], sh[kVec];
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    const int c = c0 + i;
    const float dm = sm[c] * inv_m;                      // mean - pivot
    const float mean = pv[i] + dm;
    const float var = fmaxf(sm[a.C + c] * inv_m - dm * dm, 0.f);
    const float invstd = rsqrtf(var + a.eps);
    sc[i] = a.gamma[c] * invstd;
    sh[i] = a.beta[c] - mean * sc[i];
    if (blockIdx.x == 0 && lane == 0) {
      a.save_mean[c] = mean;
      a.save_invstd[c] = invstd;
      if (a.running_mean) {
        const float unbiased = a.M > 1 ? var * static_cast<float>(a.M) / static_cast<float>(a.M - 1) : var;
        a.running_mean[c] = (1.f - a.momentum) * pv[i] + a.momentum * mean;
        a.running_var[c] = (1.f - a.momentum) * a.running_var[c] + a.momentum * unbiased;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.num_batches_tracked) *a.num_batches_tracked += 1;

  if (!active) return;
  auto finish = [&](int row, float (&f)[kVec]) {
    const size_t off = static_cast<size_t>(row) * a.C + c0;
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
  };
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    const int row = row0 + r * row_stride;
    if (row < a.M) finish(row, v[r]);
  }
  const __nv_bfloat16* xb = a.x_out ? a.x_out : a.x;
  for (int row = row0 + kRows * row_stride; row < a.M; row += row_stride) {
    float f[kVec];
    load8(xb + static_cast<size_t>(row) * a.C + c0, f);
    finish(row, f);
  }
}

// Inference: running statistics, no reduction.
__global__ void __launch_bounds__(kThreads)
bn_fwd_eval_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ xs, int nslabs, long long slab_stride,
                   const __nv_bfloat16* __restrict__ res, __nv_bfloat16* __restrict__ y, const float* __restrict__ gamma,
                   const float* __restrict__ beta, const float* __restrict__ rmean, const float* __restrict__ rvar, int M,
                   int C, int relu, float eps) {
  const int groups = C / kVec;
  const size_t total = static_cast<size_t>(M) * groups;
  for (size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
       idx += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(idx % groups) * kVec;
    const size_t off = (idx / groups) * C + c0;
    float f[kVec];
    load_row(x, xs, nslabs, static_cast<size_t>(slab_stride), off, f);
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
  const __nv_bfloat16* dy;      // [M][C] gradient w.r.t. the block output (bf16), or null when dys is given
  const float* dys;             // [S][M][C] split-K slabs of that gradient, or null
  int nslabs;
  long long slab_stride;
  const __nv_bfloat16* dy_add;  // [M][C] added to the folded slabs (residual-branch gradient), or null
  const __nv_bfloat16* y;       // [M][C] block output (ReLU mask), or null when there was no ReLU
  const __nv_bfloat16* x;       // [M][C] convolution output saved by forward
  const float* gamma;
  const float* save_mean;
  const float* save_invstd;
  __nv_bfloat16* dx;            // [M][C]
  __nv_bfloat16* dres;          // [M][C] masked gradient for the residual branch, or null
  float* dgamma;
  float* dbeta;
  BnScratch* scratch;
  int M, C;
  int wide_smem;
};

__global__ void __launch_bounds__(kThreads) bn_bwd_kernel(const BnBwdArgs a) {
  extern __shared__ float sm[];
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const uint2 bn_state = __ldcg(reinterpret_cast<const uint2*>(&a.scratch->launches));
  const int groups = a.C / kVec;
  const int lanes = kThreads / groups;
  const int g = threadIdx.x % groups;
  const int lane = threadIdx.x / groups;
  const int c0 = g * kVec;
  const int row_stride = gridDim.x * lanes;
  const int row0 = blockIdx.x * lanes + lane;
  const bool active = lane < lanes;
  const size_t ss = static_cast<size_t>(a.slab_stride);

  float mean[kVec], istd[kVec];
#pragma unroll
  for (int i = 0; i < kVec; ++i) { mean[i] = a.save_mean[c0 + i]; istd[i] = a.save_invstd[c0 + i]; }

  // masked gradient and normalised input of one row
  auto fetch = [&](int row, float (&gy)[kVec], float (&xh)[kVec]) {
    const size_t off = static_cast<size_t>(row) * a.C + c0;
    load_row(a.dy, a.dys, a.nslabs, ss, off, gy);
    if (a.dy_add) {
      float ad[kVec];
      load8(a.dy_add + off, ad);
#pragma unroll
      for (int i = 0; i < kVec; ++i) gy[i] += ad[i];
    }
    load8(a.x + off, xh);
    if (a.y) {
      float yv[kVec];
      load8(a.y + off, yv);
#pragma unroll
      for (int i = 0; i < kVec; ++i) gy[i] = yv[i] > 0.f ? gy[i] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < kVec; ++i) xh[i] = (xh[i] - mean[i]) * istd[i];
  };

  float gy[kRows][kVec], xh[kRows][kVec];
  float s1[kVec], s2[kVec];
#pragma unroll
  for (int i = 0; i < kVec; ++i) s1[i] = s2[i] = 0.f;
  if (active) {
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      const int row = row0 + r * row_stride;
      if (row < a.M) {
        fetch(row, gy[r], xh[r]);
#pragma unroll
        for (int i = 0; i < kVec; ++i) { s1[i] += gy[r][i]; s2[i] = fmaf(gy[r][i], xh[r][i], s2[i]); }
      }
    }
    for (int row = row0 + kRows * row_stride; row < a.M; row += row_stride) {
      float g2[kVec], x2[kVec];
      fetch(row, g2, x2);
#pragma unroll
      for (int i = 0; i < kVec; ++i) { s1[i] += g2[i]; s2[i] = fmaf(g2[i], x2[i], s2[i]); }
    }
  }
  combine_stats(a.scratch, bn_state, a.C, c0, active, s1, s2, sm, a.wide_smem != 0);

  if (blockIdx.x == 0 && lane == 0) {
#pragma unroll
    for (int i = 0; i < kVec; ++i) { a.dbeta[c0 + i] = sm[c0 + i]; a.dgamma[c0 + i] = sm[a.C + c0 + i]; }
  }
  if (!active) return;
  const float inv_m = 1.f / static_cast<float>(a.M);
  float k1[kVec], k2[kVec], sc[kVec];
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    sc[i] = a.gamma[c0 + i] * istd[i];
    k1[i] = sm[c0 + i] * inv_m;
    k2[i] = sm[a.C + c0 + i] * inv_m;
  }
  auto finish = [&](int row, const float (&g2)[kVec], const float (&x2)[kVec]) {
    const size_t off = static_cast<size_t>(row) * a.C + c0;
    if (a.dres) store8(a.dres + off, g2);
    float d[kVec];
#pragma unroll
    for (int i = 0; i < kVec; ++i) d[i] = sc[i] * (g2[i] - k1[i] - x2[i] * k2[i]);
    store8(a.dx + off, d);
  };
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    const int row = row0 + r * row_stride;
    if (row < a.M) finish(row, gy[r], xh[r]);
  }
  for (int row = row0 + kRows * row_stride; row < a.M; row += row_stride) {
    float g2[kVec], x2[kVec];
    fetch(row, g2, x2);
    finish(row, g2, x2);
  }
}

// Forward with statistics that the producing GEMM already accumulated (cluster split-K epilogue,
// csrc/gemm_tcgen05.cu): no reduction and no grid barrier are left, the layer is an element-wise
// pass.  stats = fp32 [2][C] (sum, sum of squares) followed by one ticket word; the last CTA to have
// read the statistics clears them for the next training step (CUDA-graph replay safe).
struct BnApplyArgs {
  const __nv_bfloat16* x;
  float* stats;
  const __nv_bfloat16* res;
  __nv_bfloat16* y;
  const float* gamma;
  const float* beta;
  float* running_mean;
  float* running_var;
  long long* num_batches_tracked;
  float* save_mean;
  float* save_invstd;
  int M, C, relu;
  float eps, momentum;
};

__global__ void __launch_bounds__(kThreads) bn_apply_train_kernel(const BnApplyArgs a) {
  extern __shared__ float sm[];                    // [C] scale, [C] shift
  __shared__ int is_last;
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const float inv_m = 1.f / static_cast<float>(a.M);
  for (int c = threadIdx.x; c < a.C; c += kThreads) {
    const float mean = __ldcg(a.stats + c) * inv_m;
    const float var = fmaxf(__ldcg(a.stats + a.C + c) * inv_m - mean * mean, 0.f);
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

  const int groups = a.C / kVec;
  const size_t total = static_cast<size_t>(a.M) * groups;
  for (size_t idx = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x; idx < total;
       idx += static_cast<size_t>(gridDim.x) * kThreads) {
    const int c0 = static_cast<int>(idx % groups) * kVec;
    const size_t off = (idx / groups) * a.C + c0;
    float f[kVec];
    load8(a.x + off, f);
#pragma unroll
    for (int i = 0; i < kVec; ++i) f[i] = fmaf(f[i], sm[c0 + i], sm[a.C + c0 + i]);
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
  // Off the critical path: the last CTA to get here (all have read the statistics long ago) clears
  // them for the next step.  atomicInc wraps to 0, so the ticket resets itself.
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned int* ticket = reinterpret_cast<unsigned int*>(a.stats + 2 * a.C);
    is_last = atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1;
  }
  __syncthreads();
  if (is_last)
    for (int c = threadIdx.x; c < 2 * a.C; c += kThreads) a.stats[c] = 0.f;
}

// Backward from the sums the producing data-gradient GEMM accumulated: g is already masked.
__global__ void __launch_bounds__(kThreads)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ x,
                    const float* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ istd,
                    float* __restrict__ stats, __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma,
                    float* __restrict__ dbeta, int M, int C) {
  extern __shared__ float sm[];                    // [C] each: k1, k2, scale, mean, invstd
  __shared__ int is_last;
  kml::pdl_launch_dependents();
  kml::pdl_wait();
  const float inv_m = 1.f / static_cast<float>(M);
  for (int c = threadIdx.x; c < C; c += kThreads) {
    const float s1 = __ldcg(stats + c), s2 = __ldcg(stats + C + c);
    const float is = istd[c];
    sm[c] = s1 * inv_m;
    sm[C + c] = s2 * inv_m;
    sm[2 * C + c] = gamma[c] * is;
    sm[3 * C + c] = mean[c];
    sm[4 * C + c] = is;
    if (blockIdx.x == 0) { dbeta[c] = s1; dgamma[c] = s2; }
  }
  __syncthreads();
  const int groups = C / kVec;
  const size_t total = static_cast<size_t>(M) * groups;
  for (size_t idx = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x; idx < total;
       idx += static_cast<size_t>(gridDim.x) * kThreads) {
    const int c0 = static_cast<int>(idx % groups) * kVec;
    const size_t off = (idx / groups) * C + c0;
    float gv[kVec], xv[kVec], d[kVec];
    load8(g + off, gv);
    load8(x + off, xv);
#pragma unroll
    for (int i = 0; i < kVec; ++i) {
      const float xh = (xv[i] - sm[3 * C + c0 + i]) * sm[4 * C + c0 + i];
      d[i] = sm[2 * C + c0 + i] * (gv[i] - sm[c0 + i] - xh * sm[C + c0 + i]);
    }
    store8(dx + off, d);
  }
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned int* ticket = reinterpret_cast<unsigned int*>(stats + 2 * C);
    is_last = atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1;
  }
  __syncthreads();
  if (is_last)
    for (int c = threadIdx.x; c < 2 * C; c += kThreads) stats[c] = 0.f;
}

int bn_grid(int M, int C, const void* kernel, size_t smem) {
  const int groups = C / kVec;
  const int lanes = kThreads / groups;
  static int occ_cache[2] = {0, 0};
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kThreads, smem);
  if (occ < 1) occ = 1;
  (void)occ_cache;
  const int cap = kml::num_sms() * (occ > 2 ? 2 : occ);
  int want = (M + lanes - 1) / lanes;                    // one row per thread if the GPU has room
  if (want > cap) want = cap;
  return want < 1 ? 1 : want;
}

}  // namespace

extern "C" {

size_t kml_bn_scratch_bytes(int C) {
  (void)C;
  return sizeof(BnScratch);
}

int kml_bn_fwd_train(const void* x, const float* x_slabs, int nslabs, long long slab_stride, void* x_out,
                     const void* res, void* y, const float* gamma, const float* beta, float* running_mean,
                     float* running_var, long long* num_batches_tracked, float* save_mean, float* save_invstd,
                     void* scratch, int M, int C, int relu, float eps, float momentum, void* stream) {
  KML_CHECK((x || x_slabs) && y && gamma && beta && save_mean && save_invstd && scratch, "kml_bn_fwd_train: null pointer");
  KML_CHECK(C % kVec == 0 && C / kVec <= kThreads && C <= kMaxC && M > 0,
            "kml_bn_fwd_train: C (%d) must be a multiple of 8 and <= %d", C, kMaxC);
  KML_CHECK(!x_slabs || (nslabs >= 1 && x_out), "kml_bn_fwd_train: slab input needs nslabs >= 1 and x_out");
  BnFwdArgs a{static_cast<const __nv_bfloat16*>(x_slabs ? nullptr : x), x_slabs, nslabs, slab_stride,
              static_cast<__nv_bfloat16*>(x_slabs ? x_out : nullptr), static_cast<const __nv_bfloat16*>(res),
              static_cast<__nv_bfloat16*>(y), gamma, beta, running_mean, running_var, num_batches_tracked,
              save_mean, save_invstd, static_cast<BnScratch*>(scratch), M, C, relu, eps, momentum, 0};
  size_t smem = 2 * C * sizeof(float);
  if (smem * (1 + kThreads / 32) <= 46 * 1024) { smem *= 1 + kThreads / 32; a.wide_smem = 1; }
  const int grid = bn_grid(M, C, reinterpret_cast<const void*>(bn_fwd_train_kernel), smem);
  KML_CUDA(kml::launch_pdl(bn_fwd_train_kernel, dim3(grid), dim3(kThreads), smem, static_cast<cudaStream_t>(stream), a));
  kml::count_launch();
  return 0;
}

size_t kml_bn_stats_bytes(int C) { return (2 * static_cast<size_t>(C) + 4) * sizeof(float); }

int kml_bn_apply_train(const void* x, float* stats, const void* res, void* y, const float* gamma, const float* beta,
                       float* running_mean, float* running_var, long long* num_batches_tracked, float* save_mean,
                       float* save_invstd, int M, int C, int relu, float eps, float momentum, void* stream) {
  KML_CHECK(x && stats && y && gamma && beta && save_mean && save_invstd, "kml_bn_apply_train: null pointer");
  KML_CHECK(C % kVec == 0 && C <= 6144 && M > 0, "kml_bn_apply_train: C (%d) must be a multiple of 8 and <= 6144 (48 KB of shared memory)", C);
  BnApplyArgs a{static_cast<const __nv_bfloat16*>(x), stats, static_cast<const __nv_bfloat16*>(res),
                static_cast<__nv_bfloat16*>(y), gamma, beta, running_mean, running_var, num_batches_tracked,
                save_mean, save_invstd, M, C, relu, eps, momentum};
  const size_t total = static_cast<size_t>(M) * (C / kVec);
  size_t want = (total + kThreads - 1) / kThreads;
  const size_t cap = static_cast<size_t>(kml::num_sms()) * 2;
  const int grid = static_cast<int>(want < cap ? (want < 1 ? 1 : want) : cap);
  KML_CUDA(kml::launch_pdl(bn_apply_train_kernel, dim3(grid), dim3(kThreads), 2 * C * sizeof(float),
                           static_cast<cudaStream_t>(stream), a));
  kml::count_launch();
  return 0;
}

int kml_bn_bwd_apply(const void* g, const void* x, const float* gamma, const float* save_mean,
                     const float* save_invstd, float* stats, void* dx, float* dgamma, float* dbeta, int M, int C,
                     void* stream) {
  KML_CHECK(g && x && gamma && save_mean && save_invstd && stats && dx && dgamma && dbeta, "kml_bn_bwd_apply: null pointer");
  KML_CHECK(C % kVec == 0 && C <= kMaxC && M > 0, "kml_bn_bwd_apply: C (%d) must be a multiple of 8 and <= %d", C, kMaxC);
  const size_t total = static_cast<size_t>(M) * (C / kVec);
  size_t want = (total + kThreads - 1) / kThreads;
  const size_t cap = static_cast<size_t>(kml::num_sms()) * 2;
  const int grid = static_cast<int>(want < cap ? (want < 1 ? 1 : want) : cap);
  KML_CUDA(kml::launch_pdl(bn_bwd_apply_kernel, dim3(grid), dim3(kThreads), 5 * C * sizeof(float),
                           static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(g),
                           static_cast<const __nv_bfloat16*>(x), gamma, save_mean, save_invstd, stats,
                           static_cast<__nv_bfloat16*>(dx), dgamma, dbeta, M, C));
  kml::count_launch();
  return 0;
}

int kml_bn_fwd_eval(const void* x, const float* x_slabs, int nslabs, long long slab_stride, const void* res, void* y,
                    const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                    int M, int C, int relu, float eps, void* stream) {
  KML_CHECK((x || x_slabs) && y && gamma && beta && running_mean && running_var, "kml_bn_fwd_eval: null pointer");
  KML_CHECK(C % kVec == 0 && M > 0, "kml_bn_fwd_eval: C must be a multiple of 8");
  const size_t total = static_cast<size_t>(M) * (C / kVec);
  size_t want = (total + kThreads - 1) / kThreads;
  const size_t cap = static_cast<size_t>(kml::num_sms()) * 8;
  const int grid = static_cast<int>(want < cap ? want : cap);
  bn_fwd_eval_kernel<<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x_slabs ? nullptr : x), x_slabs, nslabs, slab_stride,
      static_cast<const __nv_bfloat16*>(res), static_cast<__nv_bfloat16*>(y), gamma, beta, running_mean, running_var,
      M, C, relu, eps);
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

int kml_bn_bwd(const void* dy, const float* dy_slabs, int nslabs, long long slab_stride, const void* dy_add,
               const void* y, const void* x, const float* gamma, const float* save_mean, const float* save_invstd,
               void* dx, void* dres, float* dgamma, float* dbeta, void* scratch, int M, int C, void* stream) {
  KML_CHECK((dy || dy_slabs) && x && gamma && save_mean && save_invstd && dx && dgamma && dbeta && scratch,
            "kml_bn_bwd: null pointer");
  KML_CHECK(C % kVec == 0 && C / kVec <= kThreads && C <= kMaxC && M > 0,
            "kml_bn_bwd: C (%d) must be a multiple of 8 and <= %d", C, kMaxC);
  BnBwdArgs a{static_cast<const __nv_bfloat16*>(dy_slabs ? nullptr : dy), dy_slabs, nslabs, slab_stride,
              static_cast<const __nv_bfloat16*>(dy_add), static_cast<const __nv_bfloat16*>(y),
              static_cast<const __nv_bfloat16*>(x), gamma, save_mean, save_invstd, static_cast<__nv_bfloat16*>(dx),
              static_cast<__nv_bfloat16*>(dres), dgamma, dbeta, static_cast<BnScratch*>(scratch), M, C, 0};
  size_t smem = 2 * C * sizeof(float);
  if (smem * (1 + kThreads / 32) <= 46 * 1024) { smem *= 1 + kThreads / 32; a.wide_smem = 1; }
  const int grid = bn_grid(M, C, reinterpret_cast<const void*>(bn_bwd_kernel), smem);
  KML_CUDA(kml::launch_pdl(bn_bwd_kernel, dim3(grid), dim3(kThreads), smem, static_cast<cudaStream_t>(stream), a));
  kml::count_launch();
  return 0;
}

}  // extern "C"
