// Launch state shared by every kernel of one stream that synchronises its whole grid in software:
// the BatchNorm kernels (csrc/norm.cu) and the GEMMs with a fused BatchNorm epilogue
// (csrc/gemm_tcgen05.cu).  Zero-initialised once by the host, then owned by the kernels.
//
//   * `acc` is a double-buffered per-channel accumulator: launch k adds into buffer k&1 and clears
//     buffer (k+1)&1 for its successor, so nothing is zeroed by the host and a CUDA-graph replay works;
//   * `bar` counts CTA arrivals and only grows; `bar_base` is its value when the current launch began.
//     {launches, bar_base} are read as one uint2 at kernel entry (after the programmatic-dependency
//     wait) and advanced by one thread after every CTA has passed the barrier.
//
// Why a spin barrier in a regular launch is safe here (and where it is not):
//   1. The grid is sized by the host to be co-resident: <= SMs x resident CTAs of that kernel
//      (cudaOccupancyMaxActiveBlocksPerMultiprocessor / cudaOccupancyMaxActiveClusters at plan time).
//   2. Launches on one stream serialise, and a programmatic dependent launch only lets the *next*
//      kernel's CTAs become resident after EVERY CTA of this grid has started (the launch_dependents
//      trigger fires when all CTAs have executed it or exited).  So a successor can never hold the
//      slots that a not-yet-started CTA of the spinning grid needs.
//   3. Kernels of other streams (weight gradients, projection shortcuts) never wait on anything:
//      they finish and free their slots, so they can delay a spinning grid but not starve it.
//   What is NOT covered: two grids that both spin (two training streams on ONE GPU).  Each can be
//   partly resident and wait for slots the other holds.  The spin is therefore bounded: after
//   kSpinLimit cycles the kernel traps, which fails the launch ("unspecified launch failure") instead of
//   hanging the GPU.  One training stream per GPU is the supported configuration.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>

namespace kml {

constexpr int kBnMaxC = 2048;
constexpr long long kSpinLimit = 6000000000ll;     // ~3 s at 1.9 GHz

struct BnScratch {
  unsigned int launches;       // parity selects the statistics buffer of this launch
  unsigned int bar_base;       // value of `bar` when the current launch started
  unsigned int bar;            // grid barrier: arrivals, monotonically increasing
  unsigned int pad;
  float acc[2][2 * kBnMaxC];
};

// One thread of the CTA calls this after the CTA's own threads have synchronised; returns when
// `ncta` CTAs have arrived since `base`.
__device__ __forceinline__ void grid_arrive_and_wait(BnScratch* sc, unsigned base, unsigned ncta) {
  // release-arrive / acquire-poll: the release is cumulative over what this thread observed through the
  // CTA barrier before it, so no separate fences are needed on either side
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(&sc->bar), "r"(1u) : "memory");
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(&sc->bar) : "memory");
  if (v - base < ncta) {
    const long long t0 = clock64();
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(&sc->bar) : "memory");
      if (clock64() - t0 > kSpinLimit) asm volatile("trap;");   // fails the launch instead of hanging the GPU
    } while (v - base < ncta);
  }
}

}  // namespace kml
