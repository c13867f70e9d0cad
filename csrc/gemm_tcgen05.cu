// tcgen05 / TMEM / TMA GEMM family for sm_100a: plain GEMMs (linear layers, im2col'd convolutions)
// and implicit-GEMM convolution forward / data-gradient / weight-gradient.
//
// The reference has no kernels of its own: these shapes reach cuDNN / cuBLAS through
// torch 1.7.1 (ml/environment/Dockerfile:3,20; ml/experiments/kubeml/function_resnet34.py:65-75 —
// forward, F.cross_entropy, backward).  SURVEY 2.4 lists the op inventory this file covers
// (Conv2d fwd/dgrad/wgrad, Linear fwd/bwd).
//
// One CTA computes one 128 x BLOCK_N output tile (optionally one K-split of it):
//   warp 0   TMA producer: per 64-deep K chunk, cp.async.bulk.tensor boxes into a 4-stage smem ring
//   warp 1   MMA issuer: one elected lane issues 4 x tcgen05.mma (K=16) per chunk, accumulating
//            in TMEM; tcgen05.commit releases the smem slot / signals the epilogue
//   warps 2-5 epilogue: tcgen05.ld the fp32 accumulator (one TMEM lane = one output row per thread),
//            bias / ReLU / cast, store or red.add (split-K, weight gradients)
//
// Operands are always described by rank-4 tensor maps with 128B swizzle.  "K-major" operands keep
// the reduction dimension contiguous (activations NHWC over channels, weights [Cout][taps*Cin]);
// "MN-major" operands keep the M/N dimension contiguous (weight gradients reduce over pixels, so
// both dY and X are consumed channel-contiguous exactly as they sit in memory — no transposes).
// For convolutions the A box is a (64ch, BW, BH, BN) window of the NHWC tensor whose origin is
// shifted by the filter tap; out-of-bounds rows/cols are zero-filled by TMA, which *is* the padding.
#include "common.cuh"
#include "bn_scratch.cuh"
#include "driver_api.h"
#include "../include/kubeml_b200.h"
#include <cuda.h>
#include <algorithm>
#include <mutex>
#include <unordered_map>

namespace {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;            // bf16 elements = one 128-byte swizzle row
// 4 stages: split-K keeps the per-CTA K loop short, and <= 128 KB lets two CTAs share an SM so that
// independent GEMMs (weight gradient next to the data-gradient chain) overlap.
// 96 KB of pipeline per CTA either way (4 x 24 KB / 3 x 32 KB): two CTAs fit on an SM, so one CTA's
// epilogue overlaps the other's main loop (with 4 x 32 KB the 128-wide tile ran one CTA per SM)
template <int BLOCK_N> struct Stages { static constexpr int value = BLOCK_N > 128 ? 2 : (BLOCK_N > 64 ? 3 : 4); };
constexpr int kNumThreads = 192;       // producer warp, MMA warp, 4 epilogue warps
constexpr int kUmmaK = 16;

enum OperandKind : int {
  kMatK = 0,     // matrix, reduction dim contiguous:   coords (k, mn, 0, 0)
  kMatMN = 1,    // matrix, M/N dim contiguous:         coords (mn, k, 0, 0)
  kConvK = 2,    // NHWC activation, tile walks pixels, chunks walk (tap, channel block)
  kConvMN = 3,   // NHWC activation, tile walks channels, chunks walk pixel boxes (+ fixed tap shift)
  kWgtMN = 4,    // filter [Cout][taps*Cin] read as B[n=cin][k=cout] of the *flipped* tap (data gradient)
};

struct Geometry {       // how a linear index becomes a (w, h, n) box origin
  int bw, bh, bn;       // box extent
  int tiles_w, tiles_h; // boxes per row / per image
};

struct GemmArgs {
  int a_kind, b_kind;
  int m_tiles, n_tiles, splits, z_taps;  // grid = (m_tiles, n_tiles, z_taps * splits)
  int k_chunks;                          // chunks of 64 along the reduction dim (all taps included)
  // convolution description
  int kw, pad_w, pad_h, cblocks;         // filter width, padding, channel blocks per tap
  int conv_stride;                       // convolution stride (>= 1): activation boxes walk the input with this
                                         // element stride (the tensor map carries it), origins scale with it
  unsigned long long tap_mask;           // kConvK: taps that touch at least one real pixel
  Geometry tile_geo;                     // kConvK: output-tile pixel box
  Geometry chunk_geo;                    // kConvMN: reduction pixel box
  // epilogue
  void* out;
  long long ldo;                         // elements between output rows
  long long z_col_stride;                // column offset added per tap (weight gradients)
  int m_valid, n_valid;                  // rows / columns actually present
  const float* bias;                     // per output column, may be null
  const __nv_bfloat16* addend;           // bf16 [rows][ldo] added before ReLU (residual gradient), may be null
  int relu, out_f32, atomic_add;
  int wgt_taps, wgt_cin;                 // kWgtMN: taps in the filter, input channels per tap
  long long slab_stride;                 // != 0: split s stores its partial tile at out + s*slab_stride (no atomics)
  int persistent;                        // grid = resident CTAs; each walks tiles tile0, tile0+grid, ... with two
                                         // TMEM accumulators so a tile's epilogue overlaps the next tile's MMAs
  int cluster_reduce;                    // the `splits` CTAs of a tile are one cluster: partial tiles are added
                                         // through distributed shared memory, the epilogue runs once per element
  float* stats;                          // cluster mode: fp32 [2][n_valid] += column sums / sums of squares
  const __nv_bfloat16* mask_y;           // cluster mode: output *= (mask_y > 0)
  const __nv_bfloat16* bn_x;             // cluster mode: stats become {sum g, sum g * xhat}, xhat from bn_x/mean/invstd
  const float* bn_mean;
  const float* bn_invstd;
  // ---- BatchNorm fused behind the cluster reduction (kml_bn_fuse) --------------------------------
  // bn_mode 1 (forward): the reduced tile is the convolution output x; the kernel accumulates the batch
  //   statistics, synchronises the whole grid, and writes y = relu?(gamma*(x-mean)*invstd + beta + res?)
  //   plus the raw x (backward needs it) - convolution and BatchNorm in one launch.
  // bn_mode 2 (backward): the reduced tile (+ addend) masked by mask_y is the gradient g entering a
  //   BatchNorm; the kernel accumulates sum g and sum g*xhat, synchronises the grid and writes
  //   dx = gamma*invstd*(g - sum_g/M - xhat*sum_gx/M), optionally g itself (residual branch), dgamma, dbeta.
  int bn_mode;
  kml::BnScratch* bn_sc;
  const float* bn_gamma;
  const float* bn_beta;
  float* bn_rmean;
  float* bn_rvar;
  long long* bn_nbt;
  float* bn_save_mean;
  float* bn_save_invstd;
  const __nv_bfloat16* bn_res;
  __nv_bfloat16* bn_raw;
  int bn_relu;
  float bn_eps, bn_momentum;
  float* bn_dgamma;
  float* bn_dbeta;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
// Bounded wait: a pipeline bug (wrong byte count, bad descriptor) must surface as a trapped kernel,
// not as a GPU that spins until the box is reclaimed.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000ll) {
      printf("kubeml_b200 gemm: mbarrier wait timed out (block %d,%d,%d thread %d)\n", blockIdx.x, blockIdx.y,
             blockIdx.z, threadIdx.x);
      asm volatile("trap;");
    }
  }
}
__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2),
      "r"(c3) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// float4 from the shared memory of CTA `rank` of this cluster, at the same offset as local address `laddr`
__device__ __forceinline__ float4 ld_dsmem_f4(uint32_t laddr, uint32_t rank) {
  uint32_t raddr;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(laddr), "r"(rank));
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(raddr) : "memory");
  return v;
}
__device__ __forceinline__ float ld_dsmem_f32(uint32_t laddr, uint32_t rank) {
  uint32_t raddr;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(laddr), "r"(rank));
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(raddr) : "memory");
  return v;
}
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc),
      "r"(accumulate) : "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(pred)::"memory");
  return pred != 0;
}

// Shared-memory matrix descriptor, 128B swizzle (cute::UMMA::SmemDescriptor: start>>4 [0,14),
// LBO>>4 [16,30), SBO>>4 [32,46), version=1 [46,48), layout_type=2 (SWIZZLE_128B) [61,64)).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= 1ull << 46;
  d |= 2ull << 61;
  return d;
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D=f32 [4,6)=1, A=bf16 [7,10)=1, B=bf16
// [10,13)=1, a_major bit 15, b_major bit 16, N>>3 [17,23), M>>4 [24,29).
__host__ __device__ constexpr uint32_t make_idesc(int m, int n, bool a_mn, bool b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(a_mn) << 15) |
         (static_cast<uint32_t>(b_mn) << 16) | (static_cast<uint32_t>(n >> 3) << 17) |
         (static_cast<uint32_t>(m >> 4) << 24);
}

__device__ __forceinline__ void box_origin(const Geometry& g, int idx, int& w0, int& h0, int& n0) {
  const int tw = idx % g.tiles_w;
  const int t2 = idx / g.tiles_w;
  w0 = tw * g.bw;
  h0 = (t2 % g.tiles_h) * g.bh;
  n0 = (t2 / g.tiles_h) * g.bn;
}

// LINEAR: tiles are addressed by a linear, band-rasterised index (1-D grid, optionally persistent) -
// the variant for problems of many tiles.  !LINEAR: one tile per CTA straight from the 3-D block index
// (no index arithmetic at all) - the variant for the small, latency-bound launches and cluster mode.
template <int BLOCK_N, bool A_MN, bool B_MN, bool LINEAR, int BN_MODE = 0>
__global__ void __launch_bounds__(kNumThreads, 2)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                    const GemmArgs args) {
  constexpr uint32_t kABytes = kBlockM * kBlockK * 2;
  constexpr uint32_t kBBytes = BLOCK_N * kBlockK * 2;
  constexpr uint32_t kHalfBytes = 64 * kBlockK * 2;   // one 64-wide MN block of an MN-major operand
  constexpr uint32_t kTmemCols = BLOCK_N < 32 ? 32 : BLOCK_N;
  constexpr int kStages = Stages<BLOCK_N>::value;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kABytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_b + kStages * kBBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;        // [2] accumulator ready (MMA -> epilogue)
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;          // [2] accumulator drained (epilogue -> MMA)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
  int* tap_list = reinterpret_cast<int*>(tmem_slot + 1);   // up to 64 entries

  kml::pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_mn = args.m_tiles * args.n_tiles;
  const bool persistent = LINEAR && args.persistent != 0;
  const int total_tiles = LINEAR ? tiles_mn * args.z_taps * args.splits : 1;
  const int tile0 = LINEAR ? static_cast<int>(blockIdx.x) : 0;
  const int tstride = persistent ? static_cast<int>(gridDim.x) : total_tiles;
  const int per = (args.k_chunks + args.splits - 1) / args.splits;
  const uint32_t tmem_cols = persistent ? 2 * kTmemCols : kTmemCols;
  struct Tile { int mt, nt, z_tap, split, kbeg, nchunks; };
  auto decode = [&](int tile) {
    Tile t;
    int z;
    if constexpr (LINEAR) {
      // Rasterisation: walk the (m, n) tile grid in bands of kGroupM m-tiles, n fastest inside a band,
      // so that the tiles in flight at any time (~2 per SM) share few A row-blocks and few B
      // column-blocks (a 16 x 18 patch instead of a 128 x 2 stripe keeps the operands L2-resident).
      constexpr int kGroupM = 16;
      const int idx = tile % tiles_mn;
      const int band = idx / (kGroupM * args.n_tiles);
      const int band_m0 = band * kGroupM;
      const int band_rows = min(kGroupM, args.m_tiles - band_m0);
      const int in_band = idx - band * kGroupM * args.n_tiles;
      t.mt = band_m0 + in_band % band_rows;
      t.nt = in_band / band_rows;
      z = tile / tiles_mn;
    } else {
      t.mt = blockIdx.x; t.nt = blockIdx.y; z = blockIdx.z;
    }
    t.z_tap = z / args.splits;
    t.split = z % args.splits;
    t.kbeg = t.split * per;
    t.nchunks = max(0, min(args.k_chunks, t.kbeg + per) - t.kbeg);
    return t;
  };
  const Tile T_only = decode(0);      // !LINEAR: the one tile of this CTA, decoded once for all roles
  auto tile_at = [&](int tile) { return LINEAR ? decode(tile) : T_only; };

  if (threadIdx.x == 0) {
    // the descriptors are kernel parameters, not data of the previous kernel: fetch them before pdl_wait
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
    for (int s = 0; s < kStages; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    mbar_init(tmem_full_bar, 1);
    if (LINEAR) { mbar_init(tmem_full_bar + 1, 1); mbar_init(tmem_empty_bar, 4); mbar_init(tmem_empty_bar + 1, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    int n = 0;
    for (int t = 0; t < 64; ++t) if ((args.tap_mask >> t) & 1ull) tap_list[n++] = t;
  }
  if (warp == 1) {   // this warp owns the TMEM allocation
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  kml::pdl_wait();       // the set-up above overlapped the previous kernel; its results are needed from here on
  // fused BatchNorm: launch state of the software grid barrier, read early so the round trip overlaps the
  // main loop (nobody can advance it before every CTA of this grid has passed the barrier)
  uint2 gstate = make_uint2(0u, 0u);
  if constexpr (BN_MODE != 0) gstate = __ldcg(reinterpret_cast<const uint2*>(&args.bn_sc->launches));

  if (warp == 0) {
    // ===================================== TMA producer =====================================
    if (elect_one()) {
      int it = 0;                                        // chunks issued so far, across tiles
      for (int tile = tile0; tile < total_tiles; tile += tstride) {
      const Tile T = tile_at(tile);
      const int mt = T.mt, nt = T.nt, z_tap = T.z_tap, kbeg = T.kbeg, nchunks = T.nchunks;
      int tw0 = 0, th0 = 0, tn0 = 0;
      if (args.a_kind == kConvK) box_origin(args.tile_geo, mt, tw0, th0, tn0);
      const int z_dw = z_tap % max(args.kw, 1) - args.pad_w;   // kConvMN: fixed shift of this CTA's tap
      const int z_dh = z_tap / max(args.kw, 1) - args.pad_h;
      for (int i = 0; i < nchunks; ++i, ++it) {
        const int kc = kbeg + i;
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1;
        mbar_wait(empty_bar + s, ph ^ 1);
        mbar_expect_tx(full_bar + s, kABytes + kBBytes);
        uint8_t* sa = smem_a + s * kABytes;
        uint8_t* sb = smem_b + s * kBBytes;
        int cw = 0, ch = 0, cn = 0;
        if (args.a_kind == kConvMN || args.b_kind == kConvMN) box_origin(args.chunk_geo, kc, cw, ch, cn);
        // ---- A ----
        if (args.a_kind == kMatK) {
          tma_load_4d(&map_a, full_bar + s, sa, kc * kBlockK, mt * kBlockM, 0, 0);
        } else if (args.a_kind == kConvK) {
          const int tap = tap_list[kc / args.cblocks];
          const int cb = kc % args.cblocks;
          tma_load_4d(&map_a, full_bar + s, sa, cb * kBlockK, tw0 * args.conv_stride + tap % args.kw - args.pad_w,
                      th0 * args.conv_stride + tap / args.kw - args.pad_h, tn0);
        } else if (args.a_kind == kMatMN) {
#pragma unroll
          for (int hb = 0; hb < kBlockM / 64; ++hb)
            tma_load_4d(&map_a, full_bar + s, sa + hb * kHalfBytes, mt * kBlockM + hb * 64, kc * kBlockK, 0, 0);
        } else {
#pragma unroll
          for (int hb = 0; hb < kBlockM / 64; ++hb)
            tma_load_4d(&map_a, full_bar + s, sa + hb * kHalfBytes, mt * kBlockM + hb * 64, cw, ch, cn);
        }
        // ---- B ----
        if (args.b_kind == kMatK) {
          // with a convolution A operand the filter column is addressed by the *real* tap, so
          // skipped (all-padding) taps need no repacking of the weights
          const int bcol = args.a_kind == kConvK
                               ? (tap_list[kc / args.cblocks] * args.cblocks + kc % args.cblocks) * kBlockK
                               : kc * kBlockK;
          tma_load_4d(&map_b, full_bar + s, sb, bcol, nt * BLOCK_N, 0, 0);
        } else if (args.b_kind == kMatMN) {
#pragma unroll
          for (int hb = 0; hb < BLOCK_N / 64; ++hb)
            tma_load_4d(&map_b, full_bar + s, sb + hb * kHalfBytes, nt * BLOCK_N + hb * 64, kc * kBlockK, 0, 0);
        } else if (args.b_kind == kWgtMN) {
          // chunk = (tap' of the dY window, block of 64 output channels); the filter tap is the
          // point reflection of tap' (dX = dY * flip(W))
          const int tap = args.wgt_taps - 1 - tap_list[kc / args.cblocks];
          const int cb = kc % args.cblocks;
#pragma unroll
          for (int hb = 0; hb < BLOCK_N / 64; ++hb)
            tma_load_4d(&map_b, full_bar + s, sb + hb * kHalfBytes, tap * args.wgt_cin + nt * BLOCK_N + hb * 64,
                        cb * kBlockK, 0, 0);
        } else {
#pragma unroll
          for (int hb = 0; hb < BLOCK_N / 64; ++hb)
            tma_load_4d(&map_b, full_bar + s, sb + hb * kHalfBytes, nt * BLOCK_N + hb * 64, cw * args.conv_stride + z_dw,
                        ch * args.conv_stride + z_dh, cn);
        }
      }
      }   // tiles
    }
  } else if (warp == 1) {
    // ===================================== MMA issuer ========================================
    constexpr uint32_t idesc = make_idesc(kBlockM, BLOCK_N, A_MN, B_MN);
    int it = 0, j = 0;
    for (int tile = tile0; tile < total_tiles; tile += tstride, ++j) {
    const int nchunks = tile_at(tile).nchunks;
    const int acc = persistent ? (j & 1) : 0;
    const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BLOCK_N);
    // the epilogue must have drained this accumulator (tile j-2); passes at once the first time
    if (persistent) {
      mbar_wait(tmem_empty_bar + acc, ((j >> 1) & 1) ^ 1);
      tcgen05_fence_after();
    }
    if (nchunks == 0 && elect_one()) mbar_arrive(tmem_full_bar + acc);
    for (int i = 0; i < nchunks; ++i, ++it) {
      const int s = it % kStages;
      const uint32_t ph = (it / kStages) & 1;
      mbar_wait(full_bar + s, ph);
      tcgen05_fence_after();
      if (elect_one()) {
        const uint32_t a_addr = smem_u32(smem_a + s * kABytes);
        const uint32_t b_addr = smem_u32(smem_b + s * kBBytes);
#pragma unroll
        for (int k = 0; k < kBlockK / kUmmaK; ++k) {
          // K-major: 16 bf16 = 32 bytes further along the swizzled row.
          // MN-major: 16 reduction rows = two 8-row groups of 1024 bytes.
          const uint64_t adesc = A_MN ? make_smem_desc(a_addr + k * 2048, kHalfBytes, 1024)
                                      : make_smem_desc(a_addr + k * 32, 16, 1024);
          const uint64_t bdesc = B_MN ? make_smem_desc(b_addr + k * 2048, kHalfBytes, 1024)
                                      : make_smem_desc(b_addr + k * 32, 16, 1024);
          umma_bf16(d_tmem, adesc, bdesc, idesc, (i | k) != 0 ? 1u : 0u);
        }
      }
      __syncwarp();
      if (elect_one()) {
        umma_commit(empty_bar + s);                     // frees the smem slot when the MMAs retire
        if (i == nchunks - 1) umma_commit(tmem_full_bar + acc);
      }
      __syncwarp();
    }
    }   // tiles
  } else {
    // ===================================== epilogue ==========================================
    const int q = warp & 3;                              // TMEM lane quarter this warp may read
    int j = 0;
    for (int tile = tile0; tile < total_tiles; tile += tstride, ++j) {
    const Tile T = tile_at(tile);
    const int mt = T.mt, nt = T.nt, z_tap = T.z_tap, split = T.split, nchunks = T.nchunks;
    const int acc = persistent ? (j & 1) : 0;
    const int row = mt * kBlockM + q * 32 + lane;
    mbar_wait(tmem_full_bar + acc, (j >> 1) & 1);
    tcgen05_fence_after();
    const bool row_ok = row < args.m_valid;
    const long long col_base = static_cast<long long>(z_tap) * args.z_col_stride + static_cast<long long>(nt) * BLOCK_N;
#pragma unroll 1
    for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
      uint32_t v[32];
      if (nchunks > 0) {
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                               static_cast<uint32_t>(acc * BLOCK_N + c0);
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
              "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
              "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
              "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0u;
      }
      if (persistent && c0 + 32 >= BLOCK_N) {   // accumulator fully read: hand it back before the stores
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tmem_empty_bar + acc);
      }
      if (BN_MODE != 0 || args.cluster_reduce) {
        // partial tile -> own shared memory (the pipeline ring is idle now); the reduction across the
        // cluster and the real epilogue follow after the role branches
        float* part = reinterpret_cast<float*>(smem_a) + (q * 32 + lane) * (BLOCK_N + 4) + c0;
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(part + j) = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                             __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
        continue;
      }
      if constexpr (BN_MODE == 0) {      // (the fused-BatchNorm instantiations never store from here)
      if (!row_ok) continue;
      const int ncol0 = nt * BLOCK_N + c0;                // column within the (per-tap) N range
      const long long off = static_cast<long long>(row) * args.ldo + col_base + c0 +
                            static_cast<long long>(split) * args.slab_stride;
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        f[j] = __uint_as_float(v[j]);
        if (args.bias != nullptr && split == 0 && ncol0 + j < args.n_valid) f[j] += __ldg(args.bias + ncol0 + j);
      }
      if (args.addend != nullptr && split == 0) {
        const __nv_bfloat16* ad = args.addend + off;
        if (ncol0 + 32 <= args.n_valid && (off & 7) == 0) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const uint4 u = __ldg(reinterpret_cast<const uint4*>(ad + j));
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const float2 v2 = __bfloat1622float2(h[t]);
              f[j + 2 * t] += v2.x;
              f[j + 2 * t + 1] += v2.y;
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (ncol0 + j < args.n_valid) f[j] += __bfloat162float(ad[j]);
        }
      }
      if (args.relu) {
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
      }
      if (args.out_f32) {
        float* o = static_cast<float*>(args.out) + off;
        if (args.atomic_add) {
          if (ncol0 + 32 <= args.n_valid && (off & 3) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + j), "f"(f[j]), "f"(f[j + 1]),
                           "f"(f[j + 2]), "f"(f[j + 3]) : "memory");
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (ncol0 + j < args.n_valid) atomicAdd(o + j, f[j]);
          }
        } else if (ncol0 + 32 <= args.n_valid && (off & 3) == 0) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (ncol0 + j < args.n_valid) o[j] = f[j];
        }
      } else {
        __nv_bfloat16* o = static_cast<__nv_bfloat16*>(args.out) + off;
        if (ncol0 + 32 <= args.n_valid && (off & 7) == 0) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            __nv_bfloat162 p0 = __floats2bfloat162_rn(f[j], f[j + 1]);
            __nv_bfloat162 p1 = __floats2bfloat162_rn(f[j + 2], f[j + 3]);
            __nv_bfloat162 p2 = __floats2bfloat162_rn(f[j + 4], f[j + 5]);
            __nv_bfloat162 p3 = __floats2bfloat162_rn(f[j + 6], f[j + 7]);
            uint4 u;
            u.x = *reinterpret_cast<uint32_t*>(&p0); u.y = *reinterpret_cast<uint32_t*>(&p1);
            u.z = *reinterpret_cast<uint32_t*>(&p2); u.w = *reinterpret_cast<uint32_t*>(&p3);
            *reinterpret_cast<uint4*>(o + j) = u;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (ncol0 + j < args.n_valid) o[j] = __float2bfloat16_rn(f[j]);
        }
      }
      }   // BN_MODE == 0
    }
    }   // tiles
    tcgen05_fence_before();
  }

  __syncthreads();
  if constexpr (BN_MODE != 0) {
  // ===== cluster reduction + fused BatchNorm (kml_bn_fuse); instantiated only for the variants that use it =====
  // Everything here runs once per CTA on a cold instruction cache (ncu: ~12 % of this kernel's stall samples
  // were instruction fetches), so the code is kept small: one mode per instantiation, 8-byte vector accesses
  // only (the host checks N % 4 == 0 and ldo % 4 == 0, so a float4 column group is either whole or absent).
  {
    constexpr bool kFwd = BN_MODE == 1;
    const Tile T0 = tile_at(tile0);                      // cluster mode is never persistent: one tile
    const int mt = T0.mt, nt = T0.nt, split = T0.split;
    constexpr int kPartStride = BLOCK_N + 4;
    constexpr int kGroups = BLOCK_N / 4;                 // float4 groups per row
    constexpr int kRowStep = kNumThreads / kGroups;      // rows covered by one sweep of the CTA (all six warps)
    constexpr int kWarps = kNumThreads / 32;
    constexpr int kU = 4;                                // rows a thread has in flight (latency-bound pass)
    float* part = reinterpret_cast<float*>(smem_a);
    float* st_sm = part + kBlockM * kPartStride;        // [6 warps][2][BLOCK_N] per-warp column sums
    float* pv_sm = st_sm + 2 * kWarps * BLOCK_N;         // [BLOCK_N] pivots of the forward statistics
    float* co = pv_sm + BLOCK_N;                         // [3][BLOCK_N] per-column coefficients
    const int et = threadIdx.x;
    // Per-channel inputs that do not depend on the batch statistics, fetched before the barriers.  Forward:
    // pivot = running mean (shifted sums (x - pivot), (x - pivot)^2: no catastrophic cancellation when
    // |mean| >> std; read BEFORE the grid barrier, the running statistics change after it), gamma, beta.
    // Backward: gamma * invstd.
    for (int c = et; c < BLOCK_N; c += kNumThreads) {
      const int gc = nt * BLOCK_N + c;
      const bool in = gc < args.n_valid;
      if constexpr (kFwd) {
        pv_sm[c] = (args.bn_rmean != nullptr && in) ? args.bn_rmean[gc] : 0.f;
        co[c] = in ? args.bn_gamma[gc] : 0.f;
        co[BLOCK_N + c] = in ? args.bn_beta[gc] : 0.f;
      } else {
        co[c] = in ? args.bn_gamma[gc] * args.bn_invstd[gc] : 0.f;
      }
    }
    cluster_sync_all();
    const int S = args.splits;
    const int rp = (kBlockM + S - 1) / S;
    const int rbeg = split * rp;
    const int rend = min(kBlockM, rbeg + rp);
    const int c4 = (et % kGroups) * 4;
    const int ncol = nt * BLOCK_N + c4;
    const bool col_ok = ncol + 4 <= args.n_valid;
    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
    float bm[4] = {0.f, 0.f, 0.f, 0.f}, bi[4] = {0.f, 0.f, 0.f, 0.f};
    if constexpr (kFwd) {
#pragma unroll
      for (int t = 0; t < 4; ++t) bm[t] = pv_sm[c4 + t];          // forward: bm holds the pivots
    } else if (col_ok) {
#pragma unroll
      for (int t = 0; t < 4; ++t) { bm[t] = args.bn_mean[ncol + t]; bi[t] = args.bn_invstd[ncol + t]; }
    }
    const __nv_bfloat16* ad_p = args.addend;
    const __nv_bfloat16* my_p = args.mask_y;
    const __nv_bfloat16* bx_p = args.bn_x;
    auto unpack4 = [](const uint2& u, float (&v)[4]) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
      const float2 a0 = __bfloat1622float2(h[0]), a1 = __bfloat1622float2(h[1]);
      v[0] = a0.x; v[1] = a0.y; v[2] = a1.x; v[3] = a1.y;
    };
    auto pack4 = [](const float (&f)[4]) {
      const __nv_bfloat162 p0 = __floats2bfloat162_rn(f[0], f[1]), p1 = __floats2bfloat162_rn(f[2], f[3]);
      uint2 u;
      u.x = *reinterpret_cast<const uint32_t*>(&p0); u.y = *reinterpret_cast<const uint32_t*>(&p1);
      return u;
    };
    // ---- pass 1: reduce the partial tiles, (backward: + addend, mask), round, stash, column sums -------
#pragma unroll 1
    for (int rb = rbeg + et / kGroups; rb < rend; rb += kU * kRowStep) {
      long long ro[kU];
      bool ok[kU];
      uint2 uad[kU], umy[kU], ubx[kU];
      float4 acc[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {                       // the L2 round trips first
        const int r = rb + u * kRowStep;
        const int grow = mt * kBlockM + r;
        ok[u] = col_ok && r < rend && grow < args.m_valid;
        ro[u] = static_cast<long long>(grow) * args.ldo + ncol;
        uad[u] = umy[u] = ubx[u] = make_uint2(0u, 0u);
        if constexpr (!kFwd) {
          if (ok[u]) {
            if (ad_p != nullptr) uad[u] = __ldg(reinterpret_cast<const uint2*>(ad_p + ro[u]));
            if (my_p != nullptr) umy[u] = __ldg(reinterpret_cast<const uint2*>(my_p + ro[u]));
            ubx[u] = __ldg(reinterpret_cast<const uint2*>(bx_p + ro[u]));
          }
        }
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {                       // the partial tiles, kU loads in flight per peer
        const int r = rb + u * kRowStep;
        acc[u] = r < rend ? ld_dsmem_f4(smem_u32(part + r * kPartStride + c4), 0) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll 1
      for (int p = 1; p < S; ++p) {
        float4 o[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          const int r = rb + u * kRowStep;
          o[u] = r < rend ? ld_dsmem_f4(smem_u32(part + r * kPartStride + c4), static_cast<uint32_t>(p))
                          : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) { acc[u].x += o[u].x; acc[u].y += o[u].y; acc[u].z += o[u].z; acc[u].w += o[u].w; }
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        if (!ok[u]) continue;
        float f[4] = {acc[u].x, acc[u].y, acc[u].z, acc[u].w};
        if constexpr (!kFwd) {
          if (ad_p != nullptr) {
            float a[4];
            unpack4(uad[u], a);
#pragma unroll
            for (int t = 0; t < 4; ++t) f[t] += a[t];
          }
          if (my_p != nullptr) {
            float m[4];
            unpack4(umy[u], m);
#pragma unroll
            for (int t = 0; t < 4; ++t) f[t] = m[t] > 0.f ? f[t] : 0.f;
          }
        }
        // the intermediate (raw convolution output / masked gradient) is stored where a later pass needs it;
        // statistics and the stash use exactly the rounded values the consumer will read back
        const uint2 pk = pack4(f);
        if (args.bn_raw != nullptr) *reinterpret_cast<uint2*>(args.bn_raw + ro[u]) = pk;
        unpack4(pk, f);
        // Stash for the pass after the grid barrier.  In place: rows [rbeg, rend) of THIS CTA's partial tile
        // are read by this CTA only (every peer reads its own row range of it).
        *reinterpret_cast<float4*>(part + (rb + u * kRowStep) * kPartStride + c4) = make_float4(f[0], f[1], f[2], f[3]);
        if constexpr (kFwd) {
#pragma unroll
          for (int t = 0; t < 4; ++t) { const float d = f[t] - bm[t]; s1[t] += d; s2[t] = fmaf(d, d, s2[t]); }
        } else {
          float xh[4];
          unpack4(ubx[u], xh);
#pragma unroll
          for (int t = 0; t < 4; ++t) { s1[t] += f[t]; s2[t] = fmaf(f[t], (xh[t] - bm[t]) * bi[t], s2[t]); }
        }
      }
    }
    // ---- column sums: shuffle fold, one row per warp, 6-way sum per CTA --------------------------------
    kml::BnScratch* sc = args.bn_sc;
    float* gstats = sc->acc[gstate.x & 1u];
    float* cs = co + 3 * BLOCK_N;                        // [2][BLOCK_N] this CTA's column sums (cluster-local mode)
    // One 128-row tile holds every row of the layer (M <= 128: the 1x1 maps of ResNet's last stage): the cluster
    // that computes a column block already has the complete statistics of its columns.  The CTAs then exchange
    // their sums through distributed shared memory - no global atomics, no grid-wide barrier.
    const bool local_stats = args.m_tiles == 1;
    if (kGroups == 16) {
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        s1[t] += __shfl_xor_sync(0xffffffffu, s1[t], 16);
        s2[t] += __shfl_xor_sync(0xffffffffu, s2[t], 16);
      }
    }
    {
      float* wsm = st_sm + warp * 2 * BLOCK_N;
      if (lane < kGroups) {
#pragma unroll
        for (int t = 0; t < 4; ++t) { wsm[c4 + t] = s1[t]; wsm[BLOCK_N + c4 + t] = s2[t]; }
      }
    }
    __syncthreads();
    for (int i = et; i < 2 * BLOCK_N; i += kNumThreads) {
      const int c = i % BLOCK_N, which = i / BLOCK_N;
      const int gc = nt * BLOCK_N + c;
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) tot += st_sm[w * 2 * BLOCK_N + i];
      if (local_stats) cs[i] = tot;
      else if (gc < args.n_valid) atomicAdd(gstats + static_cast<size_t>(which) * args.n_valid + gc, tot);
    }
    if (local_stats) {
      cluster_sync_all();                                // every CTA's sums are in its shared memory
      if ((blockIdx.x | blockIdx.y | blockIdx.z) == 0 && et == 0 && kFwd && args.bn_nbt != nullptr) *args.bn_nbt += 1;
    } else {
      // ---- whole-grid barrier: every CTA's column sums are in, then finish the BatchNorm ------------
      const unsigned ncta = gridDim.x * gridDim.y * gridDim.z;
      if (blockIdx.x == gridDim.x - 1 && blockIdx.y == gridDim.y - 1 && blockIdx.z == gridDim.z - 1) {
        float4* nxt = reinterpret_cast<float4*>(sc->acc[(gstate.x & 1u) ^ 1u]);    // the next launch's buffer
        for (int i = et; i < 2 * kml::kBnMaxC / 4; i += kNumThreads) nxt[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      __syncthreads();
      if (et == 0) kml::grid_arrive_and_wait(sc, gstate.y, ncta);
      __syncthreads();
      if ((blockIdx.x | blockIdx.y | blockIdx.z) == 0 && et == 0) {   // everyone has arrived: publish the next launch's state
        sc->launches = gstate.x + 1u;
        sc->bar_base = gstate.y + ncta;
        if (kFwd && args.bn_nbt != nullptr) *args.bn_nbt += 1;
      }
    }
    {
      const float inv_m = 1.f / static_cast<float>(args.m_valid);
      const bool writer = mt == 0 && split == 0;         // one CTA per column block records the per-channel results
      for (int c = et; c < BLOCK_N; c += kNumThreads) {
        const int gc = nt * BLOCK_N + c;
        float c0 = 0.f, c1 = 0.f, c2 = 0.f;
        if (gc < args.n_valid) {
          float t1 = 0.f, t2 = 0.f;
          if (local_stats) {
            for (int p = 0; p < S; ++p) {
              t1 += ld_dsmem_f32(smem_u32(cs + c), static_cast<uint32_t>(p));
              t2 += ld_dsmem_f32(smem_u32(cs + BLOCK_N + c), static_cast<uint32_t>(p));
            }
          } else {
            t1 = __ldcg(gstats + gc); t2 = __ldcg(gstats + args.n_valid + gc);
          }
          if constexpr (kFwd) {
            const float d = t1 * inv_m;
            const float mean = pv_sm[c] + d;
            const float var = fmaxf(t2 * inv_m - d * d, 0.f);
            const float invstd = rsqrtf(var + args.bn_eps);
            c0 = co[c] * invstd;                         // gamma * invstd
            c1 = co[BLOCK_N + c] - mean * c0;            // beta - mean * scale
            if (writer) {
              args.bn_save_mean[gc] = mean;
              args.bn_save_invstd[gc] = invstd;
              if (args.bn_rmean != nullptr) {
                const float m = static_cast<float>(args.m_valid);
                const float unbiased = args.m_valid > 1 ? var * m / (m - 1.f) : var;
                args.bn_rmean[gc] = (1.f - args.bn_momentum) * pv_sm[c] + args.bn_momentum * mean;
                args.bn_rvar[gc] = (1.f - args.bn_momentum) * args.bn_rvar[gc] + args.bn_momentum * unbiased;
              }
            }
          } else {
            c0 = co[c];                                  // gamma * invstd
            c1 = t1 * inv_m;
            c2 = t2 * inv_m;
            if (writer) { args.bn_dbeta[gc] = t1; args.bn_dgamma[gc] = t2; }
          }
        }
        co[c] = c0; co[BLOCK_N + c] = c1; co[2 * BLOCK_N + c] = c2;
      }
    }
    __syncthreads();
    // ---- pass 2: normalise (forward: + residual, ReLU) / input gradient (backward), write the output ----
    {
      float k0[4], k1[4], k2[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) { k0[t] = co[c4 + t]; k1[t] = co[BLOCK_N + c4 + t]; k2[t] = co[2 * BLOCK_N + c4 + t]; }
      const __nv_bfloat16* aux_p = kFwd ? args.bn_res : args.bn_x;     // the one global operand of this pass
      __nv_bfloat16* outp = static_cast<__nv_bfloat16*>(args.out);
#pragma unroll 1
      for (int rb = rbeg + et / kGroups; rb < rend; rb += kU * kRowStep) {
        long long ro[kU];
        bool ok[kU];
        uint2 ua[kU];
        float4 v[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          const int r = rb + u * kRowStep;
          const int grow = mt * kBlockM + r;
          ok[u] = col_ok && r < rend && grow < args.m_valid;
          ro[u] = static_cast<long long>(grow) * args.ldo + ncol;
          ua[u] = make_uint2(0u, 0u);
          v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (ok[u]) {
            if (aux_p != nullptr) ua[u] = __ldg(reinterpret_cast<const uint2*>(aux_p + ro[u]));
            v[u] = *reinterpret_cast<const float4*>(part + r * kPartStride + c4);
          }
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          if (!ok[u]) continue;
          float f[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
          float a[4];
          unpack4(ua[u], a);
          if constexpr (kFwd) {
#pragma unroll
            for (int t = 0; t < 4; ++t) f[t] = fmaf(f[t], k0[t], k1[t]) + a[t];       // a = residual or zeros
            if (args.bn_relu) {
#pragma unroll
              for (int t = 0; t < 4; ++t) f[t] = fmaxf(f[t], 0.f);
            }
          } else {
#pragma unroll
            for (int t = 0; t < 4; ++t) f[t] = k0[t] * (f[t] - k1[t] - (a[t] - bm[t]) * bi[t] * k2[t]);
          }
          *reinterpret_cast<uint2*>(outp + ro[u]) = pack4(f);
        }
      }
    }
    cluster_sync_all();                                  // peers may still be reading this CTA's partial tile
  }
  } else if (args.cluster_reduce) {
    const Tile T0 = tile_at(tile0);                      // cluster mode is never persistent: one tile
    const int mt = T0.mt, nt = T0.nt, split = T0.split;
    // ---- split-K reduction across the cluster through distributed shared memory ----------------
    // CTA `split` owns rows [rbeg, rend) of the 128-row tile: it adds the partial tiles of all
    // peers for those rows, applies the epilogue once and writes the final output.  Column sums
    // for BatchNorm fall out of the same pass (a thread keeps the same 4 columns for all its rows).
    constexpr int kPartStride = BLOCK_N + 4;
    constexpr int kGroups = BLOCK_N / 4;                 // float4 groups per row
    float* part = reinterpret_cast<float*>(smem_a);
    float* st_sm = part + kBlockM * kPartStride;        // [4 warps][2][BLOCK_N] per-warp column sums
    cluster_sync_all();
    if (warp >= 2) {
      const int S = args.splits;
      const int rp = (kBlockM + S - 1) / S;
      const int rbeg = split * rp;
      const int rend = min(kBlockM, rbeg + rp);
      const int et = threadIdx.x - 64;                   // 0..127
      const int c4 = (et % kGroups) * 4;
      const int ncol = nt * BLOCK_N + c4;
      const bool col_ok = ncol + 4 <= args.n_valid;
      float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
      float bv[4] = {0.f, 0.f, 0.f, 0.f};
      if (args.bias != nullptr) {
#pragma unroll
        for (int t = 0; t < 4; ++t) if (ncol + t < args.n_valid) bv[t] = __ldg(args.bias + ncol + t);
      }
      float bm[4] = {0.f, 0.f, 0.f, 0.f}, bi[4] = {0.f, 0.f, 0.f, 0.f};
      if (args.bn_x != nullptr) {
#pragma unroll
        for (int t = 0; t < 4; ++t)
          if (ncol + t < args.n_valid) { bm[t] = args.bn_mean[ncol + t]; bi[t] = args.bn_invstd[ncol + t]; }
      }
      for (int r = rbeg + et / kGroups; r < rend; r += 128 / kGroups) {
        const uint32_t laddr = smem_u32(part + r * kPartStride + c4);
        float4 acc = ld_dsmem_f4(laddr, 0);
        for (int p = 1; p < S; ++p) {
          const float4 o = ld_dsmem_f4(laddr, static_cast<uint32_t>(p));
          acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w;
        }
        const int grow = mt * kBlockM + r;
        if (grow >= args.m_valid) continue;
        float f[4] = {acc.x + bv[0], acc.y + bv[1], acc.z + bv[2], acc.w + bv[3]};
        const long long ro = static_cast<long long>(grow) * args.ldo + ncol;
        if (args.addend != nullptr) {
          const __nv_bfloat16* ad = args.addend + ro;
          if (col_ok && (ro & 3) == 0) {
            const uint2 u = __ldg(reinterpret_cast<const uint2*>(ad));
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
            const float2 a0 = __bfloat1622float2(h[0]), a1 = __bfloat1622float2(h[1]);
            f[0] += a0.x; f[1] += a0.y; f[2] += a1.x; f[3] += a1.y;
          } else {
#pragma unroll
            for (int t = 0; t < 4; ++t) if (ncol + t < args.n_valid) f[t] += __bfloat162float(ad[t]);
          }
        }
        if (args.relu) {
#pragma unroll
          for (int t = 0; t < 4; ++t) f[t] = fmaxf(f[t], 0.f);
        }
        if (args.mask_y != nullptr) {
          const __nv_bfloat16* my = args.mask_y + ro;
          if (col_ok && (ro & 3) == 0) {
            const uint2 u = __ldg(reinterpret_cast<const uint2*>(my));
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
            const float2 a0 = __bfloat1622float2(h[0]), a1 = __bfloat1622float2(h[1]);
            f[0] = a0.x > 0.f ? f[0] : 0.f; f[1] = a0.y > 0.f ? f[1] : 0.f;
            f[2] = a1.x > 0.f ? f[2] : 0.f; f[3] = a1.y > 0.f ? f[3] : 0.f;
          } else {
#pragma unroll
            for (int t = 0; t < 4; ++t)
              if (ncol + t < args.n_valid) f[t] = __bfloat162float(my[t]) > 0.f ? f[t] : 0.f;
          }
        }
        if (args.out_f32) {
          float* o = static_cast<float*>(args.out) + ro;
          if (col_ok && (ro & 3) == 0) {
            *reinterpret_cast<float4*>(o) = make_float4(f[0], f[1], f[2], f[3]);
          } else {
#pragma unroll
            for (int t = 0; t < 4; ++t) if (ncol + t < args.n_valid) o[t] = f[t];
          }
        } else {
          __nv_bfloat16* o = static_cast<__nv_bfloat16*>(args.out) + ro;
          __nv_bfloat162 p0 = __floats2bfloat162_rn(f[0], f[1]);
          __nv_bfloat162 p1 = __floats2bfloat162_rn(f[2], f[3]);
          if (col_ok && (ro & 3) == 0) {
            uint2 u;
            u.x = *reinterpret_cast<uint32_t*>(&p0); u.y = *reinterpret_cast<uint32_t*>(&p1);
            *reinterpret_cast<uint2*>(o) = u;
          } else {
            const __nv_bfloat16 hv[4] = {p0.x, p0.y, p1.x, p1.y};
#pragma unroll
            for (int t = 0; t < 4; ++t) if (ncol + t < args.n_valid) o[t] = hv[t];
          }
          // statistics of exactly the values the consumer will read back
          const float2 r0 = __bfloat1622float2(p0), r1 = __bfloat1622float2(p1);
          f[0] = r0.x; f[1] = r0.y; f[2] = r1.x; f[3] = r1.y;
        }
        if (args.bn_x != nullptr) {                     // BatchNorm-backward sums: g and g * xhat
          float xh[4] = {0.f, 0.f, 0.f, 0.f};
          const __nv_bfloat16* bx = args.bn_x + ro;
          if (col_ok && (ro & 3) == 0) {
            const uint2 u = __ldg(reinterpret_cast<const uint2*>(bx));
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
            const float2 a0 = __bfloat1622float2(h[0]), a1 = __bfloat1622float2(h[1]);
            xh[0] = a0.x; xh[1] = a0.y; xh[2] = a1.x; xh[3] = a1.y;
          } else {
#pragma unroll
            for (int t = 0; t < 4; ++t) if (ncol + t < args.n_valid) xh[t] = __bfloat162float(bx[t]);
          }
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            s1[t] += f[t];
            s2[t] = fmaf(f[t], (xh[t] - bm[t]) * bi[t], s2[t]);
          }
        } else {
#pragma unroll
          for (int t = 0; t < 4; ++t) { s1[t] += f[t]; s2[t] = fmaf(f[t], f[t], s2[t]); }
        }
      }
      if (args.stats != nullptr) {
        // No shared-memory float atomics (they compile to compare-and-swap loops): lanes that share a
        // column group fold with a shuffle, each warp writes its own row of partials, then 4-way sums.
        if (kGroups == 16) {
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            s1[t] += __shfl_xor_sync(0xffffffffu, s1[t], 16);
            s2[t] += __shfl_xor_sync(0xffffffffu, s2[t], 16);
          }
        }
        float* wsm = st_sm + (warp - 2) * 2 * BLOCK_N;    // [4 warps][2][BLOCK_N]
        if (lane < kGroups) {
#pragma unroll
          for (int t = 0; t < 4; ++t) { wsm[c4 + t] = s1[t]; wsm[BLOCK_N + c4 + t] = s2[t]; }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps only
        for (int i = et; i < 2 * BLOCK_N; i += 128) {
          const int c = i % BLOCK_N, which = i / BLOCK_N;
          const int gc = nt * BLOCK_N + c;
          const float tot = st_sm[i] + st_sm[2 * BLOCK_N + i] + st_sm[4 * BLOCK_N + i] + st_sm[6 * BLOCK_N + i];
          if (gc < args.n_valid) atomicAdd(args.stats + static_cast<size_t>(which) * args.n_valid + gc, tot);
        }
      }
    }
    cluster_sync_all();                                  // peers may still be reading this CTA's partial tile
  }
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
  }
}

// --------------------------------------------------------------------------------------------
// Two-SM variant for large plain GEMMs (both operands K-major): a cluster of two CTAs computes a
// 256 x 256 tile with tcgen05.mma.cta_group::2.  Each CTA loads its own 128 rows of A and HALF of
// the B tile (128 of the 256 columns); the pair's MMA reads both halves, so a CTA moves 32 KB per
// 64-deep chunk for a 128 x 256 output instead of 48 KB - the operand traffic from L2 is what bounds
// the one-CTA kernel on large problems.  CTA 0 (the leader) issues all MMAs; TMA loads of both
// CTAs complete on the leader's "full" barrier; tcgen05.commit multicasts to the "empty" /
// "accumulator ready" barriers of both.  Each CTA then drains its own 128 accumulator lanes.
// --------------------------------------------------------------------------------------------
constexpr int k2smN = 256;
constexpr int k2smStages = 3;            // 3 x (16 + 16) KB: two CTAs (of different pairs) per SM
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // shared::cluster address of the same offset in CTA 0 of the pair

__device__ __forceinline__ void tma_load_4d_2sm(const CUtensorMap* map, uint64_t* leader_bar, void* dst, int c0, int c1,
                                                int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(leader_bar) & kPeerBitMask), "r"(c0),
      "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc),
      "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {      // arrives on `bar` in BOTH CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3)) : "memory");
}

struct Gemm2smArgs {
  int m_pairs, n_tiles, k_chunks;
  void* out;
  long long ldo;
  int m_valid, n_valid;
  const float* bias;
  int relu, out_f32;
};

__global__ void __launch_bounds__(kNumThreads, 2)
gemm2sm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const Gemm2smArgs args) {
  constexpr uint32_t kABytes = kBlockM * kBlockK * 2;            // this CTA's 128 rows of A
  constexpr uint32_t kBBytes = (k2smN / 2) * kBlockK * 2;        // this CTA's half of the B tile
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + k2smStages * kABytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_b + k2smStages * kBBytes);
  uint64_t* empty_bar = full_bar + k2smStages;
  uint64_t* tmem_full_bar = empty_bar + k2smStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  kml::pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  // Band rasterisation over the pair tiles (see the one-CTA kernel): bands of 8 pairs = 16 m-tiles,
  // n fastest inside a band.  Both CTAs of a pair compute the same mapping.
  int mt, nt;
  {
    constexpr int kGroupP = 8;
    const int lin = static_cast<int>(blockIdx.x >> 1) + args.m_pairs * static_cast<int>(blockIdx.y);
    const int band = lin / (kGroupP * args.n_tiles);
    const int band_p0 = band * kGroupP;
    const int band_rows = min(kGroupP, args.m_pairs - band_p0);
    const int in_band = lin - band * kGroupP * args.n_tiles;
    mt = 2 * (band_p0 + in_band % band_rows) + static_cast<int>(rank);   // 128-row tile of this CTA
    nt = in_band / band_rows;                                            // 256-column tile of the pair
  }
  const int nchunks = args.k_chunks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < k2smStages; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {   // the same warp of both CTAs allocates the pair's accumulator columns
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(static_cast<uint32_t>(k2smN)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  cluster_sync_all();                            // the peer's barriers exist before anything signals them
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  kml::pdl_wait();

  if (warp == 0) {
    if (elect_one()) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
      for (int i = 0; i < nchunks; ++i) {
        const int s = i % k2smStages;
        const uint32_t ph = (i / k2smStages) & 1;
        mbar_wait(empty_bar + s, ph ^ 1);
        // the leader's barrier counts the bytes of BOTH CTAs of the pair
        if (rank == 0) mbar_expect_tx(full_bar + s, 2 * (kABytes + kBBytes));
        tma_load_4d_2sm(&map_a, full_bar + s, smem_a + s * kABytes, i * kBlockK, mt * kBlockM, 0, 0);
        tma_load_4d_2sm(&map_b, full_bar + s, smem_b + s * kBBytes, i * kBlockK,
                        nt * k2smN + static_cast<int>(rank) * (k2smN / 2), 0, 0);
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc(2 * kBlockM, k2smN, false, false);
      for (int i = 0; i < nchunks; ++i) {
        const int s = i % k2smStages;
        const uint32_t ph = (i / k2smStages) & 1;
        mbar_wait(full_bar + s, ph);
        tcgen05_fence_after();
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(smem_a + s * kABytes);
          const uint32_t b_addr = smem_u32(smem_b + s * kBBytes);
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k)
            umma_bf16_2sm(tmem_base, make_smem_desc(a_addr + k * 32, 16, 1024), make_smem_desc(b_addr + k * 32, 16, 1024),
                          idesc, (i | k) != 0 ? 1u : 0u);
        }
        __syncwarp();
        if (elect_one()) {
          umma_commit_2sm(empty_bar + s);
          if (i == nchunks - 1) umma_commit_2sm(tmem_full_bar);
        }
        __syncwarp();
      }
    }
  } else {
    const int q = warp & 3;
    const int row = mt * kBlockM + q * 32 + lane;
    mbar_wait(tmem_full_bar, 0);
    tcgen05_fence_after();
    const bool row_ok = row < args.m_valid;
#pragma unroll 1
    for (int c0 = 0; c0 < k2smN; c0 += 32) {
      uint32_t v[32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c0);
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
            "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
            "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
            "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
          : "r"(taddr) : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (!row_ok) continue;
      const int ncol0 = nt * k2smN + c0;
      if (ncol0 >= args.n_valid) continue;
      const long long off = static_cast<long long>(row) * args.ldo + ncol0;
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        f[j] = __uint_as_float(v[j]);
        if (args.bias != nullptr && ncol0 + j < args.n_valid) f[j] += __ldg(args.bias + ncol0 + j);
        if (args.relu) f[j] = fmaxf(f[j], 0.f);
      }
      const bool full = ncol0 + 32 <= args.n_valid;
      if (args.out_f32) {
        float* o = static_cast<float*>(args.out) + off;
        if (full && (off & 3) == 0) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) if (ncol0 + j < args.n_valid) o[j] = f[j];
        }
      } else {
        __nv_bfloat16* o = static_cast<__nv_bfloat16*>(args.out) + off;
        if (full && (off & 7) == 0) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            __nv_bfloat162 p0 = __floats2bfloat162_rn(f[j], f[j + 1]);
            __nv_bfloat162 p1 = __floats2bfloat162_rn(f[j + 2], f[j + 3]);
            __nv_bfloat162 p2 = __floats2bfloat162_rn(f[j + 4], f[j + 5]);
            __nv_bfloat162 p3 = __floats2bfloat162_rn(f[j + 6], f[j + 7]);
            uint4 u;
            u.x = *reinterpret_cast<uint32_t*>(&p0); u.y = *reinterpret_cast<uint32_t*>(&p1);
            u.z = *reinterpret_cast<uint32_t*>(&p2); u.w = *reinterpret_cast<uint32_t*>(&p3);
            *reinterpret_cast<uint4*>(o + j) = u;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) if (ncol0 + j < args.n_valid) o[j] = __float2bfloat16_rn(f[j]);
        }
      }
    }
    tcgen05_fence_before();
  }
  __syncthreads();
  cluster_sync_all();                            // neither CTA frees the pair's TMEM while the other still reads
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"(static_cast<uint32_t>(k2smN)) : "memory");
  }
}

// --------------------------------------------------------------------------------------------
// host side
// --------------------------------------------------------------------------------------------

struct TensorSpec {          // rank-4 view, dim 0 innermost (contiguous)
  const void* ptr;
  uint64_t dim[4];
  uint64_t stride_bytes[3];  // strides of dims 1..3
  uint32_t box[4];           // extent walked in each dimension (elements of the tensor)
  uint32_t estride[4];       // traversal stride per dimension; 0 = 1.  A box of extent e with stride s delivers
                             // ceil(e / s) elements: how a strided convolution reads every s-th pixel
};

int encode_map(CUtensorMap* map, const TensorSpec& t) {
  const kml::DriverApi* d = kml::driver();
  if (!d) return 4;
  // Driver calls need a context current on *this* thread.  A thread that has not yet made a runtime
  // call (e.g. autograd's backward thread running us as its first op) has none: bind the primary one.
  CUcontext ctx = nullptr;
  if (d->cuCtxGetCurrent(&ctx) != CUDA_SUCCESS || ctx == nullptr) KML_CUDA(cudaFree(0));
  cuuint64_t dims[4] = {t.dim[0], t.dim[1], t.dim[2], t.dim[3]};
  cuuint64_t strides[3] = {t.stride_bytes[0], t.stride_bytes[1], t.stride_bytes[2]};
  cuuint32_t box[4] = {t.box[0], t.box[1], t.box[2], t.box[3]};
  cuuint32_t estr[4] = {t.estride[0] ? t.estride[0] : 1, t.estride[1] ? t.estride[1] : 1,
                        t.estride[2] ? t.estride[2] : 1, t.estride[3] ? t.estride[3] : 1};
  KML_CHECK((reinterpret_cast<uintptr_t>(t.ptr) & 15u) == 0, "tensor map base must be 16-byte aligned");
  for (int i = 0; i < 3; ++i) KML_CHECK(strides[i] % 16 == 0, "tensor map stride %d (%llu) not a multiple of 16 bytes", i, (unsigned long long)strides[i]);
  CUresult r = d->cuTensorMapEncodeTiled(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(t.ptr), dims,
                                         strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    kml::set_error("cuTensorMapEncodeTiled failed: %s (dims %llu,%llu,%llu,%llu box %u,%u,%u,%u)", kml::cu_err(r),
                   (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)dims[2],
                   (unsigned long long)dims[3], box[0], box[1], box[2], box[3]);
    return 3;
  }
  return 0;
}

template <int BLOCK_N, bool A_MN, bool B_MN, bool LINEAR, int BN_MODE = 0>
int launch_tiles(const CUtensorMap& ma, const CUtensorMap& mb, const GemmArgs& a, dim3 grid, cudaStream_t s) {
  constexpr size_t smem = Stages<BLOCK_N>::value * (kBlockM * kBlockK * 2 + BLOCK_N * kBlockK * 2) + 1024 /*align*/ + 640 /*barriers+taps*/;
  auto kernel = gemm_tcgen05_kernel<BLOCK_N, A_MN, B_MN, LINEAR, BN_MODE>;
  static kml::PerDeviceOnce configured;
  const int dev = kml::current_device();
  if (configured.need(dev)) {
    KML_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    KML_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));   // clusters of 16
    configured.mark(dev);
  }
  const dim3 cluster(1, 1, a.cluster_reduce ? a.splits : 1);
  KML_CUDA(kml::launch_pdl_cluster(kernel, grid, dim3(kNumThreads), smem, s, cluster, ma, mb, a));
  kml::count_launch();
  return 0;
}

template <int BLOCK_N, bool A_MN, bool B_MN>
int launch_variant(const CUtensorMap& ma, const CUtensorMap& mb, const GemmArgs& a, cudaStream_t s) {
  // Many tiles (more than two waves of resident CTAs): linear, band-rasterised tile index; with
  // BLOCK_N <= 128 also persistent (2 CTAs per SM stride over the tiles, double-buffered accumulator:
  // 2 x 2 x BLOCK_N <= 512 TMEM columns per SM).
  static const bool allow_persistent = [] { const char* e = getenv("KUBEML_GEMM_PERSISTENT"); return !(e && e[0] == '0'); }();
  const long long tiles = static_cast<long long>(a.m_tiles) * a.n_tiles * a.z_taps * a.splits;
  const int resident = 2 * kml::num_sms();
  if (!a.cluster_reduce && tiles > 2LL * resident) {
    GemmArgs b = a;
    dim3 grid(static_cast<unsigned>(tiles), 1, 1);
    if (allow_persistent && BLOCK_N <= 128) {
      b.persistent = 1;
      grid = dim3(resident, 1, 1);
    }
    return launch_tiles<BLOCK_N, A_MN, B_MN, true>(ma, mb, b, grid, s);
  }
  return launch_tiles<BLOCK_N, A_MN, B_MN, false>(ma, mb, a, dim3(a.m_tiles, a.n_tiles, a.z_taps * a.splits), s);
}

int launch(const TensorSpec& ta, const TensorSpec& tb, const GemmArgs& a, int block_n, cudaStream_t s) {
  CUtensorMap ma, mb;
  int rc = encode_map(&ma, ta);
  if (rc) return rc;
  rc = encode_map(&mb, tb);
  if (rc) return rc;
  const bool amn = a.a_kind == kMatMN || a.a_kind == kConvMN;
  const bool bmn = a.b_kind == kMatMN || a.b_kind == kConvMN || a.b_kind == kWgtMN;
  KML_CHECK(block_n == 64 || block_n == 128 || block_n == 256, "block_n must be 64, 128 or 256");
  if (block_n == 256) {      // 128x256 tiles: two 48 KB stages, two CTAs per SM use all 512 TMEM columns
    if (!amn && !bmn) return launch_variant<256, false, false>(ma, mb, a, s);
    if (!amn && bmn) return launch_variant<256, false, true>(ma, mb, a, s);
    return launch_variant<256, true, true>(ma, mb, a, s);
  }
  KML_CHECK(!(amn && !bmn), "A MN-major with B K-major is not instantiated");
  if (a.bn_mode != 0) {
    // The fused-BatchNorm epilogue lives in its own instantiations (forward: both operands K-major;
    // backward: the data-gradient form, B MN-major), so that every other launch keeps the lean kernel:
    // measured, one kernel carrying all epilogues cost every GEMM of the step ~1 us (168 vs 128
    // registers and a third more code to fetch on a cold instruction cache).
    const dim3 grid(a.m_tiles, a.n_tiles, a.z_taps * a.splits);
    KML_CHECK(!amn && a.cluster_reduce && ((a.bn_mode == 1 && !bmn) || (a.bn_mode == 2 && bmn)),
              "bn_fuse: forward takes K-major operands, backward the data-gradient form (B MN-major)");
    if (a.bn_mode == 1)
      return block_n == 64 ? launch_tiles<64, false, false, false, 1>(ma, mb, a, grid, s)
                           : launch_tiles<128, false, false, false, 1>(ma, mb, a, grid, s);
    return block_n == 64 ? launch_tiles<64, false, true, false, 2>(ma, mb, a, grid, s)
                         : launch_tiles<128, false, true, false, 2>(ma, mb, a, grid, s);
  }
  if (block_n == 64) {
    if (!amn && !bmn) return launch_variant<64, false, false>(ma, mb, a, s);
    if (!amn && bmn) return launch_variant<64, false, true>(ma, mb, a, s);
    return launch_variant<64, true, true>(ma, mb, a, s);
  }
  if (!amn && !bmn) return launch_variant<128, false, false>(ma, mb, a, s);
  if (!amn && bmn) return launch_variant<128, false, true>(ma, mb, a, s);
  return launch_variant<128, true, true>(ma, mb, a, s);
}

int launch_2sm(const TensorSpec& ta, const TensorSpec& tb, const Gemm2smArgs& a, cudaStream_t s) {
  CUtensorMap ma, mb;
  int rc = encode_map(&ma, ta);
  if (rc) return rc;
  rc = encode_map(&mb, tb);
  if (rc) return rc;
  constexpr size_t smem = k2smStages * (kBlockM * kBlockK * 2 + (k2smN / 2) * kBlockK * 2) + 1024 + 256;
  static kml::PerDeviceOnce configured;
  const int dev = kml::current_device();
  if (configured.need(dev)) {
    KML_CUDA(cudaFuncSetAttribute(gemm2sm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    configured.mark(dev);
  }
  KML_CUDA(kml::launch_pdl_cluster(gemm2sm_kernel, dim3(2 * a.m_pairs, a.n_tiles, 1), dim3(kNumThreads), smem, s,
                                   dim3(2, 1, 1), ma, mb, a));
  kml::count_launch();
  return 0;
}

inline int ceil_div(long long a, long long b) { return static_cast<int>((a + b - 1) / b); }

// Copies a kml_bn_fuse request into the kernel arguments (cluster mode only) and checks it.
int apply_bn_fuse(GemmArgs& g, const kml_epilogue* epi, int block_n) {
  const kml_bn_fuse* f = epi->bn_fuse;
  if (f == nullptr || f->mode == 0) return 0;
  KML_CHECK(g.cluster_reduce, "bn_fuse needs cluster mode");
  KML_CHECK(f->mode == 1 || f->mode == 2, "bn_fuse: mode must be 1 (forward) or 2 (backward)");
  KML_CHECK(!g.out_f32, "bn_fuse: bf16 output only");
  KML_CHECK(g.n_valid <= kml::kBnMaxC, "bn_fuse: at most %d channels", kml::kBnMaxC);
  KML_CHECK(g.n_valid % 4 == 0 && g.ldo % 4 == 0 && g.bias == nullptr,
            "bn_fuse: channels and the output row stride must be multiples of 4, no bias (BatchNorm absorbs it)");
  KML_CHECK(f->scratch && f->gamma && f->save_mean && f->save_invstd, "bn_fuse: null pointer");
  KML_CHECK(epi->stats == nullptr && epi->bn_x == nullptr && epi->mask_y == nullptr,
            "bn_fuse replaces stats / mask_y / bn_x of the epilogue");
  g.bn_mode = f->mode;
  g.bn_sc = static_cast<kml::BnScratch*>(f->scratch);
  g.bn_gamma = f->gamma;
  g.bn_raw = static_cast<__nv_bfloat16*>(f->raw_out);
  if (f->mode == 1) {
    KML_CHECK(f->beta && f->raw_out, "bn_fuse forward: beta and raw_out are required");
    KML_CHECK(!g.relu && g.addend == nullptr, "bn_fuse forward: ReLU / residual belong to bn_fuse, not to the GEMM epilogue");
    g.bn_beta = f->beta; g.bn_rmean = f->running_mean; g.bn_rvar = f->running_var; g.bn_nbt = f->num_batches_tracked;
    g.bn_save_mean = f->save_mean; g.bn_save_invstd = f->save_invstd;
    g.bn_res = static_cast<const __nv_bfloat16*>(f->res);
    g.bn_relu = f->relu; g.bn_eps = f->eps; g.bn_momentum = f->momentum;
  } else {
    KML_CHECK(f->bn_x && f->dgamma && f->dbeta, "bn_fuse backward: bn_x, dgamma and dbeta are required");
    g.mask_y = static_cast<const __nv_bfloat16*>(f->mask_y);
    g.bn_x = static_cast<const __nv_bfloat16*>(f->bn_x);
    g.bn_mean = f->save_mean; g.bn_invstd = f->save_invstd;
    g.bn_dgamma = f->dgamma; g.bn_dbeta = f->dbeta;
  }
  const long long ctas = static_cast<long long>(g.m_tiles) * g.n_tiles * g.z_taps * g.splits;
  const bool bmn = g.b_kind == kMatMN || g.b_kind == kConvMN || g.b_kind == kWgtMN;
  const bool amn = g.a_kind == kMatMN || g.a_kind == kConvMN;
  const int cap = kml_gemm_max_coresident(block_n, amn, bmn, g.splits);
  KML_CHECK(cap > 0 && ctas <= cap, "bn_fuse: %lld CTAs (clusters of %d) exceed the co-resident limit %d of this device",
            ctas, g.splits, cap);
  return 0;
}
inline bool pow2(int x) { return x > 0 && (x & (x - 1)) == 0; }

// Pixel box of `target` pixels over a W x H x N grid: whole rows first, then rows, then images.
bool make_geometry(int W, int H, int target, Geometry* g) {
  if (!pow2(W) || !pow2(H) || W > target) return false;
  g->bw = W;
  g->bh = H < target / W ? H : target / W;
  g->bn = target / (g->bw * g->bh);
  g->tiles_w = 1;
  g->tiles_h = H / g->bh;
  return true;
}

}  // namespace

namespace {
template <int BLOCK_N, bool B_MN, int BN_MODE>
int coresident_of(int cluster_size) {
  constexpr size_t smem = Stages<BLOCK_N>::value * (kBlockM * kBlockK * 2 + BLOCK_N * kBlockK * 2) + 1024 + 640;
  auto kernel = gemm_tcgen05_kernel<BLOCK_N, false, B_MN, false, BN_MODE>;
  if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess) return -1;
  cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  kml::note_kernel(reinterpret_cast<const void*>(kernel));
  int api_per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&api_per_sm, kernel, kNumThreads, smem) != cudaSuccess) return -1;
  if (api_per_sm < 1) api_per_sm = 1;
  // The runtime's occupancy calculator answers 1 CTA per SM for these kernels; ncu's launch__occupancy_limit_*
  // (registers 2, shared memory 2) and the hardware say 2: a 256-CTA grid of the fused kernel passes its
  // grid-wide barrier on 148 SMs (profiles/r2).  Count what the resources allow - registers in 8-register
  // units per thread, dynamic + 1 KB reserved shared memory against the 228 KB of an SM - and scale the API's
  // answer by it (KUBEML_FUSED_CAP_SCALE=1 keeps the raw API number).  Every spin is bounded: a grid that does
  // not fit fails the launch instead of hanging.
  cudaFuncAttributes fa;
  if (cudaFuncGetAttributes(&fa, kernel) != cudaSuccess) return -1;
  int dev = kml::current_device(), smem_sm = 0, regs_sm = 0;
  cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
  cudaDeviceGetAttribute(&regs_sm, cudaDevAttrMaxRegistersPerMultiprocessor, dev);
  const int regs_cta = ((fa.numRegs + 7) / 8 * 8) * kNumThreads;
  int by_res = std::min(smem_sm / static_cast<int>(smem + 1024), regs_cta > 0 ? regs_sm / regs_cta : 1);
  by_res = std::max(1, std::min(by_res, 2));
  static const int want_scale = [] { const char* e = getenv("KUBEML_FUSED_CAP_SCALE"); return e ? atoi(e) : 2; }();
  const int scale = std::max(1, std::min(want_scale, by_res / api_per_sm));
  if (cluster_size <= 1) return api_per_sm * scale * kml::num_sms();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(1, 1, cluster_size);
  cfg.blockDim = dim3(kNumThreads);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = cluster_size;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int clusters = 0;
  if (cudaOccupancyMaxActiveClusters(&clusters, kernel, &cfg) != cudaSuccess) return -1;
  return clusters * cluster_size * scale;
}
}  // namespace

extern "C" {

// Resident CTAs per SM of one instantiation as the occupancy calculator sees it (diagnostics).
int kml_gemm_blocks_per_sm(int block_n, int b_mn, int bn_mode) {
  constexpr size_t smem64 = Stages<64>::value * (kBlockM * kBlockK * 2 + 64 * kBlockK * 2) + 1024 + 640;
  int per_sm = -1;
  if (block_n != 64) return -1;
  const void* k = bn_mode == 0 ? (b_mn ? (const void*)gemm_tcgen05_kernel<64, false, true, false, 0> : (const void*)gemm_tcgen05_kernel<64, false, false, false, 0>)
                 : bn_mode == 1 ? (const void*)gemm_tcgen05_kernel<64, false, false, false, 1>
                                : (const void*)gemm_tcgen05_kernel<64, false, true, false, 2>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem64));
  kml::note_kernel(k);
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, k);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, kNumThreads, smem64);
  kml::set_error("regs %d static smem %zu max dyn %d carveout %d blocks/SM %d (dyn smem %zu)", fa.numRegs, fa.sharedSizeBytes,
                 fa.maxDynamicSharedSizeBytes, fa.preferredShmemCarveout, per_sm, smem64);
  return per_sm;
}

int kml_gemm_max_coresident(int block_n, int a_mn, int b_mn, int cluster_size) {
  // cached per (device, variant, cluster size): the occupancy queries are host-side only but not free
  static std::mutex mu;
  static std::unordered_map<long long, int> cache;
  const long long key = (static_cast<long long>(kml::current_device()) << 32) | (block_n << 8) | (a_mn << 7) | (b_mn << 6) |
                        (cluster_size & 63);
  {
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
  }
  int v = -1;
  if (!a_mn) {     // the fused-BatchNorm instantiations: forward (B K-major) and backward (B MN-major)
    if (block_n == 64) v = b_mn ? coresident_of<64, true, 2>(cluster_size) : coresident_of<64, false, 1>(cluster_size);
    else if (block_n == 128) v = b_mn ? coresident_of<128, true, 2>(cluster_size) : coresident_of<128, false, 1>(cluster_size);
  }
  if (v <= 0) { cudaGetLastError(); kml::set_error("kml_gemm_max_coresident: occupancy query failed"); }
  std::lock_guard<std::mutex> lock(mu);
  cache[key] = v;
  return v;
}

int kml_gemm_bf16(const void* A, int a_mn, long long lda, const void* B, int b_mn, long long ldb, void* out,
                  long long ldo, int M, int N, int K, const kml_epilogue* epi, void* stream) {
  const kml_epilogue dflt = {};
  if (!epi) epi = &dflt;
  const float* bias = epi->bias;
  const int relu = epi->relu, out_f32 = epi->out_f32;
  int splits = epi->splits, block_n = epi->block_n;
  KML_CHECK(A && B && out && M > 0 && N > 0 && K > 0, "kml_gemm_bf16: bad arguments");
  KML_CHECK(lda % 8 == 0 && ldb % 8 == 0, "kml_gemm_bf16: leading dimensions must be multiples of 8 elements");
  if (splits < 1) splits = 1;
  // Large plain GEMM: the two-SM kernel (256 x 256 tiles per CTA pair, tcgen05.mma.cta_group::2).
  // block_n == 512 asks for it explicitly; the default takes it from four waves of pair tiles upward.
  static const bool allow_2sm = [] { const char* e = getenv("KUBEML_GEMM_2SM"); return !(e && e[0] == '0'); }();
  const long long pair_tiles = static_cast<long long>(ceil_div(M, 256)) * ceil_div(N, 256);
  const bool plain = !a_mn && !b_mn && splits == 1 && !epi->cluster && !epi->addend_bf16 && !epi->slabs;
  if (plain && (block_n == 512 || (block_n == 0 && allow_2sm && pair_tiles >= 4LL * kml::num_sms()))) {
    TensorSpec ta{}, tb{};
    ta.ptr = A; tb.ptr = B;
    ta.dim[0] = K; ta.dim[1] = M; ta.dim[2] = ta.dim[3] = 1;
    ta.stride_bytes[0] = static_cast<uint64_t>(lda) * 2; ta.stride_bytes[1] = ta.stride_bytes[0] * M;
    ta.stride_bytes[2] = ta.stride_bytes[1];
    ta.box[0] = 64; ta.box[1] = kBlockM; ta.box[2] = ta.box[3] = 1;
    tb.dim[0] = K; tb.dim[1] = N; tb.dim[2] = tb.dim[3] = 1;
    tb.stride_bytes[0] = static_cast<uint64_t>(ldb) * 2; tb.stride_bytes[1] = tb.stride_bytes[0] * N;
    tb.stride_bytes[2] = tb.stride_bytes[1];
    tb.box[0] = 64; tb.box[1] = k2smN / 2; tb.box[2] = tb.box[3] = 1;
    Gemm2smArgs g2{};
    g2.m_pairs = ceil_div(M, 256); g2.n_tiles = ceil_div(N, k2smN); g2.k_chunks = ceil_div(K, kBlockK);
    g2.out = out; g2.ldo = ldo; g2.m_valid = M; g2.n_valid = N;
    g2.bias = bias; g2.relu = relu; g2.out_f32 = out_f32;
    return launch_2sm(ta, tb, g2, static_cast<cudaStream_t>(stream));
  }
  KML_CHECK(block_n != 512, "kml_gemm_bf16: the two-SM kernel takes K-major operands, no split-K, no addend");
  if (block_n == 0) block_n = (N > 64) ? 128 : 64;
  TensorSpec ta{}, tb{};
  ta.ptr = A; tb.ptr = B;
  // K-major [rows][K]: dim0 = K, dim1 = rows.  MN-major [K][rows]: dim0 = rows, dim1 = K.
  ta.dim[0] = a_mn ? M : K; ta.dim[1] = a_mn ? K : M; ta.dim[2] = ta.dim[3] = 1;
  ta.stride_bytes[0] = static_cast<uint64_t>(lda) * 2; ta.stride_bytes[1] = ta.stride_bytes[0] * ta.dim[1];
  ta.stride_bytes[2] = ta.stride_bytes[1];
  ta.box[0] = 64; ta.box[1] = a_mn ? kBlockK : kBlockM; ta.box[2] = ta.box[3] = 1;
  tb.dim[0] = b_mn ? N : K; tb.dim[1] = b_mn ? K : N; tb.dim[2] = tb.dim[3] = 1;
  tb.stride_bytes[0] = static_cast<uint64_t>(ldb) * 2; tb.stride_bytes[1] = tb.stride_bytes[0] * tb.dim[1];
  tb.stride_bytes[2] = tb.stride_bytes[1];
  tb.box[0] = 64; tb.box[1] = b_mn ? kBlockK : block_n; tb.box[2] = tb.box[3] = 1;
  GemmArgs g{};
  g.a_kind = a_mn ? kMatMN : kMatK;
  g.b_kind = b_mn ? kMatMN : kMatK;
  g.m_tiles = ceil_div(M, kBlockM);
  g.n_tiles = ceil_div(N, block_n);
  g.k_chunks = ceil_div(K, kBlockK);
  g.splits = splits > g.k_chunks ? g.k_chunks : splits;
  g.z_taps = 1; g.kw = 1; g.cblocks = 1; g.conv_stride = 1;
  g.out = out; g.ldo = ldo; g.m_valid = M; g.n_valid = N;
  g.bias = bias; g.relu = relu; g.out_f32 = out_f32;
  g.addend = static_cast<const __nv_bfloat16*>(epi->addend_bf16);
  g.atomic_add = g.splits > 1 && !epi->slabs;
  g.slab_stride = (g.splits > 1 && epi->slabs) ? static_cast<long long>(M) * ldo : 0;
  if (epi->cluster) {
    KML_CHECK(g.splits <= 16 && block_n <= 128, "cluster split-K: at most 16 splits, block_n <= 128");
    g.cluster_reduce = 1; g.stats = epi->stats; g.atomic_add = 0; g.slab_stride = 0;
    g.mask_y = static_cast<const __nv_bfloat16*>(epi->mask_y);
    g.bn_x = static_cast<const __nv_bfloat16*>(epi->bn_x);
    g.bn_mean = epi->bn_mean; g.bn_invstd = epi->bn_invstd;
    KML_CHECK(!g.bn_x || (g.bn_mean && g.bn_invstd && g.stats), "bn_x needs bn_mean, bn_invstd and stats");
  } else {
    KML_CHECK(epi->stats == nullptr && epi->mask_y == nullptr && epi->bn_x == nullptr,
              "column statistics / masks need cluster mode");
    KML_CHECK(g.splits == 1 || (out_f32 && !relu), "split-K needs an fp32 output and no ReLU");
  }
  if (int rc = apply_bn_fuse(g, epi, block_n)) return rc;
  return launch(ta, tb, g, block_n, static_cast<cudaStream_t>(stream));
}

static int conv_like(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                     int pad, int stride, const kml_epilogue* epi, bool dgrad, void* stream);

int kml_conv2d_fprop(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                     int pad, const kml_epilogue* epi, void* stream) {
  return conv_like(x, w, y, N, H, W, Cin, Cout, KH, KW, pad, 1, epi, false, stream);
}

// Strided convolution as the same implicit GEMM: the activation tensor map walks the input with element
// stride `stride` in W and H (cuTensorMapEncodeTiled elementStrides), the box origin is stride * (output
// pixel) + tap - pad, out-of-range pixels are zero-filled by TMA.  No im2col buffer exists.
int kml_conv2d_fprop_strided(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout, int KH,
                             int KW, int pad, int stride, const kml_epilogue* epi, void* stream) {
  KML_CHECK(stride >= 1 && stride <= 8, "kml_conv2d_fprop_strided: stride must be in 1..8");
  return conv_like(x, w, y, N, H, W, Cin, Cout, KH, KW, pad, stride, epi, false, stream);
}

// Data gradient of the stride-1 convolution y = conv(x, w, pad): dx[N,H,W,Cin] from dy[N,Ho,Wo,Cout]
// and the *forward* filter w[Cout][KH][KW][Cin] (read MN-major through the reflected tap, so no
// transposed copy of the weights exists).
int kml_conv2d_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                     int pad, const kml_epilogue* epi, void* stream) {
  const int Ho = H + 2 * pad - KH + 1, Wo = W + 2 * pad - KW + 1;
  // as a forward conv: input dy [N,Ho,Wo,Cout], output dx [N,H,W,Cin], padding K-1-pad
  KML_CHECK(KH == KW, "kml_conv2d_dgrad: square filters only");
  return conv_like(dy, w, dx, N, Ho, Wo, Cout, Cin, KH, KW, KH - 1 - pad, 1, epi, true, stream);
}

static int conv_like(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                     int pad, int stride, const kml_epilogue* epi, bool dgrad, void* stream) {
  const kml_epilogue dflt = {};
  if (!epi) epi = &dflt;
  const float* bias = epi->bias;
  const int relu = epi->relu, out_f32 = epi->out_f32;
  int splits = epi->splits, block_n = epi->block_n;
  KML_CHECK(x && w && y, "kml_conv2d_fprop: null pointer");
  KML_CHECK(Cin % 64 == 0, "kml_conv2d_fprop: Cin (%d) must be a multiple of 64 (use the im2col path otherwise)", Cin);
  KML_CHECK(Cout % 8 == 0, "kml_conv2d_fprop: Cout (%d) must be a multiple of 8", Cout);
  const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
  KML_CHECK(Ho > 0 && Wo > 0 && KH * KW <= 64, "kml_conv2d_fprop: bad geometry");
  GemmArgs g{};
  KML_CHECK(make_geometry(Wo, Ho, kBlockM, &g.tile_geo),
            "kml_conv2d_fprop: output %dx%d is not a power-of-two grid (use the im2col path)", Ho, Wo);
  if (block_n == 0) block_n = (Cout > 64) ? 128 : 64;
  if (splits < 1) splits = 1;
  TensorSpec ta{}, tb{};
  ta.ptr = x;
  ta.dim[0] = Cin; ta.dim[1] = W; ta.dim[2] = H; ta.dim[3] = N;
  ta.stride_bytes[0] = static_cast<uint64_t>(Cin) * 2; ta.stride_bytes[1] = ta.stride_bytes[0] * W;
  ta.stride_bytes[2] = ta.stride_bytes[1] * H;
  ta.box[0] = 64; ta.box[1] = g.tile_geo.bw * stride; ta.box[2] = g.tile_geo.bh * stride; ta.box[3] = g.tile_geo.bn;
  ta.estride[1] = ta.estride[2] = stride;
  KML_CHECK(ta.box[1] <= 256 && ta.box[2] <= 256, "kml_conv2d_fprop: output tile x stride exceeds the TMA box limit");
  const long long Kt = static_cast<long long>(KH) * KW * Cin;
  tb.ptr = w;
  tb.dim[0] = Kt; tb.dim[1] = Cout; tb.dim[2] = tb.dim[3] = 1;
  tb.stride_bytes[0] = static_cast<uint64_t>(Kt) * 2; tb.stride_bytes[1] = tb.stride_bytes[0] * Cout;
  tb.stride_bytes[2] = tb.stride_bytes[1];
  tb.box[0] = 64; tb.box[1] = block_n; tb.box[2] = tb.box[3] = 1;
  // Taps whose shifted window lies entirely in the zero padding contribute nothing: skip them.
  unsigned long long mask = 0;
  int ntaps = 0;
  for (int kh = 0; kh < KH; ++kh)
    for (int kw = 0; kw < KW; ++kw) {
      // some output row / column reads a real pixel through this tap: s*o + k - pad in [0, H) for an o in [0, Ho)
      auto hits = [&](int k, int extent, int outs) {
        for (int o = 0; o < outs; ++o) { const int i = stride * o + k - pad; if (i >= 0 && i < extent) return true; }
        return false;
      };
      const bool h_hit = hits(kh, H, Ho);
      const bool w_hit = hits(kw, W, Wo);
      if (h_hit && w_hit) { mask |= 1ull << (kh * KW + kw); ++ntaps; }
    }
  g.a_kind = kConvK; g.b_kind = kMatK;
  g.tap_mask = mask; g.kw = KW; g.pad_w = pad; g.pad_h = pad; g.cblocks = Cin / 64; g.conv_stride = stride;
  const long long Mtot = static_cast<long long>(N) * Ho * Wo;
  const int imgs_per_tile = g.tile_geo.bn;
  g.m_tiles = g.tile_geo.tiles_h * ceil_div(N, imgs_per_tile);
  g.n_tiles = ceil_div(Cout, block_n);
  g.k_chunks = ntaps * g.cblocks;
  g.splits = splits > g.k_chunks ? g.k_chunks : splits;
  g.z_taps = 1;
  g.out = y; g.ldo = Cout; g.m_valid = static_cast<int>(Mtot); g.n_valid = Cout;
  g.bias = bias; g.relu = relu; g.out_f32 = out_f32;
  g.atomic_add = g.splits > 1 && !epi->slabs;
  g.slab_stride = (g.splits > 1 && epi->slabs) ? Mtot * Cout : 0;
  if (epi->cluster) {
    KML_CHECK(g.splits <= 16 && block_n <= 128, "cluster split-K: at most 16 splits, block_n <= 128");
    g.cluster_reduce = 1; g.stats = epi->stats; g.atomic_add = 0; g.slab_stride = 0;
    g.mask_y = static_cast<const __nv_bfloat16*>(epi->mask_y);
    g.bn_x = static_cast<const __nv_bfloat16*>(epi->bn_x);
    g.bn_mean = epi->bn_mean; g.bn_invstd = epi->bn_invstd;
    KML_CHECK(!g.bn_x || (g.bn_mean && g.bn_invstd && g.stats), "bn_x needs bn_mean, bn_invstd and stats");
  } else {
    KML_CHECK(epi->stats == nullptr && epi->mask_y == nullptr && epi->bn_x == nullptr,
              "column statistics / masks need cluster mode");
    KML_CHECK(g.splits == 1 || (out_f32 && !relu), "split-K needs an fp32 output and no ReLU");
  }
  g.addend = static_cast<const __nv_bfloat16*>(epi->addend_bf16);
  if (dgrad) {
    // B = forward filter [rows = real Cout][cols = taps * real Cin], consumed MN-major
    tb.dim[0] = static_cast<uint64_t>(KH) * KW * Cout; tb.dim[1] = Cin;
    tb.stride_bytes[0] = tb.dim[0] * 2; tb.stride_bytes[1] = tb.stride_bytes[0] * Cin;
    tb.stride_bytes[2] = tb.stride_bytes[1];
    tb.box[0] = 64; tb.box[1] = kBlockK;
    g.b_kind = kWgtMN;
    g.wgt_taps = KH * KW;
    g.wgt_cin = Cout;
  }
  if (int rc = apply_bn_fuse(g, epi, block_n)) return rc;
  return launch(ta, tb, g, block_n, static_cast<cudaStream_t>(stream));
}

int kml_conv2d_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                     int pad, int splits, int block_n, void* stream) {
  return kml_conv2d_wgrad_strided(x, dy, dw, N, H, W, Cin, Cout, KH, KW, pad, 1, splits, block_n, stream);
}

int kml_conv2d_wgrad_strided(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int KH,
                             int KW, int pad, int stride, int splits, int block_n, void* stream) {
  KML_CHECK(x && dy && dw, "kml_conv2d_wgrad: null pointer");
  KML_CHECK(stride >= 1 && stride <= 8, "kml_conv2d_wgrad: stride must be in 1..8");
  KML_CHECK(Cin % 8 == 0 && Cout % 8 == 0, "kml_conv2d_wgrad: channel counts must be multiples of 8");
  const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
  KML_CHECK(Ho > 0 && Wo > 0, "kml_conv2d_wgrad: bad geometry");
  GemmArgs g{};
  KML_CHECK(make_geometry(Wo, Ho, kBlockK, &g.chunk_geo),
            "kml_conv2d_wgrad: output %dx%d is not a power-of-two grid (use the im2col path)", Ho, Wo);
  if (block_n == 0) block_n = (Cin > 64) ? 128 : 64;
  TensorSpec ta{}, tb{};
  ta.ptr = dy;   // A = dY^T: M = Cout (contiguous), reduction over output pixels
  ta.dim[0] = Cout; ta.dim[1] = Wo; ta.dim[2] = Ho; ta.dim[3] = N;
  ta.stride_bytes[0] = static_cast<uint64_t>(Cout) * 2; ta.stride_bytes[1] = ta.stride_bytes[0] * Wo;
  ta.stride_bytes[2] = ta.stride_bytes[1] * Ho;
  ta.box[0] = 64; ta.box[1] = g.chunk_geo.bw; ta.box[2] = g.chunk_geo.bh; ta.box[3] = g.chunk_geo.bn;
  tb.ptr = x;    // B = X shifted by the tap: N = Cin (contiguous)
  tb.dim[0] = Cin; tb.dim[1] = W; tb.dim[2] = H; tb.dim[3] = N;
  tb.stride_bytes[0] = static_cast<uint64_t>(Cin) * 2; tb.stride_bytes[1] = tb.stride_bytes[0] * W;
  tb.stride_bytes[2] = tb.stride_bytes[1] * H;
  tb.box[0] = 64; tb.box[1] = g.chunk_geo.bw * stride; tb.box[2] = g.chunk_geo.bh * stride; tb.box[3] = g.chunk_geo.bn;
  tb.estride[1] = tb.estride[2] = stride;     // the input is read at stride * (output pixel) + tap - pad
  KML_CHECK(tb.box[1] <= 256 && tb.box[2] <= 256, "kml_conv2d_wgrad: pixel box x stride exceeds the TMA box limit");
  g.a_kind = kConvMN; g.b_kind = kConvMN;
  g.kw = KW; g.pad_w = pad; g.pad_h = pad; g.cblocks = 1; g.conv_stride = stride;
  g.m_tiles = ceil_div(Cout, kBlockM);
  g.n_tiles = ceil_div(Cin, block_n);
  g.k_chunks = g.chunk_geo.tiles_h * ceil_div(N, g.chunk_geo.bn);
  if (splits < 1) splits = 1;
  g.splits = splits > g.k_chunks ? g.k_chunks : splits;
  g.z_taps = KH * KW;
  g.out = dw; g.ldo = static_cast<long long>(KH) * KW * Cin; g.z_col_stride = Cin;
  g.m_valid = Cout; g.n_valid = Cin;
  g.out_f32 = 1; g.atomic_add = g.splits > 1;
  return launch(ta, tb, g, block_n, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
