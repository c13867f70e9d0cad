// K-AVG model merge over NVLink peer memory — one kernel per rank, no NCCL.
//
// What the reference does per sync round: every worker serialises its state_dict to
// RedisAI, the Go job fetches all N copies, sums them on the CPU under a mutex
// (ml/pkg/model/model.go:249-302), divides by N (ml/pkg/model/parallelSGD.go:26-54), writes the result
// back (model.go:135-161) and every worker downloads it again (python/kubeml/kubeml/network.py:416-442)
// after an HTTP barrier (ml/pkg/train/api.go:100-126): (3N+1)x|model| bytes through one TCP endpoint.
//
// Here every replica lives in a VMM arena mapped into all peers.  Each participating rank runs
// this kernel on its own GPU:
//   phase 0  (optional) last local SGD step of the round on the own replica, fused
//   barrier  release/acquire flags in peer memory: all replicas are final
//   phase 1  the rank owns 1/P of the buffer: for each float4 it loads that element from every
//            participant (plain peer loads, or one multimem.ld_reduce through the switch), sums in
//            rank order in fp32, scales by 1/P and stores the result into *every* replica
//            (peer stores, or one multimem.st); int64 counters are summed and divided with
//            truncation like parallelSGD.go:42-43
//   phase 2  momentum buffer zeroed (the reference's optimizer reset, network.py:121-128)
//   barrier  all writes have landed before anyone runs its next local step
#include "common.cuh"
#include "../include/kubeml_b200.h"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace {

constexpr int kThreads = 512;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ float4 ld_cg_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_cg_f4(float4* p, const float4& v) {
  asm volatile("st.global.cg.v4.f32 [%0], {%1,%2,%3,%4};"
               ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float4* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f4(float4* p, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// ---------------------------------------------------------------------------------------------------
// Signal pad of a rank (64 uint32 words inside its arena, mapped by every peer).
//   ready[r]      generation for which rank r's replica is final       (written by r into every pad)
//   done[r]       generation for which rank r has written its slice    (written by r into every pad)
//   count..       local grid barrier of this GPU (monotonic counters + their values at launch)
//   go_gen/mask   decision of this GPU's CTA 0 for a generation: the effective participant mask, or abort
//   fin_*         CTAs of this GPU that finished writing (the last one publishes done[])
//   dead_from[r]  != 0: rank r takes no part from that generation on (written by r itself when its
//                 function fails - `kml_merge_tombstone` - or by a survivor that timed out on it)
// Generations only grow within a job; all comparisons are signed differences.
// ---------------------------------------------------------------------------------------------------
struct MergePad {
  uint32_t ready[KML_MAX_RANKS];
  uint32_t done[KML_MAX_RANKS];
  uint32_t count, release, count_base, release_base;
  uint32_t go_gen, go_mask, fin_count, fin_base;
  uint32_t dead_from[KML_MAX_RANKS];
  uint32_t pad0[16];
  unsigned long long wait_ns;      // words 48-49: time this GPU's merges spent waiting for peers to become ready
  unsigned long long waits;        // words 50-51: number of merges counted in wait_ns          (host resets both)
};
static_assert(sizeof(MergePad) <= 64 * sizeof(uint32_t), "signal pad is 64 words");

constexpr uint32_t kGoAbort = 0x80000000u;

__device__ __forceinline__ bool reached(uint32_t flag, uint32_t gen) { return static_cast<int32_t>(flag - gen) >= 0; }

// Reduce [begin, end) of NP source replicas (compact list, ascending rank) and store the average
// to ND destinations.  U elements per thread are loaded before any is used: NP*U 16-byte loads in
// flight per thread hide the ~2 us NVLink round trip.  Sum in list order, true division: bit-parity
// with the CPU oracle (ml/pkg/model/model.go:286-296, parallelSGD.go:31-50).
template <int NP, int U>
__device__ __forceinline__ void reduce_range(const float4* const* src, float4* const* dst, int nd, size_t begin,
                                             size_t end, size_t tid, size_t nthreads, float fn) {
  const float4* s_[NP];
#pragma unroll
  for (int j = 0; j < NP; ++j) s_[j] = src[j];
  auto finish = [&](size_t idx, const float4 (&v)[NP]) {
    float4 a = v[0];
#pragma unroll
    for (int j = 1; j < NP; ++j) { a.x += v[j].x; a.y += v[j].y; a.z += v[j].z; a.w += v[j].w; }
    a.x = a.x / fn; a.y = a.y / fn; a.z = a.z / fn; a.w = a.w / fn;
    for (int d = 0; d < nd; ++d) st_cg_f4(dst[d] + idx, a);
  };
  size_t i = begin + tid;
  for (; i + (U - 1) * nthreads < end; i += U * nthreads) {
    float4 v[U][NP];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int j = 0; j < NP; ++j) v[u][j] = ld_cg_f4(s_[j] + i + u * nthreads);
#pragma unroll
    for (int u = 0; u < U; ++u) finish(i + u * nthreads, v[u]);
  }
  for (; i < end; i += nthreads) {
    float4 v[NP];
#pragma unroll
    for (int j = 0; j < NP; ++j) v[j] = ld_cg_f4(s_[j] + i);
    finish(i, v);
  }
}

// One kernel per rank and sync round.  Structure (multi-rank case):
//   phase 0  (optional) the round's last local SGD step on the own replica        [all CTAs]
//            local grid barrier (only if phase 0 ran: the stream already orders everything else)
//   signal   CTA 0 publishes ready[me] = gen in every pad, waits until every participant is ready
//            (bounded), reads the dead table, decides the EFFECTIVE participant mask (or abort) and
//            publishes it in the own pad; the other CTAs poll that one local word.  One decision
//            point per GPU: either every CTA merges with the same mask or none touches anything.
//   phase 1  the rank owns 1/P' of the buffer: loads that slice from every effective participant
//            (peer loads, or one multimem.ld_reduce), averages, stores it into every replica
//   done     the last CTA to finish publishes done[me] = gen in every pad (no grid barrier)
//   phase 3  bf16 copy of the merged parameters, slice by slice as the owners' done flags arrive -
//            the wait for the slowest peer overlaps the conversion of the slices already there
// A survivor never blocks on a failed rank: a function that fails marks itself dead first
// (kml_merge_tombstone), a rank that just disappears costs one bounded wait, after which the kernel
// returns WITHOUT touching any replica and reports the missing ranks in `status`.
template <bool kCoop>
__global__ void __launch_bounds__(kThreads)
kavg_merge_kernel(const kml_merge_params p) {
  __shared__ const float4* s_src[KML_MAX_RANKS];
  __shared__ float4* s_dst[KML_MAX_RANKS];
  __shared__ unsigned s_cbase, s_rbase, s_fbase, s_go;
  const bool sync_ranks = p.generation != 0;
  const uint32_t gen = p.generation;
  MergePad* pad = (kCoop && !sync_ranks) ? nullptr : reinterpret_cast<MergePad*>(p.flags[p.my_rank]);
  const unsigned long long limit = p.timeout_cycles ? p.timeout_cycles : 8000000000ull;
  // A merge of this epoch already failed: later ones queued behind it return at once (the host deals
  // with the failure at its next synchronisation point).
  if (sync_ranks && p.status != nullptr && *reinterpret_cast<volatile int*>(p.status) != 0) return;
  if (threadIdx.x == 0 && pad != nullptr) {
    s_cbase = *reinterpret_cast<volatile unsigned*>(&pad->count_base);
    s_rbase = *reinterpret_cast<volatile unsigned*>(&pad->release_base);
    s_fbase = *reinterpret_cast<volatile unsigned*>(&pad->fin_base);
  }
  unsigned nbar = 0;
  auto local_barrier = [&]() {      // every CTA of this GPU
    ++nbar;
    if constexpr (kCoop) {
      cg::this_grid().sync();
    } else {
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        const unsigned arrived = atomicAdd(&pad->count, 1u) + 1u - s_cbase;
        if (arrived == nbar * gridDim.x) {
          asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&pad->release), "r"(s_rbase + nbar) : "memory");
        } else {
          unsigned v;
          const long long t0 = clock64();
          do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(&pad->release) : "memory");
            if (static_cast<unsigned long long>(clock64() - t0) > limit) { asm volatile("trap;"); }
          } while (v - s_rbase < nbar);
        }
        __threadfence();
      }
      __syncthreads();
    }
  };
  auto finish_bases = [&]() {       // leave the monotonic counters consistent for the next launch
    if (!kCoop && nbar && blockIdx.x == 0 && threadIdx.x == 0) {
      pad->count_base = s_cbase + nbar * gridDim.x;
      pad->release_base = s_rbase + nbar;
    }
  };
  const size_t tid = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t nthreads = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t n4 = p.n_f32 / 4;
  const size_t nsgd4 = p.n_sgd / 4;
  const size_t nfuse4 = (p.n_fuse != 0 && p.n_fuse < p.n_sgd ? p.n_fuse : p.n_sgd) / 4;   // phase 0 covers [0, nfuse4)

  // One participant that is also the only destination (N = 1, or a rank alone in a late round): the average
  // of one replica is that replica (x / 1.0f == x), so nothing moves.  What is left is the fused last step -
  // which then also writes the bf16 copy itself - and the optimizer reset.
  const bool solo = !sync_ranks && __popc(p.participants) == 1 && p.write_mask == p.participants;
  uint2* wb0 = (solo && p.w_bf16 != nullptr) ? reinterpret_cast<uint2*>(p.w_bf16) : nullptr;

  // ---- phase 0: fused last local step on the own replica --------------------------------------
  if (p.fuse_sgd == 2) {
    // Adam (torch.optim.Adam semantics, csrc/optim.cu): bias corrections from the device-side step count
    float4* w = reinterpret_cast<float4*>(p.replica[p.my_rank]);
    const float4* g = reinterpret_cast<const float4*>(p.grad);
    float4* m1 = reinterpret_cast<float4*>(p.momentum_buf);
    float4* m2 = reinterpret_cast<float4*>(p.state2);
    const int t = (p.step_dev ? *reinterpret_cast<volatile int*>(p.step_dev) : 0) + 1;
    const float bc1 = 1.f - powf(p.beta1, static_cast<float>(t));
    const float bc2_sqrt = sqrtf(1.f - powf(p.beta2, static_cast<float>(t)));
    const float step_size = p.lr / bc1;
    auto adam_one = [&](size_t i, float4 wv, float4 gv, float4 a, float4 b) {
      float* wp = &wv.x; float* gp = &gv.x; float* ap = &a.x; float* bp = &b.x;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float gr = fmaf(p.weight_decay, wp[k], gp[k]);
        ap[k] = fmaf(p.beta1, ap[k], (1.f - p.beta1) * gr);
        bp[k] = fmaf(p.beta2, bp[k], (1.f - p.beta2) * gr * gr);
        wp[k] -= step_size * (ap[k] / (sqrtf(bp[k]) / bc2_sqrt + p.eps));
      }
      w[i] = wv;                 // the moments are dead after the merge (optimizer reset): not written back
      if (wb0 != nullptr) {
        __nv_bfloat162 x0 = __floats2bfloat162_rn(wv.x, wv.y);
        __nv_bfloat162 x1 = __floats2bfloat162_rn(wv.z, wv.w);
        uint2 r;
        r.x = *reinterpret_cast<uint32_t*>(&x0);
        r.y = *reinterpret_cast<uint32_t*>(&x1);
        wb0[i] = r;
      }
    };
    // two elements per thread and trip: one CTA of 512 threads per SM is all that fits (90 registers), so the
    // bytes in flight come from loads issued back to back, not from more threads
    size_t i = tid;
    for (; i + nthreads < nfuse4; i += 2 * nthreads) {
      const size_t j = i + nthreads;
      const float4 w0 = w[i], g0 = __ldcs(g + i), a0 = __ldcs(m1 + i), b0 = __ldcs(m2 + i);
      const float4 w1 = w[j], g1 = __ldcs(g + j), a1 = __ldcs(m1 + j), b1 = __ldcs(m2 + j);
      adam_one(i, w0, g0, a0, b0);
      adam_one(j, w1, g1, a1, b1);
    }
    if (i < nfuse4) adam_one(i, w[i], __ldcs(g + i), __ldcs(m1 + i), __ldcs(m2 + i));
    if (!solo) local_barrier();
  } else if (p.fuse_sgd) {
    float4* w = reinterpret_cast<float4*>(p.replica[p.my_rank]);
    const float4* g = reinterpret_cast<const float4*>(p.grad);
    const float4* m = reinterpret_cast<const float4*>(p.momentum_buf);
    const bool has_m = p.momentum != 0.f && m != nullptr;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    auto sgd_one = [&](size_t i, float4 wv, float4 gv, float4 mv) {
      float4 d;
      d.x = fmaf(p.weight_decay, wv.x, gv.x); d.y = fmaf(p.weight_decay, wv.y, gv.y);
      d.z = fmaf(p.weight_decay, wv.z, gv.z); d.w = fmaf(p.weight_decay, wv.w, gv.w);
      if (has_m) {
        d.x = fmaf(p.momentum, mv.x, d.x); d.y = fmaf(p.momentum, mv.y, d.y);
        d.z = fmaf(p.momentum, mv.z, d.z); d.w = fmaf(p.momentum, mv.w, d.w);
      }
      wv.x = fmaf(-p.lr, d.x, wv.x); wv.y = fmaf(-p.lr, d.y, wv.y);
      wv.z = fmaf(-p.lr, d.z, wv.z); wv.w = fmaf(-p.lr, d.w, wv.w);
      w[i] = wv;
      if (wb0 != nullptr) {
        __nv_bfloat162 a = __floats2bfloat162_rn(wv.x, wv.y);
        __nv_bfloat162 b = __floats2bfloat162_rn(wv.z, wv.w);
        uint2 r;
        r.x = *reinterpret_cast<uint32_t*>(&a);
        r.y = *reinterpret_cast<uint32_t*>(&b);
        wb0[i] = r;
      }
    };
    size_t i = tid;                  // two elements per thread and trip (see the Adam loop above)
    for (; i + nthreads < nfuse4; i += 2 * nthreads) {
      const size_t j = i + nthreads;
      const float4 w0 = w[i], g0 = __ldcs(g + i), m0 = has_m ? __ldcs(m + i) : zero4;
      const float4 w1 = w[j], g1 = __ldcs(g + j), m1v = has_m ? __ldcs(m + j) : zero4;
      sgd_one(i, w0, g0, m0);
      sgd_one(j, w1, g1, m1v);
    }
    if (i < nfuse4) sgd_one(i, w[i], __ldcs(g + i), has_m ? __ldcs(m + i) : zero4);
    if (!solo) local_barrier();     // the whole replica is final before anyone is told so
  }
  auto reset_optimizer = [&]() {    // the reference discards optimizer state at every sync (network.py:121-128)
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.momentum_buf) {
      float4* m = reinterpret_cast<float4*>(p.momentum_buf);
      for (size_t i = tid; i < nsgd4; i += nthreads) __stcs(m + i, z);
    }
    if (p.state2) {
      float4* m = reinterpret_cast<float4*>(p.state2);
      for (size_t i = tid; i < nsgd4; i += nthreads) __stcs(m + i, z);
    }
  };
  if (solo) {
    // the step count is read by every CTA in phase 0: only the grid's last CTA may clear it
    __shared__ unsigned s_last;
    if (p.step_dev != nullptr && p.fuse_sgd == 2) {
      __syncthreads();
      if (threadIdx.x == 0) s_last = atomicInc(reinterpret_cast<unsigned*>(p.step_dev + 1), gridDim.x - 1) == gridDim.x - 1;
      __syncthreads();
      if (s_last && threadIdx.x == 0) p.step_dev[0] = 0;
    } else if (p.step_dev != nullptr && tid == 0 && p.state2 != nullptr) {
      p.step_dev[0] = 0;
    }
    reset_optimizer();
    return;
  }

  // ---- signal / decide ---------------------------------------------------------------------------
  uint32_t eff = p.participants;
  if (sync_ranks) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      // one system-scope fence, then plain flag stores to all peers at once (a release store per peer
      // would serialise a fence round trip per peer: 8 x ~2 us on the latency-bound small merges)
      fence_acq_rel_sys();
#pragma unroll
      for (int r = 0; r < KML_MAX_RANKS; ++r)
        if ((p.participants >> r) & 1u)
          st_relaxed_sys(&reinterpret_cast<MergePad*>(p.flags[r])->ready[p.my_rank], gen);
      unsigned long long ts0;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ts0));
      uint32_t missing = 0;
      const long long t0 = clock64();
#pragma unroll
      for (int r = 0; r < KML_MAX_RANKS; ++r) {
        if (!((p.participants >> r) & 1u)) continue;
        while (!reached(ld_relaxed_sys(&pad->ready[r]), gen)) {
          if (static_cast<unsigned long long>(clock64() - t0) > limit) { missing |= 1u << r; break; }
        }
      }
      fence_acq_rel_sys();                               // acquire once for all the flags just seen
      unsigned long long ts1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ts1));
      pad->wait_ns += ts1 - ts0;                         // how long the slowest peer kept this GPU waiting
      pad->waits += 1;
      uint32_t dead = 0;
#pragma unroll
      for (int r = 0; r < KML_MAX_RANKS; ++r) {
        const uint32_t d = ld_relaxed_sys(&pad->dead_from[r]);
        if (d != 0 && reached(gen, d)) dead |= 1u << r;
      }
      uint32_t decision = p.participants & ~dead;
      if ((dead >> p.my_rank) & 1u) {                   // the others gave up on this rank: it must not write
        if (p.status) atomicOr(p.status, KML_MERGE_EVICTED);
        decision = kGoAbort;
      } else if (missing & ~dead) {
        // Record the verdict where the late rank will look before it does anything: in every pad, its
        // own included.  (A rank that arrives in the microsecond between this decision and the marks
        // landing can still slip through; the timeout is seconds, the window an NVLink round trip.)
#pragma unroll
        for (int m = 0; m < KML_MAX_RANKS; ++m) {
          if (!(((missing & ~dead) >> m) & 1u)) continue;
#pragma unroll
          for (int r = 0; r < KML_MAX_RANKS; ++r)
            if (p.flags[r] != nullptr) st_release_sys(&reinterpret_cast<MergePad*>(p.flags[r])->dead_from[m], gen);
        }
        if (p.status) atomicOr(p.status, static_cast<int>(missing & ~dead));
        decision = kGoAbort;
      }
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(&pad->go_mask) = decision;
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&pad->go_gen), "r"(gen) : "memory");
    }
    if (threadIdx.x == 0) {
      unsigned v;
      const long long t0 = clock64();
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(&pad->go_gen) : "memory");
        if (static_cast<unsigned long long>(clock64() - t0) > 2 * limit + 4000000000ull) { asm volatile("trap;"); }
      } while (v != gen);
      s_go = *reinterpret_cast<volatile uint32_t*>(&pad->go_mask);
      __threadfence_system();
    }
    __syncthreads();
    eff = s_go;
    if (eff & kGoAbort) {           // nothing was touched (apart from this rank's own last local step)
      finish_bases();
      return;
    }
  }

  // ---- phase 1: reduce my slice, write everyone -----------------------------------------------
  const int nparts = __popc(eff);
  const int pos = __popc(eff & ((1u << p.my_rank) - 1u));
  const uint32_t wmask = sync_ranks ? ((p.write_mask | eff) & ~(p.participants & ~eff)) : p.write_mask;
  size_t begin = 0, end = n4;
  if (sync_ranks) {
    begin = n4 * pos / nparts;
    end = n4 * (pos + 1) / nparts;
  }
  const float inv = 1.0f / static_cast<float>(nparts);

  if (p.mcast != nullptr && sync_ranks && eff == p.participants) {
    float4* mc = reinterpret_cast<float4*>(p.mcast);
    constexpr int U = 4;                       // loads in flight per thread: NVLink round trips are ~2 us
    size_t i = begin + tid;
    for (; i + (U - 1) * nthreads < end; i += U * nthreads) {
      float4 s[U];
#pragma unroll
      for (int u = 0; u < U; ++u) s[u] = mc_ld_reduce_f4(mc + i + u * nthreads);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        s[u].x *= inv; s[u].y *= inv; s[u].z *= inv; s[u].w *= inv;
        mc_st_f4(mc + i + u * nthreads, s[u]);
      }
    }
    for (; i < end; i += nthreads) {
      float4 s = mc_ld_reduce_f4(mc + i);
      s.x *= inv; s.y *= inv; s.z *= inv; s.w *= inv;
      mc_st_f4(mc + i, s);
    }
  } else {
    const float fn = static_cast<float>(nparts);
    if (threadIdx.x == 0) {
      int ns = 0, ndl = 0;
      for (int r = 0; r < KML_MAX_RANKS; ++r) {
        if ((eff >> r) & 1u) s_src[ns++] = reinterpret_cast<const float4*>(p.replica[r]);
        if ((wmask >> r) & 1u) s_dst[ndl++] = reinterpret_cast<float4*>(p.replica[r]);
      }
    }
    const int nd = __popc(wmask);
    __syncthreads();
    switch (nparts) {
      case 1: reduce_range<1, 8>(s_src, s_dst, nd, begin, end, tid, nthreads, fn); break;
      case 2: reduce_range<2, 4>(s_src, s_dst, nd, begin, end, tid, nthreads, fn); break;
      case 3: reduce_range<3, 2>(s_src, s_dst, nd, begin, end, tid, nthreads, fn); break;
      case 4: reduce_range<4, 2>(s_src, s_dst, nd, begin, end, tid, nthreads, fn); break;
      case 5: reduce_range<5, 1>(s_src, s_dst, nd, begin, end, tid, nthreads, fn); break;
      case 6: reduce_range<6, 1>(s_src, s_dst, nd, begin, end, tid, nthreads, fn); break;
      case 7: reduce_range<7, 1>(s_src, s_dst, nd, begin, end, tid, nthreads, fn); break;
      default: reduce_range<8, 1>(s_src, s_dst, nd, begin, end, tid, nthreads, fn); break;
    }
  }

  // int64 counters: the lowest participating rank handles them all (a few dozen scalars)
  if (p.n_i64 && pos == 0) {
    for (size_t i = tid; i < p.n_i64; i += nthreads) {
      long long s = 0;
#pragma unroll
      for (int r = 0; r < KML_MAX_RANKS; ++r)
        if ((eff >> r) & 1u) s += p.ireplica[r][i];
      s = s / static_cast<long long>(nparts);  // C++ '/' truncates toward zero like Go
#pragma unroll
      for (int r = 0; r < KML_MAX_RANKS; ++r)
        if ((wmask >> r) & 1u) p.ireplica[r][i] = s;
    }
  }

  // ---- phase 2: optimizer reset -----------------------------------------------------------------
  reset_optimizer();
  if (p.step_dev != nullptr && p.state2 != nullptr && tid == 0) p.step_dev[0] = 0;   // every CTA read it before the local barrier

  uint2* wb = reinterpret_cast<uint2*>(p.w_bf16);
  const float4* wme = reinterpret_cast<const float4*>(p.replica[p.my_rank]);
  auto to_bf16 = [&](size_t lo, size_t hi) {   // this CTA's share of [lo, hi) of the merged parameters
    if (hi > nsgd4) hi = nsgd4;
    for (size_t i = lo + tid; i < hi; i += nthreads) {
      const float4 v = ld_cg_f4(wme + i);
      __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
      __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
      uint2 r;
      r.x = *reinterpret_cast<uint32_t*>(&a);
      r.y = *reinterpret_cast<uint32_t*>(&b);
      wb[i] = r;
    }
  };

  if (!sync_ranks) {
    if (wb != nullptr) { local_barrier(); to_bf16(0, nsgd4); }
    finish_bases();
    return;
  }

  // ---- done: the last CTA of this GPU to get here tells every rank that slice `pos` has landed --------
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned fin = atomicAdd(&pad->fin_count, 1u) + 1u - s_fbase;
    if (fin == gridDim.x) {
      fence_acq_rel_sys();
#pragma unroll
      for (int r = 0; r < KML_MAX_RANKS; ++r)
        if ((eff >> r) & 1u) st_relaxed_sys(&reinterpret_cast<MergePad*>(p.flags[r])->done[p.my_rank], gen);
      pad->fin_base = s_fbase + gridDim.x;              // every CTA of this launch has read it
    }
  }

  // ---- phase 3: bf16 copy, slice by slice as their owners finish (own slice first) -----------------
  // Also the exit condition: when this kernel ends every slice has landed in this replica, and every peer
  // has finished reading this replica (it published done[] after its reads), so the next local step may
  // write it.
  {
    int idx = 0;
    for (int k = 0; k < nparts; ++k) {
      const int j = (pos + k) % nparts;                 // position of the owner within eff
      int owner = 0;
      for (int r = 0, c = 0; r < KML_MAX_RANKS; ++r)
        if ((eff >> r) & 1u) { if (c == j) owner = r; ++c; }
      if (threadIdx.x == 0) {
        const long long t0 = clock64();
        while (!reached(ld_relaxed_sys(&pad->done[owner]), gen)) {
          if (static_cast<unsigned long long>(clock64() - t0) > limit) {
            if (p.status) atomicOr(p.status, KML_MERGE_LATE | (1 << owner));     // a peer died mid-merge
            break;
          }
        }
        fence_acq_rel_sys();
      }
      __syncthreads();
      if (wb != nullptr) to_bf16(n4 * j / nparts, n4 * (j + 1) / nparts);
      ++idx;
    }
    (void)idx;
  }
  finish_bases();
}

// A rank that fails in the middle of an epoch marks itself dead for the rest of it: dead_from[me] =
// `from_gen` in every pad FIRST, then ready/done = `tombstone` (a generation past every round of the
// epoch), so no survivor ever waits for it and every survivor that gets past its flag sees it dead.
__global__ void merge_tombstone_kernel(const kml_merge_params p, uint32_t from_gen, uint32_t tombstone) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  __threadfence_system();
  for (int r = 0; r < KML_MAX_RANKS; ++r) {
    if (p.flags[r] == nullptr) continue;
    MergePad* q = reinterpret_cast<MergePad*>(p.flags[r]);
    st_release_sys(&q->dead_from[p.my_rank], from_gen);
    __threadfence_system();
    if (tombstone != 0) {
      st_release_sys(&q->ready[p.my_rank], tombstone);
      st_release_sys(&q->done[p.my_rank], tombstone);
    }
  }
}

}  // namespace

// Marks rank `who` dead from generation `from_gen` on in the pads listed in p->flags (all of them when the
// failing rank calls it for itself, only the caller's own when a survivor records a rank that timed out)
// and raises that rank's flags to `tombstone`; from_gen == 0 clears the mark (a rank re-joining).
extern "C" int kml_merge_mark_dead(const kml_merge_params* pp, int who, uint32_t from_gen, uint32_t tombstone,
                                   void* stream) {
  KML_CHECK(pp != nullptr && who >= 0 && who < KML_MAX_RANKS, "kml_merge_mark_dead: bad arguments");
  kml_merge_params p = *pp;
  p.my_rank = who;
  merge_tombstone_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(p, from_gen, tombstone);
  KML_CUDA(cudaGetLastError());
  kml::count_launch();
  return 0;
}

extern "C" int kml_kavg_merge(const kml_merge_params* pp, void* stream) {
  KML_CHECK(pp != nullptr, "kml_kavg_merge: null params");
  kml_merge_params p = *pp;
  KML_CHECK(p.my_rank >= 0 && p.my_rank < KML_MAX_RANKS, "my_rank %d out of range", p.my_rank);
  KML_CHECK(p.participants != 0 && ((p.participants >> p.my_rank) & 1u),
            "rank %d launches a merge it does not take part in (mask %#x)", p.my_rank, p.participants);
  KML_CHECK(p.n_f32 % 4 == 0, "n_f32 (%zu) must be a multiple of 4", p.n_f32);
  for (int r = 0; r < KML_MAX_RANKS; ++r) {
    if (((p.participants | p.write_mask) >> r) & 1u) {
      KML_CHECK(p.replica[r] != nullptr, "replica[%d] is null", r);
      KML_CHECK((reinterpret_cast<uintptr_t>(p.replica[r]) & 15u) == 0, "replica[%d] not 16-byte aligned", r);
      KML_CHECK(p.n_i64 == 0 || p.ireplica[r] != nullptr, "ireplica[%d] is null", r);
    }
    if (p.generation != 0 && ((p.participants >> r) & 1u))
      KML_CHECK(p.flags[r] != nullptr, "flags[%d] is null", r);
  }
  if (p.fuse_sgd) KML_CHECK(p.grad != nullptr, "fuse_sgd set without grad");
  if (p.fuse_sgd == 2) KML_CHECK(p.momentum_buf != nullptr && p.state2 != nullptr, "fused Adam step needs both moment buffers");
  KML_CHECK(p.n_sgd % 4 == 0 && p.n_sgd <= p.n_f32, "n_sgd (%zu) must be a multiple of 4 and <= n_f32", p.n_sgd);
  KML_CHECK(p.n_fuse % 4 == 0 && p.n_fuse <= p.n_sgd, "n_fuse (%zu) must be a multiple of 4 and <= n_sgd", p.n_fuse);
  if (p.generation == 0) p.mcast = nullptr;

  // Cooperative launch only for the single-GPU emulation without a signal pad (tests); every real merge is
  // a regular launch whose grid is sized to be co-resident (its CTAs wait for one another on counters).
  const bool coop = p.flags[p.my_rank] == nullptr;
  KML_CHECK(!(coop && p.generation != 0), "kml_kavg_merge: a multi-rank merge needs the signal pads");
  auto kernel = coop ? kavg_merge_kernel<true> : kavg_merge_kernel<false>;
  // launch shape: tunable for experiments (KUBEML_MERGE_THREADS in {128,256,512},
  // KUBEML_MERGE_CTAS_PER_SM); the grid must stay co-resident
  static int threads = 0, ctas_per_sm = 0;
  if (threads == 0) {
    const char* e = getenv("KUBEML_MERGE_THREADS");
    threads = e ? atoi(e) : kThreads;
    if (threads != 128 && threads != 256 && threads != 512) threads = kThreads;
    const char* c = getenv("KUBEML_MERGE_CTAS_PER_SM");
    ctas_per_sm = c ? atoi(c) : 2;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
  }
  static int occ[2] = {0, 0};
  const int vi = coop ? 1 : 0;
  if (occ[vi] == 0) {
    int o = 0;
    KML_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kernel, threads, 0));
    occ[vi] = o > 0 ? o : 1;
  }
  const size_t n4 = p.n_f32 / 4;
  int max_blocks = kml::num_sms() * (occ[vi] > ctas_per_sm ? ctas_per_sm : occ[vi]);
  size_t want = (n4 + threads - 1) / threads;
  int blocks = static_cast<int>(want < static_cast<size_t>(max_blocks) ? (want ? want : 1) : max_blocks);
  if (coop) {
    void* args[] = {&p};
    KML_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kernel), dim3(blocks), dim3(threads), args, 0,
                                         static_cast<cudaStream_t>(stream)));
  } else {
    kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    KML_CUDA(cudaGetLastError());
  }
  kml::count_launch();
  return 0;
}
