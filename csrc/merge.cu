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

// Cross-rank barrier executed by one thread of the grid.  slot = phase*KML_MAX_RANKS + writer.
__device__ void rank_barrier(const kml_merge_params& p, int phase, uint32_t gen) {
  __threadfence_system();
#pragma unroll
  for (int r = 0; r < KML_MAX_RANKS; ++r)
    if ((p.participants >> r) & 1u) st_release_sys(p.flags[r] + phase * KML_MAX_RANKS + p.my_rank, gen);
  const unsigned long long limit = p.timeout_cycles ? p.timeout_cycles : 8000000000ull;
  const long long t0 = clock64();
#pragma unroll
  for (int r = 0; r < KML_MAX_RANKS; ++r) {
    if (!((p.participants >> r) & 1u)) continue;
    const uint32_t* f = p.flags[p.my_rank] + phase * KML_MAX_RANKS + r;
    // generations only grow, so ">=" tolerates a fast peer that already moved on
    while (static_cast<int32_t>(ld_acquire_sys(f) - gen) < 0) {
      if (static_cast<unsigned long long>(clock64() - t0) > limit) {
        if (p.status) atomicOr(p.status, 1 << r);         // bit r: rank r did not arrive
        break;
      }
      __nanosleep(64);
    }
  }
  __threadfence_system();
}

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

// Words 16.. of a rank's signal pad: state of the intra-GPU barrier of the regular-launch variant.
struct LocalBarrier {
  unsigned count;         // CTA arrivals, monotonically increasing across launches
  unsigned release;       // barriers completed, monotonically increasing
  unsigned count_base;    // value of `count` when the current launch started
  unsigned release_base;  // value of `release` when the current launch started
};

template <bool kMcast, bool kCoop>
__global__ void __launch_bounds__(kThreads)
kavg_merge_kernel(const kml_merge_params p) {
  __shared__ const float4* s_src[KML_MAX_RANKS];
  __shared__ float4* s_dst[KML_MAX_RANKS];
  __shared__ unsigned s_cbase, s_rbase;
  // Two ways to synchronise the grid with itself and with the other ranks:
  //  kCoop  cooperative launch: grid.sync / flag exchange by one thread / grid.sync (used when the
  //         caller provides no signal pad, i.e. the single-GPU emulation in the tests);
  //  !kCoop regular launch (a cooperative launch costs several microseconds on this part): CTAs
  //         arrive on a counter in the rank's own signal pad, the last one to arrive performs the
  //         cross-rank flag exchange on behalf of the GPU and then releases the others.  The grid
  //         is sized to be co-resident, so the spin cannot deadlock.
  LocalBarrier* lb = kCoop ? nullptr : reinterpret_cast<LocalBarrier*>(p.flags[p.my_rank] + 2 * KML_MAX_RANKS);
  if (!kCoop && threadIdx.x == 0) {
    s_cbase = *reinterpret_cast<volatile unsigned*>(&lb->count_base);
    s_rbase = *reinterpret_cast<volatile unsigned*>(&lb->release_base);
  }
  unsigned nbar = 0;
  auto full_barrier = [&](int rank_phase /* -1: this GPU only */) {
    ++nbar;
    if constexpr (kCoop) {
      cg::grid_group grid = cg::this_grid();
      grid.sync();
      if (rank_phase >= 0) {
        if (blockIdx.x == 0 && threadIdx.x == 0) rank_barrier(p, rank_phase, p.generation);
        grid.sync();
      }
    } else {
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned arrived = atomicAdd(&lb->count, 1u) + 1u - s_cbase;
        if (arrived == nbar * gridDim.x) {
          if (rank_phase >= 0) rank_barrier(p, rank_phase, p.generation);
          asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&lb->release), "r"(s_rbase + nbar) : "memory");
        } else {
          unsigned v;
          do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(&lb->release) : "memory");
          } while (v - s_rbase < nbar);
        }
        __threadfence_system();
      }
      __syncthreads();
    }
  };
  const size_t tid = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t nthreads = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t n4 = p.n_f32 / 4;
  const size_t nsgd4 = p.n_sgd / 4;
  const bool sync_ranks = p.generation != 0;

  // ---- phase 0: fused last local step on the own replica --------------------------------------
  if (p.fuse_sgd) {
    float4* w = reinterpret_cast<float4*>(p.replica[p.my_rank]);
    const float4* g = reinterpret_cast<const float4*>(p.grad);
    const float4* m = reinterpret_cast<const float4*>(p.momentum_buf);
    const bool has_m = p.momentum != 0.f && m != nullptr;
    for (size_t i = tid; i < nsgd4; i += nthreads) {
      float4 wv = w[i], gv = __ldcs(g + i);
      float4 d;
      d.x = fmaf(p.weight_decay, wv.x, gv.x); d.y = fmaf(p.weight_decay, wv.y, gv.y);
      d.z = fmaf(p.weight_decay, wv.z, gv.z); d.w = fmaf(p.weight_decay, wv.w, gv.w);
      if (has_m) {
        float4 mv = __ldcs(m + i);
        d.x = fmaf(p.momentum, mv.x, d.x); d.y = fmaf(p.momentum, mv.y, d.y);
        d.z = fmaf(p.momentum, mv.z, d.z); d.w = fmaf(p.momentum, mv.w, d.w);
      }
      wv.x = fmaf(-p.lr, d.x, wv.x); wv.y = fmaf(-p.lr, d.y, wv.y);
      wv.z = fmaf(-p.lr, d.z, wv.z); wv.w = fmaf(-p.lr, d.w, wv.w);
      w[i] = wv;
    }
  }

  if (sync_ranks) full_barrier(0);
  else if (p.fuse_sgd) full_barrier(-1);

  // A peer never arrived (rank_barrier set `status` before the grid was released): the replicas are
  // not all final, so averaging them would publish a corrupt model into every rank and throw away the
  // optimizer state.  Touch nothing; the host sees `status`, drops the missing rank from the
  // participant mask and relaunches (parallel/device_sync.py), as the reference's merger divides by
  // the notifications it actually received (ml/pkg/train/job.go:383-398).
  if (sync_ranks && p.status != nullptr && *reinterpret_cast<volatile int*>(p.status) != 0) {
    if (!kCoop && blockIdx.x == 0 && threadIdx.x == 0) {
      // every CTA has arrived `nbar` times; leave the local barrier state consistent for the relaunch
      lb->count_base = s_cbase + nbar * gridDim.x;
      lb->release_base = s_rbase + nbar;
    }
    return;
  }

  // ---- phase 1: reduce my slice, write everyone -----------------------------------------------
  int nparts = __popc(p.participants);
  int pos = __popc(p.participants & ((1u << p.my_rank) - 1u));
  size_t begin = 0, end = n4;
  if (sync_ranks) {
    begin = n4 * pos / nparts;
    end = n4 * (pos + 1) / nparts;
  }
  const float inv = 1.0f / static_cast<float>(nparts);

  if constexpr (kMcast) {
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
    int nd = 0;
    if (threadIdx.x == 0) {
      int ns = 0, ndl = 0;
      for (int r = 0; r < KML_MAX_RANKS; ++r) {
        if ((p.participants >> r) & 1u) s_src[ns++] = reinterpret_cast<const float4*>(p.replica[r]);
        if ((p.write_mask >> r) & 1u) s_dst[ndl++] = reinterpret_cast<float4*>(p.replica[r]);
      }
    }
    nd = __popc(p.write_mask);
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
        if ((p.participants >> r) & 1u) s += p.ireplica[r][i];
      s = s / static_cast<long long>(nparts);  // C++ '/' truncates toward zero like Go
#pragma unroll
      for (int r = 0; r < KML_MAX_RANKS; ++r)
        if ((p.write_mask >> r) & 1u) p.ireplica[r][i] = s;
    }
  }

  // ---- phase 2: optimizer reset -----------------------------------------------------------------
  if (p.momentum_buf) {
    float4* m = reinterpret_cast<float4*>(p.momentum_buf);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (size_t i = tid; i < nsgd4; i += nthreads) __stcs(m + i, z);
  }

  if (sync_ranks) full_barrier(1);
  else if (p.w_bf16) full_barrier(-1);

  // ---- phase 3: refresh the local bf16 tensor-core copy of the merged parameters ---------------
  if (p.w_bf16) {
    const float4* w = reinterpret_cast<const float4*>(p.replica[p.my_rank]);
    uint2* wb = reinterpret_cast<uint2*>(p.w_bf16);
    for (size_t i = tid; i < nsgd4; i += nthreads) {
      const float4 v = ld_cg_f4(w + i);
      __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
      __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
      uint2 r;
      r.x = *reinterpret_cast<uint32_t*>(&a);
      r.y = *reinterpret_cast<uint32_t*>(&b);
      wb[i] = r;
    }
  }
  if (!kCoop && nbar && blockIdx.x == 0 && threadIdx.x == 0) {
    // every CTA read the bases before its first arrival and everyone has arrived nbar times
    lb->count_base = s_cbase + nbar * gridDim.x;
    lb->release_base = s_rbase + nbar;
  }
}

}  // namespace

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
  KML_CHECK(p.n_sgd % 4 == 0 && p.n_sgd <= p.n_f32, "n_sgd (%zu) must be a multiple of 4 and <= n_f32", p.n_sgd);
  const bool use_mc = p.mcast != nullptr && p.generation != 0;

  int dev = 0;
  KML_CUDA(cudaGetDevice(&dev));
  static int force_coop = -1;
  if (force_coop < 0) {
    // measured on 2xB200: with ~300 CTAs the counter/spin barrier is slower than grid.sync
    // (LeNet-size merge 34 us vs 23 us), so the cooperative variant stays the default
    const char* e = getenv("KUBEML_MERGE_COOP");
    force_coop = (e && e[0] == '0') ? 0 : 1;
  }
  const bool coop = force_coop || p.flags[p.my_rank] == nullptr;
  auto kernel = coop ? (use_mc ? kavg_merge_kernel<true, true> : kavg_merge_kernel<false, true>)
                     : (use_mc ? kavg_merge_kernel<true, false> : kavg_merge_kernel<false, false>);
  // launch shape: tunable for experiments (KUBEML_MERGE_THREADS in {128,256,512},
  // KUBEML_MERGE_CTAS_PER_SM); the grid must stay co-resident (cooperative launch checks it)
  static int threads = 0, ctas_per_sm = 0;
  if (threads == 0) {
    const char* e = getenv("KUBEML_MERGE_THREADS");
    threads = e ? atoi(e) : kThreads;
    if (threads != 128 && threads != 256 && threads != 512) threads = kThreads;
    const char* c = getenv("KUBEML_MERGE_CTAS_PER_SM");
    ctas_per_sm = c ? atoi(c) : 2;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
  }
  static int occ[4] = {0, 0, 0, 0};
  const int vi = (use_mc ? 1 : 0) + (coop ? 2 : 0);
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
