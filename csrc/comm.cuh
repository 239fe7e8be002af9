// Gradient-synchronisation kernels over NVLink 5 / NVSwitch peer memory.  Every rank maps every other rank's
// symmetric arena (CUDA IPC) and its signal pad; kernels then read and write peers' HBM directly.
//
//   C1  gather -> mean -> scatter   root pulls all peers' gradients, reduces in rank order, scales by 1/N in registers
//                                   and pushes the mean back with peer stores (reference main_gather.py:42-59)
//   C2  one-shot all-reduce         every rank pulls the same element from all N peers, fixed rank order => bit-identical
//                                   on every rank (reference main_all_reduce.py:45-48, small messages)
//   C3  two-shot all-reduce         reduce-scatter by peer loads + all-gather by peer stores, in place; optional fused
//                                   SGD epilogue (each rank updates its slice of the parameters/momentum and pushes the
//                                   new parameters instead of the mean gradient)
//   C4  = C3 on a sub-range, launched per bucket on the DDP side stream (reference main_ddp.py:137)
//   C5  broadcast from a root       DDP's parameter / BatchNorm-buffer broadcast
//   C6  device barrier              per-CTA release/acquire flags at system scope, monotonically increasing epochs
//
// Single-GPU emulation: with EMU = true the "rank" is blockIdx.y and the peer table holds N buffers of one device; the
// launch is cooperative so all virtual ranks are co-resident.  This is how the collectives are tested on one GPU
// (B200_PROFILING.md forbids mutually-waiting kernels as separate launches on one GPU).
#pragma once
#include "ptx.cuh"

namespace dpb {

constexpr int kMaxRanks = 8;
constexpr int kCommChannels = 4;
constexpr int kCommMaxBlocks = 128;
// signal pad layout (uint32): flags[channel][block][src_rank], then epoch[channel][block] (local use only)
constexpr int kSigFlagsWords = kCommChannels * kCommMaxBlocks * kMaxRanks;
constexpr int kSigEpochWords = kCommChannels * kCommMaxBlocks;
constexpr int kSigPadBytes = (kSigFlagsWords + kSigEpochWords) * 4;

struct CommCtx {
  int rank;
  int nranks;
  int withhold_rank;         // test hook (emulated worlds only): this virtual rank never signals; -1 = none
  uint32_t* sig[kMaxRanks];  // signal pad of every rank (peer mapped)
};

struct PeerPtrs {
  float* p[kMaxRanks];
};

template <bool EMU>
__device__ __forceinline__ int comm_rank(const CommCtx& c) {
  return EMU ? (int)blockIdx.y : c.rank;
}

// All CTAs with the same blockIdx.x on all ranks meet here.  Orders every earlier global/peer write of this CTA
// before the signal, and every later read after the peers' signals.
template <bool EMU>
__device__ __forceinline__ void peer_barrier(const CommCtx& c, int channel) {
  const int rank = comm_rank<EMU>(c);
  const int blk = blockIdx.x;
  if (EMU && rank == c.withhold_rank) return;  // test hook: a rank that never shows up (its peers must time out)
  uint32_t* my_sig = c.sig[rank];
  uint32_t* epoch_p = my_sig + kSigFlagsWords + channel * kCommMaxBlocks + blk;
  __syncthreads();
  uint32_t e = 0;
  if (threadIdx.x < (unsigned)c.nranks) {
    e = *reinterpret_cast<volatile uint32_t*>(epoch_p) + 1u;
    __threadfence_system();
    const int slot = (channel * kCommMaxBlocks + blk) * kMaxRanks;
    st_release_sys(c.sig[threadIdx.x] + slot + rank, e);
    const uint32_t* waitp = my_sig + slot + threadIdx.x;
    const unsigned long long t0 = globaltimer_ns();
    uint32_t spins = 0;
    while ((int)(ld_acquire_sys(waitp) - e) < 0) {
      // wall-clock budget; the record names who waited for whom (readable on the host even after the trap)
      if ((++spins & 0xFF) == 0 && globaltimer_ns() - t0 > g_dpb_spin_timeout_ns)
        dev_fail(2, channel, (int)threadIdx.x, (int)e, rank);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) *reinterpret_cast<volatile uint32_t*>(epoch_p) = e;
}

__device__ __forceinline__ void f4_add(float4& a, const float4& b) {
  a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
}
__device__ __forceinline__ void f4_scale(float4& a, float s) {
  a.x *= s; a.y *= s; a.z *= s; a.w *= s;
}

struct SgdArgs {
  int enabled;
  float lr, momentum, wd;
  int nesterov;
};

__device__ __forceinline__ float sgd1(float& p, float& m, float g, const SgdArgs& a) {
  const float ge = fmaf(a.wd, p, g);
  m = fmaf(a.momentum, m, ge);
  const float upd = a.nesterov ? fmaf(a.momentum, m, ge) : m;
  p = fmaf(-a.lr, upd, p);
  return p;
}

// ------------------------------------------------------------------------------------------------ C3 / C4 two-shot
// data[r] + off : rank r's gradient range (n4 float4).  Slice s of the range is owned by rank s.
// With sgd.enabled: params[r] / mom (local) are indexed like the gradients; the owner updates its slice and pushes
// parameters to every rank; gradients are left untouched unless write_mean is set.
template <bool EMU, int NR, int U>
__global__ void __launch_bounds__(512)
allreduce_twoshot_kernel(CommCtx c, int channel, PeerPtrs data, long long off4, long long n4, float scale,
                         int write_mean, SgdArgs sgd, PeerPtrs params, PeerPtrs mom) {
  DPB_TRACE(20);
  const int rank = comm_rank<EMU>(c);
  const int nr = (NR > 0) ? NR : c.nranks;
  const bool weak = (write_mean & 2) != 0;  // bit 1: weak payload accesses (dpb_set_tuning key 11), bit 0: write the mean
  const bool wmean = (write_mean & 1) != 0;
  peer_barrier<EMU>(c, channel);
  const long long slice = (n4 + nr - 1) / nr;
  const long long lo = (long long)rank * slice;
  long long hi = lo + slice;
  if (hi > n4) hi = n4;
  const long long tstride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < hi; i0 += tstride * U) {
    float4 acc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (NR > 0) {
      float4 v[NR > 0 ? NR : 1][U];
#pragma unroll
      for (int r = 0; r < NR; ++r)
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long i = i0 + u * tstride;
          const float4* src = reinterpret_cast<const float4*>(data.p[r]) + off4 + i;
          v[r][u] = (i < hi) ? (weak ? ld_peer_weak_f4(src) : ld_peer_f4(src)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
      for (int r = 0; r < NR; ++r)
#pragma unroll
        for (int u = 0; u < U; ++u) f4_add(acc[u], v[r][u]);
    } else {
      for (int r = 0; r < nr; ++r)
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long i = i0 + u * tstride;
          if (i < hi) f4_add(acc[u], ld_peer_f4(reinterpret_cast<const float4*>(data.p[r]) + off4 + i));
        }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * tstride;
      if (i >= hi) continue;
      f4_scale(acc[u], scale);
      if (sgd.enabled) {
        float4* pl = reinterpret_cast<float4*>(params.p[rank]) + off4 + i;
        float4* ml = reinterpret_cast<float4*>(mom.p[rank]) + off4 + i;
        float4 pv = *pl, mv = *ml;
        sgd1(pv.x, mv.x, acc[u].x, sgd);
        sgd1(pv.y, mv.y, acc[u].y, sgd);
        sgd1(pv.z, mv.z, acc[u].z, sgd);
        sgd1(pv.w, mv.w, acc[u].w, sgd);
        *ml = mv;
        for (int r = 0; r < nr; ++r) {
          float4* dst = reinterpret_cast<float4*>(params.p[r]) + off4 + i;
          if (weak) st_peer_weak_f4(dst, pv); else st_peer_f4(dst, pv);
        }
      }
      if (!sgd.enabled || wmean) {
        for (int r = 0; r < nr; ++r) {
          float4* dst = reinterpret_cast<float4*>(data.p[r]) + off4 + i;
          if (weak) st_peer_weak_f4(dst, acc[u]); else st_peer_f4(dst, acc[u]);
        }
      }
    }
  }
  peer_barrier<EMU>(c, channel);
}

// ------------------------------------------------------------------------------------------------ C3' NVLS two-shot
// The same collective with the reduction done INSIDE the NVSwitch (NVLink SHARP): `mc` is the multicast view of the
// gradient arena (csrc/vmm.cuh).  multimem.ld_reduce on an address of that view makes the switch fetch the 16 bytes
// from every rank and return their sum; multimem.st writes 16 bytes to every rank.  Each GPU therefore receives 1/N of
// the arena already reduced and sends it out once: about S bytes per direction per GPU instead of 2 (N-1)/N S with the
// pull + push loops above, and an SM no longer pulls and pushes at the same time (the measured limit of that kernel,
// profiles/microbench/nvlink_pull_push.txt).  Every rank stores the value the switch handed to the slice owner, so
// replicas stay bit-identical.  With sgd.enabled the owner updates its slice of the LOCAL parameters / momentum and
// multicasts the new parameters through `mc_params` (reference main_all_reduce.py:45-48 + optimizer.step()).
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float4* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float4* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

template <int U>
__global__ void __launch_bounds__(512)
allreduce_nvls_kernel(CommCtx c, int channel, float* mc_grads, long long off4, long long n4, float scale, int write_mean,
                      SgdArgs sgd, float* local_params, float* mc_params, float* local_mom) {
  DPB_TRACE(21);
  const int rank = c.rank;
  const int nr = c.nranks;
  peer_barrier<false>(c, channel);  // every rank's gradients are final
  const long long slice = (n4 + nr - 1) / nr;
  const long long lo = (long long)rank * slice;
  long long hi = lo + slice;
  if (hi > n4) hi = n4;
  const long long tstride = (long long)gridDim.x * blockDim.x;
  float4* mg = reinterpret_cast<float4*>(mc_grads) + off4;
  // A multimem round trip through the switch takes a few microseconds, so the wire only fills up with ~10^5 16-byte
  // requests in flight per GPU: U independent loads per thread before the first use.
  for (long long i0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < hi; i0 += tstride * U) {
    float4 acc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * tstride;
      acc[u] = (i < hi) ? multimem_ld_reduce_f4(mg + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * tstride;
      if (i >= hi) continue;
      f4_scale(acc[u], scale);
      if (sgd.enabled) {
        float4* pl = reinterpret_cast<float4*>(local_params) + off4 + i;
        float4* ml = reinterpret_cast<float4*>(local_mom) + off4 + i;
        float4 pv = *pl, mv = *ml;
        sgd1(pv.x, mv.x, acc[u].x, sgd);
        sgd1(pv.y, mv.y, acc[u].y, sgd);
        sgd1(pv.z, mv.z, acc[u].z, sgd);
        sgd1(pv.w, mv.w, acc[u].w, sgd);
        *ml = mv;
        multimem_st_f4(reinterpret_cast<float4*>(mc_params) + off4 + i, pv);
      }
      if (!sgd.enabled || write_mean) multimem_st_f4(mg + i, acc[u]);
    }
  }
  peer_barrier<false>(c, channel);  // the results have landed everywhere
}

// ------------------------------------------------------------------------------------------------ C2 one-shot
// Every rank reduces the whole range itself.  In place: all loads complete (registers) -> barrier -> local store, so
// the range must fit one pass of the grid: n4 <= gridDim.x * blockDim.x * U.
template <bool EMU, int U>
__global__ void __launch_bounds__(512)
allreduce_oneshot_kernel(CommCtx c, int channel, PeerPtrs data, long long off4, long long n4, float scale) {
  DPB_TRACE(24);
  const int rank = comm_rank<EMU>(c);
  const int nr = c.nranks;
  peer_barrier<EMU>(c, channel);
  const long long tstride = (long long)gridDim.x * blockDim.x;
  const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  float4 acc[U];
#pragma unroll
  for (int u = 0; u < U; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int r = 0; r < nr; ++r) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * tstride;
      if (i < n4) f4_add(acc[u], ld_peer_f4(reinterpret_cast<const float4*>(data.p[r]) + off4 + i));
    }
  }
  peer_barrier<EMU>(c, channel);  // everyone has read everyone
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const long long i = i0 + u * tstride;
    if (i < n4) {
      f4_scale(acc[u], scale);
      reinterpret_cast<float4*>(data.p[rank])[off4 + i] = acc[u];
    }
  }
}

// ------------------------------------------------------------------------------------------------ C1 gather -> mean -> scatter
template <bool EMU, int U>
__global__ void __launch_bounds__(512)
gather_mean_scatter_kernel(CommCtx c, int channel, PeerPtrs data, long long off4, long long n4, float scale, int root) {
  DPB_TRACE(22);
  const int rank = comm_rank<EMU>(c);
  const int nr = c.nranks;
  peer_barrier<EMU>(c, channel);  // every rank's gradients are final
  if (rank == root) {
    const long long tstride = (long long)gridDim.x * blockDim.x;
    for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < n4; i0 += tstride * U) {
      float4 acc[U];
#pragma unroll
      for (int u = 0; u < U; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < nr; ++r) {  // rank order: mean(stack(gather_list))
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long i = i0 + u * tstride;
          if (i < n4) f4_add(acc[u], ld_peer_f4(reinterpret_cast<const float4*>(data.p[r]) + off4 + i));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * tstride;
        if (i >= n4) continue;
        f4_scale(acc[u], scale);
        for (int r = 0; r < nr; ++r) st_peer_f4(reinterpret_cast<float4*>(data.p[r]) + off4 + i, acc[u]);
      }
    }
  }
  peer_barrier<EMU>(c, channel);  // the mean has landed everywhere
}

// ------------------------------------------------------------------------------------------------ C5 broadcast (16-byte units)
template <bool EMU>
__global__ void __launch_bounds__(512)
broadcast_kernel(CommCtx c, int channel, PeerPtrs data, long long off4, long long n4, int root) {
  DPB_TRACE(23);
  const int rank = comm_rank<EMU>(c);
  const int nr = c.nranks;
  peer_barrier<EMU>(c, channel);  // receivers are done with their old contents
  if (rank == root) {
    const long long tstride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += tstride) {
      const float4 v = reinterpret_cast<const float4*>(data.p[root])[off4 + i];
      for (int r = 0; r < nr; ++r)
        if (r != root) st_peer_f4(reinterpret_cast<float4*>(data.p[r]) + off4 + i, v);
    }
  }
  peer_barrier<EMU>(c, channel);
}

// ------------------------------------------------------------------------------------------------ C6 barrier
template <bool EMU>
__global__ void barrier_kernel(CommCtx c, int channel) {
  peer_barrier<EMU>(c, channel);
}

}  // namespace dpb
