// Thin inline-PTX layer for sm_100a: mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit /
// ld / fences) and system-scope signalling for the NVLink peer kernels.
//
// Everything here is device-only and header-only.  All spin loops are bounded: a wait that exceeds its budget
// records a diagnostic in `g_dpb_dev_error` and traps, so a broken pipeline produces a CUDA error on the host
// instead of a hung GPU ("every spin needs a timeout").
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace dpb {

// Diagnostic record written by a kernel right before it gives up.  The kernel writes it twice: into device memory
// (g_dpb_dev_error) and -- because a trap leaves the CUDA context unusable, so device memory can no longer be read --
// into a record in mapped pinned HOST memory that the runtime registers per device (dpb_init_device in dpb200.cu).  The
// host reads that copy without any CUDA call (dpb_last_device_error).
struct DevError {
  int code;        // 0 = none, 1 = mbarrier timeout, 2 = peer-signal timeout, 3 = bad argument
  int where;       // call-site tag (mbarrier: pipeline slot; peer signal: channel)
  int block;       // linear block id
  int aux0, aux1;  // site specific (stage/parity; rank waited for / epoch)
  int rank;        // peer-signal timeout: the rank that was waiting
  int reserved[2];
};
__device__ DevError g_dpb_dev_error;
__device__ DevError* g_dpb_dev_error_host = nullptr;
// Budget of a peer-signal wait in nanoseconds of %globaltimer (wall time, independent of the SM clock).  Default 120 s:
// generous enough for a rank that is late because of a slow data shard or a debugger, short enough that a dead peer is
// reported instead of hanging the job (dpb_set_tuning key 12 / DPB_PEER_TIMEOUT_S).
__device__ unsigned long long g_dpb_spin_timeout_ns = 120000000000ull;

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ void dev_fail(int code, int where, int aux0, int aux1, int rank = -1) {
  DevError e;
  e.code = code;
  e.where = where;
  e.block = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  e.aux0 = aux0;
  e.aux1 = aux1;
  e.rank = rank;
  e.reserved[0] = e.reserved[1] = 0;
  g_dpb_dev_error = e;
  DevError* h = g_dpb_dev_error_host;
  if (h != nullptr) {
    volatile int* hv = reinterpret_cast<volatile int*>(h);
    hv[1] = e.where; hv[2] = e.block; hv[3] = e.aux0; hv[4] = e.aux1; hv[5] = e.rank;
    __threadfence_system();
    hv[0] = e.code;  // last: a non-zero code means the record is complete
  }
  __threadfence_system();
  __trap();
}

// ---------------------------------------------------------------- kernel timeline (experiments builds only)
// With -DDPB_EXPERIMENTS every kernel stamps %globaltimer when thread 0 of block 0 enters and leaves it, into a buffer
// registered with dpb_debug_set_trace(): [0] = number of records, then {kernel id, t_begin, t_end} triples.  That is how
// tools/overlap_timeline.py shows a bucket's collective (side stream) running under the convolution kernels of earlier
// layers inside one CUDA-graph replay.  Production builds compile all of it away.
#ifdef DPB_EXPERIMENTS
__device__ unsigned long long* g_dpb_trace = nullptr;
constexpr int kTraceMaxRecords = 4096;
struct TraceScope {
  unsigned long long* rec;
  __device__ __forceinline__ explicit TraceScope(int id) : rec(nullptr) {
    if (threadIdx.x == 0 && threadIdx.y == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
      unsigned long long* t = g_dpb_trace;
      if (t != nullptr) {
        const unsigned long long slot = atomicAdd(t, 1ull);
        if (slot < (unsigned long long)kTraceMaxRecords) {
          rec = t + 1 + 3 * slot;
          rec[0] = (unsigned long long)id;
          rec[1] = globaltimer_ns();
          rec[2] = 0;
        }
      }
    }
  }
  __device__ __forceinline__ ~TraceScope() {
    if (rec != nullptr) rec[2] = globaltimer_ns();
  }
};
#define DPB_TRACE(id) ::dpb::TraceScope dpb_trace_scope_(id)
#else
#define DPB_TRACE(id) do { } while (0)
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// First 1024-byte aligned address of the dynamic shared memory window.  Written as pointer arithmetic on the
// __shared__ array (not an integer round trip) so the compiler keeps the address space and emits LDS / STS rather
// than generic LD / ST for everything derived from it.
__device__ __forceinline__ uint8_t* smem_align_1024(uint8_t* smem_raw) {
  return smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

// ---------------------------------------------------------------- programmatic dependent launch
// Every compute kernel is launched with cudaLaunchAttributeProgrammaticStreamSerialization: it lets the NEXT kernel in
// the stream be scheduled (launch latency, prologue) while this one drains, and itself waits here until the PREVIOUS
// kernel has completed and flushed its memory before touching global memory.  Both instructions are no-ops for a
// normally launched kernel.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_prologue() {
  pdl_launch_dependents();
  pdl_wait();
}

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait.  try_wait suspends in hardware for a while per attempt, so 1<<22 attempts is seconds.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int where) {
#pragma unroll 1
  for (uint32_t it = 0; it < (1u << 22); ++it) {
    if (mbar_try_wait(bar, parity)) return;
  }
  dev_fail(1, where, (int)parity, 0);
}

// ---------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const void* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const void* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const void* map, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* dst, const void* map, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_5d(void* dst, const void* map, uint64_t* bar, int c0, int c1, int c2,
                                            int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

// ---------------------------------------------------------------- thread-block clusters
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// ---------------------------------------------------------------- tcgen05 / TMEM
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot) {
  static_assert(COLS >= 32 && COLS <= 512 && (COLS & (COLS - 1)) == 0, "TMEM columns: power of two in [32,512]");
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)),
               "n"(COLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc], TF32 inputs, FP32 accumulate.  One thread issues for the CTA.
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier once every previously issued tcgen05.mma of this thread has retired.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// ---- distributed shared memory (thread-block clusters)
// shared::cluster address of `p`'s offset inside CTA `cta` of the cluster
__device__ __forceinline__ uint32_t mapa_u32_addr(uint32_t cta_local_addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_local_addr), "r"(cta));
  return r;
}
__device__ __forceinline__ uint32_t mapa_u32(const void* p, uint32_t cta) { return mapa_u32_addr(smem_u32(p), cta); }
// loads from another CTA's shared memory (address from mapa_u32)
__device__ __forceinline__ float ld_dsmem_f32(uint32_t cluster_addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(cluster_addr) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_dsmem_f4(uint32_t cluster_addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(cluster_addr)
               : "memory");
  return v;
}
// 32 lanes x 32 consecutive fp32 columns: thread t of the warp receives lane (base_lane + t).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor.  Fields per the sm_100 UMMA descriptor format:
//   [0,14) start>>4, [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version=1, [61,64) layout type:
//   2 = SWIZZLE_128B (16-byte atom; K-major operands), 1 = SWIZZLE_128B_BASE32B (32-byte atom; the only layout the
//   tensor core accepts for MN-major tf32 operands).
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                   uint32_t layout_type) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(layout_type & 7) << 61;
  return d;
}
// Instruction descriptor, kind::tf32, fp32 accumulator, M=128.
//   [4,6) c_format=1 (F32), [7,10) a_format=2 (TF32), [10,13) b_format=2, [15] a_major, [16] b_major (1 = MN-major),
//   [17,23) N>>3, [24,29) M>>4.
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int n, bool a_mn, bool b_mn, int m = 128) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// ---------------------------------------------------------------- system-scope signalling (NVLink peers)
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
// Weak (non-"strong") variants for bulk payload that is ordered by the surrounding system-scope barriers: loads that
// bypass L1 (.cv: peer lines must never be served from a stale L1 line), plain stores.
__device__ __forceinline__ float4 ld_peer_weak_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.cv.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_weak_f4(float4* p, const float4& v) {
  asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_peer_f4(float4* p, const float4& v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

}  // namespace dpb
