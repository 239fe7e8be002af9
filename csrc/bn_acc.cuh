// BatchNorm reductions without partial rows and without a finalize kernel.
//
// Training-mode BatchNorm needs per-channel sums over the whole batch (forward: sum y, sum y^2; backward: sum g,
// sum g*xhat).  The producers of those sums -- conv epilogues, the split-K reduce, the backward reduce pass -- run as
// hundreds of CTAs; round 1 had each leave a partial row and ran a tiny finalize kernel to combine the rows (pure
// latency, 16 launches per step).  Here every CTA adds its (deterministically computed) fp32 column sums to 64-bit
// FIXED-POINT accumulators with integer atomics.  Integer addition is associative and commutative, so the total is the
// same bit pattern whatever the arrival order: results stay reproducible run to run, which floating-point atomics would
// not give.  The consumer kernel reads 2 integers per channel in its prologue and derives what it needs (scale / shift,
// or the backward coefficients) itself; the consumer's last block (ticket counter, off the critical path) clears the
// accumulators for their next use.
//
// Range / resolution: forward sums use 2^-20 (sums up to +-8.8e12), backward sums 2^-40 (up to +-8.4e6); values beyond
// saturate.  Replaces the reduction stage of aten::native_batch_norm(+_backward) as the reference reaches it through
// nn.BatchNorm2d (reference model.py:24).
#pragma once
#include "ptx.cuh"

namespace dpb {

constexpr double kBnAccScaleFwd = 1048576.0;          // 2^20
constexpr double kBnAccScaleBwd = 1099511627776.0;    // 2^40

__device__ __forceinline__ long long bn_acc_quantize(float v, double scale) {
  double d = (double)v * scale;
  d = fmin(fmax(d, -4.0e18), 4.0e18);  // saturate instead of wrapping (also maps NaN to a finite value)
  return __double2ll_rn(d);
}
__device__ __forceinline__ void bn_acc_add(unsigned long long* acc, float v, double scale = kBnAccScaleFwd) {
  atomicAdd(acc, (unsigned long long)bn_acc_quantize(v, scale));
}
__device__ __forceinline__ double bn_acc_read(const unsigned long long* acc, double scale = kBnAccScaleFwd) {
  return (double)(long long)__ldcg(acc) / scale;
}

__device__ __forceinline__ void bar_sync_n(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Every one of the `nthreads` threads (tid 0..nthreads-1, synchronised on named barrier `bar_id`) calls this when the
// block is done with the shared resource.  Returns true -- in all of them -- iff this block is the last of `expected`
// to arrive on `counter`; the counter is back at zero when the kernel ends.
__device__ __forceinline__ bool arrive_last(unsigned int* counter, unsigned int expected, int tid, int bar_id,
                                            int nthreads, volatile uint32_t* flag_smem) {
  __threadfence();
  bar_sync_n(bar_id, nthreads);
  if (tid == 0) {
    const unsigned int old = atomicAdd(counter, 1u);
    const bool last = (old == expected - 1u);
    if (last) *counter = 0u;  // nobody else touches it any more in this launch
    *flag_smem = last ? 1u : 0u;
  }
  bar_sync_n(bar_id, nthreads);
  const bool last = *flag_smem != 0u;
  if (last) __threadfence();
  return last;
}

}  // namespace dpb
