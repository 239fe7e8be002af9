// "Last block finishes the job": helpers that fold a tiny dependent kernel into its producer.
//
// The BatchNorm finalize steps (forward: batch mean / variance -> scale, shift, running statistics; backward: dgamma,
// dbeta and the three per-channel coefficients of dX) read a few hundred partial rows and write a few hundred floats.
// As kernels of their own they are pure latency (5-7 us each, 16 per training step; at 32 images per GPU that was 15 %
// of the step).  Here the block that delivers the LAST partial row of a channel group -- found with a self-resetting
// ticket counter -- performs the combine itself, in double precision and in a fixed order (row lanes walk the partial
// rows by index, lanes are then added in lane order), so the result does not depend on which block happens to be last.
//
// Replaces the separate finalize pass of aten::native_batch_norm / native_batch_norm_backward as the reference reaches
// them through nn.BatchNorm2d (reference model.py:24).
#pragma once
#include "ptx.cuh"

namespace dpb {

__device__ __forceinline__ void bar_sync_n(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Every one of the `nthreads` threads (tid 0..nthreads-1, synchronised on named barrier `bar_id`) calls this after
// writing its share of the block's partial row.  Returns true -- in all of them -- iff this block is the last of
// `expected` to arrive on `counter`; the counter is back at zero when the kernel ends.
__device__ __forceinline__ bool arrive_last(unsigned int* counter, unsigned int expected, int tid, int bar_id,
                                            int nthreads, volatile uint32_t* flag_smem) {
  __threadfence();  // my stores to the partial rows are visible device-wide ...
  bar_sync_n(bar_id, nthreads);  // ... and so are those of every thread of this block
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

// Forward finalize arguments (per layer).
struct BnFoldArgs {
  const float* gamma;
  const float* beta;
  float* running_mean;          // may be null
  float* running_var;
  long long* num_batches_tracked;  // may be null
  float momentum, eps;
  double count;                 // N*H*W
  float* scale;                 // [C] outputs
  float* shift;
  float* save_mean;
  float* save_invstd;
  unsigned int* counters;       // one per channel group; null = finalize is NOT folded (separate kernel follows)
};

// Forward finalize of channels [c0, c0 + cn) out of C, from T partial rows (part_sum / part_sq are [T][C]).
// cn % 4 == 0, nthreads % (cn / 4) == 0.  `sm`: at least 2 * (nthreads / (cn / 4)) * cn doubles of shared memory.
__device__ __forceinline__ void bn_finalize_cols(const float* part_sum, const float* part_sq, int T, int C, int c0,
                                                 int cn, const BnFoldArgs& a, bool bump_counter, int tid, int bar_id,
                                                 int nthreads, double* sm) {
  const int C4 = cn >> 2;
  const int RL = nthreads / C4;
  const int c4 = tid % C4, rl = tid / C4;
  double s[4] = {0.0, 0.0, 0.0, 0.0}, q[4] = {0.0, 0.0, 0.0, 0.0};
  if (rl < RL) {
    const float4* ps = reinterpret_cast<const float4*>(part_sum + c0) + c4;
    const float4* pq = reinterpret_cast<const float4*>(part_sq + c0) + c4;
    const size_t row4 = (size_t)(C >> 2);
    int t = rl;
    for (; t + 3 * RL < T; t += 4 * RL) {  // 8 independent 16-byte loads in flight
      float4 u[4], v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        u[k] = __ldcg(ps + (size_t)(t + k * RL) * row4);
        v[k] = __ldcg(pq + (size_t)(t + k * RL) * row4);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        s[0] += (double)u[k].x; s[1] += (double)u[k].y; s[2] += (double)u[k].z; s[3] += (double)u[k].w;
        q[0] += (double)v[k].x; q[1] += (double)v[k].y; q[2] += (double)v[k].z; q[3] += (double)v[k].w;
      }
    }
    for (; t < T; t += RL) {
      const float4 u = __ldcg(ps + (size_t)t * row4), v = __ldcg(pq + (size_t)t * row4);
      s[0] += (double)u.x; s[1] += (double)u.y; s[2] += (double)u.z; s[3] += (double)u.w;
      q[0] += (double)v.x; q[1] += (double)v.y; q[2] += (double)v.z; q[3] += (double)v.w;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      sm[(size_t)rl * cn + c4 * 4 + k] = s[k];
      sm[(size_t)(RL + rl) * cn + c4 * 4 + k] = q[k];
    }
  }
  bar_sync_n(bar_id, nthreads);
  for (int c = tid; c < cn; c += nthreads) {
    double sum = 0.0, sq = 0.0;
    for (int r = 0; r < RL; ++r) {
      sum += sm[(size_t)r * cn + c];
      sq += sm[(size_t)(RL + r) * cn + c];
    }
    const int ch = c0 + c;
    const double mean = sum / a.count;
    double var = sq / a.count - mean * mean;
    if (var < 0.0) var = 0.0;
    const float invstd = (float)(1.0 / sqrt(var + (double)a.eps));
    const float g = a.gamma[ch], bt = a.beta[ch];
    a.scale[ch] = g * invstd;
    a.shift[ch] = bt - (float)mean * g * invstd;
    a.save_mean[ch] = (float)mean;
    a.save_invstd[ch] = invstd;
    if (a.running_mean != nullptr) {
      const double unbiased = a.count > 1.0 ? var * a.count / (a.count - 1.0) : var;
      a.running_mean[ch] = (1.f - a.momentum) * a.running_mean[ch] + a.momentum * (float)mean;
      a.running_var[ch] = (1.f - a.momentum) * a.running_var[ch] + a.momentum * (float)unbiased;
    }
  }
  if (bump_counter && tid == 0 && a.num_batches_tracked != nullptr) *a.num_batches_tracked += 1;
  bar_sync_n(bar_id, nthreads);  // `sm` may be reused by the caller
}

// Largest number of partial rows a single block is allowed to combine; beyond it the separate finalize kernel
// (C / 8 blocks of 1024 threads) is faster than one block walking the rows.
constexpr int kBnFoldMaxRows = 512;

}  // namespace dpb
