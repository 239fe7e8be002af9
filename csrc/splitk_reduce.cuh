// Second phase of split-K for the tcgen05 convolution.  The conv CTAs leave fp32 partial tiles in the workspace
// ([tile][split][128][NT]); this kernel sums them in split order (deterministic), adds the bias, stores the result with
// 16-byte accesses and adds the per-channel sum / sum of squares of its rows to the BatchNorm accumulators (bn_acc.cuh).
//
// This is the fallback split-K path: GEMMs whose split factor fits a thread-block cluster (<= 16) are reduced over
// distributed shared memory inside conv_gemm_kernel and never come here.
//
// grid = (tiles, pieces): one block owns R = 128 / pieces rows x NT columns of one tile; `pieces` is chosen by the host
// so that even a layer with a handful of tiles (batch 32: one 128-row tile per 128 output channels, 36 splits) spreads
// over the whole GPU.  block = 256 threads = NT/4 float4 columns x 1024/NT row lanes (R >= row lanes); every thread
// keeps 8 independent 16-byte loads in flight.  (History: a single "last CTA" reduction took 40 us per layer at batch
// 32, a 16-block version of this kernel 35 us.)
#pragma once
#include "conv_gemm.cuh"

namespace dpb {

__device__ __forceinline__ void f4_acc(float4& a, const float4& t) {
  a.x += t.x; a.y += t.y; a.z += t.z; a.w += t.w;
}

template <int MODE>
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const ConvGemmParams p, int NT, int m_tiles, int n_tiles, int pieces) {
  DPB_TRACE(4);
  pdl_prologue();
  __shared__ float4 red[2][256];
  const int tile_id = blockIdx.x;
  const int m_tile = tile_id % m_tiles;
  const int n_tile = (tile_id / m_tiles) % n_tiles;
  const int wtap = tile_id / (m_tiles * n_tiles);
  const int R = 128 / pieces;
  const int C4 = NT >> 2;            // float4 columns per row: 16 / 32 / 64
  const int RL = 256 / C4;           // row lanes: 16 / 8 / 4
  const int c4 = threadIdx.x % C4;
  const int rl = threadIdx.x / C4;
  const int col = n_tile * NT + c4 * 4;
  const float* wsb = p.workspace + (size_t)tile_id * p.splits * (size_t)(128 * NT);
  const size_t split_stride4 = (size_t)32 * NT;  // float4 units
  float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
  if (MODE == kFprop && p.bias != nullptr) b = *reinterpret_cast<const float4*>(p.bias + col);
  float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
  const int r_end = blockIdx.y * R + R;
  int r = blockIdx.y * R + rl;
  if (p.splits <= 4) {
    // few splits: one row's loads do not cover the L2 latency, so take two rows per pass (up to 8 loads in flight)
    for (; r + RL < r_end && m_tile * 128 + r + RL < p.M_total; r += 2 * RL) {
      const float4* src0 = reinterpret_cast<const float4*>(wsb + (size_t)r * NT) + c4;
      const float4* src1 = reinterpret_cast<const float4*>(wsb + (size_t)(r + RL) * NT) + c4;
      float4 t0[4], t1[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (u < p.splits) {
          t0[u] = __ldcg(src0 + (size_t)u * split_stride4);
          t1[u] = __ldcg(src1 + (size_t)u * split_stride4);
        }
      }
      float4 a0 = b, a1 = b;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (u < p.splits) {
          f4_acc(a0, t0[u]);
          f4_acc(a1, t1[u]);
        }
      }
      const int g0 = m_tile * 128 + r, g1 = g0 + RL;
      float* d0 = (MODE == kWgrad) ? p.out + (size_t)g0 * p.out_ld + (size_t)wtap * p.CinW + col
                                   : p.out + (size_t)g0 * p.out_ld + col;
      float* d1 = (MODE == kWgrad) ? p.out + (size_t)g1 * p.out_ld + (size_t)wtap * p.CinW + col
                                   : p.out + (size_t)g1 * p.out_ld + col;
      *reinterpret_cast<float4*>(d0) = a0;
      *reinterpret_cast<float4*>(d1) = a1;
      f4_acc(s0, a0);
      s1.x += a0.x * a0.x; s1.y += a0.y * a0.y; s1.z += a0.z * a0.z; s1.w += a0.w * a0.w;
      f4_acc(s0, a1);
      s1.x += a1.x * a1.x; s1.y += a1.y * a1.y; s1.z += a1.z * a1.z; s1.w += a1.w * a1.w;
    }
  }
  for (; r < r_end; r += RL) {
    const int grow = m_tile * 128 + r;
    if (grow >= p.M_total) break;
    const float4* src = reinterpret_cast<const float4*>(wsb + (size_t)r * NT) + c4;
    float4 acc = b;
    int sp = 0;
    for (; sp + 8 <= p.splits; sp += 8) {
      float4 t[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) t[u] = __ldcg(src + (size_t)(sp + u) * split_stride4);
#pragma unroll
      for (int u = 0; u < 8; ++u) f4_acc(acc, t[u]);
    }
    for (; sp < p.splits; ++sp) f4_acc(acc, __ldcg(src + (size_t)sp * split_stride4));
    float* dst = (MODE == kWgrad) ? p.out + (size_t)grow * p.out_ld + (size_t)wtap * p.CinW + col
                                  : p.out + (size_t)grow * p.out_ld + col;
    *reinterpret_cast<float4*>(dst) = acc;
    f4_acc(s0, acc);
    s1.x += acc.x * acc.x; s1.y += acc.y * acc.y; s1.z += acc.z * acc.z; s1.w += acc.w * acc.w;
  }
  if (MODE == kFprop && p.stat_acc != nullptr) {
    red[0][threadIdx.x] = s0;
    red[1][threadIdx.x] = s1;
    __syncthreads();
    if (threadIdx.x < 2 * C4) {  // row lanes combined in lane order, then one fixed-point atomic per channel (bn_acc.cuh)
      const int which = threadIdx.x / C4, cc = threadIdx.x % C4;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int l = 0; l < RL; ++l) f4_acc(a, red[which][l * C4 + cc]);
      unsigned long long* dst = p.stat_acc + (size_t)which * p.N_total + n_tile * NT + cc * 4;
      bn_acc_add(dst + 0, a.x);
      bn_acc_add(dst + 1, a.y);
      bn_acc_add(dst + 2, a.z);
      bn_acc_add(dst + 3, a.w);
    }
  }
}

// Row pieces per tile for the reduce grid: enough blocks to cover the GPU, at least 4, at most 128 / row-lanes.
inline int splitk_pieces(int tiles_total, int m_tiles, int nt) {
  const int max_by_lanes = 128 / (1024 / nt);  // NT 64 -> 8, 128 -> 16, 256 -> 32
  int cap = m_tiles <= 64 ? 32 : 4;
  if (cap > max_by_lanes) cap = max_by_lanes;
  int pieces = 4;
  while (pieces < cap && tiles_total * pieces < 296) pieces *= 2;
  return pieces;
}

}  // namespace dpb
