// Second phase of split-K for the tcgen05 convolution.  The conv CTAs leave fp32 partial tiles in the workspace
// ([tile][split][128][NT]); this kernel sums them in split order (deterministic), adds the bias, stores the result with
// 16-byte accesses and leaves one BatchNorm partial row (sum, sum of squares per channel) per 32-row piece of a tile:
// stat row index = m_tile * 4 + piece.
//
// grid = (tiles, 4): one block owns 32 rows x NT columns of one tile.  block = 256 threads = NT/4 float4 columns x
// 1024/NT row lanes; every thread keeps (rows per thread) x 4 splits of 16-byte loads in flight.
#pragma once
#include "conv_gemm.cuh"

namespace dpb {

template <int MODE>
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const ConvGemmParams p, int NT, int m_tiles, int n_tiles) {
  __shared__ float4 red[2][256];
  const int tile_id = blockIdx.x;
  const int m_tile = tile_id % m_tiles;
  const int n_tile = (tile_id / m_tiles) % n_tiles;
  const int wtap = tile_id / (m_tiles * n_tiles);
  const int C4 = NT >> 2;            // float4 columns per row: 16 / 32 / 64
  const int RL = 256 / C4;           // row lanes: 16 / 8 / 4
  const int c4 = threadIdx.x % C4;
  const int rl = threadIdx.x / C4;
  const int col = n_tile * NT + c4 * 4;
  const float* wsb = p.workspace + (size_t)tile_id * p.splits * (size_t)(128 * NT);
  const size_t split_stride = (size_t)128 * NT;
  float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
  if (MODE == kFprop && p.bias != nullptr) b = *reinterpret_cast<const float4*>(p.bias + col);
  float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
  for (int r = blockIdx.y * 32 + rl; r < blockIdx.y * 32 + 32; r += RL) {
    const int grow = m_tile * 128 + r;
    if (grow >= p.M_total) break;
    const float4* src = reinterpret_cast<const float4*>(wsb + (size_t)r * NT) + c4;
    float4 acc = b;
    int sp = 0;
    for (; sp + 4 <= p.splits; sp += 4) {
      const float4 t0 = __ldcg(src + (size_t)(sp + 0) * (split_stride >> 2));
      const float4 t1 = __ldcg(src + (size_t)(sp + 1) * (split_stride >> 2));
      const float4 t2 = __ldcg(src + (size_t)(sp + 2) * (split_stride >> 2));
      const float4 t3 = __ldcg(src + (size_t)(sp + 3) * (split_stride >> 2));
      acc.x += t0.x; acc.y += t0.y; acc.z += t0.z; acc.w += t0.w;
      acc.x += t1.x; acc.y += t1.y; acc.z += t1.z; acc.w += t1.w;
      acc.x += t2.x; acc.y += t2.y; acc.z += t2.z; acc.w += t2.w;
      acc.x += t3.x; acc.y += t3.y; acc.z += t3.z; acc.w += t3.w;
    }
    for (; sp < p.splits; ++sp) {
      const float4 t = __ldcg(src + (size_t)sp * (split_stride >> 2));
      acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
    }
    float* dst = (MODE == kWgrad) ? p.out + (size_t)grow * p.out_ld + (size_t)wtap * p.CinW + col
                                  : p.out + (size_t)grow * p.out_ld + col;
    *reinterpret_cast<float4*>(dst) = acc;
    s0.x += acc.x; s0.y += acc.y; s0.z += acc.z; s0.w += acc.w;
    s1.x += acc.x * acc.x; s1.y += acc.y * acc.y; s1.z += acc.z * acc.z; s1.w += acc.w * acc.w;
  }
  if (MODE == kFprop && p.stat_sum != nullptr) {
    red[0][threadIdx.x] = s0;
    red[1][threadIdx.x] = s1;
    __syncthreads();
    if (threadIdx.x < 2 * C4) {
      const int which = threadIdx.x / C4, cc = threadIdx.x % C4;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int l = 0; l < RL; ++l) {
        const float4 v = red[which][l * C4 + cc];
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
      }
      float* dst = which == 0 ? p.stat_sum : p.stat_sq;
      *reinterpret_cast<float4*>(dst + (size_t)(m_tile * 4 + blockIdx.y) * p.N_total + n_tile * NT + cc * 4) = a;
    }
  }
}

}  // namespace dpb
