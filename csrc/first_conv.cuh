// First VGG layer: 3x3 conv with Cin = 3 (K = 27), too thin for a tensor-core tile and entirely bound by writing the
// 64-channel output, so it runs on the FP32 pipes.  Reads the NCHW image batch as the data loader delivers it and
// writes NHWC, which removes any layout-conversion pass.  (reference model.py:13,18-23 with in_channels = 3;
// SURVEY.md §2.6 O1 row L0, §7.4 "first layer K=27".)
//
// fprop:  Y[pix][64] = conv(x) + bias, plus per-block partial sum / sum-of-squares for BatchNorm.
// wgrad:  dW[64][3][3][3] = sum_pix dY[pix][co] * patch[pix][tap][ci]  -> per-block partials, reduced in fixed order
//         by column_reduce_kernel.
// No dgrad: the network input needs no gradient.
#pragma once
#include "ptx.cuh"

namespace dpb {

constexpr int kFirstConvCout = 64;
constexpr int kFirstConvPixPerBlock = 256;

// x: [N][3][H][W] fp32 (NCHW).  w: [64][3][3][3] (KRSC: co, r, s, ci).  y: [N*H*W][64].
__global__ void __launch_bounds__(256)
first_conv_fprop_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                        float* __restrict__ y, float* __restrict__ part_sum, float* __restrict__ part_sq, int N, int H,
                        int W) {
  // [k][channel group][16 + 4 pad]: the four channel groups of a warp read four different 16-byte bank groups
  // (with a dense [k][64] layout groups 0/2 and 1/3 collide: ncu showed 4 extra wavefronts per LDS.128)
  __shared__ __align__(16) float w_s[27 * 80];
  __shared__ float red[8][2][64];
  for (int i = threadIdx.x; i < 27 * 64; i += 256) {
    const int co = i & 63, k = i >> 6;
    w_s[k * 80 + (co >> 4) * 20 + (co & 15)] = w[co * 27 + k];
  }
  __syncthreads();
  const int cg = threadIdx.x & 3;       // 16-channel group
  const int slot = threadIdx.x >> 2;    // 64 pixel slots
  const long long total = (long long)N * H * W;
  const long long HW = (long long)H * W;
  float bsv[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) bsv[j] = bias ? bias[cg * 16 + j] : 0.f;
  float ssum[16], ssq[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) { ssum[j] = 0.f; ssq[j] = 0.f; }

#pragma unroll 1
  for (int it = 0; it < kFirstConvPixPerBlock / 64; ++it) {
    const long long pix = (long long)blockIdx.x * kFirstConvPixPerBlock + it * 64 + slot;
    if (pix < total) {
      const int n = (int)(pix / HW);
      const int hw = (int)(pix - n * HW);
      const int h = hw / W, wc = hw - h * W;
      float in[27];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          const int hh = h + r - 1, ww = wc + s - 1;
          const bool ok = (hh >= 0) && (hh < H) && (ww >= 0) && (ww < W);
#pragma unroll
          for (int ci = 0; ci < 3; ++ci)
            in[(r * 3 + s) * 3 + ci] = ok ? __ldg(x + ((long long)(n * 3 + ci) * H + hh) * W + ww) : 0.f;
        }
      }
      float acc[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) acc[j] = bsv[j];
#pragma unroll
      for (int k = 0; k < 27; ++k) {
        const float4* wr = reinterpret_cast<const float4*>(w_s + k * 80 + cg * 20);
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {
          const float4 wv = wr[j4];
          acc[j4 * 4 + 0] = fmaf(in[k], wv.x, acc[j4 * 4 + 0]);
          acc[j4 * 4 + 1] = fmaf(in[k], wv.y, acc[j4 * 4 + 1]);
          acc[j4 * 4 + 2] = fmaf(in[k], wv.z, acc[j4 * 4 + 2]);
          acc[j4 * 4 + 3] = fmaf(in[k], wv.w, acc[j4 * 4 + 3]);
        }
      }
      float4* yo = reinterpret_cast<float4*>(y + pix * 64 + cg * 16);
#pragma unroll
      for (int j4 = 0; j4 < 4; ++j4) yo[j4] = make_float4(acc[j4 * 4], acc[j4 * 4 + 1], acc[j4 * 4 + 2], acc[j4 * 4 + 3]);
#pragma unroll
      for (int j = 0; j < 16; ++j) { ssum[j] += acc[j]; ssq[j] += acc[j] * acc[j]; }
    }
  }
  if (part_sum != nullptr) {
    // lanes sharing a channel group: xor 4, 8, 16
#pragma unroll
    for (int j = 0; j < 16; ++j) {
#pragma unroll
      for (int o = 4; o < 32; o <<= 1) {
        ssum[j] += __shfl_xor_sync(0xffffffffu, ssum[j], o);
        ssq[j] += __shfl_xor_sync(0xffffffffu, ssq[j], o);
      }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane < 4) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        red[warp][0][lane * 16 + j] = ssum[j];
        red[warp][1][lane * 16 + j] = ssq[j];
      }
    }
    __syncthreads();
    if (threadIdx.x < 128) {
      const int which = threadIdx.x >> 6, c = threadIdx.x & 63;
      float a = 0.f;
#pragma unroll
      for (int wv = 0; wv < 8; ++wv) a += red[wv][which][c];
      (which ? part_sq : part_sum)[(size_t)blockIdx.x * 64 + c] = a;
    }
  }
}

// dy: [N*H*W][64], x: NCHW.  partials: [gridDim.x][64*27] laid out like dW (co*27 + tap*3 + ci).
// Register-blocked: a thread owns 4 output channels x 27 taps (108 accumulators); per pixel it loads one float4 of dY
// and the 27 patch values (the same address for the 16 threads of a pixel lane: broadcast from L1), i.e. 4 FMAs per
// load.  (The first version kept the patch in shared memory and issued one LDS per FMA: 186 us at batch 256.)
__global__ void __launch_bounds__(256)
first_conv_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ partials, int N,
                        int H, int W) {
  __shared__ float red[4][64 * 27];  // 27.6 KB
  const int cg = threadIdx.x & 15;   // output channels 4cg .. 4cg+3
  const int pl = threadIdx.x >> 4;   // 16 pixel lanes
  const long long total = (long long)N * H * W;
  const int HW = H * W;
  float acc[27][4];
#pragma unroll
  for (int k = 0; k < 27; ++k) { acc[k][0] = 0.f; acc[k][1] = 0.f; acc[k][2] = 0.f; acc[k][3] = 0.f; }
  const float4* dy4 = reinterpret_cast<const float4*>(dy);
  for (long long pix = (long long)blockIdx.x * 16 + pl; pix < total; pix += (long long)gridDim.x * 16) {
    const int n = (int)(pix / HW);
    const int hw = (int)(pix - (long long)n * HW);
    const int h = hw / W, wc = hw - h * W;
    const float4 g = __ldg(dy4 + pix * 16 + cg);
    const float* xb = x + (long long)n * 3 * HW + hw;   // channel 0, centre tap
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int hh = h + r - 1, ww = wc + s - 1;
        const bool ok = (hh >= 0) && (hh < H) && (ww >= 0) && (ww < W);
        const int off = (r - 1) * W + (s - 1);
#pragma unroll
        for (int ci = 0; ci < 3; ++ci) {
          const float xv = ok ? __ldg(xb + ci * HW + off) : 0.f;
          const int k = (r * 3 + s) * 3 + ci;
          acc[k][0] = fmaf(g.x, xv, acc[k][0]);
          acc[k][1] = fmaf(g.y, xv, acc[k][1]);
          acc[k][2] = fmaf(g.z, xv, acc[k][2]);
          acc[k][3] = fmaf(g.w, xv, acc[k][3]);
        }
      }
    }
  }
  // pixel lanes 2w and 2w+1 live in the two half-warps of warp w
#pragma unroll
  for (int k = 0; k < 27; ++k)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[k][j] += __shfl_xor_sync(0xffffffffu, acc[k][j], 16);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // fixed-order combine of the 8 warps: warps 0-3 store, warps 4-7 add on top, then the 4 slots are summed
  if (warp < 4 && lane < 16) {
#pragma unroll
    for (int k = 0; k < 27; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) red[warp][(cg * 4 + j) * 27 + k] = acc[k][j];
  }
  __syncthreads();
  if (warp >= 4 && lane < 16) {
#pragma unroll
    for (int k = 0; k < 27; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) red[warp - 4][(cg * 4 + j) * 27 + k] += acc[k][j];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 64 * 27; i += 256)
    partials[(size_t)blockIdx.x * (64 * 27) + i] = (red[0][i] + red[1][i]) + (red[2][i] + red[3][i]);
}

// out[c] = sum_t partials[t][c], fixed order.  grid = ceil(n/32) blocks of (32, 8).
__global__ void column_reduce_kernel(const float* __restrict__ partials, int T, int n, float* __restrict__ out) {
  __shared__ float s[8][33];
  const int cx = threadIdx.x, ry = threadIdx.y;
  const int c = blockIdx.x * 32 + cx;
  float a = 0.f;
  if (c < n)
    for (int t = ry; t < T; t += 8) a += partials[(size_t)t * n + c];
  s[ry][cx] = a;
  __syncthreads();
  if (ry == 0 && c < n) {
    float r = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) r += s[k][cx];
    out[c] = r;
  }
}

}  // namespace dpb
