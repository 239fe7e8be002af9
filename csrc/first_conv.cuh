// First VGG layer: 3x3 conv with Cin = 3 (K = 27), too thin for a tensor-core tile and bound by writing the 64-channel
// output, so it runs on the FP32 pipes.  Reads the NCHW image batch as the data loader delivers it and writes NHWC,
// which removes any layout-conversion pass.  (reference model.py:13,18-23 with in_channels = 3; SURVEY.md §2.6 O1 row
// L0, §7.4 "first layer K=27".)
//
// fprop:  Y[pix][64] = conv(x) + bias, plus per-block partial sum / sum-of-squares for BatchNorm.
// wgrad:  dW[64][3][3][3] = sum_pix dY[pix][co] * patch[pix][tap][ci]  -> per-block partials, reduced in fixed order
//         by column_reduce_kernel.
// No dgrad: the network input needs no gradient.
//
// Both kernels use the same register blocking: 256 threads = 16 channel groups (4 output channels each) x 16 pixel
// lanes; a thread walks pixels pix = block*16 + lane, +grid*16, ...  Per pixel it needs the 27 patch values (the same
// addresses for the 16 threads of a lane, i.e. an L1 broadcast) and does 108 FMAs on registers; the next pixel's loads
// are issued before the current pixel's FMAs (software prefetch).  History (ncu, batch 256): a version that kept the
// weights in shared memory was LDS-bandwidth bound (126 us fprop), one that kept the patch in shared memory issued one
// LDS per FMA (186 us wgrad).
#pragma once
#include "ptx.cuh"

namespace dpb {

constexpr int kFirstConvCout = 64;
constexpr int kFirstConvMaxBlocks = 296;

struct Patch27 {
  float v[27];
};

// 27 input values around pixel (n, h, wc) of an NCHW image with 3 channels; zero outside the image.
__device__ __forceinline__ void load_patch27(const float* __restrict__ x, long long pix, int H, int W, int HW,
                                             Patch27& pt) {
  const int n = (int)(pix / HW);
  const int hw = (int)(pix - (long long)n * HW);
  const int h = hw / W, wc = hw - h * W;
  const float* xb = x + (long long)n * 3 * HW + hw;  // channel 0, centre tap
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int hh = h + r - 1, ww = wc + s - 1;
      const bool ok = (hh >= 0) && (hh < H) && (ww >= 0) && (ww < W);
      const int off = (r - 1) * W + (s - 1);
#pragma unroll
      for (int ci = 0; ci < 3; ++ci) pt.v[(r * 3 + s) * 3 + ci] = ok ? __ldg(xb + ci * HW + off) : 0.f;
    }
  }
}

// x: [N][3][H][W] fp32 (NCHW).  w: [64][3][3][3] (KRSC: co, r, s, ci).  y: [N*H*W][64].
// part_sum / part_sq: [gridDim.x][64].
__global__ void __launch_bounds__(256, 1)
first_conv_fprop_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                        float* __restrict__ y, float* __restrict__ part_sum, float* __restrict__ part_sq, int N, int H,
                        int W) {
  __shared__ float red[8][2][64];
  const int cg = threadIdx.x & 15;  // output channels 4cg .. 4cg+3
  const int pl = threadIdx.x >> 4;  // 16 pixel lanes
  const long long total = (long long)N * H * W;
  const int HW = H * W;
  float wr[27][4];
#pragma unroll
  for (int k = 0; k < 27; ++k)
#pragma unroll
    for (int j = 0; j < 4; ++j) wr[k][j] = __ldg(w + (cg * 4 + j) * 27 + k);
  float bs[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) bs[j] = bias ? __ldg(bias + cg * 4 + j) : 0.f;
  float ssum[4] = {0.f, 0.f, 0.f, 0.f}, ssq[4] = {0.f, 0.f, 0.f, 0.f};

  const long long stride = (long long)gridDim.x * 16;
  long long pix = (long long)blockIdx.x * 16 + pl;
  Patch27 cur, nxt;
  if (pix < total) load_patch27(x, pix, H, W, HW, cur);
  while (pix < total) {
    const long long np = pix + stride;
    if (np < total) load_patch27(x, np, H, W, HW, nxt);
    float a0 = bs[0], a1 = bs[1], a2 = bs[2], a3 = bs[3];
#pragma unroll
    for (int k = 0; k < 27; ++k) {
      a0 = fmaf(cur.v[k], wr[k][0], a0);
      a1 = fmaf(cur.v[k], wr[k][1], a1);
      a2 = fmaf(cur.v[k], wr[k][2], a2);
      a3 = fmaf(cur.v[k], wr[k][3], a3);
    }
    __stcs(reinterpret_cast<float4*>(y + pix * 64 + cg * 4), make_float4(a0, a1, a2, a3));
    ssum[0] += a0; ssum[1] += a1; ssum[2] += a2; ssum[3] += a3;
    ssq[0] += a0 * a0; ssq[1] += a1 * a1; ssq[2] += a2 * a2; ssq[3] += a3 * a3;
    cur = nxt;
    pix = np;
  }
  if (part_sum != nullptr) {
    // pixel lanes 2w and 2w+1 are the two half-warps of warp w
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      ssum[j] += __shfl_xor_sync(0xffffffffu, ssum[j], 16);
      ssq[j] += __shfl_xor_sync(0xffffffffu, ssq[j], 16);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane < 16) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        red[warp][0][cg * 4 + j] = ssum[j];
        red[warp][1][cg * 4 + j] = ssq[j];
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
__global__ void __launch_bounds__(256, 1)
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
  const long long stride = (long long)gridDim.x * 16;
  long long pix = (long long)blockIdx.x * 16 + pl;
  Patch27 cur, nxt;
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f), gn = g;
  if (pix < total) {
    load_patch27(x, pix, H, W, HW, cur);
    g = __ldcs(dy4 + pix * 16 + cg);
  }
  while (pix < total) {
    const long long np = pix + stride;
    if (np < total) {
      load_patch27(x, np, H, W, HW, nxt);
      gn = __ldcs(dy4 + np * 16 + cg);
    }
#pragma unroll
    for (int k = 0; k < 27; ++k) {
      acc[k][0] = fmaf(g.x, cur.v[k], acc[k][0]);
      acc[k][1] = fmaf(g.y, cur.v[k], acc[k][1]);
      acc[k][2] = fmaf(g.z, cur.v[k], acc[k][2]);
      acc[k][3] = fmaf(g.w, cur.v[k], acc[k][3]);
    }
    cur = nxt;
    g = gn;
    pix = np;
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
