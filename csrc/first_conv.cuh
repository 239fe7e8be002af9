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
// Both kernels share one structure.  A block walks tiles of 4 image rows x W columns; the tile's input (3 channels,
// 6 x (W+2) with a zero halo) is staged in shared memory, double-buffered through registers so the next tile's global
// loads fly during the current tile's math.  256 threads = 16 channel groups (4 output channels, weights or gradient
// accumulators in registers) x 16 pixel lanes; a lane handles two horizontally adjacent pixels per pass, which share
// 36 instead of 54 patch values (read as shared-memory broadcasts, no bounds checks) and give 8 independent FMA chains.
// History (ncu, batch 256): weights in shared memory -> LDS-bandwidth bound, 126 us; patch in shared memory with one
// LDS per FMA -> 186 us; one pixel per thread with predicated global loads -> 370 instructions per warp-pixel at
// 12 % occupancy, 110 / 130 us.
#pragma once
#include "ptx.cuh"

namespace dpb {

constexpr int kFirstConvCout = 64;
constexpr int kFirstConvMaxBlocks = 296;
constexpr int kFcRows = 4;    // image rows per tile
constexpr int kFcMaxW = 64;   // widest image row supported by the shared-memory tile

struct FcTile {
  static constexpr int kRowsH = kFcRows + 2;
  static constexpr int kMaxElems = 3 * kRowsH * (kFcMaxW + 2);
  static constexpr int kPerThread = (kMaxElems + 255) / 256;
};

// element e of the haloed tile (ci, r, c) -> global value, zero outside the image
__device__ __forceinline__ float fc_fetch(const float* __restrict__ x, int e, int n, int h0, int H, int W) {
  const int WP = W + 2;
  const int ci = e / (FcTile::kRowsH * WP);
  const int rem = e - ci * (FcTile::kRowsH * WP);
  const int r = rem / WP, c = rem - r * WP;
  const int h = h0 - 1 + r, w = c - 1;
  if (h < 0 || h >= H || w < 0 || w >= W) return 0.f;
  return __ldg(x + ((long long)(n * 3 + ci) * H + h) * W + w);
}

// 3 rows x 4 columns x 3 channels around the pixel pair (row, col), (row, col+1) of the tile
__device__ __forceinline__ void fc_patch(const float* __restrict__ xs, int W, int row, int col, float (&v)[3][4][3]) {
  const int WP = W + 2;
#pragma unroll
  for (int ci = 0; ci < 3; ++ci)
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) v[r][c][ci] = xs[(ci * FcTile::kRowsH + row + r) * WP + col + c];
}

// x: [N][3][H][W] fp32 (NCHW).  w: [64][3][3][3] (KRSC: co, r, s, ci).  y: [N*H*W][64].
// part_sum / part_sq: [gridDim.x][64].  Requires even W <= 64.
__global__ void __launch_bounds__(256, 1)
first_conv_fprop_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                        float* __restrict__ y, float* __restrict__ part_sum, float* __restrict__ part_sq, int N, int H,
                        int W) {
  pdl_prologue();
  __shared__ float xs[2][FcTile::kMaxElems];
  __shared__ float red[8][2][64];
  const int cg = threadIdx.x & 15;  // output channels 4cg .. 4cg+3
  const int pl = threadIdx.x >> 4;  // 16 pixel lanes
  const int tiles_h = (H + kFcRows - 1) / kFcRows;
  const int tiles = N * tiles_h;
  const int elems = 3 * FcTile::kRowsH * (W + 2);
  const int pairs = kFcRows * W / 2;
  float wr[27][4];
#pragma unroll
  for (int k = 0; k < 27; ++k)
#pragma unroll
    for (int j = 0; j < 4; ++j) wr[k][j] = __ldg(w + (cg * 4 + j) * 27 + k);
  float bs[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) bs[j] = bias ? __ldg(bias + cg * 4 + j) : 0.f;
  float ssum[4] = {0.f, 0.f, 0.f, 0.f}, ssq[4] = {0.f, 0.f, 0.f, 0.f};

  float pre[FcTile::kPerThread];
  int tile = blockIdx.x;
  if (tile < tiles) {
#pragma unroll
    for (int i = 0; i < FcTile::kPerThread; ++i) {
      const int e = threadIdx.x + i * 256;
      if (e < elems) xs[0][e] = fc_fetch(x, e, tile / tiles_h, (tile % tiles_h) * kFcRows, H, W);
    }
  }
  __syncthreads();
  int buf = 0;
  while (tile < tiles) {
    const int next = tile + gridDim.x;
    if (next < tiles) {
#pragma unroll
      for (int i = 0; i < FcTile::kPerThread; ++i) {
        const int e = threadIdx.x + i * 256;
        pre[i] = (e < elems) ? fc_fetch(x, e, next / tiles_h, (next % tiles_h) * kFcRows, H, W) : 0.f;
      }
    }
    const int n = tile / tiles_h, h0 = (tile % tiles_h) * kFcRows;
    for (int pp = pl; pp < pairs; pp += 16) {
      const int pl2 = pp * 2;
      const int row = pl2 / W, col = pl2 - row * W;
      if (h0 + row >= H) break;
      float v[3][4][3];
      fc_patch(xs[buf], W, row, col, v);
      float a[2][4];
#pragma unroll
      for (int j = 0; j < 4; ++j) { a[0][j] = bs[j]; a[1][j] = bs[j]; }
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int s = 0; s < 3; ++s)
#pragma unroll
          for (int ci = 0; ci < 3; ++ci) {
            const int k = (r * 3 + s) * 3 + ci;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              a[0][j] = fmaf(v[r][s][ci], wr[k][j], a[0][j]);
              a[1][j] = fmaf(v[r][s + 1][ci], wr[k][j], a[1][j]);
            }
          }
      const long long gp = ((long long)n * H + h0 + row) * W + col;
      __stcs(reinterpret_cast<float4*>(y + gp * 64 + cg * 4), make_float4(a[0][0], a[0][1], a[0][2], a[0][3]));
      __stcs(reinterpret_cast<float4*>(y + (gp + 1) * 64 + cg * 4), make_float4(a[1][0], a[1][1], a[1][2], a[1][3]));
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        ssum[j] += a[0][j] + a[1][j];
        ssq[j] += a[0][j] * a[0][j] + a[1][j] * a[1][j];
      }
    }
    if (next < tiles) {
#pragma unroll
      for (int i = 0; i < FcTile::kPerThread; ++i) {
        const int e = threadIdx.x + i * 256;
        if (e < elems) xs[buf ^ 1][e] = pre[i];
      }
    }
    __syncthreads();
    buf ^= 1;
    tile = next;
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
      float s = 0.f;
#pragma unroll
      for (int wv = 0; wv < 8; ++wv) s += red[wv][which][c];
      (which ? part_sq : part_sum)[(size_t)blockIdx.x * 64 + c] = s;
    }
  }
}

// dy: [N*H*W][64], x: NCHW.  partials: [gridDim.x][64*27] laid out like dW (co*27 + tap*3 + ci).
__global__ void __launch_bounds__(256, 1)
first_conv_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ partials, int N,
                        int H, int W) {
  pdl_prologue();
  __shared__ float xs[2][FcTile::kMaxElems];
  __shared__ float red[4][64 * 27];  // 27.6 KB
  const int cg = threadIdx.x & 15;   // output channels 4cg .. 4cg+3
  const int pl = threadIdx.x >> 4;   // 16 pixel lanes
  const int tiles_h = (H + kFcRows - 1) / kFcRows;
  const int tiles = N * tiles_h;
  const int elems = 3 * FcTile::kRowsH * (W + 2);
  const int pairs = kFcRows * W / 2;
  float acc[27][4];
#pragma unroll
  for (int k = 0; k < 27; ++k) { acc[k][0] = 0.f; acc[k][1] = 0.f; acc[k][2] = 0.f; acc[k][3] = 0.f; }
  const float4* dy4 = reinterpret_cast<const float4*>(dy);

  float pre[FcTile::kPerThread];
  int tile = blockIdx.x;
  if (tile < tiles) {
#pragma unroll
    for (int i = 0; i < FcTile::kPerThread; ++i) {
      const int e = threadIdx.x + i * 256;
      if (e < elems) xs[0][e] = fc_fetch(x, e, tile / tiles_h, (tile % tiles_h) * kFcRows, H, W);
    }
  }
  __syncthreads();
  int buf = 0;
  while (tile < tiles) {
    const int next = tile + gridDim.x;
    if (next < tiles) {
#pragma unroll
      for (int i = 0; i < FcTile::kPerThread; ++i) {
        const int e = threadIdx.x + i * 256;
        pre[i] = (e < elems) ? fc_fetch(x, e, next / tiles_h, (next % tiles_h) * kFcRows, H, W) : 0.f;
      }
    }
    const int n = tile / tiles_h, h0 = (tile % tiles_h) * kFcRows;
    // gradient rows of the first pass; later passes are prefetched one pass ahead
    float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
    {
      const int pl2 = pl * 2;
      const int row = pl2 / W, col = pl2 - row * W;
      if (pl < pairs && h0 + row < H) {
        const long long gp = ((long long)n * H + h0 + row) * W + col;
        g0 = __ldcs(dy4 + gp * 16 + cg);
        g1 = __ldcs(dy4 + (gp + 1) * 16 + cg);
      }
    }
    for (int pp = pl; pp < pairs; pp += 16) {
      const int pl2 = pp * 2;
      const int row = pl2 / W, col = pl2 - row * W;
      if (h0 + row >= H) break;
      float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f), n1 = n0;
      {
        const int q2 = (pp + 16) * 2;
        const int qrow = q2 / W, qcol = q2 - qrow * W;
        if (pp + 16 < pairs && h0 + qrow < H) {
          const long long gq = ((long long)n * H + h0 + qrow) * W + qcol;
          n0 = __ldcs(dy4 + gq * 16 + cg);
          n1 = __ldcs(dy4 + (gq + 1) * 16 + cg);
        }
      }
      float v[3][4][3];
      fc_patch(xs[buf], W, row, col, v);
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int s = 0; s < 3; ++s)
#pragma unroll
          for (int ci = 0; ci < 3; ++ci) {
            const int k = (r * 3 + s) * 3 + ci;
            const float va = v[r][s][ci], vb = v[r][s + 1][ci];
            acc[k][0] = fmaf(g0.x, va, fmaf(g1.x, vb, acc[k][0]));
            acc[k][1] = fmaf(g0.y, va, fmaf(g1.y, vb, acc[k][1]));
            acc[k][2] = fmaf(g0.z, va, fmaf(g1.z, vb, acc[k][2]));
            acc[k][3] = fmaf(g0.w, va, fmaf(g1.w, vb, acc[k][3]));
          }
      g0 = n0;
      g1 = n1;
    }
    if (next < tiles) {
#pragma unroll
      for (int i = 0; i < FcTile::kPerThread; ++i) {
        const int e = threadIdx.x + i * 256;
        if (e < elems) xs[buf ^ 1][e] = pre[i];
      }
    }
    __syncthreads();
    buf ^= 1;
    tile = next;
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
  pdl_prologue();
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
