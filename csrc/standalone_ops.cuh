// The reference's layer types used on their own (reference model.py:16-25 builds an nn.Sequential of Conv2d,
// BatchNorm2d, ReLU and MaxPool2d; the fused block of bn_relu_pool.cuh covers the pattern VGG uses, these kernels cover
// every other arrangement of the same modules): ReLU and 2x2/stride-2 max-pool forward / backward, per-channel batch
// statistics of an arbitrary NHWC tensor (BatchNorm without a preceding conv epilogue; also the bias gradient of a lone
// conv), and BatchNorm apply / backward without the ReLU.  All NHWC fp32, float4 per thread, fixed-order reductions.
#pragma once
#include "bn_relu_pool.cuh"

namespace dpb {

constexpr int kColStatBlocks = 296;

__global__ void __launch_bounds__(256) relu_fwd_kernel(const float4* __restrict__ x, float4* __restrict__ y, long long n4) {
  pdl_prologue();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) y[i] = relu4(__ldg(x + i));
}

// dx = dout where the forward OUTPUT is positive (ReLU(inplace=True) keeps only its output)
__global__ void __launch_bounds__(256) relu_bwd_kernel(const float4* __restrict__ y, const float4* __restrict__ dout,
                                                       float4* __restrict__ dx, long long n4) {
  pdl_prologue();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 v = __ldg(y + i), d = __ldg(dout + i);
    dx[i] = make_float4(v.x > 0.f ? d.x : 0.f, v.y > 0.f ? d.y : 0.f, v.z > 0.f ? d.z : 0.f, v.w > 0.f ? d.w : 0.f);
  }
}

__device__ __forceinline__ long long pool_window_base(long long op, int H, int W) {
  const int Wo = W >> 1, Ho = H >> 1;
  const int wo = (int)(op % Wo);
  const long long t = op / Wo;
  const int ho = (int)(t % Ho);
  const long long n = t / Ho;
  return (n * H + 2 * ho) * W + 2 * wo;
}

// out[n][h/2][w/2][c] = max over the 2x2 window.  units = output pixels * C/4.
__global__ void __launch_bounds__(256) maxpool2x2_fwd_kernel(const float4* __restrict__ x, float4* __restrict__ out,
                                                             long long units, int C4, int H, int W) {
  pdl_prologue();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < units; i += stride) {
    const int c4 = (int)(i % C4);
    const long long p00 = pool_window_base(i / C4, H, W);
    const float4 a = __ldg(x + p00 * C4 + c4), b = __ldg(x + (p00 + 1) * C4 + c4);
    const float4 c = __ldg(x + (p00 + W) * C4 + c4), d = __ldg(x + (p00 + W + 1) * C4 + c4);
    out[i] = max4(max4(a, b), max4(c, d));
  }
}

__device__ __forceinline__ void pool_route_plain(const float v[4], float dout, float g[4]) {
  // first maximum in window order, strict '>' like aten::max_pool2d_with_indices
  float best = v[0];
  int idx = 0;
#pragma unroll
  for (int i = 1; i < 4; ++i)
    if (v[i] > best) { best = v[i]; idx = i; }
#pragma unroll
  for (int i = 0; i < 4; ++i) g[i] = (i == idx) ? dout : 0.f;
}

__global__ void __launch_bounds__(256) maxpool2x2_bwd_kernel(const float4* __restrict__ x, const float4* __restrict__ dout,
                                                             float4* __restrict__ dx, long long units, int C4, int H,
                                                             int W) {
  pdl_prologue();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < units; i += stride) {
    const int c4 = (int)(i % C4);
    const long long p00 = pool_window_base(i / C4, H, W);
    const long long pp[4] = {p00, p00 + 1, p00 + W, p00 + W + 1};
    const float4 d = __ldg(dout + i);
    float vx[4], vy[4], vz[4], vw[4], gx[4], gy[4], gz[4], gw[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4 v = __ldg(x + pp[k] * C4 + c4);
      vx[k] = v.x; vy[k] = v.y; vz[k] = v.z; vw[k] = v.w;
    }
    pool_route_plain(vx, d.x, gx); pool_route_plain(vy, d.y, gy);
    pool_route_plain(vz, d.z, gz); pool_route_plain(vw, d.w, gw);
#pragma unroll
    for (int k = 0; k < 4; ++k) dx[pp[k] * C4 + c4] = make_float4(gx[k], gy[k], gz[k], gw[k]);
  }
}

// Per-block partial column sums of a [M][C] tensor: part_sum[blk][c] = sum x, part_sq[blk][c] = sum x^2 (skipped
// when null).  block = 256 threads = C4 channel chunks x PL row lanes (C <= 1024).
// Feeds bn_finalize_kernel (BatchNorm on a tensor that did not come out of a conv epilogue) and column_reduce_kernel
// (bias gradient of a lone conv).
__global__ void __launch_bounds__(256) column_stats_kernel(const float4* __restrict__ x, float* __restrict__ part_sum,
                                                           float* __restrict__ part_sq, long long M, int C4) {
  pdl_prologue();
  __shared__ float4 s_red[2][256];
  const int PL = 256 / C4;
  const int c4 = threadIdx.x % C4, pl = threadIdx.x / C4;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
  if (pl < PL) {
    for (long long m = (long long)blockIdx.x * PL + pl; m < M; m += (long long)gridDim.x * PL) {
      const float4 v = __ldg(x + m * C4 + c4);
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
      b.x += v.x * v.x; b.y += v.y * v.y; b.z += v.z * v.z; b.w += v.w * v.w;
    }
  }
  s_red[0][threadIdx.x] = a;
  s_red[1][threadIdx.x] = b;
  __syncthreads();
  if (threadIdx.x < C4) {
    float4 u = make_float4(0.f, 0.f, 0.f, 0.f), v = u;
    for (int l = 0; l < PL; ++l) {
      const float4 p = s_red[0][l * C4 + threadIdx.x], q = s_red[1][l * C4 + threadIdx.x];
      u.x += p.x; u.y += p.y; u.z += p.z; u.w += p.w;
      v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
    }
    reinterpret_cast<float4*>(part_sum)[(size_t)blockIdx.x * C4 + threadIdx.x] = u;
    if (part_sq != nullptr) reinterpret_cast<float4*>(part_sq)[(size_t)blockIdx.x * C4 + threadIdx.x] = v;
  }
}

// y = x * scale + shift (BatchNorm apply without ReLU)
__global__ void __launch_bounds__(256) bn_apply_kernel(const float4* __restrict__ x, const float* __restrict__ scale,
                                                       const float* __restrict__ shift, float4* __restrict__ y,
                                                       long long n4, int C4) {
  pdl_prologue();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const int c4 = (int)(i % C4);
    y[i] = bn4(__ldg(x + i), __ldg(reinterpret_cast<const float4*>(scale) + c4),
               __ldg(reinterpret_cast<const float4*>(shift) + c4));
  }
}

// BatchNorm backward without ReLU, pass 1: part_g[blk][c] = sum dout, part_gx[blk][c] = sum dout * xhat
__global__ void __launch_bounds__(256) bn_plain_bwd_reduce_kernel(const float4* __restrict__ x, const float4* __restrict__ dout,
                                                                  const float* __restrict__ mean,
                                                                  const float* __restrict__ invstd,
                                                                  float* __restrict__ part_g, float* __restrict__ part_gx,
                                                                  long long M, int C4) {
  pdl_prologue();
  __shared__ float4 s_red[2][256];
  const int PL = 256 / C4;
  const int c4 = threadIdx.x % C4, pl = threadIdx.x / C4;
  const float4 mu = __ldg(reinterpret_cast<const float4*>(mean) + c4);
  const float4 is = __ldg(reinterpret_cast<const float4*>(invstd) + c4);
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
  if (pl < PL) {
    for (long long m = (long long)blockIdx.x * PL + pl; m < M; m += (long long)gridDim.x * PL) {
      const float4 v = __ldg(x + m * C4 + c4), d = __ldg(dout + m * C4 + c4);
      a.x += d.x; a.y += d.y; a.z += d.z; a.w += d.w;
      b.x += d.x * (v.x - mu.x) * is.x; b.y += d.y * (v.y - mu.y) * is.y;
      b.z += d.z * (v.z - mu.z) * is.z; b.w += d.w * (v.w - mu.w) * is.w;
    }
  }
  s_red[0][threadIdx.x] = a;
  s_red[1][threadIdx.x] = b;
  __syncthreads();
  if (threadIdx.x < C4) {
    float4 u = make_float4(0.f, 0.f, 0.f, 0.f), v = u;
    for (int l = 0; l < PL; ++l) {
      const float4 p = s_red[0][l * C4 + threadIdx.x], q = s_red[1][l * C4 + threadIdx.x];
      u.x += p.x; u.y += p.y; u.z += p.z; u.w += p.w;
      v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
    }
    reinterpret_cast<float4*>(part_g)[(size_t)blockIdx.x * C4 + threadIdx.x] = u;
    reinterpret_cast<float4*>(part_gx)[(size_t)blockIdx.x * C4 + threadIdx.x] = v;
  }
}

// pass 2: dx = ca * dout + k1 * x + k0 with the coefficients of bn_bwd_finalize_kernel
__global__ void __launch_bounds__(256) bn_plain_bwd_apply_kernel(const float4* __restrict__ x, const float4* __restrict__ dout,
                                                                 const float* __restrict__ coef, float4* __restrict__ dx,
                                                                 long long n4, int C) {
  pdl_prologue();
  const int C4 = C >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const int c4 = (int)(i % C4);
    const float4 ca = __ldg(reinterpret_cast<const float4*>(coef) + c4);
    const float4 k1 = __ldg(reinterpret_cast<const float4*>(coef + C) + c4);
    const float4 k0 = __ldg(reinterpret_cast<const float4*>(coef + 2 * C) + c4);
    const float4 v = __ldg(x + i), d = __ldg(dout + i);
    dx[i] = make_float4(fmaf(ca.x, d.x, fmaf(k1.x, v.x, k0.x)), fmaf(ca.y, d.y, fmaf(k1.y, v.y, k0.y)),
                        fmaf(ca.z, d.z, fmaf(k1.z, v.z, k0.z)), fmaf(ca.w, d.w, fmaf(k1.w, v.w, k0.w)));
  }
}

}  // namespace dpb
