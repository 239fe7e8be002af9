// BatchNorm (training + eval) fused with ReLU and the optional 2x2/stride-2 max-pool, forward and backward, on
// NHWC fp32 activations.  Replaces aten::native_batch_norm(+backward), relu_/threshold_backward and
// max_pool2d_with_indices(+backward) of the reference's `layers` Sequential (reference model.py:16,24,25;
// SURVEY.md §2.6 O3-O8, §8 K4-K8).
//
// Forward:  conv epilogue leaves per-tile partial sum / sum-of-squares  ->  bn_finalize (scale/shift, running stats)
//           ->  bn_relu_pool_fwd (one pass: normalise, ReLU, pool, write the next layer's input).
// Backward: nothing but the raw conv output Y is saved; the ReLU mask and the pool arg-max are recomputed from it.
//           bn_bwd_reduce (per-block partial sum(g), sum(g*xhat))  ->  bn_bwd_finalize (dgamma, dbeta, coefficients)
//           ->  bn_bwd_apply (dY = a*g + k1*y + k0).
#pragma once
#include "ptx.cuh"
#include "bn_acc.cuh"

namespace dpb {

// Finalize kernels: block = (kFinCh channels, kFinRows row lanes), grid = C / kFinCh.  Few channels per block keeps
// the per-thread chain over the partial rows short (T / 128 loads): these kernels are pure latency.
constexpr int kFinCh = 8;
constexpr int kFinRows = 128;

// ------------------------------------------------------------------------------------------------ forward finalize
// Partials are combined in double, in a fixed order.
__global__ void bn_finalize_kernel(const float* __restrict__ part_sum, const float* __restrict__ part_sq, int T, int C,
                                   double count, const float* __restrict__ gamma, const float* __restrict__ beta,
                                   float* __restrict__ running_mean, float* __restrict__ running_var,
                                   long long* __restrict__ num_batches_tracked, float momentum, float eps,
                                   float* __restrict__ scale, float* __restrict__ shift, float* __restrict__ save_mean,
                                   float* __restrict__ save_invstd) {
  pdl_prologue();
  __shared__ double s_sum[kFinRows][kFinCh + 1], s_sq[kFinRows][kFinCh + 1];
  const int cx = threadIdx.x, ry = threadIdx.y;
  const int c = blockIdx.x * kFinCh + cx;
  double a = 0.0, b = 0.0;
  if (c < C) {
#pragma unroll 4
    for (int t = ry; t < T; t += kFinRows) {
      a += (double)part_sum[(size_t)t * C + c];
      b += (double)part_sq[(size_t)t * C + c];
    }
  }
  s_sum[ry][cx] = a;
  s_sq[ry][cx] = b;
  __syncthreads();
  if (ry < 8) {  // 128 -> 8 row lanes, fixed order
    double u = 0.0, v = 0.0;
#pragma unroll
    for (int r = ry; r < kFinRows; r += 8) {
      u += s_sum[r][cx];
      v += s_sq[r][cx];
    }
    __syncwarp();
    s_sum[ry][cx] = u;
    s_sq[ry][cx] = v;
  }
  __syncthreads();
  if (ry == 0 && c < C) {
    double sum = 0.0, sq = 0.0;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      sum += s_sum[r][cx];
      sq += s_sq[r][cx];
    }
    const double mean = sum / count;
    double var = sq / count - mean * mean;
    if (var < 0.0) var = 0.0;
    const float invstd = (float)(1.0 / sqrt(var + (double)eps));
    const float g = gamma[c], bt = beta[c];
    scale[c] = g * invstd;
    shift[c] = bt - (float)mean * g * invstd;
    save_mean[c] = (float)mean;
    save_invstd[c] = invstd;
    if (running_mean != nullptr) {
      const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
      running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
  }
  if (blockIdx.x == 0 && cx == 0 && ry == 0 && num_batches_tracked != nullptr) *num_batches_tracked += 1;
}

// eval mode: scale/shift from the running statistics
__global__ void bn_eval_scale_shift_kernel(int C, const float* __restrict__ gamma, const float* __restrict__ beta,
                                           const float* __restrict__ running_mean,
                                           const float* __restrict__ running_var, float eps, float* __restrict__ scale,
                                           float* __restrict__ shift) {
  pdl_prologue();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float invstd = rsqrtf(running_var[c] + eps);
  const float s = gamma[c] * invstd;
  scale[c] = s;
  shift[c] = beta[c] - running_mean[c] * s;
}

__device__ __forceinline__ float4 bn4(const float4& y, const float4& sc, const float4& sh) {
  return make_float4(fmaf(y.x, sc.x, sh.x), fmaf(y.y, sc.y, sh.y), fmaf(y.z, sc.z, sh.z), fmaf(y.w, sc.w, sh.w));
}
__device__ __forceinline__ float4 relu4(const float4& z) {
  return make_float4(fmaxf(z.x, 0.f), fmaxf(z.y, 0.f), fmaxf(z.z, 0.f), fmaxf(z.w, 0.f));
}
__device__ __forceinline__ float4 max4(const float4& a, const float4& b) {
  return make_float4(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w));
}

// ------------------------------------------------------------------------------------------------ forward apply
// One thread per float4 of the OUTPUT.  y: [N*H*W][C];  out: POOL ? [N*(H/2)*(W/2)][C] : [N*H*W][C].
template <bool POOL>
__global__ void __launch_bounds__(256)
bn_relu_pool_fwd_kernel(const float* __restrict__ y, const float* __restrict__ scale, const float* __restrict__ shift,
                        float* __restrict__ out, long long out_units /* out pixels * C/4 */, int C4, int H, int W) {
  DPB_TRACE(5);
  pdl_prologue();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < out_units; i += stride) {
    const int c4 = (int)(i % C4);
    const long long op = i / C4;
    const float4 sc = __ldg(reinterpret_cast<const float4*>(scale) + c4);
    const float4 sh = __ldg(reinterpret_cast<const float4*>(shift) + c4);
    float4 r;
    if (POOL) {
      const int Wo = W >> 1, Ho = H >> 1;
      const int wo = (int)(op % Wo);
      const long long t = op / Wo;
      const int ho = (int)(t % Ho);
      const long long n = t / Ho;
      const long long p00 = (n * H + 2 * ho) * W + 2 * wo;
      const float4* y4 = reinterpret_cast<const float4*>(y);
      const float4 a = __ldcs(y4 + p00 * C4 + c4);
      const float4 b = __ldcs(y4 + (p00 + 1) * C4 + c4);
      const float4 c = __ldcs(y4 + (p00 + W) * C4 + c4);
      const float4 d = __ldcs(y4 + (p00 + W + 1) * C4 + c4);
      r = max4(max4(relu4(bn4(a, sc, sh)), relu4(bn4(b, sc, sh))), max4(relu4(bn4(c, sc, sh)), relu4(bn4(d, sc, sh))));
    } else {
      r = relu4(bn4(__ldcs(reinterpret_cast<const float4*>(y) + i), sc, sh));
    }
    reinterpret_cast<float4*>(out)[i] = r;
  }
}

// ------------------------------------------------------------------------------------------------ forward from accumulators
// Arguments of the training-mode forward that derives its own scale / shift from the fixed-point accumulators the conv
// kernels filled (bn_acc.cuh).  Every block recomputes the per-channel values it needs in its prologue (identical
// integers in -> identical floats out in every block); block 0 also publishes them for the backward pass and updates
// the running statistics; the block that takes the last ticket clears the accumulators for their next use.
struct BnTrainArgs {
  unsigned long long* acc;        // [2][C] sum, sum of squares (fixed point)
  unsigned int* ticket;           // zeroed, self-resetting
  const float* gamma;
  const float* beta;
  float* running_mean;            // may be null
  float* running_var;
  long long* num_batches_tracked;  // may be null
  float momentum, eps;
  double count;                   // N*H*W
  float* stats4;                  // [4][C]: scale, shift, mean, invstd (read by the backward pass)
};

constexpr int kBnMaxC = 1024;

template <bool POOL>
__global__ void __launch_bounds__(256)
bn_relu_pool_fwd_acc_kernel(const float* __restrict__ y, const BnTrainArgs a, float* __restrict__ out,
                            long long out_units /* out pixels * C/4 */, int C, int H, int W) {
  DPB_TRACE(5);
  pdl_prologue();
  __shared__ __align__(16) float s_scale[kBnMaxC];
  __shared__ __align__(16) float s_shift[kBnMaxC];
  __shared__ uint32_t s_flag;
  const int C4 = C >> 2;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const double sum = bn_acc_read(a.acc + c), sq = bn_acc_read(a.acc + C + c);
    const double mean = sum / a.count;
    double var = sq / a.count - mean * mean;
    if (var < 0.0) var = 0.0;
    const float invstd = (float)(1.0 / sqrt(var + (double)a.eps));
    const float g = __ldg(a.gamma + c), bt = __ldg(a.beta + c);
    const float sc = g * invstd, sh = bt - (float)mean * g * invstd;
    s_scale[c] = sc;
    s_shift[c] = sh;
    if (blockIdx.x == 0) {
      a.stats4[c] = sc;
      a.stats4[C + c] = sh;
      a.stats4[2 * C + c] = (float)mean;
      a.stats4[3 * C + c] = invstd;
      if (a.running_mean != nullptr) {
        const double unbiased = a.count > 1.0 ? var * a.count / (a.count - 1.0) : var;
        a.running_mean[c] = (1.f - a.momentum) * a.running_mean[c] + a.momentum * (float)mean;
        a.running_var[c] = (1.f - a.momentum) * a.running_var[c] + a.momentum * (float)unbiased;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.num_batches_tracked != nullptr) *a.num_batches_tracked += 1;
  // this block no longer needs the accumulators; the last block to say so clears them (off the critical path)
  const bool last = arrive_last(a.ticket, gridDim.x, (int)threadIdx.x, 0, 256, &s_flag);
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < out_units; i += stride) {
    const int c4 = (int)(i % C4);
    const long long op = i / C4;
    const float4 sc = *reinterpret_cast<const float4*>(s_scale + 4 * c4);
    const float4 sh = *reinterpret_cast<const float4*>(s_shift + 4 * c4);
    float4 r;
    if (POOL) {
      const int Wo = W >> 1, Ho = H >> 1;
      const int wo = (int)(op % Wo);
      const long long t = op / Wo;
      const int ho = (int)(t % Ho);
      const long long n = t / Ho;
      const long long p00 = (n * H + 2 * ho) * W + 2 * wo;
      const float4* y4 = reinterpret_cast<const float4*>(y);
      const float4 va = __ldcs(y4 + p00 * C4 + c4);
      const float4 vb = __ldcs(y4 + (p00 + 1) * C4 + c4);
      const float4 vc = __ldcs(y4 + (p00 + W) * C4 + c4);
      const float4 vd = __ldcs(y4 + (p00 + W + 1) * C4 + c4);
      r = max4(max4(relu4(bn4(va, sc, sh)), relu4(bn4(vb, sc, sh))), max4(relu4(bn4(vc, sc, sh)), relu4(bn4(vd, sc, sh))));
    } else {
      r = relu4(bn4(__ldcs(reinterpret_cast<const float4*>(y) + i), sc, sh));
    }
    reinterpret_cast<float4*>(out)[i] = r;
  }
  if (last)
    for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) a.acc[c] = 0ull;
}

// ------------------------------------------------------------------------------------------------ backward
// g for one conv-resolution element: gradient that reaches the BN output after un-pooling and the ReLU mask.
struct Win4 {
  float4 g[4];
};

__device__ __forceinline__ void pool_route(const float z[4], float dout, float g[4]) {
  // first maximum in window order of relu(z), strict '>' like aten::max_pool2d_with_indices
  float best = fmaxf(z[0], 0.f);
  int idx = 0;
#pragma unroll
  for (int i = 1; i < 4; ++i) {
    const float r = fmaxf(z[i], 0.f);
    if (r > best) { best = r; idx = i; }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) g[i] = (i == idx && z[i] > 0.f) ? dout : 0.f;
}

// Work unit = one output pixel x 4 channels.  block = 256 threads = C4 channel groups x PL pixel lanes.
// partials: part_g[blk][C], part_gx[blk][C]
template <bool POOL>
__global__ void __launch_bounds__(256)
bn_bwd_reduce_kernel(const float* __restrict__ y, const float* __restrict__ dout, const float* __restrict__ scale,
                     const float* __restrict__ shift, const float* __restrict__ mean,
                     const float* __restrict__ invstd, float* __restrict__ part_g, float* __restrict__ part_gx,
                     long long out_pixels, int C4, int H, int W) {
  pdl_prologue();
  extern __shared__ float4 s_red[];  // [2][256]
  const int PL = 256 / C4;
  const int c4 = threadIdx.x % C4;
  const int pl = threadIdx.x / C4;
  const float4 sc = __ldg(reinterpret_cast<const float4*>(scale) + c4);
  const float4 sh = __ldg(reinterpret_cast<const float4*>(shift) + c4);
  const float4 mu = __ldg(reinterpret_cast<const float4*>(mean) + c4);
  const float4 is = __ldg(reinterpret_cast<const float4*>(invstd) + c4);
  float4 sg = make_float4(0.f, 0.f, 0.f, 0.f), sgx = make_float4(0.f, 0.f, 0.f, 0.f);
  const float4* y4 = reinterpret_cast<const float4*>(y);
  const float4* d4 = reinterpret_cast<const float4*>(dout);
  if (pl < PL) {
    for (long long op = (long long)blockIdx.x * PL + pl; op < out_pixels; op += (long long)gridDim.x * PL) {
      const float4 d = __ldg(d4 + op * C4 + c4);
      if (POOL) {
        const int Wo = W >> 1, Ho = H >> 1;
        const int wo = (int)(op % Wo);
        const long long t = op / Wo;
        const int ho = (int)(t % Ho);
        const long long n = t / Ho;
        const long long p00 = (n * H + 2 * ho) * W + 2 * wo;
        const float4 yy[4] = {__ldg(y4 + p00 * C4 + c4), __ldg(y4 + (p00 + 1) * C4 + c4),
                              __ldg(y4 + (p00 + W) * C4 + c4), __ldg(y4 + (p00 + W + 1) * C4 + c4)};
        float zx[4], zy[4], zz[4], zw[4], gx[4], gy[4], gz[4], gw[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 z = bn4(yy[i], sc, sh);
          zx[i] = z.x; zy[i] = z.y; zz[i] = z.z; zw[i] = z.w;
        }
        pool_route(zx, d.x, gx); pool_route(zy, d.y, gy); pool_route(zz, d.z, gz); pool_route(zw, d.w, gw);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          sg.x += gx[i]; sg.y += gy[i]; sg.z += gz[i]; sg.w += gw[i];
          sgx.x += gx[i] * (yy[i].x - mu.x) * is.x;
          sgx.y += gy[i] * (yy[i].y - mu.y) * is.y;
          sgx.z += gz[i] * (yy[i].z - mu.z) * is.z;
          sgx.w += gw[i] * (yy[i].w - mu.w) * is.w;
        }
      } else {
        const float4 yv = __ldg(y4 + op * C4 + c4);
        const float4 z = bn4(yv, sc, sh);
        const float4 g = make_float4(z.x > 0.f ? d.x : 0.f, z.y > 0.f ? d.y : 0.f, z.z > 0.f ? d.z : 0.f,
                                     z.w > 0.f ? d.w : 0.f);
        sg.x += g.x; sg.y += g.y; sg.z += g.z; sg.w += g.w;
        sgx.x += g.x * (yv.x - mu.x) * is.x;
        sgx.y += g.y * (yv.y - mu.y) * is.y;
        sgx.z += g.z * (yv.z - mu.z) * is.z;
        sgx.w += g.w * (yv.w - mu.w) * is.w;
      }
    }
  }
  s_red[threadIdx.x] = sg;
  s_red[256 + threadIdx.x] = sgx;
  __syncthreads();
  if (threadIdx.x < C4) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    for (int l = 0; l < PL; ++l) {
      const float4 u = s_red[l * C4 + threadIdx.x];
      const float4 v = s_red[256 + l * C4 + threadIdx.x];
      a.x += u.x; a.y += u.y; a.z += u.z; a.w += u.w;
      b.x += v.x; b.y += v.y; b.z += v.z; b.w += v.w;
    }
    reinterpret_cast<float4*>(part_g)[(size_t)blockIdx.x * C4 + threadIdx.x] = a;
    reinterpret_cast<float4*>(part_gx)[(size_t)blockIdx.x * C4 + threadIdx.x] = b;
  }
}

// grid = C/32 blocks of (32, 32).  Writes dgamma, dbeta, the preceding conv's bias gradient (optionally accumulating into existing values is the caller's
// business: these are plain stores) and the three per-channel coefficients of dY = a*g + k1*y + k0.
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ part_g, const float* __restrict__ part_gx, int T,
                                       int C, double count, const float* __restrict__ gamma,
                                       const float* __restrict__ mean, const float* __restrict__ invstd,
                                       float* __restrict__ dgamma, float* __restrict__ dbeta,
                                       float* __restrict__ conv_dbias, float* __restrict__ coef /* [3][C] */) {
  pdl_prologue();
  __shared__ double s_a[kFinRows][kFinCh + 1], s_b[kFinRows][kFinCh + 1];
  const int cx = threadIdx.x, ry = threadIdx.y;
  const int c = blockIdx.x * kFinCh + cx;
  double a = 0.0, b = 0.0;
  if (c < C) {
#pragma unroll 4
    for (int t = ry; t < T; t += kFinRows) {
      a += (double)part_g[(size_t)t * C + c];
      b += (double)part_gx[(size_t)t * C + c];
    }
  }
  s_a[ry][cx] = a;
  s_b[ry][cx] = b;
  __syncthreads();
  if (ry < 8) {  // 128 -> 8 row lanes, fixed order
    double u = 0.0, v = 0.0;
#pragma unroll
    for (int r = ry; r < kFinRows; r += 8) {
      u += s_a[r][cx];
      v += s_b[r][cx];
    }
    __syncwarp();
    s_a[ry][cx] = u;
    s_b[ry][cx] = v;
  }
  __syncthreads();
  if (ry == 0 && c < C) {
    double sg = 0.0, sgx = 0.0;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      sg += s_a[r][cx];
      sgx += s_b[r][cx];
    }
    dgamma[c] = (float)sgx;
    dbeta[c] = (float)sg;
    // The bias of the conv feeding a training-mode BatchNorm has an exactly zero gradient: sum_p dY[p][c] =
    // a*(sum g - dbeta - dgamma/M * sum xhat) and sum xhat = 0.  (aten produces rounding noise ~1e-8 here.)
    if (conv_dbias != nullptr) conv_dbias[c] = 0.f;
    const double is = (double)invstd[c];
    const double ca = (double)gamma[c] * is;
    const double k1 = -ca * is * sgx / count;
    const double k0 = -ca * sg / count - k1 * (double)mean[c];
    coef[c] = (float)ca;
    coef[C + c] = (float)k1;
    coef[2 * C + c] = (float)k0;
  }
}

// One thread per output-resolution float4 (POOL: writes the 4 window positions of dY).
template <bool POOL>
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const float* __restrict__ y, const float* __restrict__ dout, const float* __restrict__ scale,
                    const float* __restrict__ shift, const float* __restrict__ coef, float* __restrict__ dy,
                    long long out_units, int C, int H, int W) {
  DPB_TRACE(7);
  pdl_prologue();
  const int C4 = C >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const float4* y4 = reinterpret_cast<const float4*>(y);
  float4* dy4 = reinterpret_cast<float4*>(dy);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < out_units; i += stride) {
    const int c4 = (int)(i % C4);
    const long long op = i / C4;
    const float4 sc = __ldg(reinterpret_cast<const float4*>(scale) + c4);
    const float4 sh = __ldg(reinterpret_cast<const float4*>(shift) + c4);
    const float4 ca = __ldg(reinterpret_cast<const float4*>(coef) + c4);
    const float4 k1 = __ldg(reinterpret_cast<const float4*>(coef + C) + c4);
    const float4 k0 = __ldg(reinterpret_cast<const float4*>(coef + 2 * C) + c4);
    const float4 d = __ldg(reinterpret_cast<const float4*>(dout) + i);
    if (POOL) {
      const int Wo = W >> 1, Ho = H >> 1;
      const int wo = (int)(op % Wo);
      const long long t = op / Wo;
      const int ho = (int)(t % Ho);
      const long long n = t / Ho;
      const long long p00 = (n * H + 2 * ho) * W + 2 * wo;
      const long long pp[4] = {p00, p00 + 1, p00 + W, p00 + W + 1};
      float4 yy[4];
      float zx[4], zy[4], zz[4], zw[4], gx[4], gy[4], gz[4], gw[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        yy[k] = __ldcs(y4 + pp[k] * C4 + c4);
        const float4 z = bn4(yy[k], sc, sh);
        zx[k] = z.x; zy[k] = z.y; zz[k] = z.z; zw[k] = z.w;
      }
      pool_route(zx, d.x, gx); pool_route(zy, d.y, gy); pool_route(zz, d.z, gz); pool_route(zw, d.w, gw);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float4 o;
        o.x = fmaf(ca.x, gx[k], fmaf(k1.x, yy[k].x, k0.x));
        o.y = fmaf(ca.y, gy[k], fmaf(k1.y, yy[k].y, k0.y));
        o.z = fmaf(ca.z, gz[k], fmaf(k1.z, yy[k].z, k0.z));
        o.w = fmaf(ca.w, gw[k], fmaf(k1.w, yy[k].w, k0.w));
        dy4[pp[k] * C4 + c4] = o;
      }
    } else {
      const float4 yv = __ldcs(y4 + i);
      const float4 z = bn4(yv, sc, sh);
      float4 o;
      o.x = fmaf(ca.x, z.x > 0.f ? d.x : 0.f, fmaf(k1.x, yv.x, k0.x));
      o.y = fmaf(ca.y, z.y > 0.f ? d.y : 0.f, fmaf(k1.y, yv.y, k0.y));
      o.z = fmaf(ca.z, z.z > 0.f ? d.z : 0.f, fmaf(k1.z, yv.z, k0.z));
      o.w = fmaf(ca.w, z.w > 0.f ? d.w : 0.f, fmaf(k1.w, yv.w, k0.w));
      dy4[i] = o;
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward, accumulators
// Pass 1: grid = (C / 32 channel groups, slices); block = 256 threads = 8 float4 channel lanes x 32 pixel lanes, so a
// block touches only its group's 128 bytes of every pixel.  The block's 32 column sums of g and g * xhat go to the
// fixed-point accumulators acc[2][C] (bn_acc.cuh, 2^-40 resolution: gradients are small numbers).
template <bool POOL>
__global__ void __launch_bounds__(256)
bn_bwd_reduce_acc_kernel(const float* __restrict__ y, const float* __restrict__ dout, const float* __restrict__ scale,
                         const float* __restrict__ shift, const float* __restrict__ mean,
                         const float* __restrict__ invstd, unsigned long long* __restrict__ acc, long long out_pixels,
                         int C, int H, int W) {
  DPB_TRACE(6);
  pdl_prologue();
  __shared__ float4 s_red[2][256];
  const int C4 = C >> 2;
  const int cg = blockIdx.x;
  const int cl = threadIdx.x & 7;   // float4 lane inside the 32-channel group
  const int pl = threadIdx.x >> 3;  // pixel lane
  const int c4 = cg * 8 + cl;
  const float4 sc = __ldg(reinterpret_cast<const float4*>(scale) + c4);
  const float4 sh = __ldg(reinterpret_cast<const float4*>(shift) + c4);
  const float4 mu = __ldg(reinterpret_cast<const float4*>(mean) + c4);
  const float4 is = __ldg(reinterpret_cast<const float4*>(invstd) + c4);
  float4 sg = make_float4(0.f, 0.f, 0.f, 0.f), sgx = make_float4(0.f, 0.f, 0.f, 0.f);
  const float4* y4 = reinterpret_cast<const float4*>(y);
  const float4* d4 = reinterpret_cast<const float4*>(dout);
  for (long long op = (long long)blockIdx.y * 32 + pl; op < out_pixels; op += (long long)gridDim.y * 32) {
    const float4 d = __ldg(d4 + op * C4 + c4);
    if (POOL) {
      const int Wo = W >> 1, Ho = H >> 1;
      const int wo = (int)(op % Wo);
      const long long t = op / Wo;
      const int ho = (int)(t % Ho);
      const long long n = t / Ho;
      const long long p00 = (n * H + 2 * ho) * W + 2 * wo;
      const float4 yy[4] = {__ldg(y4 + p00 * C4 + c4), __ldg(y4 + (p00 + 1) * C4 + c4),
                            __ldg(y4 + (p00 + W) * C4 + c4), __ldg(y4 + (p00 + W + 1) * C4 + c4)};
      float zx[4], zy[4], zz[4], zw[4], gx[4], gy[4], gz[4], gw[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 z = bn4(yy[i], sc, sh);
        zx[i] = z.x; zy[i] = z.y; zz[i] = z.z; zw[i] = z.w;
      }
      pool_route(zx, d.x, gx); pool_route(zy, d.y, gy); pool_route(zz, d.z, gz); pool_route(zw, d.w, gw);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        sg.x += gx[i]; sg.y += gy[i]; sg.z += gz[i]; sg.w += gw[i];
        sgx.x += gx[i] * (yy[i].x - mu.x) * is.x;
        sgx.y += gy[i] * (yy[i].y - mu.y) * is.y;
        sgx.z += gz[i] * (yy[i].z - mu.z) * is.z;
        sgx.w += gw[i] * (yy[i].w - mu.w) * is.w;
      }
    } else {
      const float4 yv = __ldg(y4 + op * C4 + c4);
      const float4 z = bn4(yv, sc, sh);
      const float4 g = make_float4(z.x > 0.f ? d.x : 0.f, z.y > 0.f ? d.y : 0.f, z.z > 0.f ? d.z : 0.f,
                                   z.w > 0.f ? d.w : 0.f);
      sg.x += g.x; sg.y += g.y; sg.z += g.z; sg.w += g.w;
      sgx.x += g.x * (yv.x - mu.x) * is.x;
      sgx.y += g.y * (yv.y - mu.y) * is.y;
      sgx.z += g.z * (yv.z - mu.z) * is.z;
      sgx.w += g.w * (yv.w - mu.w) * is.w;
    }
  }
  s_red[0][threadIdx.x] = sg;
  s_red[1][threadIdx.x] = sgx;
  __syncthreads();
  if (threadIdx.x < 64) {  // 2 arrays x 32 channels; pixel lanes added in lane order, one atomic per channel
    const int which = threadIdx.x >> 5, ch = threadIdx.x & 31;
    const float* col = reinterpret_cast<const float*>(&s_red[which][0]) + ch;  // element (lane l, channel ch) at l * 32 + ch
    float a = 0.f;
#pragma unroll 8
    for (int l = 0; l < 32; ++l) a += col[l * 32];
    bn_acc_add(acc + (size_t)which * C + cg * 32 + ch, a, kBnAccScaleBwd);
  }
}

// Pass 2: every block derives the three coefficients of dY = a*g + k1*y + k0 from the accumulators in its prologue;
// block 0 writes dgamma, dbeta and the (exactly zero) bias gradient of the preceding conv; the last block to finish its
// prologue clears the accumulators.
template <bool POOL>
__global__ void __launch_bounds__(256)
bn_bwd_apply_acc_kernel(const float* __restrict__ y, const float* __restrict__ dout, const float* __restrict__ scale,
                        const float* __restrict__ shift, const float* __restrict__ mean,
                        const float* __restrict__ invstd, const float* __restrict__ gamma,
                        unsigned long long* __restrict__ acc, unsigned int* __restrict__ ticket, double count,
                        float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ conv_dbias,
                        float* __restrict__ dy, long long out_units, int C, int H, int W) {
  DPB_TRACE(7);
  pdl_prologue();
  __shared__ __align__(16) float s_coef[3][kBnMaxC];
  __shared__ uint32_t s_flag;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const double sg = bn_acc_read(acc + c, kBnAccScaleBwd), sgx = bn_acc_read(acc + C + c, kBnAccScaleBwd);
    const double is = (double)__ldg(invstd + c);
    const double ca = (double)__ldg(gamma + c) * is;
    const double k1 = -ca * is * sgx / count;
    const double k0 = -ca * sg / count - k1 * (double)__ldg(mean + c);
    s_coef[0][c] = (float)ca;
    s_coef[1][c] = (float)k1;
    s_coef[2][c] = (float)k0;
    if (blockIdx.x == 0) {
      dgamma[c] = (float)sgx;
      dbeta[c] = (float)sg;
      // The bias of the conv feeding a training-mode BatchNorm has an exactly zero gradient (see bn_bwd_finalize_kernel)
      if (conv_dbias != nullptr) conv_dbias[c] = 0.f;
    }
  }
  const bool last = arrive_last(ticket, gridDim.x, (int)threadIdx.x, 0, 256, &s_flag);
  const int C4 = C >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const float4* y4 = reinterpret_cast<const float4*>(y);
  float4* dy4 = reinterpret_cast<float4*>(dy);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < out_units; i += stride) {
    const int c4 = (int)(i % C4);
    const long long op = i / C4;
    const float4 sc = __ldg(reinterpret_cast<const float4*>(scale) + c4);
    const float4 sh = __ldg(reinterpret_cast<const float4*>(shift) + c4);
    const float4 ca = *reinterpret_cast<const float4*>(&s_coef[0][4 * c4]);
    const float4 k1 = *reinterpret_cast<const float4*>(&s_coef[1][4 * c4]);
    const float4 k0 = *reinterpret_cast<const float4*>(&s_coef[2][4 * c4]);
    const float4 d = __ldg(reinterpret_cast<const float4*>(dout) + i);
    if (POOL) {
      const int Wo = W >> 1, Ho = H >> 1;
      const int wo = (int)(op % Wo);
      const long long t = op / Wo;
      const int ho = (int)(t % Ho);
      const long long n = t / Ho;
      const long long p00 = (n * H + 2 * ho) * W + 2 * wo;
      const long long pp[4] = {p00, p00 + 1, p00 + W, p00 + W + 1};
      float4 yy[4];
      float zx[4], zy[4], zz[4], zw[4], gx[4], gy[4], gz[4], gw[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        yy[k] = __ldcs(y4 + pp[k] * C4 + c4);
        const float4 z = bn4(yy[k], sc, sh);
        zx[k] = z.x; zy[k] = z.y; zz[k] = z.z; zw[k] = z.w;
      }
      pool_route(zx, d.x, gx); pool_route(zy, d.y, gy); pool_route(zz, d.z, gz); pool_route(zw, d.w, gw);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float4 o;
        o.x = fmaf(ca.x, gx[k], fmaf(k1.x, yy[k].x, k0.x));
        o.y = fmaf(ca.y, gy[k], fmaf(k1.y, yy[k].y, k0.y));
        o.z = fmaf(ca.z, gz[k], fmaf(k1.z, yy[k].z, k0.z));
        o.w = fmaf(ca.w, gw[k], fmaf(k1.w, yy[k].w, k0.w));
        dy4[pp[k] * C4 + c4] = o;
      }
    } else {
      const float4 yv = __ldcs(y4 + i);
      const float4 z = bn4(yv, sc, sh);
      float4 o;
      o.x = fmaf(ca.x, z.x > 0.f ? d.x : 0.f, fmaf(k1.x, yv.x, k0.x));
      o.y = fmaf(ca.y, z.y > 0.f ? d.y : 0.f, fmaf(k1.y, yv.y, k0.y));
      o.z = fmaf(ca.z, z.z > 0.f ? d.z : 0.f, fmaf(k1.z, yv.z, k0.z));
      o.w = fmaf(ca.w, z.w > 0.f ? d.w : 0.f, fmaf(k1.w, yv.w, k0.w));
      dy4[i] = o;
    }
  }
  if (last)
    for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) acc[c] = 0ull;
}

}  // namespace dpb
