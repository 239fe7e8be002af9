// Classifier head (Linear 512 -> 10), softmax cross-entropy (mean reduction) forward/backward, evaluation metrics and
// the fused SGD update.  Replaces aten::addmm / mm / sum (reference model.py:40,45), _log_softmax + nll_loss and their
// backward (reference main.py:34,99), the arg-max accuracy ops of test_model (reference main.py:59-61) and
// torch.optim.SGD's per-parameter add/mul_/add_ chain (reference main.py:36,103-104).  SURVEY.md §8 K9-K12.
#pragma once
#include "ptx.cuh"
#include "bn_acc.cuh"

namespace dpb {

constexpr int kMaxClasses = 16;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// logits[b][j] = sum_k x[b][k] * w[j][k] + bias[j].  One warp per sample; float4 loads, all of a k-chunk's loads are
// independent (the first version walked k serially and was latency bound: 38 us for 256 x 512 x 10).
__global__ void __launch_bounds__(128)
linear_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                  float* __restrict__ out, int B, int K, int O) {
  pdl_prologue();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= B) return;
  float acc[kMaxClasses];
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) acc[j] = 0.f;
  const float* xr = x + (size_t)warp * K;
  if ((K & 3) == 0) {
    const int K4 = K >> 2;
    const float4* x4 = reinterpret_cast<const float4*>(xr);
#pragma unroll 4
    for (int k4 = lane; k4 < K4; k4 += 32) {
      const float4 xv = x4[k4];
#pragma unroll
      for (int j = 0; j < kMaxClasses; ++j) {
        if (j < O) {
          const float4 wv = __ldg(reinterpret_cast<const float4*>(w + (size_t)j * K) + k4);
          acc[j] = fmaf(xv.x, wv.x, fmaf(xv.y, wv.y, fmaf(xv.z, wv.z, fmaf(xv.w, wv.w, acc[j]))));
        }
      }
    }
  } else {
    for (int k = lane; k < K; k += 32) {
      const float xv = xr[k];
#pragma unroll
      for (int j = 0; j < kMaxClasses; ++j)
        if (j < O) acc[j] = fmaf(xv, __ldg(w + (size_t)j * K + k), acc[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) {
    if (j < O) {
      const float s = warp_sum(acc[j]);
      if (lane == 0) out[(size_t)warp * O + j] = s + (bias ? bias[j] : 0.f);
    }
  }
}

// dx[b][k] = sum_j dz[b][j] * w[j][k]
__global__ void __launch_bounds__(256)
linear_bwd_input_kernel(const float* __restrict__ dz, const float* __restrict__ w, float* __restrict__ dx, int B, int K,
                        int O) {
  pdl_prologue();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)B * K) return;
  const int b = (int)(i / K), k = (int)(i - (long long)b * K);
  float a = 0.f;
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j)
    if (j < O) a = fmaf(__ldg(dz + (size_t)b * O + j), __ldg(w + (size_t)j * K + k), a);
  dx[i] = a;
}

// dw[j][k] = sum_b dz[b][j] * x[b][k];  db[j] = sum_b dz[b][j].   grid = K/32 blocks of (32, 8); the batch is walked
// by the 8 row lanes and combined in a fixed order.  With db, the grid has one extra block that produces only db.
__global__ void linear_bwd_weight_kernel(const float* __restrict__ dz, const float* __restrict__ x,
                                         float* __restrict__ dw, float* __restrict__ db, int B, int K, int O) {
  pdl_prologue();
  __shared__ float s[8][kMaxClasses][33];
  const int cx = threadIdx.x, ry = threadIdx.y;
  const int k = blockIdx.x * 32 + cx;
  float acc[kMaxClasses];
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) acc[j] = 0.f;
  if (k < K) {
#pragma unroll 8
    for (int b = ry; b < B; b += 8) {
      const float xv = __ldg(x + (size_t)b * K + k);
#pragma unroll
      for (int j = 0; j < kMaxClasses; ++j)
        if (j < O) acc[j] = fmaf(__ldg(dz + (size_t)b * O + j), xv, acc[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) s[ry][j][cx] = acc[j];
  __syncthreads();
  if (k < K) {
    for (int j = ry; j < O; j += 8) {
      float r = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) r += s[q][j][cx];
      dw[(size_t)j * K + k] = r;
    }
  }
  if (db != nullptr && blockIdx.x == gridDim.x - 1) {  // the extra block past the last k columns (see the launch)
    __syncthreads();
    const int t = ry * 32 + cx;  // 256 threads
    float a[kMaxClasses];
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j) a[j] = 0.f;
    for (int b = t; b < B; b += 256)
#pragma unroll
      for (int j = 0; j < kMaxClasses; ++j)
        if (j < O) a[j] += dz[(size_t)b * O + j];
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j) {
      const float v = warp_sum(a[j]);
      if (cx == 0) s[ry][j][0] = v;
    }
    __syncthreads();
    if (t < O) {
      float r = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) r += s[q][t][0];
      db[t] = r;
    }
  }
}

// Single block.  loss_out[0] = mean_i (lse_i - z_i[y_i]).  If metrics != null (eval), metrics[0] += mean loss,
// metrics[1] += number of arg-max hits (first maximum wins, like aten::max(dim)).
__global__ void __launch_bounds__(256)
cross_entropy_fwd_kernel(const float* __restrict__ logits, const long long* __restrict__ target,
                         float* __restrict__ loss_out, float* __restrict__ metrics, int B, int O) {
  pdl_prologue();
  __shared__ float s_loss[8], s_hit[8];
  float lsum = 0.f, hits = 0.f;
  for (int b = threadIdx.x; b < B; b += 256) {
    const float* z = logits + (size_t)b * O;
    float m = z[0];
    int am = 0;
    for (int j = 1; j < O; ++j)
      if (z[j] > m) { m = z[j]; am = j; }
    float se = 0.f;
    for (int j = 0; j < O; ++j) se += expf(z[j] - m);
    const int t = (int)target[b];
    lsum += (m + logf(se)) - z[t];
    hits += (am == t) ? 1.f : 0.f;
  }
  lsum = warp_sum(lsum);
  hits = warp_sum(hits);
  if ((threadIdx.x & 31) == 0) { s_loss[threadIdx.x >> 5] = lsum; s_hit[threadIdx.x >> 5] = hits; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, h = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) { l += s_loss[q]; h += s_hit[q]; }
    l /= (float)B;
    loss_out[0] = l;
    if (metrics != nullptr) { metrics[0] += l; metrics[1] += h; }
  }
}

// dz[b][j] = (softmax(z_b)[j] - [j == y_b]) * gscale / B, gscale read from device memory (autograd's grad_output).
__global__ void __launch_bounds__(256)
cross_entropy_bwd_kernel(const float* __restrict__ logits, const long long* __restrict__ target,
                         const float* __restrict__ grad_out, float* __restrict__ dz, int B, int O) {
  pdl_prologue();
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const float gs = (grad_out ? grad_out[0] : 1.f) / (float)B;
  const float* z = logits + (size_t)b * O;
  float m = z[0];
  for (int j = 1; j < O; ++j) m = fmaxf(m, z[j]);
  float se = 0.f;
  for (int j = 0; j < O; ++j) se += expf(z[j] - m);
  const float inv = 1.f / se;
  const int t = (int)target[b];
  for (int j = 0; j < O; ++j) dz[(size_t)b * O + j] = (expf(z[j] - m) * inv - (j == t ? 1.f : 0.f)) * gs;
}

// ------------------------------------------------------------------------------------------------ fused head
// Training head in two launches instead of five (linear fwd, CE fwd, CE bwd, linear bwd-input, linear bwd-weight): at 32
// images per GPU those five dependent ~4 us kernels were 7 % of the step.
//
// head_fwd_kernel: block = 8 warps = 8 samples, the O x K weight matrix staged in shared memory once per block (the
// first version of linear_fwd re-read it from L2 per sample: 16 us for 256 x 512 x 10).  LOSS = true additionally
// computes the per-sample loss / arg-max hit and dz = (softmax - onehot) / B; the block that finishes last (ticket
// counter) sums the per-sample terms in a fixed order into the mean loss (and the eval metrics).
template <bool LOSS>
__global__ void __launch_bounds__(256)
head_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                const long long* __restrict__ target, float* __restrict__ logits, float* __restrict__ dz,
                float2* __restrict__ terms /* [B] (loss_i, hit_i) */, float* __restrict__ loss_out,
                float* __restrict__ metrics, unsigned int* __restrict__ counter, int B, int K, int O) {
  DPB_TRACE(10);
  pdl_prologue();
  extern __shared__ float4 s_w4[];  // [O][K/4]
  __shared__ float s_part[2][8];
  __shared__ uint32_t s_flag;
  const int K4 = K >> 2;
  for (int i = threadIdx.x; i < O * K4; i += 256) s_w4[i] = __ldg(reinterpret_cast<const float4*>(w) + i);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 8 + warp;
  if (b < B) {
    float acc[kMaxClasses];
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j) acc[j] = 0.f;
    const float4* x4 = reinterpret_cast<const float4*>(x + (size_t)b * K);
#pragma unroll 4
    for (int k4 = lane; k4 < K4; k4 += 32) {
      const float4 xv = __ldg(x4 + k4);
#pragma unroll
      for (int j = 0; j < kMaxClasses; ++j) {
        if (j < O) {
          const float4 wv = s_w4[j * K4 + k4];
          acc[j] = fmaf(xv.x, wv.x, fmaf(xv.y, wv.y, fmaf(xv.z, wv.z, fmaf(xv.w, wv.w, acc[j]))));
        }
      }
    }
    float z[kMaxClasses];
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j) z[j] = (j < O) ? warp_sum(acc[j]) + (bias ? __ldg(bias + j) : 0.f) : 0.f;
    float mine = 0.f;  // lane j keeps logit j
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j)
      if (lane == j) mine = z[j];
    if (lane < O) logits[(size_t)b * O + lane] = mine;
    if (LOSS) {
      float m = z[0];
      int am = 0;
#pragma unroll
      for (int j = 1; j < kMaxClasses; ++j)
        if (j < O && z[j] > m) { m = z[j]; am = j; }
      float se = 0.f;
#pragma unroll
      for (int j = 0; j < kMaxClasses; ++j)
        if (j < O) se += expf(z[j] - m);
      const int t = (int)target[b];
      float zt = 0.f;
#pragma unroll
      for (int j = 0; j < kMaxClasses; ++j)
        if (j == t) zt = z[j];
      if (lane == 0) terms[b] = make_float2((m + logf(se)) - zt, am == t ? 1.f : 0.f);
      if (dz != nullptr && lane < O)
        dz[(size_t)b * O + lane] = (expf(mine - m) / se - (lane == t ? 1.f : 0.f)) / (float)B;
    }
  }
  if (LOSS) {
    if (arrive_last(counter, gridDim.x, (int)threadIdx.x, 0, 256, &s_flag)) {
      float l = 0.f, h = 0.f;
      for (int i = threadIdx.x; i < B; i += 256) {
        const float2 t = __ldcg(terms + i);
        l += t.x;
        h += t.y;
      }
      l = warp_sum(l);
      h = warp_sum(h);
      if (lane == 0) { s_part[0][warp] = l; s_part[1][warp] = h; }
      __syncthreads();
      if (threadIdx.x == 0) {
        float ls = 0.f, hs = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) { ls += s_part[0][q]; hs += s_part[1][q]; }
        ls /= (float)B;
        loss_out[0] = ls;
        if (metrics != nullptr) { metrics[0] += ls; metrics[1] += hs; }
      }
    }
  }
}

// Backward of the head in one launch: blocks [0, nx) produce dx = g * dz W, blocks [nx, nx + nw) a 32-column slice of
// dW = g * dz^T x each, the last block db = g * sum_b dz.  g = autograd's grad_output of the loss (device scalar).
__global__ void __launch_bounds__(256)
head_bwd_kernel(const float* __restrict__ dz, const float* __restrict__ x, const float* __restrict__ w,
                const float* __restrict__ grad_out, float* __restrict__ dx, float* __restrict__ dw,
                float* __restrict__ db, int B, int K, int O, int nx, int nw) {
  DPB_TRACE(11);
  pdl_prologue();
  __shared__ float s[8][kMaxClasses][33];
  const float g = grad_out ? grad_out[0] : 1.f;
  const int blk = blockIdx.x;
  if (blk < nx) {
    if (dx == nullptr) return;
    const long long i = (long long)blk * 256 + threadIdx.x;
    if (i >= (long long)B * K) return;
    const int b = (int)(i / K), k = (int)(i - (long long)b * K);
    float a = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j)
      if (j < O) a = fmaf(__ldg(dz + (size_t)b * O + j), __ldg(w + (size_t)j * K + k), a);
    dx[i] = a * g;
    return;
  }
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  if (blk < nx + nw) {
    const int k = (blk - nx) * 32 + cx;
    float acc[kMaxClasses];
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j) acc[j] = 0.f;
    if (k < K) {
#pragma unroll 8
      for (int b = ry; b < B; b += 8) {
        const float xv = __ldg(x + (size_t)b * K + k);
#pragma unroll
        for (int j = 0; j < kMaxClasses; ++j)
          if (j < O) acc[j] = fmaf(__ldg(dz + (size_t)b * O + j), xv, acc[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j) s[ry][j][cx] = acc[j];
    __syncthreads();
    if (k < K) {
      for (int j = ry; j < O; j += 8) {
        float r = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) r += s[q][j][cx];
        dw[(size_t)j * K + k] = r * g;
      }
    }
    return;
  }
  if (db == nullptr) return;
  const int t = threadIdx.x;
  float a[kMaxClasses];
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) a[j] = 0.f;
  for (int b = t; b < B; b += 256)
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j)
      if (j < O) a[j] += dz[(size_t)b * O + j];
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) {
    const float v = warp_sum(a[j]);
    if (cx == 0) s[ry][j][0] = v;
  }
  __syncthreads();
  if (t < O) {
    float r = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) r += s[q][t][0];
    db[t] = r * g;
  }
}

// Fused SGD over a flat fp32 range (torch.optim.SGD semantics, dampening 0):
//   g' = g * grad_scale + wd * p;  m = momentum * m + g'  (m starts at 0, so step 1 gives m = g');
//   p -= lr * (nesterov ? g' + momentum * m : m)
__global__ void __launch_bounds__(256)
sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, long long n4, float lr,
           float momentum, float wd, float grad_scale, int nesterov) {
  DPB_TRACE(12);
  pdl_prologue();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = __ldcs(reinterpret_cast<const float4*>(g) + i);
    float4 mv = (m != nullptr) ? reinterpret_cast<float4*>(m)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    float gg[4] = {gv.x, gv.y, gv.z, gv.w};
    float pp[4] = {pv.x, pv.y, pv.z, pv.w};
    float mm[4] = {mv.x, mv.y, mv.z, mv.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float ge = fmaf(wd, pp[k], gg[k] * grad_scale);
      float upd = ge;
      if (m != nullptr) {
        mm[k] = fmaf(momentum, mm[k], ge);
        upd = nesterov ? fmaf(momentum, mm[k], ge) : mm[k];
      }
      pp[k] = fmaf(-lr, upd, pp[k]);
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pp[0], pp[1], pp[2], pp[3]);
    if (m != nullptr) reinterpret_cast<float4*>(m)[i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
  }
}

// scalar tail / unaligned fallback
__global__ void sgd_scalar_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                  long long n, float lr, float momentum, float wd, float grad_scale, int nesterov) {
  pdl_prologue();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float ge = fmaf(wd, p[i], g[i] * grad_scale);
  float upd = ge;
  if (m != nullptr) {
    const float mv = fmaf(momentum, m[i], ge);
    m[i] = mv;
    upd = nesterov ? fmaf(momentum, mv, ge) : mv;
  }
  p[i] = fmaf(-lr, upd, p[i]);
}

}  // namespace dpb
