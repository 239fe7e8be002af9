// Small-layer BatchNorm backward as ONE thread-block-cluster kernel instead of three launches (optional path).
//
// For the deep VGG layers (and for every layer once the per-GPU batch is small) the BatchNorm work is a few MB and
// bn_bwd_reduce -> bn_bwd_finalize -> bn_bwd_apply are three dependent launches around a grid-wide reduction (the
// per-channel sums).  Here a cluster of 8 CTAs owns a group of 32 channels: every CTA reduces its share of the pixels,
// the 8 partial sums per channel are exchanged through distributed shared memory (each CTA reads its peers' partials
// in rank order, so the result is deterministic) and the same CTAs then apply dY = a*g + k1*y + k0 -- the grid-wide
// dependency has become a cluster barrier.  (reference model.py:16,24,25 via aten::native_batch_norm_backward,
// threshold_backward, max_pool2d_with_indices_backward)
//
// Measured on B200 inside the captured step it is a wash (batch 256: 0.971 vs 0.966 ms, batch 32: 0.485 vs 0.489 ms
// with 12 fewer launches): dependent small kernels cost ~1.3 us each in a CUDA graph, about what the two passes plus
// the cluster barrier cost here with only C/32 clusters in flight.  So it is off by default (dpb_set_tuning key 9 /
// DPB_TAIL_MAX_PIXELS enables it for layers up to that many pixels) and kept as a tested alternative.
//
// Work mapping: 512 threads = 8 channel chunks (float4) x 64 pixel lanes; cluster rank r owns the r-th eighth of the
// output pixels (a pooled pixel = one 2x2 window of conv pixels).  Both passes of a kernel walk the pixels in the same
// thread order, so the second pass re-reads what the same thread wrote or already pulled into L2.
#pragma once
#include "bn_relu_pool.cuh"

namespace dpb {

constexpr int kTailCL = 8;
constexpr int kTailLanes = 64;                 // pixel lanes per CTA: these layers are latency bound, so keep the
constexpr int kTailThreads = 8 * kTailLanes;   // per-thread pixel loops short (8 channel chunks x 64 lanes = 512 threads)

// Per-channel sums over the whole cluster.  a / b: this thread's partial sums for its 4 channels.  On return
// s_tot[which][ch] (which = 0 for a, 1 for b; ch = 0..31) holds the totals, identical in every CTA of the cluster.
__device__ __forceinline__ void tail_cluster_sums(const float4& a, const float4& b, float4 (&s_red)[2][kTailThreads],
                                                  float (&s_part)[2][32], double (&s_tot)[2][32]) {
  const int t = threadIdx.x;
  s_red[0][t] = a;
  s_red[1][t] = b;
  __syncthreads();
  if (t < 64) {  // pixel lanes -> one partial per channel, fixed order
    const int which = t >> 5, ch = t & 31;
    float acc = 0.f;
#pragma unroll 8
    for (int l = 0; l < kTailLanes; ++l) acc += reinterpret_cast<const float*>(&s_red[which][l * 8 + (ch >> 2)])[ch & 3];
    s_part[which][ch] = acc;
  }
  cluster_sync_all();  // every CTA's partials are published (release / acquire at cluster scope)
  if (t < 64) {
    const int which = t >> 5, ch = t & 31;
    double tot = 0.0;
#pragma unroll
    for (int r = 0; r < kTailCL; ++r) tot += (double)ld_dsmem_f32(mapa_u32(&s_part[which][ch], (uint32_t)r));
    s_tot[which][ch] = tot;
  }
  __syncthreads();
}

__device__ __forceinline__ long long tail_window_base(long long op, int H, int W) {
  const int Wo = W >> 1, Ho = H >> 1;
  const int wo = (int)(op % Wo);
  const long long t = op / Wo;
  const int ho = (int)(t % Ho);
  const long long n = t / Ho;
  return (n * H + 2 * ho) * W + 2 * wo;
}

struct TailBwdParams {
  const float* y;      // [M][C]
  const float* dout;   // [M or M/4][C]
  const float* scale;  // per-channel vectors saved by the forward pass
  const float* shift;
  const float* mean;
  const float* invstd;
  const float* gamma;
  float* dy;  // [M][C]
  float* dgamma;
  float* dbeta;
  float* conv_dbias;  // may be null; written as exact zero (see bn_bwd_finalize_kernel)
  int M, C, H, W;
};

template <bool POOL>
__global__ void __launch_bounds__(kTailThreads) tail_bwd_kernel(const TailBwdParams p) {
  pdl_prologue();
  __shared__ float4 s_red[2][kTailThreads];
  __shared__ float s_part[2][32];
  __shared__ double s_tot[2][32];
  __shared__ float s_coef[3][32];
  const int t = threadIdx.x, c8 = t & 7, pl = t >> 3;
  const int group = blockIdx.x / kTailCL;
  const int rank = (int)cluster_ctarank();
  const int C4 = p.C >> 2;
  const int chunk = group * 8 + c8;
  const int U = POOL ? p.M / 4 : p.M, upc = U / kTailCL, u0 = rank * upc;
  const float4 sc = __ldg(reinterpret_cast<const float4*>(p.scale) + chunk);
  const float4 sh = __ldg(reinterpret_cast<const float4*>(p.shift) + chunk);
  const float4 mu = __ldg(reinterpret_cast<const float4*>(p.mean) + chunk);
  const float4 is = __ldg(reinterpret_cast<const float4*>(p.invstd) + chunk);
  const float4* y4 = reinterpret_cast<const float4*>(p.y);
  const float4* d4 = reinterpret_cast<const float4*>(p.dout);

  // pass 1: sum(g), sum(g * xhat) with g = gradient after un-pooling and the ReLU mask
  float4 sg = make_float4(0.f, 0.f, 0.f, 0.f), sgx = sg;
  for (int j = pl; j < upc; j += kTailLanes) {
    const long long u = u0 + j;
    const float4 d = __ldg(d4 + u * C4 + chunk);
    if (POOL) {
      const long long p00 = tail_window_base(u, p.H, p.W);
      const float4 yy[4] = {__ldg(y4 + p00 * C4 + chunk), __ldg(y4 + (p00 + 1) * C4 + chunk),
                            __ldg(y4 + (p00 + p.W) * C4 + chunk), __ldg(y4 + (p00 + p.W + 1) * C4 + chunk)};
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
      const float4 yv = __ldg(y4 + u * C4 + chunk);
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
  tail_cluster_sums(sg, sgx, s_red, s_part, s_tot);
  if (t < 32) {
    const int c = group * 32 + t;
    const double count = (double)p.M;
    const double tg = s_tot[0][t], tgx = s_tot[1][t];
    const double isd = (double)p.invstd[c];
    const double ca = (double)p.gamma[c] * isd;
    const double k1 = -ca * isd * tgx / count;
    const double k0 = -ca * tg / count - k1 * (double)p.mean[c];
    s_coef[0][t] = (float)ca;
    s_coef[1][t] = (float)k1;
    s_coef[2][t] = (float)k0;
    if (rank == 0) {
      p.dgamma[c] = (float)tgx;
      p.dbeta[c] = (float)tg;
      if (p.conv_dbias != nullptr) p.conv_dbias[c] = 0.f;
    }
  }
  __syncthreads();
  // pass 2: dY = ca * g + k1 * y + k0
  const float4 ca = *reinterpret_cast<const float4*>(&s_coef[0][c8 * 4]);
  const float4 k1 = *reinterpret_cast<const float4*>(&s_coef[1][c8 * 4]);
  const float4 k0 = *reinterpret_cast<const float4*>(&s_coef[2][c8 * 4]);
  float4* dy4 = reinterpret_cast<float4*>(p.dy);
  for (int j = pl; j < upc; j += kTailLanes) {
    const long long u = u0 + j;
    const float4 d = __ldg(d4 + u * C4 + chunk);
    if (POOL) {
      const long long p00 = tail_window_base(u, p.H, p.W);
      const long long pp[4] = {p00, p00 + 1, p00 + p.W, p00 + p.W + 1};
      float4 yy[4];
      float zx[4], zy[4], zz[4], zw[4], gx[4], gy[4], gz[4], gw[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        yy[k] = __ldcs(y4 + pp[k] * C4 + chunk);
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
        dy4[pp[k] * C4 + chunk] = o;
      }
    } else {
      const float4 yv = __ldcs(y4 + u * C4 + chunk);
      const float4 z = bn4(yv, sc, sh);
      float4 o;
      o.x = fmaf(ca.x, z.x > 0.f ? d.x : 0.f, fmaf(k1.x, yv.x, k0.x));
      o.y = fmaf(ca.y, z.y > 0.f ? d.y : 0.f, fmaf(k1.y, yv.y, k0.y));
      o.z = fmaf(ca.z, z.z > 0.f ? d.z : 0.f, fmaf(k1.z, yv.z, k0.z));
      o.w = fmaf(ca.w, z.w > 0.f ? d.w : 0.f, fmaf(k1.w, yv.w, k0.w));
      dy4[u * C4 + chunk] = o;
    }
  }
  cluster_sync_all();
}

// A layer takes the cluster path when its channels come in groups of 32, its output pixels split evenly over the 8
// CTAs of a cluster and it is small enough that C/32 clusters beat a bandwidth-bound grid (dpb_set_tuning key 9).
inline int g_tune_tail_max_pixels = 0;  // 0 = off (see the measurement above)
inline bool tail_eligible(long long m_pixels, int c, int h, int w, bool pool) {
  if (g_tune_tail_max_pixels <= 0 || c % 32 != 0 || m_pixels > g_tune_tail_max_pixels) return false;
  if (pool && ((h | w) & 1)) return false;
  const long long units = pool ? m_pixels / 4 : m_pixels;
  return units > 0 && units % kTailCL == 0;
}

template <typename Params>
inline cudaError_t launch_tail(void (*kern)(Params), const Params& p, int c, cudaStream_t stream) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((c / 32) * kTailCL);
  cfg.blockDim = dim3(kTailThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kTailCL;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, p);
}

}  // namespace dpb
