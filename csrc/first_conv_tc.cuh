// First VGG layer on the tensor cores.  (reference model.py:13,18-23 with in_channels = 3.)
//
// The FP32-pipe kernels of first_conv.cuh are FMA-bound: 27 MACs x 64 channels per pixel is 1.8 GFMA at batch 256,
// ~55 us on 148 SMs x 128 lanes, while the layer's memory traffic (3 MB in, 67 MB out) is ~12 us.  Here the 27-deep
// reduction is padded to K = 32 and issued as four tcgen05.mma (M 128 pixels x N 64 channels x K 8, kind::tf32) per
// tile, which moves the layer onto its memory roofline.  TMA cannot gather a 3-channel NCHW halo into GEMM rows, so the
// threads build the A operand themselves: the haloed input patch goes to shared memory, then every thread writes the
// 32-float im2col row of "its" pixel straight into the SWIZZLE_128B K-major layout the MMA descriptor expects
// (16-byte chunk j of row r lives at chunk j ^ (r & 7)).  The 64 x 27 weights are laid out the same way once per CTA.
//
// A tile is 128 consecutive pixels of one image (128 / W full rows), so its NHWC output is one contiguous 32 KB block.
// A CTA walks tiles grid-stride; three CTAs per SM (64 KB shared memory, 64 TMEM columns each) overlap one another's
// patch loads, MMA round trips and stores.  BatchNorm partial sums are kept per CTA in registers across its tiles and
// written once, in a fixed order: part_sum / part_sq are [gridDim.x][64].
#pragma once
#include "ptx.cuh"
#include "bn_acc.cuh"

namespace dpb {

constexpr int kFcTcThreads = 128;
constexpr int kFcTcCtasPerSm = 3;
constexpr int kFcTcMaxBlocks = 148 * kFcTcCtasPerSm;
constexpr int kFcTcPatchFloats = 800;  // max over W in {8,16,32,64} of 3 * (128/W + 2) * (W + 2) = 792

struct FcTcSmem {
  static constexpr int kA = 0;                      // 128 pixel rows x 128 B
  static constexpr int kB = 16384;                  // 64 channel rows x 128 B
  static constexpr int kStage = kB + 8192;          // 128 x 64 fp32 output tile, XOR-swizzled 16-byte chunks
  static constexpr int kPatch = kStage + 32768;     // haloed input patch
  static constexpr int kRed = kPatch + kFcTcPatchFloats * 4;  // [4 warps][2][64]
  static constexpr int kBar = kRed + 4 * 2 * 64 * 4;
  static constexpr int kTotal = 1024 /*align slack*/ + kBar + 64;
};

__host__ __device__ inline bool first_conv_tc_supported(int h, int w) {
  return (w == 8 || w == 16 || w == 32 || w == 64) && ((h * w) % 128 == 0);
}

constexpr int kFcTcPrefetch = (kFcTcPatchFloats + kFcTcThreads - 1) / kFcTcThreads;

// Thread `tid`'s share of the haloed patch [3][R + 2][W + 2] of a tile.  The element -> (channel, row, column) split
// does not depend on the tile, so it is done once: `off` is the element's offset from the tile's first pixel in channel
// 0, `dh` its image row relative to the tile's first row (a huge negative number marks "always zero": halo column or no
// element).  Per tile that leaves one add, two compares and one load per element.
struct FcTcPatchPlan {
  int off[kFcTcPrefetch];
  int dh[kFcTcPrefetch];
};
__device__ __forceinline__ void fc_tc_plan_patch(int R, int H, int W, int tid, FcTcPatchPlan& plan) {
  const int WP = W + 2, RP = R + 2, elems = 3 * RP * WP;
#pragma unroll
  for (int i = 0; i < kFcTcPrefetch; ++i) {
    const int e = tid + i * kFcTcThreads;
    const int ci = e / (RP * WP);
    const int rem = e - ci * (RP * WP);
    const int r = rem / WP, c = rem - r * WP;
    const bool ok = e < elems && c >= 1 && c <= W;
    plan.off[i] = ci * H * W + (r - 1) * W + (c - 1);
    plan.dh[i] = ok ? r - 1 : -(1 << 28);
  }
}
// The patch of `tile` (zero outside the image) into registers.
__device__ __forceinline__ void fc_tc_fetch_patch(const float* __restrict__ x, int tile, int tiles_per_img, int R, int H,
                                                  int W, const FcTcPatchPlan& plan, float (&pre)[kFcTcPrefetch]) {
  const int n = tile / tiles_per_img, h0 = (tile - n * tiles_per_img) * R;
  const float* base = x + (long long)n * 3 * H * W + h0 * W;
#pragma unroll
  for (int i = 0; i < kFcTcPrefetch; ++i) {
    const int hh = h0 + plan.dh[i];
    pre[i] = (hh >= 0 && hh < H) ? __ldg(base + plan.off[i]) : 0.f;
  }
}

// x: [N][3][H][W] fp32 (NCHW).  w: [64][3][3][3] (KRSC).  y: [N*H*W][64].  part_sum / part_sq: [gridDim.x][64] or null.
// stat_acc (optional): [2][64] fixed-point BatchNorm accumulators (bn_acc.cuh) the CTA adds its column sums to.
__global__ void __launch_bounds__(kFcTcThreads)
first_conv_fprop_tc_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                           float* __restrict__ y, float* __restrict__ part_sum, float* __restrict__ part_sq, int N, int H,
                           int W, unsigned long long* __restrict__ stat_acc) {
  DPB_TRACE(8);
  pdl_prologue();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1024(smem_raw);
  float* sA = reinterpret_cast<float*>(smem + FcTcSmem::kA);
  float* sB = reinterpret_cast<float*>(smem + FcTcSmem::kB);
  float* stg = reinterpret_cast<float*>(smem + FcTcSmem::kStage);
  float* patch = reinterpret_cast<float*>(smem + FcTcSmem::kPatch);
  float* red = reinterpret_cast<float*>(smem + FcTcSmem::kRed);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + FcTcSmem::kBar);
  uint32_t* slot = reinterpret_cast<uint32_t*>(smem + FcTcSmem::kBar + 8);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<64>(slot);
  // weights: row co, chunk j holds k = 4j .. 4j+3 (k = (r*3+s)*3+ci, zero beyond 26)
  for (int idx = tid; idx < 64 * 8; idx += kFcTcThreads) {
    const int co = idx >> 3, j = idx & 7;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int k = j * 4 + e;
      v[e] = k < 27 ? __ldg(w + co * 27 + k) : 0.f;
    }
    *reinterpret_cast<float4*>(sB + co * 32 + ((j ^ (co & 7)) << 2)) = make_float4(v[0], v[1], v[2], v[3]);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;

  const int R = 128 / W, WP = W + 2, RP = R + 2;
  const int elems = 3 * RP * WP;
  const int tiles_per_img = H * W / 128;
  const int tiles = N * tiles_per_img;
  constexpr uint32_t idesc = umma_idesc_tf32(64, false, false);
  const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB);
  const int prow = tid / W, pcol = tid - prow * W;  // this thread's pixel inside the tile
  const int q = warp;                               // TMEM lane quarter
  uint32_t phase = 0;
  float acc_stat = 0.f;  // thread t: channel t & 63 of (t < 64 ? sum : sum of squares)
  const bool want_stats = part_sum != nullptr || stat_acc != nullptr;

  float pre[kFcTcPrefetch];  // the next tile's patch, in flight during this tile's MMA and epilogue
  FcTcPatchPlan plan;
  fc_tc_plan_patch(R, H, W, tid, plan);
  if ((int)blockIdx.x < tiles) fc_tc_fetch_patch(x, blockIdx.x, tiles_per_img, R, H, W, plan, pre);
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    // 1. haloed patch [3][RP][WP], zero outside the image
#pragma unroll
    for (int i = 0; i < kFcTcPrefetch; ++i) {
      const int e = tid + i * kFcTcThreads;
      if (e < elems) patch[e] = pre[i];
    }
    __syncthreads();
    if (tile + (int)gridDim.x < tiles) fc_tc_fetch_patch(x, tile + gridDim.x, tiles_per_img, R, H, W, plan, pre);
    // 2. im2col row of my pixel, written in the MMA's swizzled K-major layout
    {
      float v[32];
#pragma unroll
      for (int dr = 0; dr < 3; ++dr)
#pragma unroll
        for (int ds = 0; ds < 3; ++ds)
#pragma unroll
          for (int ci = 0; ci < 3; ++ci) v[(dr * 3 + ds) * 3 + ci] = patch[(ci * RP + prow + dr) * WP + pcol + ds];
#pragma unroll
      for (int k = 27; k < 32; ++k) v[k] = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(sA + tid * 32 + ((j ^ (tid & 7)) << 2)) =
            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
    fence_proxy_async();  // generic-proxy writes -> visible to the tensor core's async-proxy reads
    tc_fence_before();
    __syncthreads();
    // 3. D[128 x 64] = A[128 x 32] * B[64 x 32]^T
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < 4; ++ks)
        umma_tf32(tmem, umma_smem_desc(a_addr + ks * 32, 16, 1024, 2), umma_smem_desc(b_addr + ks * 32, 16, 1024, 2),
                  idesc, ks > 0 ? 1u : 0u);
      umma_commit(bar);
    }
    __syncwarp();
    mbar_wait(bar, phase, 700);
    phase ^= 1;
    tc_fence_after();
    // 4. epilogue: TMEM -> swizzled row-major staging -> contiguous 512-byte stores (two pixels per instruction)
    float* slab = stg + q * 32 * 64;
#pragma unroll
    for (int c0 = 0; c0 < 64; c0 += 32) {
      float v[32];
      tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int chunk = (c0 >> 2) + j;
        *reinterpret_cast<float4*>(slab + lane * 64 + ((chunk ^ (lane & 7)) << 2)) =
            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
    }
    __syncwarp();
    const int chunk = lane & 15;
    float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (bias != nullptr) b4 = __ldg(reinterpret_cast<const float4*>(bias) + chunk);
    float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f), t4 = s4;
    float4* out4 = reinterpret_cast<float4*>(y + ((size_t)tile * 128 + q * 32) * 64);
#pragma unroll 8
    for (int it = 0; it < 16; ++it) {
      const int row = it * 2 + (lane >> 4);
      float4 o = *reinterpret_cast<const float4*>(slab + row * 64 + ((chunk ^ (row & 7)) << 2));
      o.x += b4.x; o.y += b4.y; o.z += b4.z; o.w += b4.w;
      out4[it * 32 + lane] = o;
      s4.x += o.x; s4.y += o.y; s4.z += o.z; s4.w += o.w;
      t4.x += o.x * o.x; t4.y += o.y * o.y; t4.z += o.z * o.z; t4.w += o.w * o.w;
    }
    if (want_stats) {
      s4.x += __shfl_xor_sync(0xffffffffu, s4.x, 16); s4.y += __shfl_xor_sync(0xffffffffu, s4.y, 16);
      s4.z += __shfl_xor_sync(0xffffffffu, s4.z, 16); s4.w += __shfl_xor_sync(0xffffffffu, s4.w, 16);
      t4.x += __shfl_xor_sync(0xffffffffu, t4.x, 16); t4.y += __shfl_xor_sync(0xffffffffu, t4.y, 16);
      t4.z += __shfl_xor_sync(0xffffffffu, t4.z, 16); t4.w += __shfl_xor_sync(0xffffffffu, t4.w, 16);
      if (lane < 16) {
        *reinterpret_cast<float4*>(red + (q * 2 + 0) * 64 + (chunk << 2)) = s4;
        *reinterpret_cast<float4*>(red + (q * 2 + 1) * 64 + (chunk << 2)) = t4;
      }
    }
    tc_fence_before();
    __syncthreads();  // TMEM and the A tile may be overwritten by the next tile; `red` is complete
    if (want_stats) {
      const int which = tid >> 6, c = tid & 63;
#pragma unroll
      for (int ww = 0; ww < 4; ++ww) acc_stat += red[(ww * 2 + which) * 64 + c];
    }
    tc_fence_after();
  }
  if (part_sum != nullptr) (tid < 64 ? part_sum : part_sq)[(size_t)blockIdx.x * 64 + (tid & 63)] = acc_stat;
  if (stat_acc != nullptr) bn_acc_add(stat_acc + tid, acc_stat);  // tid < 64: sum of channel tid, else sum of squares
  __syncthreads();
  if (warp == 0) tmem_dealloc<64>(tmem);
}

}  // namespace dpb
