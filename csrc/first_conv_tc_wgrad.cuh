// First VGG layer, weight gradient on the tensor cores.  (reference model.py:13,18-23; autograd's conv backward.)
//
// dW[co][k] = sum over pixels of dY[pix][co] * patch[pix][k], as the GEMM  D[k][co] = P^T[k][pix] * dY[pix][co]:
// M = k (27 of 32 used), N = 64 channels, GEMM-K = pixels.  Both operands have the GEMM-K index (the pixel) as their
// slow memory axis, i.e. they are MN-major, which for tf32 means the SWIZZLE_128B_BASE32B layout: an operand atom is
// [pixel rows][128 B], 32-byte unit u of row r stored at unit u ^ (r & 3).
//   * A: the im2col row each thread builds for "its" pixel (as in the fprop kernel) IS one row of the A atom.
//   * B: dY rows (64 channels = 256 B) are two atoms; TMA (SWIZZLE_128B_ATOM_32B, box 32 channels x 128 pixels)
//     delivers them, up to two tiles ahead, so the 67 MB gradient stream does not wait for the math.
// tcgen05.mma has no M = 32, so the instruction runs at M = 128 with the three upper A atoms pointing at whatever
// follows in shared memory -- accumulator rows 32..127 are never read.  One TMEM accumulator lives across all tiles of
// the CTA (the tensor core does the reduction over pixels); every CTA writes one [64 x 27] partial, summed in a fixed
// order by column_reduce_kernel.  Two CTAs per SM: 2 A buffers + 2 dY stages = 96 KB of shared memory each (a first
// version with one CTA per SM and three stages was bound by instruction latency: one warp per scheduler).
#pragma once
#include <cuda.h>
#include "first_conv_tc.cuh"

namespace dpb {

constexpr int kFcTcWgradStages = 2;
constexpr int kFcTcWgradMaxBlocks = 296;  // two CTAs per SM (~101 KB each): two independent latency chains per SM

struct FcTcWgradSmem {
  static constexpr int kA = 0;                                    // 2 x (4 K blocks x [32 pixels][32 k])
  static constexpr int kB = 2 * 16384;                            // stages x 2 atoms x [128 pixels][32 co]
  static constexpr int kPatch = kB + kFcTcWgradStages * 32768;
  static constexpr int kBar = kPatch + kFcTcPatchFloats * 4;      // full[stages], done[2], TMEM slot
  static constexpr int kTotal = 1024 + kBar + 128;
};

// mapDy: dY as [pixels][64] fp32, box {32 channels, 128 pixels}, 32-byte-atom swizzle.  x: [N][3][H][W] (NCHW).
// partials: [gridDim.x][64 * 27].
__global__ void __launch_bounds__(kFcTcThreads, 1)
first_conv_wgrad_tc_kernel(const __grid_constant__ CUtensorMap mapDy, const float* __restrict__ x,
                           float* __restrict__ partials, int N, int H, int W) {
  DPB_TRACE(9);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1024(smem_raw);
  uint8_t* sA = smem + FcTcWgradSmem::kA;
  uint8_t* sB = smem + FcTcWgradSmem::kB;
  float* patch = reinterpret_cast<float*>(smem + FcTcWgradSmem::kPatch);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + FcTcWgradSmem::kBar);
  uint64_t* done_bar = full_bar + kFcTcWgradStages;
  uint32_t* slot = reinterpret_cast<uint32_t*>(done_bar + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    tma_prefetch_desc(&mapDy);
    for (int s = 0; s < kFcTcWgradStages; ++s) mbar_init(&full_bar[s], 1);
    mbar_init(&done_bar[0], 1);
    mbar_init(&done_bar[1], 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<64>(slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;
  pdl_prologue();  // dY is produced by the previous kernel

  const int R = 128 / W, WP = W + 2, RP = R + 2;
  const int elems = 3 * RP * WP;
  const int tiles_per_img = H * W / 128;
  const int tiles = N * tiles_per_img;
  const int my_tiles = ((int)blockIdx.x < tiles) ? (tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  constexpr uint32_t idesc = umma_idesc_tf32(64, true, true);
  const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB);
  const int prow = tid / W, pcol = tid - prow * W;
  const int kb = tid >> 5, pr = tid & 31;  // my pixel: K block and row inside it

  auto load_dy = [&](int i) {  // tile i of this CTA into stage i % stages (thread 0 only)
    const int s = i % kFcTcWgradStages;
    const int pix0 = (blockIdx.x + i * gridDim.x) * 128;
    mbar_arrive_expect_tx(&full_bar[s], 32768);
    tma_load_2d(sB + s * 32768, &mapDy, &full_bar[s], 0, pix0);
    tma_load_2d(sB + s * 32768 + 16384, &mapDy, &full_bar[s], 32, pix0);
  };
  if (tid == 0)
    for (int i = 0; i < kFcTcWgradStages && i < my_tiles; ++i) load_dy(i);
  float pre[kFcTcPrefetch];
  FcTcPatchPlan plan;
  fc_tc_plan_patch(R, H, W, tid, plan);
  if (my_tiles > 0) fc_tc_fetch_patch(x, blockIdx.x, tiles_per_img, R, H, W, plan, pre);

  for (int i = 0; i < my_tiles; ++i) {
    const int s = i % kFcTcWgradStages, ab = i & 1;
#pragma unroll
    for (int j = 0; j < kFcTcPrefetch; ++j) {
      const int e = tid + j * kFcTcThreads;
      if (e < elems) patch[e] = pre[j];
    }
    __syncthreads();
    if (i + 1 < my_tiles) fc_tc_fetch_patch(x, blockIdx.x + (i + 1) * gridDim.x, tiles_per_img, R, H, W, plan, pre);
    if (i >= 2) mbar_wait(&done_bar[ab], ((i - 2) >> 1) & 1, 720);  // the MMAs of tile i-2 have left this A buffer
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
      uint8_t* row = sA + ab * 16384 + kb * 4096 + pr * 128;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(row + ((j ^ ((pr & 3) << 1)) << 4)) =
            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      mbar_wait(&full_bar[s], (i / kFcTcWgradStages) & 1, 721);
      tc_fence_after();
#pragma unroll
      for (int b = 0; b < 4; ++b)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          umma_tf32(tmem, umma_smem_desc(a_addr + ab * 16384 + b * 4096 + ks * 1024, 4096, 512, 1),
                    umma_smem_desc(b_addr + s * 32768 + b * 4096 + ks * 1024, 16384, 512, 1), idesc,
                    (i > 0 || b > 0 || ks > 0) ? 1u : 0u);
      umma_commit(&done_bar[ab]);
      // the stage tile i-1 used is free once its MMAs retired: refill it with tile i+2
      if (i >= 1 && i - 1 + kFcTcWgradStages < my_tiles) {
        mbar_wait(&done_bar[(i - 1) & 1], ((i - 1) >> 1) & 1, 722);
        load_dy(i - 1 + kFcTcWgradStages);
      }
    }
    __syncwarp();
  }
  if (my_tiles > 0) mbar_wait(&done_bar[(my_tiles - 1) & 1], ((my_tiles - 1) >> 1) & 1, 723);
  tc_fence_after();
  if (warp == 0) {  // accumulator rows 0..31 = k, columns = co
    float* out = partials + (size_t)blockIdx.x * (64 * 27);
#pragma unroll
    for (int c0 = 0; c0 < 64; c0 += 32) {
      float v[32];
      if (my_tiles > 0) {
        tmem_ld_32x32(tmem + (uint32_t)c0, v);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.f;
      }
      if (lane < 27) {
#pragma unroll
        for (int j = 0; j < 32; ++j) out[(c0 + j) * 27 + lane] = v[j];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc<64>(tmem);
  }
}

}  // namespace dpb
