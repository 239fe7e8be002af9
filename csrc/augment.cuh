// Input pipeline on the device: gather a batch out of a uint8 dataset that is resident in HBM and apply the
// reference's transforms in one pass -- RandomCrop(32, padding=4) -> RandomHorizontalFlip -> ToTensor -> Normalize for
// the training split, ToTensor -> Normalize for the test split (reference main.py:71-82) -- writing the fp32 NCHW batch
// the first-layer kernel consumes.  Replaces the DataLoader worker processes + pin-memory thread of the reference
// (main.py:85-90): with a ~1 ms training step a host-side augmenter would be the bottleneck.
//
// The crop offset and flip bit of a sample are a hash of (seed, epoch, dataset index) -- the same arithmetic as
// augment_draws() in distributed_pytorch_b200/data/datasets.py, so the host path is a bit-exact oracle.
#pragma once
#include <cstdint>
#include "ptx.cuh"

namespace dpb {

constexpr int kAugPad = 4;

__host__ __device__ __forceinline__ uint32_t augment_hash(uint32_t idx, uint32_t seed, uint32_t epoch) {
  uint32_t h = idx * 0x9E3779B1u + seed * 0x85EBCA77u + epoch * 0xC2B2AE3Du;
  h ^= h >> 16;
  h *= 0x85EBCA6Bu;
  h ^= h >> 13;
  h *= 0xC2B2AE35u;
  h ^= h >> 16;
  return h;
}

struct AugmentNorm {
  float mean[3];
  float inv_std[3];
};

// grid = B blocks of 256 threads; a thread produces float4s (4 neighbouring x of one row and channel).
// images: [N][3][H][W] uint8, labels: [N] int64, indices: [B] int32 dataset indices.
// draws: optional [B][3] int32 (dx, dy, flip) overriding the hash (tests).
__global__ void __launch_bounds__(256)
augment_batch_kernel(const uint8_t* __restrict__ images, const long long* __restrict__ labels,
                     const int* __restrict__ indices, int H, int W, int train, uint32_t seed, uint32_t epoch,
                     const int* __restrict__ draws, AugmentNorm nrm, float* __restrict__ out_x,
                     long long* __restrict__ out_y) {
  const int b = blockIdx.x;
  const int src = indices[b];
  int dx = kAugPad, dy = kAugPad, flip = 0;  // identity crop
  if (train) {
    if (draws != nullptr) {
      dx = draws[3 * b]; dy = draws[3 * b + 1]; flip = draws[3 * b + 2];
    } else {
      const uint32_t h = augment_hash((uint32_t)src, seed, epoch);
      dx = (int)((h & 0xFFu) % (2 * kAugPad + 1));
      dy = (int)(((h >> 8) & 0xFFu) % (2 * kAugPad + 1));
      flip = (int)((h >> 16) & 1u);
    }
  }
  if (threadIdx.x == 0) out_y[b] = labels[src];
  const uint8_t* img = images + (size_t)src * 3 * H * W;
  float* dst = out_x + (size_t)b * 3 * H * W;
  const int W4 = W >> 2;
  const int units = 3 * H * W4;
  for (int u = threadIdx.x; u < units; u += blockDim.x) {
    const int x4 = u % W4;
    const int y = (u / W4) % H;
    const int c = u / (W4 * H);
    const int sy = y + dy - kAugPad;
    const float mean = nrm.mean[c], inv = nrm.inv_std[c];
    float v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int x = x4 * 4 + k;
      const int sx = (flip ? (W - 1 - x) : x) + dx - kAugPad;
      float p = 0.f;  // RandomCrop pads with black BEFORE normalisation
      if (sy >= 0 && sy < H && sx >= 0 && sx < W) p = (float)img[(c * H + sy) * W + sx] / 255.0f;
      v[k] = (p - mean) * inv;
    }
    *reinterpret_cast<float4*>(dst + (size_t)(c * H + y) * W + x4 * 4) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

}  // namespace dpb
