// Micro-benchmark behind the conv epilogue design: how fast can a CTA push a 128 x 128 fp32 accumulator tile (64 KB,
// already in shared memory) to global memory, with every CTA of the grid storing at the same time (as the lock-step
// CTAs of a one-wave GEMM do)?   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o store_bw tools/store_bw.cu
//   variant 0: per-warp 32x32 chunks through a stride-33 transpose tile, one 128-byte row per store instruction
//              (the epilogue as first written)
//   variant 1: row-major tile in smem, LDS.128 -> STG.128: one 512-byte row per store instruction
//   variant 2: bulk async copies smem -> global (cp.async.bulk, 512 B per row), issued by one warp
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int kRows = 128, kCols = 128;

template <int VARIANT>
__global__ void __launch_bounds__(128) store_kernel(float* out, int out_ld, long long* cycles, int reps) {
  extern __shared__ __align__(1024) float tile[];  // 64 KB (+ padding for variant 0)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < kRows * kCols; i += 128) tile[i] = (float)(i + blockIdx.x);
  __syncthreads();
  float* base = out + (size_t)blockIdx.x * kRows * out_ld;
  const long long t0 = clock64();
  for (int r = 0; r < reps; ++r) {
    if (VARIANT == 0) {
      float* t = tile + warp * (32 * 33);
      for (int c0 = 0; c0 < kCols; c0 += 32) {
#pragma unroll 4
        for (int i = 0; i < 32; ++i) base[(size_t)(warp * 32 + i) * out_ld + c0 + lane] = t[i * 33 + lane];
        __syncwarp();
      }
    } else if (VARIANT == 1) {
#pragma unroll 4
      for (int i = 0; i < 32; ++i) {
        const float4 v = reinterpret_cast<const float4*>(tile + (warp * 32 + i) * kCols)[lane];
        reinterpret_cast<float4*>(base + (size_t)(warp * 32 + i) * out_ld)[lane] = v;
      }
    } else {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      if (warp == 0) {
        for (int i = lane; i < kRows; i += 32) {
          const uint32_t src = (uint32_t)__cvta_generic_to_shared(tile + i * kCols);
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 512;" ::"l"(base + (size_t)i * out_ld),
                       "r"(src)
                       : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

template <int VARIANT>
void run(const char* name, int ctas, int out_ld, float* out, long long* cyc) {
  const int smem = 100 * 1024;  // two CTAs per SM, like the conv kernel
  cudaFuncSetAttribute(store_kernel<VARIANT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int reps = 1;
  store_kernel<VARIANT><<<ctas, 128, smem>>>(out, out_ld, cyc, reps);
  cudaDeviceSynchronize();
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  cudaEventRecord(a);
  for (int i = 0; i < 20; ++i) store_kernel<VARIANT><<<ctas, 128, smem>>>(out, out_ld, cyc, reps);
  cudaEventRecord(b);
  cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  long long h[1024];
  cudaMemcpy(h, cyc, sizeof(long long) * ctas, cudaMemcpyDeviceToHost);
  long long mx = 0, sum = 0;
  for (int i = 0; i < ctas; ++i) { mx = h[i] > mx ? h[i] : mx; sum += h[i]; }
  printf("%-28s ctas %4d ld %4d: store phase %6lld cyc avg %6lld max | kernel %6.2f us (incl. smem fill) | %5.2f TB/s in the store phase at 1.9 GHz\n",
         name, ctas, out_ld, sum / ctas, mx, ms * 1e3 / 20, (double)ctas * 65536 / ((double)mx / 1.9e9) / 1e12);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(e));
}

int main() {
  float* out;
  long long* cyc;
  cudaMalloc(&out, (size_t)1024 * kRows * 512 * 4);
  cudaMalloc(&cyc, 1024 * sizeof(long long));
  for (int ctas : {74, 148, 296, 592}) {
    for (int ld : {128, 256, 512}) {
      run<0>("transpose, 128 B/store", ctas, ld, out, cyc);
      run<1>("row-major, 512 B/store", ctas, ld, out, cyc);
      run<2>("bulk async copy 512 B/row", ctas, ld, out, cyc);
    }
  }
  return 0;
}
