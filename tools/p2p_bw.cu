// NVLink peer-access micro-benchmark behind the collective design (single process, GPUs 0 and 1, peer access enabled):
// how fast can SM kernels on GPU 0 pull from GPU 1 (remote loads -> local stores), push to GPU 1 (local loads -> remote
// stores), and do both at once, as a function of the grid size?  The two-shot all-reduce pulls its slice from every
// peer and pushes the result back in the same loop.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/p2p_bw.bin tools/p2p_bw.cu ; run with 2 visible GPUs
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(512) copy_kernel(const float4* __restrict__ src, float4* __restrict__ dst, long long n4) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n4; i += 4 * stride) {  // 4 independent 16-byte loads in flight per thread
    const float4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
    dst[i] = a; dst[i + stride] = b; dst[i + 2 * stride] = c; dst[i + 3 * stride] = d;
  }
  for (; i < n4; i += stride) dst[i] = src[i];
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

int main() {
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  if (ndev < 2) { printf("needs 2 GPUs (found %d)\n", ndev); return 0; }
  int can = 0;
  CK(cudaDeviceCanAccessPeer(&can, 0, 1));
  if (!can) { printf("no peer access between GPU 0 and 1\n"); return 0; }
  const size_t bytes = (size_t)256 << 20;
  const long long n4 = bytes / 16;
  float4 *l0, *l1, *r0, *r1;  // l*: on GPU 0, r*: on GPU 1
  CK(cudaSetDevice(1));
  CK(cudaDeviceEnablePeerAccess(0, 0));
  CK(cudaMalloc(&r0, bytes));
  CK(cudaMalloc(&r1, bytes));
  CK(cudaMemset(r0, 1, bytes));
  CK(cudaSetDevice(0));
  CK(cudaDeviceEnablePeerAccess(1, 0));
  CK(cudaMalloc(&l0, bytes));
  CK(cudaMalloc(&l1, bytes));
  CK(cudaMemset(l0, 1, bytes));
  cudaStream_t s1, s2;
  CK(cudaStreamCreate(&s1));
  CK(cudaStreamCreate(&s2));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  printf("256 MB per direction, kernels launched on GPU 0, 512 threads per CTA\n");
  for (int blocks : {16, 32, 64, 128, 296}) {
    float ms_pull = 0, ms_push = 0, ms_both = 0, ms_local = 0;
    for (int rep = 0; rep < 2; ++rep) {  // second repetition is the measured one
      CK(cudaEventRecord(a, s1));
      copy_kernel<<<blocks, 512, 0, s1>>>(r0, l1, n4);  // pull: remote loads
      CK(cudaEventRecord(b, s1));
      CK(cudaEventSynchronize(b));
      CK(cudaEventElapsedTime(&ms_pull, a, b));
      CK(cudaEventRecord(a, s1));
      copy_kernel<<<blocks, 512, 0, s1>>>(l0, r1, n4);  // push: remote stores
      CK(cudaEventRecord(b, s1));
      CK(cudaEventSynchronize(b));
      CK(cudaEventElapsedTime(&ms_push, a, b));
      CK(cudaEventRecord(a, s1));
      copy_kernel<<<blocks, 512, 0, s1>>>(l0, l1, n4);  // local copy for scale
      CK(cudaEventRecord(b, s1));
      CK(cudaEventSynchronize(b));
      CK(cudaEventElapsedTime(&ms_local, a, b));
      CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(a, s1));
      CK(cudaStreamWaitEvent(s2, a, 0));
      copy_kernel<<<blocks, 512, 0, s1>>>(r0, l1, n4);  // pull and push concurrently (two streams, half the SMs each
      copy_kernel<<<blocks, 512, 0, s2>>>(l0, r1, n4);  // at the larger grids)
      CK(cudaEventRecord(b, s2));
      CK(cudaStreamWaitEvent(s1, b, 0));
      CK(cudaEventRecord(b, s1));
      CK(cudaEventSynchronize(b));
      CK(cudaEventElapsedTime(&ms_both, a, b));
    }
    const double gb = bytes / 1e9;
    printf("%3d CTAs: pull %6.1f GB/s | push %6.1f GB/s | pull+push together %6.1f GB/s per direction | local copy %6.1f GB/s\n",
           blocks, gb / (ms_pull * 1e-3), gb / (ms_push * 1e-3), gb / (ms_both * 1e-3), gb / (ms_local * 1e-3));
  }
  return 0;
}
