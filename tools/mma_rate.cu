// Micro-benchmark behind the conv GEMM design: how fast does ONE CTA get tcgen05.mma kind::tf32 through the tensor
// pipe, with operands already in shared memory (no TMA in the loop), as a function of N, of the number of
// accumulators the MMAs alternate between, and of how many CTAs share the SM?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -Icsrc -o tools/mma_rate.bin tools/mma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "ptx.cuh"

using namespace dpb;

// ACCS accumulators of N columns each; `iters` groups of 4 K-steps (one 32-wide K chunk) per accumulator, round robin
template <int N, int ACCS, bool MN>
__global__ void __launch_bounds__(128, 1) mma_rate_kernel(long long* cycles, int iters, int sync_every, const float* fill_src,
                                                            int* fills_out, int fill_warps) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint64_t fbar[4];
  __shared__ uint32_t slot;
  __shared__ volatile int done;
  for (int i = threadIdx.x; i < (16384 + N * 128) / 4; i += 128) reinterpret_cast<float*>(smem)[i] = 1.0f;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    for (int j = 0; j < 4; ++j) mbar_init(&fbar[j], 1);
    done = 0;
    fence_mbar_init();
  }
  fence_proxy_async();
  constexpr int COLS = (ACCS * N <= 256) ? 256 : 512;
  if (threadIdx.x < 32) tmem_alloc<COLS>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (threadIdx.x < 32) {
    constexpr uint32_t idesc = umma_idesc_tf32(N, MN, MN);
    constexpr uint32_t lbo = MN ? 4096 : 16, sbo = MN ? 512 : 1024, lt = MN ? 1 : 2, kstep = MN ? 1024 : 32;
    const uint32_t sa = smem_u32(smem), sb = sa + 16384;
    uint32_t phase = 0;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      if (elect_one()) {
#pragma unroll
        for (int acc = 0; acc < ACCS; ++acc) {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            const uint64_t ad = umma_smem_desc(sa + ks * kstep, lbo, sbo, lt);
            const uint64_t bd = umma_smem_desc(sb + ks * kstep, lbo, sbo, lt);
            umma_tf32(tmem + acc * N, ad, bd, idesc, (i > 0 || ks > 0) ? 1u : 0u);
          }
        }
        if (sync_every > 0 && (i + 1) % sync_every == 0) umma_commit(&bar);
        if (sync_every < 0 && (i + 1) % (-sync_every) == 0) umma_commit(&fbar[3]);  // commit, nobody waits
      }
      __syncwarp();
      if (sync_every > 0 && (i + 1) % sync_every == 0) {  // a dependent wait, like the smem-slot recycling of a real pipeline
        mbar_wait(&bar, phase, 900);
        phase ^= 1;
      }
    }
    if (elect_one()) umma_commit(&bar);
    __syncwarp();
    mbar_wait(&bar, phase, 901);
    const long long t1 = clock64();
    if (threadIdx.x == 0) {
      cycles[blockIdx.x] = t1 - t0;
      done = 1;
    }
  } else if ((threadIdx.x == 32 || (threadIdx.x == 64 && fill_warps == 2)) && fill_src != nullptr) {
    const int fw = threadIdx.x / 32 - 1;  // which fill warp
    uint64_t* fb = fbar + 2 * fw;
    // a second warp keeps two 32 KB bulk copies (global/L2 -> shared) in flight while the MMAs run: the shared-memory
    // fill traffic of a real TMA pipeline, without any dependency on the MMAs
    const float* src = fill_src + (size_t)blockIdx.x * 8192;
    uint8_t* dst = smem + 65536 + fw * 65536;
    uint32_t ph[2] = {0, 0};
    int n = 0;
    for (int j = 0; j < 2; ++j) {
      mbar_arrive_expect_tx(&fb[j], 32768);
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       smem_u32(dst + j * 32768)), "l"(src), "r"(32768), "r"(smem_u32(&fb[j])) : "memory");
    }
    while (!done) {
      const int j = n & 1;
      mbar_wait(&fb[j], ph[j], 950);
      ph[j] ^= 1;
      ++n;
      mbar_arrive_expect_tx(&fb[j], 32768);
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       smem_u32(dst + j * 32768)), "l"(src), "r"(32768), "r"(smem_u32(&fb[j])) : "memory");
    }
    mbar_wait(&fb[0], ph[0], 951);
    mbar_wait(&fb[1], ph[1], 952);
    atomicAdd(&fills_out[blockIdx.x], n);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after();
    tmem_dealloc<COLS>(tmem);
  }
}

template <int N, int ACCS, bool MN>
void run(int ctas_per_sm, int sync_every, long long* cyc, const float* fill_src = nullptr, int* fills = nullptr, int fill_warps = 1) {
  if (fills != nullptr) cudaMemset(fills, 0, 296 * sizeof(int));
  const int iters = 200;
  const int smem = ctas_per_sm == 1 ? 200 * 1024 : 100 * 1024;
  cudaFuncSetAttribute(mma_rate_kernel<N, ACCS, MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int ctas = 148 * ctas_per_sm;
  mma_rate_kernel<N, ACCS, MN><<<ctas, 128, smem>>>(cyc, iters, sync_every, fill_src, fills, fill_warps);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return; }
  long long h[296];
  cudaMemcpy(h, cyc, sizeof(long long) * ctas, cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < ctas; ++i) mx = h[i] > mx ? h[i] : mx;
  const double per_mma = (double)mx / (iters * 4.0 * ACCS);
  double fill_rate = 0.0;
  if (fills != nullptr) {
    int hf[296];
    cudaMemcpy(hf, fills, sizeof(int) * ctas, cudaMemcpyDeviceToHost);
    long long tot = 0;
    for (int i = 0; i < ctas; ++i) tot += hf[i];
    fill_rate = (double)tot / ctas * 32768.0 / (double)mx * ctas_per_sm;
  }
  printf("N %3d %s accs %d, %d CTA/SM, commit+wait every %3d chunks: %6.1f cyc/MMA per CTA -> SM tensor pipe %5.1f%% of the "
         "N/2-cycle floor", N, MN ? "MN-major" : "K-major ", ACCS, ctas_per_sm, sync_every, per_mma,
         100.0 * (N / 2.0) * ctas_per_sm / per_mma);
  if (fills != nullptr) printf(" | %d fill warp(s): concurrent bulk fill %5.1f B/clk/SM (operand reads %5.1f B/clk/SM)", fill_warps, fill_rate,
                               (4096.0 + N * 32.0) / per_mma * ctas_per_sm);
  printf("\n");
}

int main() {
  long long* cyc;
  cudaMalloc(&cyc, 296 * sizeof(long long));
  // TMEM is 512 columns per SM: the 2-CTA/SM runs keep ACCS * N <= 256 (each CTA allocates 256 columns)
  float* fill_src;
  int* fills;
  cudaMalloc(&fill_src, (size_t)296 * 32768);
  cudaMemset(fill_src, 0, (size_t)296 * 32768);
  cudaMalloc(&fills, 296 * sizeof(int));
  printf("-- MMAs alone vs MMAs with a concurrent shared-memory fill (1 CTA/SM, 200 KB of shared memory)\n");
  run<128, 1, false>(1, 0, cyc);
  run<128, 1, false>(1, 0, cyc, fill_src, fills);
  run<256, 1, false>(1, 0, cyc);
  run<256, 1, false>(1, 0, cyc, fill_src, fills);
  run<64, 1, false>(1, 0, cyc, fill_src, fills);
  run<128, 1, false>(1, 0, cyc, fill_src, fills, 2);
  run<256, 1, false>(1, 0, cyc, fill_src, fills, 2);
  printf("-- tcgen05.commit every n chunks of 4 MMAs WITHOUT waiting for it (n = 1, 2, 4)\n");
  for (int n : {-1, -2, -4}) {
    run<128, 1, false>(1, n, cyc);
    run<256, 1, false>(1, n, cyc);
    run<128, 1, false>(2, n, cyc);
  }
  printf("-- issue-rate table\n");
  for (int sync_every : {0, 1, 3}) {
    run<128, 1, false>(1, sync_every, cyc);
    run<128, 2, false>(1, sync_every, cyc);
    run<256, 1, false>(1, sync_every, cyc);
    run<256, 2, false>(1, sync_every, cyc);
    run<64, 1, false>(1, sync_every, cyc);
    run<128, 1, true>(1, sync_every, cyc);
    run<256, 1, true>(1, sync_every, cyc);
    run<128, 1, false>(2, sync_every, cyc);
    run<128, 2, false>(2, sync_every, cyc);
    run<128, 1, true>(2, sync_every, cyc);
  }
  return 0;
}
