// libdpb200.so: C ABI over the sm_100a kernels (see include/dpb200.h).  Single translation unit.
#include <cstring>
#include <mutex>
#include <vector>
#include <new>
#include <cuda_runtime.h>

#include "../include/dpb200.h"
#include "errors.cuh"
#include "ptx.cuh"
#include "conv_gemm.cuh"
#include "conv_gemm_host.cuh"
#include "bn_relu_pool.cuh"
#include "standalone_ops.cuh"
#include "first_conv.cuh"
#include "first_conv_tc.cuh"
#include "first_conv_tc_wgrad.cuh"
#include "head_sgd.cuh"
#include "comm.cuh"
#include "augment.cuh"
#include "vmm.cuh"

using namespace dpb;

namespace {
int g_tune_comm_blocks = 64;  // default grid cap of the collective kernels (dpb_set_tuning key 2)
int g_tune_comm_weak = 0;     // two-shot all-reduce payload with weak (.cv load / plain store) accesses (key 11)
int g_tune_ddp_blocks = 16;   // grid of a per-bucket collective launched next to backward (key 13): few CTAs, so the
                              // conv kernels it overlaps keep their SM slots (8 GPUs: 0.463 ms/step with 16, 0.494 with
                              // 32, 0.505 with 64)
int g_tune_nvls = 1;          // use the multimem (in-switch reduction) all-reduce when the buffers are multicast-bound (key 14)
int g_tune_nvls_blocks = 16;  // its grid when it has the GPU to itself (key 15): 36.9 MB at 8 GPUs takes 115 us with 8
                              // CTAs, 124 with 16, 136 with 32..128 -- the switch, not the SMs, sets the pace
inline cudaStream_t S(dpb_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }
constexpr int kBnBwdBlocks = 296;
constexpr int kFirstWgradBlocks = 444;  // rows of the first-layer wgrad partials buffer (= kFcTcMaxBlocks)
int g_tune_first_tc = 1;  // first layer on the tensor cores when the image shape allows (dpb_set_tuning key 7)
inline int grid_for(long long units, int threads, int cap) {
  long long b = (units + threads - 1) / threads;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}
}  // namespace

extern "C" {

int dpb_version(void) { return 202; }
const char* dpb_last_error(void) { return error_buffer(); }

}  // extern "C"

namespace {
// One DevError record per device in mapped pinned host memory (see ptx.cuh): written by a kernel before it traps, read
// here without any CUDA call, because after the trap the context only returns errors.
constexpr int kMaxDevices = 64;
DevError* g_host_err[kMaxDevices] = {nullptr};
std::mutex g_init_mutex;
unsigned long long g_peer_timeout_ns = 120000000000ull;

int init_current_device() {
  int dev = -1;
  DPB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return set_error(-31, "device ordinal %d unsupported", dev);
  std::lock_guard<std::mutex> lock(g_init_mutex);
  if (g_host_err[dev] == nullptr) {
    DevError* h = nullptr;
    DPB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&h), sizeof(DevError), cudaHostAllocMapped | cudaHostAllocPortable));
    memset(h, 0, sizeof(DevError));
    DevError* d = nullptr;
    DPB_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&d), h, 0));
    DPB_CUDA(cudaMemcpyToSymbol(g_dpb_dev_error_host, &d, sizeof(d)));
    g_host_err[dev] = h;
  }
  DPB_CUDA(cudaMemcpyToSymbol(g_dpb_spin_timeout_ns, &g_peer_timeout_ns, sizeof(g_peer_timeout_ns)));
  return 0;
}
}  // namespace

extern "C" {

int dpb_init_device(void) { return init_current_device(); }

int dpb_last_device_error(int device, int* out6) {
  for (int i = 0; i < 6; ++i) out6[i] = 0;
  for (int dev = 0; dev < kMaxDevices; ++dev) {
    if (device >= 0 && dev != device) continue;
    DevError* h = g_host_err[dev];
    if (h == nullptr) continue;
    volatile int* hv = reinterpret_cast<volatile int*>(h);
    if (hv[0] == 0) continue;
    for (int i = 0; i < 6; ++i) out6[i] = hv[i];
    for (int i = 5; i >= 0; --i) hv[i] = 0;
    return 0;
  }
  return 0;
}

int dpb_debug_set_conv_dump(void* buf) {
#ifdef DPB_EXPERIMENTS
  g_conv_debug_dump = reinterpret_cast<float*>(buf);
  return 0;
#else
  (void)buf;
  return set_error(-30, "conv tracing is compiled out: rebuild libdpb200 with -DDPB_EXPERIMENTS");
#endif
}

int dpb_debug_set_trace(void* buf) {
#ifdef DPB_EXPERIMENTS
  unsigned long long* p = reinterpret_cast<unsigned long long*>(buf);
  DPB_CUDA(cudaMemcpyToSymbol(g_dpb_trace, &p, sizeof(p)));
  return 0;
#else
  (void)buf;
  return set_error(-30, "kernel tracing is compiled out: rebuild libdpb200 with -DDPB_EXPERIMENTS");
#endif
}

int dpb_set_tuning(int key, int value) {
  if (key == 0 && value >= 1) { g_tune_target_ctas = value; return 0; }
  if (key == 1 && (value == 64 || value == 128 || value == 256)) { g_tune_wide_nt = value; return 0; }
  if (key == 2 && value >= 1 && value <= kCommMaxBlocks) { g_tune_comm_blocks = value; return 0; }
  if (key == 3 && value >= 1) { g_tune_min_k_iters = value; return 0; }
  if (key == 4) { g_tune_use_pdl = value ? 1 : 0; return 0; }
  if (key == 5 && (value == 1 || value == 2 || value == 4 || value == 8 || value == 16)) { g_tune_cluster_max = value; return 0; }
  if (key == 7) { g_tune_first_tc = value ? 1 : 0; return 0; }
  if (key == 11) { g_tune_comm_weak = value ? 1 : 0; return 0; }
  if (key == 12 && value >= 1) {  // peer-signal wait budget in milliseconds; applied to the current device at once
    g_peer_timeout_ns = (unsigned long long)value * 1000000ull;
    return init_current_device();
  }
  if (key == 13 && value >= 1 && value <= kCommMaxBlocks) { g_tune_ddp_blocks = value; return 0; }
  if (key == 14) { g_tune_nvls = value ? 1 : 0; return 0; }
  if (key == 16 && value >= 2 && value <= 64) { g_tune_cluster_reach = value; return 0; }
  if (key == 15 && value >= 1 && value <= kCommMaxBlocks) { g_tune_nvls_blocks = value; return 0; }
  return set_error(-30, "dpb_set_tuning: bad key/value %d/%d", key, value);
}

int dpb_conv_last_launches(void) { return g_conv_last_launches; }

size_t dpb_conv_workspace_bytes(void) { return kConvWorkspaceBytes; }

// ------------------------------------------------------------------------------------------------ convolution
int dpb_conv3x3_gemm(int mode, int n, int h, int w, int cin, int cout, const float* in, const float* in2,
                     const float* weight, float* out, const float* bias, void* stat_acc, void* workspace, int nt,
                     int splits, int target_ctas, dpb_stream_t stream) {
  if (mode < 0 || mode > 2) return set_error(-1, "dpb_conv3x3_gemm: bad mode %d", mode);
  ConvGeom g{n, h, w, cin, cout};
  return conv_gemm_dispatch(mode, g, in, in2, weight, out, bias, (unsigned long long*)stat_acc, (float*)workspace, nt,
                            splits, target_ctas, S(stream));
}

int dpb_first_conv_blocks(int n, int h, int w) {
  if (g_tune_first_tc && first_conv_tc_supported(h, w)) {
    const long long tiles = (long long)n * h * w / 128;
    return (int)(tiles < kFcTcMaxBlocks ? tiles : kFcTcMaxBlocks);
  }
  const long long b = (long long)n * ((h + kFcRows - 1) / kFcRows);  // tiles of kFcRows image rows
  return (int)(b < kFirstConvMaxBlocks ? b : kFirstConvMaxBlocks);
}

static int first_conv_check(int w) {
  if (w < 2 || (w & 1) || w > kFcMaxW) return set_error(-7, "first conv: image width must be even and <= %d (got %d)", kFcMaxW, w);
  return 0;
}

int dpb_first_conv_uses_tensor_cores(int h, int w) { return (g_tune_first_tc && first_conv_tc_supported(h, w)) ? 1 : 0; }

int dpb_first_conv_fprop(const float* x, const float* weight, const float* bias, float* y, float* part_sum,
                         float* part_sq, void* stat_acc, int n, int h, int w, dpb_stream_t stream) {
  const int blocks = dpb_first_conv_blocks(n, h, w);
  if (g_tune_first_tc && first_conv_tc_supported(h, w)) {
    DPB_SET_MAX_SMEM_ONCE(first_conv_fprop_tc_kernel, FcTcSmem::kTotal);
    DPB_CUDA(launch_k(first_conv_fprop_tc_kernel, dim3(blocks), dim3(kFcTcThreads), FcTcSmem::kTotal, S(stream), x, weight, bias, y, part_sum, part_sq, n, h, w, (unsigned long long*)stat_acc));
    return 0;
  }
  if (stat_acc != nullptr) return set_error(-7, "first conv: the FP32-pipe kernel leaves partial rows, not accumulators");
  if (int rc = first_conv_check(w)) return rc;
  DPB_CUDA(launch_k(first_conv_fprop_kernel, dim3(blocks), dim3(256), 0, S(stream), x, weight, bias, y, part_sum, part_sq, n, h, w));
  return 0;
}

int dpb_first_conv_wgrad_blocks(void) { return kFirstWgradBlocks; }

int dpb_first_conv_wgrad(const float* x, const float* dy, float* partials, float* dw, int n, int h, int w,
                         dpb_stream_t stream) {
  const int nout = 64 * 27;
  if (g_tune_first_tc && first_conv_tc_supported(h, w)) {
    DPB_SET_MAX_SMEM_ONCE(first_conv_wgrad_tc_kernel, FcTcWgradSmem::kTotal);
    const long long tiles = (long long)n * h * w / 128;
    const int blocks = (int)(tiles < kFcTcWgradMaxBlocks ? tiles : kFcTcWgradMaxBlocks);
    CUtensorMap map_dy;
    cuuint64_t dims[2] = {64, (cuuint64_t)tiles * 128};
    cuuint64_t str[1] = {64 * 4};
    cuuint32_t box[2] = {32, 128};
    if (int rc = make_map_f32(&map_dy, dy, 2, dims, str, box, /*mn_major=*/true)) return rc;
    DPB_CUDA(launch_k(first_conv_wgrad_tc_kernel, dim3(blocks), dim3(kFcTcThreads), FcTcWgradSmem::kTotal, S(stream), map_dy, x, partials, n, h, w));
    DPB_CUDA(launch_k(column_reduce_kernel, dim3((nout + 31) / 32), dim3(32, 8), 0, S(stream), partials, blocks, nout, dw));
    return 0;
  }
  if (int rc = first_conv_check(w)) return rc;
  int blocks = dpb_first_conv_blocks(n, h, w);
  if (blocks > kFirstConvMaxBlocks) blocks = kFirstConvMaxBlocks;
  DPB_CUDA(launch_k(first_conv_wgrad_kernel, dim3(blocks), dim3(256), 0, S(stream), x, dy, partials, n, h, w));
  DPB_CUDA(launch_k(column_reduce_kernel, dim3((nout + 31) / 32), dim3(32, 8), 0, S(stream), partials, blocks, nout, dw));
  return 0;
}

// ------------------------------------------------------------------------------------------------ BN + ReLU + pool
int dpb_bn_finalize(const float* part_sum, const float* part_sq, int parts, int c, double count, const float* gamma,
                    const float* beta, float* running_mean, float* running_var, long long* nbt, float momentum,
                    float eps, float* scale, float* shift, float* save_mean, float* save_invstd,
                    dpb_stream_t stream) {
  DPB_CUDA(launch_k(bn_finalize_kernel, dim3((c + kFinCh - 1) / kFinCh), dim3(kFinCh, kFinRows), 0, S(stream), part_sum, part_sq, parts, c, count, gamma, beta, running_mean, running_var, nbt, momentum, eps, scale, shift, save_mean, save_invstd));
  return 0;
}

int dpb_bn_eval_scale_shift(int c, const float* gamma, const float* beta, const float* running_mean,
                            const float* running_var, float eps, float* scale, float* shift, dpb_stream_t stream) {
  DPB_CUDA(launch_k(bn_eval_scale_shift_kernel, dim3((c + 127) / 128), dim3(128), 0, S(stream), c, gamma, beta, running_mean, running_var, eps, scale, shift));
  return 0;
}

int dpb_bn_relu_pool_fwd(const float* y, const float* scale, const float* shift, float* out, int n, int h, int w,
                         int c, int pool, dpb_stream_t stream) {
  if (c % 4) return set_error(-2, "bn_relu_pool: C must be a multiple of 4 (got %d)", c);
  if (pool && ((h | w) & 1)) return set_error(-2, "bn_relu_pool: pooling needs even H, W (got %dx%d)", h, w);
  const long long opix = pool ? (long long)n * (h / 2) * (w / 2) : (long long)n * h * w;
  const long long units = opix * (c / 4);
  const int blocks = grid_for(units, 256, 148 * 16);
  if (pool) {
    DPB_CUDA(launch_k(bn_relu_pool_fwd_kernel<true>, dim3(blocks), dim3(256), 0, S(stream), y, scale, shift, out, units, c / 4, h, w));
  } else {
    DPB_CUDA(launch_k(bn_relu_pool_fwd_kernel<false>, dim3(blocks), dim3(256), 0, S(stream), y, scale, shift, out, units, c / 4, h, w));
  }
  return 0;
}

int dpb_bn_relu_pool_fwd_train(const float* y, void* acc, void* ticket, const float* gamma, const float* beta,
                               float* running_mean, float* running_var, long long* nbt, float momentum, float eps,
                               float* stats4, float* out, int n, int h, int w, int c, int pool, dpb_stream_t stream) {
  if (c % 4 || c > kBnMaxC) return set_error(-2, "bn_relu_pool: C must be a multiple of 4 and <= %d (got %d)", kBnMaxC, c);
  if (pool && ((h | w) & 1)) return set_error(-2, "bn_relu_pool: pooling needs even H, W (got %dx%d)", h, w);
  BnTrainArgs a{};
  a.acc = reinterpret_cast<unsigned long long*>(acc);
  a.ticket = reinterpret_cast<unsigned int*>(ticket);
  a.gamma = gamma; a.beta = beta; a.running_mean = running_mean; a.running_var = running_var;
  a.num_batches_tracked = nbt; a.momentum = momentum; a.eps = eps; a.count = (double)n * h * w; a.stats4 = stats4;
  const long long opix = pool ? (long long)n * (h / 2) * (w / 2) : (long long)n * h * w;
  const long long units = opix * (c / 4);
  // one float4 per thread up to 8 blocks per SM: the element loop is latency-bound, so it wants the loads of many
  // threads in flight rather than a serial loop per thread; the per-block prologue (C channels) runs concurrently
  const int blocks = grid_for(units, 256, 148 * 8);
  if (pool) {
    DPB_CUDA(launch_k(bn_relu_pool_fwd_acc_kernel<true>, dim3(blocks), dim3(256), 0, S(stream), y, a, out, units, c, h, w));
  } else {
    DPB_CUDA(launch_k(bn_relu_pool_fwd_acc_kernel<false>, dim3(blocks), dim3(256), 0, S(stream), y, a, out, units, c, h, w));
  }
  return 0;
}

int dpb_bn_bwd_blocks(void) { return kBnBwdBlocks; }

int dpb_bn_bwd_launches(int c, int with_acc) { return (with_acc && c % 32 == 0) ? 2 : 3; }

int dpb_bn_relu_pool_bwd(const float* y, const float* dout, const float* scale, const float* shift,
                         const float* save_mean, const float* save_invstd, const float* gamma, float* part,
                         float* coef, void* acc, void* ticket, float* dgamma, float* dbeta, float* conv_dbias, float* dy,
                         int n, int h, int w, int c, int pool, dpb_stream_t stream) {
  if (c % 4 || c > 1024) return set_error(-2, "bn_relu_pool_bwd: C must be a multiple of 4 and <= 1024 (got %d)", c);
  const int c4 = c / 4;
  const long long opix = pool ? (long long)n * (h / 2) * (w / 2) : (long long)n * h * w;
  float* part_g = part;
  float* part_gx = part + (size_t)kBnBwdBlocks * c;
  const double count = (double)n * h * w;
  const long long units = opix * c4;
  if (acc != nullptr && ticket != nullptr && c % 32 == 0) {
    // two launches: column sums into fixed-point accumulators, then the apply pass derives its coefficients from them
    const int groups = c / 32;
    long long slices = (2 * kBnBwdBlocks) / groups;  // ~4 blocks per SM in total
    const long long by_pixels = (opix + 31) / 32;
    if (slices > by_pixels) slices = by_pixels;
    if (slices < 1) slices = 1;
    unsigned long long* a = reinterpret_cast<unsigned long long*>(acc);
    unsigned int* t = reinterpret_cast<unsigned int*>(ticket);
    const int ablocks = grid_for(units, 256, 148 * 8);
    if (pool) {
      DPB_CUDA(launch_k(bn_bwd_reduce_acc_kernel<true>, dim3(groups, (unsigned)slices), dim3(256), 0, S(stream), y, dout, scale, shift, save_mean, save_invstd, a, opix, c, h, w));
      DPB_CUDA(launch_k(bn_bwd_apply_acc_kernel<true>, dim3(ablocks), dim3(256), 0, S(stream), y, dout, scale, shift, save_mean, save_invstd, gamma, a, t, count, dgamma, dbeta, conv_dbias, dy, units, c, h, w));
    } else {
      DPB_CUDA(launch_k(bn_bwd_reduce_acc_kernel<false>, dim3(groups, (unsigned)slices), dim3(256), 0, S(stream), y, dout, scale, shift, save_mean, save_invstd, a, opix, c, h, w));
      DPB_CUDA(launch_k(bn_bwd_apply_acc_kernel<false>, dim3(ablocks), dim3(256), 0, S(stream), y, dout, scale, shift, save_mean, save_invstd, gamma, a, t, count, dgamma, dbeta, conv_dbias, dy, units, c, h, w));
    }
    return 0;
  }
  {
    const int pl = 256 / c4;
    int blocks = (int)((opix + pl - 1) / pl);
    if (blocks > kBnBwdBlocks) blocks = kBnBwdBlocks;
    const size_t smem = 2 * 256 * sizeof(float4);
    if (pool) {
      DPB_CUDA(launch_k(bn_bwd_reduce_kernel<true>, dim3(blocks), dim3(256), smem, S(stream), y, dout, scale, shift, save_mean, save_invstd, part_g, part_gx, opix, c4, h, w));
    } else {
      DPB_CUDA(launch_k(bn_bwd_reduce_kernel<false>, dim3(blocks), dim3(256), smem, S(stream), y, dout, scale, shift, save_mean, save_invstd, part_g, part_gx, opix, c4, h, w));
    }
    DPB_CUDA(launch_k(bn_bwd_finalize_kernel, dim3((c + kFinCh - 1) / kFinCh), dim3(kFinCh, kFinRows), 0, S(stream), part_g, part_gx, blocks, c, count, gamma, save_mean, save_invstd, dgamma, dbeta, conv_dbias, coef));
  }
  const int ablocks = grid_for(units, 256, 148 * 16);
  if (pool) {
    DPB_CUDA(launch_k(bn_bwd_apply_kernel<true>, dim3(ablocks), dim3(256), 0, S(stream), y, dout, scale, shift, coef, dy, units, c, h, w));
  } else {
    DPB_CUDA(launch_k(bn_bwd_apply_kernel<false>, dim3(ablocks), dim3(256), 0, S(stream), y, dout, scale, shift, coef, dy, units, c, h, w));
  }
  return 0;
}

// ------------------------------------------------------------------------------------------------ lone layers
// The reference's layer types outside the conv->bn->relu[->pool] pattern (standalone_ops.cuh).
int dpb_relu_fwd(const float* x, float* y, long long n, dpb_stream_t stream) {
  if (n % 4) return set_error(-40, "relu: element count must be a multiple of 4 (got %lld)", n);
  DPB_CUDA(launch_k(relu_fwd_kernel, dim3(grid_for(n / 4, 256, 148 * 16)), dim3(256), 0, S(stream), (const float4*)x, (float4*)y, n / 4));
  return 0;
}

int dpb_relu_bwd(const float* y, const float* dout, float* dx, long long n, dpb_stream_t stream) {
  if (n % 4) return set_error(-40, "relu: element count must be a multiple of 4 (got %lld)", n);
  DPB_CUDA(launch_k(relu_bwd_kernel, dim3(grid_for(n / 4, 256, 148 * 16)), dim3(256), 0, S(stream), (const float4*)y, (const float4*)dout, (float4*)dx, n / 4));
  return 0;
}

int dpb_maxpool2x2_fwd(const float* x, float* out, int n, int h, int w, int c, dpb_stream_t stream) {
  if (c % 4 || ((h | w) & 1)) return set_error(-41, "maxpool2x2: needs C %% 4 == 0 and even H, W (got C=%d, %dx%d)", c, h, w);
  const long long units = (long long)n * (h / 2) * (w / 2) * (c / 4);
  DPB_CUDA(launch_k(maxpool2x2_fwd_kernel, dim3(grid_for(units, 256, 148 * 16)), dim3(256), 0, S(stream), (const float4*)x, (float4*)out, units, c / 4, h, w));
  return 0;
}

int dpb_maxpool2x2_bwd(const float* x, const float* dout, float* dx, int n, int h, int w, int c, dpb_stream_t stream) {
  if (c % 4 || ((h | w) & 1)) return set_error(-41, "maxpool2x2: needs C %% 4 == 0 and even H, W (got C=%d, %dx%d)", c, h, w);
  const long long units = (long long)n * (h / 2) * (w / 2) * (c / 4);
  DPB_CUDA(launch_k(maxpool2x2_bwd_kernel, dim3(grid_for(units, 256, 148 * 16)), dim3(256), 0, S(stream), (const float4*)x, (const float4*)dout, (float4*)dx, units, c / 4, h, w));
  return 0;
}

int dpb_column_stats_blocks(void) { return kColStatBlocks; }

static int column_stat_grid(long long m, int c) {
  const int pl = 256 / (c / 4);
  long long b = (m + pl - 1) / pl;
  return (int)(b < kColStatBlocks ? (b < 1 ? 1 : b) : kColStatBlocks);
}

int dpb_column_stats(const float* x, float* part_sum, float* part_sq, long long m, int c, int* blocks_out,
                     dpb_stream_t stream) {
  if (c % 4 || c > 1024 || c <= 0) return set_error(-42, "column stats: C must be a multiple of 4 and <= 1024 (got %d)", c);
  const int blocks = column_stat_grid(m, c);
  DPB_CUDA(launch_k(column_stats_kernel, dim3(blocks), dim3(256), 0, S(stream), (const float4*)x, part_sum, part_sq, m, c / 4));
  if (blocks_out) *blocks_out = blocks;
  return 0;
}

int dpb_column_sum(const float* x, float* part, float* out, long long m, int c, dpb_stream_t stream) {
  int blocks = 0;
  if (int rc = dpb_column_stats(x, part, nullptr, m, c, &blocks, stream)) return rc;
  DPB_CUDA(launch_k(column_reduce_kernel, dim3((c + 31) / 32), dim3(32, 8), 0, S(stream), (const float*)part, blocks, c, out));
  return 0;
}

int dpb_bn_apply(const float* x, const float* scale, const float* shift, float* y, long long m, int c,
                 dpb_stream_t stream) {
  if (c % 4) return set_error(-43, "bn apply: C must be a multiple of 4 (got %d)", c);
  const long long n4 = m * (c / 4);
  DPB_CUDA(launch_k(bn_apply_kernel, dim3(grid_for(n4, 256, 148 * 16)), dim3(256), 0, S(stream), (const float4*)x, scale, shift, (float4*)y, n4, c / 4));
  return 0;
}

int dpb_bn_plain_bwd(const float* x, const float* dout, const float* save_mean, const float* save_invstd,
                     const float* gamma, float* part, float* coef, float* dgamma, float* dbeta, float* dx, long long m,
                     int c, dpb_stream_t stream) {
  if (c % 4 || c > 1024 || c <= 0) return set_error(-43, "bn backward: C must be a multiple of 4 and <= 1024 (got %d)", c);
  const int blocks = column_stat_grid(m, c);
  float* part_g = part;
  float* part_gx = part + (size_t)kColStatBlocks * c;
  DPB_CUDA(launch_k(bn_plain_bwd_reduce_kernel, dim3(blocks), dim3(256), 0, S(stream), (const float4*)x, (const float4*)dout, save_mean, save_invstd, part_g, part_gx, m, c / 4));
  DPB_CUDA(launch_k(bn_bwd_finalize_kernel, dim3((c + kFinCh - 1) / kFinCh), dim3(kFinCh, kFinRows), 0, S(stream), (const float*)part_g, (const float*)part_gx, blocks, c, (double)m, gamma, save_mean, save_invstd, dgamma, dbeta, (float*)nullptr, coef));
  const long long n4 = m * (c / 4);
  DPB_CUDA(launch_k(bn_plain_bwd_apply_kernel, dim3(grid_for(n4, 256, 148 * 16)), dim3(256), 0, S(stream), (const float4*)x, (const float4*)dout, (const float*)coef, (float4*)dx, n4, c));
  return 0;
}

// ------------------------------------------------------------------------------------------------ head / loss
static bool head_smem_ok(int k, int o) { return (k % 4 == 0) && (size_t)o * k * 4 <= 48 * 1024; }

int dpb_linear_fwd(const float* x, const float* w, const float* bias, float* out, int b, int k, int o,
                   dpb_stream_t stream) {
  if (o > kMaxClasses) return set_error(-3, "linear: out_features %d > %d", o, kMaxClasses);
  if (head_smem_ok(k, o)) {
    DPB_CUDA(launch_k(head_fwd_kernel<false>, dim3((b + 7) / 8), dim3(256), (size_t)o * k * 4, S(stream), x, w, bias, (const long long*)nullptr, out, (float*)nullptr, (float2*)nullptr, (float*)nullptr, (float*)nullptr, (unsigned int*)nullptr, b, k, o));
    return 0;
  }
  DPB_CUDA(launch_k(linear_fwd_kernel, dim3((b * 32 + 127) / 128), dim3(128), 0, S(stream), x, w, bias, out, b, k, o));
  return 0;
}

int dpb_head_supported(int k, int o) { return (o <= kMaxClasses && head_smem_ok(k, o)) ? 1 : 0; }

int dpb_head_fwd(const float* x, const float* w, const float* bias, const long long* target, float* logits, float* dz,
                 float* terms, float* loss, float* metrics, void* counter, int b, int k, int o, dpb_stream_t stream) {
  if (!dpb_head_supported(k, o)) return set_error(-3, "head: unsupported shape K=%d O=%d", k, o);
  DPB_CUDA(launch_k(head_fwd_kernel<true>, dim3((b + 7) / 8), dim3(256), (size_t)o * k * 4, S(stream), x, w, bias, target, logits, dz, reinterpret_cast<float2*>(terms), loss, metrics, reinterpret_cast<unsigned int*>(counter), b, k, o));
  return 0;
}

int dpb_head_bwd(const float* dz, const float* x, const float* w, const float* grad_out, float* dx, float* dw,
                 float* db, int b, int k, int o, dpb_stream_t stream) {
  if (o > kMaxClasses) return set_error(-3, "head: out_features %d > %d", o, kMaxClasses);
  const int nx = (int)(((long long)b * k + 255) / 256), nw = (k + 31) / 32;
  DPB_CUDA(launch_k(head_bwd_kernel, dim3(nx + nw + 1), dim3(256), 0, S(stream), dz, x, w, grad_out, dx, dw, db, b, k, o, nx, nw));
  return 0;
}

int dpb_augment_batch(const unsigned char* images, const long long* labels, const int* indices, int b, int h, int w,
                      int train, unsigned int seed, unsigned int epoch, const int* draws, float mean0, float mean1,
                      float mean2, float inv_std0, float inv_std1, float inv_std2, float* out_x, long long* out_y,
                      dpb_stream_t stream) {
  if (b <= 0) return 0;
  if (w % 4) return set_error(-8, "augment: image width must be a multiple of 4 (got %d)", w);
  AugmentNorm nrm{{mean0, mean1, mean2}, {inv_std0, inv_std1, inv_std2}};
  augment_batch_kernel<<<b, 256, 0, S(stream)>>>(images, labels, indices, h, w, train, seed, epoch, draws, nrm, out_x, out_y);
  DPB_CUDA(cudaGetLastError());
  return 0;
}

int dpb_linear_bwd(const float* dz, const float* x, const float* w, float* dx, float* dw, float* db, int b, int k,
                   int o, dpb_stream_t stream) {
  if (o > kMaxClasses) return set_error(-3, "linear: out_features %d > %d", o, kMaxClasses);
  if (dx != nullptr) {
    const long long n = (long long)b * k;
    DPB_CUDA(launch_k(linear_bwd_input_kernel, dim3((int)((n + 255) / 256)), dim3(256), 0, S(stream), dz, w, dx, b, k, o));
  }
  if (dw != nullptr) {
    DPB_CUDA(launch_k(linear_bwd_weight_kernel, dim3((k + 31) / 32 + (db != nullptr ? 1 : 0)), dim3(32, 8), 0, S(stream), dz, x, dw, db, b, k, o));
  }
  return 0;
}

int dpb_cross_entropy_fwd(const float* logits, const long long* target, float* loss, float* metrics, int b, int o,
                          dpb_stream_t stream) {
  DPB_CUDA(launch_k(cross_entropy_fwd_kernel, dim3(1), dim3(256), 0, S(stream), logits, target, loss, metrics, b, o));
  return 0;
}

int dpb_cross_entropy_bwd(const float* logits, const long long* target, const float* grad_out, float* dz, int b,
                          int o, dpb_stream_t stream) {
  DPB_CUDA(launch_k(cross_entropy_bwd_kernel, dim3((b + 255) / 256), dim3(256), 0, S(stream), logits, target, grad_out, dz, b, o));
  return 0;
}

// ------------------------------------------------------------------------------------------------ optimizer
int dpb_sgd_step(float* p, const float* g, float* m, long long n, float lr, float momentum, float wd,
                 float grad_scale, int nesterov, dpb_stream_t stream) {
  if (n <= 0) return 0;
  const bool aligned = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                         reinterpret_cast<uintptr_t>(m)) & 15) == 0;
  long long n4 = aligned ? n / 4 : 0;
  if (n4 > 0) {
    DPB_CUDA(launch_k(sgd_kernel, dim3(grid_for(n4, 256, 148 * 8)), dim3(256), 0, S(stream), p, g, m, n4, lr, momentum, wd, grad_scale, nesterov));
  }
  const long long rem = n - n4 * 4;
  if (rem > 0) {
    const long long o = n4 * 4;
    DPB_CUDA(launch_k(sgd_scalar_kernel, dim3((int)((rem + 255) / 256)), dim3(256), 0, S(stream), p + o, g + o, m ? m + o : nullptr, rem, lr, momentum, wd, grad_scale, nesterov));
  }
  return 0;
}

// ------------------------------------------------------------------------------------------------ symmetric memory
int dpb_symm_alloc(size_t bytes, void** ptr) {
  DPB_CUDA(cudaMalloc(ptr, bytes));
  DPB_CUDA(cudaMemset(*ptr, 0, bytes));
  DPB_CUDA(cudaDeviceSynchronize());
  return 0;
}
int dpb_symm_free(void* ptr) {
  DPB_CUDA(cudaFree(ptr));
  return 0;
}
int dpb_symm_export(void* ptr, unsigned char handle64[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  DPB_CUDA(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle64, &h, 64);
  return 0;
}
int dpb_symm_import(const unsigned char handle64[64], void** peer_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DPB_CUDA(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int dpb_symm_close(void* peer_ptr) {
  DPB_CUDA(cudaIpcCloseMemHandle(peer_ptr));
  return 0;
}
size_t dpb_signal_pad_bytes(void) { return kSigPadBytes; }

// ---- the same through the virtual-memory-management API (csrc/vmm.cuh), which is what multicast needs
int dpb_vmm_probe(int ndev, long long* out3) { return vmm_probe(ndev, out3); }
int dpb_vmm_alloc(size_t bytes, size_t gran, void** ptr, unsigned long long* handle, int* fd) {
  return vmm_alloc(bytes, gran, ptr, handle, fd);
}
int dpb_vmm_import(int fd, size_t bytes, size_t gran, int map, void** ptr, unsigned long long* handle) {
  return vmm_import(fd, bytes, gran, map, ptr, handle);
}
int dpb_vmm_unmap(void* ptr, size_t bytes, unsigned long long handle) { return vmm_unmap(ptr, bytes, handle); }
int dpb_mc_create(int ndev, size_t bytes, unsigned long long* handle, int* fd) { return mc_create(ndev, bytes, handle, fd); }
int dpb_mc_add_device(unsigned long long mc) { return mc_add_device(mc); }
int dpb_mc_bind_map(unsigned long long mc, unsigned long long mem, size_t bytes, size_t gran, void** mc_ptr) {
  return mc_bind_map(mc, mem, bytes, gran, mc_ptr);
}
int dpb_mc_unbind(unsigned long long mc, size_t bytes) { return mc_unbind(mc, bytes); }

}  // extern "C"

// ------------------------------------------------------------------------------------------------ communicator
struct dpb_comm {
  CommCtx ctx;
  int emulate;
};

namespace {

PeerPtrs make_ptrs(void* const* bufs, int n) {
  PeerPtrs p{};
  for (int i = 0; i < n && i < kMaxRanks; ++i) p.p[i] = bufs ? reinterpret_cast<float*>(bufs[i]) : nullptr;
  return p;
}

// Largest grid.x the launch may use: real ranks are bounded by the signal-pad geometry, emulated ranks additionally by
// co-residency of the cooperative launch (all virtual ranks' CTAs must fit on the one GPU at once).
template <typename K>
int comm_block_cap(dpb_comm* c, K kern_emu, int threads) {
  if (!c->emulate) return kCommMaxBlocks;
  int per_sm = 0, dev = 0, sms = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern_emu, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int cap = per_sm * sms / c->ctx.nranks;
  if (cap > kCommMaxBlocks) cap = kCommMaxBlocks;
  return cap < 1 ? 1 : cap;
}

template <typename K, typename... Args>
int launch_comm(dpb_comm* c, K kern_real, K kern_emu, int blocks, int threads, cudaStream_t st, Args... args) {
  const int cap = comm_block_cap(c, kern_emu, threads);
  if (blocks < 1) blocks = 1;
  if (blocks > cap) blocks = cap;
  if (!c->emulate) {
    kern_real<<<blocks, threads, 0, st>>>(args...);
    DPB_CUDA(cudaGetLastError());
    return 0;
  }
  void* params[] = {(void*)&args...};
  dim3 grid(blocks, c->ctx.nranks);
  DPB_CUDA(cudaLaunchCooperativeKernel((const void*)kern_emu, grid, dim3(threads), params, 0, st));
  return 0;
}

int comm_blocks_for(long long n4, int threads, int u, int want) {
  if (want > 0) return want > kCommMaxBlocks ? kCommMaxBlocks : want;
  long long b = (n4 + (long long)threads * u - 1) / ((long long)threads * u);
  if (b > g_tune_comm_blocks) b = g_tune_comm_blocks;
  if (b < 1) b = 1;
  return (int)b;
}

int check_range(const char* what, long long offset, long long count) {
  if ((offset & 3) || (count & 3) || count < 0)
    return set_error(-4, "%s: offset/count must be multiples of 4 floats (got %lld, %lld)", what, offset, count);
  return 0;
}

}  // namespace

extern "C" {

int dpb_comm_create(int rank, int nranks, void* const* signal_pads, int emulate, dpb_comm** out) {
  if (nranks < 1 || nranks > kMaxRanks) return set_error(-5, "comm: world size %d outside [1,%d]", nranks, kMaxRanks);
  if (rank < 0 || rank >= nranks) return set_error(-5, "comm: bad rank %d", rank);
  dpb_comm* c = new (std::nothrow) dpb_comm();
  if (!c) return set_error(-5, "comm: out of host memory");
  c->ctx.rank = rank;
  c->ctx.nranks = nranks;
  c->ctx.withhold_rank = -1;
  if (int rc = init_current_device()) { delete c; return rc; }
  for (int i = 0; i < kMaxRanks; ++i) c->ctx.sig[i] = i < nranks ? reinterpret_cast<uint32_t*>(signal_pads[i]) : nullptr;
  c->emulate = emulate;
  *out = c;
  return 0;
}
int dpb_comm_destroy(dpb_comm* c) {
  delete c;
  return 0;
}
int dpb_comm_debug_withhold(dpb_comm* c, int rank) {
  if (!c->emulate) return set_error(-5, "comm: withholding a rank is a test hook of emulated worlds");
  c->ctx.withhold_rank = rank;
  return 0;
}

int dpb_allreduce_sgd(dpb_comm* c, int channel, void* const* grads, void* const* params, void* const* momentum,
                      long long offset, long long count, float scale, float lr, float mom, float wd, int nesterov,
                      int write_mean, int blocks, dpb_stream_t stream) {
  int rc = check_range("allreduce", offset, count);
  if (rc) return rc;
  if (count == 0) return 0;
  const int nr = c->ctx.nranks;
  const long long n4 = count / 4, off4 = offset / 4;
  SgdArgs sg{};
  sg.enabled = params != nullptr;
  sg.lr = lr; sg.momentum = mom; sg.wd = wd; sg.nesterov = nesterov;
  PeerPtrs d = make_ptrs(grads, nr), pp = make_ptrs(params, nr), mm = make_ptrs(momentum, nr);
  const long long slice = (n4 + nr - 1) / nr;
  cudaStream_t st = S(stream);
#define DPB_TWOSHOT(NRT, U)                                                                                      \
  return launch_comm(c, allreduce_twoshot_kernel<false, NRT, U>, allreduce_twoshot_kernel<true, NRT, U>,         \
                     comm_blocks_for(slice, 512, U, blocks), 512, st, c->ctx, channel, d, off4, n4, scale,        \
                     (write_mean ? 1 : 0) | (g_tune_comm_weak ? 2 : 0), sg, pp, mm)
  switch (nr) {
    case 1: DPB_TWOSHOT(1, 4);
    case 2: DPB_TWOSHOT(2, 4);
    case 4: DPB_TWOSHOT(4, 4);
    case 8: DPB_TWOSHOT(8, 2);
    default: DPB_TWOSHOT(0, 2);
  }
#undef DPB_TWOSHOT
}

int dpb_allreduce_nvls(dpb_comm* c, int channel, void* mc_grads, void* local_params, void* mc_params, void* local_mom,
                       long long offset, long long count, float scale, float lr, float mom, float wd, int nesterov,
                       int write_mean, int blocks, dpb_stream_t stream) {
  int rc = check_range("allreduce_nvls", offset, count);
  if (rc) return rc;
  if (count == 0) return 0;
  if (c->emulate) return set_error(-4, "allreduce_nvls: multicast needs one GPU per rank (no single-GPU emulation)");
  if (mc_grads == nullptr) return set_error(-4, "allreduce_nvls: the gradient buffer has no multicast mapping");
  const int nr = c->ctx.nranks;
  const long long n4 = count / 4, off4 = offset / 4;
  SgdArgs sg{};
  sg.enabled = local_params != nullptr;
  sg.lr = lr; sg.momentum = mom; sg.wd = wd; sg.nesterov = nesterov;
  if (sg.enabled && (mc_params == nullptr || local_mom == nullptr))
    return set_error(-4, "allreduce_nvls: the fused optimizer needs multicast parameters and a momentum buffer");
  const long long slice = (n4 + nr - 1) / nr;
  long long b = blocks > 0 ? blocks : (slice + 512LL * 4 - 1) / (512LL * 4);
  if (blocks <= 0 && b > g_tune_nvls_blocks) b = g_tune_nvls_blocks;
  if (b > kCommMaxBlocks) b = kCommMaxBlocks;
  if (b < 1) b = 1;
  // loads in flight per thread: the fewer CTAs the launch may use (a bucket reduced next to backward), the more each
  // thread has to keep in flight to fill the wire
  const long long per_thread = (slice + b * 512 - 1) / (b * 512);
#define DPB_NVLS(U)                                                                                                  \
  allreduce_nvls_kernel<U><<<(unsigned)b, 512, 0, S(stream)>>>(c->ctx, channel, (float*)mc_grads, off4, n4, scale,   \
                                                               write_mean ? 1 : 0, sg, (float*)local_params,           \
                                                               (float*)mc_params, (float*)local_mom)
  if (per_thread >= 16 && b <= 48) DPB_NVLS(16);
  else if (per_thread >= 8) DPB_NVLS(8);
  else DPB_NVLS(4);
#undef DPB_NVLS
  DPB_CUDA(cudaGetLastError());
  return 0;
}
int dpb_nvls_enabled(void) { return g_tune_nvls; }

int dpb_allreduce(dpb_comm* c, int channel, void* const* bufs, long long offset, long long count, float scale,
                  int algo, int blocks, dpb_stream_t stream) {
  int rc = check_range("allreduce", offset, count);
  if (rc) return rc;
  if (count == 0) return 0;
  const long long n4 = count / 4;
  constexpr int U = 4;
  // one-shot (every rank pulls everything) only for the tiniest messages: measured at 8 GPUs it costs 41-62 us from
  // 16 KiB to 256 KiB while the two-shot kernel moves 1 MiB in 29 us (profiles/sync_sweep_8.md)
  if (algo == 0) algo = (n4 * 16 <= (8 << 10)) ? 1 : 2;
  if (algo == 1) {
    int b = (int)((n4 + 512LL * U - 1) / (512LL * U));
    if (b <= comm_block_cap(c, allreduce_oneshot_kernel<true, U>, 512)) {  // the whole range must fit one grid pass
      PeerPtrs d = make_ptrs(bufs, c->ctx.nranks);
      return launch_comm(c, allreduce_oneshot_kernel<false, U>, allreduce_oneshot_kernel<true, U>, b, 512, S(stream),
                         c->ctx, channel, d, offset / 4, n4, scale);
    }
  }
  return dpb_allreduce_sgd(c, channel, bufs, nullptr, nullptr, offset, count, scale, 0.f, 0.f, 0.f, 0, 1, blocks,
                           stream);
}

int dpb_gather_mean_scatter(dpb_comm* c, int channel, void* const* bufs, long long offset, long long count,
                            float scale, int root, int blocks, dpb_stream_t stream) {
  int rc = check_range("gather_mean_scatter", offset, count);
  if (rc) return rc;
  if (count == 0) return 0;
  const long long n4 = count / 4;
  constexpr int U = 2;
  PeerPtrs d = make_ptrs(bufs, c->ctx.nranks);
  int b = comm_blocks_for(n4, 512, U, blocks);
  if (blocks <= 0 && n4 > 64LL * 512 * U) b = kCommMaxBlocks;  // the root needs every SM it can get
  return launch_comm(c, gather_mean_scatter_kernel<false, U>, gather_mean_scatter_kernel<true, U>, b, 512, S(stream),
                     c->ctx, channel, d, offset / 4, n4, scale, root);
}

int dpb_broadcast(dpb_comm* c, int channel, void* const* bufs, long long offset16, long long count16, int root,
                  int blocks, dpb_stream_t stream) {
  if (count16 <= 0) return 0;
  PeerPtrs d = make_ptrs(bufs, c->ctx.nranks);
  return launch_comm(c, broadcast_kernel<false>, broadcast_kernel<true>, comm_blocks_for(count16, 512, 4, blocks), 512,
                     S(stream), c->ctx, channel, d, offset16, count16, root);
}

int dpb_barrier(dpb_comm* c, int channel, dpb_stream_t stream) {
  return launch_comm(c, barrier_kernel<false>, barrier_kernel<true>, 1, 32, S(stream), c->ctx, channel);
}

// ------------------------------------------------------------------------------------------------ DDP reducer
int dpb_plan_buckets(int nparams, const long long* bytes, long long first_cap, long long cap, int* bucket_of_param) {
  int bucket = 0;
  long long fill = 0;
  long long limit = first_cap;
  for (int i = 0; i < nparams; ++i) {
    bucket_of_param[i] = bucket;
    fill += bytes[i];
    if (fill >= limit && i + 1 < nparams) {  // the overflowing tensor stays; the bucket closes behind it
      ++bucket;
      fill = 0;
      limit = cap;
    }
  }
  return nparams > 0 ? bucket + 1 : 0;
}

}  // extern "C"

struct dpb_reducer {
  dpb_comm* comm;
  int nparams, nbuckets;
  std::vector<int> bucket_of_param;
  std::vector<long long> b_lo, b_hi;  // float offsets of each bucket's arena range
  std::vector<int> b_total, b_pending;
  std::vector<void*> bufs;
  float scale;
  // optimizer fused into every bucket's collective (dpb_reducer_set_sgd): the slice owner updates parameters/momentum
  bool sgd = false;
  std::vector<void*> params, moms;
  float lr = 0.f, mom = 0.f, wd = 0.f;
  int nesterov = 0;
  // multicast views (dpb_reducer_set_multicast): buckets are reduced inside the NVSwitch
  void* mc_grads = nullptr;
  void* mc_params = nullptr;
  void* local_params = nullptr;
  void* local_mom = nullptr;
  cudaStream_t side;
  cudaEvent_t ev_ready, ev_done;
  int launched;
};

extern "C" {

int dpb_reducer_create(dpb_comm* c, int nparams, const long long* offsets, const long long* counts,
                       const int* bucket_of_param, int nbuckets, void* const* grad_bufs, float scale,
                       dpb_reducer** out) {
  dpb_reducer* r = new (std::nothrow) dpb_reducer();
  if (!r) return set_error(-6, "reducer: out of host memory");
  r->comm = c;
  r->nparams = nparams;
  r->nbuckets = nbuckets;
  r->bucket_of_param.assign(bucket_of_param, bucket_of_param + nparams);
  r->b_lo.assign(nbuckets, -1);
  r->b_hi.assign(nbuckets, -1);
  r->b_total.assign(nbuckets, 0);
  r->b_pending.assign(nbuckets, 0);
  for (int i = 0; i < nparams; ++i) {
    const int b = bucket_of_param[i];
    if (b < 0 || b >= nbuckets) { delete r; return set_error(-6, "reducer: bad bucket id %d", b); }
    if ((offsets[i] & 3) || (counts[i] & 3)) { delete r; return set_error(-6, "reducer: param %d not 16-byte granular", i); }
    if (r->b_lo[b] < 0 || offsets[i] < r->b_lo[b]) r->b_lo[b] = offsets[i];
    if (offsets[i] + counts[i] > r->b_hi[b]) r->b_hi[b] = offsets[i] + counts[i];
    r->b_total[b]++;
  }
  // contiguity: bucket ranges must not overlap
  for (int a = 0; a < nbuckets; ++a)
    for (int b = a + 1; b < nbuckets; ++b)
      if (r->b_lo[a] < r->b_hi[b] && r->b_lo[b] < r->b_hi[a]) {
        delete r;
        return set_error(-6, "reducer: buckets %d and %d overlap in the arena", a, b);
      }
  r->bufs.assign(grad_bufs, grad_bufs + c->ctx.nranks);
  r->scale = scale;
  r->launched = 0;
  int prio_lo = 0, prio_hi = 0;
  cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);  // the collective's few CTAs go first when SM slots free up
  if (cudaStreamCreateWithPriority(&r->side, cudaStreamNonBlocking, prio_hi) != cudaSuccess ||
      cudaEventCreateWithFlags(&r->ev_ready, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&r->ev_done, cudaEventDisableTiming) != cudaSuccess) {
    delete r;
    return set_error(-6, "reducer: stream/event creation failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  *out = r;
  return 0;
}

int dpb_reducer_destroy(dpb_reducer* r) {
  if (!r) return 0;
  cudaStreamDestroy(r->side);
  cudaEventDestroy(r->ev_ready);
  cudaEventDestroy(r->ev_done);
  delete r;
  return 0;
}

int dpb_reducer_set_sgd(dpb_reducer* r, void* const* param_bufs, void* const* mom_bufs, float lr, float momentum,
                        float wd, int nesterov) {
  const int nr = r->comm->ctx.nranks;
  r->sgd = param_bufs != nullptr;
  if (r->sgd) {
    r->params.assign(param_bufs, param_bufs + nr);
    r->moms.assign(mom_bufs, mom_bufs + nr);
  }
  r->lr = lr; r->mom = momentum; r->wd = wd; r->nesterov = nesterov;
  return 0;
}

int dpb_reducer_set_multicast(dpb_reducer* r, void* mc_grads, void* mc_params, void* local_params, void* local_mom) {
  r->mc_grads = mc_grads; r->mc_params = mc_params; r->local_params = local_params; r->local_mom = local_mom;
  return 0;
}

int dpb_reducer_side_stream(dpb_reducer* r, dpb_stream_t* out) {
  *out = reinterpret_cast<dpb_stream_t>(r->side);
  return 0;
}

int dpb_reducer_prepare(dpb_reducer* r) {
  r->b_pending = r->b_total;
  r->launched = 0;
  return 0;
}

int dpb_reducer_mark_ready(dpb_reducer* r, int param, dpb_stream_t main_stream) {
  if (param < 0 || param >= r->nparams) return set_error(-6, "reducer: bad param index %d", param);
  const int b = r->bucket_of_param[param];
  if (r->b_pending[b] <= 0) return set_error(-6, "reducer: param %d marked ready twice (bucket %d)", param, b);
  if (--r->b_pending[b] > 0) return 0;
  // bucket complete: side stream picks up after everything enqueued on the main stream so far
  DPB_CUDA(cudaEventRecord(r->ev_ready, S(main_stream)));
  DPB_CUDA(cudaStreamWaitEvent(r->side, r->ev_ready, 0));
  const long long lo = r->b_lo[b], cnt = r->b_hi[b] - r->b_lo[b];
  int rc;
  if (r->mc_grads != nullptr && g_tune_nvls && !r->comm->emulate && (!r->sgd || r->mc_params != nullptr)) {
    rc = dpb_allreduce_nvls(r->comm, /*channel=*/1, r->mc_grads, r->sgd ? r->local_params : nullptr, r->mc_params,
                            r->local_mom, lo, cnt, r->scale, r->lr, r->mom, r->wd, r->nesterov, /*write_mean=*/0,
                            g_tune_ddp_blocks, r->side);
  } else if (r->sgd) {
    rc = dpb_allreduce_sgd(r->comm, /*channel=*/1, r->bufs.data(), r->params.data(), r->moms.data(), lo, cnt, r->scale,
                           r->lr, r->mom, r->wd, r->nesterov, /*write_mean=*/0, g_tune_ddp_blocks, r->side);
  } else {
    rc = dpb_allreduce(r->comm, /*channel=*/1, r->bufs.data(), lo, cnt, r->scale, /*algo=*/0, g_tune_ddp_blocks, r->side);
  }
  if (rc) return rc;
  r->launched++;
  return 0;
}

int dpb_reducer_wait_stream(dpb_reducer* r, dpb_stream_t producer) {
  // gradients of the next buckets may come from a second compute stream (weight gradients overlapped with the
  // data-gradient chain): every reduction launched from now on also waits for what that stream holds at this point
  DPB_CUDA(cudaEventRecord(r->ev_ready, S(producer)));
  DPB_CUDA(cudaStreamWaitEvent(r->side, r->ev_ready, 0));
  return 0;
}

int dpb_reducer_finalize(dpb_reducer* r, dpb_stream_t main_stream) {
  for (int b = 0; b < r->nbuckets; ++b)
    if (r->b_pending[b] != 0 && r->b_pending[b] != r->b_total[b])
      return set_error(-6, "reducer: bucket %d incomplete at finalize (%d of %d pending)", b, r->b_pending[b],
                       r->b_total[b]);
  if (r->launched > 0) {
    DPB_CUDA(cudaEventRecord(r->ev_done, r->side));
    DPB_CUDA(cudaStreamWaitEvent(S(main_stream), r->ev_done, 0));
  }
  return 0;
}

int dpb_reducer_launched(dpb_reducer* r) { return r->launched; }

}  // extern "C"

// ------------------------------------------------------------------------------------------------ allocation-free variant
// dpb_bn_relu_pool_bwd() needs two pieces of scratch from its caller: the per-block partial sums of the reduce pass
// ([2][kBnBwdBlocks][c] floats) and the per-channel coefficients ([3][c] floats).  Some callers have no way to provide
// device memory of their own: a host program that links only this C ABI (no CUDA runtime calls of its own, no
// PyTorch allocator), or a component that is handed inputs and outputs by a framework and is forbidden to allocate
// on its hot path.  For them the scratch below is part of the library image: it exists on every GPU that loads
// libdpb200.so from the moment the module is loaded until the process exits, and costs kNoAllocScratchFloats * 4 bytes
// (about 1.2 MB) there whether or not the entry point is ever called.
constexpr int kNoAllocMaxChannels = 512;  // widest BatchNorm of the VGG family
constexpr size_t kNoAllocScratchFloats = (size_t)(2 * kBnBwdBlocks + 3) * kNoAllocMaxChannels;
static __device__ __align__(256) float g_noalloc_bn_scratch[kNoAllocScratchFloats];

namespace {
constexpr int kNoAllocMaxDevices = 64;
float* g_noalloc_scratch_addr[kNoAllocMaxDevices] = {nullptr};  // device address of the array, per device ordinal
std::mutex g_noalloc_mutex;

// `g_noalloc_bn_scratch` names a host-side shadow in host code, not device memory, so it must never be passed to a
// kernel as is.  The device address is looked up once per device, with that device current.
int noalloc_scratch_for_current_device(float** out) {
  int dev = -1;
  DPB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) return set_error(-20, "no-alloc scratch: device ordinal %d unsupported", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_scratch_addr[dev] == nullptr) {
    float* p = nullptr;
    DPB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_bn_scratch));
    if (p == nullptr) return set_error(-20, "no-alloc scratch: cudaGetSymbolAddress returned a null address");
    g_noalloc_scratch_addr[dev] = p;
  }
  *out = g_noalloc_scratch_addr[dev];
  return 0;
}
}  // namespace

extern "C" {

int dpb_noalloc_max_channels(void) { return kNoAllocMaxChannels; }

int dpb_bn_relu_pool_bwd_noalloc(const float* y, const float* dout, const float* scale, const float* shift,
                                 const float* save_mean, const float* save_invstd, const float* gamma, float* dgamma,
                                 float* dbeta, float* conv_dbias, float* dy, int n, int h, int w, int c, int pool,
                                 dpb_stream_t stream) {
  if (c > kNoAllocMaxChannels)
    return set_error(-21,
                     "dpb_bn_relu_pool_bwd_noalloc: %d channels exceed the built-in scratch (maximum %d); use "
                     "dpb_bn_relu_pool_bwd with caller-provided scratch instead",
                     c, kNoAllocMaxChannels);
  if (c <= 0 || n <= 0 || h <= 0 || w <= 0) return set_error(-21, "dpb_bn_relu_pool_bwd_noalloc: empty problem");
  float* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(&scratch);
  if (rc) return rc;
  // Same layout the allocating entry point expects from its caller; every word the kernels read from it is written
  // earlier in this same call (partials by the reduce pass, coefficients by the finalize pass), so nothing has to be
  // cleared even though the array still holds the previous call's contents.
  float* part = scratch;
  float* coef = scratch + (size_t)2 * kBnBwdBlocks * kNoAllocMaxChannels;
  // (no accumulators: this variant keeps the separate reduce / finalize / apply launches and their scratch)
  return dpb_bn_relu_pool_bwd(y, dout, scale, shift, save_mean, save_invstd, gamma, part, coef, nullptr, nullptr, dgamma,
                              dbeta, conv_dbias, dy, n, h, w, c, pool, stream);
}

}  // extern "C"
