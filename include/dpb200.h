/* distributed_pytorch_b200 native library: C ABI.
 *
 * libdpb200.so is plain CUDA C++ built for sm_100a (no PyTorch dependency); Python binds to it with ctypes and passes
 * raw device pointers plus the current CUDA stream.  Every function returns 0 on success; on failure the message is
 * available from dpb_last_error().  "ref:" comments name the reference code each entry point stands in for
 * (paths relative to /root/reference).
 *
 * Layout conventions: activations NHWC fp32 ([N*H*W][C]); conv weights KRSC fp32 ([Cout][3][3][Cin], i.e. a
 * channels_last view of PyTorch's OIHW tensor); all pointers 16-byte aligned.
 */
#ifndef DPB200_H_
#define DPB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* dpb_stream_t; /* cudaStream_t */

/* ---- runtime ------------------------------------------------------------------------------------------------- */
int dpb_version(void);
const char* dpb_last_error(void);
/* Diagnostic left behind by a kernel that gave up (pipeline or peer-signal timeout). out[5] = code, where, block,
 * aux0, aux1.  Clears the record. */
int dpb_last_device_error(int* out5);
/* Bring-up aid: when buf != NULL, CTA (0,0,0) of every tensor-core conv launch copies its first pipeline stage
 * (A tile then B tile, as laid out in shared memory; up to 48 KiB) to buf. NULL switches it off. */
int dpb_debug_set_conv_dump(void* buf);
/* Scratch needed by the tensor-core convolution (split-K partial tiles + arrival counters). */
size_t dpb_conv_workspace_bytes(void);
size_t dpb_conv_counter_bytes(void);

/* ---- convolution (ref: model.py:18-23 -> aten convolution / convolution_backward) ------------------------------ */
/* mode 0 fprop: in = X, w = W, out = Y (+bias, + per-M-tile BatchNorm partial sums if stat_sum != NULL)
 * mode 1 dgrad: in = dY, w = W, out = dX
 * mode 2 wgrad: in = dY, in2 = X, out = dW
 * nt / splits: 0 = heuristic.  m_tiles_out (optional) receives the number of rows written to stat_sum / stat_sq. */
int dpb_conv3x3_gemm(int mode, int n, int h, int w, int cin, int cout, const float* in, const float* in2,
                     const float* weight, float* out, const float* bias, float* stat_sum, float* stat_sq,
                     void* workspace, void* counters, int nt, int splits, int* m_tiles_out, dpb_stream_t stream);
/* Heuristic knobs (the Python loader also reads them from the environment variable in brackets).
 *  0 [DPB_CONV_TARGET_CTAS]  CTA count the split-K heuristic aims for (default 296 = two per SM)
 *  1 [DPB_CONV_NT]           preferred N tile (64 / 128 / 256, default 128) when it divides the GEMM N extent
 *  2 [DPB_COMM_BLOCKS]       default grid cap of the collective kernels (default 64, at most 128)
 *  3 [DPB_CONV_MIN_K_ITERS]  K iterations (of 32) a split must keep (default 8)
 *  4 [DPB_PDL]               programmatic dependent launch on every compute kernel (default 0)
 *  5 [DPB_CONV_CLUSTER]      TMA multicast of the B tile over a 1 / 2 / 4-CTA cluster (default 1 = none)
 *  6 [DPB_CONV_PAIR]         cta_group::2 CTA pairs, M = 256 MMAs (default 0)
 *  7 [DPB_FIRST_TC]          first (Cin = 3) layer on the tensor cores when the image tiles into 128 pixels (default 1)
 *  8 [DPB_PAIR_REDUCE]       finish two-way split-K inside the GEMM kernel instead of the reduce kernel (default 0)
 *  9 [DPB_TAIL_MAX_PIXELS]   BatchNorm backward as one cluster kernel for layers up to this many pixels (default 0 = off)
 * Defaults are the measured optimum on B200; returns non-zero for an unknown key / bad value. */
int dpb_set_tuning(int key, int value);
/* Kernels launched by the calling thread's most recent dpb_conv3x3_gemm (1, or 2 when split-K adds its reduce). */
int dpb_conv_last_launches(void);
/* First layer (Cin = 3, Cout = 64): x is NCHW.  part_* : [dpb_first_conv_blocks(n,h,w)][64]. */
int dpb_first_conv_blocks(int n, int h, int w);
int dpb_first_conv_fprop(const float* x_nchw, const float* weight, const float* bias, float* y, float* part_sum,
                         float* part_sq, int n, int h, int w, dpb_stream_t stream);
/* partials: [dpb_first_conv_wgrad_blocks()][1728] scratch; dw: [64][3][3][3]. */
int dpb_first_conv_wgrad_blocks(void);
int dpb_first_conv_wgrad(const float* x_nchw, const float* dy, float* partials, float* dw, int n, int h, int w,
                         dpb_stream_t stream);

/* ---- BatchNorm + ReLU (+ 2x2 max-pool) (ref: model.py:16,24,25) -------------------------------------------------- */
int dpb_bn_finalize(const float* part_sum, const float* part_sq, int parts, int c, double count, const float* gamma,
                    const float* beta, float* running_mean, float* running_var, long long* num_batches_tracked,
                    float momentum, float eps, float* scale, float* shift, float* save_mean, float* save_invstd,
                    dpb_stream_t stream);
int dpb_bn_eval_scale_shift(int c, const float* gamma, const float* beta, const float* running_mean,
                            const float* running_var, float eps, float* scale, float* shift, dpb_stream_t stream);
int dpb_bn_relu_pool_fwd(const float* y, const float* scale, const float* shift, float* out, int n, int h, int w,
                         int c, int pool, dpb_stream_t stream);
/* Backward of the fused block.  part: scratch [2][dpb_bn_bwd_blocks()][c]; coef: scratch [3][c].
 * conv_dbias (optional): gradient of the preceding conv's bias, exactly zero under training-mode BatchNorm. */
int dpb_bn_bwd_blocks(void);
/* Kernels dpb_bn_relu_pool_bwd launches for this shape: 1 when the layer is small enough for the cluster kernel (a
 * cluster of 8 CTAs per 32 channels exchanges the per-channel sums through distributed shared memory), else 3. */
int dpb_bn_bwd_launches(int n, int h, int w, int c, int pool);
int dpb_bn_relu_pool_bwd(const float* y, const float* dout, const float* scale, const float* shift,
                         const float* save_mean, const float* save_invstd, const float* gamma, float* part,
                         float* coef, float* dgamma, float* dbeta, float* conv_dbias, float* dy, int n, int h, int w,
                         int c, int pool, dpb_stream_t stream);

/* ---- The same layer types on their own (ref: model.py:16-25, any arrangement other than conv->bn->relu[->pool]) ---- */
int dpb_relu_fwd(const float* x, float* y, long long n, dpb_stream_t stream);
/* y: the forward OUTPUT (ReLU(inplace=True) keeps nothing else). */
int dpb_relu_bwd(const float* y, const float* dout, float* dx, long long n, dpb_stream_t stream);
int dpb_maxpool2x2_fwd(const float* x, float* out, int n, int h, int w, int c, dpb_stream_t stream);
/* Routes each gradient to the first maximum of its window (aten's tie-break). */
int dpb_maxpool2x2_bwd(const float* x, const float* dout, float* dx, int n, int h, int w, int c, dpb_stream_t stream);
/* Per-block partial column sums / sums of squares of x[m][c] for dpb_bn_finalize: part_*: [dpb_column_stats_blocks()][c],
 * *blocks_out rows are written (part_sq may be null). */
int dpb_column_stats_blocks(void);
int dpb_column_stats(const float* x, float* part_sum, float* part_sq, long long m, int c, int* blocks_out,
                     dpb_stream_t stream);
/* out[c] = sum over rows of x[m][c] (bias gradient of a lone conv); part: scratch [dpb_column_stats_blocks()][c]. */
int dpb_column_sum(const float* x, float* part, float* out, long long m, int c, dpb_stream_t stream);
/* y = x * scale + shift: BatchNorm without the ReLU. */
int dpb_bn_apply(const float* x, const float* scale, const float* shift, float* y, long long m, int c,
                 dpb_stream_t stream);
/* BatchNorm backward without ReLU.  part: scratch [2][dpb_column_stats_blocks()][c]; coef: scratch [3][c]. */
int dpb_bn_plain_bwd(const float* x, const float* dout, const float* save_mean, const float* save_invstd,
                     const float* gamma, float* part, float* coef, float* dgamma, float* dbeta, float* dx, long long m,
                     int c, dpb_stream_t stream);

/* Allocation-free variant of dpb_bn_relu_pool_bwd.  (No counterpart in the reference: it exists for callers of this
 * library that cannot obtain device memory themselves -- a host program that talks only to this C ABI, or a plug-in
 * that is given input/output buffers and must not call cudaMalloc or any framework allocator.)
 * Same arguments and results as dpb_bn_relu_pool_bwd minus `part` and `coef`: that scratch lives in a file-scope
 * __device__ array inside the library (about 1.2 MB on every GPU that loads it, held until the process exits).
 *   - limit: c <= dpb_noalloc_max_channels() (512); larger c is refused with an error, nothing is launched.
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress) with that device
 *     current; make that first call before starting any stream capture on the device.
 *   - the call itself performs no cudaMalloc / cudaFree and uses no allocating library routine.
 *   - all calls on a device share the one array, which keeps whatever the last call left in it; nothing needs
 *     clearing between calls (every word read is written earlier in the same call), but two calls whose kernels can
 *     overlap on the GPU (different streams, no ordering) race on it -- serialise them. */
int dpb_noalloc_max_channels(void);
int dpb_bn_relu_pool_bwd_noalloc(const float* y, const float* dout, const float* scale, const float* shift,
                                 const float* save_mean, const float* save_invstd, const float* gamma, float* dgamma,
                                 float* dbeta, float* conv_dbias, float* dy, int n, int h, int w, int c, int pool,
                                 dpb_stream_t stream);

/* ---- classifier head, loss, metrics (ref: model.py:40,45; main.py:34,59-61,99) ----------------------------------- */
int dpb_linear_fwd(const float* x, const float* w, const float* bias, float* out, int b, int k, int o,
                   dpb_stream_t stream);
int dpb_linear_bwd(const float* dz, const float* x, const float* w, float* dx, float* dw, float* db, int b, int k,
                   int o, dpb_stream_t stream);
int dpb_cross_entropy_fwd(const float* logits, const long long* target, float* loss, float* metrics, int b, int o,
                          dpb_stream_t stream);
int dpb_cross_entropy_bwd(const float* logits, const long long* target, const float* grad_out, float* dz, int b,
                          int o, dpb_stream_t stream);

/* ---- optimizer (ref: main.py:36,103-104 torch.optim.SGD) --------------------------------------------------------- */
int dpb_sgd_step(float* p, const float* g, float* m, long long n, float lr, float momentum, float wd,
                 float grad_scale, int nesterov, dpb_stream_t stream);

/* ---- symmetric memory (ref: c10d ProcessGroupGloo transport, main_gather.py:107) --------------------------------- */
int dpb_symm_alloc(size_t bytes, void** ptr);
int dpb_symm_free(void* ptr);
int dpb_symm_export(void* ptr, unsigned char handle64[64]);
int dpb_symm_import(const unsigned char handle64[64], void** peer_ptr);
int dpb_symm_close(void* peer_ptr);
size_t dpb_signal_pad_bytes(void);

/* Communicator: rank, world size and every rank's signal pad. emulate = 1 runs all ranks as blockIdx.y of one
 * cooperative launch on a single GPU (tests). */
typedef struct dpb_comm dpb_comm;
int dpb_comm_create(int rank, int nranks, void* const* signal_pads, int emulate, dpb_comm** out);
int dpb_comm_destroy(dpb_comm* c);

/* bufs[r] = rank r's mapping of the symmetric buffer; offset/count in floats (multiples of 4).
 * algo: 0 auto, 1 one-shot, 2 two-shot.  (ref: main_all_reduce.py:45-48) */
int dpb_allreduce(dpb_comm* c, int channel, void* const* bufs, long long offset, long long count, float scale,
                  int algo, int blocks, dpb_stream_t stream);
/* Two-shot all-reduce with the SGD update fused into the reduce-scatter epilogue: rank r owns slice r of params and
 * momentum, pushes new parameters to every rank.  write_mean also stores the mean gradient everywhere. */
int dpb_allreduce_sgd(dpb_comm* c, int channel, void* const* grads, void* const* params, void* const* momentum,
                      long long offset, long long count, float scale, float lr, float mom, float wd, int nesterov,
                      int write_mean, int blocks, dpb_stream_t stream);
/* (ref: main_gather.py:42-59) */
int dpb_gather_mean_scatter(dpb_comm* c, int channel, void* const* bufs, long long offset, long long count,
                            float scale, int root, int blocks, dpb_stream_t stream);
/* offset/count in 16-byte units.  (ref: DistributedDataParallel _sync_module_states / _sync_buffers) */
int dpb_broadcast(dpb_comm* c, int channel, void* const* bufs, long long offset16, long long count16, int root,
                  int blocks, dpb_stream_t stream);
int dpb_barrier(dpb_comm* c, int channel, dpb_stream_t stream);

/* ---- DDP reducer (ref: main_ddp.py:137 -> c10d Reducer) ----------------------------------------------------------- */
/* Bucket assignment over parameters given in gradient-ready order: first bucket capped at first_cap bytes, later
 * ones at cap bytes; a tensor that overflows the cap closes the bucket after being added.  Returns bucket count. */
int dpb_plan_buckets(int nparams, const long long* bytes_in_ready_order, long long first_cap, long long cap,
                     int* bucket_of_param);
typedef struct dpb_reducer dpb_reducer;
/* offsets/counts (floats, multiples of 4) of each parameter inside the gradient arena, in arena order;
 * bucket_of_param in the same order. Buckets must be contiguous arena ranges. */
int dpb_reducer_create(dpb_comm* c, int nparams, const long long* offsets, const long long* counts,
                       const int* bucket_of_param, int nbuckets, void* const* grad_bufs, float scale,
                       dpb_reducer** out);
int dpb_reducer_destroy(dpb_reducer* r);
int dpb_reducer_prepare(dpb_reducer* r);
/* Called from the autograd post-accumulate hook of parameter `param`; launches the bucket's all-reduce on the side
 * stream once its last gradient is ready. */
int dpb_reducer_mark_ready(dpb_reducer* r, int param, dpb_stream_t main_stream);
/* Bucket reductions launched after this call also wait for the work `producer` holds now (weight gradients computed on
 * a second compute stream).  No counterpart in the reference: its backward is single-stream. */
int dpb_reducer_wait_stream(dpb_reducer* r, dpb_stream_t producer);
/* Makes main_stream wait for every outstanding bucket. */
int dpb_reducer_finalize(dpb_reducer* r, dpb_stream_t main_stream);
int dpb_reducer_launched(dpb_reducer* r); /* buckets launched since the last prepare */

#ifdef __cplusplus
}
#endif
#endif /* DPB200_H_ */
