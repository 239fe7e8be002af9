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
int dpb_version(void); /* ABI version; the Python loader refuses a library whose version differs from its own */
const char* dpb_last_error(void);
/* Once per device, with that device current (dpb_comm_create does it too): registers the mapped host record that a
 * kernel fills in before it gives up, and applies the peer-signal time-out to the device. */
int dpb_init_device(void);
/* Diagnostic left behind by a kernel that gave up (pipeline or peer-signal time-out).  Reads host memory only, so it
 * works after the failed launch has left the CUDA context unusable.  device < 0 = any initialised device.
 * out[6] = code (0 none, 1 mbarrier, 2 peer signal), where (pipeline slot / channel), block, aux0 (peer rank waited
 * for), aux1 (epoch), rank (the waiting rank).  Clears the record. */
int dpb_last_device_error(int device, int* out6);
/* Bring-up aid, only in builds with -DDPB_EXPERIMENTS (tools/conv_timing.py): per-CTA trace of the conv kernel. */
int dpb_debug_set_conv_dump(void* buf);
/* Experiments builds: every kernel records {id, globaltimer at entry, at exit} of its block 0 into buf
 * (uint64: [0] = record count, then triples; room for 4096 records).  NULL switches it off.  tools/overlap_timeline.py */
int dpb_debug_set_trace(void* buf);
/* Scratch needed by the tensor-core convolution: split-K partial tiles of the workspace path. */
size_t dpb_conv_workspace_bytes(void);

/* ---- convolution (ref: model.py:18-23 -> aten convolution / convolution_backward) ------------------------------ */
/* mode 0 fprop: in = X, w = W, out = Y (+bias)
 * mode 1 dgrad: in = dY, w = W, out = dX
 * mode 2 wgrad: in = dY, in2 = X, out = dW
 * nt / splits / target_ctas: 0 = heuristic (target_ctas: CTA count the split-K heuristic aims for in this call, e.g.
 * one per SM for two GEMMs that share the GPU on two streams).  Split-K factors that are powers of two up to 16 are reduced inside the kernel by a
 * thread-block cluster over distributed shared memory; other factors go through `workspace` and a reduce kernel.
 * stat_acc (fprop, optional): BatchNorm batch statistics -- [2][cout] 64-bit fixed-point accumulators (2^-20 units; zero
 * on entry) to which the kernels add the per-channel sum and sum of squares of Y; dpb_bn_relu_pool_fwd_train consumes
 * and clears them.  Integer atomics make the totals independent of the arrival order. */
int dpb_conv3x3_gemm(int mode, int n, int h, int w, int cin, int cout, const float* in, const float* in2,
                     const float* weight, float* out, const float* bias, void* stat_acc, void* workspace, int nt,
                     int splits, int target_ctas, dpb_stream_t stream);
/* Heuristic knobs (the Python loader also reads them from the environment variable in brackets).
 *  0 [DPB_CONV_TARGET_CTAS]  CTA count the split-K heuristic aims for (default 296 = two per SM)
 *  1 [DPB_CONV_NT]           preferred N tile (64 / 128 / 256, default 128) when it divides the GEMM N extent
 *  2 [DPB_COMM_BLOCKS]       default grid cap of the collective kernels (default 64, at most 128)
 *  3 [DPB_CONV_MIN_K_ITERS]  K iterations (of 32) a split must keep (default 8)
 *  4 [DPB_PDL]               programmatic dependent launch on every compute kernel (default 0)
 *  5 [DPB_CONV_CLUSTER_MAX]  largest split-K factor reduced over distributed shared memory (1 = never, default 8)
 *  7 [DPB_FIRST_TC]          first (Cin = 3) layer on the tensor cores when the image tiles into 128 pixels (default 1)
 * 11 [DPB_COMM_WEAK]         two-shot all-reduce payload with weak accesses (default 0)
 * 12 [DPB_PEER_TIMEOUT_MS]   wall-clock budget of a peer-signal wait (default 120000)
 * 13 [DPB_DDP_BLOCKS]        grid of a per-bucket collective launched next to backward (default 16)
 * 14 [DPB_NVLS]              in-switch (multimem) all-reduce when the buffers are multicast-bound (default 1)
 * 15 [DPB_NVLS_BLOCKS]       its grid when it has the GPU to itself (default 16)
 * 16 [DPB_CONV_CLUSTER_REACH] a GEMM wanting up to cluster_max * reach / 2 splits still takes the cluster path (default 3)
 * Returns non-zero for an unknown key / bad value. */
int dpb_set_tuning(int key, int value);
/* Kernels launched by the calling thread's most recent dpb_conv3x3_gemm (1, or 2 when split-K adds its reduce). */
int dpb_conv_last_launches(void);
/* First layer (Cin = 3, Cout = 64): x is NCHW.  Statistics either as partial rows part_* :
 * [dpb_first_conv_blocks(n,h,w)][64] (for dpb_bn_finalize) or into stat_acc as above (tensor-core kernel only:
 * dpb_first_conv_uses_tensor_cores); any of them may be NULL. */
int dpb_first_conv_blocks(int n, int h, int w);
int dpb_first_conv_uses_tensor_cores(int h, int w);
int dpb_first_conv_fprop(const float* x_nchw, const float* weight, const float* bias, float* y, float* part_sum,
                         float* part_sq, void* stat_acc, int n, int h, int w, dpb_stream_t stream);
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
/* Training-mode forward straight from the accumulators the conv filled: every block derives scale / shift in its
 * prologue, block 0 publishes stats4 = [scale | shift | mean | invstd] ([4][c], needed by the backward pass) and updates
 * the running statistics / num_batches_tracked; the last block clears `acc`.  ticket: one zeroed uint32, self-resetting.
 * No finalize launch.  (ref: model.py:24-25,16) */
int dpb_bn_relu_pool_fwd_train(const float* y, void* acc, void* ticket, const float* gamma, const float* beta,
                               float* running_mean, float* running_var, long long* num_batches_tracked, float momentum,
                               float eps, float* stats4, float* out, int n, int h, int w, int c, int pool,
                               dpb_stream_t stream);
/* Backward of the fused block.
 * With acc ([2][c] zeroed 64-bit accumulators, cleared again by the call) and ticket (one zeroed uint32) and c a
 * multiple of 32: TWO launches -- column sums of g and g*xhat into the accumulators, then the apply pass, which derives
 * its coefficients from them.  part / coef are not used then.
 * With acc == NULL: three launches (reduce into part: scratch [2][dpb_bn_bwd_blocks()][c], finalize into coef: scratch
 * [3][c], apply).
 * conv_dbias (optional): gradient of the preceding conv's bias, exactly zero under training-mode BatchNorm. */
int dpb_bn_bwd_blocks(void);
int dpb_bn_bwd_launches(int c, int with_acc);
int dpb_bn_relu_pool_bwd(const float* y, const float* dout, const float* scale, const float* shift,
                         const float* save_mean, const float* save_invstd, const float* gamma, float* part,
                         float* coef, void* acc, void* ticket, float* dgamma, float* dbeta, float* conv_dbias, float* dy,
                         int n, int h, int w, int c, int pool, dpb_stream_t stream);

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
 * Same arguments and results as dpb_bn_relu_pool_bwd minus its scratch arguments (`part`, `coef`, and the optional
 * `acc` / `ticket`; it always takes the three-launch path): that scratch lives in a file-scope
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
/* The whole training head in two launches (ref: model.py:40,45 + main.py:34-35).  dpb_head_fwd: logits = x W^T + b,
 * mean cross-entropy -> loss[0] (and metrics[0] += loss, metrics[1] += arg-max hits if metrics != NULL),
 * dz = (softmax - onehot) / b (optional).  terms: scratch [b][2]; counter: one zeroed uint32 (self-resetting).
 * dpb_head_bwd: dx = g dz W, dw = g dz^T x, db = g sum dz with g = grad_out[0] (device scalar, NULL = 1). */
int dpb_head_supported(int k, int o);
int dpb_head_fwd(const float* x, const float* w, const float* bias, const long long* target, float* logits, float* dz,
                 float* terms, float* loss, float* metrics, void* counter, int b, int k, int o, dpb_stream_t stream);
int dpb_head_bwd(const float* dz, const float* x, const float* w, const float* grad_out, float* dx, float* dw,
                 float* db, int b, int k, int o, dpb_stream_t stream);

/* ---- input pipeline (ref: main.py:71-90 torchvision transforms + DataLoader workers) ------------------------------ */
/* One launch per batch: gather `b` images (uint8 [N][3][h][w], resident on the device) by `indices`, apply
 * RandomCrop(pad 4) + horizontal flip (train != 0; the draws are a hash of seed, epoch and the dataset index, or taken
 * from `draws` = [b][3] int32 dx, dy, flip when given) and Normalize, write fp32 NCHW + int64 labels. */
int dpb_augment_batch(const unsigned char* images, const long long* labels, const int* indices, int b, int h, int w,
                      int train, unsigned int seed, unsigned int epoch, const int* draws, float mean0, float mean1,
                      float mean2, float inv_std0, float inv_std1, float inv_std2, float* out_x, long long* out_y,
                      dpb_stream_t stream);

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
/* The same through the CUDA virtual-memory-management API, which is what an NVSwitch multicast object can be bound to.
 * The ranks exchange the POSIX file descriptors themselves (unix socket, see parallel/comm.py).
 * dpb_vmm_probe: out3 = {VMM + fd handles supported, multicast supported, granularity in bytes}.
 * dpb_vmm_alloc: bytes (a multiple of the granularity) of zeroed device memory, mapped; handle + exportable fd.
 * dpb_vmm_import: a peer's allocation (map = 1) or a multicast object (map = 0) from its fd.
 * dpb_mc_*: rank 0 creates the multicast object, every rank adds its device, then -- after a barrier -- binds its
 * memory and maps the multicast view (dpb_mc_bind_map). */
int dpb_vmm_probe(int ndev, long long* out3);
int dpb_vmm_alloc(size_t bytes, size_t gran, void** ptr, unsigned long long* handle, int* fd);
int dpb_vmm_import(int fd, size_t bytes, size_t gran, int map, void** ptr, unsigned long long* handle);
int dpb_vmm_unmap(void* ptr, size_t bytes, unsigned long long handle);
int dpb_mc_create(int ndev, size_t bytes, unsigned long long* handle, int* fd);
int dpb_mc_add_device(unsigned long long mc);
int dpb_mc_bind_map(unsigned long long mc, unsigned long long mem, size_t bytes, size_t gran, void** mc_ptr);
int dpb_mc_unbind(unsigned long long mc, size_t bytes);

/* Communicator: rank, world size and every rank's signal pad. emulate = 1 runs all ranks as blockIdx.y of one
 * cooperative launch on a single GPU (tests). */
typedef struct dpb_comm dpb_comm;
int dpb_comm_create(int rank, int nranks, void* const* signal_pads, int emulate, dpb_comm** out);
int dpb_comm_destroy(dpb_comm* c);
/* Test hook (emulated worlds): virtual rank `rank` never signals, so its peers run into the peer-signal time-out. */
int dpb_comm_debug_withhold(dpb_comm* c, int rank);

/* bufs[r] = rank r's mapping of the symmetric buffer; offset/count in floats (multiples of 4).
 * algo: 0 auto, 1 one-shot, 2 two-shot.  (ref: main_all_reduce.py:45-48) */
int dpb_allreduce(dpb_comm* c, int channel, void* const* bufs, long long offset, long long count, float scale,
                  int algo, int blocks, dpb_stream_t stream);
/* Two-shot all-reduce with the SGD update fused into the reduce-scatter epilogue: rank r owns slice r of params and
 * momentum, pushes new parameters to every rank.  write_mean also stores the mean gradient everywhere. */
int dpb_allreduce_sgd(dpb_comm* c, int channel, void* const* grads, void* const* params, void* const* momentum,
                      long long offset, long long count, float scale, float lr, float mom, float wd, int nesterov,
                      int write_mean, int blocks, dpb_stream_t stream);
/* The same collective reduced inside the NVSwitch (multimem.ld_reduce / multimem.st): mc_grads / mc_params are the
 * multicast views of the gradient / parameter arenas (dpb_mc_bind_map); local_params / local_mom non-NULL fuse the SGD
 * update as above.  Real multi-GPU worlds only.  (ref: main_all_reduce.py:45-48) */
int dpb_allreduce_nvls(dpb_comm* c, int channel, void* mc_grads, void* local_params, void* mc_params, void* local_mom,
                       long long offset, long long count, float scale, float lr, float mom, float wd, int nesterov,
                       int write_mean, int blocks, dpb_stream_t stream);
int dpb_nvls_enabled(void);
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
/* Fuse the optimizer into every bucket's collective (NULL param_bufs switches it off again). */
int dpb_reducer_set_sgd(dpb_reducer* r, void* const* param_bufs, void* const* mom_bufs, float lr, float momentum,
                        float wd, int nesterov);
/* Multicast views of the arenas: buckets are then reduced inside the NVSwitch. */
int dpb_reducer_set_multicast(dpb_reducer* r, void* mc_grads, void* mc_params, void* local_params, void* local_mom);
int dpb_reducer_side_stream(dpb_reducer* r, dpb_stream_t* out);
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
