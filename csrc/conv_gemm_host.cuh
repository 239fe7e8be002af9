// Host side of the tcgen05 implicit-GEMM convolution: TMA tensor-map construction (driver entry point fetched at
// run time, so the library links without libcuda) and launch heuristics (N tile, split-K factor).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <mutex>
#include "conv_gemm.cuh"
#include "conv_gemm_persistent.cuh"
#include "splitk_reduce.cuh"
#include "errors.cuh"

namespace dpb {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<PFN_encodeTiled>(sym);
  });
  return fn;
}

// fp32 tensor map, SWIZZLE_128B, zero fill out of bounds.  dims/strides innermost first; strides[i] is the byte pitch
// of dimension i+1.
// mn_major: the box feeds an MN-major tf32 UMMA operand, whose only legal shared-memory layout is the 128-byte
// swizzle with a 32-byte atom (UMMA layout type SWIZZLE_128B_BASE32B <-> TMA SWIZZLE_128B_ATOM_32B).
inline int make_map_f32(CUtensorMap* map, const void* base, int rank, const cuuint64_t* dims,
                        const cuuint64_t* strides_bytes, const cuuint32_t* box, bool mn_major = false) {
  PFN_encodeTiled enc = get_encode_tiled();
  if (!enc) return set_error(-10, "cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(base), dims,
                   strides_bytes, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(-11, "cuTensorMapEncodeTiled failed (rank %d, CUresult %d)", rank, (int)r);
  return 0;
}

// Launch with programmatic stream serialization (see pdl_prologue in ptx.cuh); dpb_set_tuning key 4 switches it off.
inline int g_tune_use_pdl = 0;  // measured: no gain inside CUDA graphs (1.074 vs 1.079 ms/step), so off by default
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = g_tune_use_pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// Opt a kernel in to more than 48 KB of dynamic shared memory.  The attribute is per device, so remember (per call
// site = per kernel instantiation) on which devices of this process it has been applied.
#define DPB_SET_MAX_SMEM_ONCE(kern, bytes)                                                            \
  do {                                                                                                \
    static unsigned long long seen_ = 0;                                                              \
    int dev_ = 0;                                                                                     \
    DPB_CUDA(cudaGetDevice(&dev_));                                                                   \
    const unsigned long long bit_ = 1ull << (dev_ & 63);                                              \
    if (!(seen_ & bit_)) {                                                                            \
      DPB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
      seen_ |= bit_;                                                                                  \
    }                                                                                                 \
  } while (0)

// bring-up aid, set through dpb_debug_set_conv_dump()
inline float* g_conv_debug_dump = nullptr;
// kernels launched by the most recent conv_gemm_dispatch on this thread (1, or 2 with the split-K reduce)
inline thread_local int g_conv_last_launches = 0;

struct ConvGeom {
  int n, h, w, cin, cout;
};

inline void pixel_box(int h, int w, int pixels, int* bw, int* bh, int* bn) {
  *bw = w < pixels ? w : pixels;
  int rest = pixels / *bw;
  *bh = h < rest ? h : rest;
  *bn = rest / *bh;
}

template <int MODE, int NT, int CL>
int launch_conv_gemm_cl(const CUtensorMap& mA, const CUtensorMap& mB, const ConvGemmParams& p, dim3 grid,
                        cudaStream_t stream) {
  // NT 128: 3 x 32 KB (two CTAs per SM); NT 64: 3 x 24 KB (three CTAs per SM: the N = 64 MMA is issue-limited per CTA,
  // see tools/mma_rate.cu, so more resident CTAs is what fills the tensor pipe); NT 256: 4 x 48 KB (one CTA per SM)
  constexpr int STAGES = (NT == 256) ? 4 : 3;
  using S = ConvGemmSmem<NT, STAGES>;
  auto kern = conv_gemm_kernel<MODE, NT, STAGES, CL>;
  DPB_SET_MAX_SMEM_ONCE(kern, S::kTotal);
  if (CL == 1) {
    DPB_CUDA(launch_k(kern, grid, dim3(192), S::kTotal, stream, mA, mB, p));
  } else {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(192);
    cfg.dynamicSmemBytes = S::kTotal;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CL;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    DPB_CUDA(cudaLaunchKernelEx(&cfg, kern, mA, mB, p));
  }
  const bool reduce_kernel = p.splits > 1 && !p.pair_reduce;
  g_conv_last_launches = reduce_kernel ? 2 : 1;
  if (reduce_kernel) {
    const int m_tiles = grid.x, n_tiles = grid.y;
    const int tiles = m_tiles * n_tiles * (MODE == kWgrad ? 9 : 1);
    const int pieces = splitk_pieces(tiles, m_tiles, NT);
    DPB_CUDA(launch_k(splitk_reduce_kernel<MODE>, dim3(tiles, pieces), dim3(256), 0, stream, p, NT, m_tiles, n_tiles, pieces));
  }
  return 0;
}

// CTA-pair variant (cta_group::2): cluster (2,1,1), each CTA stages half of B.  cl == kPairMode selects it.
constexpr int kPairMode = -2;
template <int MODE, int NT>
int launch_conv_gemm_pair(const CUtensorMap& mA, const CUtensorMap& mB, const ConvGemmParams& p, dim3 grid,
                          cudaStream_t stream) {
  constexpr int STAGES = 4;  // per-CTA stage: 16 KB A + NT/2 x 128 B of B (NT 128: 96 KB, NT 256: 128 KB = the staged tile)
  using S = ConvGemmSmem<NT, STAGES, 2>;
  auto kern = conv_gemm_kernel<MODE, NT, STAGES, 1, 2>;
  DPB_SET_MAX_SMEM_ONCE(kern, S::kTotal);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(192);
  cfg.dynamicSmemBytes = S::kTotal;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  DPB_CUDA(cudaLaunchKernelEx(&cfg, kern, mA, mB, p));
  const bool reduce_kernel = p.splits > 1 && !p.pair_reduce;
  g_conv_last_launches = reduce_kernel ? 2 : 1;
  if (reduce_kernel) {
    const int m_tiles = grid.x, n_tiles = grid.y;
    const int tiles = m_tiles * n_tiles * (MODE == kWgrad ? 9 : 1);
    const int pieces = splitk_pieces(tiles, m_tiles, NT);
    DPB_CUDA(launch_k(splitk_reduce_kernel<MODE>, dim3(tiles, pieces), dim3(256), 0, stream, p, NT, m_tiles, n_tiles, pieces));
  }
  return 0;
}

// Persistent variant (conv_gemm_persistent.cuh): one CTA per SM walks the work units the plain kernel would launch as
// CTAs.  cl == kPersistMode selects it.
constexpr int kPersistMode = -3;
inline int device_sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
      sms = 148;
  }
  return sms;
}
template <int MODE, int NT>
int launch_conv_gemm_persistent(const CUtensorMap& mA, const CUtensorMap& mB, const ConvGemmParams& p, dim3 grid,
                                cudaStream_t stream) {
  constexpr int STAGES = (NT == 256) ? 4 : 6;  // NT 128: 6 x 32 KB, NT 64: 6 x 24 KB, NT 256: 4 x 48 KB
  using S = ConvPersistSmem<NT, STAGES>;
  auto kern = conv_gemm_persistent_kernel<MODE, NT, STAGES>;
  DPB_SET_MAX_SMEM_ONCE(kern, S::kTotal);
  const int units = (int)(grid.x * grid.y * grid.z);
  const int ctas = units < device_sm_count() ? units : device_sm_count();
  DPB_CUDA(launch_k(kern, dim3(ctas), dim3(192), S::kTotal, stream, mA, mB, p, (int)grid.x, (int)grid.y, units));
  const bool reduce_kernel = p.splits > 1;
  g_conv_last_launches = reduce_kernel ? 2 : 1;
  if (reduce_kernel) {
    const int m_tiles = grid.x, n_tiles = grid.y;
    const int tiles = m_tiles * n_tiles * (MODE == kWgrad ? 9 : 1);
    const int pieces = splitk_pieces(tiles, m_tiles, NT);
    DPB_CUDA(launch_k(splitk_reduce_kernel<MODE>, dim3(tiles, pieces), dim3(256), 0, stream, p, NT, m_tiles, n_tiles, pieces));
  }
  return 0;
}

template <int MODE, int NT>
int launch_conv_gemm_nt(int cl, const CUtensorMap& mA, const CUtensorMap& mB, const ConvGemmParams& p, dim3 grid,
                        cudaStream_t stream) {
  if (cl == kPersistMode) return launch_conv_gemm_persistent<MODE, NT>(mA, mB, p, grid, stream);
  if (cl == kPairMode) return launch_conv_gemm_pair<MODE, NT>(mA, mB, p, grid, stream);
  if (cl == 2) return launch_conv_gemm_cl<MODE, NT, 2>(mA, mB, p, grid, stream);
  if constexpr (NT >= 128) {
    if (cl == 4) return launch_conv_gemm_cl<MODE, NT, 4>(mA, mB, p, grid, stream);
  }
  return launch_conv_gemm_cl<MODE, NT, 1>(mA, mB, p, grid, stream);
}

template <int MODE>
int launch_conv_gemm_mode(int nt, int cl, const CUtensorMap& mA, const CUtensorMap& mB, const ConvGemmParams& p,
                          dim3 grid, cudaStream_t stream) {
  switch (nt) {
    case 64: return launch_conv_gemm_nt<MODE, 64>(cl, mA, mB, p, grid, stream);
    case 128: return launch_conv_gemm_nt<MODE, 128>(cl, mA, mB, p, grid, stream);
    case 256: return launch_conv_gemm_nt<MODE, 256>(cl, mA, mB, p, grid, stream);
  }
  return set_error(-12, "unsupported N tile %d", nt);
}

// Cluster size along M for B-tile multicast: the largest of {tuned value, 2, 1} that divides the M tiles and the B atoms.
// g_tune_pair: use CTA pairs (cta_group::2) whenever the M tiles come in twos; returns kPairMode then.
inline int g_tune_cluster = 1;
inline int g_tune_pair = 0;
inline int g_tune_persistent = 0;  // bit 0: fprop, bit 1: dgrad, bit 2: wgrad run the persistent kernel
inline int pick_cluster(int m_tiles, int nt, int mode = 0) {
  if ((g_tune_persistent >> mode) & 1) return kPersistMode;
  if (g_tune_pair && m_tiles % 2 == 0 && (nt / 32) % 2 == 0) return kPairMode;
  int cl = g_tune_cluster;
  while (cl > 1 && (m_tiles % cl != 0 || (nt / 32) % cl != 0)) cl >>= 1;
  return cl < 1 ? 1 : cl;
}

constexpr size_t kConvWorkspaceBytes = (size_t)48 << 20;  // split-K partial tiles
constexpr int kConvCounterSlots = 8192;

// Number of CTAs we aim for: two waves of single-tile CTAs on 148 SMs.
// Tunables (dpb_set_tuning): CTA count the split-K heuristic aims for, and the N tile used when the GEMM N extent allows it.
inline int g_tune_target_ctas = 296;
inline int g_tune_wide_nt = 128;
inline int g_tune_min_k_iters = 8;  // k-iterations (of 32) a split must keep; fewer, fatter splits = less workspace traffic
// Two-way splits finished inside the GEMM kernel (ConvGemmParams::pair_reduce).  Off: measured 1.010 vs 0.976 ms/step
// at batch 256 although it removes 7 launches -- the second CTA's extra fence + 64 KB read + store sit on every
// tile's critical path, while the separate reduce kernel spreads the same bytes over the whole GPU.
inline int g_tune_pair_reduce = 0;

inline int pick_splits(int tiles, int k_iters, int nt, int forced) {
  int splits = 1;
  if (forced > 0) {
    splits = forced;
  } else if (tiles < g_tune_target_ctas) {
    splits = g_tune_target_ctas / tiles;
    int max_by_k = k_iters / g_tune_min_k_iters;  // keep a minimum amount of K per split
    if (max_by_k < 1) max_by_k = 1;
    if (splits > max_by_k) splits = max_by_k;
  }
  if (splits > k_iters) splits = k_iters;
  if (splits < 1) splits = 1;
  // workspace bound
  while (splits > 1 && (size_t)tiles * splits * 128 * nt * 4 > kConvWorkspaceBytes) --splits;
  return splits;
}

// mode: 0 fprop  (in = X[N,H,W,Cin], w = W[Cout,3,3,Cin], out = Y[N,H,W,Cout])
//       1 dgrad  (in = dY[N,H,W,Cout], w = W, out = dX[N,H,W,Cin])
//       2 wgrad  (in = dY[N,H,W,Cout], in2 = X[N,H,W,Cin], out = dW[Cout,3,3,Cin])
inline int conv_gemm_dispatch(int mode, const ConvGeom& g, const float* in, const float* in2, const float* w,
                              float* out, const float* bias, float* stat_sum, float* stat_sq, float* workspace,
                              unsigned int* counters, int nt_override, int splits_override, int* out_m_tiles,
                              cudaStream_t stream) {
  if (g.cin % 32 || g.cout % 32) return set_error(-13, "conv_gemm needs Cin, Cout multiples of 32 (got %d, %d)", g.cin, g.cout);
  if (g.w > 32 || (g.w & (g.w - 1)) || (g.h & (g.h - 1))) return set_error(-13, "conv_gemm needs power-of-two H, W with W <= 32 (got %d x %d)", g.h, g.w);
  const long long pixels = (long long)g.n * g.h * g.w;
  ConvGemmParams p{};
  p.H = g.h; p.W = g.w; p.NImg = g.n; p.CinW = g.cin;
  p.bias = bias; p.stat_sum = stat_sum; p.stat_sq = stat_sq; p.workspace = workspace; p.counters = counters;
  p.out = out;
  p.debug_dump = g_conv_debug_dump;
  CUtensorMap mA, mB;
  dim3 grid;
  int nt;
  int rc;
  if (mode == kFprop || mode == kDgrad) {
    const int ck = (mode == kFprop) ? g.cin : g.cout;   // GEMM K channels
    const int cn = (mode == kFprop) ? g.cout : g.cin;   // GEMM N channels
    nt = nt_override > 0 ? nt_override : (cn % g_tune_wide_nt == 0 ? g_tune_wide_nt : (cn >= 128 ? 128 : 64));
    if (cn % nt) return set_error(-13, "N tile %d does not divide %d", nt, cn);
    int bw, bh, bn;
    pixel_box(g.h, g.w, 128, &bw, &bh, &bn);
    p.box_w = bw; p.box_h = bh; p.box_n = bn;
    p.Ck = ck;
    p.M_total = (int)pixels; p.N_total = cn;
    p.k_iters = 9 * (ck / 32);
    p.out_ld = cn;
    const int m_tiles = (int)((pixels + 127) / 128);
    const int n_tiles = cn / nt;
    p.splits = pick_splits(m_tiles * n_tiles, p.k_iters, nt, splits_override);
    p.pair_reduce = (p.splits == 2 && g_tune_pair_reduce && m_tiles * n_tiles <= kConvCounterSlots) ? 1 : 0;
    const int cl = pick_cluster(m_tiles, nt, mode);
    if (cl == kPersistMode) p.pair_reduce = 0;  // the persistent kernel always leaves split-K to the reduce kernel
    {
      cuuint64_t dims[4] = {(cuuint64_t)ck, (cuuint64_t)g.w, (cuuint64_t)g.h, (cuuint64_t)g.n};
      cuuint64_t str[3] = {(cuuint64_t)ck * 4, (cuuint64_t)g.w * ck * 4, (cuuint64_t)g.h * g.w * ck * 4};
      cuuint32_t box[4] = {32, (cuuint32_t)bw, (cuuint32_t)bh, (cuuint32_t)bn};
      if ((rc = make_map_f32(&mA, in, 4, dims, str, box))) return rc;
    }
    if (mode == kFprop) {
      cuuint64_t dims[2] = {(cuuint64_t)9 * g.cin, (cuuint64_t)g.cout};
      cuuint64_t str[1] = {(cuuint64_t)9 * g.cin * 4};
      // each CTA of a multicast cluster fetches nt/cl rows; each CTA of a pair stages nt/2 rows
      cuuint32_t box[2] = {32, (cuuint32_t)(cl == kPairMode ? nt / 2 : (cl > 1 ? nt / cl : nt))};
      if ((rc = make_map_f32(&mB, w, 2, dims, str, box))) return rc;
    } else {
      cuuint64_t dims[2] = {(cuuint64_t)9 * g.cin, (cuuint64_t)g.cout};
      cuuint64_t str[1] = {(cuuint64_t)9 * g.cin * 4};
      cuuint32_t box[2] = {32, 32};
      if ((rc = make_map_f32(&mB, w, 2, dims, str, box, true))) return rc;
    }
    grid = dim3(m_tiles, n_tiles, p.splits);
    if (out_m_tiles)  // BatchNorm partial rows written
      *out_m_tiles = (p.splits > 1 && !p.pair_reduce) ? splitk_pieces(m_tiles * n_tiles, m_tiles, nt) * m_tiles : m_tiles;
    if (mode == kFprop) return launch_conv_gemm_mode<kFprop>(nt, cl, mA, mB, p, grid, stream);
    return launch_conv_gemm_mode<kDgrad>(nt, cl, mA, mB, p, grid, stream);
  }
  // wgrad
  nt = nt_override > 0 ? nt_override : (g.cin % g_tune_wide_nt == 0 ? g_tune_wide_nt : (g.cin >= 128 ? 128 : 64));
  if (g.cin % nt) return set_error(-13, "N tile %d does not divide Cin %d", nt, g.cin);
  int bw, bh, bn;
  pixel_box(g.h, g.w, 32, &bw, &bh, &bn);
  p.box_w = bw; p.box_h = bh; p.box_n = bn;
  p.Ck = 0;
  p.M_total = g.cout; p.N_total = g.cin;
  p.k_iters = (int)((pixels + 31) / 32);
  p.out_ld = (long long)9 * g.cin;
  const int m_tiles = (g.cout + 127) / 128;
  const int n_tiles = g.cin / nt;
  p.splits = pick_splits(m_tiles * n_tiles * 9, p.k_iters, nt, splits_override);
  p.pair_reduce = (p.splits == 2 && g_tune_pair_reduce && m_tiles * n_tiles * 9 <= kConvCounterSlots) ? 1 : 0;
  {
    cuuint64_t dims[2] = {(cuuint64_t)g.cout, (cuuint64_t)pixels};
    cuuint64_t str[1] = {(cuuint64_t)g.cout * 4};
    cuuint32_t box[2] = {32, 32};
    if ((rc = make_map_f32(&mA, in, 2, dims, str, box, true))) return rc;
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)g.cin, (cuuint64_t)g.w, (cuuint64_t)g.h, (cuuint64_t)g.n};
    cuuint64_t str[3] = {(cuuint64_t)g.cin * 4, (cuuint64_t)g.w * g.cin * 4, (cuuint64_t)g.h * g.w * g.cin * 4};
    cuuint32_t box[4] = {32, (cuuint32_t)bw, (cuuint32_t)bh, (cuuint32_t)bn};
    if ((rc = make_map_f32(&mB, in2, 4, dims, str, box, true))) return rc;
  }
  grid = dim3(m_tiles, n_tiles, 9 * p.splits);
  if (out_m_tiles) *out_m_tiles = m_tiles;
  return launch_conv_gemm_mode<kWgrad>(nt, pick_cluster(m_tiles, nt, kWgrad), mA, mB, p, grid, stream);
}

}  // namespace dpb
