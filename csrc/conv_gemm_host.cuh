// Host side of the tcgen05 implicit-GEMM convolution: TMA tensor-map construction (driver entry point fetched at
// run time, so the library links without libcuda) and launch heuristics (N tile, split-K factor).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <mutex>
#include "conv_gemm.cuh"
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

#ifdef DPB_EXPERIMENTS
// bring-up aid, set through dpb_debug_set_conv_dump()
inline float* g_conv_debug_dump = nullptr;
#endif
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

// cluster = true: the `p.splits` CTAs of a tile form a thread-block cluster along z and reduce over DSMEM (one launch);
// otherwise splits > 1 goes through the workspace and splitk_reduce_kernel (two launches).
template <int MODE, int NT>
int launch_conv_gemm_nt(bool cluster, const CUtensorMap& mA, const CUtensorMap& mB, const ConvGemmParams& p, dim3 grid,
                        cudaStream_t stream) {
  // NT 128: 3 x 32 KB (two CTAs per SM); NT 64: 3 x 24 KB (three CTAs per SM: the N = 64 MMA is issue-limited per CTA,
  // see tools/mma_rate.cu, so more resident CTAs is what fills the tensor pipe); NT 256: 4 x 48 KB (one CTA per SM)
  constexpr int STAGES = (NT == 256) ? 4 : 3;
  using S = ConvGemmSmem<NT, STAGES>;
  if (cluster) {
    auto kern = conv_gemm_kernel<MODE, NT, STAGES, true>;
    DPB_SET_MAX_SMEM_ONCE(kern, S::kTotal);
    if (p.splits > 8) {  // 16-CTA clusters are beyond the portable limit: opt in once per device
      static unsigned long long seen = 0;
      int dev = 0;
      DPB_CUDA(cudaGetDevice(&dev));
      if (!(seen & (1ull << (dev & 63)))) {
        DPB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        seen |= 1ull << (dev & 63);
      }
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(192);
    cfg.dynamicSmemBytes = S::kTotal;
    cfg.stream = stream;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 1;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = (unsigned)p.splits;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = g_tune_use_pdl ? 2 : 1;
    DPB_CUDA(cudaLaunchKernelEx(&cfg, kern, mA, mB, p));
    g_conv_last_launches = 1;
    return 0;
  }
  auto kern = conv_gemm_kernel<MODE, NT, STAGES, false>;
  DPB_SET_MAX_SMEM_ONCE(kern, S::kTotal);
  DPB_CUDA(launch_k(kern, grid, dim3(192), S::kTotal, stream, mA, mB, p));
  g_conv_last_launches = 1;
  if (p.splits > 1) {
    const int m_tiles = grid.x, n_tiles = grid.y;
    const int tiles = m_tiles * n_tiles * (MODE == kWgrad ? 9 : 1);
    const int pieces = splitk_pieces(tiles, m_tiles, NT);
    DPB_CUDA(launch_k(splitk_reduce_kernel<MODE>, dim3(tiles, pieces), dim3(256), 0, stream, p, NT, m_tiles, n_tiles, pieces));
    g_conv_last_launches = 2;
  }
  return 0;
}

template <int MODE>
int launch_conv_gemm_mode(int nt, bool cluster, const CUtensorMap& mA, const CUtensorMap& mB, const ConvGemmParams& p,
                          dim3 grid, cudaStream_t stream) {
  switch (nt) {
    case 64: return launch_conv_gemm_nt<MODE, 64>(cluster, mA, mB, p, grid, stream);
    case 128: return launch_conv_gemm_nt<MODE, 128>(cluster, mA, mB, p, grid, stream);
    case 256: return launch_conv_gemm_nt<MODE, 256>(cluster, mA, mB, p, grid, stream);
  }
  return set_error(-12, "unsupported N tile %d", nt);
}

constexpr size_t kConvWorkspaceBytes = (size_t)48 << 20;  // split-K partial tiles (workspace path)

// Number of CTAs we aim for: two waves of single-tile CTAs on 148 SMs.
// Tunables (dpb_set_tuning): CTA count the split-K heuristic aims for, and the N tile used when the GEMM N extent allows it.
inline int g_tune_target_ctas = 296;
inline int g_tune_wide_nt = 128;
inline int g_tune_min_k_iters = 8;  // k-iterations (of 32) a split must keep; fewer, fatter splits = less reduce traffic
// Largest split factor that is reduced inside the GEMM kernel over distributed shared memory (thread-block cluster along
// the split axis; 16 needs the non-portable cluster size).  1 sends every split-K GEMM through the workspace + reduce
// kernel instead.  8 is the measured optimum (profiles/step_variants.md): 16-CTA clusters do not pack onto the GPCs --
// of 256 CTAs 32 start a whole kernel-life late.
inline int g_tune_cluster_max = 8;
// A GEMM that wants up to cluster_max * reach / 2 splits still takes the cluster path (with cluster_max splits and
// correspondingly longer K loops); beyond that it goes through the workspace with as many splits as it wants.
inline int g_tune_cluster_reach = 3;

struct SplitPlan {
  int splits;
  bool cluster;
};

inline SplitPlan plan_splits(int tiles, int m_tiles, int k_iters, int nt, int forced, int target) {
  if (target <= 0) target = g_tune_target_ctas;
  int want = 1;
  if (forced > 0) {
    want = forced;
  } else if (tiles < target) {
    want = target / tiles;
    int max_by_k = k_iters / g_tune_min_k_iters;  // keep a minimum amount of K per split
    if (max_by_k < 1) max_by_k = 1;
    if (want > max_by_k) want = max_by_k;
  }
  if (want > k_iters) want = k_iters;
  if (want < 1) want = 1;
  SplitPlan sp{want, false};
  if (want >= 2 && g_tune_cluster_max >= 2 && nt <= 128) {
    int s = 2;
    while (s * 2 <= want && s * 2 <= g_tune_cluster_max) s *= 2;
    // heuristic choice: the power of two nearest to the wanted factor (6 -> 8, 5 -> 4), keeping >= 4 k-iterations a split
    if (forced <= 0 && want * 2 >= s * 3 && s * 2 <= g_tune_cluster_max && s * 2 * 4 <= k_iters) s *= 2;
    // forced: only an exact power of two goes to the cluster path (tests pick the path by the number they force);
    // heuristic: the cluster path unless the GEMM wants far more splits than a cluster holds (early-layer wgrad)
    const bool take = forced > 0 ? (s == want) : (want * 2 <= g_tune_cluster_max * g_tune_cluster_reach);
    if (take) {
      sp.splits = s;
      sp.cluster = true;
      return sp;
    }
  }
  // workspace bound
  while (sp.splits > 1 && (size_t)tiles * sp.splits * 128 * nt * 4 > kConvWorkspaceBytes) --sp.splits;
  return sp;
}

// mode: 0 fprop  (in = X[N,H,W,Cin], w = W[Cout,3,3,Cin], out = Y[N,H,W,Cout])
//       1 dgrad  (in = dY[N,H,W,Cout], w = W, out = dX[N,H,W,Cin])
//       2 wgrad  (in = dY[N,H,W,Cout], in2 = X[N,H,W,Cin], out = dW[Cout,3,3,Cin])
// stat_acc (fprop, optional): [2][Cout] zeroed fixed-point accumulators that receive sum / sum of squares per channel.
// target_ctas: CTA count the split-K heuristic aims for in this call (0 = the tuned default); a data-gradient and a
// weight-gradient GEMM that run side by side on two streams each ask for one CTA per SM.
inline int conv_gemm_dispatch(int mode, const ConvGeom& g, const float* in, const float* in2, const float* w,
                              float* out, const float* bias, unsigned long long* stat_acc, float* workspace,
                              int nt_override, int splits_override, int target_ctas, cudaStream_t stream) {
  if (g.cin % 32 || g.cout % 32) return set_error(-13, "conv_gemm needs Cin, Cout multiples of 32 (got %d, %d)", g.cin, g.cout);
  if (g.w > 32 || (g.w & (g.w - 1)) || (g.h & (g.h - 1))) return set_error(-13, "conv_gemm needs power-of-two H, W with W <= 32 (got %d x %d)", g.h, g.w);
  const long long pixels = (long long)g.n * g.h * g.w;
  ConvGemmParams p{};
  p.H = g.h; p.W = g.w; p.NImg = g.n; p.CinW = g.cin;
  p.bias = bias; p.stat_acc = stat_acc; p.workspace = workspace;
  p.out = out;
#ifdef DPB_EXPERIMENTS
  p.debug_dump = g_conv_debug_dump;
#endif
  CUtensorMap mA, mB;
  dim3 grid;
  int nt;
  int rc;
  if (mode == kFprop || mode == kDgrad) {
    const int ck = (mode == kFprop) ? g.cin : g.cout;   // GEMM K channels
    const int cn = (mode == kFprop) ? g.cout : g.cin;   // GEMM N channels
    nt = nt_override > 0 ? nt_override : (cn % g_tune_wide_nt == 0 ? g_tune_wide_nt : (cn >= 128 ? 128 : 64));
    if (cn % nt) return set_error(-13, "N tile %d does not divide %d", nt, cn);
    int bw, bh, bn_;
    pixel_box(g.h, g.w, 128, &bw, &bh, &bn_);
    p.box_w = bw; p.box_h = bh; p.box_n = bn_;
    p.Ck = ck;
    p.M_total = (int)pixels; p.N_total = cn;
    p.k_iters = 9 * (ck / 32);
    p.out_ld = cn;
    const int m_tiles = (int)((pixels + 127) / 128);
    const int n_tiles = cn / nt;
    const SplitPlan sp = plan_splits(m_tiles * n_tiles, m_tiles, p.k_iters, nt, splits_override, target_ctas);
    p.splits = sp.splits;
    if (!sp.cluster && sp.splits > 1 && workspace == nullptr) return set_error(-13, "conv_gemm: split-K needs a workspace");
    {
      cuuint64_t dims[4] = {(cuuint64_t)ck, (cuuint64_t)g.w, (cuuint64_t)g.h, (cuuint64_t)g.n};
      cuuint64_t str[3] = {(cuuint64_t)ck * 4, (cuuint64_t)g.w * ck * 4, (cuuint64_t)g.h * g.w * ck * 4};
      cuuint32_t box[4] = {32, (cuuint32_t)bw, (cuuint32_t)bh, (cuuint32_t)bn_};
      if ((rc = make_map_f32(&mA, in, 4, dims, str, box))) return rc;
    }
    if (mode == kFprop) {
      cuuint64_t dims[2] = {(cuuint64_t)9 * g.cin, (cuuint64_t)g.cout};
      cuuint64_t str[1] = {(cuuint64_t)9 * g.cin * 4};
      cuuint32_t box[2] = {32, (cuuint32_t)nt};
      if ((rc = make_map_f32(&mB, w, 2, dims, str, box))) return rc;
    } else {
      cuuint64_t dims[2] = {(cuuint64_t)9 * g.cin, (cuuint64_t)g.cout};
      cuuint64_t str[1] = {(cuuint64_t)9 * g.cin * 4};
      cuuint32_t box[2] = {32, 32};
      if ((rc = make_map_f32(&mB, w, 2, dims, str, box, true))) return rc;
    }
    grid = dim3(m_tiles, n_tiles, p.splits);
    if (mode == kFprop) return launch_conv_gemm_mode<kFprop>(nt, sp.cluster, mA, mB, p, grid, stream);
    return launch_conv_gemm_mode<kDgrad>(nt, sp.cluster, mA, mB, p, grid, stream);
  }
  // wgrad
  nt = nt_override > 0 ? nt_override : (g.cin % g_tune_wide_nt == 0 ? g_tune_wide_nt : (g.cin >= 128 ? 128 : 64));
  if (g.cin % nt) return set_error(-13, "N tile %d does not divide Cin %d", nt, g.cin);
  int bw, bh, bn_;
  pixel_box(g.h, g.w, 32, &bw, &bh, &bn_);
  p.box_w = bw; p.box_h = bh; p.box_n = bn_;
  p.Ck = 0;
  p.M_total = g.cout; p.N_total = g.cin;
  p.k_iters = (int)((pixels + 31) / 32);
  p.out_ld = (long long)9 * g.cin;
  const int m_tiles = (g.cout + 127) / 128;
  const int n_tiles = g.cin / nt;
  const SplitPlan sp = plan_splits(m_tiles * n_tiles * 9, m_tiles, p.k_iters, nt, splits_override, target_ctas);
  p.splits = sp.splits;
  if (!sp.cluster && sp.splits > 1 && workspace == nullptr) return set_error(-13, "conv_gemm: split-K needs a workspace");
  {
    cuuint64_t dims[2] = {(cuuint64_t)g.cout, (cuuint64_t)pixels};
    cuuint64_t str[1] = {(cuuint64_t)g.cout * 4};
    cuuint32_t box[2] = {32, 32};
    if ((rc = make_map_f32(&mA, in, 2, dims, str, box, true))) return rc;
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)g.cin, (cuuint64_t)g.w, (cuuint64_t)g.h, (cuuint64_t)g.n};
    cuuint64_t str[3] = {(cuuint64_t)g.cin * 4, (cuuint64_t)g.w * g.cin * 4, (cuuint64_t)g.h * g.w * g.cin * 4};
    cuuint32_t box[4] = {32, (cuuint32_t)bw, (cuuint32_t)bh, (cuuint32_t)bn_};
    if ((rc = make_map_f32(&mB, in2, 4, dims, str, box, true))) return rc;
  }
  grid = dim3(m_tiles, n_tiles, 9 * p.splits);
  return launch_conv_gemm_mode<kWgrad>(nt, sp.cluster, mA, mB, p, grid, stream);
}

}  // namespace dpb
