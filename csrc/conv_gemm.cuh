// 3x3 / stride 1 / pad 1 convolution as an implicit GEMM on the 5th-gen tensor cores (tcgen05, kind::tf32, fp32
// accumulation in TMEM), operands staged by TMA.  One kernel template covers the three GEMMs of a conv layer:
//
//   FPROP  D[pix, co] = sum_{tap,ci} X[pix+tap, ci] * W[co, tap, ci]      A: K-major (4-D TMA box, halo = OOB zero fill)
//                                                                         B: K-major (weights are KRSC, ci contiguous)
//   DGRAD  D[pix, ci] = sum_{tap,co} dY[pix+tap, co] * W[co, 8-tap, ci]   A: K-major, B: MN-major (same KRSC weights,
//                                                                         no repacked copy)
//   WGRAD  D[co, ci]  = sum_pix dY[pix, co] * X[pix+tap, ci]  (per tap)   A: MN-major, B: MN-major, K = pixels
//
// Replaces what the reference reaches through aten::(mkldnn_)convolution / convolution_backward
// (reference model.py:18-23).
//
// Tile: 128 (M) x NT (N) per CTA, K consumed in chunks of 32 tf32 (one 128-byte swizzle row), 4 UMMA k-steps of 8.
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + single-thread MMA issuer, warps 2..5 =
// epilogue (TMEM -> registers -> row-major XOR-swizzled tile in the drained stages -> whole 512-byte global rows).
//
// Split-K, two ways:
//   * CLUSTER = true: the `splits` CTAs of one output tile form a thread-block cluster along blockIdx.z.  Every CTA
//     stages its fp32 partial tile in its own shared memory; after a cluster barrier CTA r sums rows
//     [r * 128/S, (r+1) * 128/S) of all S partial tiles over distributed shared memory -- in split order, so the
//     result is deterministic -- adds the bias and writes those rows of the output.  No workspace, no second kernel.
//   * CLUSTER = false, splits > 1: every split writes its partial tile to a global workspace; splitk_reduce_kernel
//     (next launch) sums them.  Used when a GEMM needs more splits than a cluster can hold (early-layer wgrad).
// BatchNorm batch statistics: whichever CTA finishes output rows (unsplit, cluster, or the reduce kernel) adds its
// per-channel sum / sum of squares to fixed-point integer accumulators in global memory (bn_acc.cuh) -- integer
// addition is associative, so the totals are deterministic whatever the arrival order, and the consumer kernel
// (bn_relu_pool_fwd) derives scale / shift from them in its prologue: no partial rows, no finalize launch.
//
// TMA loads in this kernel read only this GPU's memory (activations, weights and gradients of the local replica); there
// is no cluster multicast of operand tiles.
#pragma once
#include <cuda.h>
#include <type_traits>
#include "ptx.cuh"
#include "bn_acc.cuh"

namespace dpb {

enum ConvGemmMode : int { kFprop = 0, kDgrad = 1, kWgrad = 2 };

struct ConvGemmParams {
  int H, W, NImg;        // activation geometry (input == output spatial size)
  int Ck;                // channels along GEMM-K for fprop/dgrad (Cin for fprop, Cout for dgrad); unused for wgrad
  int CinW;              // Cin of the weight tensor (row pitch of KRSC = 9*CinW)
  int M_total, N_total;  // GEMM extents: fprop/dgrad M = pixels; wgrad M = Cout.  N = channels of the output columns
  int k_iters;           // total K iterations of 32: fprop/dgrad 9*Ck/32; wgrad ceil(pixels/32)
  int splits;            // split-K factor (>=1); with CLUSTER it is the cluster size along z
  int box_w, box_h, box_n;  // pixel box of one A tile (fprop/dgrad: 128 pixels) or one K chunk (wgrad: 32 pixels)
  float* out;            // fprop: Y[pix][Cout]; dgrad: dX[pix][Cin]; wgrad: dW[Cout][9][Cin]
  long long out_ld;      // row pitch of `out` in floats
  const float* bias;     // fprop only (may be null)
  unsigned long long* stat_acc;  // fprop only (may be null): BatchNorm batch statistics, [2][N_total] fixed-point
                                 // accumulators (sum, sum of squares) that every finishing CTA adds its column sums to
  float* workspace;      // split-K partials of the workspace path: [tiles][splits][128*NT]
#ifdef DPB_EXPERIMENTS
  float* debug_dump;     // bring-up aid (tools/conv_timing.py): per-CTA globaltimer trace and issuer-thread cycle counts
#endif
};

template <int NT, int STAGES>
struct ConvGemmSmem {
  static constexpr int kABytes = 128 * 128;  // 128 rows x 32 tf32
  static constexpr int kBBytes = NT * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // The epilogue stages the whole 128 x NT fp32 tile row-major (one 32 x NT slab per warp, 16-byte chunks XOR-swizzled
  // by row so both the column-wise writes out of TMEM and the row-wise reads are bank-conflict free) and reuses the
  // pipeline stages for it: by the time the accumulator barrier fires every TMA load has landed and every MMA has
  // finished reading shared memory.  Keeping the footprint at ~100 KB (NT <= 128) lets two CTAs share an SM.
  static constexpr int kStagingBytes = 128 * NT * 4;
  static_assert(kStagingBytes <= STAGES * kStageBytes, "staging must fit in the pipeline stages");
  static constexpr int kBarBytes = 256;
  // per-warp column sums for the BatchNorm partials: also carved out of the drained stages (behind the staged tile)
  // whenever they are large enough -- with 3 x 24 KB stages that lets THREE NT = 64 CTAs share an SM
  // (the cluster reduce needs [128 / (NT/4)][2][NT] floats = 4 KB whatever NT is)
  static constexpr int kColredBytes = (2 * NT * 4 * 4) > 4096 ? (2 * NT * 4 * 4) : 4096;
  static constexpr bool kColredAliased = kStagingBytes + kColredBytes <= STAGES * kStageBytes;
  static constexpr int kColredOffset = kColredAliased ? kStagingBytes : STAGES * kStageBytes + kBarBytes;
  static constexpr int kTotal = 1024 /*align slack*/ + STAGES * kStageBytes + kBarBytes + (kColredAliased ? 0 : kColredBytes);
};

template <int MODE, int NT, int STAGES, bool CLUSTER>
__global__ void __launch_bounds__(192, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 const ConvGemmParams p) {
  constexpr bool A_MN = (MODE == kWgrad);
  constexpr bool B_MN = (MODE != kFprop);
  using S = ConvGemmSmem<NT, STAGES>;

  DPB_TRACE(1 + MODE);
  pdl_launch_dependents();  // the next kernel may start its prologue; it waits for our completion itself
#ifdef DPB_EXPERIMENTS
  const unsigned cta_linear = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  const bool trace_cta = p.debug_dump != nullptr && cta_linear < 1024u && threadIdx.x == 0;
  const long long c_entry = clock64();
  if (trace_cta) {
    unsigned long long t;
    unsigned sm;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
    reinterpret_cast<unsigned*>(p.debug_dump)[12800 + 3 * cta_linear] = (unsigned)t;
    reinterpret_cast<unsigned*>(p.debug_dump)[12800 + 3 * cta_linear + 2] = sm;
  }
#endif
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1024(smem_raw);
  uint8_t* stage_base = smem;
  float* staging = reinterpret_cast<float*>(smem);  // aliases stage 0, used only after the main loop has drained
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * S::kStageBytes);
  uint64_t* full_bar = bars;               // [STAGES]
  uint64_t* empty_bar = bars + STAGES;     // [STAGES]
  uint64_t* acc_bar = bars + 2 * STAGES;   // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 1);
  float* colred = reinterpret_cast<float*>(smem + S::kColredOffset);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int m_tile = blockIdx.x;
  const int n_tile = blockIdx.y;
  // blockIdx.z = tap * splits + split for wgrad, = split otherwise
  const int split = (MODE == kWgrad) ? (blockIdx.z % p.splits) : blockIdx.z;
  const int wtap = (MODE == kWgrad) ? (blockIdx.z / p.splits) : 0;

  // K range of this split (balanced, contiguous)
  const int kq = p.k_iters / p.splits, kr = p.k_iters % p.splits;
  const int k_begin = split * kq + min(split, kr);
  const int k_count = kq + (split < kr ? 1 : 0);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapA);
    tma_prefetch_desc(&mapB);
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(acc_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<NT>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      pdl_wait();  // inputs are produced by the previous kernel
      int a_w0 = 0, a_h0 = 0, a_n0 = 0;
      if (MODE != kWgrad) {
        // pixel box of this M tile
        const int tiles_w = p.W / p.box_w;  // box_w divides W by construction
        const int tiles_h = p.H / p.box_h;
        int t = m_tile;
        a_w0 = (t % tiles_w) * p.box_w; t /= tiles_w;
        a_h0 = (t % tiles_h) * p.box_h; t /= tiles_h;
        a_n0 = t * p.box_n;
      }
      const int cpt = (MODE != kWgrad) ? (p.Ck >> 5) : 1;  // 32-channel chunks per tap
      for (int i = 0; i < k_count; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1, 100 + s);
        uint8_t* sa = stage_base + s * S::kStageBytes;
        uint8_t* sb = sa + S::kABytes;
        const int kk = k_begin + i;
        mbar_arrive_expect_tx(&full_bar[s], S::kStageBytes);
        if (MODE == kFprop) {
          const int tap = kk / cpt, cc = kk - tap * cpt;
          const int r = tap / 3, sx = tap - 3 * r;
          tma_load_4d(sa, &mapA, &full_bar[s], cc * 32, a_w0 + sx - 1, a_h0 + r - 1, a_n0);
          tma_load_2d(sb, &mapB, &full_bar[s], tap * p.CinW + cc * 32, n_tile * NT);
        } else if (MODE == kDgrad) {
          const int tap = kk / cpt, cc = kk - tap * cpt;
          const int r = tap / 3, sx = tap - 3 * r;
          tma_load_4d(sa, &mapA, &full_bar[s], cc * 32, a_w0 + sx - 1, a_h0 + r - 1, a_n0);
          // MN-major B: one 32(co = GEMM K) x 32(ci = GEMM N) box of the KRSC weights per 32-wide N atom, flipped tap
#pragma unroll
          for (int a = 0; a < NT / 32; ++a)
            tma_load_2d(sb + a * 4096, &mapB, &full_bar[s], (8 - tap) * p.CinW + n_tile * NT + a * 32, cc * 32);
        } else {
          const int r = wtap / 3, sx = wtap - 3 * r;
          const int p0 = kk * 32;
          const int w0 = p0 % p.W;
          const int h0 = (p0 / p.W) % p.H;
          const int n0 = p0 / (p.W * p.H);
          // MN-major A: 32 pixels (GEMM K) x 32 co per M atom;  MN-major B: 32 shifted pixels x 32 ci per N atom
#pragma unroll
          for (int a = 0; a < 4; ++a)
            tma_load_2d(sa + a * 4096, &mapA, &full_bar[s], m_tile * 128 + a * 32, p0);
#pragma unroll
          for (int a = 0; a < NT / 32; ++a)
            tma_load_4d(sb + a * 4096, &mapB, &full_bar[s], n_tile * NT + a * 32, w0 + sx - 1, h0 + r - 1, n0);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = umma_idesc_tf32(NT, A_MN, B_MN, 128);
    // K-major (SWIZZLE_128B): 8-row groups 1024 B apart (SBO), LBO unused.
    // MN-major (SWIZZLE_128B_BASE32B): smem is [K rows][128 B of MN]; 4-row K groups 512 B apart (SBO), 32-element MN
    // atoms one 32-row block (4096 B) apart (LBO).
    constexpr uint32_t a_lbo = A_MN ? 4096 : 16, a_sbo = A_MN ? 512 : 1024, a_lt = A_MN ? 1 : 2;
    constexpr uint32_t b_lbo = B_MN ? 4096 : 16, b_sbo = B_MN ? 512 : 1024, b_lt = B_MN ? 1 : 2;
    constexpr uint32_t a_kstep = A_MN ? 1024 : 32;  // bytes per UMMA k-step (8 tf32)
    constexpr uint32_t b_kstep = B_MN ? 1024 : 32;
#ifdef DPB_EXPERIMENTS
    const bool timing = p.debug_dump != nullptr && blockIdx.x == gridDim.x / 2 && blockIdx.y == 0 && blockIdx.z == 0;
    long long t_wait = 0, t_begin = timing ? clock64() : 0;
#endif
    for (int i = 0; i < k_count; ++i) {
      const int s = i % STAGES;
      const uint32_t ph = (i / STAGES) & 1;
#ifdef DPB_EXPERIMENTS
      const long long tw0 = timing ? clock64() : 0;
#endif
      mbar_wait(&full_bar[s], ph, 200 + s);
#ifdef DPB_EXPERIMENTS
      if (timing) t_wait += clock64() - tw0;
#endif
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sa = smem_u32(stage_base + s * S::kStageBytes);
        const uint32_t sb = sa + S::kABytes;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t ad = umma_smem_desc(sa + ks * a_kstep, a_lbo, a_sbo, a_lt);
          const uint64_t bd = umma_smem_desc(sb + ks * b_kstep, b_lbo, b_sbo, b_lt);
          umma_tf32(tmem_acc, ad, bd, idesc, (i > 0 || ks > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);                    // frees the smem slot when these MMAs retire
        if (i == k_count - 1) umma_commit(acc_bar);    // accumulator complete
      }
      __syncwarp();
    }
#ifdef DPB_EXPERIMENTS
    if (timing && lane == 0) {  // [12288..): cycles waiting for operands, cycles in the main loop, k iterations
      p.debug_dump[12288] = (float)t_wait;
      p.debug_dump[12289] = (float)(clock64() - t_begin);
      p.debug_dump[12290] = (float)k_count;
      p.debug_dump[12293] = (float)(t_begin - c_entry);  // prologue: barriers, TMEM allocation, first block barrier
    }
#endif
  }

  // ---------------------------------------------------------------------- epilogue warps (2..5)
  constexpr int kChunks = NT / 4;  // 16-byte chunks per tile row
  const bool epi = warp >= 2;
  const int q = warp & 3;  // TMEM lane quarter this warp may read
  const int et = (int)threadIdx.x - 64;  // 0..127 inside the epilogue group
  float* tile = staging + q * (32 * NT);
  const int row_base = m_tile * 128 + q * 32;
  const int tile_id = ((MODE == kWgrad ? wtap : 0) * gridDim.y + n_tile) * gridDim.x + m_tile;
  const bool stats = MODE == kFprop && p.stat_acc != nullptr;
#ifdef DPB_EXPERIMENTS
  long long e_begin = 0, e_acc = 0;
#endif
  if (epi) {
    pdl_wait();  // outputs / workspace may still be in use by the previous kernel
#ifdef DPB_EXPERIMENTS
    e_begin = clock64();
#endif
    if (k_count > 0) {
      mbar_wait(acc_bar, 0, 300);
      tc_fence_after();
    }
#ifdef DPB_EXPERIMENTS
    e_acc = clock64();
#endif
    // TMEM -> registers (lane = accumulator row, 32 columns per load) -> this warp's 32 x NT slab, row-major with the
    // 16-byte chunk index XORed by (row & 7): a quarter warp's STS.128 (8 rows, same logical chunk) hits 8 distinct
    // bank groups, and so does a row-wise LDS.128 below.
#pragma unroll 1
    for (int c0 = 0; c0 < NT; c0 += 32) {
      float v[32];
      if (k_count > 0) {
        tmem_ld_32x32(tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.f;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int chunk = (c0 >> 2) + j;
        *reinterpret_cast<float4*>(tile + lane * NT + ((chunk ^ (lane & 7)) << 2)) =
            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
    }
    __syncwarp();

    if (!CLUSTER) {
      // slab -> global: every store instruction writes whole rows (512 B of one row at NT = 128, two rows at NT = 64)
      constexpr int kIters = 32 * kChunks / 32;                // float4 per lane
      constexpr int kSets = kChunks > 32 ? kChunks / 32 : 1;   // distinct column groups a lane meets (NT = 256: two)
      if (p.splits > 1) {
        float* ws = p.workspace + ((size_t)tile_id * p.splits + split) * (size_t)(128 * NT);
        float4* dst = reinterpret_cast<float4*>(ws + (size_t)(q * 32) * NT);
#pragma unroll 8
        for (int it = 0; it < kIters; ++it) {
          const int idx = it * 32 + lane, row = idx / kChunks, chunk = idx % kChunks;
          dst[idx] = *reinterpret_cast<const float4*>(tile + row * NT + ((chunk ^ (row & 7)) << 2));
        }
        // the partial tile is in the workspace; splitk_reduce_kernel (next launch) finishes the job
      } else {
        float4 csum[kSets], csq[kSets];
#pragma unroll
        for (int s = 0; s < kSets; ++s) csum[s] = csq[s] = make_float4(0.f, 0.f, 0.f, 0.f);
        float* obase = p.out + (size_t)row_base * p.out_ld + (size_t)(MODE == kWgrad ? wtap * p.CinW : 0) + n_tile * NT;
#pragma unroll 8
        for (int it = 0; it < kIters; ++it) {
          const int idx = it * 32 + lane, row = idx / kChunks, chunk = idx % kChunks;
          float4 y = *reinterpret_cast<const float4*>(tile + row * NT + ((chunk ^ (row & 7)) << 2));
          if (MODE == kFprop && p.bias != nullptr) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n_tile * NT) + chunk);
            y.x += b.x; y.y += b.y; y.z += b.z; y.w += b.w;
          }
          if (row_base + row < p.M_total) {
            *reinterpret_cast<float4*>(obase + (size_t)row * p.out_ld + (chunk << 2)) = y;
            if (stats) {
              float4& s = csum[it % kSets];
              float4& t = csq[it % kSets];
              s.x += y.x; s.y += y.y; s.z += y.z; s.w += y.w;
              t.x += y.x * y.x; t.y += y.y * y.y; t.z += y.z * y.z; t.w += y.w * y.w;
            }
          }
        }
        if (stats) {
          if (kChunks < 32) {  // NT = 64: lanes l and l + 16 hold the same columns (even / odd rows)
#pragma unroll
            for (int off = kChunks; off < 32; off <<= 1) {
              csum[0].x += __shfl_xor_sync(0xffffffffu, csum[0].x, off);
              csum[0].y += __shfl_xor_sync(0xffffffffu, csum[0].y, off);
              csum[0].z += __shfl_xor_sync(0xffffffffu, csum[0].z, off);
              csum[0].w += __shfl_xor_sync(0xffffffffu, csum[0].w, off);
              csq[0].x += __shfl_xor_sync(0xffffffffu, csq[0].x, off);
              csq[0].y += __shfl_xor_sync(0xffffffffu, csq[0].y, off);
              csq[0].z += __shfl_xor_sync(0xffffffffu, csq[0].z, off);
              csq[0].w += __shfl_xor_sync(0xffffffffu, csq[0].w, off);
            }
          }
#pragma unroll
          for (int s = 0; s < kSets; ++s) {
            const int chunk = s * 32 + lane;
            if (chunk < kChunks) {
              *reinterpret_cast<float4*>(colred + (q * 2 + 0) * NT + (chunk << 2)) = csum[s];
              *reinterpret_cast<float4*>(colred + (q * 2 + 1) * NT + (chunk << 2)) = csq[s];
            }
          }
          bar_sync_n(1, 128);
          for (int c = et; c < NT; c += 128) {
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int w = 0; w < 4; ++w) {
              s0 += colred[(w * 2 + 0) * NT + c];
              s1 += colred[(w * 2 + 1) * NT + c];
            }
            bn_acc_add(p.stat_acc + n_tile * NT + c, s0);
            bn_acc_add(p.stat_acc + p.N_total + n_tile * NT + c, s1);
          }
        }
      }
    }
    tc_fence_before();
  }

  if (CLUSTER) {
    // every CTA of the cluster has staged its partial tile: reduce over distributed shared memory
    __syncwarp();
    cluster_sync_all();
    if (epi) {
      const int S_ = p.splits;
      const int R = 128 / S_;                  // rows of the tile this CTA finishes
      const int crank = (int)cluster_ctarank();  // == split (cluster spans exactly the splits of one tile)
      constexpr int kRowsPerPass = 128 / kChunks;  // NT 128: 4 rows per pass of the 128 epilogue threads
      const int chunk = et % kChunks;
      const int rsub = et / kChunks;
      float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (MODE == kFprop && p.bias != nullptr) bias4 = __ldg(reinterpret_cast<const float4*>(p.bias + n_tile * NT) + chunk);
      float4 csum = make_float4(0.f, 0.f, 0.f, 0.f), csq = csum;
      float* obase = p.out + (size_t)(m_tile * 128) * p.out_ld + (size_t)(MODE == kWgrad ? wtap * p.CinW : 0) + n_tile * NT;
      // Row r of the finished tile = sum over the S staged partial tiles, in split order.  A distributed-shared-memory
      // load costs several hundred cycles, so 16 of them are in flight per thread before the first add (a dependent
      // load-add chain, and then one row pass at a time, made this loop the longest part of the kernel).
      auto reduce_rows = [&](auto s_const) {
        constexpr int SC = decltype(s_const)::value;
        constexpr int PB = 16 / SC;  // row passes per batch
        for (int r0 = 0; r0 < R; r0 += kRowsPerPass * PB) {
          float4 v[PB][SC];
          int trow[PB];
#pragma unroll
          for (int b = 0; b < PB; ++b) {
            const int rr = r0 + b * kRowsPerPass + rsub;
            trow[b] = rr < R ? crank * R + rr : -1;  // row of the 128-row tile
            const int tr = trow[b] < 0 ? 0 : trow[b];
            // address of (row, chunk) inside a staged tile: slab = row / 32, row in slab = row % 32
            const uint32_t local = smem_u32(staging + (tr >> 5) * (32 * NT) + (tr & 31) * NT + ((chunk ^ (tr & 7)) << 2));
#pragma unroll
            for (int s = 0; s < SC; ++s) v[b][s] = ld_dsmem_f4(mapa_u32_addr(local, (uint32_t)s));
          }
#pragma unroll
          for (int b = 0; b < PB; ++b) {
            float4 acc = bias4;
#pragma unroll
            for (int s = 0; s < SC; ++s) { acc.x += v[b][s].x; acc.y += v[b][s].y; acc.z += v[b][s].z; acc.w += v[b][s].w; }
            if (trow[b] >= 0 && m_tile * 128 + trow[b] < p.M_total) {
              *reinterpret_cast<float4*>(obase + (size_t)trow[b] * p.out_ld + (chunk << 2)) = acc;
              if (stats) {
                csum.x += acc.x; csum.y += acc.y; csum.z += acc.z; csum.w += acc.w;
                csq.x += acc.x * acc.x; csq.y += acc.y * acc.y; csq.z += acc.z * acc.z; csq.w += acc.w * acc.w;
              }
            }
          }
        }
      };
      switch (S_) {
        case 2: reduce_rows(std::integral_constant<int, 2>{}); break;
        case 4: reduce_rows(std::integral_constant<int, 4>{}); break;
        case 8: reduce_rows(std::integral_constant<int, 8>{}); break;
        default: reduce_rows(std::integral_constant<int, 16>{}); break;
      }
      if (stats) {
        // combine the kRowsPerPass threads that share a chunk, in rsub order; colred is [kRowsPerPass][2][NT] here
        static_assert(kRowsPerPass * 2 * NT * 4 <= S::kColredBytes, "colred too small for the cluster reduce");
        *reinterpret_cast<float4*>(colred + (rsub * 2 + 0) * NT + (chunk << 2)) = csum;
        *reinterpret_cast<float4*>(colred + (rsub * 2 + 1) * NT + (chunk << 2)) = csq;
        bar_sync_n(1, 128);
        for (int c = et; c < NT; c += 128) {
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int w = 0; w < kRowsPerPass; ++w) {
            s0 += colred[(w * 2 + 0) * NT + c];
            s1 += colred[(w * 2 + 1) * NT + c];
          }
          bn_acc_add(p.stat_acc + n_tile * NT + c, s0);
          bn_acc_add(p.stat_acc + p.N_total + n_tile * NT + c, s1);
        }
      }
    }
    // no CTA may reuse or give up its shared memory while a peer is still reading it
    __syncwarp();
    cluster_sync_all();
  } else {
    __syncwarp();  // re-converge the single-lane producer / issuer warps before the aligned barrier
    __syncthreads();
  }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<NT>(tmem_acc);
  }
#ifdef DPB_EXPERIMENTS
  if (p.debug_dump != nullptr && blockIdx.x == gridDim.x / 2 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 64) {
    p.debug_dump[12291] = (float)(e_acc - e_begin);      // cycles until the accumulator was ready
    p.debug_dump[12292] = (float)(clock64() - e_acc);    // cycles of the epilogue
  }
  if (trace_cta) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    reinterpret_cast<unsigned*>(p.debug_dump)[12800 + 3 * cta_linear + 1] = (unsigned)t;
    if (blockIdx.x == gridDim.x / 2 && blockIdx.y == 0 && blockIdx.z == 0) {
      p.debug_dump[12294] = (float)(clock64() - c_entry);  // SM cycles entry -> exit of the timed CTA
      p.debug_dump[12295] = (float)(unsigned)(t - (unsigned long long)reinterpret_cast<unsigned*>(p.debug_dump)[12800 + 3 * cta_linear]);
    }
  }
#endif
}

}  // namespace dpb
