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
// Split-K: every split writes its fp32 partial tile to a workspace; splitk_reduce_kernel then sums the partials in split
// order (deterministic) with the whole GPU, applies the bias and emits the per-channel sum / sum-of-squares partials that
// the BatchNorm finalize kernel consumes.
#pragma once
#include <cuda.h>
#include "ptx.cuh"

namespace dpb {

enum ConvGemmMode : int { kFprop = 0, kDgrad = 1, kWgrad = 2 };

struct ConvGemmParams {
  int H, W, NImg;        // activation geometry (input == output spatial size)
  int Ck;                // channels along GEMM-K for fprop/dgrad (Cin for fprop, Cout for dgrad); unused for wgrad
  int CinW;              // Cin of the weight tensor (row pitch of KRSC = 9*CinW)
  int M_total, N_total;  // GEMM extents: fprop/dgrad M = pixels; wgrad M = Cout.  N = channels of the output columns
  int k_iters;           // total K iterations of 32: fprop/dgrad 9*Ck/32; wgrad ceil(pixels/32)
  int splits;            // split-K factor (>=1)
  int pair_reduce;       // splits == 2 only: the CTA that finishes second adds its partner's partial tile itself and writes
                         // the final output (a + b == b + a, so the result does not depend on who is second); no reduce kernel
  int box_w, box_h, box_n;  // pixel box of one A tile (fprop/dgrad: 128 pixels) or one K chunk (wgrad: 32 pixels)
  float* out;            // fprop: Y[pix][Cout]; dgrad: dX[pix][Cin]; wgrad: dW[Cout][9][Cin]
  long long out_ld;      // row pitch of `out` in floats
  const float* bias;     // fprop only (may be null)
  float* stat_sum;       // fprop only: [m_tiles][N_total] partial sums (may be null)
  float* stat_sq;        // fprop only: [m_tiles][N_total]
  float* workspace;      // split-K partials: [tiles][splits][128*NT]
  unsigned int* counters;  // split-K arrival counters, one per tile, self-resetting
  float* debug_dump;       // bring-up aid: CTA (0,0,0) copies its first stage (A then B, as laid out in smem) here
};

template <int NT, int STAGES, int G = 1>
struct ConvGemmSmem {
  static constexpr int kABytes = 128 * 128;  // 128 rows x 32 tf32
  static constexpr int kBBytes = NT * 128 / G;  // CTA pair (G = 2): each CTA holds half of the B tile
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
  static constexpr int kColredBytes = 2 * NT * 4 * 4;
  static constexpr bool kColredAliased = kStagingBytes + kColredBytes <= STAGES * kStageBytes;
  static constexpr int kColredOffset = kColredAliased ? kStagingBytes : STAGES * kStageBytes + kBarBytes;
  static constexpr int kTotal = 1024 /*align slack*/ + STAGES * kStageBytes + kBarBytes + (kColredAliased ? 0 : kColredBytes);
};

// CL > 1: thread-block cluster of CL CTAs along blockIdx.x (neighbouring M tiles, same N tile / tap / split).  They
// consume the same B tile, so each CTA fetches 1/CL of it and TMA-multicasts it into every CTA's shared memory: B's
// L2 -> SM traffic drops by CL.  A stage may be refilled only when all CL CTAs have consumed it, hence the empty
// barriers count CL arrivals and every MMA commit is multicast to the whole cluster.
//
// G == 2: CTA pair (tcgen05 cta_group::2).  Two CTAs of a 2-cluster (neighbouring M tiles) run ONE 256 x NT MMA per
// k-step: each CTA stages its own 128 A rows and only HALF of the B tile, the leader CTA (cluster rank 0) issues the MMA
// which reads both CTAs' shared memory and writes both CTAs' TMEM.  Per unit of tensor work each SM ingests 25 % (NT
// 128) to 50 % (NT 256 vs two 128 tiles) fewer operand bytes -- the quantity this kernel is bound by.  Both CTAs' TMA
// loads credit the leader's full barrier; the leader's commits are multicast to both CTAs' empty / accumulator barriers.
template <int MODE, int NT, int STAGES, int CL, int G = 1>
__global__ void __launch_bounds__(192, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 const ConvGemmParams p) {
  constexpr bool A_MN = (MODE == kWgrad);
  constexpr bool B_MN = (MODE != kFprop);
  using S = ConvGemmSmem<NT, STAGES, G>;
  static_assert(CL == 1 || CL == 2 || CL == 4, "cluster size");
  static_assert(G == 1 || (G == 2 && CL == 1), "CTA pairs do not combine with B multicast");
  static_assert((NT / 32) % (CL * G) == 0, "B tile must split evenly over the cluster");
  constexpr uint16_t kClusterMask = (uint16_t)((1u << CL) - 1u);
  const uint32_t crank = (CL > 1 || G == 2) ? cluster_ctarank() : 0u;
  const bool leader = (G == 1) || (crank == 0);

  pdl_launch_dependents();  // the next kernel may start its prologue; it waits for our completion itself
  // bring-up aid: with a debug buffer every CTA logs (globaltimer at entry, at exit, SM id) at word 12800 + 3 * cta
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
      mbar_init(&full_bar[s], G);    // pair: the leader's barrier collects both CTAs' producer arrivals
      mbar_init(&empty_bar[s], CL);  // one arrival per CTA of the cluster (each CTA's MMA commit is multicast)
    }
    mbar_init(acc_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    if (G == 2) tmem_alloc_pair<NT>(tmem_slot);
    else tmem_alloc<NT>(tmem_slot);
  }
  tc_fence_before();
  if (CL > 1 || G == 2) cluster_sync_all();  // peers must see initialised barriers before anything remote lands
  else __syncthreads();
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
        if constexpr (G == 2) {
          // CTA pair: my A rows and my half of B go to MY shared memory, the bytes are credited to the LEADER's barrier
          const uint32_t lbar = mapa_u32(&full_bar[s], 0);
          if (leader) mbar_arrive_expect_tx(&full_bar[s], 2 * S::kStageBytes);
          else mbar_arrive_cluster(lbar);
          constexpr int kHalfAtoms = NT / 32 / 2;
          if (MODE == kFprop || MODE == kDgrad) {
            const int tap = kk / cpt, cc = kk - tap * cpt;
            const int r = tap / 3, sx = tap - 3 * r;
            tma_load_4d_pair(sa, &mapA, lbar, cc * 32, a_w0 + sx - 1, a_h0 + r - 1, a_n0);
            if (MODE == kFprop) {  // mapB box: NT/2 weight rows
              tma_load_2d_pair(sb, &mapB, lbar, tap * p.CinW + cc * 32, n_tile * NT + crank * (NT / 2));
            } else {
#pragma unroll
              for (int j = 0; j < kHalfAtoms; ++j)
                tma_load_2d_pair(sb + j * 4096, &mapB, lbar,
                                 (8 - tap) * p.CinW + n_tile * NT + (crank * kHalfAtoms + j) * 32, cc * 32);
            }
          } else {
            const int r = wtap / 3, sx = wtap - 3 * r;
            const int p0 = kk * 32;
            const int w0 = p0 % p.W;
            const int h0 = (p0 / p.W) % p.H;
            const int n0 = p0 / (p.W * p.H);
#pragma unroll
            for (int a = 0; a < 4; ++a) tma_load_2d_pair(sa + a * 4096, &mapA, lbar, m_tile * 128 + a * 32, p0);
#pragma unroll
            for (int j = 0; j < kHalfAtoms; ++j)
              tma_load_4d_pair(sb + j * 4096, &mapB, lbar, n_tile * NT + (crank * kHalfAtoms + j) * 32, w0 + sx - 1,
                               h0 + r - 1, n0);
          }
          continue;
        }
        mbar_arrive_expect_tx(&full_bar[s], S::kStageBytes);
        if (MODE == kFprop) {
          const int tap = kk / cpt, cc = kk - tap * cpt;
          const int r = tap / 3, sx = tap - 3 * r;
          tma_load_4d(sa, &mapA, &full_bar[s], cc * 32, a_w0 + sx - 1, a_h0 + r - 1, a_n0);
          if (CL == 1) {
            tma_load_2d(sb, &mapB, &full_bar[s], tap * p.CinW + cc * 32, n_tile * NT);
          } else {  // my NT/CL weight rows, delivered to every CTA of the cluster (mapB's box is NT/CL rows)
            tma_load_2d_mc(sb + crank * (NT / CL) * 128, &mapB, &full_bar[s], tap * p.CinW + cc * 32,
                           n_tile * NT + crank * (NT / CL), kClusterMask);
          }
        } else if (MODE == kDgrad) {
          const int tap = kk / cpt, cc = kk - tap * cpt;
          const int r = tap / 3, sx = tap - 3 * r;
          tma_load_4d(sa, &mapA, &full_bar[s], cc * 32, a_w0 + sx - 1, a_h0 + r - 1, a_n0);
          // MN-major B: one 32(co = GEMM K) x 32(ci = GEMM N) box of the KRSC weights per 32-wide N atom, flipped tap
          if (CL == 1) {
#pragma unroll
            for (int a = 0; a < NT / 32; ++a)
              tma_load_2d(sb + a * 4096, &mapB, &full_bar[s], (8 - tap) * p.CinW + n_tile * NT + a * 32, cc * 32);
          } else {
#pragma unroll
            for (int j = 0; j < NT / 32 / CL; ++j) {
              const int a = crank * (NT / 32 / CL) + j;
              tma_load_2d_mc(sb + a * 4096, &mapB, &full_bar[s], (8 - tap) * p.CinW + n_tile * NT + a * 32, cc * 32,
                             kClusterMask);
            }
          }
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
          if (CL == 1) {
#pragma unroll
            for (int a = 0; a < NT / 32; ++a)
              tma_load_4d(sb + a * 4096, &mapB, &full_bar[s], n_tile * NT + a * 32, w0 + sx - 1, h0 + r - 1, n0);
          } else {
#pragma unroll
            for (int j = 0; j < NT / 32 / CL; ++j) {
              const int a = crank * (NT / 32 / CL) + j;
              tma_load_4d_mc(sb + a * 4096, &mapB, &full_bar[s], n_tile * NT + a * 32, w0 + sx - 1, h0 + r - 1, n0,
                             kClusterMask);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = umma_idesc_tf32(NT, A_MN, B_MN, G == 2 ? 256 : 128);
    // K-major (SWIZZLE_128B): 8-row groups 1024 B apart (SBO), LBO unused.
    // MN-major (SWIZZLE_128B_BASE32B): smem is [K rows][128 B of MN]; 4-row K groups 512 B apart (SBO), 32-element MN
    // atoms one 32-row block (4096 B) apart (LBO).
    constexpr uint32_t a_lbo = A_MN ? 4096 : 16, a_sbo = A_MN ? 512 : 1024, a_lt = A_MN ? 1 : 2;
    constexpr uint32_t b_lbo = B_MN ? 4096 : 16, b_sbo = B_MN ? 512 : 1024, b_lt = B_MN ? 1 : 2;
    constexpr uint32_t a_kstep = A_MN ? 1024 : 32;  // bytes per UMMA k-step (8 tf32)
    constexpr uint32_t b_kstep = B_MN ? 1024 : 32;
    // bring-up aid: with a debug buffer, CTA (gridDim.x/2, 0, 0) reports where its issuer thread spends its cycles
    const bool timing = p.debug_dump != nullptr && blockIdx.x == gridDim.x / 2 && blockIdx.y == 0 && blockIdx.z == 0;
    long long t_wait = 0, t_begin = timing ? clock64() : 0;
    for (int i = 0; i < (leader ? k_count : 0); ++i) {  // pair: only the leader CTA issues
      const int s = i % STAGES;
      const uint32_t ph = (i / STAGES) & 1;
      const long long tw0 = timing ? clock64() : 0;
      mbar_wait(&full_bar[s], ph, 200 + s);
      if (timing) t_wait += clock64() - tw0;
      tc_fence_after();
      if (p.debug_dump != nullptr && i == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
        const float* src = reinterpret_cast<const float*>(stage_base);  // bring-up aid: raw stage 0 (A then B)
        for (int j = lane; j < S::kStageBytes / 4; j += 32) p.debug_dump[j] = src[j];
        __syncwarp();
      }
      if (elect_one()) {
        const uint32_t sa = smem_u32(stage_base + s * S::kStageBytes);
        const uint32_t sb = sa + S::kABytes;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t ad = umma_smem_desc(sa + ks * a_kstep, a_lbo, a_sbo, a_lt);
          const uint64_t bd = umma_smem_desc(sb + ks * b_kstep, b_lbo, b_sbo, b_lt);
          if (G == 2) umma_tf32_pair(tmem_acc, ad, bd, idesc, (i > 0 || ks > 0) ? 1u : 0u);
          else umma_tf32(tmem_acc, ad, bd, idesc, (i > 0 || ks > 0) ? 1u : 0u);
        }
        // frees the smem slot when these MMAs retire (in every CTA of the cluster: peers write our B slice too)
        if (G == 2) umma_commit_pair_mc(&empty_bar[s], 3);
        else if (CL == 1) umma_commit(&empty_bar[s]);
        else umma_commit_mc(&empty_bar[s], kClusterMask);
        if (i == k_count - 1) {  // accumulator complete (in both CTAs of a pair)
          if (G == 2) umma_commit_pair_mc(acc_bar, 3);
          else umma_commit(acc_bar);
        }
      }
      __syncwarp();
    }
    if (timing && lane == 0) {  // [12288..): cycles waiting for operands, cycles in the main loop, k iterations
      p.debug_dump[12288] = (float)t_wait;
      p.debug_dump[12289] = (float)(clock64() - t_begin);
      p.debug_dump[12290] = (float)k_count;
      p.debug_dump[12293] = (float)(t_begin - c_entry);  // prologue: barriers, TMEM allocation, first block barrier
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (2..5)
    pdl_wait();  // outputs / workspace may still be in use by the previous kernel
    const long long e_begin = clock64();
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    float* tile = staging + q * (32 * NT);
    const int row_base = m_tile * 128 + q * 32;
    const int tile_id = ((MODE == kWgrad ? wtap : 0) * gridDim.y + n_tile) * gridDim.x + m_tile;
    float* ws = nullptr;
    if (p.splits > 1) ws = p.workspace + ((size_t)tile_id * p.splits + split) * (size_t)(128 * NT);

    if (k_count > 0) {
      mbar_wait(acc_bar, 0, 300);
      tc_fence_after();
    }
    const long long e_acc = clock64();
    // TMEM -> registers (lane = accumulator row, 32 columns per load) -> this warp's 32 x NT slab, row-major with the
    // 16-byte chunk index XORed by (row & 7): a quarter warp's STS.128 (8 rows, same logical chunk) hits 8 distinct
    // bank groups, and so does a row-wise LDS.128 below.
    constexpr int kChunks = NT / 4;  // 16-byte chunks per tile row
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
    // slab -> global: every store instruction writes whole rows (512 B of one row at NT = 128, two rows at NT = 64)
    constexpr int kIters = 32 * kChunks / 32;                // float4 per lane
    constexpr int kSets = kChunks > 32 ? kChunks / 32 : 1;   // distinct column groups a lane meets (NT = 256: two)
    float4 csum[kSets], csq[kSets];
#pragma unroll
    for (int s = 0; s < kSets; ++s) csum[s] = csq[s] = make_float4(0.f, 0.f, 0.f, 0.f);
    bool finalize = p.splits == 1;  // this CTA writes the finished output tile (uniform over the epilogue warps)
    const float4* partner = nullptr;
    if (p.splits > 1) {
      float4* dst = reinterpret_cast<float4*>(ws + (size_t)(q * 32) * NT);
#pragma unroll 8
      for (int it = 0; it < kIters; ++it) {
        const int idx = it * 32 + lane, row = idx / kChunks, chunk = idx % kChunks;
        dst[idx] = *reinterpret_cast<const float4*>(tile + row * NT + ((chunk ^ (row & 7)) << 2));
      }
      if (p.pair_reduce) {
        // two splits: whoever takes the second ticket of this tile finds the partner's partial complete in the workspace
        // (L2) and finishes the tile from its own slab (still in shared memory) + that partial
        uint32_t* ticket = tmem_slot + 1;
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 64) {
          const unsigned old = atomicAdd(&p.counters[tile_id], 1u);
          if (old == 1u) p.counters[tile_id] = 0u;  // self-resetting for the next launch
          *ticket = old;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        finalize = *ticket == 1u;
        if (finalize) {
          __threadfence();
          partner = reinterpret_cast<const float4*>(p.workspace + ((size_t)tile_id * 2 + (1 - split)) * (size_t)(128 * NT) +
                                                    (size_t)(q * 32) * NT);
        }
      }
    }
    if (finalize) {
      const bool stats = MODE == kFprop && p.stat_sum != nullptr;
      float* obase = p.out + (size_t)row_base * p.out_ld + (size_t)(MODE == kWgrad ? wtap * p.CinW : 0) + n_tile * NT;
#pragma unroll 8
      for (int it = 0; it < kIters; ++it) {
        const int idx = it * 32 + lane, row = idx / kChunks, chunk = idx % kChunks;
        float4 y = *reinterpret_cast<const float4*>(tile + row * NT + ((chunk ^ (row & 7)) << 2));
        if (partner != nullptr) {
          const float4 o = __ldcg(partner + idx);
          y.x += o.x; y.y += o.y; y.z += o.z; y.w += o.w;
        }
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
      }
    }
    tc_fence_before();
    // named barrier over the 128 epilogue threads
    asm volatile("bar.sync 1, 128;" ::: "memory");
    const int et = threadIdx.x - 64;  // 0..127
    if (finalize) {
      if (MODE == kFprop && p.stat_sum != nullptr) {
        for (int c = et; c < NT; c += 128) {
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            s0 += colred[(w * 2 + 0) * NT + c];
            s1 += colred[(w * 2 + 1) * NT + c];
          }
          p.stat_sum[(size_t)m_tile * p.N_total + n_tile * NT + c] = s0;
          p.stat_sq[(size_t)m_tile * p.N_total + n_tile * NT + c] = s1;
        }
      }
    }
    // split-K: the partial tile is in the workspace; splitk_reduce_kernel (next launch) finishes the job
    if (p.debug_dump != nullptr && blockIdx.x == gridDim.x / 2 && blockIdx.y == 0 && blockIdx.z == 0 &&
        threadIdx.x == 64) {  // bring-up aid: cycles until the accumulator was ready, cycles of the epilogue
      p.debug_dump[12291] = (float)(e_acc - e_begin);
      p.debug_dump[12292] = (float)(clock64() - e_acc);
    }
  }

  __syncwarp();  // re-converge the single-lane producer / issuer warps before the aligned barrier
  tc_fence_before();
  // a CTA must not retire while a peer's commit can still arrive on its barriers
  if (CL > 1 || G == 2) cluster_sync_all();
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (G == 2) tmem_dealloc_pair<NT>(tmem_acc);
    else tmem_dealloc<NT>(tmem_acc);
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
}

}  // namespace dpb
