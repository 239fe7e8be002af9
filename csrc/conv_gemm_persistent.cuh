// Persistent variant of the tcgen05 implicit-GEMM convolution (same math, same operand layouts and TMA maps as
// conv_gemm.cuh, same split-K workspace protocol).
//
// Why: tools/conv_timing.py showed that a CTA of conv_gemm_kernel turns one 32 KB stage around every ~570 cycles (a
// tcgen05.commit -> mbarrier round trip is ~430 cycles, a TMA stage ~1000, and it has 3 stages) against 256 cycles of
// MMA work, and that its epilogue (store-bound, ~5 k cycles) and prologue run with the tensor pipe idle because the two
// co-resident CTAs of an SM move in lock-step.  Here ONE CTA per SM owns all of the SM's shared memory (6 stages: the
// same bytes in flight as two 3-stage CTAs), walks its work units (output tile x split-K slice [x tap]) in a static
// round-robin order, and double-buffers the accumulator in TMEM: while the four epilogue warps drain unit i
// (tcgen05.ld -> transpose tile -> global), the MMA warp is already accumulating unit i+1 into the other 128 TMEM
// columns and the TMA warp is streaming its operands, so only the last epilogue of a CTA is exposed.
//
// Barriers: full[s] / empty[s] per stage (TMA <-> MMA, running across unit boundaries), acc_full[b] (MMA -> epilogue,
// one tcgen05.commit per unit), acc_empty[b] (epilogue -> MMA, one arrival per epilogue warp after its last
// tcgen05.ld of the unit).
#pragma once
#include "conv_gemm.cuh"

namespace dpb {

template <int NT, int STAGES>
struct ConvPersistSmem {
  static constexpr int kABytes = 128 * 128;
  static constexpr int kBBytes = NT * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagingBytes = 4 * 32 * 33 * 4;  // per-warp 32 x 33 transpose tiles (the stages stay busy)
  static constexpr int kBarBytes = 256;
  static constexpr int kColredBytes = 2 * NT * 4 * 4;
  static constexpr int kTotal = 1024 + STAGES * kStageBytes + kStagingBytes + kBarBytes + kColredBytes;
  static_assert(kTotal <= 232448, "exceeds the 227 KB a CTA may own");
  static_assert((2 * STAGES + 4) * 8 + 8 <= kBarBytes, "barrier block too small");
};

struct ConvUnit {
  int m_tile, n_tile, split, wtap, k_begin, k_count, tile_id;
};

// grid_x / grid_y / grid_z: the launch grid the one-unit-per-CTA kernel would use (m tiles, n tiles, splits [x 9 taps])
template <int MODE>
__device__ __forceinline__ ConvUnit conv_unit(const ConvGemmParams& p, int u, int grid_x, int grid_y) {
  ConvUnit c;
  c.m_tile = u % grid_x;
  const int t = u / grid_x;
  c.n_tile = t % grid_y;
  const int z = t / grid_y;
  c.split = (MODE == kWgrad) ? (z % p.splits) : z;
  c.wtap = (MODE == kWgrad) ? (z / p.splits) : 0;
  const int kq = p.k_iters / p.splits, kr = p.k_iters % p.splits;
  c.k_begin = c.split * kq + min(c.split, kr);
  c.k_count = kq + (c.split < kr ? 1 : 0);
  c.tile_id = (c.wtap * grid_y + c.n_tile) * grid_x + c.m_tile;
  return c;
}

template <int MODE, int NT, int STAGES>
__global__ void __launch_bounds__(192, 1)
conv_gemm_persistent_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                            const ConvGemmParams p, const int grid_x, const int grid_y, const int units) {
  constexpr bool A_MN = (MODE == kWgrad);
  constexpr bool B_MN = (MODE != kFprop);
  using S = ConvPersistSmem<NT, STAGES>;
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_align_1024(smem_raw);
  uint8_t* stage_base = smem;
  float* staging = reinterpret_cast<float*>(smem + STAGES * S::kStageBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * S::kStageBytes + S::kStagingBytes);
  uint64_t* full_bar = bars;                    // [STAGES]
  uint64_t* empty_bar = bars + STAGES;          // [STAGES]
  uint64_t* acc_full = bars + 2 * STAGES;       // [2]
  uint64_t* acc_empty = bars + 2 * STAGES + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  float* colred = reinterpret_cast<float*>(smem + STAGES * S::kStageBytes + S::kStagingBytes + S::kBarBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int my_units = ((int)blockIdx.x < units) ? (units - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapA);
    tma_prefetch_desc(&mapB);
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&acc_full[0], 1);
    mbar_init(&acc_full[1], 1);
    mbar_init(&acc_empty[0], 4);  // one arrival per epilogue warp
    mbar_init(&acc_empty[1], 4);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<2 * NT>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      pdl_wait();
      const int cpt = (MODE != kWgrad) ? (p.Ck >> 5) : 1;
      int it = 0;  // stage counter, runs across units
      for (int i = 0; i < my_units; ++i) {
        const ConvUnit c = conv_unit<MODE>(p, blockIdx.x + i * gridDim.x, grid_x, grid_y);
        int a_w0 = 0, a_h0 = 0, a_n0 = 0;
        if (MODE != kWgrad) {
          const int tiles_w = p.W / p.box_w, tiles_h = p.H / p.box_h;
          int t = c.m_tile;
          a_w0 = (t % tiles_w) * p.box_w; t /= tiles_w;
          a_h0 = (t % tiles_h) * p.box_h; t /= tiles_h;
          a_n0 = t * p.box_n;
        }
        for (int k = 0; k < c.k_count; ++k, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&empty_bar[s], ph ^ 1, 400 + s);
          uint8_t* sa = stage_base + s * S::kStageBytes;
          uint8_t* sb = sa + S::kABytes;
          const int kk = c.k_begin + k;
          mbar_arrive_expect_tx(&full_bar[s], S::kStageBytes);
          if (MODE == kFprop) {
            const int tap = kk / cpt, cc = kk - tap * cpt;
            const int r = tap / 3, sx = tap - 3 * r;
            tma_load_4d(sa, &mapA, &full_bar[s], cc * 32, a_w0 + sx - 1, a_h0 + r - 1, a_n0);
            tma_load_2d(sb, &mapB, &full_bar[s], tap * p.CinW + cc * 32, c.n_tile * NT);
          } else if (MODE == kDgrad) {
            const int tap = kk / cpt, cc = kk - tap * cpt;
            const int r = tap / 3, sx = tap - 3 * r;
            tma_load_4d(sa, &mapA, &full_bar[s], cc * 32, a_w0 + sx - 1, a_h0 + r - 1, a_n0);
#pragma unroll
            for (int a = 0; a < NT / 32; ++a)
              tma_load_2d(sb + a * 4096, &mapB, &full_bar[s], (8 - tap) * p.CinW + c.n_tile * NT + a * 32, cc * 32);
          } else {
            const int r = c.wtap / 3, sx = c.wtap - 3 * r;
            const int p0 = kk * 32;
            const int w0 = p0 % p.W;
            const int h0 = (p0 / p.W) % p.H;
            const int n0 = p0 / (p.W * p.H);
#pragma unroll
            for (int a = 0; a < 4; ++a) tma_load_2d(sa + a * 4096, &mapA, &full_bar[s], c.m_tile * 128 + a * 32, p0);
#pragma unroll
            for (int a = 0; a < NT / 32; ++a)
              tma_load_4d(sb + a * 4096, &mapB, &full_bar[s], c.n_tile * NT + a * 32, w0 + sx - 1, h0 + r - 1, n0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = umma_idesc_tf32(NT, A_MN, B_MN);
    constexpr uint32_t a_lbo = A_MN ? 4096 : 16, a_sbo = A_MN ? 512 : 1024, a_lt = A_MN ? 1 : 2;
    constexpr uint32_t b_lbo = B_MN ? 4096 : 16, b_sbo = B_MN ? 512 : 1024, b_lt = B_MN ? 1 : 2;
    constexpr uint32_t a_kstep = A_MN ? 1024 : 32, b_kstep = B_MN ? 1024 : 32;
    int it = 0;
    // bring-up aid (tools/conv_timing.py): the middle CTA reports where its issuer thread spends its cycles
    const bool timing = p.debug_dump != nullptr && blockIdx.x == gridDim.x / 2;
    long long t_full = 0, t_acc = 0;
    const long long t_begin = timing ? clock64() : 0;
    for (int i = 0; i < my_units; ++i) {
      const ConvUnit c = conv_unit<MODE>(p, blockIdx.x + i * gridDim.x, grid_x, grid_y);
      const int b = i & 1;
      const uint32_t tmem_acc = tmem_base + (uint32_t)(b * NT);
      if (c.k_count > 0) {
        const long long t0 = timing ? clock64() : 0;
        mbar_wait(&acc_empty[b], ((i >> 1) & 1) ^ 1, 410 + b);  // the epilogue has drained this accumulator
        if (timing) t_acc += clock64() - t0;
        tc_fence_after();
      }
      for (int k = 0; k < c.k_count; ++k, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        const long long t0 = timing ? clock64() : 0;
        mbar_wait(&full_bar[s], ph, 420 + s);
        if (timing) t_full += clock64() - t0;
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(stage_base + s * S::kStageBytes);
          const uint32_t sb = sa + S::kABytes;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            umma_tf32(tmem_acc, umma_smem_desc(sa + ks * a_kstep, a_lbo, a_sbo, a_lt),
                      umma_smem_desc(sb + ks * b_kstep, b_lbo, b_sbo, b_lt), idesc, (k > 0 || ks > 0) ? 1u : 0u);
          umma_commit(&empty_bar[s]);
          if (k == c.k_count - 1) umma_commit(&acc_full[b]);
        }
        __syncwarp();
      }
    }
    if (timing && lane == 0) {
      p.debug_dump[12288] = (float)t_full;                  // cycles waiting for operands
      p.debug_dump[12289] = (float)(clock64() - t_begin);   // cycles from the first wait to the last issue
      p.debug_dump[12290] = (float)it;                      // K iterations issued
      p.debug_dump[12293] = (float)t_acc;                   // cycles waiting for a free accumulator
      p.debug_dump[12294] = (float)my_units;
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (2..5)
    pdl_wait();
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    float* tile = staging + q * (32 * 33);
    const int et = threadIdx.x - 64;  // 0..127
    const bool timing = p.debug_dump != nullptr && blockIdx.x == gridDim.x / 2 && threadIdx.x == 64;
    long long e_wait = 0;
    const long long e_begin = timing ? clock64() : 0;
    for (int i = 0; i < my_units; ++i) {
      const ConvUnit c = conv_unit<MODE>(p, blockIdx.x + i * gridDim.x, grid_x, grid_y);
      const int b = i & 1;
      const uint32_t tmem_acc = tmem_base + (uint32_t)(b * NT);
      const int row_base = c.m_tile * 128 + q * 32;
      float* ws = nullptr;
      if (p.splits > 1) ws = p.workspace + ((size_t)c.tile_id * p.splits + c.split) * (size_t)(128 * NT);
      if (c.k_count > 0) {
        const long long t0 = timing ? clock64() : 0;
        mbar_wait(&acc_full[b], (i >> 1) & 1, 430 + b);
        if (timing) e_wait += clock64() - t0;
        tc_fence_after();
      }
      const bool stats = MODE == kFprop && p.stat_sum != nullptr && p.splits == 1;
#pragma unroll 1
      for (int c0 = 0; c0 < NT; c0 += 32) {
        float v[32];
        if (c.k_count > 0) {
          tmem_ld_32x32(tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
        }
        if (c0 + 32 >= NT && c.k_count > 0) {  // last read of this accumulator: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&acc_empty[b]);
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) tile[lane * 33 + j] = v[j];
        __syncwarp();
        const int col = c.n_tile * NT + c0 + lane;
        if (p.splits > 1) {
#pragma unroll 8
          for (int r = 0; r < 32; ++r) ws[(size_t)(q * 32 + r) * NT + c0 + lane] = tile[r * 33 + lane];
        } else {
          float bv = 0.f;
          if (MODE == kFprop && p.bias != nullptr) bv = __ldg(p.bias + col);
          float* orow = p.out + (size_t)row_base * p.out_ld + (size_t)(MODE == kWgrad ? c.wtap * p.CinW : 0) + col;
          float csum = 0.f, csq = 0.f;
#pragma unroll 8
          for (int r = 0; r < 32; ++r) {
            if (row_base + r < p.M_total) {
              const float y = tile[r * 33 + lane] + bv;
              orow[(size_t)r * p.out_ld] = y;
              csum += y;
              csq += y * y;
            }
          }
          if (stats) {
            colred[(q * 2 + 0) * NT + c0 + lane] = csum;
            colred[(q * 2 + 1) * NT + c0 + lane] = csq;
          }
        }
        __syncwarp();
      }
      if (stats) {
        asm volatile("bar.sync 1, 128;" ::: "memory");
        for (int cc = et; cc < NT; cc += 128) {
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            s0 += colred[(w * 2 + 0) * NT + cc];
            s1 += colred[(w * 2 + 1) * NT + cc];
          }
          p.stat_sum[(size_t)c.m_tile * p.N_total + c.n_tile * NT + cc] = s0;
          p.stat_sq[(size_t)c.m_tile * p.N_total + c.n_tile * NT + cc] = s1;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");  // colred is rewritten by the next unit
      }
    }
    if (timing) {
      p.debug_dump[12291] = (float)e_wait;                            // cycles waiting for finished accumulators
      p.debug_dump[12292] = (float)(clock64() - e_begin - e_wait);    // cycles draining them
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<2 * NT>(tmem_base);
  }
}

}  // namespace dpb
