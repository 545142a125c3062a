// Res2NetBlock (reference: src/models/ECAPA_TDNN.py:72-129) as ONE persistent kernel per SE-Res2Net
// block instead of 7 dependent GEMM launches:
//
//   y_1 = f_1(x_1),  y_j = f_j(x_j + y_{j-1})  (j = 2..7),   f_j = BN(ReLU(conv1d_k3_dilated_reflect(.)))
//
// where x_j is channel chunk j (Cs = C/8 channels) of the tdnn1 output U and y_j goes to chunk j of R.
// The 7 steps are strictly sequential and each needs +-d frames of its input, so a launch per step is
// latency bound (49 us each, 159 TFLOP/s; profiles/r01_launches_forward_b256_bf16_v0.txt).  Here one CTA
// owns a window of 384 consecutive rows (a whole 3 s utterance with its reflect halos, or a 320-frame
// segment of a long utterance plus 7*d recomputed context frames per side) and walks the whole chain:
//
//   * step input (x_j + y_{j-1}) lives in a CTA-private scratch strip (2 x 392 rows, L2 resident);
//     the three dilated taps are row-shifted TMA loads of that strip (im2col-free, as in gemm_tc.cu);
//   * per K block the producer loads ONE weight tile and the three 128-row A tiles, the MMA thread
//     issues tcgen05.mma into three TMEM accumulators (3 x Cs columns), so weights cross L2 once per step;
//   * the epilogue applies bias/ReLU/BN in fp32, writes y_j to R (valid frames + mirrored reflect halo
//     rows) and x_{j+1} + y_j to the other scratch strip, fences the generic-proxy stores towards the
//     async proxy (fence.proxy.async.global) and releases the producer for the next step.
// Rows of the window that belong to no valid context (beyond 7*d of a segment edge, or neighbouring
// utterances) carry garbage that never reaches a stored output: a conv output row depends only on its
// own three input rows.
#include <stdlib.h>

#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

struct UttRows { long long row0; int T; int pad; };

static constexpr int RC_THREADS = 384;     // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps4-11 epilogue
static constexpr int RC_MT = 3;            // 128-row M tiles per window
static constexpr int RC_WIN = RC_MT * 128; // rows per window
static constexpr int RC_GUARD = 4;         // scratch guard rows before / after the window (>= max dilation)
static constexpr int RC_STRIP = RC_WIN + 2 * RC_GUARD;
static constexpr int RC_SEG = 320;         // output frames per segment (RC_SEG + 14 * d <= RC_WIN for d <= 4)
static constexpr int RC_STEPS = 7;
static constexpr int RC_STAGES = 3;
static constexpr int RC_A_BYTES = 128 * 64 * 2;

struct alignas(64) ChainMaps {
  CUtensorMap u;            // tdnn1 output U [rows, P*C], box [64, 128]
  CUtensorMap s;            // scratch strips [grid*2*RC_STRIP, P*Cs], box [64, 128]
  CUtensorMap w[RC_STEPS];  // packed weights [Cs, 3*P*Cs], box [64, Cs]
};

struct SegItem { int utt, seg; };

struct ChainParams {
  long long M;
  int n_items;
  const SegItem* items;
  const UttRows* utts;
  int C, P, dil, halo;
  const bf16* U; int u_ld;
  bf16* R; int r_ld;
  bf16* S; int s_ld;        // scratch base, row stride
  const float* bias[RC_STEPS];
  const float* scale[RC_STEPS];
  const float* shift[RC_STEPS];
  int w_f16;                // single-plane weights are IEEE half (b_format = F16), see pack_conv_weight
  long long* dbg;           // optional phase timestamps of CTA 0 (SVB_CHAIN_DBG=1): [step][4] clock64 values
};

// 32 fp32 values -> bf16 hi (and lo) planes of one row
template <bool SPLIT>
__device__ __forceinline__ void store_row32(bf16* dst, int plane_cols, const float (&v)[32]) {
  uint32_t hi[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) hi[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
  uint4* d = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int i = 0; i < 4; ++i) d[i] = make_uint4(hi[4 * i], hi[4 * i + 1], hi[4 * i + 2], hi[4 * i + 3]);
  if (SPLIT) {
    uint32_t lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) lo[i] = pack_bf16x2(v[2 * i] - bf16_lo_f(hi[i]), v[2 * i + 1] - bf16_hi_f(hi[i]));
    uint4* dl = reinterpret_cast<uint4*>(dst + plane_cols);
#pragma unroll
    for (int i = 0; i < 4; ++i) dl[i] = make_uint4(lo[4 * i], lo[4 * i + 1], lo[4 * i + 2], lo[4 * i + 3]);
  }
}

template <bool SPLIT>
__device__ __forceinline__ void add_row32(const bf16* src, int plane_cols, float (&v)[32]) {
#pragma unroll
  for (int pl = 0; pl < (SPLIT ? 2 : 1); ++pl) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src + pl * plane_cols);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint4 u = __ldg(s4 + i);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        v[8 * i + 2 * q] += bf16_lo_f(w[q]);
        v[8 * i + 2 * q + 1] += bf16_hi_f(w[q]);
      }
    }
  }
}

template <int CS, bool SPLIT>
__global__ void __launch_bounds__(RC_THREADS, 1)
res2net_chain_kernel(const __grid_constant__ ChainMaps tm, const ChainParams p) {
  constexpr int P = SPLIT ? 2 : 1;
  constexpr int KC = CS / 64;
  constexpr int W_BYTES = CS * 128;
  constexpr int STAGE_STRIDE = 3 * RC_A_BYTES + 16384;
  constexpr int STAGE_BYTES = 3 * RC_A_BYTES + W_BYTES;
  constexpr int NPASS = SPLIT ? 3 : 1;
  constexpr int NKB = 3 * NPASS * KC;
  constexpr int TMEM_COLS = (3 * CS > 256) ? 512 : 256;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[RC_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[RC_STAGES];
  __shared__ __align__(8) uint64_t tfull_bar, tempty_bar, sready_bar;
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int d = p.dil, H = p.halo;
  const int strip0 = blockIdx.x * 2 * RC_STRIP;   // first scratch row of this CTA (strip b at + b * RC_STRIP)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm.u);
    tma_prefetch_desc(&tm.s);
    for (int j = 0; j < RC_STEPS; ++j) tma_prefetch_desc(&tm.w[j]);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < RC_STAGES; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    mbar_init(smem_u32(&tfull_bar), 1);
    mbar_init(smem_u32(&tempty_bar), 8);
    mbar_init(smem_u32(&sready_bar), 8);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(smem_u32(&tmem_slot), TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0, nwait = 0;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        const SegItem it = p.items[item];
        const UttRows u = p.utts[it.utt];
        const int s0 = it.seg * RC_SEG;
        const int tw0 = max(s0 - RC_STEPS * d, -H);
        const int g0 = static_cast<int>(u.row0) + tw0;
        for (int j = 1; j <= RC_STEPS; ++j) {
          if (j >= 2) { mbar_wait(smem_u32(&sready_bar), nwait & 1u); ++nwait; }
          const int arow = (j == 1) ? g0 : strip0 + (j & 1) * RC_STRIP + RC_GUARD;
          const int acol0 = (j == 1) ? CS : 0;
          const int aplane_cols = (j == 1) ? p.C : CS;
          const CUtensorMap* am = (j == 1) ? &tm.u : &tm.s;
          for (int kb = 0; kb < NKB; ++kb) {
            const int tap = kb / (NPASS * KC);
            const int rem = kb - tap * (NPASS * KC);
            const int pass = rem / KC;
            const int kcx = rem - pass * KC;
            const int aplane = (pass == 2) ? 1 : 0;
            const int wplane = (pass == 1) ? 1 : 0;
            const int off = (tap - 1) * d;
            const uint32_t fb = smem_u32(&full_bar[stage]);
            mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
            mbar_arrive_expect_tx(fb, STAGE_BYTES);
            const uint32_t sa = smem_base + stage * STAGE_STRIDE;
            tma_load_2d(sa + 3 * RC_A_BYTES, &tm.w[j - 1], fb, ((tap * P + wplane) * KC + kcx) * 64, 0);
#pragma unroll
            for (int mt = 0; mt < RC_MT; ++mt)
              tma_load_2d(sa + mt * RC_A_BYTES, am, fb, aplane * aplane_cols + acol0 + kcx * 64, arow + mt * 128 + off);
            if (++stage == RC_STAGES) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, CS);
      int stage = 0;
      uint32_t phase = 0, nstep = 0;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        for (int j = 1; j <= RC_STEPS; ++j, ++nstep) {
          mbar_wait(smem_u32(&tempty_bar), (nstep & 1u) ^ 1u);
          tc_fence_after();
          for (int kb = 0; kb < NKB; ++kb) {
            mbar_wait(smem_u32(&full_bar[stage]), phase);
            tc_fence_after();
            const uint32_t sa = smem_base + stage * STAGE_STRIDE;
            const uint64_t bdesc = umma_desc_sw128(sa + 3 * RC_A_BYTES);
#pragma unroll
            for (int mt = 0; mt < RC_MT; ++mt) {
              const uint64_t adesc = umma_desc_sw128(sa + mt * RC_A_BYTES);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma_bf16(tmem_base + mt * CS, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) ? 1u : 0u);
            }
            umma_commit(smem_u32(&empty_bar[stage]));
            if (++stage == RC_STAGES) { stage = 0; phase ^= 1u; }
          }
          umma_commit(smem_u32(&tfull_bar));
        }
      }
    }
  } else if (warp >= 4) {
    // ===================================================== epilogue
    constexpr int NCOLH = CS / 2;          // columns per warp (two column halves)
    const int ew = warp - 4;
    const int quad = ew & 3;               // == warp % 4: TMEM lane quadrant
    const int half = ew >> 2;
    const uint32_t lane_sel = static_cast<uint32_t>(quad * 32) << 16;
    uint32_t nstep = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const SegItem it = p.items[item];
      const UttRows u = p.utts[it.utt];
      const int T = u.T;
      const int s0 = it.seg * RC_SEG;
      const int e0 = min(T, s0 + RC_SEG);
      const int tw0 = max(s0 - RC_STEPS * d, -H);
      const long long g0 = u.row0 + tw0;
      // per M tile: local row, targets in the scratch strip (own row unless it is a reflect-halo row, plus the
      // halo rows this frame mirrors into) and whether this frame is an output frame of the segment
      int lrow[RC_MT], mir_lo[RC_MT], mir_hi[RC_MT];
      bool own_s[RC_MT], out_r[RC_MT], xok[RC_MT];
#pragma unroll
      for (int mt = 0; mt < RC_MT; ++mt) {
        const int l = mt * 128 + quad * 32 + lane;
        const int t = tw0 + l;
        const bool in_utt = t >= 0 && t < T;
        const bool halo_row = (t >= -H && t < 0) || (t >= T && t < T + H);
        lrow[mt] = l;
        own_s[mt] = !halo_row;
        out_r[mt] = in_utt && t >= s0 && t < e0;
        mir_lo[mt] = (in_utt && t >= 1 && t <= H && l - 2 * t >= 0) ? l - 2 * t : -1;
        const int ee = T - 1 - t;
        mir_hi[mt] = (in_utt && ee >= 1 && ee <= H && l + 2 * ee < RC_WIN) ? l + 2 * ee : -1;
        xok[mt] = (g0 + l >= 0) && (g0 + l < p.M);
      }
      for (int j = 1; j <= RC_STEPS; ++j, ++nstep) {
        const float* bias = p.bias[j - 1];
        const float* scale = p.scale[j - 1];
        const float* shift = p.shift[j - 1];
        bf16* sdst = p.S + static_cast<long long>(strip0 + ((j + 1) & 1) * RC_STRIP + RC_GUARD) * p.s_ld;
        mbar_wait(smem_u32(&tfull_bar), nstep & 1u);
        tc_fence_after();
#pragma unroll
        for (int mt = 0; mt < RC_MT; ++mt) {
#pragma unroll 1
          for (int g = 0; g < NCOLH / 32; ++g) {
            const int col = half * NCOLH + g * 32;
            uint32_t r[32];
            tmem_ld32(tmem_base + mt * CS + col + lane_sel, r);
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 b = __ldg(reinterpret_cast<const float4*>(bias + col) + i);
              const float4 s = __ldg(reinterpret_cast<const float4*>(scale + col) + i);
              const float4 h = __ldg(reinterpret_cast<const float4*>(shift + col) + i);
              v[4 * i] = fmaf(fmaxf(__uint_as_float(r[4 * i]) + b.x, 0.f), s.x, h.x);
              v[4 * i + 1] = fmaf(fmaxf(__uint_as_float(r[4 * i + 1]) + b.y, 0.f), s.y, h.y);
              v[4 * i + 2] = fmaf(fmaxf(__uint_as_float(r[4 * i + 2]) + b.z, 0.f), s.z, h.z);
              v[4 * i + 3] = fmaf(fmaxf(__uint_as_float(r[4 * i + 3]) + b.w, 0.f), s.w, h.w);
            }
            const long long grow = g0 + lrow[mt];
            if (out_r[mt]) {
              bf16* rd = p.R + grow * p.r_ld + j * CS + col;
              store_row32<SPLIT>(rd, p.C, v);
              if (mir_lo[mt] >= 0) store_row32<SPLIT>(rd + static_cast<long long>(mir_lo[mt] - lrow[mt]) * p.r_ld, p.C, v);
              if (mir_hi[mt] >= 0) store_row32<SPLIT>(rd + static_cast<long long>(mir_hi[mt] - lrow[mt]) * p.r_ld, p.C, v);
            }
            if (j < RC_STEPS) {
              if (xok[mt]) add_row32<SPLIT>(p.U + grow * p.u_ld + (j + 1) * CS + col, p.C, v);
              if (own_s[mt]) store_row32<SPLIT>(sdst + static_cast<long long>(lrow[mt]) * p.s_ld + col, CS, v);
              if (mir_lo[mt] >= 0) store_row32<SPLIT>(sdst + static_cast<long long>(mir_lo[mt]) * p.s_ld + col, CS, v);
              if (mir_hi[mt] >= 0) store_row32<SPLIT>(sdst + static_cast<long long>(mir_hi[mt]) * p.s_ld + col, CS, v);
            }
          }
        }
        tc_fence_before();
        if (j < RC_STEPS) {
          __threadfence();                                              // strip stores performed at GPU scope (L2) ...
          asm volatile("fence.proxy.async.global;" ::: "memory");       // ... and ordered before the async-proxy (TMA) reads
        }
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(smem_u32(&tempty_bar));
          if (j < RC_STEPS) mbar_arrive(smem_u32(&sready_bar));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------ shared-memory resident variant (1 plane)
// 16-bit mode (single plane = IEEE half activations, fp16 or bf16 weights): the step input never leaves the SM.  The activation strip (400 rows x Cs channels, K-major,
// 128B swizzle, one plane per 64-channel K chunk) stays in shared memory for the whole chain; the epilogue
// writes x_{j+1} + y_j straight into it in the UMMA operand layout and the three dilated taps are the SAME
// strip read through descriptors whose start address is shifted by +-d rows (d * 128 bytes).  A hardware
// probe (tools/desc_shift_test.cu, profiles/r01_desc_shift_probe.txt) shows that tcgen05 applies the 128B
// swizzle to absolute shared-memory address bits, so any 128-byte-aligned start row is legal with the
// descriptor's base_offset field left 0.  Per step only the weights (3 taps) cross L2; they are loaded
// while the previous step's epilogue runs.
static constexpr int RS_ROWS = 400;                 // strip rows in smem: 4 guard + 384 window + 4 guard + 8 pad
static constexpr int RS_PLANE = RS_ROWS * 128;      // bytes per 64-channel K chunk plane
// staging tile: 64-byte rows, 16-byte chunk c of row r at ((c ^ ((r >> 1) & 3)) << 4): conflict-free both for the
// thread-per-row accesses and for the coalesced ones (4 lanes per row, 2 rows per quarter warp).  (An 80-byte pitch
// without swizzle made every coalesced access a 2-way conflict: ncu r02, 18 % of the kernel's shared wavefronts, in a
// kernel whose busiest unit is the LSU data pipe at 72 %.)
static constexpr int RS_PITCH = 64;
__device__ __forceinline__ uint32_t rs_addr(uint32_t stg, int row, int chunk) {
  return stg + row * RS_PITCH + ((chunk ^ ((row >> 1) & 3)) << 4);
}
static constexpr int RS_BOX = 200;                  // TMA box rows for the step-1 strip load (2 boxes per plane)

struct alignas(64) ChainMapsS {
  CUtensorMap u;            // U [rows, C], box [64, RS_BOX]
  CUtensorMap w[RC_STEPS];  // packed weights [Cs, 3*Cs], box [64, Cs]
};

template <int CS>
__global__ void __launch_bounds__(RC_THREADS, 1)
res2net_chain_smem_kernel(const __grid_constant__ ChainMapsS tm, const ChainParams p) {
  constexpr int KC = CS / 64;
  constexpr int W_TILE = CS * 128;
  constexpr int W_BYTES = 3 * KC * W_TILE;
  constexpr int A_BYTES = KC * RS_PLANE;
  constexpr int TMEM_COLS = (3 * CS > 256) ? 512 : 256;
  constexpr int NCOLH = CS / 2;
  constexpr int NG = NCOLH / 32;        // 32-column groups per warp per M tile
  extern __shared__ uint8_t smem_raw[];
  // per-M-tile accumulator / strip barriers: the epilogue of M tile 0 runs while the MMAs of tiles 1 and 2 are still
  // executing, and the next step's MMAs of tile mt start as soon as the strip rows they read (tiles <= mt + 1) are
  // rewritten, i.e. while the epilogue of the last tile is still running (the MMA phase, 25 % of a step, is hidden)
  __shared__ __align__(8) uint64_t afull_bar, aempty_bar, wfull_bar, wempty_bar, tfull_bar[RC_MT], tempty_bar[RC_MT],
      sready_bar[RC_MT];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_smem = smem_base;
  const uint32_t w_smem = smem_base + A_BYTES;
  const uint32_t stg = w_smem + W_BYTES + (warp >= 4 ? warp - 4 : 0) * 32 * RS_PITCH;   // per-epilogue-warp staging tile
  float* prm_gen = reinterpret_cast<float*>(smem_raw + (smem_base - smem_u32(smem_raw)) + A_BYTES + W_BYTES + 8 * 32 * RS_PITCH);
  const int d = p.dil, H = p.halo;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm.u);
    for (int j = 0; j < RC_STEPS; ++j) tma_prefetch_desc(&tm.w[j]);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(smem_u32(&afull_bar), 1);
    mbar_init(smem_u32(&aempty_bar), 1);
    mbar_init(smem_u32(&wfull_bar), 1);
    mbar_init(smem_u32(&wempty_bar), 1);
    for (int mt = 0; mt < RC_MT; ++mt) {
      mbar_init(smem_u32(&tfull_bar[mt]), 1);
      mbar_init(smem_u32(&tempty_bar[mt]), 8);
      mbar_init(smem_u32(&sready_bar[mt]), 8);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(smem_u32(&tmem_slot), TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer: strip of step 1, weights of every step
    if (elect_one()) {
      uint32_t nitem = 0, nw = 0;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++nitem) {
        const SegItem it = p.items[item];
        const UttRows u = p.utts[it.utt];
        const int tw0 = max(it.seg * RC_SEG - RC_STEPS * d, -H);
        const int g0 = static_cast<int>(u.row0) + tw0;
        mbar_wait(smem_u32(&aempty_bar), (nitem & 1u) ^ 1u);      // previous item's last MMAs have read the strip
        const uint32_t ab = smem_u32(&afull_bar);
        mbar_arrive_expect_tx(ab, KC * 2 * RS_BOX * 128);
        for (int kcx = 0; kcx < KC; ++kcx)
          for (int h = 0; h < 2; ++h)
            tma_load_2d(a_smem + kcx * RS_PLANE + h * RS_BOX * 128, &tm.u, ab, CS + kcx * 64, g0 - RC_GUARD + h * RS_BOX);
        for (int j = 1; j <= RC_STEPS; ++j, ++nw) {
          mbar_wait(smem_u32(&wempty_bar), (nw & 1u) ^ 1u);       // previous step's MMAs have read the weights
          const uint32_t wb = smem_u32(&wfull_bar);
          mbar_arrive_expect_tx(wb, W_BYTES);
          for (int tap = 0; tap < 3; ++tap)
            for (int kcx = 0; kcx < KC; ++kcx)
              tma_load_2d(w_smem + (tap * KC + kcx) * W_TILE, &tm.w[j - 1], wb, (tap * KC + kcx) * 64, 0);
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_f16(128, CS, true, p.w_f16 != 0);      // activations of the 1-plane path are IEEE half
      uint32_t nitem = 0, nstep = 0, nsr = 0;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++nitem) {
        for (int j = 1; j <= RC_STEPS; ++j, ++nstep) {
          mbar_wait(smem_u32(&wfull_bar), nstep & 1u);
          if (j == 1) mbar_wait(smem_u32(&afull_bar), nitem & 1u);
          if (p.dbg != nullptr && blockIdx.x == 0 && nstep < 64) p.dbg[nstep * 4 + 0] = clock64();
#pragma unroll
          for (int mt = 0; mt < RC_MT; ++mt) {
            // strip rows [mt*128 - d, mt*128 + 128 + d) of this step's input: written by the previous step's epilogue
            // of tiles <= mt + 1 (epilogue tiles complete in order)
            if (j > 1) mbar_wait(smem_u32(&sready_bar[mt + 1 < RC_MT ? mt + 1 : RC_MT - 1]), nsr & 1u);
            mbar_wait(smem_u32(&tempty_bar[mt]), (nstep & 1u) ^ 1u);
            tc_fence_after();
            uint32_t accf = 0;
#pragma unroll
            for (int tap = 0; tap < 3; ++tap) {
#pragma unroll
              for (int kcx = 0; kcx < KC; ++kcx) {
                const uint64_t adesc = umma_desc_sw128(a_smem + kcx * RS_PLANE + (RC_GUARD + mt * 128 + (tap - 1) * d) * 128);
                const uint64_t bdesc = umma_desc_sw128(w_smem + (tap * KC + kcx) * W_TILE);
#pragma unroll
                for (int k = 0; k < 4; ++k) { umma_bf16(tmem_base + mt * CS, adesc + 2 * k, bdesc + 2 * k, idesc, accf); accf = 1; }
              }
            }
            umma_commit(smem_u32(&tfull_bar[mt]));
          }
          if (j > 1) ++nsr;
          umma_commit(smem_u32(&wempty_bar));
          if (j == RC_STEPS) umma_commit(smem_u32(&aempty_bar));
          if (p.dbg != nullptr && blockIdx.x == 0 && nstep < 64) p.dbg[nstep * 4 + 1] = clock64();
        }
      }
    }
  } else if (warp >= 4) {
    // ===================================================== epilogue
    const int ew = warp - 4;
    const int quad = ew & 3;               // == warp % 4: TMEM lane quadrant
    const int half = ew >> 2;
    const uint32_t lane_sel = static_cast<uint32_t>(quad * 32) << 16;
    uint32_t nstep = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const SegItem it = p.items[item];
      const UttRows u = p.utts[it.utt];
      const int T = u.T;
      const int s0 = it.seg * RC_SEG;
      const int e0 = min(T, s0 + RC_SEG);
      const int tw0 = max(s0 - RC_STEPS * d, -H);
      const long long g0 = u.row0 + tw0;
      int lrow[RC_MT], mir_lo[RC_MT], mir_hi[RC_MT];
      bool own_s[RC_MT], out_r[RC_MT], xok[RC_MT];
#pragma unroll
      for (int mt = 0; mt < RC_MT; ++mt) {
        const int l = mt * 128 + quad * 32 + lane;
        const int t = tw0 + l;
        const bool in_utt = t >= 0 && t < T;
        const bool halo_row = (t >= -H && t < 0) || (t >= T && t < T + H);
        lrow[mt] = l;
        own_s[mt] = !halo_row;
        out_r[mt] = in_utt && t >= s0 && t < e0;
        mir_lo[mt] = (in_utt && t >= 1 && t <= H && l - 2 * t >= 0) ? l - 2 * t : -1;
        const int ee = T - 1 - t;
        mir_hi[mt] = (in_utt && ee >= 1 && ee <= H && l + 2 * ee < RC_WIN) ? l + 2 * ee : -1;
        xok[mt] = (g0 + l >= 0) && (g0 + l < p.M);
      }
      // L1 handles one 128-byte line per cycle: a warp instruction in which every lane touches its own row costs
      // 32 tag cycles, and an epilogue built from such accesses (thread = TMEM lane = row) ran 21 k cycles per step.
      // All global traffic below is therefore issued in a "coalesced" mapping (instruction i covers rows 8i + lane/4,
      // 16-byte chunk lane%4 of a 64-byte row segment: 8 lines per instruction) and transposed to / from the
      // thread-per-row mapping through a small per-warp staging tile in shared memory.
      uint32_t m_out[RC_MT];
#pragma unroll
      for (int mt = 0; mt < RC_MT; ++mt) m_out[mt] = __ballot_sync(0xffffffffu, out_r[mt]);
      const int crow = lane >> 2, cchk = lane & 3;
      const long long gq = g0 + quad * 32;     // global row of this warp's first row in M tile 0
      // per-item 64-bit bases; everything inside the step loops is 32-bit offset arithmetic from them
      const bf16* xbase = p.U + (gq + crow) * static_cast<long long>(p.u_ld) + half * NCOLH + cchk * 8;
      bf16* rbase0 = p.R + (gq + crow) * static_cast<long long>(p.r_ld) + half * NCOLH + cchk * 8;
      uint32_t xmask = 0;                      // bit mt * 4 + i: row gq + mt * 128 + 8 i + crow lies inside the buffer
#pragma unroll
      for (int q = 0; q < 4 * RC_MT; ++q) {
        const long long row = gq + (q >> 2) * 128 + 8 * (q & 3) + crow;
        if (row >= 0 && row < p.M) xmask |= 1u << q;
      }
      auto load_x = [&](int mt, int g, int j, uint4 (&xc)[4]) {
        const int off = mt * 128 * p.u_ld + (j + 1) * CS + g * 32;
        const uint32_t vm = xmask >> (mt * 4);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          xc[i] = ((vm >> i) & 1u) ? __ldg(reinterpret_cast<const uint4*>(xbase + (off + 8 * i * p.u_ld))) : make_uint4(0, 0, 0, 0);
      };
      for (int j = 1; j <= RC_STEPS; ++j, ++nstep) {
        // this step's bias / BN scale / BN shift -> shared memory (double buffered by step parity)
        float* prm = prm_gen + (nstep & 1u) * 3 * CS;
        for (int i = ew * 32 + lane; i < 3 * CS; i += 256) {
          const int arr = i / CS, c = i - arr * CS;
          const float* src = arr == 0 ? p.bias[j - 1] : (arr == 1 ? p.scale[j - 1] : p.shift[j - 1]);
          prm[i] = __ldg(src + c);
        }
        uint4 xc[4];
        if (j < RC_STEPS) {
          // x_{j+1} does not depend on this step's MMAs: pull it towards L2 and fetch the first group while they run
#pragma unroll
          for (int mt = 0; mt < RC_MT; ++mt)
            if (xok[mt])
              asm volatile("prefetch.global.L2 [%0];" ::"l"(p.U + (g0 + lrow[mt]) * p.u_ld + (j + 1) * CS + half * NCOLH));
          load_x(0, 0, j, xc);
        }
        named_bar_sync(1, 256);
        bool all_mma_done = false;
#pragma unroll
        for (int mt = 0; mt < RC_MT; ++mt) {
          mbar_wait(smem_u32(&tfull_bar[mt]), nstep & 1u);
          tc_fence_after();
          if (mt == 0 && p.dbg != nullptr && blockIdx.x == 0 && ew == 0 && lane == 0 && nstep < 64) p.dbg[nstep * 4 + 2] = clock64();
#pragma unroll 1
          for (int g = 0; g < NG; ++g) {
            const int col = half * NCOLH + g * 32;
            uint4 xrow[4];
            if (j < RC_STEPS) {
              // transpose the x tile: coalesced registers -> staging -> own row
#pragma unroll
              for (int i = 0; i < 4; ++i)
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rs_addr(stg, 8 * i + crow, cchk)),
                             "r"(xc[i].x), "r"(xc[i].y), "r"(xc[i].z), "r"(xc[i].w) : "memory");
              __syncwarp();
#pragma unroll
              for (int i = 0; i < 4; ++i)
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(xrow[i].x), "=r"(xrow[i].y), "=r"(xrow[i].z),
                             "=r"(xrow[i].w) : "r"(rs_addr(stg, lane, i)) : "memory");
              __syncwarp();
              const int gn = (g + 1 < NG) ? g + 1 : 0, mn = (g + 1 < NG) ? mt : mt + 1;
              if (mn < RC_MT) load_x(mn, gn, j, xc);
            }
            uint32_t r[32];
            tmem_ld32(tmem_base + mt * CS + col + lane_sel, r);
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 b = *reinterpret_cast<const float4*>(prm + col + 4 * i);
              const float4 sc = *reinterpret_cast<const float4*>(prm + CS + col + 4 * i);
              const float4 h = *reinterpret_cast<const float4*>(prm + 2 * CS + col + 4 * i);
              v[4 * i] = fmaf(fmaxf(__uint_as_float(r[4 * i]) + b.x, 0.f), sc.x, h.x);
              v[4 * i + 1] = fmaf(fmaxf(__uint_as_float(r[4 * i + 1]) + b.y, 0.f), sc.y, h.y);
              v[4 * i + 2] = fmaf(fmaxf(__uint_as_float(r[4 * i + 2]) + b.z, 0.f), sc.z, h.z);
              v[4 * i + 3] = fmaf(fmaxf(__uint_as_float(r[4 * i + 3]) + b.w, 0.f), sc.w, h.w);
            }
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) pk[i] = pack_f16x2(v[2 * i], v[2 * i + 1]);
            // y_j -> R: own row into staging, then coalesced stores of the segment's output rows
#pragma unroll
            for (int i = 0; i < 4; ++i)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rs_addr(stg, lane, i)), "r"(pk[4 * i]),
                           "r"(pk[4 * i + 1]), "r"(pk[4 * i + 2]), "r"(pk[4 * i + 3]) : "memory");
            __syncwarp();
            bf16* rbase = rbase0 + (mt * 128 * p.r_ld + j * CS + g * 32);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int rr = 8 * i + crow;
              if ((m_out[mt] >> rr) & 1u) {
                uint4 o;
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(o.x), "=r"(o.y), "=r"(o.z), "=r"(o.w)
                             : "r"(rs_addr(stg, rr, cchk)) : "memory");
                *reinterpret_cast<uint4*>(rbase + 8 * i * p.r_ld) = o;
              }
            }
            if (out_r[mt] && (mir_lo[mt] >= 0 || mir_hi[mt] >= 0)) {
              // reflect-halo copies of the first / last H frames (a handful of rows per utterance): direct row stores
              bf16* rd = p.R + (g0 + lrow[mt]) * p.r_ld + j * CS + col;
              if (mir_lo[mt] >= 0) {
                uint4* dd = reinterpret_cast<uint4*>(rd + static_cast<long long>(mir_lo[mt] - lrow[mt]) * p.r_ld);
#pragma unroll
                for (int i = 0; i < 4; ++i) dd[i] = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
              }
              if (mir_hi[mt] >= 0) {
                uint4* dd = reinterpret_cast<uint4*>(rd + static_cast<long long>(mir_hi[mt] - lrow[mt]) * p.r_ld);
#pragma unroll
                for (int i = 0; i < 4; ++i) dd[i] = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
              }
            }
            __syncwarp();
            if (j < RC_STEPS) {
              // next step's input x_{j+1} + y_j (single rounding from fp32) straight into the operand strip
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint32_t w[4] = {xrow[i].x, xrow[i].y, xrow[i].z, xrow[i].w};
#pragma unroll
                for (int q = 0; q < 4; ++q)
                  pk[4 * i + q] = pack_f16x2(v[8 * i + 2 * q] + f16_lo_f(w[q]), v[8 * i + 2 * q + 1] + f16_hi_f(w[q]));
              }
              if (!all_mma_done) {
                // the strip is single buffered: this step's MMAs of the later tiles read rows next to (and, through the
                // reflect mirrors, inside) the rows rewritten here -- all of them must have completed first
                mbar_wait(smem_u32(&tfull_bar[RC_MT - 1]), nstep & 1u);
                all_mma_done = true;
              }
              const uint32_t plane = a_smem + (col >> 6) * RS_PLANE;
              const int ci0 = (col & 63) >> 3;
              const int tg[3] = {own_s[mt] ? lrow[mt] : -1, mir_lo[mt], mir_hi[mt]};
#pragma unroll
              for (int q = 0; q < 3; ++q) {
                if (tg[q] < 0) continue;
                const int rho = RC_GUARD + tg[q];
                const uint32_t rowaddr = plane + rho * 128;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + (((ci0 + i) ^ (rho & 7)) << 4)),
                               "r"(pk[4 * i]), "r"(pk[4 * i + 1]), "r"(pk[4 * i + 2]), "r"(pk[4 * i + 3])
                               : "memory");
              }
            }
          }
          // tile mt done: accumulator free, strip rows (and the mirrors they feed) in place
          tc_fence_before();
          if (j < RC_STEPS) fence_proxy_async();   // strip stores (generic proxy) -> visible to the tensor core (async proxy)
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(smem_u32(&tempty_bar[mt]));
            if (j < RC_STEPS) mbar_arrive(smem_u32(&sready_bar[mt]));
          }
        }
        if (p.dbg != nullptr && blockIdx.x == 0 && ew == 0 && lane == 0 && nstep < 64) p.dbg[nstep * 4 + 3] = clock64();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, TMEM_COLS);
}

template <int CS>
static int launch_chain_smem(const ChainMapsS& tm, const ChainParams& p, cudaStream_t stream) {
  constexpr int SMEM = (CS / 64) * RS_PLANE + 3 * (CS / 64) * CS * 128 + 8 * 32 * RS_PITCH + 2 * 3 * CS * 4 + 1024;
  static unsigned long long attr_devs = 0;   // per device: kernel attributes do not carry over
  if (first_on_device(&attr_devs)) {
    SVB_CUDA_OK(cudaFuncSetAttribute(res2net_chain_smem_kernel<CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  }
  const int grid = p.n_items < num_sms() ? p.n_items : num_sms();
  res2net_chain_smem_kernel<CS><<<grid, RC_THREADS, SMEM, stream>>>(tm, p);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------ host side
size_t res2net_scratch_bytes(int Cs, int planes) {
  return static_cast<size_t>(num_sms()) * 2 * RC_STRIP * planes * Cs * sizeof(bf16);
}
int res2net_seg_frames() { return RC_SEG; }

template <int CS, bool SPLIT>
static int launch_chain(const ChainMaps& tm, const ChainParams& p, cudaStream_t stream) {
  constexpr int SMEM = RC_STAGES * (3 * RC_A_BYTES + 16384) + 1024;
  static unsigned long long attr_devs = 0;   // per device: kernel attributes do not carry over
  if (first_on_device(&attr_devs)) {
    SVB_CUDA_OK(cudaFuncSetAttribute(res2net_chain_kernel<CS, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  }
  const int grid = p.n_items < num_sms() ? p.n_items : num_sms();
  res2net_chain_kernel<CS, SPLIT><<<grid, RC_THREADS, SMEM, stream>>>(tm, p);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// U, R: [rows, planes*C] bf16 (hi | lo at +C).  w[j] packed [Cs, 3*planes*Cs]; items: device table of (utt, segment)
// with segments of res2net_seg_frames() frames; scratch: res2net_scratch_bytes(Cs, planes) bytes.
int res2net_chain(const bf16* U, bf16* R, int C, int planes, long long rows, const bf16* const* w, const float* const* bias,
                  const float* const* scale, const float* const* shift, int dil, int halo, const void* items, int n_items,
                  const void* utts, void* scratch, cudaStream_t stream, int w_f16) {
  const int Cs = C / 8;
  SVB_CHECK(!w_f16 || planes == 1, "fp16 weights are single plane");
  SVB_CHECK(Cs == 64 || Cs == 128, "the fused Res2Net chain supports C = 512 or 1024 (Res2Net chunks of 64 / 128 channels)");
  SVB_CHECK(dil >= 1 && dil <= RC_GUARD && halo >= dil && RC_SEG + 2 * RC_STEPS * dil <= RC_WIN, "unsupported dilation %d", dil);
  if (planes == 1) {
    ChainMapsS ts;
    if (make_map(&ts.u, U, rows, C, C, RS_BOX)) return -1;
    for (int j = 0; j < RC_STEPS; ++j)
      if (make_map(&ts.w[j], w[j], Cs, 3 * Cs, 3 * Cs, Cs)) return -1;
    ChainParams q{};
    q.M = rows; q.n_items = n_items; q.items = static_cast<const SegItem*>(items);
    q.utts = static_cast<const UttRows*>(utts);
    q.C = C; q.P = 1; q.dil = dil; q.halo = halo; q.w_f16 = w_f16;
    q.U = U; q.u_ld = C; q.R = R; q.r_ld = C;
    for (int j = 0; j < RC_STEPS; ++j) { q.bias[j] = bias[j]; q.scale[j] = scale[j]; q.shift[j] = shift[j]; }
    if (getenv("SVB_CHAIN_DBG") != nullptr) {
      // diagnostics: per-step phase timestamps of CTA 0 (MMA start / MMA issued / epilogue start / epilogue end)
      long long* d_dbg = nullptr;
      SVB_CUDA_OK(cudaMalloc(&d_dbg, 64 * 4 * sizeof(long long)));
      SVB_CUDA_OK(cudaMemsetAsync(d_dbg, 0, 64 * 4 * sizeof(long long), stream));
      q.dbg = d_dbg;
      const int rc = Cs == 128 ? launch_chain_smem<128>(ts, q, stream) : launch_chain_smem<64>(ts, q, stream);
      long long h[64 * 4];
      SVB_CUDA_OK(cudaStreamSynchronize(stream));
      SVB_CUDA_OK(cudaMemcpy(h, d_dbg, sizeof(h), cudaMemcpyDeviceToHost));
      cudaFree(d_dbg);
      for (int i = 0; i < 14 && h[i * 4] != 0; ++i)
        fprintf(stderr, "chain step %2d: mma_wait_end +%6lld | issue %5lld | mma->epi_start %6lld | epilogue %6lld clk\n", i,
                i ? h[i * 4] - h[(i - 1) * 4 + 3] : 0LL, h[i * 4 + 1] - h[i * 4], h[i * 4 + 2] - h[i * 4], h[i * 4 + 3] - h[i * 4 + 2]);
      return rc;
    }
    return Cs == 128 ? launch_chain_smem<128>(ts, q, stream) : launch_chain_smem<64>(ts, q, stream);
  }
  ChainMaps tm;
  if (make_map(&tm.u, U, rows, static_cast<long long>(planes) * C, static_cast<long long>(planes) * C, 128)) return -1;
  const long long srows = static_cast<long long>(num_sms()) * 2 * RC_STRIP;
  if (make_map(&tm.s, scratch, srows, planes * Cs, planes * Cs, 128)) return -1;
  for (int j = 0; j < RC_STEPS; ++j)
    if (make_map(&tm.w[j], w[j], Cs, 3 * planes * Cs, 3 * planes * Cs, Cs)) return -1;
  ChainParams p{};
  p.M = rows; p.n_items = n_items; p.items = static_cast<const SegItem*>(items);
  p.utts = static_cast<const UttRows*>(utts);
  p.C = C; p.P = planes; p.dil = dil; p.halo = halo;
  p.U = U; p.u_ld = planes * C; p.R = R; p.r_ld = planes * C;
  p.S = static_cast<bf16*>(scratch); p.s_ld = planes * Cs;
  for (int j = 0; j < RC_STEPS; ++j) { p.bias[j] = bias[j]; p.scale[j] = scale[j]; p.shift[j] = shift[j]; }
  if (Cs == 128) return planes == 2 ? launch_chain<128, true>(tm, p, stream) : launch_chain<128, false>(tm, p, stream);
  return planes == 2 ? launch_chain<64, true>(tm, p, stream) : launch_chain<64, false>(tm, p, stream);
}

}  // namespace svb
