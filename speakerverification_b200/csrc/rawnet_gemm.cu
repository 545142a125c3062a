// tcgen05 implicit-GEMM kernels for the RawNet2 raw-waveform encoder (reference:
// src/models/RawNet2_custom.py:161-227, src/models/RawNet_baseline.py:181-361).
//
// Same persistent warp-specialised skeleton as gemm_tc.cu (warp 0 TMA producer, warp 1 tcgen05.mma issuer,
// fp32 accumulators double buffered in TMEM, 8 epilogue warps); what is specific to RawNet2 is the operand view
// of the first layer and the epilogues:
//
//  EPI_SINC   SincConv_fast (128 band-pass FIRs x 251 taps, stride 1, RawNet_baseline.py:339-361) + abs +
//             max_pool1d(3) + first_bn + LeakyReLU(0.3) (RawNet2_custom.py:171-176) and layer1's bn1 + LeakyReLU
//             (RawNet_baseline.py:221).  The conv is a GEMM over a SLIDING-WINDOW view of the normalised
//             waveform: A row m = samples [48 m, 48 m + 320) -- a TMA map whose row stride (96 B) is smaller
//             than its row extent (profiles/r01_tma_overlap_probe.txt) -- against 48 shifted copies of every
//             filter (W'[(jg, f, ph), k'] = h[f, k' - (3 jg + ph)]), so one accumulator row holds outputs
//             t = 48 m .. 48 m + 47 of all filters, ordered so that the three phases of a pooling window are
//             adjacent columns: the max-pool is a 3-column max inside one thread.  No im2col, no [B,128,63750]
//             intermediate in HBM (the reference writes 32.6 MB per utterance there).
//  EPI_AFF_LRELU   conv1 of a RawNetBasicBlock: lrelu(bn2(conv1(a)))  (RawNet_baseline.py:223-224)
//  EPI_ADD         conv2 + shortcut                                    (RawNet_baseline.py:224-225)
//  EPI_PLAIN       1x1 shortcut conv                                    (RawNet_baseline.py:213-217,222)
//
// Activations are time-major rows [utt][1 zero row | T rows | 1 zero row][C] bf16: the zero rows ARE the conv's
// zero padding, so the k=3 taps are row-shifted TMA loads.  Every tensor is one plane of IEEE half in the 16-bit mode
// and bf16 hi|lo planes in fp32-grade mode (operand and residual-carrier plane counts are separate parameters).  T is
// the same for every utterance (the model's LayerNorm weights fix the input length, RawNet_baseline.py:13-24), so
// utterance / frame of a row are computed arithmetically.
//
// 16-bit mode variants (template WRES / runtime flags, each with an environment switch for A/B runs):
//   * sinc: the W boxes of the CTA's n-tile resident, A-only stage ring, outputs through staging boxes and 4-D TMA stores;
//   * 128-channel conv1 / conv2: all of W resident, one 136-row A box per K chunk for the three taps, the residual
//     through an identity MMA, outputs through staging boxes and TMA stores, and (conv2 of the down-sampling blocks)
//     max_pool1d(3) fused into the epilogue on per-utterance tiles of 42 pooling windows;
//   * 256 / 512-channel layers: generic mainloop, outputs through 32 x 32 staging boxes and TMA stores, residual
//     requested before the accumulator wait.
#include <stdlib.h>

#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

static constexpr int BLOCK_M = 128;
static constexpr int BLOCK_K = 64;
static constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;
enum { EPI_SINC = 0, EPI_AFF_LRELU = 1, EPI_ADD = 2, EPI_PLAIN = 3 };

// warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, then 8 epilogue warps -- 16 for the sinc layer, whose epilogue (96
// accumulators -> abs, max-pool, two affine + LeakyReLU, two stores per thread and tile) was latency bound with 8:
// ncu r02, tensor pipe 34 %, issue 30 %, DRAM 18 % busy at 3.2 us per tile against 1.0 us of MMAs.
static constexpr int SINC_EPI_WARPS = 16;
// (the W-resident sinc variant stores through TMA boxes of 32 rows x 32 filters per warp and runs 8 epilogue warps)
template <int EPI, bool WRES = false> struct RnThreads {
  static constexpr int N = (EPI == EPI_SINC && !WRES) ? 128 + 32 * SINC_EPI_WARPS : 384;
};

template <int BN>
struct RnCfg {
  static constexpr int B_STAGE_BYTES = BN * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int STAGES = (BN >= 192) ? 4 : 6;
  static constexpr int OUT_BYTES = 8 * 2048;                      // one 32-row x 32-column store box per epilogue warp
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + OUT_BYTES + 1024;
  static constexpr int TMEM_COLS = (2 * BN <= 256) ? 256 : 512;   // power of two
};

// max of two packed IEEE halves (same result as decoding, fmaxf and re-encoding)
__device__ __forceinline__ uint32_t hmax2_u32(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ float lrelu03(float x) { return fmaxf(x, 0.3f * x); }   // slope < 1: max picks the right branch

// 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// n consecutive fp32 values -> bf16 hi and lo = v - hi planes at dst / dst + plane_cols (planes == 2), or ONE plane of
// IEEE half (planes == 1, the 16-bit mode); N multiple of 8
template <int N>
__device__ __forceinline__ void store_planes(bf16* dst, int planes, int plane_cols, const float (&v)[N]) {
  uint32_t hi[N / 2];
#pragma unroll
  for (int i = 0; i < N / 2; ++i)
    hi[i] = planes == 1 ? pack_f16x2(v[2 * i], v[2 * i + 1]) : pack_bf16x2(v[2 * i], v[2 * i + 1]);
  uint4* d = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int i = 0; i < N / 8; ++i) d[i] = make_uint4(hi[4 * i], hi[4 * i + 1], hi[4 * i + 2], hi[4 * i + 3]);
  if (planes == 2) {
    uint32_t lo[N / 2];
#pragma unroll
    for (int i = 0; i < N / 2; ++i)
      lo[i] = pack_bf16x2(v[2 * i] - bf16_lo_f(hi[i]), v[2 * i + 1] - bf16_hi_f(hi[i]));
    uint4* dl = reinterpret_cast<uint4*>(dst + plane_cols);
#pragma unroll
    for (int i = 0; i < N / 8; ++i) dl[i] = make_uint4(lo[4 * i], lo[4 * i + 1], lo[4 * i + 2], lo[4 * i + 3]);
  }
}

// WRES ("weights resident, taps shared"): the k = 3 convolutions of the 128-channel layers (layer1 / layer2: 21 250 and
// 7 083 frames per utterance, 37 % of the RawNet2 step).  With N = K = 128 these GEMMs have the arithmetic intensity of
// a copy: the generic mainloop re-loads the 16 KB W tile with every A tile and loads every A row three times (one
// row-shifted box per tap), 96 KB from L2 per 128 rows x 64 channels, and ran at 55 % of the HBM peak.  Here all six W
// tiles (3 taps x 2 K chunks = 96 KB) are loaded ONCE per CTA, and one 136-row A box per K chunk serves the three taps
// through shared-memory descriptors whose start address is shifted by 0 / 1 / 2 rows (tcgen05 applies the 128B swizzle
// to absolute address bits, so any 128-byte aligned start row is legal: profiles/r01_desc_shift_probe.txt) -- 17 KB from
// L2 per 128 rows x 64 channels.
//
// conv2 (EPI_ADD) of these layers also takes its RESIDUAL through the tensor core: the shortcut tile [128 rows x 128 ch]
// is two more TMA boxes per tile, multiplied by an identity matrix that sits in shared memory next to W (fp16 x 1.0 in
// an fp32 accumulator is exact).  The epilogue then has no global load at all.  Before: 8 epilogue warps each waited for
// two dependent DRAM round trips per tile (ncu: issue active 11 %, DRAM 38 %, 3.08 ms for layer1's conv2).
static constexpr int WRES_A_ROWS = 136;
static constexpr int WRES_A_BYTES = WRES_A_ROWS * 128;      // 17 408 = 17 x 1024
static constexpr int WRES_W_BYTES = 6 * 128 * 128;          // 3 taps x 2 K chunks of [128 rows x 64 ch]
static constexpr int WRES_I_BYTES = 64 * 128;               // identity [64 x 64] (EPI_ADD only): one N = 64 MMA per column half
// Outputs of the W-resident variants leave through shared memory and TMA stores (one 32-row box per epilogue warp): a
// thread-per-row STG.128 touches 32 lines per instruction, and the LSU data pipe was the busiest unit of these kernels
// (ncu r02: l1tex__data_pipe_lsu_wavefronts 71 % for the sinc layer).
static constexpr int WRES_OUT_BYTES = 8 * 4096;
static constexpr int WRES_R_BYTES = 128 * 128;              // one residual box [128 rows x 64 ch]
// W-resident sinc variant (16-bit mode: one pass, K = 320 = 5 k-blocks): the five W boxes of the CTA's current n-tile stay
// in shared memory, the stage ring carries A boxes only.  The generic mainloop moved 40 KB per k-block from L2 (ncu
// r02: 40.6 GB L2->SM per launch = 9.4 TB/s, the limiter at 34 % tensor-pipe and 18 % DRAM utilisation); this one 16 KB.
static constexpr int WSINC_KB = 5;
static constexpr int WSINC_W_BYTES = WSINC_KB * 192 * 128;  // 122 880
static constexpr int WSINC_STAGES = 4;
static constexpr int WSINC_SMEM_BYTES = WSINC_W_BYTES + WSINC_STAGES * A_STAGE_BYTES + WRES_OUT_BYTES + 1024;
template <int EPI> struct WresCfg {
  static constexpr bool RES = EPI == 2;                     // EPI_ADD
  static constexpr int STAGES = 5;
  static constexpr int FIXED_BYTES = WRES_W_BYTES + (RES ? WRES_I_BYTES : 0);
  static constexpr int SMEM_BYTES = FIXED_BYTES + STAGES * WRES_A_BYTES + WRES_OUT_BYTES + 1024;
};

template <int BN, int EPI, bool WRES = false>
__global__ void __launch_bounds__(RnThreads<EPI, WRES>::N, 1)
rn_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
               const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmO,
               const __grid_constant__ CUtensorMap tmO2, const RnGemmParams p) {
  using Cfg = RnCfg<BN>;
  using WC = WresCfg<EPI>;
  constexpr int STAGES = WRES ? (EPI == EPI_SINC ? WSINC_STAGES : WC::STAGES) : Cfg::STAGES;
  constexpr bool RESMMA = WRES && WC::RES;                  // residual added by an identity MMA
  constexpr int AFF_MAX = 512;                              // affine parameters cached in shared memory up to this N
  constexpr int NUM_THREADS = RnThreads<EPI, WRES>::N;
  constexpr int EPI_WARPS = NUM_THREADS / 32 - 4;
  constexpr bool WCONV = WRES && EPI != EPI_SINC;           // 128-channel k = 3 layers, all of W resident
  constexpr bool WSINC = WRES && EPI == EPI_SINC;           // sinc layer, the W tile of ONE n-tile resident
  static_assert(!WCONV || BN == 128, "the W-resident conv variant serves the 128-channel layers");
  static_assert(!WSINC || BN == 192, "the W-resident sinc variant uses 192-column tiles");
  __shared__ __align__(8) uint64_t wfull_bar;
  __shared__ __align__(8) uint64_t wfree_bar;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[STAGES];
  __shared__ __align__(8) uint64_t empty_bar[STAGES];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float4 s_aff[EPI == 0 ? 128 : (EPI == 1 ? AFF_MAX / 2 : 1)];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;

  // affine parameters of the epilogue, once per CTA: sinc {s1, h1, s2, h2} per filter; conv1 {s[2i], h[2i], s[2i+1], h[2i+1]}
  if (EPI == 0) {
    for (int i = threadIdx.x; i < 128; i += NUM_THREADS) s_aff[i] = make_float4(p.s1[i], p.h1[i], p.s2[i], p.h2[i]);
  } else if (EPI == 1) {
    if (p.N <= AFF_MAX)
      for (int i = threadIdx.x; i < p.N / 2; i += NUM_THREADS)
        s_aff[i] = make_float4(p.s1[2 * i], p.h1[2 * i], p.s1[2 * i + 1], p.h1[2 * i + 1]);
  }
  if (RESMMA) {
    // 64 x 64 identity in the K-major 128B-swizzled operand layout: element (n, n) at byte
    // n * 128 + ((n / 8) ^ (n & 7)) * 16 + (n % 8) * 2
    const uint32_t i_smem = smem_base + WRES_W_BYTES;
    for (int i = threadIdx.x; i < WRES_I_BYTES / 16; i += NUM_THREADS)
      asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(i_smem + i * 16), "r"(0) : "memory");
    __syncthreads();
    if (threadIdx.x < 64) {
      const int n = threadIdx.x;
      const uint32_t addr = i_smem + n * 128 + (((n >> 3) ^ (n & 7)) << 4) + (n & 7) * 2;
      asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(static_cast<unsigned short>(0x3c00)) : "memory");   // 1.0 (IEEE half)
    }
    fence_proxy_async();
  }

  // pool3 (W-resident conv2 with the max-pool fused in): tiles never span utterances; tile i of an utterance starts at
  // its frame 126 i (row utt * Tp + 1 + 126 i) and holds 42 whole pooling windows -- the last two rows of a tile are
  // recomputed by the next one and the last tile of an utterance runs into rows it discards (1.6 % + < 1.5 % more MMAs)
  const bool pool3 = WRES && EPI == EPI_ADD && p.pool3 != 0;
  const int tpu = pool3 ? (p.Tout + 41) / 42 : 1;                 // tiles per utterance
  const int num_m_tiles = pool3 ? p.B * tpu : (p.M + BLOCK_M - 1) / BLOCK_M;
  auto tile_row0 = [&](int mt) { return pool3 ? (mt / tpu) * p.Tp + 1 + (mt % tpu) * 126 : mt * BLOCK_M; };
  const int num_n_tiles = (p.N + BN - 1) / BN;
  const int num_tiles = num_m_tiles * num_n_tiles;
  const int kc = p.kc_per_tap;
  const int kb_per_tap = p.num_pass * kc;
  const int num_kb = p.num_taps * kb_per_tap;
  // tile schedule: interleaved over the CTAs with n fastest (consecutive CTAs share the A tile in L2) -- except WSINC,
  // where every CTA takes ONE contiguous run of the n-major tile order so that its W tile changes at most twice
  const int t_per = WSINC ? (num_tiles + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x) : 0;
  const int t_begin = WSINC ? static_cast<int>(blockIdx.x) * t_per : static_cast<int>(blockIdx.x);
  const int t_end = WSINC ? min(num_tiles, t_begin + t_per) : num_tiles;
  const int t_step = WSINC ? 1 : static_cast<int>(gridDim.x);
  auto tile_m = [&](int tile) { return WSINC ? tile % num_m_tiles : tile / num_n_tiles; };
  auto tile_n = [&](int tile) { return WSINC ? tile / num_m_tiles : tile % num_n_tiles; };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&tfull_bar[a]), 1);
      mbar_init(smem_u32(&tempty_bar[a]), EPI_WARPS);
    }
    mbar_init(smem_u32(&wfull_bar), 1);
    mbar_init(smem_u32(&wfree_bar), 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(smem_u32(&tmem_slot), Cfg::TMEM_COLS);
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
      uint32_t phase = 0;
      if (WSINC) {
        const uint32_t wb = smem_u32(&wfull_bar);
        int cur_nt = -1;
        uint32_t wfree_phase = 0;
        for (int tile = t_begin; tile < t_end; ++tile) {
          const int m0 = tile_m(tile) * BLOCK_M, nt = tile_n(tile);
          if (nt != cur_nt) {
            if (cur_nt >= 0) { mbar_wait(smem_u32(&wfree_bar), wfree_phase); wfree_phase ^= 1u; }   // old W no longer read
            mbar_arrive_expect_tx(wb, WSINC_W_BYTES);
            for (int kb = 0; kb < WSINC_KB; ++kb) tma_load_2d(smem_base + kb * Cfg::B_STAGE_BYTES, &tmW, wb, kb * BLOCK_K, nt * BN);
            cur_nt = nt;
          }
          for (int kb = 0; kb < WSINC_KB; ++kb) {
            const uint32_t fb = smem_u32(&full_bar[stage]);
            mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
            mbar_arrive_expect_tx(fb, A_STAGE_BYTES);
            tma_load_2d(smem_base + WSINC_W_BYTES + stage * A_STAGE_BYTES, &tmA, fb, kb * BLOCK_K, m0 + p.tap_off[0]);
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          }
        }
      }
      if (WCONV) {
        const uint32_t wb = smem_u32(&wfull_bar);
        mbar_arrive_expect_tx(wb, WRES_W_BYTES);
        for (int j = 0; j < 6; ++j) tma_load_2d(smem_base + j * 16384, &tmW, wb, j * BLOCK_K, 0);    // j = tap * 2 + kcx
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
          const int m0 = tile_row0(tile);
          for (int kcx = 0; kcx < 2; ++kcx) {
            const uint32_t fb = smem_u32(&full_bar[stage]);
            mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
            mbar_arrive_expect_tx(fb, WRES_A_BYTES);
            tma_load_2d(smem_base + WC::FIXED_BYTES + stage * WRES_A_BYTES, &tmA, fb, kcx * BLOCK_K, m0 - 1);
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          }
          for (int kcx = 0; kcx < 2 && RESMMA; ++kcx) {
            const uint32_t fb = smem_u32(&full_bar[stage]);
            mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
            mbar_arrive_expect_tx(fb, WRES_R_BYTES);
            tma_load_2d(smem_base + WC::FIXED_BYTES + stage * WRES_A_BYTES, &tmR, fb, kcx * BLOCK_K, m0);
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          }
        }
      }
      for (int tile = blockIdx.x; tile < num_tiles && !WRES; tile += gridDim.x) {
        // n fastest: consecutive CTAs share the A tile (L2), W is small and stays cache resident
        const int m0 = tile_m(tile) * BLOCK_M;
        const int n0 = tile_n(tile) * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          const int tap = kb / kb_per_tap;
          const int rem = kb - tap * kb_per_tap;
          const int pass_i = rem / kc;
          const int kcx = rem - pass_i * kc;
          // 0: hi*hi   1: hi(A)*lo(W)   2: lo(A)*hi(W); two passes = split activations, bf16 weights (0 and 2)
          const int pass = (p.num_pass == 2 && pass_i == 1) ? 2 : pass_i;
          const int aplane = (pass == 2) ? 1 : 0;
          const int wplane = (pass == 1) ? 1 : 0;
          const int a_col = aplane * p.a_plane_cols + kcx * BLOCK_K;
          const int a_row = m0 + p.tap_off[tap];
          const int w_col = ((tap * p.w_planes + wplane) * kc + kcx) * BLOCK_K;
          const uint32_t fb = smem_u32(&full_bar[stage]);
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
          mbar_arrive_expect_tx(fb, Cfg::STAGE_BYTES);
          const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
          tma_load_2d(sa, &tmA, fb, a_col, a_row);
          tma_load_2d(sa + A_STAGE_BYTES, &tmW, fb, w_col, n0);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (elect_one()) {
      // single plane = IEEE half on both operands (a kind::f16 UMMA cannot mix formats); split mode = bf16 planes
      const uint32_t idesc = umma_idesc_f16(BLOCK_M, BN, p.w_planes == 1, p.w_planes == 1);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      if (WSINC) {
        int cur_nt = -1;
        uint32_t wfull_phase = 0;
        for (int tile = t_begin; tile < t_end; ++tile, ++it) {
          const int nt = tile_n(tile);
          if (nt != cur_nt) { mbar_wait(smem_u32(&wfull_bar), wfull_phase); wfull_phase ^= 1u; cur_nt = nt; }
          const int acc = it & 1;
          const uint32_t acc_phase = (it >> 1) & 1u;
          mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1u);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * BN;
          for (int kb = 0; kb < WSINC_KB; ++kb) {
            mbar_wait(smem_u32(&full_bar[stage]), phase);
            tc_fence_after();
            const uint64_t adesc = umma_desc_sw128(smem_base + WSINC_W_BYTES + stage * A_STAGE_BYTES);
            const uint64_t bdesc = umma_desc_sw128(smem_base + kb * Cfg::B_STAGE_BYTES);
            // the shifted filters end at column 250 + 47 = 297: the last 16 of the 320 K columns are zero in every W row
            const int ksteps = kb == WSINC_KB - 1 ? BLOCK_K / 16 - 1 : BLOCK_K / 16;
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k)
              if (k < ksteps) umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) ? 1u : 0u);
            umma_commit(smem_u32(&empty_bar[stage]));
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          }
          umma_commit(smem_u32(&tfull_bar[acc]));
          if (tile + 1 < t_end && tile_n(tile + 1) != cur_nt) umma_commit(smem_u32(&wfree_bar));
        }
      }
      if (WCONV) {
        mbar_wait(smem_u32(&wfull_bar), 0);
        tc_fence_after();
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
          const int acc = it & 1;
          const uint32_t acc_phase = (it >> 1) & 1u;
          mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1u);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * BN;
          uint32_t accf = 0;
          for (int kcx = 0; kcx < 2; ++kcx) {
            mbar_wait(smem_u32(&full_bar[stage]), phase);
            tc_fence_after();
            const uint32_t sa = smem_base + WC::FIXED_BYTES + stage * WRES_A_BYTES;
#pragma unroll
            for (int tap = 0; tap < 3; ++tap) {
              const uint64_t adesc = umma_desc_sw128(sa + tap * 128);                      // rows m0 - 1 + tap ...
              const uint64_t bdesc = umma_desc_sw128(smem_base + (tap * 2 + kcx) * 16384);
#pragma unroll
              for (int k = 0; k < BLOCK_K / 16; ++k) { umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, accf); accf = 1; }
            }
            umma_commit(smem_u32(&empty_bar[stage]));
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          }
          for (int kcx = 0; kcx < 2 && RESMMA; ++kcx) {
            mbar_wait(smem_u32(&full_bar[stage]), phase);
            tc_fence_after();
            const uint64_t adesc = umma_desc_sw128(smem_base + WC::FIXED_BYTES + stage * WRES_A_BYTES);
            // channels 64 kcx .. 64 kcx + 63 of the residual land in the same accumulator columns: N = 64, D shifted
            const uint64_t bdesc = umma_desc_sw128(smem_base + WRES_W_BYTES);
            const uint32_t idesc64 = umma_idesc_f16(BLOCK_M, 64, 1, 1);
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k) umma_bf16(d_tmem + kcx * 64, adesc + 2 * k, bdesc + 2 * k, idesc64, 1u);
            umma_commit(smem_u32(&empty_bar[stage]));
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          }
          umma_commit(smem_u32(&tfull_bar[acc]));
        }
      }
      for (int tile = blockIdx.x; tile < num_tiles && !WRES; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1u;
        mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
          const uint64_t adesc = umma_desc_sw128(sa);
          const uint64_t bdesc = umma_desc_sw128(sa + A_STAGE_BYTES);
#pragma unroll
          for (int k = 0; k < BLOCK_K / 16; ++k)
            umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) ? 1u : 0u);
          umma_commit(smem_u32(&empty_bar[stage]));
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit(smem_u32(&tfull_bar[acc]));
      }
    }
  } else if (warp >= 4) {
    // ===================================================== epilogue
    const int ew = warp - 4;
    const int quad = ew & 3;   // TMEM lane quadrant of this warp
    const int half = ew >> 2;  // column half of the tile (sinc: column group of the warp)
    // this warp's staging box behind the stage ring: 4 KB in the W-resident variants, 2 KB (32 rows x 32 columns, 64B
    // swizzle) in the generic ones
    const uint32_t out_box = WRES ? smem_base + (WSINC ? WSINC_W_BYTES + STAGES * A_STAGE_BYTES : WC::FIXED_BYTES + STAGES * WRES_A_BYTES) + ew * 4096
                                  : smem_base + STAGES * Cfg::STAGE_BYTES + (ew & 7) * 2048;
    // generic instances, 16-bit mode: 32 columns of this thread's row -> the warp's box -> one TMA store (all lanes call)
    auto tma_chunk = [&](int gcol, int grow, const float (&v)[32]) {
      uint32_t pk[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) pk[i] = pack_f16x2(v[2 * i], v[2 * i + 1]);
      if (lane == 0) bulk_wait_read<0>();
      __syncwarp();
      const uint32_t my_row = out_box + lane * 64;
      const uint32_t sw = (lane >> 1) & 3;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + ((j ^ sw) << 4)), "r"(pk[4 * j]),
                     "r"(pk[4 * j + 1]), "r"(pk[4 * j + 2]), "r"(pk[4 * j + 3]) : "memory");
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        tma_store_2d(&tmO, out_box, gcol, grow);
        bulk_commit();
      }
    };
    int it = 0;
    for (int tile = t_begin; tile < t_end; tile += t_step, ++it) {
      const int m0 = tile_row0(tile_m(tile));
      const int nt = tile_n(tile);
      const int n0 = nt * BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1u;
      const long long row = m0 + quad * 32 + lane;
      const uint32_t tbase = tmem_base + acc * BN + (static_cast<uint32_t>(quad * 32) << 16);

      if (EPI == EPI_SINC) {
        // row = window index; this tile = pooled output jg of 64 filters; this warp half = 32 of them
        const int utt = static_cast<int>(row / p.Mu);
        const int ml = static_cast<int>(row - static_cast<long long>(utt) * p.Mu);
        const int jg = nt >> 1;
        const int t_out = 16 * ml + jg;
        const bool valid = row < p.M && utt < p.B && t_out < p.T;
        const long long orow = static_cast<long long>(utt) * p.Tp + 1 + t_out;
        mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
        tc_fence_after();
        // the tile's 192 columns are 4 pieces of 48 = 16 filters x 3 pooling phases; a warp owns 4 / (EPI_WARPS / 4) of them
        constexpr int PPW = 16 / EPI_WARPS;
        if (WSINC) {
          // 8 warps: 32 rows x 32 filters per warp and output, staged as 64-byte rows (64B swizzle) and stored by TMA
          // through the (channel, jg, ml, utt) view; frames of the last, partial ml of an utterance go out directly
          uint32_t xk[16], ak[16];
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            uint32_t ra[32];
            uint32_t rb[16];
            const uint32_t col = (half * 2 + h2) * 48;
            tmem_ld32(tbase + col, ra);
            tmem_ld16(tbase + col + 32, rb);
            tmem_ld_wait();
            if (h2 == 1) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
            }
            const int f = (nt & 1) * 64 + (half * 2 + h2) * 16;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              float xv[2], av[2];
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                const int j = 3 * (2 * i + e);
                const float c0 = fabsf(__uint_as_float(j < 32 ? ra[j < 32 ? j : 0] : rb[j >= 32 ? j - 32 : 0]));
                const float c1 = fabsf(__uint_as_float(j + 1 < 32 ? ra[j + 1 < 32 ? j + 1 : 0] : rb[j + 1 >= 32 ? j + 1 - 32 : 0]));
                const float c2 = fabsf(__uint_as_float(j + 2 < 32 ? ra[j + 2 < 32 ? j + 2 : 0] : rb[j + 2 >= 32 ? j + 2 - 32 : 0]));
                const float4 af = s_aff[f + 2 * i + e];
                xv[e] = lrelu03(fmaf(fmaxf(c0, fmaxf(c1, c2)), af.x, af.y));
                av[e] = lrelu03(fmaf(xv[e], af.z, af.w));
              }
              xk[h2 * 8 + i] = pack_f16x2(xv[0], xv[1]);
              ak[h2 * 8 + i] = pack_f16x2(av[0], av[1]);
            }
          }
          // the TMA box covers the rows of the warp's FIRST utterance with ml < T / 16 (the rest is clipped by the view's
          // extents: a negative start coordinate for the next utterance's rows faults on store, measured); the last,
          // partial ml of an utterance and a warp's rows in the next utterance are stored directly
          const int ml_full = p.T >> 4;
          const int utt_w = __shfl_sync(0xffffffffu, utt, 0);
          if (valid && (ml >= ml_full || utt != utt_w)) {
            const int f = (nt & 1) * 64 + half * 32;
            uint4* dx = reinterpret_cast<uint4*>(p.out + orow * p.out_ld + f);
            uint4* da = reinterpret_cast<uint4*>(p.out2 + orow * p.out2_ld + f);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              dx[i] = make_uint4(xk[4 * i], xk[4 * i + 1], xk[4 * i + 2], xk[4 * i + 3]);
              da[i] = make_uint4(ak[4 * i], ak[4 * i + 1], ak[4 * i + 2], ak[4 * i + 3]);
            }
          }
          if (lane == 0) bulk_wait_read<0>();
          __syncwarp();
          const uint32_t xrow = out_box + lane * 64, arow = xrow + 2048;
          const uint32_t sw = (lane >> 1) & 3;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(xrow + ((j ^ sw) << 4)), "r"(xk[4 * j]),
                         "r"(xk[4 * j + 1]), "r"(xk[4 * j + 2]), "r"(xk[4 * j + 3]) : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(arow + ((j ^ sw) << 4)), "r"(ak[4 * j]),
                         "r"(ak[4 * j + 1]), "r"(ak[4 * j + 2]), "r"(ak[4 * j + 3]) : "memory");
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            const long long rw0 = static_cast<long long>(m0) + quad * 32;
            const int utt0 = static_cast<int>(rw0 / p.Mu);
            const int ml0 = static_cast<int>(rw0 - static_cast<long long>(utt0) * p.Mu);
            const int c0 = (nt & 1) * 64 + half * 32;
            tma_store_4d(&tmO, out_box, c0, jg, ml0, utt0);
            tma_store_4d(&tmO2, out_box + 2048, c0, jg, ml0, utt0);
            bulk_commit();
          }
          continue;
        }
#pragma unroll
        for (int h2 = 0; h2 < PPW; ++h2) {
          const int piece = half * PPW + h2;
          uint32_t ra[32];
          uint32_t rb[16];
          const uint32_t col = piece * 48;
          tmem_ld32(tbase + col, ra);
          tmem_ld16(tbase + col + 32, rb);
          tmem_ld_wait();
          if (h2 == PPW - 1) {          // the accumulator is in registers: hand the TMEM buffer back before the arithmetic
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
          }
          if (valid) {
            float x0[16], a0[16];
            const int f = (nt & 1) * 64 + piece * 16;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              float c[3];
#pragma unroll
              for (int q = 0; q < 3; ++q) {
                const int j = 3 * i + q;
                c[q] = fabsf(__uint_as_float(j < 32 ? ra[j < 32 ? j : 0] : rb[j >= 32 ? j - 32 : 0]));
              }
              const float mx = fmaxf(c[0], fmaxf(c[1], c[2]));
              const float4 af = s_aff[f + i];
              const float x = lrelu03(fmaf(mx, af.x, af.y));
              x0[i] = x;
              a0[i] = lrelu03(fmaf(x, af.z, af.w));
            }
            store_planes<16>(p.out + orow * p.out_ld + f, p.out_planes, p.out_plane_cols, x0);
            store_planes<16>(p.out2 + orow * p.out2_ld + f, p.out2_planes, p.out2_plane_cols, a0);
          }
        }
      } else {
        const int utt = static_cast<int>(row / p.Tp);
        const int t = static_cast<int>(row - static_cast<long long>(utt) * p.Tp) - 1;
        const bool in_m = row < p.M;
        const bool valid = in_m && t >= 0 && t < p.T;
        if (EPI == EPI_ADD && !RESMMA) {
          // the residual's hi plane (the only one in the 16-bit mode) is requested BEFORE the accumulator is waited for:
          // its DRAM latency hides under the tile's MMAs instead of adding one round trip per 32-column chunk
          uint4 res[BN / 64][4];
          const bf16* ap0 = p.add + row * p.add_ld + n0 + half * (BN / 2);
#pragma unroll
          for (int c = 0; c < BN / 64; ++c)
#pragma unroll
            for (int i = 0; i < 4; ++i)
              res[c][i] = (valid && n0 + half * (BN / 2) + c * 32 < p.N) ? __ldg(reinterpret_cast<const uint4*>(ap0 + c * 32) + i)
                                                                         : make_uint4(0u, 0u, 0u, 0u);
          mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < BN / 64; ++c) {
            const int col0 = half * (BN / 2) + c * 32;
            uint32_t r[32];
            tmem_ld32(tbase + col0, r);
            tmem_ld_wait();
            const int n = n0 + col0;
            if (n >= p.N || (!valid && !p.tma_out)) continue;
            float v[32];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t w[4] = {res[c][i].x, res[c][i].y, res[c][i].z, res[c][i].w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                v[8 * i + 2 * q] = __uint_as_float(r[8 * i + 2 * q]) + (p.add_planes == 1 ? f16_lo_f(w[q]) : bf16_lo_f(w[q]));
                v[8 * i + 2 * q + 1] = __uint_as_float(r[8 * i + 2 * q + 1]) + (p.add_planes == 1 ? f16_hi_f(w[q]) : bf16_hi_f(w[q]));
              }
            }
            if (p.add_planes == 2) {
              const bf16* ap = p.add + row * p.add_ld + n;
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 u = __ldg(reinterpret_cast<const uint4*>(ap + p.add_plane_cols) + i);
                const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  v[8 * i + 2 * q] += bf16_lo_f(w[q]);
                  v[8 * i + 2 * q + 1] += bf16_hi_f(w[q]);
                }
              }
            }
            if (p.tma_out) {                    // single plane; padding rows are written as zeros
              if (!valid) {
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = 0.f;
              }
              tma_chunk(n, m0 + quad * 32, v);
            } else {
              store_planes<32>(p.out + row * p.out_ld + n, p.out_planes, p.out_plane_cols, v);
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
          continue;
        }
        mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
        tc_fence_after();
        // one 32-column chunk of this thread's row: affine + LeakyReLU (conv1) or plain (shortcut conv, and conv2 whose
        // residual the identity MMA already added), packed and stored
        auto emit = [&](int col0, const uint32_t (&r)[32]) {
          const int n = n0 + col0;
          if (n >= p.N || (!in_m && !p.tma_out)) return;
          float v[32];
          if (EPI == EPI_AFF_LRELU) {
            // zero rows stay zero: they are the next conv's padding
            if (p.N <= AFF_MAX) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const float4 af = s_aff[(n >> 1) + i];
                v[2 * i] = valid ? lrelu03(fmaf(__uint_as_float(r[2 * i]), af.x, af.y)) : 0.f;
                v[2 * i + 1] = valid ? lrelu03(fmaf(__uint_as_float(r[2 * i + 1]), af.z, af.w)) : 0.f;
              }
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                v[i] = valid ? lrelu03(fmaf(__uint_as_float(r[i]), __ldg(p.s1 + n + i), __ldg(p.h1 + n + i))) : 0.f;
            }
          } else {
            if (!valid && !p.tma_out) return;
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = valid ? __uint_as_float(r[i]) : 0.f;
          }
          if (p.tma_out) tma_chunk(n, m0 + quad * 32, v);
          else store_planes<32>(p.out + row * p.out_ld + n, p.out_planes, p.out_plane_cols, v);
        };
        if (WCONV) {
          // 64 columns of this thread's row into registers, TMEM handed back, affine + LeakyReLU (conv1) or nothing
          // (conv2: the identity MMA already added the residual), packed into the warp's 32 x 64 box (128-byte rows,
          // 128B swizzle) and stored by TMA.  Padding rows are written as zeros, rows beyond M are clipped by the map.
          uint32_t r0[32], r1[32];
          tmem_ld32(tbase + half * 64, r0);
          tmem_ld32(tbase + half * 64 + 32, r1);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
          uint32_t pk[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            float v0 = __uint_as_float(i < 16 ? r0[2 * (i & 15)] : r1[2 * (i & 15)]);
            float v1 = __uint_as_float(i < 16 ? r0[2 * (i & 15) + 1] : r1[2 * (i & 15) + 1]);
            if (EPI == EPI_AFF_LRELU) {
              const float4 af = s_aff[half * 32 + i];
              v0 = lrelu03(fmaf(v0, af.x, af.y));
              v1 = lrelu03(fmaf(v1, af.z, af.w));
            }
            pk[i] = (valid || pool3) ? pack_f16x2(v0, v1) : 0u;
          }
          const uint32_t my_row = out_box + lane * 128;
          const uint32_t sw = lane & 7;
          if (pool3) {
            // the four warps of this column half share one 128-row x 64-column staging tile (their boxes are adjacent):
            // rows in, barrier, then (window, 16-byte chunk) items: three rows -> max -> one 16-byte store of the pooled
            // row (8 lanes cover a 128-byte line).  Windows of the padding triple at the end of an utterance are skipped.
            named_bar_sync(1 + half, 128);                     // everybody is done reading the previous tile's rows
#pragma unroll
            for (int j = 0; j < 8; ++j)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + ((j ^ sw) << 4)), "r"(pk[4 * j]),
                           "r"(pk[4 * j + 1]), "r"(pk[4 * j + 2]), "r"(pk[4 * j + 3]) : "memory");
            named_bar_sync(1 + half, 128);
            const uint32_t tile_s = out_box - quad * 4096;     // first row of this half's staging tile
            const int mt = tile_m(tile);
            const int utt_w = mt / tpu, j_first = (mt - utt_w * tpu) * 42;
#pragma unroll
            for (int k2 = 0; k2 < 3; ++k2) {
              const int idx = quad * 32 + lane + 128 * k2;
              const int wl = idx >> 3, c = idx & 7;
              if (wl >= 42) break;
              uint4 q[3];
#pragma unroll
              for (int r3 = 0; r3 < 3; ++r3) {
                const int rr = 3 * wl + r3;
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(q[r3].x), "=r"(q[r3].y), "=r"(q[r3].z), "=r"(q[r3].w)
                             : "r"(tile_s + rr * 128 + ((c ^ (rr & 7)) << 4)) : "memory");
              }
              const int jw = j_first + wl;
              if (jw < p.Tout) {
                uint4 o;
                o.x = hmax2_u32(q[0].x, hmax2_u32(q[1].x, q[2].x));
                o.y = hmax2_u32(q[0].y, hmax2_u32(q[1].y, q[2].y));
                o.z = hmax2_u32(q[0].z, hmax2_u32(q[1].z, q[2].z));
                o.w = hmax2_u32(q[0].w, hmax2_u32(q[1].w, q[2].w));
                bf16* dst = p.out + (static_cast<long long>(utt_w) * (p.Tout + 2) + 1 + jw) * p.out_ld + half * 64 + c * 8;
                *reinterpret_cast<uint4*>(dst) = o;
              }
            }
            continue;
          }
          if (lane == 0) bulk_wait_read<0>();
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + ((j ^ sw) << 4)), "r"(pk[4 * j]),
                         "r"(pk[4 * j + 1]), "r"(pk[4 * j + 2]), "r"(pk[4 * j + 3]) : "memory");
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmO, out_box, half * 64, m0 + quad * 32);
            bulk_commit();
          }
          continue;
        }
        if (BN == 128) {
          // both chunks into registers, TMEM buffer handed back, then the arithmetic and the stores
          uint32_t r0[32], r1[32];
          tmem_ld32(tbase + half * 64, r0);
          tmem_ld32(tbase + half * 64 + 32, r1);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
          emit(half * 64, r0);
          emit(half * 64 + 32, r1);
          continue;
        }
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c) {
          const int col0 = half * (BN / 2) + c * 32;
          uint32_t r[32];
          tmem_ld32(tbase + col0, r);
          tmem_ld_wait();
          emit(col0, r);
        }
      }
      if (EPI == EPI_SINC) continue;            // released inside
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
    }
    if (lane == 0) bulk_wait<0>();               // the last TMA stores read shared memory of this CTA
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ------------------------------------------------------------------ host side
template <int BN, int EPI, bool WRES = false>
static int launch_rn_t(const CUtensorMap& a, const CUtensorMap& w, const CUtensorMap& r, const CUtensorMap& o,
                       const CUtensorMap& o2, const RnGemmParams& p, cudaStream_t stream) {
  using Cfg = RnCfg<BN>;
  constexpr int SMEM = WRES ? (EPI == EPI_SINC ? WSINC_SMEM_BYTES : WresCfg<EPI>::SMEM_BYTES) : Cfg::SMEM_BYTES;
  static unsigned long long attr_devs = 0;   // per device: kernel attributes do not carry over
  if (first_on_device(&attr_devs)) {
    SVB_CUDA_OK(cudaFuncSetAttribute(rn_gemm_kernel<BN, EPI, WRES>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  }
  const int tiles = p.pool3 ? p.B * ((p.Tout + 41) / 42) : ((p.M + BLOCK_M - 1) / BLOCK_M) * ((p.N + BN - 1) / BN);
  const int grid = tiles < num_sms() ? tiles : num_sms();
  rn_gemm_kernel<BN, EPI, WRES><<<grid, RnThreads<EPI, WRES>::N, SMEM, stream>>>(a, w, r, o, o2, p);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// conv2 + max_pool1d(3) in one kernel: 128 channels, 16-bit mode
bool rn_conv2_pool_supported(int channels, int planes, int xplanes, int T) {
  static const bool on = [] {
    const char* w = getenv("SVB_RN_WRES");
    const char* f = getenv("SVB_RN_POOLFUSE");
    return (w == nullptr || w[0] != '0') && (f == nullptr || f[0] != '0');
  }();
  return on && channels == 128 && planes == 1 && xplanes == 1 && T >= 3;
}

// a: activation matrix (or the sliding-window waveform view when epi == EPI_SINC: a_ld = 48, a_cols = extent)
int launch_rn_gemm(int epi, const bf16* a, long long a_rows, long long a_cols, long long a_ld, const bf16* w,
                   int w_rows, int w_cols, const RnGemmParams& p, cudaStream_t stream) {
  SVB_CHECK(p.num_pass >= 1 && p.num_pass <= 3, "num_pass must be 1, 2 or 3");
  CUtensorMap ma, mw;
  if (make_map(&ma, a, a_rows, a_cols, a_ld, BLOCK_M)) return -1;
  if (epi == EPI_SINC) {
    SVB_CHECK(p.N % 192 == 0, "sinc GEMM: N must be a multiple of 192");
    if (make_map(&mw, w, w_rows, w_cols, w_cols, 192)) return -1;
    static const bool wsinc_on = [] { const char* e = getenv("SVB_RN_WSINC"); return e == nullptr || e[0] != '0'; }();
    if (wsinc_on && p.num_pass == 1 && p.num_taps == 1 && p.kc_per_tap == WSINC_KB && p.out_planes == 1 &&
        p.out2_planes == 1 && p.T >= 16) {
      // store views (channel, jg, ml, utt) of the two outputs, from row 1 (frame 0) of utterance 0
      CUtensorMap mo, mo2;
      if (make_map_sinc_out(&mo, p.out + p.out_ld, 128, p.out_ld, p.T / 16, p.Tp, p.B)) return -1;
      if (make_map_sinc_out(&mo2, p.out2 + p.out2_ld, 128, p.out2_ld, p.T / 16, p.Tp, p.B)) return -1;
      return launch_rn_t<192, EPI_SINC, true>(ma, mw, ma, mo, mo2, p, stream);
    }
    return launch_rn_t<192, EPI_SINC>(ma, mw, ma, ma, ma, p, stream);
  }
  SVB_CHECK(p.N % 128 == 0, "RawNet conv GEMM: N must be a multiple of 128");
  const int bn = (p.N % 256 == 0) ? 256 : 128;
  if (make_map(&mw, w, w_rows, w_cols, w_cols, bn)) return -1;
  // 128-channel k = 3 layers in the 16-bit mode: W-resident, tap-shared mainloop (SVB_RN_WRES=0: generic mainloop)
  static const bool wres_on = [] { const char* e = getenv("SVB_RN_WRES"); return e == nullptr || e[0] != '0'; }();
  if (wres_on && p.N == 128 && p.num_taps == 3 && p.kc_per_tap == 2 && p.num_pass == 1 && p.w_planes == 1 &&
      p.tap_off[0] == -1 && p.tap_off[1] == 0 && p.tap_off[2] == 1 &&
      p.out_planes == 1 && (epi == EPI_AFF_LRELU || (epi == EPI_ADD && p.add_planes == 1 && p.add_ld == 128))) {
    CUtensorMap ma136, mr, mo;
    if (make_map(&ma136, a, a_rows, a_cols, a_ld, WRES_A_ROWS)) return -1;
    if (p.pool3) {
      SVB_CHECK(epi == EPI_ADD && p.Tout == p.T / 3 && p.Tout >= 1 && p.B >= 1, "fused max-pool: conv2 with Tout = T / 3 pooled frames");
      mo = ma136;                                                             // pooled rows are stored directly
    } else if (make_map(&mo, p.out, p.M, 128, p.out_ld, 32)) return -1;       // store boxes: 32 rows x 64 channels
    if (epi == EPI_AFF_LRELU) return launch_rn_t<128, EPI_AFF_LRELU, true>(ma136, mw, ma136, mo, mo, p, stream);
    if (make_map(&mr, p.add, a_rows, 128, p.add_ld, BLOCK_M)) return -1;      // the shortcut has the rows of the output
    return launch_rn_t<128, EPI_ADD, true>(ma136, mw, mr, mo, mo, p, stream);
  }
  // generic mainloop; in the 16-bit mode (one output plane) the tile leaves through 32 x 32 TMA store boxes
  static const bool tma_on = [] { const char* e = getenv("SVB_RN_TMA_OUT"); return e == nullptr || e[0] != '0'; }();
  RnGemmParams q = p;
  CUtensorMap mo = ma;
  q.tma_out = (tma_on && p.out_planes == 1) ? 1 : 0;
  if (q.tma_out && make_map_sw64(&mo, p.out, p.M, p.N, p.out_ld, 32)) return -1;
  SVB_CHECK(!p.pool3, "fused max-pool needs the W-resident conv2 (rn_conv2_pool_supported)");
  if (bn == 256) {
    if (epi == EPI_AFF_LRELU) return launch_rn_t<256, EPI_AFF_LRELU>(ma, mw, ma, mo, mo, q, stream);
    if (epi == EPI_ADD) return launch_rn_t<256, EPI_ADD>(ma, mw, ma, mo, mo, q, stream);
    return launch_rn_t<256, EPI_PLAIN>(ma, mw, ma, mo, mo, q, stream);
  }
  if (epi == EPI_AFF_LRELU) return launch_rn_t<128, EPI_AFF_LRELU>(ma, mw, ma, mo, mo, q, stream);
  if (epi == EPI_ADD) return launch_rn_t<128, EPI_ADD>(ma, mw, ma, mo, mo, q, stream);
  return launch_rn_t<128, EPI_PLAIN>(ma, mw, ma, mo, mo, q, stream);
}

}  // namespace svb
