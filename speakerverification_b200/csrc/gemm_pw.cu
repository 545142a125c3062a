// Pointwise (k = 1) TDNN layers of ECAPA-TDNN -- tdnn1 / tdnn2 / mfa / asp.tdnn -- as a persistent
// warp-specialised tcgen05 GEMM with a lean epilogue (reference: TDNNBlock = norm(activation(conv(x))),
// src/models/ECAPA_TDNN.py:27-69; 92 % of the encoder's FLOPs are these layers).
//
//   out[r, n] = BN( act( sum_c A[r, c] * W[n, c] + bias[n] (+ rowbias[utt(r), n]) ) )  (+ tanh)
//
// Same mainloop as gemm_tc.cu (TMA -> 128B-swizzled smem ring -> tcgen05.mma, fp32 accumulators double
// buffered in TMEM).  What differs is the epilogue, which was the bottleneck of the K=1024 layers
// (profiles/r01_ncu_tdnn_v0.md: 41 instructions per output element, IPC 1.6, tensor pipe 32 % busy):
//   * activation / output kind / row-bias are template parameters (no per-element runtime branches);
//   * GELU(erf) is gelu_fast (11 FP32 + 2 MUFU, abs err 3.3e-7) instead of erff (~25 instructions);
//   * results are packed to bf16 in registers, staged per warp in a 4 KB 128B-swizzled smem box and
//     written with ONE TMA store per 32 x 64 block (coalesced, asynchronous) instead of 32 uncoalesced
//     16-byte row stores per warp instruction.
// Output storage: HILO instances write bf16 hi|lo planes (fp32-grade mode); the single-plane instances write IEEE half
// (16-bit mode, OUT_F16) -- there is no single-plane bf16 variant of this kernel (gemm_tc.cu keeps one).
// Pointwise layers compute halo rows in place (the input halo rows mirror valid rows and the op is
// per-row), so every row of a tile is stored and no mirror logic is needed.
#include <stdlib.h>

#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

static constexpr int BLOCK_M = 128;
static constexpr int BLOCK_K = 64;
static constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;
// warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps 4.. epilogue.  EW = 8 epilogue warps (two per TMEM lane
// quadrant, 128B-swizzled 32 x 64 staging boxes) or, for the GELU instances with BN = 256, EW = 16 (four per quadrant,
// 64B-swizzled 32 x 32 boxes): with K = 1024 the epilogue of 8 warps took 1.5x the MMA time of a tile and is latency
// bound (2 warps per scheduler); with 16 warps it idles 25 % of the time (profiles/r01_gemm_pw.md), which is the
// headroom the fused column statistics use.  (Measured and rejected: cp.async.bulk.prefetch.tensor.L2 of the next
// tile's A rows, burst or interleaved with the loads -- 7-9 % SLOWER on every shape.)
static constexpr int OUT_BOX_BYTES = 32 * 64 * 2;   // EW = 8: one warp's 32 rows x 64 bf16 columns
static constexpr int OUT_BOX64_BYTES = 32 * 32 * 2; // EW = 16: one warp's 32 rows x 32 bf16 columns
static constexpr int ACT_RAW_F32 = 3;               // epilogue variant: no bias/act/BN, fp32 accumulators stored as is
static constexpr int ACT_FILTER = 5;                // epilogue variant: nothing is stored but the accumulators >= a per-row threshold
static constexpr int ACT_GELU_ERF = 4;              // GELU through erff: the k > 1 layer (block 0) keeps gemm_tc's exact numerics
__device__ __forceinline__ float gelu_erf_exact(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }

template <int BN, int EW = 8, bool TWO = false>
struct PwCfg {
  static constexpr int NUM_THREADS = 128 + EW * 32;
  static constexpr int B_STAGE_BYTES = (TWO ? BN / 2 : BN) * BLOCK_K * 2;   // TWO: this CTA's half of the W tile
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int STAGES = (BN == 256 && !TWO) ? 4 : 6;
  static constexpr int OUT_BYTES = 8 * OUT_BOX_BYTES;   // EW = 16: one 32 x 64 box per warp PAIR
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + OUT_BYTES + 1024;
  static constexpr int TMEM_COLS = 2 * BN;
};

struct PwParams {
  int M, N;
  int kc;                // 64-column K chunks per plane
  int num_pass;          // 1: hi*hi ; 3: hi*hi + hi*lo + lo*hi
  int a_plane_cols, a2_kc, w_planes;
  const float* bias;
  const float* scale;
  const float* shift;
  const float* rowbias;  // [utterances, rowbias_ld] (RBT instances only)
  int rowbias_ld;
  const RowMeta* row_meta;
  int out_plane_cols;    // column offset of the lo plane in the output (HILO instances only)
  int num_taps;          // k > 1: the K loop runs tap-major over row-shifted A loads (same order as gemm_tc.cu)
  int tap_off[5];
  long long* colsum;     // STATS instances: fixed-point per-utterance column sums of the bf16 outputs (valid rows)
  long long* colsq;      // STATS == 2: and of their squares
  int stats_ld;
  int w_f16;             // the W operand is IEEE half (b_format = F16) instead of bf16
  int a_f16;             // the A operand is IEEE half (a_format = F16): activations of the 16-bit mode
  int a_plane_rows;      // > 0: the lo plane of A row r is row r + a_plane_rows of the same map (K-packed block 0)
  // ACT_FILTER (AS-norm cohort scores, score.cu): accumulators >= filt_thr[row] of columns < filt_ncols are appended
  // to the row's candidate list cand[row * filt_cap + slot], slot = atomicAdd(&filt_cnt[row], hits)
  const float* filt_thr; unsigned* filt_cnt; float* filt_cand; int filt_cap, filt_ncols;
};

__device__ __forceinline__ float tanh_exact(float x) {
  // tanh(x) = 1 - 2 / (exp(2x) + 1); ex2/rcp approximations are ~1e-7 relative, far inside the fp32-mode budget
  const float e = fast_ex2(x * 2.88539008178f);
  return 1.0f - 2.0f * fast_rcp(e + 1.0f);
}

// MC: launched as clusters of two CTAs that work on the two M tiles of an (M-tile pair, N tile) unit at the same time.
// They need the same W tile, so each loads HALF of it and TMA-multicasts the half into both CTAs: L2 -> SM traffic per
// CTA and k-block drops from 48 KB to 32 KB.  A stage may be refilled only when BOTH CTAs have consumed it (the peer's
// multicast lands in it too), so the `empty` barriers count two multicast tcgen05.commit arrivals.
// TWO (implies the cluster schedule of MC): the pair runs ONE tcgen05.mma.cta_group::2 of M = 256 per k-step.  Each CTA
// stages its own 128 x 64 A tile and only its HALF of the W tile (128 of the 256 N rows), so a stage is 32 KB instead of
// 48 KB (6 stages instead of 4, a third less L2 -> SM traffic, half the shared-memory operand reads per CTA).  The
// leader (cluster rank 0) owns the `full` barriers (both producers arrive.expect_tx on them, both CTAs' TMA loads
// complete_tx on them), issues every MMA and multicasts its commits to the `empty` / `tfull` barriers of both CTAs; the
// peer's epilogue warps release the accumulator on the leader's `tempty` barrier.
template <int BN, int ACT, bool RBT, bool HILO, int EW, int STATS, bool MC, bool TWO = false>
__global__ void __launch_bounds__(128 + EW * 32, 1)
gemm_pw_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
               const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmOut,
               const PwParams p) {
  using Cfg = PwCfg<BN, EW, TWO>;
  constexpr int STAGES = Cfg::STAGES;
  static_assert(!TWO || (MC && BN == 256), "the 2-SM MMA instances use the cluster schedule");
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[STAGES];
  __shared__ __align__(8) uint64_t empty_bar[STAGES];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t out_base = smem_base + STAGES * Cfg::STAGE_BYTES;

  const int num_m_tiles = (p.M + BLOCK_M - 1) / BLOCK_M;
  const int num_n_tiles = (p.N + BN - 1) / BN;
  const uint32_t crank = MC ? cluster_ctarank() : 0u;
  const int num_tiles = MC ? ((num_m_tiles + 1) >> 1) * num_n_tiles : num_m_tiles * num_n_tiles;   // schedule units
  const int sched_first = MC ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int sched_stride = MC ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  auto tile_m0 = [&](int t) { const int mt = t / num_n_tiles; return (MC ? 2 * mt + static_cast<int>(crank) : mt) * BLOCK_M; };
  const int kc = p.kc;
  const int kb_per_tap = p.num_pass * kc;
  const int num_kb = p.num_taps * kb_per_tap;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmA2);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmOut);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&full_bar[s]), TWO ? 2 : 1);             // TWO: one arrive.expect_tx per producer of the pair
      mbar_init(smem_u32(&empty_bar[s]), (MC && !TWO) ? 2 : 1);   // TWO: one multicast commit from the leader
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&tfull_bar[a]), 1);
      mbar_init(smem_u32(&tempty_bar[a]), TWO ? 2 * EW : EW);   // TWO: the epilogue warps of both CTAs
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    if (TWO) { tmem_alloc_2sm(smem_u32(&tmem_slot), Cfg::TMEM_COLS); tmem_relinquish_2sm(); }
    else { tmem_alloc(smem_u32(&tmem_slot), Cfg::TMEM_COLS); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if (MC) cluster_sync_all();      // the peer's barriers are initialised before anything is multicast into this CTA
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = sched_first; tile < num_tiles; tile += sched_stride) {
        const int m0 = tile_m0(tile);
        const int n0 = (tile % num_n_tiles) * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          const int tap = kb / kb_per_tap;
          const int rem = kb - tap * kb_per_tap;
          const int pass = rem / kc;
          const int kcx = rem - pass * kc;
          const int aplane = (pass == 2) ? 1 : 0;
          const int wplane = (pass == 1) ? 1 : 0;
          const int a_col = aplane * p.a_plane_cols + kcx * BLOCK_K;
          const int w_col = ((tap * p.w_planes + wplane) * kc + kcx) * BLOCK_K;
          const int a_row = m0 + p.tap_off[tap] + aplane * p.a_plane_rows;
          const uint32_t fb = smem_u32(&full_bar[stage]);
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
          const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
          if (TWO) {
            const uint32_t lfb = fb & 0xFEFFFFFFu;          // the leader's barrier (bit 24 = rank inside the CTA pair)
            mbar_arrive_expect_tx_cluster(lfb, Cfg::STAGE_BYTES);
            tma_load_2d_2sm(sa, (kcx < p.a2_kc) ? &tmA2 : &tmA, lfb, a_col, a_row);
            tma_load_2d_2sm(sa + A_STAGE_BYTES, &tmW, lfb, w_col, n0 + static_cast<int>(crank) * (BN / 2));
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            continue;
          }
          mbar_arrive_expect_tx(fb, Cfg::STAGE_BYTES);
          tma_load_2d(sa, (kcx < p.a2_kc) ? &tmA2 : &tmA, fb, a_col, a_row);
          if (MC)
            tma_load_2d_mc(sa + A_STAGE_BYTES + crank * (Cfg::B_STAGE_BYTES / 2), &tmW, fb, w_col,
                           n0 + static_cast<int>(crank) * (BN / 2), 0x3);
          else
            tma_load_2d(sa + A_STAGE_BYTES, &tmW, fb, w_col, n0);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1 && (!TWO || crank == 0)) {
    // ===================================================== MMA issuer (TWO: the leader CTA only, M = 256 over the pair)
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_f16(TWO ? 2 * BLOCK_M : BLOCK_M, BN, p.a_f16 != 0, p.w_f16 != 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = sched_first; tile < num_tiles; tile += sched_stride, ++it) {
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
          for (int k = 0; k < BLOCK_K / 16; ++k) {
            if (TWO) umma_bf16_2sm(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) ? 1u : 0u);
            else umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) ? 1u : 0u);
          }
          if (TWO) umma_commit_2sm(smem_u32(&empty_bar[stage]), 0x3);
          else if (MC) umma_commit_mc(smem_u32(&empty_bar[stage]), 0x3);
          else umma_commit(smem_u32(&empty_bar[stage]));
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        if (TWO) umma_commit_2sm(smem_u32(&tfull_bar[acc]), 0x3);
        else umma_commit(smem_u32(&tfull_bar[acc]));
      }
    }
  } else if (warp >= 4 && EW == 16 && ACT != ACT_FILTER) {
    // ===================================================== epilogue, 16 warps: quadrant = warp % 4; the two warps of a
    // PAIR (same quadrant, column groups 2P and 2P+1) cover 128 columns and fill, per unit, the left / right half of ONE
    // 32 x 64 box with 128-byte rows (128B swizzle), stored by one TMA instruction: 128-byte store requests instead of
    // the 64-byte rows of per-warp 32 x 32 boxes (the L2 path of these kernels is request bound: every extra request
    // stream tried, e.g. the L2 prefetch, cost 8-10 %).  TMEM -> regs -> bias/GELU/BN -> bf16 -> box -> TMA store.
    static_assert(EW != 16 || ACT == ACT_FILTER || (BN == 256 && (ACT == ACT_GELU || ACT == ACT_GELU_ERF) && !RBT),
                  "the 16-warp epilogue serves the GELU BN=256 instances");
    static_assert(!STATS || (EW == 16 && !HILO), "fused column statistics: 16-warp epilogue, single bf16 plane");
    const int ew = warp - 4;
    const int quad = ew & 3;
    const int cgroup = ew >> 2;
    const int pr = cgroup >> 1, pos = cgroup & 1;          // pair of column groups, position inside the pair
    const int pair = quad + 4 * pr;                        // 0..7: named barrier 1 + pair, box `pair`
    const bool issuer = pos == 0;
    const uint32_t my_box = out_base + pair * OUT_BOX_BYTES;
    const uint32_t my_row = my_box + lane * 128;
    const int sw = lane & 7;
    int it = 0;
    for (int tile = sched_first; tile < num_tiles; tile += sched_stride, ++it) {
      const int m0 = tile_m0(tile);
      const int n0 = (tile % num_n_tiles) * BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1u;
      int my_utt = -1;        // utterance of this lane's row if it is a valid (non-halo) frame
      unsigned stat_rows = 0;
      if (STATS) {
        const long long row = static_cast<long long>(m0) + quad * 32 + lane;
        if (row < p.M) {
          const RowMeta rm = p.row_meta[row];
          if (rm.t >= 0 && rm.t < rm.T) my_utt = rm.utt;
        }
        stat_rows = __ballot_sync(0xffffffffu, my_utt >= 0);
      }
      mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
      tc_fence_after();
#pragma unroll 1
      for (int u = 0; u < 2; ++u) {
        const int col0 = pr * 128 + u * 64 + pos * 32;
        const int n = n0 + col0;
        const int n_box = n0 + pr * 128 + u * 64;          // first column of the pair's box (N % 64 == 0)
        if (n_box >= p.N) break;                           // uniform over the pair
        uint32_t r[32];
        tmem_ld32(tmem_base + acc * BN + col0 + (static_cast<uint32_t>(quad * 32) << 16), r);
        tmem_ld_wait();
        uint32_t hi[16];
        uint32_t lo[16];   // dead (eliminated) unless HILO
        const float4* b4 = reinterpret_cast<const float4*>(p.bias + n);
        const float4* s4 = reinterpret_cast<const float4*>(p.scale + n);
        const float4* h4 = reinterpret_cast<const float4*>(p.shift + n);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 b = __ldg(b4 + i), s = __ldg(s4 + i), h = __ldg(h4 + i);
          float x0 = __uint_as_float(r[4 * i]) + b.x, x1 = __uint_as_float(r[4 * i + 1]) + b.y;
          float x2 = __uint_as_float(r[4 * i + 2]) + b.z, x3 = __uint_as_float(r[4 * i + 3]) + b.w;
          if (ACT == ACT_GELU_ERF) {
            x0 = gelu_erf_exact(x0); x1 = gelu_erf_exact(x1); x2 = gelu_erf_exact(x2); x3 = gelu_erf_exact(x3);
          } else {
            x0 = gelu_fast(x0); x1 = gelu_fast(x1); x2 = gelu_fast(x2); x3 = gelu_fast(x3);
          }
          x0 = fmaf(x0, s.x, h.x); x1 = fmaf(x1, s.y, h.y);
          x2 = fmaf(x2, s.z, h.z); x3 = fmaf(x3, s.w, h.w);
          const uint32_t p0 = pack16x2<!HILO>(x0, x1), p1 = pack16x2<!HILO>(x2, x3);
          hi[2 * i] = p0;
          hi[2 * i + 1] = p1;
          if (HILO) {
            lo[2 * i] = pack_bf16x2(x0 - bf16_lo_f(p0), x1 - bf16_hi_f(p0));
            lo[2 * i + 1] = pack_bf16x2(x2 - bf16_lo_f(p1), x3 - bf16_hi_f(p1));
          }
        }
        if (issuer && lane == 0) bulk_wait_read<0>();      // the previous store has finished reading the box
        named_bar_sync(1 + pair, 64);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + (((pos * 4 + j) ^ sw) << 4)), "r"(hi[4 * j]),
                       "r"(hi[4 * j + 1]), "r"(hi[4 * j + 2]), "r"(hi[4 * j + 3])
                       : "memory");
        fence_proxy_async();
        named_bar_sync(1 + pair, 64);
        if (issuer && lane == 0) {
          tma_store_2d(&tmOut, my_box, n_box, m0 + quad * 32);
          bulk_commit();
        }
        if (STATS) {
          // column sums of the staged bf16 box over the valid rows of each utterance it touches (usually one): lane =
          // (row parity h, column pair cp); 16 conflict-free LDS.32 per lane, one shuffle, int64 fixed-point atomics
          const int h = lane >> 4, cp = lane & 15;
          const uint32_t col_off = static_cast<uint32_t>((cp & 3) * 4);
          unsigned todo = stat_rows;
          while (todo) {
            const int src = __ffs(todo) - 1;
            const int utt = __shfl_sync(0xffffffffu, my_utt, src);
            const unsigned m = __ballot_sync(0xffffffffu, my_utt == utt);
            todo &= ~m;
            float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int r = 2 * i + h;
              if ((m >> r) & 1u) {
                uint32_t w;
                asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w)
                             : "r"(my_box + r * 128 + (((pos * 4 + (cp >> 2)) ^ (r & 7)) << 4) + col_off));
                const float a = f16_lo_f(w), b = f16_hi_f(w);      // STATS instances are single plane = IEEE half
                s0 += a; s1 += b;
                if (STATS == 2) { q0 = fmaf(a, a, q0); q1 = fmaf(b, b, q1); }
              }
            }
            s0 += __shfl_xor_sync(0xffffffffu, s0, 16);
            s1 += __shfl_xor_sync(0xffffffffu, s1, 16);
            if (STATS == 2) {
              q0 += __shfl_xor_sync(0xffffffffu, q0, 16);
              q1 += __shfl_xor_sync(0xffffffffu, q1, 16);
            }
            if (lane < 16) {
              const long long o = static_cast<long long>(utt) * p.stats_ld + n + 2 * cp;
              atomicAdd(reinterpret_cast<unsigned long long*>(p.colsum + o),
                        static_cast<unsigned long long>(__float2ll_rn(s0 * 4294967296.f)));
              atomicAdd(reinterpret_cast<unsigned long long*>(p.colsum + o + 1),
                        static_cast<unsigned long long>(__float2ll_rn(s1 * 4294967296.f)));
              if (STATS == 2) {
                atomicAdd(reinterpret_cast<unsigned long long*>(p.colsq + o),
                          static_cast<unsigned long long>(__float2ll_rn(q0 * 4294967296.f)));
                atomicAdd(reinterpret_cast<unsigned long long*>(p.colsq + o + 1),
                          static_cast<unsigned long long>(__float2ll_rn(q1 * 4294967296.f)));
              }
            }
          }
        }
        if (HILO) {
          if (issuer && lane == 0) bulk_wait_read<0>();
          named_bar_sync(1 + pair, 64);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + (((pos * 4 + j) ^ sw) << 4)), "r"(lo[4 * j]),
                         "r"(lo[4 * j + 1]), "r"(lo[4 * j + 2]), "r"(lo[4 * j + 3])
                         : "memory");
          fence_proxy_async();
          named_bar_sync(1 + pair, 64);
          if (issuer && lane == 0) {
            tma_store_2d(&tmOut, my_box, p.out_plane_cols + n_box, m0 + quad * 32);
            bulk_commit();
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (TWO) mbar_arrive_cluster(smem_u32(&tempty_bar[acc]) & 0xFEFFFFFFu);   // the leader issues the pair's MMAs
        else mbar_arrive(smem_u32(&tempty_bar[acc]));
      }
    }
    if (lane == 0) bulk_wait<0>();
  } else if (warp >= 4) {
    // ===================================================== epilogue: TMEM -> regs -> swizzled smem -> TMA store
    const int ew = warp - 4;
    const int quad = ew & 3;   // == warp % 4: the TMEM lane quadrant this warp may read
    const int half = ew >> 2;  // column half of the tile (ACT_FILTER with EW = 16: column quarter)
    constexpr int PARTS = EW / 4;      // warps per lane quadrant = column parts of a tile
    const uint32_t my_box = out_base + ew * OUT_BOX_BYTES;
    const uint32_t my_row = my_box + lane * 128;
    const int sw = lane & 7;
    int it = 0;
    for (int tile = sched_first; tile < num_tiles; tile += sched_stride, ++it) {
      const int m0 = tile_m0(tile);
      const int n0 = (tile % num_n_tiles) * BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1u;
      const float* rb = nullptr;
      if (RBT) {
        const long long row = m0 + quad * 32 + lane;
        const int utt = (row < p.M) ? p.row_meta[row].utt : -1;
        rb = p.rowbias + static_cast<long long>(utt < 0 ? 0 : utt) * p.rowbias_ld;
      }
      float f_thr = INFINITY;      // ACT_FILTER: this lane's row threshold (rows past M never pass)
      long long f_row = 0;
      if (ACT == ACT_FILTER) {
        f_row = static_cast<long long>(m0) + quad * 32 + lane;
        if (f_row < p.M) f_thr = __ldg(p.filt_thr + f_row);
      }
      mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
      tc_fence_after();
      if (ACT == ACT_FILTER) {
        // Fused top-K pre-filter: the score tile never leaves the SM.  Pass 1 reads this thread's BN / 2 accumulators
        // of its row and keeps only hit MASKS; one atomic per (row, tile half) reserves the slots; pass 2 re-reads the
        // strips from TMEM (on chip, cheap) and stores the ~2.5 % of the scores that pass.  A single atomic round trip
        // per tile hides under the next tile's MMAs; one per 32-column strip (first version) made the epilogue the
        // bottleneck (622 TFLOP/s against 890 for the plain store epilogue).
        constexpr int NSTRIP = BN / PARTS / 32;              // 32-column strips per thread
        uint32_t masks[NSTRIP];
        int total = 0;
#pragma unroll
        for (int st = 0; st < NSTRIP; ++st) {
          const int col0 = half * (BN / PARTS) + st * 32;
          uint32_t r[32];
          tmem_ld32(tmem_base + acc * BN + col0 + (static_cast<uint32_t>(quad * 32) << 16), r);
          tmem_ld_wait();
          uint32_t mask = 0;
#pragma unroll
          for (int i = 0; i < 32; ++i) mask |= (__uint_as_float(r[i]) >= f_thr ? 1u : 0u) << i;
          const int ncols = p.filt_ncols - (n0 + col0);      // valid (non-padding) columns of this strip
          if (ncols < 32) mask &= ncols <= 0 ? 0u : ((1u << ncols) - 1u);
          masks[st] = mask;
          total += __popc(mask);
        }
        if (__any_sync(0xffffffffu, total > 0)) {            // tcgen05.ld is warp collective: uniform branch
          uint32_t slot = 0;
          if (total > 0) slot = atomicAdd(p.filt_cnt + f_row, static_cast<unsigned>(total));
          float* dst = p.filt_cand + f_row * p.filt_cap;
#pragma unroll
          for (int st = 0; st < NSTRIP; ++st) {
            if (!__any_sync(0xffffffffu, masks[st] != 0u)) continue;
            const int col0 = half * (BN / PARTS) + st * 32;
            uint32_t r[32];
            tmem_ld32(tmem_base + acc * BN + col0 + (static_cast<uint32_t>(quad * 32) << 16), r);
            tmem_ld_wait();
            const uint32_t mask = masks[st];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              if ((mask >> i) & 1u) {
                if (slot < static_cast<uint32_t>(p.filt_cap)) dst[slot] = __uint_as_float(r[i]);
                ++slot;
              }
            }
          }
        }
      }
#pragma unroll 1
      for (int c = 0; c < BN / 128 && ACT != ACT_FILTER; ++c) {
        const int col0 = half * (BN / 2) + c * 64;
        const int n = n0 + col0;
        if (n >= p.N) break;
        if (ACT == ACT_RAW_F32) {
          // raw fp32 accumulators (cohort scores): two 32 x 32 fp32 boxes per 64 columns, same swizzled staging
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            uint32_t r[32];
            tmem_ld32(tmem_base + acc * BN + col0 + hh * 32 + (static_cast<uint32_t>(quad * 32) << 16), r);
            tmem_ld_wait();
            if (lane == 0) bulk_wait_read<0>();
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; ++j)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + ((j ^ sw) << 4)), "r"(r[4 * j]),
                           "r"(r[4 * j + 1]), "r"(r[4 * j + 2]), "r"(r[4 * j + 3])
                           : "memory");
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmOut, my_box, n + hh * 32, m0 + quad * 32);
              bulk_commit();
            }
          }
          continue;
        }
        uint32_t hi[32];
        uint32_t lo[32];   // dead (eliminated) unless HILO
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint32_t r[32];
          tmem_ld32(tmem_base + acc * BN + col0 + hh * 32 + (static_cast<uint32_t>(quad * 32) << 16), r);
          tmem_ld_wait();
          const float4* b4 = reinterpret_cast<const float4*>(p.bias + n + hh * 32);
          const float4* s4 = reinterpret_cast<const float4*>(p.scale + n + hh * 32);
          const float4* h4 = reinterpret_cast<const float4*>(p.shift + n + hh * 32);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 b = __ldg(b4 + i), s = __ldg(s4 + i), h = __ldg(h4 + i);
            float x0 = __uint_as_float(r[4 * i]) + b.x, x1 = __uint_as_float(r[4 * i + 1]) + b.y;
            float x2 = __uint_as_float(r[4 * i + 2]) + b.z, x3 = __uint_as_float(r[4 * i + 3]) + b.w;
            if (RBT) {
              const float4 q = __ldg(reinterpret_cast<const float4*>(rb + n + hh * 32) + i);
              x0 += q.x; x1 += q.y; x2 += q.z; x3 += q.w;
            }
            if (ACT == ACT_GELU) {
              x0 = gelu_fast(x0); x1 = gelu_fast(x1); x2 = gelu_fast(x2); x3 = gelu_fast(x3);
            } else if (ACT == ACT_RELU) {
              x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); x2 = fmaxf(x2, 0.f); x3 = fmaxf(x3, 0.f);
            }
            x0 = fmaf(x0, s.x, h.x); x1 = fmaf(x1, s.y, h.y);
            x2 = fmaf(x2, s.z, h.z); x3 = fmaf(x3, s.w, h.w);
            if (RBT) { x0 = tanh_exact(x0); x1 = tanh_exact(x1); x2 = tanh_exact(x2); x3 = tanh_exact(x3); }
            const uint32_t p0 = pack16x2<!HILO>(x0, x1), p1 = pack16x2<!HILO>(x2, x3);
            hi[hh * 16 + 2 * i] = p0;
            hi[hh * 16 + 2 * i + 1] = p1;
            if (HILO) {
              lo[hh * 16 + 2 * i] = pack_bf16x2(x0 - bf16_lo_f(p0), x1 - bf16_hi_f(p0));
              lo[hh * 16 + 2 * i + 1] = pack_bf16x2(x2 - bf16_lo_f(p1), x3 - bf16_hi_f(p1));
            }
          }
        }
        // stage the 32 x 64 block (row = lane, 128 bytes, 16-byte chunks XOR-swizzled by row & 7) and store it
        if (lane == 0) bulk_wait_read<0>();
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + ((j ^ sw) << 4)), "r"(hi[4 * j]),
                       "r"(hi[4 * j + 1]), "r"(hi[4 * j + 2]), "r"(hi[4 * j + 3])
                       : "memory");
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmOut, my_box, n, m0 + quad * 32);
          bulk_commit();
        }
        if (HILO) {
          if (lane == 0) bulk_wait_read<0>();
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + ((j ^ sw) << 4)),
                         "r"(lo[4 * j]), "r"(lo[4 * j + 1]), "r"(lo[4 * j + 2]), "r"(lo[4 * j + 3])
                         : "memory");
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmOut, my_box, p.out_plane_cols + n, m0 + quad * 32);
            bulk_commit();
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (TWO) mbar_arrive_cluster(smem_u32(&tempty_bar[acc]) & 0xFEFFFFFFu);   // the leader issues the pair's MMAs
        else mbar_arrive(smem_u32(&tempty_bar[acc]));
      }
    }
    if (lane == 0) bulk_wait<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (MC) cluster_sync_all();      // no CTA exits while its peer may still multicast data or arrivals into it
  if (warp == 2) {
    if (TWO) tmem_dealloc_2sm(tmem_base, Cfg::TMEM_COLS);
    else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------ host side
template <int BN, int ACT, bool RBT, bool HILO, int EW = 8, int STATS = 0, bool MC = false, bool TWO = false>
static int launch_pw_t(const CUtensorMap& a, const CUtensorMap& a2, const CUtensorMap& w, const CUtensorMap& o,
                       const PwParams& p, cudaStream_t stream) {
  using Cfg = PwCfg<BN, EW, TWO>;
  static unsigned long long attr_devs = 0;   // per device: kernel attributes do not carry over
  if (first_on_device(&attr_devs)) {
    SVB_CUDA_OK(cudaFuncSetAttribute(gemm_pw_kernel<BN, ACT, RBT, HILO, EW, STATS, MC, TWO>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     Cfg::SMEM_BYTES));
  }
  const int m_tiles = (p.M + BLOCK_M - 1) / BLOCK_M, n_tiles = (p.N + BN - 1) / BN;
  if (MC) {
    const int units = ((m_tiles + 1) / 2) * n_tiles;
    const int clusters = units < num_sms() / 2 ? units : num_sms() / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * clusters);
    cfg.blockDim = dim3(Cfg::NUM_THREADS);
    cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    SVB_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_pw_kernel<BN, ACT, RBT, HILO, EW, STATS, MC, TWO>, a, a2, w, o, p));
  } else {
    const int tiles = m_tiles * n_tiles;
    const int grid = tiles < num_sms() ? tiles : num_sms();
    gemm_pw_kernel<BN, ACT, RBT, HILO, EW, STATS, MC, TWO><<<grid, Cfg::NUM_THREADS, Cfg::SMEM_BYTES, stream>>>(a, a2, w, o, p);
  }
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// Pointwise TDNN layer.  `g` carries the operand description shared with launch_gemm (taps must be 1,
// store_mode STORE_ALL, bf16 output).  Supported: GELU layers with N % 256-tileable outputs (tdnn1/tdnn2/mfa)
// and the ReLU + row-bias + tanh attention layer (asp.tdnn).
int launch_gemm_pw(const GemmOperands& ops, const GemmParams& g, cudaStream_t stream) {
  SVB_CHECK(g.store_mode == STORE_ALL && g.num_taps >= 1 && g.num_taps <= 5 &&
                (g.num_taps > 1 || g.tap_off[0] == 0 || g.a_plane_rows > 0),
            "launch_gemm_pw stores every row (k > 1 layers need a halo fix-up by the caller)");
  SVB_CHECK(g.a_plane_rows == 0 || (g.num_taps == 1 && g.a_plane_cols == 0 && ops.a2 == nullptr),
            "row-stacked A planes come with the single-tap K-packed window view");
  SVB_CHECK(g.out_kind == OUT_F16 || g.out_kind == OUT_BF16_HILO,
            "launch_gemm_pw writes one plane of IEEE half or bf16 hi|lo planes");
  SVB_CHECK(g.N % 64 == 0 && g.out2 == nullptr, "launch_gemm_pw: N must be a multiple of 64 and there is no second output");
  SVB_CHECK(g.num_pass == 1 || g.num_pass == 3, "num_pass must be 1 or 3");
  const bool rbt = g.rowbias != nullptr;
  SVB_CHECK(rbt == (g.post_tanh != 0), "row-bias and tanh come together (asp.tdnn)");
  SVB_CHECK(!rbt || g.row_meta != nullptr, "row-bias needs row_meta");
  const bool hilo = g.out_kind == OUT_BF16_HILO;
  const int bn = (g.N >= 256 && !rbt) ? 256 : 128;
  CUtensorMap ma, ma2, mw, mo;
  // (a_cols may exceed lda: overlapping rows, the K-packed window view of a dense activation matrix)
  if (make_map(&ma, ops.a, static_cast<long long>(g.M) + g.a_plane_rows, ops.a_cols, ops.lda, BLOCK_M)) return -1;
  if (ops.a2 != nullptr) {
    if (make_map(&ma2, ops.a2, g.M, ops.a2_cols, ops.lda2, BLOCK_M)) return -1;
  } else {
    ma2 = ma;
  }
  // cluster pairs with W-tile multicast for the GELU BN=256 layers (SVB_PW_MC=0 selects the single-CTA instances)
  static const bool mc_enabled = [] { const char* e = getenv("SVB_PW_MC"); return e == nullptr || e[0] != '0'; }();
  const bool mc = mc_enabled && bn == 256 && !rbt;
  if (make_map(&mw, ops.w, ops.w_rows, ops.w_cols, ops.w_cols, mc ? bn / 2 : bn)) return -1;
  const long long out_cols = hilo ? static_cast<long long>(g.out_plane_cols) + g.N : g.N;
  if (make_map(&mo, g.out, g.M, out_cols, g.out_ld, 32)) return -1;   // 32 x 64 boxes, 128B swizzle (8 and 16 warps)
  PwParams p{};
  p.M = g.M; p.N = g.N; p.kc = g.kc_per_tap; p.num_pass = g.num_pass; p.a_plane_cols = g.a_plane_cols;
  p.a2_kc = g.a2_kc; p.w_planes = g.w_planes; p.bias = g.bias; p.scale = g.scale; p.shift = g.shift;
  p.rowbias = g.rowbias; p.rowbias_ld = g.rowbias_ld; p.row_meta = g.row_meta; p.out_plane_cols = g.out_plane_cols;
  p.num_taps = g.num_taps;
  for (int j = 0; j < g.num_taps; ++j) p.tap_off[j] = g.tap_off[j];
  p.colsum = g.colsum; p.colsq = g.colsq; p.stats_ld = g.stats_ld;
  p.w_f16 = ops.w_f16; p.a_f16 = ops.a_f16; p.a_plane_rows = g.a_plane_rows;
  SVB_CHECK(ops.a_f16 == ops.w_f16, "activations and weights of a layer share one 16-bit format: a kind::f16 UMMA "
            "with a_format != b_format is an illegal instruction on sm_100a");
  SVB_CHECK(!ops.a_f16 || g.num_pass == 1, "fp16 activations are single plane / single pass");
  SVB_CHECK(!ops.w_f16 || (g.num_pass == 1 && g.w_planes == 1), "fp16 weights are single plane / single pass");
  const int stats = g.colsum == nullptr ? 0 : (g.colsq == nullptr ? 1 : 2);
  SVB_CHECK(!stats || (bn == 256 && !hilo && !rbt && g.row_meta != nullptr && g.stats_ld % 2 == 0),
            "fused column statistics need the BN=256 single-plane GELU instance and row_meta");
  if (rbt) {
    SVB_CHECK(g.act == ACT_RELU, "the row-bias instance is ReLU + BN + tanh");
    return hilo ? launch_pw_t<128, ACT_RELU, true, true>(ma, ma2, mw, mo, p, stream)
                : launch_pw_t<128, ACT_RELU, true, false>(ma, ma2, mw, mo, p, stream);
  }
  SVB_CHECK(g.act == ACT_GELU, "the plain pointwise instance is GELU + BN");
  SVB_CHECK(g.num_taps == 1 || (bn == 256 && !stats && g.a2_kc == 0), "k > 1 runs on the BN=256 instances without statistics");
  if (g.num_taps > 1) {   // block 0: exact-erf GELU, bit-identical to the gemm_tc path it replaces
    static const bool two_b0 = [] { const char* e = getenv("SVB_PW_2SM"); return e == nullptr || e[0] != '0'; }();
    if (mc && two_b0 && !hilo)
      return launch_pw_t<256, ACT_GELU_ERF, false, false, 16, 0, true, true>(ma, ma2, mw, mo, p, stream);
    if (mc)
      return hilo ? launch_pw_t<256, ACT_GELU_ERF, false, true, 16, 0, true>(ma, ma2, mw, mo, p, stream)
                  : launch_pw_t<256, ACT_GELU_ERF, false, false, 16, 0, true>(ma, ma2, mw, mo, p, stream);
    return hilo ? launch_pw_t<256, ACT_GELU_ERF, false, true, 16>(ma, ma2, mw, mo, p, stream)
                : launch_pw_t<256, ACT_GELU_ERF, false, false, 16>(ma, ma2, mw, mo, p, stream);
  }
  if (bn == 256)
  {
    static const bool two_enabled = [] { const char* e = getenv("SVB_PW_2SM"); return e == nullptr || e[0] != '0'; }();
    if (mc && two_enabled && !hilo)
      return stats == 2 ? launch_pw_t<256, ACT_GELU, false, false, 16, 2, true, true>(ma, ma2, mw, mo, p, stream)
           : stats == 1 ? launch_pw_t<256, ACT_GELU, false, false, 16, 1, true, true>(ma, ma2, mw, mo, p, stream)
                        : launch_pw_t<256, ACT_GELU, false, false, 16, 0, true, true>(ma, ma2, mw, mo, p, stream);
    if (mc)
      return hilo         ? launch_pw_t<256, ACT_GELU, false, true, 16, 0, true>(ma, ma2, mw, mo, p, stream)
             : stats == 2 ? launch_pw_t<256, ACT_GELU, false, false, 16, 2, true>(ma, ma2, mw, mo, p, stream)
             : stats == 1 ? launch_pw_t<256, ACT_GELU, false, false, 16, 1, true>(ma, ma2, mw, mo, p, stream)
                          : launch_pw_t<256, ACT_GELU, false, false, 16, 0, true>(ma, ma2, mw, mo, p, stream);
    return hilo         ? launch_pw_t<256, ACT_GELU, false, true, 16>(ma, ma2, mw, mo, p, stream)
           : stats == 2 ? launch_pw_t<256, ACT_GELU, false, false, 16, 2>(ma, ma2, mw, mo, p, stream)
           : stats == 1 ? launch_pw_t<256, ACT_GELU, false, false, 16, 1>(ma, ma2, mw, mo, p, stream)
                        : launch_pw_t<256, ACT_GELU, false, false, 16>(ma, ma2, mw, mo, p, stream);
  }
  return hilo ? launch_pw_t<128, ACT_GELU, false, true>(ma, ma2, mw, mo, p, stream)
              : launch_pw_t<128, ACT_GELU, false, false>(ma, ma2, mw, mo, p, stream);
}

// Plain fp32-output GEMM  out[M, N] = A * W^T  (no bias / activation / BN): the cohort-score matrix of AS-norm.
// out_ld in floats; N a multiple of 64.
int launch_gemm_f32(const GemmOperands& ops, const GemmParams& g, cudaStream_t stream) {
  SVB_CHECK(g.num_taps == 1 && g.tap_off[0] == 0 && g.N % 64 == 0, "launch_gemm_f32: pointwise, N % 64 == 0");
  SVB_CHECK((reinterpret_cast<uintptr_t>(g.out) & 15) == 0 && (g.out_ld * 4) % 16 == 0, "fp32 output must be 16-byte aligned");
  // 256-column tiles unless they would leave SMs idle (the GRU's recurrent GEMM: 592 x 3072 is 60 tiles of 256
  // columns on 148 SMs, 120 of 128 columns: 23.9 -> 20.0 us per launch)
  const long long tiles256 = static_cast<long long>((g.M + BLOCK_M - 1) / BLOCK_M) * ((g.N + 255) / 256);
  const int bn = (g.N >= 256 && tiles256 >= num_sms()) ? 256 : 128;
  CUtensorMap ma, mw, mo;
  if (make_map(&ma, ops.a, g.M, ops.a_cols, ops.lda, BLOCK_M)) return -1;
  if (make_map(&mw, ops.w, ops.w_rows, ops.w_cols, ops.w_cols, bn)) return -1;
  if (make_map_f32(&mo, g.out, g.M, g.N, g.out_ld, 32)) return -1;
  PwParams p{};
  p.M = g.M; p.N = g.N; p.kc = g.kc_per_tap; p.num_pass = g.num_pass; p.a_plane_cols = g.a_plane_cols;
  p.a2_kc = 0; p.w_planes = g.w_planes; p.num_taps = 1; p.tap_off[0] = 0;
  return bn == 256 ? launch_pw_t<256, ACT_RAW_F32, false, false>(ma, ma, mw, mo, p, stream)
                   : launch_pw_t<128, ACT_RAW_F32, false, false>(ma, ma, mw, mo, p, stream);
}

// Cohort-score GEMM with the top-K pre-filter fused into the epilogue (score.cu): scores[M, N] = A * W^T are compared
// with thr[row] as they leave TMEM; survivors of columns < n_valid go to cand[row * cap + ...], cnt[row] counts them
// (it may exceed cap: the caller treats that row as unproven).  No score is stored anywhere else.
int launch_gemm_filter(const GemmOperands& ops, const GemmParams& g, const float* thr, unsigned* cnt, float* cand, int cap,
                       int n_valid, cudaStream_t stream) {
  SVB_CHECK(g.num_taps == 1 && g.tap_off[0] == 0 && g.N % 64 == 0, "launch_gemm_filter: pointwise, N % 64 == 0");
  SVB_CHECK(thr != nullptr && cnt != nullptr && cand != nullptr && cap > 0, "launch_gemm_filter: null filter buffers");
  SVB_CHECK(ops.a_f16 == ops.w_f16, "activations and weights share one 16-bit format");
  const int bn = g.N >= 256 ? 256 : 128;
  CUtensorMap ma, mw;
  if (make_map(&ma, ops.a, g.M, ops.a_cols, ops.lda, BLOCK_M)) return -1;
  if (make_map(&mw, ops.w, ops.w_rows, ops.w_cols, ops.w_cols, bn)) return -1;
  PwParams p{};
  p.M = g.M; p.N = g.N; p.kc = g.kc_per_tap; p.num_pass = g.num_pass; p.a_plane_cols = g.a_plane_cols;
  p.a2_kc = 0; p.w_planes = g.w_planes; p.num_taps = 1; p.tap_off[0] = 0;
  p.w_f16 = ops.w_f16; p.a_f16 = ops.a_f16;
  p.filt_thr = thr; p.filt_cnt = cnt; p.filt_cand = cand; p.filt_cap = cap; p.filt_ncols = n_valid;
  // 16 epilogue warps: each thread owns 64 columns of its row (two strips), half the per-warp epilogue time of EW = 8.
  // Default: the cta_group::2 instance (M = 256 over a CTA pair, each CTA stages half of the W tile): 158-160 against
  // 153-156 M trials/s for the 1-SM instance in one call (SVB_FILTER_2SM=0 selects the latter).
  static const bool two = [] { const char* e = getenv("SVB_FILTER_2SM"); return e == nullptr || e[0] != '0'; }();
  if (bn == 256 && two) {
    CUtensorMap mw2;
    if (make_map(&mw2, ops.w, ops.w_rows, ops.w_cols, ops.w_cols, bn / 2)) return -1;
    return launch_pw_t<256, ACT_FILTER, false, false, 16, 0, true, true>(ma, ma, mw2, ma, p, stream);
  }
  return bn == 256 ? launch_pw_t<256, ACT_FILTER, false, false, 16>(ma, ma, mw, ma, p, stream)
                   : launch_pw_t<128, ACT_FILTER, false, false, 16>(ma, ma, mw, ma, p, stream);
}

}  // namespace svb
