// Attentive statistics pooling tail (reference: AttentiveStatisticsPooling.forward,
// src/models/ECAPA_TDNN.py:247-260): logits = asp.conv(att) [128 -> 3C], softmax over the frames of
// each utterance, attention-weighted mean and std of the MFA output x -- in ONE kernel that never
// writes the [frames, 3C] logits (972 MB fp32 per 256-utterance launch in the unfused version).
//
// The GEMM is issued TRANSPOSED: D[channel, frame] = Wc[channel, :] . att[frame, :], so channels
// sit on the 128 TMEM lanes and frames on the columns.  Every epilogue thread then owns ONE channel
// and reduces over time serially in registers (online softmax: running max, sum e, sum e*x,
// sum e*x^2) -- no cross-lane reductions at all.  The x values arrive in the same (lane = channel,
// column = frame) layout through a second MMA with an identity matrix as the A operand
// (D2 = I . Y_tile^T: a tensor-core transpose, exact for bf16 inputs with fp32 accumulation), so
// the epilogue reads logits and x with the same tcgen05.ld pattern and touches no shared memory.
//
// Work item = (utterance, block of 128 channels); frames are streamed in chunks of CH_N columns
// through a 2-stage TMA ring and a double-buffered TMEM slot pair; the two epilogue warpgroups take
// alternate chunks and merge their partial (max, sums) once per item.  The conv bias is constant over
// time and cancels in the softmax.  Variance is accumulated around the utterance's global mean g
// (already needed for the ASP context): var = sum a (x-g)^2 - (sum a (x-g))^2, which equals the
// reference's sum a (x - mean)^2 without the second pass and without cancellation trouble.
#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

struct UttRows { long long row0; int T; int pad; };

static constexpr int ASP_THREADS = 384;   // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps4-11 epilogue
static constexpr int ASP_KC = 2;          // attention channels = 128 = 2 K chunks of 64
static constexpr float LOG2E = 1.4426950408889634f;

template <bool SPLIT>
struct AspCfg {
  static constexpr int P = SPLIT ? 2 : 1;
  static constexpr int CH_N = SPLIT ? 32 : 128;                 // frames per chunk (split mode: smem holds 2 planes of everything)
  static constexpr int BOX_BYTES = CH_N * 128;                  // one [CH_N rows x 64 cols] bf16 box
  static constexpr int STAGE_BYTES = P * 4 * BOX_BYTES;         // att (2 K chunks) + Y (2 K chunks), per plane
  static constexpr int STAGES = 2;
  static constexpr int W_BYTES = P * 2 * 16384;                 // Wc block [128 x 128] per plane
  static constexpr int I_BYTES = 2 * 16384;                     // identity [128 x 128]
  static constexpr int SMEM_BYTES = W_BYTES + I_BYTES + STAGES * STAGE_BYTES + 1024;
  static constexpr int TMEM_COLS = 4 * CH_N;                    // 2 slots x (logits + x)
};

struct AspParams {
  int B, C3, n_items;
  const UttRows* utts;
  const float* gmean; int gmean_ld;
  float* pooled; int pooled_ld;
  float eps;
  int y_plane_cols;      // column offset of the lo plane of Y
  const float* bound;    // FIXMAX: per-channel upper bound of logit * log2(e) (att = tanh(.) lies in [-1, 1])
};

// FIXMAX: the softmax is shift invariant and the logits are bounded by the L1 norm of their weight row (|att| <= 1),
// so that bound replaces the running maximum: no chunk-maximum pass over the logits (one TMEM read and one FMNMX per
// element less), no rescaling of the partial sums, and both warpgroups accumulate on the same scale.  The host picks
// this instance only when every bound is small enough for exp2(logit - bound) to stay far from fp32 underflow.
template <bool SPLIT, bool FIXMAX>
__global__ void __launch_bounds__(ASP_THREADS, 1)
asp_fused_kernel(const __grid_constant__ CUtensorMap tmAtt, const __grid_constant__ CUtensorMap tmY,
                 const __grid_constant__ CUtensorMap tmW, const AspParams p) {
  using Cfg = AspCfg<SPLIT>;
  constexpr int P = Cfg::P, CH_N = Cfg::CH_N, STAGES = Cfg::STAGES, BOX = Cfg::BOX_BYTES;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[STAGES];
  __shared__ __align__(8) uint64_t empty_bar[STAGES];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ __align__(8) uint64_t wfull_bar, wempty_bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float4 merge_buf[2][128];
  __shared__ __align__(8) uint64_t mfull_bar[2], mempty_bar[2];   // hand-over of warpgroup 1's partial statistics

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t w_smem = smem_base;
  const uint32_t i_smem = w_smem + Cfg::W_BYTES;
  const uint32_t st_smem = i_smem + Cfg::I_BYTES;

  // contiguous item range of this CTA; item = cb * B + b (channel block major, so Wc rarely changes)
  const int per = (p.n_items + gridDim.x - 1) / gridDim.x;
  const int item_begin = blockIdx.x * per;
  const int item_end = min(p.n_items, item_begin + per);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmAtt);
    tma_prefetch_desc(&tmY);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&tfull_bar[a]), 1);
      mbar_init(smem_u32(&tempty_bar[a]), 4);
    }
    mbar_init(smem_u32(&wfull_bar), 1);
    mbar_init(smem_u32(&wempty_bar), 1);
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&mfull_bar[a]), 4);
      mbar_init(smem_u32(&mempty_bar[a]), 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(smem_u32(&tmem_slot), Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  // identity matrix in the K-major 128B-swizzled operand layout: element (r, c) lives in K chunk c / 64 at
  // byte r * 128 + (((c % 64) / 8) ^ (r & 7)) * 16 + (c % 8) * 2
  for (int i = threadIdx.x; i < Cfg::I_BYTES / 16; i += ASP_THREADS)
    asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(i_smem + i * 16), "r"(0) : "memory");
  __syncthreads();
  if (threadIdx.x < 128) {
    const int r = threadIdx.x, c = r;
    const uint32_t addr = i_smem + (c >> 6) * 16384 + r * 128 + ((((c & 63) >> 3) ^ (r & 7)) << 4) + (c & 7) * 2;
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(static_cast<unsigned short>(SPLIT ? 0x3f80 : 0x3c00)) : "memory");   // 1.0 as bf16 / as IEEE half
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer
    if (elect_one()) {
      int stage = 0, cur_cb = -1;
      uint32_t phase = 0, wchg = 0;
      for (int item = item_begin; item < item_end; ++item) {
        const int cb = item / p.B, b = item - cb * p.B;
        const UttRows u = p.utts[b];
        if (cb != cur_cb) {
          // new channel block: (re)load Wc once every MMA reading the old block has completed
          mbar_wait(smem_u32(&wempty_bar), (wchg & 1u) ^ 1u);
          const uint32_t wb = smem_u32(&wfull_bar);
          mbar_arrive_expect_tx(wb, Cfg::W_BYTES);
          for (int pl = 0; pl < P; ++pl)
            for (int kcx = 0; kcx < ASP_KC; ++kcx)
              tma_load_2d(w_smem + (pl * 2 + kcx) * 16384, &tmW, wb, pl * 128 + kcx * 64, cb * 128);
          cur_cb = cb;
          ++wchg;
        }
        const int nch = (u.T + CH_N - 1) / CH_N;
        for (int ch = 0; ch < nch; ++ch) {
          const int row = static_cast<int>(u.row0) + ch * CH_N;
          const uint32_t fb = smem_u32(&full_bar[stage]);
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
          mbar_arrive_expect_tx(fb, Cfg::STAGE_BYTES);
          const uint32_t sa = st_smem + stage * Cfg::STAGE_BYTES;
          for (int pl = 0; pl < P; ++pl)
            for (int kcx = 0; kcx < ASP_KC; ++kcx) {
              tma_load_2d(sa + (pl * 2 + kcx) * BOX, &tmAtt, fb, pl * 128 + kcx * 64, row);
              tma_load_2d(sa + (P * 2 + pl * 2 + kcx) * BOX, &tmY, fb, pl * p.y_plane_cols + cb * 128 + kcx * 64, row);
            }
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (elect_one()) {
      // split mode: every operand is a bf16 plane; 16-bit mode: every operand (asp.conv weights, the identity, att, Y)
      // is IEEE half.  A kind::f16 UMMA cannot mix the two: a_format != b_format raises an illegal-instruction error
      // on sm_100a (measured, profiles/r02_precision.md), although the descriptor has two fields.
      constexpr uint32_t idesc = umma_idesc_f16(128, CH_N, !SPLIT, !SPLIT);
      int stage = 0, cur_cb = -1;
      uint32_t phase = 0, wchg = 0, g = 0;
      for (int item = item_begin; item < item_end; ++item) {
        const int cb = item / p.B, b = item - cb * p.B;
        const UttRows u = p.utts[b];
        if (cb != cur_cb) {
          mbar_wait(smem_u32(&wfull_bar), wchg & 1u);
          tc_fence_after();
          cur_cb = cb;
          ++wchg;
        }
        const int nch = (u.T + CH_N - 1) / CH_N;
        for (int ch = 0; ch < nch; ++ch, ++g) {
          const uint32_t slot = g & 1u;
          mbar_wait(smem_u32(&tempty_bar[slot]), ((g >> 1) & 1u) ^ 1u);
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          const uint32_t sa = st_smem + stage * Cfg::STAGE_BYTES;
          const uint32_t d_log = tmem_base + slot * 2 * CH_N;
          const uint32_t d_x = d_log + CH_N;
          // logits^T = Wc . att^T  (split mode: hi*hi + lo*hi + hi*lo)
          uint32_t accf = 0;
#pragma unroll
          for (int pass = 0; pass < (SPLIT ? 3 : 1); ++pass) {
            const int wpl = (pass == 1) ? 1 : 0, apl = (pass == 2) ? 1 : 0;
#pragma unroll
            for (int kcx = 0; kcx < ASP_KC; ++kcx) {
              const uint64_t adesc = umma_desc_sw128(w_smem + (wpl * 2 + kcx) * 16384);
              const uint64_t bdesc = umma_desc_sw128(sa + (apl * 2 + kcx) * BOX);
#pragma unroll
              for (int k = 0; k < 4; ++k) { umma_bf16(d_log, adesc + 2 * k, bdesc + 2 * k, idesc, accf); accf = 1; }
            }
          }
          // x^T = I . Y_tile^T  (tensor-core transpose; hi + lo planes accumulate)
          accf = 0;
#pragma unroll
          for (int pl = 0; pl < P; ++pl) {
#pragma unroll
            for (int kcx = 0; kcx < ASP_KC; ++kcx) {
              const uint64_t adesc = umma_desc_sw128(i_smem + kcx * 16384);
              const uint64_t bdesc = umma_desc_sw128(sa + (P * 2 + pl * 2 + kcx) * BOX);
#pragma unroll
              for (int k = 0; k < 4; ++k) { umma_bf16(d_x, adesc + 2 * k, bdesc + 2 * k, idesc, accf); accf = 1; }
            }
          }
          umma_commit(smem_u32(&empty_bar[stage]));
          umma_commit(smem_u32(&tfull_bar[slot]));
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        const int next_cb = (item + 1 < item_end) ? (item + 1) / p.B : -1;
        if (next_cb != cb) umma_commit(smem_u32(&wempty_bar));
      }
    }
  } else if (warp >= 4) {
    // ===================================================== epilogue: per-channel online softmax statistics
    const int ew = warp - 4;
    const int wg = ew >> 2;      // warpgroup 0 / 1 takes even / odd chunks (TMEM slot == wg)
    const int quad = ew & 3;     // == warp % 4: TMEM lane quadrant
    const int ch_local = quad * 32 + lane;
    const uint32_t lane_sel = static_cast<uint32_t>(quad * 32) << 16;
    uint32_t g = 0, nproc = 0;
    for (int item = item_begin; item < item_end; ++item) {
      const int cb = item / p.B, b = item - cb * p.B;
      const UttRows u = p.utts[b];
      const int c = cb * 128 + ch_local;
      const float gm = __ldg(p.gmean + static_cast<long long>(b) * p.gmean_ld + c);
      float m = FIXMAX ? __ldg(p.bound + c) : -INFINITY;   // m is kept in log2 units (logit * log2 e)
      float l = 0.f, s1 = 0.f, s2 = 0.f;
      const int nch = (u.T + CH_N - 1) / CH_N;
      for (int ch = 0; ch < nch; ++ch, ++g) {
        if ((g & 1u) != static_cast<uint32_t>(wg)) continue;
        const int nvalid = min(CH_N, u.T - ch * CH_N);
        mbar_wait(smem_u32(&tfull_bar[wg]), nproc & 1u);
        tc_fence_after();
        const uint32_t t_log = tmem_base + wg * 2 * CH_N + lane_sel;
        const uint32_t t_x = t_log + CH_N;
        const int ngrp = (nvalid + 31) >> 5;
        if (!FIXMAX) {
          // pass 1: chunk maximum
          float cmax = -INFINITY;
          for (int grp = 0; grp < ngrp; ++grp) {
            uint32_t r[32];
            tmem_ld32(t_log + grp * 32, r);
            tmem_ld_wait();
            const int lim = nvalid - grp * 32;
            if (lim >= 32) {
#pragma unroll
              for (int i = 0; i < 32; ++i) cmax = fmaxf(cmax, __uint_as_float(r[i]));
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) cmax = fmaxf(cmax, i < lim ? __uint_as_float(r[i]) : -INFINITY);
            }
          }
          const float mz = fmaxf(m, cmax * LOG2E);
          const float corr = fast_ex2(m - mz);
          l *= corr; s1 *= corr; s2 *= corr;
          m = mz;
        }
        // pass 2: e = exp(logit - max); accumulate sum e, sum e (x - g), sum e (x - g)^2
        for (int grp = 0; grp < ngrp; ++grp) {
          uint32_t r[32], xr[32];
          tmem_ld32(t_log + grp * 32, r);
          tmem_ld32(t_x + grp * 32, xr);
          tmem_ld_wait();
          const int lim = nvalid - grp * 32;
          if (lim >= 32) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float e = fast_ex2(fmaf(__uint_as_float(r[i]), LOG2E, -m));
              const float xs = __uint_as_float(xr[i]) - gm;
              const float pe = e * xs;
              l += e; s1 += pe; s2 = fmaf(pe, xs, s2);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float e = i < lim ? fast_ex2(fmaf(__uint_as_float(r[i]), LOG2E, -m)) : 0.f;
              const float xs = __uint_as_float(xr[i]) - gm;
              const float pe = e * xs;
              l += e; s1 += pe; s2 = fmaf(pe, xs, s2);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[wg]));
        ++nproc;
      }
      // merge the two warpgroups' partial statistics; warpgroup 0 writes the pooled mean / std.  The hand-over is a
      // two-slot mailbox (mbarriers), not a barrier over both warpgroups: the chunk counter alternates which of them
      // gets the item's extra chunk (3 chunks per 3 s crop), so over two items they do the same work -- but only if the
      // one that finishes an item first may start the next (a 256-thread barrier here made it wait: 10 % of the
      // kernel's stall samples, ncu r02).
      const int li = item - item_begin;
      const int mslot = li & 1;
      const uint32_t mphase = static_cast<uint32_t>(li >> 1) & 1u;
      float4* mb = merge_buf[mslot];
      if (wg == 1) {
        mbar_wait(smem_u32(&mempty_bar[mslot]), mphase ^ 1u);       // the slot's previous content has been read
        mb[ch_local] = make_float4(m, l, s1, s2);
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&mfull_bar[mslot]));
      } else {
        mbar_wait(smem_u32(&mfull_bar[mslot]), mphase);
      }
      if (wg == 0) {
        const float4 o = mb[ch_local];
        const float M = fmaxf(m, o.x);
        const float f0 = l > 0.f ? fast_ex2(m - M) : 0.f;
        const float f1 = o.y > 0.f ? fast_ex2(o.x - M) : 0.f;
        const float L = l * f0 + o.y * f1;
        const float S1 = s1 * f0 + o.z * f1;
        const float S2 = s2 * f0 + o.w * f1;
        const float d = S1 / L;
        const float var = S2 / L - d * d;
        float* po = p.pooled + static_cast<long long>(b) * p.pooled_ld;
        po[c] = gm + d;
        po[p.C3 + c] = sqrtf(fmaxf(var, p.eps));
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&mempty_bar[mslot]));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ------------------------------------------------------------------ host side
template <bool SPLIT, bool FIXMAX>
static int launch_asp(const CUtensorMap& ma, const CUtensorMap& my, const CUtensorMap& mw, const AspParams& p,
                      cudaStream_t stream) {
  using Cfg = AspCfg<SPLIT>;
  static unsigned long long attr_devs = 0;   // per device: kernel attributes do not carry over
  if (first_on_device(&attr_devs)) {
    SVB_CUDA_OK(cudaFuncSetAttribute(asp_fused_kernel<SPLIT, FIXMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
  }
  const int grid = p.n_items < num_sms() ? p.n_items : num_sms();
  asp_fused_kernel<SPLIT, FIXMAX><<<grid, ASP_THREADS, Cfg::SMEM_BYTES, stream>>>(ma, my, mw, p);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// att [rows, planes*128] bf16 (tanh output of asp.tdnn), y [rows, y_ld] bf16 MFA output (hi | lo at +y_plane_cols),
// bound [C3] or nullptr (see FIXMAX);
// wc packed asp.conv weight [C3, planes*128]; gmean [B, gmean_ld] global time means; pooled [B, pooled_ld] = [mean | std].
int asp_fused(const bf16* att, const bf16* y, int y_ld, int y_plane_cols, const bf16* wc, int planes, long long rows,
              const void* utts, int B, int C3, const float* gmean, int gmean_ld, float eps, float* pooled, int pooled_ld,
              const float* bound, cudaStream_t stream) {
  SVB_CHECK(C3 % 128 == 0, "ASP channels must be a multiple of 128 (got %d)", C3);
  SVB_CHECK(planes == 1 || planes == 2, "planes must be 1 or 2");
  const int chn = planes == 2 ? 32 : 128;
  CUtensorMap ma, my, mw;
  if (make_map(&ma, att, rows, planes * 128, planes * 128, chn)) return -1;
  if (make_map(&my, y, rows, planes == 2 ? y_plane_cols + C3 : C3, y_ld, chn)) return -1;
  if (make_map(&mw, wc, C3, planes * 128, planes * 128, 128)) return -1;
  AspParams p{};
  p.B = B; p.C3 = C3; p.n_items = (C3 / 128) * B;
  p.utts = static_cast<const UttRows*>(utts);
  p.gmean = gmean; p.gmean_ld = gmean_ld; p.pooled = pooled; p.pooled_ld = pooled_ld; p.eps = eps;
  p.y_plane_cols = y_plane_cols;
  p.bound = bound;
  if (bound != nullptr)
    return planes == 2 ? launch_asp<true, true>(ma, my, mw, p, stream) : launch_asp<false, true>(ma, my, mw, p, stream);
  return planes == 2 ? launch_asp<true, false>(ma, my, mw, p, stream) : launch_asp<false, false>(ma, my, mw, p, stream);
}

}  // namespace svb
