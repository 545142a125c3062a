// Persistent warp-specialised implicit-GEMM kernel for the TDNN / Conv1d layers of ECAPA-TDNN
// (reference: TDNNBlock = norm(activation(conv(x))), src/models/ECAPA_TDNN.py:27-69, with the
// reflect "same" padding of src/models/layers/cnn.py:104-157).
//
//   out[r, n] = BN( act( sum_tap sum_c A[r + off_tap, c] * W[n, tap, c] + bias[n] (+ rowbias) ) )
//
// * A (time-major activations, bf16) and W (packed weights, bf16) are staged by TMA into
//   128B-swizzled shared memory; tcgen05.mma (cta_group::1, M=128, N=BN, K=16) accumulates fp32
//   in TMEM; 8 epilogue warps read TMEM with tcgen05.ld and apply bias/activation/BN in fp32.
// * Dilated taps are row-shifted TMA loads of the same matrix (im2col-free).  Reflect padding is
//   realised on the STORE side: every utterance owns `halo` mirrored rows before and after its
//   valid rows, written by the producing layer (STORE_MIRROR) or computed in place by pointwise
//   layers (STORE_ALL), so loads never need boundary handling.
// * fp32-grade mode: operands are split hi+lo bf16 and the K loop runs hi*hi + hi*lo + lo*hi
//   (num_pass 3); num_pass 2 splits only the activations (hi*hi + lo*hi: bf16 weights, fp32-grade input).
#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

static constexpr int BLOCK_M = 128;
static constexpr int BLOCK_K = 64;
static constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;
static constexpr int NUM_THREADS = 384;  // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps4-11 epilogue

template <int BN>
struct GemmCfg {
  static constexpr int B_STAGE_BYTES = BN * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int STAGES = (BN == 256) ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024;
  static constexpr int TMEM_COLS = 2 * BN;  // double-buffered accumulator (power of two >= 32)
};

__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

struct RowTargets {
  long long r[3];
  int n;
};

template <int KIND>
__device__ __forceinline__ void store32(void* base, int ld, int plane_cols, long long row, int col,
                                        const float (&v)[32]) {
  if (KIND == OUT_F32) {
    float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(base) + row * ld + col);
#pragma unroll
    for (int i = 0; i < 8; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  } else {
    bf16* p = reinterpret_cast<bf16*>(base) + row * ld + col;
    uint32_t hi[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) hi[i] = pack16x2<KIND == OUT_F16>(v[2 * i], v[2 * i + 1]);
    uint4* dst = reinterpret_cast<uint4*>(p);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = make_uint4(hi[4 * i], hi[4 * i + 1], hi[4 * i + 2], hi[4 * i + 3]);
    if (KIND == OUT_BF16_HILO) {
      uint32_t lo[16];
#pragma unroll
      for (int i = 0; i < 16; ++i)
        lo[i] = pack_bf16x2(v[2 * i] - bf16_lo_f(hi[i]), v[2 * i + 1] - bf16_hi_f(hi[i]));
      uint4* dl = reinterpret_cast<uint4*>(p + plane_cols);
#pragma unroll
      for (int i = 0; i < 4; ++i) dl[i] = make_uint4(lo[4 * i], lo[4 * i + 1], lo[4 * i + 2], lo[4 * i + 3]);
    }
  }
}

template <int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
               const __grid_constant__ CUtensorMap tmW, const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[STAGES];
  __shared__ __align__(8) uint64_t empty_bar[STAGES];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;

  const int num_m_tiles = (p.M + BLOCK_M - 1) / BLOCK_M;
  const int num_n_tiles = (p.N + BN - 1) / BN;
  const int num_tiles = num_m_tiles * num_n_tiles;
  const int kc = p.kc_per_tap;
  const int kb_per_tap = p.num_pass * kc;
  const int num_kb = p.num_taps * kb_per_tap;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmA2);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&tfull_bar[a]), 1);
      mbar_init(smem_u32(&tempty_bar[a]), 8);
    }
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
    // ===================================================== TMA producer (one elected lane)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / num_n_tiles) * BLOCK_M;
        const int n0 = (tile % num_n_tiles) * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          const int tap = kb / kb_per_tap;
          const int rem = kb - tap * kb_per_tap;
          const int pass_i = rem / kc;
          const int kcx = rem - pass_i * kc;
          const int pass = (p.num_pass == 2 && pass_i == 1) ? 2 : pass_i;   // 2 passes: hi*hi + lo(A)*hi(W)
          const int aplane = (pass == 2) ? 1 : 0;
          const int wplane = (pass == 1) ? 1 : 0;
          const int a_col = (p.a_plane_rows ? 0 : aplane * p.a_plane_cols) + kcx * BLOCK_K;
          const int a_row = m0 + p.tap_off[tap] + aplane * p.a_plane_rows;
          const int w_col = ((tap * p.w_planes + wplane) * kc + kcx) * BLOCK_K;
          const uint32_t fb = smem_u32(&full_bar[stage]);
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
          mbar_arrive_expect_tx(fb, Cfg::STAGE_BYTES);
          const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
          tma_load_2d(sa, (kcx < p.a2_kc) ? &tmA2 : &tmA, fb, a_col, a_row);
          tma_load_2d(sa + A_STAGE_BYTES, &tmW, fb, w_col, n0);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer (one elected lane)
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_f16(BLOCK_M, BN, p.a_f16 != 0, p.w_f16 != 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
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
            // +32 bytes (16 bf16) along K inside the 128B swizzle row: start-address field += 2
            umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) ? 1u : 0u);
          }
          umma_commit(smem_u32(&empty_bar[stage]));
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit(smem_u32(&tfull_bar[acc]));
      }
    }
  } else if (warp >= 4) {
    // ===================================================== epilogue: TMEM -> regs -> global
    const int ew = warp - 4;
    const int quad = ew & 3;   // == warp % 4: the TMEM lane quadrant this warp may read
    const int half = ew >> 2;  // column half of the tile
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int m0 = (tile / num_n_tiles) * BLOCK_M;
      const int n0 = (tile % num_n_tiles) * BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1u;
      const long long row = m0 + quad * 32 + lane;

      RowTargets tg;
      tg.n = 0;
      int utt = -1;
      if (row < p.M) {
        if (p.row_meta != nullptr) {
          const int4 mt = *reinterpret_cast<const int4*>(p.row_meta + row);
          utt = mt.x;
          if (p.store_mode == STORE_MIRROR) {
            const int t = mt.y, T = mt.z;
            if (utt >= 0 && t >= 0 && t < T) {
              tg.r[tg.n++] = row;
              if (t >= 1 && t <= p.halo) tg.r[tg.n++] = row - 2 * t;
              const int e = T - 1 - t;
              if (e >= 1 && e <= p.halo) tg.r[tg.n++] = row + 2 * e;
            }
          } else {
            tg.r[tg.n++] = row;
          }
        } else {
          tg.r[tg.n++] = row;
        }
      }

      mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < BN / 64; ++c) {
        const int col0 = half * (BN / 2) + c * 32;
        uint32_t r[32];
        tmem_ld32(tmem_base + acc * BN + col0 + (static_cast<uint32_t>(quad * 32) << 16), r);
        tmem_ld_wait();
        const int n = n0 + col0;
        if (n < p.N && tg.n > 0) {
          float v[32];
          const float4* b4 = reinterpret_cast<const float4*>(p.bias + n);
          const float4* s4 = reinterpret_cast<const float4*>(p.scale + n);
          const float4* h4 = reinterpret_cast<const float4*>(p.shift + n);
          const float4* rb4 = (p.rowbias != nullptr && utt >= 0)
                                  ? reinterpret_cast<const float4*>(p.rowbias + (long long)utt * p.rowbias_ld + n)
                                  : nullptr;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 b = __ldg(b4 + i), s = __ldg(s4 + i), h = __ldg(h4 + i);
            float x0 = __uint_as_float(r[4 * i]) + b.x, x1 = __uint_as_float(r[4 * i + 1]) + b.y;
            float x2 = __uint_as_float(r[4 * i + 2]) + b.z, x3 = __uint_as_float(r[4 * i + 3]) + b.w;
            if (rb4 != nullptr) {
              const float4 rb = __ldg(rb4 + i);
              x0 += rb.x; x1 += rb.y; x2 += rb.z; x3 += rb.w;
            }
            if (p.act == ACT_GELU) {
              x0 = gelu_erf(x0); x1 = gelu_erf(x1); x2 = gelu_erf(x2); x3 = gelu_erf(x3);
            } else if (p.act == ACT_RELU) {
              x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); x2 = fmaxf(x2, 0.f); x3 = fmaxf(x3, 0.f);
            }
            x0 = fmaf(x0, s.x, h.x); x1 = fmaf(x1, s.y, h.y);
            x2 = fmaf(x2, s.z, h.z); x3 = fmaf(x3, s.w, h.w);
            if (p.post_tanh) { x0 = tanhf(x0); x1 = tanhf(x1); x2 = tanhf(x2); x3 = tanhf(x3); }
            v[4 * i] = x0; v[4 * i + 1] = x1; v[4 * i + 2] = x2; v[4 * i + 3] = x3;
          }
          for (int j = 0; j < tg.n; ++j) {
            if (p.out_kind == OUT_BF16) store32<OUT_BF16>(p.out, p.out_ld, 0, tg.r[j], n, v);
            else if (p.out_kind == OUT_F16) store32<OUT_F16>(p.out, p.out_ld, 0, tg.r[j], n, v);
            else if (p.out_kind == OUT_BF16_HILO) store32<OUT_BF16_HILO>(p.out, p.out_ld, p.out_plane_cols, tg.r[j], n, v);
            else store32<OUT_F32>(p.out, p.out_ld, 0, tg.r[j], n, v);
          }
          if (p.out2 != nullptr) {
            // second output: this layer's result plus the next Res2Net chunk (x_{i+1} + y_i,
            // src/models/ECAPA_TDNN.py:126), stored with the same mirrored-halo rule.
            const bf16* ap = p.add_src + row * p.add_ld + n;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint4 u = __ldg(reinterpret_cast<const uint4*>(ap) + i);
              const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                v[8 * i + 2 * q] += p.a_f16 ? f16_lo_f(w[q]) : bf16_lo_f(w[q]);      // add_src shares A's format
                v[8 * i + 2 * q + 1] += p.a_f16 ? f16_hi_f(w[q]) : bf16_hi_f(w[q]);
              }
            }
            if (p.add_planes == 2) {
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
            for (int j = 0; j < tg.n; ++j) {
              if (p.out_kind == OUT_BF16) store32<OUT_BF16>(p.out2, p.out2_ld, 0, tg.r[j], n, v);
              else if (p.out_kind == OUT_F16) store32<OUT_F16>(p.out2, p.out2_ld, 0, tg.r[j], n, v);
              else store32<OUT_BF16_HILO>(p.out2, p.out2_ld, p.out2_plane_cols, tg.r[j], n, v);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  return fn;
}

// 2D bf16 row-major matrix [rows, ld] exposing `cols` columns; box = [64 cols, box_rows], 128B swizzle.
int make_map(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld, int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  SVB_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  SVB_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld * 2) % 16 == 0,
            "TMA operand must be 16-byte aligned (base %p, ld %lld)", base, ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {BLOCK_K, static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SVB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with code %d", static_cast<int>(r));
  return 0;
}

// Store view of the sinc layer's output (rawnet_gemm.cu): activations [utt][1 + t][ld] with t = 16 * ml + jg seen as a
// 4-D tensor (channel, jg, ml, utt) so that one box = 32 consecutive ml of one (jg, utt) x 32 channels (64 B rows, 64B
// swizzle).  `base` = row 1 of utterance 0; ml_extent = T / 16 (every jg is a valid frame below it).
int make_map_sinc_out(CUtensorMap* m, const void* base, int channels, long long ld, int ml_extent, long long rows_per_utt,
                      int utts) {
  EncodeTiledFn fn = get_encode_fn();
  SVB_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  SVB_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld * 2) % 16 == 0 && ml_extent >= 1,
            "TMA store view must be 16-byte aligned (base %p, ld %lld)", base, ld);
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(channels), 16, static_cast<cuuint64_t>(ml_extent), static_cast<cuuint64_t>(utts)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(ld) * 2, static_cast<cuuint64_t>(ld) * 32,
                           static_cast<cuuint64_t>(rows_per_utt) * ld * 2};
  cuuint32_t box[4] = {32, 1, 32, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SVB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (sinc store view) failed with code %d", static_cast<int>(r));
  return 0;
}

// Same matrix with box = [32 cols (64 B), box_rows] and 64B swizzle: the 2 KB store boxes of gemm_pw's 16-warp epilogue.
int make_map_sw64(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld, int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  SVB_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  SVB_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld * 2) % 16 == 0,
            "TMA operand must be 16-byte aligned (base %p, ld %lld)", base, ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {32, static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SVB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (64B swizzle) failed with code %d", static_cast<int>(r));
  return 0;
}

// 2D fp32 row-major matrix [rows, ld] exposing `cols` columns; box = [32 cols (128 B), box_rows], 128B swizzle.
int make_map_f32(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld, int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  SVB_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 4};
  cuuint32_t box[2] = {32, static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SVB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (fp32) failed with code %d", static_cast<int>(r));
  return 0;
}

template <int BN>
static int launch_t(const CUtensorMap& a, const CUtensorMap& a2, const CUtensorMap& w, const GemmParams& p,
                    cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static unsigned long long attr_devs = 0;   // per device: kernel attributes do not carry over
  if (first_on_device(&attr_devs)) {
    SVB_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     Cfg::SMEM_BYTES));
  }
  const int tiles = ((p.M + BLOCK_M - 1) / BLOCK_M) * ((p.N + BN - 1) / BN);
  const int grid = tiles < num_sms() ? tiles : num_sms();
  gemm_tc_kernel<BN><<<grid, NUM_THREADS, Cfg::SMEM_BYTES, stream>>>(a, a2, w, p);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_gemm(const GemmOperands& ops, const GemmParams& p_in, int block_n, cudaStream_t stream) {
  GemmParams p = p_in;
  p.w_f16 = ops.w_f16; p.a_f16 = ops.a_f16;
  SVB_CHECK(ops.a_f16 == ops.w_f16, "activations and weights of a layer share one 16-bit format: a kind::f16 UMMA "
            "with a_format != b_format is an illegal instruction on sm_100a");
  SVB_CHECK(!p.a_f16 || (p.num_pass == 1 && p.add_planes <= 1), "fp16 activations are single plane / single pass");
  SVB_CHECK(!p.w_f16 || (p.num_pass == 1 && p.w_planes == 1), "fp16 weights are single plane / single pass");
  SVB_CHECK(p.M > 0 && p.N > 0 && p.kc_per_tap > 0 && p.num_taps >= 1 && p.num_taps <= 5, "bad GEMM shape");
  SVB_CHECK(p.num_pass >= 1 && p.num_pass <= 3, "num_pass must be 1, 2 or 3");
  SVB_CHECK(p.N % 32 == 0, "N must be a multiple of 32 (got %d)", p.N);
  CUtensorMap ma, ma2, mw;
  // (a_cols may exceed lda: overlapping rows, the K-packed window view of a dense activation matrix)
  if (make_map(&ma, ops.a, static_cast<long long>(p.M) + p.a_plane_rows, ops.a_cols, ops.lda, BLOCK_M)) return -1;
  if (ops.a2 != nullptr) {
    if (make_map(&ma2, ops.a2, p.M, ops.a2_cols, ops.lda2, BLOCK_M)) return -1;
  } else {
    ma2 = ma;
  }
  if (make_map(&mw, ops.w, ops.w_rows, ops.w_cols, ops.w_cols, block_n)) return -1;
  switch (block_n) {
    case 256: return launch_t<256>(ma, ma2, mw, p, stream);
    case 128: return launch_t<128>(ma, ma2, mw, p, stream);
    case 64: return launch_t<64>(ma, ma2, mw, p, stream);
    default: set_error("unsupported block_n %d", block_n); return -1;
  }
}

}  // namespace svb
