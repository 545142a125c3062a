// Per-utterance reductions and small dense layers of ECAPA-TDNN that are not GEMM-shaped:
//   SEBlock            src/models/ECAPA_TDNN.py:164-177  (time mean -> 2 tiny convs -> gate * x [+ residual :336])
//   AttentiveStatisticsPooling  src/models/ECAPA_TDNN.py:213-260 (global mean/std, softmax over time,
//                               attention-weighted mean/std)
//   asp_bn + fc head   src/models/ECAPA_TDNN.py:496-500
// All statistics are fp32; activations are read from bf16 hi|lo planes (planes == 2, fp32-grade mode) or from ONE plane
// of IEEE half (planes == 1, 16-bit mode).
#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

struct UttRows { long long row0; int T; int pad; };  // first VALID row and frame count of an utterance

__device__ __forceinline__ void load8(const bf16* p, int planes, int plane_cols, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
  if (planes == 1) {
#pragma unroll
    for (int q = 0; q < 4; ++q) { v[2 * q] = f16_lo_f(w[q]); v[2 * q + 1] = f16_hi_f(w[q]); }
    return;
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) { v[2 * q] = bf16_lo_f(w[q]); v[2 * q + 1] = bf16_hi_f(w[q]); }
  if (planes == 2) {
    const uint4 l = *reinterpret_cast<const uint4*>(p + plane_cols);
    const uint32_t z[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) { v[2 * q] += bf16_lo_f(z[q]); v[2 * q + 1] += bf16_hi_f(z[q]); }
  }
}

__device__ __forceinline__ float load1(const bf16* p, int planes, int plane_cols) {
  if (planes == 1) return __half2float(*reinterpret_cast<const __half*>(p));
  return __bfloat162float(p[0]) + __bfloat162float(p[plane_cols]);
}

// grid (C/256, B), block 256: warp w sweeps rows w, w+8, ...; lane owns 8 consecutive channels.
// mode 0: mean only (SE squeeze).  mode 1: mean and std = sqrt(clamp(mean((x-mean)^2), eps)) (ASP global context),
// in ONE pass over x: sums are taken around the first frame s = x[0] (sum (x-s), sum (x-s)^2), so
// var = E[(x-s)^2] - (E[x-s])^2 has no large-offset cancellation, and x (486 MB for the MFA output of a
// 256-utterance launch, far beyond L2) is read from HBM once instead of twice.
__global__ void __launch_bounds__(256)
col_stats_kernel(const bf16* __restrict__ x, int ld, int planes, int plane_cols, const UttRows* __restrict__ utts,
                 int C, int mode, float eps, float* __restrict__ mean_out, float* __restrict__ std_out, int out_ld) {
  __shared__ float red[8][256];
  __shared__ float red2[8][256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c0 = blockIdx.x * 256 + lane * 8;
  const UttRows u = utts[blockIdx.y];
  const bool active = c0 < C;
  float acc[8], acc2[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { acc[i] = 0.f; acc2[i] = 0.f; sh[i] = 0.f; }
  if (active) {
    if (mode == 1) load8(x + u.row0 * ld + c0, planes, plane_cols, sh);
    for (int t = warp; t < u.T; t += 8) {
      float v[8];
      load8(x + (u.row0 + t) * ld + c0, planes, plane_cols, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) { const float d = v[i] - sh[i]; acc[i] += d; acc2[i] = fmaf(d, d, acc2[i]); }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) { red[warp][lane * 8 + i] = acc[i]; red2[warp][lane * 8 + i] = acc2[i]; }
  __syncthreads();
  float s = 0.f, q = 0.f;
  for (int w = 0; w < 8; ++w) { s += red[w][threadIdx.x]; q += red2[w][threadIdx.x]; }
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  const float inv = 1.f / static_cast<float>(u.T);
  const float d = s * inv;
  float shift = 0.f;
  if (mode == 1) {
    shift = load1(x + u.row0 * ld + c, planes, plane_cols);
  }
  mean_out[static_cast<long long>(blockIdx.y) * out_ld + c] = shift + d;
  if (mode == 1) std_out[static_cast<long long>(blockIdx.y) * out_ld + c] = sqrtf(fmaxf(q * inv - d * d, eps));
}

// out[b, j] = act( sum_c W[j, c] * (in[b, c] * in_scale[c] + in_shift[c]) + bias[j] ), fp32 SIMT.
// The per-utterance dense layers (SE squeeze/excite, ASP context bias, asp_bn + fc head) are skinny:
// B utterances x J outputs x K up to 6C.  64 x 64 output tiles, 4 x 4 per thread, K split across
// gridDim.z so that ~148 CTAs are busy; partial sums are reduced in a FIXED order by a second tiny
// kernel (no atomics: results are bit-reproducible run to run).
enum { LIN_NONE = 0, LIN_RELU = 1, LIN_SIGMOID = 2 };
static constexpr int SL_TILE = 64, SL_KSTEP = 32, SL_PAD = 68;

__device__ __forceinline__ float lin_act(float v, int act) {
  if (act == LIN_RELU) return fmaxf(v, 0.f);
  if (act == LIN_SIGMOID) return 1.f / (1.f + expf(-v));
  return v;
}

__global__ void __launch_bounds__(256)
skinny_linear_kernel(const float* __restrict__ W, int w_ld, const float* __restrict__ bias, const float* __restrict__ in,
                     int in_ld, const float* __restrict__ in_scale, const float* __restrict__ in_shift, int B, int J,
                     int K, int slice, int act, float* __restrict__ out, int out_ld, float* __restrict__ part) {
  __shared__ __align__(16) float As[SL_KSTEP][SL_PAD];
  __shared__ __align__(16) float Ws[SL_KSTEP][SL_PAD];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int j0 = blockIdx.x * SL_TILE, b0 = blockIdx.y * SL_TILE;
  const int k_begin = blockIdx.z * slice, k_end = min(K, k_begin + slice);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int kk = k_begin; kk < k_end; kk += SL_KSTEP) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * 256;
      const int row = idx >> 3, kq = (idx & 7) * 4, k = kk + kq;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), w = a;
      if (k < k_end) {
        if (b0 + row < B) {
          a = *reinterpret_cast<const float4*>(in + static_cast<long long>(b0 + row) * in_ld + k);
          if (in_scale != nullptr) {
            const float4 sc = __ldg(reinterpret_cast<const float4*>(in_scale + k));
            const float4 sh = __ldg(reinterpret_cast<const float4*>(in_shift + k));
            a.x = fmaf(a.x, sc.x, sh.x); a.y = fmaf(a.y, sc.y, sh.y); a.z = fmaf(a.z, sc.z, sh.z); a.w = fmaf(a.w, sc.w, sh.w);
          }
        }
        if (j0 + row < J) w = __ldg(reinterpret_cast<const float4*>(W + static_cast<long long>(j0 + row) * w_ld + k));
      }
      As[kq][row] = a.x; As[kq + 1][row] = a.y; As[kq + 2][row] = a.z; As[kq + 3][row] = a.w;
      Ws[kq][row] = w.x; Ws[kq + 1][row] = w.y; Ws[kq + 2][row] = w.z; Ws[kq + 3][row] = w.w;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < SL_KSTEP; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 w = *reinterpret_cast<const float4*>(&Ws[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int b = b0 + ty * 4 + i;
    if (b >= B) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int jj = j0 + tx * 4 + j;
      if (jj >= J) continue;
      if (part != nullptr) part[(static_cast<long long>(blockIdx.z) * B + b) * J + jj] = acc[i][j];
      else out[static_cast<long long>(b) * out_ld + jj] = lin_act(acc[i][j] + (bias != nullptr ? bias[jj] : 0.f), act);
    }
  }
}

__global__ void skinny_reduce_kernel(const float* __restrict__ part, int S, int B, int J, const float* __restrict__ bias,
                                     int act, float* __restrict__ out, int out_ld) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * J) return;
  const int b = i / J, j = i - b * J;
  float v = 0.f;
  for (int s = 0; s < S; ++s) v += part[static_cast<long long>(s) * B * J + i];
  out[static_cast<long long>(b) * out_ld + j] = lin_act(v + (bias != nullptr ? bias[j] : 0.f), act);
}

// out[row, c] = gate[utt, c] * x[row, c] + resid[row, c] over every row of an utterance (halo rows too).
__global__ void __launch_bounds__(256)
se_apply_kernel(const bf16* __restrict__ x, int x_ld, int x_pc, const bf16* __restrict__ resid, int r_ld, int r_pc,
                const float* __restrict__ gate, int gate_ld, const RowMeta* __restrict__ meta, long long rows, int C,
                int planes, bf16* __restrict__ out, int o_ld, int o_pc) {
  const int per_row = C / 8;
  const long long total = rows * per_row;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long row = i / per_row;
    const int c = static_cast<int>(i - row * per_row) * 8;
    const int utt = meta[row].utt;
    if (utt < 0) continue;
    // all global loads of the iteration are issued before the first decode (the kernel is a pure HBM stream: with the
    // loads of V and of the residual serialised behind each other's decode it ran at 56 % instead of 90 % of the copy peak)
    const uint4 xv = __ldg(reinterpret_cast<const uint4*>(x + row * x_ld + c));
    const uint4 rv = __ldg(reinterpret_cast<const uint4*>(resid + row * r_ld + c));
    uint4 xl = make_uint4(0, 0, 0, 0), rl = make_uint4(0, 0, 0, 0);
    if (planes == 2) {
      xl = __ldg(reinterpret_cast<const uint4*>(x + row * x_ld + x_pc + c));
      rl = __ldg(reinterpret_cast<const uint4*>(resid + row * r_ld + r_pc + c));
    }
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(gate + static_cast<long long>(utt) * gate_ld + c));
    const float4 g1 = __ldg(reinterpret_cast<const float4*>(gate + static_cast<long long>(utt) * gate_ld + c + 4));
    const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w}, rw[4] = {rv.x, rv.y, rv.z, rv.w};
    const uint32_t xlw[4] = {xl.x, xl.y, xl.z, xl.w}, rlw[4] = {rl.x, rl.y, rl.z, rl.w};
    float v[8], r[8];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (planes == 1) {
        v[2 * q] = f16_lo_f(xw[q]); v[2 * q + 1] = f16_hi_f(xw[q]);
        r[2 * q] = f16_lo_f(rw[q]); r[2 * q + 1] = f16_hi_f(rw[q]);
      } else {
        v[2 * q] = bf16_lo_f(xw[q]) + bf16_lo_f(xlw[q]); v[2 * q + 1] = bf16_hi_f(xw[q]) + bf16_hi_f(xlw[q]);
        r[2 * q] = bf16_lo_f(rw[q]) + bf16_lo_f(rlw[q]); r[2 * q + 1] = bf16_hi_f(rw[q]) + bf16_hi_f(rlw[q]);
      }
    }
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float a = fmaf(g[2 * q], v[2 * q], r[2 * q]);
      const float b = fmaf(g[2 * q + 1], v[2 * q + 1], r[2 * q + 1]);
      if (planes == 1) { hi[q] = pack_f16x2(a, b); continue; }
      hi[q] = pack_bf16x2(a, b);
      lo[q] = pack_bf16x2(a - bf16_lo_f(hi[q]), b - bf16_hi_f(hi[q]));
    }
    *reinterpret_cast<uint4*>(out + row * o_ld + c) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    if (planes == 2) *reinterpret_cast<uint4*>(out + row * o_ld + o_pc + c) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
}

// Attention-weighted statistics: softmax over the valid frames of each (utterance, channel),
// mean = sum a*x, std = sqrt(clamp(sum a*(x-mean)^2, eps)).  grid (C/64, B), block 256 = 64 channels x 4 time slices.
__global__ void __launch_bounds__(256)
asp_pool_kernel(const float* __restrict__ logits, int l_ld, const bf16* __restrict__ x, int x_ld, int planes,
                int plane_cols, const UttRows* __restrict__ utts, int C, float eps, float* __restrict__ pooled,
                int pooled_ld) {
  __shared__ float sm[4][64], sl[4][64], ss[4][64];
  __shared__ float gmax[64], gsum[64], gmean[64];
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int c = blockIdx.x * 64 + tx;
  const UttRows u = utts[blockIdx.y];
  float m = -INFINITY, l = 0.f, s1 = 0.f;
  for (int t = ty; t < u.T; t += 4) {
    const long long row = u.row0 + t;
    const float lg = logits[row * l_ld + c];
    const float xv = load1(x + row * x_ld + c, planes, plane_cols);
    const float mn = fmaxf(m, lg);
    const float corr = expf(m - mn), e = expf(lg - mn);
    l = l * corr + e;
    s1 = s1 * corr + e * xv;
    m = mn;
  }
  sm[ty][tx] = m; sl[ty][tx] = l; ss[ty][tx] = s1;
  __syncthreads();
  if (ty == 0) {
    float M = sm[0][tx];
    for (int i = 1; i < 4; ++i) M = fmaxf(M, sm[i][tx]);
    float L = 0.f, S = 0.f;
    for (int i = 0; i < 4; ++i) {
      const float f = (sl[i][tx] > 0.f) ? expf(sm[i][tx] - M) : 0.f;
      L += sl[i][tx] * f;
      S += ss[i][tx] * f;
    }
    gmax[tx] = M; gsum[tx] = L; gmean[tx] = S / L;
  }
  __syncthreads();
  const float M = gmax[tx], mu = gmean[tx];
  float s2 = 0.f;
  for (int t = ty; t < u.T; t += 4) {
    const long long row = u.row0 + t;
    const float lg = logits[row * l_ld + c];
    const float xv = load1(x + row * x_ld + c, planes, plane_cols);
    const float d = xv - mu;
    s2 = fmaf(expf(lg - M), d * d, s2);
  }
  ss[ty][tx] = s2;
  __syncthreads();
  if (ty == 0) {
    const float S2 = (ss[0][tx] + ss[1][tx]) + (ss[2][tx] + ss[3][tx]);
    float* o = pooled + static_cast<long long>(blockIdx.y) * pooled_ld;
    o[c] = mu;
    o[C + c] = sqrtf(fmaxf(S2 / gsum[tx], eps));
  }
}

// rows of [B, D] -> unit L2 norm (F.normalize(p=2, dim=1), src/model.py:422-423); one warp per row.
__global__ void l2norm_rows_kernel(const float* __restrict__ in, int in_ld, float* __restrict__ out, int out_ld, int B,
                                   int D, float eps) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= B) return;
  float s = 0.f;
  for (int c = lane; c < D; c += 32) { const float v = in[static_cast<long long>(row) * in_ld + c]; s = fmaf(v, v, s); }
  s = warp_sum(s);
  const float inv = 1.f / fmaxf(sqrtf(s), eps);
  for (int c = lane; c < D; c += 32) out[static_cast<long long>(row) * out_ld + c] = in[static_cast<long long>(row) * in_ld + c] * inv;
}

// ------------------------------------------------------------------ launchers
// Finalises the fixed-point column sums a GELU GEMM epilogue accumulated (gemm_pw.cu, STATS instances):
// mean = sum / T, std = sqrt(clamp(sumsq / T - mean^2, eps)) in fp64 (the sums are exact to 2^-32 per 32-row box, so
// the one-pass variance has no cancellation problem).  Same outputs as col_stats_kernel.
__global__ void colsum_finalize_kernel(const long long* __restrict__ sum, const long long* __restrict__ sq, int stats_ld,
                                       const UttRows* __restrict__ utts, int B, int C, float eps,
                                       float* __restrict__ mean_out, float* __restrict__ std_out, int out_ld) {
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= static_cast<long long>(B) * C) return;
  const int b = static_cast<int>(i / C), c = static_cast<int>(i - static_cast<long long>(b) * C);
  const double inv = 1.0 / (4294967296.0 * static_cast<double>(utts[b].T));
  const double mean = static_cast<double>(sum[static_cast<long long>(b) * stats_ld + c]) * inv;
  mean_out[static_cast<long long>(b) * out_ld + c] = static_cast<float>(mean);
  if (sq != nullptr) {
    const double ex2 = static_cast<double>(sq[static_cast<long long>(b) * stats_ld + c]) * inv;
    std_out[static_cast<long long>(b) * out_ld + c] = sqrtf(fmaxf(static_cast<float>(ex2 - mean * mean), eps));
  }
}

int colsum_finalize(const long long* sum, const long long* sq, int stats_ld, const void* utts, int B, int C, float eps,
                    float* mean_out, float* std_out, int out_ld, cudaStream_t stream) {
  const long long n = static_cast<long long>(B) * C;
  colsum_finalize_kernel<<<static_cast<int>((n + 255) / 256), 256, 0, stream>>>(
      sum, sq, stats_ld, static_cast<const UttRows*>(utts), B, C, eps, mean_out, std_out, out_ld);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------- SE gate in one launch
// gate[u, :] = sigmoid(W2 . relu(W1 . mean[u, :] + b1) + b2)  (SEBlock, src/models/ECAPA_TDNN.py:181-203), mean = the
// per-utterance time mean of the block output, taken either from the fixed-point column sums of the GEMM epilogue
// (colsum_finalize's conversion) or from a float buffer.  Replaces colsum_finalize + two split-K linears + a reduce
// (4 launches, ~70 us per block at 592 utterances: each of them latency bound) by one kernel of SE_UPC utterances per
// CTA.  Weights come TRANSPOSED (w1t [C][128], w2t [128][C]) so that a warp reads whole rows.  Hidden width 128.
// Phase 1: warp w owns C / 8 input channels, lane 4 hidden units; fixed-order reduction over the 8 warps (run-to-run
// bit reproducible).  Phase 2: a thread owns 4 adjacent output channels per 1024.
static constexpr int SE_UPC = 4, SE_H = 128;

__global__ void __launch_bounds__(512)
se_gate_kernel(const long long* __restrict__ sum, int sum_ld, const float* __restrict__ mean_in, int mean_ld,
               const UttRows* __restrict__ utts, int B, int C, const float* __restrict__ w1t,
               const float* __restrict__ b1, const float* __restrict__ w2t, const float* __restrict__ b2,
               float* __restrict__ gate, int gate_ld) {
  extern __shared__ __align__(16) float se_smem[];
  float* mean = se_smem;                                  // [SE_UPC][C]
  float* part = mean + SE_UPC * C;                        // [16][SE_UPC][SE_H]; phase 2: [256][SE_UPC * 4]
  float* hid = part + 16 * SE_UPC * SE_H;                 // [SE_UPC][SE_H]
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int u0 = blockIdx.x * SE_UPC;
  for (int i = tid; i < SE_UPC * C; i += 512) {
    const int u = i / C, c = i - u * C;
    float m = 0.f;
    if (u0 + u < B) {
      if (sum != nullptr) {
        const double inv = 1.0 / (4294967296.0 * static_cast<double>(utts[u0 + u].T));
        m = static_cast<float>(static_cast<double>(sum[static_cast<long long>(u0 + u) * sum_ld + c]) * inv);
      } else {
        m = mean_in[static_cast<long long>(u0 + u) * mean_ld + c];
      }
    }
    mean[i] = m;
  }
  __syncthreads();
  {
    // phase 1: warp w owns input channels [w C / 16, (w + 1) C / 16), lane owns hidden units 4 lane .. 4 lane + 3
    float acc[SE_UPC][4];
#pragma unroll
    for (int u = 0; u < SE_UPC; ++u)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[u][j] = 0.f;
    const int kper = C / 16, k0 = wid * kper;
    const float4* wp = reinterpret_cast<const float4*>(w1t + static_cast<long long>(k0) * SE_H) + lane;
#pragma unroll 16
    for (int k = 0; k < kper; ++k) {
      const float4 w = __ldg(wp + static_cast<long long>(k) * (SE_H / 4));
#pragma unroll
      for (int u = 0; u < SE_UPC; ++u) {
        const float m = mean[u * C + k0 + k];
        acc[u][0] = fmaf(m, w.x, acc[u][0]); acc[u][1] = fmaf(m, w.y, acc[u][1]);
        acc[u][2] = fmaf(m, w.z, acc[u][2]); acc[u][3] = fmaf(m, w.w, acc[u][3]);
      }
    }
#pragma unroll
    for (int u = 0; u < SE_UPC; ++u)
      *reinterpret_cast<float4*>(part + (wid * SE_UPC + u) * SE_H + lane * 4) = make_float4(acc[u][0], acc[u][1], acc[u][2], acc[u][3]);
  }
  __syncthreads();
  for (int i = tid; i < SE_UPC * SE_H; i += 512) {
    const int u = i / SE_H, o = i - u * SE_H;
    float s = __ldg(b1 + o);
#pragma unroll
    for (int w = 0; w < 16; ++w) s += part[(w * SE_UPC + u) * SE_H + o];      // fixed order: bit reproducible
    hid[i] = fmaxf(s, 0.f);
  }
  __syncthreads();
  // phase 2: thread (t & 255) owns output channels 4 t .. 4 t + 3 (+ 1024 i); the two thread halves split the hidden
  // units and the upper half hands its partial sums over through shared memory
  const int th = tid & 255, jh = tid >> 8;
  for (int o4 = th * 4; o4 < C; o4 += 1024) {
    float acc[SE_UPC][4];
#pragma unroll
    for (int u = 0; u < SE_UPC; ++u)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[u][j] = 0.f;
    const float4* wp = reinterpret_cast<const float4*>(w2t + static_cast<long long>(jh) * (SE_H / 2) * C + o4);
#pragma unroll 16
    for (int j = 0; j < SE_H / 2; ++j) {
      const float4 w = __ldg(wp + static_cast<long long>(j) * (C / 4));
#pragma unroll
      for (int u = 0; u < SE_UPC; ++u) {
        const float h = hid[u * SE_H + jh * (SE_H / 2) + j];
        acc[u][0] = fmaf(h, w.x, acc[u][0]); acc[u][1] = fmaf(h, w.y, acc[u][1]);
        acc[u][2] = fmaf(h, w.z, acc[u][2]); acc[u][3] = fmaf(h, w.w, acc[u][3]);
      }
    }
    __syncthreads();                                     // (uniform: C is the same for every thread) part is free again
    if (jh == 1) {
#pragma unroll
      for (int u = 0; u < SE_UPC; ++u)
        *reinterpret_cast<float4*>(part + (th * SE_UPC + u) * 4) = make_float4(acc[u][0], acc[u][1], acc[u][2], acc[u][3]);
    }
    __syncthreads();
    if (jh == 0) {
      const float4 bb = __ldg(reinterpret_cast<const float4*>(b2 + o4));
#pragma unroll
      for (int u = 0; u < SE_UPC; ++u) {
        if (u0 + u >= B) break;
        const float4 q = *reinterpret_cast<const float4*>(part + (th * SE_UPC + u) * 4);
        float4 g;
        g.x = 1.f / (1.f + expf(-(acc[u][0] + q.x + bb.x))); g.y = 1.f / (1.f + expf(-(acc[u][1] + q.y + bb.y)));
        g.z = 1.f / (1.f + expf(-(acc[u][2] + q.z + bb.z))); g.w = 1.f / (1.f + expf(-(acc[u][3] + q.w + bb.w)));
        *reinterpret_cast<float4*>(gate + static_cast<long long>(u0 + u) * gate_ld + o4) = g;
      }
    }
  }
}

bool se_gate_supported(int C, int se_ch) { return se_ch == SE_H && C % 1024 == 0 && C <= 4096; }

int se_gate(const long long* sum, int sum_ld, const float* mean_in, int mean_ld, const void* utts, int B, int C,
            const float* w1t, const float* b1, const float* w2t, const float* b2, float* gate, int gate_ld,
            cudaStream_t stream) {
  SVB_CHECK(se_gate_supported(C, SE_H) && gate_ld % 4 == 0 && (sum != nullptr || mean_in != nullptr), "se_gate: unsupported geometry");
  const size_t smem = (static_cast<size_t>(SE_UPC) * C + 17 * SE_UPC * SE_H) * sizeof(float);
  static unsigned long long attr_devs = 0;   // per device: kernel attributes do not carry over
  if (first_on_device(&attr_devs))
    SVB_CUDA_OK(cudaFuncSetAttribute(se_gate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  SVB_CHECK(smem <= 96 * 1024, "se_gate: %zu bytes of shared memory", smem);
  se_gate_kernel<<<(B + SE_UPC - 1) / SE_UPC, 512, smem, stream>>>(sum, sum_ld, mean_in, mean_ld,
                                                                  static_cast<const UttRows*>(utts), B, C, w1t, b1, w2t, b2,
                                                                  gate, gate_ld);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

__global__ void transpose_f32_kernel(const float* __restrict__ in, int R, int C, float* __restrict__ out) {
  const long long total = static_cast<long long>(R) * C;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / C), c = static_cast<int>(i - static_cast<long long>(r) * C);
    out[static_cast<long long>(c) * R + r] = in[i];
  }
}

// out [C][R] = in [R][C]^T (engine creation only)
int transpose_f32(const float* in, int R, int C, float* out, cudaStream_t stream) {
  transpose_f32_kernel<<<256, 256, 0, stream>>>(in, R, C, out);
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// Reflect-halo fix-up after a k > 1 layer that stored every row in place (gemm_pw): halo row -h of every utterance
// becomes a copy of frame h, halo row T-1+h a copy of frame T-1-h (F.pad(..., 'reflect'), src/models/layers/cnn.py:157),
// exactly what gemm_tc's STORE_MIRROR epilogue writes.  grid (utterances, 2 * halo), 16-byte copies of the whole row.
__global__ void mirror_halo_rows_kernel(bf16* __restrict__ x, int ld, int row_bytes, const UttRows* __restrict__ utts,
                                        int halo) {
  const UttRows u = utts[blockIdx.x];
  const int h = static_cast<int>(blockIdx.y % halo) + 1;
  const bool tail = static_cast<int>(blockIdx.y) >= halo;
  const long long dst = tail ? u.row0 + u.T - 1 + h : u.row0 - h;
  const long long src = tail ? u.row0 + u.T - 1 - h : u.row0 + h;
  const uint4* s4 = reinterpret_cast<const uint4*>(x + src * ld);
  uint4* d4 = reinterpret_cast<uint4*>(x + dst * ld);
  for (int i = threadIdx.x; i < row_bytes / 16; i += blockDim.x) d4[i] = s4[i];
}

int mirror_halo_rows(bf16* x, int ld, int cols, const void* utts, int B, int halo, cudaStream_t stream) {
  SVB_CHECK((cols * 2) % 16 == 0 && (ld * 2) % 16 == 0, "mirror_halo_rows copies 16-byte chunks");
  if (halo == 0) return 0;
  mirror_halo_rows_kernel<<<dim3(B, 2 * halo), 256, 0, stream>>>(x, ld, cols * 2, static_cast<const UttRows*>(utts), halo);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

int col_stats(const bf16* x, int ld, int planes, int plane_cols, const void* utts, int B, int C, int mode, float eps,
              float* mean_out, float* std_out, int out_ld, cudaStream_t stream) {
  SVB_CHECK(C % 8 == 0, "C must be a multiple of 8");
  dim3 grid((C + 255) / 256, B);
  col_stats_kernel<<<grid, 256, 0, stream>>>(x, ld, planes, plane_cols, static_cast<const UttRows*>(utts), C, mode, eps,
                                             mean_out, std_out, out_ld);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// part_ws: scratch for the split-K partial sums, at least small_linear_ws_floats(B, J, K) floats
static int sl_splits(int K, int* slice_out) {
  int slice = (K / 16 + SL_KSTEP - 1) / SL_KSTEP * SL_KSTEP;
  if (slice < 128) slice = 128;
  if (slice_out) *slice_out = slice;
  return (K + slice - 1) / slice;
}
size_t small_linear_ws_floats(int B, int J, int K) {
  const int S = sl_splits(K, nullptr);
  return S > 1 ? static_cast<size_t>(S) * B * J : 0;
}

int small_linear(const float* W, int w_ld, const float* bias, const float* in, int in_ld, const float* in_scale,
                 const float* in_shift, int B, int J, int K, int act, float* out, int out_ld, float* part_ws,
                 cudaStream_t stream) {
  SVB_CHECK(K % 4 == 0 && in_ld % 4 == 0 && w_ld % 4 == 0, "small_linear needs 16-byte aligned rows");
  int slice = 0;
  const int S = sl_splits(K, &slice);
  SVB_CHECK(S == 1 || part_ws != nullptr, "small_linear: split-K scratch missing");
  dim3 grid((J + SL_TILE - 1) / SL_TILE, (B + SL_TILE - 1) / SL_TILE, S);
  skinny_linear_kernel<<<grid, 256, 0, stream>>>(W, w_ld, bias, in, in_ld, in_scale, in_shift, B, J, K, slice, act, out,
                                                 out_ld, S > 1 ? part_ws : nullptr);
  count_launch();
  if (S > 1) {
    skinny_reduce_kernel<<<(B * J + 255) / 256, 256, 0, stream>>>(part_ws, S, B, J, bias, act, out, out_ld);
    count_launch();
  }
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

int se_apply(const bf16* x, int x_ld, int x_pc, const bf16* resid, int r_ld, int r_pc, const float* gate, int gate_ld,
             const RowMeta* meta, long long rows, int C, int planes, bf16* out, int o_ld, int o_pc, cudaStream_t stream) {
  const long long total = rows * (C / 8);
  const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, 148LL * 32));
  se_apply_kernel<<<grid, 256, 0, stream>>>(x, x_ld, x_pc, resid, r_ld, r_pc, gate, gate_ld, meta, rows, C, planes, out,
                                            o_ld, o_pc);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

int asp_pool(const float* logits, int l_ld, const bf16* x, int x_ld, int planes, int plane_cols, const void* utts, int B,
             int C, float eps, float* pooled, int pooled_ld, cudaStream_t stream) {
  SVB_CHECK(C % 64 == 0, "channels must be a multiple of 64");
  dim3 grid(C / 64, B);
  asp_pool_kernel<<<grid, 256, 0, stream>>>(logits, l_ld, x, x_ld, planes, plane_cols, static_cast<const UttRows*>(utts),
                                            C, eps, pooled, pooled_ld);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

int l2norm_rows(const float* in, int in_ld, float* out, int out_ld, int B, int D, float eps, cudaStream_t stream) {
  l2norm_rows_kernel<<<(B + 7) / 8, 256, 0, stream>>>(in, in_ld, out, out_ld, B, D, eps);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace svb
