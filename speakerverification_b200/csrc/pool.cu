// Per-utterance reductions and small dense layers of ECAPA-TDNN that are not GEMM-shaped:
//   SEBlock            src/models/ECAPA_TDNN.py:164-177  (time mean -> 2 tiny convs -> gate * x [+ residual :336])
//   AttentiveStatisticsPooling  src/models/ECAPA_TDNN.py:213-260 (global mean/std, softmax over time,
//                               attention-weighted mean/std)
//   asp_bn + fc head   src/models/ECAPA_TDNN.py:496-500
// All statistics are fp32; activations are read from bf16 hi(+lo) planes.
#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

struct UttRows { long long row0; int T; int pad; };  // first VALID row and frame count of an utterance

__device__ __forceinline__ void load8(const bf16* p, int planes, int plane_cols, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int q = 0; q < 4; ++q) { v[2 * q] = bf16_lo_f(w[q]); v[2 * q + 1] = bf16_hi_f(w[q]); }
  if (planes == 2) {
    const uint4 l = *reinterpret_cast<const uint4*>(p + plane_cols);
    const uint32_t z[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) { v[2 * q] += bf16_lo_f(z[q]); v[2 * q + 1] += bf16_hi_f(z[q]); }
  }
}

// grid (C/256, B), block 256: warp w sweeps rows w, w+8, ...; lane owns 8 consecutive channels.
// mode 0: mean only (SE squeeze).  mode 1: mean and std = sqrt(clamp(mean((x-mean)^2), eps)) (ASP global context).
__global__ void __launch_bounds__(256)
col_stats_kernel(const bf16* __restrict__ x, int ld, int planes, int plane_cols, const UttRows* __restrict__ utts,
                 int C, int mode, float eps, float* __restrict__ mean_out, float* __restrict__ std_out, int out_ld) {
  __shared__ float red[8][256];
  __shared__ float mean_s[256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c0 = blockIdx.x * 256 + lane * 8;
  const UttRows u = utts[blockIdx.y];
  const bool active = c0 < C;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  if (active)
    for (int t = warp; t < u.T; t += 8) {
      float v[8];
      load8(x + (u.row0 + t) * ld + c0, planes, plane_cols, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] += v[i];
    }
#pragma unroll
  for (int i = 0; i < 8; ++i) red[warp][lane * 8 + i] = acc[i];
  __syncthreads();
  {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x];
    const float m = s / static_cast<float>(u.T);
    mean_s[threadIdx.x] = m;
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c < C) mean_out[static_cast<long long>(blockIdx.y) * out_ld + c] = m;
  }
  if (mode == 0) return;
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  if (active)
    for (int t = warp; t < u.T; t += 8) {
      float v[8];
      load8(x + (u.row0 + t) * ld + c0, planes, plane_cols, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) { const float d = v[i] - mean_s[lane * 8 + i]; acc[i] = fmaf(d, d, acc[i]); }
    }
#pragma unroll
  for (int i = 0; i < 8; ++i) red[warp][lane * 8 + i] = acc[i];
  __syncthreads();
  {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x];
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c < C) std_out[static_cast<long long>(blockIdx.y) * out_ld + c] = sqrtf(fmaxf(s / static_cast<float>(u.T), eps));
  }
}

// out[b, j] = act( sum_c W[j, c] * (in[b, c] * in_scale[c] + in_shift[c]) + bias[j] ), fp32.
// grid (ceil(J/8), ceil(B/8)), block 256: warp -> output j, 8 utterances share each weight read.
enum { LIN_NONE = 0, LIN_RELU = 1, LIN_SIGMOID = 2 };
__global__ void __launch_bounds__(256)
small_linear_kernel(const float* __restrict__ W, int w_ld, const float* __restrict__ bias, const float* __restrict__ in,
                    int in_ld, const float* __restrict__ in_scale, const float* __restrict__ in_shift, int B, int J,
                    int K, int act, float* __restrict__ out, int out_ld) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * 8 + warp;
  const int b0 = blockIdx.y * 8;
  if (j >= J) return;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  const float* wr = W + static_cast<long long>(j) * w_ld;
  for (int c = lane * 4; c < K; c += 128) {
    const float4 w4 = __ldg(reinterpret_cast<const float4*>(wr + c));
    float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
    if (in_scale != nullptr) {
      sc = __ldg(reinterpret_cast<const float4*>(in_scale + c));
      sh = __ldg(reinterpret_cast<const float4*>(in_shift + c));
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (b0 + i < B) {
        const float4 x4 = *reinterpret_cast<const float4*>(in + static_cast<long long>(b0 + i) * in_ld + c);
        acc[i] = fmaf(w4.x, fmaf(x4.x, sc.x, sh.x), acc[i]);
        acc[i] = fmaf(w4.y, fmaf(x4.y, sc.y, sh.y), acc[i]);
        acc[i] = fmaf(w4.z, fmaf(x4.z, sc.z, sh.z), acc[i]);
        acc[i] = fmaf(w4.w, fmaf(x4.w, sc.w, sh.w), acc[i]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = warp_sum(acc[i]);
  if (lane == 0) {
    const float bj = bias != nullptr ? bias[j] : 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (b0 + i < B) {
        float v = acc[i] + bj;
        if (act == LIN_RELU) v = fmaxf(v, 0.f);
        else if (act == LIN_SIGMOID) v = 1.f / (1.f + expf(-v));
        out[static_cast<long long>(b0 + i) * out_ld + j] = v;
      }
  }
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
    float v[8], r[8];
    load8(x + row * x_ld + c, planes, x_pc, v);
    load8(resid + row * r_ld + c, planes, r_pc, r);
    const float4 g0 = *reinterpret_cast<const float4*>(gate + static_cast<long long>(utt) * gate_ld + c);
    const float4 g1 = *reinterpret_cast<const float4*>(gate + static_cast<long long>(utt) * gate_ld + c + 4);
    const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float a = fmaf(g[2 * q], v[2 * q], r[2 * q]);
      const float b = fmaf(g[2 * q + 1], v[2 * q + 1], r[2 * q + 1]);
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
    float xv = __bfloat162float(x[row * x_ld + c]);
    if (planes == 2) xv += __bfloat162float(x[row * x_ld + plane_cols + c]);
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
    float xv = __bfloat162float(x[row * x_ld + c]);
    if (planes == 2) xv += __bfloat162float(x[row * x_ld + plane_cols + c]);
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

int small_linear(const float* W, int w_ld, const float* bias, const float* in, int in_ld, const float* in_scale,
                 const float* in_shift, int B, int J, int K, int act, float* out, int out_ld, cudaStream_t stream) {
  SVB_CHECK(K % 4 == 0 && in_ld % 4 == 0 && w_ld % 4 == 0, "small_linear needs 16-byte aligned rows");
  dim3 grid((J + 7) / 8, (B + 7) / 8);
  small_linear_kernel<<<grid, 256, 0, stream>>>(W, w_ld, bias, in, in_ld, in_scale, in_shift, B, J, K, act, out, out_ld);
  count_launch();
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
