// RawNet2 raw-waveform encoder forward (reference: models/RawNet2_custom.py:161-243 with the default
// front_proc='sinc', aggregate='gru'; layers models/RawNet_baseline.py:13-24, 48-68, 181-361) as a fixed sequence of
// libsvb kernels.  Weights arrive in the reference's checkpoint layout (fp32 device pointers keyed by state-dict
// name) and are repacked once at create time.
//
//   K_ln      per-utterance LayerNorm over samples  -> bf16 hi|lo signal planes (zero padded to a 48-sample grid)
//   K_sinc    rn_gemm<192, EPI_SINC>: sinc conv + abs + maxpool3 + first_bn + lrelu (+ layer1 bn1 + lrelu)
//   per RawNetBasicBlock:  rn_gemm conv1 (bn2 + lrelu fused), [rn_gemm 1x1 shortcut], rn_gemm conv2 (+ shortcut),
//             K_pool (maxpool3 + AFMS time sums), K_fc (AFMS sigmoid(fc(mean))), K_apply ((x + alpha) * y and the
//             NEXT block's bn1 + lrelu, or bn_before_gru + lrelu after the last block)
//   GRU       input projection for all 29 steps as one GEMM; per step one [B,1024]x[1024,3072] split-bf16 GEMM and
//             a gate kernel; fc_after_gru as K_fc without the sigmoid.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include <cuda_fp16.h>

#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

int fold_bn(const float*, const float*, const float*, const float*, int, float, float*, float*, cudaStream_t);
int split_bf16(const float*, long long, int, int, bf16*, int, int, int, cudaStream_t);
int l2norm_rows(const float*, int, float*, int, int, int, float, cudaStream_t);
int launch_rn_gemm(int, const bf16*, long long, long long, long long, const bf16*, int, int, const RnGemmParams&,
                   cudaStream_t);
bool rn_conv2_pool_supported(int channels, int planes, int xplanes, int T);
enum { EPI_SINC = 0, EPI_AFF_LRELU = 1, EPI_ADD = 2, EPI_PLAIN = 3 };

static constexpr int SINC_STRIDE = 48;     // samples per GEMM row (= 16 pooled outputs)
static constexpr int SINC_K = 320;         // 251 taps + 47 shifts, padded to 5 x 64
static constexpr int SINC_TAPS = 251;
static constexpr int NB = 8;               // residual blocks
static const int RN_CIN[NB] = {128, 128, 128, 256, 256, 256, 512, 512};
static const int RN_COUT[NB] = {128, 128, 256, 256, 256, 512, 512, 512};
static const int RN_DOWN[NB] = {1, 1, 1, 0, 1, 1, 0, 1};
static const char* RN_NAME[NB] = {"layer1.0", "layer2.0", "layer3.0", "layer4.0", "layer4.1", "layer5.0", "layer6.0",
                                  "layer6.1"};

// ------------------------------------------------------------------ elementwise kernels
__device__ __forceinline__ float lrelu03(float x) { return x > 0.f ? x : 0.3f * x; }

// 8 consecutive channels of one row -> fp32: bf16 hi + lo planes (planes == 2, fp32-grade mode) or ONE plane of IEEE
// half (planes == 1, 16-bit mode; the CPU model of the data flow puts its trial-score deviation at 1.4e-3 against
// 1.0e-2 for bf16 operands, profiles/r02_precision.md)
__device__ __forceinline__ void load8(const bf16* p, int planes, int plane_cols, float (&v)[8]) {
  const uint4 u = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
  if (planes == 1) {
#pragma unroll
    for (int q = 0; q < 4; ++q) { v[2 * q] = f16_lo_f(w[q]); v[2 * q + 1] = f16_hi_f(w[q]); }
    return;
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) { v[2 * q] = bf16_lo_f(w[q]); v[2 * q + 1] = bf16_hi_f(w[q]); }
  if (planes == 2) {
    const uint4 l = __ldg(reinterpret_cast<const uint4*>(p + plane_cols));
    const uint32_t x[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) { v[2 * q] += bf16_lo_f(x[q]); v[2 * q + 1] += bf16_hi_f(x[q]); }
  }
}
__device__ __forceinline__ void store8(bf16* p, int planes, int plane_cols, const float (&v)[8]) {
  uint32_t hi[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) hi[q] = planes == 1 ? pack_f16x2(v[2 * q], v[2 * q + 1]) : pack_bf16x2(v[2 * q], v[2 * q + 1]);
  *reinterpret_cast<uint4*>(p) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
  if (planes == 2) {
    uint32_t lo[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) lo[q] = pack_bf16x2(v[2 * q] - bf16_lo_f(hi[q]), v[2 * q + 1] - bf16_hi_f(hi[q]));
    *reinterpret_cast<uint4*>(p + plane_cols) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
}

// raw 16-byte loads now, decode later: the elementwise kernels below are pure HBM streams, and a thread that decodes
// (or stores) before it has issued its next load keeps one 16-byte request in flight -- rn_apply / rn_pool ran at 30-45 %
// of the copy peak that way.  They issue the loads of four rows (three for the pooling window) before touching any.
struct Raw8 { uint4 hi, lo; };
__device__ __forceinline__ Raw8 ldraw8(const bf16* p, int planes, int plane_cols) {
  Raw8 r;
  r.hi = __ldg(reinterpret_cast<const uint4*>(p));
  r.lo = planes == 2 ? __ldg(reinterpret_cast<const uint4*>(p + plane_cols)) : make_uint4(0, 0, 0, 0);
  return r;
}
__device__ __forceinline__ void decode8(const Raw8& r, int planes, float (&v)[8]) {
  const uint32_t w[4] = {r.hi.x, r.hi.y, r.hi.z, r.hi.w}, x[4] = {r.lo.x, r.lo.y, r.lo.z, r.lo.w};
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    if (planes == 1) { v[2 * q] = f16_lo_f(w[q]); v[2 * q + 1] = f16_hi_f(w[q]); }
    else { v[2 * q] = bf16_lo_f(w[q]) + bf16_lo_f(x[q]); v[2 * q + 1] = bf16_hi_f(w[q]) + bf16_hi_f(x[q]); }
  }
}

// K_ln: LayerNorm over the samples of one utterance (RawNet_baseline.py:20-24: unbiased std, eps added to std).
// One CTA per utterance.  Output: bf16 hi plane at xn[utt * Lp + i], lo plane at + plane_dist; [L, Lp) zero.
__global__ void __launch_bounds__(1024)
rn_ln_kernel(const float* __restrict__ wav, long long wav_ld, int L, int Lp, const float* __restrict__ gamma,
             const float* __restrict__ beta, float eps, bf16* __restrict__ xn, long long plane_dist, int planes) {
  __shared__ double red[32];
  __shared__ float stat[2];
  const int utt = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const float* x = wav + static_cast<long long>(utt) * wav_ld;
  double s = 0.0;
  for (int i = tid; i < L; i += 1024) s += x[i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) red[wid] = s;
  __syncthreads();
  if (tid == 0) { double t = 0.0; for (int w = 0; w < 32; ++w) t += red[w]; stat[0] = static_cast<float>(t / L); }
  __syncthreads();
  const float mean = stat[0];
  double q = 0.0;
  for (int i = tid; i < L; i += 1024) { const float d = x[i] - mean; q += static_cast<double>(d) * d; }
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  __syncthreads();
  if (lane == 0) red[wid] = q;
  __syncthreads();
  if (tid == 0) { double t = 0.0; for (int w = 0; w < 32; ++w) t += red[w]; stat[1] = static_cast<float>(sqrt(t / (L - 1))); }
  __syncthreads();
  const float inv = 1.0f / (stat[1] + eps);
  bf16* hi = xn + static_cast<long long>(utt) * Lp;
  for (int i = tid; i < Lp; i += 1024) {
    float v = 0.f;
    if (i < L) v = gamma[i] * (x[i] - mean) * inv + beta[i];
    if (planes == 1) { reinterpret_cast<__half*>(hi)[i] = __float2half_rn(v); continue; }
    const bf16 h = __float2bfloat16_rn(v);
    hi[i] = h;
    hi[i + plane_dist] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

// K_pool: max_pool1d(3) over time (floor) and per-chunk time sums of the pooled rows for AFMS.
// z rows [B * (Tin + 2)], zp rows [B * (Tout + 2)] (valid rows written); psum [B, nchunk, C].
// grid (nchunk, B); thread = 8 channels x a strided set of output frames.  pool == 1: no zp, sums only.
__global__ void __launch_bounds__(256)
rn_pool_kernel(const bf16* __restrict__ z, int z_ld, int planes, int plane_cols, int Tin, int pool, bf16* __restrict__ zp,
               int zp_ld, int zp_plane_cols, int Tout, int C, int frames_per_chunk, float* __restrict__ psum) {
  extern __shared__ float acc_s[];   // [groups_t][C]
  const int utt = blockIdx.y, chunk = blockIdx.x;
  const int cg = C / 8;                    // channel groups
  const int tid = threadIdx.x;
  const int g = tid % cg, tl = tid / cg, tstep = 256 / cg;
  const int t0 = chunk * frames_per_chunk, t1 = min(Tout, t0 + frames_per_chunk);
  float sum[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  // six row loads are issued before the first decode: two output frames x three pooled rows, or (pool == 1: sums only)
  // six frames
  const int per = pool == 3 ? 2 : 6;
  for (int tb = t0 + tl; tb < t1; tb += per * tstep) {
    Raw8 raw[2][3];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int t = min(tb + (pool == 3 ? k : 3 * k + j) * tstep, t1 - 1);
        const long long r = static_cast<long long>(utt) * (Tin + 2) + 1 + (pool == 3 ? static_cast<long long>(t) * 3 + j : t);
        raw[k][j] = ldraw8(z + r * z_ld + g * 8, planes, plane_cols);
      }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      if (pool == 3) {
        const int t = tb + k * tstep;
        if (t >= t1) break;
        float v[8], a[8], b[8];
        decode8(raw[k][0], planes, v);
        decode8(raw[k][1], planes, a);
        decode8(raw[k][2], planes, b);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], fmaxf(a[i], b[i]));
        const long long ro = static_cast<long long>(utt) * (Tout + 2) + 1 + t;
        store8(zp + ro * zp_ld + g * 8, planes, zp_plane_cols, v);
#pragma unroll
        for (int i = 0; i < 8; ++i) sum[i] += v[i];
      } else {
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          if (tb + (3 * k + j) * tstep >= t1) break;
          float v[8];
          decode8(raw[k][j], planes, v);
#pragma unroll
          for (int i = 0; i < 8; ++i) sum[i] += v[i];
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) acc_s[tl * C + g * 8 + i] = sum[i];
  __syncthreads();
  for (int c = tid; c < C; c += 256) {
    float s = 0.f;
    for (int k = 0; k < tstep; ++k) s += acc_s[k * C + c];
    psum[(static_cast<long long>(utt) * gridDim.x + chunk) * C + c] = s;
  }
}

// K_fc: out[u, o] = act( sum_j W[o, j] * in[u, j] + b[o] ), in = (sum over nchunk partial rows) * in_scale.
// Used for AFMS (sigmoid, RawNet_baseline.py:61-62) and fc_after_gru (no activation, RawNet2_custom.py:198).
// grid (ceil(Cout / 32), ceil(B / 4)); 8 warps; each warp 4 outputs x 4 utterances.
__global__ void __launch_bounds__(256)
rn_fc_kernel(const float* __restrict__ in, int nchunk, float in_scale, int Cin, const float* __restrict__ W,
             const float* __restrict__ b, int Cout, int B, int sigmoid, float* __restrict__ out, int out_ld) {
  extern __shared__ float xin[];   // [4][Cin]
  const int u0 = blockIdx.y * 4;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (int i = tid; i < 4 * Cin; i += 256) {
    const int uu = i / Cin, j = i - uu * Cin;
    float s = 0.f;
    if (u0 + uu < B)
      for (int k = 0; k < nchunk; ++k) s += in[(static_cast<long long>(u0 + uu) * nchunk + k) * Cin + j];
    xin[i] = s * in_scale;
  }
  __syncthreads();
  for (int oo = 0; oo < 4; ++oo) {
    const int o = blockIdx.x * 32 + wid * 4 + oo;
    if (o >= Cout) break;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    const float* wr = W + static_cast<long long>(o) * Cin;
    for (int j = lane; j < Cin; j += 32) {
      const float w = __ldg(wr + j);
      a0 = fmaf(w, xin[j], a0); a1 = fmaf(w, xin[Cin + j], a1);
      a2 = fmaf(w, xin[2 * Cin + j], a2); a3 = fmaf(w, xin[3 * Cin + j], a3);
    }
    a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2); a3 = warp_sum(a3);
    if (lane < 4) {
      float v = lane == 0 ? a0 : (lane == 1 ? a1 : (lane == 2 ? a2 : a3));
      v += b[o];
      if (sigmoid) v = 1.0f / (1.0f + __expf(-v));
      if (u0 + lane < B) out[static_cast<long long>(u0 + lane) * out_ld + o] = v;
    }
  }
}

// K_apply: AFMS scaling (RawNet_baseline.py:64-66) and the NEXT pre-activation (bn1 / bn_before_gru + LeakyReLU 0.3).
//   x = (zp + alpha[c]) * gate[u, c]  -> xo (valid rows; only when the next block's shortcut is the identity)
//   a = lrelu(s[c] * x + h[c])        -> ao (zero on the padding rows)
// grid (row chunks, utterances); a thread owns 8 channels (per-channel constants in registers) and walks rows.
__global__ void __launch_bounds__(256)
rn_apply_kernel(const bf16* __restrict__ zp, int zp_ld, int planes, int zp_plane_cols, int T, int C,
                const float* __restrict__ alpha, const float* __restrict__ gate, const float* __restrict__ s,
                const float* __restrict__ h, bf16* __restrict__ xo, int xo_ld, int xo_plane_cols, bf16* __restrict__ ao,
                int ao_ld, int ao_planes, int ao_plane_cols, int rows_per_cta) {
  const int cg = C >> 3;
  const int utt = blockIdx.y;
  const int g = threadIdx.x % cg, tl = threadIdx.x / cg, tstep = 256 / cg;
  const int c = g * 8;
  float al[8], gt[8], sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    al[i] = __ldg(alpha + c + i);
    gt[i] = __ldg(gate + static_cast<long long>(utt) * C + c + i);
    sc[i] = __ldg(s + c + i);
    sh[i] = __ldg(h + c + i);
  }
  const int r0 = blockIdx.x * rows_per_cta;                 // row inside the utterance: 0 .. T + 1 (t = r - 1)
  const int r1 = min(T + 2, r0 + rows_per_cta);
  const long long base = static_cast<long long>(utt) * (T + 2);
  for (int rb = r0 + tl; rb < r1; rb += 4 * tstep) {
    Raw8 raw[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int r = rb + k * tstep;
      if (r >= 1 && r <= T && r < r1) raw[k] = ldraw8(zp + (base + r) * zp_ld + c, planes, zp_plane_cols);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int r = rb + k * tstep;
      if (r >= r1) break;
      const long long row = base + r;
      float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (r >= 1 && r <= T) {
        float v[8];
        decode8(raw[k], planes, v);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float x = (v[i] + al[i]) * gt[i];
          v[i] = x;
          a[i] = lrelu03(fmaf(x, sc[i], sh[i]));
        }
        if (xo != nullptr) store8(xo + row * xo_ld + c, planes, xo_plane_cols, v);
      }
      store8(ao + row * ao_ld + c, ao_planes, ao_plane_cols, a);
    }
  }
}

// GRU gates (torch.nn.GRU cell): r = sig(gi_r + gh_r), z = sig(gi_z + gh_z), n = tanh(gi_n + r * gh_n),
// h' = (1 - z) * n + z * h.  gi already holds b_ih; gh (may be null at t = 0: h = 0) gets b_hh here.
__global__ void __launch_bounds__(256)
rn_gru_gate_kernel(const float* __restrict__ gi, long long gi_row_stride, long long gi_utt_stride, const float* __restrict__ gh,
                   int gh_ld, const float* __restrict__ b_hh, int H, int B, float* __restrict__ h, bf16* __restrict__ h16,
                   int h16_ld) {
  const long long idx = blockIdx.x * 256LL + threadIdx.x;
  if (idx >= static_cast<long long>(B) * H) return;
  const int u = static_cast<int>(idx / H), j = static_cast<int>(idx - static_cast<long long>(u) * H);
  const float* g = gi + u * gi_utt_stride;
  (void)gi_row_stride;
  float hr = b_hh[j], hz = b_hh[H + j], hn = b_hh[2 * H + j];
  if (gh != nullptr) {
    const float* q = gh + static_cast<long long>(u) * gh_ld;
    hr += q[j]; hz += q[H + j]; hn += q[2 * H + j];
  }
  const float r = 1.0f / (1.0f + expf(-(g[j] + hr)));
  const float z = 1.0f / (1.0f + expf(-(g[H + j] + hz)));
  const float n = tanhf(g[2 * H + j] + r * hn);
  const float hp = h[idx];
  const float hnew = (1.0f - z) * n + z * hp;
  h[idx] = hnew;
  const bf16 hi = __float2bfloat16_rn(hnew);
  h16[static_cast<long long>(u) * h16_ld + j] = hi;
  h16[static_cast<long long>(u) * h16_ld + H + j] = __float2bfloat16_rn(hnew - __bfloat162float(hi));
}

// zero padding rows (first and last row of every utterance) of a [B * (T + 2), ld] matrix
__global__ void rn_zero_halo_kernel(bf16* __restrict__ a, int ld, int T, int B) {
  const int u = blockIdx.x >> 1, which = blockIdx.x & 1;
  if (u >= B) return;
  uint4* row = reinterpret_cast<uint4*>(a + (static_cast<long long>(u) * (T + 2) + (which ? T + 1 : 0)) * ld);
  for (int i = threadIdx.x; i < ld / 8; i += blockDim.x) row[i] = make_uint4(0, 0, 0, 0);
}

__global__ void rn_fill_kernel(float* p, long long n, float v) {
  const long long i = blockIdx.x * 256LL + threadIdx.x;
  if (i < n) p[i] = v;
}

// ------------------------------------------------------------------ model
struct RnBlock {
  int cin = 0, cout = 0, down = 0;
  bf16 *w1 = nullptr, *w2 = nullptr, *wsc = nullptr;      // packed [cout, 3, planes, cin] / [cout, 1, planes, cin]
  float *s1 = nullptr, *h1 = nullptr;                     // bn1 (applied by the PREVIOUS stage's epilogue)
  float *s2 = nullptr, *h2 = nullptr;                     // bn2 (conv1 epilogue)
  float *alpha = nullptr, *fc_w = nullptr, *fc_b = nullptr;
};

// ------------------------------------------------------------------ attentive aggregation heads (aggregate = 'asp' / 'sap')
// RawNet2_custom.py:200-224 on the 29 frames that leave the last residual block; x = lrelu(bn_before_agg(.)) arrives
// pre-activated from the last rn_apply_kernel.  The heads are 0.06 % of the network's FLOPs: one CTA per utterance,
// fp32 SIMT, x staged once in shared memory, weights read coalesced from their TRANSPOSED copies (made at create).
//   asp: w = softmax_T(conv2(BN(LeakyReLU_0.01(conv1(x)))));  m = sum_t x w;  s = sqrt(clamp(sum_t x^2 w - m^2, 1e-5))
//   sap: a_t = tanh(v . (W x_t + b));  w = softmax_T(a);  mean = (sum_t w_t x_t) / T;  var = sum_t w_t x_t^2 - mean^2
//   out = fc(cat(m, s))   (fc: Linear(2 * 512, code_dim), shared by both heads)
enum { AGG_GRU = 0, AGG_SAP = 1, AGG_ASP = 2 };
static constexpr int AGG_C = 512;
static constexpr int AGG_TMAX = 32;

__global__ void transpose_kernel(const float* __restrict__ in, int R, int C, float* __restrict__ out) {
  const long long total = static_cast<long long>(R) * C;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / C), c = static_cast<int>(i - static_cast<long long>(r) * C);
    out[static_cast<long long>(c) * R + r] = in[i];
  }
}

// grid B, block 256.  dynamic smem: x_s [T][512] + h_s [T][att] (asp) / a_s [T] (sap) + pooled [1024]
__global__ void __launch_bounds__(256)
rn_agg_kernel(const bf16* __restrict__ a, int a_ld, int planes, int T, int Tp, int mode, int att,
              const float* __restrict__ w1t, const float* __restrict__ b1, const float* __restrict__ bn_s,
              const float* __restrict__ bn_h, const float* __restrict__ w2t, const float* __restrict__ b2,
              const float* __restrict__ sap_v, const float* __restrict__ fc_w, const float* __restrict__ fc_b, int n_out,
              float* __restrict__ emb, int emb_ld) {
  extern __shared__ float sm[];
  float* x_s = sm;                                   // [T][512]
  float* h_s = x_s + T * AGG_C;                      // asp: [T][att]; sap: [T] attention logits (first T floats)
  float* pooled = h_s + T * (mode == AGG_ASP ? att : 1) + 32;
  const int u = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  // stage x: rows u * Tp + 1 + t (valid rows), 8 channels per thread and step
  for (int i = tid; i < T * (AGG_C / 8); i += 256) {
    const int t = i / (AGG_C / 8), c = (i - t * (AGG_C / 8)) * 8;
    float v[8];
    load8(a + (static_cast<long long>(u) * Tp + 1 + t) * a_ld + c, planes, AGG_C, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) x_s[t * AGG_C + c + k] = v[k];
  }
  __syncthreads();
  if (mode == AGG_ASP) {
    // conv1 (512 -> att) + LeakyReLU(0.01) + BN: thread = (att channel, half of the frames)
    const int half_t = (T + 1) / 2;
    for (int o = tid; o < 2 * att; o += 256) {
      const int ac = o % att, t0 = (o / att) * half_t, t1 = min(T, t0 + half_t);
      float acc[AGG_TMAX / 2];
#pragma unroll
      for (int t = 0; t < AGG_TMAX / 2; ++t) acc[t] = 0.f;
      for (int c = 0; c < AGG_C; ++c) {
        const float w = __ldg(w1t + static_cast<long long>(c) * att + ac);
#pragma unroll
        for (int t = 0; t < AGG_TMAX / 2; ++t)
          if (t0 + t < t1) acc[t] = fmaf(w, x_s[(t0 + t) * AGG_C + c], acc[t]);
      }
#pragma unroll
      for (int t = 0; t < AGG_TMAX / 2; ++t)
        if (t0 + t < t1) {
          float v = acc[t] + b1[ac];
          v = v > 0.f ? v : 0.01f * v;
          h_s[(t0 + t) * att + ac] = fmaf(v, bn_s[ac], bn_h[ac]);
        }
    }
    __syncthreads();
    // conv2 (att -> 512), softmax over time, weighted statistics: thread = channel
    for (int c = tid; c < AGG_C; c += 256) {
      float lg[AGG_TMAX];
#pragma unroll
      for (int t = 0; t < AGG_TMAX; ++t) lg[t] = 0.f;
      for (int k = 0; k < att; ++k) {
        const float w = __ldg(w2t + static_cast<long long>(k) * AGG_C + c);
#pragma unroll
        for (int t = 0; t < AGG_TMAX; ++t)
          if (t < T) lg[t] = fmaf(w, h_s[t * att + k], lg[t]);
      }
      float mx = -INFINITY;
#pragma unroll
      for (int t = 0; t < AGG_TMAX; ++t) if (t < T) { lg[t] += b2[c]; mx = fmaxf(mx, lg[t]); }
      float den = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int t = 0; t < AGG_TMAX; ++t)
        if (t < T) {
          const float e = expf(lg[t] - mx), xv = x_s[t * AGG_C + c];
          den += e; s1 = fmaf(e, xv, s1); s2 = fmaf(e * xv, xv, s2);
        }
      const float m = s1 / den;
      pooled[c] = m;
      pooled[AGG_C + c] = sqrtf(fmaxf(s2 / den - m * m, 1e-5f));
    }
  } else {
    // sap: a_t = tanh(v . (W x_t + b)) = tanh(sum_o v_o (b_o + sum_c W[o, c] x_t[c])); thread = (o, frame half)
    const int half_t = (T + 1) / 2;
    float part[AGG_TMAX / 2];
#pragma unroll
    for (int t = 0; t < AGG_TMAX / 2; ++t) part[t] = 0.f;
    const int grp = tid >> 7;                         // frame half of this thread
    const int t0 = grp * half_t, t1 = min(T, t0 + half_t);
    for (int o = tid & 127; o < AGG_C; o += 128) {
      float acc[AGG_TMAX / 2];
#pragma unroll
      for (int t = 0; t < AGG_TMAX / 2; ++t) acc[t] = 0.f;
      for (int c = 0; c < AGG_C; ++c) {
        const float w = __ldg(w1t + static_cast<long long>(c) * AGG_C + o);
#pragma unroll
        for (int t = 0; t < AGG_TMAX / 2; ++t)
          if (t0 + t < t1) acc[t] = fmaf(w, x_s[(t0 + t) * AGG_C + c], acc[t]);
      }
      const float vo = sap_v[o], bo = b1[o];
#pragma unroll
      for (int t = 0; t < AGG_TMAX / 2; ++t) part[t] = fmaf(vo, acc[t] + bo, part[t]);
    }
    // reduce the per-thread partial dot products over the 128 threads of each frame half (fixed order: deterministic)
    float* red = pooled;                              // [256][T/2] scratch, reused before `pooled` is written
#pragma unroll
    for (int t = 0; t < AGG_TMAX / 2; ++t) if (t < half_t) red[tid * half_t + t] = part[t];
    __syncthreads();
    if (tid < T) {
      const int g2 = tid / half_t, tt = tid - g2 * half_t;
      float s = 0.f;
      for (int k = 0; k < 128; ++k) s += red[(g2 * 128 + k) * half_t + tt];
      h_s[tid] = tanhf(s);
    }
    __syncthreads();
    float mx = -INFINITY;
    for (int t = 0; t < T; ++t) mx = fmaxf(mx, h_s[t]);
    float den = 0.f;
    for (int t = 0; t < T; ++t) den += expf(h_s[t] - mx);
    __syncthreads();                                  // everyone has read `red` before pooled overwrites it
    for (int c = tid; c < AGG_C; c += 256) {
      float s1 = 0.f, s2 = 0.f;
      for (int t = 0; t < T; ++t) {
        const float w = expf(h_s[t] - mx) / den, xv = x_s[t * AGG_C + c];
        s1 = fmaf(w, xv, s1); s2 = fmaf(w * xv, xv, s2);
      }
      const float mean = s1 / static_cast<float>(T);  // torch.mean over time of the weighted frames (:153)
      pooled[c] = mean;
      pooled[AGG_C + c] = s2 - mean * mean;           // weighted_sd (:146-150): a variance, no square root
    }
  }
  __syncthreads();
  // fc: Linear(1024 -> n_out), one warp per output (coalesced weight rows)
  for (int o = wid; o < n_out; o += 8) {
    const float* wr = fc_w + static_cast<long long>(o) * 2 * AGG_C;
    float acc = 0.f;
    for (int k = lane; k < 2 * AGG_C; k += 32) acc = fmaf(__ldg(wr + k), pooled[k], acc);
    acc = warp_sum(acc);
    if (lane == 0) emb[static_cast<long long>(u) * emb_ld + o] = acc + fc_b[o];
  }
}

struct Rawnet2 {
  int L = 0, Lp = 0, Mu = 0, T1 = 0, n_out = 0, H = 1024, sr = 16000;
  // bf16 planes of the tensor-core operands (1: bf16 mode, 2: hi|lo fp32 grade) and of the residual carriers.
  // Measured (tools/rawnet_bf16_err.py): hi|lo carriers in bf16 mode do not lower the error (it is dominated by the
  // operand rounding of the first layers, amplified ~1.4x per block by the network) and cost 30 % throughput.
  int planes = 1, xplanes = 1;
  int T[NB + 1] = {};            // frames entering block i; T[NB] = frames entering the GRU
  float *ln_g = nullptr, *ln_b = nullptr;
  bf16* sinc_w = nullptr;        // [6144, planes * 320]
  float *fbn_s = nullptr, *fbn_h = nullptr;
  RnBlock blk[NB];
  float *gbn_s = nullptr, *gbn_h = nullptr;
  bf16 *w_ih = nullptr, *w_hh = nullptr;                 // packed [3H, planes, 512] / [3H, 2, H]
  float *b_ih = nullptr, *b_hh = nullptr, *ones = nullptr, *zeros = nullptr;
  float *fcg_w = nullptr, *fcg_b = nullptr;
  // aggregate = 'sap' / 'asp' heads (RawNet2_custom.py:97-112,200-224); gbn_s / gbn_h then hold bn_before_agg
  int aggregate = AGG_GRU, att = 128;
  float *ag_w1t = nullptr, *ag_b1 = nullptr, *ag_bn_s = nullptr, *ag_bn_h = nullptr, *ag_w2t = nullptr, *ag_b2 = nullptr;
  float *ag_v = nullptr, *ag_fc_w = nullptr, *ag_fc_b = nullptr;
  std::vector<void*> owned;
};

typedef std::map<std::string, const float*> TensorMap;
static const float* find(const TensorMap& tm, const std::string& key) {
  auto it = tm.find(key);
  if (it == tm.end()) { set_error("missing tensor '%s'", key.c_str()); return nullptr; }
  return it->second;
}
template <typename T>
static T* dalloc(Rawnet2* e, size_t n) {
  void* p = nullptr;
  if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
  e->owned.push_back(p);
  return static_cast<T*>(p);
}
static float* dcopy(Rawnet2* e, const float* src, size_t n) {
  float* d = dalloc<float>(e, n);
  if (d == nullptr) return nullptr;
  cudaMemcpy(d, src, n * sizeof(float), cudaMemcpyDeviceToDevice);
  return d;
}
static int make_bn(Rawnet2* e, const TensorMap& tm, const std::string& p, int c, float** s, float** h) {
  const float *g = find(tm, p + "weight"), *b = find(tm, p + "bias"), *m = find(tm, p + "running_mean"),
              *v = find(tm, p + "running_var");
  if (!g || !b || !m || !v) return -1;
  *s = dalloc<float>(e, c); *h = dalloc<float>(e, c);
  SVB_CHECK(*s && *h, "out of device memory");
  return fold_bn(g, b, m, v, c, 1e-5f, *s, *h, 0);
}
static bf16* make_conv(Rawnet2* e, const TensorMap& tm, const std::string& key, int cout, int cin, int k) {
  const float* w = find(tm, key);
  if (!w) return nullptr;
  bf16* d = dalloc<bf16>(e, static_cast<size_t>(cout) * k * e->planes * cin);
  if (d == nullptr) { set_error("out of device memory"); return nullptr; }
  if (pack_conv_weight(w, cout, cin, k, cin, e->planes, d, 0, e->planes == 1 ? 1 : 0)) return nullptr;
  return d;
}

// Band-pass FIRs of SincConv_fast.forward (RawNet_baseline.py:339-357) from (low_hz_, band_hz_), computed on the
// host in fp32 with the reference's operation order, then expanded into the 48-shift GEMM operand
//   W'[((jg * 128 + f) * 3 + ph), k'] = h[f, k' - (3 jg + ph)]   (0 outside the 251 taps), bf16 hi | lo planes.
static int build_sinc(Rawnet2* e, const float* low_d, const float* band_d, int F) {
  std::vector<float> low(F), band(F);
  SVB_CUDA_OK(cudaMemcpy(low.data(), low_d, F * sizeof(float), cudaMemcpyDeviceToHost));
  SVB_CUDA_OK(cudaMemcpy(band.data(), band_d, F * sizeof(float), cudaMemcpyDeviceToHost));
  const int K = SINC_TAPS, half = K / 2;       // 125
  const float pi = 3.14159265358979323846f;
  std::vector<float> win(half), n_(half);
  {
    // torch.linspace(0, K/2 - 1, steps = 125) (symmetric evaluation) and 2*pi*arange(-125, 0)/sr, in fp32
    const float start = 0.f, end = static_cast<float>(K / 2.0 - 1.0);
    const float step = (end - start) / static_cast<float>(half - 1);
    for (int i = 0; i < half; ++i) {
      const float lin = (i < half / 2) ? start + step * i : end - step * (half - 1 - i);
      win[i] = 0.54f - 0.46f * cosf(2.0f * pi * lin / static_cast<float>(K));
      n_[i] = 2.0f * pi * static_cast<float>(i - half) / static_cast<float>(e->sr);
    }
  }
  std::vector<float> h(static_cast<size_t>(F) * K);
  for (int f = 0; f < F; ++f) {
    const float lo = 50.0f + fabsf(low[f]);
    float hi = lo + 50.0f + fabsf(band[f]);
    hi = fminf(fmaxf(hi, 50.0f), static_cast<float>(e->sr) / 2.0f);
    const float bw = hi - lo;
    for (int i = 0; i < half; ++i) {
      const float fl = lo * n_[i], fh = hi * n_[i];
      const float left = ((sinf(fh) - sinf(fl)) / (n_[i] / 2.0f)) * win[i];
      const float v = left / (2.0f * bw);
      h[static_cast<size_t>(f) * K + i] = v;
      h[static_cast<size_t>(f) * K + (K - 1 - i)] = v;
    }
    h[static_cast<size_t>(f) * K + half] = (2.0f * bw) / (2.0f * bw);
  }
  const int N = 16 * F * 3, planes = e->planes;
  std::vector<bf16> w(static_cast<size_t>(N) * planes * SINC_K, __float2bfloat16(0.f));
  for (int jg = 0; jg < 16; ++jg)
    for (int f = 0; f < F; ++f)
      for (int ph = 0; ph < 3; ++ph) {
        const size_t n = (static_cast<size_t>(jg) * F + f) * 3 + ph;
        const int j = 3 * jg + ph;
        for (int k = 0; k < K; ++k) {
          const float v = h[static_cast<size_t>(f) * K + k];
          if (planes == 1) {       // 16-bit mode: IEEE half (shares the activations' format)
            const __half hv = __float2half(v);
            memcpy(&w[n * planes * SINC_K + j + k], &hv, sizeof(hv));
            continue;
          }
          const bf16 hi = __float2bfloat16(v);
          w[n * planes * SINC_K + j + k] = hi;
          w[n * planes * SINC_K + SINC_K + j + k] = __float2bfloat16(v - __bfloat162float(hi));
        }
      }
  e->sinc_w = dalloc<bf16>(e, w.size());
  SVB_CHECK(e->sinc_w != nullptr, "out of device memory");
  SVB_CUDA_OK(cudaMemcpy(e->sinc_w, w.data(), w.size() * sizeof(bf16), cudaMemcpyHostToDevice));
  return 0;
}

static float* dtranspose(Rawnet2* e, const float* src, int R, int C) {
  float* d = dalloc<float>(e, static_cast<size_t>(R) * C);
  if (d == nullptr) return nullptr;
  transpose_kernel<<<std::min((R * C + 255) / 256, 148 * 8), 256>>>(src, R, C, d);
  return d;
}

int rawnet2_create(int n_samples, int n_out, int gru_node, int sample_rate, int fp32_mode, int aggregate, int att_dim,
                   const char* const* names, const void* const* ptrs, int n, Rawnet2** out) {
  SVB_CHECK(n_samples > 1000 && n_out >= 1 && gru_node == 1024, "unsupported RawNet2 configuration (gru_node must be 1024)");
  SVB_CHECK(aggregate == AGG_GRU || aggregate == AGG_SAP || aggregate == AGG_ASP, "aggregate must be 0 (gru), 1 (sap) or 2 (asp)");
  SVB_CHECK(aggregate != AGG_ASP || (att_dim >= 1 && att_dim <= 512), "att_dim must be 1..512");
  TensorMap tm;
  for (int i = 0; i < n; ++i) tm[names[i]] = static_cast<const float*>(ptrs[i]);
  Rawnet2* e = new Rawnet2();
  e->L = n_samples; e->n_out = n_out; e->H = gru_node; e->sr = sample_rate; e->planes = fp32_mode ? 2 : 1; e->xplanes = e->planes;
  e->aggregate = aggregate; e->att = att_dim;
  const int T0 = n_samples - (SINC_TAPS - 1);
  e->T1 = T0 / 3;
  e->Mu = (n_samples + SINC_K + SINC_STRIDE - 1) / SINC_STRIDE;
  e->Lp = e->Mu * SINC_STRIDE;
  e->T[0] = e->T1;
  for (int i = 0; i < NB; ++i) e->T[i + 1] = RN_DOWN[i] ? e->T[i] / 3 : e->T[i];
  int rc = -1;
  do {
    const float *g = find(tm, "ln.gamma"), *b = find(tm, "ln.beta");
    const float *lo = find(tm, "first_conv.low_hz_"), *bd = find(tm, "first_conv.band_hz_");
    if (!g || !b || !lo || !bd) break;
    e->ln_g = dcopy(e, g, n_samples); e->ln_b = dcopy(e, b, n_samples);
    if (build_sinc(e, lo, bd, 128)) break;
    if (make_bn(e, tm, "first_bn.", 128, &e->fbn_s, &e->fbn_h)) break;
    bool ok = true;
    for (int i = 0; i < NB && ok; ++i) {
      RnBlock& k = e->blk[i];
      k.cin = RN_CIN[i]; k.cout = RN_COUT[i]; k.down = RN_DOWN[i];
      const std::string p = std::string(RN_NAME[i]) + ".";
      ok = ok && !make_bn(e, tm, p + "bn1.", k.cin, &k.s1, &k.h1);
      ok = ok && !make_bn(e, tm, p + "bn2.", k.cout, &k.s2, &k.h2);
      ok = ok && (k.w1 = make_conv(e, tm, p + "conv1.weight", k.cout, k.cin, 3)) != nullptr;
      ok = ok && (k.w2 = make_conv(e, tm, p + "conv2.weight", k.cout, k.cout, 3)) != nullptr;
      if (k.cin != k.cout) ok = ok && (k.wsc = make_conv(e, tm, p + "shortcut.0.weight", k.cout, k.cin, 1)) != nullptr;
      const float *al = find(tm, p + "afms.alpha"), *fw = find(tm, p + "afms.fc.weight"), *fb = find(tm, p + "afms.fc.bias");
      ok = ok && al && fw && fb;
      if (ok) {
        k.alpha = dcopy(e, al, k.cout); k.fc_w = dcopy(e, fw, static_cast<size_t>(k.cout) * k.cout);
        k.fc_b = dcopy(e, fb, k.cout);
      }
    }
    if (!ok) break;
    if (aggregate != AGG_GRU) {
      if (make_bn(e, tm, "bn_before_agg.", 512, &e->gbn_s, &e->gbn_h)) break;
      const float *fw = find(tm, "fc.weight"), *fb = find(tm, "fc.bias");
      if (!fw || !fb) break;
      e->ag_fc_w = dcopy(e, fw, static_cast<size_t>(n_out) * 1024); e->ag_fc_b = dcopy(e, fb, n_out);
      if (aggregate == AGG_SAP) {
        const float *lw = find(tm, "attention_sap.lin_proj.weight"), *lb = find(tm, "attention_sap.lin_proj.bias");
        const float* v = find(tm, "attention_sap.v");
        if (!lw || !lb || !v) break;
        e->ag_w1t = dtranspose(e, lw, 512, 512); e->ag_b1 = dcopy(e, lb, 512); e->ag_v = dcopy(e, v, 512);
      } else {
        const float *w1 = find(tm, "attention.0.weight"), *b1 = find(tm, "attention.0.bias");
        const float *w2 = find(tm, "attention.3.weight"), *b2 = find(tm, "attention.3.bias");
        if (!w1 || !b1 || !w2 || !b2) break;
        if (make_bn(e, tm, "attention.2.", att_dim, &e->ag_bn_s, &e->ag_bn_h)) break;
        e->ag_w1t = dtranspose(e, w1, att_dim, 512); e->ag_b1 = dcopy(e, b1, att_dim);      // Conv1d weight [att, 512, 1]
        e->ag_w2t = dtranspose(e, w2, 512, att_dim); e->ag_b2 = dcopy(e, b2, 512);          // Conv1d weight [512, att, 1]
      }
      if (!e->ag_fc_w || !e->ag_fc_b || !e->ag_w1t || !e->ag_b1) { set_error("out of device memory"); break; }
      if (e->T[NB] > AGG_TMAX) { set_error("the attentive heads handle up to %d frames (got %d)", AGG_TMAX, e->T[NB]); break; }
      if (cudaDeviceSynchronize() != cudaSuccess) { set_error("RawNet2 weight preparation failed: %s", cudaGetErrorString(cudaGetLastError())); break; }
      rc = 0;
      break;
    }
    if (make_bn(e, tm, "bn_before_gru.", 512, &e->gbn_s, &e->gbn_h)) break;
    const float *wih = find(tm, "gru.weight_ih_l0"), *whh = find(tm, "gru.weight_hh_l0");
    const float *bih = find(tm, "gru.bias_ih_l0"), *bhh = find(tm, "gru.bias_hh_l0");
    const float *fw = find(tm, "fc_after_gru.weight"), *fb = find(tm, "fc_after_gru.bias");
    if (!wih || !whh || !bih || !bhh || !fw || !fb) break;
    const int H = e->H;
    e->w_ih = dalloc<bf16>(e, static_cast<size_t>(3 * H) * e->planes * 512);
    e->w_hh = dalloc<bf16>(e, static_cast<size_t>(3 * H) * 2 * H);
    if (!e->w_ih || !e->w_hh) { set_error("out of device memory"); break; }
    if (pack_conv_weight(wih, 3 * H, 512, 1, 512, e->planes, e->w_ih, 0, e->planes == 1 ? 1 : 0)) break;
    if (pack_conv_weight(whh, 3 * H, H, 1, H, 2, e->w_hh, 0)) break;     // the recurrence is always fp32 grade
    e->b_ih = dcopy(e, bih, 3 * H); e->b_hh = dcopy(e, bhh, 3 * H);
    e->fcg_w = dcopy(e, fw, static_cast<size_t>(n_out) * H); e->fcg_b = dcopy(e, fb, n_out);
    e->ones = dalloc<float>(e, 3 * H); e->zeros = dalloc<float>(e, 3 * H);
    rn_fill_kernel<<<(3 * H + 255) / 256, 256>>>(e->ones, 3 * H, 1.f);
    rn_fill_kernel<<<(3 * H + 255) / 256, 256>>>(e->zeros, 3 * H, 0.f);
    if (cudaDeviceSynchronize() != cudaSuccess) { set_error("RawNet2 weight preparation failed: %s", cudaGetErrorString(cudaGetLastError())); break; }
    rc = 0;
  } while (false);
  if (rc) {
    for (void* p : e->owned) cudaFree(p);
    delete e;
    return rc;
  }
  *out = e;
  return 0;
}

void rawnet2_destroy(Rawnet2* e) {
  if (e == nullptr) return;
  for (void* p : e->owned) cudaFree(p);
  delete e;
}

// ------------------------------------------------------------------ workspace
static size_t al256(size_t x) { return (x + 255) / 256 * 256; }

struct RnPlan {
  size_t xn, x[2], a[2], hbuf, z, sc, zp, psum, gate, gi, gh, hstate, h16, total;
  size_t act_bytes_max;
};

static size_t rows_bytes(int planes, int B, int T, int C) {
  return al256(static_cast<size_t>(B) * (T + 2) * C * planes * 2);
}
static int pool_chunks(int Tout) { return (Tout + 255) / 256; }

static RnPlan make_plan(const Rawnet2* e, int B) {
  RnPlan p{};
  size_t o = 0;
  auto take = [&](size_t n) { size_t r = o; o += al256(n); return r; };
  p.xn = take((static_cast<size_t>(B) * e->Lp * 2 + 4096) * 2);   // the signal is always split hi|lo
  // ping-pong x / a: index 0 holds block-even inputs.  Largest: front output (T1 x 128)
  // operands (a, h): e->planes planes; residual carriers (x, z, sc, zp): e->xplanes
  const int P = e->planes, PX = e->xplanes;
  size_t hmax = 0, zmax = 0, midx = 0, mida = 0, sc = 0, zp = 0;
  for (int i = 0; i < NB; ++i) {
    hmax = std::max(hmax, rows_bytes(P, B, e->T[i], e->blk[i].cout));
    zmax = std::max(zmax, rows_bytes(PX, B, e->T[i], e->blk[i].cout));
    if (i > 0) midx = std::max(midx, rows_bytes(PX, B, e->T[i], e->blk[i].cin));
    if (i > 0) mida = std::max(mida, rows_bytes(P, B, e->T[i], e->blk[i].cin));
    if (e->blk[i].cin != e->blk[i].cout) sc = std::max(sc, rows_bytes(PX, B, e->T[i], e->blk[i].cout));
    if (e->blk[i].down) zp = std::max(zp, rows_bytes(PX, B, e->T[i + 1], e->blk[i].cout));
  }
  mida = std::max(mida, rows_bytes(P, B, e->T[NB], 512));
  p.x[0] = take(rows_bytes(PX, B, e->T[0], 128)); p.a[0] = take(rows_bytes(P, B, e->T[0], 128));
  p.x[1] = take(midx); p.a[1] = take(mida);
  p.hbuf = take(hmax); p.z = take(zmax);
  p.sc = take(sc); p.zp = take(zp);
  p.psum = take(static_cast<size_t>(B) * pool_chunks(e->T[1]) * 512 * 4);
  p.gate = take(static_cast<size_t>(B) * 512 * 4);
  p.gi = take(static_cast<size_t>(B) * (e->T[NB] + 2) * 3 * e->H * 4);
  const size_t Bp = (static_cast<size_t>(B) + 127) / 128 * 128;
  p.gh = take(Bp * 3 * e->H * 4);
  p.hstate = take(static_cast<size_t>(B) * e->H * 4);
  p.h16 = take(Bp * 2 * e->H * 2);
  p.total = o;
  return p;
}

size_t rawnet2_workspace_bytes(const Rawnet2* e, int B) { return make_plan(e, B).total; }

// byte offset of a named intermediate inside the workspace (tests): "front" = a/x entering layer1, "x<i>" = the
// block-i input x (i = 1..8 -> output of block i-1), "gru" = final hidden state (fp32 [B, H])
long long rawnet2_ws_offset(const Rawnet2* e, int B, const char* name) {
  const RnPlan p = make_plan(e, B);
  if (!strcmp(name, "xn")) return static_cast<long long>(p.xn);
  if (!strcmp(name, "x0")) return static_cast<long long>(p.x[0]);
  if (!strcmp(name, "a0")) return static_cast<long long>(p.a[0]);
  if (!strcmp(name, "x1")) return static_cast<long long>(p.x[1]);
  if (!strcmp(name, "a1")) return static_cast<long long>(p.a[1]);
  if (!strcmp(name, "gru")) return static_cast<long long>(p.hstate);
  if (!strcmp(name, "gi")) return static_cast<long long>(p.gi);
  return -1;
}
int rawnet2_frames(const Rawnet2* e, int level) { return (level >= 0 && level <= NB) ? e->T[level] : -1; }

// ------------------------------------------------------------------ forward
static int conv_gemm(const Rawnet2* e, int epi, const bf16* a, int B, int T, int cin, const bf16* w, int cout, int k,
                     const float* s, const float* h, const bf16* add, bf16* out, int out_planes, cudaStream_t st,
                     int pool_tout = 0) {
  RnGemmParams g{};
  const int P = e->planes;   // operand planes (a / h and the packed weights)
  g.M = B * (T + 2); g.N = cout; g.kc_per_tap = cin / 64; g.num_taps = k;
  g.tap_off[0] = k == 3 ? -1 : 0; g.tap_off[1] = 0; g.tap_off[2] = 1;
  g.num_pass = P == 2 ? 3 : 1; g.a_plane_cols = cin; g.w_planes = P;
  g.T = T; g.Tp = T + 2; g.s1 = s; g.h1 = h; g.B = B;
  if (pool_tout > 0) { g.pool3 = 1; g.Tout = pool_tout; }   // `out` is the pooled tensor [B][pool_tout + 2][cout]
  g.add = add; g.add_ld = e->xplanes * cout; g.add_planes = e->xplanes; g.add_plane_cols = cout;
  g.out = out; g.out_ld = out_planes * cout; g.out_planes = out_planes; g.out_plane_cols = cout;
  return launch_rn_gemm(epi, a, g.M, static_cast<long long>(P) * cin, static_cast<long long>(P) * cin, w, cout,
                        k * P * cin, g, st);
}

int rawnet2_forward(const Rawnet2* e, const float* wav, long long wav_ld, int B, void* ws, size_t ws_bytes, float* emb,
                    int emb_ld, int l2norm, cudaStream_t st) {
  const RnPlan pl = make_plan(e, B);
  SVB_CHECK(ws != nullptr && ws_bytes >= pl.total, "RawNet2 workspace too small: need %zu bytes, got %zu", pl.total, ws_bytes);
  SVB_CHECK((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "workspace must be 256-byte aligned");
  uint8_t* base = static_cast<uint8_t*>(ws);
  const int P = e->planes, PX = e->xplanes, H = e->H;
  bf16* xn = reinterpret_cast<bf16*>(base + pl.xn);
  bf16* xb[2] = {reinterpret_cast<bf16*>(base + pl.x[0]), reinterpret_cast<bf16*>(base + pl.x[1])};
  bf16* ab[2] = {reinterpret_cast<bf16*>(base + pl.a[0]), reinterpret_cast<bf16*>(base + pl.a[1])};
  bf16* hb = reinterpret_cast<bf16*>(base + pl.hbuf);
  bf16* zb = reinterpret_cast<bf16*>(base + pl.z);
  bf16* scb = reinterpret_cast<bf16*>(base + pl.sc);
  bf16* zpb = reinterpret_cast<bf16*>(base + pl.zp);
  float* psum = reinterpret_cast<float*>(base + pl.psum);
  float* gate = reinterpret_cast<float*>(base + pl.gate);
  float* gi = reinterpret_cast<float*>(base + pl.gi);
  float* gh = reinterpret_cast<float*>(base + pl.gh);
  float* hs = reinterpret_cast<float*>(base + pl.hstate);
  bf16* h16 = reinterpret_cast<bf16*>(base + pl.h16);

  // ---- front: LayerNorm -> sinc GEMM (+abs, maxpool3, first_bn, lrelu, layer1.bn1, lrelu)
  const long long plane_dist = (static_cast<long long>(B) * e->Lp + 63) / 64 * 64;
  // The normalised signal is written as bf16 hi|lo planes (fp32-grade mode, three split passes) or as one plane of
  // IEEE half (16-bit mode, one pass).
  SVB_CUDA_OK(cudaMemsetAsync(xn, 0, (static_cast<size_t>(B) * e->Lp * 2 + 4096) * 2, st));
  rn_ln_kernel<<<B, 1024, 0, st>>>(wav, wav_ld, e->L, e->Lp, e->ln_g, e->ln_b, 1e-6f, xn, plane_dist, P);
  count_launch();
  // the first block's padding rows must be zero: the sinc epilogue writes valid rows only
  rn_zero_halo_kernel<<<2 * B, 64, 0, st>>>(ab[0], P * 128, e->T[0], B);
  count_launch();
  {
    RnGemmParams g{};
    g.M = B * e->Mu; g.N = 16 * 128 * 3; g.kc_per_tap = SINC_K / 64; g.num_taps = 1; g.tap_off[0] = 0;
    g.num_pass = P == 2 ? 3 : 1; g.a_plane_cols = static_cast<int>(plane_dist); g.w_planes = P;
    g.T = e->T[0]; g.Tp = e->T[0] + 2; g.Mu = e->Mu; g.B = B;
    g.s1 = e->fbn_s; g.h1 = e->fbn_h; g.s2 = e->blk[0].s1; g.h2 = e->blk[0].h1;
    g.out = xb[0]; g.out_ld = PX * 128; g.out_planes = PX; g.out_plane_cols = 128;
    g.out2 = ab[0]; g.out2_ld = P * 128; g.out2_planes = P; g.out2_plane_cols = 128;
    SVB_CHECK(plane_dist < (1LL << 31) - 1024, "batch too large for the sinc operand view");
    const long long a_cols = plane_dist + SINC_K + 64;
    if (launch_rn_gemm(EPI_SINC, xn, static_cast<long long>(B) * e->Mu, a_cols, SINC_STRIDE, e->sinc_w, g.N,
                       P * SINC_K, g, st)) return -1;
  }
  // debugging aid for the stage tests: SVB_RAWNET_STOP=<i> returns after residual block i (-1: after the front)
  const char* stop_env = getenv("SVB_RAWNET_STOP");
  const int stop_after = stop_env ? atoi(stop_env) : 1000;
  if (stop_after < 0) return 0;
  // ---- residual blocks
  int cur = 0;
  for (int i = 0; i < NB; ++i) {
    const RnBlock& k = e->blk[i];
    const int Tin = e->T[i], Tout = e->T[i + 1];
    const bf16* a_in = ab[cur];
    const bf16* x_in = xb[cur];
    if (conv_gemm(e, EPI_AFF_LRELU, a_in, B, Tin, k.cin, k.w1, k.cout, 3, k.s2, k.h2, nullptr, hb, P, st)) return -1;
    const bf16* shortcut = x_in;
    if (k.wsc != nullptr) {
      if (conv_gemm(e, EPI_PLAIN, a_in, B, Tin, k.cin, k.wsc, k.cout, 1, nullptr, nullptr, nullptr, scb, PX, st)) return -1;
      shortcut = scb;
    }
    const int pool = k.down ? 3 : 1;
    const int nchunk = pool_chunks(Tout);
    const int cg = k.cout / 8, tstep = 256 / cg;
    if (k.down && k.wsc == nullptr && rn_conv2_pool_supported(k.cout, P, PX, Tin)) {
      // layers 1 and 2 in the 16-bit mode: conv2 pools its own output (z, 5.4 / 1.8 MB per utterance, is never written) and a
      // sums-only pass over the pooled tensor (a third of z) feeds the AFMS gate
      if (conv_gemm(e, EPI_ADD, hb, B, Tin, k.cout, k.w2, k.cout, 3, nullptr, nullptr, shortcut, zpb, PX, st, Tout)) return -1;
      rn_pool_kernel<<<dim3(nchunk, B), 256, static_cast<size_t>(tstep) * k.cout * 4, st>>>(
          zpb, PX * k.cout, PX, k.cout, Tout, 1, nullptr, PX * k.cout, k.cout, Tout, k.cout, 256, psum);
      count_launch();
    } else {
      if (conv_gemm(e, EPI_ADD, hb, B, Tin, k.cout, k.w2, k.cout, 3, nullptr, nullptr, shortcut, zb, PX, st)) return -1;
      rn_pool_kernel<<<dim3(nchunk, B), 256, static_cast<size_t>(tstep) * k.cout * 4, st>>>(
          zb, PX * k.cout, PX, k.cout, Tin, pool, zpb, PX * k.cout, k.cout, Tout, k.cout, 256, psum);
      count_launch();
    }
    rn_fc_kernel<<<dim3((k.cout + 31) / 32, (B + 3) / 4), 256, 4 * k.cout * 4, st>>>(
        psum, nchunk, 1.0f / static_cast<float>(Tout), k.cout, k.fc_w, k.fc_b, k.cout, B, 1, gate, k.cout);
    count_launch();
    const int nxt = cur ^ 1;
    const bool last = i == NB - 1;
    const float* ns = last ? e->gbn_s : e->blk[i + 1].s1;
    const float* nh = last ? e->gbn_h : e->blk[i + 1].h1;
    const bool need_x = !last && e->blk[i + 1].wsc == nullptr;
    const bf16* src = k.down ? zpb : zb;
    const int rows_per_cta = 128;
    const int PA = P;
    rn_apply_kernel<<<dim3((Tout + 2 + rows_per_cta - 1) / rows_per_cta, B), 256, 0, st>>>(
        src, PX * k.cout, PX, k.cout, Tout, k.cout, k.alpha, gate, ns, nh, need_x ? xb[nxt] : nullptr, PX * k.cout, k.cout,
        ab[nxt], PA * k.cout, PA, k.cout, rows_per_cta);
    count_launch();
    SVB_CUDA_OK(cudaGetLastError());
    cur = nxt;
    if (stop_after == i) return 0;
  }
  const int Tg = e->T[NB];
  if (e->aggregate != AGG_GRU) {
    // ---- attentive heads (RawNet2_custom.py:200-224): one CTA per utterance on the 29 pre-activated frames
    const int att = e->aggregate == AGG_ASP ? e->att : 1;
    const size_t smem = (static_cast<size_t>(Tg) * AGG_C + static_cast<size_t>(Tg) * att + 32 + 4096) * sizeof(float);
    static unsigned long long attr_devs = 0;   // per device: kernel attributes do not carry over
    if (first_on_device(&attr_devs)) {
      SVB_CUDA_OK(cudaFuncSetAttribute(rn_agg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    }
    SVB_CHECK(smem <= 200 * 1024, "attentive head: %zu bytes of shared memory", smem);
    float* dst = l2norm ? hs : emb;                       // hs ([B, H] fp32, H >= n_out rows' worth) as the un-normalised buffer
    const int dst_ld = l2norm ? e->n_out : emb_ld;
    SVB_CHECK(!l2norm || e->n_out <= H, "nOut above gru_node with l2norm");
    rn_agg_kernel<<<B, 256, smem, st>>>(ab[cur], P * 512, P, Tg, Tg + 2, e->aggregate, e->att, e->ag_w1t, e->ag_b1, e->ag_bn_s,
                                        e->ag_bn_h, e->ag_w2t, e->ag_b2, e->ag_v, e->ag_fc_w, e->ag_fc_b, e->n_out, dst, dst_ld);
    count_launch();
    SVB_CUDA_OK(cudaGetLastError());
    if (l2norm) return l2norm_rows(dst, dst_ld, emb, emb_ld, B, e->n_out, 1e-12f, st);
    return 0;
  }
  // ---- GRU (RawNet2_custom.py:192-198): input projection for all steps, then T sequential steps
  {
    GemmOperands ops{};
    ops.a = ab[cur]; ops.lda = P * 512; ops.a_cols = P * 512;
    ops.w = e->w_ih; ops.w_rows = 3 * H; ops.w_cols = P * 512;
    ops.a_f16 = ops.w_f16 = (P == 1) ? 1 : 0;
    GemmParams g{};
    g.M = B * (Tg + 2); g.N = 3 * H; g.kc_per_tap = 512 / 64; g.num_taps = 1; g.tap_off[0] = 0;
    g.num_pass = P == 2 ? 3 : 1; g.a_plane_cols = 512; g.w_planes = P;
    g.bias = e->b_ih; g.scale = e->ones; g.shift = e->zeros; g.act = ACT_NONE; g.store_mode = STORE_ALL;
    g.out = gi; g.out_ld = 3 * H; g.out_kind = OUT_F32;
    if (launch_gemm(ops, g, 256, st)) return -1;
  }
  SVB_CUDA_OK(cudaMemsetAsync(hs, 0, static_cast<size_t>(B) * H * 4, st));
  SVB_CUDA_OK(cudaMemsetAsync(h16, 0, (static_cast<size_t>(B) + 127) / 128 * 128 * 2 * H * 2, st));
  for (int t = 0; t < Tg; ++t) {
    if (t > 0) {
      GemmOperands ops{};
      ops.a = h16; ops.lda = 2 * H; ops.a_cols = 2 * H;
      ops.w = e->w_hh; ops.w_rows = 3 * H; ops.w_cols = 2 * H;
      GemmParams g{};
      g.M = B; g.N = 3 * H; g.kc_per_tap = H / 64; g.num_taps = 1; g.tap_off[0] = 0; g.num_pass = 3;
      g.a_plane_cols = H; g.w_planes = 2; g.out = gh; g.out_ld = 3 * H;
      if (launch_gemm_f32(ops, g, st)) return -1;
    }
    rn_gru_gate_kernel<<<(B * H + 255) / 256, 256, 0, st>>>(gi + static_cast<long long>(1 + t) * 3 * H, 0,
                                                             static_cast<long long>(Tg + 2) * 3 * H, t > 0 ? gh : nullptr,
                                                             3 * H, e->b_hh, H, B, hs, h16, 2 * H);
    count_launch();
  }
  // ---- fc_after_gru (+ optional F.normalize, src/model.py:422-423)
  rn_fc_kernel<<<dim3((e->n_out + 31) / 32, (B + 3) / 4), 256, 4 * H * 4, st>>>(hs, 1, 1.0f, H, e->fcg_w, e->fcg_b,
                                                                                  e->n_out, B, 0, emb, emb_ld);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  if (l2norm) return l2norm_rows(emb, emb_ld, emb, emb_ld, B, e->n_out, 1e-12f, st);
  return 0;
}

}  // namespace svb
