// Trial scoring: cosine and adaptive S-norm (AS-norm) with a cohort bank.
//
// Reference call sites replaced (paths in the reference tree):
//   cosine_similarity    src/utils.py:163-164   mean_i |cos(ref_i, com_i)| over index-paired crops
//   ZT_norm_similarity   src/utils.py:135-160   per trial: 2 x (cohort GEMM, full sort, mean/std of top-K)
//   per-trial Python loop src/model.py:415-448
// AS-norm is linear in the crop rows, so the cohort statistics are computed ONCE per unique
// utterance from its mean embedding: S = E_mean * cohort^T on tensor cores
// (split-bf16, fp32 grade), exact top-K mean / population-std by radix select, then one gather
// kernel over all trials.
#include <math.h>

#include <algorithm>

#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

int split_bf16(const float*, long long, int, int, bf16*, int, int, int, cudaStream_t);
int pack_conv_weight(const float*, int, int, int, int, int, bf16*, cudaStream_t);

// ---------------------------------------------------------------- cosine trials
// emb [U, n, D] fp32; one warp per trial.
__global__ void cosine_trials_kernel(const float* __restrict__ emb, int n, int D, const int32_t* __restrict__ ia,
                                     const int32_t* __restrict__ ib, long long ntrials, float eps,
                                     float* __restrict__ out) {
  const long long tr = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (tr >= ntrials) return;
  const float* a = emb + static_cast<long long>(ia[tr]) * n * D;
  const float* b = emb + static_cast<long long>(ib[tr]) * n * D;
  float acc = 0.f;
  for (int i = 0; i < n; ++i) {
    float ab = 0.f, aa = 0.f, bb = 0.f;
    for (int c = lane; c < D; c += 32) {
      const float x = a[i * D + c], y = b[i * D + c];
      ab = fmaf(x, y, ab); aa = fmaf(x, x, aa); bb = fmaf(y, y, bb);
    }
    ab = warp_sum(ab); aa = warp_sum(aa); bb = warp_sum(bb);
    acc += fabsf(ab / (fmaxf(sqrtf(aa), eps) * fmaxf(sqrtf(bb), eps)));
  }
  if (lane == 0) out[tr] = acc / static_cast<float>(n);
}

// mean over the n crop rows of each utterance, optionally L2-normalising every row first
// (test_normalize, src/model.py:421-423).  One warp per utterance.
__global__ void mean_rows_kernel(const float* __restrict__ emb, int U, int n, int D, int normalize,
                                 float* __restrict__ out) {
  const int u = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (u >= U) return;
  for (int c = lane; c < D; c += 32) out[static_cast<long long>(u) * D + c] = 0.f;
  for (int i = 0; i < n; ++i) {
    const float* r = emb + (static_cast<long long>(u) * n + i) * D;
    float inv = 1.f;
    if (normalize) {
      float s = 0.f;
      for (int c = lane; c < D; c += 32) s = fmaf(r[c], r[c], s);
      s = warp_sum(s);
      inv = 1.f / fmaxf(sqrtf(s), 1e-12f);
    }
    for (int c = lane; c < D; c += 32) out[static_cast<long long>(u) * D + c] += r[c] * inv;
  }
  for (int c = lane; c < D; c += 32) out[static_cast<long long>(u) * D + c] /= static_cast<float>(n);
}

// ---------------------------------------------------------------- exact top-K mean / std per row
__device__ __forceinline__ uint32_t f2key(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);   // monotone: larger float <-> larger key
}
__device__ __forceinline__ float key2f(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// One CTA per row of S [rows, ld]; selects the K largest of the first N entries by a 3-pass
// (11 + 11 + 10 bit) radix select, then accumulates sum and sum of squares in fp64.
__global__ void __launch_bounds__(256)
topk_stats_kernel(const float* __restrict__ S, int ld, int N, int K, float* __restrict__ mu,
                  float* __restrict__ sigma) {
  __shared__ uint32_t hist[2048];
  __shared__ uint32_t sel_prefix, sel_remaining;
  __shared__ double red_s[8], red_q[8];
  const float* row = S + static_cast<long long>(blockIdx.x) * ld;
  const int tid = threadIdx.x;
  uint32_t prefix = 0, prefix_mask = 0, remaining = K;
  const int shifts[3] = {21, 10, 0};
  const int bits[3] = {11, 11, 10};
  for (int pass = 0; pass < 3; ++pass) {
    const int nb = 1 << bits[pass];
    for (int i = tid; i < nb; i += 256) hist[i] = 0;
    __syncthreads();
    for (int i = tid; i < N; i += 256) {
      const uint32_t k = f2key(row[i]);
      if ((k & prefix_mask) == prefix) atomicAdd(&hist[(k >> shifts[pass]) & (nb - 1)], 1u);
    }
    __syncthreads();
    if (tid < 32) {
      // warp 0: walk bins from the top; each lane owns a contiguous slice
      const int per = nb / 32;
      uint32_t mine = 0;
      for (int j = 0; j < per; ++j) mine += hist[nb - 1 - (tid * per + j)];
      uint32_t incl = mine;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (tid >= o) incl += v;
      }
      const uint32_t excl = incl - mine;
      if (excl < remaining && incl >= remaining) {
        uint32_t run = excl;
        for (int j = 0; j < per; ++j) {
          const int bin = nb - 1 - (tid * per + j);
          const uint32_t c = hist[bin];
          if (run + c >= remaining) {
            sel_prefix = prefix | (static_cast<uint32_t>(bin) << shifts[pass]);
            sel_remaining = remaining - run;   // how many still to take inside this bin
            break;
          }
          run += c;
        }
      }
    }
    __syncthreads();
    prefix = sel_prefix;
    remaining = sel_remaining;
    prefix_mask |= static_cast<uint32_t>(nb - 1) << shifts[pass];
    __syncthreads();
  }
  // prefix == key of the K-th largest value; `remaining` copies of it belong to the top-K set
  const uint32_t kth = prefix;
  double s = 0.0, q = 0.0;
  for (int i = tid; i < N; i += 256) {
    const float v = row[i];
    if (f2key(v) > kth) { s += v; q += static_cast<double>(v) * v; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if ((tid & 31) == 0) { red_s[tid >> 5] = s; red_q[tid >> 5] = q; }
  __syncthreads();
  if (tid == 0) {
    double ts = 0.0, tq = 0.0;
    for (int w = 0; w < 8; ++w) { ts += red_s[w]; tq += red_q[w]; }
    const double kv = key2f(kth);
    ts += remaining * kv;
    tq += remaining * kv * kv;
    const double m = ts / K;
    const double var = fmax(tq / K - m * m, 0.0);
    mu[blockIdx.x] = static_cast<float>(m);
    sigma[blockIdx.x] = static_cast<float>(sqrt(var));
  }
}

// ---------------------------------------------------------------- AS-norm trial scores
// emean [U, D]; out = 0.5 * ((s - mu_a)/sigma_a + (s - mu_b)/sigma_b), s = <emean_a, emean_b>.
__global__ void asnorm_trials_kernel(const float* __restrict__ emean, int D, const int32_t* __restrict__ ia,
                                     const int32_t* __restrict__ ib, long long ntrials, const float* __restrict__ mu,
                                     const float* __restrict__ sigma, float* __restrict__ out) {
  const long long tr = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (tr >= ntrials) return;
  const int a = ia[tr], b = ib[tr];
  float s = 0.f;
  for (int c = lane; c < D; c += 32) s = fmaf(emean[static_cast<long long>(a) * D + c], emean[static_cast<long long>(b) * D + c], s);
  s = warp_sum(s);
  if (lane == 0) out[tr] = 0.5f * ((s - mu[a]) / sigma[a] + (s - mu[b]) / sigma[b]);
}

__global__ void fill_kernel(float* p, int n, float v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// ---------------------------------------------------------------- launchers
int cosine_trials(const float* emb, int n, int D, const int32_t* ia, const int32_t* ib, long long ntrials, float* out,
                  cudaStream_t stream) {
  if (ntrials == 0) return 0;
  cosine_trials_kernel<<<static_cast<int>((ntrials + 7) / 8), 256, 0, stream>>>(emb, n, D, ia, ib, ntrials, 1e-5f, out);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

int mean_rows(const float* emb, int U, int n, int D, int normalize, float* out, cudaStream_t stream) {
  mean_rows_kernel<<<(U + 7) / 8, 256, 0, stream>>>(emb, U, n, D, normalize, out);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

static size_t align256(size_t x) { return (x + 255) / 256 * 256; }

size_t asnorm_workspace_bytes(int U, int N, int D, int* chunk_out) {
  const int Dp = (D + 63) / 64 * 64, Np = (N + 31) / 32 * 32;
  const size_t fixed = align256(static_cast<size_t>(N) * 2 * Dp * 2) + 3 * align256(static_cast<size_t>(Np) * 4);
  long long chunk = std::min<long long>(U, std::max<long long>(128, (1LL << 30) / (static_cast<long long>(Np) * 4)));
  chunk = (chunk + 127) / 128 * 128;
  if (chunk_out) *chunk_out = static_cast<int>(chunk);
  return fixed + align256(static_cast<size_t>(chunk) * 2 * Dp * 2) + align256(static_cast<size_t>(chunk) * Np * 4);
}

// mu/sigma of the top-K cohort scores of every row of emean [U, D].
int asnorm_stats(const float* emean, int U, const float* cohort, int N, int D, int topk, float* mu, float* sigma,
                 void* ws, size_t ws_bytes, cudaStream_t stream) {
  int K = topk;
  if (K < 0) K = N + K;            // numpy slice semantics of S[::-1][:top] (top=-1 drops the smallest)
  if (K > N) K = N;
  SVB_CHECK(K >= 1, "top-K must select at least one cohort score (N=%d, top=%d)", N, topk);
  int chunk = 0;
  const size_t need = asnorm_workspace_bytes(U, N, D, &chunk);
  SVB_CHECK(ws != nullptr && ws_bytes >= need, "AS-norm workspace too small: need %zu bytes, got %zu", need, ws_bytes);
  const int Dp = (D + 63) / 64 * 64, Np = (N + 31) / 32 * 32;
  uint8_t* p = static_cast<uint8_t*>(ws);
  bf16* wpk = reinterpret_cast<bf16*>(p); p += align256(static_cast<size_t>(N) * 2 * Dp * 2);
  float* zeros = reinterpret_cast<float*>(p); p += align256(static_cast<size_t>(Np) * 4);
  float* ones = reinterpret_cast<float*>(p); p += align256(static_cast<size_t>(Np) * 4);
  p += align256(static_cast<size_t>(Np) * 4);
  bf16* a16 = reinterpret_cast<bf16*>(p); p += align256(static_cast<size_t>(chunk) * 2 * Dp * 2);
  float* S = reinterpret_cast<float*>(p);
  // cohort [N, D] is a k=1 "conv weight" [cout=N, cin=D, 1]
  if (pack_conv_weight(cohort, N, D, 1, Dp, 2, wpk, stream)) return -1;
  fill_kernel<<<(Np + 255) / 256, 256, 0, stream>>>(zeros, Np, 0.f);
  fill_kernel<<<(Np + 255) / 256, 256, 0, stream>>>(ones, Np, 1.f);
  count_launch(2);
  for (int u0 = 0; u0 < U; u0 += chunk) {
    const int rows = std::min(chunk, U - u0);
    SVB_CUDA_OK(cudaMemsetAsync(a16, 0, static_cast<size_t>(rows) * 2 * Dp * 2, stream));
    if (split_bf16(emean + static_cast<long long>(u0) * D, rows, D, D, a16, 2 * Dp, 2, Dp, stream)) return -1;
    GemmOperands ops{};
    ops.a = a16; ops.lda = 2 * Dp; ops.a_cols = 2 * Dp;
    ops.w = wpk; ops.w_rows = N; ops.w_cols = 2 * Dp;
    GemmParams g{};
    g.M = rows; g.N = Np; g.kc_per_tap = Dp / 64; g.num_taps = 1; g.tap_off[0] = 0; g.num_pass = 3;
    g.a_plane_cols = Dp; g.w_planes = 2; g.bias = zeros; g.scale = ones; g.shift = zeros;
    g.act = ACT_NONE; g.store_mode = STORE_ALL; g.out = S; g.out_ld = Np; g.out_kind = OUT_F32;
    if (launch_gemm(ops, g, Np >= 256 ? 256 : (Np >= 128 ? 128 : 64), stream)) return -1;
    topk_stats_kernel<<<rows, 256, 0, stream>>>(S, Np, N, K, mu + u0, sigma + u0);
    count_launch();
    SVB_CUDA_OK(cudaGetLastError());
  }
  return 0;
}

int asnorm_trials(const float* emean, int D, const int32_t* ia, const int32_t* ib, long long ntrials, const float* mu,
                  const float* sigma, float* out, cudaStream_t stream) {
  if (ntrials == 0) return 0;
  asnorm_trials_kernel<<<static_cast<int>((ntrials + 7) / 8), 256, 0, stream>>>(emean, D, ia, ib, ntrials, mu, sigma, out);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace svb
