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

#include <stdlib.h>

#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

int split_bf16(const float*, long long, int, int, bf16*, int, int, int, cudaStream_t);

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

// pnorm trials (utils.pnorm_similarity, src/utils.py:167-169): mean_i || a_i - b_i + eps ||_p over index-paired
// crops (F.pairwise_distance adds eps to the DIFFERENCE), rows optionally L2-normalised first (test_normalize,
// src/model.py:421-423).  pmode 0: general p (powf), 1: p = 1, 2: p = 2, 3: p = inf.  One warp per trial.
__global__ void pnorm_trials_kernel(const float* __restrict__ emb, int n, int D, const int32_t* __restrict__ ia,
                                    const int32_t* __restrict__ ib, long long ntrials, float pw, int pmode, float eps,
                                    int normalize, float* __restrict__ out) {
  const long long tr = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (tr >= ntrials) return;
  const float* a = emb + static_cast<long long>(ia[tr]) * n * D;
  const float* b = emb + static_cast<long long>(ib[tr]) * n * D;
  float acc = 0.f;
  for (int i = 0; i < n; ++i) {
    float ia_ = 1.f, ib_ = 1.f;
    if (normalize) {
      float aa = 0.f, bb = 0.f;
      for (int c = lane; c < D; c += 32) { const float x = a[i * D + c], y = b[i * D + c]; aa = fmaf(x, x, aa); bb = fmaf(y, y, bb); }
      aa = warp_sum(aa); bb = warp_sum(bb);
      ia_ = 1.f / fmaxf(sqrtf(aa), 1e-12f); ib_ = 1.f / fmaxf(sqrtf(bb), 1e-12f);
    }
    float s = 0.f;
    for (int c = lane; c < D; c += 32) {
      const float d = fabsf(a[i * D + c] * ia_ - b[i * D + c] * ib_ + eps);
      if (pmode == 2) s = fmaf(d, d, s);
      else if (pmode == 1) s += d;
      else if (pmode == 3) s = fmaxf(s, d);
      else s += powf(d, pw);
    }
    s = (pmode == 3) ? warp_max(s) : warp_sum(s);
    acc += (pmode == 2) ? sqrtf(s) : ((pmode == 1 || pmode == 3) ? s : powf(s, 1.f / pw));
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

// Row cached in shared memory, value-linear first level.  S is read ONCE (float4; from L2 when the GEMM chunk
// that produced it is still resident).  Cohort scores of one utterance cluster in a handful of float exponents,
// so a radix pass on the key's top bits would serialise thousands of shared-memory atomics on a few bins;
// instead level 0 histograms the values LINEARLY over the row's [min, max] (2048 bins, conflicts rare), which
// leaves only the few candidates of the threshold bin for the exact key-radix select.
//   1. load row -> smem, block min/max            2. linear histogram, find the bin holding the K-th largest
//   3. sweep: bins above -> fp64 sum / sum-sq, threshold bin -> candidate list (cap TOPK_CAND)
//   4. exact 3-pass key radix among the candidates (or, on overflow of the list, over the whole row with the
//      bin as a filter) -> K-th key and the number of copies of it that belong to the top-K set.
// Dynamic smem: N floats.
#define TOPK_CAND 2048
__device__ __forceinline__ int lin_bin(float v, float lo, float scale) {
  return min(2047, __float2int_rz((v - lo) * scale));   // saturating; -inf maps below bin 0
}

// All NT threads: walks `nb` (multiple of NT) histogram bins from the top and finds the bin in which the
// cumulative count reaches `remaining`.  Writes sel[0] = prefix | bin << shift, sel[1] = how many elements
// of that bin are still needed.  Ends with a barrier.
template <int NT>
__device__ __forceinline__ void find_bin(const uint32_t* hist, int nb, uint32_t remaining, uint32_t prefix, int shift,
                                         uint32_t* sel, uint32_t* wtot) {
  static_assert(NT >= 512, "find_bin covers at most 4 bins per thread: 2048 bins need >= 512 threads");
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  constexpr int NW = NT / 32;
  // nb is NT or 2 * NT or 4 * NT (1024 / 2048 bins); unused slots count zero
  const int per = nb / NT;
  const int b0 = nb - 1 - tid * per;
  const uint32_t c0 = hist[b0];
  const uint32_t c1 = per > 1 ? hist[b0 - 1] : 0u;
  const uint32_t c2 = per > 2 ? hist[b0 - 2] : 0u;
  const uint32_t c3 = per > 2 ? hist[b0 - 3] : 0u;
  const uint32_t mine = c0 + c1 + c2 + c3;
  uint32_t incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) wtot[wid] = incl;
  __syncthreads();
  if (wid == 0) {      // exclusive scan of the warp totals by one warp
    const uint32_t t = lane < NW ? wtot[lane] : 0u;
    uint32_t x = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += v;
    }
    if (lane < NW) wtot[lane] = x - t;
  }
  __syncthreads();
  incl += wtot[wid];
  const uint32_t excl = incl - mine;
  if (excl < remaining && incl >= remaining) {
    uint32_t run = excl;
    int j = 0;
    if (run + c0 < remaining) { run += c0; j = 1;
      if (run + c1 < remaining) { run += c1; j = 2;
        if (run + c2 < remaining) { run += c2; j = 3; } } }
    sel[0] = prefix | (static_cast<uint32_t>(b0 - j) << shift);
    sel[1] = remaining - run;
  }
  __syncthreads();
}

template <bool FILTER, int NT>
__device__ __forceinline__ void radix_kth(const float* __restrict__ src, int n, float lo, float scale, int bin_sel,
                                          uint32_t need, uint32_t* hist, uint32_t* sel, uint32_t* wtot,
                                          uint32_t& kth_out, uint32_t& copies_out) {
  const int tid = threadIdx.x;
  uint32_t prefix = 0, prefix_mask = 0, remaining = need;
  const int shifts[3] = {21, 10, 0};
  const int bits[3] = {11, 11, 10};
  for (int pass = 0; pass < 3; ++pass) {
    const int nb = 1 << bits[pass];
    for (int i = tid; i < nb; i += NT) hist[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += NT) {
      const float v = src[i];
      if (FILTER && lin_bin(v, lo, scale) != bin_sel) continue;
      const uint32_t k = f2key(v);
      if ((k & prefix_mask) == prefix) atomicAdd(&hist[(k >> shifts[pass]) & (nb - 1)], 1u);
    }
    __syncthreads();
    find_bin<NT>(hist, nb, remaining, prefix, shifts[pass], sel, wtot);
    prefix = sel[0];
    remaining = sel[1];
    prefix_mask |= static_cast<uint32_t>(nb - 1) << shifts[pass];
    __syncthreads();
  }
  kth_out = prefix;
  copies_out = remaining;
}

__global__ void __launch_bounds__(512)
topk_stats_smem_kernel(const float* __restrict__ S, int ld, int N, int K, float* __restrict__ mu,
                       float* __restrict__ sigma) {
  extern __shared__ float vals[];
  __shared__ uint32_t hist[2048];
  __shared__ float cand[TOPK_CAND];
  __shared__ uint32_t sel[2], ncand_s, wtot[16];
  __shared__ float red_lo[16], red_hi[16];
  __shared__ double red_s[16], red_q[16];
  const float* row = S + static_cast<long long>(blockIdx.x) * ld;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (int i = tid; i < 2048; i += 512) hist[i] = 0;
  if (tid == 0) ncand_s = 0;
  // 1. load + min/max
  float lo = INFINITY, hi = -INFINITY;
  const int n4 = N >> 2;
  const float4* row4 = reinterpret_cast<const float4*>(row);
  for (int i = tid; i < n4; i += 512) {
    const float4 v = row4[i];
    reinterpret_cast<float4*>(vals)[i] = v;
    lo = fminf(fminf(lo, v.x), fminf(v.y, fminf(v.z, v.w)));
    hi = fmaxf(fmaxf(hi, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
  }
  for (int i = (n4 << 2) + tid; i < N; i += 512) {
    const float v = row[i];
    vals[i] = v;
    lo = fminf(lo, v); hi = fmaxf(hi, v);
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if (lane == 0) { red_lo[wid] = lo; red_hi[wid] = hi; }
  __syncthreads();
  lo = red_lo[0]; hi = red_hi[0];
  for (int w = 1; w < 16; ++w) { lo = fminf(lo, red_lo[w]); hi = fmaxf(hi, red_hi[w]); }
  const float range = hi - lo;
  // a constant (or non-finite-range) row goes through the key radix alone: every element lands in bin 0
  const float scale = (range > 0.f && range < INFINITY) ? 2047.f / range : 0.f;
  // 2. linear histogram, threshold bin
  for (int i = tid; i < N; i += 512) atomicAdd(&hist[lin_bin(vals[i], lo, scale)], 1u);
  __syncthreads();
  find_bin<512>(hist, 2048, static_cast<uint32_t>(K), 0u, 0, sel, wtot);
  const int bin_sel = static_cast<int>(sel[0]);
  const uint32_t need = sel[1];          // how many of the threshold bin's elements belong to the top-K set
  __syncthreads();
  // 3. sweep
  double s = 0.0, q = 0.0;
  for (int i = tid; i < N; i += 512) {
    const float v = vals[i];
    const int b = lin_bin(v, lo, scale);
    if (b > bin_sel) { s += v; q += static_cast<double>(v) * v; }
    else if (b == bin_sel) {
      const uint32_t slot = atomicAdd(&ncand_s, 1u);
      if (slot < TOPK_CAND) cand[slot] = v;
    }
  }
  __syncthreads();
  const uint32_t ncand = ncand_s;
  // 4. exact select inside the threshold bin
  uint32_t kth, copies;
  if (ncand <= TOPK_CAND) {
    radix_kth<false, 512>(cand, static_cast<int>(ncand), lo, scale, bin_sel, need, hist, sel, wtot, kth, copies);
    for (int i = tid; i < static_cast<int>(ncand); i += 512) {
      const float v = cand[i];
      if (f2key(v) > kth) { s += v; q += static_cast<double>(v) * v; }
    }
  } else {
    radix_kth<true, 512>(vals, N, lo, scale, bin_sel, need, hist, sel, wtot, kth, copies);
    for (int i = tid; i < N; i += 512) {
      const float v = vals[i];
      if (lin_bin(v, lo, scale) == bin_sel && f2key(v) > kth) { s += v; q += static_cast<double>(v) * v; }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane == 0) { red_s[wid] = s; red_q[wid] = q; }
  __syncthreads();
  if (tid == 0) {
    double ts = 0.0, tq = 0.0;
    for (int w = 0; w < 16; ++w) { ts += red_s[w]; tq += red_q[w]; }
    if (copies > 0) {
      const double kv = key2f(kth);
      ts += copies * kv;
      tq += copies * kv * kv;
    }
    const double m = ts / K;
    const double var = fmax(tq / K - m * m, 0.0);
    mu[blockIdx.x] = static_cast<float>(m);
    sigma[blockIdx.x] = static_cast<float>(sqrt(var));
  }
}

// Register-resident variant for N <= NT * 4 * PER4: each thread keeps PER4 float4 of the row in registers (all
// loads in flight at once, one L2 round trip), so the kernel needs no dynamic shared memory and four CTAs share
// an SM.  Same algorithm as topk_stats_smem_kernel.
// FIXUP instances walk all `rows` with a small grid and redo only the rows the sampled kernel below flagged
// (sigma < 0): the exact path stays the arbiter for every row the fast path cannot prove.
template <int PER4, int NT, bool FIXUP>
__global__ void __launch_bounds__(NT, 2048 / NT / 2 > 0 ? 2048 / NT / 2 : 1)
topk_stats_reg_kernel(const float* __restrict__ S, int ld, int N, int K, float* __restrict__ mu,
                      float* __restrict__ sigma, int rows) {
 if (FIXUP) {
  // usually nothing is flagged: all threads look at the CTA's rows at once and leave
  int any = 0;
  for (int r = blockIdx.x + threadIdx.x * gridDim.x; r < rows; r += NT * gridDim.x) any |= !(sigma[r] >= 0.f);
  if (!__syncthreads_or(any)) return;
 }
 for (int rowi = blockIdx.x; rowi < rows; rowi += gridDim.x) {
  if (FIXUP) {
    if (sigma[rowi] >= 0.f) continue;      // uniform over the CTA
    __syncthreads();                       // shared tables of the previous row are no longer read
  }
  __shared__ uint32_t hist[2048];
  __shared__ float cand[TOPK_CAND];
  __shared__ uint32_t sel[2], ncand_s, wtot[NT / 32];
  __shared__ float red_lo[NT / 32], red_hi[NT / 32];
  __shared__ double red_s[NT / 32], red_q[NT / 32];
  const float* row = S + static_cast<long long>(rowi) * ld;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (int i = tid; i < 2048; i += NT) hist[i] = 0;
  if (tid == 0) ncand_s = 0;
  // 1. load (rows are padded to a multiple of 64 floats, so whole float4 reads stay inside the row); slots past
  //    N hold -inf: never counted, never above the threshold bin
  float v[PER4 * 4];
  const float4* row4 = reinterpret_cast<const float4*>(row);
  const int n4 = (N + 3) >> 2;
  const float NINF = -INFINITY;
#pragma unroll
  for (int j = 0; j < PER4; ++j) {
    const int i4 = tid + NT * j;
    float4 q = make_float4(NINF, NINF, NINF, NINF);
    if (i4 < n4) {
      q = row4[i4];
      if (4 * i4 + 1 >= N) q.y = NINF;
      if (4 * i4 + 2 >= N) q.z = NINF;
      if (4 * i4 + 3 >= N) q.w = NINF;
    }
    v[4 * j] = q.x; v[4 * j + 1] = q.y; v[4 * j + 2] = q.z; v[4 * j + 3] = q.w;
  }
  float lo = INFINITY, hi = NINF;
#pragma unroll
  for (int j = 0; j < PER4 * 4; ++j) {
    hi = fmaxf(hi, v[j]);
    lo = fminf(lo, v[j] == NINF ? INFINITY : v[j]);
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if (lane == 0) { red_lo[wid] = lo; red_hi[wid] = hi; }
  __syncthreads();
  lo = red_lo[0]; hi = red_hi[0];
  for (int w = 1; w < NT / 32; ++w) { lo = fminf(lo, red_lo[w]); hi = fmaxf(hi, red_hi[w]); }
  const float range = hi - lo;
  const float scale = (range > 0.f && range < INFINITY) ? 2047.f / range : 0.f;
  // 2. linear histogram (predicated shared-memory atomics, no branches)
#pragma unroll
  for (int j = 0; j < PER4 * 4; ++j)
    if (v[j] != NINF) atomicAdd(&hist[lin_bin(v[j], lo, scale)], 1u);
  __syncthreads();
  find_bin<NT>(hist, 2048, static_cast<uint32_t>(K), 0u, 0, sel, wtot);
  const int bin_sel = static_cast<int>(sel[0]);
  const uint32_t need = sel[1];
  __syncthreads();
  // 3. sweep: one rarely-taken branch per element
  double s = 0.0, q = 0.0;
#pragma unroll
  for (int j = 0; j < PER4 * 4; ++j) {
    const int b = lin_bin(v[j], lo, scale);          // -inf -> negative bin
    if (b >= bin_sel) {
      if (b > bin_sel) { s += v[j]; q += static_cast<double>(v[j]) * v[j]; }
      else {
        const uint32_t slot = atomicAdd(&ncand_s, 1u);
        if (slot < TOPK_CAND) cand[slot] = v[j];
      }
    }
  }
  __syncthreads();
  const uint32_t ncand = ncand_s;
  uint32_t kth = 0, copies = 0;
  if (ncand <= NT) {
    // few candidates (the usual case): thread t ranks candidate t by counting; the `need` best are summed
    // directly (ties broken by slot, which cannot change a sum of equal values)
    if (tid < static_cast<int>(ncand)) {
      const float mine = cand[tid];
      uint32_t rank = 0;
      for (int i = 0; i < static_cast<int>(ncand); ++i) {
        const float c = cand[i];
        rank += (c > mine || (c == mine && i < tid)) ? 1u : 0u;
      }
      if (rank < need) { s += mine; q += static_cast<double>(mine) * mine; }
    }
  } else if (ncand <= TOPK_CAND) {
    radix_kth<false, NT>(cand, static_cast<int>(ncand), lo, scale, bin_sel, need, hist, sel, wtot, kth, copies);
    for (int i = tid; i < static_cast<int>(ncand); i += NT) {
      const float c = cand[i];
      if (f2key(c) > kth) { s += c; q += static_cast<double>(c) * c; }
    }
  } else {
    radix_kth<true, NT>(row, N, lo, scale, bin_sel, need, hist, sel, wtot, kth, copies);   // degenerate rows: re-read
#pragma unroll
    for (int j = 0; j < PER4 * 4; ++j) {
      if (lin_bin(v[j], lo, scale) == bin_sel && f2key(v[j]) > kth) { s += v[j]; q += static_cast<double>(v[j]) * v[j]; }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane == 0) { red_s[wid] = s; red_q[wid] = q; }
  __syncthreads();
  if (tid == 0) {
    double ts = 0.0, tq = 0.0;
    for (int w = 0; w < NT / 32; ++w) { ts += red_s[w]; tq += red_q[w]; }
    if (copies > 0) {
      const double kv = key2f(kth);
      ts += copies * kv;
      tq += copies * kv * kv;
    }
    const double m = ts / K;
    const double var = fmax(tq / K - m * m, 0.0);
    mu[rowi] = static_cast<float>(m);
    sigma[rowi] = static_cast<float>(sqrt(var));
  }
 }
}


// Sampled pre-filter + exact selection among the survivors.  The exact kernel above spends ~20 instructions per
// score, most of them on the 2048-bin shared-memory histogram of ALL N scores.  Only the top K matter: a histogram of
// every 4th score (the x components of the float4 the thread already holds) estimates a threshold that keeps about
// K + 8 sqrt(K) scores, the full row is then swept with ONE compare per score, and the exact algorithm (linear
// histogram, threshold bin, rank by counting) runs on the few hundred survivors in shared memory.  Exactness: the
// survivors are {v >= thr}, so they contain the K largest whenever there are at least K of them; rows where the
// estimate fails (fewer than K or more than TOPK_L1 survivors, degenerate ranges, too many ties) are flagged with
// sigma = -1 and redone by the exact kernel (FIXUP instance).
#define TOPK_L1 2048
// The row is NOT kept in registers here (the exact kernel's 24 values per thread fill the register file with one CTA
// per SM, and a row is a chain of ~10 barrier-separated phases: 8.5 us of latency per row with nothing to overlap it).
// Only the sample lives in registers; the sweep re-reads the row from L2.  512 threads, 32 registers: four CTAs per SM.
template <int SAMP4, int NT>
__global__ void __launch_bounds__(NT, 4)
topk_stats_sampled_kernel(const float* __restrict__ S, int ld, int N, int K, float* __restrict__ mu,
                          float* __restrict__ sigma) {
  __shared__ uint32_t hist[2048];
  __shared__ float l1[TOPK_L1];
  __shared__ float cand[NT];
  __shared__ uint32_t sel[2], n1_s, ncand_s, wtot[NT / 32];
  __shared__ float red_lo[NT / 32], red_hi[NT / 32];
  __shared__ double red_s[NT / 32], red_q[NT / 32];
  const float* row = S + static_cast<long long>(blockIdx.x) * ld;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (int i = tid; i < 2048; i += NT) hist[i] = 0;
  if (tid == 0) { n1_s = 0; ncand_s = 0; }
  const float4* row4 = reinterpret_cast<const float4*>(row);
  const int n4 = N >> 2;                       // whole float4 groups; the (< 4) tail elements are handled in the sweep
  const float NINF = -INFINITY;
  // A. sample = every 4th float4 of the row (kept in registers), its range, its histogram
  float v[SAMP4 * 4];
  int nsamp = 0;
#pragma unroll
  for (int j = 0; j < SAMP4; ++j) {
    const int i4 = 4 * (tid + NT * j);
    float4 q = make_float4(NINF, NINF, NINF, NINF);
    if (i4 < n4) { q = row4[i4]; nsamp += 4; }
    v[4 * j] = q.x; v[4 * j + 1] = q.y; v[4 * j + 2] = q.z; v[4 * j + 3] = q.w;
  }
  float lo = INFINITY, hi = NINF;
#pragma unroll
  for (int j = 0; j < SAMP4 * 4; ++j) {
    hi = fmaxf(hi, v[j]);
    lo = fminf(lo, v[j] == NINF ? INFINITY : v[j]);
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    nsamp += __shfl_xor_sync(0xffffffffu, nsamp, o);
  }
  if (lane == 0) { red_lo[wid] = lo; red_hi[wid] = hi; wtot[wid] = static_cast<uint32_t>(nsamp); }
  __syncthreads();
  lo = red_lo[0]; hi = red_hi[0];
  uint32_t total_samp = wtot[0];
  for (int w = 1; w < NT / 32; ++w) { lo = fminf(lo, red_lo[w]); hi = fmaxf(hi, red_hi[w]); total_samp += wtot[w]; }
  __syncthreads();                             // wtot is reused by find_bin
  const float range = hi - lo;
  const float scale = (range > 0.f && range < INFINITY) ? 2047.f / range : 0.f;
#pragma unroll
  for (int j = 0; j < SAMP4 * 4; ++j)
    if (v[j] != NINF) atomicAdd(&hist[lin_bin(v[j], lo, scale)], 1u);
  __syncthreads();
  const int ks4 = (K + 3) / 4;
  const uint32_t want = static_cast<uint32_t>(ks4 + 6 * static_cast<int>(sqrtf(static_cast<float>(ks4)) + 1.f) + 8);
  const bool sample_ok = scale > 0.f && want <= total_samp;
  float thr = lo;
  if (sample_ok) {
    find_bin<NT>(hist, 2048, want, 0u, 0, sel, wtot);
    thr = lo + static_cast<float>(sel[0]) / scale;          // about the lower edge of the threshold bin
  }
  __syncthreads();
  // B. sweep of the whole row (from L2): one compare per score, survivors to shared memory
  if (sample_ok) {
    for (int i4 = tid; i4 < n4; i4 += NT) {
      const float4 q = row4[i4];
      const float x[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (x[c] >= thr) {
          const uint32_t slot = atomicAdd(&n1_s, 1u);
          if (slot < TOPK_L1) l1[slot] = x[c];
        }
      }
    }
    if (tid < (N & 3)) {
      const float x = row[4 * n4 + tid];
      if (x >= thr) {
        const uint32_t slot = atomicAdd(&n1_s, 1u);
        if (slot < TOPK_L1) l1[slot] = x;
      }
    }
  }
  for (int i = tid; i < 2048; i += NT) hist[i] = 0;
  __syncthreads();
  const uint32_t n1 = n1_s;
  if (!sample_ok || n1 < static_cast<uint32_t>(K) || n1 > TOPK_L1) {
    if (tid == 0) sigma[blockIdx.x] = -1.f;                 // exact kernel redoes this row
    return;
  }
  // C. exact top-K among the survivors: linear histogram over [thr, max], threshold bin, rank inside it
  float hi1 = NINF;
  for (int i = tid; i < static_cast<int>(n1); i += NT) hi1 = fmaxf(hi1, l1[i]);
  for (int o = 16; o > 0; o >>= 1) hi1 = fmaxf(hi1, __shfl_xor_sync(0xffffffffu, hi1, o));
  if (lane == 0) red_hi[wid] = hi1;
  __syncthreads();
  hi1 = red_hi[0];
  for (int w = 1; w < NT / 32; ++w) hi1 = fmaxf(hi1, red_hi[w]);
  const float range1 = hi1 - thr;
  const float scale1 = (range1 > 0.f && range1 < INFINITY) ? 2047.f / range1 : 0.f;
  for (int i = tid; i < static_cast<int>(n1); i += NT) atomicAdd(&hist[max(0, lin_bin(l1[i], thr, scale1))], 1u);
  __syncthreads();
  find_bin<NT>(hist, 2048, static_cast<uint32_t>(K), 0u, 0, sel, wtot);
  const int bin_sel = static_cast<int>(sel[0]);
  const uint32_t need = sel[1];
  __syncthreads();
  double s = 0.0, q = 0.0;
  for (int i = tid; i < static_cast<int>(n1); i += NT) {
    const float x = l1[i];
    const int b = max(0, lin_bin(x, thr, scale1));
    if (b > bin_sel) { s += x; q += static_cast<double>(x) * x; }
    else if (b == bin_sel) {
      const uint32_t slot = atomicAdd(&ncand_s, 1u);
      if (slot < NT) cand[slot] = x;
    }
  }
  __syncthreads();
  const uint32_t ncand = ncand_s;
  if (ncand > NT) {                                          // heavy ties in the threshold bin: exact kernel
    if (tid == 0) sigma[blockIdx.x] = -1.f;
    return;
  }
  if (tid < static_cast<int>(ncand)) {
    const float mine = cand[tid];
    uint32_t rank = 0;
    for (int i = 0; i < static_cast<int>(ncand); ++i) {
      const float c = cand[i];
      rank += (c > mine || (c == mine && i < tid)) ? 1u : 0u;
    }
    if (rank < need) { s += mine; q += static_cast<double>(mine) * mine; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane == 0) { red_s[wid] = s; red_q[wid] = q; }
  __syncthreads();
  if (tid == 0) {
    double ts = 0.0, tq = 0.0;
    for (int w = 0; w < NT / 32; ++w) { ts += red_s[w]; tq += red_q[w]; }
    const double m = ts / K;
    const double var = fmax(tq / K - m * m, 0.0);
    mu[blockIdx.x] = static_cast<float>(m);
    sigma[blockIdx.x] = static_cast<float>(sqrt(var));
  }
}

// ---------------------------------------------------------------- fused pre-filter path (threshold -> filter GEMM -> select)
// The cohort-score matrix never reaches memory on this path.  Per chunk of utterance rows:
//   1. a SAMPLE GEMM scores the rows against the first Ns cohort vectors (cohort rows are exchangeable, so any fixed
//      subset is a sample) -> S1 [rows, Ns] (L2 resident);
//   2. thr_warp_kernel turns each row's sample into a threshold that about K + 4.5 sqrt(K * N / Ns) of the N
//      scores will pass (histogram quantile of the sample), and zeroes the row's candidate counter;
//   3. the FILTER GEMM (gemm_pw.cu, ACT_FILTER) recomputes all N scores and appends those >= thr to the row's
//      candidate list straight from TMEM;
//   4. cand_select_warp_kernel does the exact top-K statistics among the candidates (they contain the K largest whenever
//      there are at least K of them).  Rows it cannot prove -- fewer than K or more than FILT_CAP candidates, heavy ties
//      -- are appended to a list and redone exactly by fixup_scores_kernel + fixup_rows_kernel (direct fp32 scores
//      + 3-pass radix select).
// Extra work: Ns / N of the GEMM FLOPs (1/8); saved: the 4-byte store and the re-read of every score.
#define FILT_CAP 1024

// cohort [N, D] fp32 -> packed split-bf16 rows [N, 2, Dp] in a PERMUTED order: packed row j holds cohort row
// (j * stride) mod N, stride coprime to N (about 0.618 N).  The order of the cohort is irrelevant to the top-K
// statistics, and with this order any prefix of the packed rows is spread evenly over the cohort, so the sample GEMM's
// first Ns rows stay representative when the caller's cohort is sorted (by speaker, by language, by score ...).
__global__ void pack_cohort_permuted_kernel(const float* __restrict__ c, int N, int D, int Dp, long long stride,
                                            bf16* __restrict__ out) {
  const long long total = static_cast<long long>(N) * Dp;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int d = static_cast<int>(i % Dp);
    const long long j = i / Dp;
    const long long src = (j * stride) % N;
    const float v = d < D ? c[src * D + d] : 0.f;
    const bf16 h = __float2bfloat16_rn(v);
    out[j * 2 * Dp + d] = h;
    out[j * 2 * Dp + Dp + d] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

static long long coprime_stride(int N) {
  long long s = static_cast<long long>(0.6180339887 * N);
  if (s < 1) s = 1;
  auto gcd = [](long long a, long long b) { while (b) { const long long t = a % b; a = b; b = t; } return a; };
  while (gcd(s, N) != 1) ++s;
  return s;
}

// Both per-row kernels run ONE WARP per row (no block barriers: a row is 2560 sample scores / ~550 candidates, far too
// little work to amortise the ~10 barrier-separated phases of a CTA-per-row kernel -- a CTA-per-row first version spent
// 65 + 57 us per 8192 rows here, 25 % of the fused path).
//
// Threshold from the row's sample: 256-bin linear histogram over the sample's range in this warp's shared-memory slice,
// scan from the top, thr = lower edge of the bin that holds the want-th largest sample score.  The bin adds at most a
// handful of sample scores (~5 at cfg5's geometry), i.e. a few dozen extra candidates.
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
thr_warp_kernel(const float* __restrict__ S1, int ld, int Ns, int want, float* __restrict__ thr,
                unsigned* __restrict__ cnt, int rows) {
  __shared__ uint32_t hist[WARPS][256];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int r = blockIdx.x * WARPS + wid;
  if (r >= rows) return;
  const float* row = S1 + static_cast<long long>(r) * ld;
  const float4* row4 = reinterpret_cast<const float4*>(row);
  const int n4 = Ns >> 2;                               // ld and Ns are multiples of 4 (Ns a multiple of 256 or = N)
  // the sample row is read ONCE: up to 20 float4 per lane (Ns <= 2560, i.e. N <= 20 480) stay in registers between the
  // range pass and the histogram pass (the chunk's sample matrix is 335 MB at 32 768 rows: not L2 resident); longer
  // samples re-read the tail from memory
  constexpr int REG4 = 20;
  float4 keep[REG4];
  float lo = INFINITY, hi = -INFINITY;
#pragma unroll
  for (int j = 0; j < REG4; ++j) {
    const int i = lane + 32 * j;
    keep[j] = i < n4 ? row4[i] : make_float4(INFINITY, INFINITY, INFINITY, INFINITY);
  }
#pragma unroll
  for (int j = 0; j < REG4; ++j) {
    if (lane + 32 * j < n4) {
      const float4 v = keep[j];
      lo = fminf(fminf(lo, v.x), fminf(v.y, fminf(v.z, v.w)));
      hi = fmaxf(fmaxf(hi, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
    }
  }
  for (int i = lane + 32 * REG4; i < n4; i += 32) {
    const float4 v = row4[i];
    lo = fminf(fminf(lo, v.x), fminf(v.y, fminf(v.z, v.w)));
    hi = fmaxf(fmaxf(hi, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
  }
  for (int i = (n4 << 2) + lane; i < Ns; i += 32) { lo = fminf(lo, row[i]); hi = fmaxf(hi, row[i]); }
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  const float range = hi - lo;
  const float scale = (range > 0.f && range < INFINITY) ? 255.f / range : 0.f;
  if (scale == 0.f || want > Ns) {          // constant / non-finite sample: everything passes, the row goes to the fix-up
    if (lane == 0) { thr[r] = -INFINITY; cnt[r] = 0u; }
    return;
  }
  uint32_t* h = hist[wid];
#pragma unroll
  for (int j = 0; j < 8; ++j) h[lane * 8 + j] = 0;
  __syncwarp();
  auto bin = [&](float v) { return min(255, __float2int_rz((v - lo) * scale)); };
#pragma unroll
  for (int j = 0; j < REG4; ++j) {
    if (lane + 32 * j < n4) {
      const float4 v = keep[j];
      atomicAdd(&h[bin(v.x)], 1u); atomicAdd(&h[bin(v.y)], 1u); atomicAdd(&h[bin(v.z)], 1u); atomicAdd(&h[bin(v.w)], 1u);
    }
  }
  for (int i = lane + 32 * REG4; i < n4; i += 32) {
    const float4 v = row4[i];
    atomicAdd(&h[bin(v.x)], 1u); atomicAdd(&h[bin(v.y)], 1u); atomicAdd(&h[bin(v.z)], 1u); atomicAdd(&h[bin(v.w)], 1u);
  }
  for (int i = (n4 << 2) + lane; i < Ns; i += 32) atomicAdd(&h[bin(row[i])], 1u);
  __syncwarp();
  // lane l owns bins 255 - 8 l ... 255 - 8 l - 7 (descending); inclusive prefix over lanes = count from the top
  uint32_t c[8], mine = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) { c[j] = h[255 - 8 * lane - j]; mine += c[j]; }
  uint32_t incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  const uint32_t excl = incl - mine;
  if (excl < static_cast<uint32_t>(want) && incl >= static_cast<uint32_t>(want)) {     // exactly one lane
    uint32_t run = excl;
    int j = 0;
#pragma unroll
    for (int q = 0; q < 7; ++q) if (j == q && run + c[q] < static_cast<uint32_t>(want)) { run += c[q]; j = q + 1; }
    thr[r] = lo + static_cast<float>(255 - 8 * lane - j) / scale;
    cnt[r] = 0u;
  }
}

// Exact top-K statistics among the candidates of one row, one warp per row: 256-bin linear histogram over the
// candidates' range, the bin holding the K-th largest, fp64 sums of everything above it, rank-by-counting inside it.
// Unproven rows (fewer than K or more than FILT_CAP candidates, more than 32 values in the K-th value's bin: heavy
// ties) go to the fix-up list.
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
cand_select_warp_kernel(const float* __restrict__ cand, const unsigned* __restrict__ cnt, int K, float* __restrict__ mu,
                        float* __restrict__ sigma, int row0, int rows, unsigned* __restrict__ nflag,
                        int* __restrict__ flag_list) {
  __shared__ uint32_t hist[WARPS][256];
  __shared__ float tie[WARPS][32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int r = blockIdx.x * WARPS + wid;
  if (r >= rows) return;
  const unsigned n = cnt[r];
  if (n < static_cast<unsigned>(K) || n > FILT_CAP) {
    if (lane == 0) flag_list[atomicAdd(nflag, 1u)] = row0 + r;
    return;
  }
  const float* c = cand + static_cast<long long>(r) * FILT_CAP;
  const int nj = (static_cast<int>(n) + 31) >> 5;        // warp-uniform trip count (<= FILT_CAP / 32)
  float v[FILT_CAP / 32];
  float lo = INFINITY, hi = -INFINITY;
#pragma unroll
  for (int j = 0; j < FILT_CAP / 32; ++j) {
    v[j] = -INFINITY;
    if (j < nj) {
      const int i = lane + 32 * j;
      if (i < static_cast<int>(n)) { v[j] = c[i]; hi = fmaxf(hi, v[j]); lo = fminf(lo, v[j]); }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  const float range = hi - lo;
  const float scale = (range > 0.f && range < INFINITY) ? 255.f / range : 0.f;      // 0: all equal -> one bin
  uint32_t* h = hist[wid];
#pragma unroll
  for (int j = 0; j < 8; ++j) h[lane + 32 * j] = 0;
  __syncwarp();
  auto bin = [&](float x) { return min(255, __float2int_rz((x - lo) * scale)); };
#pragma unroll
  for (int j = 0; j < FILT_CAP / 32; ++j)
    if (j < nj && v[j] != -INFINITY) atomicAdd(&h[bin(v[j])], 1u);
  __syncwarp();
  // lane l owns bins 255 - 8 l ... 255 - 8 l - 7 (descending); inclusive prefix over lanes = count from the top
  uint32_t cb[8], mine = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) { cb[j] = h[255 - 8 * lane - j]; mine += cb[j]; }
  uint32_t incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const uint32_t excl = incl - mine;
  const bool owner = excl < static_cast<uint32_t>(K) && incl >= static_cast<uint32_t>(K);
  int bin_sel = 0;
  uint32_t need = 0;
  if (owner) {
    uint32_t run = excl;
    int j = 0;
#pragma unroll
    for (int q = 0; q < 7; ++q) if (j == q && run + cb[q] < static_cast<uint32_t>(K)) { run += cb[q]; j = q + 1; }
    bin_sel = 255 - 8 * lane - j;
    need = static_cast<uint32_t>(K) - run;          // how many of the threshold bin's values belong to the top-K set
  }
  const unsigned om = __ballot_sync(0xffffffffu, owner);
  const int src = __ffs(om) - 1;
  bin_sel = __shfl_sync(0xffffffffu, bin_sel, src);
  need = __shfl_sync(0xffffffffu, need, src);
  double s = 0.0, q = 0.0;
  uint32_t ntie = 0;
#pragma unroll
  for (int j = 0; j < FILT_CAP / 32; ++j) {
    if (j >= nj) break;                                  // uniform
    const bool live = v[j] != -INFINITY;
    const int b = live ? bin(v[j]) : -1;
    if (b > bin_sel) { s += v[j]; q += static_cast<double>(v[j]) * v[j]; }
    const unsigned tm = __ballot_sync(0xffffffffu, b == bin_sel);
    if (b == bin_sel) {
      const uint32_t pos = ntie + __popc(tm & ((1u << lane) - 1u));
      if (pos < 32) tie[wid][pos] = v[j];
    }
    ntie += __popc(tm);
  }
  __syncwarp();
  if (ntie > 32) {                                   // heavy ties in the threshold bin: exact fix-up
    if (lane == 0) flag_list[atomicAdd(nflag, 1u)] = row0 + r;
    return;
  }
  if (lane < static_cast<int>(ntie)) {
    const float me = tie[wid][lane];
    uint32_t rank = 0;
    for (int i = 0; i < static_cast<int>(ntie); ++i) {
      const float t = tie[wid][i];
      rank += (t > me || (t == me && i < lane)) ? 1u : 0u;
    }
    if (rank < need) { s += me; q += static_cast<double>(me) * me; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane == 0) {
    const double m = s / K;
    const double var = fmax(q / K - m * m, 0.0);
    mu[row0 + r] = static_cast<float>(m);
    sigma[row0 + r] = static_cast<float>(sqrt(var));
  }
}

// Exact fall-back for the rows the fused path could not prove (flag_list[0 .. *nflag)): the row's N cohort scores are
// recomputed directly in fp32 (one warp per cohort vector, coalesced reads, the cohort stays in L2) and selected with
// the 3-pass key radix.  A sweep of random embeddings flags ~3e-6 of its rows (the 4.5-sigma margin of the sampled
// threshold), i.e. one row in every third 100 000-row call, so the common case is ONE flagged row: its score row is
// spread over the whole GPU (fixup_scores_kernel, 128 cohort vectors per work item, ~10 us) instead of being walked by
// one CTA (2.6 ms measured: 2 500 dependent L2 round trips per warp -- two of eight ranks paid it in the round-2
// 8-GPU run).  Flagged rows beyond the first FIXUP_CTAS are computed by the CTA that selects them (slow, exact).
static constexpr int FIXUP_SLICE = 128;

// scores of cohort vectors [j0, j1) against e_s into srow; warps of the CTA interleave groups of four vectors so that
// 4 * D / 32 independent loads are in flight per lane.  Same summation order per score as a one-vector loop.
__device__ __forceinline__ void fixup_scores_span(const float* __restrict__ e_s, const float* __restrict__ cohort, int D,
                                                  int j0, int j1, float* __restrict__ srow, int wid, int lane, int nw) {
  for (int j = j0 + wid * 4; j < j1; j += nw * 4) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    const float* c0 = cohort + static_cast<long long>(j) * D;
    const float* c1 = cohort + static_cast<long long>(min(j + 1, j1 - 1)) * D;
    const float* c2 = cohort + static_cast<long long>(min(j + 2, j1 - 1)) * D;
    const float* c3 = cohort + static_cast<long long>(min(j + 3, j1 - 1)) * D;
    for (int c = lane; c < D; c += 32) {
      const float e = e_s[c];
      a0 = fmaf(e, __ldg(c0 + c), a0);
      a1 = fmaf(e, __ldg(c1 + c), a1);
      a2 = fmaf(e, __ldg(c2 + c), a2);
      a3 = fmaf(e, __ldg(c3 + c), a3);
    }
    a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2); a3 = warp_sum(a3);
    const float v = lane == 0 ? a0 : lane == 1 ? a1 : lane == 2 ? a2 : a3;
    if (lane < 4 && j + lane < j1) srow[j + lane] = v;
  }
}

// Work item = (flagged row f < min(*nflag, max_rows), slice of FIXUP_SLICE cohort vectors); scores go to scratch row f.
__global__ void __launch_bounds__(256)
fixup_scores_kernel(const float* __restrict__ emean, const float* __restrict__ cohort, int N, int D,
                    const unsigned* __restrict__ nflag, const int* __restrict__ flag_list, float* __restrict__ scratch,
                    int scratch_ld, int max_rows) {
  extern __shared__ float e_s[];          // [D]
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const unsigned rows = min(*nflag, static_cast<unsigned>(max_rows));
  const unsigned nslices = static_cast<unsigned>((N + FIXUP_SLICE - 1) / FIXUP_SLICE);
  const unsigned long long total = static_cast<unsigned long long>(rows) * nslices;
  int cur = -1;
  for (unsigned long long it = blockIdx.x; it < total; it += gridDim.x) {
    const int f = static_cast<int>(it / nslices), sl = static_cast<int>(it % nslices);
    if (f != cur) {
      __syncthreads();
      const int u = flag_list[f];
      for (int c = tid; c < D; c += 256) e_s[c] = emean[static_cast<long long>(u) * D + c];
      __syncthreads();
      cur = f;
    }
    fixup_scores_span(e_s, cohort, D, sl * FIXUP_SLICE, min(N, (sl + 1) * FIXUP_SLICE),
                      scratch + static_cast<long long>(f) * scratch_ld, wid, lane, 8);
  }
}

// Selection: CTA b takes flagged rows b, b + gridDim.x, ...; the first one's scores are already in scratch row b
// (fixup_scores_kernel, gridDim.x == its max_rows), later ones are computed here into the same row.
__global__ void __launch_bounds__(256)
fixup_rows_kernel(const float* __restrict__ emean, const float* __restrict__ cohort, int N, int D, int K,
                  const unsigned* __restrict__ nflag, const int* __restrict__ flag_list, float* __restrict__ scratch,
                  int scratch_ld, float* __restrict__ mu, float* __restrict__ sigma) {
  extern __shared__ float e_s[];          // [D]
  __shared__ uint32_t hist[2048];
  __shared__ uint32_t sel_prefix, sel_remaining;
  __shared__ double red_s[8], red_q[8];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  float* srow = scratch + static_cast<long long>(blockIdx.x) * scratch_ld;
  const unsigned total = *nflag;
  for (unsigned it = blockIdx.x; it < total; it += gridDim.x) {
    const int u = flag_list[it];
    __syncthreads();
    if (it != blockIdx.x) {
      for (int c = tid; c < D; c += 256) e_s[c] = emean[static_cast<long long>(u) * D + c];
      __syncthreads();
      fixup_scores_span(e_s, cohort, D, 0, N, srow, wid, lane, 8);
      __syncthreads();
    }
    // 3-pass (11 + 11 + 10 bit) key radix select of the K largest (same algorithm as topk_stats_kernel)
    uint32_t prefix = 0, prefix_mask = 0, remaining = K;
    const int shifts[3] = {21, 10, 0};
    const int bits[3] = {11, 11, 10};
    for (int pass = 0; pass < 3; ++pass) {
      const int nb = 1 << bits[pass];
      for (int i = tid; i < nb; i += 256) hist[i] = 0;
      __syncthreads();
      for (int i = tid; i < N; i += 256) {
        const uint32_t k = f2key(srow[i]);
        if ((k & prefix_mask) == prefix) atomicAdd(&hist[(k >> shifts[pass]) & (nb - 1)], 1u);
      }
      __syncthreads();
      if (tid < 32) {
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
              sel_remaining = remaining - run;
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
    const uint32_t kth = prefix;
    double s = 0.0, q = 0.0;
    for (int i = tid; i < N; i += 256) {
      const float v = srow[i];
      if (f2key(v) > kth) { s += v; q += static_cast<double>(v) * v; }
    }
    for (int o = 16; o > 0; o >>= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, o);
      q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if (lane == 0) { red_s[wid] = s; red_q[wid] = q; }
    __syncthreads();
    if (tid == 0) {
      double ts = 0.0, tq = 0.0;
      for (int w = 0; w < 8; ++w) { ts += red_s[w]; tq += red_q[w]; }
      const double kv = key2f(kth);
      ts += remaining * kv;
      tq += remaining * kv * kv;
      const double m = ts / K;
      const double var = fmax(tq / K - m * m, 0.0);
      mu[u] = static_cast<float>(m);
      sigma[u] = static_cast<float>(sqrt(var));
    }
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

// ---------------------------------------------------------------- nearest-enrolment search (--predict)
// The statistic of src/inference.py:284-294: dist[s] = mean_i || q_i - e_(i,s) + eps ||_2 over the n index-paired crops
// (torch's pairwise_distance adds eps to the difference), enrolment bank in the reference's saved layout
// [n, D, S] (src/model.py:634-641: per-speaker mean crop embeddings concatenated on the LAST axis).  Thread = speaker
// (coalesced bank reads), ENROL_QT queries per block share every bank element from a register; the crop's query
// vectors sit in shared memory.  fp32 throughout, direct differences (no ||a||^2 + ||b||^2 - 2ab cancellation).
static constexpr int ENROL_QT = 8;
static constexpr int ENROL_NT = 128;
__global__ void __launch_bounds__(ENROL_NT) enrol_dist_kernel(const float* __restrict__ query, const float* __restrict__ bank,
                                                              int Q, int n, int D, int S, float eps,
                                                              float* __restrict__ dist) {
  extern __shared__ float qs[];   // [ENROL_QT][D]
  const int s = blockIdx.x * ENROL_NT + threadIdx.x;
  const int q0 = blockIdx.y * ENROL_QT;
  float tot[ENROL_QT];
#pragma unroll
  for (int q = 0; q < ENROL_QT; ++q) tot[q] = 0.f;
  for (int i = 0; i < n; ++i) {
    __syncthreads();
    for (int k = threadIdx.x; k < ENROL_QT * D; k += ENROL_NT) {
      const int q = q0 + k / D;
      qs[k] = q < Q ? query[(static_cast<long long>(q) * n + i) * D + (k % D)] + eps : 0.f;
    }
    __syncthreads();
    if (s < S) {
      float acc[ENROL_QT];
#pragma unroll
      for (int q = 0; q < ENROL_QT; ++q) acc[q] = 0.f;
      const float* b = bank + static_cast<long long>(i) * D * S + s;
      for (int d = 0; d < D; ++d) {
        const float e = __ldg(b + static_cast<long long>(d) * S);
#pragma unroll
        for (int q = 0; q < ENROL_QT; ++q) {
          const float df = qs[q * D + d] - e;
          acc[q] = fmaf(df, df, acc[q]);
        }
      }
#pragma unroll
      for (int q = 0; q < ENROL_QT; ++q) tot[q] += sqrtf(acc[q]);
    }
  }
  if (s < S) {
#pragma unroll
    for (int q = 0; q < ENROL_QT; ++q)
      if (q0 + q < Q) dist[static_cast<long long>(q0 + q) * S + s] = tot[q] / static_cast<float>(n);
  }
}

// The k nearest speakers of one query (k <= 8; np.argsort(dist)[:k], ties to the lower index): k rounds of a block
// arg-min over the query's distance row, each round excluding everything at or before the previous winner in
// (distance, index) order.
__global__ void __launch_bounds__(256) enrol_topk_kernel(const float* __restrict__ dist, int S, int k,
                                                         float* __restrict__ out_d, int32_t* __restrict__ out_i) {
  __shared__ float rv[8];
  __shared__ int ri[8];
  const float* row = dist + static_cast<long long>(blockIdx.x) * S;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float pv = -INFINITY;
  int pi = -1;
  for (int r = 0; r < k; ++r) {
    float bv = INFINITY;
    int bi = 0x7fffffff;
    for (int s = threadIdx.x; s < S; s += 256) {
      const float v = row[s];
      const bool after = v > pv || (v == pv && s > pi);
      if (after && (v < bv || (v == bv && s < bi))) { bv = v; bi = s; }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) { rv[warp] = bv; ri[warp] = bi; }
    __syncthreads();
    bv = rv[0]; bi = ri[0];
    for (int w = 1; w < 8; ++w)
      if (rv[w] < bv || (rv[w] == bv && ri[w] < bi)) { bv = rv[w]; bi = ri[w]; }
    __syncthreads();
    if (threadIdx.x == 0) {
      out_d[static_cast<long long>(blockIdx.x) * k + r] = bv;
      out_i[static_cast<long long>(blockIdx.x) * k + r] = bi == 0x7fffffff ? -1 : bi;
    }
    pv = bv; pi = bi;
  }
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

int pnorm_trials(const float* emb, int n, int D, const int32_t* ia, const int32_t* ib, long long ntrials, float pw,
                 float eps, int normalize, float* out, cudaStream_t stream) {
  if (ntrials == 0) return 0;
  SVB_CHECK(pw > 0.f, "pnorm needs p > 0 (got %g)", static_cast<double>(pw));
  const int pmode = isinf(pw) ? 3 : (pw == 2.f ? 2 : (pw == 1.f ? 1 : 0));
  pnorm_trials_kernel<<<static_cast<int>((ntrials + 7) / 8), 256, 0, stream>>>(emb, n, D, ia, ib, ntrials, pw, pmode, eps,
                                                                             normalize, out);
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

int enrol_nearest(const float* query, const float* bank, int Q, int n, int D, int S, float eps, int k, float* dist,
                  float* out_d, int32_t* out_i, cudaStream_t stream) {
  if (Q == 0) return 0;
  const size_t smem = static_cast<size_t>(ENROL_QT) * D * sizeof(float);
  if (smem > 48 * 1024) { set_error("enrol_nearest: embedding size above 1536"); return -1; }
  dim3 grid((S + ENROL_NT - 1) / ENROL_NT, (Q + ENROL_QT - 1) / ENROL_QT);
  enrol_dist_kernel<<<grid, ENROL_NT, smem, stream>>>(query, bank, Q, n, D, S, eps, dist);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  if (k > 0) {
    enrol_topk_kernel<<<Q, 256, 0, stream>>>(dist, S, k, out_d, out_i);
    count_launch();
    SVB_CUDA_OK(cudaGetLastError());
  }
  return 0;
}

static size_t align256(size_t x) { return (x + 255) / 256 * 256; }

// Workspace plan.  The U x N score matrix is never materialised as a whole: it is produced in row chunks of at
// most S_CHUNK_BYTES (about one L2's worth, so most of a chunk is still cache resident when the selection
// kernel reads it back, once).  Measured (profiles/r01_scoring.md): 40 MB chunks keep S fully in L2 but cost
// more in per-launch fill/drain of the 9-k-block GEMM tiles than the DRAM round trip of 133 MB chunks.
static const size_t S_CHUNK_BYTES = 144u << 20;
static const int TOPK_SMEM_MAX_N = 50000;   // row cached in smem (<= 200 KB)

// Fused path (see "fused pre-filter path" above): rows per chunk and sample size, buffers overlaid on the S chunk
// Rows per chunk of the fused path.  Measured on one B200 (200 k rows x 20 k cohort): 8192 rows (sample matrix L2
// resident) 147 M trials/s, 16 384: 148 M, 32 768: 154 M with a third of the launches.  (The two slow ranks of the round-2
// 8-GPU runs were NOT a launch-rate effect: they had one flagged row each, see fixup_scores_kernel.)  SVB_TOPK_CHUNK overrides.
static int fused_chunk_rows() {
  static const int v = [] { const char* e = getenv("SVB_TOPK_CHUNK"); const int x = e ? atoi(e) : 0; return x >= 128 ? (x + 127) / 128 * 128 : 32768; }();
  return v;
}
static const int FIXUP_CTAS = 296;
struct AsnormPlan {
  int Dp, Np, chunk;
  size_t o_w, o_zeros, o_ones, o_a, o_s, total;
  int Ns, fch;                                           // fused path: sample columns, rows per chunk
  size_t o_s1, o_thr, o_cnt, o_cand, o_nflag, o_flag, o_fix;
};

static bool fused_enabled() {
  static const bool on = [] { const char* e = getenv("SVB_TOPK_FUSED"); return e == nullptr || e[0] != '0'; }();
  return on;
}
// the pre-filter pays when K << N and the sample (N / 8 cohort vectors) is large enough for a stable quantile
static bool fused_applies(int N, int K) { return fused_enabled() && N >= 8192 && K >= 16 && 8 * K <= N && K <= FILT_CAP / 2; }

static AsnormPlan asnorm_plan(int U, int N, int D) {
  AsnormPlan p{};
  p.Dp = (D + 63) / 64 * 64;
  p.Np = (N + 63) / 64 * 64;
  // rows per chunk: a multiple of 128 that keeps S within S_CHUNK_BYTES and fills whole waves of the
  // persistent GEMM (tiles / (ceil(tiles / SMs) * SMs) maximal)
  const long long cap = std::max<long long>(1, static_cast<long long>(S_CHUNK_BYTES / (static_cast<size_t>(p.Np) * 4)) / 128);
  const long long all = (static_cast<long long>(U) + 127) / 128;
  const int n_tiles = (p.Np + 255) / 256, sms = std::max(1, num_sms());
  long long best = std::min(cap, all);
  if (all > cap) {
    double best_eff = 0.0;
    for (long long mt = std::max<long long>(1, cap / 2); mt <= cap; ++mt) {
      const long long tiles = mt * n_tiles;
      const double eff = static_cast<double>(tiles) / (static_cast<double>((tiles + sms - 1) / sms) * sms);
      if (eff >= best_eff) { best_eff = eff; best = mt; }
    }
  }
  p.chunk = static_cast<int>(best * 128);
  const size_t Upad = (static_cast<size_t>(U) + 127) / 128 * 128;
  size_t o = 0;
  p.o_w = o; o += align256(static_cast<size_t>(N) * 2 * p.Dp * 2);
  p.o_zeros = o; o += align256(static_cast<size_t>(p.Np) * 4);
  p.o_ones = o; o += align256(static_cast<size_t>(p.Np) * 4);
  p.o_a = o; o += align256(Upad * 2 * p.Dp * 2);
  p.o_s = o;
  const size_t old_end = o + align256(static_cast<size_t>(p.chunk) * p.Np * 4);
  // fused-path buffers share the S-chunk region (only one of the two paths runs in a call)
  p.Ns = std::min(p.Np, (N / 8 + 255) / 256 * 256);
  p.fch = static_cast<int>(std::min<long long>(fused_chunk_rows(), all * 128));
  p.o_s1 = o; o += align256(static_cast<size_t>(p.fch) * p.Ns * 4);
  p.o_thr = o; o += align256(static_cast<size_t>(p.fch) * 4);
  p.o_cnt = o; o += align256(static_cast<size_t>(p.fch) * 4);
  p.o_cand = o; o += align256(static_cast<size_t>(p.fch) * FILT_CAP * 4);
  p.o_nflag = o; o += 256;
  p.o_flag = o; o += align256(static_cast<size_t>(U) * 4 + 4);
  p.o_fix = o; o += align256(static_cast<size_t>(FIXUP_CTAS) * p.Np * 4);
  p.total = std::max(old_end, o);
  return p;
}

size_t asnorm_workspace_bytes(int U, int N, int D, int* chunk_out) {
  const AsnormPlan p = asnorm_plan(U, N, D);
  if (chunk_out) *chunk_out = p.chunk;
  return p.total;
}

// Diagnostic (synchronous): how many rows of the LAST asnorm_stats call on this workspace went through the exact
// fix-up kernel (fused pre-filter path only; -1 if the call used the store-and-select path).
int asnorm_fixup_rows(const void* ws, int U, int N, int D, int topk) {
  int K = topk < 0 ? N + topk : topk;
  if (K > N) K = N;
  if (!fused_applies(N, K)) return -1;
  const AsnormPlan pl = asnorm_plan(U, N, D);
  unsigned n = 0;
  if (cudaMemcpy(&n, static_cast<const uint8_t*>(ws) + pl.o_nflag, sizeof(n), cudaMemcpyDeviceToHost) != cudaSuccess) return -2;
  return static_cast<int>(n);
}

// mu/sigma of the top-K cohort scores of every row of emean [U, D].
int asnorm_stats(const float* emean, int U, const float* cohort, int N, int D, int topk, float* mu, float* sigma,
                 void* ws, size_t ws_bytes, cudaStream_t stream) {
  int K = topk;
  if (K < 0) K = N + K;            // numpy slice semantics of S[::-1][:top] (top=-1 drops the smallest)
  if (K > N) K = N;
  SVB_CHECK(K >= 1, "top-K must select at least one cohort score (N=%d, top=%d)", N, topk);
  const AsnormPlan pl = asnorm_plan(U, N, D);
  SVB_CHECK(ws != nullptr && ws_bytes >= pl.total, "AS-norm workspace too small: need %zu bytes, got %zu", pl.total,
            ws_bytes);
  const int Dp = pl.Dp, Np = pl.Np, chunk = pl.chunk;
  uint8_t* base = static_cast<uint8_t*>(ws);
  bf16* wpk = reinterpret_cast<bf16*>(base + pl.o_w);
  float* zeros = reinterpret_cast<float*>(base + pl.o_zeros);
  float* ones = reinterpret_cast<float*>(base + pl.o_ones);
  bf16* a16 = reinterpret_cast<bf16*>(base + pl.o_a);
  float* S = reinterpret_cast<float*>(base + pl.o_s);
  // cohort [N, D] is a k=1 "conv weight" [cout=N, cin=D, 1]; embeddings are split into hi|lo planes once
  const bool fused = fused_applies(N, K);
  if (fused) {
    const long long total = static_cast<long long>(N) * Dp;
    pack_cohort_permuted_kernel<<<static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16)), 256, 0, stream>>>(
        cohort, N, D, Dp, coprime_stride(N), wpk);
    count_launch();
    SVB_CUDA_OK(cudaGetLastError());
  } else if (pack_conv_weight(cohort, N, D, 1, Dp, 2, wpk, stream)) {
    return -1;
  }
  fill_kernel<<<(Np + 255) / 256, 256, 0, stream>>>(zeros, Np, 0.f);
  fill_kernel<<<(Np + 255) / 256, 256, 0, stream>>>(ones, Np, 1.f);
  count_launch(2);
  const size_t Upad = (static_cast<size_t>(U) + 127) / 128 * 128;
  SVB_CUDA_OK(cudaMemsetAsync(a16, 0, Upad * 2 * Dp * 2, stream));
  if (split_bf16(emean, U, D, D, a16, 2 * Dp, 2, Dp, stream)) return -1;
  if (fused) {
    // ---- fused pre-filter path: no score is ever stored except the 1/8 sample and the ~2.5 % that pass the threshold
    float* S1 = reinterpret_cast<float*>(base + pl.o_s1);
    float* thr = reinterpret_cast<float*>(base + pl.o_thr);
    unsigned* cnt = reinterpret_cast<unsigned*>(base + pl.o_cnt);
    float* cand = reinterpret_cast<float*>(base + pl.o_cand);
    unsigned* nflag = reinterpret_cast<unsigned*>(base + pl.o_nflag);
    int* flag_list = reinterpret_cast<int*>(base + pl.o_flag);
    float* fix = reinterpret_cast<float*>(base + pl.o_fix);
    const int Ns = pl.Ns;
    const double kf = static_cast<double>(K) * Ns / N;                       // expected sample hits of the K-th score
    const int want = static_cast<int>(kf + 4.5 * sqrt(kf) + 4.0);
    SVB_CUDA_OK(cudaMemsetAsync(nflag, 0, sizeof(unsigned), stream));
    for (int u0 = 0; u0 < U; u0 += pl.fch) {
      const int rows = std::min(pl.fch, U - u0);
      GemmOperands ops{};
      ops.a = a16 + static_cast<size_t>(u0) * 2 * Dp; ops.lda = 2 * Dp; ops.a_cols = 2 * Dp;
      ops.w = wpk; ops.w_rows = std::min(N, Ns); ops.w_cols = 2 * Dp;
      GemmParams g{};
      // the SAMPLE scores only place a histogram bin edge (bin width ~2e-3): the hi x hi pass alone (error ~3e-4) is
      // enough for them; the filter GEMM below recomputes every score with all three split passes
      g.M = rows; g.N = Ns; g.kc_per_tap = Dp / 64; g.num_taps = 1; g.tap_off[0] = 0; g.num_pass = 1;
      g.a_plane_cols = Dp; g.w_planes = 2; g.act = ACT_NONE; g.store_mode = STORE_ALL; g.out = S1; g.out_ld = Ns;
      g.out_kind = OUT_F32;
      if (launch_gemm_f32(ops, g, stream)) return -1;
      g.num_pass = 3;
      thr_warp_kernel<16><<<(rows + 15) / 16, 512, 0, stream>>>(S1, Ns, std::min(N, Ns), want, thr, cnt, rows);
      count_launch();
      ops.w_rows = N;
      g.N = Np; g.out = nullptr;
      if (launch_gemm_filter(ops, g, thr, cnt, cand, FILT_CAP, N, stream)) return -1;
      cand_select_warp_kernel<16><<<(rows + 15) / 16, 512, 0, stream>>>(cand, cnt, K, mu, sigma, u0, rows, nflag, flag_list);
      count_launch();
      SVB_CUDA_OK(cudaGetLastError());
    }
    fixup_scores_kernel<<<FIXUP_CTAS * 2, 256, static_cast<size_t>(D) * 4, stream>>>(emean, cohort, N, D, nflag, flag_list, fix,
                                                                                Np, FIXUP_CTAS);
    fixup_rows_kernel<<<FIXUP_CTAS, 256, static_cast<size_t>(D) * 4, stream>>>(emean, cohort, N, D, K, nflag, flag_list, fix, Np,
                                                                          mu, sigma);
    count_launch();
    count_launch();
    SVB_CUDA_OK(cudaGetLastError());
    return 0;
  }
  const bool smem_rows = Np <= TOPK_SMEM_MAX_N;   // the launch asks for Np * 4 bytes (N rounded up to 64)
  if (smem_rows) {
    static unsigned long long attr_devs = 0;   // per device: kernel attributes do not carry over
    if (first_on_device(&attr_devs)) {
      SVB_CUDA_OK(cudaFuncSetAttribute(topk_stats_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       TOPK_SMEM_MAX_N * 4));
    }
  }
  for (int u0 = 0; u0 < U; u0 += chunk) {
    const int rows = std::min(chunk, U - u0);
    GemmOperands ops{};
    ops.a = a16 + static_cast<size_t>(u0) * 2 * Dp; ops.lda = 2 * Dp; ops.a_cols = 2 * Dp;
    ops.w = wpk; ops.w_rows = N; ops.w_cols = 2 * Dp;
    GemmParams g{};
    g.M = rows; g.N = Np; g.kc_per_tap = Dp / 64; g.num_taps = 1; g.tap_off[0] = 0; g.num_pass = 3;
    g.a_plane_cols = Dp; g.w_planes = 2; g.bias = zeros; g.scale = ones; g.shift = zeros;
    g.act = ACT_NONE; g.store_mode = STORE_ALL; g.out = S; g.out_ld = Np; g.out_kind = OUT_F32;
    if (launch_gemm_f32(ops, g, stream)) return -1;
    static const bool sampled = [] { const char* e = getenv("SVB_TOPK_SAMPLED"); return e == nullptr || e[0] != '0'; }();
    const bool fast = sampled && K >= 16 && 8 * K <= N;        // the pre-filter pays when K << N
    if (N <= 512 * 16) {
      if (fast) {
        topk_stats_sampled_kernel<1, 512><<<rows, 512, 0, stream>>>(S, Np, N, K, mu + u0, sigma + u0);
        topk_stats_reg_kernel<4, 512, true><<<std::min(rows, 2 * num_sms()), 512, 0, stream>>>(S, Np, N, K, mu + u0,
                                                                                              sigma + u0, rows);
        count_launch();
      } else {
        topk_stats_reg_kernel<4, 512, false><<<rows, 512, 0, stream>>>(S, Np, N, K, mu + u0, sigma + u0, rows);
      }
    } else if (N <= 1024 * 24) {
      if (fast) {
        topk_stats_sampled_kernel<3, 512><<<rows, 512, 0, stream>>>(S, Np, N, K, mu + u0, sigma + u0);
        topk_stats_reg_kernel<6, 1024, true><<<std::min(rows, num_sms()), 1024, 0, stream>>>(S, Np, N, K, mu + u0,
                                                                                             sigma + u0, rows);
        count_launch();
      } else {
        topk_stats_reg_kernel<6, 1024, false><<<rows, 1024, 0, stream>>>(S, Np, N, K, mu + u0, sigma + u0, rows);
      }
    }
    else if (smem_rows)
      topk_stats_smem_kernel<<<rows, 512, static_cast<size_t>(Np) * 4, stream>>>(S, Np, N, K, mu + u0, sigma + u0);
    else
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

// ---------------------------------------------------------------- AS-norm statistics without any allocation
// Added (no counterpart in the reference): callers that cannot allocate -- code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc -- get the top-K cohort statistics with
// the workspace of asnorm_stats in a file-scope device array instead of a caller buffer.  Every GPU that loads the
// library holds ASNORM_STATIC_BYTES of device memory until the process exits.  ALL calls on a device share the array:
// calls that overlap on the GPU (two streams, two host threads) race; the array keeps the previous call's contents, and
// asnorm_stats itself clears what must start at zero (operand padding, counters) at the start of every call.
static constexpr size_t ASNORM_STATIC_BYTES = 16u << 20;          // 16 MiB, e.g. 1024 utterances x 2048 cohort vectors x 192
static __device__ __align__(256) unsigned char g_asnorm_scratch[ASNORM_STATIC_BYTES];

size_t asnorm_static_bytes() { return ASNORM_STATIC_BYTES; }

int asnorm_stats_static(const float* emean, int U, const float* cohort, int N, int D, int topk, float* mu, float* sigma,
                        cudaStream_t stream) {
  const size_t need = asnorm_workspace_bytes(U, N, D, nullptr);
  SVB_CHECK(need <= ASNORM_STATIC_BYTES,
            "svb_asnorm_stats_static: U=%d, N=%d, D=%d needs %zu bytes of scratch, the built-in array holds %zu; use "
            "svb_asnorm_stats with a caller workspace", U, N, D, need, static_cast<size_t>(ASNORM_STATIC_BYTES));
  // g_asnorm_scratch is a host-side stand-in in this function: the device address comes from the runtime, once per
  // device, with that device current
  static constexpr int MAX_DEV = 64;
  static void* dev_ptr[MAX_DEV] = {};
  int dev = 0;
  SVB_CUDA_OK(cudaGetDevice(&dev));
  SVB_CHECK(dev >= 0 && dev < MAX_DEV, "device ordinal %d out of range", dev);
  void* ws = __atomic_load_n(&dev_ptr[dev], __ATOMIC_ACQUIRE);
  if (ws == nullptr) {
    SVB_CUDA_OK(cudaGetSymbolAddress(&ws, g_asnorm_scratch));
    SVB_CHECK(ws != nullptr && (reinterpret_cast<uintptr_t>(ws) & 255) == 0, "static scratch array is not 256-byte aligned");
    __atomic_store_n(&dev_ptr[dev], ws, __ATOMIC_RELEASE);
  }
  return asnorm_stats(emean, U, cohort, N, D, topk, mu, sigma, ws, ASNORM_STATIC_BYTES, stream);
}

}  // namespace svb
