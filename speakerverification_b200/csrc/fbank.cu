// Fused log-mel front-end: pre-emphasis -> centred reflect-padded STFT frames -> hamming window ->
// 512-point real FFT -> power -> sparse slaney mel -> log(x + 1e-6) -> per-utterance mean
// normalisation, written straight into the time-major bf16 (hi|lo) layout block 0 consumes.
//
// Reference call sites replaced (paths in the reference tree):
//   PreEmphasis.forward                    src/utils.py:63-71
//   melspectrogram(lib='torchaudio')       src/models/FeatureExtraction/feature.py:84-94
//   log + mean normalisation               src/models/ECAPA_TDNN.py:473-476
// The power spectrum [B,257,T] the reference materialises in HBM never leaves shared memory.
#include <math.h>

#include <algorithm>
#include <vector>

#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

static constexpr int NFFT = 512;
static constexpr int NBIN = NFFT / 2 + 1;
static constexpr int FRAMES_PER_TILE = 32;
static constexpr int FB_ROW_SLACK = 128;   // slack rows after each plane of the stacked layout (= the workspace plan's row slack)
static constexpr int FB_WARPS = 4;
static constexpr int MAX_MELS = 128;
static constexpr int MAX_NNZ = 4096;

struct FbankTables {
  float window[NFFT];     // win_length taps centred in n_fft, zeros elsewhere
  float2 tw256[256];      // exp(-2 pi i k / 256)
  float2 tw512[NBIN];     // exp(-2 pi i k / 512)
  int mel_start[MAX_MELS];
  int mel_count[MAX_MELS];
  int mel_off[MAX_MELS];
  float mel_w[MAX_NNZ];
  int pad_[2];             // sizeof % 16 == 0 (copied to shared memory in 16-byte pieces)
};
static_assert(sizeof(FbankTables) % 16 == 0, "FbankTables must be a multiple of 16 bytes");

struct Fbank {
  FbankTables* d_tab;
  int win_length, hop, n_mels, nnz, max_count;
  float preemph, log_eps;
};

// utt_tab[b] = {wav_off, n_samples, n_frames, row_base, tile_base, n_tiles}
struct UttDesc { long long wav_off; int nsamp; int T; long long row_base; int tile_base; int ntiles; };

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 mul_mi(float2 a) { return make_float2(a.y, -a.x); }   // a * (-i)

// forward 4-point DFT (W4 = -i), in place
__device__ __forceinline__ void dft4(float2& a0, float2& a1, float2& a2, float2& a3) {
  const float2 s02 = cadd(a0, a2), d02 = csub(a0, a2), s13 = cadd(a1, a3), d13 = mul_mi(csub(a1, a3));
  a0 = cadd(s02, s13); a1 = cadd(d02, d13); a2 = csub(s02, s13); a3 = csub(d02, d13);
}

// forward 16-point DFT in registers, natural order in and out: A[k] = sum_n a[n] W16^{nk}.
// n = na + 4 nb, k = kb + 4 ka: 4-point DFTs over nb, twiddle W16^{na kb}, 4-point DFTs over na.
__device__ __forceinline__ void fft16(float2 (&a)[16]) {
  constexpr float C1 = 0.92387953251128674f, S1 = 0.38268343236508977f, R2 = 0.70710678118654752f;
#pragma unroll
  for (int na = 0; na < 4; ++na) dft4(a[na], a[na + 4], a[na + 8], a[na + 12]);   // a[na + 4 kb] = B[na][kb]
  // twiddles W16^{na * kb}
  a[5] = cmul(a[5], make_float2(C1, -S1));      // na 1 kb 1: W^1
  a[9] = cmul(a[9], make_float2(R2, -R2));      // na 1 kb 2: W^2
  a[13] = cmul(a[13], make_float2(S1, -C1));    // na 1 kb 3: W^3
  a[6] = cmul(a[6], make_float2(R2, -R2));      // na 2 kb 1: W^2
  a[10] = mul_mi(a[10]);                        // na 2 kb 2: W^4 = -i
  a[14] = cmul(a[14], make_float2(-R2, -R2));   // na 2 kb 3: W^6
  a[7] = cmul(a[7], make_float2(S1, -C1));      // na 3 kb 1: W^3
  a[11] = cmul(a[11], make_float2(-R2, -R2));   // na 3 kb 2: W^6
  a[15] = cmul(a[15], make_float2(-C1, S1));    // na 3 kb 3: W^9
#pragma unroll
  for (int kb = 0; kb < 4; ++kb) dft4(a[4 * kb], a[4 * kb + 1], a[4 * kb + 2], a[4 * kb + 3]);  // over na -> ka
  // now a[4 kb + ka] = A[kb + 4 ka]: transpose the 4 x 4 index grid to natural order
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = i + 1; j < 4; ++j) { const float2 t = a[4 * i + j]; a[4 * i + j] = a[4 * j + i]; a[4 * j + i] = t; }
}

// One CTA per 32-frame tile, 4 warps.  A frame is transformed by a HALF warp (16 threads): the 512-point real
// FFT is a 256-point complex FFT of the even/odd packed, windowed frame, factored 16 x 16: each thread runs a
// 16-point FFT in registers, the 16 x 16 transpose goes once through shared memory, a second 16-point FFT in
// registers, and the real-FFT split pairs bins k and 256-k with two warp shuffles per bin.  Window and twiddle
// factors live in registers (fixed per thread), so per frame shared memory is touched for the samples, one
// transpose and the 257-bin power spectrum feeding the sparse mel sums.
static constexpr int EX_PITCH = 18;   // float2 per transpose row (144 B: 16-byte aligned, conflict-free 128-bit reads)
static constexpr int EX_FRAME = 16 * EX_PITCH;   // float2 per frame buffer

// 4 CTAs per SM (<= 128 registers, 49 KB of shared memory each): the kernel is latency / instruction-fetch bound
// (no_inst 28 %, wait 19 % of the stall samples at 3 CTAs per SM), so the fourth CTA is worth the few spilled values
__global__ void __launch_bounds__(FB_WARPS * 32, 4)
fbank_logmel_kernel(const float* __restrict__ wav, const UttDesc* __restrict__ utts,
                    const int2* __restrict__ tiles, const FbankTables* __restrict__ tab, int hop, int n_mels, int nnz,
                    float preemph, float log_eps, int halo, int raw_power, float* __restrict__ logmel,
                    float* __restrict__ tile_sums) {
  extern __shared__ __align__(16) uint8_t fb_smem[];
  float2* exch = reinterpret_cast<float2*>(fb_smem);                                   // [warps][2][EX_FRAME]
  float* sums = reinterpret_cast<float*>(exch + FB_WARPS * 2 * EX_FRAME);              // [warps*2][MAX_MELS]
  int* mstart = reinterpret_cast<int*>(sums + FB_WARPS * 2 * MAX_MELS);                // [MAX_MELS]
  int* mcount = mstart + MAX_MELS;
  int* moff = mcount + MAX_MELS;
  float* melw = reinterpret_cast<float*>(moff + MAX_MELS);                             // [nnz]
  float* samp = melw + ((nnz + 3) & ~3);                                               // staged samples

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n1 = lane & 15, fi = lane >> 4;
  const int2 tile = tiles[blockIdx.x];
  const UttDesc u = utts[tile.x];
  const int t0 = tile.y;
  const int nf = min(FRAMES_PER_TILE, u.T - t0);
  const int L = u.nsamp;
  const float* x = wav + u.wav_off;

  for (int i = threadIdx.x; i < MAX_MELS; i += blockDim.x) { mstart[i] = tab->mel_start[i]; mcount[i] = tab->mel_count[i]; moff[i] = tab->mel_off[i]; }
  for (int i = threadIdx.x; i < nnz; i += blockDim.x) melw[i] = tab->mel_w[i];
  // pre-emphasised, reflect-padded samples of this tile: padded index i <-> sample j = i0 + i, samp[i] = x[j] - c x[j-1].
  const int nsmp = hop * (nf - 1) + NFFT;
  const int i0 = hop * t0 - NFFT / 2;
  // Interior tiles (no reflection at either end of the utterance: all but the first and last two tiles of a 3 s crop)
  // stage their contiguous sample span [i0 - 1, i0 + nsmp) with ONE bulk copy of the TMA engine (cp.async.bulk, 22 KB)
  // into `samp` itself, then apply the pre-emphasis in place: every thread pulls its strided elements and their left
  // neighbours into registers, barrier, writes back.  The copy starts at the 16-byte aligned address below the span
  // (crop offsets are arbitrary), `sh` floats early, and is rounded up to 16 bytes: hence the + 8 floats of slack in the
  // shared-memory plan and the + 3 in the interior test (never reads past the utterance's last sample).
  const bool interior = i0 >= 1 && i0 + nsmp + 3 <= L - 1;
  __shared__ __align__(8) uint64_t stage_bar;
  if (interior) {
    const float* src = x + (i0 - 1);
    const int sh = static_cast<int>((reinterpret_cast<uintptr_t>(src) & 15) >> 2);
    const uint32_t bytes = static_cast<uint32_t>((sh + nsmp + 1) * 4 + 15) & ~15u;
    if (threadIdx.x == 0) {
      mbar_init(smem_u32(&stage_bar), 1);
      fence_barrier_init();
      mbar_arrive_expect_tx(smem_u32(&stage_bar), bytes);
      bulk_load_1d(smem_u32(samp), src - sh, bytes, smem_u32(&stage_bar));
    }
    __syncthreads();                                    // the barrier is initialised before anyone polls it
    mbar_wait(smem_u32(&stage_bar), 0);
    constexpr int PER = (160 * (FRAMES_PER_TILE - 1) + NFFT + FB_WARPS * 32 - 1) / (FB_WARPS * 32);   // hop <= 160
    if (hop <= 160) {
      float cur[PER], prv[PER];
#pragma unroll
      for (int k = 0; k < PER; ++k) {
        const int i = k * FB_WARPS * 32 + threadIdx.x;
        const int ii = i < nsmp ? i : nsmp - 1;
        cur[k] = samp[sh + 1 + ii];
        prv[k] = samp[sh + ii];
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < PER; ++k) {
        const int i = k * FB_WARPS * 32 + threadIdx.x;
        if (i < nsmp) samp[i] = cur[k] - preemph * prv[k];
      }
    } else {                                            // long hops: element-serial in-place pass by one warp
      __syncthreads();
      if (warp == 0) {
        for (int base = 0; base < nsmp; base += 32) {
          const int i = base + lane;
          float c = 0.f, pv = 0.f;
          if (i < nsmp) { c = samp[sh + 1 + i]; pv = samp[sh + i]; }
          __syncwarp();
          if (i < nsmp) samp[i] = c - preemph * pv;     // writes index i <= the lowest index still to be read (i + sh)
          __syncwarp();
        }
      }
    }
  }
  // edge tiles: reflect padding of the centred STFT (and of the pre-emphasis) sample by sample.
  // Eight independent samples per thread and iteration: all 16 loads are in flight before the first store
  // (the one-sample-per-iteration version spent 44 % of the kernel's stall samples waiting here).
  for (int ib = 0; ib < nsmp && !interior; ib += 8 * FB_WARPS * 32) {
    float cur[8], prv[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int i = ib + k * FB_WARPS * 32 + threadIdx.x;
      int j = i0 + i;
      if (j < 0) j = -j;
      if (j >= L) j = 2 * (L - 1) - j;
      j = min(max(j, 0), L - 1);                            // only reached by the (unused) tail of the last iteration
      cur[k] = __ldg(x + j);
      prv[k] = __ldg(x + (j > 0 ? j - 1 : 1));              // reflect pad of one sample on the left
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int i = ib + k * FB_WARPS * 32 + threadIdx.x;
      if (i < nsmp) samp[i] = cur[k] - preemph * prv[k];
    }
  }
  // per-thread constants: window taps of its 16 packed points, W256^{n1 k2}, W512^{k2' + 16 k1} (k2' = n1 as well)
  float2 win[16], tw2[16], tw5[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) {
    const int n = n1 + 16 * q;
    win[q] = make_float2(tab->window[2 * n], tab->window[2 * n + 1]);
    tw2[q] = tab->tw256[n1 * q];
    tw5[q] = tab->tw512[n1 + 16 * q];
  }
  __syncthreads();

  float2* ex = exch + (warp * 2 + fi) * EX_FRAME;
  float* pw = reinterpret_cast<float*>(ex);
  float acc[MAX_MELS / 16];
#pragma unroll
  for (int q = 0; q < MAX_MELS / 16; ++q) acc[q] = 0.f;
  const int src_lane = (lane & 16) | ((16 - n1) & 15);

#pragma unroll 1   // the body is ~2 k instructions: unrolled copies only add instruction-cache misses (no_inst was 24 % of the stalls)
  for (int fb = 0; fb < nf; fb += FB_WARPS * 2) {
    const int fr = fb + warp * 2 + fi;
    const bool live = fr < nf;
    const int f = live ? fr : nf - 1;
    const float2* s2 = reinterpret_cast<const float2*>(samp + hop * f);   // hop is even: 8-byte aligned
    float2 a[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) {
      const float2 v = s2[n1 + 16 * q];
      a[q] = make_float2(v.x * win[q].x, v.y * win[q].y);
    }
    fft16(a);                                               // over n2 -> a[k2] = Y[n1, k2]
#pragma unroll
    for (int k2 = 1; k2 < 16; ++k2) a[k2] = cmul(a[k2], tw2[k2]);
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) ex[k2 * EX_PITCH + n1] = a[k2];
    __syncwarp();
    {
      const float4* row = reinterpret_cast<const float4*>(ex + n1 * EX_PITCH);   // thread k2 = n1 reads Y[:, k2]
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float4 v = row[q];
        a[2 * q] = make_float2(v.x, v.y);
        a[2 * q + 1] = make_float2(v.z, v.w);
      }
    }
    fft16(a);                                               // over n1 -> a[k1] = Z[k2 + 16 k1]
    __syncwarp();
    // real-FFT split: X[k] = E + W512^k O with E, O from Z[k] and Z[256 - k]; power spectrum to shared memory
#pragma unroll
    for (int k1 = 0; k1 < 16; ++k1) {
      float pr = __shfl_sync(0xffffffffu, a[15 - k1].x, src_lane);
      float pi = __shfl_sync(0xffffffffu, a[15 - k1].y, src_lane);
      if (n1 == 0) { pr = a[(16 - k1) & 15].x; pi = a[(16 - k1) & 15].y; }
      const float2 e = make_float2(0.5f * (a[k1].x + pr), 0.5f * (a[k1].y - pi));
      const float2 o = make_float2(0.5f * (a[k1].y + pi), -0.5f * (a[k1].x - pr));
      const float2 wo = cmul(tw5[k1], o);
      const float re = e.x + wo.x, im = e.y + wo.y;
      pw[n1 + 16 * k1] = re * re + im * im;
    }
    if (n1 == 0) { const float dd = a[0].x - a[0].y; pw[256] = dd * dd; }
    __syncwarp();
    const long long row = u.row_base + halo + t0 + f;
#pragma unroll
    for (int q = 0; q < MAX_MELS / 16; ++q) {
      const int m = n1 + 16 * q;
      if (m < n_mels) {
        const int b0 = mstart[m], cnt = mcount[m];
        const float* wv = melw + moff[m];
        float sacc = 0.f;
        for (int i = 0; i < cnt; ++i) sacc = fmaf(pw[b0 + i], wv[i], sacc);
        const float lm = raw_power ? sacc : logf(sacc + log_eps);
        if (live) {
          logmel[row * n_mels + m] = lm;
          acc[q] += lm;
        }
      }
    }
    __syncwarp();
  }
  // deterministic per-tile column sums (fixed slot order)
#pragma unroll
  for (int q = 0; q < MAX_MELS / 16; ++q) sums[(warp * 2 + fi) * MAX_MELS + n1 + 16 * q] = acc[q];
  __syncthreads();
  for (int m = threadIdx.x; m < n_mels; m += blockDim.x) {
    float s = 0.f;
    for (int w = 0; w < FB_WARPS * 2; ++w) s += sums[w * MAX_MELS + m];
    tile_sums[static_cast<long long>(blockIdx.x) * n_mels + m] = s;
  }
}

__global__ void fbank_mean_kernel(const UttDesc* __restrict__ utts, const float* __restrict__ tile_sums, int n_mels,
                                  float* __restrict__ mean) {
  const UttDesc u = utts[blockIdx.x];
  for (int m = threadIdx.x; m < n_mels; m += blockDim.x) {
    float s = 0.f;
    for (int i = 0; i < u.ntiles; ++i) s += tile_sums[static_cast<long long>(u.tile_base + i) * n_mels + m];
    mean[blockIdx.x * n_mels + m] = s / static_cast<float>(u.T);
  }
}

// One warp per padded row: subtract the utterance mean, mirror into halo rows, split to hi|lo bf16
// and zero the channel padding (columns n_mels..plane_cols-1) so K can be tiled in 64-column chunks.
// lo_off = element distance hi -> lo of the same (row, column): plane_cols when the planes sit side by side in a
// row, (rows + FB_ROW_SLACK) * out_ld when they are stacked by rows (dense layout for the K-packed block 0, where
// out_ld == n_mels and the 5-tap window of a frame is ONE contiguous run of 5 * n_mels elements).  Rows
// [rows, rows + 8) are zeroed: the K padding of the packed window reads a few elements past the last row.
// (inorm_scale / inorm_bias != nullptr: input_norm=True, v = (x - mean) * scale[utt, m] + bias[m], see fbank_inorm_kernel)
__global__ void fbank_cmn_pack_kernel(const float* __restrict__ logmel, const float* __restrict__ mean,
                                      const float* __restrict__ inorm_scale, const float* __restrict__ inorm_bias,
                                      const RowMeta* __restrict__ meta, long long rows, int n_mels,
                                      bf16* __restrict__ out, int out_ld, int plane_cols, int planes, long long lo_off,
                                      long long total_rows) {
  const long long row = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= total_rows) return;
  RowMeta mt{-1, 0, 0, 0};
  if (row < rows) mt = meta[row];
  for (int m = lane; m < plane_cols; m += 32) {
    float v = 0.f;
    if (mt.utt >= 0 && m < n_mels) {
      int t = mt.t;
      if (t < 0) t = -t;
      if (t >= mt.T) t = 2 * (mt.T - 1) - t;
      const long long src = row + (t - mt.t);
      v = logmel[src * n_mels + m] - mean[mt.utt * n_mels + m];
      if (inorm_scale != nullptr) v = fmaf(v, inorm_scale[mt.utt * n_mels + m], inorm_bias[m]);
    }
    const bf16 h = __float2bfloat16_rn(v);
    out[row * out_ld + m] = h;
    if (planes == 2) out[row * out_ld + lo_off + m] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

// input_norm=True of ECAPA_TDNN (nn.InstanceNorm1d(n_mels, affine=True, eps, track_running_stats=False) applied to the
// mean-normalised log-mel, src/models/ECAPA_TDNN.py:406-408,477-478): per (utterance, mel) scale = weight / sqrt(var +
// eps) with the biased variance over the utterance's frames (the mean is the CMN mean).  One block per utterance; the
// fp32 log-mel rows it re-reads (96 KB per 3 s crop) are still L2 resident.
__global__ void fbank_inorm_kernel(const float* __restrict__ logmel, const float* __restrict__ mean,
                                   const UttDesc* __restrict__ utts, int halo, int n_mels, const float* __restrict__ weight,
                                   float eps, float* __restrict__ scale) {
  const UttDesc u = utts[blockIdx.x];
  for (int m = threadIdx.x; m < n_mels; m += blockDim.x) {
    const float mu = mean[blockIdx.x * n_mels + m];
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int t = 0;
    for (; t + 3 < u.T; t += 4) {
      const float* r = logmel + (u.row_base + halo + t) * n_mels + m;
      const float d0 = r[0] - mu, d1 = r[n_mels] - mu, d2 = r[2 * n_mels] - mu, d3 = r[3 * n_mels] - mu;
      s0 = fmaf(d0, d0, s0); s1 = fmaf(d1, d1, s1); s2 = fmaf(d2, d2, s2); s3 = fmaf(d3, d3, s3);
    }
    for (; t < u.T; ++t) { const float d = logmel[(u.row_base + halo + t) * n_mels + m] - mu; s0 = fmaf(d, d, s0); }
    const float var = ((s0 + s1) + (s2 + s3)) / static_cast<float>(u.T);
    scale[blockIdx.x * n_mels + m] = weight[m] / sqrtf(var + eps);
  }
}

// Encoder-plugin entry (features computed elsewhere): power-mel rows -> log(x + eps) rows.
__global__ void log_rows_kernel(const float* __restrict__ power, long long n, float log_eps, float* __restrict__ out) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = logf(power[i] + log_eps);
}
// per-utterance column mean over the valid rows of an fp32 [rows, n_mels] matrix
__global__ void utt_mean_f32_kernel(const float* __restrict__ x, const UttDesc* __restrict__ utts, int halo, int n_mels,
                                    float* __restrict__ mean) {
  const UttDesc u = utts[blockIdx.x];
  for (int m = threadIdx.x; m < n_mels; m += blockDim.x) {
    float s = 0.f;
    for (int t = 0; t < u.T; ++t) s += x[(u.row_base + halo + t) * n_mels + m];
    mean[blockIdx.x * n_mels + m] = s / static_cast<float>(u.T);
  }
}

// ------------------------------------------------------------------ host
int fbank_create(const float* window_host, int win_length, int hop, int n_fft, const float* fb_host, int n_freqs,
                 int n_mels, float preemph, float log_eps, Fbank** out) {
  SVB_CHECK(n_fft == NFFT && n_freqs == NBIN, "fbank kernel is specialised for n_fft=512 (got %d)", n_fft);
  SVB_CHECK(win_length <= NFFT && n_mels <= MAX_MELS && hop > 0 && hop % 2 == 0, "unsupported fbank config");
  std::vector<uint8_t> raw(sizeof(FbankTables), 0);
  FbankTables* t = reinterpret_cast<FbankTables*>(raw.data());
  const int lo = (NFFT - win_length) / 2;
  for (int i = 0; i < win_length; ++i) t->window[lo + i] = window_host[i];
  const double PI = 3.14159265358979323846;
  for (int k = 0; k < 256; ++k)
    t->tw256[k] = make_float2(static_cast<float>(cos(-2.0 * PI * k / 256.0)), static_cast<float>(sin(-2.0 * PI * k / 256.0)));
  for (int k = 0; k < NBIN; ++k)
    t->tw512[k] = make_float2(static_cast<float>(cos(-2.0 * PI * k / 512.0)), static_cast<float>(sin(-2.0 * PI * k / 512.0)));
  int nnz = 0, max_count = 0;
  for (int m = 0; m < n_mels; ++m) {
    int first = -1, last = -1;
    for (int k = 0; k < n_freqs; ++k)
      if (fb_host[k * n_mels + m] != 0.f) { if (first < 0) first = k; last = k; }
    t->mel_start[m] = first < 0 ? 0 : first;
    t->mel_count[m] = first < 0 ? 0 : last - first + 1;
    t->mel_off[m] = nnz;
    SVB_CHECK(nnz + t->mel_count[m] <= MAX_NNZ, "mel filterbank too dense for the sparse table");
    for (int i = 0; i < t->mel_count[m]; ++i) t->mel_w[nnz++] = fb_host[(first + i) * n_mels + m];
    if (t->mel_count[m] > max_count) max_count = t->mel_count[m];
  }
  Fbank* fbk = new Fbank();
  fbk->win_length = win_length; fbk->hop = hop; fbk->n_mels = n_mels; fbk->nnz = nnz; fbk->max_count = max_count;
  fbk->preemph = preemph; fbk->log_eps = log_eps;
  if (cudaMalloc(&fbk->d_tab, sizeof(FbankTables)) != cudaSuccess) { delete fbk; set_error("cudaMalloc failed"); return -2; }
  SVB_CUDA_OK(cudaMemcpy(fbk->d_tab, t, sizeof(FbankTables), cudaMemcpyHostToDevice));
  *out = fbk;
  return 0;
}

void fbank_destroy(Fbank* f) {
  if (f != nullptr) { cudaFree(f->d_tab); delete f; }
}

int fbank_n_mels(const Fbank* f) { return f->n_mels; }
int fbank_hop(const Fbank* f) { return f->hop; }

size_t fbank_smem_bytes(const Fbank* f) {
  return FB_WARPS * 2 * EX_FRAME * sizeof(float2) + FB_WARPS * 2 * MAX_MELS * sizeof(float) + 3 * MAX_MELS * sizeof(int) +
         static_cast<size_t>((f->nnz + 3) & ~3) * sizeof(float) +
         (static_cast<size_t>(f->hop) * (FRAMES_PER_TILE - 1) + NFFT + 8) * sizeof(float);   // + 8: bulk-copy alignment slack
}

struct InputNorm { const float* weight; const float* bias; float eps; float* scale_ws; };   // weight == nullptr: off

static int pack_rows(const Fbank* f, const RowMeta* meta, long long rows, const float* logmel_ws, const float* mean_ws,
                     bf16* out, int out_ld, int plane_cols, int planes, cudaStream_t stream, const void* utt_tab = nullptr,
                     int B = 0, int halo = 0, const InputNorm* in = nullptr) {
  const float *in_scale = nullptr, *in_bias = nullptr;
  if (in != nullptr && in->weight != nullptr) {
    SVB_CHECK(in->bias != nullptr && in->scale_ws != nullptr && utt_tab != nullptr, "input_norm needs weight, bias and scratch");
    fbank_inorm_kernel<<<B, 128, 0, stream>>>(logmel_ws, mean_ws, static_cast<const UttDesc*>(utt_tab), halo, f->n_mels,
                                             in->weight, in->eps, in->scale_ws);
    count_launch();
    SVB_CUDA_OK(cudaGetLastError());
    in_scale = in->scale_ws; in_bias = in->bias;
  }
  const int rows_per_block = 8;
  // planes stacked by rows when they do not fit side by side (see the kernel comment); the caller's buffer then holds
  // 2 * (rows + FB_ROW_SLACK) rows of out_ld elements
  const bool stacked = planes == 2 && out_ld < 2 * plane_cols;
  const long long lo_off = stacked ? (rows + FB_ROW_SLACK) * out_ld : plane_cols;
  const long long nrows = stacked ? rows + 8 : rows;
  fbank_cmn_pack_kernel<<<static_cast<int>((nrows + rows_per_block - 1) / rows_per_block), rows_per_block * 32, 0,
                          stream>>>(logmel_ws, mean_ws, in_scale, in_bias, meta, rows, f->n_mels, out, out_ld, plane_cols, planes,
                                    lo_off, nrows);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// power-mel rows [rows, n_mels] (time-major, valid rows filled) -> packed mean-normalised log-mel
int fbank_from_power(const Fbank* f, const float* power_rows, const void* utt_tab, int B, const RowMeta* meta,
                     long long rows, int halo, float* logmel_ws, float* mean_ws, bf16* out, int out_ld, int plane_cols,
                     int planes, cudaStream_t stream, const float* in_w, const float* in_b, float in_eps, float* in_scale_ws) {
  SVB_CHECK(plane_cols >= f->n_mels, "plane_cols must be >= n_mels");
  const long long n = rows * f->n_mels;
  log_rows_kernel<<<static_cast<int>(std::min<long long>((n + 255) / 256, 148 * 16)), 256, 0, stream>>>(
      power_rows, n, f->log_eps, logmel_ws);
  utt_mean_f32_kernel<<<B, 128, 0, stream>>>(logmel_ws, static_cast<const UttDesc*>(utt_tab), halo, f->n_mels, mean_ws);
  count_launch(2);
  SVB_CUDA_OK(cudaGetLastError());
  const InputNorm in{in_w, in_b, in_eps, in_scale_ws};
  return pack_rows(f, meta, rows, logmel_ws, mean_ws, out, out_ld, plane_cols, planes, stream, utt_tab, B, halo, &in);
}

// mode 0: full front-end (log, mean normalisation, bf16 pack); mode 1: power-mel rows only (logmel_ws receives them)
int fbank_forward(const Fbank* f, const float* wav, const void* utt_tab, const int32_t* tile_tab, int B, int ntiles,
                  const RowMeta* meta, long long rows, int halo, float* logmel_ws, float* tile_sums_ws,
                  float* mean_ws, bf16* out, int out_ld, int plane_cols, int planes, int mode, cudaStream_t stream,
                  const float* in_w, const float* in_b, float in_eps, float* in_scale_ws) {
  SVB_CHECK(mode == 1 || plane_cols >= f->n_mels, "plane_cols must be >= n_mels");
  const size_t smem = fbank_smem_bytes(f);
  static unsigned long long attr_devs = 0;   // per device: kernel attributes do not carry over
  if (first_on_device(&attr_devs)) {
    SVB_CUDA_OK(cudaFuncSetAttribute(fbank_logmel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(smem)));
    SVB_CUDA_OK(cudaFuncSetAttribute(fbank_logmel_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  }
  fbank_logmel_kernel<<<ntiles, FB_WARPS * 32, smem, stream>>>(
      wav, static_cast<const UttDesc*>(utt_tab), reinterpret_cast<const int2*>(tile_tab), f->d_tab, f->hop,
      f->n_mels, f->nnz, f->preemph, f->log_eps, halo, mode, logmel_ws, tile_sums_ws);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  if (mode == 1) return 0;
  fbank_mean_kernel<<<B, 128, 0, stream>>>(static_cast<const UttDesc*>(utt_tab), tile_sums_ws, f->n_mels, mean_ws);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  const InputNorm in{in_w, in_b, in_eps, in_scale_ws};
  return pack_rows(f, meta, rows, logmel_ws, mean_ws, out, out_ld, plane_cols, planes, stream, utt_tab, B, halo, &in);
}

}  // namespace svb
