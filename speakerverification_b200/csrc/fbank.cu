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

__global__ void __launch_bounds__(FB_WARPS * 32)
fbank_logmel_kernel(const float* __restrict__ wav, const UttDesc* __restrict__ utts,
                    const int2* __restrict__ tiles, const FbankTables* __restrict__ tab, int hop, int n_mels,
                    float preemph, float log_eps, int halo, int raw_power, float* __restrict__ logmel,
                    float* __restrict__ tile_sums) {
  extern __shared__ __align__(16) uint8_t fb_smem[];
  FbankTables* st = reinterpret_cast<FbankTables*>(fb_smem);
  float2* fftbuf = reinterpret_cast<float2*>(fb_smem + sizeof(FbankTables));      // [warps][2][256]
  float* sums = reinterpret_cast<float*>(fftbuf + FB_WARPS * 512);                // [warps][MAX_MELS]
  float* samp = sums + FB_WARPS * MAX_MELS;                                       // staged samples

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int2 tile = tiles[blockIdx.x];
  const UttDesc u = utts[tile.x];
  const int t0 = tile.y;
  const int nf = min(FRAMES_PER_TILE, u.T - t0);
  const int L = u.nsamp;
  const float* x = wav + u.wav_off;

  // tables -> smem (coalesced 16B copies)
  {
    const uint4* src = reinterpret_cast<const uint4*>(tab);
    uint4* dst = reinterpret_cast<uint4*>(st);
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(FbankTables) / 16); i += blockDim.x) dst[i] = __ldg(src + i);
  }
  // pre-emphasised, reflect-padded samples of this tile: padded index i <-> sample j = i - 256
  const int nsmp = hop * (nf - 1) + NFFT;
  const long long i0 = static_cast<long long>(hop) * t0;
  for (int i = threadIdx.x; i < nsmp; i += blockDim.x) {
    long long j = i0 + i - NFFT / 2;
    if (j < 0) j = -j;
    if (j >= L) j = 2LL * (L - 1) - j;
    const float cur = __ldg(x + j);
    const float prev = __ldg(x + (j > 0 ? j - 1 : 1));   // reflect pad of one sample on the left
    samp[i] = cur - preemph * prev;
  }
  __syncthreads();

  float2* A = fftbuf + warp * 512;
  float2* Bf = A + 256;
  float acc[MAX_MELS / 32];
#pragma unroll
  for (int q = 0; q < MAX_MELS / 32; ++q) acc[q] = 0.f;

  for (int f = warp; f < nf; f += FB_WARPS) {
    const float* s = samp + hop * f;
    // pack even/odd samples into a 256-point complex sequence
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int n = lane + 32 * j;
      A[n] = make_float2(st->window[2 * n] * s[2 * n], st->window[2 * n + 1] * s[2 * n + 1]);
    }
    __syncwarp();
    // Stockham radix-4, 4 passes (Ns = 1, 4, 16, 64)
    float2* src = A;
    float2* dst = Bf;
#pragma unroll
    for (int pass = 0; pass < 4; ++pass) {
      const int Ns = 1 << (2 * pass);
      const int tstep = 64 >> (2 * pass);   // 256 / (4 * Ns)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int j = lane + 32 * h;
        const int jm = j & (Ns - 1);
        float2 v0 = src[j], v1 = src[j + 64], v2 = src[j + 128], v3 = src[j + 192];
        if (pass > 0) {
          v1 = cmul(v1, st->tw256[jm * tstep]);
          v2 = cmul(v2, st->tw256[2 * jm * tstep]);
          v3 = cmul(v3, st->tw256[3 * jm * tstep]);
        }
        const float2 a0 = make_float2(v0.x + v2.x, v0.y + v2.y);
        const float2 a1 = make_float2(v0.x - v2.x, v0.y - v2.y);
        const float2 a2 = make_float2(v1.x + v3.x, v1.y + v3.y);
        const float2 a3 = make_float2(v1.y - v3.y, v3.x - v1.x);   // -i * (v1 - v3)
        const int j0 = ((j - jm) << 2) + jm;
        dst[j0] = make_float2(a0.x + a2.x, a0.y + a2.y);
        dst[j0 + Ns] = make_float2(a1.x + a3.x, a1.y + a3.y);
        dst[j0 + 2 * Ns] = make_float2(a0.x - a2.x, a0.y - a2.y);
        dst[j0 + 3 * Ns] = make_float2(a1.x - a3.x, a1.y - a3.y);
      }
      __syncwarp();
      float2* tmp = src; src = dst; dst = tmp;
    }
    // src == A holds Z = FFT256(z).  Real-FFT split and power spectrum -> pw (aliases Bf).
    float* pw = reinterpret_cast<float*>(Bf);
    for (int k = lane; k < NBIN; k += 32) {
      const float2 a = A[k & 255];
      const float2 bq = A[(256 - k) & 255];
      const float2 e = make_float2(0.5f * (a.x + bq.x), 0.5f * (a.y - bq.y));
      const float2 o = make_float2(0.5f * (a.y + bq.y), -0.5f * (a.x - bq.x));
      const float2 wo = cmul(st->tw512[k], o);
      const float re = e.x + wo.x, im = e.y + wo.y;
      pw[k] = re * re + im * im;
    }
    __syncwarp();
    const long long row = u.row_base + halo + t0 + f;
#pragma unroll
    for (int q = 0; q < MAX_MELS / 32; ++q) {
      const int m = lane + 32 * q;
      if (m < n_mels) {
        const int b0 = st->mel_start[m], cnt = st->mel_count[m];
        const float* wv = st->mel_w + st->mel_off[m];
        float sacc = 0.f;
        for (int i = 0; i < cnt; ++i) sacc = fmaf(pw[b0 + i], wv[i], sacc);
        const float lm = raw_power ? sacc : logf(sacc + log_eps);
        logmel[row * n_mels + m] = lm;
        acc[q] += lm;
      }
    }
    __syncwarp();
  }
  // deterministic per-tile column sums (fixed warp order)
#pragma unroll
  for (int q = 0; q < MAX_MELS / 32; ++q) sums[warp * MAX_MELS + lane + 32 * q] = acc[q];
  __syncthreads();
  for (int m = threadIdx.x; m < n_mels; m += blockDim.x) {
    float s = 0.f;
    for (int w = 0; w < FB_WARPS; ++w) s += sums[w * MAX_MELS + m];
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
__global__ void fbank_cmn_pack_kernel(const float* __restrict__ logmel, const float* __restrict__ mean,
                                      const RowMeta* __restrict__ meta, long long rows, int n_mels,
                                      bf16* __restrict__ out, int out_ld, int plane_cols, int planes) {
  const long long row = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const RowMeta mt = meta[row];
  for (int m = lane; m < plane_cols; m += 32) {
    float v = 0.f;
    if (mt.utt >= 0 && m < n_mels) {
      int t = mt.t;
      if (t < 0) t = -t;
      if (t >= mt.T) t = 2 * (mt.T - 1) - t;
      const long long src = row + (t - mt.t);
      v = logmel[src * n_mels + m] - mean[mt.utt * n_mels + m];
    }
    const bf16 h = __float2bfloat16_rn(v);
    out[row * out_ld + m] = h;
    if (planes == 2) out[row * out_ld + plane_cols + m] = __float2bfloat16_rn(v - __bfloat162float(h));
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
  return sizeof(FbankTables) + FB_WARPS * 512 * sizeof(float2) + FB_WARPS * MAX_MELS * sizeof(float) +
         (static_cast<size_t>(f->hop) * (FRAMES_PER_TILE - 1) + NFFT) * sizeof(float);
}

static int pack_rows(const Fbank* f, const RowMeta* meta, long long rows, const float* logmel_ws, const float* mean_ws,
                     bf16* out, int out_ld, int plane_cols, int planes, cudaStream_t stream) {
  const int rows_per_block = 8;
  fbank_cmn_pack_kernel<<<static_cast<int>((rows + rows_per_block - 1) / rows_per_block), rows_per_block * 32, 0,
                          stream>>>(logmel_ws, mean_ws, meta, rows, f->n_mels, out, out_ld, plane_cols, planes);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// power-mel rows [rows, n_mels] (time-major, valid rows filled) -> packed mean-normalised log-mel
int fbank_from_power(const Fbank* f, const float* power_rows, const void* utt_tab, int B, const RowMeta* meta,
                     long long rows, int halo, float* logmel_ws, float* mean_ws, bf16* out, int out_ld, int plane_cols,
                     int planes, cudaStream_t stream) {
  SVB_CHECK(plane_cols >= f->n_mels, "plane_cols must be >= n_mels");
  const long long n = rows * f->n_mels;
  log_rows_kernel<<<static_cast<int>(std::min<long long>((n + 255) / 256, 148 * 16)), 256, 0, stream>>>(
      power_rows, n, f->log_eps, logmel_ws);
  utt_mean_f32_kernel<<<B, 128, 0, stream>>>(logmel_ws, static_cast<const UttDesc*>(utt_tab), halo, f->n_mels, mean_ws);
  count_launch(2);
  SVB_CUDA_OK(cudaGetLastError());
  return pack_rows(f, meta, rows, logmel_ws, mean_ws, out, out_ld, plane_cols, planes, stream);
}

// mode 0: full front-end (log, mean normalisation, bf16 pack); mode 1: power-mel rows only (logmel_ws receives them)
int fbank_forward(const Fbank* f, const float* wav, const void* utt_tab, const int32_t* tile_tab, int B, int ntiles,
                  const RowMeta* meta, long long rows, int halo, float* logmel_ws, float* tile_sums_ws,
                  float* mean_ws, bf16* out, int out_ld, int plane_cols, int planes, int mode, cudaStream_t stream) {
  SVB_CHECK(mode == 1 || plane_cols >= f->n_mels, "plane_cols must be >= n_mels");
  const size_t smem = fbank_smem_bytes(f);
  static bool attr_set = false;
  if (!attr_set) {
    SVB_CUDA_OK(cudaFuncSetAttribute(fbank_logmel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(smem)));
    attr_set = true;
  }
  fbank_logmel_kernel<<<ntiles, FB_WARPS * 32, smem, stream>>>(
      wav, static_cast<const UttDesc*>(utt_tab), reinterpret_cast<const int2*>(tile_tab), f->d_tab, f->hop,
      f->n_mels, f->preemph, f->log_eps, halo, mode, logmel_ws, tile_sums_ws);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  if (mode == 1) return 0;
  fbank_mean_kernel<<<B, 128, 0, stream>>>(static_cast<const UttDesc*>(utt_tab), tile_sums_ws, f->n_mels, mean_ws);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return pack_rows(f, meta, rows, logmel_ws, mean_ws, out, out_ld, plane_cols, planes, stream);
}

}  // namespace svb
