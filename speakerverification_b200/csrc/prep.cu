// Operand preparation kernels: fp32 -> bf16 hi/lo planes, Conv1d weight packing, BN folding,
// and the row bookkeeping table of the time-major-with-halos layout.
#include <mutex>
#include <cuda_fp16.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "svb_internal.h"
#include "svb_ptx.cuh"

namespace svb {

__global__ void split_bf16_kernel(const float* __restrict__ x, long long rows, int cols, int ld_in,
                                  bf16* __restrict__ out, int ld_out, int planes, int plane_cols) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const int c = static_cast<int>(i - r * cols);
    const float v = x[r * ld_in + c];
    const bf16 h = __float2bfloat16_rn(v);
    out[r * ld_out + c] = h;
    if (planes == 2) out[r * ld_out + plane_cols + c] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

__global__ void to_f16_kernel(const float* __restrict__ x, long long rows, int cols, int ld_in, __half* __restrict__ out,
                              int ld_out) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const int c = static_cast<int>(i - r * cols);
    const uint32_t pk = pack_f16x2(x[r * ld_in + c], 0.f);
    out[r * ld_out + c] = __ushort_as_half(static_cast<unsigned short>(pk & 0xffffu));
  }
}

int to_f16(const float* x, long long rows, int cols, int ld_in, void* out, int ld_out, cudaStream_t stream) {
  const long long total = rows * cols;
  if (total == 0) return 0;
  const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16));
  to_f16_kernel<<<grid, 256, 0, stream>>>(x, rows, cols, ld_in, static_cast<__half*>(out), ld_out);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// w [cout, cin, k] -> out [cout, k, planes, cin_pad].  F16: single-plane IEEE half instead of bf16 (the B operand of a
// kind::f16 UMMA may be .f16 while A stays .bf16: three more mantissa bits for the weights at the same MMA rate).
template <bool F16>
__global__ void pack_conv_weight_kernel(const float* __restrict__ w, int cout, int cin, int k, int cin_pad,
                                        int planes, bf16* __restrict__ out) {
  const long long total = (long long)cout * k * cin_pad;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = static_cast<int>(i % cin_pad);
    const int j = static_cast<int>((i / cin_pad) % k);
    const int o = static_cast<int>(i / ((long long)cin_pad * k));
    const float v = (c < cin) ? w[((long long)o * cin + c) * k + j] : 0.f;
    bf16* dst = out + ((long long)(o * k + j) * planes) * cin_pad + c;
    if (F16) {
      *reinterpret_cast<__half*>(dst) = __float2half_rn(v);
      continue;
    }
    const bf16 h = __float2bfloat16_rn(v);
    dst[0] = h;
    if (planes == 2) dst[cin_pad] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

// max |x| over n floats (engine creation: does a weight tensor fit the fp16 range?)
__global__ void absmax_kernel(const float* __restrict__ x, long long n, unsigned* __restrict__ out) {
  float m = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float a = fabsf(x[i]);
    m = (a > m || a != a) ? a : m;         // NaN propagates
  }
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) atomicMax(out, __float_as_uint(m));   // non-negative floats order like their bit patterns
}

int absmax_host(const float* x, long long n, float* result_host) {
  unsigned* d = nullptr;
  SVB_CUDA_OK(cudaMalloc(reinterpret_cast<void**>(&d), sizeof(unsigned)));
  SVB_CUDA_OK(cudaMemset(d, 0, sizeof(unsigned)));
  const int grid = static_cast<int>(std::min<long long>((n + 255) / 256, 148 * 8));
  absmax_kernel<<<grid < 1 ? 1 : grid, 256>>>(x, n, d);
  unsigned h = 0;
  const cudaError_t e = cudaMemcpy(&h, d, sizeof(unsigned), cudaMemcpyDeviceToHost);
  cudaFree(d);
  SVB_CUDA_OK(e);
  memcpy(result_host, &h, sizeof(float));
  return 0;
}

// w [cout, cin, k] -> out [cout, planes, kpad] with K index j * cin + c (tap-major, channels contiguous): the layout of a
// "K-packed" conv whose A operand row is the contiguous k * cin window of a dense [rows, cin] activation matrix.
__global__ void pack_conv_weight_kpacked_kernel(const float* __restrict__ w, int cout, int cin, int k, int kpad, int planes,
                                                bf16* __restrict__ out) {
  const long long total = (long long)cout * kpad;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int kk = static_cast<int>(i % kpad);
    const int o = static_cast<int>(i / kpad);
    const int j = kk / cin, c = kk - j * cin;
    const float v = (kk < cin * k) ? w[((long long)o * cin + c) * k + j] : 0.f;
    const bf16 h = __float2bfloat16_rn(v);
    bf16* dst = out + ((long long)o * planes) * kpad + kk;
    dst[0] = h;
    if (planes == 2) dst[kpad] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

__global__ void fold_bn_kernel(const float* g, const float* b, const float* m, const float* v, int n, float eps,
                               float* scale, float* shift) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const float s = g[i] / sqrtf(v[i] + eps);
    scale[i] = s;
    shift[i] = b[i] - m[i] * s;
  }
}

// Crop/stack stage of the evaluation loader (loadWAV eval mode, src/processing/audio_loader.py:110-150):
// utterance u's PCM (int16 or fp32, src_off/src_len) is converted to fp32 (int16 / 32768 = soundfile's
// float32 scaling) and laid down at dst_off with np.pad(..., 'wrap') semantics for dst_len > src_len.
// The crops themselves are never materialised: a crop is a row offset into this staged buffer (svb_batch_create).
template <typename T>
__global__ void pcm_stage_kernel(const T* __restrict__ pcm, const PcmSeg* __restrict__ segs, float* __restrict__ out) {
  const PcmSeg s = segs[blockIdx.y];
  const T* src = pcm + s.src_off;
  float* dst = out + s.dst_off;
  const float scale = (sizeof(T) == 2) ? (1.0f / 32768.0f) : 1.0f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < s.dst_len; i += gridDim.x * blockDim.x) {
    const int j = (i < s.src_len) ? i : (i % s.src_len);
    dst[i] = static_cast<float>(src[j]) * scale;
  }
}

int pcm_stage(const void* pcm, int sample_kind, const PcmSeg* segs, int U, int max_len, float* out,
              cudaStream_t stream) {
  SVB_CHECK(sample_kind == 0 || sample_kind == 1, "sample_kind must be 0 (fp32) or 1 (int16)");
  if (U == 0) return 0;
  SVB_CHECK(U <= 65535, "at most 65535 utterances per staging call");
  int gx = (max_len + 256 * 8 - 1) / (256 * 8);
  gx = gx < 1 ? 1 : (gx > 64 ? 64 : gx);
  const dim3 grid(gx, U);
  if (sample_kind == 1)
    pcm_stage_kernel<int16_t><<<grid, 256, 0, stream>>>(static_cast<const int16_t*>(pcm), segs, out);
  else
    pcm_stage_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(pcm), segs, out);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

// row_meta from per-utterance frame counts without any stream synchronisation: the (tiny) prefix table is written to a
// slot of a library-owned ring of page-locked buffers and uploaded / expanded on the stream; a slot is reused only
// after the event recorded behind its last kernel has completed (4 calls later: in practice never a wait).
__global__ void row_meta_kernel(const long long* __restrict__ row0, int B, int halo, RowMeta* __restrict__ meta,
                                long long total_rows) {
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < total_rows;
       r += (long long)gridDim.x * blockDim.x) {
    int lo = 0, hi = B;                    // largest b with row0[b] <= r  (row0[B] = rows in use)
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (row0[mid] <= r) lo = mid; else hi = mid;
    }
    if (r >= row0[B]) { meta[r] = RowMeta{-1, 0, 0, 0}; continue; }
    const int T = static_cast<int>(row0[lo + 1] - row0[lo]) - 2 * halo;
    meta[r] = RowMeta{lo, static_cast<int>(r - row0[lo]) - halo, T, 0};
  }
}

namespace {
struct MetaSlot { long long* host = nullptr; long long* dev = nullptr; size_t cap = 0; cudaEvent_t done = nullptr; bool used = false; };
constexpr int META_DEVS = 16;
MetaSlot g_meta_ring[META_DEVS][4];        // per device: the staging buffers and the event live on one GPU
int g_meta_pos[META_DEVS] = {};
std::mutex g_meta_mutex;                 // the ring is shared by every host thread of the process
}  // namespace

int build_row_meta(const int32_t* frames_host, int B, int halo, int32_t* row_meta, long long total_rows,
                   cudaStream_t stream) {
  SVB_CHECK(B > 0 && halo >= 0, "bad row_meta arguments");
  std::lock_guard<std::mutex> lock(g_meta_mutex);
  int dev = 0;
  SVB_CUDA_OK(cudaGetDevice(&dev));
  SVB_CHECK(dev >= 0 && dev < META_DEVS, "device ordinal %d out of range", dev);
  MetaSlot& sl = g_meta_ring[dev][g_meta_pos[dev]];
  g_meta_pos[dev] = (g_meta_pos[dev] + 1) & 3;
  if (sl.used) SVB_CUDA_OK(cudaEventSynchronize(sl.done));     // its upload of 4 calls ago has executed
  const size_t need = static_cast<size_t>(B) + 1;
  if (sl.cap < need) {
    if (sl.host) cudaFreeHost(sl.host);
    if (sl.dev) cudaFree(sl.dev);
    sl.cap = need < 1024 ? 1024 : 2 * need;
    SVB_CUDA_OK(cudaMallocHost(reinterpret_cast<void**>(&sl.host), sl.cap * sizeof(long long)));
    SVB_CUDA_OK(cudaMalloc(reinterpret_cast<void**>(&sl.dev), sl.cap * sizeof(long long)));
  }
  if (sl.done == nullptr) SVB_CUDA_OK(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
  long long r = 0;
  for (int b = 0; b < B; ++b) {
    const int T = frames_host[b];
    SVB_CHECK(T > 2 * halo, "utterance %d has %d frames; need more than 2*halo = %d", b, T, 2 * halo);
    sl.host[b] = r;
    r += T + 2 * halo;
  }
  sl.host[B] = r;
  SVB_CHECK(r <= total_rows, "row_meta buffer too small");
  SVB_CUDA_OK(cudaMemcpyAsync(sl.dev, sl.host, need * sizeof(long long), cudaMemcpyHostToDevice, stream));
  const int grid = static_cast<int>(std::min<long long>((total_rows + 255) / 256, 148 * 8));
  row_meta_kernel<<<grid < 1 ? 1 : grid, 256, 0, stream>>>(sl.dev, B, halo, reinterpret_cast<RowMeta*>(row_meta), total_rows);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  SVB_CUDA_OK(cudaEventRecord(sl.done, stream));
  sl.used = true;
  return 0;
}

int split_bf16(const float* x, long long rows, int cols, int ld_in, bf16* out, int ld_out, int planes,
               int plane_cols, cudaStream_t stream) {
  const long long total = rows * cols;
  if (total == 0) return 0;
  const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16));
  split_bf16_kernel<<<grid, 256, 0, stream>>>(x, rows, cols, ld_in, out, ld_out, planes, plane_cols);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

int pack_conv_weight(const float* w, int cout, int cin, int k, int cin_pad, int planes, bf16* out,
                     cudaStream_t stream, int f16) {
  SVB_CHECK(cin_pad % 64 == 0 && cin_pad >= cin, "cin_pad must be a multiple of 64 and >= cin");
  SVB_CHECK(!f16 || planes == 1, "fp16 weights are single plane");
  const long long total = (long long)cout * k * cin_pad;
  const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16));
  if (f16) pack_conv_weight_kernel<true><<<grid, 256, 0, stream>>>(w, cout, cin, k, cin_pad, planes, out);
  else pack_conv_weight_kernel<false><<<grid, 256, 0, stream>>>(w, cout, cin, k, cin_pad, planes, out);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

int pack_conv_weight_kpacked(const float* w, int cout, int cin, int k, int kpad, int planes, bf16* out, cudaStream_t stream) {
  SVB_CHECK(kpad % 64 == 0 && kpad >= cin * k, "kpad must be a multiple of 64 and >= cin * k");
  const long long total = (long long)cout * kpad;
  const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16));
  pack_conv_weight_kpacked_kernel<<<grid, 256, 0, stream>>>(w, cout, cin, k, kpad, planes, out);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

int fold_bn(const float* g, const float* b, const float* m, const float* v, int n, float eps, float* scale,
            float* shift, cudaStream_t stream) {
  fold_bn_kernel<<<(n + 255) / 256, 256, 0, stream>>>(g, b, m, v, n, eps, scale, shift);
  count_launch();
  SVB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace svb
