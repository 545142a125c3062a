/* libsvb C ABI -- B200 (sm_100a) speaker-embedding and trial-scoring hot path.
 *
 * Drop-in boundary for the inference hot path of hiimmuc/SpeakerVerification (a pure-Python
 * reference with no FFI of its own): each entry point replaces the library kernels PyTorch eager
 * launches behind one reference call site, cited per function (paths relative to the reference
 * tree).  Plain pointers and sizes only; no torch types cross this boundary.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless its name ends in _host;
 *  - `stream` is a cudaStream_t passed as void*; all work is enqueued asynchronously on it;
 *  - return value 0 = ok, negative = error (message from svb_last_error(), thread local);
 *  - there is no CPU fallback: a missing CUDA device or a non-sm_100 GPU is an error.
 *
 * Row layout ("time-major with mirrored halos").  An activation matrix holds, for every
 * utterance b, H halo rows, T_b valid rows and H halo rows, back to back; row r of utterance b
 * with frame index t lives at row_base[b] + H + t.  Halo rows mirror frames 1..H and T-2..T-1-H
 * (reflect padding, src/models/layers/cnn.py:104-157), so dilated taps are plain row-shifted reads.
 */
#ifndef SVB_H_
#define SVB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVB_VERSION 100

int svb_version(void);
const char* svb_last_error(void);
/* kernels launched by this library since the last reset (bench.py's gpu_launches) */
long long svb_launch_count(void);
void svb_reset_launch_count(void);

/* ------------------------------------------------------------------ row bookkeeping */
/* Fills row_meta[total_rows] = {utt, t, T, 0} (int32 x4) from per-utterance frame counts.
 * frames_host[B]; rows of utterance b start at sum_{i<b}(frames[i] + 2*halo). */
int svb_build_row_meta(const int32_t* frames_host, int B, int halo, int32_t* row_meta,
                       int64_t total_rows, void* stream);

/* ------------------------------------------------------------------ operand preparation */
/* fp32 [rows, cols] (ld_in) -> bf16 planes: hi at out[r*ld_out + c], optional lo at +plane_cols. */
int svb_split_bf16(const float* x, int64_t rows, int cols, int ld_in, void* out, int ld_out,
                   int planes, int plane_cols, void* stream);
/* Conv1d weight [cout, cin, k] fp32 (nn.Conv1d layout, src/models/layers/cnn.py:79-88) ->
 * packed bf16 [cout, k, planes, cin_pad] (cin_pad multiple of 64, zero filled; planes: hi[, lo]). */
int svb_pack_conv_weight(const float* w, int cout, int cin, int k, int cin_pad, int planes,
                         void* out, void* stream);
/* BatchNorm1d(eval) -> per-channel scale/shift (src/models/layers/normalization.py:62-68, eps 1e-5). */
int svb_fold_bn(const float* gamma, const float* beta, const float* mean, const float* var, int n,
                float eps, float* scale, float* shift, void* stream);

/* ------------------------------------------------------------------ TDNN layer (tensor cores)
 * One TDNNBlock / Conv1d of ECAPA-TDNN as an implicit GEMM on tcgen05 tensor cores:
 *   out = BN(act(conv_reflect_same(x) + bias))          src/models/ECAPA_TDNN.py:68-69
 * replaces F.pad(reflect) + nn.Conv1d + activation + BatchNorm1d (4 library kernels). */
enum { SVB_ACT_NONE = 0, SVB_ACT_RELU = 1, SVB_ACT_GELU = 2 };
enum { SVB_OUT_BF16 = 0, SVB_OUT_BF16_HILO = 1, SVB_OUT_F32 = 2 };
enum { SVB_STORE_ALL = 0, SVB_STORE_MIRROR = 1 };

typedef struct SvbTdnnArgs {
  int64_t rows;            /* rows of x / out (valid + halo rows) */
  const void* x;           /* bf16 [rows, x_ld]; hi plane at column 0, lo plane at +x_plane_cols */
  int32_t x_ld, x_cols, x_plane_cols;
  const void* x2;          /* optional: 64-col chunks [0, x2_chunks) of K come from x2 instead of x */
  int32_t x2_ld, x2_cols, x2_chunks;
  const void* w;           /* packed by svb_pack_conv_weight */
  int32_t cout, cin_pad, k, dilation, w_planes;
  int32_t passes;          /* 1 = bf16 ; 3 = split-bf16 (fp32-grade: hi*hi + hi*lo + lo*hi) */
  const float* bias;       /* [cout] */
  const float* bn_scale;   /* [cout] */
  const float* bn_shift;   /* [cout] */
  const float* row_bias;   /* optional [utterances, row_bias_ld], added before the activation */
  int32_t row_bias_ld;
  int32_t act, post_tanh;
  const int32_t* row_meta; /* int32 x4 per row, required for k > 1 or row_bias */
  int32_t store_mode, halo;
  void* out; int32_t out_ld, out_kind, out_plane_cols;
  const void* add_src;     /* optional second output out2 = out_value + add_src (bf16 planes) */
  int32_t add_ld, add_planes, add_plane_cols;
  void* out2; int32_t out2_ld, out2_plane_cols;
  int32_t block_n;         /* 0 = auto */
} SvbTdnnArgs;

int svb_tdnn_layer(const SvbTdnnArgs* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SVB_H_ */
