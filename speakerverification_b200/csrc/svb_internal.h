// Internal (non-ABI) declarations shared by the .cu translation units of libsvb.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>

namespace svb {

typedef __nv_bfloat16 bf16;

// ------------------------------------------------------------------ errors
void set_error(const char* fmt, ...);
#define SVB_CUDA_OK(expr)                                                                   \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      svb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return -2;                                                                            \
    }                                                                                       \
  } while (0)
#define SVB_CHECK(cond, ...)        \
  do {                              \
    if (!(cond)) {                  \
      svb::set_error(__VA_ARGS__);  \
      return -1;                    \
    }                               \
  } while (0)

int num_sms();
bool first_on_device(unsigned long long* seen);   // true once per device ordinal (per-device kernel attributes)
int count_launch(int n = 1);  // running count of kernel launches (reported by bench.py)

// ------------------------------------------------------------------ rows
// Time-major activation matrices hold, per utterance, `halo` mirrored rows, T valid rows and
// `halo` mirrored rows again.  row_meta[r] = {utterance, t (may be <0 or >=T on halo rows), T, 0};
// utterance == -1 on rows that belong to no utterance.
struct RowMeta { int utt, t, T, pad; };
// one utterance of the crop/stack stage (svb_pcm_stage): same layout as SvbPcmSeg in include/svb.h
struct PcmSeg { long long src_off, dst_off; int src_len, dst_len; };

// ------------------------------------------------------------------ tensor-core GEMM
enum { ACT_NONE = 0, ACT_RELU = 1, ACT_GELU = 2 };
// Activation storage: two bf16 planes hi|lo (fp32-grade mode), ONE plane of IEEE half (16-bit mode: OUT_F16; fp16
// operands run at the bf16 tensor-core rate with 3 more mantissa bits -- the score deviation of the 16-bit mode drops
// from ~2e-3 to ~2e-4), fp32, or one plane of bf16 (generic gemm_tc path, RawNet2).
enum { OUT_BF16 = 0, OUT_BF16_HILO = 1, OUT_F32 = 2, OUT_F16 = 3 };
enum { STORE_ALL = 0, STORE_MIRROR = 1 };

struct GemmOperands {
  // A: activations [M, lda] bf16 (K-major); optional second source for k-chunks < a2_kc.
  const bf16* a; int lda; int a_cols;          // a_cols: columns addressable through this map
  const bf16* a2; int lda2; int a2_cols; int a2_kc;
  int a_plane_cols;                            // column distance hi-plane -> lo-plane in A (and A2)
  const bf16* w; int w_rows; int w_cols;       // packed weights [N, taps*planes*kc*64]
  int w_f16;                                   // weights are IEEE half (single plane) instead of bf16
  int a_f16;                                   // A (and A2, add_src) are IEEE half (single plane) instead of bf16
};

struct GemmParams {
  int M, N;
  int kc_per_tap;        // 64-column chunks per tap per plane
  int num_taps;
  int tap_off[5];        // row offset of each tap
  int num_pass;          // 1: hi*hi ; 3: hi*hi + hi*lo + lo*hi (fp32-grade split-bf16)
  int a_plane_cols;
  int a2_kc;
  int w_planes;
  const float* bias;     // [N_pad] conv bias
  const float* scale;    // [N_pad] BN scale (applied after the activation)
  const float* shift;    // [N_pad] BN shift
  const float* rowbias;  // optional [utterances, rowbias_ld] added before the activation
  int rowbias_ld;
  int act, post_tanh;
  const RowMeta* row_meta;
  int store_mode, halo;
  void* out; int out_ld; int out_kind; int out_plane_cols;
  const bf16* add_src; int add_ld; int add_planes; int add_plane_cols;
  void* out2; int out2_ld; int out2_plane_cols;
  // optional fused per-utterance column statistics of the OUTPUT (pointwise GELU layers, bf16 mode): fixed-point
  // (2^-32) int64 sums [utterances, stats_ld] of the valid rows' bf16 outputs (and of their squares), accumulated with
  // integer atomics -- order independent, so results stay bit-reproducible.  Buffers must be zeroed by the caller.
  long long* colsum; long long* colsq; int stats_ld;
  // A planes stacked by rows (K-packed block 0): the lo plane of row r is row r + a_plane_rows of the same map
  int a_plane_rows;
  int w_f16, a_f16;      // filled by launch_gemm from GemmOperands (b_format / a_format of the UMMA)
};

// RawNet2 GEMMs (rawnet_gemm.cu): fixed-length utterances, rows [utt][1 zero | T | 1 zero]
struct RnGemmParams {
  int M, N;
  int kc_per_tap, num_taps, tap_off[3], num_pass, a_plane_cols, w_planes;
  int T, Tp;               // valid rows per utterance / rows per utterance of the OUTPUT layout
  int Mu, B;               // sinc: GEMM rows (48-sample windows) per utterance, utterances
  int pool3, Tout;         // W-resident conv2 only: max_pool1d(3) fused into the epilogue, `out` is the POOLED tensor [B][Tout + 2][N]
  int tma_out;             // set by launch_rn_gemm: single-plane output leaves through shared memory + TMA stores
  const float *s1, *h1;    // affine applied to the accumulator (first_bn / bn2)
  const float *s2, *h2;    // sinc: second affine (layer1 bn1) for the pre-activated copy
  const bf16* add; int add_ld, add_planes, add_plane_cols;
  bf16* out; int out_ld, out_planes, out_plane_cols;
  bf16* out2; int out2_ld, out2_planes, out2_plane_cols;
};

// Conv1d weight [cout, cin, k] fp32 -> packed [cout, k, planes, cin_pad] bf16 hi(|lo) planes, or (f16 != 0, one plane)
// IEEE half: weights are the B operand of kind::f16 UMMAs, whose a_format / b_format are independent, so bf16
// activations x fp16 weights run at the bf16 rate with 11 instead of 8 significant weight bits.
int pack_conv_weight(const float* w, int cout, int cin, int k, int cin_pad, int planes, bf16* out, cudaStream_t stream,
                     int f16 = 0);
int absmax_host(const float* x, long long n, float* result_host);   // synchronous; engine creation only

// 2D bf16 row-major matrix [rows, ld] exposing `cols` columns as a TMA map; box = [64 cols, box_rows], 128B swizzle.
int make_map(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld, int box_rows);

int make_map_sw64(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld, int box_rows);
int make_map_sinc_out(CUtensorMap* m, const void* base, int channels, long long ld, int ml_extent, long long rows_per_utt,
                      int utts);
int make_map_f32(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld, int box_rows);
// Plain fp32-output GEMM out[M, N] = A * W^T through the TMA-store epilogue of gemm_pw.cu (N % 64 == 0).
int launch_gemm_f32(const GemmOperands& ops, const GemmParams& p, cudaStream_t stream);

// Cohort-score GEMM whose epilogue keeps only the scores >= thr[row] (fused top-K pre-filter, gemm_pw.cu / score.cu).
int launch_gemm_filter(const GemmOperands& ops, const GemmParams& p, const float* thr, unsigned* cnt, float* cand, int cap,
                       int n_valid, cudaStream_t stream);

// Launches the persistent tcgen05 GEMM.  block_n in {64, 128, 256}.
int launch_gemm(const GemmOperands& ops, const GemmParams& p, int block_n, cudaStream_t stream);

// Pointwise (k = 1, STORE_ALL, bf16 out) layers through the lean-epilogue kernel of gemm_pw.cu.
int launch_gemm_pw(const GemmOperands& ops, const GemmParams& p, cudaStream_t stream);

}  // namespace svb
