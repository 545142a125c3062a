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
/* fp32 [rows, cols] (ld_in) -> one plane of IEEE half (saturating), the activation format of the 16-bit mode. */
int svb_to_f16(const float* x, int64_t rows, int cols, int ld_in, void* out, int ld_out, void* stream);
/* Conv1d weight [cout, cin, k] fp32 (nn.Conv1d layout, src/models/layers/cnn.py:79-88) ->
 * packed bf16 [cout, k, planes, cin_pad] (cin_pad multiple of 64, zero filled; planes: hi[, lo]). */
int svb_pack_conv_weight(const float* w, int cout, int cin, int k, int cin_pad, int planes,
                         void* out, void* stream);
/* Same layout with the weights stored as IEEE half (one plane): the weight operand of a tensor-core layer may be fp16
 * while the activations stay bf16 (SvbTdnnArgs.w_f16 = 1, passes = 1) -- same MMA rate, 11 instead of 8 significant
 * weight bits.  This is what bf16 mode of the ECAPA / RawNet2 engines uses for every single-plane layer. */
int svb_pack_conv_weight_f16(const float* w, int cout, int cin, int k, int cin_pad, void* out, void* stream);
/* BatchNorm1d(eval) -> per-channel scale/shift (src/models/layers/normalization.py:62-68, eps 1e-5). */
int svb_fold_bn(const float* gamma, const float* beta, const float* mean, const float* var, int n,
                float eps, float* scale, float* shift, void* stream);

/* ------------------------------------------------------------------ TDNN layer (tensor cores)
 * One TDNNBlock / Conv1d of ECAPA-TDNN as an implicit GEMM on tcgen05 tensor cores:
 *   out = BN(act(conv_reflect_same(x) + bias))          src/models/ECAPA_TDNN.py:68-69
 * replaces F.pad(reflect) + nn.Conv1d + activation + BatchNorm1d (4 library kernels). */
enum { SVB_ACT_NONE = 0, SVB_ACT_RELU = 1, SVB_ACT_GELU = 2 };
/* output storage: one bf16 plane, bf16 hi|lo planes (fp32 grade), fp32, or one plane of IEEE half (16-bit mode) */
enum { SVB_OUT_BF16 = 0, SVB_OUT_BF16_HILO = 1, SVB_OUT_F32 = 2, SVB_OUT_F16 = 3 };
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
  /* optional fused per-utterance column statistics of the OUTPUT (pointwise GELU layers, cout % 256 == 0, bf16 output,
   * row_meta required): col_sum[utt * stats_ld + c] += round(2^32 * sum over the utterance's valid rows of out[r, c]),
   * col_sq likewise for out^2.  int64 fixed point, integer atomics (order independent, bit reproducible); the caller
   * zeroes the buffers.  Replaces the separate time-mean passes of SEBlock (src/models/ECAPA_TDNN.py:171-172) and of
   * AttentiveStatisticsPooling's global context (src/models/ECAPA_TDNN.py:222-227). */
  int64_t* col_sum; int64_t* col_sq; int32_t stats_ld;
  int32_t w_f16;           /* w was packed by svb_pack_conv_weight_f16 (w_planes = 1, passes = 1) */
  int32_t x_f16;           /* x (x2, add_src) hold ONE plane of IEEE half instead of bf16 (passes = 1): the activation
                            * format of the 16-bit mode, i.e. what an SVB_OUT_F16 layer wrote */
} SvbTdnnArgs;

int svb_tdnn_layer(const SvbTdnnArgs* args, void* stream);


/* ------------------------------------------------------------------ batch plan
 * Host-built device tables for one batch of utterances (row_meta, frame tiles, per-utterance
 * offsets).  wav_off_host[b] = offset (in samples) of utterance b inside the wav buffer given to
 * the forward calls; nsamp_host[b] = its length; frames T_b = 1 + nsamp/hop (centred STFT). */
typedef struct SvbBatch SvbBatch;
int svb_batch_create(const int64_t* wav_off_host, const int32_t* nsamp_host, int B, int hop, int halo,
                     SvbBatch** out);
/* Same plan with CALLER-OWNED tables: built in blob_pinned_host (page-locked host memory) and uploaded with one
 * stream-ordered copy into blob_dev (256-byte aligned device memory), both of svb_batch_blob_bytes() bytes.  No
 * allocation and no synchronisation: used for the per-launch plans of the crop/stack pipeline.  The blobs must outlive
 * the plan and the host blob must not be rewritten before the copy has executed. */
size_t svb_batch_blob_bytes(const int32_t* nsamp_host, int B, int hop, int halo);
int svb_batch_create_into(const int64_t* wav_off_host, const int32_t* nsamp_host, int B, int hop, int halo,
                          void* blob_pinned_host, void* blob_dev, size_t blob_bytes, void* stream, SvbBatch** out);
void svb_batch_destroy(SvbBatch* b);
int64_t svb_batch_rows(const SvbBatch* b);    /* sum_b (T_b + 2*halo) */
int64_t svb_batch_frames(const SvbBatch* b);  /* sum_b T_b */
int32_t svb_batch_tiles(const SvbBatch* b);   /* 32-frame front-end tiles */
const int32_t* svb_batch_row_meta(const SvbBatch* b); /* device, int32 x4 per row */

/* ------------------------------------------------------------------ crop / stack stage
 * Replaces the evaluation branch of loadWAV (src/processing/audio_loader.py:110-150: wrap-pad of short
 * signals :126-129, num_eval evenly spaced crops :134,147-148) and the stacking done by test_data_loader
 * (src/dataloader.py:275-292).  Each utterance is uploaded ONCE as PCM (int16 as stored in the wav file, or
 * fp32) and staged on the device: out[dst_off + i] = float(pcm[src_off + i mod src_len]) (* 1/32768 for int16,
 * soundfile's float32 scaling) for i < dst_len, i.e. np.pad(audio, (0, dst_len - src_len), 'wrap').
 * Crops are not materialised: crop j of utterance u is the row (wav_off = dst_off[u] + int(linspace(0,
 * dst_len - L, n)[j]), nsamp = L) handed to svb_batch_create.  segs is a DEVICE array of U records. */
typedef struct SvbPcmSeg {
  int64_t src_off, dst_off;   /* in samples */
  int32_t src_len, dst_len;
} SvbPcmSeg;
int svb_pcm_stage(const void* pcm, int sample_kind /* 0 = fp32, 1 = int16 */, const SvbPcmSeg* segs, int U,
                  int max_dst_len, float* out, void* stream);

/* ------------------------------------------------------------------ log-mel front-end (K1)
 * Replaces PreEmphasis (src/utils.py:63-71) + melspectrogram(lib='torchaudio')
 * (src/models/FeatureExtraction/feature.py:84-94) + log/mean-norm (src/models/ECAPA_TDNN.py:473-476).
 * window_host[win_length] and fb_host[n_freqs, n_mels] are HOST tables built from the config keys. */
typedef struct SvbFbank SvbFbank;
int svb_fbank_create(const float* window_host, int win_length, int hop, int n_fft, const float* fb_host,
                     int n_freqs, int n_mels, float preemph, float log_eps, SvbFbank** out);
void svb_fbank_destroy(SvbFbank* f);
/* wav fp32 -> out bf16 [rows, out_ld]: hi plane at column 0, lo plane (planes==2) at +plane_cols;
 * columns n_mels..plane_cols-1 are zero; halo rows mirrored.  Scratch: logmel_ws [rows*n_mels] fp32,
 * tile_sums_ws [tiles*n_mels] fp32, mean_ws [B*n_mels] fp32 (logmel_ws keeps the un-normalised log-mel). */
int svb_fbank_forward(const SvbFbank* f, const SvbBatch* b, const float* wav, void* out, int out_ld,
                      int plane_cols, int planes, float* logmel_ws, float* tile_sums_ws, float* mean_ws,
                      void* stream);

/* Feature-plugin entry: power-mel only (what compute_features(x) returns at src/model.py:112-113),
 * written time-major into mel_rows [rows*n_mels] fp32 (valid rows; halo rows untouched). */
int svb_fbank_power(const SvbFbank* f, const SvbBatch* b, const float* wav, float* mel_rows,
                    float* tile_sums_ws, void* stream);

/* ------------------------------------------------------------------ ECAPA-TDNN encoder (K2-K7)
 * Replaces models.ECAPA_TDNN.MainModel.forward (src/models/ECAPA_TDNN.py:460-502) behind
 * SpeakerEncoder.forward (src/model.py:104-129).  Weights: fp32 DEVICE pointers in the reference
 * checkpoint layout, keyed by state-dict name without the "__S__." prefix
 * (e.g. "blocks.1.tdnn1.conv.conv.weight"); they are repacked into library-owned buffers. */
typedef struct SvbEcapa SvbEcapa;
typedef struct SvbEcapaCfg {
  int32_t channels;        /* C (channels[0..3]); channels[4] must be 3*C */
  int32_t n_mels, n_out, attention_channels, se_channels, res2net_scale;
  int32_t kernel_sizes[5], dilations[5];
  int32_t fp32_mode;       /* 0: 16-bit mode (IEEE-half activations and weights, fp32 accumulation and epilogues);
                            * 1: split-bf16 (fp32-grade parity mode, 3 tensor-core passes) */
  int32_t input_norm;      /* input_norm=True of src/models/ECAPA_TDNN.py:406-408,477-478 (the ECAPA branch of the
                            * Raw3_ECAPA hybrid, src/models/Raw3_ECAPA.py:21-24): InstanceNorm1d(n_mels, affine) after the
                            * log / mean normalisation; needs the tensors "instance_norm.weight" / ".bias" */
  int32_t no_global_context; /* global_context=False (src/models/ECAPA_TDNN.py:204-207,238-248): asp.tdnn has 3C inputs,
                            * the attention does not see the utterance's global mean / std */
} SvbEcapaCfg;
int svb_ecapa_create(const SvbEcapaCfg* cfg, const char* const* names, const void* const* ptrs, int n,
                     SvbEcapa** out);
void svb_ecapa_destroy(SvbEcapa* e);
int32_t svb_ecapa_halo(const SvbEcapa* e);
size_t svb_ecapa_workspace_bytes(const SvbEcapa* e, const SvbBatch* b);
/* byte offset of a named intermediate inside the workspace (tests / profiling), -1 if unknown */
int64_t svb_ecapa_ws_offset(const SvbEcapa* e, const SvbBatch* b, const char* name);
/* wav (device fp32) -> emb [B, emb_ld] fp32; l2norm != 0 applies F.normalize (src/model.py:422-423) */
int svb_ecapa_forward(const SvbEcapa* e, const SvbFbank* f, const SvbBatch* b, const float* wav, void* ws,
                      size_t ws_bytes, float* emb, int emb_ld, int l2norm, void* stream);

/* Encoder-plugin entry (MainModel.forward(x) with x = power-mel, src/models/ECAPA_TDNN.py:460-502):
 * mel_rows [rows, n_mels] fp32 power-mel in the time-major layout (valid rows filled). */
int svb_ecapa_forward_mel(const SvbEcapa* e, const SvbFbank* f, const SvbBatch* b, const float* mel_rows,
                          void* ws, size_t ws_bytes, float* emb, int emb_ld, int l2norm, void* stream);

/* ------------------------------------------------------------------ RawNet2 raw-waveform encoder
 * Replaces models.RawNet2_custom.MainModel.forward (src/models/RawNet2_custom.py:161-227, defaults
 * front_proc='sinc', aggregate='gru', layers [1,1,1,2,1,2], filters [128,128,256,256,512,512], :230-243) behind
 * SpeakerEncoder.forward with features == 'raw' (src/model.py:104-129): LayerNorm over samples, SincConv_fast +
 * |.| + max_pool1d(3) + first_bn + LeakyReLU(0.3), eight RawNetBasicBlocks with AFMS
 * (src/models/RawNet_baseline.py:13-24, 48-68, 181-361), bn_before_gru, GRU(512 -> gru_node) last state,
 * fc_after_gru; or, aggregate = 'sap' / 'asp', bn_before_agg + the attentive statistics heads of :200-224 + fc.  Weights: fp32 DEVICE pointers in the reference checkpoint layout keyed by state-dict name
 * without the "__S__." prefix ("ln.gamma", "first_conv.low_hz_", "layer3.0.shortcut.0.weight", ...).
 * Every utterance has exactly n_samples samples (ln.gamma / ln.beta have that length in the reference). */
typedef struct SvbRawnet2 SvbRawnet2;
typedef struct SvbRawnet2Cfg {
  int32_t n_samples;       /* audio_spec.sentence_len * sample_rate (64 000 for 4 s at 16 kHz) */
  int32_t n_out;           /* code_dim / nOut */
  int32_t gru_node;        /* 1024 */
  int32_t sample_rate;
  int32_t fp32_mode;       /* 0: 16-bit mode (IEEE-half operands); 1: split-bf16 (fp32-grade parity mode) */
  int32_t aggregate;       /* 0 = 'gru' (default head), 1 = 'sap', 2 = 'asp' (src/models/RawNet2_custom.py:84-112,188-224) */
  int32_t att_dim;         /* 'asp': channels of the attention bottleneck (128) */
} SvbRawnet2Cfg;
int svb_rawnet2_create(const SvbRawnet2Cfg* cfg, const char* const* names, const void* const* ptrs, int n,
                       SvbRawnet2** out);
void svb_rawnet2_destroy(SvbRawnet2* e);
size_t svb_rawnet2_workspace_bytes(const SvbRawnet2* e, int B);
/* byte offset of a named intermediate inside the workspace (tests / profiling), -1 if unknown */
int64_t svb_rawnet2_ws_offset(const SvbRawnet2* e, int B, const char* name);
/* frames entering residual block `level` (0..7); level 8 = frames entering the GRU */
int32_t svb_rawnet2_frames(const SvbRawnet2* e, int level);
/* wav [B, wav_ld] fp32 (device) -> emb [B, emb_ld] fp32; l2norm != 0 applies F.normalize (src/model.py:422-423) */
int svb_rawnet2_forward(const SvbRawnet2* e, const float* wav, int64_t wav_ld, int B, void* ws, size_t ws_bytes,
                        float* emb, int emb_ld, int l2norm, void* stream);

/* ------------------------------------------------------------------ trial scoring (K8/K9)
 * emb: [U, n_per_utt, D] fp32 crop embeddings of the U unique utterances; idx_a/idx_b: int32 indices
 * of the enrol / test utterance of every trial.
 * svb_cosine_trials     = utils.cosine_similarity (src/utils.py:163-164) for all trials at once:
 *                         mean_i |cos(a_i, b_i)| over index-paired crops, eps 1e-5.
 * svb_mean_embeddings   = per-utterance mean of (optionally L2-normalised, src/model.py:421-423) rows.
 * svb_asnorm_stats      = per-utterance mean / population-std of the top-K cohort scores
 *                         (np.inner + np.sort()[::-1][:top] + np.mean/np.std, src/utils.py:144-147);
 *                         topk < 0 follows the slice semantics of the reference (top=-1 drops one).
 * svb_asnorm_trials     = S-norm combination (src/utils.py:148-156) for all trials at once. */
int svb_cosine_trials(const float* emb, int n_per_utt, int D, const int32_t* idx_a, const int32_t* idx_b,
                      int64_t n_trials, float* out, void* stream);
/* svb_pnorm_trials      = utils.pnorm_similarity (src/utils.py:167-169; dispatched as 'pnorm' at :129-130, called with
 *                         p=2 from src/model.py:442-444,546-548) for all trials at once:
 *                         mean_i || a_i - b_i + eps ||_p over index-paired crops (F.pairwise_distance, eps 1e-6 added to
 *                         the difference); normalize != 0 L2-normalises every crop row first (src/model.py:421-423).
 *                         p = INFINITY selects the max norm. */
int svb_pnorm_trials(const float* emb, int n_per_utt, int D, const int32_t* idx_a, const int32_t* idx_b,
                     int64_t n_trials, float p, float eps, int normalize, float* out, void* stream);
int svb_mean_embeddings(const float* emb, int U, int n_per_utt, int D, int normalize, float* out,
                        void* stream);
size_t svb_asnorm_workspace_bytes(int U, int N, int D);
int svb_asnorm_stats(const float* emean, int U, const float* cohort, int N, int D, int topk, float* mu,
                     float* sigma, void* ws, size_t ws_bytes, void* stream);
/* svb_asnorm_stats without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  Same inputs, outputs and results as svb_asnorm_stats; the workspace is a file-scope device array
 * of svb_asnorm_static_bytes() = 16 MiB inside the library (every GPU that loads libsvb holds it until the process
 * exits).  Geometries with svb_asnorm_workspace_bytes(U, N, D) above that limit are rejected with an error message
 * (e.g. 1024 utterances x 2048 cohort vectors x 192 fits).  The call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine).  All calls on a device SHARE the array: calls that overlap on the GPU (different
 * streams or host threads) race -- serialise them; the array keeps the previous call's contents and the call itself
 * clears what must start at zero. */
size_t svb_asnorm_static_bytes(void);
int svb_asnorm_stats_static(const float* emean, int U, const float* cohort, int N, int D, int topk, float* mu,
                            float* sigma, void* stream);
/* Diagnostic (profiling): mainloop of a pointwise fp16 layer [M, K] x [N, K]^T with a store-free epilogue; thr_inf[M]
 * must hold +inf, cnt[M] / cand[M] are scratch. */
int svb_debug_gemm_noepi(const void* a, const void* w, int M, int N, int K, const float* thr_inf, unsigned* cnt,
                         float* cand, void* stream);
/* Diagnostic, synchronous: rows of the last svb_asnorm_stats call on `ws` that the fused pre-filter could not prove and
 * the exact fix-up kernel redid (-1: the call used the store-and-select path). */
int svb_asnorm_fixup_rows(const void* ws, int U, int N, int D, int topk);
int svb_asnorm_trials(const float* emean, int D, const int32_t* idx_a, const int32_t* idx_b,
                      int64_t n_trials, const float* mu, const float* sigma, float* out, void* stream);

/* ------------------------------------------------------------------ nearest-enrolment search
 * The --predict statistic of src/inference.py:284-294 for Q query utterances at once:
 *   dist[q, s] = mean_i || query[q, i, :] - bank[i, :, s] + eps ||_2   (F.pairwise_distance over the feature axis)
 * query: [Q, n_per_utt, D] fp32 crop embeddings (embed_utterance, src/model.py:675-704); bank: [n_per_utt, D, S] fp32,
 * the layout prepare('embed') saves as embeds.pt (src/model.py:634-641).  dist: [Q, S] out.  topk (0..8) > 0
 * additionally returns np.argsort(dist[q])[:topk] (ties to the lower index) in top_idx [Q, topk] with the distances
 * in top_dist [Q, topk]; the reference's score is 1 - dist^2 / 2 (host side). */
int svb_enrol_nearest(const float* query, const float* bank, int Q, int n_per_utt, int D, int S, float eps,
                      int topk, float* dist, float* top_dist, int32_t* top_idx, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SVB_H_ */
