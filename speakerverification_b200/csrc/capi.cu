// extern "C" surface of libsvb (declared in include/svb.h).
#include "../../include/svb.h"

#include "svb_internal.h"

namespace svb {
const char* last_error();
long long launches();
void reset_launches();
int build_row_meta(const int32_t*, int, int, int32_t*, long long, cudaStream_t);
int to_f16(const float*, long long, int, int, void*, int, cudaStream_t);
int split_bf16(const float*, long long, int, int, bf16*, int, int, int, cudaStream_t);
int fold_bn(const float*, const float*, const float*, const float*, int, float, float*, float*, cudaStream_t);
int pcm_stage(const void*, int, const PcmSeg*, int, int, float*, cudaStream_t);
struct Fbank; struct Batch; struct Ecapa;
int fbank_create(const float*, int, int, int, const float*, int, int, float, float, Fbank**);
void fbank_destroy(Fbank*);
int fbank_forward(const Fbank*, const float*, const void*, const int32_t*, int, int, const RowMeta*, long long, int,
                  float*, float*, float*, bf16*, int, int, int, int, cudaStream_t, const float*, const float*, float, float*);
int batch_create(const int64_t*, const int32_t*, int, int, int, Batch**);
int batch_create_into(const int64_t*, const int32_t*, int, int, int, void*, void*, size_t, cudaStream_t, Batch**);
size_t batch_blob_bytes(const int32_t*, int, int, int);
void batch_destroy(Batch*);
long long batch_rows(const Batch*);
long long batch_frames(const Batch*);
const int32_t* batch_row_meta(const Batch*);
int batch_tiles(const Batch*);
int batch_B(const Batch*);
int batch_halo(const Batch*);
const void* batch_utt_tab(const Batch*);
const int32_t* batch_tile_tab(const Batch*);
int ecapa_create(int, int, int, int, int, int, const int*, const int*, int, int, int, const char* const*, const void* const*,
                 int, Ecapa**);
void ecapa_destroy(Ecapa*);
int ecapa_halo(const Ecapa*);
size_t ecapa_workspace_bytes(const Ecapa*, const Batch*);
long long ecapa_ws_offset(const Ecapa*, const Batch*, const char*);
int ecapa_forward(const Ecapa*, const Fbank*, const Batch*, const float*, int, void*, size_t, float*, int, int,
                  cudaStream_t);
int cosine_trials(const float*, int, int, const int32_t*, const int32_t*, long long, float*, cudaStream_t);
int asnorm_fixup_rows(const void*, int, int, int, int);
int pnorm_trials(const float*, int, int, const int32_t*, const int32_t*, long long, float, float, int, float*, cudaStream_t);
int mean_rows(const float*, int, int, int, int, float*, cudaStream_t);
int enrol_nearest(const float*, const float*, int, int, int, int, float, int, float*, float*, int32_t*, cudaStream_t);
size_t asnorm_workspace_bytes(int, int, int, int*);
int asnorm_stats(const float*, int, const float*, int, int, int, float*, float*, void*, size_t, cudaStream_t);
int asnorm_stats_static(const float*, int, const float*, int, int, int, float*, float*, cudaStream_t);
size_t asnorm_static_bytes();
int asnorm_trials(const float*, int, const int32_t*, const int32_t*, long long, const float*, const float*, float*,
                  cudaStream_t);
struct Rawnet2;
int rawnet2_create(int, int, int, int, int, int, int, const char* const*, const void* const*, int, Rawnet2**);
void rawnet2_destroy(Rawnet2*);
size_t rawnet2_workspace_bytes(const Rawnet2*, int);
long long rawnet2_ws_offset(const Rawnet2*, int, const char*);
int rawnet2_frames(const Rawnet2*, int);
int rawnet2_forward(const Rawnet2*, const float*, long long, int, void*, size_t, float*, int, int, cudaStream_t);
}  // namespace svb

using namespace svb;

static int check_device() {
  static int ok = 0;
  if (ok) return 0;
  int dev = 0;
  cudaDeviceProp prop;
  SVB_CUDA_OK(cudaGetDevice(&dev));
  SVB_CUDA_OK(cudaGetDeviceProperties(&prop, dev));
  SVB_CHECK(prop.major == 10, "libsvb needs an sm_100 (Blackwell B200) GPU, found sm_%d%d; there is no fallback",
            prop.major, prop.minor);
  ok = 1;
  return 0;
}

extern "C" {

int svb_version(void) { return SVB_VERSION; }
const char* svb_last_error(void) { return last_error(); }
long long svb_launch_count(void) { return launches(); }
void svb_reset_launch_count(void) { reset_launches(); }

int svb_build_row_meta(const int32_t* frames_host, int B, int halo, int32_t* row_meta, int64_t total_rows,
                       void* stream) {
  if (check_device()) return -1;
  return build_row_meta(frames_host, B, halo, row_meta, total_rows, static_cast<cudaStream_t>(stream));
}

int svb_pcm_stage(const void* pcm, int sample_kind, const SvbPcmSeg* segs, int U, int max_dst_len, float* out,
                  void* stream) {
  if (check_device()) return -1;
  static_assert(sizeof(SvbPcmSeg) == sizeof(PcmSeg), "SvbPcmSeg layout");
  return pcm_stage(pcm, sample_kind, reinterpret_cast<const PcmSeg*>(segs), U, max_dst_len, out,
                   static_cast<cudaStream_t>(stream));
}

int svb_split_bf16(const float* x, int64_t rows, int cols, int ld_in, void* out, int ld_out, int planes,
                   int plane_cols, void* stream) {
  if (check_device()) return -1;
  return split_bf16(x, rows, cols, ld_in, static_cast<bf16*>(out), ld_out, planes, plane_cols,
                    static_cast<cudaStream_t>(stream));
}

int svb_pack_conv_weight(const float* w, int cout, int cin, int k, int cin_pad, int planes, void* out,
                         void* stream) {
  if (check_device()) return -1;
  return pack_conv_weight(w, cout, cin, k, cin_pad, planes, static_cast<bf16*>(out),
                          static_cast<cudaStream_t>(stream));
}

int svb_to_f16(const float* x, int64_t rows, int cols, int ld_in, void* out, int ld_out, void* stream) {
  if (check_device()) return -1;
  return to_f16(x, rows, cols, ld_in, out, ld_out, static_cast<cudaStream_t>(stream));
}

int svb_pack_conv_weight_f16(const float* w, int cout, int cin, int k, int cin_pad, void* out, void* stream) {
  if (check_device()) return -1;
  return pack_conv_weight(w, cout, cin, k, cin_pad, 1, static_cast<bf16*>(out), static_cast<cudaStream_t>(stream), 1);
}

int svb_fold_bn(const float* gamma, const float* beta, const float* mean, const float* var, int n, float eps,
                float* scale, float* shift, void* stream) {
  if (check_device()) return -1;
  return fold_bn(gamma, beta, mean, var, n, eps, scale, shift, static_cast<cudaStream_t>(stream));
}

int svb_tdnn_layer(const SvbTdnnArgs* a, void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(a != nullptr && a->x != nullptr && a->w != nullptr && a->out != nullptr, "null argument");
  SVB_CHECK(a->k == 1 || a->k == 3 || a->k == 5, "kernel size must be 1, 3 or 5 (got %d)", a->k);
  SVB_CHECK(a->cin_pad % 64 == 0, "cin_pad must be a multiple of 64");
  SVB_CHECK(a->k == 1 || (a->row_meta != nullptr && a->store_mode == SVB_STORE_MIRROR),
            "k > 1 needs row_meta and SVB_STORE_MIRROR (reflect padding lives in the halo rows)");
  SVB_CHECK(a->k == 1 || a->dilation * (a->k - 1) / 2 <= a->halo, "halo %d too small for k=%d dilation=%d",
            a->halo, a->k, a->dilation);
  GemmOperands ops{};
  ops.a = static_cast<const bf16*>(a->x);
  ops.lda = a->x_ld;
  ops.a_cols = a->x_cols;
  ops.a2 = static_cast<const bf16*>(a->x2);
  ops.lda2 = a->x2_ld;
  ops.a2_cols = a->x2_cols;
  ops.w = static_cast<const bf16*>(a->w);
  ops.w_rows = a->cout;
  ops.w_cols = a->k * a->w_planes * a->cin_pad;
  ops.w_f16 = a->w_f16;
  ops.a_f16 = a->x_f16;
  SVB_CHECK(!a->x_f16 || (a->passes == 1 && (a->add_src == nullptr || a->add_planes == 1)),
            "fp16 activations (x_f16) are single plane: passes = 1");
  GemmParams p{};
  p.M = static_cast<int>(a->rows);
  p.N = a->cout;
  p.kc_per_tap = a->cin_pad / 64;
  p.num_taps = a->k;
  for (int j = 0; j < a->k; ++j) p.tap_off[j] = (j - (a->k - 1) / 2) * a->dilation;
  p.num_pass = a->passes;
  p.a_plane_cols = a->x_plane_cols;
  p.a2_kc = (a->x2 != nullptr) ? a->x2_chunks : 0;
  p.w_planes = a->w_planes;
  SVB_CHECK(a->passes == 1 || a->w_planes == 2, "split mode needs hi/lo weight planes");
  p.bias = a->bias;
  p.scale = a->bn_scale;
  p.shift = a->bn_shift;
  p.rowbias = a->row_bias;
  p.rowbias_ld = a->row_bias_ld;
  p.act = a->act;
  p.post_tanh = a->post_tanh;
  p.row_meta = reinterpret_cast<const RowMeta*>(a->row_meta);
  p.store_mode = a->store_mode;
  p.halo = a->halo;
  p.out = a->out;
  p.out_ld = a->out_ld;
  p.out_kind = a->out_kind;
  p.out_plane_cols = a->out_plane_cols;
  p.add_src = static_cast<const bf16*>(a->add_src);
  p.add_ld = a->add_ld;
  p.add_planes = a->add_planes;
  p.add_plane_cols = a->add_plane_cols;
  p.out2 = (a->add_src != nullptr) ? a->out2 : nullptr;
  p.out2_ld = a->out2_ld;
  p.out2_plane_cols = a->out2_plane_cols;
  p.colsum = reinterpret_cast<long long*>(a->col_sum);
  p.colsq = reinterpret_cast<long long*>(a->col_sq);
  p.stats_ld = a->stats_ld;
  SVB_CHECK(a->col_sq == nullptr || a->col_sum != nullptr, "col_sq needs col_sum");
  int bn = a->block_n;
  // pointwise GELU / (ReLU + row-bias + tanh) layers with fp16 or bf16 hi|lo output take the lean-epilogue kernel
  // (gemm_pw.cu); single-plane bf16 output stays on the generic kernel
  const bool pointwise = bn == 0 && a->k == 1 && a->store_mode == SVB_STORE_ALL &&
                         (a->out_kind == SVB_OUT_F16 || a->out_kind == SVB_OUT_BF16_HILO) &&
                         p.out2 == nullptr && a->cout % 64 == 0 && p.num_pass != 2 &&
                         ((a->act == SVB_ACT_GELU && a->row_bias == nullptr && !a->post_tanh) ||
                          (a->act == SVB_ACT_RELU && a->row_bias != nullptr && a->post_tanh && a->row_meta != nullptr));
  if (pointwise) return launch_gemm_pw(ops, p, static_cast<cudaStream_t>(stream));
  SVB_CHECK(a->col_sum == nullptr, "fused column statistics are an option of the pointwise GELU layers");
  if (bn == 0) bn = (a->cout >= 256) ? 256 : (a->cout >= 128 ? 128 : 64);
  return launch_gemm(ops, p, bn, static_cast<cudaStream_t>(stream));
}


// ---------------------------------------------------------------- batch / fbank / ecapa handles
int svb_batch_create(const int64_t* wav_off_host, const int32_t* nsamp_host, int B, int hop, int halo, SvbBatch** out) {
  if (check_device()) return -1;
  SVB_CHECK(wav_off_host && nsamp_host && out, "null argument");
  return batch_create(wav_off_host, nsamp_host, B, hop, halo, reinterpret_cast<Batch**>(out));
}
size_t svb_batch_blob_bytes(const int32_t* nsamp_host, int B, int hop, int halo) {
  if (!nsamp_host) return 0;
  return batch_blob_bytes(nsamp_host, B, hop, halo);
}
int svb_batch_create_into(const int64_t* wav_off_host, const int32_t* nsamp_host, int B, int hop, int halo,
                          void* blob_pinned_host, void* blob_dev, size_t blob_bytes, void* stream, SvbBatch** out) {
  if (check_device()) return -1;
  SVB_CHECK(wav_off_host && nsamp_host && out, "null argument");
  return batch_create_into(wav_off_host, nsamp_host, B, hop, halo, blob_pinned_host, blob_dev, blob_bytes,
                           static_cast<cudaStream_t>(stream), reinterpret_cast<Batch**>(out));
}
void svb_batch_destroy(SvbBatch* b) { batch_destroy(reinterpret_cast<Batch*>(b)); }
int64_t svb_batch_rows(const SvbBatch* b) { return batch_rows(reinterpret_cast<const Batch*>(b)); }
int64_t svb_batch_frames(const SvbBatch* b) { return batch_frames(reinterpret_cast<const Batch*>(b)); }
int32_t svb_batch_tiles(const SvbBatch* b) { return batch_tiles(reinterpret_cast<const Batch*>(b)); }
const int32_t* svb_batch_row_meta(const SvbBatch* b) { return batch_row_meta(reinterpret_cast<const Batch*>(b)); }

int svb_fbank_create(const float* window_host, int win_length, int hop, int n_fft, const float* fb_host, int n_freqs,
                     int n_mels, float preemph, float log_eps, SvbFbank** out) {
  if (check_device()) return -1;
  SVB_CHECK(window_host && fb_host && out, "null argument");
  return fbank_create(window_host, win_length, hop, n_fft, fb_host, n_freqs, n_mels, preemph, log_eps,
                      reinterpret_cast<Fbank**>(out));
}
void svb_fbank_destroy(SvbFbank* f) { fbank_destroy(reinterpret_cast<Fbank*>(f)); }

int svb_fbank_forward(const SvbFbank* f, const SvbBatch* b, const float* wav, void* out, int out_ld, int plane_cols,
                      int planes, float* logmel_ws, float* tile_sums_ws, float* mean_ws, void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(f && b && wav && out && logmel_ws && tile_sums_ws && mean_ws, "null argument");
  const Batch* bt = reinterpret_cast<const Batch*>(b);
  return fbank_forward(reinterpret_cast<const Fbank*>(f), wav, batch_utt_tab(bt), batch_tile_tab(bt), batch_B(bt),
                       batch_tiles(bt), reinterpret_cast<const RowMeta*>(batch_row_meta(bt)), batch_rows(bt),
                       batch_halo(bt), logmel_ws, tile_sums_ws, mean_ws, static_cast<bf16*>(out), out_ld, plane_cols,
                       planes, 0, static_cast<cudaStream_t>(stream), nullptr, nullptr, 0.f, nullptr);
}

int svb_fbank_power(const SvbFbank* f, const SvbBatch* b, const float* wav, float* mel_rows, float* tile_sums_ws,
                    void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(f && b && wav && mel_rows && tile_sums_ws, "null argument");
  const Batch* bt = reinterpret_cast<const Batch*>(b);
  return fbank_forward(reinterpret_cast<const Fbank*>(f), wav, batch_utt_tab(bt), batch_tile_tab(bt), batch_B(bt),
                       batch_tiles(bt), reinterpret_cast<const RowMeta*>(batch_row_meta(bt)), batch_rows(bt),
                       batch_halo(bt), mel_rows, tile_sums_ws, nullptr, nullptr, 0, 0, 0, 1,
                       static_cast<cudaStream_t>(stream), nullptr, nullptr, 0.f, nullptr);
}

int svb_ecapa_create(const SvbEcapaCfg* cfg, const char* const* names, const void* const* ptrs, int n, SvbEcapa** out) {
  if (check_device()) return -1;
  SVB_CHECK(cfg && names && ptrs && out, "null argument");
  return ecapa_create(cfg->channels, cfg->n_mels, cfg->n_out, cfg->attention_channels, cfg->se_channels,
                      cfg->res2net_scale, cfg->kernel_sizes, cfg->dilations, cfg->fp32_mode, cfg->input_norm, !cfg->no_global_context, names, ptrs, n,
                      reinterpret_cast<Ecapa**>(out));
}
void svb_ecapa_destroy(SvbEcapa* e) { ecapa_destroy(reinterpret_cast<Ecapa*>(e)); }
int32_t svb_ecapa_halo(const SvbEcapa* e) { return ecapa_halo(reinterpret_cast<const Ecapa*>(e)); }
size_t svb_ecapa_workspace_bytes(const SvbEcapa* e, const SvbBatch* b) {
  return ecapa_workspace_bytes(reinterpret_cast<const Ecapa*>(e), reinterpret_cast<const Batch*>(b));
}
int64_t svb_ecapa_ws_offset(const SvbEcapa* e, const SvbBatch* b, const char* name) {
  return ecapa_ws_offset(reinterpret_cast<const Ecapa*>(e), reinterpret_cast<const Batch*>(b), name);
}
int svb_ecapa_forward(const SvbEcapa* e, const SvbFbank* f, const SvbBatch* b, const float* wav, void* ws,
                      size_t ws_bytes, float* emb, int emb_ld, int l2norm, void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(e && f && b && wav && emb, "null argument");
  return ecapa_forward(reinterpret_cast<const Ecapa*>(e), reinterpret_cast<const Fbank*>(f),
                       reinterpret_cast<const Batch*>(b), wav, 0, ws, ws_bytes, emb, emb_ld, l2norm,
                       static_cast<cudaStream_t>(stream));
}
int svb_ecapa_forward_mel(const SvbEcapa* e, const SvbFbank* f, const SvbBatch* b, const float* mel_rows, void* ws,
                          size_t ws_bytes, float* emb, int emb_ld, int l2norm, void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(e && f && b && mel_rows && emb, "null argument");
  return ecapa_forward(reinterpret_cast<const Ecapa*>(e), reinterpret_cast<const Fbank*>(f),
                       reinterpret_cast<const Batch*>(b), mel_rows, 1, ws, ws_bytes, emb, emb_ld, l2norm,
                       static_cast<cudaStream_t>(stream));
}


// ---------------------------------------------------------------- scoring
int svb_pnorm_trials(const float* emb, int n_per_utt, int D, const int32_t* idx_a, const int32_t* idx_b, int64_t n_trials,
                     float p, float eps, int normalize, float* out, void* stream) {
  if (check_device()) return -1;
  return pnorm_trials(emb, n_per_utt, D, idx_a, idx_b, n_trials, p, eps, normalize, out, static_cast<cudaStream_t>(stream));
}

int svb_cosine_trials(const float* emb, int n_per_utt, int D, const int32_t* idx_a, const int32_t* idx_b,
                      int64_t n_trials, float* out, void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(emb && idx_a && idx_b && out && n_per_utt >= 1 && D >= 1, "bad argument");
  return cosine_trials(emb, n_per_utt, D, idx_a, idx_b, n_trials, out, static_cast<cudaStream_t>(stream));
}
int svb_mean_embeddings(const float* emb, int U, int n_per_utt, int D, int normalize, float* out, void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(emb && out && U >= 1 && n_per_utt >= 1, "bad argument");
  return mean_rows(emb, U, n_per_utt, D, normalize, out, static_cast<cudaStream_t>(stream));
}
int svb_enrol_nearest(const float* query, const float* bank, int Q, int n_per_utt, int D, int S, float eps, int topk,
                      float* dist, float* top_dist, int32_t* top_idx, void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(query && bank && dist && Q >= 0 && n_per_utt >= 1 && D >= 1 && S >= 1, "bad argument");
  SVB_CHECK(Q <= 8 * 65535, "at most %d query utterances per call", 8 * 65535);   // grid.y of the distance kernel
  SVB_CHECK(topk >= 0 && topk <= 8 && topk <= S && (topk == 0 || (top_dist && top_idx)), "topk must be 0..min(8, S)");
  return enrol_nearest(query, bank, Q, n_per_utt, D, S, eps, topk, dist, top_dist, top_idx,
                       static_cast<cudaStream_t>(stream));
}
size_t svb_asnorm_workspace_bytes(int U, int N, int D) { return asnorm_workspace_bytes(U, N, D, nullptr); }
int svb_asnorm_stats(const float* emean, int U, const float* cohort, int N, int D, int topk, float* mu, float* sigma,
                     void* ws, size_t ws_bytes, void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(emean && cohort && mu && sigma && U >= 1 && N >= 1 && D >= 1, "bad argument");
  return asnorm_stats(emean, U, cohort, N, D, topk, mu, sigma, ws, ws_bytes, static_cast<cudaStream_t>(stream));
}
// No-allocation variant (no reference counterpart): workspace in a 16 MiB file-scope device array
size_t svb_asnorm_static_bytes(void) { return asnorm_static_bytes(); }
int svb_asnorm_stats_static(const float* emean, int U, const float* cohort, int N, int D, int topk, float* mu, float* sigma,
                            void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(emean && cohort && mu && sigma && U >= 1 && N >= 1 && D >= 1, "bad argument");
  return asnorm_stats_static(emean, U, cohort, N, D, topk, mu, sigma, static_cast<cudaStream_t>(stream));
}
// Diagnostic: the GEMM mainloop of a pointwise layer with the epilogue reduced to TMEM reads and compares (nothing
// stored: thr = +inf everywhere) -- separates operand-feed time from epilogue time (tools/pw_alone.py).
int svb_debug_gemm_noepi(const void* a, const void* w, int M, int N, int K, const float* thr_inf, unsigned* cnt, float* cand,
                         void* stream) {
  if (check_device()) return -1;
  GemmOperands ops{};
  ops.a = static_cast<const bf16*>(a); ops.lda = K; ops.a_cols = K;
  ops.w = static_cast<const bf16*>(w); ops.w_rows = N; ops.w_cols = K; ops.a_f16 = ops.w_f16 = 1;
  GemmParams g{};
  g.M = M; g.N = N; g.kc_per_tap = K / 64; g.num_taps = 1; g.tap_off[0] = 0; g.num_pass = 1; g.w_planes = 1;
  return launch_gemm_filter(ops, g, thr_inf, cnt, cand, 1, N, static_cast<cudaStream_t>(stream));
}

int svb_asnorm_fixup_rows(const void* ws, int U, int N, int D, int topk) { return asnorm_fixup_rows(ws, U, N, D, topk); }

int svb_asnorm_trials(const float* emean, int D, const int32_t* idx_a, const int32_t* idx_b, int64_t n_trials,
                      const float* mu, const float* sigma, float* out, void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(emean && idx_a && idx_b && mu && sigma && out, "bad argument");
  return asnorm_trials(emean, D, idx_a, idx_b, n_trials, mu, sigma, out, static_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------- RawNet2
int svb_rawnet2_create(const SvbRawnet2Cfg* cfg, const char* const* names, const void* const* ptrs, int n,
                       SvbRawnet2** out) {
  if (check_device()) return -1;
  SVB_CHECK(cfg && names && ptrs && out && n > 0, "bad argument");
  return rawnet2_create(cfg->n_samples, cfg->n_out, cfg->gru_node, cfg->sample_rate, cfg->fp32_mode, cfg->aggregate,
                        cfg->att_dim, names, ptrs, n,
                        reinterpret_cast<Rawnet2**>(out));
}
void svb_rawnet2_destroy(SvbRawnet2* e) { rawnet2_destroy(reinterpret_cast<Rawnet2*>(e)); }
size_t svb_rawnet2_workspace_bytes(const SvbRawnet2* e, int B) {
  return (e && B > 0) ? rawnet2_workspace_bytes(reinterpret_cast<const Rawnet2*>(e), B) : 0;
}
int64_t svb_rawnet2_ws_offset(const SvbRawnet2* e, int B, const char* name) {
  return (e && name) ? rawnet2_ws_offset(reinterpret_cast<const Rawnet2*>(e), B, name) : -1;
}
int32_t svb_rawnet2_frames(const SvbRawnet2* e, int level) {
  return e ? rawnet2_frames(reinterpret_cast<const Rawnet2*>(e), level) : -1;
}
int svb_rawnet2_forward(const SvbRawnet2* e, const float* wav, int64_t wav_ld, int B, void* ws, size_t ws_bytes,
                        float* emb, int emb_ld, int l2norm, void* stream) {
  if (check_device()) return -1;
  SVB_CHECK(e && wav && ws && emb && B > 0, "bad argument");
  return rawnet2_forward(reinterpret_cast<const Rawnet2*>(e), wav, wav_ld, B, ws, ws_bytes, emb, emb_ld, l2norm,
                         static_cast<cudaStream_t>(stream));
}

}  // extern "C"
