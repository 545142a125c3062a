// extern "C" surface of libsvb (declared in include/svb.h).
#include "../../include/svb.h"

#include "svb_internal.h"

namespace svb {
const char* last_error();
long long launches();
void reset_launches();
int build_row_meta(const int32_t*, int, int, int32_t*, long long, cudaStream_t);
int split_bf16(const float*, long long, int, int, bf16*, int, int, int, cudaStream_t);
int pack_conv_weight(const float*, int, int, int, int, int, bf16*, cudaStream_t);
int fold_bn(const float*, const float*, const float*, const float*, int, float, float*, float*, cudaStream_t);
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
  int bn = a->block_n;
  if (bn == 0) bn = (a->cout >= 256) ? 256 : (a->cout >= 128 ? 128 : 64);
  return launch_gemm(ops, p, bn, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
