"""ctypes binding of libsvb.so (the C ABI declared in include/svb.h).

PyTorch is used for device memory and streams only: tensors are passed as
``data_ptr()`` and the current CUDA stream as a raw handle.  There is no
fallback: a missing library or a non-CUDA tensor raises.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SVB_LIB", os.path.join(_HERE, "libsvb.so"))   # SVB_LIB: A/B builds of the same ABI

ACT_NONE, ACT_RELU, ACT_GELU = 0, 1, 2
OUT_BF16, OUT_BF16_HILO, OUT_F32, OUT_F16 = 0, 1, 2, 3
STORE_ALL, STORE_MIRROR = 0, 1

_c = ctypes
_i32, _i64, _vp, _f32 = _c.c_int32, _c.c_int64, _c.c_void_p, _c.c_float


class SvbTdnnArgs(_c.Structure):
    _fields_ = [
        ("rows", _i64),
        ("x", _vp), ("x_ld", _i32), ("x_cols", _i32), ("x_plane_cols", _i32),
        ("x2", _vp), ("x2_ld", _i32), ("x2_cols", _i32), ("x2_chunks", _i32),
        ("w", _vp), ("cout", _i32), ("cin_pad", _i32), ("k", _i32), ("dilation", _i32), ("w_planes", _i32),
        ("passes", _i32),
        ("bias", _vp), ("bn_scale", _vp), ("bn_shift", _vp),
        ("row_bias", _vp), ("row_bias_ld", _i32),
        ("act", _i32), ("post_tanh", _i32),
        ("row_meta", _vp), ("store_mode", _i32), ("halo", _i32),
        ("out", _vp), ("out_ld", _i32), ("out_kind", _i32), ("out_plane_cols", _i32),
        ("add_src", _vp), ("add_ld", _i32), ("add_planes", _i32), ("add_plane_cols", _i32),
        ("out2", _vp), ("out2_ld", _i32), ("out2_plane_cols", _i32),
        ("block_n", _i32),
        ("col_sum", _vp), ("col_sq", _vp), ("stats_ld", _i32),
        ("w_f16", _i32), ("x_f16", _i32),
    ]


_lib = None


def lib():
    """Loads libsvb.so once; raises if it is missing (no CPU / eager fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libsvb.so not found at %s: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(the CUDA library is the only implementation of this path)" % LIB_PATH)
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.svb_last_error.restype = _c.c_char_p
        _lib.svb_launch_count.restype = _c.c_longlong
        _declare(_lib)
    return _lib


def _declare(L):
    L.svb_build_row_meta.argtypes = [_vp, _i32, _i32, _vp, _i64, _vp]
    L.svb_split_bf16.argtypes = [_vp, _i64, _i32, _i32, _vp, _i32, _i32, _i32, _vp]
    L.svb_pack_conv_weight.argtypes = [_vp, _i32, _i32, _i32, _i32, _i32, _vp, _vp]
    L.svb_to_f16.argtypes = [_vp, _i64, _i32, _i32, _vp, _i32, _vp]
    L.svb_pack_conv_weight_f16.argtypes = [_vp, _i32, _i32, _i32, _i32, _vp, _vp]
    L.svb_fold_bn.argtypes = [_vp, _vp, _vp, _vp, _i32, _f32, _vp, _vp, _vp]
    L.svb_tdnn_layer.argtypes = [_c.POINTER(SvbTdnnArgs), _vp]
    for name, (args, res) in _EXTRA_SIGS.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = res


class SvbEcapaCfg(_c.Structure):
    _fields_ = [("channels", _i32), ("n_mels", _i32), ("n_out", _i32), ("attention_channels", _i32),
                ("se_channels", _i32), ("res2net_scale", _i32), ("kernel_sizes", _i32 * 5),
                ("dilations", _i32 * 5), ("fp32_mode", _i32), ("input_norm", _i32), ("no_global_context", _i32)]


class SvbRawnet2Cfg(_c.Structure):
    _fields_ = [("n_samples", _i32), ("n_out", _i32), ("gru_node", _i32), ("sample_rate", _i32), ("fp32_mode", _i32),
                ("aggregate", _i32), ("att_dim", _i32)]


_pp = _c.POINTER(_vp)
_EXTRA_SIGS = {
    "svb_batch_create": ([_vp, _vp, _i32, _i32, _i32, _pp], _i32),
    "svb_batch_blob_bytes": ([_vp, _i32, _i32, _i32], _c.c_size_t),
    "svb_batch_create_into": ([_vp, _vp, _i32, _i32, _i32, _vp, _vp, _c.c_size_t, _vp, _pp], _i32),
    "svb_batch_destroy": ([_vp], None),
    "svb_batch_rows": ([_vp], _i64),
    "svb_batch_frames": ([_vp], _i64),
    "svb_batch_tiles": ([_vp], _i32),
    "svb_batch_row_meta": ([_vp], _vp),
    "svb_pcm_stage": ([_vp, _i32, _vp, _i32, _i32, _vp, _vp], _i32),
    "svb_fbank_create": ([_vp, _i32, _i32, _i32, _vp, _i32, _i32, _f32, _f32, _pp], _i32),
    "svb_fbank_destroy": ([_vp], None),
    "svb_fbank_forward": ([_vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp], _i32),
    "svb_ecapa_create": ([_c.POINTER(SvbEcapaCfg), _c.POINTER(_c.c_char_p), _c.POINTER(_vp), _i32, _pp], _i32),
    "svb_ecapa_destroy": ([_vp], None),
    "svb_ecapa_halo": ([_vp], _i32),
    "svb_ecapa_workspace_bytes": ([_vp, _vp], _c.c_size_t),
    "svb_ecapa_ws_offset": ([_vp, _vp, _c.c_char_p], _i64),
    "svb_ecapa_forward": ([_vp, _vp, _vp, _vp, _vp, _c.c_size_t, _vp, _i32, _i32, _vp], _i32),
    "svb_fbank_power": ([_vp, _vp, _vp, _vp, _vp, _vp], _i32),
    "svb_ecapa_forward_mel": ([_vp, _vp, _vp, _vp, _vp, _c.c_size_t, _vp, _i32, _i32, _vp], _i32),
    "svb_cosine_trials": ([_vp, _i32, _i32, _vp, _vp, _i64, _vp, _vp], _i32),
    "svb_pnorm_trials": ([_vp, _i32, _i32, _vp, _vp, _i64, _c.c_float, _c.c_float, _i32, _vp, _vp], _i32),
    "svb_mean_embeddings": ([_vp, _i32, _i32, _i32, _i32, _vp, _vp], _i32),
    "svb_enrol_nearest": ([_vp, _vp, _i32, _i32, _i32, _i32, _c.c_float, _i32, _vp, _vp, _vp, _vp], _i32),
    "svb_asnorm_workspace_bytes": ([_i32, _i32, _i32], _c.c_size_t),
    "svb_asnorm_stats": ([_vp, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _c.c_size_t, _vp], _i32),
    "svb_debug_gemm_noepi": ([_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp], _i32),
    "svb_asnorm_fixup_rows": ([_vp, _i32, _i32, _i32, _i32], _i32),
    "svb_asnorm_static_bytes": ([], _c.c_size_t),
    "svb_asnorm_stats_static": ([_vp, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _vp], _i32),
    "svb_asnorm_trials": ([_vp, _i32, _vp, _vp, _i64, _vp, _vp, _vp, _vp], _i32),
    "svb_rawnet2_create": ([_c.POINTER(SvbRawnet2Cfg), _c.POINTER(_c.c_char_p), _c.POINTER(_vp), _i32, _pp], _i32),
    "svb_rawnet2_destroy": ([_vp], None),
    "svb_rawnet2_workspace_bytes": ([_vp, _i32], _c.c_size_t),
    "svb_rawnet2_ws_offset": ([_vp, _i32, _c.c_char_p], _i64),
    "svb_rawnet2_frames": ([_vp, _i32], _i32),
    "svb_rawnet2_forward": ([_vp, _vp, _i64, _i32, _vp, _c.c_size_t, _vp, _i32, _i32, _vp], _i32),
    "svb_version": ([], _i32),
    "svb_reset_launch_count": ([], None),
}


def check(rc):
    if rc != 0:
        raise RuntimeError("libsvb: %s (code %d)" % (lib().svb_last_error().decode(), rc))


def stream_ptr(device=None):
    """Raw handle of torch's current stream on `device` (default: the current device).  libsvb launches on the
    CURRENT CUDA context, so a call for another device must run inside ``on_device``."""
    return _vp(torch.cuda.current_stream(device).cuda_stream)


def on_device(device):
    """Context manager that makes `device` the current CUDA device for the library calls inside it: libsvb never
    calls cudaSetDevice itself (its launches, cudaMalloc and TMA maps use the current context), so handles built for
    cuda:N must be driven with cuda:N current (src/model.py:55-56,171-172: the reference's gpu=N selects cuda:N)."""
    return torch.cuda.device(device)


def device_method(fn):
    """Decorator for methods of objects that carry ``self.device``: runs the body with that device current."""
    import functools

    @functools.wraps(fn)
    def wrapped(self, *a, **k):
        with torch.cuda.device(self.device):
            return fn(self, *a, **k)
    return wrapped


def require_current(device):
    """Raises unless `device` is the current CUDA device (a launch on another device would run device-N pointers on
    device 0's context)."""
    dev = torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    if dev.type != "cuda" or idx != torch.cuda.current_device():
        raise RuntimeError("libsvb call for %s while cuda:%d is current; wrap the call in torch.cuda.device(%s)"
                           % (dev, torch.cuda.current_device(), idx))


def ptr(t):
    """Device pointer of a CUDA tensor (None -> NULL); refuses CPU tensors."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("libsvb operates on CUDA tensors only (got a %s tensor); there is no CPU path" % t.device)
    return _vp(t.data_ptr())


# ------------------------------------------------------------------ thin op wrappers
def build_row_meta(frames, halo, device):
    """int32 [rows, 4] table {utt, t, T, 0} for utterances of `frames` valid rows each."""
    frames = [int(f) for f in frames]
    total = sum(f + 2 * halo for f in frames)
    meta = torch.empty((total, 4), dtype=torch.int32, device=device)
    arr = (_i32 * len(frames))(*frames)
    check(lib().svb_build_row_meta(arr, len(frames), halo, ptr(meta), total, stream_ptr()))
    return meta


def split_bf16(x, planes=1, out=None, plane_cols=None):
    """fp32 [rows, cols] -> bf16 [rows, planes*plane_cols] (hi | lo)."""
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    rows, cols = x.shape
    if plane_cols is None:
        plane_cols = cols
    if out is None:
        out = torch.zeros((rows, planes * plane_cols), dtype=torch.bfloat16, device=x.device)
    check(lib().svb_split_bf16(ptr(x), rows, cols, x.stride(0), ptr(out), out.stride(0), planes, plane_cols,
                               stream_ptr()))
    return out


def to_f16(x):
    """fp32 [rows, cols] -> one plane of IEEE half (saturating): the activation format of the 16-bit mode."""
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    out = torch.empty(x.shape, dtype=torch.float16, device=x.device)
    check(lib().svb_to_f16(ptr(x), x.shape[0], x.shape[1], x.stride(0), ptr(out), out.stride(0), stream_ptr()))
    return out


def pack_conv_weight(w, planes, cin_pad=None, f16=False):
    """nn.Conv1d weight [cout, cin, k] fp32 -> packed bf16 [cout, k*planes*cin_pad] (f16: one plane of IEEE half)."""
    w = w.contiguous().float()
    cout, cin, k = w.shape
    if cin_pad is None:
        cin_pad = (cin + 63) // 64 * 64
    if f16:
        assert planes == 1
        out = torch.empty((cout, k * cin_pad), dtype=torch.float16, device=w.device)
        check(lib().svb_pack_conv_weight_f16(ptr(w), cout, cin, k, cin_pad, ptr(out), stream_ptr()))
        return out
    out = torch.empty((cout, k * planes * cin_pad), dtype=torch.bfloat16, device=w.device)
    check(lib().svb_pack_conv_weight(ptr(w), cout, cin, k, cin_pad, planes, ptr(out), stream_ptr()))
    return out


def fold_bn(gamma, beta, mean, var, eps=1e-5):
    n = gamma.numel()
    scale = torch.empty(n, dtype=torch.float32, device=gamma.device)
    shift = torch.empty(n, dtype=torch.float32, device=gamma.device)
    check(lib().svb_fold_bn(ptr(gamma.contiguous()), ptr(beta.contiguous()), ptr(mean.contiguous()),
                            ptr(var.contiguous()), n, eps, ptr(scale), ptr(shift), stream_ptr()))
    return scale, shift


def tdnn_layer(x, w_packed, cout, cin_pad, k, dilation, bias, bn_scale, bn_shift, out, *, act=ACT_NONE,
               passes=1, w_planes=1, x_plane_cols=0, x_cols=None, x2=None, x2_chunks=0, row_bias=None,
               post_tanh=False, row_meta=None, store_mode=STORE_ALL, halo=0, out_kind=OUT_BF16,
               out_plane_cols=0, add_src=None, add_planes=1, add_plane_cols=0, out2=None,
               out2_plane_cols=0, block_n=0, rows=None, col_sum=None, col_sq=None, w_f16=False, x_f16=False):
    a = SvbTdnnArgs()
    a.w_f16, a.x_f16 = int(w_f16), int(x_f16)
    a.rows = x.shape[0] if rows is None else rows
    a.x, a.x_ld = ptr(x), x.stride(0)
    a.x_cols = x.shape[1] if x_cols is None else x_cols
    a.x_plane_cols = x_plane_cols
    if x2 is not None:
        a.x2, a.x2_ld, a.x2_cols, a.x2_chunks = ptr(x2), x2.stride(0), x2.shape[1], x2_chunks
    a.w, a.cout, a.cin_pad, a.k, a.dilation, a.w_planes = ptr(w_packed), cout, cin_pad, k, dilation, w_planes
    a.passes = passes
    a.bias, a.bn_scale, a.bn_shift = ptr(bias), ptr(bn_scale), ptr(bn_shift)
    if row_bias is not None:
        a.row_bias, a.row_bias_ld = ptr(row_bias), row_bias.stride(0)
    a.act, a.post_tanh = act, int(post_tanh)
    a.row_meta = ptr(row_meta)
    a.store_mode, a.halo = store_mode, halo
    a.out, a.out_ld, a.out_kind, a.out_plane_cols = ptr(out), out.stride(0), out_kind, out_plane_cols
    if add_src is not None:
        a.add_src, a.add_ld, a.add_planes, a.add_plane_cols = ptr(add_src), add_src.stride(0), add_planes, add_plane_cols
        a.out2, a.out2_ld, a.out2_plane_cols = ptr(out2), out2.stride(0), out2_plane_cols
    a.block_n = block_n
    if col_sum is not None:
        a.col_sum, a.stats_ld = ptr(col_sum), col_sum.stride(0)
        if col_sq is not None:
            a.col_sq = ptr(col_sq)
    check(lib().svb_tdnn_layer(ctypes.byref(a), stream_ptr()))
    return out
