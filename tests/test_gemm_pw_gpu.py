"""The pointwise-layer kernel of the ECAPA path (gemm_pw.cu through svb_tdnn_layer, block_n = 0): 16-warp GELU epilogue,
2-CTA clusters with W-tile multicast (odd M-tile counts, a single tile), the 16-bit mode (IEEE-half activations and
weights in, IEEE half out), split-bf16 mode, mixed bf16 x fp16 operands, and the fused per-utterance column statistics
(SE squeeze / ASP global mean and std), against plain PyTorch fp64 on the same rounded operands."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
H = 4


def _nv():
    from speakerverification_b200 import _native as nv
    nv.lib()
    return nv


def _layer(M, K, N, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    a = torch.randn(M, K, generator=g).cuda()
    w = (torch.randn(N, K, 1, generator=g) / K ** 0.5).cuda()
    bias = (torch.randn(N, generator=g) * 0.1).cuda()
    scale = (torch.rand(N, generator=g) + 0.5).cuda()
    shift = (torch.randn(N, generator=g) * 0.1).cuda()
    return a, w, bias, scale, shift


@pytest.mark.parametrize("M,K,N", [(1000, 256, 256), (4133, 1024, 1024), (33, 128, 512), (129, 1024, 3072), (2048, 3072, 256)])
def test_gelu_layer_f16(M, K, N):
    """16-bit mode: fp16 activations x fp16 weights -> fp32 accumulate -> bias / GELU / BN in fp32 -> fp16."""
    nv = _nv()
    a, w, bias, scale, shift = _layer(M, K, N, M + K + N)
    a16 = nv.to_f16(a)
    assert torch.equal(a16, a.half())
    wp = nv.pack_conv_weight(w, planes=1, f16=True)
    out = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
    nv.tdnn_layer(a16, wp, N, K, 1, 1, bias, scale, shift, out, act=nv.ACT_GELU, out_kind=nv.OUT_F16, w_f16=True, x_f16=True)
    torch.cuda.synchronize()
    ref = F.gelu(a16.double() @ w[:, :, 0].half().double().t() + bias.double()) * scale.double() + shift.double()
    assert torch.isfinite(out.float()).all()
    err = (out.double() - ref).abs()
    assert (err <= ref.abs() * 2 ** -11 + 3e-5).all(), float(err.max())


def test_f16_store_saturates():
    """Stores saturate at the largest finite half instead of producing inf (cvt.rn.satfinite)."""
    nv = _nv()
    M, K, N = 256, 128, 256
    a, w, bias, scale, shift = _layer(M, K, N, 3)
    out = torch.zeros((M, N), dtype=torch.float16, device="cuda")
    nv.tdnn_layer(nv.to_f16(a), nv.pack_conv_weight(w, planes=1, f16=True), N, K, 1, 1, bias, scale * 1e6, shift, out,
                  act=nv.ACT_GELU, out_kind=nv.OUT_F16, w_f16=True, x_f16=True)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all() and out.float().abs().max().item() == 65504.0


def test_gelu_layer_split_mode():
    """fp32-grade mode: hi|lo planes on both operands (3 passes), hi|lo output planes."""
    nv = _nv()
    M, K, N = 777, 512, 512
    a, w, bias, scale, shift = _layer(M, K, N, 5)
    a2 = nv.split_bf16(a, planes=2)
    wp = nv.pack_conv_weight(w, planes=2)
    out = torch.zeros((M, 2 * N), dtype=torch.bfloat16, device="cuda")
    nv.tdnn_layer(a2, wp, N, K, 1, 1, bias, scale, shift, out, act=nv.ACT_GELU, passes=3, w_planes=2, x_plane_cols=K,
                  x_cols=2 * K, out_kind=nv.OUT_BF16_HILO, out_plane_cols=N)
    torch.cuda.synchronize()
    got = out[:, :N].float() + out[:, N:].float()
    ref = F.gelu(a.double() @ w[:, :, 0].double().t() + bias.double()) * scale.double() + shift.double()
    assert (got.double() - ref).abs().max().item() < 2e-4


@pytest.mark.parametrize("with_sq", [False, True])
def test_fused_column_statistics(with_sq):
    """Fixed-point column sums of the stored bf16 outputs over the VALID rows of every utterance (halo rows and rows
    of other utterances in the same 32-row box excluded), bit-identical from run to run."""
    nv = _nv()
    frames = [301, 9, 40, 700, 128, 33, 17, 64]          # boxes with 1, 2 and 3 utterances, utterances spanning many boxes
    K, N = 256, 512
    M = sum(t + 2 * H for t in frames)
    a, w, bias, scale, shift = _layer(M, K, N, 11)
    a16 = nv.to_f16(a)
    wp = nv.pack_conv_weight(w, planes=1, f16=True)
    meta = nv.build_row_meta(frames, H, "cuda")
    runs = []
    for _ in range(2):
        out = torch.zeros((M, N), dtype=torch.float16, device="cuda")
        s = torch.zeros((len(frames), N), dtype=torch.int64, device="cuda")
        q = torch.zeros((len(frames), N), dtype=torch.int64, device="cuda") if with_sq else None
        nv.tdnn_layer(a16, wp, N, K, 1, 1, bias, scale, shift, out, act=nv.ACT_GELU, out_kind=nv.OUT_F16, row_meta=meta,
                      halo=H, col_sum=s, col_sq=q, w_f16=True, x_f16=True)
        torch.cuda.synchronize()
        runs.append((out, s, q))
    out, s, q = runs[0]
    assert torch.equal(s, runs[1][1]) and (q is None or torch.equal(q, runs[1][2]))      # order-independent atomics
    r0 = 0
    for u, T in enumerate(frames):
        v = out[r0 + H:r0 + H + T].double()
        want = v.sum(0)
        got = s[u].double() / 2.0 ** 32
        assert (got - want).abs().max().item() <= 1e-5 * max(1.0, want.abs().max().item()), u
        if with_sq:
            want2 = (v * v).sum(0)
            got2 = q[u].double() / 2.0 ** 32
            assert (got2 - want2).abs().max().item() <= 2e-5 * max(1.0, want2.abs().max().item()), u
        r0 += T + 2 * H


def test_mixed_operand_formats_are_refused():
    """The instruction descriptor of a kind::f16 UMMA has separate a_format / b_format fields, but bf16 activations x
    fp16 weights raise 'an illegal instruction was encountered' on sm_100a (measured in round 2): the library refuses
    the combination up front, on both GEMM kernels."""
    nv = _nv()
    M, K, N = 256, 128, 256
    a, w, bias, scale, shift = _layer(M, K, N, 9)
    for block_n, kind, dt in ((0, nv.OUT_F16, torch.float16), (256, nv.OUT_BF16, torch.bfloat16)):
        out = torch.zeros((M, N), dtype=dt, device="cuda")
        with pytest.raises(RuntimeError, match="share one 16-bit format"):
            nv.tdnn_layer(nv.split_bf16(a), nv.pack_conv_weight(w, planes=1, f16=True), N, K, 1, 1, bias, scale, shift, out,
                          act=nv.ACT_GELU, out_kind=kind, w_f16=True, block_n=block_n)
    # fp16 x fp16 on the generic kernel (gemm_tc) with IEEE-half output
    out = torch.zeros((M, N), dtype=torch.float16, device="cuda")
    nv.tdnn_layer(nv.to_f16(a), nv.pack_conv_weight(w, planes=1, f16=True), N, K, 1, 1, bias, scale, shift, out,
                  act=nv.ACT_GELU, out_kind=nv.OUT_F16, w_f16=True, x_f16=True, block_n=256)
    torch.cuda.synchronize()
    ref = F.gelu(a.half().double() @ w[:, :, 0].half().double().t() + bias.double()) * scale.double() + shift.double()
    assert ((out.double() - ref).abs() <= ref.abs() * 2 ** -11 + 1e-4).all()


def test_row_meta_table():
    nv = _nv()
    frames = [301, 9, 40, 700]
    for _ in range(6):                                    # more calls than ring slots
        meta = nv.build_row_meta(frames, H, "cuda")
    torch.cuda.synchronize()
    m = meta.cpu()
    r = 0
    for u, T in enumerate(frames):
        for t in (-H, 0, T - 1, T + H - 1):
            assert m[r + H + t].tolist() == [u, t, T, 0]
        r += T + 2 * H
