"""The pointwise-layer kernel of the ECAPA path (gemm_pw.cu through svb_tdnn_layer, block_n = 0): 16-warp GELU epilogue,
2-CTA clusters with W-tile multicast (odd M-tile counts, a single tile), split-bf16 mode, and the fused per-utterance
column statistics (SE squeeze / ASP global mean and std), against plain PyTorch on the same bf16-rounded operands."""
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
def test_gelu_layer_bf16(M, K, N):
    nv = _nv()
    a, w, bias, scale, shift = _layer(M, K, N, M + K + N)
    a16 = nv.split_bf16(a)
    wp = nv.pack_conv_weight(w, planes=1)
    out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device="cuda")
    nv.tdnn_layer(a16, wp, N, K, 1, 1, bias, scale, shift, out, act=nv.ACT_GELU, out_kind=nv.OUT_BF16)
    torch.cuda.synchronize()
    ref = F.gelu(a16.float() @ w[:, :, 0].bfloat16().float().t() + bias) * scale + shift
    assert torch.isfinite(out.float()).all()
    err = (out.float() - ref).abs()
    assert (err <= ref.abs() * 2 ** -8 + 1e-3).all(), float(err.max())


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
    a16 = nv.split_bf16(a)
    wp = nv.pack_conv_weight(w, planes=1)
    meta = nv.build_row_meta(frames, H, "cuda")
    runs = []
    for _ in range(2):
        out = torch.zeros((M, N), dtype=torch.bfloat16, device="cuda")
        s = torch.zeros((len(frames), N), dtype=torch.int64, device="cuda")
        q = torch.zeros((len(frames), N), dtype=torch.int64, device="cuda") if with_sq else None
        nv.tdnn_layer(a16, wp, N, K, 1, 1, bias, scale, shift, out, act=nv.ACT_GELU, out_kind=nv.OUT_BF16, row_meta=meta,
                      halo=H, col_sum=s, col_sq=q)
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
