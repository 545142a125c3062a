"""GPU parity of the tcgen05 TDNN/Conv1d layer (C ABI svb_tdnn_layer) against plain PyTorch fp32."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

H = 4


def _native():
    from speakerverification_b200 import _native as nv
    nv.lib()
    return nv


def _padded_rows(x_list, halo):
    """list of [T_b, C] -> [sum(T_b + 2*halo), C] with mirrored halos."""
    rows = []
    for x in x_list:
        T = x.shape[0]
        lo = x[1:halo + 1].flip(0)
        hi = x[T - 1 - halo:T - 1].flip(0)
        rows += [lo, x, hi]
    return torch.cat(rows, 0)


@pytest.mark.parametrize("M,K,N,bn", [(1000, 256, 512, 256), (300, 64, 128, 128), (4133, 1024, 1024, 256),
                                      (700, 128, 64, 64)])
def test_pointwise_bf16(M, K, N, bn):
    nv = _native()
    dev = "cuda"
    g = torch.Generator(device="cpu").manual_seed(M + K)
    a = torch.randn(M, K, generator=g).to(dev)
    w = (torch.randn(N, K, 1, generator=g) / K ** 0.5).to(dev)
    bias = torch.randn(N, generator=g).to(dev) * 0.1
    scale = (torch.rand(N, generator=g) + 0.5).to(dev)
    shift = torch.randn(N, generator=g).to(dev) * 0.1
    a16 = nv.split_bf16(a)
    wp = nv.pack_conv_weight(w, planes=1)
    out = torch.full((M, N), float("nan"), device=dev)
    nv.tdnn_layer(a16, wp, N, K, 1, 1, bias, scale, shift, out, act=nv.ACT_GELU, out_kind=nv.OUT_F32, block_n=bn)
    torch.cuda.synchronize()
    ref = F.gelu(a16.float() @ w[:, :, 0].bfloat16().float().t() + bias) * scale + shift
    assert torch.isfinite(out).all()
    assert (out - ref).abs().max().item() < 2e-4
    # bf16 output
    out16 = torch.zeros((M, N), dtype=torch.bfloat16, device=dev)
    nv.tdnn_layer(a16, wp, N, K, 1, 1, bias, scale, shift, out16, act=nv.ACT_RELU, out_kind=nv.OUT_BF16, block_n=bn)
    torch.cuda.synchronize()
    ref = F.relu(a16.float() @ w[:, :, 0].bfloat16().float().t() + bias) * scale + shift
    assert (out16.float() - ref).abs().max().item() < 3e-2


@pytest.mark.parametrize("k,d,C,N", [(3, 2, 128, 128), (3, 4, 128, 128), (5, 1, 128, 512), (3, 3, 64, 64)])
def test_dilated_taps_reflect_split(k, d, C, N):
    """k>1: row-shifted taps over mirrored halos == F.pad(reflect) + conv1d; split-bf16 (fp32 grade)."""
    nv = _native()
    dev = "cuda"
    frames = [50, 301, 77, 9]
    g = torch.Generator(device="cpu").manual_seed(k * 10 + d)
    xs = [torch.randn(T, C, generator=g).to(dev) for T in frames]
    w = (torch.randn(N, C, k, generator=g) / (C * k) ** 0.5).to(dev)
    bias = torch.randn(N, generator=g).to(dev) * 0.1
    scale = (torch.rand(N, generator=g) + 0.5).to(dev)
    shift = torch.randn(N, generator=g).to(dev) * 0.1
    xp = _padded_rows(xs, H)
    M = xp.shape[0]
    meta = nv.build_row_meta(frames, H, dev)
    assert meta.shape[0] == M
    x16 = nv.split_bf16(xp, planes=2)
    wp = nv.pack_conv_weight(w, planes=2)
    out = torch.full((M, 2 * N), float("nan"), dtype=torch.bfloat16, device=dev)
    nv.tdnn_layer(x16, wp, N, C, k, d, bias, scale, shift, out, act=nv.ACT_RELU, passes=3, w_planes=2,
                  x_plane_cols=C, x_cols=2 * C, row_meta=meta, store_mode=nv.STORE_MIRROR, halo=H,
                  out_kind=nv.OUT_BF16_HILO, out_plane_cols=N)
    torch.cuda.synchronize()
    got = out[:, :N].float() + out[:, N:].float()
    refs = []
    for x in xs:
        xt = x.double().t().unsqueeze(0)
        pad = d * (k - 1) // 2
        y = F.conv1d(F.pad(xt, (pad, pad), mode="reflect"), w.double(), bias.double(), dilation=d)
        y = F.relu(y) * scale.double()[None, :, None] + shift.double()[None, :, None]
        refs.append(y[0].t().float())
    ref_rows = _padded_rows(refs, H)
    assert torch.isfinite(got).all()
    err = (got - ref_rows).abs().max().item()
    assert err < 5e-5, err


def test_second_source_and_sum_output():
    """x2 (first K chunks from another matrix) and out2 = out + add_src (Res2Net running sum)."""
    nv = _native()
    dev = "cuda"
    frames = [120, 64]
    C, N, k, d = 128, 128, 3, 2
    g = torch.Generator(device="cpu").manual_seed(5)
    xs = [torch.randn(T, C, generator=g).to(dev) for T in frames]
    adds = [torch.randn(T, N, generator=g).to(dev) for T in frames]
    w = (torch.randn(N, C, k, generator=g) / (C * k) ** 0.5).to(dev)
    zero = torch.zeros(N, device=dev)
    one = torch.ones(N, device=dev)
    xp = _padded_rows(xs, H)
    ap = _padded_rows(adds, H)
    M = xp.shape[0]
    meta = nv.build_row_meta(frames, H, dev)
    x16 = nv.split_bf16(xp)
    a16 = nv.split_bf16(ap)
    x2 = x16.clone()
    x_main = x16.clone()
    x_main[:, :64] = 7.0  # must be ignored: chunk 0 comes from x2
    wp = nv.pack_conv_weight(w, planes=1)
    out = torch.zeros((M, N), dtype=torch.bfloat16, device=dev)
    out2 = torch.zeros((M, N), dtype=torch.bfloat16, device=dev)
    nv.tdnn_layer(x_main, wp, N, C, k, d, zero, one, zero, out, act=nv.ACT_NONE, x2=x2, x2_chunks=1,
                  row_meta=meta, store_mode=nv.STORE_MIRROR, halo=H, add_src=a16, out2=out2)
    torch.cuda.synchronize()
    refs = []
    for x in xs:
        xt = x.bfloat16().float().t().unsqueeze(0)
        y = F.conv1d(F.pad(xt, (d, d), mode="reflect"), w.bfloat16().float(), None, dilation=d)
        refs.append(y[0].t())
    ref = _padded_rows(refs, H)
    assert (out.float() - ref).abs().max().item() < 3e-2
    ref2 = _padded_rows([r + a.bfloat16().float() for r, a in zip(refs, adds)], H)
    assert (out2.float() - ref2).abs().max().item() < 5e-2
