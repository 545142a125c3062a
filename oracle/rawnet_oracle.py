"""CPU fp32 restatement of the reference's RawNet2 raw-waveform encoder (test infrastructure).

Functional PyTorch code (state dict in, tensors out); every function cites the reference
file:line it restates (paths under /root/reference/src/models).  Pinned against the reference's
own outputs by tests/test_rawnet_oracle_cpu.py through tests/golden/cfg4_rawnet2.npz.
Covers the default configuration of RawNet2_custom.MainModel (front_proc='sinc', aggregate='gru',
layers [1,1,1,2,1,2], filters [128,128,256,256,512,512], RawNet2_custom.py:230-243).
"""
import math

import torch
import torch.nn.functional as F

LAYERS = [1, 1, 1, 2, 1, 2]
FILTERS = [128, 128, 256, 256, 512, 512]


def layer_norm_samples(x, gamma, beta, eps=1e-6):
    """RawNet_baseline.py:13-24: per-utterance normalisation over SAMPLES, unbiased std, eps added to std."""
    mean = x.mean(-1, keepdim=True)
    std = x.std(-1, keepdim=True)
    return gamma * (x - mean) / (std + eps) + beta


def sinc_filters(low_hz_, band_hz_, kernel_size=251, sample_rate=16000, min_low_hz=50, min_band_hz=50):
    """RawNet_baseline.py:309-357: [C, kernel_size] band-pass FIRs rebuilt from (low_hz_, band_hz_)."""
    n_lin = torch.linspace(0, (kernel_size / 2) - 1, steps=int(kernel_size / 2))
    window = 0.54 - 0.46 * torch.cos(2 * math.pi * n_lin / kernel_size)
    n = (kernel_size - 1) / 2.0
    n_ = 2 * math.pi * torch.arange(-n, 0).view(1, -1) / sample_rate
    low = min_low_hz + torch.abs(low_hz_)
    high = torch.clamp(low + min_band_hz + torch.abs(band_hz_), min_low_hz, sample_rate / 2)
    band = (high - low)[:, 0]
    f_low = torch.matmul(low, n_)
    f_high = torch.matmul(high, n_)
    left = ((torch.sin(f_high) - torch.sin(f_low)) / (n_ / 2)) * window
    center = 2 * band.view(-1, 1)
    right = torch.flip(left, dims=[1])
    bp = torch.cat([left, center, right], dim=1)
    return bp / (2 * band[:, None])


def _bn(x, sd, p, eps=1e-5):
    scale = sd[p + "weight"] / torch.sqrt(sd[p + "running_var"] + eps)
    shift = sd[p + "bias"] - sd[p + "running_mean"] * scale
    return x * scale[None, :, None] + shift[None, :, None]


def afms(x, sd, p):
    """RawNet_baseline.py:48-68: y = sigmoid(fc(mean_T x)); (x + alpha) * y."""
    y = x.mean(dim=2)
    y = torch.sigmoid(F.linear(y, sd[p + "fc.weight"], sd[p + "fc.bias"])).unsqueeze(-1)
    return (x + sd[p + "alpha"]) * y


def basic_block(x, sd, p, downsample):
    """RawNet_baseline.py:181-232 (pre-activation block, zero padding, no conv bias)."""
    out = F.leaky_relu(_bn(x, sd, p + "bn1."), 0.3)
    if (p + "shortcut.0.weight") in sd:
        shortcut = F.conv1d(out, sd[p + "shortcut.0.weight"])
    else:
        shortcut = x
    out = F.conv1d(out, sd[p + "conv1.weight"], padding=1)
    out = F.conv1d(F.leaky_relu(_bn(out, sd, p + "bn2."), 0.3), sd[p + "conv2.weight"], padding=1)
    out = out + shortcut
    if downsample:
        out = F.max_pool1d(out, 3)
    return afms(out, sd, p + "afms.")


def gru_last(x, sd, p):
    """nn.GRU(batch_first=True, one layer) last hidden state (RawNet2_custom.py:188-197).  x [B, T, I]."""
    w_ih, w_hh = sd[p + "weight_ih_l0"], sd[p + "weight_hh_l0"]
    b_ih, b_hh = sd[p + "bias_ih_l0"], sd[p + "bias_hh_l0"]
    H = w_hh.shape[1]
    h = torch.zeros(x.shape[0], H, dtype=x.dtype)
    for t in range(x.shape[1]):
        gi = F.linear(x[:, t], w_ih, b_ih)
        gh = F.linear(h, w_hh, b_hh)
        r = torch.sigmoid(gi[:, :H] + gh[:, :H])
        z = torch.sigmoid(gi[:, H:2 * H] + gh[:, H:2 * H])
        n = torch.tanh(gi[:, 2 * H:] + r * gh[:, 2 * H:])
        h = (1 - z) * n + z * h
    return h


def rawnet2_forward(wave, sd, prefix="__S__.", stages=None, sample_rate=16000):
    """RawNet2.forward (RawNet2_custom.py:161-227), sinc front-end + GRU aggregation.  wave [B, L] -> [B, nOut]."""
    sd = {k: torch.as_tensor(v) for k, v in sd.items()}
    x = torch.as_tensor(wave, dtype=torch.float32)
    P = prefix
    x = layer_norm_samples(x, sd[P + "ln.gamma"], sd[P + "ln.beta"])                       # :171
    filt = sinc_filters(sd[P + "first_conv.low_hz_"], sd[P + "first_conv.band_hz_"], sample_rate=sample_rate)
    x = F.conv1d(x.unsqueeze(1), filt.unsqueeze(1))                                       # :172-174
    x = F.max_pool1d(torch.abs(x), 3)
    x = F.leaky_relu(_bn(x, sd, P + "first_bn."), 0.3)                                    # :175-176
    if stages is not None:
        stages["front"] = x
    for li, nb in enumerate(LAYERS):                                                      # :181-186
        for bi in range(nb):
            x = basic_block(x, sd, P + "layer%d.%d." % (li + 1, bi), downsample=(bi == nb - 1))
        if stages is not None:
            stages["layer%d" % (li + 1)] = x
    if P + "bn_before_agg.weight" in sd:
        return attentive_head(x, sd, P).squeeze()                                         # :200-224
    x = F.leaky_relu(_bn(x, sd, P + "bn_before_gru."), 0.3)                               # :192-193
    h = gru_last(x.permute(0, 2, 1), sd, P + "gru.")                                      # :194-197
    if stages is not None:
        stages["gru"] = h
    out = F.linear(h, sd[P + "fc_after_gru.weight"], sd[P + "fc_after_gru.bias"])         # :198
    return out.squeeze()


def attentive_head(x, sd, P):
    """aggregate = 'sap' (src/models/RawNet2_custom.py:200-207 with Classic_Attention, RawNet_baseline.py:423-436, and
    stat_attn_pool / weighted_sd :146-157) or 'asp' (:210-224), selected by the checkpoint's keys; x [B, 512, T]."""
    x = F.leaky_relu(_bn(x, sd, P + "bn_before_agg."), 0.3)
    if P + "attention_sap.v" in sd:
        xt = x.permute(0, 2, 1)                                                            # [B, T, C]
        lin = F.linear(xt, sd[P + "attention_sap.lin_proj.weight"], sd[P + "attention_sap.lin_proj.bias"])
        w = F.softmax(torch.tanh(lin @ sd[P + "attention_sap.v"]), dim=1)                  # [B, T]
        el = xt * w.unsqueeze(2)
        mean = el.mean(dim=1)                                                              # torch.mean over time (:153)
        var = (xt * el).sum(dim=1) - mean * mean                                           # weighted_sd (:146-150)
        pooled = torch.cat([mean, var], dim=1)
    else:
        h = F.conv1d(x, sd[P + "attention.0.weight"], sd[P + "attention.0.bias"])
        h = _bn(F.leaky_relu(h, 0.01), sd, P + "attention.2.")
        w = F.softmax(F.conv1d(h, sd[P + "attention.3.weight"], sd[P + "attention.3.bias"]), dim=-1)
        m = (x * w).sum(dim=-1)
        s = torch.sqrt((((x ** 2) * w).sum(dim=-1) - m ** 2).clamp(min=1e-5))
        pooled = torch.cat([m, s], dim=1)
    return F.linear(pooled, sd[P + "fc.weight"], sd[P + "fc.bias"])


# ------------------------------------------------------------------ 16-bit-operand emulation
def _r(t):
    return t.to(torch.float16).to(torch.float32)


def rawnet2_forward_f16emu(wave, sd, prefix="__S__.", sample_rate=16000):
    """The same network with every stored activation and every tensor-core weight rounded to IEEE half exactly
    where libsvb's 16-bit mode rounds (fp32 accumulation, fp32 BN / AFMS / GRU-gate arithmetic, fp32-grade GRU
    recurrence).  It separates two things the fp32 comparison mixes: whether
    the kernels compute what bf16 operands imply (this function, tight tolerance) and how far bf16 operands drift
    from fp32 on a given checkpoint (a property of the checkpoint, loose tolerance)."""
    sd = {k: torch.as_tensor(v) for k, v in sd.items()}
    P = prefix
    x = torch.as_tensor(wave, dtype=torch.float32)
    x = layer_norm_samples(x, sd[P + "ln.gamma"], sd[P + "ln.beta"])
    x = _r(x)
    filt = _r(sinc_filters(sd[P + "first_conv.low_hz_"], sd[P + "first_conv.band_hz_"], sample_rate=sample_rate))
    v = F.max_pool1d(torch.abs(F.conv1d(x.unsqueeze(1), filt.unsqueeze(1))), 3)
    v = F.leaky_relu(_bn(v, sd, P + "first_bn."), 0.3)
    names = [(li + 1, bi, bi == nb - 1) for li, nb in enumerate(LAYERS) for bi in range(nb)]
    pre = [P + "layer%d.%d.bn1." % (l, b) for l, b, _ in names] + [P + "bn_before_gru."]
    x_r, a_r = _r(v), _r(F.leaky_relu(_bn(v, sd, pre[0]), 0.3))
    for i, (l, b, down) in enumerate(names):
        p = P + "layer%d.%d." % (l, b)
        h = _r(F.leaky_relu(_bn(F.conv1d(a_r, _r(sd[p + "conv1.weight"]), padding=1), sd, p + "bn2."), 0.3))
        sc = _r(F.conv1d(a_r, _r(sd[p + "shortcut.0.weight"]))) if (p + "shortcut.0.weight") in sd else x_r
        z = _r(F.conv1d(h, _r(sd[p + "conv2.weight"]), padding=1) + sc)
        if down:
            z = F.max_pool1d(z, 3)
        y = torch.sigmoid(F.linear(z.mean(dim=2), sd[p + "afms.fc.weight"], sd[p + "afms.fc.bias"])).unsqueeze(-1)
        v = (z + sd[p + "afms.alpha"]) * y
        x_r, a_r = _r(v), _r(F.leaky_relu(_bn(v, sd, pre[i + 1]), 0.3))
    gsd = dict(sd)
    gsd[P + "gru.weight_ih_l0"] = _r(sd[P + "gru.weight_ih_l0"])
    h = gru_last(a_r.permute(0, 2, 1), gsd, P + "gru.")
    return F.linear(h, sd[P + "fc_after_gru.weight"], sd[P + "fc_after_gru.bias"]).squeeze()
