"""CPU fp32 restatement of the reference's embedding path (test infrastructure).

Functional (state-dict in, tensors out) PyTorch code; each function cites the
reference file:line it restates.  Pinned against the reference's own outputs
by tests/test_oracle_cpu.py through the fixtures in tests/golden/.
"""
import math

import torch
import torch.nn.functional as F


# ------------------------------------------------------------------ front-end
def hz_to_mel_slaney(f):
    # torchaudio.functional.melscale_fbanks(mel_scale='slaney'), selected at
    # src/models/FeatureExtraction/feature.py:86-93
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = math.log(6.4) / 27.0
    if f >= min_log_hz:
        return min_log_mel + math.log(f / min_log_hz) / logstep
    return f / f_sp


def mel_to_hz_slaney(m):
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = math.log(6.4) / 27.0
    hz = f_sp * m
    log_t = m >= min_log_mel
    hz = torch.where(log_t, min_log_hz * torch.exp(logstep * (m - min_log_mel)), hz)
    return hz


def mel_filterbank(n_freqs=257, f_min=0.0, f_max=8000.0, n_mels=80, sr=16000):
    """[n_freqs, n_mels] slaney-scale, slaney-normalised triangles (feature.py:86-93)."""
    all_freqs = torch.linspace(0, sr // 2, n_freqs)
    m_pts = torch.linspace(hz_to_mel_slaney(f_min), hz_to_mel_slaney(f_max), n_mels + 2)
    f_pts = mel_to_hz_slaney(m_pts)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts.unsqueeze(0) - all_freqs.unsqueeze(1)
    down = (-1.0 * slopes[:, :-2]) / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    fb = torch.clamp(torch.min(down, up), min=0.0)
    enorm = 2.0 / (f_pts[2:n_mels + 2] - f_pts[:n_mels])
    return fb * enorm.unsqueeze(0)


def preemphasis(x, coef=0.97):
    """src/utils.py:63-71: reflect-pad one sample on the left, 2-tap FIR [-coef, 1]."""
    prev = torch.cat([x[:, 1:2], x[:, :-1]], dim=1)
    return x - coef * prev


def power_mel(x, sr=16000, n_fft=512, win_length=400, hop_length=160, n_mels=80,
              f_min=0.0, f_max=None, pre_emphasis=True, fb=None):
    """melspectrogram(lib='torchaudio') of feature.py:66-94 -> [B, n_mels, T] power-mel."""
    if pre_emphasis:
        x = preemphasis(x)
    pad = n_fft // 2
    xp = F.pad(x.unsqueeze(1), (pad, pad), mode="reflect").squeeze(1)
    frames = xp.unfold(-1, n_fft, hop_length)                       # [B, T, n_fft]
    win = torch.zeros(n_fft, device=x.device)
    lo = (n_fft - win_length) // 2
    win[lo:lo + win_length] = torch.hamming_window(win_length, periodic=True).to(x.device)
    spec = torch.fft.rfft(frames * win, dim=-1)
    power = spec.real ** 2 + spec.imag ** 2                         # [B, T, n_fft/2+1]
    if fb is None:
        fb = mel_filterbank(n_fft // 2 + 1, f_min, f_max if f_max else sr / 2.0, n_mels, sr)
    return torch.matmul(power, fb.to(x.device)).transpose(1, 2).contiguous()


def log_cmn(mel):
    """src/models/ECAPA_TDNN.py:473-476."""
    x = (mel + 1e-6).log()
    return x - x.mean(dim=-1, keepdim=True)


# ------------------------------------------------------------------ encoder
def input_instance_norm(x, weight, bias, eps=1e-5):
    """nn.InstanceNorm1d(affine=True, track_running_stats=False) on [B, C, T] (src/models/ECAPA_TDNN.py:406-408)."""
    m = x.mean(dim=2, keepdim=True)
    v = ((x - m) ** 2).mean(dim=2, keepdim=True)
    return (x - m) / torch.sqrt(v + eps) * weight[None, :, None] + bias[None, :, None]


def _conv_same_reflect(x, w, b, dilation):
    # src/models/layers/cnn.py:104-108,157 + get_padding_elem :787-805 (stride 1)
    k = w.shape[-1]
    pad = dilation * (k - 1) // 2
    if pad:
        x = F.pad(x, (pad, pad), mode="reflect")
    return F.conv1d(x, w, b, dilation=dilation)


def _bn(x, sd, p, eps=1e-5):
    # src/models/layers/normalization.py:62-68 (eval mode)
    scale = sd[p + "norm.weight"] / torch.sqrt(sd[p + "norm.running_var"] + eps)
    shift = sd[p + "norm.bias"] - sd[p + "norm.running_mean"] * scale
    return x * scale[None, :, None] + shift[None, :, None]


def tdnn_block(x, sd, p, dilation, act):
    # src/models/ECAPA_TDNN.py:68-69 : norm(activation(conv(x)))
    y = _conv_same_reflect(x, sd[p + "conv.conv.weight"], sd[p + "conv.conv.bias"], dilation)
    y = F.gelu(y) if act == "gelu" else F.relu(y)
    return _bn(y, sd, p + "norm.")


def res2net(x, sd, p, dilation, scale=8):
    # src/models/ECAPA_TDNN.py:118-129
    outs, prev = [], None
    for i, xi in enumerate(torch.chunk(x, scale, dim=1)):
        if i == 0:
            yi = xi
        elif i == 1:
            yi = tdnn_block(xi, sd, p + "blocks.%d." % (i - 1), dilation, "relu")
        else:
            yi = tdnn_block(xi + prev, sd, p + "blocks.%d." % (i - 1), dilation, "relu")
        prev = yi
        outs.append(yi)
    return torch.cat(outs, dim=1)


def se_block(x, sd, p):
    # src/models/ECAPA_TDNN.py:164-177 with lengths=None
    s = x.mean(dim=2, keepdim=True)
    s = F.relu(F.conv1d(s, sd[p + "conv1.conv.weight"], sd[p + "conv1.conv.bias"]))
    s = torch.sigmoid(F.conv1d(s, sd[p + "conv2.conv.weight"], sd[p + "conv2.conv.bias"]))
    return s * x


def se_res2net_block(x, sd, p, dilation):
    # src/models/ECAPA_TDNN.py:326-336 (no shortcut conv: in == out channels)
    y = tdnn_block(x, sd, p + "tdnn1.", 1, "gelu")
    y = res2net(y, sd, p + "res2net_block.", dilation)
    y = tdnn_block(y, sd, p + "tdnn2.", 1, "gelu")
    y = se_block(y, sd, p + "se_block.")
    return y + x


def attentive_stats_pool(x, sd, p, eps=1e-12):
    # src/models/ECAPA_TDNN.py:213-260 with lengths=None (mask of ones)
    L = x.shape[-1]

    def stats(x, m):
        mean = (m * x).sum(2)
        std = torch.sqrt((m * (x - mean.unsqueeze(2)).pow(2)).sum(2).clamp(eps))
        return mean, std

    if sd[p + "tdnn.conv.conv.weight"].shape[1] == x.shape[1]:
        attn = x                                       # global_context=False (:204-207, :247-248)
    else:
        m0 = torch.full((x.shape[0], 1, L), 1.0 / L, device=x.device, dtype=x.dtype)
        mean, std = stats(x, m0)
        attn = torch.cat([x, mean.unsqueeze(2).expand(-1, -1, L), std.unsqueeze(2).expand(-1, -1, L)], 1)
    attn = tdnn_block(attn, sd, p + "tdnn.", 1, "relu")
    attn = F.conv1d(torch.tanh(attn), sd[p + "conv.conv.weight"], sd[p + "conv.conv.bias"])
    attn = F.softmax(attn, dim=2)
    mean, std = stats(x, attn)
    return torch.cat([mean, std], dim=1).unsqueeze(2)


def ecapa_forward(feat, sd, dilations=(1, 2, 3, 4, 1), prefix="__S__.", stages=None):
    """src/models/ECAPA_TDNN.py:460-502: [B, n_mels, T] power-mel -> [B, nOut]."""
    sd = {k: torch.as_tensor(v) for k, v in sd.items()}
    x = log_cmn(feat)
    if prefix + "instance_norm.weight" in sd:
        # input_norm=True (src/models/ECAPA_TDNN.py:406-408,477-478): nn.InstanceNorm1d(n_mels, affine=True, eps=1e-5,
        # track_running_stats=False) = per (utterance, mel) biased statistics over time
        x = input_instance_norm(x, sd[prefix + "instance_norm.weight"], sd[prefix + "instance_norm.bias"])
    if stages is not None:
        stages["logmel"] = x
    xl = []
    x = tdnn_block(x, sd, prefix + "blocks.0.", dilations[0], "gelu")
    xl.append(x)
    for i in (1, 2, 3):
        x = se_res2net_block(x, sd, prefix + "blocks.%d." % i, dilations[i])
        xl.append(x)
    if stages is not None:
        for i, t in enumerate(xl):
            stages["block%d" % i] = t
    x = torch.cat(xl[1:], dim=1)
    x = tdnn_block(x, sd, prefix + "mfa.", dilations[4], "gelu")
    if stages is not None:
        stages["mfa"] = x
    x = attentive_stats_pool(x, sd, prefix + "asp.")
    if stages is not None:
        stages["asp"] = x
    x = _bn(x, sd, prefix + "asp_bn.")
    x = F.conv1d(x, sd[prefix + "fc.conv.weight"], sd[prefix + "fc.conv.bias"])
    return x.squeeze(-1)


def embed(wave, sd, stages=None, **fb_kw):
    """src/model.py:104-129: waveform rows [B, L] -> embeddings [B, nOut] (fp32; on the device `wave` and `sd` live on:
    the CPU for parity work, a CUDA device for bench.py's library-kernel baseline leg)."""
    with torch.no_grad():
        wave = torch.as_tensor(wave, dtype=torch.float32)
        mel = power_mel(wave, **fb_kw)
        if stages is not None:
            stages["mel"] = mel
        return ecapa_forward(mel, sd, stages=stages)


# ------------------------------------------------------------------ loader
def load_wav_eval(audio, sentence_len_s, sr, num_eval):
    """Eval-mode cropping of src/processing/audio_loader.py:110-150 on an in-memory signal."""
    import numpy as np
    max_audio = int(sentence_len_s * sr)
    n = audio.shape[0]
    if n <= max_audio:
        audio = np.pad(audio, (0, max_audio - n + 1), "wrap")
        n = audio.shape[0]
    if num_eval == 0:
        return audio[None].astype(np.float32)
    starts = np.linspace(0, n - max_audio, num=num_eval)
    return np.stack([audio[int(s):int(s) + max_audio] for s in starts]).astype(np.float32)
