"""Deterministic synthetic inputs for parity tests and the bench (test infrastructure).

Everything is driven by numpy's PCG64 (platform independent), so the GPU box
regenerates bit-identical waveforms / weights / trial lists from seeds alone and
only the reference OUTPUTS need to be committed under tests/golden/.
Follows SURVEY.md section 8d (speaker-like harmonic generator, seeded trials).
"""
import numpy as np

SR = 16000


def speaker_bank(n_speakers, seed=0):
    """Per-speaker pitch and 40-harmonic spectral envelope."""
    rng = np.random.default_rng(seed)
    f0 = rng.uniform(80.0, 300.0, size=n_speakers)
    env = rng.uniform(0.05, 1.0, size=(n_speakers, 40))
    return f0, env


def utterance(f0, env, n_samples, seed):
    """One 'speaker-like' utterance: jittered harmonic stack with vibrato, AM and noise."""
    rng = np.random.default_rng(seed)
    t = np.arange(n_samples, dtype=np.float64) / SR
    f = f0 * (1.0 + 0.03 * rng.standard_normal())
    vib_hz = rng.uniform(3.0, 7.0)
    am_hz = rng.uniform(1.0, 4.0)
    am_ph = rng.uniform(0, 2 * np.pi)
    phase = 2 * np.pi * f * t + (0.01 * f / vib_hz) * np.sin(2 * np.pi * vib_hz * t)
    sig = np.zeros(n_samples, dtype=np.float64)
    ph0 = rng.uniform(0, 2 * np.pi, size=env.shape[0])
    for h in range(env.shape[0]):
        if (h + 1) * f >= 7600.0:
            break
        sig += env[h] * np.sin((h + 1) * phase + ph0[h])
    sig *= 1.0 + 0.5 * np.sin(2 * np.pi * am_hz * t + am_ph)
    sig *= 0.3 / np.max(np.abs(sig))
    sig += 0.01 * rng.standard_normal(n_samples)
    return np.clip(sig, -1.0, 1.0).astype(np.float32)


def utterance_set(n_speakers, utts_per_speaker, n_samples, spk_seed=0, utt_seed0=1000):
    """[n_speakers*utts_per_speaker, n_samples] float32 and the speaker id of each row."""
    f0, env = speaker_bank(n_speakers, spk_seed)
    waves, spk = [], []
    for s in range(n_speakers):
        for u in range(utts_per_speaker):
            i = s * utts_per_speaker + u
            waves.append(utterance(f0[s], env[s], n_samples, utt_seed0 + i))
            spk.append(s)
    return np.stack(waves), np.asarray(spk)


def ragged_lengths(n, lo_s=2.0, hi_s=20.0, seed=2):
    rng = np.random.default_rng(seed)
    return (rng.uniform(lo_s, hi_s, size=n) * SR).astype(np.int64)


def white_noise(n, n_samples, seed=7):
    rng = np.random.default_rng(seed)
    return (0.1 * rng.standard_normal((n, n_samples))).astype(np.float32)


def trial_list(spk, n_trials, seed=11):
    """Seeded sample of unordered utterance pairs: (label, i, j) int arrays."""
    n = len(spk)
    iu, ju = np.triu_indices(n, k=1)
    rng = np.random.default_rng(seed)
    pick = rng.permutation(len(iu))[:n_trials]
    i, j = iu[pick], ju[pick]
    lab = (spk[i] == spk[j]).astype(np.int64)
    return lab, i, j


# --------------------------------------------------------------------------------------
# ECAPA-TDNN state dict in the reference's checkpoint layout
# --------------------------------------------------------------------------------------

def ecapa_param_shapes(channels, n_out, n_mels=80, scale=8, se_ch=128, att_ch=128,
                       kernel_sizes=(5, 3, 3, 3, 1), prefix="__S__."):
    """Ordered (key, shape, kind) list matching the reference ECAPA_TDNN state-dict."""
    out = []

    def conv(p, co, ci, k):
        out.append((p + "conv.weight", (co, ci, k), "w"))
        out.append((p + "conv.bias", (co,), "b"))

    def bn(p, c):
        out.append((p + "norm.weight", (c,), "g"))
        out.append((p + "norm.bias", (c,), "beta"))
        out.append((p + "norm.running_mean", (c,), "rm"))
        out.append((p + "norm.running_var", (c,), "rv"))
        out.append((p + "norm.num_batches_tracked", (), "nbt"))

    def tdnn(p, co, ci, k):
        conv(p + "conv.", co, ci, k)
        bn(p + "norm.", co)

    C = channels
    tdnn(prefix + "blocks.0.", C[0], n_mels, kernel_sizes[0])
    for i in range(1, len(C) - 1):
        p = prefix + "blocks.%d." % i
        tdnn(p + "tdnn1.", C[i], C[i - 1], 1)
        for j in range(scale - 1):
            tdnn(p + "res2net_block.blocks.%d." % j, C[i] // scale, C[i] // scale, kernel_sizes[i])
        tdnn(p + "tdnn2.", C[i], C[i], 1)
        conv(p + "se_block.conv1.", se_ch, C[i], 1)
        conv(p + "se_block.conv2.", C[i], se_ch, 1)
    tdnn(prefix + "mfa.", C[-1], C[-1], kernel_sizes[-1])
    tdnn(prefix + "asp.tdnn.", att_ch, 3 * C[-1], 1)
    conv(prefix + "asp.conv.", C[-1], att_ch, 1)
    bn(prefix + "asp_bn.", 2 * C[-1])
    conv(prefix + "fc.", n_out, 2 * C[-1], 1)
    return out


def ecapa_state_dict(channels, n_out, seed=1234, bn_stats=None, **kw):
    """Random ECAPA weights (numpy, float32) under the reference's key names.

    ``bn_stats`` (key -> array) overrides running_mean / running_var (the
    BN-calibrated statistics committed with the golden fixtures).
    """
    rng = np.random.default_rng(seed)
    sd = {}
    for key, shape, kind in ecapa_param_shapes(channels, n_out, **kw):
        if kind == "w":
            fan_in = shape[1] * shape[2]
            v = rng.uniform(-1.0, 1.0, size=shape) * np.sqrt(3.0 / fan_in)
        elif kind == "b":
            v = rng.uniform(-0.1, 0.1, size=shape)
        elif kind == "g":
            v = rng.uniform(0.5, 1.5, size=shape)
        elif kind == "beta":
            v = 0.1 * rng.standard_normal(shape)
        elif kind == "rm":
            v = 0.1 * rng.standard_normal(shape)
        elif kind == "rv":
            v = rng.uniform(0.5, 1.5, size=shape)
        else:
            sd[key] = np.asarray(1, dtype=np.int64)
            continue
        sd[key] = v.astype(np.float32)
    if bn_stats:
        for k, v in bn_stats.items():
            assert k in sd and sd[k].shape == v.shape, k
            sd[k] = np.asarray(v, dtype=np.float32)
    return sd


def scoring_embeddings(n_utts, n_speakers, dim=192, noise=0.7, seed=3):
    """cfg5-style unit embeddings around speaker centroids."""
    rng = np.random.default_rng(seed)
    cent = rng.standard_normal((n_speakers, dim))
    cent /= np.linalg.norm(cent, axis=1, keepdims=True)
    spk = rng.integers(0, n_speakers, size=n_utts)
    e = cent[spk] + noise * rng.standard_normal((n_utts, dim)) / np.sqrt(dim)
    e /= np.linalg.norm(e, axis=1, keepdims=True)
    return e.astype(np.float32), spk


# --------------------------------------------------------------------------------------
# RawNet2 state dict in the reference's checkpoint layout
# --------------------------------------------------------------------------------------
RAWNET_LAYERS = [1, 1, 1, 2, 1, 2]
RAWNET_FILTERS = [128, 128, 256, 256, 512, 512]


def rawnet2_param_shapes(n_out=512, n_samples=64000, gru_node=1024, prefix="__S__.", aggregate="gru", att_dim=128):
    """Ordered (key, shape, kind) list matching RawNet2_custom.MainModel's state-dict (144 entries)."""
    out = []

    def bn(p, c):
        out.append((p + "weight", (c,), "g"))
        out.append((p + "bias", (c,), "beta"))
        out.append((p + "running_mean", (c,), "rm"))
        out.append((p + "running_var", (c,), "rv"))
        out.append((p + "num_batches_tracked", (), "nbt"))

    P = prefix
    out.append((P + "ln.gamma", (n_samples,), "g"))
    out.append((P + "ln.beta", (n_samples,), "beta"))
    out.append((P + "first_conv.low_hz_", (RAWNET_FILTERS[0], 1), "sinc_low"))
    out.append((P + "first_conv.band_hz_", (RAWNET_FILTERS[0], 1), "sinc_band"))
    bn(P + "first_bn.", RAWNET_FILTERS[0])
    inpl = RAWNET_FILTERS[0]
    for li, (nb, planes) in enumerate(zip(RAWNET_LAYERS, RAWNET_FILTERS)):
        for bi in range(nb):
            p = P + "layer%d.%d." % (li + 1, bi)
            bn(p + "bn1.", inpl)
            out.append((p + "conv1.weight", (planes, inpl, 3), "w"))
            bn(p + "bn2.", planes)
            out.append((p + "conv2.weight", (planes, planes, 3), "w"))
            out.append((p + "afms.alpha", (planes, 1), "g"))
            out.append((p + "afms.fc.weight", (planes, planes), "lin"))
            out.append((p + "afms.fc.bias", (planes,), "b"))
            if inpl != planes:
                out.append((p + "shortcut.0.weight", (planes, inpl, 1), "w"))
            inpl = planes
    C5 = RAWNET_FILTERS[5]
    if aggregate == "gru":
        bn(P + "bn_before_gru.", C5)
        out.append((P + "gru.weight_ih_l0", (3 * gru_node, C5), "gru"))
        out.append((P + "gru.weight_hh_l0", (3 * gru_node, gru_node), "gru"))
        out.append((P + "gru.bias_ih_l0", (3 * gru_node,), "gru"))
        out.append((P + "gru.bias_hh_l0", (3 * gru_node,), "gru"))
        out.append((P + "fc_after_gru.weight", (n_out, gru_node), "lin"))
        out.append((P + "fc_after_gru.bias", (n_out,), "b"))
    elif aggregate == "sap":                                  # RawNet2_custom.py:97-100, RawNet_baseline.py:423-429
        bn(P + "bn_before_agg.", C5)
        out.append((P + "attention_sap.v", (C5,), "vec"))
        out.append((P + "attention_sap.lin_proj.weight", (C5, C5), "lin"))
        out.append((P + "attention_sap.lin_proj.bias", (C5,), "b"))
    else:                                                     # 'asp', RawNet2_custom.py:102-111
        bn(P + "bn_before_agg.", C5)
        out.append((P + "attention.0.weight", (att_dim, C5, 1), "w"))
        out.append((P + "attention.0.bias", (att_dim,), "b"))
        bn(P + "attention.2.", att_dim)
        out.append((P + "attention.3.weight", (C5, att_dim, 1), "w"))
        out.append((P + "attention.3.bias", (C5,), "b"))
    out.append((P + "fc.weight", (n_out, 2 * RAWNET_FILTERS[5]), "lin"))
    out.append((P + "fc.bias", (n_out,), "b"))
    return out


def rawnet2_state_dict(n_out=512, n_samples=64000, gru_node=1024, seed=4321, bn_stats=None, sample_rate=16000,
                       aggregate="gru", att_dim=128):
    """Random RawNet2 weights (numpy float32) under the reference's key names; sinc bands start from the
    reference's mel-spaced initialisation (RawNet_baseline.py:296-309) with a seeded perturbation."""
    rng = np.random.default_rng(seed)
    sd = {}
    to_mel = lambda hz: 2595 * np.log10(1 + hz / 700)
    to_hz = lambda mel: 700 * (10 ** (mel / 2595) - 1)
    hz = to_hz(np.linspace(to_mel(10), to_mel(sample_rate / 2 - 100), RAWNET_FILTERS[0] + 1))
    for key, shape, kind in rawnet2_param_shapes(n_out, n_samples, gru_node, aggregate=aggregate, att_dim=att_dim):
        if kind == "w":
            fan_in = shape[1] * shape[2]
            v = rng.uniform(-1.0, 1.0, size=shape) * np.sqrt(4.5 / fan_in)
        elif kind == "lin":
            v = rng.uniform(-1.0, 1.0, size=shape) * np.sqrt(3.0 / shape[1])
        elif kind == "gru":
            v = rng.uniform(-1.0, 1.0, size=shape) / np.sqrt(gru_node)
        elif kind == "vec":
            v = rng.standard_normal(shape) / np.sqrt(shape[0]) * 4.0
        elif kind == "b":
            v = rng.uniform(-0.1, 0.1, size=shape)
        elif kind == "g":
            v = rng.uniform(0.5, 1.5, size=shape)
        elif kind == "beta":
            v = 0.1 * rng.standard_normal(shape)
        elif kind == "rm":
            v = 0.1 * rng.standard_normal(shape)
        elif kind == "rv":
            v = rng.uniform(0.5, 1.5, size=shape)
        elif kind == "sinc_low":
            v = (hz[:-1] * rng.uniform(0.9, 1.1, size=hz[:-1].shape)).reshape(shape)
        elif kind == "sinc_band":
            v = (np.diff(hz) * rng.uniform(0.8, 1.2, size=hz[:-1].shape)).reshape(shape)
        else:
            sd[key] = np.asarray(1, dtype=np.int64)
            continue
        sd[key] = v.astype(np.float32)
    if bn_stats:
        for k, v in bn_stats.items():
            assert k in sd and sd[k].shape == v.shape, k
            sd[k] = np.asarray(v, dtype=np.float32)
    return sd
