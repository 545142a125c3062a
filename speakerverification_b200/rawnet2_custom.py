"""Encoder plugin: host-side mirror of the reference's ``models.RawNet2_custom.MainModel``
(src/models/RawNet2_custom.py:18-243, layers in src/models/RawNet_baseline.py) for the default
configuration ``front_proc='sinc'``, ``aggregate='gru'``.

The module tree exposes parameters and buffers under the reference's names and shapes (144 state-dict
entries: ``ln.gamma``, ``first_conv.low_hz_``, ``layer3.0.shortcut.0.weight``, ``gru.weight_hh_l0`` ...)
so checkpoints load and save unchanged.  None of the ``nn`` holders is ever called: the forward pass is
libsvb (sliding-window tcgen05 GEMM for the sinc layer, implicit-GEMM convs, fused pooling / AFMS / GRU
kernels).  Input: raw waveform ``[B, L]`` with ``L = sentence_len * sample_rate`` exactly (the reference's
LayerNorm weights have that length); ``features`` must be ``'raw'`` (src/model.py:111-112).
"""
import ctypes

import numpy as np
import torch
import torch.nn as nn

from . import _native as nv

_vp = ctypes.c_void_p
LAYERS = [1, 1, 1, 2, 1, 2]
FILTERS = [128, 128, 256, 256, 512, 512]


class LayerNorm(nn.Module):          # RawNet_baseline.py:13-24 (parameter holder)
    def __init__(self, features, eps=1e-6):
        super().__init__()
        self.gamma = nn.Parameter(torch.ones(features))
        self.beta = nn.Parameter(torch.zeros(features))
        self.eps = eps


class SincConv_fast(nn.Module):      # RawNet_baseline.py:235-361 (parameter holder, mel-spaced initialisation)
    def __init__(self, out_channels, kernel_size, sample_rate=16000, min_low_hz=50, min_band_hz=50):
        super().__init__()
        if kernel_size % 2 == 0:
            kernel_size += 1
        self.out_channels, self.kernel_size, self.sample_rate = out_channels, kernel_size, sample_rate
        to_mel = lambda hz: 2595 * np.log10(1 + hz / 700)
        to_hz = lambda mel: 700 * (10 ** (mel / 2595) - 1)
        hz = to_hz(np.linspace(to_mel(10), to_mel(sample_rate / 2 - (min_low_hz + min_band_hz)), out_channels + 1))
        self.low_hz_ = nn.Parameter(torch.Tensor(hz[:-1]).view(-1, 1))
        self.band_hz_ = nn.Parameter(torch.Tensor(np.diff(hz)).view(-1, 1))


class AFMS(nn.Module):               # RawNet_baseline.py:48-68
    def __init__(self, nb_dim):
        super().__init__()
        self.alpha = nn.Parameter(torch.ones((nb_dim, 1)))
        self.fc = nn.Linear(nb_dim, nb_dim)


class RawNetBasicBlock(nn.Module):   # RawNet_baseline.py:181-232
    def __init__(self, inplanes, planes, downsample=None):
        super().__init__()
        self.downsample = downsample
        self.bn1 = nn.BatchNorm1d(inplanes)
        self.conv1 = nn.Conv1d(inplanes, planes, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn2 = nn.BatchNorm1d(planes)
        self.conv2 = nn.Conv1d(planes, planes, kernel_size=3, stride=1, padding=1, bias=False)
        self.afms = AFMS(planes)
        if inplanes != planes:
            self.shortcut = nn.Sequential(nn.Conv1d(inplanes, planes, kernel_size=1, stride=1, bias=False))


class Classic_Attention(nn.Module):   # RawNet_baseline.py:423-436 (parameter holder of the 'sap' head)
    def __init__(self, input_dim, embed_dim):
        super().__init__()
        self.lin_proj = nn.Linear(input_dim, embed_dim)
        self.v = nn.Parameter(torch.randn(embed_dim))


class RawNet2Engine:
    """Packed RawNet2 weights + forward (svb_rawnet2_*).  `state` maps reference state-dict names (without
    the '__S__.' prefix) to fp32 CUDA tensors."""

    AGGREGATES = {"gru": 0, "sap": 1, "asp": 2}

    def __init__(self, state, n_samples, n_out, gru_node=1024, sample_rate=16000, precision="f16", aggregate="gru",
                 att_dim=128):
        from .engine import normalize_precision
        precision = normalize_precision(precision)
        self.n_samples, self.n_out, self.precision = int(n_samples), int(n_out), precision
        cfg = nv.SvbRawnet2Cfg(self.n_samples, self.n_out, int(gru_node), int(sample_rate), 1 if precision == "fp32" else 0,
                               self.AGGREGATES[aggregate], int(att_dim))
        keep, names, ptrs = [], [], []
        for k, v in state.items():
            if not torch.is_tensor(v) or not v.is_floating_point():
                continue
            t = v.detach().to(dtype=torch.float32).contiguous()
            if not t.is_cuda:
                raise RuntimeError("RawNet2 weights must live on the GPU (tensor '%s' is on %s)" % (k, t.device))
            keep.append(t)
            names.append(k.encode())
            ptrs.append(t.data_ptr())
        self.device = keep[0].device
        h = _vp()
        torch.cuda.synchronize(self.device)
        with nv.on_device(self.device):      # the library allocates / launches on the current device
            nv.check(nv.lib().svb_rawnet2_create(ctypes.byref(cfg), (ctypes.c_char_p * len(names))(*names),
                                                 (_vp * len(ptrs))(*ptrs), len(names), ctypes.byref(h)))
        self.handle = h
        self.planes = 2 if precision == "fp32" else 1
        self._ws = None

    def workspace(self, B):
        need = nv.lib().svb_rawnet2_workspace_bytes(self.handle, B)
        if self._ws is None or self._ws.numel() < need:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._ws

    def forward_flat(self, *a, **k):
        raise NotImplementedError("RawNet2 embeds fixed-length crops only (ln.gamma / ln.beta have n_samples entries, "
                                  "src/models/RawNet_baseline.py:13-24): whole-utterance embedding (num_eval=0) has no "
                                  "meaning for this encoder")

    @nv.device_method
    def forward(self, wav, l2norm=False, out=None):
        """wav: [B, n_samples] fp32 CUDA tensor -> [B, n_out] fp32."""
        if not wav.is_cuda or wav.dtype != torch.float32:
            raise RuntimeError("wave input must be an fp32 CUDA tensor")
        if wav.dim() != 2 or wav.shape[1] != self.n_samples:
            raise RuntimeError("RawNet2 needs crops of exactly %d samples (ln.gamma / ln.beta have that length); got %s"
                               % (self.n_samples, tuple(wav.shape)))
        if wav.stride(1) != 1:
            wav = wav.contiguous()
        B = wav.shape[0]
        ws = self.workspace(B)
        if out is None:
            out = torch.empty((B, self.n_out), dtype=torch.float32, device=self.device)
        nv.check(nv.lib().svb_rawnet2_forward(self.handle, nv.ptr(wav), wav.stride(0), B, nv.ptr(ws), ws.numel(),
                                              nv.ptr(out), out.stride(0), int(l2norm), nv.stream_ptr()))
        return out

    def frames(self, level):
        return int(nv.lib().svb_rawnet2_frames(self.handle, level))

    def stage(self, B, name, level, channels):
        """fp32 copy of a bf16 (hi|lo) activation inside the workspace: [B, T, C] (tests, profiling)."""
        off = nv.lib().svb_rawnet2_ws_offset(self.handle, B, name.encode())
        if off < 0:
            raise KeyError(name)
        T = self.frames(level)
        planes = self.planes
        n = B * (T + 2) * planes * channels
        dt = torch.bfloat16 if planes == 2 else torch.float16        # bf16 hi|lo planes / one plane of IEEE half
        v = self._ws[off:off + 2 * n].view(dt).view(B, T + 2, planes, channels).float().sum(2)
        return v[:, 1:T + 1]

    def __del__(self):
        if getattr(self, "handle", None) and nv is not None and getattr(nv, "_lib", None) is not None:   # nv is None at interpreter shutdown
            nv._lib.svb_rawnet2_destroy(self.handle)
            self.handle = None


class RawNet2(nn.Module):
    def __init__(self, block, layers, nb_filters, audio_spec, front_proc="sinc", aggregate="gru", att_dim=128,
                 code_dim=512, in_channels=1, log_input=True, nb_gru_layers=1, gru_node=1024, first_conv_size=251,
                 precision=None, **kwargs):
        super().__init__()
        if front_proc != "sinc" or aggregate not in ("gru", "sap", "asp") or nb_gru_layers != 1 or in_channels != 1:
            raise NotImplementedError("the libsvb RawNet2 path covers front_proc='sinc' with aggregate 'gru' / 'sap' / 'asp'")
        if str(kwargs.get("features", "raw")).strip().lower() != "raw":
            raise NotImplementedError("RawNet2 consumes raw waveforms: set features: raw")
        if list(layers) != LAYERS or list(nb_filters) != FILTERS or first_conv_size != 251:
            raise NotImplementedError("only the MainModel layout %s / %s is built" % (LAYERS, FILTERS))
        sample_rate = audio_spec["sample_rate"]
        self.sample_rate = int(sample_rate)
        self.n_samples = int(audio_spec["sentence_len"] * sample_rate)
        self.n_out, self.gru_node = code_dim, gru_node
        self.inplanes = nb_filters[0]
        self.ln = LayerNorm(self.n_samples)
        self.first_conv = SincConv_fast(out_channels=nb_filters[0], kernel_size=first_conv_size,
                                        sample_rate=self.sample_rate)
        self.first_bn = nn.BatchNorm1d(nb_filters[0])
        for i in range(6):
            setattr(self, "layer%d" % (i + 1), self._make_layer(block, nb_filters[i], layers[i]))
        self.aggregate, self.att_dim = aggregate, att_dim
        if aggregate == "gru":                                # src/models/RawNet2_custom.py:84-112
            self.bn_before_gru = nn.BatchNorm1d(nb_filters[5])
            self.gru = nn.GRU(input_size=nb_filters[5], hidden_size=gru_node, num_layers=1, batch_first=True)
            self.fc_after_gru = nn.Linear(gru_node, code_dim)
        elif aggregate == "sap":
            self.bn_before_agg = nn.BatchNorm1d(nb_filters[5])
            self.attention_sap = Classic_Attention(nb_filters[5], nb_filters[5])
        else:
            self.bn_before_agg = nn.BatchNorm1d(nb_filters[5])
            self.attention = nn.Sequential(nn.Conv1d(nb_filters[5], att_dim, kernel_size=1), nn.LeakyReLU(),
                                           nn.BatchNorm1d(att_dim), nn.Conv1d(att_dim, nb_filters[5], kernel_size=1),
                                           nn.Softmax(dim=-1))
        self.fc = nn.Linear(nb_filters[5] * 2, code_dim)      # the sap / asp embedding layer; unused by the gru head
        if precision is None:
            # the reference's `mixedprec` only autocasts the TRAINING step (src/model.py:256); evaluateFromList /
            # testFromList always run fp32, so fp32-grade is the default and bf16 is opt-in through svb_precision
            precision = kwargs.get("svb_precision") or "fp32"
        from .engine import normalize_precision
        self.precision = normalize_precision(precision)
        self._engine, self._fingerprint = None, None
        for p in self.parameters():
            p.requires_grad_(False)

    def _make_layer(self, block, planes, nb_layer):
        layers = []
        for d in [False] * (nb_layer - 1) + [True]:
            layers.append(block(self.inplanes, planes, downsample=d))
            self.inplanes = planes
        return nn.Sequential(*layers)

    def _state_fingerprint(self):
        return tuple((t.data_ptr(), t._version) for t in self.state_dict().values())

    def engine(self, frontend=None):
        """Packed-weight handle; rebuilt when parameters were modified (e.g. loadParameters)."""
        fp = self._state_fingerprint()
        if self._engine is None or fp != self._fingerprint:
            self._engine = RawNet2Engine(dict(self.state_dict()), self.n_samples, self.n_out, self.gru_node,
                                         self.sample_rate, precision=self.precision, aggregate=self.aggregate,
                                         att_dim=self.att_dim)
            self._fingerprint = fp
        return self._engine

    def set_precision(self, precision):
        from .engine import normalize_precision
        precision = normalize_precision(precision)
        if precision != self.precision:
            self.precision, self._engine = precision, None

    def forward(self, x):
        """[B, L] raw waveform -> [B, nOut] (squeezed like src/models/RawNet2_custom.py:226)."""
        if not x.is_cuda:
            raise RuntimeError("the libsvb encoder runs on CUDA tensors only")
        return self.engine().forward(x.float()).squeeze()


def MainModel(nOut=512, **kwargs):
    return RawNet2(RawNetBasicBlock, layers=LAYERS, nb_filters=FILTERS, code_dim=nOut, **kwargs)
