"""Encoder plugin: host-side mirror of the reference's ``models.ECAPA_TDNN.MainModel``
(src/models/ECAPA_TDNN.py:339-507).

The module tree exists to expose parameters and buffers under the reference's names and
shapes (``blocks.1.tdnn1.conv.conv.weight`` ...; 231 state-dict entries) so checkpoints load
and save unchanged.  None of these ``nn.Conv1d`` / ``nn.BatchNorm1d`` holders is ever called:
the forward pass is libsvb (tcgen05 implicit GEMMs + fused pooling kernels).
"""
import torch
import torch.nn as nn

from . import _native as nv
from .engine import EcapaEngine, Frontend, normalize_precision


class _Conv(nn.Module):          # key: <p>.conv.{weight,bias}          (layers/cnn.py Conv1d wrapper)
    def __init__(self, cin, cout, k):
        super().__init__()
        self.conv = nn.Conv1d(cin, cout, k)


class _Norm(nn.Module):          # key: <p>.norm.{weight,bias,running_*} (layers/normalization.py wrapper)
    def __init__(self, c):
        super().__init__()
        self.norm = nn.BatchNorm1d(c)


class _TDNN(nn.Module):          # TDNNBlock: conv -> activation -> norm
    def __init__(self, cin, cout, k):
        super().__init__()
        self.conv = _Conv(cin, cout, k)
        self.norm = _Norm(cout)


class _Res2Net(nn.Module):
    def __init__(self, c, scale, k):
        super().__init__()
        self.blocks = nn.ModuleList([_TDNN(c // scale, c // scale, k) for _ in range(scale - 1)])


class _SE(nn.Module):
    def __init__(self, c, se):
        super().__init__()
        self.conv1 = _Conv(c, se, 1)
        self.conv2 = _Conv(se, c, 1)


class _SERes2Net(nn.Module):
    def __init__(self, c, scale, se, k):
        super().__init__()
        self.tdnn1 = _TDNN(c, c, 1)
        self.res2net_block = _Res2Net(c, scale, k)
        self.tdnn2 = _TDNN(c, c, 1)
        self.se_block = _SE(c, se)


class _ASP(nn.Module):
    def __init__(self, c, att, global_context=True):
        super().__init__()
        self.tdnn = _TDNN((3 if global_context else 1) * c, att, 1)      # ECAPA_TDNN.py:204-207
        self.conv = _Conv(att, c, 1)


class ECAPA_TDNN(nn.Module):
    def __init__(self, input_size=80, lin_neurons=192, channels=(1024, 1024, 1024, 1024, 3072),
                 kernel_sizes=(5, 3, 3, 3, 1), dilations=(1, 2, 3, 4, 1), attention_channels=128, res2net_scale=8,
                 se_channels=128, input_norm=False, global_context=True, precision=None, activation=None, **kwargs):
        super().__init__()
        # the reference's `activation` (default torch.nn.GELU, src/models/ECAPA_TDNN.py:377) is fused into the GEMM
        # epilogues as exact-erf GELU: any other choice has no kernel and must not run silently as GELU
        if activation is not None and activation is not nn.GELU and not isinstance(activation, nn.GELU):
            raise NotImplementedError("ECAPA_TDNN activation %r: the libsvb encoder implements the default torch.nn.GELU"
                                      % (activation,))
        assert len(channels) == len(kernel_sizes) == len(dilations)
        assert input_size == kwargs["n_mels"], "inappropriate input size, should equal feature_dim"
        if str(kwargs.get("features", "melspectrogram")).strip() != "melspectrogram":
            raise NotImplementedError("the libsvb encoder consumes mel-spectrogram features only")
        self.channels = list(channels)
        self.kernel_sizes, self.dilations = list(kernel_sizes), list(dilations)
        self.n_mels, self.n_out = input_size, lin_neurons
        self.att, self.se, self.scale = attention_channels, se_channels, res2net_scale
        self.kwargs = kwargs
        self.input_norm, self.global_context = bool(input_norm), bool(global_context)
        if self.input_norm:      # src/models/ECAPA_TDNN.py:406-408: keys instance_norm.{weight,bias}
            self.instance_norm = nn.InstanceNorm1d(input_size, affine=True, eps=1e-05, momentum=0.1,
                                                   track_running_stats=False)
        if precision is None:
            # the reference's `mixedprec` only autocasts the TRAINING step (src/model.py:256); evaluateFromList /
            # testFromList always run fp32, so fp32-grade is the default and bf16 is opt-in through svb_precision
            precision = kwargs.get("svb_precision") or "fp32"
        self.precision = normalize_precision(precision)
        C = channels
        self.blocks = nn.ModuleList([_TDNN(input_size, C[0], kernel_sizes[0])])
        for i in range(1, len(C) - 1):
            assert C[i] == C[i - 1], "ECAPA_TDNN uses equal channel counts in the SE-Res2Net blocks"
            self.blocks.append(_SERes2Net(C[i], res2net_scale, se_channels, kernel_sizes[i]))
        self.mfa = _TDNN(C[-1], C[-1], kernel_sizes[-1])
        self.asp = _ASP(C[-1], attention_channels, global_context)
        self.asp_bn = _Norm(2 * C[-1])
        self.fc = _Conv(2 * C[-1], lin_neurons, 1)
        self._engine, self._fingerprint, self._frontend = None, None, None
        for p in self.parameters():
            p.requires_grad_(False)

    # ------------------------------------------------------------------ libsvb handle
    def _state_fingerprint(self):
        return tuple((t.data_ptr(), t._version) for t in self.state_dict().values())

    def frontend(self):
        if self._frontend is None:
            kw = self.kwargs
            self._frontend = Frontend(sr=kw.get("sr", 8000), n_fft=kw.get("n_fft", 512),
                                      win_length=kw.get("win_length", 200), hop_length=kw.get("hop_length", 80),
                                      n_mels=self.n_mels, window=kw.get("window", "hamming"),
                                      fmin=kw.get("fmin", 0.0), fmax=kw.get("fmax", None),
                                      pre_emphasis=kw.get("pre_emphasis", True))
        return self._frontend

    def engine(self, frontend=None):
        """Packed-weight handle; rebuilt when parameters were modified (e.g. loadParameters)."""
        fp = self._state_fingerprint()
        if self._engine is None or fp != self._fingerprint:
            if frontend is not None:
                self._frontend = frontend
            state = {k: v for k, v in self.state_dict().items()}
            self._engine = EcapaEngine(state, self.channels, self.n_mels, self.n_out, self.frontend(),
                                       attention_channels=self.att, se_channels=self.se, res2net_scale=self.scale,
                                       kernel_sizes=self.kernel_sizes, dilations=self.dilations,
                                       precision=self.precision, input_norm=self.input_norm,
                                       global_context=self.global_context)
            self._fingerprint = fp
        return self._engine

    def set_precision(self, precision):
        precision = normalize_precision(precision)
        if precision != self.precision:
            self.precision, self._engine = precision, None

    # ------------------------------------------------------------------ plugin forward
    def forward(self, x, lengths=None):
        """[B, n_mels, T] power-mel (what the feature plugin returns) -> [B, nOut] (squeezed like
        src/models/ECAPA_TDNN.py:500)."""
        if lengths is not None:
            raise NotImplementedError("relative-length masking is never used on the inference path")
        if not x.is_cuda:
            raise RuntimeError("the libsvb encoder runs on CUDA tensors only")
        eng = self.engine()
        B, M, T = x.shape
        assert M == self.n_mels
        hop = eng.frontend.hop
        plan = eng.plan([(T - 1) * hop + hop // 2] * B)
        rows = torch.zeros((plan.rows, M), dtype=torch.float32, device=x.device)
        rows[plan.valid_row_index(x.device)] = x.float().transpose(1, 2).reshape(-1, M)
        ws = eng.workspace(plan)
        out = torch.empty((B, self.n_out), dtype=torch.float32, device=x.device)
        with nv.on_device(x.device):
            nv.check(nv.lib().svb_ecapa_forward_mel(eng.handle, eng.frontend.handle, plan.handle, nv.ptr(rows), nv.ptr(ws),
                                                    ws.numel(), nv.ptr(out), out.stride(0), 0, nv.stream_ptr()))
        return out.squeeze()


def MainModel(nOut=512, **kwargs):
    return ECAPA_TDNN(lin_neurons=nOut, **kwargs)
