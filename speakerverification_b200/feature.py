"""Feature plugin: host-side mirror of the reference's
``models.FeatureExtraction.feature.melspectrogram`` factory
(src/models/FeatureExtraction/feature.py:66-94) over the fused libsvb front-end.

Same factory name, keyword arguments and defaults; the returned module keeps the
reference's ``Sequential(PreEmphasis(), MelSpectrogram)`` layout and buffer names
(``0.flipped_filter``, ``1.spectrogram.window``, ``1.mel_scale.fb``) so checkpoints
round-trip, but the arithmetic is one CUDA kernel (no cuFFT / cuBLAS / ATen calls).
"""
import torch
import torch.nn as nn

from . import _native as nv
from .engine import BatchPlan, Frontend, slaney_mel_filterbank


class PreEmphasis(nn.Module):
    """Parameter holder for src/utils.py:53-71; the filter itself runs inside the fbank kernel."""

    def __init__(self, coef: float = 0.97):
        super().__init__()
        self.coef = coef
        self.register_buffer("flipped_filter", torch.FloatTensor([-coef, 1.0]).unsqueeze(0).unsqueeze(0))

    def forward(self, x):
        raise NotImplementedError("pre-emphasis is fused into the libsvb front-end; call the enclosing module")


class _Window(nn.Module):
    def __init__(self, window):
        super().__init__()
        self.register_buffer("window", window)


class _MelScale(nn.Module):
    def __init__(self, fb):
        super().__init__()
        self.register_buffer("fb", fb)


class MelSpectrogram(nn.Module):
    """Buffer holder mirroring torchaudio.transforms.MelSpectrogram's state-dict keys."""

    def __init__(self, window, fb):
        super().__init__()
        self.spectrogram = _Window(window)
        self.mel_scale = _MelScale(fb)

    def forward(self, x):
        raise NotImplementedError("the STFT/mel projection is fused into the libsvb front-end")


class MelFrontend(nn.Sequential):
    """[B, L] waveform -> [B, n_mels, T] power-mel, T = 1 + L // hop (centred, reflect padded)."""

    def __init__(self, sr, n_fft, win_length, n_mels, hop_length, window, fmin, fmax, pre_emphasis):
        win_fn = torch.hamming_window if window == "hamming" else torch.hann_window
        w = win_fn(win_length, periodic=True, dtype=torch.float32)
        fb = slaney_mel_filterbank(n_fft // 2 + 1, float(fmin), float(fmax) if fmax else sr / 2.0, n_mels, sr)
        mods = [PreEmphasis()] if pre_emphasis else []
        super().__init__(*mods, MelSpectrogram(w, fb))
        self.cfg = dict(sr=sr, n_fft=n_fft, win_length=win_length, hop_length=hop_length, n_mels=n_mels,
                        window=window, fmin=fmin, fmax=fmax, pre_emphasis=pre_emphasis)
        self._frontend = None

    def frontend(self):
        if self._frontend is None:
            # tables always come from the config keys, never from checkpoint buffers
            self._frontend = Frontend(**self.cfg)
        return self._frontend

    def forward(self, x):
        if x.dim() != 2:
            raise AssertionError("The number of dimensions of input tensor must be 2, got %s" % (tuple(x.shape),))
        if not x.is_cuda:
            raise RuntimeError("the libsvb front-end runs on CUDA tensors only")
        fe = self.frontend()
        x = x.contiguous().float()
        B, L = x.shape
        with nv.on_device(x.device):
            plan = BatchPlan([L] * B, fe.hop, 0)
        mel = torch.empty((plan.rows, fe.n_mels), dtype=torch.float32, device=x.device)
        tsum = torch.empty((plan.tiles, fe.n_mels), dtype=torch.float32, device=x.device)
        with nv.on_device(x.device):
            nv.check(nv.lib().svb_fbank_power(fe.handle, plan.handle, nv.ptr(x), nv.ptr(mel), nv.ptr(tsum),
                                              nv.stream_ptr()))
        return mel.view(B, plan.T[0], fe.n_mels).transpose(1, 2)


def melspectrogram(lib="nnaudio", sr=8000, n_fft=512, win_length=200, n_mels=80, hop_length=80, window="hamming",
                   fmin=0.0, fmax=None, verbose=False, pre_emphasis=True, **kwargs):
    """Same signature/defaults as the reference factory.  `lib` is accepted and ignored: nnAudio and
    torchaudio compute the same transform (hamming window centred in n_fft, reflect-centre pad,
    |STFT|^2, slaney mel basis); there is exactly one implementation here."""
    if str(lib).lower() == "librosa":
        raise ValueError("lib='librosa' (a host-side numpy path in the reference) is out of scope")
    return MelFrontend(sr, n_fft, win_length, n_mels, hop_length, window, fmin, fmax, pre_emphasis)
