"""Encoder plugin: host-side mirror of the reference's ``models.Raw3_ECAPA.MainModel`` (src/models/Raw3_ECAPA.py:10-59),
the hybrid the reference's default YAML selects (yaml/configuration.yaml:75-78):

    out = cat([ECAPA_TDNN(nOut=192, channels=[512, 512, 512, 512, 1536], input_norm=True)(melspectrogram(x)),
               RawNet3(nOut=nOut - 192, ...)(x)], dim=-1)                       (:44-47)

The ECAPA branch (front-end, InstanceNorm1d input norm, encoder) runs on libsvb and is pinned against the reference
(tests/golden/cfg1_c512_inorm.npz).  The RawNet3 branch imports ``asteroid_filterbanks.ParamSincFB``
(src/models/RawNet3.py:5), a dependency that is absent from this image and from the reference tree, so neither the
reference module nor an oracle of it can be run here: **parity of that branch is unpinned** and no kernel is built for
it.  The hybrid therefore takes the second branch as a caller-supplied ``torch.nn.Module`` (``svb_second_encoder``:
at a site that has asteroid_filterbanks this is the reference's own ``models.RawNet3.MainModel(...)``, run in eager
PyTorch on the same device) and does the plumbing the reference does: same raw waveform into both branches, outputs
concatenated, parameters under the reference's names (``ECAPA_TDNN.*``, ``rawnet.*``, ``compute_features.*``).
"""
import torch
import torch.nn as nn

from . import ecapa_tdnn, feature

ECAPA_OUT = 192
ECAPA_CHANNELS = [512, 512, 512, 512, 1536]


class HybridEngine:
    """forward(wav [B, L] fp32 CUDA) -> [B, 192 + second] : libsvb ECAPA branch | caller's second branch."""

    def __init__(self, ecapa_engine, second, n_out):
        self.ecapa, self.second, self.n_out = ecapa_engine, second, n_out
        self.device = ecapa_engine.device

    def forward(self, wav, l2norm=False, out=None):
        if l2norm:
            raise NotImplementedError("the hybrid returns raw embeddings (F.normalize is applied by the caller)")
        if out is None:
            out = torch.empty((wav.shape[0], self.n_out), dtype=torch.float32, device=self.device)
        self.ecapa.forward(wav, out=out[:, :ECAPA_OUT])
        with torch.no_grad():
            second = self.second(wav)
        out[:, ECAPA_OUT:] = second.reshape(wav.shape[0], -1).to(torch.float32)
        return out

    def forward_flat(self, *a, **k):
        raise NotImplementedError("the Raw3_ECAPA hybrid embeds fixed-length crops (num_eval > 0)")


class Raw_ECAPA(nn.Module):
    def __init__(self, nOut=512, svb_second_encoder=None, **kwargs):
        super().__init__()
        if nOut <= ECAPA_OUT:
            raise ValueError("Raw3_ECAPA needs nOut > %d (the ECAPA branch takes %d)" % (ECAPA_OUT, ECAPA_OUT))
        if svb_second_encoder is None:
            raise NotImplementedError(
                "Raw3_ECAPA: the RawNet3 branch needs asteroid_filterbanks (src/models/RawNet3.py:5), which is not "
                "available here; pass svb_second_encoder=<nn.Module mapping [B, L] -> [B, nOut - 192]> (e.g. the "
                "reference's models.RawNet3.MainModel) -- its parity is unpinned in this repository")
        self.n_out = nOut
        kw = dict(kwargs)
        kw.pop("channels", None)
        kw.pop("input_norm", None)
        kw["features"] = "melspectrogram"                      # src/models/Raw3_ECAPA.py:33 hard-codes the feature plugin
        self.ECAPA_TDNN = ecapa_tdnn.MainModel(nOut=ECAPA_OUT, channels=ECAPA_CHANNELS, input_norm=True, **kw)
        self.rawnet = svb_second_encoder
        fkw = {k: v for k, v in kwargs.items() if k != "features"}
        self.compute_features = feature.melspectrogram(**fkw)
        self._hybrid = None

    @property
    def precision(self):
        return self.ECAPA_TDNN.precision

    def engine(self, frontend=None):
        eng = self.ECAPA_TDNN.engine(self.compute_features.frontend())
        if self._hybrid is None or self._hybrid.ecapa is not eng:
            self._hybrid = HybridEngine(eng, self.rawnet, self.n_out)
        return self._hybrid

    def forward(self, x):
        """[B, L] raw waveform -> [B, nOut] (src/models/Raw3_ECAPA.py:36-47)."""
        if not x.is_cuda:
            raise RuntimeError("the libsvb encoder runs on CUDA tensors only")
        return self.engine().forward(x.float().contiguous())


def MainModel(nOut=512, **kwargs):
    return Raw_ECAPA(nOut=nOut, **kwargs)
