"""Host-side driver of the libsvb handles: front-end tables, packed ECAPA weights, batch plans
and workspaces.  PyTorch supplies device memory and the stream; all arithmetic is in libsvb."""
import ctypes
import math

import numpy as np
import torch

from . import _native as nv

_vp = ctypes.c_void_p


def slaney_mel_filterbank(n_freqs, f_min, f_max, n_mels, sample_rate):
    """[n_freqs, n_mels] fp32 triangles, slaney scale + slaney norm: the basis
    torchaudio.transforms.MelSpectrogram(norm='slaney', mel_scale='slaney') builds at
    src/models/FeatureExtraction/feature.py:86-93, derived from the config keys."""
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = math.log(6.4) / 27.0

    def to_mel(f):
        return min_log_mel + math.log(f / min_log_hz) / logstep if f >= min_log_hz else f / f_sp

    all_freqs = torch.linspace(0, sample_rate // 2, n_freqs)
    m = torch.linspace(to_mel(f_min), to_mel(f_max), n_mels + 2)
    f_pts = torch.where(m >= min_log_mel, min_log_hz * torch.exp(logstep * (m - min_log_mel)), f_sp * m)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts.unsqueeze(0) - all_freqs.unsqueeze(1)
    fb = torch.clamp(torch.min(-slopes[:, :-2] / f_diff[:-1], slopes[:, 2:] / f_diff[1:]), min=0.0)
    return (fb * (2.0 / (f_pts[2:n_mels + 2] - f_pts[:n_mels])).unsqueeze(0)).contiguous()


class Frontend:
    """svb fbank handle for one (sr, n_fft, win_length, hop_length, window, fmin, fmax, n_mels) config."""

    def __init__(self, sr=8000, n_fft=512, win_length=200, hop_length=80, n_mels=80, window="hamming",
                 fmin=0.0, fmax=None, pre_emphasis=True, coef=0.97, fb=None, **_):
        self.sr, self.n_fft, self.win_length, self.hop, self.n_mels = sr, n_fft, win_length, hop_length, n_mels
        win_fn = torch.hamming_window if window == "hamming" else torch.hann_window
        self.window = win_fn(win_length, periodic=True, dtype=torch.float32)
        self.fb = fb if fb is not None else slaney_mel_filterbank(
            n_fft // 2 + 1, float(fmin), float(fmax) if fmax else sr / 2.0, n_mels, sr)
        self.preemph = float(coef) if pre_emphasis else 0.0
        self._win_np = np.ascontiguousarray(self.window.numpy(), dtype=np.float32)
        self._fb_np = np.ascontiguousarray(self.fb.cpu().numpy(), dtype=np.float32)
        self._handles = {}                   # device ordinal -> SvbFbank* (its tables live on that GPU)
        _ = self.handle                      # fail early (bad geometry) on the device that is current now

    @property
    def handle(self):
        """The SvbFbank handle of the CURRENT device: engines read it inside their own device context, so one
        Frontend serves engines on several GPUs of one process."""
        dev = torch.cuda.current_device()
        h = self._handles.get(dev)
        if h is None:
            h = _vp()
            nv.check(nv.lib().svb_fbank_create(self._win_np.ctypes.data_as(_vp), self.win_length, self.hop, self.n_fft,
                                               self._fb_np.ctypes.data_as(_vp), self.n_fft // 2 + 1, self.n_mels,
                                               self.preemph, 1e-6, ctypes.byref(h)))
            self._handles[dev] = h
        return h

    def __del__(self):
        if nv is None or getattr(nv, "_lib", None) is None:          # interpreter shutdown
            return
        for h in getattr(self, "_handles", {}).values():
            nv._lib.svb_fbank_destroy(h)
        self._handles = {}


class BatchPlan:
    """Device tables for a batch of utterances laid out back to back in one wav buffer."""

    def __init__(self, nsamp, hop, halo, wav_off=None, blobs=None):
        """`blobs` = (pinned host uint8 tensor, device uint8 tensor): caller-owned tables uploaded on the current
        stream without any synchronisation (svb_batch_create_into); otherwise the library allocates them."""
        ns_np = np.ascontiguousarray(nsamp, dtype=np.int32)
        if wav_off is None:
            wav_off = np.concatenate([[0], np.cumsum(ns_np, dtype=np.int64)[:-1]])
        off_np = np.ascontiguousarray(wav_off, dtype=np.int64)
        self.B = len(ns_np)
        self.nsamp = ns_np.tolist()
        nsamp = self.nsamp
        off, ns = off_np.ctypes.data_as(_vp), ns_np.ctypes.data_as(_vp)
        h = _vp()
        if blobs is None:
            nv.check(nv.lib().svb_batch_create(off, ns, self.B, hop, halo, ctypes.byref(h)))
        else:
            need = nv.lib().svb_batch_blob_bytes(ns, self.B, hop, halo)
            if need == 0 or blobs[0].numel() < need or blobs[1].numel() < need:
                raise RuntimeError("plan blobs too small: need %d bytes" % need)
            nv.check(nv.lib().svb_batch_create_into(off, ns, self.B, hop, halo, blobs[0].data_ptr(), blobs[1].data_ptr(),
                                                    min(blobs[0].numel(), blobs[1].numel()), nv.stream_ptr(),
                                                    ctypes.byref(h)))
            self._blobs = blobs
        self.handle = h
        self.rows = nv.lib().svb_batch_rows(h)
        self.frames = nv.lib().svb_batch_frames(h)
        self.tiles = nv.lib().svb_batch_tiles(h)
        self.halo = halo
        self.T = [1 + n // hop for n in nsamp]

    def valid_row_index(self, device):
        idx, r = [], 0
        for T in self.T:
            idx.append(torch.arange(r + self.halo, r + self.halo + T, device=device))
            r += T + 2 * self.halo
        return torch.cat(idx)

    def __del__(self):
        if getattr(self, "handle", None) and nv is not None and getattr(nv, "_lib", None) is not None:   # nv is None at interpreter shutdown
            nv._lib.svb_batch_destroy(self.handle)
            self.handle = None


def normalize_precision(precision):
    """'fp32' = split-bf16 operands, 3 tensor-core passes (fp32-grade parity mode); 'f16' = the 16-bit throughput mode:
    IEEE-half activations and weights, fp32 accumulation and epilogues (same tcgen05 kind::f16 rate as bf16, 11 instead
    of 8 significant bits).  'bf16' is accepted as a name for the 16-bit mode (BASELINE.json calls the config that)."""
    if precision in ("f16", "fp16", "bf16", "half"):
        return "f16"
    if precision == "fp32":
        return "fp32"
    raise ValueError("precision must be 'fp32' or 'f16' (alias 'bf16'); got %r" % (precision,))


class EcapaEngine:
    """Packed ECAPA-TDNN weights + forward (svb_ecapa_*).  `state` maps reference state-dict names
    (without the '__S__.' prefix) to fp32 CUDA tensors."""

    def __init__(self, state, channels, n_mels, n_out, frontend, attention_channels=128, se_channels=128,
                 res2net_scale=8, kernel_sizes=(5, 3, 3, 3, 1), dilations=(1, 2, 3, 4, 1), precision="f16",
                 input_norm=False, global_context=True):
        precision = normalize_precision(precision)
        C = int(channels[0])
        if list(channels[:4]) != [C] * 4 or int(channels[4]) != 3 * C:
            raise ValueError("ECAPA channels must be [C, C, C, C, 3C]; got %s" % (list(channels),))
        self.frontend, self.C, self.n_out, self.precision = frontend, C, n_out, precision
        cfg = nv.SvbEcapaCfg()
        cfg.channels, cfg.n_mels, cfg.n_out = C, n_mels, n_out
        cfg.attention_channels, cfg.se_channels, cfg.res2net_scale = attention_channels, se_channels, res2net_scale
        cfg.kernel_sizes = (ctypes.c_int32 * 5)(*kernel_sizes)
        cfg.dilations = (ctypes.c_int32 * 5)(*dilations)
        cfg.fp32_mode = 1 if precision == "fp32" else 0
        cfg.input_norm = 1 if input_norm else 0
        cfg.no_global_context = 0 if global_context else 1
        keep, names, ptrs = [], [], []
        for k, v in state.items():
            if not torch.is_tensor(v) or not v.is_floating_point():
                continue
            t = v.detach().to(dtype=torch.float32).contiguous()
            if not t.is_cuda:
                raise RuntimeError("ECAPA weights must live on the GPU (tensor '%s' is on %s)" % (k, t.device))
            keep.append(t)
            names.append(k.encode())
            ptrs.append(t.data_ptr())
        self.device = keep[0].device
        arr_n = (ctypes.c_char_p * len(names))(*names)
        arr_p = (_vp * len(ptrs))(*ptrs)
        h = _vp()
        torch.cuda.synchronize(self.device)
        with nv.on_device(self.device):      # the library allocates / launches on the current device
            nv.check(nv.lib().svb_ecapa_create(ctypes.byref(cfg), arr_n, arr_p, len(names), ctypes.byref(h)))
        self.handle = h
        self.halo = nv.lib().svb_ecapa_halo(h)
        self._plans = {}
        self._row_plans = []
        self._ws = None

    @nv.device_method
    def plan(self, nsamp):
        key = tuple(int(n) for n in nsamp)
        p = self._plans.get(key)
        if p is None:
            if len(self._plans) > 64:
                self._plans.clear()
            p = self._plans[key] = BatchPlan(key, self.frontend.hop, self.halo)
        return p

    def workspace(self, plan):
        need = nv.lib().svb_ecapa_workspace_bytes(self.handle, plan.handle)
        if self._ws is None or self._ws.numel() < need:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._ws

    @nv.device_method
    def forward_flat(self, wav_flat, nsamp, l2norm=False, out=None):
        """wav_flat: 1-D fp32 CUDA tensor holding the utterances back to back; returns [B, n_out] fp32."""
        if not wav_flat.is_cuda or wav_flat.dtype != torch.float32:
            raise RuntimeError("wave input must be an fp32 CUDA tensor")
        plan = self.plan(nsamp)
        assert wav_flat.numel() >= sum(plan.nsamp)
        ws = self.workspace(plan)
        if out is None:
            out = torch.empty((plan.B, self.n_out), dtype=torch.float32, device=self.device)
        nv.check(nv.lib().svb_ecapa_forward(self.handle, self.frontend.handle, plan.handle, nv.ptr(wav_flat),
                                            nv.ptr(ws), ws.numel(), nv.ptr(out), out.stride(0), int(l2norm),
                                            nv.stream_ptr()))
        return out

    @nv.device_method
    def forward_rows(self, wav_flat, wav_off, nsamp, l2norm=False, out=None):
        """Rows given as (offset, length) views into one staged fp32 CUDA buffer (crops of the same utterance may
        overlap; nothing is copied).  The plan is specific to the offsets and is not cached."""
        if not wav_flat.is_cuda or wav_flat.dtype != torch.float32:
            raise RuntimeError("wave input must be an fp32 CUDA tensor")
        # per-launch plan in caller-owned blobs (ring of 4 pinned + device pairs): no cudaMalloc / cudaFree / blocking
        # copy on the critical path, so the copy stream keeps running ahead of the encoder
        ns_np = np.ascontiguousarray(nsamp, dtype=np.int32)
        need = int(nv.lib().svb_batch_blob_bytes(ns_np.ctypes.data_as(_vp), len(ns_np), self.frontend.hop, self.halo))
        if need == 0:
            nv.check(-1)
        k = self._ring_pos = (getattr(self, "_ring_pos", -1) + 1) % 4
        if len(self._row_plans) <= k:
            self._row_plans.append(None)
        slot = self._row_plans[k]
        if slot is not None:
            slot["done"].synchronize()                       # its upload (4 launches ago) has executed
            slot["plan"] = None
        if slot is None or slot["host"].numel() < need:
            size = max(need, 1 << 21)
            slot = self._row_plans[k] = {"host": torch.empty(size, dtype=torch.uint8).pin_memory(),
                                         "dev": torch.empty(size, dtype=torch.uint8, device=self.device),
                                         "done": torch.cuda.Event(), "plan": None}
        plan = slot["plan"] = BatchPlan(ns_np, self.frontend.hop, self.halo, wav_off=wav_off,
                                        blobs=(slot["host"], slot["dev"]))
        slot["done"].record()
        assert int((np.asarray(wav_off, dtype=np.int64) + ns_np).max()) <= wav_flat.numel()
        ws = self.workspace(plan)
        if out is None:
            out = torch.empty((plan.B, self.n_out), dtype=torch.float32, device=self.device)
        nv.check(nv.lib().svb_ecapa_forward(self.handle, self.frontend.handle, plan.handle, nv.ptr(wav_flat),
                                            nv.ptr(ws), ws.numel(), nv.ptr(out), out.stride(0), int(l2norm),
                                            nv.stream_ptr()))
        return out

    def forward(self, wav, l2norm=False, out=None):
        """wav: [B, L] fp32 CUDA tensor of equal-length crops."""
        wav = wav.contiguous()
        return self.forward_flat(wav.view(-1), [wav.shape[1]] * wav.shape[0], l2norm=l2norm, out=out)

    def stage(self, plan, name, cols, dtype=None, planes=1):
        """View of a named intermediate inside the workspace: [rows, planes*cols] (tests, profiling); bf16 hi|lo planes
        in fp32 mode, one plane of IEEE half in the 16-bit mode."""
        if dtype is None:
            dtype = torch.bfloat16 if planes == 2 else torch.float16
        off = nv.lib().svb_ecapa_ws_offset(self.handle, plan.handle, name.encode())
        if off < 0:
            raise KeyError(name)
        esz = torch.empty((), dtype=dtype).element_size()
        n = plan.rows * planes * cols
        return self._ws[off:off + n * esz].view(dtype).view(plan.rows, planes * cols)

    def __del__(self):
        if getattr(self, "handle", None) and nv is not None and getattr(nv, "_lib", None) is not None:   # nv is None at interpreter shutdown
            nv._lib.svb_ecapa_destroy(self.handle)
            self.handle = None


def fbank_logmel(frontend, wav, halo=0, planes=2, plane_cols=128):
    """Standalone K1: [B, L] fp32 CUDA waveform -> (packed bf16 [rows, planes*plane_cols], raw log-mel
    fp32 [rows, n_mels] before mean normalisation, BatchPlan)."""
    wav = wav.contiguous()
    B, L = wav.shape
    plan = BatchPlan([L] * B, frontend.hop, halo)
    dev = wav.device
    out = torch.empty((plan.rows, planes * plane_cols), dtype=torch.bfloat16, device=dev)
    logmel = torch.zeros((plan.rows, frontend.n_mels), dtype=torch.float32, device=dev)
    tsum = torch.empty((plan.tiles, frontend.n_mels), dtype=torch.float32, device=dev)
    mean = torch.empty((B, frontend.n_mels), dtype=torch.float32, device=dev)
    nv.check(nv.lib().svb_fbank_forward(frontend.handle, plan.handle, nv.ptr(wav), nv.ptr(out), out.stride(0),
                                        plane_cols, planes, nv.ptr(logmel), nv.ptr(tsum), nv.ptr(mean),
                                        nv.stream_ptr()))
    return out, logmel, mean, plan
