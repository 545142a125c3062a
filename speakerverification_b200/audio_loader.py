"""Host-side mirror of the reference's evaluation-mode waveform loader
(``loadWAV``, src/processing/audio_loader.py:53-152; ``normalize_audio_amp``,
src/processing/wav_conversion.py:35-41).  Training augmentation branches are out of scope."""
from pathlib import Path

import numpy as np


def _read_wav(path):
    try:
        import soundfile as sf
        audio, sr = sf.read(path, dtype="float32", always_2d=False)
        return audio, sr
    except ImportError:
        from scipy.io import wavfile
        sr, data = wavfile.read(path)
        if np.issubdtype(data.dtype, np.integer):
            data = data.astype(np.float32) / float(-np.iinfo(data.dtype).min)   # soundfile's float32 scaling
        return data.astype(np.float32), sr


def read_pcm(path):
    """Samples of a wav file as stored: int16 for PCM-16 files (uploaded as is, scaled by 1/32768 on the
    device = soundfile's float32 convention), float32 otherwise.  Feeds the GPU crop/stack stage."""
    try:
        import soundfile as sf
        if sf.info(path).subtype == "PCM_16":
            return sf.read(path, dtype="int16", always_2d=False)[0]
        return sf.read(path, dtype="float32", always_2d=False)[0]
    except ImportError:
        from scipy.io import wavfile
        _, data = wavfile.read(path)
        if data.dtype == np.int16:
            return data
        if np.issubdtype(data.dtype, np.integer):
            return data.astype(np.float32) / float(-np.iinfo(data.dtype).min)
        return data.astype(np.float32)


def crop_plan(src_len, audio_spec, num_eval):
    """Index arithmetic of loadWAV's evaluation branch (src/processing/audio_loader.py:110-150) without touching
    samples.  Returns (dst_len [U] int64: length after the wrap-pad of :126-129, starts: list of U int64 arrays with the
    crop start offsets int(linspace(0, dst_len - L, num_eval)) of :134,147-148 ([0] and whole signal when
    num_eval == 0), L)."""
    L = int(audio_spec["sentence_len"] * audio_spec["sample_rate"])
    src_len = np.asarray(src_len, dtype=np.int64)
    dst_len = np.where(src_len <= L, L + 1, src_len)
    if num_eval == 0:
        return dst_len, list(np.zeros((len(dst_len), 1), dtype=np.int64)), L
    # np.linspace with an array-valued stop evaluates the same expression per column as the scalar call of the
    # reference (arange * (stop / div), last point = stop); int() truncation = astype on the non-negative values
    starts = np.linspace(0, dst_len - L, num=num_eval, axis=1).astype(np.int64)
    return dst_len, list(starts), L


def normalize_audio_amp(signal):
    try:
        info = np.iinfo(signal.dtype)
        return signal / max(info.max, -info.min)
    except ValueError:
        return signal / max(signal.max(), -signal.min())


def loadWAV(audio_source, audio_spec=None, evalmode=True, num_eval=10, augment=False, augment_options=None,
            target_db=None, read_mode="sf", random_chunk=True, load_all=False, dtype=np.float32, **kwargs):
    """path or ndarray -> [num_eval, sentence_len*sr] evenly spaced crops (whole signal if num_eval == 0)."""
    if not audio_spec:
        audio_spec = {"sample_rate": 8000, "channels": 1, "sentence_len": 2.0, "win_len": 0.025, "hop_len": 0.01}
    if augment or not evalmode:
        raise NotImplementedError("training-mode loading / augmentation is outside the inference hot path")
    if isinstance(audio_source, (str, Path)):
        audio, _ = _read_wav(str(Path(audio_source)))
    elif isinstance(audio_source, np.ndarray):
        audio = normalize_audio_amp(audio_source)
    else:
        raise TypeError("Invalid format of audio source, available: str, ndarray")
    if load_all:
        return np.expand_dims(audio, 0)
    sr = audio_spec["sample_rate"]
    max_audio = int(audio_spec["sentence_len"] * sr)
    n = audio.shape[0]
    if n <= max_audio:
        audio = np.pad(audio, (0, max_audio - n + 1), "wrap")
        n = audio.shape[0]
    if num_eval == 0:
        return np.stack([audio], axis=0).astype(dtype)
    starts = np.linspace(0, n - max_audio, num=num_eval)
    return np.stack([audio[int(s):int(s) + max_audio] for s in starts], axis=0).astype(dtype)
