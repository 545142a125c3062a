"""Golden vectors of the reference's evaluation-mode loader (test infrastructure; build container only).

Runs the UNMODIFIED ``loadWAV`` (src/processing/audio_loader.py:53-152) through ``oracle/ref_loader`` on seeded
int16-valued signals served as in-memory "files" (soundfile.read stub returns the float32 = int16 / 32768 array, the
same values sf.read yields for a PCM-16 file) and commits the crops to tests/golden/loader.npz.

    python -m oracle.make_golden_loader
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

SPEC = {"sample_rate": 1000, "channels": 1, "sentence_len": 0.5, "win_len": 0.025, "hop_len": 0.01}   # L = 500
LENGTHS = [7, 249, 499, 500, 501, 502, 777, 1000, 1501, 4096]
NUM_EVAL = [0, 1, 2, 5, 10]


def signals():
    rng = np.random.default_rng(77)
    return [rng.integers(-32768, 32768, n).astype(np.int16) for n in LENGTHS]


def main():
    ref_loader.load()
    import processing.audio_loader as ref_al
    out = {"lengths": np.array(LENGTHS), "num_eval": np.array(NUM_EVAL)}
    for i, pcm in enumerate(signals()):
        ref_loader.WAVS["/mem/%d.wav" % i] = pcm.astype(np.float32) / 32768.0
        for n in NUM_EVAL:
            crops = ref_al.loadWAV("/mem/%d.wav" % i, SPEC, evalmode=True, augment=False, augment_options=[],
                                   num_eval=n, random_chunk=False)
            out["crops_%d_%d" % (i, n)] = crops.astype(np.float32)
    path = os.path.join(ROOT, "tests", "golden", "loader.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
