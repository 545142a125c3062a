"""Golden vectors for the 'pnorm' scoring mode (test infrastructure; build container only).

The UNMODIFIED reference function ``utils.pnorm_similarity`` (src/utils.py:167-169, reached through
``similarity_measure('pnorm', ref_feat, com_feat, p=2)`` at src/model.py:442-444) is run on the reference's own crop
embeddings of the cfg1 synthetic set (tests/golden/cfg1_c512.npz ``emb_list`` [64, 2, 192], L2-normalised per crop as
src/model.py:421-423 does for test_normalize losses) for the 1000 trials of the cfg1 list, for p = 2 (the call sites'
value), p = 1 and p = 3.

    python -m oracle.make_golden_pnorm
"""
import os

import numpy as np
import torch
import torch.nn.functional as F

from . import ref_loader, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    _, ref_utils = ref_loader.load()
    g = np.load(os.path.join(ROOT, "tests", "golden", "cfg1_c512.npz"))
    emb = torch.from_numpy(g["emb_list"])                       # [64, 2, 192]
    _, spk = synth.utterance_set(16, 4, 16)
    lab, ti, tj = synth.trial_list(spk, 1000)
    out = {}
    for p in (2, 1, 3):
        sc = []
        for i, j in zip(ti, tj):
            ref_feat = F.normalize(emb[i], p=2, dim=1)
            com_feat = F.normalize(emb[j], p=2, dim=1)
            sc.append(ref_utils.similarity_measure("pnorm", ref_feat, com_feat, p=p))
        out["scores_p%d" % p] = np.asarray(sc, dtype=np.float32)
    # un-normalised rows too (test_normalize = False losses)
    out["scores_p2_raw"] = np.asarray([ref_utils.pnorm_similarity(emb[i], emb[j], p=2) for i, j in zip(ti, tj)], np.float32)
    path = os.path.join(ROOT, "tests", "golden", "pnorm.npz")
    np.savez_compressed(path, ti=ti.astype(np.int32), tj=tj.astype(np.int32), **out)
    print("wrote", path, os.path.getsize(path), "bytes", {k: float(v.mean()) for k, v in out.items()})


if __name__ == "__main__":
    main()
