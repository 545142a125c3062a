"""Golden vectors for the enrolment bank / nearest-enrolment search (test infrastructure; build container only).

Inputs are the REFERENCE's own crop embeddings of the cfg1 synthetic set (tests/golden/cfg1_c512.npz ``emb_list``
[64, 2, 192], produced by the unmodified ``ModelHandling.embed_utterance``); the bank and the distances are computed
with the torch calls of src/model.py:617-641 and src/inference.py:290-292 (``F.normalize`` as embed_utterance applies
it, ``torch.cat`` / ``torch.mean``, ``F.pairwise_distance`` on the feature axis).

    python -m oracle.make_golden_predict
"""
import os

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    g = np.load(os.path.join(ROOT, "tests", "golden", "cfg1_c512.npz"))
    emb = torch.from_numpy(g["emb_list"])                       # [64, 2, 192]; speaker = index // 4
    per_spk, n_spk = 4, 16
    enrol_files = [s * per_spk + j for s in range(n_spk) for j in range(3)]     # 3 enrolment files per speaker
    embeds = None
    for s in range(n_spk):
        mean_embed = None
        for f in enrol_files[3 * s:3 * s + 3]:
            e = F.normalize(emb[f], p=2, dim=1)
            mean_embed = e.unsqueeze(0) if mean_embed is None else torch.cat((mean_embed, e.unsqueeze(0)), 0)
        mean_embed = torch.mean(mean_embed, dim=0)
        embeds = mean_embed.unsqueeze(-1) if embeds is None else torch.cat((embeds, mean_embed.unsqueeze(-1)), -1)
    dist = []
    for q in range(emb.shape[0]):
        e = F.normalize(emb[q], p=2, dim=1)
        d = F.pairwise_distance(e.unsqueeze(1), embeds.permute(0, 2, 1)).numpy()   # [n, S], norm over features
        dist.append(np.mean(d, axis=0))
    dist = np.stack(dist)
    path = os.path.join(ROOT, "tests", "golden", "predict.npz")
    np.savez_compressed(path, enrol_files=np.array(enrol_files), bank=embeds.numpy(), dist=dist.astype(np.float32),
                        nearest=np.argmin(dist, axis=1).astype(np.int32))
    print("wrote", path, os.path.getsize(path), "bytes; top-1 accuracy",
          float(np.mean(np.argmin(dist, axis=1) == np.arange(64) // 4)))


if __name__ == "__main__":
    main()
