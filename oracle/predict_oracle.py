"""CPU restatement of the reference's enrolment bank and nearest-enrolment search (test infrastructure).

numpy fp32; cites src/model.py:610-645 (``prepare('embed')``) and src/inference.py:254-327 (``--predict``).
Pinned by tests/golden/predict.npz (torch ``F.pairwise_distance`` / ``F.normalize`` on reference embeddings,
``oracle/make_golden_predict.py``).

Reference quirk (SURVEY app. B, item 10 and this file): as shipped, ``--predict`` first raises ``TypeError``
(``eval_frames=`` passed to ``embed_utterance``), and its ``F.pairwise_distance(embed[n, D, 1], embeds[n, D, S])``
reduces the INNERMOST axis, i.e. the speakers, giving ``[n, D]`` instead of ``[n, S]``.  The statistic the
surrounding code consumes (``np.argmin(dist)`` indexes ``classes``, ``score = 1 - dist**2 / 2`` = cosine of unit
vectors) is the distance over the feature axis; that is what is restated here and what libsvb computes.
"""
import numpy as np


def enrol_bank(file_emb, owner, n_spk, normalize=True, eps=1e-12):
    """src/model.py:617-641: per speaker, the mean over its files of the [n, D] crop embeddings
    (``embed_utterance(..., normalize=test_normalize)``), concatenated on the last axis -> [n, D, S]."""
    e = np.asarray(file_emb, np.float32)
    if normalize:
        e = e / np.maximum(np.sqrt((e * e).sum(-1, keepdims=True)), eps)
    owner = np.asarray(owner)
    bank = np.stack([e[owner == s].mean(axis=0) for s in range(n_spk)], axis=-1)
    return bank.astype(np.float32)


def enrol_dist(query, bank, eps=1e-6):
    """src/inference.py:290-292: ``np.mean(F.pairwise_distance(embed, embeds), axis=0)`` per query utterance, with the
    norm over the feature axis.  query [Q, n, D], bank [n, D, S] -> [Q, S]."""
    q = np.asarray(query, np.float32)
    b = np.asarray(bank, np.float32)
    diff = q[:, :, :, None] - b[None] + np.float32(eps)          # [Q, n, D, S]
    return np.sqrt((diff * diff).sum(axis=2)).mean(axis=1).astype(np.float32)


def nearest(dist, k):
    """``np.argsort(dist)[:k]`` (:296, :315) with ties to the lower index; returns (distances, indices)."""
    idx = np.argsort(dist, axis=1, kind="stable")[:, :k]
    return np.take_along_axis(dist, idx, axis=1), idx.astype(np.int32)


def predict_score(d):
    """src/inference.py:293: ``1 - min(dist)**2 / 2``."""
    return 1.0 - np.asarray(d, np.float32) ** 2 / 2
