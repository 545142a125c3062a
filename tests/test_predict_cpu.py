"""Oracle of the enrolment bank / nearest-enrolment search against the torch-generated golden vectors
(tests/golden/predict.npz, oracle/make_golden_predict.py; inputs are the reference's own embeddings)."""
import os

import numpy as np

from oracle import predict_oracle as po

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_bank_and_distances_match_torch_golden():
    g = np.load(os.path.join(GOLD, "predict.npz"))
    emb = np.load(os.path.join(GOLD, "cfg1_c512.npz"))["emb_list"]
    files = g["enrol_files"]
    bank = po.enrol_bank(emb[files], np.arange(len(files)) // 3, 16, normalize=True)
    assert bank.shape == (2, 192, 16)
    assert np.abs(bank - g["bank"]).max() <= 1e-6
    q = emb / np.linalg.norm(emb, axis=-1, keepdims=True)
    dist = po.enrol_dist(q, g["bank"])
    assert np.abs(dist - g["dist"]).max() <= 2e-6
    td, ti = po.nearest(dist, 3)
    assert np.array_equal(ti[:, 0], g["nearest"])
    assert np.all(np.diff(td, axis=1) >= 0)
    # unit vectors: 1 - d^2 / 2 is the cosine (src/inference.py:293), for one crop exactly
    one = po.enrol_dist(q[:, :1], g["bank"][:1] / np.linalg.norm(g["bank"][:1], axis=1, keepdims=True), eps=0.0)
    cos = np.einsum("qd,ds->qs", q[:, 0], g["bank"][0] / np.linalg.norm(g["bank"][0], axis=0, keepdims=True))
    assert np.abs(po.predict_score(one) - cos).max() <= 1e-5


def test_nearest_ties_go_to_the_lower_index():
    d = np.array([[0.5, 0.25, 0.25, 0.75], [1.0, 1.0, 1.0, 1.0]], np.float32)
    td, ti = po.nearest(d, 3)
    assert ti.tolist() == [[1, 2, 0], [0, 1, 2]]
