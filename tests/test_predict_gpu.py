"""Nearest-enrolment search on the device (svb_enrol_nearest) against the oracle and the torch golden; the
reference-facing calls ``prepare('embed', <dir>)`` + ``predict`` against the golden bank built from the reference's
own embeddings."""
import os

import numpy as np
import pytest
import torch

from test_api_gpu import _handle, workdir, L  # noqa: F401  (module fixture reuse)

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("Q,n,D,S,k", [(64, 2, 192, 16, 3), (1, 1, 192, 1, 1), (13, 10, 192, 1000, 8), (37, 3, 512, 130, 5),
                                       (9, 4, 48, 257, 2)])
def test_enrol_nearest_matches_oracle(Q, n, D, S, k):
    from oracle import predict_oracle as po
    from speakerverification_b200 import utils
    rng = np.random.default_rng(Q * 1000 + S)
    q = rng.standard_normal((Q, n, D)).astype(np.float32)
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    bank = rng.standard_normal((n, D, S)).astype(np.float32)
    bank /= np.linalg.norm(bank, axis=1, keepdims=True)
    dist, td, ti = utils.enrol_nearest(torch.from_numpy(q).cuda(), torch.from_numpy(bank).cuda(), topk=k)
    ref = po.enrol_dist(q, bank)
    assert np.abs(dist.cpu().numpy() - ref).max() <= 2e-6
    # selection is exact on the device's own distances (bit-exact index work)
    rd, ri = po.nearest(dist.cpu().numpy(), k)
    assert np.array_equal(ti.cpu().numpy(), ri)
    assert np.array_equal(td.cpu().numpy(), rd)


def test_enrol_nearest_ties_and_golden():
    from speakerverification_b200 import utils
    g = np.load(os.path.join(GOLD, "predict.npz"))
    emb = np.load(os.path.join(GOLD, "cfg1_c512.npz"))["emb_list"]
    q = emb / np.linalg.norm(emb, axis=-1, keepdims=True)
    dist, td, ti = utils.enrol_nearest(torch.from_numpy(q.astype(np.float32)).cuda(), torch.from_numpy(g["bank"]).cuda(), topk=3)
    assert np.abs(dist.cpu().numpy() - g["dist"]).max() <= 2e-6
    assert np.array_equal(ti.cpu().numpy()[:, 0], g["nearest"])
    # duplicated speakers: equal distances, the lower index wins every round
    bank = torch.from_numpy(np.concatenate([g["bank"], g["bank"]], axis=2)).cuda()
    _, td2, ti2 = utils.enrol_nearest(torch.from_numpy(q.astype(np.float32)).cuda(), bank, topk=4)
    ti2 = ti2.cpu().numpy()
    assert np.array_equal(ti2[:, 0], g["nearest"]) and np.array_equal(ti2[:, 1], g["nearest"] + 16)
    assert np.array_equal(td2.cpu().numpy()[:, 0], td2.cpu().numpy()[:, 1])
    with pytest.raises(RuntimeError):
        utils.enrol_nearest(torch.zeros((2, 2, 192)).cuda(), torch.zeros((2, 192, 16)).cuda(), topk=9)   # the C-ABI serves topk <= 8


def test_prepare_embed_dir_and_predict(workdir):  # noqa: F811
    g = np.load(os.path.join(GOLD, "predict.npz"))
    net, h, args = _handle(workdir, "fp32")
    root = workdir["tmp"] / "enrol"
    for s in range(16):
        d = root / ("spk%02d" % s)
        d.mkdir(parents=True, exist_ok=True)
        for f in g["enrol_files"][3 * s:3 * s + 3]:
            dst = d / os.path.basename(workdir["names"][f])
            if not dst.exists():
                os.symlink(workdir["names"][f], dst)
    out = workdir["tmp"] / "enrol_out"
    out.mkdir(exist_ok=True)
    assert h.prepare(save_path=str(out), prepare_type="embed", num_eval=2, source=str(root)) is True
    bank = torch.load(out / "embeds.pt")
    classes = np.load(str(out / "classes.npy"), allow_pickle=True).item()
    assert tuple(bank.shape) == (2, 192, 16) and classes == {s: "spk%02d" % s for s in range(16)}
    assert np.abs(bank.numpy() - g["bank"]).max() <= 1e-4
    res = h.predict(workdir["names"], out / "embeds.pt", str(out / "classes.npy"), num_eval=2, topk=3)
    assert res["dist"].shape == (64, 16) and np.abs(res["dist"] - g["dist"]).max() <= 1e-4
    order = np.sort(g["dist"], axis=1)
    clear = (order[:, 1] - order[:, 0]) > 1e-3                     # decisions away from a tie must agree
    assert clear.sum() >= 60 and np.array_equal(res["top_idx"][clear, 0], g["nearest"][clear])
    assert np.abs(res["score"][:, 0] - (1 - order[:, 0] ** 2 / 2)).max() <= 1e-4
    assert res["names"][0][0] == classes[int(res["top_idx"][0, 0])]
