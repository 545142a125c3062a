"""GPU parity of the scoring kernels (C ABI svb_cosine_trials / svb_asnorm_*) against the reference's
own scores (tests/golden) and the numpy oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _trials():
    from oracle import synth
    spk = np.repeat(np.arange(16), 4)
    return synth.trial_list(spk, 1000)


def test_cosine_trials_match_reference(golden_c512):
    from speakerverification_b200 import utils as U
    lab, ti, tj = _trials()
    emb = torch.from_numpy(golden_c512["emb_list"]).cuda()
    ia = torch.from_numpy(ti.astype(np.int32)).cuda()
    ib = torch.from_numpy(tj.astype(np.int32)).cuda()
    got = U.cosine_trials(emb, ia, ib).cpu().numpy()
    assert np.abs(got - golden_c512["scores_cos"]).max() <= 1e-5
    # per-trial signature of the reference (utils.similarity_measure)
    s = U.similarity_measure("cosine", emb[ti[0]], emb[tj[0]])
    assert abs(s - golden_c512["scores_cos"][0]) <= 1e-5


@pytest.mark.parametrize("K", [200, 300])
def test_asnorm_matches_reference(golden_c512, K):
    from speakerverification_b200 import utils as U
    lab, ti, tj = _trials()
    emb = torch.from_numpy(golden_c512["emb_list"]).cuda()
    ia = torch.from_numpy(ti.astype(np.int32)).cuda()
    ib = torch.from_numpy(tj.astype(np.int32)).cuda()
    got = U.score_trials(emb, ia, ib, "norm", golden_c512["cohort"], top=K, normalize=True).cpu().numpy()
    ref = golden_c512["scores_asnorm%d" % K]
    assert np.abs(got - ref).max() <= 1e-4, np.abs(got - ref).max()
    # reference per-trial call with the cohort passed positionally (src/model.py:434-438)
    en = torch.nn.functional.normalize(emb, dim=2)
    s = U.similarity_measure("zt_norm", en[ti[3]], en[tj[3]], golden_c512["cohort"], top=K)
    assert abs(s - ref[3]) <= 1e-4


def test_asnorm_metrics_identical(golden_c512):
    from oracle import scoring_oracle as so
    from speakerverification_b200 import utils as U
    lab, ti, tj = _trials()
    emb = torch.from_numpy(golden_c512["emb_list"]).cuda()
    ia = torch.from_numpy(ti.astype(np.int32)).cuda()
    ib = torch.from_numpy(tj.astype(np.int32)).cuda()
    for mode, name in (("cosine", "scores_cos"), ("norm", "scores_asnorm300")):
        got = U.score_trials(emb, ia, ib, mode, golden_c512["cohort"], top=300).cpu().numpy()
        res = U.tuneThresholdfromScore(got, lab, [1, 0.1])
        fn, fp, th = U.ComputeErrorRates(got, lab)
        dcf, _ = U.ComputeMinDcf(fn, fp, th, 0.05, 1, 1)
        assert round(float(res["roc"][1]), 4) == round(float(golden_c512["eer/" + name]), 4)
        assert round(dcf, 4) == round(float(golden_c512["mindcf/" + name]), 4)
        assert round(so.eer_percent(got, lab)[0], 4) == round(float(golden_c512["eer/" + name]), 4)


@pytest.mark.parametrize("U_,N,D,K", [(37, 300, 192, 300), (64, 1000, 192, -1), (130, 20000, 192, 300),
                                      (5, 6000, 192, 300), (9, 257, 64, 1)])
def test_topk_stats_vs_numpy(U_, N, D, K):
    from speakerverification_b200 import utils as U
    rng = np.random.default_rng(N + D)
    e = rng.standard_normal((U_, D)).astype(np.float32)
    e /= np.linalg.norm(e, axis=1, keepdims=True)
    c = rng.standard_normal((N, D)).astype(np.float32)
    c /= np.linalg.norm(c, axis=1, keepdims=True)
    c[N // 2:N // 2 + 7] = c[0]                       # exact ties inside the cohort
    mu, sg = U.asnorm_stats(torch.from_numpy(e).cuda(), torch.from_numpy(c).cuda(), K)
    S = np.sort(np.inner(e.astype(np.float64), c.astype(np.float64)), axis=1)[:, ::-1]
    S = S[:, :K] if K != -1 else S[:, :-1]
    np.testing.assert_allclose(mu.cpu().numpy(), S.mean(1), atol=2e-6)
    np.testing.assert_allclose(sg.cpu().numpy(), S.std(1), atol=2e-6)


def test_scoring_sweep_properties():
    """cfg5-style synthetic sweep at reduced size: symmetry and invariance properties that hold at any size."""
    from oracle import synth
    from speakerverification_b200 import utils as U
    e, spk = synth.scoring_embeddings(4096, 100)
    c, _ = synth.scoring_embeddings(2048, 64, seed=9)
    emb = torch.from_numpy(e).cuda().unsqueeze(1)
    rng = np.random.default_rng(0)
    ia = torch.from_numpy(rng.integers(0, 4096, 20000).astype(np.int32)).cuda()
    ib = torch.from_numpy(rng.integers(0, 4096, 20000).astype(np.int32)).cuda()
    s_ab = U.score_trials(emb, ia, ib, "norm", c, top=300)
    s_ba = U.score_trials(emb, ib, ia, "norm", c, top=300)
    assert torch.equal(s_ab, s_ba)                                   # S-norm is symmetric
    perm = torch.randperm(2048)
    s_perm = U.score_trials(emb, ia, ib, "norm", c[perm.numpy()], top=300)
    assert (s_ab - s_perm).abs().max().item() < 1e-4                 # cohort order is irrelevant
    same = U.score_trials(emb, ia, ia, "cosine")
    assert (same - 1).abs().max().item() < 1e-6


@pytest.mark.parametrize("N,K", [(20000, 300), (6000, 300), (8000, 100)])
def test_topk_sampled_prefilter_adversarial_rows(N, K):
    """The sampled pre-filter select against numpy's full sort on rows built to break its estimate: one-hot embeddings
    pick cohort COLUMNS as score rows.  Ties everywhere, the top-K hidden from / concentrated in the sampled positions
    (every 4th float4), thousands of copies of the maximum, bimodal, sorted and random rows: every row must come out
    exact, through the fast path or through the exact-kernel fix-up."""
    from speakerverification_b200 import utils as U
    rng = np.random.default_rng(N + K)
    D = 64
    cols = []
    idx = np.arange(N)
    sampled = (idx // 4) % 4 == 0
    base = 0.01 * rng.standard_normal(N)
    cols.append(np.full(N, 0.25))                                            # all equal
    c1 = base.copy(); c1[np.flatnonzero(sampled)[:K]] = 0.5 + 0.1 * rng.random(K); cols.append(c1)     # top-K only in the sample
    c2 = base.copy(); c2[np.flatnonzero(~sampled)[:K]] = 0.5 + 0.1 * rng.random(K); cols.append(c2)    # top-K hidden from the sample
    c3 = base.copy(); c3[rng.permutation(N)[:5000]] = 0.75; cols.append(c3)  # 5000 copies of the maximum
    c4 = np.full(N, -0.125); c4[rng.permutation(N)[:K - 1]] = 0.5 + 0.1 * rng.random(K - 1); cols.append(c4)   # bimodal, K-th is a tie
    cols.append(rng.standard_normal(N) * 0.1)                                # plain
    cols.append(np.linspace(-0.5, 0.5, N))                                   # sorted ascending
    cols.append(np.linspace(0.5, -0.5, N))                                   # sorted descending
    c8 = base.copy(); c8[:2 * K] = 1e3 * (1 + rng.random(2 * K)); cols.append(c8)    # huge dynamic range
    c = np.zeros((N, D), dtype=np.float32)
    for j, col in enumerate(cols):
        c[:, j] = col.astype(np.float32)
    e = np.zeros((len(cols), D), dtype=np.float32)
    e[np.arange(len(cols)), np.arange(len(cols))] = 1.0
    mu, sg = U.asnorm_stats(torch.from_numpy(e).cuda(), torch.from_numpy(c).cuda(), K)
    S = np.sort(c[:, :len(cols)].T.astype(np.float64), axis=1)[:, ::-1][:, :K]
    scale = np.maximum(1.0, np.abs(S).max(1))
    assert (np.abs(mu.cpu().numpy() - S.mean(1)) <= 4e-6 * scale).all(), (mu.cpu().numpy(), S.mean(1))
    assert (np.abs(sg.cpu().numpy() - S.std(1)) <= 4e-6 * scale).all(), (sg.cpu().numpy(), S.std(1))


def test_pnorm_trials_match_reference(golden_c512):
    """'pnorm' scoring (utils.pnorm_similarity, src/utils.py:167-169; the third branch of src/model.py:433-444) for all
    trials at once against the reference function's own outputs (tests/golden/pnorm.npz), p = 2 (the call sites'
    value), 1 and 3, normalised and raw rows, plus the per-trial reference signatures."""
    import os
    from speakerverification_b200 import utils as U
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "pnorm.npz"))
    emb = torch.from_numpy(golden_c512["emb_list"]).cuda()
    ia, ib = torch.from_numpy(g["ti"]).cuda(), torch.from_numpy(g["tj"]).cuda()
    for p in (2, 1, 3):
        got = U.score_trials(emb, ia, ib, "pnorm", normalize=True, p=p).cpu().numpy()
        ref = g["scores_p%d" % p]
        assert np.abs(got - ref).max() <= 1e-5 * max(1.0, float(ref.max())), (p, np.abs(got - ref).max())
    raw = U.pnorm_trials(emb, ia, ib, p=2, normalize=False).cpu().numpy()
    assert np.abs(raw - g["scores_p2_raw"]).max() <= 1e-5 * float(g["scores_p2_raw"].max())
    en = torch.nn.functional.normalize(emb, dim=2)
    s = U.similarity_measure("pnorm", en[int(g["ti"][0])], en[int(g["tj"][0])], p=2)
    assert abs(s - float(g["scores_p2"][0])) <= 1e-5
    assert abs(U.pnorm_similarity(en[int(g["ti"][1])].cpu(), en[int(g["tj"][1])].cpu()) - float(g["scores_p2"][1])) <= 1e-5
    assert U.similarity_measure("no_such_method", en[0], en[1]) is None        # src/utils.py:126-132 falls through


@pytest.mark.parametrize("order", ["random", "sorted_by_first_score"])
def test_fused_prefilter_path_cfg5_like(order):
    """BASELINE configs[4] geometry at reduced U (4096 utterances vs a 20 000-vector cohort, top-300): the fused path
    (sample GEMM -> threshold -> filter GEMM epilogue -> candidate select; the score matrix is never stored) against
    numpy's full sort, and the share of rows that needed the exact fix-up stays tiny -- also when the caller's cohort
    is SORTED (the packed cohort is stride-permuted, so the sample stays representative)."""
    from oracle import synth
    from speakerverification_b200 import utils as U
    e, _ = synth.scoring_embeddings(4096, 200, seed=3)
    c, _ = synth.scoring_embeddings(20000, 2000, seed=103)
    if order != "random":
        c = c[np.argsort(c @ e[0])]                       # ascending in the first utterance's score: worst case for a prefix sample
    mu, sg, nfix = U.asnorm_stats(torch.from_numpy(e).cuda(), torch.from_numpy(c).cuda(), 300, return_fixups=True)
    S = -np.sort(-(e.astype(np.float64) @ c.astype(np.float64).T), axis=1)[:, :300]
    assert np.abs(mu.cpu().numpy() - S.mean(1)).max() <= 2e-6
    assert np.abs(sg.cpu().numpy() - S.std(1)).max() <= 2e-6
    print("fused AS-norm path, cohort order", order, ": fix-up rows", nfix, "of 4096")
    assert 0 <= nfix <= 40


def test_score_trials_from_host_tensors():
    """utils.score_trials with the embeddings still on the host (pinned): pieces are uploaded under the cohort GEMM of
    the previous piece; the scores equal the device-resident call bit for bit (per-utterance statistics do not depend
    on the piece boundaries)."""
    from oracle import synth
    from speakerverification_b200 import utils as U
    e, _ = synth.scoring_embeddings(70000, 300, seed=5)
    c, _ = synth.scoring_embeddings(20000, 2000, seed=105)
    rng = np.random.default_rng(1)
    ia = torch.from_numpy(rng.integers(0, 70000, 50000).astype(np.int32)).cuda()
    ib = torch.from_numpy(rng.integers(0, 70000, 50000).astype(np.int32)).cuda()
    host = torch.from_numpy(e).unsqueeze(1).pin_memory()
    want = U.score_trials(host.cuda(), ia, ib, "norm", c, top=300, normalize=False)
    got = U.score_trials(host, ia, ib, "norm", c, top=300, normalize=False)
    assert got.is_cuda and torch.equal(got, want)


def test_fused_path_stays_inside_its_buffers():
    """Canary check of the fused AS-norm path through the raw C ABI (compute-sanitizer is not available on the pool):
    workspace, mu and sigma are carved out of larger buffers whose guard bytes must survive; odd sizes (U not a multiple
    of 128, N not a multiple of 64)."""
    import ctypes
    from oracle import synth
    from speakerverification_b200 import _native as nv
    U_, N, D, K = 4133, 20011, 192, 300
    e, _ = synth.scoring_embeddings(U_, 100, seed=7)
    c, _ = synth.scoring_embeddings(N, 500, seed=107)
    emb, coh = torch.from_numpy(e).cuda(), torch.from_numpy(c).cuda()
    need = int(nv.lib().svb_asnorm_workspace_bytes(U_, N, D))
    G = 1 << 16
    big = torch.full((need + 2 * G,), 0xA5, dtype=torch.uint8, device="cuda")
    out = torch.full((2, U_ + 512), 7.25, dtype=torch.float32, device="cuda")
    ws = big[G:G + need]
    assert ws.data_ptr() % 256 == 0
    nv.check(nv.lib().svb_asnorm_stats(nv.ptr(emb), U_, nv.ptr(coh), N, D, K, ctypes.c_void_p(out[0, 256:].data_ptr()),
                                       ctypes.c_void_p(out[1, 256:].data_ptr()), nv.ptr(ws), need, nv.stream_ptr()))
    torch.cuda.synchronize()
    assert bool((big[:G] == 0xA5).all()) and bool((big[G + need:] == 0xA5).all())
    assert bool((out[:, :256] == 7.25).all()) and bool((out[:, 256 + U_:] == 7.25).all())
    S = -np.sort(-(e.astype(np.float64) @ c.astype(np.float64).T), axis=1)[:, :K]
    assert np.abs(out[0, 256:256 + U_].cpu().numpy() - S.mean(1)).max() <= 2e-6
    assert np.abs(out[1, 256:256 + U_].cpu().numpy() - S.std(1)).max() <= 2e-6


def test_fixup_rows_beyond_the_spread_capacity():
    """700 of 1000 rows are forced to the exact fix-up (5000 tied maxima: more candidates than the filter keeps), more
    than the 296 score rows fixup_scores_kernel spreads over the GPU, so both the spread rows and the rows a CTA
    recomputes itself are covered; the other 300 rows are ordinary.  All against numpy's full sort."""
    from speakerverification_b200 import utils as U
    rng = np.random.default_rng(11)
    N, D, K, R = 20000, 64, 300, 1000
    c = (0.05 * rng.standard_normal((N, D))).astype(np.float32)
    c[:, 0] = 0.01 * rng.standard_normal(N)
    c[rng.permutation(N)[:5000], 0] = 0.75
    e = (0.2 * rng.standard_normal((R, D))).astype(np.float32)
    tied = rng.permutation(R)[:700]
    e[tied] = 0.0
    e[tied, 0] = (0.5 + rng.random(700)).astype(np.float32)
    mu, sg, nfix = U.asnorm_stats(torch.from_numpy(e).cuda(), torch.from_numpy(c).cuda(), K, return_fixups=True)
    assert nfix >= 700
    S = -np.sort(-(e.astype(np.float64) @ c.astype(np.float64).T), axis=1)[:, :K]
    assert np.abs(mu.cpu().numpy() - S.mean(1)).max() <= 2e-6
    assert np.abs(sg.cpu().numpy() - S.std(1)).max() <= 2e-6
