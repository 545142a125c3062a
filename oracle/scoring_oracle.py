"""CPU restatement of the reference's trial scoring and metrics (test infrastructure).

numpy fp32 like the reference; cites src/utils.py.  Pinned via tests/golden/.
"""
import numpy as np


def l2_normalize(e, eps=1e-12):
    """F.normalize(p=2, dim=1) as used at src/model.py:422-423."""
    n = np.sqrt((e.astype(np.float32) ** 2).sum(axis=1, keepdims=True))
    return e / np.maximum(n, eps)


def cosine_score(ref, com, eps=1e-5):
    """src/utils.py:163-164: mean_i |cos(ref_i, com_i)| over index-paired crops."""
    ref = np.asarray(ref, np.float32)
    com = np.asarray(com, np.float32)
    num = (ref * com).sum(-1)
    den = np.maximum(np.sqrt((ref * ref).sum(-1)), eps) * np.maximum(np.sqrt((com * com).sum(-1)), eps)
    return float(np.mean(np.abs(num / den)))


def pnorm_score(ref, com, p=2, eps=1e-6):
    """src/utils.py:167-169: np.mean(F.pairwise_distance(ref, com, p=p, eps=1e-06, keepdim=True)) -- the eps is added
    to the DIFFERENCE (torch.nn.functional.pairwise_distance), index-paired crops."""
    d = np.abs(np.asarray(ref, np.float32) - np.asarray(com, np.float32) + np.float32(eps)).astype(np.float64)
    if np.isinf(p):
        n = d.max(-1)
    else:
        n = (d ** p).sum(-1) ** (1.0 / p)
    return float(np.mean(n.astype(np.float32)))


def _zt(ref, com, cohorts, top):
    # src/utils.py:140-150
    S = np.mean(np.inner(cohorts, ref), axis=1)
    S = np.sort(S, axis=0)[::-1][:top]
    score = np.mean(np.inner(ref, com))
    return (score - np.mean(S)) / np.std(S)


def asnorm_score(ref, com, cohorts, top):
    """src/utils.py:135-160 (ZT_norm_similarity = adaptive S-norm)."""
    ref = np.asarray(ref, np.float32)
    com = np.asarray(com, np.float32)
    return float((_zt(ref, com, cohorts, top) + _zt(com, ref, cohorts, top)) / 2)


def asnorm_stats(e_mean, cohorts, top):
    """Per-utterance (mu, sigma) of the top-`top` cohort scores (exact by linearity)."""
    S = np.inner(e_mean.astype(np.float32), cohorts.astype(np.float32))     # [U, N]
    S = -np.sort(-S, axis=1)[:, :top]
    return S.mean(axis=1), S.std(axis=1)


def eer_percent(scores, labels):
    """EER branch of src/utils.py:74-105 (sklearn roc_curve, idxE = nanargmin |fnr - fpr|)."""
    from sklearn import metrics
    labels = np.nan_to_num(labels)
    scores = np.nan_to_num(scores)
    fpr, tpr, thr = metrics.roc_curve(labels, scores, pos_label=1)
    fnr = (1 - tpr) * 100
    fpr = fpr * 100
    idx = np.nanargmin(np.absolute(fnr - fpr))
    return float(np.mean([fpr[idx], fnr[idx]])), float(thr[idx])


def min_dcf(scores, labels, p_target=0.05, c_miss=1, c_fa=1):
    """src/utils.py:221-275 (ComputeErrorRates + ComputeMinDcf), vectorised."""
    scores = np.asarray(scores, dtype=np.float64)
    labels = np.asarray(labels, dtype=np.float64)
    order = np.argsort(scores, kind="stable")
    lab = labels[order]
    fnrs = np.cumsum(lab) / lab.sum()
    fprs = 1 - np.cumsum(1 - lab) / (len(lab) - lab.sum())
    c_det = c_miss * fnrs * p_target + c_fa * fprs * (1 - p_target)
    i = int(np.argmin(c_det))
    return float(c_det[i] / min(c_miss * p_target, c_fa * (1 - p_target))), float(scores[order][i])
