"""Host-side mirror of the scoring and metric functions of the reference's ``src/utils.py``:
``similarity_measure`` / ``cosine_similarity`` / ``ZT_norm_similarity`` (:126-169) run on libsvb
kernels; the EER / minDCF metrics (:74-121, :221-275) are host fp64 like the reference.
Batched entry points (``score_trials``) replace the per-trial Python loop of src/model.py:415-448."""
import numpy as np
import torch

from . import _native as nv


def _dev(t, device=None):
    if isinstance(t, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(t))
    t = t.to(dtype=torch.float32)
    if not t.is_cuda:
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        t = t.to(device)
    return t.contiguous()


def _on_arg_device(fn):
    """Runs a scoring op with its first tensor argument's device current (libsvb launches on the current device)."""
    import functools

    @functools.wraps(fn)
    def wrapped(t, *a, **k):
        with torch.cuda.device(t.device):
            return fn(t, *a, **k)
    return wrapped


# ------------------------------------------------------------------ batched device scoring
@_on_arg_device
def cosine_trials(emb, idx_a, idx_b):
    """emb [U, n, D] CUDA fp32; int32 index tensors -> [trials] fp32 (mean_i |cos(a_i, b_i)|)."""
    U, n, D = emb.shape
    out = torch.empty(idx_a.numel(), dtype=torch.float32, device=emb.device)
    nv.check(nv.lib().svb_cosine_trials(nv.ptr(emb), n, D, nv.ptr(idx_a), nv.ptr(idx_b), idx_a.numel(), nv.ptr(out),
                                        nv.stream_ptr()))
    return out


@_on_arg_device
def pnorm_trials(emb, idx_a, idx_b, p=2, normalize=False, eps=1e-6):
    """emb [U, n, D] CUDA fp32 -> [trials] fp32: mean_i ||a_i - b_i + eps||_p (utils.pnorm_similarity, src/utils.py:167-169),
    rows L2-normalised first when `normalize` (test_normalize, src/model.py:421-423)."""
    U, n, D = emb.shape
    out = torch.empty(idx_a.numel(), dtype=torch.float32, device=emb.device)
    nv.check(nv.lib().svb_pnorm_trials(nv.ptr(emb), n, D, nv.ptr(idx_a), nv.ptr(idx_b), idx_a.numel(), float(p), float(eps),
                                       int(normalize), nv.ptr(out), nv.stream_ptr()))
    return out


@_on_arg_device
def mean_embeddings(emb, normalize):
    U, n, D = emb.shape
    out = torch.empty((U, D), dtype=torch.float32, device=emb.device)
    nv.check(nv.lib().svb_mean_embeddings(nv.ptr(emb), U, n, D, int(normalize), nv.ptr(out), nv.stream_ptr()))
    return out


@_on_arg_device
def asnorm_stats(emean, cohorts, top, return_fixups=False):
    """Per-utterance (mu, sigma) of the `top` largest cohort scores; emean [U, D], cohorts [N, D]."""
    U, D = emean.shape
    N = cohorts.shape[0]
    need = nv.lib().svb_asnorm_workspace_bytes(U, N, D)
    ws = torch.empty(need, dtype=torch.uint8, device=emean.device)
    mu = torch.empty(U, dtype=torch.float32, device=emean.device)
    sg = torch.empty(U, dtype=torch.float32, device=emean.device)
    nv.check(nv.lib().svb_asnorm_stats(nv.ptr(emean), U, nv.ptr(cohorts), N, D, int(top), nv.ptr(mu), nv.ptr(sg),
                                       nv.ptr(ws), need, nv.stream_ptr()))
    if return_fixups:      # diagnostic (synchronises): rows the fused pre-filter sent to the exact fix-up kernel
        return mu, sg, int(nv.lib().svb_asnorm_fixup_rows(nv.ptr(ws), U, N, D, int(top)))
    return mu, sg


@_on_arg_device
def asnorm_trials(emean, idx_a, idx_b, mu, sigma):
    out = torch.empty(idx_a.numel(), dtype=torch.float32, device=emean.device)
    nv.check(nv.lib().svb_asnorm_trials(nv.ptr(emean), emean.shape[1], nv.ptr(idx_a), nv.ptr(idx_b), idx_a.numel(),
                                        nv.ptr(mu), nv.ptr(sigma), nv.ptr(out), nv.stream_ptr()))
    return out


def enrol_nearest(query, bank, topk=3, eps=1e-6):
    """The --predict statistic of src/inference.py:284-294 for all query utterances at once: query [Q, n, D] crop
    embeddings, bank [n, D, S] (embeds.pt layout of src/model.py:634-641) -> (dist [Q, S] = mean over crops of the
    L2 distance with torch's pairwise_distance eps, top_dist [Q, k], top_idx [Q, k] = np.argsort(dist)[:k])."""
    query, bank = _dev(query), _dev(bank, query.device if torch.is_tensor(query) and query.is_cuda else None)
    Q, n, D = query.shape
    if tuple(bank.shape[:2]) != (n, D):
        raise ValueError("bank %s does not match query crops %s" % (tuple(bank.shape), tuple(query.shape)))
    S = bank.shape[2]
    k = min(int(topk), S)
    dist = torch.empty((Q, S), dtype=torch.float32, device=query.device)
    td = torch.empty((Q, k), dtype=torch.float32, device=query.device)
    ti = torch.empty((Q, k), dtype=torch.int32, device=query.device)
    with torch.cuda.device(query.device):
        nv.check(nv.lib().svb_enrol_nearest(nv.ptr(query), nv.ptr(bank), Q, n, D, S, float(eps), k, nv.ptr(dist),
                                            nv.ptr(td), nv.ptr(ti), nv.stream_ptr()))
    return dist, td, ti


def score_trials(emb, idx_a, idx_b, scoring_mode="cosine", cohorts=None, top=200, normalize=True, p=2):
    """All trials at once.  emb [U, n, D] (raw crop embeddings); returns a CUDA fp32 vector.
    'cosine' = utils.cosine_similarity per trial; 'norm' = utils.ZT_norm_similarity(top); 'pnorm' =
    utils.pnorm_similarity(p) (the three branches of src/model.py:433-444)."""
    if scoring_mode == "pnorm":
        return pnorm_trials(emb, idx_a, idx_b, p=p, normalize=normalize)
    if scoring_mode == "cosine":
        return cosine_trials(emb, idx_a, idx_b)      # cosine is scale invariant: normalisation is a no-op
    if scoring_mode == "norm":
        if cohorts is None:
            raise ValueError("scoring_mode='norm' needs a cohort bank")
        if torch.is_tensor(emb) and not emb.is_cuda:
            return _score_norm_from_host(emb, idx_a, idx_b, cohorts, top, normalize)
        em = mean_embeddings(emb, normalize)
        mu, sg = asnorm_stats(em, _dev(cohorts, emb.device), top)
        return asnorm_trials(em, idx_a, idx_b, mu, sg)
    raise ValueError("scoring_mode %r: the reference knows 'cosine', 'norm' and 'pnorm' (src/model.py:433-444)"
                     % (scoring_mode,))


def _score_norm_from_host(emb, idx_a, idx_b, cohorts, top, normalize, rows_per_piece=32768, device=None):
    """AS-norm with the crop embeddings still on the HOST (pinned for full overlap): the cohort statistics of an
    utterance depend on that utterance alone, so the [U, n, D] tensor is uploaded in pieces on a side stream while the
    cohort GEMM of the previous piece runs; the trial kernel starts when the last piece's statistics are in."""
    if device is None:
        device = idx_a.device if torch.is_tensor(idx_a) and idx_a.is_cuda else torch.device("cuda", torch.cuda.current_device())
    with torch.cuda.device(device):
        U, n, D = emb.shape
        coh = _dev(cohorts, device)
        ia = idx_a.to(device, non_blocking=True) if torch.is_tensor(idx_a) else torch.as_tensor(idx_a, device=device)
        ib = idx_b.to(device, non_blocking=True) if torch.is_tensor(idx_b) else torch.as_tensor(idx_b, device=device)
        main = torch.cuda.current_stream(device)
        copy = torch.cuda.Stream(device=device)
        d_emb = torch.empty((U, n, D), dtype=torch.float32, device=device)
        em = torch.empty((U, D), dtype=torch.float32, device=device)
        mu = torch.empty(U, dtype=torch.float32, device=device)
        sg = torch.empty(U, dtype=torch.float32, device=device)
        copy.wait_stream(main)
        events = []
        for r0 in range(0, U, rows_per_piece):
            r1 = min(U, r0 + rows_per_piece)
            with torch.cuda.stream(copy):
                d_emb[r0:r1].copy_(emb[r0:r1], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy)
            events.append((r0, r1, ev))
        for r0, r1, ev in events:
            main.wait_event(ev)
            em[r0:r1] = mean_embeddings(d_emb[r0:r1], normalize)
            m, s_ = asnorm_stats(em[r0:r1], coh, top)
            mu[r0:r1], sg[r0:r1] = m, s_
        d_emb.record_stream(main)
        return asnorm_trials(em, ia.to(torch.int32), ib.to(torch.int32), mu, sg)


# ------------------------------------------------------------------ per-trial reference signatures
def cosine_similarity(ref, com, **kwargs):
    ref, com = _dev(ref), _dev(com, None)
    emb = torch.stack([ref.view(-1, ref.shape[-1]), com.view(-1, com.shape[-1])])
    i0 = torch.zeros(1, dtype=torch.int32, device=emb.device)
    return float(cosine_trials(emb, i0, i0 + 1).item())


def ZT_norm_similarity(ref, com, cohorts, top=-1):
    ref, com = _dev(ref), _dev(com)
    emb = torch.stack([ref.view(-1, ref.shape[-1]), com.view(-1, com.shape[-1])])
    i0 = torch.zeros(1, dtype=torch.int32, device=emb.device)
    em = mean_embeddings(emb, False)
    mu, sg = asnorm_stats(em, _dev(cohorts, emb.device), top)
    return float(asnorm_trials(em, i0, i0 + 1, mu, sg).item())


def pnorm_similarity(ref, com, p=2, **kwargs):
    """src/utils.py:167-169: np.mean(F.pairwise_distance(ref, com, p=p, eps=1e-06, keepdim=True))."""
    ref, com = _dev(ref), _dev(com)
    emb = torch.stack([ref.view(-1, ref.shape[-1]), com.view(-1, com.shape[-1])])
    i0 = torch.zeros(1, dtype=torch.int32, device=emb.device)
    return float(pnorm_trials(emb, i0, i0 + 1, p=p, normalize=False).item())


def similarity_measure(method="cosine", ref=None, com=None, *args, **kwargs):
    """Same dispatch as src/utils.py:126-132.  Unlike the reference, extra positionals are accepted,
    so the call site src/model.py:434-438 (zt_norm with the cohort passed positionally) works."""
    if method == "cosine":
        return cosine_similarity(ref, com, **kwargs)
    if method == "zt_norm":
        return ZT_norm_similarity(ref, com, *args, **kwargs)
    if method == "pnorm":
        return pnorm_similarity(ref, com, **kwargs)
    return None      # src/utils.py:126-132 falls through (returns None) for an unknown method


# ------------------------------------------------------------------ metrics (host, fp64)
def tuneThresholdfromScore(scores, labels, target_fa, target_fr=None):
    """src/utils.py:74-121 (sklearn roc_curve / precision_recall_curve); same result dict."""
    from sklearn import metrics
    labels = np.nan_to_num(labels)
    scores = np.nan_to_num(scores)
    fpr, tpr, thresholds = metrics.roc_curve(labels, scores, pos_label=1)
    gmean = np.sqrt(tpr * (1 - fpr))
    idxG = int(np.argmax(gmean))
    fnr = (1 - tpr) * 100
    fpr = fpr * 100
    tuned = []
    if target_fr:
        for tfr in target_fr:
            idx = np.nanargmin(np.absolute(tfr - fnr))
            tuned.append([thresholds[idx], fpr[idx], fnr[idx]])
    for tfa in target_fa:
        idx = np.nanargmin(np.absolute(tfa - fpr))
        tuned.append([thresholds[idx], fpr[idx], fnr[idx]])
    idxE = np.nanargmin(np.absolute(fnr - fpr))
    eer = np.mean([fpr[idxE], fnr[idxE]])
    precision, recall, thr_pr = metrics.precision_recall_curve(labels, scores, pos_label=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        fscore = (2 * precision * recall) / (precision + recall)
    ix = int(np.nanargmax(fscore))
    return {"gmean": [idxG, gmean[idxG], thresholds[idxG]],
            "roc": [tuned, eer, metrics.auc(fpr, tpr), thresholds[idxE]],
            "prec_recall": [precision, recall, fscore[ix], thr_pr[min(ix, len(thr_pr) - 1)]]}


def ComputeErrorRates(scores, labels):
    """src/utils.py:221-256 vectorised: (fnrs, fprs, thresholds) over ascending scores."""
    scores = np.asarray(scores, dtype=np.float64)
    labels = np.asarray(labels, dtype=np.float64)
    order = np.argsort(scores, kind="stable")
    lab = labels[order]
    fnrs = np.cumsum(lab) / lab.sum()
    fprs = 1 - np.cumsum(1 - lab) / (len(lab) - lab.sum())
    return list(fnrs), list(fprs), list(scores[order])


def ComputeMinDcf(fnrs, fprs, thresholds, p_target, c_miss, c_fa):
    """src/utils.py:262-275."""
    c_det = c_miss * np.asarray(fnrs) * p_target + c_fa * np.asarray(fprs) * (1 - p_target)
    i = int(np.argmin(c_det))
    return float(c_det[i] / min(c_miss * p_target, c_fa * (1 - p_target))), thresholds[i]
