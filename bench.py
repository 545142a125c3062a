#!/usr/bin/env python
"""Throughput bench for the B200 speaker-embedding hot path (contract: task statement / DESIGN.md).

  python bench.py [--gpus N --steps K --warmup W] [--impl reference]

Headline workload (BASELINE.json configs[1]): ECAPA-TDNN C=1024, 80-mel fbank, 10 000 synthetic 3 s 16 kHz
utterances per GPU, 16-bit tensor-core operands (IEEE half, fp32 accumulation: BASELINE's "bf16" config at the same
tcgen05 rate, see profiles/r02_precision.md).  One step = one pass of the hot path (waveform -> embedding) over all
utterances.  `value` is device-timed with the waveforms resident in HBM; `e2e` goes through the public API
(ModelHandling.embed_batches) from pinned HOST buffers with the H2D copies and the D2H read of the embeddings inside
the timed region.  The default run (--workload all) adds the other BASELINE configs as sub-objects of the same line:
`eval` (configs[2], ragged 2-20 s + AS-norm), `rawnet2` (configs[3]), `scoring` (configs[4]), at every N; at N = 1 also
`gpu_library_baseline` (the reference's forward as PyTorch eager runs it on this GPU: cuFFT / cuBLAS / cuDNN, fp32 and
autocast-bf16) and `cpu_baseline` (the oracle port on the host cores); at N > 1 per-rank step times, one strong-scaling
line and the distributed API check (evaluateFromList(distributed=True) against the single-GPU call).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CH = [1024, 1024, 1024, 1024, 3072]
N_OUT = 192
L = 48000
FLOP_PER_UTT = 11.29e9          # 37.50 MFLOP/frame x 301 frames (C=1024, 3 s crop)
MFA_MAC_PER_FRAME = 9437184     # SURVEY.md 8a GEMM table, mfa row (3C x 3C)
# dram__bytes_read.sum + dram__bytes_write.sum of one MFA launch at 592 rows (ncu --set full, profiles/r01_ncu_final.md:
# 1.215 GB + 1.132 GB; fp16 operands have the same footprint as bf16) against 2.267 GB algorithmic
MFA_DRAM_BYTES = 2.487e9


def model_args(device, precision):
    return dict(device=device, gpu=int(os.environ.get("LOCAL_RANK", 0)), seed=0, save_folder="/tmp/svb_bench",
                model={"name": "ECAPA_TDNN", "nOut": N_OUT}, channels=CH,
                criterion={"name": "AAmSoftmaxAP", "margin": 0.2, "scale": 30}, nClasses=10,
                classifier={"input_size": N_OUT, "out_neurons": 10},
                optimizer={"name": "adam", "weight_decay": 2e-5, "lr_decay": 0.95}, lr=1e-3,
                callbacks={"name": "steplr"}, step_size=10, features="melspectrogram", lib="torchaudio", n_mels=80,
                n_mfcc=80, sr=16000, n_fft=512, win_length=400, hop_length=160,
                audio_spec={"sample_rate": 16000, "channels": 1, "sentence_len": 3.0, "win_len": 0.025, "hop_len": 0.01},
                dataloader_options={"nPerSpeaker": 2, "num_workers": 0, "batch_size": 4, "shuffle": False},
                augment=False, augment_options={"augment_chain": []}, mixedprec=False, svb_precision=precision)


def rawnet_args(device, precision):
    a = model_args(device, precision)
    a.update(model={"name": "RawNet2_custom", "nOut": 512}, features="raw",
             audio_spec={"sample_rate": 16000, "channels": 1, "sentence_len": 4.0, "win_len": 0.025, "hop_len": 0.01},
             classifier={"input_size": 512, "out_neurons": 10})
    return a


def cpu_rawnet_rate(min_seconds, batch=8):
    """Reference CPU path of RawNet2_custom (oracle port, fp32, all host threads): embeddings/s."""
    from oracle import rawnet_oracle as ro, synth
    torch.set_num_threads(os.cpu_count())
    sd = synth.rawnet2_state_dict()
    rng = np.random.default_rng(0)
    waves = (0.1 * rng.standard_normal((batch, 64000))).astype(np.float32)
    with torch.no_grad():
        ro.rawnet2_forward(waves[:2], sd)
        t0 = time.perf_counter()
        done = 0
        while time.perf_counter() - t0 < min_seconds or done == 0:
            ro.rawnet2_forward(waves, sd)
            done += batch
    dt = time.perf_counter() - t0
    return done / dt, done, dt


_REAL_STDOUT = None


def quiet_stdout():
    """Everything the libraries print (model banner, NCCL version line) goes to stderr; stdout carries the ONE JSON line."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(line + "\n")
    out.flush()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], d.get("bf16_tflops_sustained", d["bf16_tflops"]), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def synth_waves(n, device, seed, length=None):
    """n distinct utterances of `length` samples (default 3 s): speaker-like harmonic signals (synth.py) mixed, shifted
    and scaled on the GPU."""
    from speakerverification_b200 import synth
    L = length or globals()["L"]
    pool, _ = synth.utterance_set(8, 4, L)
    pool = torch.from_numpy(pool).to(device)
    g = torch.Generator(device=device).manual_seed(seed)
    out = torch.empty((n, L), dtype=torch.float32, device=device)
    for i0 in range(0, n, 1000):
        m = min(1000, n - i0)
        a = torch.randint(0, pool.shape[0], (m,), generator=g, device=device)
        b = torch.randint(0, pool.shape[0], (m,), generator=g, device=device)
        w = torch.rand((m, 1), generator=g, device=device)
        gain = 0.2 + 0.8 * torch.rand((m, 1), generator=g, device=device)
        x = w * pool[a] + (1 - w) * torch.roll(pool[b], 7919, dims=1)
        out[i0:i0 + m] = gain * x + 0.01 * torch.randn((m, L), generator=g, device=device)
    return out


def cpu_reference_rate(n_utts, batch=16, min_seconds=0.0):
    """Reference CPU path (oracle port of SpeakerEncoder.forward, fp32, all host threads): embeddings/s.
    Runs at least n_utts utterances and at least min_seconds of CPU work."""
    from oracle import ecapa_oracle as eo, synth
    torch.set_num_threads(os.cpu_count())
    sd = synth.ecapa_state_dict(CH, N_OUT)
    sd = {k: torch.as_tensor(v) for k, v in sd.items()}
    rng = np.random.default_rng(0)
    waves = (0.1 * rng.standard_normal((batch, L))).astype(np.float32)
    eo.embed(waves, sd)                                  # warm-up
    t0 = time.perf_counter()
    done = 0
    while done < n_utts or time.perf_counter() - t0 < min_seconds:
        eo.embed(waves, sd)
        done += batch
    dt = time.perf_counter() - t0
    return done / dt, done, dt


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port; the Python reference tree
    itself cannot travel to the GPU box), timed on the host cores."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    if args.workload == "scoring":
        U, N, NT = min(args.score_utts, 20000), args.score_cohort, 4096
        e, c, ia, ib = scoring_inputs(U, N, NT)
        cpu_scoring_rate(e, c, ia, ib, 300, 0.5)
        t0 = time.perf_counter()
        n = 0
        for _ in range(args.steps):
            _, done, _ = cpu_scoring_rate(e, c, ia, ib, 300, 3.0)
            n += done
        dt = time.perf_counter() - t0
        val = n / dt
        emit(json.dumps({
            "impl": "reference", "metric": "AS-norm trials/sec", "value": val, "unit": "trials/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(args.steps, 1),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "scoring sweep (configs[4]): per-trial AS-norm vs %d-cohort, top-300, D=%d" % (N, N_OUT),
                       "sample": "3 s of trials per step"},
            "cpu_baseline": {"value": val, "unit": "trials/s", "cores": os.cpu_count(), "kind": "port",
                             "sample": "%d trials, per-trial ZT_norm_similarity loop (numpy)" % n},
            "e2e": {"value": val, "unit": "trials/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    if args.workload == "eval":
        from speakerverification_b200 import synth
        U, NT = args.eval_utts, args.eval_trials
        lengths = synth.ragged_lengths(U, seed=2)
        signals = eval_signals(lengths[:64], seed=50)
        cohort, _ = synth.scoring_embeddings(6000, 2000, seed=103)
        t0 = time.perf_counter()
        acc = []
        for _ in range(args.steps):
            acc.append(cpu_eval_rate(signals, cohort, 300, 6.0))
        dt = time.perf_counter() - t0
        t_emb = float(np.mean([a[0] / (a[4] / a[2]) for a in acc])) * float(lengths.sum()) / 16000.0 / U
        t_trial = float(np.mean([a[1] for a in acc]))
        val = 1.0 / (t_emb + t_trial * NT / U)
        emit(json.dumps({
            "impl": "reference", "metric": "full-length eval utterances/sec (2-20 s utt, fbank + ECAPA + AS-norm)",
            "value": val, "unit": "utterances/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dt / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "full-length eval (configs[2]): ragged U(2,20) s utterances, ECAPA-TDNN C=1024, AS-norm "
                                   "vs 6000-cohort top-300, %d trials per %d utterances" % (NT, U),
                       "sample": "about 8 s of CPU work per step, composed per utterance"},
            "cpu_baseline": {"value": val, "unit": "utterances/s", "cores": os.cpu_count(), "kind": "port",
                             "sample": "%d whole utterances + %d trials per step, oracle port" % (acc[0][2], acc[0][3])},
            "e2e": {"value": val, "unit": "utterances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    if args.workload == "rawnet2":
        t0 = time.perf_counter()
        n = 0
        for _ in range(args.steps):
            _, done, _ = cpu_rawnet_rate(3.0)
            n += done
        dt = time.perf_counter() - t0
        val = n / dt
        emit(json.dumps({
            "impl": "reference", "metric": "RawNet2 embeddings/sec (4 s utt)", "value": val, "unit": "embeddings/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(args.steps, 1),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "RawNet2_custom (sinc + GRU) nOut=512, 4 s 16 kHz utterances (configs[3])",
                       "sample": "3 s of batches of 8 per step"},
            "cpu_baseline": {"value": val, "unit": "embeddings/s", "cores": os.cpu_count(), "kind": "port",
                             "sample": "%d x 4 s utterances, batch 8, torch fp32 oracle port" % n},
            "e2e": {"value": val, "unit": "embeddings/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    per_step = 128
    for _ in range(min(args.warmup, 1)):
        cpu_reference_rate(16)
    t0 = time.perf_counter()
    n = 0
    for _ in range(args.steps):
        _, done, _ = cpu_reference_rate(per_step)
        n += done
    dt = time.perf_counter() - t0
    val = n / dt
    emit(json.dumps({
        "impl": "reference", "metric": "ECAPA embeddings/sec (3 s utt)", "value": val, "unit": "embeddings/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "ECAPA-TDNN C=1024, 80-mel fbank, 3 s 16 kHz utterances (configs[1])",
                   "sample": "%d utterances per step, batch 16" % per_step},
        "cpu_baseline": {"value": val, "unit": "embeddings/s", "cores": os.cpu_count(), "kind": "port",
                         "sample": "%d x 3 s utterances per step, batch 16, torch fp32 oracle port" % per_step},
        "e2e": {"value": val, "unit": "embeddings/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------ scoring sweep (BASELINE.json configs[4])
def scoring_inputs(U, N, n_trials, seed=3):
    """cfg5: unit embeddings around 2 000 speaker centroids, disjoint cohort, seeded trial pairs (10 % target)."""
    from speakerverification_b200 import synth
    e, spk = synth.scoring_embeddings(U, 2000, seed=seed)
    c, _ = synth.scoring_embeddings(N, 2000, seed=seed + 100)
    rng = np.random.default_rng(seed + 1)
    ia = rng.integers(0, U, n_trials)
    ib = rng.integers(0, U, n_trials)
    order = np.argsort(spk, kind="stable")
    pos = rng.integers(0, U - 1, n_trials // 10)
    nxt = np.where(spk[order[pos]] == spk[order[pos + 1]], pos + 1, np.maximum(pos - 1, 0))
    ia[:len(pos)], ib[:len(pos)] = order[pos], order[nxt]
    return e, c, ia.astype(np.int32), ib.astype(np.int32)


def cpu_scoring_rate(e, c, ia, ib, top, min_seconds):
    """The reference's own per-trial formulation (utils.ZT_norm_similarity in the loop of model.py:415-448;
    oracle port, numpy on all host threads): two [N, D] x [D, n] products and two full sorts per trial."""
    from oracle import scoring_oracle as so
    t0 = time.perf_counter()
    n = 0
    while time.perf_counter() - t0 < min_seconds or n < 8:
        so.asnorm_score(e[ia[n % len(ia)]][None], e[ib[n % len(ib)]][None], c, top)
        n += 1
    dt = time.perf_counter() - t0
    return n / dt, n, dt


def run_scoring(args, device, rank, world, dist, steps, full_line=True):
    """AS-norm trials/sec: per step, cohort statistics (top-K mean/std against N cohort vectors) of this rank's
    U unique utterances and the S-norm scores of its n_trials trials (trial- and utterance-sharded, weak)."""
    from speakerverification_b200 import _native as nv
    from speakerverification_b200 import utils as SU
    U, N, K, NT, D = args.score_utts, args.score_cohort, 300, args.score_trials, N_OUT
    e, c, ia, ib = scoring_inputs(U, N, NT, seed=3 + rank)
    emb = torch.from_numpy(e).to(device)
    coh = torch.from_numpy(c).to(device)
    tia, tib = torch.from_numpy(ia).to(device), torch.from_numpy(ib).to(device)

    def step():
        mu, sg = SU.asnorm_stats(emb, coh, K)
        return SU.asnorm_trials(emb, tia, tib, mu, sg)

    for _ in range(3):
        out = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    nv.lib().svb_reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    launches = int(nv.lib().svb_launch_count())
    ms, per_rank = rank_stats(e0.elapsed_time(e1), {"device": device, "world": world, "dist": dist})
    value = world * NT * steps / (ms * 1e-3)
    # end to end: host embeddings / cohort / trial indices in, host scores out, through utils.score_trials
    he = torch.from_numpy(e).unsqueeze(1).pin_memory()
    hia, hib = torch.from_numpy(ia).pin_memory(), torch.from_numpy(ib).pin_memory()
    hc = torch.from_numpy(c).pin_memory()

    def e2e_step():
        # host tensors straight into the public call: it uploads the embeddings in pieces under the cohort GEMM
        return SU.score_trials(he, hia.to(device, non_blocking=True), hib.to(device, non_blocking=True), "norm",
                               hc.to(device, non_blocking=True), top=K, normalize=False).cpu()
    e2e_step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        hs = e2e_step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    assert bool(torch.isfinite(hs).all()) and torch.allclose(hs, out.cpu(), atol=1e-5)
    res = {"metric": "AS-norm trials/sec", "value": value, "unit": "trials/s", "n_gpus": world, "steps": steps,
           "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "bf16x3 (split, fp32 grade) GEMM, fp64 top-K statistics", "data": "synthetic",
           "config": {"workload": "scoring sweep (configs[4]): %d trials over %d unique embeddings vs %d-cohort, "
                                  "top-%d AS-norm, D=%d, per GPU" % (NT, U, N, K, D),
                      "l2": "the score matrix is never stored (top-K pre-filter in the GEMM epilogue); embeddings + "
                            "cohort %.0f MB per step, every step re-packs them from fp32" % ((U + N) * D * 4 / 1e6)},
           "e2e": {"value": world * NT * steps / dt, "unit": "trials/s",
                   "h2d_bytes_per_step": (U + N) * D * 4 + 2 * NT * 4, "d2h_bytes_per_step": NT * 4,
                   "api": "utils.score_trials(emb, idx_a, idx_b, 'norm', cohorts, top) from pinned host buffers"},
           "gpu_launches": launches}
    if per_rank is not None:
        res["per_rank_ms_per_step"] = {k: (v / steps if not isinstance(v, list) else [x / steps for x in v])
                                       for k, v in per_rank.items()}
    if rank == 0:
        hbm, tf_burst, tf_sust, src = peaks()
        flops = 2.0 * U * N * D
        res["roofline"] = {"kernel": "svb_asnorm_stats, whole phase: sample GEMM + threshold + FILTER GEMM "
                                     "(gemm_pw_kernel<256, ACT_FILTER>, 63 % of the phase) + candidate select",
                           "bound": "tensor", "achieved": flops * steps / (ms * 1e-3) / 1e12, "peak": tf_sust / 3.0,
                           "unit": "TFLOP/s", "frac": flops * steps / (ms * 1e-3) / 1e12 / (tf_sust / 3.0),
                           # dram__bytes_read.sum + dram__bytes_write.sum of ONE filter-GEMM launch (32 768 utterance rows x
                           # 20 032 cohort columns, ncu --set full, profiles/r02_ncu.md): 41.4 MB + 27.5 MB (the write side is
                           # the candidate lists) against (32768 + 20000) * 192 * 4 B = 40.5 MB of operands -- the 2.6 GB
                           # fp32 score tile of the chunk never leaves the SM
                           "traffic": 6.882e7 if (N == 20000 and U >= 32768) else None,
                           "traffic_unit": "bytes per filter-GEMM launch (32768 rows); operands 4.05e7",
                           "note": "algorithmic 2*U*N*D FLOP per step over the step time; fp32-grade scores cost 3 "
                                   "bf16 passes, so the bound is the sustained bf16 peak / 3 (" + src + "); the fused "
                                   "path executes 1.128x the algorithmic FLOPs (the 1/8 sample GEMM)"}
        if world == 1 and full_line:
            rate, n, dtc = cpu_scoring_rate(e, c, ia, ib, K, 10.0)
            res["cpu_baseline"] = {"value": rate, "unit": "trials/s", "cores": os.cpu_count(), "kind": "port",
                                   "sample": "%d trials of the same sweep, per-trial ZT_norm_similarity loop "
                                             "(numpy, %.1f s)" % (n, dtc)}
    return res


# ------------------------------------------------------------------ full-length eval (BASELINE.json configs[2])
def eval_signals(lengths, seed):
    """int16 PCM utterances of the given lengths (speaker-like pool, per-utterance shift and gain)."""
    from speakerverification_b200 import synth
    pool, _ = synth.utterance_set(8, 2, 320000)
    rng = np.random.default_rng(seed)
    out = []
    for n in lengths:
        a, sh, g = rng.integers(0, len(pool)), rng.integers(0, 320000), rng.uniform(0.3, 1.0)
        x = np.roll(pool[a], int(sh))[:int(n)]
        out.append(np.round(g * x / 0.3 * 20000).astype(np.int16))
    return out


def cpu_eval_rate(signals, cohort, top, min_seconds):
    """Reference CPU path of the full-length eval (oracle port, all host threads): one forward per whole utterance
    (src/model.py:380-396) + per-trial AS-norm loop (src/model.py:415-448), on a bounded sample."""
    from oracle import ecapa_oracle as eo, scoring_oracle as so
    from speakerverification_b200 import synth
    torch.set_num_threads(os.cpu_count())
    sd = {k: torch.as_tensor(v) for k, v in synth.ecapa_state_dict(CH, N_OUT).items()}
    t0 = time.perf_counter()
    embs, secs = [], 0.0
    for x in signals:
        embs.append(eo.embed((x.astype(np.float32) / 32768.0)[None], sd).numpy())
        secs += len(x) / 16000.0
        if time.perf_counter() - t0 > min_seconds and len(embs) >= 2:
            break
    t_emb = (time.perf_counter() - t0) / len(embs)
    t1 = time.perf_counter()
    nt = 0
    while time.perf_counter() - t1 < 2.0:
        so.asnorm_score(embs[nt % len(embs)], embs[(nt + 1) % len(embs)], cohort, top)
        nt += 1
    t_trial = (time.perf_counter() - t1) / nt
    cpu_eval_rate.last_embeddings = np.concatenate(embs, axis=0)      # [n, D]: reused as a parity probe by run_eval
    return t_emb, t_trial, len(embs), nt, secs


def run_eval(args, device, rank, world, dist, steps):
    """cfg3: ragged 2-20 s utterances -> whole-utterance embeddings (num_eval = 0) -> AS-norm (6 000 cohort, top-300)
    -> trial scores.  Utterances are sharded length-balanced (ModelHandling.shard_balanced); the exchange steps are the
    all-gathers of the embedding shards, of the (mu, sigma) shards and of the score slices."""
    from speakerverification_b200 import _native as nv
    from speakerverification_b200 import synth
    from speakerverification_b200 import utils as SU
    from speakerverification_b200.model import ModelHandling, SpeakerEncoder, WrappedModel
    U, N, K, NT = args.eval_utts, 6000, 300, args.eval_trials
    margs = model_args("cuda", args.precision)
    net = WrappedModel(SpeakerEncoder(**margs))
    # end-to-end launches are smaller than the device-resident ones: finer double buffering of the H2D copies
    handle = ModelHandling(net, **dict(margs, T_max=10, svb_batch_rows=4096,
                                       svb_batch_samples=args.e2e_rows * L))
    sd = net.module.state_dict()
    for k, v in synth.ecapa_state_dict(CH, N_OUT).items():
        sd[k].copy_(torch.as_tensor(v))
    net.eval()
    eng = net.module.engine()
    lengths = synth.ragged_lengths(world * U, seed=2)
    parts = ModelHandling.shard_balanced(lengths, world)
    mine = parts[rank]
    signals = eval_signals(lengths[mine], seed=50 + rank)
    cohort, _ = synth.scoring_embeddings(N, 2000, seed=103)
    coh = torch.from_numpy(cohort).to(device)
    rng = np.random.default_rng(9)
    ia_all = torch.from_numpy(rng.integers(0, world * U, world * NT).astype(np.int32)).to(device)
    ib_all = torch.from_numpy(rng.integers(0, world * U, world * NT).astype(np.int32)).to(device)
    ia, ib = ia_all[rank * NT:(rank + 1) * NT], ib_all[rank * NT:(rank + 1) * NT]
    # device-resident staging for `value`: fp32 samples back to back, launches of <= batch_samples
    # (signals of <= 3 s are wrap-padded to L + 1 samples as loadWAV does, src/processing/audio_loader.py:126-129)
    staged = [np.pad(x, (0, L + 1 - len(x)), "wrap") if len(x) <= L else x for x in signals]
    flat = torch.from_numpy(np.concatenate(staged).astype(np.float32) / 32768.0).to(device)
    groups, i, off = [], 0, 0
    while i < len(staged):
        j, tot = i, 0
        while j < len(staged) and (j == i or tot + len(staged[j]) <= args.batch_rows * L):
            tot += len(staged[j])
            j += 1
        groups.append((i, j, off, tot))
        off += tot
        i = j
    emb = torch.empty((len(signals), N_OUT), dtype=torch.float32, device=device)

    def score(e_local):
        full = ModelHandling.gather_assigned(e_local, parts, rank, world) if world > 1 else e_local
        mu, sg = SU.asnorm_stats(e_local, coh, K)
        if world > 1:
            st = ModelHandling.gather_assigned(torch.stack([mu, sg], 1), parts, rank, world)
            mu, sg = st[:, 0].contiguous(), st[:, 1].contiguous()
        sc = SU.asnorm_trials(full, ia, ib, mu, sg)
        if world > 1:
            allsc = torch.empty(world * NT, dtype=torch.float32, device=device)
            dist.all_gather_into_tensor(allsc, sc)
            return allsc
        return sc

    def step():
        for (a, b, o, tot) in groups:
            eng.forward_flat(flat[o:o + tot], [len(x) for x in staged[a:b]], l2norm=True, out=emb[a:b])   # F.normalize fused
        return score(emb)

    for _ in range(max(args.warmup, 3)):
        out = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", 0)))
    if rank == 0:
        sampler.start()
    nv.lib().svb_reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    launches = int(nv.lib().svb_launch_count())
    clocks = sampler.stop() if rank == 0 else None
    ms, per_rank = rank_stats(e0.elapsed_time(e1), {"device": device, "world": world, "dist": dist})
    # end to end: int16 PCM on the host -> GPU crop/stack stage -> encoder -> gathers -> scores on the host
    handle.embed_signals(signals[:64], 0)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_emb = torch.nn.functional.normalize(handle.embed_signals(signals, 0).view(len(signals), N_OUT), dim=1)
        hs = score(e2e_emb).cpu()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    cos = torch.nn.functional.cosine_similarity(e2e_emb, emb, dim=1)
    assert bool(torch.isfinite(hs).all()) and float(cos.min()) > 0.9999, (float(cos.min()), float((hs - out.cpu()).abs().max()))
    if rank != 0:
        return
    audio_s = float(lengths.sum()) / 16000.0
    frames = float((1 + lengths // 160).sum())
    hbm, tf_burst, tf_sust, src = peaks()
    tfl = frames * 37.50e6 * steps / (ms * 1e-3) / 1e12 / world
    line = {"metric": "full-length eval utterances/sec (2-20 s utt, fbank + ECAPA + AS-norm)",
            "value": world * U * steps / (ms * 1e-3), "unit": "utterances/s", "n_gpus": world, "steps": steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None,
            "dtype": "f16 (IEEE-half tensor-core operands, fp32 accumulate)" if args.precision != "fp32" else "bf16x3 (split, fp32 grade)",
            "data": "synthetic",
            "config": {"workload": "full-length eval (configs[2]): %d ragged U(2,20) s utterances per GPU, whole-utterance "
                                   "ECAPA-TDNN C=1024 embeddings, AS-norm vs 6000-cohort top-300, %d trials per GPU"
                                   % (U, NT),
                       "parallelism": "utterances length-balanced over %d GPU(s); all-gather of embeddings, (mu, sigma) "
                                      "and score slices" % world,
                       "audio_seconds_per_s": audio_s * steps / (ms * 1e-3),
                       "equiv_3s_embeddings_per_s": frames / 301.0 * steps / (ms * 1e-3),
                       "l2": "inputs larger than L2 (%.2f GB of samples per GPU per step)" % (flat.numel() * 4 / 1e9)},
            "e2e": {"value": world * U * steps / dt, "unit": "utterances/s",
                    "h2d_bytes_per_step": int(sum(len(x) for x in signals)) * 2, "d2h_bytes_per_step": world * NT * 4,
                    "api": "ModelHandling.embed_signals(int16 PCM list, num_eval=0) -> gathers -> utils.asnorm_* -> .cpu()"},
            "gpu_launches": launches, "clocks": clocks,
            "roofline": {"kernel": "whole step (tensor-bound encoder), algorithmic 37.50 MFLOP per frame", "bound": "tensor",
                         "achieved": tfl, "peak": tf_sust, "unit": "TFLOP/s", "frac": tfl / tf_sust, "traffic": None,
                         "peak_source": src + " (sustained)"}}
    if per_rank is not None:
        line["per_rank_ms_per_step"] = {k: (v / steps if not isinstance(v, list) else [x / steps for x in v])
                                        for k, v in per_rank.items()}
    if world == 1:
        t_emb, t_trial, ne, nt, secs = cpu_eval_rate(signals, cohort, K, 12.0 if args.workload == "eval" else 6.0)
        per_utt = t_emb * (audio_s / U) / (secs / ne) + t_trial * NT / U
        # the oracle embeddings the CPU leg just produced double as a parity probe of the timed GPU pass (same signals,
        # same random-init weights): whole-utterance embeddings of the first `ne` utterances longer than 3 s
        ref = torch.nn.functional.normalize(torch.from_numpy(cpu_eval_rate.last_embeddings), dim=1)
        keep = [i for i in range(ne) if len(signals[i]) > L]
        if keep:
            cs = (ref[keep] * emb[:ne].cpu()[keep]).sum(1)
            line["parity_check"] = {"against": "oracle.ecapa_oracle.embed (CPU fp32) on the same int16 signals",
                                    "utterances": len(keep), "min_cos": float(cs.min()),
                                    "max_pair_score_dev": float((ref[keep] @ ref[keep].T - emb[:ne].cpu()[keep] @ emb[:ne].cpu()[keep].T).abs().max())}
        line["cpu_baseline"] = {"value": 1.0 / per_utt, "unit": "utterances/s", "cores": os.cpu_count(), "kind": "port",
                                "sample": "%d whole utterances (%.0f s of audio) + %d AS-norm trials, torch/numpy oracle "
                                          "port; composed per utterance at the workload's mean length and trials per "
                                          "utterance" % (ne, secs, nt)}
    return line


def gemm_roofline(device, batch_rows, halo=4):
    """Times the dominant kernel (gemm_pw_kernel<256, GELU>) alone on the MFA shape of one encoder launch."""
    from speakerverification_b200 import _native as nv
    C3 = CH[4]
    T = 1 + L // 160
    M = batch_rows * (T + 2 * halo)
    a = torch.randn((M, C3), device=device).to(torch.float16)
    w = torch.randn((C3, C3, 1), device=device) / C3 ** 0.5
    wp = nv.pack_conv_weight(w, planes=1, f16=True)
    z = torch.zeros(C3, device=device)
    o = torch.ones(C3, device=device)
    out = torch.empty((M, C3), dtype=torch.float16, device=device)
    run = lambda: nv.tdnn_layer(a, wp, C3, C3, 1, 1, z, o, z, out, act=nv.ACT_GELU, out_kind=nv.OUT_F16, w_f16=True,
                                x_f16=True)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    reps = 10
    ev[0].record()
    for _ in range(reps):
        run()
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / reps
    flops = 2.0 * MFA_MAC_PER_FRAME * batch_rows * T       # algorithmic: valid frames only
    return flops / (ms * 1e-3) / 1e12, ms


def rank_stats(ms_local, ctx):
    """(max over ranks, {min, median, max, per_rank}) of a per-rank device time; every rank calls it."""
    if ctx["world"] == 1:
        return ms_local, None
    t = torch.tensor([ms_local], device=ctx["device"])
    allt = [torch.empty_like(t) for _ in range(ctx["world"])]
    ctx["dist"].all_gather(allt, t)
    v = sorted(float(x.item()) for x in allt)
    return v[-1], {"min": v[0], "median": float(np.median(v)), "max": v[-1], "per_rank": [float(x.item()) for x in allt]}


def bench_encoder(args, ctx, rawnet):
    """Embedding extraction (configs[1] ECAPA-TDNN C=1024 3 s crops, or configs[3] RawNet2 4 s crops): `value` with the
    waveforms resident in HBM, `e2e` through ModelHandling.embed_batches from pinned host memory.  Returns the JSON line
    (rank 0) or None."""
    from speakerverification_b200 import _native as nv
    from speakerverification_b200.model import ModelHandling, SpeakerEncoder, WrappedModel
    from speakerverification_b200 import synth
    device, rank, world, dist = ctx["device"], ctx["rank"], ctx["world"], ctx["dist"]
    Lw, n_out, flop_per_utt = (64000, 512, 13.96e9) if rawnet else (L, N_OUT, FLOP_PER_UTT)   # SURVEY.md 8a (a13, a26)
    steps = args.steps if not rawnet or args.workload == "rawnet2" else max(10, min(args.steps, 10))
    margs = rawnet_args("cuda", args.precision) if rawnet else model_args("cuda", args.precision)
    net = WrappedModel(SpeakerEncoder(**margs))
    handle = ModelHandling(net, **dict(margs, T_max=10, svb_batch_rows=args.batch_rows))
    sd = net.module.state_dict()
    weights = synth.rawnet2_state_dict() if rawnet else synth.ecapa_state_dict(CH, N_OUT)
    for k, v in weights.items():                                # random-init weights of the named architecture
        sd[k].copy_(torch.as_tensor(v))
    net.eval()
    eng = net.module.engine()

    U = args.utts
    waves = synth_waves(U, device, seed=100 + rank, length=Lw)
    emb = torch.empty((U, n_out), dtype=torch.float32, device=device)
    gathered = torch.empty((world * U, n_out), dtype=torch.float32, device=device) if world > 1 else None

    def step(n_utts=U, gather=True):
        for r0 in range(0, n_utts, args.batch_rows):
            r1 = min(n_utts, r0 + args.batch_rows)
            eng.forward(waves[r0:r1], out=emb[r0:r1])
        if world > 1 and gather:
            dist.all_gather_into_tensor(gathered, emb)          # the path's one exchange step

    def timed(n_steps, **kw):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_steps):
            step(**kw)
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        return e0.elapsed_time(e1)

    for _ in range(max(args.warmup, 3)):
        step()
    sampler = ClockSampler(ctx["local"])
    if rank == 0:
        sampler.start()
    nv.lib().svb_reset_launch_count()
    ms_local = timed(steps)
    launches = int(nv.lib().svb_launch_count())
    clocks = sampler.stop() if rank == 0 else None
    ms, per_rank = rank_stats(ms_local, ctx)
    value = world * U * steps / (ms * 1e-3)

    # strong scaling: ONE set of U utterances split over the ranks (weak scaling is the headline)
    strong = None
    if world > 1:
        share = (U + world - 1) // world
        emb_s = torch.empty((world * share, n_out), dtype=torch.float32, device=device)

        def strong_step():
            for r0 in range(0, share, args.batch_rows):
                r1 = min(share, r0 + args.batch_rows)
                eng.forward(waves[r0:r1], out=emb[r0:r1])
            dist.all_gather_into_tensor(emb_s, emb[:share])
        strong_step()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            strong_step()
        e1.record()
        torch.cuda.synchronize()
        ms_s, _ = rank_stats(e0.elapsed_time(e1), ctx)
        strong = {"value": world * share * steps / (ms_s * 1e-3), "unit": "embeddings/s", "ms_per_step": ms_s / steps,
                  "utterances_total": world * share, "scaling": "strong"}

    # end to end through the public API from pinned host buffers
    e2e = None
    if not args.no_e2e:
        host = torch.empty((U, Lw), dtype=torch.float32).pin_memory()
        host.copy_(waves)
        torch.cuda.synchronize()
        handle.embed_batches(host[:4 * args.batch_rows]).cpu()  # warm-up
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            out = handle.embed_batches(host).cpu()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {"value": world * U * steps / dt, "unit": "embeddings/s", "h2d_bytes_per_step": U * Lw * 4,
               "d2h_bytes_per_step": U * n_out * 4, "api": "ModelHandling.embed_batches(pinned host [U, L]) -> .cpu()"}
        # the e2e embeddings are the device-resident pass's embeddings (same kernels, same inputs)
        cos = torch.nn.functional.cosine_similarity(out.to(device), emb, dim=1)
        assert out.shape == (U, n_out) and bool(torch.isfinite(out).all()) and float(cos.min()) > 0.99999, float(cos.min())
        del host
    if rank != 0:
        return None
    hbm, tf_burst, tf_sust, src = peaks()
    dtype = "f16 (IEEE-half tensor-core operands, fp32 accumulate)" if args.precision != "fp32" else "bf16x3 (split, fp32 grade)"
    line = {"metric": "RawNet2 embeddings/sec (4 s utt)" if rawnet else "ECAPA embeddings/sec (3 s utt)",
            "value": value, "unit": "embeddings/s", "n_gpus": world, "steps": steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dtype,
            "data": "synthetic", "e2e": e2e, "gpu_launches": launches, "clocks": clocks}
    if per_rank is not None:
        line["per_rank_ms_per_step"] = {k: (v / steps if not isinstance(v, list) else [x / steps for x in v])
                                        for k, v in per_rank.items()}
    if strong is not None:
        line["strong_scaling"] = strong
    line["step_tensor"] = {"achieved_tflops": value / world * flop_per_utt / 1e12, "peak": tf_sust,
                           "frac": value / world * flop_per_utt / 1e12 / tf_sust,
                           "note": "whole step, algorithmic %.2f GFLOP per utterance vs sustained bf16 peak" % (flop_per_utt / 1e9)}
    if rawnet:
        # dominant kernel: the sliding-window sinc GEMM (front end = LayerNorm + sinc GEMM, timed alone)
        os.environ["SVB_RAWNET_STOP"] = "-1"
        wv = waves[:args.batch_rows]
        for _ in range(3):
            eng.forward(wv)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        for _ in range(10):
            eng.forward(wv)
        ev[1].record()
        torch.cuda.synchronize()
        del os.environ["SVB_RAWNET_STOP"]
        front_ms = ev[0].elapsed_time(ev[1]) / 10
        ach = 4.10e9 * wv.shape[0] / (front_ms * 1e-3) / 1e12
        line["config"] = {"workload": "RawNet2_custom (sinc front-end + 8 residual blocks + GRU) nOut=512, %d synthetic 4 s "
                                      "16 kHz utterances per GPU (configs[3])" % U,
                          "batch_rows": args.batch_rows, "parallelism": "utterance-sharded x%d" % world,
                          "l2": "inputs larger than L2 (%.2f GB of waveforms per step)" % (U * Lw * 4 / 1e9)}
        line["roofline"] = {"kernel": "rn_gemm_kernel<EPI_SINC> (+ LayerNorm kernel) of one %d-utterance launch" % wv.shape[0],
                            "bound": "tensor", "achieved": ach, "peak": tf_burst, "unit": "TFLOP/s", "frac": ach / tf_burst,
                            "traffic": None, "peak_source": src + " (burst, kernel timed alone)", "kernel_ms": front_ms,
                            "note": "algorithmic 4.10 GFLOP per utterance (SURVEY 8a a23)"}
        if world == 1:
            rate, n, dt = cpu_rawnet_rate(12.0 if args.workload == "rawnet2" else 6.0)
            line["cpu_baseline"] = {"value": rate, "unit": "embeddings/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": "%d x 4 s utterances, batch 8, torch fp32 oracle port (%.1f s)" % (n, dt)}
    else:
        if world == 1:
            # baseline leg: the CPU oracle as a CHECKER of the timed pass (16 of its utterances, same random-init weights)
            from oracle import ecapa_oracle as eo
            osd = {k: torch.as_tensor(v) for k, v in weights.items()}
            ref = torch.nn.functional.normalize(eo.embed(waves[:16].cpu(), osd), dim=1)
            got = torch.nn.functional.normalize(emb[:16].cpu(), dim=1)
            line["parity_check"] = {"against": "oracle.ecapa_oracle.embed (CPU fp32) on 16 utterances of the timed pass",
                                    "min_cos": float((ref * got).sum(1).min()),
                                    "max_pair_score_dev": float((ref @ ref.T - got @ got.T).abs().max())}
        ach, gemm_ms = gemm_roofline(device, args.batch_rows)
        line["config"] = {"workload": "ECAPA-TDNN C=1024 nOut=192, 80-mel fbank (16 kHz, n_fft 512, win 400, hop 160), "
                                      "%d synthetic 3 s utterances per GPU (configs[1])" % U,
                          "batch_rows": args.batch_rows, "parallelism": "utterance-sharded x%d" % world,
                          "l2": "inputs larger than L2 (%.2f GB of waveforms per step)" % (U * Lw * 4 / 1e9)}
        line["roofline"] = {"kernel": "gemm_pw_kernel<256,GELU> (MFA layer 3C x 3C of one %d-utterance launch)" % args.batch_rows,
                            "bound": "tensor", "achieved": ach, "peak": tf_burst, "unit": "TFLOP/s", "frac": ach / tf_burst,
                            "traffic": MFA_DRAM_BYTES if args.batch_rows == 592 else None,
                            "traffic_unit": "bytes per launch: dram__bytes_read.sum + dram__bytes_write.sum of the ncu --set "
                                            "full capture of this kernel at this shape (profiles/); algorithmic 2.267e9",
                            "peak_source": src + " (burst, kernel timed alone)", "kernel_ms": gemm_ms}
    del waves, emb, eng, handle, net
    torch.cuda.empty_cache()
    return line


def gpu_library_baseline(device, batch=592, reps=3):
    """SURVEY 8d 'secondary baseline (the library kernels to beat)': the reference's forward as PyTorch eager launches it
    on this same B200 -- oracle.ecapa_oracle.embed (the functional restatement of SpeakerEncoder.forward, pinned to the
    reference to <= 1e-5) moved to CUDA: cuFFT + cuBLAS + cuDNN kernels, fp32 and under torch.autocast(bfloat16).
    Baseline leg only (nothing under speakerverification_b200/ can import it).  Same model (C=1024), same 3 s crops,
    same launch size; CUDA-event timed."""
    from oracle import ecapa_oracle as eo
    from speakerverification_b200 import synth
    sd = {k: torch.as_tensor(v).to(device) for k, v in synth.ecapa_state_dict(CH, N_OUT).items()}
    wav = synth_waves(batch, device, seed=7)
    out = {"batch": batch, "model": "ECAPA-TDNN C=1024, 3 s crops, torch %s eager (cuFFT / cuBLAS / cuDNN)" % torch.__version__}
    for name, ctxmgr in (("fp32", None), ("autocast_bf16", torch.autocast("cuda", dtype=torch.bfloat16))):
        def run():
            if ctxmgr is None:
                return eo.embed(wav, sd)
            with ctxmgr:
                return eo.embed(wav, sd)
        try:
            for _ in range(2):
                e = run()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                e = run()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            out[name] = {"value": batch / (ms * 1e-3), "unit": "embeddings/s", "ms_per_launch": ms,
                         "finite": bool(torch.isfinite(e.float()).all())}
        except Exception as ex:                                  # e.g. out of memory on a shared box: report, do not fail
            out[name] = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:200])}
        torch.cuda.empty_cache()
    return out


def api_distributed_check(ctx, precision):
    """evaluateFromList(distributed=True) on `world` ranks returns, on rank 0, the scores of the single-GPU call and
    ([], [], []) on the other ranks, as src/model.py:400-450 prescribes.  C=512 model, 24 synthetic wav files, 200
    trials, 300-row cohort.  The sharded run embeds every utterance at another position of another launch; the fused
    SE / ASP column statistics sum fp32 partials per 32-row box of the launch (gemm_pw.cu), so an embedding moves by
    fp32 summation-order noise (~1e-7 relative: measured 9.5e-7 on cosine scores, run-to-run bit reproducible) --
    the same class of difference cuDNN / cuBLAS show across batch compositions in the reference.  Criteria: cosine
    <= 5e-6; AS-norm <= 5e-6 / sigma_min (the normalisation divides by the cohort sigma)."""
    from scipy.io import wavfile
    from speakerverification_b200 import synth
    from speakerverification_b200.model import ModelHandling, SpeakerEncoder, WrappedModel
    rank, world, dist, device = ctx["rank"], ctx["world"], ctx["dist"], ctx["device"]
    root = "/tmp/svb_bench_api"
    ch = [512, 512, 512, 512, 1536]
    margs = dict(model_args("cuda", precision), channels=ch, save_folder=root)
    waves, spk = synth.utterance_set(6, 4, L + 1)
    lab, ti, tj = synth.trial_list(spk, 200)
    names = [os.path.join(root, "utt%03d.wav" % i) for i in range(len(waves))]
    if rank == 0:
        os.makedirs(root, exist_ok=True)
        for n, w in zip(names, waves):
            wavfile.write(n, 16000, w)
        with open(os.path.join(root, "trials.txt"), "w") as f:
            for a, i, j in zip(lab, ti, tj):
                f.write("%d %s %s\n" % (a, names[i], names[j]))
        coh, _ = synth.scoring_embeddings(300, 50, seed=5)
        np.save(os.path.join(root, "cohorts.npy"), coh)
    dist.barrier()
    net = WrappedModel(SpeakerEncoder(**margs))
    h = ModelHandling(net, **dict(margs, T_max=10))
    sd = net.module.state_dict()
    for k, v in synth.ecapa_state_dict(ch, N_OUT).items():
        sd[k].copy_(torch.as_tensor(v))
    net.eval()
    res = {}
    ok_empty = True
    for mode in ("cosine", "norm"):
        sc, lb, tr = h.evaluateFromList(os.path.join(root, "trials.txt"), distributed=True, dataloader_options={"num_workers": 0},
                                        cohorts_path=os.path.join(root, "cohorts.npy"), num_eval=2, scoring_mode=mode, top=200)
        if rank == 0:
            s1, l1, t1 = h.evaluateFromList(os.path.join(root, "trials.txt"), distributed=False,
                                            dataloader_options={"num_workers": 0},
                                            cohorts_path=os.path.join(root, "cohorts.npy"), num_eval=2, scoring_mode=mode, top=200)
            d = float(np.abs(np.asarray(sc) - np.asarray(s1)).max())
            res[mode] = {"max_abs_dev_vs_single_gpu": d, "bit_equal": bool(np.array_equal(np.asarray(sc), np.asarray(s1))),
                         "labels_trials_equal": lb == l1 and tr == t1, "n_trials": len(sc)}
            if mode == "norm":      # sigma of the top-200 cohort scores, from the single-GPU embeddings
                from speakerverification_b200 import utils as SU
                files = sorted(set(names))
                em = SU.mean_embeddings(h._embed_files(files, 2), True)
                _, sg = SU.asnorm_stats(em, torch.from_numpy(np.load(os.path.join(root, "cohorts.npy"))).to(device), 200)
                res[mode]["sigma_min"] = float(sg.min().item())
        else:
            ok_empty = ok_empty and (sc, lb, tr) == ([], [], [])
    flag = torch.tensor([1 if ok_empty else 0], device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank != 0:
        return None
    res["other_ranks_returned_empty"] = bool(flag.item())
    res["passed"] = bool(res["cosine"]["max_abs_dev_vs_single_gpu"] <= 5e-6 and
                         res["norm"]["max_abs_dev_vs_single_gpu"] <= 5e-6 / res["norm"]["sigma_min"] and
                         res["cosine"]["labels_trials_equal"] and res["norm"]["labels_trials_equal"] and
                         res["other_ranks_returned_empty"])
    return res


SUB_KEYS = ("metric", "value", "unit", "steps", "ms_per_step", "dtype", "config", "e2e", "roofline", "gpu_launches",
            "per_rank_ms_per_step", "strong_scaling", "step_tensor", "cpu_baseline", "parity_check")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="svb", choices=["svb", "reference"])
    ap.add_argument("--utts", type=int, default=10000)
    ap.add_argument("--batch-rows", type=int, default=592)   # 4 x 148: whole waves of the per-utterance chain kernel
    ap.add_argument("--precision", default="f16", choices=["f16", "bf16", "fp32"])   # 'bf16' = the 16-bit mode's alias
    ap.add_argument("--no-e2e", action="store_true")
    # 'all' (default): the headline line is configs[1] (embed) and carries configs[2] / [3] / [4] as sub-objects
    ap.add_argument("--workload", default="all", choices=["all", "embed", "scoring", "rawnet2", "eval"])
    ap.add_argument("--e2e-rows", type=int, default=296)
    ap.add_argument("--eval-utts", type=int, default=2000)
    ap.add_argument("--eval-trials", type=int, default=100000)
    ap.add_argument("--score-utts", type=int, default=200000)
    ap.add_argument("--score-cohort", type=int, default=20000)
    ap.add_argument("--score-trials", type=int, default=1000000)
    ap.add_argument("--no-library-baseline", action="store_true")
    args = ap.parse_args()
    if args.precision == "bf16":
        args.precision = "f16"
    quiet_stdout()
    if args.impl == "reference":
        if args.workload == "all":
            args.workload = "embed"
        return run_reference(args)

    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    ctx = {"device": device, "rank": rank, "world": world, "local": local, "dist": dist}
    sub_steps = max(10, min(args.steps, 10))

    def finish(line):
        if rank == 0 and line is not None:
            emit(json.dumps(line))
        if world > 1:
            dist.destroy_process_group()

    if args.workload == "scoring":
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        res = run_scoring(args, device, rank, world, dist, args.steps)
        if rank == 0:
            res["warmup"] = 3
            res["clocks"] = sampler.stop()
        return finish(res if rank == 0 else None)
    if args.workload == "eval":
        return finish(run_eval(args, device, rank, world, dist, args.steps))
    if args.workload in ("embed", "rawnet2"):
        line = bench_encoder(args, ctx, rawnet=args.workload == "rawnet2")
        if line is not None and world == 1 and args.workload == "embed":
            rate, n, dt = cpu_reference_rate(64, min_seconds=12.0)
            line["cpu_baseline"] = {"value": rate, "unit": "embeddings/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": "%d x 3 s utterances, batch 16, torch fp32 oracle port (%.1f s)" % (n, dt)}
        return finish(line)

    # ---- default: every BASELINE.json config in one driver-observed line
    line = bench_encoder(args, ctx, rawnet=False)                       # configs[1]: the headline keys
    torch.cuda.empty_cache()
    ev = run_eval(args, device, rank, world, dist, sub_steps)           # configs[2]
    torch.cuda.empty_cache()
    rn = bench_encoder(args, ctx, rawnet=True)                          # configs[3]
    torch.cuda.empty_cache()
    sc = run_scoring(args, device, rank, world, dist, sub_steps, full_line=True)   # configs[4]
    torch.cuda.empty_cache()
    api = api_distributed_check(ctx, args.precision) if world > 1 else None
    if rank == 0:
        line["eval"] = {k: ev[k] for k in SUB_KEYS if k in ev}
        line["rawnet2"] = {k: rn[k] for k in SUB_KEYS if k in rn}
        line["scoring"] = {k: sc[k] for k in SUB_KEYS if k in sc}
        if api is not None:
            line["api_distributed_check"] = api
        if world == 1:
            if not args.no_library_baseline:
                lib = gpu_library_baseline(device, args.batch_rows)
                for k in ("fp32", "autocast_bf16"):
                    if "value" in lib.get(k, {}):
                        lib[k]["svb_speedup"] = line["value"] / lib[k]["value"]
                line["gpu_library_baseline"] = lib
            rate, n, dt = cpu_reference_rate(64, min_seconds=12.0)
            line["cpu_baseline"] = {"value": rate, "unit": "embeddings/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": "%d x 3 s utterances, batch 16, torch fp32 oracle port (%.1f s)" % (n, dt)}
    finish(line)


if __name__ == "__main__":
    main()
