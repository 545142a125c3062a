"""Drop-in API parity: the host-side mirror of src/model.py driven exactly like the reference drives
ModelHandling (trial list files in, scores / labels / trials out), checked against the outputs the
reference itself produced for the same synthetic wavs, trial list and checkpoint (tests/golden)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

L = 48000
CH512 = [512, 512, 512, 512, 1536]


CH1024 = [1024, 1024, 1024, 1024, 3072]


def _args(tmp, precision, channels=CH512):
    return dict(device="cuda", gpu=0, seed=0, save_folder=str(tmp),
                model={"name": "ECAPA_TDNN", "nOut": 192}, channels=channels,
                criterion={"name": "AAmSoftmaxAP", "margin": 0.2, "scale": 30}, nClasses=10,
                classifier={"input_size": 192, "out_neurons": 10},
                optimizer={"name": "adam", "weight_decay": 2e-5, "lr_decay": 0.95}, lr=1e-3,
                callbacks={"name": "steplr"}, step_size=10,
                features="melspectrogram", lib="torchaudio", n_mels=80, n_mfcc=80,
                sr=16000, n_fft=512, win_length=400, hop_length=160,
                audio_spec={"sample_rate": 16000, "channels": 1, "sentence_len": 3.0, "win_len": 0.025, "hop_len": 0.01},
                dataloader_options={"nPerSpeaker": 2, "num_workers": 4, "batch_size": 4, "shuffle": False},
                augment=False, augment_options={"augment_chain": []}, mixedprec=False, svb_precision=precision)


def _make_workdir(tmp, golden, channels):
    from scipy.io import wavfile
    from oracle import synth
    waves, spk = synth.utterance_set(16, 4, L + 1)
    names = []
    for i, w in enumerate(waves):
        p = str(tmp / ("utt%03d.wav" % i))
        wavfile.write(p, 16000, w)
        names.append(p)
    lab, ti, tj = synth.trial_list(spk, 1000)
    trial_txt = str(tmp / "trials.txt")
    with open(trial_txt, "w") as f:
        for a, i, j in zip(lab, ti, tj):
            f.write("%d %s %s\n" % (a, names[i], names[j]))
    # checkpoint in the reference layout, with a 'module.' prefixed key and loss weights
    bn = {k[3:]: v for k, v in golden.items() if k.startswith("bn/")}
    sd = {k: torch.as_tensor(v) for k, v in synth.ecapa_state_dict(channels, 192, bn_stats=bn).items()}
    sd["module.__S__.fc.conv.bias"] = sd.pop("__S__.fc.conv.bias")
    sd["__L__.aamsoftmax.weight"] = torch.arange(12.0).view(3, 4)
    sd["compute_features.1.some_nnaudio_buffer"] = torch.zeros(3)
    ckpt = str(tmp / "best_state.pt")
    torch.save(sd, ckpt)
    np.save(str(tmp / "cohorts.npy"), golden["cohort"])
    return dict(tmp=tmp, names=names, waves=waves, trials=trial_txt, lab=lab, ti=ti, tj=tj, ckpt=ckpt,
                cohorts=str(tmp / "cohorts.npy"), channels=channels)


@pytest.fixture(scope="module")
def workdir(tmp_path_factory, golden_c512):
    return _make_workdir(tmp_path_factory.mktemp("svb_api"), golden_c512, CH512)


@pytest.fixture(scope="module")
def workdir1024(tmp_path_factory, golden_c1024):
    return _make_workdir(tmp_path_factory.mktemp("svb_api1024"), golden_c1024, CH1024)


def _handle(workdir, precision):
    from speakerverification_b200.model import ModelHandling, SpeakerEncoder, WrappedModel
    args = _args(workdir["tmp"], precision, workdir["channels"])
    net = WrappedModel(SpeakerEncoder(**args))
    h = ModelHandling(net, **dict(args, T_max=10))
    h.loadParameters(workdir["ckpt"], show_error=False)
    net.eval()
    return net, h, args


def test_evaluate_from_list_fp32(workdir, golden_c512):
    from speakerverification_b200 import utils as U
    net, h, args = _handle(workdir, "fp32")
    sc, lab, trials = h.evaluateFromList(workdir["trials"], distributed=False,
                                         dataloader_options=args["dataloader_options"],
                                         cohorts_path=workdir["cohorts"], num_eval=2, scoring_mode="cosine")
    assert lab == list(workdir["lab"]) and len(trials) == 1000
    assert trials[0] == workdir["names"][workdir["ti"][0]] + " " + workdir["names"][workdir["tj"][0]]
    dev = np.abs(np.asarray(sc) - golden_c512["scores_cos"]).max()
    print("fp32 list-API cosine score max dev", dev)
    assert dev <= 1e-4
    res = U.tuneThresholdfromScore(sc, lab, [1, 0.1])
    fn, fp, th = U.ComputeErrorRates(sc, lab)
    dcf, _ = U.ComputeMinDcf(fn, fp, th, 0.05, 1, 1)
    assert round(float(res["roc"][1]), 4) == round(float(golden_c512["eer/scores_cos"]), 4)
    assert round(dcf, 4) == round(float(golden_c512["mindcf/scores_cos"]), 4)
    # AS-norm through the API (the reference's own call site raises TypeError)
    sn, _, _ = h.evaluateFromList(workdir["trials"], distributed=False, dataloader_options=args["dataloader_options"],
                                  cohorts_path=workdir["cohorts"], num_eval=2, scoring_mode="norm", top=300)
    devn = np.abs(np.asarray(sn) - golden_c512["scores_asnorm300"]).max()
    print("fp32 list-API AS-norm score max dev", devn)
    # normalised scores amplify embedding error by 1/sigma (sigma ~ 0.03-0.1): 1e-4 holds for the scoring
    # kernels on identical embeddings (test_scoring_gpu.py); end to end the bound is 1e-4 / sigma_min.
    # Measured on B200 (round 2): 2.5e-4 with a raw-cosine deviation of 4.9e-5; asserted at 2x.
    assert devn <= 5e-4


def test_evaluate_from_list_f16_c1024(workdir1024, golden_c1024):
    """BASELINE.json configs[1] (C=1024, 16-bit mode = what bench.py times): trial scores agree with the reference fp32
    path to <= 2e-3 (north star) -- asserted at 5e-4, the CPU model of the data flow says 2.0e-4 -- and EER / minDCF
    equal the reference's at four decimals.  (The decisive score gap at the EER threshold of this list is 7e-5, trials
    936 / 501: bf16 operands, deviation ~3e-4 there, flipped the EER by one ROC step; see profiles/r02_precision.md.)"""
    from speakerverification_b200 import utils as U
    net, h, args = _handle(workdir1024, "f16")
    sc, lab, _ = h.evaluateFromList(workdir1024["trials"], distributed=False,
                                    dataloader_options=args["dataloader_options"],
                                    cohorts_path=workdir1024["cohorts"], num_eval=2, scoring_mode="cosine")
    d = np.abs(np.asarray(sc) - golden_c1024["scores_cos"])
    print("f16 C=1024 list-API cosine score dev: max", d.max(), "mean", d.mean())
    assert d.max() <= 5e-4
    # EER / minDCF at the benchmark's precision equal the reference's at four decimals (north star)
    res = U.tuneThresholdfromScore(sc, lab, [1, 0.1])
    fn, fp, th = U.ComputeErrorRates(sc, lab)
    dcf, _ = U.ComputeMinDcf(fn, fp, th, 0.05, 1, 1)
    print("f16 C=1024 EER", res["roc"][1], "reference", float(golden_c1024["eer/scores_cos"]),
          "minDCF", dcf, "reference", float(golden_c1024["mindcf/scores_cos"]))
    assert round(float(res["roc"][1]), 4) == round(float(golden_c1024["eer/scores_cos"]), 4)
    assert round(dcf, 4) == round(float(golden_c1024["mindcf/scores_cos"]), 4)


def test_evaluate_from_list_f16_c512(workdir, golden_c512):
    """C=512 in the 16-bit mode: the north-star score tolerance with margin, plus EER / minDCF at four decimals."""
    from speakerverification_b200 import utils as U
    net, h, args = _handle(workdir, "f16")
    sc, lab, _ = h.evaluateFromList(workdir["trials"], distributed=False, dataloader_options=args["dataloader_options"],
                                    cohorts_path=workdir["cohorts"], num_eval=2, scoring_mode="cosine")
    d = np.abs(np.asarray(sc) - golden_c512["scores_cos"])
    print("f16 C=512 list-API cosine score dev: max", d.max(), "mean", d.mean())
    assert d.max() <= 5e-4
    res = U.tuneThresholdfromScore(sc, lab, [1, 0.1])
    fn, fp, th = U.ComputeErrorRates(sc, lab)
    dcf, _ = U.ComputeMinDcf(fn, fp, th, 0.05, 1, 1)
    assert round(float(res["roc"][1]), 4) == round(float(golden_c512["eer/scores_cos"]), 4)
    assert round(dcf, 4) == round(float(golden_c512["mindcf/scores_cos"]), 4)


def test_embed_utterance_prepare_and_checkpoint(workdir, golden_c512):
    from oracle import synth
    net, h, args = _handle(workdir, "fp32")
    e = h.embed_utterance(workdir["names"][3], num_eval=2, normalize=False)
    assert e.device.type == "cpu" and tuple(e.shape) == (2, 192)
    ref = golden_c512["emb_list"][3]
    cos = torch.nn.functional.cosine_similarity(e, torch.from_numpy(ref), dim=1)
    assert cos.min().item() >= 0.9999
    # direct batch call the reference exposes: net(tensor[B, L]) with a HOST tensor
    emb = net(torch.from_numpy(workdir["waves"][:8, :L]))
    assert emb.is_cuda and tuple(emb.shape) == (8, 192)
    c = torch.nn.functional.cosine_similarity(emb.cpu(), torch.from_numpy(golden_c512["emb_batch"][:8]), dim=1)
    assert c.min().item() >= 0.9999
    # cohort preparation (src/model.py:578-609) for the first 12 cohort speakers
    from scipy.io import wavfile
    cf0, cenv = synth.speaker_bank(300, seed=77)
    meta = str(workdir["tmp"] / "cohort_meta.txt")
    with open(meta, "w") as f:
        for s in range(12):
            p = str(workdir["tmp"] / ("coh%04d.wav" % s))
            wavfile.write(p, 16000, synth.utterance(cf0[s], cenv[s], L + 1, 50000 + s))
            f.write("spk%d %s\n" % (s, p))
    out = str(workdir["tmp"] / "cohort_out.npy")
    assert h.prepare(save_path=out, prepare_type="cohorts", num_eval=2, source=meta) is True
    got = np.load(out)
    assert got.shape == (12, 192)
    assert np.abs(got - golden_c512["cohort"][:12]).max() <= 1e-4
    # checkpoint round trip keeps the reference key set, including adopted loss weights
    path = str(workdir["tmp"] / "resaved.pt")
    h.saveParameters(path)
    sd = torch.load(path)
    assert "__L__.aamsoftmax.weight" in sd and torch.equal(sd["__L__.aamsoftmax.weight"].cpu(), torch.arange(12.0).view(3, 4))
    keys = [k for k in sd if k.startswith("__S__.")]
    assert len(keys) == 231
    assert {"compute_features.0.flipped_filter", "compute_features.1.spectrogram.window",
            "compute_features.1.mel_scale.fb"} <= set(sd)


def test_plugin_interfaces_and_test_from_list(workdir, golden_c512):
    net, h, args = _handle(workdir, "fp32")
    enc = net.module
    wav = torch.from_numpy(workdir["waves"][:2, :L]).cuda()
    mel = enc.compute_features(wav)                                   # feature plugin: [B, n_mels, T]
    assert tuple(mel.shape) == (2, 80, 301)
    ref = torch.from_numpy(golden_c512["stage/mel"])
    rel = ((mel[0].cpu() - ref).abs() / (ref.abs() + 1e-6)).max().item()
    assert rel < 1e-3
    emb = enc.__S__(mel)                                              # encoder plugin: features -> embedding
    c = torch.nn.functional.cosine_similarity(emb.cpu(), torch.from_numpy(golden_c512["emb_batch"][:2]), dim=1)
    assert c.min().item() >= 0.9999
    # testFromList: CSV with header in, result CSV + strings out
    csv_in = str(workdir["tmp"] / "pairs.csv")
    with open(csv_in, "w") as f:
        f.write("audio_1,audio_2\n")
        for i, j in zip(workdir["ti"][:50], workdir["tj"][:50]):
            f.write("%s,%s\n" % (workdir["names"][i], workdir["names"][j]))
    out_csv = str(workdir["tmp"] / "pairs_result.csv")
    res = h.testFromList(csv_in, thresh_score=0.5, cohorts_path=workdir["cohorts"], num_eval=2,
                         scoring_mode="cosine", output_file=out_csv)
    assert len(res) == 50 and os.path.exists(out_csv)
    a, b, s = res[0].split(",")
    assert a == "utt%03d.wav" % workdir["ti"][0] and abs(float(s) - golden_c512["scores_cos"][0]) <= 1e-4
    with open(out_csv) as f:
        rows = f.read().splitlines()
    assert rows[0] == "audio_1,audio_2,pred_label,score" and len(rows) == 51


def test_benchmark_consumer_flow(workdir, golden_c512):
    """The call sequence of the reference's integration consumer src/benchmark/benmark_model.py:55-125:
    evaluateFromList -> tuneThresholdfromScore over 50 target-FA points -> best sum rate / EER / g-mean threshold ->
    evaluation_results.txt -> testFromList at that threshold; every decision checked against the same flow run on the
    reference's own scores (tests/golden)."""
    import csv
    from speakerverification_b200 import utils as U
    net, h, args = _handle(workdir, "fp32")
    sc, lab, trials = h.evaluateFromList(listfilename=workdir["trials"], distributed=False,
                                         dataloader_options=args["dataloader_options"], cohorts_path=workdir["cohorts"],
                                         num_eval=2, scoring_mode="cosine")
    target_fa = np.linspace(5, 0, num=50)
    res = U.tuneThresholdfromScore(sc, lab, target_fa)
    ref = U.tuneThresholdfromScore(list(golden_c512["scores_cos"]), lab, target_fa)
    assert set(res) == {"gmean", "roc", "prec_recall"} and len(res["roc"][0]) == 50

    def best(r):
        rates = [row[1] + row[2] for row in r["roc"][0]]
        return int(np.argmin(rates)), min(rates)
    assert abs(best(res)[1] - best(ref)[1]) <= 0.25       # best FA + FR sum rate in %, one trial = 0.1-1 %
    assert round(float(res["roc"][1]), 4) == round(float(ref["roc"][1]), 4)            # EER, 4 decimals
    assert abs(res["roc"][2] - ref["roc"][2]) <= 1e-2     # AUC (x100 scale): scores within 1e-4 may swap neighbouring ranks
    assert res["gmean"][0] == ref["gmean"][0] and abs(res["gmean"][2] - ref["gmean"][2]) <= 1e-4
    thr = res["gmean"][-1]
    out = workdir["tmp"] / "evaluation_results.txt"
    with open(out, "w", newline="") as wf:
        w = csv.writer(wf, delimiter=",")
        w.writerow(["audio_1", "audio_2", "label", "predict_label", "score"])
        for score, label, pair in zip(sc, lab, trials):
            com, refp = pair.strip().split(" ")
            w.writerow([com, refp, label, int(score >= thr), score])
    preds = np.asarray(sc) >= thr
    gold = golden_c512["scores_cos"]
    clear = np.abs(gold - ref["gmean"][-1]) > 2e-4
    assert clear.sum() >= 990 and np.array_equal(preds[clear], (gold >= ref["gmean"][-1])[clear])
    # the verification list goes through testFromList at the tuned threshold
    csv_in = str(workdir["tmp"] / "verify.csv")
    with open(csv_in, "w") as f:
        f.write("audio_1,audio_2\n")
        for pair in trials[:200]:
            a, b = pair.split(" ")
            f.write("%s,%s\n" % (a, b))
    out_csv = str(workdir["tmp"] / "verify_result.csv")
    got = h.testFromList(csv_in, thresh_score=thr, distributed=False, dataloader_options=args["dataloader_options"],
                         cohorts_path=workdir["cohorts"], num_eval=2, scoring_mode="cosine", output_file=out_csv)
    assert len(got) == 200
    with open(out_csv) as f:
        rows = list(csv.reader(f))[1:]
    for k, row in enumerate(rows):
        assert abs(float(row[3]) - gold[k]) <= 1e-4
        if clear[k]:
            assert row[2] == ("1" if gold[k] >= ref["gmean"][-1] else "0")


def test_no_cpu_path(workdir):
    from speakerverification_b200.model import SpeakerEncoder
    with pytest.raises(RuntimeError):
        SpeakerEncoder(**dict(_args(workdir["tmp"], "fp32"), device="cpu"))


@pytest.mark.parametrize("precision", ["fp32", "f16"])
def test_ragged_eval_asnorm_scores_vs_oracle(workdir, golden_c512, precision):
    """BASELINE.json configs[2] at test size, score level: 16 ragged 2-20 s utterances -> embed_signals(num_eval=0)
    (whole-utterance embeddings, int16 PCM in) -> AS-norm against a 6000-row cohort, top-300 -> all 120 pair scores,
    against oracle.ecapa_oracle.embed + oracle.scoring_oracle (cosine_score / asnorm_score) on the same signals.
    The cohort bank is an INPUT of the scoring stage (6000 one-second synthetic utterances embedded once, fp32 mode)
    and is handed to both sides.  Tolerances: raw cosine <= 1e-4 (fp32) / 2e-3 north star, asserted 5e-4 (16-bit mode);
    AS-norm divides by the cohort sigma, so its bound is the cosine bound / sigma_min (computed by the oracle)."""
    from oracle import ecapa_oracle as eo, scoring_oracle as so, synth
    from speakerverification_b200 import utils as U
    net, h, args = _handle(workdir, precision)
    sd = synth.ecapa_state_dict(CH512, 192, bn_stats={k[3:]: v for k, v in golden_c512.items() if k.startswith("bn/")})
    lengths = synth.ragged_lengths(16, seed=21)
    lengths[0], lengths[1] = 32000, 320000                       # both ends of the configs[2] range
    f0, env = synth.speaker_bank(6, seed=31)
    sig = [synth.utterance(f0[i % 6], env[i % 6], int(n), 700 + i) for i, n in enumerate(lengths)]
    pcm = [np.round(s / np.abs(s).max() * 20000).astype(np.int16) for s in sig]
    emb = h.embed_signals(pcm, 0)                                # [16, 1, 192]
    # cohort: 6000 one-second utterances of 300 cohort speakers, embedded in fp32 mode (input data for both sides)
    hc = h if precision == "fp32" else _handle(workdir, "fp32")[1]
    hc.audio_spec = dict(hc.audio_spec, sentence_len=1.0)
    cf0, cenv = synth.speaker_bank(300, seed=77)
    rng = np.random.default_rng(8)
    base = [synth.utterance(cf0[s], cenv[s], 16001, 60000 + s) for s in range(300)]
    coh_sig = [np.round((base[i % 300] * rng.uniform(0.3, 1.0) + 0.02 * rng.standard_normal(16001)) * 30000).astype(np.int16)
               for i in range(6000)]
    cohort = torch.nn.functional.normalize(hc.embed_signals(coh_sig, 1)[:, 0], dim=1).cpu().numpy()
    hc.audio_spec = dict(hc.audio_spec, sentence_len=3.0)
    ti, tj = np.triu_indices(16, 1)
    ia, ib = torch.from_numpy(ti.astype(np.int32)).cuda(), torch.from_numpy(tj.astype(np.int32)).cuda()
    got_cos = U.score_trials(emb, ia, ib, "cosine").cpu().numpy()
    got_as = U.score_trials(emb, ia, ib, "norm", cohort, top=300, normalize=True).cpu().numpy()
    # oracle: one forward per whole utterance (src/model.py:380-396), then the per-trial scoring functions
    # (loadWAV wrap-pads signals of <= sentence_len to sentence_len + 1 sample, src/processing/audio_loader.py:126-129)
    whole = [eo.load_wav_eval(p.astype(np.float32) / 32768.0, 3.0, 16000, 0)[0] for p in pcm]
    ref = np.stack([eo.embed(np.stack([x, x]), sd).numpy()[0] for x in whole])
    rn = so.l2_normalize(ref)
    want_cos = np.array([so.cosine_score(ref[i][None], ref[j][None]) for i, j in zip(ti, tj)])
    want_as = np.array([so.asnorm_score(rn[i][None], rn[j][None], cohort, 300) for i, j in zip(ti, tj)])
    _, sigma = so.asnorm_stats(rn, cohort, 300)
    tol = 1e-4 if precision == "fp32" else 5e-4
    dc, da = np.abs(got_cos - want_cos).max(), np.abs(got_as - want_as).max()
    print("ragged eval", precision, "cosine dev", dc, "AS-norm dev", da, "sigma_min", sigma.min(), "bound", 2 * tol / sigma.min())
    assert dc <= tol, dc
    assert da <= 2 * tol / sigma.min(), (da, sigma.min())
