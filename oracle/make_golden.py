"""Generate tests/golden/*.npz by running the UNMODIFIED reference on CPU.

Run in the build container only (needs /root/reference):
    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden
Inputs (waveforms, weights, trials) are regenerated from seeds by oracle/synth.py;
only BN running statistics and reference OUTPUTS are stored.
"""
import os
import sys

import numpy as np
import torch

from . import ref_loader, synth

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
L = 48000


def _load_weights(net, sd):
    tgt = net.module.state_dict()
    for k, v in sd.items():
        assert k in tgt, k
        tgt[k].copy_(torch.as_tensor(v))


def _calibrate_bn(net, waves, batch=32):
    """cumulative-average BN statistics from one train-mode pass."""
    bns = [m for m in net.modules() if isinstance(m, torch.nn.BatchNorm1d)]
    for m in bns:
        m.reset_running_stats()
        m.momentum = None
    net.train()
    with torch.no_grad():
        for i in range(0, len(waves), batch):
            net(torch.from_numpy(waves[i:i + batch]))
    net.eval()
    out = {}
    for k, v in net.module.state_dict().items():
        if k.startswith("__S__.") and (k.endswith("running_mean") or k.endswith("running_var")):
            out[k] = v.detach().numpy().copy()
    return out


def _stage_hooks(net, store):
    S = net.module.__S__
    hs = [net.module.compute_features.register_forward_hook(lambda m, i, o: store.__setitem__("mel", o.detach()))]
    for i in range(4):
        hs.append(S.blocks[i].register_forward_hook(
            lambda m, inp, o, i=i: store.__setitem__("block%d" % i, o.detach())))
    hs.append(S.mfa.register_forward_hook(lambda m, i, o: store.__setitem__("mfa", o.detach())))
    hs.append(S.asp.register_forward_hook(lambda m, i, o: store.__setitem__("asp", o.detach())))
    return hs


def make(channels, n_out, tag, n_spk=16, per_spk=4, n_trials=1000, n_cohort=300, full=True):
    ref_model, ref_utils = ref_loader.load()
    net, handle, args = ref_loader.build(channels, n_out)
    sd = synth.ecapa_state_dict(channels, n_out, seed=1234)
    _load_weights(net, sd)
    waves, spk = synth.utterance_set(n_spk, per_spk, L + 1)
    bn = _calibrate_bn(net, waves[:, :L])
    out = {"bn/" + k: v for k, v in bn.items()}

    # direct batch forward + per-stage dumps for utterance 0 (every 16th channel)
    store = {}
    hooks = _stage_hooks(net, store)
    with torch.no_grad():
        emb = net(torch.from_numpy(waves[:, :L]))
    for h in hooks:
        h.remove()
    out["emb_batch"] = emb.numpy()
    out["stage/mel"] = store["mel"][0].numpy()
    for k in ["block0", "block1", "block2", "block3", "mfa"]:
        out["stage/" + k] = store[k][0, ::16].numpy()
    out["stage/asp"] = store["asp"][0, :, 0].numpy()

    if full:
        # list API (src/model.py:306-450), cosine, num_eval=2
        names = ["utt%03d.wav" % i for i in range(len(waves))]
        for n, w in zip(names, waves):
            ref_loader.WAVS[n] = w
        lab, ti, tj = synth.trial_list(spk, n_trials)
        trial_txt = "/tmp/svb_trials_%s.txt" % tag
        with open(trial_txt, "w") as f:
            for a, i, j in zip(lab, ti, tj):
                f.write("%d %s %s\n" % (a, names[i], names[j]))
        sc, labs, _ = handle.evaluateFromList(trial_txt, distributed=False,
                                              dataloader_options=args["dataloader_options"],
                                              cohorts_path="unused.npy", num_eval=2, scoring_mode="cosine")
        out["scores_cos"] = np.asarray(sc, np.float64)
        assert list(labs) == list(lab)
        emb_list = np.stack([handle.embed_utterance(n, num_eval=2, normalize=False).numpy() for n in names])
        out["emb_list"] = emb_list                                            # [U, 2, D]

        # cohort bank as prepare('cohorts') builds it (src/model.py:598-606), one file per speaker
        cf0, cenv = synth.speaker_bank(n_cohort, seed=77)
        cohort = []
        for s in range(n_cohort):
            ref_loader.WAVS["coh%04d.wav" % s] = synth.utterance(cf0[s], cenv[s], L + 1, 50000 + s)
            e = handle.embed_utterance("coh%04d.wav" % s, num_eval=2, normalize=True)
            cohort.append(np.mean(e.numpy(), axis=0, keepdims=True))
        cohort = np.vstack(cohort)
        out["cohort"] = cohort
        en = torch.nn.functional.normalize(torch.from_numpy(emb_list), p=2, dim=2)
        for K in (200, 300):
            s = [ref_utils.ZT_norm_similarity(en[i], en[j], cohort, top=K) for i, j in zip(ti, tj)]
            out["scores_asnorm%d" % K] = np.asarray(s, np.float64)

        # metrics (src/utils.py:74-121, 221-275) on both score sets
        for name in ("scores_cos", "scores_asnorm300"):
            res = ref_utils.tuneThresholdfromScore(out[name], lab, [1, 0.1])
            fnrs, fprs, thr = ref_utils.ComputeErrorRates(out[name], lab)
            dcf, _ = ref_utils.ComputeMinDcf(fnrs, fprs, thr, 0.05, 1, 1)
            out["eer/" + name] = np.float64(res["roc"][1])
            out["mindcf/" + name] = np.float64(dcf)
            print(tag, name, "EER %.4f minDCF %.4f" % (res["roc"][1], dcf))
    os.makedirs(GOLD, exist_ok=True)
    path = os.path.join(GOLD, "%s.npz" % tag)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


def make_nogc(tag="cfg1_c512_nogc", channels=(512, 512, 512, 512, 1536), n_out=192):
    """ECAPA-TDNN with global_context=False (src/models/ECAPA_TDNN.py:204-207,238-248: the attention of the pooling
    layer sees x only; asp.tdnn has 3C instead of 9C inputs)."""
    ref_loader.load()
    net, handle, args = ref_loader.build(list(channels), n_out, global_context=False)
    sd = synth.ecapa_state_dict(list(channels), n_out, seed=1234)
    C3 = channels[-1]
    sd["__S__.asp.tdnn.conv.conv.weight"] = np.ascontiguousarray(sd["__S__.asp.tdnn.conv.conv.weight"][:, :C3]) * np.float32(1.7)
    _load_weights(net, sd)
    waves, spk = synth.utterance_set(16, 4, L + 1)
    bn = _calibrate_bn(net, waves[:, :L])
    out = {"bn/" + k: v for k, v in bn.items()}
    with torch.no_grad():
        emb = net(torch.from_numpy(waves[:, :L]))
    out["emb_batch"] = emb.numpy()
    _, ref_utils = ref_loader.load()
    lab, ti, tj = synth.trial_list(spk, 1000)
    out["scores_cos"] = np.asarray([float(ref_utils.cosine_similarity(emb[i:i + 1], emb[j:j + 1])) for i, j in zip(ti, tj)])
    path = os.path.join(GOLD, "%s.npz" % tag)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


def make_inorm(tag="cfg1_c512_inorm", channels=(512, 512, 512, 512, 1536), n_out=192):
    """ECAPA-TDNN with input_norm=True (nn.InstanceNorm1d(n_mels, affine=True) after the log / mean normalisation,
    src/models/ECAPA_TDNN.py:406-408,477-478) -- the configuration the Raw3_ECAPA hybrid builds its ECAPA branch
    with (src/models/Raw3_ECAPA.py:21-24).  The affine parameters are seeded non-trivial values."""
    ref_loader.load()
    net, handle, args = ref_loader.build(list(channels), n_out, input_norm=True)
    sd = synth.ecapa_state_dict(list(channels), n_out, seed=1234)
    rng = np.random.default_rng(4321)
    sd["__S__.instance_norm.weight"] = (1.0 + 0.3 * rng.standard_normal(80)).astype(np.float32)
    sd["__S__.instance_norm.bias"] = (0.2 * rng.standard_normal(80)).astype(np.float32)
    _load_weights(net, sd)
    waves, spk = synth.utterance_set(16, 4, L + 1)
    bn = _calibrate_bn(net, waves[:, :L])
    out = {"bn/" + k: v for k, v in bn.items()}
    out["instance_norm.weight"], out["instance_norm.bias"] = sd["__S__.instance_norm.weight"], sd["__S__.instance_norm.bias"]
    store = {}
    hooks = _stage_hooks(net, store)
    with torch.no_grad():
        emb = net(torch.from_numpy(waves[:, :L]))
    for h in hooks:
        h.remove()
    out["emb_batch"] = emb.numpy()
    out["stage/block0"] = store["block0"][0, ::16].numpy()
    _, ref_utils = ref_loader.load()
    lab, ti, tj = synth.trial_list(spk, 1000)
    out["scores_cos"] = np.asarray([float(ref_utils.cosine_similarity(emb[i:i + 1], emb[j:j + 1])) for i, j in zip(ti, tj)])
    path = os.path.join(GOLD, "%s.npz" % tag)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


def make_rawnet(tag="cfg4_rawnet2", n_spk=4, per_spk=4, n_out=512, L4=64000):
    """RawNet2_custom (sinc + GRU, 4 s crops): embeddings through the reference's
    SpeakerEncoder(features='raw') and strided per-stage dumps for utterance 0."""
    ref_loader.load()
    net, handle, args = ref_loader.build([512] * 4 + [1536], n_out, sentence_len=4.0, features="raw",
                                         model_name="RawNet2_custom")
    sd = synth.rawnet2_state_dict(n_out, L4)
    _load_weights(net, sd)
    waves, spk = synth.utterance_set(n_spk, per_spk, L4)
    bn = _calibrate_bn(net, waves, batch=8)
    out = {"bn/" + k: v for k, v in bn.items()}
    S = net.module.__S__
    store = {}
    hs = [S.first_bn.register_forward_hook(lambda m, i, o: store.__setitem__("front", torch.nn.functional.leaky_relu(o.detach(), 0.3)))]
    for li in range(6):
        hs.append(getattr(S, "layer%d" % (li + 1)).register_forward_hook(
            lambda m, i, o, li=li: store.__setitem__("layer%d" % (li + 1), o.detach())))
    hs.append(S.gru.register_forward_hook(lambda m, i, o: store.__setitem__("gru", o[0][:, -1, :].detach())))
    with torch.no_grad():
        emb = net(torch.from_numpy(waves))
    for h in hs:
        h.remove()
    out["emb_batch"] = emb.numpy()
    out["stage/front"] = store["front"][0, ::8, ::25].numpy()
    for li in range(6):
        k = "layer%d" % (li + 1)
        step = max(1, store[k].shape[2] // 400)
        out["stage/" + k] = store[k][0, ::8, ::step].numpy()
    out["stage/gru"] = store["gru"].numpy()
    # cosine trial scores over all pairs (src/utils.py:163-164, n = 1 crop per utterance)
    _, ref_utils = ref_loader.load()
    lab, ti, tj = synth.trial_list(spk, 120)
    out["scores_cos"] = np.asarray([float(ref_utils.cosine_similarity(emb[i:i + 1], emb[j:j + 1])) for i, j in zip(ti, tj)])
    path = os.path.join(GOLD, "%s.npz" % tag)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


def make_rawnet_heads(tag="cfg4_rawnet2_heads", n_out=512, L4=64000):
    """RawNet2_custom with the attentive heads aggregate='sap' and aggregate='asp' (src/models/RawNet2_custom.py:97-112,
    200-224): embeddings of 16 synthetic 4 s utterances through the reference's SpeakerEncoder(features='raw')."""
    ref_loader.load()
    out = {}
    waves, spk = synth.utterance_set(4, 4, L4)
    _, ref_utils = ref_loader.load()
    lab, ti, tj = synth.trial_list(spk, 120)
    for agg in ("sap", "asp"):
        net, handle, args = ref_loader.build([512] * 4 + [1536], n_out, sentence_len=4.0, features="raw",
                                             model_name="RawNet2_custom", aggregate=agg)
        sd = synth.rawnet2_state_dict(n_out, L4, aggregate=agg)
        _load_weights(net, sd)
        bn = _calibrate_bn(net, waves, batch=8)
        for k, v in bn.items():
            out["%s/bn/%s" % (agg, k)] = v
        with torch.no_grad():
            emb = net(torch.from_numpy(waves))
        out["%s/emb_batch" % agg] = emb.numpy()
        out["%s/scores_cos" % agg] = np.asarray([float(ref_utils.cosine_similarity(emb[i:i + 1], emb[j:j + 1]))
                                                 for i, j in zip(ti, tj)])
        print(agg, "emb", emb.shape, "scores", out["%s/scores_cos" % agg][:4])
    path = os.path.join(GOLD, "%s.npz" % tag)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count())
    which = sys.argv[1:] or ["cfg1", "cfg2", "cfg4"]
    if "cfg1" in which:
        make([512, 512, 512, 512, 1536], 192, "cfg1_c512")
    if "cfg2" in which:
        make([1024, 1024, 1024, 1024, 3072], 192, "cfg2_c1024")
    if "cfg4" in which:
        make_rawnet()
    if "cfg1i" in which:
        make_inorm()
    if "cfg1g" in which:
        make_nogc()
    if "cfg4h" in which:
        make_rawnet_heads()
