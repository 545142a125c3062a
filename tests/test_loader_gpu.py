"""GPU crop/stack stage (svb_pcm_stage + row-offset crops) against the reference's loadWAV crops
(tests/golden/loader.npz) -- bit exact -- and the embedding pipeline built on it against the host-crop path."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "loader.npz"))


def _stage(signals, dst_len, kind):
    from speakerverification_b200 import _native as nv
    seg_dt = np.dtype([("src_off", "<i8"), ("dst_off", "<i8"), ("src_len", "<i4"), ("dst_len", "<i4")])
    segs = np.zeros(len(signals), dtype=seg_dt)
    segs["src_len"] = [len(s) for s in signals]
    segs["dst_len"] = dst_len
    segs["src_off"] = np.concatenate([[0], np.cumsum(segs["src_len"])[:-1]])
    segs["dst_off"] = np.concatenate([[0], np.cumsum(segs["dst_len"])[:-1]])
    pcm = torch.from_numpy(np.concatenate(signals)).cuda()
    dseg = torch.from_numpy(segs.view(np.uint8).copy()).cuda()
    out = torch.full((int(segs["dst_len"].sum()),), float("nan"), device="cuda")
    nv.check(nv.lib().svb_pcm_stage(nv.ptr(pcm), kind, nv.ptr(dseg), len(signals), int(max(dst_len)), nv.ptr(out),
                                    nv.stream_ptr()))
    torch.cuda.synchronize()
    return out.cpu().numpy(), segs


@pytest.mark.parametrize("kind", [1, 0])
def test_pcm_stage_bit_exact_vs_reference_loader(kind):
    from oracle.make_golden_loader import NUM_EVAL, SPEC, signals
    from speakerverification_b200.audio_loader import crop_plan
    sigs = signals()
    if kind == 0:
        sigs = [s.astype(np.float32) / np.float32(32768.0) for s in sigs]
    for n in NUM_EVAL:
        dst_len, starts, L = crop_plan([len(s) for s in sigs], SPEC, n)
        buf, segs = _stage(sigs, dst_len, kind)
        assert not np.isnan(buf).any()
        for i in range(len(sigs)):
            base = int(segs["dst_off"][i])
            want = GOLD["crops_%d_%d" % (i, n)]
            for j, st in enumerate(starts[i]):
                ln = int(dst_len[i]) if n == 0 else L
                assert np.array_equal(buf[base + int(st):base + int(st) + ln], want[j]), (kind, n, i, j)


def test_embed_signals_matches_host_crop_path(golden_c512):
    """int16 utterances of ragged length (shorter than, equal to and longer than the crop) through the GPU crop
    pipeline == crops cut on the host by the loadWAV mirror and embedded as plain [R, L] rows."""
    from oracle import synth
    from test_api_gpu import CH512, _args
    from speakerverification_b200.model import ModelHandling, SpeakerEncoder, WrappedModel
    args = _args("/tmp/svb_loader", "fp32", CH512)
    net = WrappedModel(SpeakerEncoder(**args))
    h = ModelHandling(net, **dict(args, T_max=10, svb_batch_rows=24))     # several launches: exercises the double buffering
    bn = {k[3:]: v for k, v in golden_c512.items() if k.startswith("bn/")}
    sd = net.module.state_dict()
    for k, v in synth.ecapa_state_dict(CH512, 192, bn_stats=bn).items():
        sd[k].copy_(torch.as_tensor(v))
    net.eval()
    waves, _ = synth.utterance_set(3, 4, 80000)
    lens = [20000, 47999, 48000, 48001, 48002, 60000, 80000, 31000, 52345, 48000, 79999, 70001]
    pcm = [np.round(w[:n] * 32767).astype(np.int16) for w, n in zip(waves, lens)]
    for num_eval in (5, 0):
        got = h.embed_signals(pcm, num_eval).cpu().numpy()
        for i, p in enumerate(pcm):
            audio = p.astype(np.float32) / np.float32(32768.0)
            # loadWAV(ndarray) would amplitude-normalise; cut the crops with the same rule on the raw samples instead
            L = 48000
            a = np.pad(audio, (0, L - len(audio) + 1), "wrap") if len(audio) <= L else audio
            if num_eval == 0:
                crops = a[None]
            else:
                crops = np.stack([a[int(s):int(s) + L] for s in np.linspace(0, len(a) - L, num=num_eval)])
            want = net.module.engine().forward(torch.from_numpy(np.ascontiguousarray(crops)).cuda()).cpu().numpy()
            assert got[i].shape == want.shape
            assert np.abs(got[i] - want).max() <= 1e-5 * max(1.0, np.abs(want).max()), (num_eval, i)


def test_embed_signals_back_to_back_calls():
    """Two embed_signals calls without a synchronisation in between (>= 4 launch groups each) reuse the persistent pinned
    / device staging buffers: the second call must wait for the copies and kernels of the first (the first
    groups of a call used to skip the ready / freed events of the previous call's last groups)."""
    import torch
    from oracle import synth
    from speakerverification_b200.model import ModelHandling, SpeakerEncoder, WrappedModel
    CH = [512, 512, 512, 512, 1536]
    args = dict(device="cuda", gpu=0, save_folder="/tmp/svb_t", model={"name": "ECAPA_TDNN", "nOut": 192}, channels=CH,
                criterion={"name": "AAmSoftmaxAP"}, features="melspectrogram", n_mels=80, sr=16000, n_fft=512,
                win_length=400, hop_length=160,
                audio_spec={"sample_rate": 16000, "sentence_len": 1.0, "win_len": 0.025, "hop_len": 0.01},
                dataloader_options={"nPerSpeaker": 2}, svb_precision="f16", svb_batch_rows=8)
    net = WrappedModel(SpeakerEncoder(**args))
    h = ModelHandling(net, **dict(args, T_max=10))
    tgt = net.module.state_dict()
    for k, v in synth.ecapa_state_dict(CH, 192).items():
        tgt[k].copy_(torch.as_tensor(v))
    rng = np.random.default_rng(5)
    sigs_a = [np.round(3000 * rng.standard_normal(int(n))).astype(np.int16) for n in rng.integers(16000, 40000, 20)]
    sigs_b = [np.round(3000 * rng.standard_normal(int(n))).astype(np.int16) for n in rng.integers(16000, 40000, 20)]
    want_a = h.embed_signals(sigs_a, 2).clone()
    torch.cuda.synchronize()
    want_b = h.embed_signals(sigs_b, 2).clone()
    torch.cuda.synchronize()
    got_a = h.embed_signals(sigs_a, 2)          # 5 groups of 4 utterances x 2 crops; no sync before the next call
    got_b = h.embed_signals(sigs_b, 2)
    got_a2 = h.embed_signals(sigs_a, 2)
    torch.cuda.synchronize()
    assert torch.equal(got_a, want_a) and torch.equal(got_b, want_b) and torch.equal(got_a2, want_a)
