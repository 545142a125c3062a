"""One process driving two GPUs (an engine per device): kernel attributes (dynamic shared memory opt-in), TMA maps and the
library's own staging buffers are per device.  Skipped on a single-GPU box; run with `gpurun --gpus 2`."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

L = 48000
CH512 = [512, 512, 512, 512, 1536]


def _need_two():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one process")


def test_ecapa_and_scoring_on_a_second_device():
    _need_two()
    import os
    from oracle import synth
    from speakerverification_b200 import _native as nv, utils as U
    from speakerverification_b200.engine import EcapaEngine, Frontend
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "cfg1_c512.npz"))
    bn = {k[3:]: v for k, v in g.items() if k.startswith("bn/")}
    sd = synth.ecapa_state_dict(CH512, 192, bn_stats=bn)
    waves, _ = synth.utterance_set(3, 2, L)
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        state = {k[len("__S__."):]: torch.as_tensor(v).to(dev) for k, v in sd.items()}
        fe = Frontend(sr=16000, n_fft=512, win_length=400, hop_length=160, n_mels=80)
        for precision in ("fp32", "f16"):
            eng = EcapaEngine(state, CH512, 80, 192, fe, precision=precision)
            assert str(eng.device) == dev
            outs.append(eng.forward(torch.from_numpy(waves).to(dev)).cpu())
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    assert torch.equal(outs[0], outs[2]) and torch.equal(outs[1], outs[3])     # same kernels, same order of sums
    # scoring: both AS-norm paths (store-and-select, fused pre-filter) and the static-scratch entry point on cuda:1
    e, _ = synth.scoring_embeddings(2048, 100, seed=5)
    c_small, _ = synth.scoring_embeddings(2000, 200, seed=105)
    c_big, _ = synth.scoring_embeddings(9000, 500, seed=106)
    for c, top in ((c_small, 300), (c_big, 300)):
        ref = U.asnorm_stats(torch.from_numpy(e).to("cuda:0"), torch.from_numpy(c).to("cuda:0"), top)
        got = U.asnorm_stats(torch.from_numpy(e).to("cuda:1"), torch.from_numpy(c).to("cuda:1"), top)
        assert got[0].device.index == 1
        assert torch.equal(ref[0].cpu(), got[0].cpu()) and torch.equal(ref[1].cpu(), got[1].cpu())
    with nv.on_device("cuda:1"):
        emb, coh = torch.from_numpy(e[:512]).to("cuda:1"), torch.from_numpy(c_small).to("cuda:1")
        out = torch.empty((2, 512), device="cuda:1")
        nv.check(nv.lib().svb_asnorm_stats_static(nv.ptr(emb), 512, nv.ptr(coh), 2000, 192, 300, nv.ptr(out[0]),
                                                  nv.ptr(out[1]), nv.stream_ptr()))
        torch.cuda.synchronize(1)
    ref = U.asnorm_stats(torch.from_numpy(e[:512]).to("cuda:0"), torch.from_numpy(c_small).to("cuda:0"), 300)
    assert torch.equal(out[0].cpu(), ref[0].cpu()) and torch.equal(out[1].cpu(), ref[1].cpu())


def test_rawnet2_on_a_second_device():
    _need_two()
    from speakerverification_b200 import synth
    from speakerverification_b200.rawnet2_custom import RawNet2Engine
    sd = synth.rawnet2_state_dict()
    wav = 0.1 * torch.randn(3, 64000)
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        state = {k[len("__S__."):]: torch.as_tensor(v).to(dev) for k, v in sd.items()}
        for precision in ("fp32", "f16"):
            eng = RawNet2Engine(state, 64000, 512, precision=precision)
            with torch.cuda.device(dev):
                outs.append(eng.forward(wav.to(dev)).cpu())
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    assert torch.equal(outs[0], outs[2]) and torch.equal(outs[1], outs[3])


def test_product_api_with_gpu_1(tmp_path, golden_c512):
    """The reference's own switch: gpu=N selects cuda:N (src/model.py:55-56,171-172) while cuda:0 stays torch's
    current device -- list evaluation (cosine and AS-norm) gives the scores of the gpu=0 run."""
    _need_two()
    import test_api_gpu as api
    from speakerverification_b200.model import ModelHandling, SpeakerEncoder, WrappedModel
    wd = api._make_workdir(tmp_path, golden_c512, api.CH512)
    scores = []
    for gpu in (0, 1):
        args = dict(api._args(wd["tmp"], "f16", wd["channels"]), gpu=gpu)
        net = WrappedModel(SpeakerEncoder(**args))
        h = ModelHandling(net, **dict(args, T_max=10))
        h.loadParameters(wd["ckpt"], show_error=False)
        net.eval()
        assert torch.cuda.current_device() == 0
        for mode in ("cosine", "norm"):
            sc, lab, _ = h.evaluateFromList(wd["trials"], distributed=False, dataloader_options=args["dataloader_options"],
                                            cohorts_path=wd["cohorts"], num_eval=2, scoring_mode=mode, top=300)
            scores.append(np.asarray(sc))
    assert np.array_equal(scores[0], scores[2]) and np.array_equal(scores[1], scores[3])
