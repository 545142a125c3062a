"""Size-independent properties at BASELINE.json's full embedding workload (configs[1]: C=1024, bf16, 10 000 x 3 s):
the embedding of an utterance does not depend on where it sits in the 10 000-utterance pass, on the launch size, or on
the run; every embedding is finite."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_embedding_pass_10k_position_and_launch_invariance():
    import bench
    from speakerverification_b200 import synth
    from speakerverification_b200.engine import EcapaEngine, Frontend
    CH = [1024, 1024, 1024, 1024, 3072]
    sd = synth.ecapa_state_dict(CH, 192)
    state = {k[len("__S__."):]: torch.as_tensor(v).cuda() for k, v in sd.items()}
    fe = Frontend(sr=16000, n_fft=512, win_length=400, hop_length=160, n_mels=80)
    eng = EcapaEngine(state, CH, 80, 192, fe, precision="bf16")
    U = 10000
    waves = bench.synth_waves(U, torch.device("cuda"), seed=100)
    probes = [0, 591, 592, 5000, 9999]
    for i in probes[1:]:
        waves[i] = waves[0]                                   # the same utterance at five positions of the pass
    emb = torch.empty((U, 192), device="cuda")
    for r0 in range(0, U, 592):
        eng.forward(waves[r0:r0 + 592], out=emb[r0:r0 + 592])
    torch.cuda.synchronize()
    assert bool(torch.isfinite(emb).all())
    e0 = emb[0]
    for i in probes[1:]:
        cos = float(torch.nn.functional.cosine_similarity(e0, emb[i], dim=0))
        assert cos >= 1.0 - 1e-6, (i, cos)
    # a second pass reproduces the first bit for bit (fixed-order / integer reductions only)
    emb2 = torch.empty_like(emb[:1184])
    for r0 in range(0, 1184, 592):
        eng.forward(waves[r0:r0 + 592], out=emb2[r0:r0 + 592])
    torch.cuda.synchronize()
    assert torch.equal(emb2, emb[:1184])
    # launch-size invariance: the same utterances in launches of 7 and of 256
    small = torch.cat([eng.forward(waves[r0:r0 + 7]).clone() for r0 in range(0, 21, 7)])
    mid = eng.forward(waves[:256]).clone()
    torch.cuda.synchronize()
    c1 = torch.nn.functional.cosine_similarity(small, emb[:21], dim=1).min().item()
    c2 = torch.nn.functional.cosine_similarity(mid, emb[:256], dim=1).min().item()
    assert c1 >= 1.0 - 1e-5 and c2 >= 1.0 - 1e-5, (c1, c2)
    # distinct utterances stay distinct (no row aliasing)
    assert float((emb[1] - emb[2]).abs().max()) > 0
