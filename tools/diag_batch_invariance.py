"""bf16-mode batch-composition invariance: the same ragged utterances embedded in two different launch groupings."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speakerverification_b200 import synth
from speakerverification_b200.engine import EcapaEngine, Frontend
CH = [1024, 1024, 1024, 1024, 3072]
gold = dict(np.load("tests/golden/cfg2_c1024.npz"))
bn = {k[3:]: v for k, v in gold.items() if k.startswith("bn/")}
lens = synth.ragged_lengths(48, seed=5)
pool, _ = synth.utterance_set(8, 2, 320000)
sig = [pool[i % 16][:n].copy() for i, n in enumerate(lens)]
fe = Frontend(sr=16000, n_fft=512, win_length=400, hop_length=160, n_mels=80)
for tag, stats in (("uncalibrated", None), ("calibrated", bn)):
    sd = synth.ecapa_state_dict(CH, 192, bn_stats=stats)
    state = {k[len("__S__."):]: torch.as_tensor(v).cuda() for k, v in sd.items()}
    for prec in ("bf16", "fp32"):
        eng = EcapaEngine(state, CH, 80, 192, fe, precision=prec)
        def run(groups):
            out = []
            for g in groups:
                flat = torch.from_numpy(np.concatenate([sig[i] for i in g])).cuda()
                out.append(eng.forward_flat(flat, [len(sig[i]) for i in g]).clone())
            return torch.cat(out)
        a = run([list(range(0, 48))])
        a2 = run([list(range(0, 48))])
        b = run([list(range(0, 7)), list(range(7, 30)), list(range(30, 48))])
        c = run([[i] for i in range(48)])
        cos = lambda x, y: float(torch.nn.functional.cosine_similarity(x, y, dim=1).min())
        xc = torch.nn.functional.normalize(a, dim=1)
        cross = float((xc @ xc.T).min())
        print(tag, prec, "same-grouping maxdiff %.3g" % float((a - a2).abs().max()), "regroup cos_min %.7f" % cos(a, b),
              "single cos_min %.7f" % cos(a, c), "| min cos between different utterances %.4f" % cross)
