"""Host-side profile of ModelHandling.embed_signals on the cfg3 workload (cProfile, top cumulative entries)."""
import cProfile, os, pstats, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from speakerverification_b200 import synth
from speakerverification_b200.model import ModelHandling, SpeakerEncoder, WrappedModel
margs = bench.model_args("cuda", "bf16")
net = WrappedModel(SpeakerEncoder(**margs))
h = ModelHandling(net, **dict(margs, T_max=10, svb_batch_rows=592, svb_batch_samples=296 * 48000))
net.eval()
lengths = synth.ragged_lengths(2000, seed=2)
signals = bench.eval_signals(lengths, seed=50)
h.embed_signals(signals[:64], 0); torch.cuda.synchronize()
for _ in range(2):
    t0 = time.perf_counter(); h.embed_signals(signals, 0); torch.cuda.synchronize(); print("embed_signals %.1f ms" % (1e3 * (time.perf_counter() - t0)))
pr = cProfile.Profile(); pr.enable(); h.embed_signals(signals, 0); torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
