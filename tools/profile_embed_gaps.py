"""GPU-side timeline of ModelHandling.embed_signals on the cfg3 workload: per-launch busy time and the idle gaps between
consecutive launches on the compute stream (CUDA events recorded around svb_pcm_stage ... forward_rows)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from speakerverification_b200 import synth, engine as E
from speakerverification_b200.model import ModelHandling, SpeakerEncoder, WrappedModel
margs = bench.model_args("cuda", "bf16")
net = WrappedModel(SpeakerEncoder(**margs))
h = ModelHandling(net, **dict(margs, T_max=10))
net.eval()
signals = bench.eval_signals(synth.ragged_lengths(2000, seed=2), seed=50)
h.embed_signals(signals[:64], 0); torch.cuda.synchronize()
h.embed_signals(signals, 0); torch.cuda.synchronize()
ev = []
orig = E.EcapaEngine.forward_rows
def wrapped(self, *a, **k):
    r = orig(self, *a, **k)
    e = torch.cuda.Event(enable_timing=True); e.record(); ev.append(e)
    return r
E.EcapaEngine.forward_rows = wrapped
h.stage_trace = []
s0 = torch.cuda.Event(enable_timing=True); s0.record()
t0 = time.perf_counter(); h.embed_signals(signals, 0); torch.cuda.synchronize(); wall = 1e3 * (time.perf_counter() - t0)
ends = [s0.elapsed_time(e) for e in ev]
d = np.diff([0.0] + ends)
import io
out = open("gpurun_out/embed_gaps.txt", "w")
print("wall %.1f ms, launches %d, end-to-end on stream %.1f ms" % (wall, len(ev), ends[-1]), file=out)
print("per-launch interval ms:", np.round(d, 2).tolist(), file=out)
print("stage (group, wait, fill, enqueue) ms:", [tuple(round(x, 2) for x in t) for t in h.stage_trace], file=out)
out.close()
