"""Per-group host staging times of ModelHandling.embed_signals on the cfg3 workload (stage_trace: wait for the pinned
slot, parallel fill, enqueue) against the device time of the same pass."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from speakerverification_b200 import synth
from speakerverification_b200.model import ModelHandling, SpeakerEncoder, WrappedModel
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 296
margs = bench.model_args("cuda", "f16")
net = WrappedModel(SpeakerEncoder(**margs))
h = ModelHandling(net, **dict(margs, T_max=10, svb_batch_rows=4096, svb_batch_samples=rows * 48000))
net.eval()
lengths = synth.ragged_lengths(2000, seed=2)
signals = bench.eval_signals(lengths, seed=50)
h.embed_signals(signals[:64], 0); torch.cuda.synchronize()
for rep in range(3):
    h.stage_trace = []
    t0 = time.perf_counter(); h.embed_signals(signals, 0); torch.cuda.synchronize(); dt = 1e3 * (time.perf_counter() - t0)
    tr = np.array([t[1:] for t in sorted(h.stage_trace)])
    print("rows %d: embed_signals %.1f ms, groups %d | per group ms: wait %.2f fill %.2f enqueue %.2f | sums: wait %.1f fill %.1f enqueue %.1f"
          % (rows, dt, len(tr), *tr.mean(0), *tr.sum(0)))
