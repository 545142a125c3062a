"""Where the scoring step of bench.py goes: svb_asnorm_stats and svb_asnorm_trials timed separately on the bench's own
inputs (100 000 utterances, 20 000-vector cohort, 1 M trials, top-300)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from speakerverification_b200 import utils as U

e, c, ia, ib = bench.scoring_inputs(100000, 20000, 1000000, seed=3)
emb, coh = torch.from_numpy(e).cuda(), torch.from_numpy(c).cuda()
tia, tib = torch.from_numpy(ia).cuda(), torch.from_numpy(ib).cuda()


def timed(fn, n=10):
    for _ in range(3):
        r = fn()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(n):
        r = fn()
    t1.record(); torch.cuda.synchronize()
    return t0.elapsed_time(t1) / n, r


ms_stats, (mu, sg) = timed(lambda: U.asnorm_stats(emb, coh, 300))
ms_trials, _ = timed(lambda: U.asnorm_trials(emb, tia, tib, mu, sg))
ms_both, _ = timed(lambda: U.asnorm_trials(emb, tia, tib, *U.asnorm_stats(emb, coh, 300)))
print("asnorm_stats %.3f ms  asnorm_trials %.3f ms  both %.3f ms" % (ms_stats, ms_trials, ms_both))
