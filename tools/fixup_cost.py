"""Cost of the exact fix-up of the fused AS-norm path: cfg5 geometry (100 000 x 20 000, top-300) with 0, 1, 8 and 400
rows forced to the fix-up (all-zero embeddings: every score ties).  Prints ms per svb_asnorm_stats call."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import synth
from speakerverification_b200 import utils as U

e, _ = synth.scoring_embeddings(100000, 1000, seed=5)
c, _ = synth.scoring_embeddings(20000, 2000, seed=105)
coh = torch.from_numpy(c).cuda()
for forced in (0, 1, 8, 400):
    x = e.copy()
    x[np.linspace(0, 99999, forced).astype(int)] = 0.0 if forced else x[0]
    emb = torch.from_numpy(x).cuda()
    for _ in range(2):
        mu, sg, nfix = U.asnorm_stats(emb, coh, 300, return_fixups=True)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(5):
        U.asnorm_stats(emb, coh, 300)
    t1.record(); torch.cuda.synchronize()
    print("forced %4d  fix-up rows %4d  %.3f ms per call" % (forced, nfix, t0.elapsed_time(t1) / 5))
