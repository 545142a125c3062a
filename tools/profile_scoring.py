"""Two AS-norm statistics passes (U x 20k cohort, top-300) for ncu launch lists."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speakerverification_b200 import synth
from speakerverification_b200 import utils as SU
U = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
e, _ = synth.scoring_embeddings(U, 2000)
c, _ = synth.scoring_embeddings(N, 2000, seed=103)
e, c = torch.from_numpy(e).cuda(), torch.from_numpy(c).cuda()
for _ in range(2):
    mu, sg = SU.asnorm_stats(e, c, 300)
torch.cuda.synchronize()
print("ok", float(mu.mean()), float(sg.mean()))
