"""Encoder launches of the RawNet2 configuration (cfg4: 4 s crops) for ncu launch lists / quick timing."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speakerverification_b200 import synth
from speakerverification_b200.rawnet2_custom import RawNet2Engine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
prec = sys.argv[2] if len(sys.argv) > 2 else "bf16"
n_fwd = int(sys.argv[3]) if len(sys.argv) > 3 else 2
sd = synth.rawnet2_state_dict()
state = {k[len("__S__."):]: torch.as_tensor(v).cuda() for k, v in sd.items()}
eng = RawNet2Engine(state, 64000, 512, precision=prec)
wav = 0.1 * torch.randn(B, 64000, device="cuda")
torch.cuda.synchronize()
for _ in range(n_fwd):
    out = eng.forward(wav)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n_fwd):
    out = eng.forward(wav)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n_fwd
print("ok", float(out.abs().mean()), "ms/forward %.3f  utt/s %.0f  ws GB %.2f" % (ms, B / ms * 1e3, eng._ws.numel() / 1e9))
