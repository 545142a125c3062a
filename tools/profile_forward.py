"""Two encoder launches of the bench configuration (C=1024, bf16, 256 x 3 s) for ncu launch lists."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speakerverification_b200 import synth
from speakerverification_b200.engine import EcapaEngine, Frontend

CH = [1024, 1024, 1024, 1024, 3072]
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
prec = sys.argv[2] if len(sys.argv) > 2 else "bf16"
n_fwd = int(sys.argv[3]) if len(sys.argv) > 3 else 2
sd = synth.ecapa_state_dict(CH, 192)
state = {k[len("__S__."):]: torch.as_tensor(v).cuda() for k, v in sd.items()}
fe = Frontend(sr=16000, n_fft=512, win_length=400, hop_length=160, n_mels=80)
eng = EcapaEngine(state, CH, 80, 192, fe, precision=prec)
wav = 0.1 * torch.randn(B, 48000, device="cuda")
torch.cuda.synchronize()
for _ in range(n_fwd):
    out = eng.forward(wav)
torch.cuda.synchronize()
print("ok", float(out.abs().mean()))
