"""Stage-wise comparison of the RawNet2 engine (fp32-grade mode) against the CPU oracle at a non-default crop length."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import rawnet_oracle as ro, synth
from speakerverification_b200.rawnet2_custom import RawNet2Engine
n_samples = int(sys.argv[1]) if len(sys.argv) > 1 else 48000
g = np.load("tests/golden/cfg4_rawnet2.npz")
bn = {k[3:]: v for k, v in g.items() if k.startswith("bn/")}
sd = synth.rawnet2_state_dict(n_samples=n_samples, bn_stats=bn)
waves, _ = synth.utterance_set(3, 2, n_samples); waves = waves[:2]
if len(sys.argv) > 2: waves = waves[[1, 1]]      # the second utterance twice
torch.set_num_threads(os.cpu_count())
st = {}
with torch.no_grad():
    want = ro.rawnet2_forward(waves, sd, stages=st).numpy()
state = {k[len("__S__."):]: torch.as_tensor(v).cuda() for k, v in sd.items()}
tsd = {k: torch.as_tensor(v) for k, v in sd.items()}
eng = RawNet2Engine(state, n_samples, 512, precision="fp32")
w = torch.from_numpy(waves).cuda()
names = ["layer1.0", "layer2.0", "layer3.0", "layer4.0", "layer4.1", "layer5.0", "layer6.0", "layer6.1"]
down = [1, 1, 1, 0, 1, 1, 0, 1]
NEXT_BN = ["layer2.0.bn1.", "layer3.0.bn1.", "layer4.0.bn1.", "layer4.1.bn1.", "layer5.0.bn1.", "layer6.0.bn1.", "layer6.1.bn1.", "bn_before_gru."]
COUT = [128, 128, 256, 256, 256, 512, 512, 512]
os.environ["SVB_RAWNET_STOP"] = "-1"
eng.forward(w); torch.cuda.synchronize()
front = st["front"].permute(0, 2, 1).numpy()
got = eng.stage(2, "x0", 0, 128).cpu().numpy()
print("front", got.shape, front.shape, np.abs(got - front).max() / np.abs(front).max())
x = st["front"]
for b in range(8):
    os.environ["SVB_RAWNET_STOP"] = str(b)
    eng.forward(w); torch.cuda.synchronize()
    with torch.no_grad():
        x = ro.basic_block(x, tsd, "__S__." + names[b] + ".", bool(down[b]))
        pre = torch.nn.functional.leaky_relu(ro._bn(x, tsd, "__S__." + NEXT_BN[b]), 0.3).permute(0, 2, 1).numpy()
    got = eng.stage(2, "a1" if b % 2 == 0 else "a0", b + 1, COUT[b]).cpu().numpy()
    d = np.abs(got - pre)
    print("block", b, got.shape, pre.shape, "err/scale %.2e" % (d.max() / np.abs(pre).max()), "argmax", np.unravel_index(d.argmax(), d.shape))
    if b == 5:
        pf = d.max(2) / np.abs(pre).max()
        print("  per-frame err utt0:", np.array2string(pf[0], precision=1, max_line_width=250))
        print("  per-frame err utt1:", " ".join("%.0e" % v for v in pf[1]))
        pc = d.max(1) / np.abs(pre).max()
        print("  per-channel err > 1e-4 (utt1):", np.flatnonzero(pc[1] > 1e-4)[:40], "count", (pc[1] > 1e-4).sum(), "max per ch", " ".join("%.0e" % v for v in np.sort(pc[1])[-8:]))
del os.environ["SVB_RAWNET_STOP"]
got = eng.forward(w).cpu().numpy()
print("emb cos", (got * want).sum(1) / np.linalg.norm(got, axis=1) / np.linalg.norm(want, axis=1))
