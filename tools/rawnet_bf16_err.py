"""Per-stage relative error of the bf16 RawNet2 path against the CPU oracle (diagnostic)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import rawnet_oracle as ro, synth
from speakerverification_b200.rawnet2_custom import RawNet2Engine
g = dict(np.load("tests/golden/cfg4_rawnet2.npz"))
bn = {k[3:]: v for k, v in g.items() if k.startswith("bn/")}
sd = synth.rawnet2_state_dict(bn_stats=bn)
waves, _ = synth.utterance_set(4, 4, 64000)
tsd = {k: torch.as_tensor(v) for k, v in sd.items()}
st = {}
with torch.no_grad():
    emb = ro.rawnet2_forward(waves[:2], sd, stages=st)
state = {k[len("__S__."):]: torch.as_tensor(v).cuda() for k, v in sd.items()}
names = ["layer1.0", "layer2.0", "layer3.0", "layer4.0", "layer4.1", "layer5.0", "layer6.0", "layer6.1"]
down = [1, 1, 1, 0, 1, 1, 0, 1]
nbn = ["layer1.0.bn1."] + [n + ".bn1." for n in names[1:]] + ["bn_before_gru."]
cout = [128, 128, 128, 256, 256, 256, 512, 512, 512]
for prec in sys.argv[1:] or ["bf16"]:
    eng = RawNet2Engine(state, 64000, 512, precision=prec)
    x = st["front"]
    for blk in range(-1, 8):
        os.environ["SVB_RAWNET_STOP"] = str(blk)
        eng.forward(torch.from_numpy(waves[:2]).cuda()); torch.cuda.synchronize()
        if blk >= 0:
            with torch.no_grad():
                x = ro.basic_block(x, tsd, "__S__." + names[blk] + ".", bool(down[blk]))
        want = torch.nn.functional.leaky_relu(ro._bn(x, tsd, "__S__." + nbn[blk + 1]), 0.3).permute(0, 2, 1).numpy()
        buf = "a0" if (blk + 1) % 2 == 0 else "a1"
        got = eng.stage(2, buf, blk + 1, cout[blk + 1]).cpu().numpy()
        d = got - want
        print(prec, "after block %2d: rms rel %.2e  max rel %.2e" % (blk, np.sqrt((d ** 2).mean() / (want ** 2).mean()), np.abs(d).max() / np.abs(want).max()))
    del os.environ["SVB_RAWNET_STOP"]
    out = eng.forward(torch.from_numpy(waves[:2]).cuda()).cpu().numpy()
    e = emb.numpy()
    print(prec, "emb cos", (out * e).sum(1) / np.linalg.norm(out, axis=1) / np.linalg.norm(e, axis=1))
    off = eng.stage  # gru state
    import ctypes
    from speakerverification_b200 import _native as nv
    o = nv.lib().svb_rawnet2_ws_offset(eng.handle, 2, b"gru")
    h = eng._ws[o:o + 2 * 1024 * 4].view(torch.float32).view(2, 1024).cpu().numpy()
    hw = st["gru"].numpy()
    print(prec, "gru state rms rel %.2e" % np.sqrt(((h - hw) ** 2).mean() / (hw ** 2).mean()))
