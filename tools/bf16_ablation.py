"""Which bf16 rounding point costs the score accuracy?  CPU numerical model of the bf16 data flow (tools/emulate_bf16.py)
with one rounding site at a time switched back to fp32, C = 1024, 64 utterances, all-pairs cosine scores against the
reference golden embeddings (tests/golden/cfg2_c1024.npz).  Design aid (next steps), not product.

    python tools/bf16_ablation.py            # ~ minutes on the container's CPU cores
"""
import sys, time
import numpy as np, torch, torch.nn.functional as F
sys.path.insert(0, '.')
from oracle import ecapa_oracle as eo, synth

SITES = ["weights", "tdnn1_out", "chain", "tdnn2_out", "block_out", "mfa_out", "att"]
WGROUPS = ["w:tdnn1", "w:res2net", "w:tdnn2", "w:mfa", "w:asp"]      # weight rounding, one layer family at a time


def run(waves, sd, keep_fp32=()):
    sd = {k: torch.as_tensor(v) for k, v in sd.items()}
    P = "__S__."

    def q(x, site):
        return x if site in keep_fp32 else x.to(torch.bfloat16).to(torch.float32)

    def tdnn(x, p, d, act, qw=True):
        w = sd[p + "conv.conv.weight"]; b = sd[p + "conv.conv.bias"]
        if qw:
            fam = "w:" + ("tdnn1" if "tdnn1" in p else "res2net" if "res2net" in p else "tdnn2" if "tdnn2" in p else "mfa")
            if fam not in keep_fp32: w = q(w, "weights")
        k = w.shape[-1]; pad = d * (k - 1) // 2
        if pad: x = F.pad(x, (pad, pad), mode='reflect')
        y = F.conv1d(x, w, b, dilation=d)
        y = F.gelu(y) if act == 'gelu' else F.relu(y)
        return eo._bn(y, sd, p + "norm.")
    with torch.no_grad():
        x = eo.log_cmn(eo.power_mel(torch.as_tensor(waves)))
        xs = q(tdnn(x, P + "blocks.0.", 1, 'gelu', qw=False), "block_out")      # block 0 is split (fp32 grade)
        outs = []
        for i, d in ((1, 2), (2, 3), (3, 4)):
            p = P + "blocks.%d." % i
            u = q(tdnn(xs, p + "tdnn1.", 1, 'gelu'), "tdnn1_out")
            ch = torch.chunk(u, 8, 1); ys = [ch[0]]; prev = None
            for j in range(1, 8):
                inp = ch[j] if j == 1 else q(ch[j] + prev, "chain")
                prev = tdnn(inp, p + "res2net_block.blocks.%d." % (j - 1), d, 'relu')
                ys.append(q(prev, "chain"))
            v = q(tdnn(torch.cat(ys, 1), p + "tdnn2.", 1, 'gelu'), "tdnn2_out")
            xs = q(eo.se_block(v, sd, p + "se_block.") + xs, "block_out")
            outs.append(xs)
        y = q(tdnn(torch.cat(outs, 1), P + "mfa.", 1, 'gelu'), "mfa_out")
        mean = y.mean(2); std = torch.sqrt(((y - mean[:, :, None]) ** 2).mean(2).clamp(1e-12))
        wa = sd[P + "asp.tdnn.conv.conv.weight"][:, :, 0]; C3 = y.shape[1]
        bias = (mean @ wa[:, C3:2 * C3].T + std @ wa[:, 2 * C3:].T) + sd[P + "asp.tdnn.conv.conv.bias"]
        qa = (lambda t: t) if "w:asp" in keep_fp32 else (lambda t: q(t, "weights"))
        h = torch.einsum('oc,bct->bot', qa(wa[:, :C3]), y) + bias[:, :, None]
        h = q(torch.tanh(eo._bn(F.relu(h), sd, P + "asp.tdnn.norm.")), "att")
        lg = F.conv1d(h, qa(sd[P + "asp.conv.conv.weight"]), sd[P + "asp.conv.conv.bias"])
        a = F.softmax(lg, 2)
        m = (a * y).sum(2); s = torch.sqrt((a * (y - m[:, :, None]) ** 2).sum(2).clamp(1e-12))
        pooled = eo._bn(torch.cat([m, s], 1).unsqueeze(2), sd, P + "asp_bn.")
        return F.conv1d(pooled, sd[P + "fc.conv.weight"], sd[P + "fc.conv.bias"]).squeeze(-1)


if __name__ == "__main__":
    g = dict(np.load('tests/golden/cfg2_c1024.npz'))
    bn = {k[3:]: v for k, v in g.items() if k.startswith('bn/')}
    sd = synth.ecapa_state_dict([1024, 1024, 1024, 1024, 3072], 192, bn_stats=bn)
    waves, _ = synth.utterance_set(16, 4, 48001)
    rn = F.normalize(torch.as_tensor(g['emb_batch']), dim=1); S0 = rn @ rn.T
    print("| fp32 kept at | max all-pairs score dev | mean | min cos |\n|---|---|---|---|")
    for keep in [()] + [(s,) for s in SITES] + [(w,) for w in WGROUPS] + [("w:tdnn1", "w:tdnn2"), tuple(SITES)]:
        t = time.time()
        en = F.normalize(run(waves[:, :48000], sd, keep), dim=1); S = en @ en.T
        print("| %s | %.3e | %.3e | %.6f |  (%.0f s)" % ("+".join(keep) if keep else "nothing (all bf16 sites)", (S - S0).abs().max(),
                                                        (S - S0).abs().mean(), (en * rn).sum(1).min(), time.time() - t), flush=True)
