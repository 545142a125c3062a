"""CPU numerical model of the planned bf16 data flow (design aid, not product)."""
import sys, numpy as np, torch, torch.nn.functional as F
sys.path.insert(0, '.')
from oracle import ecapa_oracle as eo, synth

def q(x): return x.to(torch.bfloat16).to(torch.float32)

def run(waves, sd, resid_hilo=False, asp_logits_bf16=True):
    sd = {k: torch.as_tensor(v) for k, v in sd.items()}
    P = "__S__."
    def tdnn(x, p, d, act, qin=True):
        w = sd[p+"conv.conv.weight"]; b = sd[p+"conv.conv.bias"]
        if qin: x = q(x); w = q(w)
        k = w.shape[-1]; pad = d*(k-1)//2
        if pad: x = F.pad(x, (pad,pad), mode='reflect')
        y = F.conv1d(x, w, b, dilation=d)
        y = F.gelu(y) if act=='gelu' else F.relu(y)
        return eo._bn(y, sd, p+"norm.")
    with torch.no_grad():
        mel = eo.power_mel(torch.as_tensor(waves))
        x = eo.log_cmn(mel)
        x = tdnn(x, P+"blocks.0.", 1, 'gelu', qin=False)   # split (fp32-grade) block0
        xs = q(x); xfull = x if resid_hilo else xs
        outs = []
        for i,d in ((1,2),(2,3),(3,4)):
            p = P+"blocks.%d."%i
            u = q(tdnn(xs, p+"tdnn1.", 1, 'gelu'))
            ch = torch.chunk(u, 8, 1); ys=[ch[0]]; prev=None
            for j in range(1,8):
                inp = ch[j] if j==1 else q(ch[j]+prev)
                prev = tdnn(inp, p+"res2net_block.blocks.%d."%(j-1), d, 'relu')
                ys.append(q(prev))
            r = torch.cat(ys,1)
            v = q(tdnn(r, p+"tdnn2.", 1, 'gelu'))
            o = eo.se_block(v, sd, p+"se_block.") + xfull
            xs = q(o); xfull = o if resid_hilo else xs
            outs.append(xs)
        y = q(tdnn(torch.cat(outs,1), P+"mfa.", 1, 'gelu'))
        # ASP (fp32 statistics, bf16 GEMM operands)
        L = y.shape[-1]
        mean = y.mean(2); std = torch.sqrt(((y-mean[:,:,None])**2).mean(2).clamp(1e-12))
        wa = sd[P+"asp.tdnn.conv.conv.weight"][:,:,0]; C3 = y.shape[1]
        bias = (mean @ wa[:,C3:2*C3].T + std @ wa[:,2*C3:].T) + sd[P+"asp.tdnn.conv.conv.bias"]
        h = torch.einsum('oc,bct->bot', q(wa[:,:C3]), y) + bias[:,:,None]
        h = eo._bn(F.relu(h), sd, P+"asp.tdnn.norm.")
        h = q(torch.tanh(h))
        lg = F.conv1d(h, q(sd[P+"asp.conv.conv.weight"]), sd[P+"asp.conv.conv.bias"])
        a = F.softmax(lg, 2)
        m = (a*y).sum(2); s = torch.sqrt((a*(y-m[:,:,None])**2).sum(2).clamp(1e-12))
        pooled = torch.cat([m,s],1).unsqueeze(2)
        pooled = eo._bn(pooled, sd, P+"asp_bn.")
        return F.conv1d(pooled, sd[P+"fc.conv.weight"], sd[P+"fc.conv.bias"]).squeeze(-1)

if __name__ == "__main__":
    g = dict(np.load('tests/golden/cfg1_c512.npz'))
    bn = {k[3:]: v for k,v in g.items() if k.startswith('bn/')}
    sd = synth.ecapa_state_dict([512,512,512,512,1536], 192, bn_stats=bn)
    waves,_ = synth.utterance_set(16,4,48001)
    ref = torch.as_tensor(g['emb_batch']); rn = F.normalize(ref, dim=1); S0 = rn@rn.T
    for hilo in (False, True):
        e = run(waves[:, :48000], sd, resid_hilo=hilo)
        en = F.normalize(e, dim=1); S = en@en.T
        cos = (en*rn).sum(1)
        print("resid_hilo", hilo, "min cos %.6f  score dev max %.5f mean %.5f" % (cos.min(), (S-S0).abs().max(), (S-S0).abs().mean()))
