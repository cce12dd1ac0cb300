"""CPU emulation of reduced-pass fp16-pair schemes for the decoder (precision study, fp64 accumulation).
Schemes: fwd products {hh, hl, lh, ll} subsets; bwd products likewise.  Prints SDF and Jacobian-row errors
against the fp64 oracle."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth
from oracle import gn_oracle as G
torch.set_num_threads(8)
st = synth.make_decoder_state(0)
fd = G.FoldedDecoder(st, torch.float64)
rng = np.random.default_rng(5)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
x = torch.from_numpy(synth._ball_points(rng, n)).double()
z = torch.from_numpy((rng.standard_normal(64) * 0.1).astype(np.float32)).double()
s_ref, g_ref = G.decoder_forward_jacobian(fd, z, x)

def hi(t): return t.half().double()
def lo(t): return (t - hi(t)).half().double()

def mm(a, w, scheme):
    """a @ w.T with the listed partial products; a, w fp32-valued doubles"""
    a = a.float().double(); w = w.float().double()
    ah, al, wh, wl = hi(a), lo(a), hi(w), lo(w)
    out = 0
    if "hh" in scheme: out = out + ah @ wh.t()
    if "lh" in scheme: out = out + al @ wh.t()
    if "hl" in scheme: out = out + ah @ wl.t()
    if "ll" in scheme: out = out + al @ wl.t()
    return out

def run(fs, bs):
    inp = torch.cat([z.expand(n, -1), x], 1)
    a = inp; masks = []
    for l in range(8):
        if l == 4: a = torch.cat([a, inp], 1)
        a = mm(a, fd.W[l], fs) + fd.b[l]
        masks.append(a > 0); a = torch.relu(a)
    s = torch.tanh(a.float().double() @ fd.W[8][0] + fd.b[8][0])
    g = ((1 - s * s)[:, None] * fd.W[8]) * masks[7]
    skip = None
    for l in range(7, 0, -1):
        g = mm(g, fd.W[l].t(), bs)
        if l == 4: skip = g[:, -67:]; g = g[:, :-67]
        g = g * masks[l - 1]
    g = mm(g, fd.W[0].t(), bs) + skip
    return s, g

for fs, bs in [(("hh","lh","hl","ll"),)*2, (("hh","lh","hl"),)*2, (("hh","lh","hl"),("hh","lh")), (("hh","lh","hl"),("hh","hl")), (("hh","lh","hl"),("hh",)), (("hh",),("hh",)), (("hh","lh"),("hh","lh"))]:
    s, g = run(fs, bs)
    es = (s - s_ref).abs()
    rel = (g - g_ref).norm(dim=1) / g_ref.norm(dim=1)
    ok = rel < 0.05   # rows with a flipped ReLU are a different story
    print(f"fwd {'+'.join(fs):12s} bwd {'+'.join(bs):12s}: sdf err max {es.max():.2e} rms {es.pow(2).mean().sqrt():.2e} | "
          f"jac row rel err median {rel.median():.2e} p99 {rel.quantile(0.99):.2e} max(<5%) {rel[ok].max():.2e} rows>1e-3 {(rel>1e-3).sum().item()} flipped {(~ok).sum().item()}")
