"""Per-tensor gradient error of the fp16 mode against the fp32 oracle (diagnostic): python tools/grad_diag.py H L n"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import inr_for_audio_b200 as P
from oracle import inr_oracle as O
H, Lh, n = (int(x) for x in sys.argv[1:4])
spec = P.SirenSpec(1, H, Lh, 1, True); ospec = O.SirenSpec(1, H, Lh, 1, True)
ps = O.siren_init(ospec, seed=H + Lh)
flat = torch.cat([p.reshape(-1) for p in ps])
gt = torch.from_numpy(P.utils.synth_waveform(max(n, 2), seed=3)[:n].copy())
coords = O.get_coord(n, 1).view(-1, 1)
loss_o, _, grads_o, out_o = O.siren_loss_and_grads(ospec, ps, coords, gt.view(-1, 1))
f = P.SirenFitter(spec, flat, gt.cuda(), precision="fp16", use_graph=False)
loss, grads, out = f.loss_and_grads()
print("loss", loss, float(loss_o))
o = 0
g = grads.cpu().numpy().astype(np.float64)
for i, t in enumerate(grads_o):
    m = t.numel(); a = g[o:o + m]; b = t.numpy().ravel().astype(np.float64); o += m
    print("tensor %d %-12s rel err %.3e  |g| %.3e" % (i, tuple(t.shape), np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30), np.linalg.norm(b)))
