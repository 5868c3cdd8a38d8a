"""Diagnostic (test infrastructure, imports the oracle): loss trajectories of our fitter vs the oracle loop run
with torch eager on the same GPU, for the converge.npz configuration."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import inr_for_audio_b200 as P
from oracle import inr_oracle as O
from scipy.io import wavfile

idx = [0, 1, 2, 10, 100, 300, 500, 700, 1000, 1200, 1499]
sr, data = wavfile.read(os.path.join(ROOT, "data", "gt_bach.wav"))
data = data.astype(np.float32)[44100:44100 + 8192]
gtn = data / np.max(np.abs(data))
dev = torch.device("cuda:0")
g = np.load(os.path.join(ROOT, "tests/golden/converge.npz"))
print("ref t1", g["losses_t1"], g["lr_final_t1"], g["snr_t1"])
print("ref t8", g["losses_t8"], g["lr_final_t8"], g["snr_t8"])
prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
for rep in range(3):
    torch.manual_seed(0)
    model = P.Siren(1, 128, 2, 1, True)
    f = P.SirenFitter(model.spec, model.flat_parameters(), torch.from_numpy(gtn).to(dev), use_graph=(rep != 2), precision=prec)
    f.step(1500)
    l = f.losses()
    st = f.opt_state()
    snr = P.calculate_snr(gtn, f.forward().cpu().numpy())
    print("ours rep%d" % rep, " ".join("%.3e" % l[i] for i in idx), "lr %.2e best %.3e nbad %d step %d snr %.2f min %.3e" % (st.lr, st.best, st.num_bad, st.step, snr, l.min()))
# oracle loop with torch eager on the GPU
for rep in range(2):
    spec = O.SirenSpec(1, 128, 2, 1, True)
    ps = [p.to(dev) for p in O.siren_init(spec, seed=0)]
    coords = O.get_coord(8192, 1).view(1, -1, 1).to(dev)
    gt = torch.from_numpy(gtn).view(1, -1, 1).to(dev)
    losses, st, sched = O.siren_fit(spec, ps, coords, gt, 1500)
    out = O.siren_forward(spec, ps, coords).cpu().numpy().ravel()
    print("eager-gpu rep%d" % rep, " ".join("%.3e" % losses[i] for i in idx), "lr %.2e snr %.2f" % (st.lr, O.calculate_snr(gtn, out)))
