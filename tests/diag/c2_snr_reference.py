"""SNR after 200 / 1000 steps of the reference's own loop (oracle restatement of run.py:156-190,
pinned against the unmodified reference) in its real mode -- PyTorch eager fp32 on the same GPU -- for
Siren(1,512,5,1,True,30,30) on synth(2 646 000, seed 0).  One JSON line; bench.py quotes it from profiles/r02_snr_reference.json."""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import inr_for_audio_b200 as P  # noqa: E402
from oracle import inr_oracle as O  # noqa: E402

total = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
dev = torch.device("cuda:0")
T = 2646000
gtn = P.utils.synth_waveform(T, seed=0)
spec = O.SirenSpec(1, 512, 5, 1, True, 30.0, 30.0)
ps = [p.to(dev) for p in O.siren_init(spec, seed=0)]
coords = O.get_coord(T, 1).view(1, -1, 1).to(dev)
gt = torch.from_numpy(gtn).view(1, -1, 1).to(dev)
snr = {}


def rec(step, lv, params, grads, st):
    if step in (200,):
        with torch.no_grad():
            snr[str(step)] = float(O.calculate_snr(gtn, O.siren_forward(spec, params, coords).cpu().numpy().ravel()))


t0 = time.time()
losses, st, _ = O.siren_fit(spec, ps, coords, gt, total, record=rec)
with torch.no_grad():
    snr[str(total)] = float(O.calculate_snr(gtn, O.siren_forward(spec, ps, coords).cpu().numpy().ravel()))
torch.cuda.synchronize()
print(json.dumps({"config": "C2 synth 60 s, SIREN 5x512, seed 0", "arm": "reference loop, torch eager fp32 on B200", "snr_db": snr,
                  "loss_first_last": [losses[0], losses[-1]], "lr": st.lr, "wall_s": time.time() - t0, "ms_per_step": (time.time() - t0) / total * 1e3}))
