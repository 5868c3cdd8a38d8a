"""Diagnostic: a few un-graphed C2 steps in one arithmetic mode, per-kernel event times printed (and the process ncu attaches to).
    python tests/diag/c2_step_probe.py PREC [steps] [unused] [T]"""
import ctypes as C, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import inr_for_audio_b200 as P
from inr_for_audio_b200 import _lib as L
prec = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
pf = int(sys.argv[3]) if len(sys.argv) > 3 else 0
T = int(sys.argv[4]) if len(sys.argv) > 4 else 2646000
H, Lh = int(os.environ.get("PROBE_H", "512")), int(os.environ.get("PROBE_L", "5"))
dev = torch.device("cuda:0")
gt = torch.from_numpy(P.utils.synth_waveform(T, seed=0)).to(dev)
torch.manual_seed(0)
m = P.Siren(1, H, Lh, 1, True)
f = P.SirenFitter(m.spec, m.flat_parameters(), gt, None, precision=prec, use_graph=False)
f.step(1)
torch.cuda.synchronize()
L.profile_begin()
f.step(steps)
prof = L.profile_end()
print(json.dumps({"prec": prec, "pf": pf, "T": T, "H": H, "L": Lh, "ms_per_step": {k: round(v[1] / steps, 4) for k, v in sorted(prof.items())},
                  "total": round(sum(v[1] for v in prof.values()) / steps, 3)}))
