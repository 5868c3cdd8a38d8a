"""fp16 convergence-envelope probe (the tests/test_siren_gpu.py case) for one arithmetic mode: python tools/envelope_probe.py [fp16s|fp16|fp32] [runs]."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import inr_for_audio_b200 as P
from scipy.io import wavfile
sr, data = wavfile.read(os.path.join(ROOT, "data", "gt_bach.wav"))
data = data.astype(np.float32)[44100:44100 + 8192]
gtn = data / np.max(np.abs(data))
sig = float(np.mean(gtn.astype(np.float64) ** 2))
prec = sys.argv[1] if len(sys.argv) > 1 else "fp16"
for rep in range(int(sys.argv[2]) if len(sys.argv) > 2 else 3):
    torch.manual_seed(0)
    model = P.Siren(1, 128, 2, 1, True)
    f = P.SirenFitter(model.spec, model.flat_parameters(), torch.from_numpy(gtn).cuda(), precision=prec)
    f.step(1500)
    losses = f.losses().astype(np.float64)
    env = 10 * np.log10(sig / np.minimum.accumulate(losses))
    print(prec, " ".join("%d:%.2f" % (i + 1, env[i]) for i in (9, 49, 99, 299, 599, 999, 1499)), "loss[:3]", losses[:3])
