"""C1 (gt_bach 6 s, SIREN 3x256, seed 0, 20 000 steps): the fit started in the fp16 tcgen05 mode and finished in the fp32
mode (SirenFitter.set_precision), for several switch points; SNR = calculate_snr(target, model(coords)) (utils.py:77-97).
One JSON line per arm.   python tests/diag/c1_hybrid_probe.py [switch_step ...]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import inr_for_audio_b200 as P  # noqa: E402
from oracle import inr_oracle as O  # noqa: E402

TOTAL = 20000
MARKS = [2000, 5000, 10000, 14000, 16000, 18000, 19000, 20000]


def main():
    switches = [int(a) for a in sys.argv[1:] if a != "auto"] or ([] if "auto" in sys.argv[1:] else [20000, 18000, 16000, 10000])
    dev = torch.device("cuda:0")
    gtn = P.WaveformFitting(os.path.join(ROOT, "data", "gt_bach.wav"), duration=6, verbose=False).normalised()
    flat0 = torch.cat([p.reshape(-1) for p in O.siren_init(O.SirenSpec(1, 256, 3, 1, True, 30.0, 30.0), seed=0)])
    spec = P.SirenSpec(1, 256, 3, 1, True, 30.0, 30.0)
    if "auto" in sys.argv[1:]:
        f = P.SirenFitter(spec, flat0.clone(), torch.from_numpy(gtn).to(dev), None, precision="fp16", hist_cap=TOTAL)
        torch.cuda.synchronize()
        t0 = time.time()
        sw = f.step_auto(TOTAL)
        torch.cuda.synchronize()
        print(json.dumps({"arm": "auto (step_auto: noise ratio > 0.05)", "switched_at": sw,
                          "snr_db": {TOTAL: round(float(P.calculate_snr(gtn, f.forward().cpu().numpy())), 3)},
                          "last_loss": float(f.losses()[-1]), "final_lr": f.last_lr(), "wall_s": round(time.time() - t0, 2)}),
              flush=True)
    for sw in switches:
        f = P.SirenFitter(spec, flat0.clone(), torch.from_numpy(gtn).to(dev), None, precision="fp16", hist_cap=TOTAL)
        torch.cuda.synchronize()
        t0 = time.time()
        snr, lrs, done = {}, {}, 0
        for m in sorted(set(MARKS + [sw])):
            if m > TOTAL:
                break
            f.step(m - done)
            done = m
            snr[m] = round(float(P.calculate_snr(gtn, f.forward().cpu().numpy())), 3)
            lrs[m] = f.last_lr()
            if m == sw and sw < TOTAL:
                f.set_precision("fp32")
        torch.cuda.synchronize()
        print(json.dumps({"arm": "fp16 -> fp32 at step %d" % sw if sw < TOTAL else "fp16 only", "snr_db": snr, "lr": lrs,
                          "last_loss": float(f.losses()[-1]), "wall_s": round(time.time() - t0, 2)}), flush=True)


if __name__ == "__main__":
    main()
