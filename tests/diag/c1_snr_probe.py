"""SNR (utils.py:77-97 calculate_snr(target, model(coords))) of the gt_bach 6 s fit with
Siren(1,256,3,1,True,30,30), seed 0, mse, Adam lr 1e-3, at 500 / 2 000 / 20 000 steps -- this library's fp32 and fp16
modes beside the reference's own loop run in its real mode (PyTorch eager on the same GPU: the oracle's restatement of
run.py:156-190, pinned against the unmodified reference).  One JSON line per arm.

    python tests/diag/c1_snr_probe.py [fp32] [fp16] [eager] [--steps 20000]
"""
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

MARKS = [500, 2000, 20000]


def main():
    known = ("fp32", "fp16", "fp16s", "eager", "eager_tf32")
    args = [a for a in sys.argv[1:] if a in known] or ["fp32", "fp16s", "eager"]
    total = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 20000
    seed = int(sys.argv[sys.argv.index("--seed") + 1]) if "--seed" in sys.argv else 0
    marks = [m for m in MARKS if m <= total]
    dev = torch.device("cuda:0")
    ds = P.WaveformFitting(os.path.join(ROOT, "data", "gt_bach.wav"), duration=6, verbose=False)
    gtn = ds.normalised()
    T = len(gtn)
    ospec = O.SirenSpec(1, 256, 3, 1, True, 30.0, 30.0)
    p0 = O.siren_init(ospec, seed=seed)
    flat0 = torch.cat([p.reshape(-1) for p in p0])
    for arm in args:
        t0 = time.time()
        snr, snr32 = {}, {}
        if arm in ("eager", "eager_tf32"):
            # the reference loop in torch eager; eager_tf32 lets cuBLAS use TF32 tensor cores for the fp32 matmuls
            # (10 explicit mantissa bits: the "fp32 mode on tensor cores" DESIGN.md §3 weighs against fp16 hi+lo)
            torch.backends.cuda.matmul.allow_tf32 = (arm == "eager_tf32")
            ps = [p.clone().to(dev) for p in p0]
            coords = O.get_coord(T, 1).view(1, -1, 1).to(dev)
            gt = torch.from_numpy(gtn).view(1, -1, 1).to(dev)

            def rec(step, lv, params, grads, st):
                if step in marks:
                    snr[step] = float(O.calculate_snr(gtn, O.siren_forward(ospec, params, coords).cpu().numpy().ravel()))
            losses, st, _ = O.siren_fit(ospec, ps, coords, gt, total, record=rec)
            snr[total] = float(O.calculate_snr(gtn, O.siren_forward(ospec, ps, coords).cpu().numpy().ravel()))
            lr, last = st.lr, losses[-1]
        else:
            spec = P.SirenSpec(1, 256, 3, 1, True, 30.0, 30.0)
            f = P.SirenFitter(spec, flat0.clone(), torch.from_numpy(gtn).to(dev), None, precision=arm, hist_cap=total)
            done = 0
            ws32 = P.engine.Workspace(dev)
            for m in marks:
                f.step(m - done)
                done = m
                snr[m] = float(P.calculate_snr(gtn, f.forward().cpu().numpy()))
                # the same parameters evaluated by the fp32 forward: separates the fit from the evaluation's own rounding
                o32 = P.engine.siren_forward(spec.desc("fp32"), f.params, None, T, ws32, 0, T, 1.0)
                snr32[m] = float(P.calculate_snr(gtn, o32.cpu().numpy()))
            lr, last = f.last_lr(), float(f.losses()[-1])
        torch.cuda.synchronize()
        print(json.dumps({"arm": arm, "config": "C1 gt_bach 6 s, SIREN 3x256", "T": T, "seed": seed, "snr_db": snr, "snr_db_fp32_eval": snr32, "final_lr": lr,
                          "last_loss": last, "wall_s": time.time() - t0}), flush=True)


if __name__ == "__main__":
    main()
