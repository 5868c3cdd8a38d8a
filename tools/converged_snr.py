"""Converged SNR of one fit on the gt_bach 6 s clip (T = 264 600, utils.py:77-97 calculate_snr(target, model(coords)))
for network families beyond C1's 3x256, in this library's modes and in the reference's own loop (the oracle's pinned
restatement of run.py:156-190 run in torch eager on the same GPU).  One JSON line per arm (append to a .jsonl).

    python tools/converged_snr.py --net h512|c1snake|c1 [--seed S] [--steps 20000] fp16s [fp32] [fp16] [eager]

  h512     Siren(1, 512, 5, 1, True, 30, 30): the layer-major H = 512 kernels the C2 headline times
           (k_tc_layer2<512,.,SPLIT>, k_tc_dw2<512>, k_tc_last2<512>), on a clip that converges in 20 000 steps
  c1snake  run.py:30's default architecture: SirenWithSnakeTanh(num_sine=2, num_snake=2, hidden 256, omega 22000,
           hidden_omega 30, a_initial 0.5): the layer-major Snake kernels
  c1       Siren(1, 256, 3, 1, True, 30, 30) (BASELINE.json configs[0]; same as tests/diag/c1_snr_probe.py)

This is a measurement tool: it imports oracle/ as the comparison arm, so it lives outside the product package.
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import inr_for_audio_b200 as P  # noqa: E402
from oracle import inr_oracle as O  # noqa: E402

NETS = {
    "h512": dict(hidden_features=512, hidden_layers=5, first_omega_0=30.0, hidden_omega_0=30.0),
    "c1snake": dict(hidden_features=256, hidden_layers=2, num_snake=2, first_omega_0=22000.0, hidden_omega_0=30.0),
    "c1": dict(hidden_features=256, hidden_layers=3, first_omega_0=30.0, hidden_omega_0=30.0),
}
A_INITIAL = {"c1snake": 0.5}


def _opt(name, default, cast):
    return cast(sys.argv[sys.argv.index(name) + 1]) if name in sys.argv else default


def main():
    net = _opt("--net", "h512", str)
    seed = _opt("--seed", 0, int)
    total = _opt("--steps", 20000, int)
    arms = [a for a in sys.argv[1:] if a in ("fp32", "fp16", "fp16s", "eager")] or ["fp16s"]
    marks = [m for m in (500, 2000, 5000, 10000, 20000) if m <= total]
    if total not in marks:
        marks.append(total)
    dev = torch.device("cuda:0")
    ds = P.WaveformFitting(os.path.join(ROOT, "data", "gt_bach.wav"), duration=6, verbose=False)
    gtn = ds.normalised()
    T = len(gtn)
    kw = NETS[net]
    ospec = O.SirenSpec(1, kw["hidden_features"], kw["hidden_layers"], 1, True, kw["first_omega_0"], kw["hidden_omega_0"],
                        num_snake=kw.get("num_snake", 0), a_initial=A_INITIAL.get(net, 50.0))
    spec = P.SirenSpec(1, kw["hidden_features"], kw["hidden_layers"], 1, True, kw["first_omega_0"], kw["hidden_omega_0"],
                       num_snake=kw.get("num_snake", 0))
    p0 = O.siren_init(ospec, seed=seed)
    flat0 = torch.cat([p.reshape(-1) for p in p0])
    for arm in arms:
        t0 = time.time()
        snr, lrs, sat = {}, {}, {}
        if arm == "eager":
            torch.backends.cuda.matmul.allow_tf32 = False
            ps = [p.clone().to(dev) for p in p0]
            coords = O.get_coord(T, 1).view(1, -1, 1).to(dev)
            gt = torch.from_numpy(gtn).view(1, -1, 1).to(dev)

            def rec(step, lv, params, grads, st):
                if step in marks:
                    with torch.no_grad():
                        snr[step] = float(O.calculate_snr(gtn, O.siren_forward(ospec, params, coords).cpu().numpy().ravel()))
                    lrs[step] = float(st.lr)
            losses, st, _ = O.siren_fit(ospec, ps, coords, gt, total, record=rec)
            with torch.no_grad():
                snr[total] = float(O.calculate_snr(gtn, O.siren_forward(ospec, ps, coords).cpu().numpy().ravel()))
            lr, last = float(st.lr), float(losses[-1])
        else:
            f = P.SirenFitter(spec, flat0.clone(), torch.from_numpy(gtn).to(dev), None, precision=arm, hist_cap=total)
            done = 0
            for m in marks:
                f.step(max(m - done, 0))
                done = m
                snr[m] = float(P.calculate_snr(gtn, f.forward().cpu().numpy()))
                lrs[m] = float(f.last_lr())
                if arm in ("fp16", "fp16s"):
                    f.step(1)                      # the forward above reused the workspace: look right after a training step
                    done += 1
                    sat[m] = f.dz_saturated()
            lr, last = float(f.last_lr()), float(f.losses()[-1])
            f.close()
        torch.cuda.synchronize()
        wall = time.time() - t0
        print(json.dumps({"net": net, "arm": arm, "T": T, "seed": seed, "steps": total, "snr_db": snr, "lr_at": lrs, "dz_saturated_at": sat,
                          "final_lr": lr, "last_loss": last, "wall_s": round(wall, 1),
                          "ms_per_step": round(1e3 * wall / total, 3)}), flush=True)


if __name__ == "__main__":
    main()
