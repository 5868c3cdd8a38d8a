"""Diagnostic (imports the oracle): where does the fp16 tensor-core mode's C1 fit leave the fp32 fit, and what is its
forward rounding noise made of?  (1) SNR / lr / loss of the fit at a ladder of step counts, own forward and fp32 forward;
(2) at the final parameters, torch emulation of the forward with only S_l, only W_l, or both rounded to fp16, against
the exact fp32 forward.

    python tests/diag/c1_noise_probe.py PREC TOTAL [seed]
"""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import inr_for_audio_b200 as P  # noqa: E402
from oracle import inr_oracle as O  # noqa: E402


def emul_forward(ps, x, L, om0, om, rs, rw):
    r16 = lambda t: t.half().float()
    S = torch.sin(om0 * (x @ ps[0].t() + ps[1]))
    if rs:
        S = r16(S)
    for l in range(1, L + 1):
        W = r16(ps[2 * l]) if rw else ps[2 * l]
        S = torch.sin(om * ((S.double() @ W.double().t()).float() + ps[2 * l + 1]))
        if rs:
            S = r16(S)
    return (S.double() @ ps[2 * (L + 1)].double().t()).float() + ps[2 * (L + 1) + 1]


def main():
    prec, total = sys.argv[1], int(sys.argv[2])
    seed = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    dev = torch.device("cuda:0")
    ds = P.WaveformFitting(os.path.join(ROOT, "data", "gt_bach.wav"), duration=6, verbose=False)
    gtn = ds.normalised()
    T = len(gtn)
    ospec = O.SirenSpec(1, 256, 3, 1, True, 30.0, 30.0)
    p0 = O.siren_init(ospec, seed=seed)
    flat0 = torch.cat([p.reshape(-1) for p in p0])
    spec = P.SirenSpec(1, 256, 3, 1, True, 30.0, 30.0)
    f = P.SirenFitter(spec, flat0.clone(), torch.from_numpy(gtn).to(dev), None, precision=prec, hist_cap=total)
    ws32 = P.engine.Workspace(dev)
    marks = [m for m in (1000, 2000, 3000, 4000, 6000, 8000, 10000, 12000, 15000, 20000) if m <= total]
    done = 0
    t0 = time.time()
    for m in marks:
        f.step(m - done)
        done = m
        own = float(P.calculate_snr(gtn, f.forward().cpu().numpy()))
        o32 = P.engine.siren_forward(spec.desc("fp32"), f.params, None, T, ws32, 0, T, 1.0)
        st = f.opt_state()
        l = f.losses()
        print(json.dumps({"prec": prec, "step": m, "snr_own": own, "snr_fp32eval": float(P.calculate_snr(gtn, o32.cpu().numpy())),
                          "lr": st.lr, "loss": float(l[-1]), "loss_min": float(l.min()), "gmax_prev": st.gmax_prev,
                          "wall_s": time.time() - t0}), flush=True)
    # decomposition of the forward rounding noise at the final parameters
    sizes = spec.param_sizes()
    shapes = [(256, 1), (256,)] + [(256, 256), (256,)] * 3 + [(1, 256), (1,)]
    ps, o = [], 0
    for n, sh in zip(sizes, shapes):
        ps.append(f.params[o:o + n].view(sh).clone())
        o += n
    x = O.get_coord(T, 1).view(-1, 1).to(dev)
    gt = torch.from_numpy(gtn).to(dev)
    ref = emul_forward(ps, x, 3, 30.0, 30.0, False, False).view(-1)
    sig = float((gt.double() ** 2).mean())
    rec = {"prec": prec, "fit_err_db": 10 * np.log10(float(((ref - gt).double() ** 2).mean()) / sig)}
    for name, rs, rw in (("S", True, False), ("W", False, True), ("SW", True, True)):
        o = emul_forward(ps, x, 3, 30.0, 30.0, rs, rw).view(-1)
        rec["noise_%s_db" % name] = 10 * np.log10(float(((o - ref).double() ** 2).mean()) / sig)
    own = f.forward()
    rec["noise_own_db"] = 10 * np.log10(float(((own - ref).double() ** 2).mean()) / sig)
    rec["w_last_rms"] = float(ps[-2].pow(2).mean().sqrt())
    rec["w_rms"] = [float(ps[2 * l].pow(2).mean().sqrt()) for l in range(0, 5)]
    rec["w_absmax"] = [float(ps[2 * l].abs().max()) for l in range(0, 5)]
    print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
