"""Diagnostic (test infrastructure, imports the oracle): which 16-bit operand format lets the tensor-core SIREN step
converge like the fp32 reference loop?  Emulates the tcgen05 dataflow on the CPU with torch: fp32 accumulation, but the
GEMM operands S_l, C_l, W_l, dZ_l rounded to a chosen format (the first layer, sin/cos arguments, loss and Adam in fp32,
exactly like csrc/siren_tc.cu).  Toy of tests/golden/converge*.npz (gt_bach slice, T = 8192, Siren(1,128,2,1)).

    python tests/diag/opfmt_emul.py ARM STEPS [seed]       ARM in fp32 | bf16 | fp16 | fp16_wsplit | fp16_s_only | ...
"""
import os, sys, json, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import inr_oracle as O
from scipy.io import wavfile


def rnd(x, fmt):
    if fmt == "fp32":
        return x
    if fmt == "bf16":
        return x.bfloat16().float()
    if fmt == "fp16":
        return x.half().float()
    raise ValueError(fmt)


def step_emul(spec, ps, x, gt, fmtS, fmtC, fmtW, fmtWb, fmtdZ, gscale):
    """One loss + gradient evaluation with rounded GEMM operands; returns loss, grads (state_dict order)."""
    L = spec.hidden_layers
    om0, om = spec.first_omega_0, spec.hidden_omega_0
    W = [ps[2 * i] for i in range(L + 2)]
    b = [ps[2 * i + 1] for i in range(L + 2)]
    z0 = om0 * (x @ W[0].t() + b[0])
    S = [rnd(torch.sin(z0), fmtS)]
    C = [rnd(torch.cos(z0), fmtC)]
    for l in range(1, L + 1):
        z = om * (S[l - 1] @ rnd(W[l], fmtW).t() + b[l])
        S.append(rnd(torch.sin(z), fmtS))
        C.append(rnd(torch.cos(z), fmtC))
    out = S[L] @ W[L + 1].t() + b[L + 1]
    diff = out - gt
    T = x.shape[0]
    loss = float((diff.double() ** 2).mean())
    g = (2.0 / T) * diff                               # (T,1)
    grads = [None] * len(ps)
    grads[2 * (L + 1)] = g.t() @ S[L]
    grads[2 * (L + 1) + 1] = g.sum(0)
    dZ = rnd((g * gscale) @ W[L + 1] * om * C[L], fmtdZ)     # scaled by gscale (power of two)
    for l in range(L, 0, -1):
        grads[2 * l] = (dZ.t() @ S[l - 1]) / gscale
        grads[2 * l + 1] = dZ.sum(0) / gscale
        omp = om0 if l - 1 == 0 else om
        dZ = rnd((dZ @ rnd(W[l], fmtWb)) * omp * C[l - 1], fmtdZ)
    grads[0] = (dZ.t() @ x) / gscale
    grads[1] = dZ.sum(0) / gscale
    return loss, grads


ARMS = {  # S, C, W(fwd), W(bwd), dZ, dynamic grad scale
    "fp32": ("fp32",) * 5 + (False,),
    "bf16": ("bf16",) * 5 + (False,),
    "fp16": ("fp16",) * 5 + (True,),
    "fp16_noscale": ("fp16",) * 5 + (False,),
    "fp16_wsplit": ("fp16", "fp16", "fp32", "fp16", "fp16", True),      # W = hi + lo in the forward GEMM
    "fp16_fwd32": ("fp32", "fp16", "fp32", "fp16", "fp16", True),       # what an exact forward would give
    "fp16_s_only": ("fp16", "fp32", "fp32", "fp32", "fp32", True),
    "fp16_w_only": ("fp32", "fp32", "fp16", "fp32", "fp32", True),
    "fp16_bwd_only": ("fp32", "fp16", "fp32", "fp16", "fp16", True),
    "bf16_bwd_only": ("fp32", "bf16", "fp32", "bf16", "bf16", False),
}

if __name__ == "__main__":
    arm = sys.argv[1]
    steps = int(sys.argv[2])
    seed = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    H = int(os.environ.get("EMUL_H", "128")); Lh = int(os.environ.get("EMUL_L", "2")); T = int(os.environ.get("EMUL_T", "8192"))
    torch.set_num_threads(int(os.environ.get("EMUL_THREADS", "1")))
    sr, data = wavfile.read(os.path.join(ROOT, "data", "gt_bach.wav"))
    data = data.astype(np.float32)[44100:44100 + T]
    gtn = data / np.max(np.abs(data))
    spec = O.SirenSpec(1, H, Lh, 1, True)
    ps = O.siren_init(spec, seed=seed)
    x = O.get_coord(T, 1).view(-1, 1)
    gt = torch.from_numpy(gtn).view(-1, 1)
    fS, fC, fW, fWb, fdZ, dyn = ARMS[arm]
    st = O.AdamState(ps, lr=1e-3)
    sched = O.PlateauState(min_lr=1e-6)
    marks = sorted(set([100, 300, 500, 1000, 1500, 2000, 3000, 4000, 6000, 8000, 10000, 15000, 20000, steps]))
    rec = {"arm": arm, "seed": seed, "H": H, "L": Lh, "T": T, "snr": {}, "loss": {}}
    t0 = time.time()
    gscale = 1.0
    with torch.no_grad():
        for s in range(steps):
            if dyn:
                # loss scale from the previous step's loss: residual rms ~ sqrt(loss); g = 2 diff / T
                lprev = rec.get("_last", 1.0)
                gscale = 2.0 ** round(np.log2(T / max(np.sqrt(lprev), 1e-6)) + 4)
            loss, grads = step_emul(spec, ps, x, gt, fS, fC, fW, fWb, fdZ, gscale)
            rec["_last"] = loss
            O.adam_step(ps, grads, st)
            sched.step(loss, st)
            if (s + 1) in marks:
                out = O.siren_forward(spec, ps, x)
                rec["snr"][s + 1] = round(O.calculate_snr(gtn, out.numpy().ravel()), 3)
                rec["loss"][s + 1] = loss
    rec.pop("_last", None)
    rec["lr"] = st.lr
    rec["wall_s"] = round(time.time() - t0, 1)
    print(json.dumps(rec), flush=True)
