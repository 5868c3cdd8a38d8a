"""Informative competitor of SURVEY.md §8(d): the reference's own execution mode -- PyTorch eager on the same B200 --
for the fit step (forward + loss + backward + Adam + plateau scheduler), timed with CUDA events after warm-up.

The loop is the oracle's restatement of run.py:156-190 (oracle/inr_oracle.py siren_fit, pinned against the unmodified
reference) with every tensor on cuda:0; torch defaults (fp32 matmul, TF32 off).  Lives under tests/ because only tests/,
smoke() and bench.py's baseline legs may execute oracle/.  Prints one JSON line per workload.

    python tests/diag/eager_gpu_probe.py [c1] [c2q]      (c2q = SIREN 5x512 on a quarter of C2's coordinates)
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import inr_oracle as O  # noqa: E402

WORK = {"c1": dict(T=264600, H=256, L=3, seed=0), "c2q": dict(T=2646000 // 4, H=512, L=5, seed=0),
        "c3": dict(T=1323000, H=256, L=3, seed=1)}


def run(name, steps=10, warmup=3):
    w = WORK[name]
    dev = torch.device("cuda:0")
    spec = O.SirenSpec(1, w["H"], w["L"], 1, True, 30.0, 30.0)
    ps = [p.to(dev) for p in O.siren_init(spec, seed=w["seed"])]
    coords = O.get_coord(w["T"], 1).view(1, -1, 1).to(dev)
    gt = torch.from_numpy(O.synth(w["T"], seed=w["seed"]).astype(np.float32)).view(1, -1, 1).to(dev)
    O.siren_fit(spec, ps, coords, gt, warmup)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    losses, st, _ = O.siren_fit(spec, ps, coords, gt, steps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print(json.dumps({"impl": "torch-eager-cuda (oracle loop)", "workload": name, "T": w["T"], "H": w["H"], "L": w["L"],
                      "ms_per_step": ms, "coord_samples_per_s": w["T"] / ms * 1e3, "steps": steps, "warmup": warmup,
                      "torch": torch.__version__, "tf32_matmul": torch.backends.cuda.matmul.allow_tf32,
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30, "loss_last": float(losses[-1])}))


if __name__ == "__main__":
    for n in (sys.argv[1:] or ["c1", "c2q"]):
        run(n)
