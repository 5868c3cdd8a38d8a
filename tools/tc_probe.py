"""One tcgen05 kernel case in its own process (a trapped kernel kills the CUDA context):
    python tools/tc_probe.py layer H mode variant rows      |     python tools/tc_probe.py dw H variant rows
Prints 'RESULT ... max_err=... ok=...'.  Used by tests/test_tc_gpu.py and for the descriptor-convention probe."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inr_for_audio_b200 import _lib as L  # noqa: E402


def bf(x):
    """round to the operand format of the SIREN tensor-core path (IEEE fp16)"""
    return x.to(torch.float16).to(torch.float32)


def run_layer(H, mode, variant, rows):
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(H + mode + rows)
    A = (torch.rand(rows, H, generator=g) * 2 - 1).to(dev)
    W = ((torch.rand(H, H, generator=g) * 2 - 1) * (6.0 / H) ** 0.5 / 30.0).to(dev)
    bias = ((torch.rand(H, generator=g) * 2 - 1) / H ** 0.5).to(dev)
    Cin = (torch.rand(rows, H, generator=g) * 2 - 1).to(dev)
    omega = 30.0
    out0 = torch.full((rows, H), float("nan"), device=dev)
    out1 = torch.full((rows, H), float("nan"), device=dev)
    nb = L.lib().inr_dbg_tc_scratch_bytes(H, rows)
    scratch = torch.empty(nb, dtype=torch.uint8, device=dev)
    L.check(L.lib().inr_dbg_tc_layer(H, mode, L.ptr(A), L.ptr(W), L.ptr(bias), omega, L.ptr(Cin), L.ptr(out0), L.ptr(out1),
                                     rows, variant, L.ptr(scratch), nb, L.stream_ptr()))
    torch.cuda.synchronize()
    if mode == 0:
        z = bf(A).double() @ bf(W).double().t()
        t = omega * z + (omega * bias).double()
        e0 = (out0.double() - torch.sin(t)).abs().max().item()
        e1 = (out1.double() - torch.cos(t)).abs().max().item()
        err, tol = max(e0, e1), 2e-3
    else:
        z = bf(A).double() @ bf(W).double()
        ref = z * omega * bf(Cin).double()
        err = ((out0.double() - ref).abs() / (ref.abs() + 0.05)).max().item()
        tol = 2e-3
    return err, tol


def run_dw(H, variant, rows):
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(H + rows)
    dZ = (torch.rand(rows, H, generator=g) * 2 - 1).to(dev)
    S = (torch.rand(rows, H, generator=g) * 2 - 1).to(dev)
    gW = torch.zeros(H, H, device=dev)
    nb = L.lib().inr_dbg_tc_scratch_bytes(H, rows)
    scratch = torch.empty(nb, dtype=torch.uint8, device=dev)
    L.check(L.lib().inr_dbg_tc_dw(H, L.ptr(dZ), L.ptr(S), L.ptr(gW), rows, variant, L.ptr(scratch), nb, L.stream_ptr()))
    torch.cuda.synchronize()
    ref = bf(dZ).double().t() @ bf(S).double()
    err = (gW.double() - ref).abs().max().item() / (rows ** 0.5)
    return err, 2e-4


if __name__ == "__main__":
    kind = sys.argv[1]
    if kind == "layer":
        H, mode, variant, rows = map(int, sys.argv[2:6])
        err, tol = run_layer(H, mode, variant, rows)
        print("RESULT layer H=%d mode=%d variant=%d rows=%d max_err=%.3e ok=%d" % (H, mode, variant, rows, err, err <= tol))
    else:
        H, variant, rows = map(int, sys.argv[2:5])
        err, tol = run_dw(H, variant, rows)
        print("RESULT dw H=%d variant=%d rows=%d max_err=%.3e ok=%d" % (H, variant, rows, err, err <= tol))
