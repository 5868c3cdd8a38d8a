"""inr_siren_forward_static -- the forward for callers that cannot allocate device memory -- against inr_siren_forward.

Why the entry point exists: a real-time audio render callback may not call cudaMalloc/cudaFree (driver-wide locks and
possible device synchronisation: unbounded latency), a process whose HBM is already owned by another allocator gets an
out-of-memory from a late cudaMalloc, and a plain-C host may have no CUDA binding to allocate with.  The static variant
keeps the activations in one INR_STATIC_SCRATCH_BYTES array inside the library instead of a caller workspace.

What is checked, through the raw C ABI: on several calls in a row -- different networks, so each call finds the
previous call's bytes in the shared array -- the output equals the ordinary entry point's within the forward tolerances
of tests/test_siren_gpu.py (3e-6 fp32) and tests/test_tc_gpu.py (2e-5 fp16s); a request beyond the array is refused
with INR_E_WORKSPACE; a warmed-up call leaves the device's free memory unchanged.

This file is named to be collected LAST: under `pytest -x` a failure here must not hide the other tests.  It passes
alone in a fresh process (its first call does the per-device set-up) and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _net(L, H, layers, precision, seed, in_features=1):
    """descriptor + SIREN-initialised flat parameters (models.py:94-115,258-262 ranges) on the device"""
    d = L.SirenDesc(in_features, H, layers, 1, 1, 30.0, 30.0, L.PRECISIONS[precision], 0, 0, 0)
    P = L.lib().inr_siren_num_params(C.byref(d))
    rng = np.random.default_rng(seed)
    parts = []
    fan = in_features
    for l in range(layers + 1):
        bound = 1.0 / fan if l == 0 else np.sqrt(6.0 / fan) / 30.0
        parts += [rng.uniform(-bound, bound, H * fan), rng.uniform(-1, 1, H) / np.sqrt(fan)]
        fan = H
    parts += [rng.uniform(-1, 1, H) * np.sqrt(6.0 / H) / 30.0, rng.uniform(-1, 1, 1) / np.sqrt(H)]
    p = np.concatenate(parts).astype(np.float32)
    assert p.size == P
    return d, torch.from_numpy(p).cuda()


def _forward_ws(L, d, params, coords, n, T):
    """the existing entry point, with a torch-allocated workspace"""
    ws_bytes = L.lib().inr_siren_workspace_bytes(C.byref(d), n, 0)
    assert ws_bytes > 0
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    out = torch.empty(n, device="cuda")
    L.check(L.lib().inr_siren_forward(C.byref(d), L.ptr(params), L.ptr(coords), n, 0, T, 1.0, L.ptr(out), L.ptr(ws),
                                      ws_bytes, L.stream_ptr()))
    torch.cuda.synchronize()
    return out


def _forward_static(L, d, params, coords, n, T, out=None):
    out = torch.empty(n, device="cuda") if out is None else out
    L.check(L.lib().inr_siren_forward_static(C.byref(d), L.ptr(params), L.ptr(coords), n, 0, T, 1.0, L.ptr(out),
                                             L.stream_ptr()))
    torch.cuda.synchronize()
    return out


def test_static_scratch_forward_matches_workspace_forward():
    from inr_for_audio_b200 import _lib as L
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)

    # (network, rows, explicit coordinates?, tolerance): consecutive calls differ in width, depth, row count and mode
    cases = [
        (_net(L, 256, 3, "fp32", 0), 10000, False, 3e-6),
        (_net(L, 64, 2, "fp32", 1), 4096, True, 3e-6),
        (_net(L, 256, 3, "fp16s", 2), 5000, False, 2e-5),
        (_net(L, 512, 2, "fp32", 3), 8000, False, 3e-6),
        (_net(L, 256, 3, "fp32", 0), 10000, False, 3e-6),      # the first case again, over everybody's leftovers
    ]
    for (d, params), n, explicit, tol in cases:
        T = 3 * n + 1
        coords = torch.linspace(-1, 1, n, device="cuda") if explicit else None
        ref = _forward_ws(L, d, params, coords, n, T)
        assert torch.isfinite(ref).all() and ref.abs().max() > 1e-4
        for _ in range(2):                                     # two calls in a row on the same inputs
            got = _forward_static(L, d, params, coords, n, T)
            np.testing.assert_allclose(got.cpu().numpy(), ref.cpu().numpy(), rtol=0, atol=tol)

    # a warmed-up call allocates nothing: the device's free memory is the same before and after
    (d, params), n = cases[0][0], cases[0][1]
    out = torch.empty(n, device="cuda")
    _forward_static(L, d, params, None, n, 3 * n + 1, out)
    free0 = torch.cuda.mem_get_info()[0]
    _forward_static(L, d, params, None, n, 3 * n + 1, out)
    assert torch.cuda.mem_get_info()[0] == free0


def test_static_scratch_rejects_oversized_request():
    from inr_for_audio_b200 import _lib as L
    d, params = _net(L, 256, 3, "fp32", 0)
    n = 1 << 20                                                # 2 KiB of activations per row: far beyond 32 MiB
    assert L.lib().inr_siren_workspace_bytes(C.byref(d), n, 0) > L.STATIC_SCRATCH_BYTES
    out = torch.empty(n, device="cuda")
    rc = L.lib().inr_siren_forward_static(C.byref(d), L.ptr(params), None, n, 0, n, 1.0, L.ptr(out), L.stream_ptr())
    assert rc == L.INR_E_WORKSPACE
    assert b"INR_STATIC_SCRATCH_BYTES" in L.lib().inr_last_error()
    # the largest request that fits is accepted
    per_row = 2 * 256 * 4
    n_ok = L.STATIC_SCRATCH_BYTES // per_row
    assert L.lib().inr_siren_workspace_bytes(C.byref(d), n_ok, 0) <= L.STATIC_SCRATCH_BYTES
    L.check(L.lib().inr_siren_forward_static(C.byref(d), L.ptr(params), None, n_ok, 0, n, 1.0, L.ptr(out),
                                             L.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.isfinite(out[:n_ok]).all()
