"""KAN fit path (csrc/kan.cu through the C ABI) against golden vectors made by the reference's kan.py and against the
oracle on seeded inputs.  Tolerances: b_splines bit-exact (same operation order as kan.py:93-104, incl. the exact
knots and the half-open edges); forward 1e-5 abs; per-step loss 1e-3 relative (observed ~1e-6); gradients 1e-4 of norm.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def pkg():
    import inr_for_audio_b200 as P
    assert torch.cuda.is_available()
    return P


def dev():
    return torch.device("cuda:0")


def _golden_model(pkg, g):
    torch.manual_seed(0)
    k = pkg.KAN([1, 16, 16, 1])
    sd = {key: torch.from_numpy(g["sd_" + key]) for key in g["sd_keys"]}
    k.load_state_dict(sd)      # exactly the reference's tensors (lstsq is not bit-reproducible run to run)
    return k.cuda()


def test_b_splines_bit_exact_with_reference(pkg, golden):
    g = golden("kan")
    k = _golden_model(pkg, g)
    x = torch.from_numpy(g["x"]).view(-1, 1).cuda()
    got = k.layers[0].b_splines(x).cpu().numpy()
    assert got.shape == g["bsp"].shape
    assert np.array_equal(got, g["bsp"])
    # zero outside [-2.2, 2.2): only the SiLU path contributes there (kan.py:94 half-open test)
    far = k.layers[0].b_splines(torch.tensor([[2.2000000477], [-2.3], [5.0]], device=dev()))
    assert torch.count_nonzero(far) == 0


def test_forward_and_gradients_match_reference(pkg, golden):
    g = golden("kan")
    k = _golden_model(pkg, g)
    x = torch.from_numpy(g["x"]).view(1, -1, 1).cuda()
    y = k(x)
    assert y.shape == (1, x.shape[1], 1)
    np.testing.assert_allclose(y.detach().cpu().numpy().ravel(), g["y"], atol=1e-5)
    loss = torch.nn.functional.mse_loss(k(x), torch.sin(3 * x))
    assert abs(loss.item() - float(g["loss"])) <= 1e-5 * float(g["loss"])
    loss.backward()
    for n, p in k.named_parameters():
        ref = g["grad_" + n]
        err = np.linalg.norm(p.grad.cpu().numpy() - ref) / max(np.linalg.norm(ref), 1e-30)
        assert err <= 1e-4, (n, err)


@pytest.mark.parametrize("widths,T", [([1, 256, 256, 1], 5000), ([1, 64, 1], 3001), ([1, 32, 48, 16, 1], 777), ([1, 8, 1], 1)])
def test_loss_and_grads_against_oracle(pkg, widths, T):
    from oracle import inr_oracle as O
    layers = O.kan_init(widths, seed=3)
    gt = pkg.utils.synth_waveform(max(T, 2), seed=2)[:T]
    coords = O.get_coord(T, 1) if T > 1 else torch.zeros(1, 1)
    ps = []
    for l in layers:
        for key in ("base_weight", "spline_weight", "spline_scaler"):
            l[key] = l[key].clone().requires_grad_(True)
            ps.append(l[key])
    out_o = O.kan_forward(coords.view(1, -1, 1), layers)
    lo = torch.mean((out_o.view(-1) - torch.from_numpy(gt)) ** 2)
    go = torch.cat([x.reshape(-1) for x in torch.autograd.grad(lo, ps)]).numpy().astype(np.float64)
    spec = pkg.KanSpec(tuple(widths))
    flat = torch.cat([p.detach().reshape(-1) for p in ps])
    grids = torch.cat([l["grid"].reshape(-1) for l in layers])
    f = pkg.KanFitter(spec, flat, grids, torch.from_numpy(gt).to(dev()), None if T > 1 else coords.to(dev()), use_graph=False)
    loss, grads, out = f.loss_and_grads()
    np.testing.assert_allclose(out.cpu().numpy(), out_o.detach().numpy().ravel(), atol=2e-5)
    assert abs(loss - float(lo)) <= 1e-4 * float(lo) + 1e-12, (loss, float(lo))
    gerr = np.linalg.norm(grads.cpu().numpy() - go) / np.linalg.norm(go)
    assert gerr <= 2e-4, gerr


def test_fit_steps_teacher_free_run(pkg):
    """First steps of the device loop (forward + loss + backward + Adam + plateau) against the oracle loop."""
    from oracle import inr_oracle as O
    widths, T = [1, 64, 64, 1], 4096
    layers = O.kan_init(widths, seed=1)
    gt = torch.from_numpy(pkg.utils.synth_waveform(T, seed=4))
    coords = O.get_coord(T, 1).view(1, -1, 1)
    names = ("base_weight", "spline_weight", "spline_scaler")
    ps = [l[k].clone() for l in layers for k in names]
    st = O.AdamState(ps, lr=1e-3)
    sched = O.PlateauState()
    ref_losses = []
    for step in range(5):
        req = [p.detach().clone().requires_grad_(True) for p in ps]
        ls = [dict(zip(names, req[3 * i:3 * i + 3]), grid=layers[i]["grid"]) for i in range(len(layers))]
        loss = torch.mean((O.kan_forward(coords, ls).view(-1) - gt) ** 2)
        grads = torch.autograd.grad(loss, req)
        ref_losses.append(float(loss))
        O.adam_step(ps, grads, st)
        sched.step(float(loss), st)
    flat = torch.cat([l[k].reshape(-1) for l in layers for k in names])
    grids = torch.cat([l["grid"].reshape(-1) for l in layers])
    f = pkg.KanFitter(pkg.KanSpec(tuple(widths)), flat, grids, gt.to(dev()))
    f.step(5)
    np.testing.assert_allclose(f.losses(), ref_losses, rtol=1e-3)
    np.testing.assert_allclose(f.params.cpu().numpy(), torch.cat([p.reshape(-1) for p in ps]).numpy(), atol=2e-5)
    f.step(45)
    assert f.losses()[-1] < f.losses()[0]


def test_kan_bad_arguments(pkg):
    from inr_for_audio_b200 import _lib as L
    with pytest.raises(L.InrLibraryError):
        pkg.KAN([1, 8, 1])(torch.zeros(1, 4, 1))            # model on the CPU: no fallback
    with pytest.raises(L.InrLibraryError):
        pkg.KanSpec((1, 8, 1), spline_order=2).desc() and L.check(
            L.lib().inr_kan_num_params(pkg.KanSpec((1, 8, 1), spline_order=2).desc()))


@pytest.mark.parametrize("widths,T", [([1, 256, 256, 1], 5000), ([1, 256, 256, 1], 257), ([1, 256, 256, 256, 1], 1000),
                                      ([1, 32, 256, 1], 300), ([1, 256, 1], 64)])
def test_bf16_tensor_core_mode_against_oracle(pkg, widths, T):
    """INR_PREC_BF16: the fused basis + contraction kernels (kan_tc.cuh).  bf16 operands, fp32 accumulation: loss within
    2e-3 relative, gradient within 3e-2 of its norm (the SIREN bf16 mode's bars), ragged row counts included."""
    from oracle import inr_oracle as O
    layers = O.kan_init(widths, seed=5)
    gt = pkg.utils.synth_waveform(max(T, 2), seed=2)[:T]
    coords = O.get_coord(T, 1)
    ps = []
    for l in layers:
        for key in ("base_weight", "spline_weight", "spline_scaler"):
            l[key] = l[key].clone().requires_grad_(True)
            ps.append(l[key])
    out_o = O.kan_forward(coords.view(1, -1, 1), layers)
    lo = torch.mean((out_o.view(-1) - torch.from_numpy(gt)) ** 2)
    go = torch.cat([x.reshape(-1) for x in torch.autograd.grad(lo, ps)]).numpy().astype(np.float64)
    spec = pkg.KanSpec(tuple(widths))
    flat = torch.cat([p.detach().reshape(-1) for p in ps])
    grids = torch.cat([l["grid"].reshape(-1) for l in layers])
    f = pkg.KanFitter(spec, flat, grids, torch.from_numpy(gt).to(dev()), precision="bf16", use_graph=False)
    loss, grads, out = f.loss_and_grads()
    np.testing.assert_allclose(out.cpu().numpy(), out_o.detach().numpy().ravel(), atol=5e-3)
    np.testing.assert_allclose(f.forward().cpu().numpy(), out.cpu().numpy(), atol=1e-6)       # inference path == training path
    assert abs(loss - float(lo)) <= 2e-3 * float(lo), (loss, float(lo))
    gerr = np.linalg.norm(grads.cpu().numpy() - go) / np.linalg.norm(go)
    assert gerr <= 3e-2, gerr
    # per-tensor check so a wrong (small) block cannot hide behind the norm of the big one
    o = 0
    for t in ps:
        n = t.numel()
        a, b = grads[o:o + n].cpu().numpy(), go[o:o + n]
        assert np.linalg.norm(a - b) <= 5e-2 * np.linalg.norm(b) + 1e-9, (o, np.linalg.norm(a - b) / np.linalg.norm(b))
        o += n
    # the graphed loop tracks the fp32 mode's trajectory over the first steps (before chaos separates them)
    f2 = pkg.KanFitter(spec, flat, grids, torch.from_numpy(gt).to(dev()), precision="bf16")
    f3 = pkg.KanFitter(spec, flat, grids, torch.from_numpy(gt).to(dev()), precision="fp32")
    f2.step(30)
    f3.step(30)
    ls = f2.losses()
    assert np.all(np.isfinite(ls))
    np.testing.assert_allclose(ls[:4], f3.losses()[:4], rtol=1e-2)
    assert abs(ls[-1] - f3.losses()[-1]) <= 0.1 * f3.losses()[-1]


def test_bf16_needs_uniform_grid(pkg):
    from inr_for_audio_b200 import _lib as L
    torch.manual_seed(0)
    k = pkg.KAN([1, 256, 256, 1])
    grids = k.flat_grids()
    grids[5] += 0.05
    with pytest.raises(L.InrLibraryError):
        pkg.KanFitter(k.spec, k.flat_parameters(), grids, torch.zeros(100, device=dev()), precision="bf16")
