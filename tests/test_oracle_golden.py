"""The oracle (oracle/inr_oracle.py) against the golden vectors produced by the reference itself
(tests/golden/make_golden.py).  CPU only."""
import hashlib

import numpy as np
import pytest
import torch

from oracle import inr_oracle as O


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_linspace_bit_exact(golden):
    g = golden("coords")
    for T in g["Ts"]:
        T = int(T)
        c = O.linspace_f32(T)
        assert sha(c) == str(g[f"sha_{T}"]), T
        assert np.array_equal(c[g[f"idx_{T}"]].view(np.uint32), g[f"bits_{T}"])
    assert np.array_equal(O.get_coord(5, 2).numpy(), g["coord2d_5"])
    assert np.array_equal(O.get_coord(17, 1, 100).numpy(), g["coord_scale100_17"])


def _split(spec, flat):
    ps, o = [], 0
    for fin, fout in spec.layer_dims():
        ps.append(torch.from_numpy(flat[o:o + fin * fout].copy()).view(fout, fin)); o += fin * fout
        ps.append(torch.from_numpy(flat[o:o + fout].copy())); o += fout
    assert o == flat.size
    return ps


def _flat(ps):
    return np.concatenate([p.detach().numpy().ravel() for p in ps])


def test_siren_init_matches_reference(golden):
    g = golden("siren")
    spec = O.SirenSpec(1, 64, 2, 1, True)
    assert np.array_equal(_flat(O.siren_init(spec, seed=0)), g["small_p0"])
    h = hashlib.sha256()
    spec = O.SirenSpec(1, 256, 3, 1, True)
    ps = O.siren_init(spec, seed=0)
    for k, p in zip(g["sd_keys_256x3"], ps):
        h.update(str(k).encode()); h.update(p.numpy().tobytes())
    assert h.hexdigest() == str(g["sd_hash_256x3_seed0"]) == str(g["sd_hash_swst_256x3_seed0"])
    assert spec.num_params() == int(g["nparams_256x3"]) == 198145
    assert O.SirenSpec(1, 512, 5, 1, True).num_params() == int(g["nparams_512x5"])
    spec3 = O.SirenSpec(1, 64, 2, 1, True, 22000, 30)
    assert np.array_equal(_flat(O.siren_init(spec3, seed=4)), g["m3_p0"])
    spec2 = O.SirenSpec(1, 32, 1, 1, False, 3000, 20.0)
    assert np.array_equal(_flat(O.siren_init(spec2, seed=3)), g["m2_p0"])


def test_siren_forward_and_trajectory(golden):
    g = golden("siren")
    torch.set_num_threads(8)
    spec = O.SirenSpec(1, 64, 2, 1, True)
    coords = O.get_coord(4096, 1).view(1, -1, 1)
    gt = torch.from_numpy(g["small_gt"]).view(1, -1, 1)
    ps = _split(spec, g["small_p0"])
    rec = {}

    def record(step, lv, params, grads, st):
        rec[step] = (_flat(params), _flat(grads))
    losses, st, sched = O.siren_fit(spec, ps, coords, gt, 12, record=record)
    # free-running for the first 3 steps, then teacher-forced checks below
    np.testing.assert_allclose(losses[:3], g["small_losses"][:3], rtol=2e-6)
    np.testing.assert_allclose(rec[0][1], g["small_g_0"], rtol=1e-4, atol=1e-9)
    # teacher-forced at steps 2, 5, 10: reference params + Adam state in -> loss, grads, params' out
    for k in (2, 5, 10):
        ps = _split(spec, g[f"small_p_{k}"])
        loss, _, grads, out = O.siren_loss_and_grads(spec, ps, coords, gt)
        assert abs(float(loss) - g["small_losses"][k]) <= 2e-6 * g["small_losses"][k]
        np.testing.assert_allclose(out.numpy().ravel(), g[f"small_out_{k}"], atol=2e-6)
        gref = g[f"small_g_{k}"]
        assert np.linalg.norm(_flat(grads) - gref) <= 1e-5 * np.linalg.norm(gref)
        a = O.AdamState(ps)
        a.t = k
        a.m = _split(spec, g[f"small_m_{k}"]); a.v = _split(spec, g[f"small_v_{k}"])
        O.adam_step(ps, _split(spec, gref), a)
        np.testing.assert_allclose(_flat(ps), g[f"small_pnext_{k}"], rtol=0, atol=1e-7)


def test_siren_variants(golden):
    g = golden("siren")
    coords = O.get_coord(4096, 1).view(1, -1, 1)
    gt = torch.from_numpy(g["small_gt"]).view(1, -1, 1)
    spec2 = O.SirenSpec(1, 32, 1, 1, False, 3000, 20.0)
    out = O.siren_forward(spec2, _split(spec2, g["m2_p0"]), coords)
    np.testing.assert_allclose(out.numpy().ravel(), g["m2_out"], atol=2e-6)
    spec3 = O.SirenSpec(1, 64, 2, 1, True, 22000, 30)
    ps = _split(spec3, g["m3_p0"])
    np.testing.assert_allclose(O.siren_forward(spec3, ps, coords).numpy().ravel(), g["m3_out"], atol=2e-6)
    loss, _, grads, _ = O.siren_loss_and_grads(spec3, ps, coords, gt, "mae")
    assert abs(float(loss) - g["m3_mae_losses"][0]) < 1e-6
    assert np.linalg.norm(_flat(grads) - g["m3_mae_g0"]) <= 1e-4 * np.linalg.norm(g["m3_mae_g0"])


def test_adam_and_plateau_against_torch():
    torch.manual_seed(1)
    p0 = [torch.randn(7, 5), torch.randn(5)]
    ps = [p.clone() for p in p0]
    tp = [torch.nn.Parameter(p.clone()) for p in p0]
    opt = torch.optim.Adam(tp, lr=1e-2)
    sch = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode="min", factor=0.8, patience=3, min_lr=5e-3)
    st = O.AdamState(ps, lr=1e-2)
    sc = O.PlateauState(patience=3, min_lr=5e-3)
    for it in range(40):
        gs = [torch.randn_like(p) for p in ps]
        for p, g_ in zip(tp, gs):
            p.grad = g_.clone()
        opt.step()
        O.adam_step(ps, gs, st)
        loss = 1.0 if it % 7 else 1.0 / (it + 1)
        sch.step(loss)
        sc.step(loss, st)
        assert st.lr == opt.param_groups[0]["lr"]
        for a, b in zip(ps, tp):
            assert torch.equal(a, b.detach())


def test_kbd_mdct_stmdct(golden):
    g = golden("mdct")
    np.testing.assert_array_equal(O.kbd_window(2048), g["kbd_2048"])
    np.testing.assert_array_equal(O.kbd_window(256), g["kbd_256"])
    w = O.kbd_window(2048)
    assert np.max(np.abs(w[:1024] ** 2 + w[1024:] ** 2 - 1)) < 1e-14  # Princen-Bradley
    np.testing.assert_array_equal(O.mdct_fft(g["frame"], 1024, 1024), g["mdct_frame"])
    np.testing.assert_allclose(O.mdct_direct(g["frame"], 1024, 1024), g["mdct_frame"], atol=1e-12)
    np.testing.assert_allclose(O.imdct_direct(g["mdct_frame"], 1024, 1024), g["imdct_frame"], atol=1e-11)
    np.testing.assert_array_equal(O.stmdct(g["sig"], 2048), g["stmdct_2048"])
    np.testing.assert_array_equal(O.stmdct(g["sig"][:5000], 256), g["stmdct_256"])
    np.testing.assert_array_equal(O.stmdct(g["sig"][:5120], 256), g["stmdct_256_div"])
    np.testing.assert_allclose(O.istmdct(g["stmdct_256"], 256), g["istmdct_256"], atol=1e-12)
    for L, F_ in zip(g["frame_lens"], g["frame_counts_2048"]):
        assert O.stmdct_frames(int(L), 2048)[2] == int(F_)
    assert O.stmdct_frames(1323000, 2048) == (1024, 8, 1291)
    # differentiable restatement == numpy STMDCT
    x = torch.from_numpy(g["sig"])
    np.testing.assert_allclose(O.stmdct_torch(x, 2048).numpy(), g["stmdct_2048"], atol=1e-13)
    # TDAC round trip (interior)
    rec = O.istmdct(O.stmdct(g["sig"], 256), 256)
    assert np.max(np.abs(rec[128:9856] - g["sig"][128:9856])) < 1e-12
    # 2-D coordinate grid of MDCTFitting
    H, W = [int(v) for v in g["mf_shape"]]
    c = O.get_coord_2d(H, W).numpy()
    assert sha(c) == str(g["mf_coords_sha"])


def test_kan(golden):
    g = golden("kan")
    layers = O.kan_init([1, 16, 16, 1], seed=0)
    for i, l in enumerate(layers):
        for k in ("base_weight", "spline_weight", "spline_scaler", "grid"):
            # torch.linalg.lstsq (gelsy) is not run-to-run deterministic at the 1e-8 level even inside
            # the reference itself, so spline_weight is pinned to 1e-7 abs; everything else bit-exact.
            np.testing.assert_allclose(l[k].numpy(), g[f"sd_layers.{i}.{k}"], rtol=0,
                                       atol=1e-7 if k == "spline_weight" else 0)
    x = torch.from_numpy(g["x"]).view(1, -1, 1)
    np.testing.assert_allclose(O.b_splines(x.view(-1, 1), layers[0]["grid"]).numpy(), g["bsp"], atol=0)
    y = O.kan_forward(x, layers)
    np.testing.assert_allclose(y.numpy().ravel(), g["y"], atol=1e-6)
    b = O.b_splines(torch.linspace(-1, 1, 101).view(-1, 1), layers[0]["grid"])
    assert torch.allclose(b.sum(-1), torch.ones(101, 1), atol=1e-6)          # partition of unity on [-1,1]
    assert torch.count_nonzero(O.b_splines(torch.tensor([[2.2], [-2.3], [5.0]]), layers[0]["grid"])) == 0
    ps = []
    for l in layers:
        for k in ("base_weight", "spline_weight", "spline_scaler"):
            l[k] = l[k].clone().requires_grad_(True); ps.append(l[k])
    loss = torch.nn.functional.mse_loss(O.kan_forward(x, layers), torch.sin(3 * x))
    assert abs(loss.item() - float(g["loss"])) < 1e-6
    grads = torch.autograd.grad(loss, ps)
    for n, gr in zip(g["pnames"], grads):
        np.testing.assert_allclose(gr.numpy(), g["grad_" + str(n)], rtol=1e-4, atol=1e-7)


def test_stft_snr_losses_selfconsistent():
    """auraloss is absent: parity UNPINNED.  Only self-consistency of the restatement is checked."""
    torch.manual_seed(0)
    x = torch.randn(1, 4096, 1) * 0.1
    assert float(O.stft_loss(x, x)) == pytest.approx(0.0, abs=1e-6)
    y = x + 0.01 * torch.randn_like(x)
    assert float(O.stft_loss(y, x)) > 0
    assert O.stft_mag(x.view(1, -1)).shape == (1, 513, 17)
    s = float(O.snr_loss(y.view(1, 1, -1), x.view(1, 1, -1)))
    assert -25 < s < -15


SNAKE_SPECS = {
    # the oracle's SirenSpec for the cases of tests/golden/make_golden.py make_snake
    "a": (O.SirenSpec(1, 64, 1, 1, True, 30.0, 30.0, num_snake=2, num_tanh=1, a_initial=0.5), "mse"),
    "b": (O.SirenSpec(1, 40, 2, 1, True, 22000.0, 30.0, num_snake=2, num_tanh=0, a_initial=0.5), "mse"),
    "c": (O.SirenSpec(1, 32, 0, 1, False, 30.0, 30.0, num_snake=1, num_tanh=2, first_linear=True, a_initial=50.0), "mae"),
    "d": (O.SirenSpec(1, 24, 0, 1, True, 30.0, 30.0, num_snake=1, num_tanh=1, a_initial=2.0), "mse"),
}


def _split_sizes(flat, sizes_like):
    out, o = [], 0
    for t in sizes_like:
        n = t.numel()
        out.append(torch.from_numpy(np.asarray(flat[o:o + n])).view(t.shape).clone())
        o += n
    assert o == len(flat)
    return out


@pytest.mark.parametrize("name", sorted(SNAKE_SPECS))
def test_snake_tanh_stacks_against_reference(golden, name):
    """models.py:185-241,306-394: seeded init (bit-exact), forward, per-step loss + gradients (teacher-forced from the
    reference's recorded parameters) and the first free-running losses, against outputs of the reference itself."""
    g = golden("snake")
    spec, loss_mode = SNAKE_SPECS[name]
    ps = O.siren_init(spec, seed=7)
    assert np.array_equal(_flat(ps), g[name + "_p0"])
    assert spec.num_params() == g[name + "_p0"].size
    coords = O.get_coord(4096, 1).view(1, -1, 1)
    gt = torch.from_numpy(g["gt"]).view(1, -1, 1)
    tol = 5e-4 if spec.first_omega_0 > 1000 else 2e-5       # omega=22000: sin argument ~4e4 rad, see test_siren_variants
    np.testing.assert_allclose(O.siren_forward(spec, ps, coords).numpy().ravel(), g[name + "_out0"], atol=tol)
    for k in (0, 1, 5):
        pk = _split_sizes(g[f"{name}_p_{k}"], ps)
        loss, _, grads, out = O.siren_loss_and_grads(spec, pk, coords, gt, loss_mode)
        assert abs(float(loss) - g[name + "_losses"][k]) <= 1e-5 * g[name + "_losses"][k]
        gref = g[f"{name}_g_{k}"].astype(np.float64)
        assert np.linalg.norm(_flat(grads) - gref) <= (1e-3 if loss_mode == "mae" else 1e-4) * np.linalg.norm(gref)
    losses, _, _ = O.siren_fit(spec, ps, coords, gt, 3, loss_mode=loss_mode)
    np.testing.assert_allclose(losses, g[name + "_losses"][:3], rtol=1e-4)


# ------------------------------------------------------------------ encoded first layer
ENC_SPECS = {
    # the oracle's SirenSpec for the cases of tests/golden/make_golden.py make_encoding (in_features = 16 = 2*num_freq)
    "s": (O.SirenSpec(16, 48, 2, 1, True, 30.0, 30.0), "mse"),
    "w": (O.SirenSpec(16, 40, 1, 1, True, 30.0, 30.0, num_snake=1, num_tanh=0, a_initial=0.5), "mse"),
    "l": (O.SirenSpec(16, 32, 1, 1, True, 30.0, 30.0, num_snake=0, num_tanh=1, first_linear=True, a_initial=2.0), "mae"),
}


def test_pos_encoding_nerf_against_reference(golden):
    """models.py:14-59: layout [x, sin, cos per frequency and feature], the nyquist frequency-count rule."""
    g = golden("encoding")
    c1 = O.get_coord(1000, 1)
    assert np.array_equal(O.pos_encoding_nerf(c1, 6).numpy(), g["nerf1"])
    assert int(g["nerf1_nyq_nf"]) == 10
    assert np.array_equal(O.pos_encoding_nerf(c1, 10).numpy(), g["nerf1_nyq"])
    c2 = O.get_coord(24, 2)
    assert np.array_equal(O.pos_encoding_nerf(c2, 3, scale=1.5).numpy(), g["nerf2"])


def test_gaussian_encoding_definition():
    """rff GaussianEncoding (third-party, absent: parity unpinned): shape, cos|sin halves, unit modulus, b = 0 -> (1, 0)."""
    torch.manual_seed(0)
    b = torch.randn(8, 1) * 10.0
    v = O.get_coord(100, 1)
    e = O.gaussian_encoding(v, b)
    assert e.shape == (100, 16)
    np.testing.assert_allclose((e[:, :8] ** 2 + e[:, 8:] ** 2).numpy(), 1.0, atol=1e-6)
    vp = (2 * np.pi * v.double()) @ b.double().T
    np.testing.assert_allclose(e[:, :8].numpy(), np.cos(vp.numpy()), atol=2e-5)
    np.testing.assert_allclose(e[:, 8:].numpy(), np.sin(vp.numpy()), atol=2e-5)
    z = O.gaussian_encoding(v, torch.zeros(3, 1))
    assert torch.equal(z[:, :3], torch.ones(100, 3)) and torch.equal(z[:, 3:], torch.zeros(100, 3))


@pytest.mark.parametrize("name", sorted(ENC_SPECS))
def test_encoded_first_layer_against_reference(golden, name):
    """run.py:81-82: the network on an encoded coordinate (in_features = 2*num_freq): seeded init, forward, teacher-forced
    loss + gradients, first free-running losses, against outputs of the reference itself."""
    g = golden("encoding")
    spec, loss_mode = ENC_SPECS[name]
    ps = O.siren_init(spec, seed=11)
    assert np.array_equal(_flat(ps), g[name + "_p0"])
    x = torch.from_numpy(g["x"]).view(1, 4096, 16)
    gt = torch.from_numpy(g["gt"]).view(1, -1, 1)
    np.testing.assert_allclose(O.siren_forward(spec, ps, x).numpy().ravel(), g[name + "_out0"], atol=2e-5)
    for k in (0, 1, 5):
        pk = _split_sizes(g[f"{name}_p_{k}"], ps)
        loss, _, grads, out = O.siren_loss_and_grads(spec, pk, x, gt, loss_mode)
        assert abs(float(loss) - g[name + "_losses"][k]) <= 1e-5 * g[name + "_losses"][k]
        gref = g[f"{name}_g_{k}"].astype(np.float64)
        assert np.linalg.norm(_flat(grads) - gref) <= (1e-3 if loss_mode == "mae" else 1e-4) * np.linalg.norm(gref)
    losses, _, _ = O.siren_fit(spec, ps, x, gt, 3, loss_mode=loss_mode)
    np.testing.assert_allclose(losses, g[name + "_losses"][:3], rtol=1e-4)
