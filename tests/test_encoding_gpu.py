"""Encoded first layer (run.py:81-82,141-144 `num_freq`; models.py:14-59 PosEncodingNeRF) on the CUDA
path: the encoding kernels against the reference's outputs / the oracle, the network with in_features = 2*num_freq
against golden vectors made by the reference (tests/golden/encoding.npz), and run.train(num_freq=...).  fp32 mode."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ENC_CASES = {
    # name: (class, ctor kwargs, loss_mode) -- tests/golden/make_golden.py make_encoding
    "s": ("Siren", dict(in_features=16, hidden_features=48, hidden_layers=2, out_features=1, outermost_linear=True,
                        first_omega_0=30, hidden_omega_0=30.0), "mse"),
    "w": ("SirenWithSnakeTanh", dict(in_features=16, out_features=1, hidden_features=40, num_sine=1, num_snake=1,
                                     num_tanh=0, first_omega_0=30, hidden_omega_0=30.0, a_initial=0.5), "mse"),
    "l": ("SirenWithSnakeTanh", dict(in_features=16, out_features=1, hidden_features=32, num_sine=1, num_snake=0,
                                     num_tanh=1, first_linear=True, a_initial=2.0), "mae"),
}


@pytest.fixture(scope="module")
def pkg():
    import inr_for_audio_b200 as P
    assert torch.cuda.is_available()
    return P


def dev():
    return torch.device("cuda:0")


def test_pos_encoding_nerf_kernel(pkg, golden):
    """models.py:14-59 against the reference's own outputs: exact layout, values within fp32 sin/cos rounding."""
    g = golden("encoding")
    c1 = pkg.get_coord(1000, 1, device=dev())
    pe = pkg.PosEncodingNeRF(in_features=1, num_frequencies=6)
    assert pe.out_dim == 13
    out = pe(c1)
    assert out.shape == (1000, 13)
    assert np.array_equal(out[:, 0].cpu().numpy(), g["nerf1"][:, 0])
    np.testing.assert_allclose(out.cpu().numpy(), g["nerf1"], atol=2e-6)
    # coordinates generated inside the kernel == the coordinates passed in, bit for bit; shards concatenate
    assert torch.equal(pe(n=1000), out)
    assert torch.equal(torch.cat([pe(n=400, i0=0, T_total=1000), pe(n=600, i0=400, T_total=1000)]), out)
    pn = pkg.PosEncodingNeRF(in_features=1, fn_samples=4096)
    assert pn.num_frequencies == int(g["nerf1_nyq_nf"]) == 10
    np.testing.assert_allclose(pn(c1).cpu().numpy(), g["nerf1_nyq"], atol=2e-4)   # s^9 pi x ~ 1.6e3 rad in fp32
    c2 = pkg.get_coord(24, 2, device=dev())
    p2 = pkg.PosEncodingNeRF(in_features=2, num_frequencies=3, scale=1.5)
    np.testing.assert_allclose(p2(c2).cpu().numpy(), g["nerf2"], atol=2e-6)
    assert p2(c2.view(1, 576, 2)).shape == (1, 576, 14)


def test_gaussian_encoding_kernel(pkg):
    """rff.layers.GaussianEncoding (third-party, absent: parity unpinned) against the oracle's restatement."""
    from oracle import inr_oracle as O
    torch.manual_seed(3)
    enc = pkg.GaussianEncoding(10.0, 1, 64)
    assert enc.b.shape == (64, 1) and not enc.b.requires_grad
    torch.manual_seed(3)
    assert torch.equal(enc.b, torch.randn(64, 1) * 10.0)                     # rff.functional.sample_b
    with pytest.raises(Exception):
        enc(O.get_coord(8, 1))                                               # no CPU fallback
    b = enc.b.detach().clone()
    enc = enc.cuda()
    v = O.get_coord(5000, 1)
    ref = O.gaussian_encoding(v, b).numpy()
    out = enc(v.cuda())
    assert out.shape == (5000, 128)
    np.testing.assert_allclose(out.cpu().numpy(), ref, atol=3e-5)            # arguments up to ~2 pi 10 3 = 190 rad
    assert torch.equal(enc(n=5000), out)
    assert torch.equal(torch.cat([enc(n=2048, i0=0, T_total=5000), enc(n=2952, i0=2048, T_total=5000)]), out)
    assert enc(v.cuda().view(1, 5000, 1)).shape == (1, 5000, 128)
    # 2-D input
    torch.manual_seed(4)
    e2 = pkg.GaussianEncoding(2.0, 2, 16).cuda()
    v2 = O.get_coord(20, 2)
    np.testing.assert_allclose(e2(v2.cuda()).cpu().numpy(), O.gaussian_encoding(v2, e2.b.detach().cpu()).numpy(), atol=2e-5)
    with pytest.raises(ValueError):
        pkg.GaussianEncoding(1.0, 1, 4, b=torch.zeros(4, 1))


def _model(pkg, name):
    cls, kw, loss_mode = ENC_CASES[name]
    torch.manual_seed(11)
    return getattr(pkg, cls)(**kw), loss_mode


@pytest.mark.parametrize("name", sorted(ENC_CASES))
def test_encoded_network_against_reference(pkg, golden, name):
    """in_features = 16: seeded init bit-exact, forward, teacher-forced loss / gradients / Adam update against the
    reference's recorded tensors.  Tolerances as in test_snake_gpu (north star: per-step loss within 1e-3 relative)."""
    g = golden("encoding")
    m, loss_mode = _model(pkg, name)
    p0 = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).numpy()
    assert np.array_equal(p0, g[name + "_p0"])
    x = torch.from_numpy(g["x"]).to(dev())
    gt = torch.from_numpy(g["gt"]).to(dev())
    f = pkg.SirenFitter(m.spec, torch.from_numpy(g[name + "_p0"]), gt, x, loss_mode=loss_mode, use_graph=False)
    np.testing.assert_allclose(f.forward().cpu().numpy(), g[name + "_out0"], atol=1e-5)
    for k in (0, 1, 5):
        f = pkg.SirenFitter(m.spec, torch.from_numpy(g[f"{name}_p_{k}"]), gt, x, loss_mode=loss_mode, use_graph=False)
        loss, grads, out = f.loss_and_grads()
        ref_loss = float(g[name + "_losses"][k])
        assert abs(loss - ref_loss) <= 1e-5 * ref_loss
        gref = g[f"{name}_g_{k}"].astype(np.float64)
        gtol = 2e-2 if loss_mode == "mae" else 5e-5
        assert np.linalg.norm(grads.cpu().numpy() - gref) <= gtol * np.linalg.norm(gref)
        # the first-layer weight gradient (H x 16) on its own
        n0 = m.spec.hidden_features * 16
        assert np.linalg.norm(grads[:n0].cpu().numpy() - gref[:n0]) <= gtol * np.linalg.norm(gref[:n0]) + 1e-12
        if k > 0:
            f.m.copy_(torch.from_numpy(g[f"{name}_m_{k}"]).to(dev()))
            f.v.copy_(torch.from_numpy(g[f"{name}_v_{k}"]).to(dev()))
            f.set_opt(step=k)
        f.step(1)
        pref = g[f"{name}_pnext_{k}"]
        upd = np.abs(pref - g[f"{name}_p_{k}"]).max()
        assert np.abs(f.params.cpu().numpy() - pref).max() <= (0.1 if loss_mode == "mae" else 2e-3) * upd + 1e-9
    f = pkg.SirenFitter(m.spec, torch.from_numpy(g[name + "_p0"]), gt, x, loss_mode=loss_mode)
    f.step(6)
    np.testing.assert_allclose(f.losses()[:3], g[name + "_losses"][:3], rtol=1e-3)
    np.testing.assert_allclose(f.losses(), g[name + "_losses"], rtol=2e-2)


def test_module_route_and_fp16_refusal(pkg, golden):
    """model(encoded coords) + loss.backward() (run.py:156-187) with in_features = 16; the fp16 tcgen05 mode refuses
    encoded inputs loudly instead of falling back."""
    g = golden("encoding")
    m, _ = _model(pkg, "s")
    m.cuda()
    x = torch.from_numpy(g["x"]).view(1, 4096, 16).cuda()
    gt = torch.from_numpy(g["gt"]).view(1, -1, 1).cuda()
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    losses = []
    for _ in range(3):
        out = m(x)
        loss = torch.nn.functional.mse_loss(out, gt)
        losses.append(loss.item())
        opt.zero_grad()
        loss.backward()
        opt.step()
    np.testing.assert_allclose(losses, g["s_losses"][:3], rtol=2e-5)
    spec = pkg.SirenSpec(16, 64, 2, 1, True, 30.0, 30.0)            # H = 64 is not a tcgen05 width
    with pytest.raises(Exception, match="in_features"):
        pkg.SirenFitter(spec, torch.zeros(spec.num_params()), gt.view(-1), x.view(-1, 16), precision="fp16").step(1)


@pytest.mark.parametrize("K,H,ns,nk,fl,n", [(16, 128, 2, 0, False, 4099), (64, 256, 1, 1, False, 6000), (100, 128, 1, 0, True, 2000),
                                            (128, 128, 1, 0, False, 1000), (34, 512, 1, 0, False, 2500)])
@pytest.mark.parametrize("prec", ["fp16", "fp16s"])
def test_fp16_mode_wide_inputs_against_oracle(pkg, K, H, ns, nk, fl, n, prec):
    """Encoded coordinates in the tcgen05 mode: the rows become a zero-padded H-wide tile image, layer 0 one more layer GEMM,
    its weight gradient one more dW GEMM cut back to (H x in).  Loss within 1e-3 relative, gradients (and the first-layer
    weight gradient on its own) within fp16 rounding noise of the fp32 oracle."""
    from oracle import inr_oracle as O
    ospec = O.SirenSpec(K, H, ns, 1, True, 30.0, 30.0, num_snake=nk, num_tanh=0, first_linear=fl, a_initial=0.5)
    spec = pkg.SirenSpec(K, H, ns, 1, True, 30.0, 30.0, nk, 0, fl)
    ps = O.siren_init(ospec, seed=K + H)
    flat = torch.cat([p.reshape(-1) for p in ps])
    assert flat.numel() == spec.num_params()
    torch.manual_seed(K)
    x = torch.rand(n, K) * 2 - 1
    gt = torch.sin(torch.linspace(0, 20, n))
    loss_o, _, grads_o, out_o = O.siren_loss_and_grads(ospec, ps, x, gt.view(-1, 1))
    f = pkg.SirenFitter(spec, flat, gt.to(dev()), x.to(dev()), precision=prec, use_graph=False)
    loss, grads, out = f.loss_and_grads()
    oerr = float(np.abs(out.cpu().numpy() - out_o.numpy().ravel()).max())
    ltol, otol = (1e-3, 3e-2) if prec == "fp16" else (5e-5, 5e-4)     # fp16s: the input image X is split too
    assert abs(loss - float(loss_o)) <= ltol * abs(float(loss_o)) + 1e-9, (loss, float(loss_o), oerr)
    assert oerr <= otol, oerr
    go = torch.cat([t.reshape(-1) for t in grads_o]).numpy().astype(np.float64)
    gn = grads.cpu().numpy().astype(np.float64)
    gerr = np.linalg.norm(gn - go) / np.linalg.norm(go)
    n0 = H * K
    g0err = np.linalg.norm(gn[:n0] - go[:n0]) / np.linalg.norm(go[:n0])
    gb0err = np.linalg.norm(gn[n0:n0 + H] - go[n0:n0 + H]) / np.linalg.norm(go[n0:n0 + H])
    print("%s wide K=%d H=%d: loss rel err %.2e, out max err %.2e, grad rel err %.2e, gW0 %.2e, gb0 %.2e" % (
        prec, K, H, abs(loss - float(loss_o)) / float(loss_o), oerr, gerr, g0err, gb0err))
    assert gerr <= 1e-2 and g0err <= 1e-2 and gb0err <= 1e-2
    # forward-only entry and the graph-captured step
    np.testing.assert_allclose(f.forward().cpu().numpy(), out.cpu().numpy(), atol=1e-6)
    f2 = pkg.SirenFitter(spec, flat, gt.to(dev()), x.to(dev()), precision=prec)
    f2.step(5)
    ls = f2.losses()
    assert np.all(np.isfinite(ls)) and ls.min() < 2 * ls[0]


@pytest.mark.parametrize("K,H,n", [(3, 40, 1000), (64, 136, 4099), (130, 64, 257), (512, 256, 3000)])
def test_wide_inputs_ragged_against_oracle(pkg, K, H, n):
    """in_features not a multiple of the GEMM tile, rows ragged: loss / output / gradients against the oracle."""
    from oracle import inr_oracle as O
    ospec = O.SirenSpec(K, H, 1, 1, True, 30.0, 30.0)
    spec = pkg.SirenSpec(K, H, 1, 1, True, 30.0, 30.0)
    ps = O.siren_init(ospec, seed=K)
    flat = torch.cat([p.reshape(-1) for p in ps])
    assert flat.numel() == spec.num_params()
    torch.manual_seed(K + 1)
    x = torch.rand(n, K) * 2 - 1
    gt = torch.sin(torch.linspace(0, 20, n))
    loss_o, _, grads_o, out_o = O.siren_loss_and_grads(ospec, ps, x, gt.view(-1, 1))
    f = pkg.SirenFitter(spec, flat, gt.to(dev()), x.to(dev()), use_graph=False)
    loss, grads, out = f.loss_and_grads()
    assert abs(loss - float(loss_o)) <= 2e-5 * abs(float(loss_o)) + 1e-12
    np.testing.assert_allclose(out.cpu().numpy(), out_o.numpy().ravel(), atol=3e-5)
    go = torch.cat([t.reshape(-1) for t in grads_o]).numpy().astype(np.float64)
    assert np.linalg.norm(grads.cpu().numpy() - go) <= 1e-4 * np.linalg.norm(go) + 1e-12


def test_train_dropin_num_freq(pkg, tmp_path):
    """run.py:30 train(num_freq=...): rff encoding of in-kernel coordinates, in_features = 2*num_freq, checkpoint hand-off
    (prev_ckpt_path re-uses the stored b), bwe evaluation through the same encoding."""
    import json
    import os
    from scipy.io import wavfile
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    kw = dict(inst="gt_bach", duration=1, method="wave", num_hidden_features=64, num_sine=2, num_snake=0, num_tanh=0,
              num_freq=32, omega=30, hidden_omega=30, learning_rate=1e-3, data_dir=os.path.join(root, "data"), seed=0,
              verbose=False)
    ck1 = pkg.train(str(tmp_path), "rff", total_steps=150, **kw)
    c1 = torch.load(ck1, map_location="cpu")
    assert c1["model_state_dict"]["net.0.linear.weight"].shape == (64, 64)      # in_features = 2*num_freq
    assert c1["encoding_b"].shape == (32, 1)
    p1 = json.load(open(os.path.join(os.path.dirname(ck1), "parameters.json")))
    assert p1["num_freq"] == 32 and np.isfinite(p1["SNR"]) and p1["SNR"] > 3.0
    # coarse-to-fine hand-off (run.py:84-106,485-495): decimated fit resumed from the checkpoint, evaluated at the
    # original-rate grid through the same encoding (run.py:131,137,143-144,251-253)
    ck2 = pkg.train(str(tmp_path), "rff2", total_steps=50, prev_ckpt_path=ck1, decimation=2, bwe=True, **kw)
    c2 = torch.load(ck2, map_location="cpu")
    assert torch.equal(c2["encoding_b"], c1["encoding_b"])
    sr, rec = wavfile.read(os.path.join(os.path.dirname(ck2), "output.wav"))
    assert sr == 44100 and rec.shape == (44100,) and np.all(np.isfinite(rec))
    assert np.isfinite(json.load(open(os.path.join(os.path.dirname(ck2), "parameters.json")))["final_loss_dB"])
    with pytest.raises(Exception):                                  # H = 64 is not a tcgen05 width: loud refusal, no fallback
        pkg.train(str(tmp_path), "rff3", total_steps=1, precision="fp16", **kw)
    kw.update(num_hidden_features=128)
    ck3 = pkg.train(str(tmp_path), "rff4", total_steps=100, precision="fp16", **kw)   # 2*num_freq = 64 <= H = 128
    p3 = json.load(open(os.path.join(os.path.dirname(ck3), "parameters.json")))
    assert np.isfinite(p3["SNR"]) and p3["SNR"] > 3.0
