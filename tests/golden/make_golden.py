"""Manufacture the golden vectors under tests/golden/ by RUNNING THE REFERENCE ITSELF.

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
The reference has no tests or fixtures of its own, so every pin is an output of the
reference's unmodified modules (imported through oracle/ref_import.py), of torch.optim for the optimiser,
plus gt_bach.wav.  The oracle (oracle/inr_oracle.py) and the CUDA path are both tested against these.
"""
import hashlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402

ref = ref_import.load()
torch.set_num_threads(8)


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def sd_hash(sd) -> str:
    h = hashlib.sha256()
    for k, v in sd.items():
        h.update(k.encode())
        h.update(np.ascontiguousarray(v.detach().cpu().numpy()).tobytes())
    return h.hexdigest()


# ---------------------------------------------------------------- coords (utils.py:99-109)
def make_coords():
    out = {}
    Ts = [2, 3, 1001, 4096, 88200, 264600, 441000, 1323000, 2646000]
    out["Ts"] = np.array(Ts)
    for T in Ts:
        c = ref.utils.get_coord(T, 1).numpy().reshape(-1)
        out[f"sha_{T}"] = np.array(sha(c))
        idx = np.unique(np.clip(np.concatenate([np.arange(8), T // 2 + np.arange(-4, 4), T - 1 - np.arange(8)]), 0, T - 1))
        out[f"idx_{T}"] = idx
        out[f"bits_{T}"] = c[idx].view(np.uint32)
    c2 = ref.utils.get_coord(5, 2).numpy()
    out["coord2d_5"] = c2
    out["coord_scale100_17"] = ref.utils.get_coord(17, 1, 100).numpy()
    np.savez_compressed(os.path.join(HERE, "coords.npz"), **out)



# ---------------------------------------------------------------- Snake / Tanh stacks (models.py:185-241,306-394)
def make_snake():
    """SirenWithSnakeTanh with Snake / Tanh layers -- run.py's actual default architecture (run.py:30: num_sine=2,
    num_snake=2, a_initial=0.5).  Small widths, full tensors: init, forward, teacher-forced steps, free-run losses."""
    out = {}
    sr, data = gt_bach(4096, 44100)
    gt = torch.from_numpy(data / np.max(np.abs(data))).view(1, -1, 1)
    coords = ref.utils.get_coord(4096, 1).view(1, -1, 1)
    out["gt"] = gt.numpy().ravel()
    cases = {
        # name: (ctor kwargs, loss_mode)
        "a": (dict(in_features=1, out_features=1, hidden_features=64, num_sine=1, num_snake=2, num_tanh=1,
                   first_omega_0=30, hidden_omega_0=30.0, a_initial=0.5), "mse"),
        # run.py's defaults (run.py:30,112-113) at a small width: omega=22000, two sine + two snake layers
        "b": (dict(in_features=1, out_features=1, hidden_features=40, num_sine=2, num_snake=2, num_tanh=0,
                   first_omega_0=22000, hidden_omega_0=30, a_initial=0.5), "mse"),
        # first_linear (Linear+Snake as layer 0), class-default a_initial=50, final SineLayer, mae loss
        "c": (dict(in_features=1, out_features=1, hidden_features=32, num_sine=0, num_snake=1, num_tanh=2,
                   first_linear=True, last_linear=False, hidden_omega_0=30.0), "mae"),
        # the last hidden layer is a Tanh layer; no sine layer besides the first
        "d": (dict(in_features=1, out_features=1, hidden_features=24, num_sine=0, num_snake=1, num_tanh=1,
                   a_initial=2.0), "mse"),
    }
    for name, (kw, loss_mode) in cases.items():
        torch.manual_seed(7)
        m = ref.models.SirenWithSnakeTanh(**kw)
        out[name + "_keys"] = np.array(list(m.state_dict().keys()))
        out[name + "_p0"] = np.concatenate([p.detach().numpy().ravel() for p in m.parameters()])
        out[name + "_out0"] = m(coords).detach().numpy().ravel()
        losses, lrs, rec = ref_fit(m, coords, gt, 8, loss_mode=loss_mode, record_steps=(0, 1, 5))
        out[name + "_losses"] = np.array(losses, dtype=np.float64)
        for k, v in rec.items():
            out[name + "_" + k] = v
    np.savez_compressed(os.path.join(HERE, "snake.npz"), **out)


def make_encoding():
    """PosEncodingNeRF (models.py:14-59) outputs, and a reference Siren whose first layer takes an encoded coordinate
    (in_features = 2*num_freq as run.py:81-82 builds it; the encoding itself comes from the absent third-party `rff`, so
    the encoded input here is PosEncodingNeRF's, which the reference does carry): forward + teacher-forced steps."""
    out = {}
    c1 = ref.utils.get_coord(1000, 1)
    pe1 = ref.models.PosEncodingNeRF(in_features=1, num_frequencies=6)
    out["nerf1"] = pe1(c1).numpy()
    pe1n = ref.models.PosEncodingNeRF(in_features=1, fn_samples=4096)          # nyquist rule: floor(log2(4096/4)) = 10
    out["nerf1_nyq_nf"] = np.array(pe1n.num_frequencies)
    out["nerf1_nyq"] = pe1n(c1).numpy()
    c2 = ref.utils.get_coord(24, 2)
    pe2 = ref.models.PosEncodingNeRF(in_features=2, num_frequencies=3, scale=1.5)
    out["nerf2"] = pe2(c2).numpy()
    sr, data = gt_bach(4096, 44100)
    gt = torch.from_numpy(data / np.max(np.abs(data))).view(1, -1, 1)
    out["gt"] = gt.numpy().ravel()
    enc = ref.models.PosEncodingNeRF(in_features=1, num_frequencies=8)(ref.utils.get_coord(4096, 1))   # (4096, 17)
    x = enc[:, 1:].contiguous().view(1, 4096, 16)                                                 # 16 = 2*num_freq inputs
    out["x"] = x.numpy().reshape(4096, 16)
    cases = {
        "s": (lambda: ref.models.Siren(16, 48, 2, 1, outermost_linear=True, first_omega_0=30, hidden_omega_0=30.0), "mse"),
        "w": (lambda: ref.models.SirenWithSnakeTanh(in_features=16, out_features=1, hidden_features=40, num_sine=1,
                                                     num_snake=1, num_tanh=0, first_omega_0=30, hidden_omega_0=30.0,
                                                     a_initial=0.5), "mse"),
        "l": (lambda: ref.models.SirenWithSnakeTanh(in_features=16, out_features=1, hidden_features=32, num_sine=1,
                                                     num_snake=0, num_tanh=1, first_linear=True, a_initial=2.0), "mae"),
    }
    for name, (ctor, loss_mode) in cases.items():
        torch.manual_seed(11)
        m = ctor()
        out[name + "_p0"] = np.concatenate([p.detach().numpy().ravel() for p in m.parameters()])
        out[name + "_out0"] = m(x).detach().numpy().ravel()
        losses, lrs, rec = ref_fit(m, x, gt, 6, loss_mode=loss_mode, record_steps=(0, 1, 5))
        out[name + "_losses"] = np.array(losses, dtype=np.float64)
        for k, v in rec.items():
            out[name + "_" + k] = v
    np.savez_compressed(os.path.join(HERE, "encoding.npz"), **out)


# ---------------------------------------------------------------- SIREN (models.py)
def gt_bach(n=None, start=0):
    from scipy.io import wavfile
    sr, data = wavfile.read(os.path.join(ROOT, "data", "gt_bach.wav"))
    data = data.astype(np.float32)
    if n is not None:
        data = data[start:start + n]
    return sr, data


def ref_fit(model, coords, gt, steps, lr=1e-3, min_lr=1e-6, loss_mode="mse", record_steps=()):
    """run.py:116-117,124-125,156-190 restated around the REFERENCE model + torch.optim (alpha=0)."""
    opt = torch.optim.Adam(model.parameters(), lr=lr)
    sched = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode="min", factor=0.8, patience=200, min_lr=min_lr)
    mae, mse = torch.nn.L1Loss(), torch.nn.MSELoss()
    losses, lrs, rec = [], [], {}
    for step in range(steps):
        out = model(coords)
        loss = mae(out, gt) if loss_mode == "mae" else mse(out, gt)
        losses.append(loss.item())
        opt.zero_grad()
        loss.backward()
        if step in record_steps:
            rec[f"p_{step}"] = np.concatenate([p.detach().numpy().ravel() for p in model.parameters()])
            rec[f"g_{step}"] = np.concatenate([p.grad.detach().numpy().ravel() for p in model.parameters()])
            st = opt.state
            if len(st):
                rec[f"m_{step}"] = np.concatenate([st[p]["exp_avg"].numpy().ravel() for p in model.parameters()])
                rec[f"v_{step}"] = np.concatenate([st[p]["exp_avg_sq"].numpy().ravel() for p in model.parameters()])
            rec[f"out_{step}"] = out.detach().numpy().ravel().copy()
        opt.step()
        sched.step(loss)
        lrs.append(sched.get_last_lr()[0])
        if step in record_steps:
            rec[f"pnext_{step}"] = np.concatenate([p.detach().numpy().ravel() for p in model.parameters()])
    return losses, lrs, rec


def make_siren():
    out = {}
    # structural pins
    torch.manual_seed(0)
    big = ref.models.Siren(1, 256, 3, 1, True)
    out["sd_keys_256x3"] = np.array(list(big.state_dict().keys()))
    out["sd_hash_256x3_seed0"] = np.array(sd_hash(big.state_dict()))
    torch.manual_seed(0)
    big2 = ref.models.SirenWithSnakeTanh(1, 1, 256, 3, 0, 0)
    out["sd_hash_swst_256x3_seed0"] = np.array(sd_hash(big2.state_dict()))
    torch.manual_seed(0)
    c2 = ref.models.Siren(1, 512, 5, 1, True)
    out["sd_hash_512x5_seed0"] = np.array(sd_hash(c2.state_dict()))
    out["nparams_256x3"] = np.array(sum(p.numel() for p in big.parameters()))
    out["nparams_512x5"] = np.array(sum(p.numel() for p in c2.parameters()))

    # small model: full tensors.  H=64, L=2, T=4096 samples of gt_bach starting at 1 s.
    sr, data = gt_bach(4096, 44100)
    gt = torch.from_numpy(data / np.max(np.abs(data))).view(1, -1, 1)
    coords = ref.utils.get_coord(4096, 1).view(1, -1, 1)
    torch.manual_seed(0)
    m = ref.models.Siren(1, 64, 2, 1, True, first_omega_0=30, hidden_omega_0=30.0)
    out["small_p0"] = np.concatenate([p.detach().numpy().ravel() for p in m.parameters()])
    out["small_gt"] = gt.numpy().ravel()
    losses, lrs, rec = ref_fit(m, coords, gt, 12, record_steps=(0, 1, 2, 5, 10))
    out["small_losses"] = np.array(losses, dtype=np.float64)
    for k, v in rec.items():
        out["small_" + k] = v

    # small model, non-default omegas, mae loss, not outermost-linear checked through forward only
    torch.manual_seed(3)
    m2 = ref.models.Siren(1, 32, 1, 1, False, first_omega_0=3000, hidden_omega_0=20.0)
    out["m2_p0"] = np.concatenate([p.detach().numpy().ravel() for p in m2.parameters()])
    out["m2_out"] = m2(coords).detach().numpy().ravel()
    torch.manual_seed(4)
    m3 = ref.models.SirenWithSnakeTanh(1, 1, 64, 2, 0, 0, first_omega_0=22000, hidden_omega_0=30)
    out["m3_p0"] = np.concatenate([p.detach().numpy().ravel() for p in m3.parameters()])
    out["m3_out"] = m3(coords).detach().numpy().ravel()
    losses, _, rec = ref_fit(m3, coords, gt, 2, loss_mode="mae", record_steps=(0,))
    out["m3_mae_losses"] = np.array(losses)
    out["m3_mae_g0"] = rec["g_0"]

    # C1: Siren 3x256 on gt_bach 6 s, T = 264 600: losses at steps 0..2 and the step-0 gradient norm
    sr, data = gt_bach()
    data = data[:6 * sr]
    gt = torch.from_numpy(data / np.max(np.abs(data))).view(1, -1, 1)
    coords = ref.utils.get_coord(len(data), 1).view(1, -1, 1)
    torch.manual_seed(0)
    m = ref.models.Siren(1, 256, 3, 1, True)
    losses, _, rec = ref_fit(m, coords, gt, 3, record_steps=(0,))
    out["c1_T"] = np.array(len(data))
    out["c1_losses"] = np.array(losses, dtype=np.float64)
    out["c1_g0_norm"] = np.array(np.linalg.norm(rec["g_0"].astype(np.float64)))
    out["c1_g0_head"] = rec["g_0"][:512]
    out["c1_g0_tail"] = rec["g_0"][-257:]
    out["c1_out0_head"] = rec["out_0"][:256]
    np.savez_compressed(os.path.join(HERE, "siren.npz"), **out)


def make_converge():
    """Converged-SNR pins with the reference's own run-to-run (thread-count) spread."""
    out = {}
    sr, data = gt_bach(8192, 44100)
    gtn = data / np.max(np.abs(data))
    gt = torch.from_numpy(gtn).view(1, -1, 1)
    coords = ref.utils.get_coord(8192, 1).view(1, -1, 1)
    snrs = {}
    for nt in (1, 8):
        torch.set_num_threads(nt)
        torch.manual_seed(0)
        m = ref.models.Siren(1, 128, 2, 1, True)
        steps = 1500
        losses, lrs, _ = ref_fit(m, coords, gt, steps)
        y = m(coords).detach().numpy().ravel()
        snrs[nt] = ref.utils.calculate_snr(gtn, y)
        out[f"losses_t{nt}"] = np.array(losses)[[0, 1, 2, 10, 100, 500, 1000, 1499]]
        out[f"all_losses_t{nt}"] = np.array(losses, dtype=np.float32)  # Adam at lr 1e-3 spikes: keep the whole curve
        out[f"lr_final_t{nt}"] = np.array(lrs[-1])
    torch.set_num_threads(8)
    out["steps"] = np.array(1500)
    out["snr_t1"] = np.array(snrs[1])
    out["snr_t8"] = np.array(snrs[8])
    print("converged SNR (1 thread / 8 threads):", snrs)
    np.savez_compressed(os.path.join(HERE, "converge.npz"), **out)


def make_converge_ens():
    """The reference's own run-to-run distribution of the running-minimum loss envelope: the same seed-0 fit as
    make_converge, run with 1..8 intra-op threads (each thread count changes only the summation order of the
    reference's GEMM/reductions, which is enough to decorrelate the trajectories after ~20 steps)."""
    sr, data = gt_bach(8192, 44100)
    gtn = data / np.max(np.abs(data))
    sig = float(np.mean(gtn.astype(np.float64) ** 2))
    gt = torch.from_numpy(gtn).view(1, -1, 1)
    coords = ref.utils.get_coord(8192, 1).view(1, -1, 1)
    steps = 1500
    marks = np.array([99, 299, 599, 999, 1499])
    env, final_snr = [], []
    for nt in range(1, 9):
        torch.set_num_threads(nt)
        torch.manual_seed(0)
        m = ref.models.Siren(1, 128, 2, 1, True)
        losses, _, _ = ref_fit(m, coords, gt, steps)
        e = 10 * np.log10(sig / np.minimum.accumulate(np.asarray(losses, dtype=np.float64)))
        env.append(e[marks])
        final_snr.append(ref.utils.calculate_snr(gtn, m(coords).detach().numpy().ravel()))
        print("threads", nt, "envelope dB", np.round(e[marks], 2), "final-step SNR %.2f" % final_snr[-1], flush=True)
    torch.set_num_threads(8)
    np.savez_compressed(os.path.join(HERE, "converge_ens.npz"), marks=marks, env_db=np.array(env),
                        final_snr=np.array(final_snr), threads=np.arange(1, 9), steps=np.array(steps))


# ---------------------------------------------------------------- MDCT / window (mdct.py, window.py)
def make_mdct():
    out = {}
    rng = np.random.default_rng(0)
    out["kbd_2048"] = ref.window.KBDWindow(np.ones(2048))
    out["kbd_256"] = ref.window.KBDWindow(np.ones(256))
    x = rng.standard_normal(2048)
    out["frame"] = x
    out["mdct_frame"] = ref.mdct.MDCT(x, 1024, 1024)
    out["imdct_frame"] = ref.mdct.IMDCT(out["mdct_frame"], 1024, 1024)
    sig = rng.uniform(-1, 1, 10000)
    out["sig"] = sig
    out["stmdct_2048"] = ref.mdct.STMDCT(sig, 2048)
    out["stmdct_256"] = ref.mdct.STMDCT(sig[:5000], 256)
    out["stmdct_256_div"] = ref.mdct.STMDCT(sig[:5120], 256)  # len % halfN == 0 -> full extra half frame
    out["istmdct_256"] = ref.mdct.ISTMDCT(out["stmdct_256"], 256)
    # framing table
    lens = [2048, 5000, 5120, 220500, 264600, 1323000]
    out["frame_lens"] = np.array(lens)
    out["frame_counts_2048"] = np.array([ref.mdct.STMDCT(np.zeros(L), 2048).shape[1] if L < 20000 else L // 1024 for L in lens])
    # MDCTFitting on gt_bach 2 s: normalised coefficient image + 2-D coords (utils.py:313-414)
    ds = ref.utils.MDCTFitting(os.path.join(ROOT, "data", "gt_bach.wav"), duration=2, N=2048, takelog=False)
    out["mf_shape"] = np.array([ds.height, ds.width])
    out["mf_mean"] = np.array(ds.mean)
    out["mf_scale"] = np.array(ds.scale)
    out["mf_pixels_sha"] = np.array(sha(ds.pixels))
    out["mf_pixels_head"] = ds.pixels[:64].ravel()
    out["mf_coords_sha"] = np.array(sha(ds.coords.numpy()))
    out["mf_coords_head"] = ds.coords.numpy()[:200]
    np.savez_compressed(os.path.join(HERE, "mdct.npz"), **out)


# ---------------------------------------------------------------- KAN (kan.py)
def make_kan():
    out = {}
    torch.manual_seed(0)
    k = ref.kan.KAN([1, 16, 16, 1])
    sd = k.state_dict()
    for key, v in sd.items():
        out["sd_" + key] = v.numpy()
    out["sd_keys"] = np.array(list(sd.keys()))
    x = torch.linspace(-2.5, 2.5, 301).view(1, -1, 1)
    # include exact knots and the half-open edges
    xk = torch.cat([x.view(-1), k.layers[0].grid[0]]).view(1, -1, 1)
    out["x"] = xk.numpy().ravel()
    out["y"] = k(xk).detach().numpy().ravel()
    out["bsp"] = k.layers[0].b_splines(xk.view(-1, 1)).detach().numpy()
    gt = torch.sin(3 * xk)
    loss = torch.nn.functional.mse_loss(k(xk), gt)
    grads = torch.autograd.grad(loss, list(k.parameters()))
    out["loss"] = np.array(loss.item())
    out["pnames"] = np.array([n for n, _ in k.named_parameters()])
    for (n, _), g in zip(k.named_parameters(), grads):
        out["grad_" + n] = g.numpy()
    torch.manual_seed(0)
    big = ref.kan.KAN([1, 256, 256, 1])
    out["sd_hash_256_seed0"] = np.array(sd_hash(big.state_dict()))
    out["nparams_256"] = np.array(sum(p.numel() for p in big.parameters()))
    np.savez_compressed(os.path.join(HERE, "kan.npz"), **out)


# ---------------------------------------------------------------- in_features = 2: the MDCT-image fit (utils.py:382-403, run.py:67-76)
def make_mdct2d():
    """reference MDCTFitting (gt_bach, 1 s -> 1024 x 43 coefficient image) fitted by the reference Siren(2, 128, 2, 1):
    init, forward, and teacher-forced steps 0/1/2 of run.py's loop."""
    out = {}
    ds = ref.utils.MDCTFitting(os.path.join(ROOT, "data", "gt_bach.wav"), duration=1, N=2048, takelog=False)
    coords, pixels = ds[0]
    coords = torch.as_tensor(coords).reshape(1, -1, 2).float()
    gt = torch.as_tensor(np.asarray(pixels)).reshape(1, -1, 1).float()          # run.py:75 DataLoader adds the batch dim
    out["shape"] = np.array([ds.height, ds.width])
    out["coords_sha"] = np.array(sha(coords.numpy()))
    out["gt"] = gt.numpy().ravel()
    torch.manual_seed(5)
    m = ref.models.Siren(2, 128, 2, 1, True, first_omega_0=30, hidden_omega_0=30.0)
    out["p0"] = np.concatenate([p.detach().numpy().ravel() for p in m.parameters()])
    losses, lrs, rec = ref_fit(m, coords, gt, 4, record_steps=(0, 1, 2))
    out["losses"] = np.array(losses, dtype=np.float64)
    for k, v in rec.items():
        out[k] = v
    np.savez_compressed(os.path.join(HERE, "mdct2d.npz"), **out)


# ---------------------------------------------------------------- KAN convergence (run.py:109-110 + the loop)
def make_kan_converge():
    """reference kan.KAN([1, 256, 256, 1]) (BASELINE.json config 4's network) fitted to 8192 samples of gt_bach by
    run.py's loop for 500 steps, at 8 threads and at 1 thread (the reference's own run-to-run band): SNR at marks."""
    sr, data = gt_bach(8192, 44100)
    gtn = data / np.max(np.abs(data))
    gt = torch.from_numpy(gtn).view(1, -1, 1)
    coords = ref.utils.get_coord(8192, 1).view(1, -1, 1)
    marks = [50, 100, 200, 300, 500]
    out = {"gt": gtn, "marks": np.array(marks)}
    for nt in (8, 1):
        torch.set_num_threads(nt)
        torch.manual_seed(2)
        m = ref.kan.KAN([1, 256, 256, 1])
        if nt == 8:
            out["p0"] = np.concatenate([p.detach().numpy().ravel() for p in m.parameters()])
        opt = torch.optim.Adam(m.parameters(), lr=1e-3)
        sched = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode="min", factor=0.8, patience=200, min_lr=1e-6)
        mse = torch.nn.MSELoss()
        losses, snr = [], []
        for step in range(500):
            o = m(coords)
            loss = mse(o, gt)
            losses.append(loss.item())
            opt.zero_grad()
            loss.backward()
            opt.step()
            sched.step(loss)
            if step + 1 in marks:
                with torch.no_grad():
                    snr.append(ref.utils.calculate_snr(gtn, m(coords).numpy().ravel()))
                print("kan t%d step %d loss %.3e snr %.2f" % (nt, step + 1, losses[-1], snr[-1]), flush=True)
        out["losses_t%d" % nt] = np.array(losses)
        out["snr_t%d" % nt] = np.array(snr)
    torch.set_num_threads(8)
    np.savez_compressed(os.path.join(HERE, "kan_converge.npz"), **out)


if __name__ == "__main__":
    which = sys.argv[1:] or ["coords", "siren", "snake", "encoding", "mdct", "kan", "converge"]
    for w in which:
        print("making", w)
        globals()["make_" + w]()
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))
