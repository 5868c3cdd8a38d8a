"""Device-side windowed-transform losses (csrc/spectral.cu) behind small Python objects.

MdctLoss / StftLoss / SnrLoss hold the target-side precomputation (coefficient image, magnitudes) in device memory
and expose `accumulate(pred, gout, opt)`: add weight * loss to the optimiser state's loss accumulator and
d(weight * loss)/d pred to `gout` -- the upstream gradient that inr_siren_train_backward consumes.  Mixing follows
run.py:163-169: loss = (1 - alpha) * time_loss + alpha * spectral_loss.

  MdctLoss   mdct.py:51-72 STMDCT (N=2048, KBD alpha=4) of output and target, L1 ('l1') or MSE ('mse') of the
             coefficient images (BASELINE.json config 3; the reference has no such loss: definition in DESIGN.md).
  StftLoss   auraloss.freq.STFTLoss() defaults (run.py:128,160) -- parity unpinned (third-party, un-vendored).
  SnrLoss    auraloss.time.SNRLoss() (run.py:126,165)           -- parity unpinned.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib as L
from .window import kbd_table


def stmdct_device(x: torch.Tensor, N: int = 2048, window: torch.Tensor | None = None) -> torch.Tensor:
    """mdct.py:51-72 on the device: x (T,) fp32 cuda -> (F, N/2) fp32 FRAME-major (the reference returns (N/2, F))."""
    if not x.is_cuda:
        raise L.InrLibraryError("stmdct_device: input must be a CUDA tensor (no CPU fallback; use mdct.STMDCT on the host)")
    x = x.reshape(-1).contiguous().float()
    T = x.numel()
    if window is None:
        window = torch.from_numpy(kbd_table(N).astype(np.float32)).to(x.device)
    F = int(L.check(L.lib().inr_stmdct_num_frames(T, N)))
    coef = torch.empty((F, N // 2), dtype=torch.float32, device=x.device)
    if F == 0:
        return coef
    L.check(L.lib().inr_stmdct_forward(L.ptr(x), T, N, L.ptr(window), L.ptr(coef), L.stream_ptr()))
    return coef


def istmdct_device(coef: torch.Tensor, N: int = 2048, window: torch.Tensor | None = None) -> torch.Tensor:
    """mdct.py:74-113 ISTMDCT on the device: coef (F, N/2) fp32 cuda, FRAME-major (i.e. the transpose of what the
    reference takes) -> (F * N/2,) fp32: IMDCT per frame, window, overlap-add at hop N/2, last half frame dropped."""
    if not coef.is_cuda:
        raise L.InrLibraryError("istmdct_device: input must be a CUDA tensor (no CPU fallback; use mdct.ISTMDCT on the host)")
    coef = coef.contiguous().float()
    F = coef.shape[0]
    assert coef.dim() == 2 and coef.shape[1] == N // 2, coef.shape
    if window is None:
        window = torch.from_numpy(kbd_table(N).astype(np.float32)).to(coef.device)
    out = torch.empty(F * (N // 2), dtype=torch.float32, device=coef.device)
    if F:
        L.check(L.lib().inr_istmdct(L.ptr(coef), F, N, L.ptr(window), L.ptr(out), L.stream_ptr()))
    return out


class MdctLoss:
    def __init__(self, target: torch.Tensor, N: int = 2048, kind: str = "l1", kbd_alpha: float = 4.0):
        if not target.is_cuda:
            raise L.InrLibraryError("MdctLoss: target must be a CUDA tensor")
        self.N = int(N)
        self.T = int(target.numel())
        self.kind = L.LOSS_MAE if kind in ("l1", "mae") else L.LOSS_MSE
        self.window = torch.from_numpy(kbd_table(self.N, kbd_alpha).astype(np.float32)).to(target.device)
        self.coef = stmdct_device(target, self.N, self.window)
        self.count = max(self.coef.numel(), 1)
        # algorithmic HBM bytes of one evaluation: read pred, read target coefficients, read-modify-write gout
        self.bytes_per_eval = 4 * self.T + 4 * self.coef.numel() + 8 * self.T

    def accumulate(self, pred: torch.Tensor, gout: torch.Tensor, opt: torch.Tensor, weight: float):
        L.check(L.lib().inr_mdct_loss(L.ptr(pred), L.ptr(self.coef), self.T, self.N, L.ptr(self.window), self.kind,
                                      float(weight) / self.count, L.ptr(gout), L.ptr(opt), L.stream_ptr()))


class StftLoss:
    def __init__(self, target: torch.Tensor, n_fft: int = 1024, hop: int = 256, eps: float = 1e-8):
        if not target.is_cuda:
            raise L.InrLibraryError("StftLoss: target must be a CUDA tensor")
        self.n_fft, self.hop, self.eps = int(n_fft), int(hop), float(eps)
        self.T = int(target.numel())
        dev = target.device
        self.window = torch.hann_window(self.n_fft, dtype=torch.float32).to(dev)   # periodic, as torch.stft / auraloss
        F = int(L.check(L.lib().inr_stft_num_frames(self.T, self.hop)))
        self.mag = torch.empty((F, self.n_fft // 2 + 1), dtype=torch.float32, device=dev)
        sumsq = torch.zeros(1, dtype=torch.float64, device=dev)
        t = target.reshape(-1).contiguous().float()
        L.check(L.lib().inr_stft_mag(L.ptr(t), self.T, self.n_fft, self.hop, L.ptr(self.window), self.eps, L.ptr(self.mag),
                                     L.ptr(sumsq), L.stream_ptr()))
        self.sumsq = float(sumsq.item())
        self.bytes_per_eval = 2 * (4 * self.T * self.n_fft // self.hop + 4 * self.mag.numel()) + 8 * self.T

    def accumulate(self, pred: torch.Tensor, gout: torch.Tensor, opt: torch.Tensor, weight: float):
        L.check(L.lib().inr_stft_loss(L.ptr(pred), L.ptr(self.mag), C.c_double(self.sumsq), self.T, self.n_fft, self.hop,
                                      L.ptr(self.window), self.eps, float(weight), L.ptr(gout), L.ptr(opt),
                                      L.stream_ptr()))


class SnrLoss:
    def __init__(self, target: torch.Tensor, eps: float = 1e-8):
        if not target.is_cuda:
            raise L.InrLibraryError("SnrLoss: target must be a CUDA tensor")
        self.target = target.reshape(-1).contiguous().float()
        self.T = self.target.numel()
        self.eps = float(eps)
        self.scratch = torch.zeros(2, dtype=torch.float64, device=target.device)
        self.bytes_per_eval = 2 * 8 * self.T + 8 * self.T

    def accumulate(self, pred: torch.Tensor, gout: torch.Tensor, opt: torch.Tensor, weight: float):
        L.check(L.lib().inr_snr_loss(L.ptr(pred), L.ptr(self.target), self.T, self.eps, float(weight), L.ptr(gout),
                                     L.ptr(opt), L.ptr(self.scratch), L.stream_ptr()))


def make_spectral_loss(kind, target: torch.Tensor):
    """'mdct' | 'mdct_mse' | 'stft' | None -> loss object."""
    if kind is None or kind is False:
        return None
    if kind == "mdct":
        return MdctLoss(target, 2048, "l1")
    if kind == "mdct_mse":
        return MdctLoss(target, 2048, "mse")
    if kind == "stft":
        return StftLoss(target)
    raise ValueError("unknown spectral loss %r (mdct, mdct_mse, stft)" % (kind,))
