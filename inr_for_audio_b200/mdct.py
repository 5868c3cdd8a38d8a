"""Host-side MDCT / IMDCT / STMDCT / ISTMDCT with the reference's signatures (mdct.py:16-113).

Data preparation and reconstruction only (once per fit, float64 numpy, as in the reference).  Instead of the
reference's per-frame Python loop, frames are gathered with a strided view and transformed by one batched
FFT; the framing rule (tail-only zero pad of halfN - len % halfN samples, hop halfN, len // halfN frames,
mdct.py:54-58) is reproduced exactly -- `frame_plan` is the single source of truth for it and is what the
device loss kernels use too.
"""
import numpy as np
from numpy.lib.stride_tricks import sliding_window_view
from scipy.fft import fft, ifft

from .window import KBDWindow, kbd_table  # noqa: F401  (the reference's mdct.py re-exports window.*)


def frame_plan(length: int, N: int):
    """(halfN, pad_length, num_frames) of mdct.py:54-58."""
    halfN = N // 2
    return halfN, halfN - (length % halfN), length // halfN


def _twiddles(N: int, n0: float):
    n = np.arange(N)
    k = np.arange(N // 2)
    return n, k


def MDCT(data, a, b, isInverse=False):
    """mdct.py:16-42 for one frame or a batch of frames along the last axis."""
    data = np.asarray(data)
    N = a + b
    n0 = (b + 1) / 2
    n, k = _twiddles(N, n0)
    if isInverse:
        spec = data * np.exp(2j * np.pi * n0 * k / N)
        t = ifft(spec, N, axis=-1) * N
        return 2 * (t * np.exp(1j * np.pi * (n + n0) / N)).real
    t = fft(data * np.exp(-1j * np.pi * n / N), axis=-1)[..., :N // 2]
    return (2 / N) * (t * np.exp(-2j * np.pi * n0 * (k + 0.5) / N)).real


def IMDCT(data, a, b):
    """mdct.py:46-48."""
    return MDCT(data, a, b, True)


def STMDCT(data, N=1024):
    """mdct.py:51-72 -> (N/2, num_frames) float64."""
    data = np.asarray(data)
    halfN, pad, F = frame_plan(len(data), N)
    padded = np.concatenate([data.astype(np.float64, copy=False), np.zeros(pad)])
    if F == 0:
        return np.zeros((halfN, 0))
    frames = sliding_window_view(padded, N)[::halfN][:F]          # (F, N), frame i = padded[i*halfN : i*halfN+N]
    return MDCT(frames * kbd_table(N), halfN, halfN).T.copy()


def ISTMDCT(mdct_coefficients, N=1024):
    """mdct.py:74-113: IMDCT, window, overlap-add at hop N/2, drop the last half frame."""
    C = np.asarray(mdct_coefficients)
    halfN = N // 2
    F = C.shape[1]
    frames = IMDCT(C.T, halfN, halfN) * kbd_table(N)              # (F, N)
    rec = np.zeros(halfN * (F + 1))
    left = frames[:, :halfN].reshape(-1)
    right = frames[:, halfN:].reshape(-1)
    rec[:halfN * F] += left
    rec[halfN:] += right
    return rec[:halfN * F]
