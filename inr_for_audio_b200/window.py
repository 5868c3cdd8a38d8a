"""Host-side window functions with the reference's signatures (window.py:16-64).

Data-preparation code (runs once per fit, float64 numpy like the reference); the per-step MDCT-domain loss uses
the device kernels, which take the window table produced by `kbd_table` here.
"""
import functools

import numpy as np
from scipy.special import i0


@functools.lru_cache(maxsize=16)
def _kbd_cached(N: int, alpha: float) -> np.ndarray:
    # window.py:52-58: Kaiser kernel sampled at (2n+1)/(N/2+1) - 1 (note N/2+1, not N/2), normalised by the sum
    # over N/2+1 samples while the running sum covers the first N/2 of them; right half mirrors the left.
    half = N // 2
    pos = (2.0 * np.arange(half + 1) + 1.0) / (N / 2 + 1) - 1.0
    kaiser = i0(np.pi * alpha * np.sqrt(1.0 - pos ** 2)) / i0(np.pi * alpha)
    left = np.sqrt(np.cumsum(kaiser[:half]) / np.sum(kaiser))
    w = np.concatenate([left, left[::-1]])
    w.setflags(write=False)
    return w


def kbd_table(N: int, alpha: float = 4.0) -> np.ndarray:
    """The length-N KBD window itself (float64), cached: the reference recomputes it for every frame."""
    return _kbd_cached(int(N), float(alpha))


def KBDWindow(dataSampleArray, alpha=4.):
    """window.py:42-61."""
    x = np.asarray(dataSampleArray)
    return kbd_table(len(x), alpha) * x


def SineWindow(dataSampleArray):
    """window.py:16-26."""
    x = np.asarray(dataSampleArray)
    N = len(x)
    return np.sin(np.pi * (np.arange(N) + 0.5) / N) * x


def HanningWindow(dataSampleArray):
    """window.py:29-39."""
    x = np.asarray(dataSampleArray)
    N = len(x)
    return 0.5 * (1.0 - np.cos(2.0 * np.pi * (np.arange(N) + 0.5) / N)) * x


def RECTWindow(dataSampleArray):
    """window.py:63-64 (note the reference's 0.2 gain)."""
    x = np.asarray(dataSampleArray)
    return np.ones(len(x)) * x * 0.2
