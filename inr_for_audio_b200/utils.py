"""Drop-in for the hot-path part of the reference's utils.py: get_coord, WaveformFitting, MDCTFitting,
calculate_snr (utils.py:77-149, 312-414).  Plotting / filter helpers are out of scope.

Like the reference's `from utils import *` contract (run.py:8) this module re-exports os, np, mdct, decimate.
"""
import ctypes as C
import os  # noqa: F401

import numpy as np
import torch
from scipy.io import wavfile
from scipy.signal import decimate  # noqa: F401
from torch.utils.data import Dataset

from . import mdct  # noqa: F401
from . import _lib as L


def calculate_snr(original_signal, noisy_signal):
    """utils.py:77-97: 10 log10(mean(x^2) / mean((y-x)^2))."""
    x = np.asarray(original_signal)
    y = np.asarray(noisy_signal)
    err = y - x
    return 10 * np.log10(np.mean(x ** 2) / np.mean(err ** 2))


def get_coord(sidelen, dim=2, scale=1, device=None):
    """utils.py:99-109.  On a CUDA `device` the axis comes from the bit-exact device generator
    (inr_coords_linspace); on the CPU (default, as the reference) from torch.linspace itself."""
    if device is not None and torch.device(device).type == "cuda":
        axis = torch.empty(sidelen, dtype=torch.float32, device=device)
        with torch.cuda.device(axis.device):
            L.check(L.lib().inr_coords_linspace(L.ptr(axis), 0, sidelen, sidelen, float(scale), L.stream_ptr()))
    else:
        axis = torch.linspace(-1 * scale, 1 * scale, steps=sidelen)
    grids = torch.meshgrid(*([axis] * dim), indexing="ij")
    return torch.stack(grids, dim=-1).reshape(-1, dim)


class WaveformFitting(Dataset):
    """utils.py:111-149: channel 0, float32, first duration*sr samples, optional scipy decimate; item =
    (coords (T,1), amplitude / max|amplitude| (T,1))."""

    def __init__(self, filename, duration, decimation=1, verbose=True):
        self.sample_rate, data = wavfile.read(filename)
        if data.ndim > 1:
            data = data[:, 0]
        self.data = data.astype(np.float32)[0:duration * self.sample_rate]
        self.original_sample_rate = self.sample_rate
        if decimation > 1:
            q = int(decimation)
            self.data = decimate(self.data, q=q)
            self.sample_rate = self.sample_rate // q
        self.height = len(self.data)
        self.width = 1
        self.coord = get_coord(len(self.data), 1)
        if verbose:
            print('Original sample rate: ', self.original_sample_rate)
            print('Training sample rate:', self.sample_rate)
            print('Training coord shape: ', self.coord.shape)

    def get_num_samples(self):
        return self.coord.shape[0]

    def __len__(self):
        return 1

    def normalised(self) -> np.ndarray:
        """amplitude / max|amplitude| as a flat float32 array (what __getitem__ wraps)."""
        amp = self.data
        return (amp / np.max(np.abs(amp))).astype(np.float32, copy=False)

    def __getitem__(self, idx):
        amp = self.data
        amp = amp / np.max(np.abs(amp))
        return self.coord, torch.Tensor(amp).view(-1, 1)


class MDCTFitting(Dataset):
    """utils.py:313-414: channel 1 of a stereo file, normalised by the max of the WHOLE file, STMDCT image,
    optional log, mean/scale normalisation, 2-D 'ij' coordinate grid flattened frequency-major."""

    def __init__(self, filename, duration, N=1024, highpass=False, takelog=False, verbose=True):
        super().__init__()
        self.sample_rate, data = wavfile.read(filename)
        self.original_sample_rate = self.sample_rate
        if data.ndim > 1:
            data = data[:, 1]
        if highpass:
            raise NotImplementedError("hpfilter (utils.py:49-52) is outside the hot-path scope")
        peak = np.max(np.abs(data))
        self.data = torch.from_numpy(data.astype(np.float32)[:duration * self.sample_rate] / peak)
        self.mdct = mdct.STMDCT(self.data.numpy(), N=N).astype(np.float32)
        self.shift = 0.0
        if takelog:
            self.shift = np.abs(np.min(self.mdct)) + 1e-8
            self.mdct = np.log(self.mdct + self.shift)
        self.mean = np.mean(self.mdct)
        self.mdct = self.mdct - self.mean
        self.scale = np.max(np.abs(self.mdct))
        self.mdct = self.mdct / self.scale
        self.height, self.width = self.mdct.shape
        if verbose:
            print("sample rate: ", self.sample_rate)
            print("mdct scale: ", self.scale)
            print("mdct mean: ", self.mean)
            print("height: ", self.height)
            print("width: ", self.width)
        h = torch.linspace(-1, 1, steps=self.height)
        w = torch.linspace(-1, 1, steps=self.width)
        hg, wg = torch.meshgrid(h, w, indexing='ij')
        self.coords = torch.stack((hg, wg), dim=-1).reshape(self.height * self.width, -1)
        self.pixels = self.mdct.reshape(-1, 1)

    def __len__(self):
        return 1

    def __getitem__(self, idx):
        if idx > 0:
            raise IndexError
        return self.coords, self.pixels


def synth_waveform(T: int, sr: int = 44100, seed: int = 0) -> np.ndarray:
    """Synthetic benchmark waveform: 32 tremolo'd partials, log-uniform in [55, 8000] Hz, plus
    white noise (sigma 1e-3); float64 synthesis, peak-normalised to 1.0 like utils.py:144-148, returned as fp32."""
    rng = np.random.default_rng(seed)
    K = 32
    t = np.arange(T, dtype=np.float64) / sr
    f = np.exp(rng.uniform(np.log(55.0), np.log(8000.0), K))
    amp = rng.uniform(0.5, 1.0, K) / (np.arange(1, K + 1) ** 0.7)
    ph = rng.uniform(0, 2 * np.pi, K)
    r = rng.uniform(0.1, 2.0, K)
    phi = rng.uniform(0, 2 * np.pi, K)
    x = np.zeros(T, dtype=np.float64)
    for k in range(K):
        x += amp[k] * 0.5 * (1 + np.sin(2 * np.pi * r[k] * t + phi[k])) * np.sin(2 * np.pi * f[k] * t + ph[k])
    x += rng.normal(0.0, 1e-3, T)
    x /= np.max(np.abs(x))
    return x.astype(np.float32)
