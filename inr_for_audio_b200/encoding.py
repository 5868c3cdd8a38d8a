"""Input encodings in front of the first layer, on the device through the C ABI.

GaussianEncoding mirrors `rff.layers.GaussianEncoding(sigma, input_size, encoded_size, b=None)`, the third-party module
the reference applies to the coordinates once before the loop when `num_freq` is set (run.py:141-144): the package is
not vendored and not installed here, so this follows its published definition (PARITY UNPINNED):
    b ~ sigma * randn(encoded_size, input_size)   (drawn at construction, a frozen nn.Parameter)
    forward(v) = cat(cos(vp), sin(vp)), vp = 2*pi*v @ b.T
PosEncodingNeRF mirrors models.py:14-59 (pinned against the reference class in tests/golden).

Both produce the encoded coordinates with one kernel launch (inr_encode_gaussian / inr_encode_nerf); with `n=` instead
of a tensor they encode linspace(-scale, scale, T_total)[i0:i0+n] generated in the kernel, which is what a coordinate
shard of the multi-GPU fit uses.  No CPU fallback: a CPU tensor raises.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
from torch import nn

from . import _lib as L


def _prep(v: Optional[torch.Tensor], in_features: int, n, i0, T_total, dev):
    if v is not None:
        if not v.is_cuda:
            raise L.InrLibraryError("encoding: coordinates are on %s; the INR path runs only on a CUDA device" % v.device)
        lead = v.shape[:-1]
        v2 = v.detach().reshape(-1, in_features).contiguous().float()
        return v2, lead, v2.shape[0], 0, v2.shape[0]
    if in_features != 1:
        raise ValueError("in-kernel coordinates need in_features == 1")
    n = int(n)
    return None, (n,), n, int(i0 or 0), int(T_total or n)


class GaussianEncoding(nn.Module):
    """rff.layers.GaussianEncoding: (..., input_size) -> (..., 2*encoded_size)."""

    def __init__(self, sigma: Optional[float] = None, input_size: Optional[int] = None,
                 encoded_size: Optional[int] = None, b: Optional[torch.Tensor] = None):
        super().__init__()
        if b is None:
            if sigma is None or input_size is None or encoded_size is None:
                raise ValueError('Arguments "sigma," "input_size," and "encoded_size" are required.')
            b = torch.randn((encoded_size, input_size)) * sigma
        elif sigma is not None or input_size is not None or encoded_size is not None:
            raise ValueError('Only specify the "b" argument when using it.')
        self.b = nn.parameter.Parameter(b, requires_grad=False)

    @property
    def encoded_size(self) -> int:
        return self.b.shape[0]

    def forward(self, v: Optional[torch.Tensor] = None, *, n: Optional[int] = None, i0: int = 0,
                T_total: Optional[int] = None, scale: float = 1.0) -> torch.Tensor:
        F, in_features = self.b.shape
        if not self.b.is_cuda:
            raise L.InrLibraryError("GaussianEncoding: call .cuda() first (run.py:142); there is no CPU path")
        v2, lead, n, i0, T_total = _prep(v, in_features, n, i0, T_total, self.b.device)
        out = torch.empty((n, 2 * F), dtype=torch.float32, device=self.b.device)
        b = self.b.detach().contiguous().float()
        L.check(L.lib().inr_encode_gaussian(L.ptr(v2), in_features, n, i0, T_total, float(scale), L.ptr(b), F, L.ptr(out),
                                            L.stream_ptr()))
        return out.view(*lead, 2 * F)


class PosEncodingNeRF(nn.Module):
    """models.py:14-59: (..., in) -> (..., in + 2*in*num_frequencies), [x, sin(s^i pi x_j), cos(s^i pi x_j), ...]."""

    def __init__(self, in_features, sidelength=None, fn_samples=None, use_nyquist=True, num_frequencies=None, scale=2):
        super().__init__()
        self.in_features = in_features
        self.scale = scale
        self.sidelength = sidelength
        if num_frequencies is None:
            if self.in_features == 3:
                self.num_frequencies = 10
            elif self.in_features == 2:
                assert sidelength is not None
                if isinstance(sidelength, int):
                    sidelength = (sidelength, sidelength)
                self.num_frequencies = 4
                if use_nyquist:
                    self.num_frequencies = self.get_num_frequencies_nyquist(min(sidelength[0], sidelength[1]))
            elif self.in_features == 1:
                assert fn_samples is not None
                self.num_frequencies = 4
                if use_nyquist:
                    self.num_frequencies = self.get_num_frequencies_nyquist(fn_samples)
        else:
            self.num_frequencies = num_frequencies
        self.out_dim = in_features + in_features * 2 * self.num_frequencies

    def get_num_frequencies_nyquist(self, samples):
        nyquist_rate = 1 / (2 * (2 * 1 / samples))
        return int(np.floor(np.log2(nyquist_rate)))

    def forward(self, coords: Optional[torch.Tensor] = None, *, n: Optional[int] = None, i0: int = 0,
                T_total: Optional[int] = None, scale: float = 1.0, device=None) -> torch.Tensor:
        dev = coords.device if coords is not None else torch.device(device or "cuda")
        v2, lead, n, i0, T_total = _prep(coords, self.in_features, n, i0, T_total, dev)
        out = torch.empty((n, self.out_dim), dtype=torch.float32, device=dev)
        L.check(L.lib().inr_encode_nerf(L.ptr(v2), self.in_features, n, i0, T_total, float(scale),
                                        int(self.num_frequencies), float(self.scale), L.ptr(out), L.stream_ptr()))
        return out.view(*lead, self.out_dim)
