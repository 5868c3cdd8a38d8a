"""Drop-in for the reference's kan.py (the vendored efficient-kan): KANLinear, KAN (kan.py:6-285).

Same constructor arguments, parameter names / shapes / state_dict keys (`layers.{i}.base_weight`,
`.spline_weight`, `.spline_scaler`, buffer `.grid`) and the same CPU-RNG consumption at construction, so
`torch.manual_seed(s); KAN([1, 256, 256, 1])` gives the reference's parameters bit for bit (the model is built on
the CPU and then moved, as run.py:110-115 does: `curve2coeff` uses the CPU lstsq driver).

`forward` runs the CUDA path of csrc/kan.cu through the C ABI (inr_kan_forward / inr_kan_loss_backward); there is no
PyTorch fallback for it.  `update_grid` and `regularization_loss` are never called by run.py (run.py:158 calls
model(x), i.e. update_grid=False); they are kept for API parity as plain host-side torch code.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import List, Optional

import torch
from torch import nn

from . import _lib as L
from .engine import SirenFitter, Workspace


@dataclass
class KanSpec:
    layers_hidden: tuple
    grid_size: int = 5
    spline_order: int = 3

    @property
    def in_features(self):
        return self.layers_hidden[0]

    @property
    def nb(self):
        return self.grid_size + self.spline_order

    def desc(self, precision="fp32") -> L.KanDesc:
        prec = L.PRECISIONS[precision] if isinstance(precision, str) else int(precision)
        if len(self.layers_hidden) - 1 > L.KAN_MAX_LAYERS:
            raise L.InrLibraryError("KAN: at most %d layers on the device path" % L.KAN_MAX_LAYERS)
        w = (C.c_int32 * (L.KAN_MAX_LAYERS + 1))(*self.layers_hidden)
        return L.KanDesc(len(self.layers_hidden) - 1, w, self.grid_size, self.spline_order, prec)

    def param_sizes(self):
        """flat sizes in state_dict order: base_weight, spline_weight, spline_scaler per layer."""
        out = []
        for i, o in zip(self.layers_hidden, self.layers_hidden[1:]):
            out += [o * i, o * i * self.nb, o * i]
        return out

    def num_params(self) -> int:
        return sum(self.param_sizes())

    def flops_per_coord(self) -> int:
        """dense-equivalent fwd + dPhi + dW flops per coordinate; the first layer has no dPhi."""
        tot = 0
        for l, (i, o) in enumerate(zip(self.layers_hidden, self.layers_hidden[1:])):
            tot += (3 if l > 0 else 2) * 2 * i * (self.nb + 1) * o
        return tot


class KANLinear(nn.Module):
    """kan.py:6-237."""

    def __init__(self, in_features, out_features, grid_size=5, spline_order=3, scale_noise=0.1, scale_base=1.0,
                 scale_spline=1.0, enable_standalone_scale_spline=True, base_activation=nn.SiLU, grid_eps=0.02,
                 grid_range=[-1, 1]):
        super().__init__()
        if not enable_standalone_scale_spline:
            raise NotImplementedError("enable_standalone_scale_spline=False is never used by run.py; not on the device path")
        if base_activation is not nn.SiLU:
            raise NotImplementedError("the device path implements the default base activation SiLU (kan.py:17)")
        self.in_features, self.out_features = in_features, out_features
        self.grid_size, self.spline_order = grid_size, spline_order
        self.scale_noise, self.scale_base, self.scale_spline = scale_noise, scale_base, scale_spline
        self.enable_standalone_scale_spline = True
        self.base_activation = base_activation()
        self.grid_eps = grid_eps
        step = (grid_range[1] - grid_range[0]) / grid_size
        knots = torch.arange(-spline_order, grid_size + spline_order + 1) * step + grid_range[0]      # kan.py:26-35
        self.register_buffer("grid", knots.expand(in_features, -1).contiguous())
        nb = grid_size + spline_order
        self.base_weight = nn.Parameter(torch.empty(out_features, in_features))
        self.spline_weight = nn.Parameter(torch.empty(out_features, in_features, nb))
        self.spline_scaler = nn.Parameter(torch.empty(out_features, in_features))
        self.reset_parameters()

    # -- init: same torch calls in the same order as kan.py:56-76 (RNG parity) -----------------------------------
    def reset_parameters(self):
        nn.init.kaiming_uniform_(self.base_weight, a=math.sqrt(5) * self.scale_base)
        with torch.no_grad():
            noise = (torch.rand(self.grid_size + 1, self.in_features, self.out_features) - 1 / 2) * self.scale_noise / self.grid_size
            self.spline_weight.data.copy_(self.curve2coeff(self.grid.T[self.spline_order:-self.spline_order], noise))
            nn.init.kaiming_uniform_(self.spline_scaler, a=math.sqrt(5) * self.scale_spline)

    def b_splines(self, x: torch.Tensor) -> torch.Tensor:
        """kan.py:78-111: (batch, in) -> (batch, in, grid_size + spline_order).  On a CUDA tensor this is the device
        kernel (inr_kan_b_splines); on the CPU (construction-time curve2coeff only) the recursion runs in torch."""
        assert x.dim() == 2 and x.size(1) == self.in_features
        if x.is_cuda:
            x = x.contiguous().float()
            out = torch.empty(x.size(0), self.in_features, self.grid_size + self.spline_order, device=x.device)
            L.check(L.lib().inr_kan_b_splines(L.ptr(x), L.ptr(self.grid.contiguous()), x.size(0), self.in_features,
                                              self.grid_size, self.spline_order, L.ptr(out), L.stream_ptr()))
            return out
        g = self.grid
        xe = x.unsqueeze(-1)
        b = ((xe >= g[:, :-1]) & (xe < g[:, 1:])).to(x.dtype)
        for k in range(1, self.spline_order + 1):
            left = (xe - g[:, :-(k + 1)]) / (g[:, k:-1] - g[:, :-(k + 1)])
            right = (g[:, k + 1:] - xe) / (g[:, k + 1:] - g[:, 1:(-k)])
            b = left * b[:, :, :-1] + right * b[:, :, 1:]
        return b.contiguous()

    def curve2coeff(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """kan.py:113-143: least-squares spline coefficients, (out, in, grid_size + spline_order)."""
        assert x.dim() == 2 and x.size(1) == self.in_features
        assert y.size() == (x.size(0), self.in_features, self.out_features)
        sol = torch.linalg.lstsq(self.b_splines(x).transpose(0, 1), y.transpose(0, 1)).solution
        return sol.permute(2, 0, 1).contiguous()

    @property
    def scaled_spline_weight(self):
        return self.spline_weight * self.spline_scaler.unsqueeze(-1)

    def forward(self, x: torch.Tensor):
        """kan.py:153-166 -- as a single-layer KAN on the device."""
        return _kan_apply([self], x)

    @torch.no_grad()
    def update_grid(self, x: torch.Tensor, margin=0.01):
        """kan.py:168-215 (host-side torch; not on the fit path)."""
        assert x.dim() == 2 and x.size(1) == self.in_features
        n = x.size(0)
        per_input = torch.einsum("bik,oik->bio", self.b_splines(x), self.scaled_spline_weight)
        xs = torch.sort(x, dim=0)[0]
        pick = torch.linspace(0, n - 1, self.grid_size + 1, dtype=torch.int64, device=x.device)
        adaptive = xs[pick]
        step = (xs[-1] - xs[0] + 2 * margin) / self.grid_size
        uniform = torch.arange(self.grid_size + 1, dtype=torch.float32, device=x.device).unsqueeze(1) * step + xs[0] - margin
        core = self.grid_eps * uniform + (1 - self.grid_eps) * adaptive
        k = self.spline_order
        lo = core[:1] - step * torch.arange(k, 0, -1, device=x.device).unsqueeze(1)
        hi = core[-1:] + step * torch.arange(1, k + 1, device=x.device).unsqueeze(1)
        self.grid.copy_(torch.cat([lo, core, hi], dim=0).T)
        self.spline_weight.data.copy_(self.curve2coeff(x, per_input))

    def regularization_loss(self, regularize_activation=1.0, regularize_entropy=1.0):
        """kan.py:217-237."""
        l1 = self.spline_weight.abs().mean(-1)
        act = l1.sum()
        p = l1 / act
        return regularize_activation * act + regularize_entropy * (-(p * p.log()).sum())


class _KanFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, flat, grids, x2d, spec, ws):
        n = x2d.shape[0]
        desc = spec.desc("fp32")
        out = torch.empty(n, spec.layers_hidden[-1], dtype=torch.float32, device=flat.device)
        buf = ws.get(desc, n, False)
        L.check(L.lib().inr_kan_forward(C.byref(desc), L.ptr(flat), L.ptr(grids), L.ptr(x2d), n, 0, n, 1.0, L.ptr(out),
                                        L.ptr(buf), buf.numel(), L.stream_ptr()))
        ctx.spec, ctx.ws = spec, ws
        ctx.save_for_backward(flat, grids, x2d)
        return out

    @staticmethod
    def backward(ctx, gout):
        flat, grids, x2d = ctx.saved_tensors
        spec, ws = ctx.spec, ctx.ws
        if spec.layers_hidden[-1] != 1:
            raise L.InrLibraryError("KAN backward on the device path needs layers_hidden[-1] == 1 (run.py:110)")
        n = x2d.shape[0]
        desc = spec.desc("fp32")
        grads = torch.zeros_like(flat)
        opt = torch.zeros(L.OPT_STATE_BYTES + 8, dtype=torch.uint8, device=flat.device)
        buf = ws.get(desc, n, True)
        g = gout.reshape(-1).contiguous().float()
        L.check(L.lib().inr_kan_loss_backward(C.byref(desc), L.ptr(flat), L.ptr(grids), L.ptr(x2d), None, n, 0, n, 1.0,
                                              L.LOSS_MSE, 0.0, L.ptr(g), L.ptr(grads), L.ptr(opt), None, L.ptr(buf),
                                              buf.numel(), L.stream_ptr()))
        return grads, None, None, None, None


_WS = {}


def _kan_apply(layers: List[KANLinear], x: torch.Tensor) -> torch.Tensor:
    ps = [p for l in layers for p in (l.base_weight, l.spline_weight, l.spline_scaler)]
    if not ps[0].is_cuda:
        raise L.InrLibraryError("KAN.forward: the model is on %s; the INR hot path runs only on a CUDA device "
                                "(no CPU fallback) -- call .cuda() first as run.py:115 does" % ps[0].device)
    spec = KanSpec(tuple([layers[0].in_features] + [l.out_features for l in layers]), layers[0].grid_size,
                   layers[0].spline_order)
    flat = torch.cat([p.reshape(-1) for p in ps])
    grids = torch.cat([l.grid.reshape(-1) for l in layers]).contiguous().float()
    if x.requires_grad and torch.is_grad_enabled():
        # the device path returns parameter gradients only (run.py never reads d out / d coords);
        # a differentiable INPUT -- stacked KANLinear modules, update_grid=True inside a training step -- would silently lose
        # the gradient to earlier layers, so it is refused instead
        raise L.InrLibraryError("KAN device path: gradients with respect to the input are not computed (the reference's "
                                "KANLinear is differentiable in x); call the whole KAN as one module, or detach() the input")
    x2d = x.detach().reshape(-1, spec.in_features).contiguous().float()
    ws = _WS.setdefault(flat.device, Workspace(flat.device))
    out = _KanFn.apply(flat, grids, x2d, spec, ws)
    return out.view(*x.shape[:-1], spec.layers_hidden[-1])


class KAN(nn.Module):
    """kan.py:240-285."""

    def __init__(self, layers_hidden, grid_size=5, spline_order=3, scale_noise=0.1, scale_base=1.0, scale_spline=1.0,
                 base_activation=nn.SiLU, grid_eps=0.02, grid_range=[-1, 1]):
        super().__init__()
        self.grid_size, self.spline_order = grid_size, spline_order
        self.layers = nn.ModuleList(
            KANLinear(i, o, grid_size=grid_size, spline_order=spline_order, scale_noise=scale_noise, scale_base=scale_base,
                      scale_spline=scale_spline, base_activation=base_activation, grid_eps=grid_eps, grid_range=grid_range)
            for i, o in zip(layers_hidden, layers_hidden[1:]))
        self.spec = KanSpec(tuple(layers_hidden), grid_size, spline_order)

    def forward(self, x: torch.Tensor, update_grid=False):
        if update_grid:
            for layer in self.layers:        # host-side path of the reference (kan.py:275-278); not used by run.py
                layer.update_grid(x.reshape(-1, layer.in_features))
                x = layer(x)
            return x
        return _kan_apply(list(self.layers), x)

    def regularization_loss(self, regularize_activation=1.0, regularize_entropy=1.0):
        return sum(l.regularization_loss(regularize_activation, regularize_entropy) for l in self.layers)

    def flat_parameters(self) -> torch.Tensor:
        """flat fp32 parameter vector in state_dict order (a copy; load_flat writes it back)."""
        return torch.cat([p.detach().reshape(-1) for p in self.parameters()])

    def flat_grids(self) -> torch.Tensor:
        return torch.cat([l.grid.reshape(-1) for l in self.layers]).contiguous().float()

    @torch.no_grad()
    def load_flat(self, flat: torch.Tensor):
        o = 0
        for p in self.parameters():
            n = p.numel()
            p.copy_(flat[o:o + n].view(p.shape))
            o += n


def _require_uniform_grids(spec: KanSpec, grids: torch.Tensor):
    """The bf16 tensor-core mode evaluates the cubic B-spline in closed form on a uniform knot vector shared by all
    inputs of a layer (what KANLinear.__init__ builds, kan.py:26-35); grids refined by update_grid need precision='fp32'."""
    g = grids.detach().reshape(-1).float().cpu()
    nk = spec.grid_size + 2 * spec.spline_order + 1
    o = 0
    for fin in spec.layers_hidden[:-1]:
        blk = g[o:o + fin * nk].view(fin, nk)
        o += fin * nk
        step = (blk[0, -1] - blk[0, 0]) / (nk - 1)
        ideal = blk[0, 0] + step * torch.arange(nk)
        if not (torch.allclose(blk, ideal.expand(fin, nk), rtol=0, atol=1e-5 * float(abs(step)) + 1e-7)):
            raise L.InrLibraryError("KAN precision='bf16' needs uniform, input-independent knot vectors; use precision='fp32'")


class KanFitter(SirenFitter):
    """Device-resident fit of a KAN to one waveform: run.py:156-190 with arch='kan' (run.py:110).  Same engine as
    SirenFitter (flat arena, device Adam + plateau scheduler, CUDA graph, sharded allreduce); only the forward /
    backward entry points differ."""

    def __init__(self, spec: KanSpec, params0: torch.Tensor, grids: torch.Tensor, target: torch.Tensor,
                 coords: Optional[torch.Tensor] = None, **kw):
        self.grids = grids.reshape(-1).contiguous().float().to(target.device)
        kw.setdefault("precision", "fp32")
        if kw["precision"] == "bf16":
            _require_uniform_grids(spec, grids)
        if kw.get("spectral") is not None or kw.get("loss_mode", "mse") == "snr":
            raise NotImplementedError("KAN fits use the waveform mse/mae losses (run.py:110 is waveform-only)")
        super().__init__(spec, params0, target, coords, **kw)

    def _enqueue_backward(self, extra_gout=None, out=None):
        buf = self.ws.get(self.desc, self.n, True)
        L.check(L.lib().inr_kan_loss_backward(
            C.byref(self.desc), L.ptr(self.params), L.ptr(self.grids), L.ptr(self.coords), L.ptr(self.target), self.n,
            self.i0, self.T_total, self.coord_scale, self.loss_kind, self._w() / float(self.T_total), L.ptr(extra_gout),
            L.ptr(self.grads), L.ptr(self.opt), L.ptr(out), L.ptr(buf), buf.numel(), L.stream_ptr()))

    def forward(self, coords: Optional[torch.Tensor] = None, n: Optional[int] = None, i0: Optional[int] = None,
                T_total: Optional[int] = None) -> torch.Tensor:
        if coords is None and n is None:
            coords, n, i0, T_total = self.coords, self.n, self.i0, self.T_total
        elif coords is not None:
            coords = coords.reshape(-1, self.spec.in_features).contiguous().float()
            n, i0, T_total = coords.shape[0], 0, coords.shape[0]
        out = torch.empty(int(n), dtype=torch.float32, device=self.device)
        buf = self.ws.get(self.desc, int(n), False)
        L.check(L.lib().inr_kan_forward(C.byref(self.desc), L.ptr(self.params), L.ptr(self.grids), L.ptr(coords), int(n),
                                        int(i0 or 0), int(T_total or n), self.coord_scale, L.ptr(out), L.ptr(buf),
                                        buf.numel(), L.stream_ptr()))
        return out
