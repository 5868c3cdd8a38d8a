"""Drop-in for the reference's models.py on the fit path: SineLayer, Snake, Siren, SirenWithSnakeTanh.

Same constructor arguments, module tree, state_dict keys and CPU-RNG consumption as the reference
(models.py:84-120, 243-275, 306-394), so `torch.manual_seed(s); Siren(...)` yields bit-identical parameters and
checkpoints interchange.  `forward(coords)` does not run nn.Sequential: it sends the whole network to the
fused CUDA path (inr_siren_forward / inr_siren_loss_backward) through a torch.autograd.Function, so
`loss.backward()` + a stock `torch.optim.Adam(model.parameters())` still work (the compatible, slower route);
the fast route is engine.SirenFitter / run.train, which keep the optimiser on the device too.

After construction every parameter is a view into ONE flat fp32 buffer in state_dict order -- the layout the
kernels consume (include/inr_b200.h "flat params").
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
from torch import nn

from . import _lib as L
from .encoding import GaussianEncoding, PosEncodingNeRF  # noqa: F401  (models.py:14-59; rff.layers.GaussianEncoding)
from .engine import SirenSpec, Workspace


class SineLayer(nn.Module):
    """models.py:84-120: sin(omega_0 * linear(x)); W ~ U(+-1/in) if first else U(+-sqrt(6/in)/omega_0)."""

    def __init__(self, in_features, out_features, bias=True, is_first=False, omega_0=30):
        super().__init__()
        self.omega_0 = omega_0
        self.is_first = is_first
        self.in_features = in_features
        self.linear = nn.Linear(in_features, out_features, bias=bias)  # draws default W and b first (RNG parity)
        self.init_weights()

    def init_weights(self):
        with torch.no_grad():
            bound = 1 / self.in_features if self.is_first else np.sqrt(6 / self.in_features) / self.omega_0
            self.linear.weight.uniform_(-bound, bound)


class _SirenFn(torch.autograd.Function):
    """forward(coords) and its backward over the C ABI; gradients flow to the flat parameter buffer only
    (d/d coords is computed by the reference, models.py:272, but nothing on the path reads it)."""

    @staticmethod
    def forward(ctx, flat, coords2d, model):
        n = coords2d.shape[0]
        out = torch.empty(n, dtype=torch.float32, device=flat.device)
        desc = model._desc()
        buf = model._ws.get(desc, n, False)
        L.check(L.lib().inr_siren_forward(C.byref(desc), L.ptr(flat), L.ptr(coords2d), n, 0, n, 1.0, L.ptr(out),
                                          L.ptr(buf), buf.numel(), L.stream_ptr()))
        ctx.model = model
        ctx.save_for_backward(flat, coords2d)
        return out

    @staticmethod
    def backward(ctx, gout):
        flat, coords2d = ctx.saved_tensors
        model = ctx.model
        n = coords2d.shape[0]
        desc = model._desc()
        grads = torch.zeros_like(flat)
        opt = torch.zeros(L.OPT_STATE_BYTES + 8, dtype=torch.uint8, device=flat.device)
        buf = model._ws.get(desc, n, True)
        g = gout.reshape(-1).contiguous().float()
        L.check(L.lib().inr_siren_loss_backward(C.byref(desc), L.ptr(flat), L.ptr(coords2d), None, n, 0, n, 1.0,
                                                L.LOSS_MSE, 0.0, L.ptr(g), L.ptr(grads), L.ptr(opt), None, L.ptr(buf),
                                                buf.numel(), L.stream_ptr()))
        return grads, None, None


class _FlatSirenBase(nn.Module):
    """Shared machinery: flat parameter buffer + C-ABI forward."""

    precision = "fp32"

    def _finish_init(self, spec: SirenSpec):
        self.spec = spec
        self._ws = None
        self._flatten()

    def _param_list(self):
        return [p for _, p in self.named_parameters()]

    def _flatten(self):
        """Re-home every parameter as a view of one flat buffer (state_dict order)."""
        ps = self._param_list()
        dev = ps[0].device
        flat = torch.empty(sum(p.numel() for p in ps), dtype=torch.float32, device=dev)
        o = 0
        with torch.no_grad():
            for p in ps:
                n = p.numel()
                flat[o:o + n].copy_(p.detach().reshape(-1))
                p.data = flat[o:o + n].view(p.shape)
                o += n
        self._flat = flat
        self._ws = Workspace(dev) if dev.type == "cuda" else None

    def _apply(self, fn, *a, **k):
        r = super()._apply(fn, *a, **k)
        self._flatten()  # .cuda()/.to() re-allocate each parameter separately: restore the single arena
        return r

    def flat_parameters(self) -> torch.Tensor:
        """The flat fp32 parameter vector (a live view: writing to it updates the module)."""
        ps = self._param_list()
        base = self._flat
        if any(p.data_ptr() < base.data_ptr() or p.data_ptr() >= base.data_ptr() + base.numel() * 4 for p in ps):
            self._flatten()  # someone re-assigned .data (e.g. load_state_dict(assign=True))
        return self._flat

    def _desc(self) -> L.SirenDesc:
        return self.spec.desc(self.precision)

    def forward(self, coords):
        flat = self.flat_parameters()
        if not flat.is_cuda:
            raise L.InrLibraryError("Siren.forward: the model is on %s; the INR hot path runs only on a CUDA device "
                                    "(no CPU fallback) -- call .cuda() first as run.py:115 does" % flat.device)
        coords2d = coords.detach().reshape(-1, self.spec.in_features).contiguous().float()
        if torch.is_grad_enabled() and any(p.requires_grad for p in self._param_list()):
            out = self._fn_with_param_grads(flat, coords2d)
        else:
            out = _SirenFn.apply(flat, coords2d, self)
        return out.view(*coords.shape[:-1], self.spec.out_features)

    def _fn_with_param_grads(self, flat, coords2d):
        # autograd sees the individual nn.Parameters: rebuild the flat vector from them so .grad lands on each
        ps = self._param_list()
        flat_g = torch.cat([p.reshape(-1) for p in ps])
        return _SirenFn.apply(flat_g, coords2d, self)


class Siren(_FlatSirenBase):
    """models.py:243-275."""

    def __init__(self, in_features, hidden_features, hidden_layers, out_features, outermost_linear=False,
                 first_omega_0=30, hidden_omega_0=30.):
        super().__init__()
        net = [SineLayer(in_features, hidden_features, is_first=True, omega_0=first_omega_0)]
        for _ in range(hidden_layers):
            net.append(SineLayer(hidden_features, hidden_features, is_first=False, omega_0=hidden_omega_0))
        if outermost_linear:
            final_linear = nn.Linear(hidden_features, out_features)
            with torch.no_grad():
                bound = np.sqrt(6 / hidden_features) / hidden_omega_0
                final_linear.weight.uniform_(-bound, bound)
            net.append(final_linear)
        else:
            net.append(SineLayer(hidden_features, out_features, is_first=False, omega_0=hidden_omega_0))
        self.net = nn.Sequential(*net)
        self._finish_init(SirenSpec(in_features, hidden_features, hidden_layers, out_features, bool(outermost_linear),
                                    float(first_omega_0), float(hidden_omega_0)))


class Snake(nn.Module):
    """models.py:185-241: y = x + sin^2(a x)/a with a per-feature `a`, always trainable (the reference sets a
    misspelt `requiresGrad`, models.py:233, so `a` stays a trainable nn.Parameter).  Holds the parameter only:
    the activation itself runs inside the fused CUDA path (forward epilogue of the layer GEMM)."""

    def __init__(self, in_features, a=None, trainable=True):
        super().__init__()
        self.in_features = in_features if isinstance(in_features, list) else [in_features]
        if a is not None:
            self.a = nn.Parameter(torch.ones(self.in_features) * a)
        else:  # models.py:229-231: random init = mix of frequencies
            m = torch.distributions.exponential.Exponential(torch.tensor([0.1]))
            self.a = nn.Parameter((m.rsample(self.in_features)).squeeze())
        self.a.requiresGrad = trainable


class SirenWithSnakeTanh(_FlatSirenBase):
    """models.py:306-394 -- the class run.py:112 instantiates: layer 0 (SineLayer, or nn.Linear + Snake when
    first_linear), num_sine SineLayers, num_snake x (nn.Linear + Snake), num_tanh x (nn.Linear + Tanh), then the
    final nn.Linear (last_linear) or SineLayer.  Same module tree, state_dict keys and RNG consumption as the
    reference.  Sine-only stacks are identical to Siren(outermost_linear=last_linear); every layer mix runs in either
    precision (fp16: H in {128, 256, 512}, in_features <= 2)."""

    def __init__(self, in_features, out_features, hidden_features, num_sine, num_snake, num_tanh, first_linear=False,
                 last_linear=True, first_omega_0=30, hidden_omega_0=30., a_initial=50, num_freq=None, scale=2.0):
        super().__init__()
        net = []
        if first_linear:
            net.append(nn.Linear(in_features, hidden_features))
            net.append(Snake(hidden_features, a=a_initial))
        else:
            net.append(SineLayer(in_features, hidden_features, is_first=True, omega_0=first_omega_0))
        for _ in range(num_sine):
            net.append(SineLayer(hidden_features, hidden_features, is_first=False, omega_0=hidden_omega_0))
        for _ in range(num_snake):
            net.append(nn.Linear(hidden_features, hidden_features))
            net.append(Snake(hidden_features, a=a_initial))
        for _ in range(num_tanh):
            net.append(nn.Linear(hidden_features, hidden_features))
            net.append(nn.Tanh())
        if last_linear:
            final_linear = nn.Linear(hidden_features, out_features)
            with torch.no_grad():
                bound = np.sqrt(6 / hidden_features) / hidden_omega_0
                final_linear.weight.uniform_(-bound, bound)
            net.append(final_linear)
        else:
            net.append(SineLayer(hidden_features, out_features, is_first=False, omega_0=hidden_omega_0))
        self.net = nn.Sequential(*net)
        self._finish_init(SirenSpec(in_features, hidden_features, num_sine, out_features, bool(last_linear),
                                    float(first_omega_0), float(hidden_omega_0), int(num_snake), int(num_tanh),
                                    bool(first_linear)))
