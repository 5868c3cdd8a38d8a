"""Host-side engine of the SIREN fit path: owns the flat device buffers and drives libinr_b200.so.

SirenFitter replaces the reference's hot loop (run.py:156-190): one `step()` = forward + loss + backward +
Adam + ReduceLROnPlateau, entirely on the device (the loss history is read back on demand, not per step).
Single-GPU steps are captured once into a CUDA graph and replayed; with torch.distributed initialised and
`shard=True` the coordinates are sharded by rank and one sum-allreduce of [grads | loss] runs between the
backward and the optimiser.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from . import _lib as L


@dataclass
class SirenSpec:
    """Shape of the MLP: models.py:244-269 Siren, or the general models.py:310-394 SirenWithSnakeTanh stack
    (layer 0, `hidden_layers` sine layers, `num_snake` Linear+Snake layers, `num_tanh` Linear+Tanh layers, last)."""
    in_features: int = 1
    hidden_features: int = 256
    hidden_layers: int = 3          # number of HxH SineLayers (num_sine)
    out_features: int = 1
    outermost_linear: bool = True
    first_omega_0: float = 30.0
    hidden_omega_0: float = 30.0
    num_snake: int = 0
    num_tanh: int = 0
    first_linear: bool = False

    def desc(self, precision="fp32") -> L.SirenDesc:
        prec = L.PRECISIONS[precision] if isinstance(precision, str) else int(precision)
        return L.SirenDesc(self.in_features, self.hidden_features, self.hidden_layers, self.out_features,
                           int(bool(self.outermost_linear)), float(self.first_omega_0), float(self.hidden_omega_0), prec,
                           int(self.num_snake), int(self.num_tanh), int(bool(self.first_linear)))

    @property
    def total_hidden(self) -> int:
        return self.hidden_layers + self.num_snake + self.num_tanh

    def sine_only(self) -> bool:
        return self.num_snake == 0 and self.num_tanh == 0 and not self.first_linear

    def layer_kinds(self):
        """activation of layers 0..L: 'sine' | 'snake' | 'tanh'."""
        return (["snake" if self.first_linear else "sine"] + ["sine"] * self.hidden_layers +
                ["snake"] * self.num_snake + ["tanh"] * self.num_tanh)

    def layer_shapes(self):
        H = self.hidden_features
        return [(H, self.in_features)] + [(H, H)] * self.total_hidden + [(self.out_features, H)]

    def param_sizes(self):
        """flat sizes in state_dict order: weight, bias per layer, a Snake layer's `a` right after its bias."""
        H = self.hidden_features
        kinds = self.layer_kinds() + [None]
        out = []
        for (o, i), k in zip(self.layer_shapes(), kinds):
            out += [o * i, o]
            if k == "snake":
                out.append(H)
        return out

    def num_params(self) -> int:
        return sum(self.param_sizes())

    def flops_per_coord(self) -> int:
        """fwd + bwd dense flops per coordinate, SURVEY.md §8(d): 6 L H^2 + 10 H."""
        return 6 * self.total_hidden * self.hidden_features ** 2 + 10 * self.hidden_features


def shard_range(T: int, rank: int, world: int, tile: int = 128):
    """Coordinate sharding of ONE fit over `world` GPUs: rank r owns the contiguous global index range
    [i0, i0 + n), every shard a multiple of the 128-row tile except the last; coordinates are generated in-kernel from
    the GLOBAL index and the GLOBAL length, so they are bit-identical to the single-GPU run.  -> (i0, n, shard_size)"""
    per = (T + world - 1) // world
    per = (per + tile - 1) // tile * tile
    i0 = min(rank * per, T)
    return i0, max(0, min(per, T - i0)), per


def allreduce_step_state(grads: torch.Tensor, loss_acc: torch.Tensor, pg, world: int, whole_output_loss: bool):
    """The one exchange step of a sharded fit: sum-allreduce of the flat gradient and of the loss accumulators
    [time-loss numerator, spectral loss].  A whole-output loss was evaluated redundantly (identically) on every rank,
    so its sum is divided by the world size; the division also makes the value bit-identical on all ranks, which keeps
    the replicated plateau schedulers in lock-step."""
    if world <= 1:
        return
    torch.distributed.all_reduce(grads, group=pg)
    torch.distributed.all_reduce(loss_acc[0:2], group=pg)
    if whole_output_loss:
        loss_acc[1:2].div_(world)


def allreduce_fused(gbuf: torch.Tensor, P: int, loss_acc: torch.Tensor, pg, world: int, whole_output_loss: bool):
    """The same exchange as ONE collective (a single fused buffer, capturable in a CUDA graph): the
    two float64 loss accumulators ride behind the P gradient floats of `gbuf` (fp32, at least P + 4 long) as fp32
    hi + lo pairs (the reduced sum is fp32-grade, ~1e-7 relative: the loss the scheduler sees is a float32) -- so a step issues one sum-allreduce
    of P + 4 floats.  The result is bitwise identical on every rank, which keeps the replicas in lock-step."""
    if world <= 1:
        return
    tail = gbuf[P:P + 4]
    hi = loss_acc[0:2].to(torch.float32)
    tail[0:2].copy_(hi)
    tail[2:4].copy_((loss_acc[0:2] - hi.to(torch.float64)).to(torch.float32))
    torch.distributed.all_reduce(gbuf[:P + 4], group=pg)
    loss_acc[0:2].copy_(tail[0:2].to(torch.float64) + tail[2:4].to(torch.float64))
    tail.zero_()
    if whole_output_loss:
        loss_acc[1:2].div_(world)


def _require_cuda(t: torch.Tensor, what: str):
    if not t.is_cuda:
        raise L.InrLibraryError("%s must live on a CUDA device: the INR hot path has no CPU fallback" % what)


class Workspace:
    """Grow-only scratch buffer handed to the C ABI."""

    def __init__(self, device):
        self.device = device
        self.buf: Optional[torch.Tensor] = None

    def get(self, desc, n_rows: int, training: bool) -> torch.Tensor:
        fn = L.lib().inr_kan_workspace_bytes if isinstance(desc, L.KanDesc) else L.lib().inr_siren_workspace_bytes
        need = L.check(fn(C.byref(desc), n_rows, int(training)))
        if self.buf is None or self.buf.numel() < need:
            # zero-filled once, so that dz_saturated() below never reads bytes no kernel wrote
            self.buf = torch.zeros(max(int(need), 256), dtype=torch.uint8, device=self.device)
        return self.buf

    def count_fp16_saturated(self, chunk_bytes: int = 1 << 28) -> int:
        """Number of 16-bit words in the buffer that are +-65504, the value cvt.rn.satfinite clips an fp16 conversion to.
        In the tcgen05 modes the buffer holds fp16 images of sin/cos values (|x| <= 1), packed weights and the loss-scaled
        dZ_l, so after a training step any hit is a CLIPPED gradient element (csrc/tc_common.cuh pack_op2): the loss scale
        of csrc/siren_tc.cu gscale_compute was too large for that layer.  Diagnostic; synchronises."""
        if self.buf is None:
            return 0
        n = self.buf.numel() // 2 * 2
        total = 0
        for o in range(0, n, chunk_bytes):
            v = self.buf[o:min(o + chunk_bytes, n)].view(torch.int16)
            total += int(((v & 0x7FFF) == 0x7BFF).sum())
        return total


def siren_forward(desc: L.SirenDesc, params: torch.Tensor, coords: Optional[torch.Tensor], n: int, ws: Workspace,
                  i0: int = 0, T_total: int = 0, scale: float = 1.0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """models.py:271-275 through inr_siren_forward.  coords None -> generated in-kernel (in_features == 1)."""
    _require_cuda(params, "params")
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=params.device)
    buf = ws.get(desc, n, False)
    L.check(L.lib().inr_siren_forward(C.byref(desc), L.ptr(params), L.ptr(coords), n, i0, T_total, scale, L.ptr(out),
                                      L.ptr(buf), buf.numel(), L.stream_ptr()))
    return out


class SirenFitter:
    """Device-resident fit of one SIREN to one target signal.

    params0: flat fp32 parameter vector in state_dict order (CPU or CUDA; copied).
    target:  (T_local,) fp32 on the device -- this rank's shard when sharded.
    coords:  None -> linspace(-scale, scale, T_total)[i0 : i0+T_local] generated in the kernels (bit-exact with
             utils.py:99-109), else (T_local, in_features) on the device.
    """

    def __init__(self, spec: SirenSpec, params0: torch.Tensor, target: torch.Tensor, coords: Optional[torch.Tensor] = None,
                 *, i0: int = 0, T_total: Optional[int] = None, coord_scale: float = 1.0, loss_mode: str = "mse",
                 lr: float = 1e-3, min_lr: float = 1e-6, alpha: float = 0.0, precision: str = "fp32",
                 hist_cap: int = 65536, process_group=None, use_graph: bool = True, spectral=None,
                 spectral_target: Optional[torch.Tensor] = None, workspace: Optional["Workspace"] = None):
        _require_cuda(target, "target")
        self.spec = spec
        self.device = target.device
        self.desc = spec.desc(precision)
        self.precision = precision
        self.P = spec.num_params()
        assert params0.numel() == self.P, (params0.numel(), self.P)
        self.target = target.reshape(-1).contiguous().float()
        self.n = self.target.numel()
        self.i0 = int(i0)
        self.T_total = int(T_total if T_total is not None else self.n)
        self.coord_scale = float(coord_scale)
        self.coords = None if coords is None else coords.reshape(self.n, spec.in_features).contiguous().float()
        if loss_mode not in ("mse", "mae", "snr"):
            raise ValueError("loss_mode %r: the device path implements 'mse', 'mae' and 'snr' (run.py:161-169)" % loss_mode)
        self.loss_mode = loss_mode
        self.loss_kind = L.LOSS_MAE if loss_mode == "mae" else L.LOSS_MSE
        self.alpha = float(alpha)
        self.pg = process_group
        self.world = torch.distributed.get_world_size(process_group) if process_group is not None else 1
        self.rank = torch.distributed.get_rank(process_group) if process_group is not None else 0
        # losses that need the whole output (run.py:160,165): evaluated between the forward and backward halves on
        # the full signal; with coordinate sharding every rank all-gathers the outputs and evaluates them redundantly
        # (they cost microseconds), then back-propagates its own slice of the gradient.
        from . import spectral as SP
        full_t = self.target if spectral_target is None else spectral_target.reshape(-1).contiguous().float()
        if (spectral is not None or loss_mode == "snr") and self.world > 1 and spectral_target is None:
            raise ValueError("sharded fits with a spectral / SNR loss need spectral_target = the FULL target signal")
        self.spectral = SP.make_spectral_loss(spectral, full_t) if isinstance(spectral, (str, type(None))) else spectral
        self.snr = SP.SnrLoss(full_t) if loss_mode == "snr" else None
        self.split = self.spectral is not None or self.snr is not None
        if self.split:
            self.out = torch.empty(self.n, dtype=torch.float32, device=self.device)
            self.full_T = full_t.numel()
            self.gout_full = torch.zeros(self.full_T, dtype=torch.float32, device=self.device)
            self.out_full = self.out if self.world == 1 else torch.empty(self.full_T, dtype=torch.float32, device=self.device)
            if self.world > 1:
                # equal shard sizes are what all_gather_into_tensor needs; the bench/train shard that way
                self._shard = shard_range(self.full_T, 0, self.world)[2]
        dev = self.device
        # one arena: [params | grads | m | v], each padded to 64 floats so every block is 256-byte aligned
        self.Ppad = (self.P + 4 + 63) // 64 * 64        # + 4: the loss accumulators ride behind the gradients (allreduce_fused)
        self.arena = torch.zeros(4 * self.Ppad, dtype=torch.float32, device=dev)
        self.params = self.arena[0:self.P]
        self.grads = self.arena[self.Ppad:self.Ppad + self.P]
        self.m = self.arena[2 * self.Ppad:2 * self.Ppad + self.P]
        self.v = self.arena[3 * self.Ppad:3 * self.Ppad + self.P]
        self.params.copy_(params0.reshape(-1).to(dev, torch.float32))
        self.hist_cap = int(hist_cap)
        self.loss_hist = torch.zeros(self.hist_cap, dtype=torch.float32, device=dev)
        self.opt = torch.zeros((L.OPT_STATE_BYTES + 7) // 8 * 8, dtype=torch.uint8, device=dev)
        self.loss_acc = self.opt[L.OPT_LOSS_ACC_OFFSET:L.OPT_LOSS_ACC_OFFSET + 32].view(torch.float64)
        L.check(L.lib().inr_opt_init(L.ptr(self.opt), lr, min_lr, self.hist_cap, L.stream_ptr()))
        self.ws = workspace if workspace is not None else Workspace(dev)
        self.ws.get(self.desc, self.n, True)
        self.steps_done = 0
        # sharded fits are captured too: the one NCCL allreduce of a step is a graph node like the kernels around it
        self.use_graph = bool(use_graph)
        if self.world > 1:
            self._allreduce()                         # communicator set-up outside any capture (all-zero buffers)
        self._graph = None
        self.launches_per_step = None

    def dz_saturated(self) -> int:
        """fp16 gradient-image elements the LAST step clipped at +-65504 (0 = the loss scale had head-room everywhere);
        meaningful in the 'fp16' / 'fp16s' modes right after step() or loss_and_grads()."""
        return self.ws.count_fp16_saturated()

    def close(self):
        """Drop the captured CUDA graph and wait for the device.  A sharded fitter's graph holds the NCCL allreduce it
        captured: release every fitter (close() or del) BEFORE torch.distributed.destroy_process_group(), otherwise the
        communicator tear-down can hang on the live graph.  The fitter stays usable: the next step() re-captures."""
        self._graph = None
        if torch.cuda.is_available():
            torch.cuda.synchronize(self.device)

    def __del__(self):
        try:
            self._graph = None
        except Exception:
            pass

    def set_precision(self, precision: str):
        """Continue the same fit (parameters, Adam moments, step count, plateau scheduler, loss history) in another
        arithmetic mode: 'fp16' (plain fp16 tcgen05 operands: fastest, fit floor ~31 dB on gt_bach), 'fp16s' (fp16 hi+lo
        forward operands: fp32-grade forward, reaches the reference's converged SNR), 'fp32' (CUDA cores)."""
        if precision == self.precision:
            return
        self.desc = self.spec.desc(precision)
        self.precision = precision
        self.ws.get(self.desc, self.n, True)
        self._graph = None

    # -- one step -------------------------------------------------------------------------------
    def _w(self):
        return 1.0 - self.alpha

    def _enqueue_backward(self, extra_gout=None, out=None):
        buf = self.ws.get(self.desc, self.n, True)
        L.check(L.lib().inr_siren_loss_backward(
            C.byref(self.desc), L.ptr(self.params), L.ptr(self.coords), L.ptr(self.target), self.n, self.i0,
            self.T_total, self.coord_scale, self.loss_kind, self._w() / float(self.T_total), L.ptr(extra_gout),
            L.ptr(self.grads), L.ptr(self.opt), L.ptr(out), L.ptr(buf), buf.numel(), L.stream_ptr()))

    def _enqueue_adam(self):
        L.check(L.lib().inr_adam_step(L.ptr(self.params), L.ptr(self.grads), L.ptr(self.m), L.ptr(self.v), self.P,
                                      L.ptr(self.opt), self._w() / float(self.T_total), L.ptr(self.loss_hist),
                                      L.stream_ptr()))

    def _enqueue_split_backward(self):
        """forward (activations kept) -> whole-output losses -> backward with their gradient as the upstream term."""
        buf = self.ws.get(self.desc, self.n, True)
        lib = L.lib()
        L.check(lib.inr_siren_train_forward(C.byref(self.desc), L.ptr(self.params), L.ptr(self.coords), self.n, self.i0,
                                            self.T_total, self.coord_scale, L.ptr(self.out), L.ptr(buf), buf.numel(),
                                            L.stream_ptr()))
        if self.world > 1:
            self._gather_out()
        self.gout_full.zero_()
        if self.spectral is not None:
            self.spectral.accumulate(self.out_full, self.gout_full, self.opt, self.alpha)
        if self.snr is not None:
            self.snr.accumulate(self.out_full, self.gout_full, self.opt, self._w())
        gslice = self.gout_full[self.i0:self.i0 + self.n]
        tgt = None if self.snr is not None else self.target
        L.check(lib.inr_siren_train_backward(
            C.byref(self.desc), L.ptr(self.params), L.ptr(self.coords), L.ptr(tgt), self.n, self.i0, self.T_total,
            self.coord_scale, self.loss_kind, self._w() / float(self.T_total), L.ptr(gslice), L.ptr(self.grads),
            L.ptr(self.opt), L.ptr(buf), buf.numel(), L.stream_ptr()))

    def _gather_out(self):
        """all-gather the per-rank output shards into out_full (ranks own equal contiguous ranges, the last may be short)."""
        sh = self._shard
        pad = torch.zeros(sh, dtype=torch.float32, device=self.device)
        pad[:self.n].copy_(self.out)
        gathered = torch.empty(sh * self.world, dtype=torch.float32, device=self.device)
        torch.distributed.all_gather_into_tensor(gathered, pad, group=self.pg)
        self.out_full.copy_(gathered[:self.full_T])

    def _allreduce(self):
        """the one exchange of a sharded step: [grads | loss hi, lo] in a single NCCL sum-allreduce (allreduce_fused
        is the same arithmetic on torch tensors, used where there is no device: tests/test_sharding_cpu.py)"""
        if self.world <= 1:
            return
        gbuf = self.arena[self.Ppad:self.Ppad + self.P + 4]
        tail = gbuf[self.P:]
        L.check(L.lib().inr_opt_loss_to_f32(L.ptr(self.opt), L.ptr(tail), L.stream_ptr()))
        torch.distributed.all_reduce(gbuf, group=self.pg)
        L.check(L.lib().inr_opt_loss_from_f32(L.ptr(self.opt), L.ptr(tail), 1.0 / self.world if self.split else 1.0,
                                              L.stream_ptr()))

    def _enqueue_step(self):
        if self.split:
            self._enqueue_split_backward()
        else:
            self._enqueue_backward()
        self._allreduce()
        self._enqueue_adam()

    def step(self, n_steps: int = 1):
        """run.py:158-189, n_steps times, no host synchronisation."""
        if self.use_graph and self._graph is None:
            # warm the allocator / module load outside capture, on a side stream as torch requires
            s = torch.cuda.Stream(device=self.device)
            s.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(s):
                g = torch.cuda.CUDAGraph()
                lc0 = L.lib().inr_launch_count()
                with torch.cuda.graph(g, stream=s):
                    self._enqueue_step()
                # kernels of this library per replay (the capture ran the same host code once)
                self.launches_per_step = int(L.lib().inr_launch_count() - lc0)
            torch.cuda.current_stream(self.device).wait_stream(s)
            self._graph = g
        for _ in range(n_steps):
            if self._graph is not None:
                self._graph.replay()
            else:
                self._enqueue_step()
        self.steps_done += n_steps

    # -- state hand-off ---------------------------------------------------------------------------
    def set_opt(self, **fields):
        """Overwrite fields of the device optimiser state (lr, step, num_bad, best ...); the bias-correction
        scalars for the next step are recomputed in float64 exactly as the device does."""
        st = self.opt_state()
        for k, v in fields.items():
            setattr(st, k, v)
        if st.hist_base > st.step:          # the loss ring is indexed by step - hist_base: never let it go negative
            st.hist_base = st.step
        # (float)(1.0 - beta) formed in float64 as torch passes them to its kernels; k_adam reads these, not beta itself
        st.omb1 = 1.0 - float(st.beta1)
        st.omb2 = 1.0 - float(st.beta2)
        t = st.step + 1
        st.step_size = float(st.lr) / (1.0 - float(st.beta1) ** t)
        st.bc2_sqrt = math.sqrt(1.0 - float(st.beta2) ** t)
        raw = np.frombuffer(bytes(st), dtype=np.uint8).copy()
        self.opt[:raw.size].copy_(torch.from_numpy(raw).to(self.device))

    def load_adam_state(self, osd: dict):
        """Load a torch.optim.Adam.state_dict() (run.py:84-106 resume): exp_avg / exp_avg_sq / step / lr."""
        o, step = 0, 0
        sizes = self.spec.param_sizes()
        for i, n in enumerate(sizes):
            ent = osd["state"].get(i)
            if ent is not None:
                self.m[o:o + n].copy_(ent["exp_avg"].reshape(-1).to(self.device, torch.float32))
                self.v[o:o + n].copy_(ent["exp_avg_sq"].reshape(-1).to(self.device, torch.float32))
                step = int(ent["step"])
            o += n
        lr = float(osd["param_groups"][0]["lr"]) if osd.get("param_groups") else self.last_lr()
        # hist_base: the loss ring restarts at this fitter's first step, whatever Adam's step count is
        self.set_opt(step=step, lr=lr, hist_base=step)

    # -- read-backs (synchronising) -------------------------------------------------------------
    def opt_state(self) -> L.OptState:
        raw = self.opt.cpu().numpy().tobytes()
        return L.OptState.from_buffer_copy(raw[:L.OPT_STATE_BYTES])

    def losses(self) -> np.ndarray:
        """Per-step losses of the steps done so far (the last hist_cap of them), as run.py:171 would print."""
        n = self.steps_done
        h = self.loss_hist.cpu().numpy()
        if n <= self.hist_cap:
            return h[:n].copy()
        idx = np.arange(n - self.hist_cap, n) % self.hist_cap
        return h[idx]

    def last_lr(self) -> float:
        return float(self.opt_state().lr)

    def forward(self, coords: Optional[torch.Tensor] = None, n: Optional[int] = None, i0: Optional[int] = None,
                T_total: Optional[int] = None) -> torch.Tensor:
        """Model output at this fitter's own coordinates (default) or at the given ones."""
        if coords is None and n is None:
            coords, n, i0, T_total = self.coords, self.n, self.i0, self.T_total
        elif coords is not None:
            coords = coords.reshape(-1, self.spec.in_features).contiguous().float()
            n, i0, T_total = coords.shape[0], 0, coords.shape[0]
        return siren_forward(self.desc, self.params, coords, int(n), self.ws, int(i0 or 0), int(T_total or n),
                             self.coord_scale)

    def rounding_noise_ratio(self) -> float:
        """How much of the current fit error is the plain fp16 mode's own forward rounding: mean((out_fp16 - out_fp16s)^2)
        / mean((out_fp16s - target)^2) at the current parameters (two forward passes; summed over the ranks of a sharded
        fit).  Far below 1 the fp16 mode fits as well as the split mode; near 1 the fit sits on the fp16 noise floor
        (profiles/r02_c1_noise_fp16.jsonl: forward noise -31 dB of the signal, W rounding -33 dB, S rounding -36 dB)."""
        buf0 = self.ws.buf
        o16 = siren_forward(self.spec.desc("fp16"), self.params, self.coords, self.n, self.ws, self.i0, self.T_total,
                            self.coord_scale)
        o32 = siren_forward(self.spec.desc("fp16s"), self.params, self.coords, self.n, self.ws, self.i0, self.T_total,
                            self.coord_scale)
        sums = torch.stack([((o16 - o32).double() ** 2).sum(), ((o32 - self.target).double() ** 2).sum()])
        if self.pg is not None and self.world > 1:
            torch.distributed.all_reduce(sums, group=self.pg)
        self.ws.get(self.desc, self.n, True)      # back to this mode's training workspace
        if self.ws.buf is not buf0:
            self._graph = None                    # the scratch buffer grew: the captured step holds the old pointer
        num, den = (float(x) for x in sums.cpu())
        return num / max(den, 1e-300)

    def step_auto(self, n_steps: int, check_every: int = 250, threshold: float = 0.05, to: str = "fp16s"):
        """n_steps steps that start in the plain fp16 mode and move to the split mode `to` once the fp16 forward rounding
        noise is more than `threshold` of the fit error (checked every `check_every` steps: two forward passes, one host
        sync).  Returns the step index at which the fit switched (None if it stayed in fp16)."""
        switched = None
        done = 0
        while done < n_steps:
            k = min(check_every, n_steps - done) if self.precision == "fp16" else n_steps - done
            self.step(k)
            done += k
            if self.precision == "fp16" and done < n_steps and self.rounding_noise_ratio() > threshold:
                self.set_precision(to)
                switched = self.steps_done
        return switched

    def loss_and_grads(self, extra_gout: Optional[torch.Tensor] = None):
        """One forward+backward without the optimiser: (loss, flat grads clone, output) -- teacher-forced tests."""
        self.grads.zero_()
        self.loss_acc.zero_()
        if self.split:
            assert extra_gout is None
            self._enqueue_split_backward()
            out = self.out.clone()
        else:
            out = torch.empty(self.n, dtype=torch.float32, device=self.device)
            self._enqueue_backward(extra_gout, out)
        self._allreduce()
        acc = self.loss_acc.cpu().numpy()
        loss = float(acc[0]) * self._w() / float(self.T_total) + float(acc[1])
        g = self.grads.clone()
        self.grads.zero_()
        self.loss_acc.zero_()
        return loss, g, out


class ClipBatchFitter:
    """Clip-parallel batch fit (BASELINE.json config 5): K independent clips, each with its OWN SIREN,
    parameters, Adam state and plateau scheduler, fitted concurrently on one GPU with no communication.

    Every clip is a SirenFitter state (3.2 MB of arena at 3x256); the clips are dealt round-robin onto `n_streams`
    CUDA streams, each stream owning one activation workspace that its clips reuse in turn, and ONE CUDA graph holds
    a whole batch step (all clips, all streams, fork/join) so a step costs one launch from the host.  Across GPUs the
    clips are simply partitioned (`clip_range`): there is no collective on this path.
    """

    def __init__(self, spec: SirenSpec, params0: torch.Tensor, targets: torch.Tensor, *, precision: str = "fp32",
                 loss_mode: str = "mse", lr: float = 1e-3, min_lr: float = 1e-6, n_streams: int = 8,
                 hist_cap: int = 4096, use_graph: bool = True, sm_partition: bool = True):
        _require_cuda(targets, "targets")
        assert targets.dim() == 2 and params0.dim() == 2 and params0.shape[0] == targets.shape[0]
        # sm_partition: every stream's kernels are sized for 1/n_streams of the SMs (inr_set_sm_limit), so the streams'
        # clips run side by side on disjoint SMs (longer tile loops per CTA, launch gaps overlapped) instead of
        # time-slicing the whole GPU with one short persistent kernel after the other
        self.sm_partition = bool(sm_partition)
        self.K, self.T = targets.shape
        self.device = targets.device
        self.spec = spec
        self.n_streams = max(1, min(int(n_streams), self.K))
        self.workspaces = [Workspace(self.device) for _ in range(self.n_streams)]
        self.fitters = [SirenFitter(spec, params0[k], targets[k], None, loss_mode=loss_mode, lr=lr, min_lr=min_lr,
                                    precision=precision, hist_cap=hist_cap, use_graph=False,
                                    workspace=self.workspaces[k % self.n_streams]) for k in range(self.K)]
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(self.n_streams)]
        self.use_graph = bool(use_graph)
        self._graph = None
        self.steps_done = 0
        self.launches_per_step = None

    @staticmethod
    def clip_range(num_clips: int, rank: int, world: int):
        """Contiguous block partition of the clips over ranks (the first num_clips % world ranks get one more)."""
        base, rem = divmod(num_clips, world)
        lo = rank * base + min(rank, rem)
        return lo, lo + base + (1 if rank < rem else 0)

    def _enqueue_batch_step(self, main: torch.cuda.Stream):
        for s in self.streams:
            s.wait_stream(main)
        if self.sm_partition and self.n_streams > 1:
            sms = torch.cuda.get_device_properties(self.device).multi_processor_count
            L.check(L.lib().inr_set_sm_limit(max(8, sms // self.n_streams)))
        try:
            for si, s in enumerate(self.streams):
                with torch.cuda.stream(s):
                    for k in range(si, self.K, self.n_streams):
                        self.fitters[k]._enqueue_step()
        finally:
            L.check(L.lib().inr_set_sm_limit(0))
        for s in self.streams:
            main.wait_stream(s)

    def close(self):
        """Drop the captured graph (and the per-clip fitters' graphs) and wait for the device; see SirenFitter.close."""
        self._graph = None
        for f in self.fitters:
            f.close()

    def step(self, n_steps: int = 1):
        cur = torch.cuda.current_stream(self.device)
        if self.use_graph and self._graph is None:
            cap = torch.cuda.Stream(device=self.device)
            cap.wait_stream(cur)
            with torch.cuda.stream(cap):
                g = torch.cuda.CUDAGraph()
                lc0 = L.lib().inr_launch_count()
                with torch.cuda.graph(g, stream=cap):
                    self._enqueue_batch_step(cap)
                self.launches_per_step = int(L.lib().inr_launch_count() - lc0)
            cur.wait_stream(cap)
            self._graph = g
        for _ in range(n_steps):
            if self._graph is not None:
                self._graph.replay()
            else:
                self._enqueue_batch_step(cur)
        self.steps_done += n_steps
        for f in self.fitters:
            f.steps_done += n_steps

    def losses(self) -> np.ndarray:
        """(K, steps) per-clip loss histories."""
        return np.stack([f.losses() for f in self.fitters])

    def snr_db(self, targets: torch.Tensor) -> np.ndarray:
        """utils.py:77-97 calculate_snr(target, model output) per clip, evaluated on the device."""
        out = np.empty(self.K)
        for k, f in enumerate(self.fitters):
            y = f.forward().double()
            t = targets[k].double()
            out[k] = float(10.0 * torch.log10((t * t).mean() / ((y - t) ** 2).mean()))
        return out
