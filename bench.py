#!/usr/bin/env python
"""bench.py -- coord-samples/sec of the SIREN fit step (fwd + loss + bwd + Adam [+ allreduce]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c5clip]
                    [--precision bf16|fp32]

N > 1 is launched by the driver under torch.distributed.run (one rank per GPU, NCCL); the coordinates of the ONE
waveform are sharded by rank and every step has one sum-allreduce of [grads | loss]  => "scaling": "strong".
Prints ONE JSON line (rank 0).  See the task contract for the keys; additions: roofline, cpu_baseline, e2e,
gpu_launches, clocks, kernels (per-kernel ms/step from the event profiler of libinr_b200).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[1]: SIREN 5x512 on a synthetic 60 s 44.1 kHz waveform -- the config the metric is quoted on
    "c2": dict(T=2646000, H=512, L=5, seed=0, desc="SIREN 5x512, synthetic 60 s @ 44.1 kHz (T=2646000), mse, Adam lr 1e-3"),
    # configs[0] model shape on a synthetic clip of gt_bach's length (parity case; not a bench line by default)
    "c1": dict(T=264600, H=256, L=3, seed=0, desc="SIREN 3x256, synthetic 6 s @ 44.1 kHz (T=264600), mse, Adam lr 1e-3"),
    "c5clip": dict(T=88200, H=256, L=3, seed=0, desc="SIREN 3x256, one synthetic 2 s clip (T=88200)"),
}
METRIC = "coord-samples/sec (fwd+bwd+Adam) SIREN fit"
UNIT = "coord-samples/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--precision", default=os.environ.get("INR_BENCH_PRECISION", "auto"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=20.0)
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=float(p["hbm_gbs"]), bf16=float(p["bf16_tflops"]), bf16_sus=float(p["bf16_tflops_sustained"]),
                    src="measured (MEASURED_PEAKS.json)")
    except Exception:
        return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, src="fallback (B200_PROFILING.md)")


# ------------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi sampled every 200 ms during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
                except ValueError:
                    continue
                for nm, v in zip(names, f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            os.unlink(self.path)
        except Exception:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------- reference arm
def oracle_fit_rate(wl, rows, steps, warmup, threads):
    """The reference's own algorithm for the path on the host cores: oracle/inr_oracle.py (a line-for-line
    restatement of run.py:116-117,156-190 around models.Siren, pinned against the reference's outputs) on the
    first `rows` coordinates of the workload.  Returns (coord-samples/s, seconds per step)."""
    import torch
    from oracle import inr_oracle as O
    from inr_for_audio_b200.utils import synth_waveform
    torch.set_num_threads(threads)
    spec = O.SirenSpec(1, wl["H"], wl["L"], 1, True)
    params = O.siren_init(spec, seed=wl["seed"])
    gt_full = synth_waveform(wl["T"], seed=wl["seed"])
    coords = O.get_coord(wl["T"], 1)[:rows].view(1, -1, 1)
    gt = torch.from_numpy(gt_full[:rows]).view(1, -1, 1)
    st = O.AdamState(params, lr=1e-3)
    sched = O.PlateauState()
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        loss, _, grads, _ = O.siren_loss_and_grads(spec, params, coords, gt)
        O.adam_step(params, grads, st)
        sched.step(float(loss), st)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    per = float(np.mean(times))
    return rows / per, per


def pick_rows(wl, budget_s, n_steps, threads):
    """Rows per step so that n_steps take about budget_s on this host (probe with a small slice first)."""
    probe = min(wl["T"], 8192)
    _, per = oracle_fit_rate(wl, probe, 1, 1, threads)
    rows = int(probe * budget_s / max(per * n_steps, 1e-9))
    rows = max(4096, min(wl["T"], rows))
    return rows // 1024 * 1024 if rows < wl["T"] else rows


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n_steps = args.steps + args.warmup
    rows = pick_rows(wl, min(150.0, 8.0 * n_steps), n_steps, threads)
    rate, per = oracle_fit_rate(wl, rows, args.steps, args.warmup, threads)
    sample = "first %d of %d coordinates per step (full model), %d timed steps" % (rows, wl["T"], args.steps)
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": per * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "note": "CPU reference arm: oracle port of run.py's loop on host cores"},
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------- our arm
def run_ours(args, wl):
    import torch
    import torch.distributed as dist
    import inr_for_audio_b200 as P
    from inr_for_audio_b200 import _lib as L

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    pg = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        pg = dist.group.WORLD
    T, H, Lh = wl["T"], wl["H"], wl["L"]

    prec = args.precision
    if prec == "auto":
        try:
            L.check(L.lib().inr_siren_workspace_bytes(P.SirenSpec(1, H, Lh, 1, True).desc("bf16"), 1024, 1))
            prec = "bf16"
        except L.InrLibraryError:
            prec = "fp32"

    # data + model: synthetic waveform, reference-seeded init (random-init weights of the named architecture)
    gt_np = P.utils.synth_waveform(T, seed=wl["seed"])
    torch.manual_seed(wl["seed"])
    model = P.Siren(1, H, Lh, 1, True)
    spec = model.spec
    # contiguous shard per rank, rounded to the 128-row tile
    per = (T + world - 1) // world
    per = (per + 127) // 128 * 128
    i0 = min(rank * per, T)
    n = max(0, min(per, T - i0))
    host_target = torch.from_numpy(gt_np[i0:i0 + n]).pin_memory()
    target = host_target.to(dev, non_blocking=True)

    def make(use_graph):
        return P.SirenFitter(spec, model.flat_parameters(), target, None, i0=i0, T_total=T, precision=prec,
                             process_group=pg, use_graph=use_graph, hist_cap=4096)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(k)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- device-resident throughput ("value")
    fit = make(True)
    fit.step(args.warmup)
    lc0 = L.lib().inr_launch_count()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(fit.step, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    value = T * args.steps / (ms * 1e-3)
    losses = fit.losses()

    # ---- launches per step (counted on an un-graphed step: graph replays bypass the counter)
    fit_ng = make(False)
    fit_ng.step(1)
    torch.cuda.synchronize(dev)
    lc0 = L.lib().inr_launch_count()
    fit_ng.step(1)
    launches_per_step = int(L.lib().inr_launch_count() - lc0)

    # ---- per-kernel device time with the library's event profiler (separate pass, same steps, no graph)
    kern = {}
    prof_steps = max(1, min(args.steps, 5))
    fit_ng.step(2)
    barrier()
    L.profile_begin()
    fit_ng.step(prof_steps)
    prof = L.profile_end()
    for tag, (cnt, tot) in prof.items():
        kern[tag] = {"launches_per_step": cnt / prof_steps, "ms_per_step": tot / prof_steps}
    gemm_tags = [t for t in kern if ("sgemm" in t or t.startswith("k_tc_"))]
    gemm_ms = sum(kern[t]["ms_per_step"] for t in gemm_tags)
    gemm_launches = sum(kern[t]["launches_per_step"] for t in gemm_tags)
    pk = peaks()
    hidden_flops = 6.0 * Lh * H * H * n                   # this rank's rows: fwd + dX + dW of the L hidden layers
    achieved = hidden_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    roofline = {"bound": "tensor", "achieved": achieved, "peak": pk["bf16_sus"], "unit": "TFLOP/s",
                "frac": achieved / pk["bf16_sus"], "traffic": None,
                "kernel": "+".join(sorted(gemm_tags)), "flops_per_launch": hidden_flops / max(gemm_launches, 1),
                "avg_launch_ms": gemm_ms / max(gemm_launches, 1), "gemm_share_of_step": gemm_ms / (ms / args.steps),
                "peak_source": pk["src"] + ", sustained bf16 (kernel timed inside a long step)",
                "mma_kind": "tcgen05 kind::f16 (bf16)" if prec == "bf16" else "none: fp32 FFMA on CUDA cores (parity mode)"}

    # ---- end to end through the public API with HOST buffers: H2D of the step's target, D2H of its loss
    fit_e = make(True)
    fit_e.step(args.warmup)
    torch.cuda.synchronize(dev)

    def e2e_steps(k):
        for _ in range(k):
            fit_e.target.copy_(host_target, non_blocking=True)
            fit_e.step(1)
            _ = fit_e.loss_hist[(fit_e.steps_done - 1) % fit_e.hist_cap].item()
    e2e_steps(1)
    barrier()
    t0 = time.perf_counter()
    e2e_steps(args.steps)
    barrier()
    dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e = {"value": T * args.steps / float(dt.item()), "unit": UNIT, "h2d_bytes_per_step": int(n * 4),
           "d2h_bytes_per_step": 4}

    # ---- CPU baseline beside it (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        rows = pick_rows(wl, args.cpu_budget_s, 4, threads)
        rate, per_s = oracle_fit_rate(wl, rows, 3, 1, threads)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "oracle (torch-CPU fp32 restatement of run.py's loop) on the first %d of %d coordinates, "
                         "3 timed steps after 1 warm-up, %.2f s/step" % (rows, T, per_s)}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "bf16" if prec == "bf16" else "f32", "data": "synthetic",
                "config": {"workload": wl["desc"], "precision": prec, "parallelism": "coord-shard x%d" % world,
                           "rows_per_rank": n, "l2": "working set (activation workspace, GBs) >> 126 MB L2; no flush needed",
                           "cuda_graph": True},
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_step * args.steps,
                "roofline": roofline, "cpu_baseline": cpu, "kernels": kern,
                "loss_first_last": [float(losses[0]), float(losses[-1])] if len(losses) else None}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
