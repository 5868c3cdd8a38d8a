"""Diagnostic (run under torchrun, N >= 2): where does the time of a SMALL sharded step go?  C1-shaped fit (3x256, T = 264 600)
sharded over the ranks: step time with the graph-captured exchange, un-graphed, and with the allreduce skipped (replicas
diverge: timing only).  Rank 0 prints one JSON line."""
import json, os, sys, time
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import inr_for_audio_b200 as P
from inr_for_audio_b200.engine import shard_range

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
T = 264600
gt = torch.from_numpy(P.utils.synth_waveform(T, seed=0)).to(dev)
i0, n, _ = shard_range(T, rank, world)
torch.manual_seed(0)
m = P.Siren(1, 256, 3, 1, True)


def run(tag, use_graph, skip_ar, steps=300):
    f = P.SirenFitter(m.spec, m.flat_parameters(), gt[i0:i0 + n], i0=i0, T_total=T, precision="fp16s",
                      process_group=dist.group.WORLD, use_graph=use_graph)
    if skip_ar:
        f._allreduce = lambda: None
    f.step(20)
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); f.step(steps); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    del f
    return ms


out = {"world": world, "rows_per_rank": n}
out["graph"] = run("graph", True, False)
out["no_graph"] = run("nograph", False, False)
out["graph_no_allreduce"] = run("graph_noar", True, True)
# the collective alone, back to back, on a buffer of the same size
buf = torch.zeros(m.spec.num_params() + 4, device=dev)
for _ in range(20):
    dist.all_reduce(buf)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200):
    dist.all_reduce(buf)
e1.record(); torch.cuda.synchronize()
out["allreduce_alone_ms"] = e0.elapsed_time(e1) / 200
if rank == 0:
    print(json.dumps(out), flush=True)
dist.barrier()
import gc, threading
t = threading.Timer(30.0, lambda: os._exit(0)); t.daemon = True; t.start()
gc.collect(); torch.cuda.synchronize()
dist.destroy_process_group()
