#!/usr/bin/env python
"""Print a one-screen summary of bench.py JSON lines: python tools/bench_summary.py gpurun_out/a.json [...]"""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as ex:  # noqa: BLE001
        print(f, "ERR", ex)
        continue
    ck = d.get("clocks") or {}
    print("%s value=%.4g ms/step=%.3f e2e=%.4g frac=%.3f sm=%s %s launches=%s" % (
        f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], ck.get("sm_mhz"), ck.get("reasons"),
        d.get("gpu_launches")))
    for k, v in sorted(d.get("kernels", {}).items(), key=lambda kv: -kv[1]["ms_per_step"])[:9]:
        print("   %-24s %5.1f x %.3f ms" % (k, v["launches_per_step"], v["ms_per_step"]))
