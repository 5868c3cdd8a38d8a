#!/usr/bin/env python
"""Summarise an `ncu --page raw --csv` export into a markdown table + a per-kernel DRAM-traffic JSON.

    python tools/ncu_summary.py gpurun_out/r01b_full_raw.csv profiles/r01b_ncu_full_summary.md profiles/r01b_traffic.json
"""
import csv
import json
import re
import sys

COLS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU (MUFU) %"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %"),
    ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2->SM bytes"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs"),
    ("launch__shared_mem_per_block_dynamic", "dyn smem"),
    ("sm__cycles_elapsed.max", "cycles"),
]


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*", "", name)
    return name.replace("inr::", "").replace("tc::", "")


def main():
    src, md, js = sys.argv[1:4]
    rows = list(csv.reader(open(src)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    idx = [(hdr.index(c), lab) for c, lab in COLS if c in hdr]
    out = ["| kernel | " + " | ".join("%s (%s)" % (lab, units[i]) if units[i] else lab for i, lab in idx) + " |",
           "|---|" + "---|" * len(idx)]
    agg = {}
    for d in data:
        k = short(d[ki])
        vals = []
        for i, lab in idx:
            try:
                vals.append("%.4g" % float(d[i]))
            except ValueError:
                vals.append(d[i])
        out.append("| `%s` | " % k + " | ".join(vals) + " |")
        a = agg.setdefault(k, {"launches": 0, "dram_bytes": 0.0, "ms": 0.0})
        a["launches"] += 1
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
        for c in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            j = hdr.index(c)
            a["dram_bytes"] += float(d[j]) * scale.get(units[j], 1.0)
        j = hdr.index("gpu__time_duration.sum")
        a["ms"] += float(d[j]) * {"ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}.get(units[j], 1.0)
    for k, a in agg.items():
        a["dram_bytes_per_launch"] = a.pop("dram_bytes") / a["launches"]
        a["ms_per_launch_under_ncu"] = a.pop("ms") / a["launches"]
    open(md, "a").write("\n".join(out) + "\n")
    json.dump(agg, open(js, "w"), indent=1, sort_keys=True)
    print(json.dumps(agg, indent=1))


if __name__ == "__main__":
    main()
