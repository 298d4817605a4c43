#!/usr/bin/env python
"""Turn the ncu outputs of scripts/gpu_round.sh into the small text summaries committed under profiles/.

  python scripts/summarize_ncu.py list gpurun_out/launches.csv           > profiles/rNN_launches.txt
  python scripts/summarize_ncu.py full gpurun_out/prof_attn_tc.ncu-rep   > profiles/rNN_ncu_attn_tc.txt
"""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict

KEYS = [
    "gpu__time_duration.sum", "sm__cycles_active.avg", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tensor", "sm__pipe_tensor", "pipe_xu", "pipe_fma", "pipe_alu", "issue_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__occupancy_limit",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__cycles_active.avg",
    "warp_issue_stalled", "l1tex__data_bank_conflicts", "shared_mem", "sm__inst_executed.avg.per_cycle",
]


def short(name):
    name = re.sub(r"\(.*", "", name)
    name = re.sub(r"^void\s+", "", name)
    return name[:90]


def do_list(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr, rows = rows[0], rows[1:]
    ik, iv, im = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name")
    agg = OrderedDict()
    total = 0.0
    for r in rows:
        if r[im] != "gpu__time_duration.sum":
            continue
        t = float(r[iv].replace(",", "")) / 1e3  # ns -> us
        k = short(r[ik])
        c = agg.setdefault(k, [0, 0.0])
        c[0] += 1
        c[1] += t
        total += t
    print(f"# {path}: {sum(c[0] for c in agg.values())} launches, {total / 1e3:.2f} ms serialised (cold-cache, under ncu)")
    print(f"{'share':>7} {'ms':>9} {'count':>6} {'us/launch':>10}  kernel")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{100 * t / total:6.2f}% {t / 1e3:9.3f} {n:6d} {t / n:10.1f}  {k}")


def do_full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, rows = rows[0], rows[1], rows[2:]
    ik = hdr.index("Kernel Name")
    ig = hdr.index("Grid Size")
    for r in rows:
        print(f"## {short(r[ik])} grid {r[ig]}")
        for i, h in enumerate(hdr):
            if any(k in h for k in KEYS) and "device__attribute" not in h and ".min." not in h and ".max." not in h and r[i] not in ("", "no data"):
                print(f"  {h:<95} {r[i]:>16} {units[i]}")


if __name__ == "__main__":
    {"list": do_list, "full": do_full}[sys.argv[1]](sys.argv[2])
