#!/usr/bin/env python
"""Summarise ncu output (read on the CPU box) into small committed files under profiles/.

  python profiles/summarize_ncu.py launches <launches.csv> <out.md>      # per-kernel share of the step
  python profiles/summarize_ncu.py full <report.ncu-rep> <out.md> [--traffic-json profiles/traffic.json]
"""
import csv
import io
import json
import subprocess
import sys
from collections import defaultdict

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_write.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum",
    "lts__t_sectors_op_write.sum", "lts__t_sectors_op_read.sum", "smsp__inst_executed.sum",
    "smsp__average_warp_latency_issue_stalled_barrier.ratio", "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_membar.ratio",
]


def short(name):
    """'void tile_build_kernel<0, 1>(const uint4 *, ...)' -> 'tile_build_kernel' (the name mv3d_kernel_name reports)."""
    name = name.split("(")[0].split("<")[0].strip()
    return name[5:] if name.startswith("void ") else name


def launches(path, out):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    per = defaultdict(list)
    for r in rows[1:]:
        per[short(r[ik])].append(float(r[iv].replace(",", "")))
    total = sum(sum(v) for v in per.values())
    with open(out, "w") as f:
        f.write(f"# ncu launch list ({path})\n\n`ncu --metrics gpu__time_duration.sum --clock-control none` -- cold-cache, "
                "serialised launches: compare SHARES, not absolutes.\n\n| kernel | launches | mean us | share of all launches |\n|---|---|---|---|\n")
        for k, v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
            f.write(f"| {k} | {len(v)} | {sum(v) / len(v) / 1e3:.1f} | {sum(v) / total * 100:.1f}% |\n")
    print(open(out).read())


def full(path, out, traffic_json=None):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw[raw.index('"ID"'):])))
    hdr, units = rows[0], rows[1]
    ik = hdr.index("Kernel Name")
    traffic = {}
    with open(out, "w") as f:
        f.write(f"# ncu --set full summary ({path})\n\n")
        for r in rows[2:]:
            f.write(f"## {short(r[ik])} (launch id {r[0]})\n\n| metric | value | unit |\n|---|---|---|\n")
            vals = {}
            for k in KEYS:
                if k in hdr:
                    i = hdr.index(k)
                    vals[k] = r[i]
                    f.write(f"| {k} | {r[i]} | {units[i]} |\n")
            f.write("\n")

            def to_bytes(k):
                i = hdr.index(k)
                mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
                return float(r[i].replace(",", "")) * mult
            traffic.setdefault(short(r[ik]), []).append(to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum"))
    print(open(out).read())
    if traffic_json:
        try:
            cur = json.load(open(traffic_json))
        except Exception:
            cur = {}
        cur.update({k: sum(v) / len(v) for k, v in traffic.items()})
        json.dump(cur, open(traffic_json, "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    else:
        tj = sys.argv[sys.argv.index("--traffic-json") + 1] if "--traffic-json" in sys.argv else None
        full(sys.argv[2], sys.argv[3], tj)
