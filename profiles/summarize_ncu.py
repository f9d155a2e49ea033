#!/usr/bin/env python
"""Summarise ncu output (read on the CPU box) into small committed files under profiles/.

  python profiles/summarize_ncu.py launches <launches.csv> <out.md>      # per-kernel share of the step
  python profiles/summarize_ncu.py full <report.ncu-rep> <out.md> [--traffic-json profiles/traffic.json]
  python profiles/summarize_ncu.py strategies <report.ncu-rep> <out.md> [--traffic-json profiles/traffic.json]
      # the report of tools/strategy_profile.py: one table per (config, strategy) with the counters DESIGN.md 5.1 cites
"""
import csv
import io
import json
import os
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
    "lts__t_sectors_srcunit_tex_op_atom.sum", "lts__t_sectors_srcunit_tex_op_red.sum",
    "lts__d_atomic_input_cycles_active.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed_op_shared_atom.sum",
    "lts__t_sectors_op_write.sum", "lts__t_sectors_op_read.sum", "smsp__inst_executed.sum",
    "smsp__average_warp_latency_issue_stalled_barrier.ratio", "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_membar.ratio",
]


def short(name):
    """'void tile_build_kernel<0, 1>(const uint4 *, ...)' -> 'tile_build_kernel' (the name mv3d_kernel_name reports)."""
    name = name.split("(")[0].split("<")[0].strip()
    name = name[5:] if name.startswith("void ") else name
    return name[6:] if name.startswith("mv3d::") else name


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
        commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
        for k, v in traffic.items():  # same record as the strategies mode writes: bench.py prints commit and capture
            cur[k] = {"bytes": sum(v) / len(v), "commit": commit, "ncu_file": os.path.basename(path),
                      "workload": "config 2, one launch of 64 sweeps (bench.py's step)", "strategy": "tile"}
        json.dump(cur, open(traffic_json, "w"), indent=1)


def strategies(path, out, traffic_json=None):
    """tools/strategy_profile.py runs config 2 then config 5, each through tile, atomic, sort, one launch of every
    kernel per pair. Groups are recovered from the launch order: a tile group opens with classify_kernel, a sort group
    with sort_keys_kernel, an atomic group with the first fill_zero_kernel that does not belong to a sort group."""
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw[raw.index('"ID"'):])))
    hdr, units = rows[0], rows[1]
    ik = hdr.index("Kernel Name")

    def num(r, k, scale=True):
        if k not in hdr:
            return 0.0
        i = hdr.index(k)
        mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(units[i], 1) if scale else 1
        try:
            return float(r[i].replace(",", "")) * mult
        except ValueError:
            return 0.0

    groups, cur, kind = [], None, None
    for r in rows[2:]:
        name = short(r[ik])
        start = None
        if name == "classify_kernel":
            start = "tile"
        elif name == "sort_keys_kernel":
            start = "sort"
        elif name == "fill_zero_kernel" and kind != "sort" and not (cur and short(cur[-1][ik]) == "fill_zero_kernel"):
            start = "atomic"
        if start:
            cur, kind = [], start
            groups.append((start, cur))
        if cur is not None:
            cur.append(r)
    per_config = len(groups) // 2
    commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    traffic = {}
    with open(out, "w") as f:
        f.write(f"# ncu --set full, one launch of every kernel of the three reduction strategies ({path})\n\n"
                f"Captured by `tools/strategy_profile.py` at commit {commit}; `--clock-control none`, launches serialised and "
                "cold-cache (compare counters and shares, not absolute times). L2 atom / red sectors = `lts__t_sectors_srcunit_tex_op_atom.sum` / "
                "`lts__t_sectors_srcunit_tex_op_red.sum` (L2 sectors touched by atomics that return a value / by reductions; this ncu "
                "names them `srcunit_tex`, older ones `lts__t_sectors_op_atom/red`), L2 atomic unit busy = "
                "`lts__d_atomic_input_cycles_active.avg.pct_of_peak_sustained_elapsed`, shared-memory atomic instr = "
                "`smsp__inst_executed_op_shared_atom.sum`, bank = "
                "`l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum`, dram% = `gpu__dram_throughput...pct_of_peak`, "
                "sm% = `sm__throughput...pct_of_peak`.\n\n")
        for gi, (kind, g) in enumerate(groups):
            config = "config 2 (64 sweeps x 120 k points, BEV 704x800x27 + FV)" if gi < per_config else "config 5 (8 clouds x 2 M points, BEV 1408x1600x10)"
            f.write(f"## {config} -- {kind.upper()}\n\n| kernel | us | dram read MB | dram write MB | dram GB/s | L2 atom sectors | L2 red sectors | L2 atomic unit busy % | shared-memory atomic instr | bank conflicts | dram% | sm% | regs |\n|---|---|---|---|---|---|---|---|---|---|---|---|---|\n")
            tot_us = tot_r = tot_w = tot_atom = tot_red = 0.0
            for r in g:
                us = num(r, "gpu__time_duration.sum")
                rd, wr = num(r, "dram__bytes_read.sum"), num(r, "dram__bytes_write.sum")
                atom = num(r, "lts__t_sectors_srcunit_tex_op_atom.sum") or num(r, "lts__t_sectors_op_atom.sum")
                red = num(r, "lts__t_sectors_srcunit_tex_op_red.sum") or num(r, "lts__t_sectors_op_red.sum")
                l2a = num(r, "lts__d_atomic_input_cycles_active.avg.pct_of_peak_sustained_elapsed", False)
                satom = num(r, "smsp__inst_executed_op_shared_atom.sum")
                bank = num(r, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")
                f.write(f"| {short(r[ik])} | {us:.1f} | {rd / 1e6:.1f} | {wr / 1e6:.1f} | {(rd + wr) / us / 1e3 if us else 0:.0f} | {atom:.3g} | {red:.3g} | {l2a:.1f} | {satom:.3g} | {bank:.3g} | "
                        f"{num(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', False):.0f} | {num(r, 'sm__throughput.avg.pct_of_peak_sustained_elapsed', False):.0f} | {num(r, 'launch__registers_per_thread', False):.0f} |\n")
                tot_us += us; tot_r += rd; tot_w += wr; tot_atom += atom; tot_red += red
                if gi < per_config:
                    traffic[short(r[ik])] = {"bytes": rd + wr, "commit": commit, "ncu_file": os.path.basename(path),
                                             "workload": "config 2, one launch of 64 sweeps", "strategy": kind}
            f.write(f"| **all** | **{tot_us:.1f}** | {tot_r / 1e6:.1f} | {tot_w / 1e6:.1f} | {(tot_r + tot_w) / tot_us / 1e3:.0f} | {tot_atom:.3g} | {tot_red:.3g} | | | | | | |\n\n")
    print(open(out).read())
    if traffic_json:
        try:
            cur = json.load(open(traffic_json))
        except Exception:
            cur = {}
        cur = {k: v for k, v in cur.items() if isinstance(v, dict)}
        # keep the tile strategy's entry where two strategies share a kernel name (fill_zero_kernel, front_finalize)
        cur.update(traffic)
        json.dump(cur, open(traffic_json, "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    elif sys.argv[1] == "strategies":
        tj = sys.argv[sys.argv.index("--traffic-json") + 1] if "--traffic-json" in sys.argv else None
        strategies(sys.argv[2], sys.argv[3], tj)
    else:
        tj = sys.argv[sys.argv.index("--traffic-json") + 1] if "--traffic-json" in sys.argv else None
        full(sys.argv[2], sys.argv[3], tj)
