#!/usr/bin/env python
"""bench.py -- sweeps/sec of the LiDAR view encoder (BEV + FV, 120k-point KITTI-shaped sweeps) on N B200s.

  python bench.py --gpus N --steps K --warmup W            # our arm (CUDA, through the C ABI)
  python bench.py --impl reference --gpus N --steps K ...   # the reference's own C routine on the host cores

A step = one pass of the hot path over one batch of 64 synthetic sweeps per GPU (BASELINE.json configs[1]):
BEV 704x800x27 + FV 64x512x3 for every sweep. Frames are sharded round-robin over ranks (frame i -> rank
i mod N), no collective on the data path, so scaling is weak: every rank encodes 64 sweeps per step.

  value     sweeps/s with inputs already resident in HBM (CUDA events on the launching stream, max over ranks)
  e2e       sweeps/s through mv3d_encode_batch_host with pinned HOST buffers: H2D of the points, encode, D2H of
            both encoded tensors, every step
  roofline  the dominant kernel (tile_build_kernel) against measured HBM bandwidth, timed live with CUDA
            events recorded by the library around each launch inside the timed region
  cpu_baseline  the reference's unmodified C code (oracle/_ref) on this box's host cores, rank 0, N=1 only

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH = 64  # sweeps per GPU per step
METRIC = "sweeps/sec (BEV+FV encode, 120k pts)"
WORKLOAD = "configs[1]: batch of 64 KITTI-shaped sweeps (120k pts), BEV 704x800x27 + FV 64x512x3"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--strategy", default="auto", choices=["auto", "tile", "patch", "atomic", "sort"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiling runs)")
    ap.add_argument("--no-also", action="store_true", help="skip the short legs for the widened rows (profiling runs)")
    return ap.parse_args()


def algorithmic_bytes(top, front, n_points):
    """every input point read once, every output element written once."""
    bev = 4 * top.xn * top.yn * (top.zn + 2)
    fv = 4 * front.rn * front.cn * front.fn
    return {"points": 16 * n_points, "bev": bev, "fv": fv, "total": 16 * n_points + bev + fv}


def measured_hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel):
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(kernel)
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms while the timed region runs."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); smax.append(float(parts[1])); power.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------------
# CPU legs (the only places bench.py executes oracle/): cpu_baseline and --impl reference
# ------------------------------------------------------------------------------------------------------
def cpu_baseline(top, front):
    """Bounded sample of the same workload on the host cores: every core encodes 8 config-1 sweeps."""
    from oracle.cpu_bench import CpuEncoderPool
    pool = CpuEncoderPool(top, front)
    try:
        warm = list(range(pool.cores * 2))
        pool.run([f % BATCH for f in warm])
        frames = [f % BATCH for f in range(pool.cores * 8)]
        wall, per_call = pool.run(frames)
        # SURVEY.md 8(d), CPU baseline (1): one process at a time on an otherwise idle host, median of 10 calls
        alone = sorted(pool.run([f])[1][0] for f in range(10))
    finally:
        pool.close()
    per_call.sort()
    return {"value": len(frames) / wall, "unit": "sweeps/s", "cores": pool.cores, "kind": pool.kind,
            "sample": f"{len(frames)} sweeps (8 per core) of the config-1 sweep set, BEV+FV via "
                      f"createTopAndFrontMaps, multiprocessing.Pool({pool.cores}); median {per_call[len(per_call) // 2] * 1e3:.1f} ms/call/core",
            "ms_per_call_median": per_call[len(per_call) // 2] * 1e3,
            "single_process_ms_per_call_median": alone[len(alone) // 2] * 1e3}


def run_reference(args):
    """The reference's own CPU implementation of the path on this box's host cores (rank 0 only)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from mv3d_b200.params import KITTI_FRONT_64x512, KITTI_PAPER_TOP
    from oracle.cpu_bench import CpuEncoderPool
    pool = CpuEncoderPool(KITTI_PAPER_TOP, KITTI_FRONT_64x512)
    sample = BATCH if pool.cores >= 4 else 16
    frames = list(range(sample))
    try:
        for _ in range(max(args.warmup, 1)):
            pool.run(frames)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.run(frames)
        elapsed = time.perf_counter() - t0
    finally:
        pool.close()
    value = sample * args.steps / elapsed
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "sweeps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": elapsed / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch_per_step": sample, "runs_on": "host CPU"},
        "cpu_baseline": {"value": value, "unit": "sweeps/s", "cores": pool.cores, "kind": pool.kind,
                         "sample": f"each step = {sample} sweeps of the workload's batch through "
                                   f"createTopAndFrontMaps on multiprocessing.Pool({pool.cores})"},
        "e2e": {"value": value, "unit": "sweeps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------
# short legs for the rows built around the path (not the headline metric; device-resident, same 64 sweeps)
# ------------------------------------------------------------------------------------------------------
def widened_rows(d_pts, offs, device, steps):
    import torch
    from mv3d_b200.anchors import AnchorFilter, make_anchors
    from mv3d_b200.encoder import ViewEncoder
    from mv3d_b200.params import PYTWIN_KITTI, REFERENCE_KITTI_TOP
    from mv3d_b200.pytwin import PyTwinEncoder

    def rate(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return BATCH * steps / (e0.elapsed_time(e1) * 1e-3)

    out = {"unit": "sweeps/s", "steps": steps, "batch": BATCH}
    py = PyTwinEncoder(device)
    top = torch.empty((BATCH, *PYTWIN_KITTI.top_shape), dtype=torch.float32, device=d_pts.device)
    front = torch.empty((BATCH, *PYTWIN_KITTI.front_shape), dtype=torch.float32, device=d_pts.device)
    out["python_twin_encode"] = {"value": rate(lambda: py.encode_device(d_pts, offs, PYTWIN_KITTI, top_out=top, front_out=front)),
                                 "what": "lidar_to_top + lidar_to_front_fast semantics, BEV 800x600x27 + FV 1500x100x3"}
    py.close()
    del front
    enc = ViewEncoder(device, semantics="top")
    filt = AnchorFilter(enc, anchors=make_anchors(REFERENCE_KITTI_TOP.shape[:2], stride=8))
    out["encode_and_filter_fused"] = {
        "value": rate(lambda: filt.encode_and_filter(d_pts, offs, REFERENCE_KITTI_TOP, top_out=top)),
        "what": "createTopMaps (BEV 800x600x27) + remove_empty_anchor over 30000 anchors, one call, nothing leaves the device"}
    out["filter_resident_views"] = {"value": rate(lambda: filt.filter_device(top)), "unit": "frames/s",
                                    "what": "remove_empty_anchor alone on 64 resident views"}
    enc.close()
    return out


# ------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------
def run_ours(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    from mv3d_b200.params import KITTI_FRONT_64x512, KITTI_PAPER_TOP
    top_p, front_p = KITTI_PAPER_TOP, KITTI_FRONT_64x512

    # CPU baseline first: it forks worker processes, which must happen before CUDA is initialised here.
    cpu = None
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        cpu = cpu_baseline(top_p, front_p)

    import numpy as np
    import torch
    import torch.distributed as dist
    from mv3d_b200.encoder import ViewEncoder, pinned_empty
    from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the encoder has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        from mv3d_b200.shard import init_process_group_quiet
        init_process_group_quiet("nccl", dev)  # keeps NCCL's version banner off stdout (one JSON line only)

    def barrier():
        if world > 1:
            dist.barrier()

    # frame i -> rank i mod world (BASELINE.json config 4); every rank holds BATCH sweeps
    sweeps = [hdl64_sweep(rank + world * i) for i in range(BATCH)]
    pts, offs = pack_sweeps(sweeps)
    n_points = int(offs[-1])
    d_pts = torch.from_numpy(pts).to(dev)
    top_out = torch.empty((BATCH, *top_p.shape), dtype=torch.float32, device=dev)
    front_out = torch.empty((BATCH, *front_p.shape), dtype=torch.float32, device=dev)
    enc = ViewEncoder(local_rank, semantics="topfront", strategy=args.strategy)
    strategy = enc.resolved_strategy(top_p)

    def step():
        enc.encode_device(d_pts, offs, top_p, front_p, top_out=top_out, front_out=front_out)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()

    # ---- timed region: inputs resident in HBM --------------------------------------------------------
    sampler = ClockSampler(local_rank)
    enc.kernel_times(reset=True)
    enc.set_profiling(True)
    launches0 = enc.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    if rank == 0:
        sampler.start()
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps - 1)]  # between steps: no synchronisation
    ev0.record()
    for k in range(args.steps):
        step()
        if k < args.steps - 1:
            marks[k].record()
    ev1.record()
    torch.cuda.synchronize()
    barrier()
    edges = [ev0] + marks + [ev1]
    step_times = sorted(edges[k].elapsed_time(edges[k + 1]) for k in range(args.steps))

    def pct(q):
        return step_times[min(len(step_times) - 1, max(0, int(round(q * (len(step_times) - 1)))))]
    clocks = sampler.stop() if rank == 0 else None
    elapsed_ms = ev0.elapsed_time(ev1)
    launches = enc.launch_count - launches0
    enc.set_profiling(False)
    ktimes = enc.kernel_times(reset=True)
    if world > 1:
        t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())
    value = world * BATCH * args.steps / (elapsed_ms * 1e-3)

    # ---- e2e: pinned host buffers through mv3d_encode_batch_host, H2D + encode + D2H every step --------
    e2e = None
    if not args.no_e2e:
        h_pts = pinned_empty(pts.shape)
        h_pts[...] = pts
        h_top = pinned_empty((BATCH, *top_p.shape))
        h_front = pinned_empty((BATCH, *front_p.shape))
        enc.encode_host(h_pts, offs, top_p, front_p, top_out=h_top, front_out=h_front)  # warm-up (allocates slots)
        barrier()
        torch.cuda.synchronize()
        enc.transfer_bytes(reset=True)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            enc.encode_host(h_pts, offs, top_p, front_p, top_out=h_top, front_out=h_front)
        e2e_s = time.perf_counter() - t0
        h2d_total, d2h_total = enc.transfer_bytes()
        checksum = float(h_top[0, :, :, top_p.zn + 1].sum()) + float(h_front[0].sum())  # touch the result
        if world > 1:
            t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_s = float(t.item())
        e2e = {"value": world * BATCH * args.steps / e2e_s, "unit": "sweeps/s",
               "h2d_bytes_per_step": h2d_total // args.steps, "d2h_bytes_per_step": d2h_total // args.steps,
               "host_result_bytes_per_step": int(h_top.nbytes + h_front.nbytes),
               "readback": os.environ.get("MV3D_HOST_READBACK", "compact"),
               "ms_per_step": e2e_s / args.steps * 1e3, "api": "mv3d_encode_batch_host (pinned host buffers)",
               "result_checksum": checksum}

    # ---- also (N=1 only, outside the timed region above): the rows built around the path, SURVEY.md 8(f) -------
    also = None
    if world == 1 and not args.no_also:
        also = widened_rows(d_pts, offs, local_rank, max(3, min(args.steps, 10)))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ---------------------------------------------------------------
    ab = algorithmic_bytes(top_p, front_p, n_points // BATCH)
    peak, peak_src = measured_hbm_peak()
    dominant = max(ktimes, key=lambda k: ktimes[k][0])
    dom_ms, dom_launches = ktimes[dominant]
    step_kernel_ms = sum(v[0] for v in ktimes.values())
    # bytes the dominant kernel is responsible for, per launch
    # (a step may launch it several times: one launch per pipelined sub-batch)
    per_step = {"tile_build_kernel": ab["bev"] * BATCH, "fill_tma_kernel": ab["bev"] * BATCH, "fill_zero_kernel": ab["bev"] * BATCH,
                "ce_zero_fill": ab["bev"] * BATCH}.get(dominant, ab["total"] * BATCH)
    per_launch = per_step * args.steps / max(dom_launches, 1)
    achieved = per_launch / (dom_ms / max(dom_launches, 1) * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    step_achieved = ab["total"] * BATCH * args.steps / (elapsed_ms * 1e-3) / 1e9
    traffic = ncu_traffic(dominant)
    roofline = {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": per_launch,
                "kernel_ms_per_launch": dom_ms / max(dom_launches, 1),
                "kernel_launches_per_step": dom_launches / args.steps,
                "kernel_share_of_step": dom_ms / step_kernel_ms if step_kernel_ms else None,
                "step_achieved": step_achieved, "step_frac": step_achieved / peak,
                "algorithmic_bytes_per_sweep": ab["total"],
                "kernels_ms_per_step": {k: v[0] / args.steps for k, v in ktimes.items() if v[1]}}

    line = {
        "metric": METRIC, "value": value, "unit": "sweeps/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch_per_gpu": BATCH, "points_per_sweep": n_points // BATCH,
                   "strategy": strategy, "sharding": f"frame i -> rank i mod {world}, no collective",
                   "l2": "each step writes 3.9 GB of output per GPU (>> 126 MB L2), so no step starts with a warm L2"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
        "step_ms": {"median": pct(0.5), "p5": pct(0.05), "p95": pct(0.95), "min": step_times[0], "max": step_times[-1],
                    "what": "rank 0, CUDA events between the steps of the timed region"},
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
    if also is not None:
        line["also"] = also
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
