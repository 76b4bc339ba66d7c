#!/usr/bin/env python
"""bench.py -- throughput of the n-variate regular-grid interpolation path.

Default workload (BASELINE.json: the configuration the metric's target is
quoted on): bivariate bilinear, float64 0.25 degree global Grid2D (1440 x 720,
circular longitude), 1e9 uniformly random query points per GPU, device
resident.  One "step" = one pass of the hot path over all points of the rank
= ONE kernel launch.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--points P]
                    [--workload bilinear|bicubic|trivariate|quadrivariate]
                    [--impl b200|reference]

Prints one JSON line (rank 0).  `value` is device-resident throughput over all
ranks; `e2e` goes through the public Python API / C-ABI with pinned HOST
buffers (H2D + kernel + D2H inside the timed region); `roofline` is the
algorithmic bytes of the dominant kernel over its CUDA-event duration against
the measured HBM copy bandwidth; `cpu_baseline` is the reference's CPU path
(oracle/_ref when built, else the C restatement) timed on this box's cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BYTES_PER_POINT = {"bilinear": 24, "bicubic": 24, "trivariate": 60, "quadrivariate": 100,
                   # K6: streams only; the windows are compute / L2 work
                   "c_spline2d": 24, "bicubic3d": 28}
# windowed workloads: fitting model of the public API
WINDOWED = {"bicubic": "bicubic", "c_spline2d": "c_spline", "bicubic3d": "bicubic"}
METRIC = "Mpoints/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--points", type=float, default=None, help="query points per GPU per step")
    ap.add_argument("--e2e-points", type=float, default=1e8)
    ap.add_argument("--workload", default="bilinear", choices=sorted(BYTES_PER_POINT))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU-baseline sample time")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def default_points(workload: str) -> int:
    return {"bilinear": 1_000_000_000, "bicubic": 100_000_000, "trivariate": 100_000_000,
            "quadrivariate": 100_000_000, "c_spline2d": 20_000_000, "bicubic3d": 20_000_000}[workload]


def ncu_traffic(workload, n):
    """DRAM bytes per launch from the committed `ncu --set full` capture of the
    workload's kernel (profiles/roofline_traffic.json: bytes per point measured
    under ncu), scaled to this launch's point count; None when no capture."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        with open(path) as f:
            e = json.load(f)[workload]
        return {"bytes_per_launch": e["dram_bytes_per_point"] * n, "dram_bytes_per_point": e["dram_bytes_per_point"],
                "source": e["source"], "secondary": e.get("secondary")}
    except Exception:
        return None


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------
# workloads
# ---------------------------------------------------------------------------
def build_host_case(workload: str):
    """Axis values + grid array (numpy) of the workload."""
    syn = __import__("importlib").import_module("pangeo-pyinterp_b200.synthetic")
    lon, lat = syn.lon_axis_values(), syn.lat_axis_values()
    if workload in ("bilinear", "bicubic", "c_spline2d"):
        return dict(nd=2, lon=lon, lat=lat, grid=syn.field(lon, lat), dtype="f64")
    if workload in ("trivariate", "bicubic3d"):
        t = syn.time_axis_values(365)
        return dict(nd=3, lon=lon, lat=lat, time=t, grid=syn.field(lon, lat, nz=365, dtype="float32", noise=0.0),
                    dtype="f32")
    t = syn.time_axis_values(12, 30)
    depth = syn.depth_axis_values(50)
    return dict(nd=4, lon=lon, lat=lat, time=t, depth=depth,
                grid=syn.field(lon, lat, nz=12, nu=50, dtype="float32", noise=0.0), dtype="f32")


def host_queries(case, n, seed):
    syn = __import__("importlib").import_module("pangeo-pyinterp_b200.synthetic")
    return syn.queries(n, seed, ndim=case["nd"], time_values=case.get("time"), nan_fraction=0.0)


def workload_name(workload, n):
    shape = {"bilinear": "bivariate bilinear f64 Grid2D 1440x720 circular lon",
             "bicubic": "bivariate bicubic (6x6 window) f64 Grid2D 1440x720 circular lon",
             "trivariate": "trivariate bilinear f32 Grid3D 1440x720x365 int64 TemporalAxis",
             "c_spline2d": "bivariate windowed c_spline (6x6 window) f64 Grid2D 1440x720 circular lon",
             "bicubic3d": "trivariate windowed bicubic (6x6 window, linear in time) f32 Grid3D 1440x720x365",
             "quadrivariate": "quadrivariate bilinear f32 Grid4D 1440x720x12x50 time on z, irregular depth on u"}
    return f"{shape[workload]}, {n:.3g} uniform random points per GPU"


# ---------------------------------------------------------------------------
# CPU baseline (the reference's own CPU path; test infrastructure, timed only)
# ---------------------------------------------------------------------------
def cpu_run(case, workload, q, threads=0):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    kind = "reference" if orc.available("reference") else "port"
    o = orc.Oracle(kind)
    axes = [o.axis(case["lon"], period=360.0), o.axis(case["lat"])]
    if case["nd"] >= 3:
        axes.append(o.axis(case["time"].astype("datetime64[ns]").astype(np.int64)))
    if case["nd"] >= 4:
        axes.append(o.axis(case["depth"]))

    def run(qq):
        t0 = time.perf_counter()
        if workload == "bicubic":
            out, _ = o.windowed2d(axes, case["grid"], qq[0], qq[1], num_threads=threads, stateless=False)
        elif workload in WINDOWED:
            out, _ = o.windowed_nd(axes, case["grid"], *qq, method=WINDOWED[workload], num_threads=threads,
                                   stateless=kind != "reference")  # the C port has the stateless form only
        else:
            out, _ = o.geometric(axes, case["grid"], *qq, num_threads=threads)
        return out, time.perf_counter() - t0

    return kind, run


def cpu_baseline(case, workload, seconds):
    cores = os.cpu_count() or 1
    kind, run = cpu_run(case, workload, None)
    n = 2_000_000
    q = host_queries(case, n, 77)
    _, dt = run(q)  # warm-up + calibration
    n = int(min(1e8, max(n, n * seconds / max(dt, 1e-3))))
    q = host_queries(case, n, 78)
    _, dt = run(q)
    return {"value": n / dt / 1e6, "unit": METRIC, "cores": cores, "kind": kind,
            "sample": f"{n} points of the same workload, one call, num_threads=0 ({cores} threads), {dt:.2f} s"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    case = build_host_case(args.workload)
    cores = os.cpu_count() or 1
    kind, run = cpu_run(case, args.workload, None)
    n = 2_000_000
    q = host_queries(case, n, 77)
    _, dt = run(q)
    # each step: a bounded sample sized for ~seconds/steps of CPU work
    per_step = max(1.0, args.cpu_seconds * 3 / max(1, args.steps + args.warmup))
    n = int(min(1e8, max(1e6, n * per_step / max(dt, 1e-3))))
    q = host_queries(case, n, 78)
    for _ in range(args.warmup):
        run(q)
    t = 0.0
    for _ in range(args.steps):
        t += run(q)[1]
    value = n * args.steps / t / 1e6
    sample = f"{n} points per step, num_threads=0 ({cores} threads)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": case["dtype"],
        "data": "synthetic", "config": {"workload": workload_name(args.workload, n)},
        "cpu_baseline": {"value": value, "unit": METRIC, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock and throttle reasons DURING the timed region: NVML in
    this process every 5 ms (nvidia-smi as a subprocess is too slow for a
    70 ms region); falls back to nvidia-smi when NVML is unavailable."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.index, self.samples, self.stop = index, [], threading.Event()
        self.max_mhz, self.nvml, self.handle = None, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if index < len(ids) and ids[index].isdigit():
                    phys = int(ids[index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None
        self.thread = threading.Thread(target=self._loop, daemon=True)

    def _sample_nvml(self):
        n = self.nvml
        mhz = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
        r = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)) if hasattr(
            n, "nvmlDeviceGetCurrentClocksEventReasons") else int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
        flags = [bool(r & 0x8), bool(r & 0x40), bool(r & 0x20), bool(r & 0x4)]
        self.samples.append((mhz, flags))

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                              "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
        f = [s.strip() for s in out.strip().split(",")]
        if len(f) >= 6:
            self.max_mhz = float(f[1])
            self.samples.append((float(f[0]), [v.lower().startswith("active") for v in f[2:6]]))

    def _loop(self):
        while not self.stop.is_set():
            try:
                if self.nvml is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self.stop.wait(0.005 if self.nvml is not None else 0.2)

    def __enter__(self):
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.thread.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        mhz = sorted(s[0] for s in self.samples)
        reasons = [n for k, n in enumerate(self.NAMES) if any(s[1][k] for s in self.samples)]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(mhz), "source": "nvml" if self.nvml is not None else "nvidia-smi"}


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
def main():
    args = parse()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    import pyinterp_b200 as pyi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL's banner / debug lines go to stderr: stdout carries the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    n = int(args.points or default_points(args.workload))
    case = build_host_case(args.workload)
    nd = case["nd"]

    # grid: generated on rank 0, replicated once over NVLink (one NCCL broadcast)
    t_arr = torch.from_numpy(case["grid"]).to(dev) if rank == 0 else \
        torch.empty(case["grid"].shape, dtype=torch.from_numpy(case["grid"][:1]).dtype, device=dev)
    bcast_ms = None
    if world > 1:
        # communicator set-up is a one-off of the process group, not of the grid
        warm = torch.zeros(1, device=dev)
        dist.all_reduce(warm)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pyi.dist.broadcast_grid_array(t_arr)
        torch.cuda.synchronize()
        bcast_ms = 1e3 * (time.perf_counter() - t0)
    axes = [pyi.Axis(case["lon"], period=360.0), pyi.Axis(case["lat"])]
    if nd >= 3:
        axes.append(pyi.TemporalAxis(case["time"]))
    if nd >= 4:
        axes.append(pyi.Axis(case["depth"]))
    grid = pyi.Grid(*axes, t_arr)

    # parity guard: a fixed 1e6-point sub-chunk against the CPU oracle before any number is recorded
    if rank == 0 and not args.no_cpu:
        q = host_queries(case, 1_000_000, 1001)
        _, run = cpu_run(case, args.workload, None)
        if args.workload in WINDOWED:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import orc
            o = orc.Oracle("reference" if orc.available("reference") else "port")
            oaxes = [o.axis(case["lon"], period=360.0), o.axis(case["lat"])]
            qq = list(q)
            if nd >= 3:
                oaxes.append(o.axis(case["time"].astype("datetime64[ns]").astype(np.int64)))
                qq[2] = qq[2].view("datetime64[ns]")
            if args.workload == "bicubic":
                want, _ = o.windowed2d(oaxes, case["grid"], q[0], q[1], num_threads=0, stateless=True)
            else:
                want, _ = o.windowed_nd(oaxes, case["grid"], *q, method=WINDOWED[args.workload], num_threads=0,
                                        stateless=True)
            got = {2: pyi.bivariate, 3: pyi.trivariate}[nd](grid, *qq, WINDOWED[args.workload])
        else:
            want, _ = run(q)
            fn = {2: pyi.bivariate, 3: pyi.trivariate, 4: pyi.quadrivariate}[nd]
            qq = list(q)
            if nd >= 3:
                qq[2] = qq[2].view("datetime64[ns]")
            got = fn(grid, *qq)
        tol = 1e-12 if got.dtype == np.float64 else 1e-5
        nan_ok = np.array_equal(np.isnan(got), np.isnan(want))
        m = ~np.isnan(want)
        err = np.abs(got[m].astype(np.float64) - want[m].astype(np.float64))
        if not nan_ok or np.any(err > tol * np.maximum(np.abs(want[m]), 1.0)):
            raise SystemExit("bench.py: parity check against the CPU oracle FAILED; no number recorded")

    # device-resident query points (synthetic, generated on the device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1005 + rank)
    x = torch.rand(n, dtype=torch.float64, device=dev, generator=gen).mul_(720.0).sub_(180.0)
    y = torch.rand(n, dtype=torch.float64, device=dev, generator=gen).mul_(180.0).sub_(90.0)
    coords = [x, y]
    if nd >= 3:
        tv = case["time"].astype("datetime64[ns]").astype(np.int64)
        day = 86_400_000_000_000
        lo, hi = int(tv.min()) - day, int(tv.max()) + day
        z = torch.rand(n, dtype=torch.float64, device=dev, generator=gen).mul_(float(hi - lo)).to(torch.int64).add_(lo)
        coords.append(z)
    if nd >= 4:
        coords.append(torch.rand(n, dtype=torch.float64, device=dev, generator=gen).mul_(5610.0).sub_(10.0))
    fn = {2: pyi.core.bivariate, 3: pyi.core.trivariate, 4: pyi.core.quadrivariate}[nd]
    if args.workload in WINDOWED:
        cfg = getattr({2: pyi.windowed.Bivariate, 3: pyi.windowed.Trivariate}[nd], WINDOWED[args.workload])()
    else:
        cfg = {2: pyi.geometric.Bivariate, 3: pyi.geometric.Trivariate, 4: pyi.geometric.Quadrivariate}[nd].bilinear()

    def step():
        return fn(grid, *coords, cfg)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        out = step()
    del out
    barrier()
    launches0 = pyi.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local) as clk:
        t_all0 = torch.cuda.Event(enable_timing=True)
        t_all1 = torch.cuda.Event(enable_timing=True)
        t_all0.record()
        for a, b in evs:
            a.record()
            out = step()
            b.record()
        t_all1.record()
        barrier()
    launches = pyi.launch_count() - launches0
    total_ms = t_all0.elapsed_time(t_all1)
    kernel_ms = sorted(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = world * n * args.steps / (total_ms / 1e3) / 1e6
    finite = float(torch.isfinite(out).double().mean().item())
    del out

    # end to end: pinned host buffers -> public API (H2D + kernel + D2H per step)
    ne = int(min(args.e2e_points, n))
    hq = []
    for c in coords:
        h = torch.empty(ne, dtype=c.dtype, pin_memory=True)
        h.copy_(c[:ne])
        hq.append(h.numpy())
    if nd >= 3:
        hq[2] = hq[2].view("datetime64[ns]")
    api = {2: pyi.bivariate, 3: pyi.trivariate, 4: pyi.quadrivariate}[nd]
    method = WINDOWED.get(args.workload, "bilinear")
    grid._resident = True
    for _ in range(2):
        r = api(grid, *hq, method)
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(2, min(args.steps, 5))
    for _ in range(e2e_steps):
        r = api(grid, *hq, method)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * ne * e2e_steps / float(t.item()) / 1e6
    h2d = sum(int(h.nbytes) for h in hq)
    d2h = int(r.nbytes)

    if rank == 0:
        peak, peak_src = peaks()
        bpp = BYTES_PER_POINT[args.workload]
        k_ms = sum(kernel_ms) / len(kernel_ms)
        achieved = n * bpp / (k_ms / 1e3) / 1e9
        tr = ncu_traffic(args.workload, n)
        line = {
            "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": case["dtype"], "data": "synthetic",
            "config": {"workload": workload_name(args.workload, n), "points_per_gpu": n,
                       "l2": "inputs (x, y, out streams) are far larger than L2; no flush needed",
                       "grid_broadcast_ms": bcast_ms, "finite_fraction": finite,
                       "e2e_points_per_step": ne, "kernel_ms_min": kernel_ms[0], "kernel_ms_median": kernel_ms[len(kernel_ms) // 2]},
            "e2e": {"value": e2e_value, "unit": METRIC, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": (tr or {}).get("bytes_per_launch"), "traffic_detail": tr,
                         "peak_source": peak_src,
                         "bytes_per_point": bpp, "kernel_ms": k_ms,
                         "secondary": (tr or {}).get("secondary")},
            "clocks": clk.summary(),
        }
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(case, args.workload, args.cpu_seconds)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
