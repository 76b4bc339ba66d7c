#!/usr/bin/env python
"""bench.py -- throughput of the n-variate regular-grid interpolation path.

Default workload (BASELINE.json: the configuration the metric's target is
quoted on): bivariate bilinear, float64 0.25 degree global Grid2D (1440 x 720,
circular longitude), 1e9 uniformly random query points per GPU, device
resident.  One "step" = one pass of the hot path over all points of the rank
= ONE kernel launch.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--points P]
                    [--workload bilinear|bilinear_irrlat|bicubic|trivariate|quadrivariate]
                    [--no-secondary]
                    [--impl b200|reference]

Prints one JSON line (rank 0).  `value` is device-resident throughput over all
ranks; `e2e` goes through the public Python API / C-ABI with pinned HOST
buffers (H2D + kernel + D2H inside the timed region); `roofline` is the
algorithmic bytes of the dominant kernel over its CUDA-event duration against
the measured HBM copy bandwidth; `cpu_baseline` is the reference's CPU path
(oracle/_ref when built, else the C restatement) timed on this box's cores.
With the default workload the line also carries `secondary`: short sub-runs
(3 warm-up + 5 timed passes each, parity-checked against the oracle on every
rank) of the other BASELINE configurations -- bicubic, trivariate,
quadrivariate and bilinear on an irregular latitude axis with its point-count
sweep -- each with its own roofline entry.
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

BYTES_PER_POINT = {"bilinear": 24, "bilinear_irrlat": 24, "bicubic": 24, "trivariate": 60, "quadrivariate": 100,
                   # the other half of the geometric matrix: idw on the 2-D grid, bilinear on a float64 3-D grid
                   # (x, y, z 24 + out 8 + 8 corners x 8 = 96)
                   "idw": 24, "trivariate_f64": 96,
                   # the same 2-D grid stored as float32 (float32 results: 16 in + 4 out)
                   "bilinear_f32": 20,
                   # K6: streams only; the windows are compute / L2 work
                   "c_spline2d": 24, "akima2d": 24, "steffen2d": 24, "polynomial2d": 24, "bicubic3d": 28, "c_spline3d": 28}
# windowed workloads: fitting model of the public API
WINDOWED = {"bicubic": "bicubic", "c_spline2d": "c_spline", "akima2d": "akima", "steffen2d": "steffen", "polynomial2d": "polynomial",
            "bicubic3d": "bicubic", "c_spline3d": "c_spline"}
# geometric workloads: spatial method of the public API (default bilinear)
GEOMETRIC = {"idw": "idw"}
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
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the short sub-runs of the other BASELINE configurations")
    return ap.parse_args()


def default_points(workload: str) -> int:
    return {"bilinear": 1_000_000_000, "bilinear_irrlat": 1_000_000_000, "bilinear_f32": 1_000_000_000,
            "idw": 100_000_000,
            "trivariate_f64": 100_000_000, "bicubic": 100_000_000, "trivariate": 100_000_000,
            "quadrivariate": 100_000_000, "c_spline2d": 20_000_000, "akima2d": 20_000_000, "steffen2d": 20_000_000, "polynomial2d": 20_000_000,
            "bicubic3d": 20_000_000, "c_spline3d": 20_000_000}[workload]


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
    if workload == "bilinear_irrlat":  # BASELINE configs[4]: irregular latitude axis
        lat = syn.irregular_lat_values()
    if workload == "bilinear_f32":
        return dict(nd=2, lon=lon, lat=lat, grid=syn.field(lon, lat).astype(np.float32), dtype="f32")
    if workload in ("bilinear", "bilinear_irrlat", "bicubic", "c_spline2d", "akima2d", "steffen2d", "polynomial2d", "idw"):
        return dict(nd=2, lon=lon, lat=lat, grid=syn.field(lon, lat), dtype="f64")
    if workload == "trivariate_f64":
        t = syn.time_axis_values(120)
        return dict(nd=3, lon=lon, lat=lat, time=t, grid=syn.field(lon, lat, nz=120, dtype="float64", noise=0.0),
                    dtype="f64")
    if workload in ("trivariate", "bicubic3d", "c_spline3d"):
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
             "bilinear_irrlat": "bivariate bilinear f64 Grid2D 1440x720 circular lon, irregular (Gaussian-like) latitude",
             "idw": "bivariate inverse distance weighting (exp 2) f64 Grid2D 1440x720 circular lon",
             "bilinear_f32": "bivariate bilinear f32 Grid2D 1440x720 circular lon (float32 results)",
             "trivariate_f64": "trivariate bilinear f64 Grid3D 1440x720x120 int64 TemporalAxis",
             "bicubic": "bivariate bicubic (6x6 window) f64 Grid2D 1440x720 circular lon",
             "trivariate": "trivariate bilinear f32 Grid3D 1440x720x365 int64 TemporalAxis",
             "c_spline2d": "bivariate windowed c_spline (6x6 window) f64 Grid2D 1440x720 circular lon",
             "bicubic3d": "trivariate windowed bicubic (6x6 window, linear in time) f32 Grid3D 1440x720x365",
             "akima2d": "bivariate windowed akima (6x6 window) f64 Grid2D 1440x720 circular lon",
             "steffen2d": "bivariate windowed steffen (6x6 window) f64 Grid2D 1440x720 circular lon",
             "polynomial2d": "bivariate windowed polynomial (6x6 window) f64 Grid2D 1440x720 circular lon",
             "c_spline3d": "trivariate windowed c_spline (6x6 window, linear in time) f32 Grid3D 1440x720x365",
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
            out, _ = o.geometric(axes, case["grid"], *qq, method={"idw": 1}.get(workload, 0), num_threads=threads)
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
class Env:
    """Process-wide state of the GPU arm: rank, device, torch.distributed."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            # NCCL's banner / debug lines go to stderr: stdout carries the one JSON line
            # (the banner is written to file descriptor 1 by the library itself)
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
            sys.stdout.flush()
            saved = os.dup(1)
            os.dup2(2, 1)
            try:
                dist.init_process_group("nccl", device_id=self.dev)
                warm = torch.zeros(1, device=self.dev)  # communicator set-up is a one-off of the process group
                dist.all_reduce(warm)
                torch.cuda.synchronize()
            finally:
                os.dup2(saved, 1)
                os.close(saved)
            # a CPU-side group: ranks that only wait must not park an NCCL kernel on their GPU
            self.cpu_group = dist.new_group(backend="gloo")

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def cpu_barrier(self):
        """Barrier on the host only (gloo): a rank waiting here leaves its GPU idle."""
        if self.world > 1:
            self.dist.barrier(group=self.cpu_group)

    def max_over_ranks(self, v: float) -> float:
        if self.world == 1:
            return float(v)
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def min_over_ranks(self, v: float) -> float:
        return -self.max_over_ranks(-v)


def make_grid(env, pyi, case):
    """Grid generated on rank 0 and replicated once over NVLink (one NCCL
    broadcast per grid handle); returns (grid, broadcast ms)."""
    torch = env.torch
    t_arr = torch.from_numpy(case["grid"]).to(env.dev) if env.rank == 0 else \
        torch.empty(case["grid"].shape, dtype=torch.from_numpy(case["grid"][:1]).dtype, device=env.dev)
    bcast_ms = None
    if env.world > 1:
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pyi.dist.broadcast_grid_array(t_arr)
        torch.cuda.synchronize()
        bcast_ms = 1e3 * (time.perf_counter() - t0)
    axes = [pyi.Axis(case["lon"], period=360.0), pyi.Axis(case["lat"])]
    if case["nd"] >= 3:
        axes.append(pyi.TemporalAxis(case["time"]))
    if case["nd"] >= 4:
        axes.append(pyi.Axis(case["depth"]))
    return pyi.Grid(*axes, t_arr), bcast_ms


def oracle_guard(env, pyi, case, grid, workload):
    """Parity guard, on EVERY rank: a fixed 1e6-point sub-chunk of the workload
    (seed 1001 + rank) against the CPU oracle before any number is recorded."""
    nd = case["nd"]
    # 2-D bilinear: from 2^22 points per launch the default value mode runs the tolerance kernel -- the
    # kernel the timed region runs -- so the guard is that large and goes through the device entry
    # (one launch; the host path would cut it into chunks below the threshold)
    big = workload in ("bilinear", "bilinear_irrlat", "bilinear_f32")
    n_guard = 4_200_000 if big else 1_000_000
    q = host_queries(case, n_guard, 1001 + env.rank)
    kind, run = cpu_run(case, workload, None)
    if workload in WINDOWED:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import orc
        o = orc.Oracle(kind)
        oaxes = [o.axis(case["lon"], period=360.0), o.axis(case["lat"])]
        if nd >= 3:
            oaxes.append(o.axis(case["time"].astype("datetime64[ns]").astype(np.int64)))
        if workload == "bicubic":
            want, _ = o.windowed2d(oaxes, case["grid"], q[0], q[1], num_threads=0, stateless=True)
        else:
            want, _ = o.windowed_nd(oaxes, case["grid"], *q, method=WINDOWED[workload], num_threads=0,
                                    stateless=True)
    else:
        want, _ = run(q)
    qq = list(q)
    if nd >= 3:
        qq[2] = qq[2].view("datetime64[ns]")
    api = {2: pyi.bivariate, 3: pyi.trivariate, 4: pyi.quadrivariate}[nd]
    if big:
        dq = [env.torch.from_numpy(np.ascontiguousarray(a)).to(env.dev) for a in q]
        got = api(grid, *dq, "bilinear").cpu().numpy()
    else:
        got = api(grid, *qq, WINDOWED.get(workload, GEOMETRIC.get(workload, "bilinear")))
    tol = 1e-12 if got.dtype == np.float64 else 1e-5
    if workload in GEOMETRIC:
        # idw goes through sin / cos / pow, correctly rounded on neither side: the slack of tests/test_gpu_parity.py
        tol *= 20
    nan_ok = np.array_equal(np.isnan(got), np.isnan(want))
    m = ~np.isnan(want)
    err = np.abs(got[m].astype(np.float64) - want[m].astype(np.float64))
    ok = bool(nan_ok and not np.any(err > tol * np.maximum(np.abs(want[m]), 1.0)))
    if env.min_over_ranks(1.0 if ok else 0.0) < 1.0:
        raise SystemExit(f"bench.py: parity check of '{workload}' against the CPU oracle FAILED on a rank; "
                         "no number recorded")
    return {"checked": "ok", "points": n_guard, "ranks": env.world, "oracle": kind,
            "bit_identical_fraction": float((err == 0).mean()) if err.size else 1.0,
            "max_abs_err": float(err.max()) if err.size else 0.0}


def device_queries(env, case, n):
    """Device-resident query points (synthetic, generated on the device)."""
    torch = env.torch
    gen = torch.Generator(device=env.dev)
    gen.manual_seed(1005 + env.rank)
    x = torch.rand(n, dtype=torch.float64, device=env.dev, generator=gen).mul_(720.0).sub_(180.0)
    y = torch.rand(n, dtype=torch.float64, device=env.dev, generator=gen).mul_(180.0).sub_(90.0)
    coords = [x, y]
    if case["nd"] >= 3:
        tv = case["time"].astype("datetime64[ns]").astype(np.int64)
        day = 86_400_000_000_000
        lo, hi = int(tv.min()) - day, int(tv.max()) + day
        z = torch.rand(n, dtype=torch.float64, device=env.dev, generator=gen).mul_(float(hi - lo)).to(torch.int64).add_(lo)
        coords.append(z)
    if case["nd"] >= 4:
        coords.append(torch.rand(n, dtype=torch.float64, device=env.dev, generator=gen).mul_(5610.0).sub_(10.0))
    return coords


def make_config(pyi, case, workload):
    nd = case["nd"]
    if workload in WINDOWED:
        return getattr({2: pyi.windowed.Bivariate, 3: pyi.windowed.Trivariate}[nd], WINDOWED[workload])()
    cls = {2: pyi.geometric.Bivariate, 3: pyi.geometric.Trivariate, 4: pyi.geometric.Quadrivariate}[nd]
    return getattr(cls, GEOMETRIC.get(workload, "bilinear"))()


def time_device(env, pyi, grid, coords, cfg, steps, warmup, sampler=None):
    """`steps` timed passes over device-resident points: CUDA events on the
    launching stream around every pass and around the whole region, barrier +
    synchronize on both sides, max over ranks."""
    torch = env.torch
    fn = {2: pyi.core.bivariate, 3: pyi.core.trivariate, 4: pyi.core.quadrivariate}[grid.ndim]
    out = None
    for _ in range(max(3, warmup)):
        out = fn(grid, *coords, cfg)
    del out
    env.barrier()
    launches0 = pyi.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    t_all0, t_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def region():
        t_all0.record()
        o = None
        for a, b in evs:
            a.record()
            o = fn(grid, *coords, cfg)
            b.record()
        t_all1.record()
        env.barrier()
        return o

    if sampler is not None:
        with sampler:
            out = region()
    else:
        out = region()
    launches = pyi.launch_count() - launches0
    total_ms = env.max_over_ranks(t_all0.elapsed_time(t_all1))
    kernel_ms = sorted(a.elapsed_time(b) for a, b in evs)
    finite = float(torch.isfinite(out).double().mean().item())
    del out
    return total_ms, kernel_ms, int(launches), finite


def live_copy_probe(torch, n_gpus, in_bytes_per_point, out_bytes_per_point, points):
    """Copy-only ceiling of the host-fed path ON THIS BOX, measured in this run: the bytes one step of the
    end-to-end call moves (inputs up, results down), `points` split over the first n_gpus devices, 8M-point
    chunks on three streams per device, page-locked buffers, no kernel.  Returns GB/s (best of 3)."""
    per = -(-points // n_gpus)
    chunk = 1 << 23
    state = []
    for d in range(n_gpus):
        with torch.cuda.device(d):
            hin = torch.empty(per * in_bytes_per_point, dtype=torch.uint8, pin_memory=True)
            hout = torch.empty(per * out_bytes_per_point, dtype=torch.uint8, pin_memory=True)
            streams = [torch.cuda.Stream(device=d) for _ in range(3)]
            dev = [torch.empty(chunk * (in_bytes_per_point + out_bytes_per_point), dtype=torch.uint8,
                               device=f"cuda:{d}") for _ in range(3)]
            state.append((hin, hout, streams, dev))
    best = float("inf")
    for _ in range(4):
        for d in range(n_gpus):
            torch.cuda.synchronize(d)
        t0 = time.perf_counter()
        for d, (hin, hout, streams, dev) in enumerate(state):
            k = 0
            for b in range(0, per, chunk):
                c = min(chunk, per - b)
                s = streams[k % 3]
                with torch.cuda.stream(s):
                    dev[k % 3][: c * in_bytes_per_point].copy_(hin[b * in_bytes_per_point:(b + c) * in_bytes_per_point],
                                                               non_blocking=True)
                    o0 = chunk * in_bytes_per_point
                    hout[b * out_bytes_per_point:(b + c) * out_bytes_per_point].copy_(
                        dev[k % 3][o0:o0 + c * out_bytes_per_point], non_blocking=True)
                k += 1
        for d in range(n_gpus):
            torch.cuda.synchronize(d)
        best = min(best, time.perf_counter() - t0)
    del state
    return per * n_gpus * (in_bytes_per_point + out_bytes_per_point) / best / 1e9


def e2e_bound(n_gpus, gbs):
    """Names the limiter of the host-fed path from the committed copy-only probe
    (tools/pcie_probe.cu -> profiles/r2_pcie_probe.json: the best aggregate
    GB/s of 16 B in + 8 B out per point with no kernel at this GPU count)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_pcie_probe.json")) as f:
            probe = json.load(f)
        e = probe["gpus"][str(n_gpus)]
        return {"bound": e["bound"], "probe_ceiling_gbs": e["gbs"], "frac_of_probe": gbs / e["gbs"],
                "probe_source": probe["source"]}
    except Exception:
        return {"bound": "pcie", "probe_ceiling_gbs": None, "frac_of_probe": None, "probe_source": None}


def roofline_entry(workload, n, kernel_ms):
    peak, peak_src = peaks()
    bpp = BYTES_PER_POINT[workload]
    k_ms = sum(kernel_ms) / len(kernel_ms)
    achieved = n * bpp / (k_ms / 1e3) / 1e9
    tr = ncu_traffic(workload, n)
    return {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": (tr or {}).get("bytes_per_launch"), "traffic_detail": tr, "peak_source": peak_src,
            "bytes_per_point": bpp, "kernel_ms": k_ms, "secondary": (tr or {}).get("secondary")}


def run_secondary(env, pyi, workload, n, steps, sweep=()):
    """One short sub-run (3 warm-up + `steps` timed passes) of another BASELINE
    configuration, parity-checked on every rank first."""
    torch = env.torch
    case = build_host_case(workload)
    grid, bcast_ms = make_grid(env, pyi, case)
    parity = oracle_guard(env, pyi, case, grid, workload)
    coords = device_queries(env, case, n)
    cfg = make_config(pyi, case, workload)
    total_ms, kernel_ms, launches, finite = time_device(env, pyi, grid, coords, cfg, steps, 3)
    entry = {"workload": workload_name(workload, n), "value": env.world * n * steps / (total_ms / 1e3) / 1e6,
             "unit": METRIC, "ms_per_step": total_ms / steps, "steps": steps, "warmup": 3,
             "gpu_launches": launches, "dtype": case["dtype"], "finite_fraction": finite,
             "grid_broadcast_ms": bcast_ms, "roofline": roofline_entry(workload, n, kernel_ms), "parity": parity}
    if sweep:
        # point-count sweep (BASELINE configs[4]): same grid, the first m points
        pts = []
        for m in sweep:
            if m > n:
                continue
            sub = [c[:m] for c in coords]
            t_ms, k_ms, _, _ = time_device(env, pyi, grid, sub, cfg, steps, 3)
            pts.append({"points_per_gpu": int(m), "value": env.world * m * steps / (t_ms / 1e3) / 1e6,
                        "ms_per_step": t_ms / steps, "kernel_ms_min": k_ms[0]})
        entry["sweep"] = pts
    del coords, grid
    torch.cuda.empty_cache()
    return entry


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference_arm(args)
    import pyinterp_b200 as pyi
    env = Env()
    torch = env.torch
    world, rank = env.world, env.rank
    n = int(args.points or default_points(args.workload))
    case = build_host_case(args.workload)
    nd = case["nd"]
    grid, bcast_ms = make_grid(env, pyi, case)
    parity = None if args.no_cpu else oracle_guard(env, pyi, case, grid, args.workload)
    coords = device_queries(env, case, n)
    cfg = make_config(pyi, case, args.workload)
    clk = ClockSampler(env.local)
    total_ms, kernel_ms, launches, finite = time_device(env, pyi, grid, coords, cfg, args.steps, args.warmup, clk)
    ms_per_step = total_ms / args.steps
    value = world * n * args.steps / (total_ms / 1e3) / 1e6

    # end to end: HOST buffers -> public API (H2D + kernel + D2H per step).
    # N > 1: ONE process (rank 0) hands ONE call of N x ne points to the library,
    # which shards it over the N devices (pyi_dist_init) -- the shape of the
    # reference's drop-in caller, one process whose parallel_for splits a call's
    # points; the other ranks idle at the barrier.  `ranks` is the other way to
    # feed N GPUs: every torchrun rank calling with its own ne points.
    api = {2: pyi.bivariate, 3: pyi.trivariate, 4: pyi.quadrivariate}[nd]
    method = WINDOWED.get(args.workload, GEOMETRIC.get(args.workload, "bilinear"))
    grid._resident = True
    ne = int(min(args.e2e_points if world == 1 else args.e2e_points / 2, n))
    e2e_steps = max(3, min(args.steps, 10))

    def pinned_copy(count):
        hq = []
        for c in coords:
            h = torch.empty(count, dtype=c.dtype, pin_memory=True)
            h.copy_(c[:count])
            hq.append(h.numpy())
        if nd >= 3:
            hq[2] = hq[2].view("datetime64[ns]")
        return hq

    def time_calls(host_arrays, steps, active=True):
        """Wall clock of `steps` API calls between two barriers, max over ranks."""
        r = None
        if active:
            for _ in range(2):
                r = api(grid, *host_arrays, method)
        env.cpu_barrier()
        env.barrier()
        t0 = time.perf_counter()
        if active:
            for _ in range(steps):
                r = api(grid, *host_arrays, method)
            torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        # idle ranks wait on the host: an NCCL collective entered early would
        # spin on THEIR device, which the active rank's call is also using
        env.cpu_barrier()
        return env.max_over_ranks(dt), r

    e2e_ranks = None
    if world > 1:
        hq = pinned_copy(ne)
        dt, r = time_calls(hq, e2e_steps)
        e2e_ranks = {"value": world * ne * e2e_steps / dt / 1e6, "unit": METRIC,
                     "how": f"{world} torchrun ranks, each calling with its own {ne} points"}
        del hq, r
    total = ne * world
    lead = rank == 0
    if lead:
        pyi.dist.init(world)
        hq = pinned_copy(min(total, n))
        total = hq[0].size
    else:
        hq = None
    dt, r = time_calls(hq, e2e_steps, active=lead)
    e2e_value = total * e2e_steps / dt / 1e6
    h2d = sum(int(h.nbytes) for h in hq) if lead else 0
    d2h = int(r.nbytes) if lead else 0
    e2e_gbs = (h2d + d2h) * e2e_steps / dt / 1e9
    # the same call with plain (pageable) numpy inputs, as an unmodified caller passes them
    pageable = [np.array(h, copy=True) for h in hq] if lead else None
    dt, _ = time_calls(pageable, e2e_steps, active=lead)
    e2e_pageable = total * e2e_steps / dt / 1e6
    live_probe = None
    if lead:
        pyi.dist.init(1)
        try:
            live_probe = live_copy_probe(torch, world, h2d // total, d2h // total, total)
        except Exception as e:  # the probe is evidence, not the measurement
            live_probe = f"{type(e).__name__}: {e}"
    env.cpu_barrier()
    del pageable, hq, r, coords
    torch.cuda.empty_cache()

    secondary = {}
    if args.workload == "bilinear" and not args.no_secondary:
        del grid
        torch.cuda.empty_cache()
        sub = [("bicubic", 100_000_000, ()), ("c_spline2d", 100_000_000, ()), ("akima2d", 50_000_000, ()),
               ("trivariate", 100_000_000, ()),
               ("quadrivariate", 100_000_000, ()),
               ("idw", 100_000_000, ()), ("trivariate_f64", 100_000_000, ()), ("bilinear_f32", 400_000_000, ()),
               ("bilinear_irrlat", 1_000_000_000, (1_000_000, 10_000_000, 100_000_000, 1_000_000_000))]
        if world >= 8 and not args.points:
            # configs[4] sweeps to 1e10 points in total: 1.25e9 per GPU on eight of them (30 GB of streams each)
            top = -(-10_000_000_000 // world)
            sub[-1] = ("bilinear_irrlat", top, (1_000_000, 10_000_000, 100_000_000, 1_000_000_000, top))
        for name, m, sweep in sub:
            m = int(min(m, args.points)) if args.points else m
            try:
                secondary[name] = run_secondary(env, pyi, name, m, 5, sweep)
            except SystemExit:
                raise
            except Exception as e:  # a failed sub-run must not hide the headline
                secondary[name] = {"error": f"{type(e).__name__}: {e}"}
                torch.cuda.empty_cache()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": case["dtype"], "data": "synthetic",
            "config": {"workload": workload_name(args.workload, n), "points_per_gpu": n,
                       "l2": "inputs (x, y, out streams) are far larger than L2; no flush needed",
                       "grid_broadcast_ms": bcast_ms, "finite_fraction": finite, "parity": parity,
                       "e2e_points_per_step": total, "kernel_ms_min": kernel_ms[0],
                       "kernel_ms_median": kernel_ms[len(kernel_ms) // 2]},
            "e2e": {"value": e2e_value, "unit": METRIC, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "points_per_step": total, "input": "page-locked host arrays",
                    "how": f"one process, one API call per step, sharded over {world} device(s) inside the C-ABI "
                           "(pyi_dist_init)",
                    "host_link_gbs": e2e_gbs, **e2e_bound(world, e2e_gbs),
                    "live_probe_gbs": live_probe,
                    "frac_of_live_probe": (e2e_gbs / live_probe) if isinstance(live_probe, float) else None,
                    "pageable": {"value": e2e_pageable, "unit": METRIC,
                                 "input": "plain numpy arrays, staged through the library's pinned slots"},
                    "ranks": e2e_ranks},
            "gpu_launches": int(launches),
            "roofline": roofline_entry(args.workload, n, kernel_ms),
            "clocks": clk.summary(),
        }
        if secondary:
            line["secondary"] = secondary
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(case, args.workload, args.cpu_seconds)
        print(json.dumps(line), flush=True)
    if world > 1:
        env.dist.destroy_process_group()


if __name__ == "__main__":
    main()
