"""Device throughput of the parity-surface kernels of the path (a1-a5 Axis::find_indexes /
find_index, a25 dateutils::convert) and of Grid1D univariate: device-resident inputs, CUDA events, median
of 7, achieved GB/s against MEASURED_PEAKS.json.  Usage: python tools/aux_bench.py [n]"""
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pyi = importlib.import_module("pangeo-pyinterp_b200")
s = pyi.synthetic


def timed(fn, reps=7):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 200_000_000
    peak = 6550.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    rng = np.random.default_rng(3)
    lon, lat = s.lon_axis_values(), s.lat_axis_values()
    time_values = s.time_axis_values(365)
    t64 = time_values.astype("datetime64[ns]").astype(np.int64)
    cases = [
        ("lon regular periodic f64", pyi.Axis(lon, period=360.0), torch.from_numpy(rng.uniform(-180.0, 540.0, n))),
        ("lat regular f64", pyi.Axis(lat), torch.from_numpy(rng.uniform(lat[0], lat[-1], n))),
        ("lat irregular f64", pyi.Axis(s.irregular_lat_values()), torch.from_numpy(rng.uniform(-89.0, 89.0, n))),
        ("time regular int64 (daily ns)", pyi.TemporalAxis(time_values),
         torch.from_numpy(rng.integers(t64[0], t64[-1], n))),
    ]
    print(f"n = {n}, peak = {peak:.1f} GB/s")
    for name, axis, c in cases:
        cd = c.cuda()
        ms = timed(lambda: axis.find_indexes(cd))
        print(f"find_indexes {name:32s} {n / ms / 1e6:8.2f} Gpts/s  {n * 24 / ms / 1e6:8.1f} GB/s "
              f"({n * 24 / ms / 1e6 / peak:.3f} of peak; 8 B in + 16 B out)")
        ms = timed(lambda: axis.find_index(cd, bounded=True))
        print(f"find_index   {name:32s} {n / ms / 1e6:8.2f} Gpts/s  {n * 16 / ms / 1e6:8.1f} GB/s "
              f"({n * 16 / ms / 1e6 / peak:.3f} of peak; 8 B in + 8 B out)")
        del cd
    v = torch.from_numpy(rng.integers(0, 2_000_000_000, n)).cuda()
    for src, dst in (("s", "ns"), ("ns", "D"), ("h", "us")):
        ms = timed(lambda: pyi.core.convert_resolution(v, src, dst))
        print(f"datetime convert {src:>2s} -> {dst:2s}              {n / ms / 1e6:8.2f} Gpts/s  {n * 16 / ms / 1e6:8.1f} GB/s "
              f"({n * 16 / ms / 1e6 / peak:.3f} of peak; 8 B in + 8 B out)")
    del v
    # Grid1D univariate (K6, 6-node window) on a regular axis
    xa = np.arange(4096) * 0.25
    g1 = pyi.Grid1D(pyi.Axis(xa), np.sin(xa))
    m = min(n, 50_000_000)
    xq = torch.from_numpy(rng.uniform(1.0, xa[-1] - 1.0, m)).cuda()
    for model in ("c_spline", "akima", "linear"):
        cfg = getattr(pyi.windowed.Univariate, model)()
        ms = timed(lambda: pyi.core.univariate(g1, xq, cfg), reps=3)
        print(f"univariate {model:10s} (Grid1D, 6 nodes)        {m / ms / 1e6:8.2f} Gpts/s")


if __name__ == "__main__":
    main()
