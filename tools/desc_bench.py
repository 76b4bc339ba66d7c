"""Throughput of geometric bilinear on a DESCENDING regular latitude (90 -> -90,
the ERA5 convention) against the ascending one: device-resident queries, CUDA
events, 0.25 degree grid.  Usage: python tools/desc_bench.py [n_points]"""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pyi = importlib.import_module("pangeo-pyinterp_b200")
s = pyi.synthetic


def timed(fn, reps=5):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
    lon, lat = s.lon_axis_values(), s.lat_axis_values()
    g = s.field(lon, lat)
    x, y = s.queries(n, 1001)
    xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    rows = []
    for name, la, gg in (("ascending", lat, g), ("descending", lat[::-1].copy(), g[:, ::-1].copy())):
        grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(la), gg, resident=True)
        for mode in ("tolerance", "exact"):
            os.environ["PYINTERP_B200_BILINEAR"] = mode
            ms = timed(lambda: pyi.bivariate(grid, xd, yd))
            rows.append((f"bivariate f64 {name} {mode}", n / ms / 1e6))
        os.environ.pop("PYINTERP_B200_BILINEAR", None)
    # the two grids describe the same field: same values to rounding
    ga = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g, resident=True)
    gd = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat[::-1].copy()), g[:, ::-1].copy(), resident=True)
    a = pyi.bivariate(ga, xd[:4_000_000], yd[:4_000_000]).cpu().numpy()
    d = pyi.bivariate(gd, xd[:4_000_000], yd[:4_000_000]).cpu().numpy()
    m = ~np.isnan(a) & ~np.isnan(d)
    rows.append(("max |ascending - descending| (same field)", float(np.abs(a[m] - d[m]).max())))
    # 3-D float32, int64 time, descending latitude
    m3 = min(n, 50_000_000)
    t = s.time_axis_values(48)
    g3 = s.field(lon, lat, nz=48, dtype="float32", noise=0.0)
    q = s.queries(m3, 1003, ndim=3, time_values=t)
    qd = [torch.from_numpy(v).cuda() for v in q]
    for name, la, gg in (("ascending", lat, g3), ("descending", lat[::-1].copy(), g3[:, ::-1].copy())):
        grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(la), pyi.TemporalAxis(t), gg, resident=True)
        ms = timed(lambda: pyi.core.trivariate(grid, qd[0], qd[1], qd[2], pyi.geometric.Trivariate.bilinear()))
        rows.append((f"trivariate f32 {name}", m3 / ms / 1e6))
    for k, v in rows:
        print(f"{k:50s} {v:12.4g}" + (" Gpts/s" if "max" not in k else ""))


if __name__ == "__main__":
    main()
