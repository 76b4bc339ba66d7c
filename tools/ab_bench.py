"""A/B timing of one library build: bilinear (tolerance, exact; float64 and float32 grids,
regular and irregular latitude), bicubic, trivariate / quadrivariate.  Device-resident queries,
CUDA events, median of 7.  PYINTERP_B200_LIB selects the build.  Usage: python tools/ab_bench.py [n]"""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if os.environ.get("PYINTERP_B200_LIB"):  # an older build may lack the newest entry points
    import ctypes
    _abi = importlib.import_module("pangeo-pyinterp_b200._abi")
    _probe = ctypes.CDLL(os.environ["PYINTERP_B200_LIB"])
    for _name in [k for k in _abi.SIGNATURES if not hasattr(_probe, k)]:
        del _abi.SIGNATURES[_name]
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
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
    print("library:", os.environ.get("PYINTERP_B200_LIB", "(in tree)"))
    lon, lat = s.lon_axis_values(), s.lat_axis_values()
    g = s.field(lon, lat)
    x, y = s.queries(n, 1001)
    xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    for name, la, gg in (("f64 regular", lat, g), ("f32 regular", lat, g.astype(np.float32)),
                         ("i16 regular", lat, np.round(g * 1000).astype(np.int16)),
                         ("f32 descending", lat[::-1].copy(), g[:, ::-1].astype(np.float32)),
                         ("f64 irregular", s.irregular_lat_values(), g)):
        grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(la), gg, resident=True)
        for mode in ("tolerance", "exact"):
            os.environ["PYINTERP_B200_BILINEAR"] = mode
            ms = timed(lambda: pyi.bivariate(grid, xd, yd))
            print(f"bilinear {name:14s} {mode:10s} {n / ms / 1e6:8.2f} Gpts/s")
        os.environ.pop("PYINTERP_B200_BILINEAR", None)
        if "regular" in name and "irregular" not in name:
            m = n // 2
            for mode in (None, "exact"):
                if mode:
                    os.environ["PYINTERP_B200_BICUBIC"] = mode
                ms = timed(lambda: pyi.bivariate(grid, xd[:m], yd[:m], "bicubic"))
                print(f"bicubic  {name:14s} {mode or 'default':10s} {m / ms / 1e6:8.2f} Gpts/s")
                os.environ.pop("PYINTERP_B200_BICUBIC", None)
    m3 = n // 2
    t = s.time_axis_values(48)
    g3 = s.field(lon, lat, nz=48, dtype="float32", noise=0.0)
    q = s.queries(m3, 1003, ndim=3, time_values=t)
    qd = [torch.from_numpy(v).cuda() for v in q]
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), pyi.TemporalAxis(t), g3, resident=True)
    ms = timed(lambda: pyi.core.trivariate(grid, qd[0], qd[1], qd[2], pyi.geometric.Trivariate.bilinear()))
    print(f"trivariate f32 (48 planes)           {m3 / ms / 1e6:8.2f} Gpts/s")


if __name__ == "__main__":
    main()
