"""Throughput of windowed bicubic (6 x 6) on the 0.25 degree grid stored as
float64 / float32: device-resident queries, CUDA events.
Usage: python tools/bicubic_dtype_bench.py [n_points]"""
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
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 50_000_000
    lon, lat = s.lon_axis_values(), s.lat_axis_values()
    g = s.field(lon, lat)
    x, y = s.queries(n, 1002)
    xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    for dtype in ("float64", "float32"):
        grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g.astype(dtype), resident=True)
        for mode in ("separable (default)", "weights", "exact"):
            if mode.startswith("sep"):
                os.environ.pop("PYINTERP_B200_BICUBIC", None)
            else:
                os.environ["PYINTERP_B200_BICUBIC"] = mode
            ms = timed(lambda: pyi.bivariate(grid, xd, yd, "bicubic"))
            print(f"bicubic {dtype} {mode:20s} {n / ms / 1e6:8.2f} Gpts/s")
    os.environ.pop("PYINTERP_B200_BICUBIC", None)


if __name__ == "__main__":
    main()
