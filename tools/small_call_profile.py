"""Where a small device-resident call spends its time: wall time per call at several sizes, and a
cProfile of 2000 calls at n = 1000 (pure overhead)."""
import cProfile
import os
import pstats
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pyinterp_b200 as pyi  # noqa: E402

syn = pyi.synthetic
lon, lat = syn.lon_axis_values(), syn.lat_axis_values()
grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), syn.field(lon, lat), resident=True)
cfg = pyi.geometric.Bivariate.bilinear()
for n in (1_000, 100_000, 1_000_000, 4_000_000, 8_000_000):
    x, y = syn.queries(n, 5)
    xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    for fn, name in ((lambda: pyi.bivariate(grid, xd, yd), "pyi.bivariate(str)"),
                     (lambda: pyi.core.bivariate(grid, xd, yd, cfg), "core.bivariate(cfg)")):
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
        reps = 300
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        print(f"n={n:9d} {name:20s} host {1e6 * (t1 - t0) / reps:7.1f} us/call, with sync {1e6 * (t2 - t0) / reps:7.1f} us/call "
              f"-> {n * reps / (t2 - t0) / 1e9:6.2f} Gpts/s", flush=True)
x, y = syn.queries(1000, 5)
xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
pr = cProfile.Profile()
pr.enable()
for _ in range(2000):
    pyi.core.bivariate(grid, xd, yd, cfg)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
