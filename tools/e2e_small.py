#!/usr/bin/env python
"""Host-array call time against the point count (fixed overhead of the
pipelined host path), page-locked and pageable inputs, one device."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pyinterp_b200 as pyi  # noqa: E402

syn = pyi.synthetic
lon, lat = syn.lon_axis_values(), syn.lat_axis_values()
grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), syn.field(lon, lat), resident=True)
for n in (100_000, 1_000_000, 6_000_000, 24_000_000, 100_000_000):
    x, y = syn.queries(n, 5)
    xp, yp = (torch.from_numpy(a).pin_memory().numpy() for a in (x, y))
    for name, a, b in (("pinned", xp, yp), ("pageable", x, y)):
        for _ in range(3):
            pyi.bivariate(grid, a, b)
        t0 = time.perf_counter()
        reps = 10
        for _ in range(reps):
            pyi.bivariate(grid, a, b)
        dt = (time.perf_counter() - t0) / reps
        print(f"n={n:10d} {name:8s} {dt * 1e3:8.3f} ms  {n / dt / 1e9:6.3f} Gpts/s  {n * 24 / dt / 1e9:6.1f} GB/s", flush=True)
