#!/usr/bin/env python
"""Pageable-input call time against the number of staging threads
(PYINTERP_B200_COPY_THREADS is read once per process: run once per value)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pyinterp_b200 as pyi  # noqa: E402

syn = pyi.synthetic
lon, lat = syn.lon_axis_values(), syn.lat_axis_values()
grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), syn.field(lon, lat), resident=True)
for n in (6_000_000, 100_000_000):
    x, y = syn.queries(n, 5)
    for _ in range(3):
        pyi.bivariate(grid, x, y)
    t0 = time.perf_counter()
    reps = 8
    for _ in range(reps):
        pyi.bivariate(grid, x, y)
    dt = (time.perf_counter() - t0) / reps
    print(f"threads={os.environ.get('PYINTERP_B200_COPY_THREADS', 'default')} cpus={os.cpu_count()} n={n:10d} "
          f"{dt * 1e3:8.3f} ms  {n / dt / 1e9:6.3f} Gpts/s  {n * 24 / dt / 1e9:6.1f} GB/s", flush=True)
