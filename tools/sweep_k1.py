#!/usr/bin/env python
"""Sweep of the K1 kernel variants on one GPU: exact kernel (PYI_FAST_VARIANT)
and tolerance kernel (PYI_TOL_VARIANT = <points per thread><blocks per SM>),
regular and irregular latitude, 1e9 device-resident points, CUDA events.
Prints one line per variant: Gpts/s and fraction of the measured HBM peak."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pyinterp_b200 as pyi  # noqa: E402

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
peak = 6550.0
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
syn = pyi.synthetic
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev)
gen.manual_seed(1005)
x = torch.rand(n, dtype=torch.float64, device=dev, generator=gen).mul_(720.0).sub_(180.0)
y = torch.rand(n, dtype=torch.float64, device=dev, generator=gen).mul_(180.0).sub_(90.0)
lon = syn.lon_axis_values()


def run(grid, cfg, steps=8):
    for _ in range(3):
        out = pyi.core.bivariate(grid, x, y, cfg)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        out = pyi.core.bivariate(grid, x, y, cfg)
    b.record()
    torch.cuda.synchronize()
    del out
    return a.elapsed_time(b) / steps


for lat_name, lat in (("regular", syn.lat_axis_values()), ("irregular", syn.irregular_lat_values())):
    g = torch.from_numpy(syn.field(lon, lat)).to(dev)
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g)
    exact = pyi.geometric.Bivariate.bilinear().with_value_mode("exact")
    tol = pyi.geometric.Bivariate.bilinear().with_value_mode("tolerance")
    ms = run(grid, exact)
    print(f"{lat_name:9s} exact   default  {ms:8.3f} ms {n / ms / 1e6:8.1f} Gpts/s  frac {n * 24 / ms / 1e6 / peak:.4f}", flush=True)
    variants = ["23", "23n", "22n", "24", "24n", "25", "25n", "42n", "43n", "44n"] if lat_name == "regular" else ["23", "23n", "24n", "25n"]
    for v in variants:
        os.environ["PYI_TOL_VARIANT"] = v
        ms = run(grid, tol)
        print(f"{lat_name:9s} tol     {v:8s} {ms:8.3f} ms {n / ms / 1e6:8.1f} Gpts/s  frac {n * 24 / ms / 1e6 / peak:.4f}", flush=True)
    os.environ.pop("PYI_TOL_VARIANT", None)
    del grid, g
