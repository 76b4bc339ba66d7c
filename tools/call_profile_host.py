import sys, time, os
sys.path.insert(0, '/root/repo')
import numpy as np
import pyinterp_b200 as pyi
rng = np.random.default_rng(0)
xa = np.linspace(0, 360, 720); ya = np.linspace(-90, 90, 360)
data = rng.normal(size=(720, 360))
for resident in (False, True):
    grid = pyi.Grid(pyi.Axis(xa, period=360.0), pyi.Axis(ya), data, resident=resident)
    for n in (10_000, 100_000, 1_000_000):
        x = rng.uniform(0, 360, n); y = rng.uniform(-90, 90, n)
        cfg = pyi.geometric.Bivariate.bilinear()
        for _ in range(5): pyi.core.bivariate(grid, x, y, cfg)
        t0 = time.perf_counter()
        for _ in range(20): pyi.core.bivariate(grid, x, y, cfg)
        dt = (time.perf_counter() - t0) / 20
        print(f"resident={resident} n={n} {dt*1e3:.3f} ms", flush=True)
import cProfile, pstats
grid = pyi.Grid(pyi.Axis(xa, period=360.0), pyi.Axis(ya), data)
n = 1_000_000
x = rng.uniform(0, 360, n); y = rng.uniform(-90, 90, n)
pr = cProfile.Profile(); pr.enable()
for _ in range(50): pyi.core.bivariate(grid, x, y, cfg)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(8)
