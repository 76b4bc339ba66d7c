#!/usr/bin/env python
"""The reference's own benchmark harness, interpolation part, on this library
(/root/reference/scripts/benchmark_suite.py:98-166 grids, :170-212 points, :229-256 protocol,
:305-360 / :414-460 / :520-560 method lists, :862-882 sizes): the same synthetic grids (float64,
linspace axes with period 360 on x), the same point generator, the same methods with their default
configurations, host numpy arrays in and numpy arrays out (`core.bivariate(grid, x, y, config)`, the
call the reference's harness times), 1 warm-up + 5 timed runs, mean time and Mpts/s.  Beside every line:
the reference's CPU implementation (oracle/_ref: the reference's headers, num_threads=0 = all host
threads) on the same arrays, its result compared with the GPU's (NaN positions identical, values within
the tolerance of the tests).  The oracle is the checker and the baseline here, as in bench.py.

usage: python tools/reference_suite.py [--quick] [--iterations 5] [--no-cpu]
"""
import argparse
import os
import statistics
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pyinterp_b200 as pyi  # noqa: E402

RNG = np.random.default_rng(42)


def axis_values(start, stop, num):
    return np.linspace(start, stop, num, dtype=np.float64)


def make_case(shape):
    nd = len(shape)
    vals = [axis_values(0, 360, shape[0]), axis_values(-90, 90, shape[1])]
    kw = [dict(period=360.0), {}]
    if nd >= 3:
        vals.append(axis_values(0, 1000, shape[2]))
        kw.append({})
    if nd >= 4:
        vals.append(axis_values(0, 100, shape[3]))
        kw.append({})
    if nd == 2:
        xx, yy = np.meshgrid(np.linspace(0, 360, shape[0]), np.linspace(-90, 90, shape[1]))
        data = np.ascontiguousarray((np.sin(np.radians(xx)) * np.cos(np.radians(yy))).T.astype(np.float64))
    else:
        data = RNG.standard_normal(shape).astype(np.float64)
    return vals, kw, data


def points(nd, n):
    ranges = [(0, 360), (-90, 90), (0, 1000), (0, 100)][:nd]
    return [RNG.uniform(lo, hi, n).astype(np.float64) for lo, hi in ranges]


def timed(fn, iterations):
    fn()
    ts = []
    for _ in range(iterations):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return statistics.mean(ts), (statistics.stdev(ts) if len(ts) > 1 else 0.0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--iterations", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if pyi.device_count() < 1:
        raise SystemExit("reference_suite.py: no CUDA device (the library has no CPU fallback)")
    o = None
    if not args.no_cpu:
        import orc
        o = orc.Oracle("reference" if orc.available("reference") else "port")
    if args.quick:
        sizes = {2: [(180, 90)], 3: [(90, 45, 20)], 4: [(45, 23, 10, 5)]}
        counts = [10_000, 100_000]
    else:
        sizes = {2: [(180, 90), (360, 180), (720, 360)], 3: [(180, 90, 50), (360, 180, 100)], 4: [(90, 45, 20, 10)]}
        counts = [10_000, 100_000, 1_000_000]
    geometric = ["bilinear", "nearest", "idw"]
    windowed = {2: ["bicubic", "linear", "c_spline", "akima", "steffen", "polynomial"],
                3: ["bicubic", "linear", "c_spline", "akima"], 4: ["bicubic", "linear"]}
    g_cls = {2: pyi.geometric.Bivariate, 3: pyi.geometric.Trivariate, 4: pyi.geometric.Quadrivariate}
    w_cls = {2: pyi.windowed.Bivariate, 3: pyi.windowed.Trivariate, 4: pyi.windowed.Quadrivariate}
    api = {2: pyi.core.bivariate, 3: pyi.core.trivariate, 4: pyi.core.quadrivariate}
    cores = os.cpu_count()
    print(f"reference benchmark suite on pyinterp_b200; iterations {args.iterations}; host threads {cores}; "
          f"CPU column: oracle kind = {o.kind if o else 'off'} (num_threads=0)")
    worst = 0.0
    for nd in (2, 3, 4):
        for shape in sizes[nd]:
            vals, kw, data = make_case(shape)
            grid = pyi.Grid(*[pyi.Axis(v, **k) for v, k in zip(vals, kw)], data)
            oaxes = [o.axis(v, **k) for v, k in zip(vals, kw)] if o else None
            print(f"\n--- {nd}-D grid {' x '.join(map(str, shape))} ---")
            for n in counts:
                q = points(nd, n)
                print(f"  points: {n:,}")
                for kind, names in (("geometric", geometric), ("windowed", windowed[nd])):
                    for name in names:
                        cfg = getattr(g_cls[nd] if kind == "geometric" else w_cls[nd], name)()
                        mean, std = timed(lambda: api[nd](grid, *q, cfg), args.iterations)
                        line = (f"    {kind[:4]} {name:11s}: {mean * 1e3:8.3f} ms (+-{std * 1e3:.3f}) | "
                                f"{n / mean / 1e6:9.2f} Mpts/s")
                        if o:
                            if kind == "geometric":
                                m = {"bilinear": 0, "idw": 1, "nearest": 2}[name]
                                zm = 1 if name == "nearest" else 0  # config/geometric.hpp:127-139: tied to the method
                                cpu = lambda: o.geometric(oaxes, data, *q, method=m, z_method=zm, u_method=zm,  # noqa: E731
                                                          num_threads=0)[0]
                            else:
                                cpu = lambda: o.windowed_nd(oaxes, data, *q, method=name, num_threads=0,  # noqa: E731
                                                            stateless=o.kind != "reference")[0]
                            cmean, _ = timed(cpu, max(2, args.iterations // 2))
                            got, want = api[nd](grid, *q, cfg), cpu()
                            same_nan = np.array_equal(np.isnan(got), np.isnan(want))
                            ok = ~np.isnan(want)
                            err = float(np.max(np.abs(got[ok] - want[ok]) / np.maximum(np.abs(want[ok]), 1.0))) if ok.any() else 0.0
                            worst = max(worst, err) if kind == "windowed" or name == "bilinear" else worst
                            line += (f" | CPU {cmean * 1e3:9.3f} ms {n / cmean / 1e6:8.2f} Mpts/s | x{cmean / mean:7.1f}"
                                     f" | NaN {'same' if same_nan else 'DIFFER'} err {err:.1e}")
                        print(line, flush=True)
    print(f"\nworst relative difference to the reference (bilinear and windowed models): {worst:.2e}")
    # gap filling (benchmark_suite.py:673-706): Gauss-Seidel, 100 iterations at most, epsilon 1e-6, 10 % of the
    # cells undefined.  The harness fills ONE buffer in its warm-up run and times the following runs on that
    # (already filled) buffer; both that protocol and a fresh copy per run are timed here.
    print("\n--- gap filling: fill.gauss_seidel(max_iterations=100, epsilon=1e-6) ---")
    for nx, ny in ([(180, 90)] if args.quick else [(360, 180), (720, 360)]):
        data = RNG.standard_normal((nx, ny)).astype(np.float64)
        data[RNG.random((nx, ny)) < 0.1] = np.nan
        buf = data.copy()
        mean_h, _ = timed(lambda: pyi.fill.gauss_seidel(buf, max_iterations=100, epsilon=1e-6), args.iterations)
        fresh = [data.copy() for _ in range(args.iterations + 1)]
        it = iter(fresh)
        mean_f, _ = timed(lambda: pyi.fill.gauss_seidel(next(it), max_iterations=100, epsilon=1e-6), args.iterations)
        line = f"  {nx} x {ny}: harness protocol {mean_h * 1e3:8.3f} ms | fresh copy per run {mean_f * 1e3:8.3f} ms"
        if o:
            cbuf = data.copy()
            cmean_h, _ = timed(lambda: o.fill_gauss_seidel(cbuf, max_iterations=100, epsilon=1e-6, num_threads=0),
                               max(2, args.iterations // 2))
            cfresh = [data.copy() for _ in range(4)]
            cit = iter(cfresh)
            cmean_f, _ = timed(lambda: o.fill_gauss_seidel(next(cit), max_iterations=100, epsilon=1e-6,
                                                           num_threads=0), 3)
            same = np.array_equal(fresh[0], cfresh[0])
            line += (f" | CPU {cmean_h * 1e3:8.3f} / {cmean_f * 1e3:8.3f} ms | x{cmean_h / mean_h:6.1f} / x{cmean_f / mean_f:6.1f}"
                     f" | filled grids {'bit-identical' if same else 'DIFFER'}")
        print(line, flush=True)


if __name__ == "__main__":
    main()
