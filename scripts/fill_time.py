import sys, time, importlib, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests'); sys.path.insert(0, '/root/repo/oracle')
import golden_util as gu, orc, torch
pyi = importlib.import_module("pangeo-pyinterp_b200")
o = orc.Oracle("reference") if orc.available("reference") else orc.Oracle("port")
for dtype in ("float64", "float32"):
    g0 = gu.fill_field((1440, 720), dtype, 7, fraction=0.05, blobs=12)
    for its in (100, 500):
        w = g0.copy(); t0 = time.perf_counter(); rw = o.fill_gauss_seidel(w, "zonal_average", True, its, 0.0, 1.0, num_threads=0); tc = time.perf_counter() - t0
        g = g0.copy(); pyi.fill.gauss_seidel(g.copy(), max_iterations=10)
        t0 = time.perf_counter(); rg = pyi.fill.gauss_seidel(g, max_iterations=its, epsilon=0.0); th = time.perf_counter() - t0
        t = torch.from_numpy(g0.copy()).cuda(); torch.cuda.synchronize()
        t0 = time.perf_counter(); rd = pyi.fill.gauss_seidel(t, max_iterations=its, epsilon=0.0); torch.cuda.synchronize(); td = time.perf_counter() - t0
        print(dtype, its, "cpu(ref,%s) %.1f ms" % (o.kind, tc * 1e3), "gpu host %.2f ms" % (th * 1e3), "gpu device %.2f ms (%.2f us/iter)" % (td * 1e3, td * 1e6 / its), rw == rg == rd, np.array_equal(w, g), np.array_equal(w, t.cpu().numpy()))

for dtype in ("float64", "float32"):
    g0 = gu.fill_field((1440, 720), dtype, 7, fraction=0.05, blobs=12)
    for its in (5, 20):
        w = g0.copy(); t0 = time.perf_counter(); rw = o.fill_multigrid(w, "zonal_average", True, its, 0.0, 2, 2, num_threads=0); tc = time.perf_counter() - t0
        g = g0.copy(); pyi.fill.multigrid(g.copy(), max_iterations=2)
        t0 = time.perf_counter(); rg = pyi.fill.multigrid(g, max_iterations=its, epsilon=0.0); th = time.perf_counter() - t0
        t = torch.from_numpy(g0.copy()).cuda(); torch.cuda.synchronize()
        t0 = time.perf_counter(); rd = pyi.fill.multigrid(t, max_iterations=its, epsilon=0.0); torch.cuda.synchronize(); td = time.perf_counter() - t0
        print("multigrid", dtype, its, "cpu(ref,%s) %.1f ms" % (o.kind, tc * 1e3), "gpu host %.2f ms" % (th * 1e3), "gpu device %.2f ms (%.1f us/cycle)" % (td * 1e3, td * 1e6 / its), rw == rg == rd, np.array_equal(w, g), np.array_equal(w, t.cpu().numpy()))
