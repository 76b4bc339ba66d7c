"""Diagnostic: product fill vs the oracle per shape / option (GPU box)."""
import sys, importlib, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests'); sys.path.insert(0, '/root/repo/oracle')
import golden_util as gu, orc
pyi = importlib.import_module("pangeo-pyinterp_b200")
o = orc.Oracle("reference") if orc.available("reference") else orc.Oracle("port")
for shape in [(40, 30), (37, 23), (37, 24), (36, 23), (5, 64), (3, 3), (129, 67)]:
    for per in (True, False):
        for fg in ("zero", "zonal_average"):
            g0 = gu.fill_field(shape, "float64", 201, blobs=2)
            row = []
            for its in (1, 2, 3, 10, 42):
                w = g0.copy(); rw = o.fill_gauss_seidel(w, fg, per, its, 0.0, 1.0)
                g = g0.copy(); rg = pyi.fill.gauss_seidel(g, first_guess=fg, is_periodic=per, max_iterations=its, epsilon=0.0)
                bad = np.argwhere(w != g)
                row.append((its, rw == rg, len(bad), bad[:3].tolist()))
            print(shape, per, fg, row)
