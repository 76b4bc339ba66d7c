#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/g_tests.log 2>&1; echo "tests rc=$?"; tail -6 gpurun_out/g_tests.log
python - <<'PY' > gpurun_out/g_loess_time.txt 2>&1
import time, numpy as np, sys
sys.path.insert(0, '.'); sys.path.insert(0, 'oracle')
import pyinterp_b200 as pyi, orc, torch
s = pyi.synthetic
g = s.field(s.lon_axis_values(), s.lat_axis_values())
rng = np.random.default_rng(4); g[rng.random(g.shape) < 0.3] = np.nan
o = orc.Oracle("reference") if orc.available("reference") else orc.Oracle("port")
d = torch.from_numpy(g).cuda()
for its in (1, 10):
    pyi.fill.loess(d, nx=5, ny=5, max_iterations=its, epsilon=0.0); torch.cuda.synchronize()
    t0 = time.perf_counter(); pyi.fill.loess(d, nx=5, ny=5, max_iterations=its, epsilon=0.0); torch.cuda.synchronize(); tg = time.perf_counter() - t0
    t0 = time.perf_counter(); o.fill_loess(g, 5, 5, "undefined", "zonal_average", True, its, 0.0, num_threads=0); tc = time.perf_counter() - t0
    print(f"loess 1440x720 f64, 30% undefined, 11x11 window, max_iterations={its}: device {tg*1e3:.2f} ms, reference CPU (all threads) {tc*1e3:.1f} ms")
PY
cat gpurun_out/g_loess_time.txt
