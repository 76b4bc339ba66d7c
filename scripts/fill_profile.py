"""One device-resident fill call (profiling target): python scripts/fill_profile.py [method] [iterations]"""
import sys, importlib
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import golden_util as gu, torch
pyi = importlib.import_module("pangeo-pyinterp_b200")
method = sys.argv[1] if len(sys.argv) > 1 else "gauss_seidel"
its = int(sys.argv[2]) if len(sys.argv) > 2 else 50
g0 = gu.fill_field((1440, 720), "float64", 7, fraction=0.05, blobs=12)
t = torch.from_numpy(g0).cuda()
fn = getattr(pyi.fill, method)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for rep in range(3):
    t.copy_(torch.from_numpy(g0))
    torch.cuda.synchronize()
    ev[0].record()
    r = fn(t, max_iterations=its, epsilon=0.0)
    ev[1].record()
    torch.cuda.synchronize()
    print(method, r, "%.3f ms" % ev[0].elapsed_time(ev[1]), "launches", pyi.launch_count())
