"""Randomised grid shapes against the CPU oracle: small and odd sizes, periodic
or open longitude, ascending or descending latitude, float64 / float32 grids
with NaN cells at the corners of the array -- the places where the fast
kernels' aligned vector loads, seam cells and "cell 0" fallbacks live.  Dense
queries (every cell is hit many times, edges and nodes included).

The same file is what `tools/gpu_run_check.sh` runs on the bounds-checked
build (csrc/Makefile `check`)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = {np.dtype("float64"): 1e-12, np.dtype("float32"): 1e-5}


@pytest.fixture(scope="module")
def orc_best():
    import orc
    return orc.Oracle("reference") if orc.available("reference") else orc.Oracle("port")


def _close(got, want, what):
    assert got.dtype == want.dtype, what
    ng, nw = np.isnan(got), np.isnan(want)
    assert np.array_equal(ng, nw), f"{what}: NaN positions differ at {np.flatnonzero(ng != nw)[:6]}"
    g, w = got[~ng].astype(np.float64), want[~ng].astype(np.float64)
    err = np.abs(g - w)
    bound = TOL[got.dtype] * np.maximum(np.maximum(np.abs(g), np.abs(w)), 1.0)
    assert np.all(err <= bound), f"{what}: worst {err.max():.3e}"
    return float((err == 0).mean()) if err.size else 1.0


def _case(seed):
    rng = np.random.default_rng(seed)
    nx = int(rng.integers(6, 40))
    ny = int(rng.integers(6, 40))
    periodic = bool(rng.integers(0, 2))
    step_x = float(rng.choice([0.25, 0.5, 1.0, 0.3, 360.0 / nx]))
    if periodic:
        step_x = 360.0 / nx  # the axis closes its period (is_periodic)
    x0 = float(rng.choice([0.0, -180.0, 10.0]))
    lon = x0 + np.arange(nx) * step_x
    step_y = float(rng.choice([0.25, 0.5, 1.0, 0.3]))
    lat = float(rng.choice([-5.0, 0.0, -89.0 + 0.1])) + np.arange(ny) * step_y
    if rng.integers(0, 3) == 0:
        lat = lat[::-1].copy()
    dtype = str(rng.choice(["float64", "float32"]))
    g = rng.normal(size=(nx, ny)) * 10.0
    g = g.astype(dtype)
    # NaN at the four corners of the array and a few inside
    for i, j in ((0, 0), (0, ny - 1), (nx - 1, 0), (nx - 1, ny - 1)):
        if rng.integers(0, 2):
            g[i, j] = np.nan
    g[rng.random(g.shape) < 0.01] = np.nan
    n = 30_000
    span = lon[-1] - lon[0] + step_x
    x = rng.uniform(lon[0] - 1.5 * span, lon[-1] + 1.5 * span, n)
    y = rng.uniform(lat.min() - 2 * step_y, lat.max() + 2 * step_y, n)
    x[0::17] = lon[rng.integers(0, nx, x[0::17].size)]
    y[1::19] = lat[rng.integers(0, ny, y[1::19].size)]
    x[2::23] = lon[rng.integers(0, nx, x[2::23].size)] + (360.0 if periodic else 0.0)
    y[3::29] = np.nextafter(lat[rng.integers(0, ny, y[3::29].size)], 1e9)
    x[4::997] = np.nan
    y[5::991] = np.nan
    return dict(lon=lon, lat=lat, periodic=periodic, g=g, x=x, y=y, dtype=dtype)


@pytest.mark.parametrize("seed", range(24))
def test_random_2d_shapes(pyi, orc_best, seed, monkeypatch):
    c = _case(seed)
    kw = dict(period=360.0) if c["periodic"] else {}
    oaxes = [orc_best.axis(c["lon"], **kw), orc_best.axis(c["lat"])]
    grid = pyi.Grid(pyi.Axis(c["lon"], **kw), pyi.Axis(c["lat"]), c["g"])
    x, y = c["x"], c["y"]
    what = f"seed {seed} {c['g'].shape} {c['dtype']} periodic={c['periodic']}"
    # geometric bilinear: tolerance kernel forced (float64 results), then bit-identical
    want, _ = orc_best.geometric(oaxes, c["g"], x, y, method=0, num_threads=0)
    monkeypatch.setenv("PYINTERP_B200_BILINEAR", "tolerance")
    _close(pyi.bivariate(grid, x, y), want, what + " bilinear tolerance")
    monkeypatch.setenv("PYINTERP_B200_BILINEAR", "exact")
    assert _close(pyi.bivariate(grid, x, y), want, what + " bilinear exact") == 1.0
    # windowed bicubic, both boundary modes: separable / weights kernels, then bit-identical
    for boundary in ("undef", "shrink"):
        want, _ = orc_best.windowed2d(oaxes, c["g"], x, y, boundary=boundary, num_threads=0, stateless=True)
        for mode in ("separable", "weights", "exact"):
            if mode == "separable":
                monkeypatch.delenv("PYINTERP_B200_BICUBIC", raising=False)
            else:
                monkeypatch.setenv("PYINTERP_B200_BICUBIC", mode)
            got = pyi.bivariate(grid, x, y, "bicubic", boundary_mode=boundary)
            same = _close(got, want, f"{what} bicubic {boundary} {mode}")
            if mode == "exact":
                assert same == 1.0, f"{what} bicubic {boundary} exact: {same}"


@pytest.mark.parametrize("seed", range(8))
def test_random_3d_4d_shapes(pyi, orc_best, seed):
    rng = np.random.default_rng(100 + seed)
    c = _case(200 + seed)
    nz, nu = int(rng.integers(2, 7)), int(rng.integers(2, 6))
    z = np.cumsum(rng.uniform(0.5, 2.0, nz)) if rng.integers(0, 2) else np.arange(nz) * 2.5
    if rng.integers(0, 3) == 0 and np.allclose(np.diff(z), z[1] - z[0]):
        z = z[::-1].copy()  # a descending regular level axis
    u = np.arange(nu) * 10.0
    kw = dict(period=360.0) if c["periodic"] else {}
    n = c["x"].size
    qz = rng.uniform(z.min() - 1.0, z.max() + 1.0, n)
    qu = rng.uniform(-5.0, u[-1] + 5.0, n)
    qz[0::13] = z[rng.integers(0, nz, qz[0::13].size)]
    for nd in (3, 4):
        shape = c["g"].shape + ((nz,) if nd == 3 else (nz, nu))
        g = (rng.normal(size=shape) * 10.0).astype(c["dtype"])
        g[rng.random(shape) < 0.01] = np.nan
        g.reshape(-1)[0] = np.nan
        g.reshape(-1)[-1] = np.nan
        oaxes = [orc_best.axis(c["lon"], **kw), orc_best.axis(c["lat"]), orc_best.axis(z)]
        axes = [pyi.Axis(c["lon"], **kw), pyi.Axis(c["lat"]), pyi.Axis(z)]
        q = [c["x"], c["y"], qz]
        if nd == 4:
            oaxes.append(orc_best.axis(u))
            axes.append(pyi.Axis(u))
            q.append(qu)
        grid = pyi.Grid(*axes, g)
        want, _ = orc_best.geometric(oaxes, g, *q, method=0, num_threads=0)
        fn = pyi.core.trivariate if nd == 3 else pyi.core.quadrivariate
        cls = pyi.geometric.Trivariate if nd == 3 else pyi.geometric.Quadrivariate
        got = fn(grid, *q, cls.bilinear())
        assert _close(got, want, f"seed {seed} {shape} {c['dtype']}") == 1.0
