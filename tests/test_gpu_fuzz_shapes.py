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


@pytest.mark.parametrize("seed", range(60))
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
    same = _close(pyi.bivariate(grid, x, y), want, what + " bilinear tolerance")
    assert same == 1.0 or want.dtype == np.float64, what  # float32 results: bit-identical in every mode (K1f)
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


@pytest.mark.parametrize("seed", range(24))
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


MODELS = ["bicubic", "bilinear", "akima", "akima_periodic", "c_spline", "c_spline_not_a_knot", "c_spline_periodic",
          "linear", "steffen", "polynomial"]


def _rand_axis(rng, n, kind):
    """(values, kwargs) of a random axis: regular with an inexact step, irregular, descending."""
    if kind == "periodic":
        return float(rng.choice([0.0, 10.0, -180.0])) + np.arange(n) * (360.0 / n), dict(period=360.0)
    step = float(rng.choice([0.25, 0.3, 1.0 / 3.0, 2.5]))
    v = float(rng.choice([0.0, -7.1, 100.3])) + np.arange(n) * step
    if kind == "irregular":
        v = v + rng.uniform(-0.3, 0.3, n) * step
    if kind == "descending":
        v = v[::-1].copy()
    return v, {}


@pytest.mark.parametrize("seed", range(90))
def test_random_windowed_models(pyi, orc_best, seed):
    """Every fitting model of the windowed path on random 2-D .. 4-D shapes, window
    sizes, boundary modes and axis kinds (tests/test_gpu_windowed_nd.py holds the
    fixed-shape cases)."""
    rng = np.random.default_rng(500 + seed)
    nd = int(rng.choice([2, 2, 3, 4]))
    model = MODELS[seed % len(MODELS)]
    boundary = str(rng.choice(["undef", "shrink"]))
    # smallest window a model can fit (fewer nodes: NaN everywhere, nothing to compare)
    least = {"akima": 3, "akima_periodic": 3, "bilinear": 1, "linear": 1, "polynomial": 1}.get(model, 2)
    half = (int(rng.integers(least, 5)), int(rng.integers(least, 5)))
    if seed % 2 == 0:
        half = (3, 3)  # the fast kernels' window (bicubic; the stencil also serves the linear-in-values models)
    if model == "c_spline_periodic":
        half, boundary = (half[0], half[0]), "undef"  # DESIGN.md, deviations: equal, unshrunken windows only
    nx, ny = int(rng.integers(9, 30)), int(rng.integers(9, 30))
    xv, xkw = _rand_axis(rng, nx, str(rng.choice(["periodic", "periodic", "regular", "irregular"])))
    yv, ykw = _rand_axis(rng, ny, str(rng.choice(["regular", "irregular", "descending"])))
    axes_v = [(xv, xkw), (yv, ykw)]
    n = 12_000
    q = [rng.uniform(xv.min() - 400.0 if xkw else xv.min() - 1.0, xv.max() + 400.0 if xkw else xv.max() + 1.0, n),
         rng.uniform(yv.min() - 0.5, yv.max() + 0.5, n)]
    q[0][::13] = xv[rng.integers(0, nx, q[0][::13].size)]
    q[1][1::17] = yv[rng.integers(0, ny, q[1][1::17].size)]
    q[0][2::499] = np.nan
    zm = um = 0
    if nd >= 3:
        zv, zkw = _rand_axis(rng, int(rng.integers(2, 6)), str(rng.choice(["regular", "irregular", "descending"])))
        axes_v.append((zv, zkw))
        q.append(rng.uniform(zv.min() - 0.3, zv.max() + 0.3, n))
        zm = int(rng.integers(0, 2))
    if nd >= 4:
        uv, ukw = _rand_axis(rng, int(rng.integers(2, 5)), str(rng.choice(["regular", "irregular"])))
        axes_v.append((uv, ukw))
        q.append(rng.uniform(uv.min() - 0.3, uv.max() + 0.3, n))
        um = int(rng.integers(0, 2))
    dtype = str(rng.choice(["float64", "float32"]))
    shape = tuple(len(a[0]) for a in axes_v)
    g = (rng.normal(size=shape) * 30.0).astype(dtype)
    g[rng.random(shape) < 0.0005] = np.nan
    oaxes = [orc_best.axis(v, **kw) for v, kw in axes_v]
    grid = pyi.Grid(*[pyi.Axis(v, **kw) for v, kw in axes_v], g)
    cls = {2: pyi.windowed.Bivariate, 3: pyi.windowed.Trivariate, 4: pyi.windowed.Quadrivariate}[nd]
    cfg = getattr(cls, model)().with_boundary_mode(getattr(pyi.windowed.BoundaryConfig, boundary)())
    cfg = cfg.with_half_window_size_x(half[0]).with_half_window_size_y(half[1])
    if nd >= 3:
        cfg = cfg.with_third_axis(getattr(pyi.windowed.AxisConfig, "nearest" if zm else "linear")())
    if nd >= 4:
        cfg = cfg.with_fourth_axis(getattr(pyi.windowed.AxisConfig, "nearest" if um else "linear")())
    fn = {2: pyi.core.bivariate, 3: pyi.core.trivariate, 4: pyi.core.quadrivariate}[nd]
    got = fn(grid, *q, cfg)
    want, _ = orc_best.windowed_nd(oaxes, g, *q, method=model, half_x=half[0], half_y=half[1], boundary=boundary,
                                   num_threads=0, z_method=zm, u_method=um)
    _close(got, want, f"seed {seed} {nd}-D {shape} {model} {half} {boundary} {dtype}")


def _irregular(rng, n, lo, hi):
    """Random irregular ascending nodes: uniform, clustered or geometric spacing."""
    style = int(rng.integers(0, 3))
    if style == 0:
        v = np.sort(rng.uniform(lo, hi, n))
    elif style == 1:  # clusters: many nodes inside a few bins of the lookup table
        centres = rng.uniform(lo, hi, max(2, n // 40))
        v = np.sort(centres[rng.integers(0, centres.size, n)] + rng.normal(0.0, (hi - lo) * 1e-4, n))
    else:
        v = lo + (hi - lo) * (np.geomspace(1.0, 1000.0, n) - 1.0) / 999.0
    v = np.unique(v)
    return v


@pytest.mark.parametrize("seed", range(200))
def test_random_geometric_axes(pyi, orc_best, seed, monkeypatch):
    """Geometric bilinear (2-D and 3-D) on random axes: periods other than 360,
    inexact regular steps, descending and irregular (uniform / clustered /
    geometric) latitudes of up to 3000 nodes, int64 time axes with large values,
    every grid dtype; queries many periods away and on nodes."""
    rng = np.random.default_rng(9000 + seed)
    period = float(rng.choice([360.0, 2.0 * np.pi, 24.0, 1.0, 1000.0]))
    nx = int(rng.integers(8, 300))
    if rng.integers(0, 4) == 0:  # open regular x
        xv, xkw = float(rng.uniform(-50, 50)) + np.arange(nx) * float(rng.uniform(0.01, 3.0)), {}
    else:
        xv, xkw = float(rng.choice([0.0, -period / 2, period * 0.1])) + np.arange(nx) * (period / nx), dict(period=period)
    ykind = int(rng.integers(0, 4))
    if ykind == 0:
        yv = _irregular(rng, int(rng.integers(3, 3000)), -80.0, 80.0)
    elif ykind == 1:
        yv = float(rng.uniform(-90, 0)) + np.arange(int(rng.integers(2, 400))) * float(rng.uniform(0.05, 0.7))
    elif ykind == 2:
        yv = (float(rng.uniform(-90, 0)) + np.arange(int(rng.integers(2, 400))) * float(rng.choice([0.25, 0.5, 0.1])))[::-1].copy()
    else:
        yv = -89.875 + np.arange(int(rng.integers(2, 200))) * 0.25
    ny = yv.size
    dtype = str(rng.choice(["float64", "float64", "float32", "int16", "uint8", "int64"]))
    nd = 2 if seed % 3 else 3
    shape = (nx, ny) if nd == 2 else (nx, ny, int(rng.integers(2, 6)))
    if nx * ny > 200_000:
        nd, shape = 2, (nx, ny)
    g = rng.normal(size=shape) * 40.0
    if np.dtype(dtype).kind == "f":
        g = g.astype(dtype)
        g[rng.random(shape) < 0.002] = np.nan
    else:
        info = np.iinfo(dtype)
        g = np.clip(np.round(g), info.min, info.max).astype(dtype)
    n = 40_000
    span = (period if xkw else xv[-1] - xv[0])
    far = float(rng.choice([2.0, 50.0, 1e5])) if xkw else 0.05
    x = rng.uniform(xv[0] - far * span, xv[-1] + far * span, n)
    y = rng.uniform(yv.min() - 0.01 * (yv.max() - yv.min() + 1), yv.max() + 0.01 * (yv.max() - yv.min() + 1), n)
    x[::11] = xv[rng.integers(0, nx, x[::11].size)] + (period * rng.integers(-3, 4, x[::11].size) if xkw else 0.0)
    y[1::13] = yv[rng.integers(0, ny, y[1::13].size)]
    y[2::17] = np.nextafter(yv[rng.integers(0, ny, y[2::17].size)], rng.choice([-1e9, 1e9]))
    x[3::997] = np.nan
    y[4::991] = np.inf
    oaxes = [orc_best.axis(xv, **xkw), orc_best.axis(yv)]
    axes = [pyi.Axis(xv, **xkw), pyi.Axis(yv)]
    q = [x, y]
    if nd == 3:
        t0 = int(rng.choice([0, 1_609_459_200_000_000_000, -4_000_000_000_000_000_000]))
        dt = int(rng.choice([1, 3_600_000_000_000, 86_400_000_000_000, 7]))
        tv = t0 + np.arange(shape[2], dtype=np.int64) * dt
        oaxes.append(orc_best.axis(tv))
        axes.append(pyi.TemporalAxis(tv.view("datetime64[ns]")))
        z = rng.integers(tv[0] - dt, tv[-1] + dt + 1, n, dtype=np.int64)
        z[5::19] = tv[rng.integers(0, tv.size, z[5::19].size)]
        z[6::983] = np.iinfo(np.int64).min
        q.append(z)
    grid = pyi.Grid(*axes, g)
    want, _ = orc_best.geometric(oaxes, g, *q, method=0, num_threads=0)
    what = f"seed {seed} {shape} {dtype} period={period if xkw else None} ykind={ykind}"
    args = list(q)
    if nd == 3:
        args[2] = args[2].view("datetime64[ns]")
    fn = pyi.core.bivariate if nd == 2 else pyi.core.trivariate
    cfg = (pyi.geometric.Bivariate if nd == 2 else pyi.geometric.Trivariate).bilinear()
    monkeypatch.setenv("PYINTERP_B200_BILINEAR", "tolerance")
    same = _close(fn(grid, *args, cfg), want, what + " tolerance")
    assert same == 1.0 or (nd == 2 and want.dtype == np.float64), what  # only K1t (float64, 2-D) rounds differently
    monkeypatch.setenv("PYINTERP_B200_BILINEAR", "exact")
    assert _close(fn(grid, *args, cfg), want, what + " exact") == 1.0, what


@pytest.mark.parametrize("seed", range(120))
def test_random_axis_lookups(pyi, orc_best, seed):
    """find_indexes / find_index on random axes -- float64 or int64, regular or
    irregular, ascending or descending, with or without a period (which may be
    longer than the axis) -- on adversarial coordinates: bit-exact."""
    import axes_zoo
    rng = np.random.default_rng(4000 + seed)
    n = int(rng.integers(1, 400))
    integer = bool(rng.integers(0, 3) == 0)
    irregular = bool(rng.integers(0, 2)) and n >= 3
    period = None
    if integer:
        step = int(rng.choice([1, 7, 3600, 86_400_000_000_000]))
        start = int(rng.choice([0, -31, 1_609_459_200_000_000_000]))
        v = start + np.arange(n, dtype=np.int64) * step
        if irregular:
            v = start + np.cumsum(rng.integers(1, 9, n)).astype(np.int64) * step
        if rng.integers(0, 3) == 0 and n >= 2:
            period = int((v[-1] - v[0]) + step * int(rng.integers(1, 4)))
    else:
        step = float(rng.choice([0.25, 0.1, 1.0 / 3.0, 17.142857142857142, 1e-3]))
        start = float(rng.choice([0.0, -180.0, 1e6, -89.875]))
        v = start + np.arange(n) * step
        if irregular:
            v = start + np.cumsum(rng.uniform(0.2, 3.0, n)) * step
        if rng.integers(0, 3) == 0 and n >= 2:
            period = float((v[-1] - v[0]) + step * float(rng.choice([1.0, 2.5, 1.0 + 1e-9])))
    flip = bool(rng.integers(0, 3) == 0) and n >= 2
    kw = {} if period is None else dict(period=period)
    try:
        c = orc_best.axis(v, **kw)
    except Exception:
        pytest.skip("the reference rejects this axis")
    if integer:
        per = kw.get("period")
        g = pyi.TemporalAxis(v.view("datetime64[ns]"), period=None if per is None else np.timedelta64(per, "ns"))
    else:
        g = pyi.Axis(v, **kw)
    if flip:
        g.flip(inplace=True)
        c.flip()
    d = c.describe()
    gd = g._desc()
    assert (gd.kind, gd.size, gd.is_ascending, gd.is_periodic) == (d["kind"], d["size"], d["is_ascending"], d["is_periodic"])
    coords = axes_zoo.coords_for(d, 20_000, seed)
    got = g.find_indexes(coords)
    want = c.find_indexes(coords)
    assert np.array_equal(got, want), (seed, np.flatnonzero((got != want).any(1))[:5])
    for bounded in (False, True):
        assert np.array_equal(g.find_index(coords, bounded), c.find_index(coords, bounded)), (seed, bounded)


@pytest.mark.parametrize("seed", range(40))
def test_random_fills(pyi, orc_best, seed):
    """gauss_seidel / multigrid / loess on random shapes (1 x n, n x 1, odd, tiny)
    and parameters: grid, iteration count and residual bit for bit."""
    import golden_util as gu
    rng = np.random.default_rng(7000 + seed)
    shape = (int(rng.integers(1, 90)), int(rng.integers(1, 90)))
    if seed % 7 == 0:
        shape = (1, shape[1]) if seed % 2 else (shape[0], 1)
    dtype = str(rng.choice(["float64", "float32"]))
    g0 = gu.fill_field(shape, dtype, 7100 + seed, blobs=int(rng.integers(1, 4)))
    fg = str(rng.choice(["zonal_average", "zero"]))
    per = bool(rng.integers(0, 2))
    its = int(rng.integers(0, 40))
    eps = float(rng.choice([0.0, 1e-4, 1e-2]))
    what = (seed, shape, dtype, fg, per, its, eps)
    relax = float(rng.choice([1.0, 1.3, 1.9]))
    want = g0.copy()
    r_want = orc_best.fill_gauss_seidel(want, fg, per, its, eps, relax, num_threads=3)
    got = g0.copy()
    assert pyi.fill.gauss_seidel(got, first_guess=fg, is_periodic=per, max_iterations=its, epsilon=eps,
                                 relaxation=relax) == r_want, what
    assert np.array_equal(got, want, equal_nan=True), what
    pre, post = int(rng.integers(0, 4)), int(rng.integers(1, 4))
    want = g0.copy()
    r_want = orc_best.fill_multigrid(want, fg, per, min(its, 6), eps, pre, post, num_threads=3)
    got = g0.copy()
    assert pyi.fill.multigrid(got, first_guess=fg, is_periodic=per, max_iterations=min(its, 6), epsilon=eps,
                              pre_smooth=pre, post_smooth=post) == r_want, what
    assert np.array_equal(got, want, equal_nan=True), what
    hx, hy = int(rng.integers(1, 7)), int(rng.integers(1, 7))
    vt = str(rng.choice(["undefined", "defined", "all"]))
    want = orc_best.fill_loess(g0, hx, hy, vt, fg, per, min(its, 12), eps)
    got = pyi.fill.loess(g0, nx=hx, ny=hy, value_type=vt, first_guess=fg, is_periodic=per,
                         max_iterations=min(its, 12), epsilon=eps)
    assert np.array_equal(got, want, equal_nan=True), what + (hx, hy, vt)


def test_ragged_sizes_through_the_host_pipeline(pyi, orc_best, monkeypatch):
    """Empty, tiny and ragged point counts (around the pipeline's chunk sizes and
    the vector widths of the kernels), host and device entry, every path of the
    0.25 degree grid: identical to each other and to the oracle."""
    import torch
    s = pyi.synthetic
    lon, lat = s.lon_axis_values(), s.lat_axis_values()
    g = s.field(lon, lat)
    g[::37, ::11] = np.nan
    t = s.time_axis_values(6)
    g3 = s.field(lon, lat, nz=6, dtype="float32", noise=0.0)
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g, resident=True)
    grid3 = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), pyi.TemporalAxis(t), g3, resident=True)
    o2 = [orc_best.axis(lon, period=360.0), orc_best.axis(lat)]
    o3 = o2 + [orc_best.axis(t.astype("datetime64[ns]").astype(np.int64))]
    nmax = (1 << 23) + 3
    x, y, z = s.queries(nmax, 77, ndim=3, time_values=t)
    sizes = [0, 1, 2, 3, 5, 31, 32, 33, 127, 129, 1000, (1 << 17) - 1, (1 << 17) + 1, (1 << 20) - 1, (1 << 20) + 1,
             3 * (1 << 20) + 5, (1 << 22) - 1, (1 << 22), (1 << 22) + 7, nmax]
    want_bil, _ = orc_best.geometric(o2, g, x, y, method=0, num_threads=0)
    want_tri, _ = orc_best.geometric(o3, g3, x, y, z, method=0, num_threads=0)
    want_cub, _ = orc_best.windowed2d(o2, g, x[:1 << 21], y[:1 << 21], num_threads=0, stateless=True)
    xd, yd, zd = (torch.from_numpy(v).cuda() for v in (x, y, z))
    for n in sizes:
        for mode in ("tolerance", "exact"):
            monkeypatch.setenv("PYINTERP_B200_BILINEAR", mode)
            host = pyi.bivariate(grid, x[:n], y[:n])
            dev = pyi.bivariate(grid, xd[:n], yd[:n]).cpu().numpy()
            assert host.shape == (n,) and np.array_equal(host, dev, equal_nan=True), (n, mode)
            if mode == "exact":
                assert np.array_equal(host, want_bil[:n], equal_nan=True), n
            else:
                _close(host, want_bil[:n], f"bilinear n={n}")
        host = pyi.trivariate(grid3, x[:n], y[:n], z[:n].view("datetime64[ns]"))
        dev = pyi.trivariate(grid3, xd[:n], yd[:n], zd[:n]).cpu().numpy()
        assert host.dtype == np.float32 and np.array_equal(host, dev, equal_nan=True), n
        assert np.array_equal(host, want_tri[:n], equal_nan=True), n
        m = min(n, 1 << 21)
        host = pyi.bivariate(grid, x[:m], y[:m], "bicubic")
        dev = pyi.bivariate(grid, xd[:m], yd[:m], "bicubic").cpu().numpy()
        assert np.array_equal(host, dev, equal_nan=True), n
        _close(host, want_cub[:m], f"bicubic n={m}")
