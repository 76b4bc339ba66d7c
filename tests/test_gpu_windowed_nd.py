"""GPU parity of the widened windowed path: every fitting
model on 1-D .. 4-D grids through the public API / C-ABI (pyi_windowed)
against the CPU oracle on identical seeded inputs: the reference build
(oracle/_ref: the reference's own InterpolationCache, factory(), Univariate /
BivariateBase classes) when present, else the plain-C restatement pinned
bit-for-bit against it (tests/test_oracle_vs_ref.py); the committed golden
vectors (test_gpu_golden.py, cases wnd*) are checked as well.

Bar: NaN positions identical, values within 1e-12 (float64) / 1e-5 (float32)
relative with the field-scale guard; the fraction of bit-identical values is
asserted too where the operation sequence is the reference's own.
"""
import numpy as np
import pytest

import axes_zoo
from test_gpu_parity import close, _mk_axis, _queries, orc_best  # noqa: F401

pytestmark = pytest.mark.gpu

MODELS = ["akima", "akima_periodic", "c_spline", "c_spline_not_a_knot", "c_spline_periodic", "linear",
          "steffen", "polynomial"]
SPATIAL = ["bicubic", "bilinear"] + MODELS
# models that are linear in the window values and cubic on a cell: on a 6 x 6 window of
# a 2-D float grid with regular ascending axes the tolerance mode computes them with the
# separable stencil (K4s, pyi_bicubics.cuh), values within the tolerance, not bit-identical
STENCIL = ("bicubic", "bilinear", "linear", "c_spline", "c_spline_not_a_knot", "c_spline_periodic")
# akima / akima_periodic / steffen have a 6 x 6 tolerance kernel too (K4m, pyi_winmodels.cuh, 2-D .. 4-D),
# which also serves polynomial and, on 3-D / 4-D grids, the other linear-in-values models
FAST6 = STENCIL + ("akima", "akima_periodic", "steffen", "polynomial")


def _axes(pyi, o, specs):
    pairs = [_mk_axis(pyi, o, s) for s in specs]
    return [p[0] for p in pairs], [p[1] for p in pairs]


def _field(shape, dtype, seed=5, nan=0.0):
    rng = np.random.default_rng(seed)
    g = rng.normal(size=shape) * 40.0
    if np.dtype(dtype).kind == "f":
        g = g.astype(dtype)
        if nan:
            g[rng.random(shape) < nan] = np.nan
    else:
        info = np.iinfo(dtype)
        g = np.clip(np.round(g), info.min, info.max).astype(dtype)
    return g


def _cfg(pyi, nd, model, half, boundary, z_method="linear", u_method="linear"):
    cls = {1: pyi.windowed.Univariate, 2: pyi.windowed.Bivariate, 3: pyi.windowed.Trivariate,
           4: pyi.windowed.Quadrivariate}[nd]
    cfg = getattr(cls, model)().with_boundary_mode(getattr(pyi.windowed.BoundaryConfig, boundary)())
    if nd == 1:
        return cfg.with_half_window_size(half[0])
    cfg = cfg.with_half_window_size_x(half[0]).with_half_window_size_y(half[1])
    if nd >= 3:
        cfg = cfg.with_third_axis(getattr(pyi.windowed.AxisConfig, z_method)())
    if nd >= 4:
        cfg = cfg.with_fourth_axis(getattr(pyi.windowed.AxisConfig, u_method)())
    return cfg


@pytest.mark.parametrize("model", MODELS)
def test_univariate_models(pyi, orc_best, model):
    ref = orc_best
    """core.univariate / univariate_derivative on regular periodic, irregular
    and descending 1-D grids, window sizes 2 .. 10, both boundary modes."""
    fz = axes_zoo.f64_axes()
    specs = {"lon": (np.arange(72) * 5.0, dict(period=360.0), False), "irr": fz["lat_irregular"],
             "desc": fz["lat_desc"]}
    for aname, spec in specs.items():
        (ga,), (ca,) = _axes(pyi, ref, [spec])
        d = ca.describe()
        g = _field((d["size"],), "float64", nan=0.001 if aname == "irr" else 0.0)
        grid = pyi.Grid1D(ga, g)
        rng = np.random.default_rng(11)
        x = rng.uniform(d["min_value"] - 30, d["max_value"] + 30, 20_000)
        x[::37] = np.nan
        for half in (1, 2, 3, 5):
            for boundary in ("undef", "shrink"):
                cfg = _cfg(pyi, 1, model, (half,), boundary)
                for deriv in (False, True):
                    fn = pyi.core.univariate_derivative if deriv else pyi.core.univariate
                    got = fn(grid, x, cfg)
                    want, _ = ref.windowed_nd([ca], g, x, method=model, half_x=half, boundary=boundary,
                                              num_threads=0, derivative=deriv)
                    exact = close(got, want, 50.0, f"{model} {aname} half={half} {boundary} d={deriv}")
                    assert exact == 1.0, (model, aname, half, boundary, deriv, exact)


@pytest.mark.parametrize("model", SPATIAL)
@pytest.mark.parametrize("boundary", ["undef", "shrink"])
def test_bivariate_models(pyi, orc_best, bicubic_mode, model, boundary):
    ref = orc_best
    fz = axes_zoo.f64_axes()
    for dtype, nan, xs, ys in (("float64", 0.0, (np.arange(72) * 5.0, dict(period=360.0), False),
                                (-87.5 + np.arange(36) * 5.0, {}, False)),
                               ("float32", 0.002, fz["lon_180"], fz["lat_irregular"]),
                               ("int16", 0.0, (np.arange(72) * 5.0, {}, False), fz["small_irregular"])):
        (gx, gy), caxes = _axes(pyi, ref, [xs, ys])
        ds = [c.describe() for c in caxes]
        g = _field((ds[0]["size"], ds[1]["size"]), dtype, nan=nan)
        grid = pyi.Grid(gx, gy, g)
        q = _queries(2, 20_000, 21, xr=(ds[0]["min_value"] - 200, ds[0]["max_value"] + 200),
                     yr=(ds[1]["min_value"] - 3, ds[1]["max_value"] + 3))
        for half in ((3, 3), (2, 4), (1, 1), (5, 2)):
            if model == "c_spline_periodic" and (half[0] != half[1] or boundary == "shrink"):
                # The reference's CSplinePeriodic only re-sizes its system when it
                # grows (cspline_periodic.hpp:134-146): after a larger solve (rows
                # vs columns of unequal windows) it solves a stale larger system,
                # history dependent.  The product solves every system afresh
                # (deviations); only equal windows are comparable.
                continue
            cfg = _cfg(pyi, 2, model, half, boundary)
            got = pyi.core.bivariate(grid, q[0], q[1], cfg)
            want, _ = ref.windowed_nd(caxes, g, q[0], q[1], method=model, half_x=half[0], half_y=half[1],
                                      boundary=boundary, num_threads=0)
            exact = close(got, want, 50.0, f"{model} {boundary} {half} {dtype}")
            assert exact == 1.0 or (bicubic_mode == "weights" and model in FAST6 and half == (3, 3)), \
                (model, boundary, half, dtype, exact)


@pytest.mark.parametrize("model", ["bilinear", "linear", "c_spline", "c_spline_not_a_knot", "c_spline_periodic",
                                   "akima", "akima_periodic", "steffen", "polynomial"])
@pytest.mark.parametrize("dtype", ["float64", "float32"])
def test_stencil_models_on_the_quarter_degree_grid(pyi, synthetic, orc_best, bicubic_mode, model, dtype):
    """The fitting models that are linear in the window values on the default 6 x 6 window over the
    0.25 degree grid (circular longitude): the separable stencil with the model's own
    cubics in tolerance mode, the generic kernel bit for bit in exact mode; both
    boundary modes, NaN cells, queries on nodes / outside / non-finite, the
    bounds-error report, host and device memory (bivariate/spline.hpp:96-118)."""
    import torch
    ref = orc_best
    lon, lat = synthetic.lon_axis_values(1440), synthetic.lat_axis_values(720)
    g = synthetic.field(lon, lat).astype(dtype)
    rng = np.random.default_rng(77)
    g[rng.random(g.shape) < 0.0005] = np.nan
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g)
    caxes = [ref.axis(lon, period=360.0), ref.axis(lat)]
    n = 300_000
    x = rng.uniform(-400.0, 800.0, n)
    y = rng.uniform(lat[0] - 1.0, lat[-1] + 1.0, n)
    x[::11] = lon[rng.integers(0, lon.size, x[::11].size)]
    y[1::13] = lat[rng.integers(0, lat.size, y[1::13].size)]
    x[2::997] = np.nan
    y[3::991] = np.inf
    # c_spline_periodic: unshrunken windows only (deviations: the reference solves a
    # stale larger system after a window shrinks)
    periodic = model == "c_spline_periodic"
    for boundary in ("undef",) if periodic else ("undef", "shrink"):
        cfg = _cfg(pyi, 2, model, (3, 3), boundary)
        want, _ = ref.windowed_nd(caxes, g, x, y, method=model, half_x=3, half_y=3, boundary=boundary,
                                  num_threads=0)
        got = pyi.core.bivariate(grid, x, y, cfg)
        exact = close(got, want, 1.0, f"{model} {dtype} {boundary} host")
        assert exact == 1.0 or bicubic_mode == "weights", (model, dtype, boundary, exact)
        dev = pyi.core.bivariate(grid, torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), cfg)
        assert np.array_equal(dev.cpu().numpy(), got, equal_nan=True), (model, dtype, boundary)
        # explicit value modes override the process default
        ex = pyi.core.bivariate(grid, x, y, cfg.with_value_mode("exact"))
        assert close(ex, want, 1.0, f"{model} {dtype} {boundary} exact mode") == 1.0
    inside = np.isfinite(x) & np.isfinite(y) & (y >= lat[0]) & (y <= lat[-1])
    edge = "undef" if periodic else "shrink"
    cfg = _cfg(pyi, 2, model, (3, 3), edge).with_bounds_error(True)
    got = pyi.core.bivariate(grid, x[inside], y[inside], cfg)
    want, _ = ref.windowed_nd(caxes, g, x[inside], y[inside], method=model, half_x=3, half_y=3,
                              boundary=edge, num_threads=0)
    close(got, want, 1.0, f"{model} {dtype} bounds_error on, all inside")
    xs, ys = x[inside][:50_000].copy(), y[inside][:50_000].copy()
    ys[40_000] = lat[-1] + 2.0
    oob = ref.windowed_nd(caxes, g, xs, ys, method=model, half_x=3, half_y=3, boundary=edge,
                          bounds_error=True, num_threads=0)[1]
    assert oob is not None and oob[0] == 40_000
    with pytest.raises(IndexError) as ei:
        pyi.core.bivariate(grid, xs, ys, cfg)
    assert str(ei.value) == oob[2]


@pytest.mark.parametrize("model", ["bicubic", "bilinear", "c_spline", "akima", "steffen", "polynomial"])
@pytest.mark.parametrize("z_kind,z_method", [("monthly30", "linear"), ("monthly30", "nearest"),
                                             ("depth", "linear")])
def test_trivariate_models(pyi, orc_best, bicubic_mode, model, z_kind, z_method):
    ref = orc_best
    fz, iz = axes_zoo.f64_axes(), axes_zoo.i64_axes()
    zs = iz[z_kind] if z_kind in iz else fz[z_kind]
    specs = [(np.arange(72) * 5.0, dict(period=360.0), False), (-87.5 + np.arange(36) * 5.0, {}, False), zs]
    gaxes, caxes = _axes(pyi, ref, specs)
    ds = [c.describe() for c in caxes]
    for dtype, nan in (("float32", 0.001), ("float64", 0.0)):
        g = _field(tuple(d["size"] for d in ds), dtype, nan=nan)
        grid = pyi.Grid(*gaxes, g)
        q = _queries(3, 15_000, 5, z_f64=z_kind not in iz)
        for boundary in ("undef", "shrink"):
            cfg = _cfg(pyi, 3, model, (3, 3), boundary, z_method=z_method)
            z = q[2].view("datetime64[ns]") if grid.has_temporal_axis else q[2]
            got = pyi.core.trivariate(grid, q[0], q[1], z, cfg)
            want, _ = ref.windowed_nd(caxes, g, q[0], q[1], q[2], method=model, boundary=boundary,
                                      num_threads=0, z_method=0 if z_method == "linear" else 1)
            exact = close(got, want, 50.0, f"{model} {z_kind} {z_method} {boundary} {dtype}")
            assert exact == 1.0 or (bicubic_mode == "weights" and model in FAST6), \
                (model, z_kind, z_method, boundary, dtype, exact)


@pytest.mark.parametrize("model", ["bicubic", "bilinear", "c_spline_not_a_knot", "linear", "akima_periodic"])
@pytest.mark.parametrize("methods", [("linear", "linear"), ("nearest", "linear"), ("linear", "nearest")])
def test_quadrivariate_models(pyi, orc_best, model, methods):
    ref = orc_best
    iz = axes_zoo.i64_axes()
    specs = [(np.arange(36) * 10.0, dict(period=360.0), False), (-85.0 + np.arange(18) * 10.0, {}, False),
             iz["monthly30"], (np.round(5500.0 * (np.arange(9) / 8) ** 2.2, 1), {}, False)]
    gaxes, caxes = _axes(pyi, ref, specs)
    ds = [c.describe() for c in caxes]
    for dtype in ("float32", "float64"):
        g = _field(tuple(d["size"] for d in ds), dtype, nan=0.0005)
        grid = pyi.Grid(*gaxes, g)
        q = _queries(4, 10_000, 6)
        cfg = _cfg(pyi, 4, model, (3, 2), "shrink", z_method=methods[0], u_method=methods[1])
        got = pyi.core.quadrivariate(grid, q[0], q[1], q[2].view("datetime64[ns]"), q[3], cfg)
        want, _ = ref.windowed_nd(caxes, g, q[0], q[1], q[2], q[3], method=model, half_x=3, half_y=2,
                                  boundary="shrink", num_threads=0, z_method=0 if methods[0] == "linear" else 1,
                                  u_method=0 if methods[1] == "linear" else 1)
        exact = close(got, want, 50.0, f"{model} {methods} {dtype}")
        assert exact == 1.0, (model, methods, dtype, exact)


@pytest.mark.parametrize("model", ["c_spline", "linear", "akima", "akima_periodic", "steffen", "polynomial"])
def test_quadrivariate_models_default_window(pyi, orc_best, bicubic_mode, model):
    """4-D grids, the default 6 x 6 window: the tolerance pass over the plane-major copy
    (pyi_winmodels.cuh) against the reference, both boundary modes, every z / u method."""
    ref = orc_best
    iz = axes_zoo.i64_axes()
    specs = [(np.arange(36) * 10.0, dict(period=360.0), False), (-85.0 + np.arange(18) * 10.0, {}, False),
             iz["monthly30"], (np.round(5500.0 * (np.arange(9) / 8) ** 2.2, 1), {}, False)]
    gaxes, caxes = _axes(pyi, ref, specs)
    ds = [c.describe() for c in caxes]
    for dtype in ("float32", "float64"):
        g = _field(tuple(d["size"] for d in ds), dtype, nan=0.0005)
        grid = pyi.Grid(*gaxes, g)
        q = _queries(4, 10_000, 6)
        for boundary, methods in (("undef", ("linear", "linear")), ("shrink", ("nearest", "linear")),
                                  ("undef", ("linear", "nearest"))):
            cfg = _cfg(pyi, 4, model, (3, 3), boundary, z_method=methods[0], u_method=methods[1])
            got = pyi.core.quadrivariate(grid, q[0], q[1], q[2].view("datetime64[ns]"), q[3], cfg)
            want, _ = ref.windowed_nd(caxes, g, q[0], q[1], q[2], q[3], method=model, half_x=3, half_y=3,
                                      boundary=boundary, num_threads=0,
                                      z_method=0 if methods[0] == "linear" else 1,
                                      u_method=0 if methods[1] == "linear" else 1)
            exact = close(got, want, 50.0, f"{model} {methods} {boundary} {dtype}")
            assert exact == 1.0 or bicubic_mode == "weights", (model, methods, boundary, dtype, exact)


def test_windowed_nd_bounds_error_and_device_tensors(pyi, orc_best):
    ref = orc_best
    import torch
    iz = axes_zoo.i64_axes()
    specs = [(np.arange(72) * 5.0, {}, False), (-87.5 + np.arange(36) * 5.0, {}, False), iz["monthly30"]]
    gaxes, caxes = _axes(pyi, ref, specs)
    ds = [c.describe() for c in caxes]
    g = _field(tuple(d["size"] for d in ds), "float64")
    grid = pyi.Grid(*gaxes, g)
    q = _queries(3, 4000, 8)
    want = ref.windowed_nd(caxes, g, q[0], q[1], q[2], method="steffen", bounds_error=True, num_threads=3)[1]
    assert want is not None
    cfg = pyi.windowed.Trivariate.steffen().with_bounds_error(True)
    with pytest.raises(IndexError) as ei:
        pyi.core.trivariate(grid, q[0], q[1], q[2].view("datetime64[ns]"), cfg)
    assert str(ei.value) == want[2]
    # z out of bounds only
    x, y = np.full(10, 100.0), np.full(10, 0.0)
    z = np.full(10, ds[2]["min_value"] + 5, dtype=np.int64)
    z[6] = ds[2]["max_value"] + 10
    want = ref.windowed_nd(caxes, g, x, y, z, method="bicubic", bounds_error=True)[1]
    assert want[0] == 6 and want[1] == 2
    with pytest.raises(IndexError) as ei:
        pyi.trivariate(grid, x, y, z.view("datetime64[ns]"), "bicubic", bounds_error=True)
    # the oracle drives math::Axis<int64> (integer repr); the product formats the
    # coordinate like the reference's TemporalAxis (math/temporal_axis.hpp:80-83)
    assert str(ei.value) == ("2021-11-27T00:00:00.000000010 is out of bounds for axis z "
                             "[2021-01-01T00:00:00.000000000, ..., 2021-11-27T00:00:00.000000000]")
    # device tensors == host arrays
    host = pyi.trivariate(grid, q[0], q[1], q[2].view("datetime64[ns]"), "akima", third_axis="nearest")
    dev = pyi.trivariate(grid, torch.from_numpy(q[0]).cuda(), torch.from_numpy(q[1]).cuda(),
                         torch.from_numpy(q[2]).cuda(), "akima", third_axis="nearest")
    # (a small host call on a large tracked grid may run the bit-identical generic kernel on the plain
    # array where the device call runs the 6 x 6 tolerance kernel: equal within the tolerance)
    assert dev.is_cuda
    close(host, dev.cpu().numpy(), 50.0, "host arrays vs device tensors")


def test_windowed_nd_api_errors(pyi):
    lon, lat = np.arange(72) * 5.0, -87.5 + np.arange(36) * 5.0
    grid2 = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), np.zeros((72, 36)))
    grid1 = pyi.Grid1D(pyi.Axis(lon), np.zeros(72, dtype=np.float32))
    assert pyi.univariate(grid1, np.array([12.0])).dtype == np.float32
    with pytest.raises(ValueError, match="univariate requires 1D grid, got 2D"):
        pyi.univariate(grid2, np.zeros(3))
    with pytest.raises(ValueError, match="Unknown windowed method"):
        pyi.univariate(grid1, np.zeros(3), "bicubic2")
    with pytest.raises(TypeError, match="not valid for geometric method"):
        pyi.bivariate(grid2, np.zeros(3), np.zeros(3), "bilinear", half_window_size_x=2)
    with pytest.raises(AttributeError):
        pyi.windowed.Univariate.bicubic()
    out = pyi.bivariate(grid2, np.zeros(0), np.zeros(0), "steffen")
    assert out.shape == (0,) and out.dtype == np.float64


UNITS = ["Y", "M", "W", "D", "h", "m", "s", "ms", "us", "ns"]


def test_datetime_resolution_conversion(pyi, orc_best):
    ref = orc_best
    """dateutils::convert on the device (pyi_datetime_convert) against the
    reference's own function, every pair of numpy units, negative and large
    values, host arrays and device tensors."""
    import torch
    rng = np.random.default_rng(3)
    overflowed = 0
    for src in UNITS:
        span = {"Y": 3000, "M": 40_000, "W": 200_000, "D": 1_500_000, "h": 30_000_000, "m": 2_000_000_000}.get(
            src, 2 ** 45)
        v = rng.integers(-span, span, 20_000, dtype=np.int64)
        v[:8] = [0, -1, 1, -span, span, 59, -61, 86_399]
        for dst in UNITS:
            try:
                want = ref.datetime_convert(v, src, dst)
            except OverflowError as e:  # e.g. +-3000 years in nanoseconds
                with pytest.raises(OverflowError, match=str(e)):
                    pyi.core.convert_resolution(v, src, dst)
                overflowed += 1
                continue
            got = pyi.core.convert_resolution(v, src, dst)
            assert got.dtype == np.int64 and np.array_equal(got, want), (src, dst)
            dev = pyi.core.convert_resolution(torch.from_numpy(v).cuda(), src, dst)
            assert dev.is_cuda and np.array_equal(dev.cpu().numpy(), want), (src, dst)
    assert 0 < overflowed < 30
    # numpy agrees wherever it defines the conversion without ambiguity (finer units)
    d = rng.integers(-100_000, 100_000, 1000).astype("datetime64[D]")
    assert np.array_equal(pyi.core.convert_resolution(d.view(np.int64), "D", "ns"),
                          d.astype("datetime64[ns]").view(np.int64))


def test_datetime_conversion_overflow(pyi, orc_best):
    ref = orc_best
    v = np.array([5, 2 ** 40, -3, 2 ** 62], dtype=np.int64)
    with pytest.raises(OverflowError) as want:
        ref.datetime_convert(v, "s", "ns")
    with pytest.raises(OverflowError) as got:
        pyi.core.convert_resolution(v, "s", "ns")
    assert str(want.value) in str(got.value)
    # NaT is converted as the integer it is (no special case, dateutils.hpp:845-933)
    nat = np.array([np.iinfo(np.int64).min], dtype=np.int64)
    assert np.array_equal(pyi.core.convert_resolution(nat, "ns", "s"), ref.datetime_convert(nat, "ns", "s"))
    with pytest.raises(OverflowError):
        pyi.core.convert_resolution(nat, "s", "ns")


def test_datetime_conversion_extremes(pyi, orc_best):
    """The closed form of the fixed-length units (value * F / floor(value / G), pyi_datetime.cu) hands
    over to the checked sequence where an intermediate of the reference could overflow: values around
    that hand-over, around the true overflow limit and at the ends of int64, one by one against the
    reference (its result, or its overflow error)."""
    ref = orc_best
    fixed = ["D", "h", "m", "s", "ms", "us", "ns", "ps"]
    length = dict(zip(fixed, [86400 * 10 ** 12, 3600 * 10 ** 12, 60 * 10 ** 12, 10 ** 12, 10 ** 9, 10 ** 6, 10 ** 3, 1]))
    top, low = 2 ** 63 - 1, -2 ** 63
    rng = np.random.default_rng(8)
    checked = raised = 0
    for src in fixed:
        for dst in fixed + ["W", "fs"]:
            if src == dst:
                continue
            f = max(1, length[src] // length.get(dst, 1)) if dst in length else 1
            mag = 10 ** 12 // length[dst] if dst in length and length[dst] <= 10 ** 12 else 1
            edges = {0, 1, -1, top, low, low + 1, 2 ** 62, -2 ** 62, top // f, -(top // f)}
            for centre in (top // f, (top - 3 * 86400 * mag) // f, top - 3 * 86400 * mag):
                for d in (-2, -1, 0, 1, 2, 86400 * mag, -86400 * mag):
                    edges.update({centre + d, -(centre + d)})
            edges.update(int(v) for v in rng.integers(low, top, 12, dtype=np.int64))
            vals = np.array(sorted(v for v in edges if low <= v <= top), dtype=np.int64)
            want = np.zeros_like(vals)
            bad = np.zeros(vals.size, dtype=bool)
            for k in range(vals.size):
                try:
                    want[k] = ref.datetime_convert(vals[k:k + 1], src, dst)[0]
                except OverflowError:
                    bad[k] = True
            good = vals[~bad]
            if good.size:
                got = pyi.core.convert_resolution(good, src, dst)
                assert np.array_equal(got, want[~bad]), (src, dst, good[got != want[~bad]][:4])
                checked += good.size
            if bad.any():  # the whole array: same error as the reference (first failing value)
                with pytest.raises(OverflowError) as e_ref:
                    ref.datetime_convert(vals, src, dst)
                with pytest.raises(OverflowError) as e_got:
                    pyi.core.convert_resolution(vals, src, dst)
                assert str(e_ref.value) in str(e_got.value), (src, dst)
                for v in vals[bad][:6]:
                    with pytest.raises(OverflowError):
                        pyi.core.convert_resolution(np.array([v], dtype=np.int64), src, dst)
                raised += 1
    assert checked > 1500 and raised > 20


def test_temporal_axis_casts_on_device(pyi, orc_best):
    ref = orc_best
    """trivariate with datetime64[s] / [D] coordinates on a datetime64[ns]
    axis: the pre-step of grid_dispatch.hpp:190-195 runs on the GPU."""
    t = np.datetime64("2021-01-01", "ns") + np.arange(20) * np.timedelta64(1, "D")
    lon, lat = np.arange(72) * 5.0, -87.5 + np.arange(36) * 5.0
    g = _field((72, 36, 20), "float64")
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), pyi.TemporalAxis(t), g)
    rng = np.random.default_rng(1)
    x, y = rng.uniform(0, 360, 5000), rng.uniform(-80, 80, 5000)
    zs = (np.datetime64("2021-01-01", "s") + rng.integers(0, 19 * 86400, 5000).astype("timedelta64[s]"))
    got = pyi.trivariate(grid, x, y, zs)
    caxes = [ref.axis(lon, period=360.0), ref.axis(lat), ref.axis(t.astype(np.int64))]
    want, _ = ref.geometric(caxes, g, x, y, ref.datetime_convert(zs.view(np.int64), "s", "ns"), num_threads=0)
    assert np.array_equal(got, want, equal_nan=True) and np.isfinite(got).all()
    assert np.array_equal(grid.z.cast_to_int64(zs), zs.astype("datetime64[ns]").view(np.int64))
    with pytest.raises(ValueError, match="must be of type datetime64"):
        grid.z.cast_to_int64(np.arange(3).astype("timedelta64[s]"))


# ---- bicubic fast path on 3-D / 4-D grids (pyi_fastwin_nd.cuh) -----------------
def _generic(fn):
    """Runs fn with the windowed fast paths disabled (generic K6 kernel)."""
    import os
    os.environ["PYINTERP_B200_GENERIC"] = "1"
    try:
        return fn()
    finally:
        os.environ.pop("PYINTERP_B200_GENERIC", None)


@pytest.mark.parametrize("dtype,nan", [("float64", 0.0), ("float64", 0.002), ("float32", 0.001), ("int16", 0.0),
                                       ("uint64", 0.0)])
@pytest.mark.parametrize("y_kind", ["regular", "irregular"])
def test_bicubic_nd_fast_path_trivariate(pyi, orc_best, bicubic_mode, dtype, nan, y_kind):
    """Fast pass + redo pass == generic kernel == reference, bit for bit: regular
    axes (uniform spline constants) and an irregular latitude, NaN cells, integer
    grids, edges, out-of-range and NaN coordinates, temporal z, both z methods."""
    ref = orc_best
    iz = axes_zoo.i64_axes()
    lat = -87.5 + np.arange(36) * 5.0
    if y_kind == "irregular":
        lat = lat + 1.7 * np.sin(np.arange(36) * 0.9)
    specs = [(np.arange(72) * 5.0, dict(period=360.0), False), (lat, {}, False), iz["monthly30"]]
    gaxes, caxes = _axes(pyi, ref, specs)
    ds = [c.describe() for c in caxes]
    g = _field(tuple(d["size"] for d in ds), dtype, seed=31, nan=nan)
    grid = pyi.Grid(*gaxes, g)
    q = _queries(3, 60_000, 33)
    z = q[2].view("datetime64[ns]")
    for z_method in ("linear", "nearest"):
        for boundary in ("undef", "shrink"):
            cfg = _cfg(pyi, 3, "bicubic", (3, 3), boundary, z_method=z_method)
            got = pyi.core.trivariate(grid, q[0], q[1], z, cfg)
            gen = _generic(lambda: pyi.core.trivariate(grid, q[0], q[1], z, cfg))
            want, _ = ref.windowed_nd(caxes, g, q[0], q[1], q[2], method="bicubic", boundary=boundary,
                                      num_threads=0, z_method=0 if z_method == "linear" else 1)
            assert np.array_equal(gen, want, equal_nan=True)
            exact = close(got, want, 50.0, f"{dtype} {y_kind} {z_method} {boundary}")
            assert exact == 1.0 or bicubic_mode == "weights", (z_method, boundary, exact)
    assert np.isfinite(got).mean() > 0.3  # the fast pass did most of the work


@pytest.mark.parametrize("dtype", ["float64", "float32", "int32"])
def test_bicubic_nd_fast_path_quadrivariate(pyi, orc_best, bicubic_mode, dtype):
    ref = orc_best
    iz = axes_zoo.i64_axes()
    specs = [(np.arange(36) * 10.0, dict(period=360.0), False), (-85.0 + np.arange(18) * 10.0, {}, False),
             iz["monthly30"], (np.round(5500.0 * (np.arange(9) / 8) ** 2.2, 1), {}, False)]
    gaxes, caxes = _axes(pyi, ref, specs)
    ds = [c.describe() for c in caxes]
    g = _field(tuple(d["size"] for d in ds), dtype, seed=41, nan=0.0005)
    grid = pyi.Grid(*gaxes, g)
    q = _queries(4, 40_000, 43)
    z = q[2].view("datetime64[ns]")
    for methods in (("linear", "linear"), ("nearest", "linear"), ("linear", "nearest")):
        cfg = _cfg(pyi, 4, "bicubic", (3, 3), "undef", z_method=methods[0], u_method=methods[1])
        got = pyi.core.quadrivariate(grid, q[0], q[1], z, q[3], cfg)
        gen = _generic(lambda: pyi.core.quadrivariate(grid, q[0], q[1], z, q[3], cfg))
        want, _ = ref.windowed_nd(caxes, g, q[0], q[1], q[2], q[3], method="bicubic", boundary="undef",
                                  num_threads=0, z_method=0 if methods[0] == "linear" else 1,
                                  u_method=0 if methods[1] == "linear" else 1)
        assert np.array_equal(gen, want, equal_nan=True)
        exact = close(got, want, 50.0, f"{dtype} {methods}")
        assert exact == 1.0 or bicubic_mode == "weights", (methods, exact)


def test_bicubic_nd_fast_path_refresh_bounds_and_device(pyi, orc_best, bicubic_mode):
    """The plane-major copy follows the caller's array (non-resident grids are
    refreshed before every call); bounds errors come from the redo pass; device
    tensors give the same bits as host arrays; float64 depth axis."""
    import torch
    ref = orc_best
    fz = axes_zoo.f64_axes()
    specs = [(np.arange(72) * 5.0, dict(period=360.0), False), (-87.5 + np.arange(36) * 5.0, {}, False), fz["depth"]]
    gaxes, caxes = _axes(pyi, ref, specs)
    ds = [c.describe() for c in caxes]
    g = _field(tuple(d["size"] for d in ds), "float64", seed=51)
    grid = pyi.Grid(*gaxes, g)
    q = _queries(3, 30_000, 53, z_f64=True)
    first = pyi.trivariate(grid, q[0], q[1], q[2], "bicubic")
    g *= -2.0  # the reference borrows the buffer: the next call sees the new values
    second = pyi.trivariate(grid, q[0], q[1], q[2], "bicubic")
    want, _ = ref.windowed_nd(caxes, g, q[0], q[1], q[2], method="bicubic", num_threads=0)
    exact = close(second, want, 50.0, "refreshed")
    assert exact == 1.0 or bicubic_mode == "weights"
    assert not np.array_equal(first, second, equal_nan=True)
    dev = pyi.trivariate(grid, *(torch.from_numpy(c).cuda() for c in q), "bicubic")
    assert np.array_equal(dev.cpu().numpy(), second, equal_nan=True)
    want_err = ref.windowed_nd(caxes, g, q[0], q[1], q[2], method="bicubic", bounds_error=True, num_threads=0)[1]
    with pytest.raises(IndexError) as ei:
        pyi.trivariate(grid, q[0], q[1], q[2], "bicubic", bounds_error=True)
    assert str(ei.value) == want_err[2]
