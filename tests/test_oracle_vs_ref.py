"""Pins oracle/liboracle.so (plain-C restatement) against the reference's own
headers compiled in place (oracle/_ref/liboracle_ref.so): bit for bit."""
import numpy as np

import golden_util as gu
import pytest

import axes_zoo


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    assert a.dtype == b.dtype and a.shape == b.shape
    assert np.array_equal(a.view(np.uint8), b.view(np.uint8)) or np.array_equal(a, b, equal_nan=True)


def _desc_equal(d0, d1):
    assert d0.keys() == d1.keys()
    for k in d0:
        if k == "points":
            assert np.array_equal(d0[k], d1[k])
        else:
            assert d0[k] == d1[k] or (d0[k] != d0[k] and d1[k] != d1[k]), k


def _build(o, spec):
    values, kw, flip = spec
    a = o.axis(values, **kw)
    return a.flip() if flip else a


@pytest.mark.parametrize("name", sorted(axes_zoo.f64_axes()))
def test_axis_f64(port, ref, name):
    spec = axes_zoo.f64_axes()[name]
    a, b = _build(port, spec), _build(ref, spec)
    _desc_equal(a.describe(), b.describe())
    c = axes_zoo.coords_for(b.describe(), 20000, 11)
    assert np.array_equal(a.find_indexes(c), b.find_indexes(c))
    for bounded in (False, True):
        assert np.array_equal(a.find_index(c, bounded), b.find_index(c, bounded))
    for half in (1, 2, 3, 5):
        for boundary in ("undef", "shrink", "expand", "wrap", "sym"):
            for cc in c[:: max(1, c.size // 300)]:
                assert a.find_window(cc, half, boundary) == b.find_window(cc, half, boundary)


@pytest.mark.parametrize("name", sorted(axes_zoo.i64_axes()))
def test_axis_i64(port, ref, name):
    spec = axes_zoo.i64_axes()[name]
    a, b = _build(port, spec), _build(ref, spec)
    _desc_equal(a.describe(), b.describe())
    c = axes_zoo.coords_for(b.describe(), 20000, 12)
    assert np.array_equal(a.find_indexes(c), b.find_indexes(c))
    for bounded in (False, True):
        assert np.array_equal(a.find_index(c, bounded), b.find_index(c, bounded))
    for half in (1, 3):
        for boundary in ("undef", "shrink"):
            for cc in c[:: max(1, c.size // 200)]:
                assert a.find_window(cc, half, boundary) == b.find_window(cc, half, boundary)


def test_axis_errors(port, ref):
    for o in (port, ref):
        with pytest.raises(ValueError, match="monotonic"):
            o.axis([0.0, 2.0, 1.0, 5.0])
        with pytest.raises(ValueError, match="monotonic"):
            o.axis([0.0, 1.0, 1.0, 5.0])
        with pytest.raises(ValueError):
            o.axis(np.array([], dtype=np.float64))


def _grid_axes(o, nd, x_kind="lon_periodic", y_kind="lat", flip_z=False):
    fz, iz = axes_zoo.f64_axes(), axes_zoo.i64_axes()
    lon = np.arange(72) * 5.0
    lat = -87.5 + np.arange(36) * 5.0
    axes = [o.axis(lon, period=360.0) if x_kind == "lon_periodic" else _build(o, fz[x_kind]),
            o.axis(lat) if y_kind == "lat" else _build(o, fz[y_kind])]
    if nd >= 3:
        axes.append(_build(o, iz["monthly30"]))
        if flip_z:
            axes[-1].flip()
    if nd >= 4:
        axes.append(o.axis(np.round(5500.0 * (np.arange(9) / 8) ** 2.2, 1)))
    return axes


def _field(shape, dtype, seed=3, nan=0.0):
    rng = np.random.default_rng(seed)
    g = rng.normal(size=shape) * 50.0
    if np.dtype(dtype).kind == "f":
        g = g.astype(dtype)
        if nan:
            g[rng.random(shape) < nan] = np.nan
        return g
    info = np.iinfo(dtype)
    return np.clip(np.round(g), info.min, info.max).astype(dtype)


def _queries(nd, n, seed):
    rng = np.random.default_rng(seed)
    q = dict(x=rng.uniform(-400, 800, n), y=rng.uniform(-95, 95, n))
    q["x"][::97] = np.nan
    q["x"][5::101] = 5.0 * rng.integers(0, 72, q["x"][5::101].size)  # exact nodes
    q["y"][7::89] = -87.5 + 5.0 * rng.integers(0, 36, q["y"][7::89].size)
    if nd >= 3:
        q["z"] = rng.integers(axes_zoo.T0 - 10 * axes_zoo.DAY, axes_zoo.T0 + 340 * axes_zoo.DAY, n, dtype=np.int64)
        q["z"][::113] = np.iinfo(np.int64).min
        q["z"][3::57] = axes_zoo.T0 + 30 * axes_zoo.DAY * rng.integers(0, 12, q["z"][3::57].size)
    if nd >= 4:
        q["u"] = rng.uniform(-50, 5600, n)
    return q


@pytest.mark.parametrize("nd", [2, 3, 4])
@pytest.mark.parametrize("dtype", ["float64", "float32", "int16", "int64", "uint8", "uint64", "int32"])
@pytest.mark.parametrize("method", [0, 1, 2])
def test_geometric(port, ref, nd, dtype, method):
    n = 4000
    q = _queries(nd, n, 100 + nd)
    shape = [72, 36, 12, 9][:nd]
    g = _field(shape, dtype, nan=0.02)
    res = []
    for o in (port, ref):
        axes = _grid_axes(o, nd)
        out, oob = o.geometric(axes, g, q["x"], q["y"], q.get("z"), q.get("u"), method=method,
                               z_method=method == 2, u_method=method == 2, num_threads=3)
        assert oob is None
        res.append(out)
    assert res[0].dtype == res[1].dtype
    assert np.array_equal(res[0], res[1], equal_nan=True)
    assert np.isfinite(res[0]).sum() > n // 8


@pytest.mark.parametrize("nd", [2, 3, 4])
def test_geometric_mixed_axis_methods_and_axes(port, ref, nd):
    q = _queries(nd, 3000, 7)
    shape = [72, 36, 12, 9][:nd]
    g = _field(shape, "float32")
    for x_kind, y_kind, flip_z in [("lon_periodic", "lat", True), ("lon_partial_period", "small_irregular_desc", False)]:
        res = []
        for o in (port, ref):
            axes = _grid_axes(o, nd, x_kind, y_kind, flip_z)
            shp = [a.describe()["size"] for a in axes]
            gg = _field(shp, "float32")
            qq = dict(q)
            if y_kind != "lat":
                qq["y"] = np.random.default_rng(1).uniform(-5, 55, 3000)
                qq["x"] = np.random.default_rng(2).uniform(-400, 500, 3000)
            out, oob = o.geometric(axes, gg, qq["x"], qq["y"], qq.get("z"), qq.get("u"), method=0,
                                   z_method=1, u_method=0, num_threads=2)
            res.append(out)
        assert np.array_equal(res[0], res[1], equal_nan=True)
        assert np.isfinite(res[0]).any()


@pytest.mark.parametrize("nd", [2, 3, 4])
def test_geometric_bounds_error(port, ref, nd):
    q = _queries(nd, 5000, 9)
    shape = [72, 36, 12, 9][:nd]
    g = _field(shape, "float64")
    res = []
    for o in (port, ref):
        axes = _grid_axes(o, nd)
        for threads in (1, 4):
            out, oob = o.geometric(axes, g, q["x"], q["y"], q.get("z"), q.get("u"), bounds_error=True,
                                   num_threads=threads)
            assert oob is not None
            res.append(oob[:2])
            if o is ref and oob[1] != 2:
                pmsg = port.geometric(_grid_axes(port, nd), g, q["x"], q["y"], q.get("z"), q.get("u"),
                                      bounds_error=True, num_threads=threads)[1][2]
                assert pmsg == oob[2]
    assert len(set(res)) == 1


@pytest.mark.parametrize("method", [0, 1])
@pytest.mark.parametrize("boundary", ["undef", "shrink"])
@pytest.mark.parametrize("half", [(3, 3), (2, 4), (1, 1), (5, 2)])
def test_windowed2d(port, ref, method, boundary, half):
    n = 3000
    q = _queries(2, n, 21)
    for dtype, nan in (("float64", 0.0), ("float32", 0.002), ("int16", 0.0)):
        g = _field([72, 36], dtype, nan=nan)
        res = []
        for o in (port, ref):
            axes = _grid_axes(o, 2)
            out, oob = o.windowed2d(axes, g, q["x"], q["y"], method=method, half_x=half[0], half_y=half[1],
                                    boundary=boundary, num_threads=2, stateless=True)
            assert oob is None
            res.append(out)
        assert np.array_equal(res[0], res[1], equal_nan=True)
        if method == 0 and min(half) < 2:
            # natural cubic spline needs >= 4 nodes (cspline.hpp:29): all NaN
            assert not np.isfinite(res[0]).any()
        else:
            assert np.isfinite(res[0]).sum() > n // 10


def test_windowed2d_stateful_and_axes(port, ref):
    """The reference's per-thread cache (history dependent) is restated too."""
    rng = np.random.default_rng(5)
    x = np.cumsum(rng.uniform(-0.5, 1.5, 4000)) % 720 - 180
    y = np.clip(np.cumsum(rng.uniform(-1, 1, 4000)), -95, 95)
    g = _field([72, 36], "float64", nan=0.003)
    for x_kind, y_kind in [("lon_periodic", "lat"), ("lon_180", "lat_desc"), ("lon_partial_period", "lat_irregular")]:
        res = []
        for o in (port, ref):
            axes = _grid_axes(o, 2, x_kind, y_kind)
            shp = [a.describe()["size"] for a in axes]
            gg = _field(shp, "float64", nan=0.001)
            for stateless in (True, False):
                out, _ = o.windowed2d(axes, gg, x, y, num_threads=1, stateless=stateless)
                res.append(out)
        assert np.array_equal(res[0], res[2], equal_nan=True)
        assert np.array_equal(res[1], res[3], equal_nan=True)


def test_windowed2d_bounds_error(port, ref):
    q = _queries(2, 2000, 33)
    g = _field([72, 36], "float64")
    res = []
    for o in (port, ref):
        axes = _grid_axes(o, 2)
        out, oob = o.windowed2d(axes, g, q["x"], q["y"], bounds_error=True, num_threads=3)
        assert oob is not None
        res.append(oob)
    assert res[0] == res[1]
    # a window that does not fit is NOT an error (cache_loader.hpp:69-80)
    for o in (port, ref):
        axes = _grid_axes(o, 2)
        out, oob = o.windowed2d(axes, g, np.array([10.0]), np.array([-86.0]), bounds_error=True)
        assert oob is None and np.isnan(out[0])


def test_scalar_probes(port, ref):
    rng = np.random.default_rng(0)
    L, R = port.lib, ref.lib
    for _ in range(300):
        a = rng.uniform(-10, 10, 10)
        for m in (0, 1, 2):
            for cart in (0, 1):
                args = (m, 2 + (_ % 3), cart, *a)
                assert L.orc_eval2d_f64(*args) == R.orc_eval2d_f64(*args) or np.isnan(L.orc_eval2d_f64(*args))
                assert L.orc_eval2d_f32(*args) == R.orc_eval2d_f32(*args) or np.isnan(L.orc_eval2d_f32(*args))
        zi = [int(v) for v in rng.integers(-10**18, 10**18, 3)]
        for m in (0, 1):
            assert L.orc_axis_interp_i64(m, *zi, a[0], a[1]) == R.orc_axis_interp_i64(m, *zi, a[0], a[1])
            assert L.orc_axis_interp_i64_f32(m, *zi, a[0], a[1]) == R.orc_axis_interp_i64_f32(m, *zi, a[0], a[1])
            assert L.orc_axis_interp_f64(m, *a[:5]) == R.orc_axis_interp_f64(m, *a[:5])
    xa = np.sort(rng.uniform(0, 10, 9))
    ya = rng.normal(size=9)
    for x in np.r_[xa, rng.uniform(-1, 11, 50)]:
        for d in (False, True):
            p, r = port.cspline(xa, ya, x, d), ref.cspline(xa, ya, x, d)
            assert p == r or (np.isnan(p) and np.isnan(r))


# ---- the widened windowed path and dateutils::convert ----
WND_MODELS = ["akima", "akima_periodic", "c_spline", "c_spline_not_a_knot", "c_spline_periodic", "linear",
              "steffen", "polynomial"]


@pytest.mark.parametrize("model", WND_MODELS)
def test_univariate_models_port_vs_ref(port, ref, model):
    """Every model on explicit nodes, 2 .. 16 of them, value and derivative
    (covers the Eigen packet traversals restated for polynomial / periodic)."""
    rng = np.random.default_rng(21)
    for n in range(2, 17):
        for trial in range(30):
            xa = np.sort(rng.uniform(-5, 5, n))
            if trial % 7 == 3 and n > 3:
                xa[2] = xa[1]  # duplicate abscissa
            ya = rng.normal(size=n) * 10
            for x in (rng.uniform(xa[0], xa[-1]), xa[n // 2], xa[0], xa[-1], xa[-1] + 1.0):
                for der in (False, True):
                    a, b = port.univariate(model, xa, ya, x, der), ref.univariate(model, xa, ya, x, der)
                    assert (a == b) or (a != a and b != b), (model, n, trial, x, der, a, b)


@pytest.mark.parametrize("nd", [1, 2, 3, 4])
@pytest.mark.parametrize("boundary", ["undef", "shrink"])
def test_windowed_nd_port_vs_ref(port, ref, nd, boundary):
    rng = np.random.default_rng(nd)
    T0, DAY = axes_zoo.T0, axes_zoo.DAY
    specs = [(np.arange(36) * 10.0, dict(period=360.0)), (np.degrees(np.arcsin(np.linspace(-0.99, 0.99, 20))), {}),
             (T0 + np.arange(6, dtype=np.int64) * DAY, {}), (np.array([0.0, 5.0, 20.0, 100.0]), {})][:nd]
    shape = tuple(len(s[0]) for s in specs)
    n = 1500
    q = [rng.uniform(-200, 500, n), rng.uniform(-95, 95, n),
         rng.integers(T0 - DAY, T0 + 6 * DAY, n, dtype=np.int64), rng.uniform(-5, 105, n)][:nd]
    q[0][::50] = np.nan
    models = WND_MODELS if nd == 1 else ["bicubic", "bilinear"] + WND_MODELS
    for dtype in ("float64", "float32", "int16"):
        g = (rng.normal(size=shape) * 30).astype(dtype)
        if dtype == "float32":
            g[rng.random(shape) < 0.003] = np.nan
        for model in models:
            for half in ((3, 3), (2, 4)):
                if model == "c_spline_periodic" and nd >= 2 and (half[0] != half[1] or boundary == "shrink"):
                    continue  # reference solves a stale larger system (deviation)
                kw = dict(method=model, half_x=half[0], half_y=half[1], boundary=boundary, z_method=nd % 2,
                          u_method=(nd + 1) % 2, derivative=(nd == 1 and half[0] == 2))
                outs = []
                for o in (port, ref):
                    axes = [o.axis(v, **k) for v, k in specs]
                    outs.append(o.windowed_nd(axes, g, *q, num_threads=2, **kw)[0])
                _same(outs[0], outs[1])
    # bounds error: same point, axis and message
    errs = []
    for o in (port, ref):
        axes = [o.axis(v, **k) for v, k in specs]
        errs.append(o.windowed_nd(axes, g, *q, method=models[-1], bounds_error=True, num_threads=3)[1])
    if nd >= 2:
        assert errs[0] is not None and errs[0] == errs[1]


def test_datetime_convert_port_vs_ref(port, ref):
    units = ["Y", "M", "W", "D", "h", "m", "s", "ms", "us", "ns", "ps"]
    rng = np.random.default_rng(2)
    for src in units:
        span = {"Y": 2000, "M": 20_000, "W": 100_000, "D": 700_000}.get(src, 2 ** 40)
        v = rng.integers(-span, span, 5000, dtype=np.int64)
        v[:4] = [0, -1, 1, np.iinfo(np.int64).min if src in ("ns", "ps") else 7]
        for dst in units:
            res = []
            for o in (port, ref):
                try:
                    res.append(o.datetime_convert(v, src, dst))
                except OverflowError as e:
                    res.append(str(e))
            if isinstance(res[1], str):
                assert res[0] == res[1], (src, dst)
            else:
                assert np.array_equal(res[0], res[1]), (src, dst)


@pytest.mark.parametrize("dtype", ["float64", "float32"])
def test_fill_gauss_seidel_port_vs_ref(port, ref, dtype):
    """fill::gauss_seidel: filled grid, iteration count and residual of the C
    restatement equal the reference build's, serial and threaded (the red /
    black sweeps are order independent; the port sweeps black first like the
    g++ build of the reference)."""
    for k, shape in enumerate([(40, 30), (37, 23), (1, 9), (9, 1), (2, 2), (3, 3), (5, 64), (64, 5), (1, 1)]):
        g0 = gu.fill_field(shape, dtype, 100 + k, blobs=2)
        for fg in ("zonal_average", "zero"):
            for per in (True, False):
                for relax, its, eps in ((1.0, 60, 1e-4), (1.5, 7, 0.0)):
                    a, b = g0.copy(), g0.copy()
                    ra = ref.fill_gauss_seidel(a, fg, per, its, eps, relax, num_threads=3)
                    rb = port.fill_gauss_seidel(b, fg, per, its, eps, relax)
                    assert ra == rb, (shape, fg, per, relax)
                    assert np.array_equal(a, b, equal_nan=True), (shape, fg, per, relax)
    # no NaN: nothing happens (gauss_seidel.hpp:175-177); all NaN: zeros
    for o in (port, ref):
        g = np.arange(12, dtype=dtype).reshape(3, 4)
        assert o.fill_gauss_seidel(g.copy()) == (0, 0.0)
        g = np.full((6, 5), np.nan, dtype=dtype)
        its, res = o.fill_gauss_seidel(g, max_iterations=10)
        assert (its, res) == (1, 0.0) and not g.any()


@pytest.mark.parametrize("dtype", ["float64", "float32"])
def test_fill_multigrid_port_vs_ref(port, ref, dtype):
    """fill::multigrid: V-cycles of the C restatement equal the reference
    build's bit for bit on even / odd / degenerate sizes."""
    for k, shape in enumerate([(40, 30), (37, 23), (77, 41), (9, 9), (5, 5), (4, 50), (64, 5), (6, 7), (1, 1)]):
        g0 = gu.fill_field(shape, dtype, 300 + k, blobs=2)
        for fg, per, pre, post, its, eps in (("zonal_average", True, 2, 2, 8, 1e-6), ("zero", False, 1, 3, 3, 0.0),
                                             ("zero", True, 0, 1, 2, 0.0)):
            a, b = g0.copy(), g0.copy()
            ra = ref.fill_multigrid(a, fg, per, its, eps, pre, post, num_threads=2)
            rb = port.fill_multigrid(b, fg, per, its, eps, pre, post)
            assert ra == rb, (shape, fg, per)
            assert np.array_equal(a, b, equal_nan=True), (shape, fg, per)
    for o in (port, ref):
        g = np.arange(12, dtype=dtype).reshape(3, 4)
        assert o.fill_multigrid(g.copy()) == (0, 0.0)


@pytest.mark.parametrize("dtype", ["float64", "float32"])
def test_fill_loess_port_vs_ref(port, ref, dtype):
    """fill::loess (fill/loess.hpp:207-253): the C restatement's processed grid
    equals the reference build's bit for bit -- single pass and iterated, the
    three value types, periodic or mirrored x, both first guesses (the zonal
    average is a SEQUENTIAL sum in storage order in the reference build),
    windows larger than the grid, threaded or not."""
    for k, shape in enumerate([(40, 30), (37, 23), (1, 9), (9, 1), (2, 2), (5, 64), (64, 5), (1, 1), (130, 67)]):
        g0 = gu.fill_field(shape, dtype, 500 + k, blobs=2)
        for vt in ("undefined", "defined", "all"):
            for fg, per, its, eps, hx, hy in (("zonal_average", True, 1, 1e-4, 3, 3), ("zero", False, 4, 0.0, 2, 5),
                                              ("zonal_average", False, 30, 1e-3, 4, 1), ("zonal_average", True, 0, 1e-4, 3, 3)):
                a = ref.fill_loess(g0, hx, hy, vt, fg, per, its, eps, num_threads=3)
                b = port.fill_loess(g0, hx, hy, vt, fg, per, its, eps)
                assert np.array_equal(a, b, equal_nan=True), (shape, vt, fg, per, its)
    for o in (port, ref):
        g = np.full((6, 5), np.nan, dtype=dtype)
        assert np.isnan(o.fill_loess(g, max_iterations=1)).all()      # nothing to average: cells keep their value
        assert not o.fill_loess(g, max_iterations=3).any()            # iterative: the first guess (0) is kept
