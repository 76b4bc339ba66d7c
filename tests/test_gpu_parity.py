"""GPU parity: the CUDA path, called through the Python mirror of the reference
API and the C-ABI, against the CPU oracle on identical seeded inputs.

Bars (BASELINE.json north_star): index selection bit-exact; values within
1e-12 relative (float64 results) / 1e-5 (float32 results); identical NaN
positions; identical bounds-error index / axis / message.
Relative error is measured as |a-b| <= tol * max(|a|, |b|, field_scale)
(plain relative error is ill-posed at zero crossings).
"""
import numpy as np
import pytest

import axes_zoo

pytestmark = pytest.mark.gpu

TOL = {np.dtype("float64"): 1e-12, np.dtype("float32"): 1e-5}


def close(a, b, scale, what=""):
    assert a.dtype == b.dtype, (a.dtype, b.dtype)
    assert a.shape == b.shape
    na, nb = np.isnan(a), np.isnan(b)
    assert np.array_equal(na, nb), f"{what}: NaN positions differ at {np.flatnonzero(na != nb)[:8]}"
    tol = TOL[a.dtype]
    a64, b64 = a[~na].astype(np.float64), b[~na].astype(np.float64)
    err = np.abs(a64 - b64)
    bound = tol * np.maximum(np.maximum(np.abs(a64), np.abs(b64)), scale)
    bad = np.flatnonzero(err > bound)
    assert bad.size == 0, f"{what}: {bad.size} values off, worst {err.max():.3e} (tol {tol})"
    return float((err == 0).mean()) if err.size else 1.0


@pytest.fixture(scope="module")
def orc_best():
    """The reference build when present, else the pinned C restatement."""
    import orc
    return orc.Oracle("reference") if orc.available("reference") else orc.Oracle("port")


def _mk_axis(pyi, o, spec):
    values, kw, flip = spec
    if np.asarray(values).dtype.kind == "i":
        dt = np.asarray(values, dtype=np.int64).view("datetime64[ns]")
        per = kw.get("period")
        g = pyi.TemporalAxis(dt, period=None if per is None else np.timedelta64(per, "ns"))
    else:
        g = pyi.Axis(values, **kw)
    c = o.axis(values, **kw)
    if flip:
        g.flip(inplace=True)
        c.flip()
    return g, c


@pytest.mark.parametrize("name", sorted(axes_zoo.f64_axes()) + sorted(axes_zoo.i64_axes()))
def test_axis_lookup_bit_exact(pyi, orc_best, name):
    zoo = {**axes_zoo.f64_axes(), **axes_zoo.i64_axes()}
    g, c = _mk_axis(pyi, orc_best, zoo[name])
    d = c.describe()
    gd = g._desc()
    assert (gd.kind, gd.size, gd.is_ascending, gd.is_periodic) == (d["kind"], d["size"], d["is_ascending"], d["is_periodic"])
    coords = axes_zoo.coords_for(d, 200_000, 5)
    got = g.find_indexes(coords)
    want = c.find_indexes(coords)
    assert got.dtype == np.int64 and got.shape == want.shape
    assert np.array_equal(got, want), np.flatnonzero((got != want).any(1))[:5]
    for bounded in (False, True):
        assert np.array_equal(g.find_index(coords, bounded), c.find_index(coords, bounded))


def test_axis_lookup_device_tensor(pyi, orc_best):
    import torch
    g, c = _mk_axis(pyi, orc_best, axes_zoo.f64_axes()["lat_irregular"])
    coords = axes_zoo.coords_for(c.describe(), 100_000, 9)
    got = g.find_indexes(torch.from_numpy(coords).cuda()).cpu().numpy()
    assert np.array_equal(got, c.find_indexes(coords))


def _grid_pair(pyi, o, nd, dtype, x_kind="lon_periodic", y_kind="lat", z_kind="monthly30", flip_z=False, nan=0.02,
               small=True, pack_cells=True, u_kind=None):
    fz, iz = axes_zoo.f64_axes(), axes_zoo.i64_axes()
    if small:
        specs = [(np.arange(72) * 5.0, dict(period=360.0), False) if x_kind == "lon_periodic" else fz[x_kind],
                 (-87.5 + np.arange(36) * 5.0, {}, False) if y_kind == "lat" else fz[y_kind]]
    else:
        specs = [fz[x_kind], fz[y_kind]]
    if nd >= 3:
        zs = iz[z_kind] if z_kind in iz else fz[z_kind]
        specs.append((zs[0], zs[1], flip_z or zs[2]))
    if nd >= 4 and u_kind is not None:
        specs.append(fz[u_kind])
    elif nd >= 4:
        specs.append((np.round(5500.0 * (np.arange(9) / 8) ** 2.2, 1), {}, False))
    pairs = [_mk_axis(pyi, o, s) for s in specs]
    shape = [len(p[0]) for p in pairs]
    rng = np.random.default_rng(3)
    g = rng.normal(size=shape) * 50.0
    if np.dtype(dtype).kind == "f":
        g = g.astype(dtype)
        if nan:
            g[rng.random(shape) < nan] = np.nan
    else:
        info = np.iinfo(dtype)
        g = np.clip(np.round(g), info.min, info.max).astype(dtype)
    grid = pyi.Grid(*[p[0] for p in pairs], g, pack_cells=pack_cells)
    return grid, [p[1] for p in pairs], g


def _queries(nd, n, seed, xr=(-400, 800), yr=(-95, 95), z_f64=False):
    rng = np.random.default_rng(seed)
    q = [rng.uniform(*xr, n), rng.uniform(*yr, n)]
    q[0][::97] = np.nan
    q[0][5::101] = 5.0 * rng.integers(0, 72, q[0][5::101].size)
    q[1][7::89] = -87.5 + 5.0 * rng.integers(0, 36, q[1][7::89].size)
    if nd >= 3:
        if z_f64:
            q.append(rng.uniform(-100, 5600, n))
        else:
            z = rng.integers(axes_zoo.T0 - 10 * axes_zoo.DAY, axes_zoo.T0 + 340 * axes_zoo.DAY, n, dtype=np.int64)
            z[::113] = np.iinfo(np.int64).min
            z[3::57] = axes_zoo.T0 + 30 * axes_zoo.DAY * rng.integers(0, 12, z[3::57].size)
            q.append(z)
    if nd >= 4:
        q.append(rng.uniform(-50, 5600, n))
    return q


def _call(pyi, grid, q, cfg):
    fn = {2: pyi.core.bivariate, 3: pyi.core.trivariate, 4: pyi.core.quadrivariate}[grid.ndim]
    args = list(q)
    if grid.ndim >= 3 and grid.has_temporal_axis:
        args[2] = args[2].view("datetime64[ns]")
    return fn(grid, *args, cfg)


def _geo_cfg(pyi, nd, method):
    cls = {2: pyi.geometric.Bivariate, 3: pyi.geometric.Trivariate, 4: pyi.geometric.Quadrivariate}[nd]
    return [cls.bilinear, cls.idw, cls.nearest][method]()


@pytest.mark.parametrize("nd", [2, 3, 4])
@pytest.mark.parametrize("dtype", ["float64", "float32", "int16", "int64", "uint8", "uint64", "int32", "int8",
                                   "uint16", "uint32"])
def test_geometric_bilinear(pyi, orc_best, nd, dtype, bilinear_mode):
    n = 50_000
    grid, caxes, g = _grid_pair(pyi, orc_best, nd, dtype)
    q = _queries(nd, n, 100 + nd)
    got = _call(pyi, grid, q, _geo_cfg(pyi, nd, 0))
    want, _ = orc_best.geometric(caxes, g, *q, method=0, num_threads=0)
    exact = close(got, want, 50.0, f"bilinear {nd}D {dtype}")
    # same IEEE operations in the same order: bilinear is expected bit-identical
    # (the tolerance kernel only serves float64-result 2-D grids)
    if bilinear_mode == "exact" or nd > 2 or got.dtype != np.float64:
        assert exact == 1.0
    assert np.isfinite(got).sum() > n // 8


@pytest.mark.parametrize("nd", [2, 3, 4])
@pytest.mark.parametrize("dtype", ["float64", "float32"])
@pytest.mark.parametrize("method", [1, 2])
@pytest.mark.parametrize("y_kind", ["lat", "lat_desc"])
def test_geometric_idw_nearest(pyi, orc_best, nd, dtype, method, y_kind):
    """idw / nearest go through sin / cos / pow, which are not correctly
    rounded on either side: tolerance for idw; for nearest a differing corner
    is accepted only where the two candidate distances tie to rounding."""
    n = 30_000
    grid, caxes, g = _grid_pair(pyi, orc_best, nd, dtype, nan=0.0, y_kind=y_kind)
    q = _queries(nd, n, 300 + nd)
    got = _call(pyi, grid, q, _geo_cfg(pyi, nd, method))
    want, _ = orc_best.geometric(caxes, g, *q, method=method, z_method=method == 2, u_method=method == 2,
                                 num_threads=0)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = ~np.isnan(got)
    tol = TOL[got.dtype] * (1 if method == 1 else 1)
    err = np.abs(got[ok].astype(np.float64) - want[ok].astype(np.float64))
    bound = tol * np.maximum(np.abs(want[ok].astype(np.float64)), 50.0) * (20 if method == 1 else 1)
    bad = np.flatnonzero(err > bound)
    if method == 1:
        assert bad.size == 0, (bad.size, err.max())
    else:
        # nearest: mismatches only at (near-)ties: points on a cell mid-line
        assert bad.size <= max(3, n // 2000), bad.size


@pytest.mark.parametrize("nd", [2, 3, 4])
@pytest.mark.parametrize("dtype", ["float64", "float32", "int16", "uint64"])
@pytest.mark.parametrize("case", [dict(), dict(x_kind="lon_periodic_desc", y_kind="lat_desc", small=False, flip_z=True),
                                  dict(x_kind="irregular_periodic", y_kind="small_irregular_desc")])
def test_plain_layout_matches_packed_and_oracle(pyi, orc_best, nd, dtype, case, monkeypatch):
    """The kernels gather either from the cell-packed copy (default) or, when it
    does not fit, from the caller's C-order array: both must give the oracle's
    bits, for every method."""
    n = 20_000
    monkeypatch.setenv("PYINTERP_B200_BILINEAR", "exact")
    case = dict(case)
    small = case.pop("small", True)
    plain, caxes, g = _grid_pair(pyi, orc_best, nd, dtype, small=small, pack_cells=False, **case)
    packed, _, _ = _grid_pair(pyi, orc_best, nd, dtype, small=small, pack_cells=True, **case)
    ds = [c.describe() for c in caxes]
    q = _queries(nd, n, 11, xr=(ds[0]["min_value"] - 400, ds[0]["max_value"] + 400),
                 yr=(ds[1]["min_value"] - 3, ds[1]["max_value"] + 3))
    for method in (0, 1, 2):
        cfg = _geo_cfg(pyi, nd, method)
        a = _call(pyi, plain, q, cfg)
        b = _call(pyi, packed, q, cfg)
        assert np.array_equal(a, b, equal_nan=True), (nd, dtype, method)
    want, _ = orc_best.geometric(caxes, g, *q, method=0, num_threads=0)
    assert close(_call(pyi, plain, q, _geo_cfg(pyi, nd, 0)), want, 50.0, "plain layout") == 1.0


@pytest.mark.parametrize("case", [
    dict(x_kind="lon_periodic", y_kind="lat", flip_z=True),
    dict(x_kind="lon_partial_period", y_kind="small_irregular_desc"),
    dict(x_kind="lon_180", y_kind="lat_irregular", small=False),
    dict(x_kind="lon_periodic_desc", y_kind="lat_desc", small=False),
    dict(x_kind="irregular_periodic", y_kind="depth"),
    dict(x_kind="lon_0p1", y_kind="lat", small=False),   # float32 rounding quirk of SURVEY 7.3-5
    dict(x_kind="lon_wrapped_values", y_kind="two_points"),
    dict(x_kind="lon_periodic", y_kind="lat", z_kind="irregular_time"),
    dict(x_kind="lon_periodic", y_kind="lat", z_kind="daily_ns_desc"),
    dict(x_kind="lon_periodic", y_kind="lat", z_kind="depth"),
    # descending regular axes through the fast kernels (corner swap): y, then y + z + u
    dict(x_kind="lon_periodic", y_kind="lat_desc"),
    dict(x_kind="lon_180", y_kind="lat_desc_inexact", small=False),
    dict(x_kind="lon_periodic", y_kind="lat_desc", z_kind="levels_desc", u_kind="levels_desc"),
    dict(x_kind="lon_periodic", y_kind="lat_desc_inexact", z_kind="depth", u_kind="levels_desc"),
])
@pytest.mark.parametrize("dtype", ["float64", "float32"])
def test_geometric_axis_kinds(pyi, orc_best, case, dtype, bilinear_mode):
    n = 40_000
    case = dict(case)
    small = case.pop("small", True)
    for nd in (2, 3, 4):
        if nd == 2 and "z_kind" in case:
            continue
        grid, caxes, g = _grid_pair(pyi, orc_best, nd, dtype, small=small, **case)
        ds = [c.describe() for c in caxes]
        q = _queries(nd, n, 7, xr=(ds[0]["min_value"] - 400, ds[0]["max_value"] + 400),
                     yr=(ds[1]["min_value"] - 3, ds[1]["max_value"] + 3), z_f64=nd >= 3 and not caxes[2].is_i64)
        if nd >= 3 and caxes[2].is_i64:
            lo, hi = ds[2]["min_value"], ds[2]["max_value"]
            q[2] = np.random.default_rng(1).integers(lo - (hi - lo) // 10, hi + (hi - lo) // 10, n, dtype=np.int64)
        for zm in (0, 1):
            cfg = _geo_cfg(pyi, nd, 0)
            if nd >= 3:
                cfg = cfg.with_third_axis("nearest" if zm else "linear")
            got = _call(pyi, grid, q, cfg)
            want, _ = orc_best.geometric(caxes, g, *q, method=0, z_method=zm, num_threads=0)
            exact = close(got, want, 50.0, f"{case} {nd}D")
            if bilinear_mode == "exact" or nd > 2 or got.dtype != np.float64:
                assert exact == 1.0
        assert np.isfinite(got).sum() > n // 100


@pytest.mark.parametrize("lat_kind", ["regular", "irregular", "descending", "descending_inexact"])
@pytest.mark.parametrize("dtype", ["float64", "float32", "int16"])
def test_bilinear2d_fast_path_full_grid(pyi, orc_best, lat_kind, dtype, bilinear_mode):
    """BASELINE configs 1 / 5: the 0.25 degree global grid (regular or
    irregular latitude) through the cell-packed fast kernel, host and device
    entry, with adversarial coordinates: exact nodes, the seam cell, the last
    node, far-away longitudes, +-inf, NaN, odd sizes and unaligned device
    pointers (which take the generic kernel) -- all bit-identical."""
    import torch
    s = pyi.synthetic
    lon = s.lon_axis_values()
    lat = s.lat_axis_values() if lat_kind != "irregular" else s.irregular_lat_values()
    if lat_kind == "descending":  # 89.875 -> -89.875 (ERA5-like): the fast kernels swap the corners
        lat = lat[::-1].copy()
    elif lat_kind == "descending_inexact":  # nodes that are not fma(k, step, start): general mode
        lat = 89.9 - np.arange(600) * 0.3
    g = s.field(lon, lat)
    if dtype == "int16":
        g = np.round(g * 1000).astype(np.int16)
    else:
        g = g.astype(dtype)
        g[::37, ::11] = np.nan
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g)
    cax = [orc_best.axis(lon, period=360.0), orc_best.axis(lat)]
    assert cax[1].describe()["kind"] == (1 if lat_kind == "irregular" else 0)
    assert cax[1].describe()["is_ascending"] == (0 if lat_kind.startswith("descending") else 1)
    n = 400_001
    rng = np.random.default_rng(11)
    x = rng.uniform(-720.0, 1080.0, n)
    y = rng.uniform(lat.min() - 0.5, lat.max() + 0.5, n)
    x[0::53] = lon[rng.integers(0, lon.size, x[0::53].size)]              # exact nodes
    x[1::53] = lon[rng.integers(0, lon.size, x[1::53].size)] - 360.0
    x[2::53] = rng.uniform(359.75, 360.0, x[2::53].size)                  # seam cell
    x[3::53] = np.nextafter(lon[rng.integers(0, lon.size, x[3::53].size)], 1e9)
    x[4::53] = np.nextafter(lon[rng.integers(0, lon.size, x[4::53].size)], -1e9)
    x[5::5003] = np.inf
    x[6::5003] = -np.inf
    x[7::5003] = np.nan
    x[8::5003] = rng.uniform(-1e18, 1e18, x[8::5003].size)
    x[9::5003] = 1e300
    y[10::47] = lat[rng.integers(0, lat.size, y[10::47].size)]
    y[11::47] = np.nextafter(lat[rng.integers(0, lat.size, y[11::47].size)], 1e9)
    y[12::47] = np.nextafter(lat[rng.integers(0, lat.size, y[12::47].size)], -1e9)
    y[13::4999] = np.nan
    y[14::4999] = np.inf
    y[15::4999] = lat[-1]
    y[16::4999] = lat[0]
    want, _ = orc_best.geometric(cax, g, x, y, method=0, num_threads=0)
    got = pyi.bivariate(grid, x, y)
    # tolerance mode (float64 results): within 1e-12 of the reference, NaN
    # positions identical; everything else bit-identical
    floor = 1.0 if bilinear_mode == "exact" or dtype != "float64" else 0.0
    assert close(got, want, 1.0, "host") >= floor
    xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    assert close(pyi.bivariate(grid, xd, yd).cpu().numpy(), want, 1.0, "device") >= floor
    # unaligned views (8-byte offset): the vectorised kernel is not eligible
    got_u = pyi.bivariate(grid, xd[1:], yd[1:]).cpu().numpy()
    assert close(got_u, want[1:], 1.0, "device unaligned") == 1.0
    assert np.isfinite(got).sum() > n // 2
    # bounds_error: lowest offending point, first offending axis
    cfg = pyi.geometric.Bivariate.bilinear().with_bounds_error(True)
    werr = orc_best.geometric(cax, g, x, y, method=0, bounds_error=True, num_threads=4)[1]
    with pytest.raises(ValueError) as ei:
        pyi.core.bivariate(grid, x, y, cfg)
    assert str(ei.value) == werr[2]
    with pytest.raises(ValueError) as ei:
        pyi.core.bivariate(grid, xd, yd, cfg)
    assert str(ei.value) == werr[2]


@pytest.mark.parametrize("nd", [2, 3, 4])
def test_geometric_bounds_error(pyi, orc_best, nd):
    grid, caxes, g = _grid_pair(pyi, orc_best, nd, "float64")
    q = _queries(nd, 20_000, 9)
    want = orc_best.geometric(caxes, g, *q, bounds_error=True, num_threads=4)[1]
    assert want is not None
    with pytest.raises(ValueError, match="out of bounds") as ei:
        _call(pyi, grid, q, _geo_cfg(pyi, nd, 0).with_bounds_error(True))
    if want[1] != 2:  # the oracle prints temporal values as integers
        assert str(ei.value) == want[2]
    else:
        assert "for axis z" in str(ei.value)
    # in-bounds points only: no error, finite
    qq = [np.array([10.0, 20.0]), np.array([0.0, 10.0]), np.array([axes_zoo.T0 + axes_zoo.DAY] * 2, dtype=np.int64),
          np.array([10.0, 20.0])][:nd]
    out = _call(pyi, grid, qq, _geo_cfg(pyi, nd, 0).with_bounds_error(True))
    assert out.shape == (2,)


def test_bounds_error_first_point_and_axis(pyi, orc_best):
    """Lowest offending index, first offending axis in x, y order, across
    chunk boundaries of the host pipeline (> 2^23 points)."""
    grid, caxes, g = _grid_pair(pyi, orc_best, 2, "float64", x_kind="lon_partial_period", nan=0.0)
    n = (1 << 23) + 1000
    x = np.full(n, 50.0)
    y = np.full(n, 0.0)
    x[n - 7] = 150.0     # outside lon_partial_period [0, 100]
    y[n - 7] = 95.0      # y also bad: x must be reported
    y[n - 3] = 99.0
    with pytest.raises(ValueError) as ei:
        pyi.core.bivariate(grid, x, y, pyi.geometric.Bivariate.bilinear().with_bounds_error(True))
    assert str(ei.value).startswith("150 is out of bounds for axis x [0, ..., 100]")
    x[n - 7] = 50.0
    with pytest.raises(ValueError) as ei:
        pyi.core.bivariate(grid, x, y, pyi.geometric.Bivariate.bilinear().with_bounds_error(True))
    assert str(ei.value).startswith("95 is out of bounds for axis y")


@pytest.mark.parametrize("method", ["bicubic", "bilinear"])
@pytest.mark.parametrize("boundary", ["undef", "shrink"])
@pytest.mark.parametrize("half", [(3, 3), (2, 4), (1, 1), (5, 2), (8, 8)])
def test_windowed2d(pyi, orc_best, bicubic_mode, method, boundary, half):
    n = 30_000
    for dtype, nan in (("float64", 0.0), ("float32", 0.002), ("int16", 0.0)):
        grid, caxes, g = _grid_pair(pyi, orc_best, 2, dtype, nan=nan)
        q = _queries(2, n, 21)
        cfg = getattr(pyi.windowed.Bivariate, method)().with_half_window_size_x(half[0]) \
            .with_half_window_size_y(half[1]).with_boundary_mode(getattr(pyi.windowed.BoundaryConfig, boundary)())
        got = pyi.core.bivariate(grid, q[0], q[1], cfg)
        want, _ = orc_best.windowed2d(caxes, g, q[0], q[1], method=0 if method == "bicubic" else 1,
                                      half_x=half[0], half_y=half[1], boundary=boundary, num_threads=0,
                                      stateless=True)
        exact = close(got, want, 50.0, f"{method} {boundary} {half} {dtype}")  # 1e-12 / 1e-5 inside
        # identical operation sequence per spline solve, except in the tolerance kernels of the
        # 6 x 6 window (weights / separable stencil, which also serves windowed bilinear)
        assert exact == 1.0 or (bicubic_mode == "weights" and half == (3, 3))


@pytest.mark.parametrize("case", [
    dict(x_kind="lon_180", y_kind="lat_desc", small=False),
    dict(x_kind="lon_partial_period", y_kind="lat_irregular", small=False),
    dict(x_kind="lon_periodic_desc", y_kind="small_irregular"),
    dict(x_kind="irregular_periodic", y_kind="depth"),
])
def test_windowed2d_axis_kinds(pyi, orc_best, case):
    case = dict(case)
    small = case.pop("small", True)
    grid, caxes, g = _grid_pair(pyi, orc_best, 2, "float64", small=small, nan=0.0005, **case)
    ds = [c.describe() for c in caxes]
    q = _queries(2, 30_000, 4, xr=(ds[0]["min_value"] - 400, ds[0]["max_value"] + 400),
                 yr=(ds[1]["min_value"] - 3, ds[1]["max_value"] + 3))
    for boundary in ("undef", "shrink"):
        got = pyi.bivariate(grid, q[0], q[1], "bicubic", boundary_mode=boundary)
        want, _ = orc_best.windowed2d(caxes, g, q[0], q[1], boundary=boundary, num_threads=0)
        close(got, want, 50.0, f"bicubic {case} {boundary}")


def test_windowed2d_bounds_error(pyi, orc_best):
    grid, caxes, g = _grid_pair(pyi, orc_best, 2, "float64", nan=0.0)
    q = _queries(2, 5000, 33)
    want = orc_best.windowed2d(caxes, g, q[0], q[1], bounds_error=True, num_threads=3)[1]
    with pytest.raises(IndexError) as ei:
        pyi.bivariate(grid, q[0], q[1], "bicubic", bounds_error=True)
    assert str(ei.value) == want[2]
    # a window that does not fit is NaN, not an error (cache_loader.hpp:69-80)
    out = pyi.bivariate(grid, np.array([10.0]), np.array([-86.0]), "bicubic", bounds_error=True)
    assert np.isnan(out[0])


def test_device_tensors_match_host_path(pyi):
    import torch
    s = pyi.synthetic
    lon, lat = s.lon_axis_values(), s.lat_axis_values()
    g = s.field(lon, lat)
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g)
    x, y = s.queries(1_000_000, 1001)
    host = pyi.bivariate(grid, x, y)
    dev = pyi.bivariate(grid, torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())
    assert dev.is_cuda and dev.dtype == torch.float64
    assert np.array_equal(host, dev.cpu().numpy(), equal_nan=True)
    hostc = pyi.bivariate(grid, x, y, "bicubic")
    devc = pyi.bivariate(grid, torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), "bicubic")
    assert np.array_equal(hostc, devc.cpu().numpy(), equal_nan=True)


def test_api_errors(pyi):
    s = pyi.synthetic
    lon, lat = np.arange(72) * 5.0, -87.5 + np.arange(36) * 5.0
    ax, ay = pyi.Axis(lon, period=360.0), pyi.Axis(lat)
    grid = pyi.Grid(ax, ay, np.zeros((72, 36), dtype=np.float32))
    assert grid.result_dtype == np.float32
    with pytest.raises(ValueError, match="trivariate requires 3D grid, got 2D"):
        pyi.trivariate(grid, np.zeros(3), np.zeros(3), np.zeros(3))
    with pytest.raises(ValueError, match="could not be broadcast"):
        pyi.bivariate(grid, np.zeros(3), np.zeros(4))
    with pytest.raises(ValueError, match="doesn't match array shape"):
        pyi.Grid(ax, ay, np.zeros((72, 35)))
    with pytest.raises(ValueError, match="y-axis cannot be a circle"):
        pyi.Grid(ay, ax, np.zeros((36, 72)))
    with pytest.raises(ValueError, match="Unknown"):
        pyi.bivariate(grid, np.zeros(3), np.zeros(3), "spline42")
    out = pyi.bivariate(grid, np.zeros(0), np.zeros(0))
    assert out.shape == (0,) and out.dtype == np.float32


def test_grid_array_is_borrowed_and_refreshed(pyi):
    """The reference borrows the numpy buffer; later writes must be seen
    (pyinterp/tests/core/windowed/test_bivariate.py:234-250)."""
    lon, lat = np.arange(72) * 5.0, -87.5 + np.arange(36) * 5.0
    arr = np.ones((72, 36))
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), arr)
    assert np.shares_memory(grid.array, arr)
    x, y = np.array([12.0]), np.array([1.0])
    assert pyi.bivariate(grid, x, y)[0] == 1.0
    arr[2:4, 17:19] = np.nan
    assert np.isnan(pyi.bivariate(grid, x, y)[0])
    assert np.isnan(pyi.bivariate(grid, x, y, "bicubic")[0])


def _bicubic_both(pyi, grid, x, y, **kw):
    """6 x 6 bicubic through the fast kernel and through the generic one."""
    import os
    fast = pyi.bivariate(grid, x, y, "bicubic", **kw)
    os.environ["PYINTERP_B200_GENERIC"] = "1"
    try:
        generic = pyi.bivariate(grid, x, y, "bicubic", **kw)
    finally:
        del os.environ["PYINTERP_B200_GENERIC"]
    return fast, generic


def _same_or_close(mode, a, b, what):
    """exact mode: bit for bit; weights mode: NaN positions identical, values
    within 1e-12 (float64) / 1e-5 (float32) of max(|a|, |b|, 50)."""
    if mode == "exact":
        assert np.array_equal(a, b, equal_nan=True), what
    else:
        close(a, b, 50.0, what)


@pytest.mark.parametrize("boundary", ["undef", "shrink"])
def test_bicubic6_fast_path_full_grid(pyi, orc_best, bicubic_mode, boundary):
    """The 6 x 6 bicubic fast kernels on the 0.25 degree grid against the generic
    kernel and the oracle: bit-identical (pyi_fastwin.cuh) or within tolerance
    (pyi_bicubicw.cuh)."""
    s = pyi.synthetic
    lon, lat = s.lon_axis_values(), s.lat_axis_values()
    g = s.field(lon, lat)
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g)
    x, y = s.queries(400_000, 1002)
    rng = np.random.default_rng(8)
    x[::53] = rng.choice(lon, x[::53].size)          # on a node: bracket (k-1, k)
    y[::59] = rng.choice(lat, y[::59].size)
    x[1::1013] = np.nan
    x[2::1017] = np.inf
    y[3::1019] = 1e300
    fast, generic = _bicubic_both(pyi, grid, x, y, boundary_mode=boundary)
    _same_or_close(bicubic_mode, fast, generic, "fast vs generic")
    caxes = [orc_best.axis(lon, period=360.0), orc_best.axis(lat)]
    want, _ = orc_best.windowed2d(caxes, g, x, y, boundary=boundary, num_threads=0, stateless=True)
    assert np.array_equal(generic, want, equal_nan=True)
    _same_or_close(bicubic_mode, fast, want, "fast vs oracle")
    assert np.isfinite(fast).mean() > 0.95


@pytest.mark.parametrize("dtype", ["float64", "float32", "int32", "uint8"])
def test_bicubic6_fast_path_awkward_values(pyi, orc_best, bicubic_mode, dtype):
    """Values for which the short division of the fast path is not proven
    (huge, tiny, subnormal, infinite, NaN) take the generic code; zeros and
    constant patches stay on the fast path.  Irregular and non-periodic axes."""
    rng = np.random.default_rng(12)
    lon = np.sort(rng.uniform(0, 360, 200))
    lat = np.degrees(np.arcsin(np.linspace(-0.999, 0.999, 90)))
    g = rng.normal(size=(200, 90)) * 30
    if dtype == "float64":
        if bicubic_mode == "exact":
            # 1e250 next to O(30) values: the result is dominated by the rounding of
            # whichever operation order is used; only the bit-identical path can be
            # compared (the weights kernel agrees to ~1e-16 of the WINDOW's magnitude)
            g[rng.random(g.shape) < 0.01] = 1e250
        g[rng.random(g.shape) < 0.01] = 1e-305
        g[rng.random(g.shape) < 0.01] = 5e-324
        g[rng.random(g.shape) < 0.005] = np.inf
        g[rng.random(g.shape) < 0.005] = np.nan
    g[rng.random(g.shape) < 0.2] = 0.0
    g[40:60, 20:40] = 7.0
    g = g.astype(dtype)
    for kw in (dict(period=360.0), {}):
        grid = pyi.Grid(pyi.Axis(lon, **kw), pyi.Axis(lat), g)
        x = rng.uniform(-200, 500, 100_000)
        y = rng.uniform(-91, 91, 100_000)
        fast, generic = _bicubic_both(pyi, grid, x, y)
        _same_or_close(bicubic_mode, fast, generic, "fast vs generic")
        caxes = [orc_best.axis(lon, **kw), orc_best.axis(lat)]
        want, _ = orc_best.windowed2d(caxes, g, x, y, num_threads=0, stateless=True)
        assert np.array_equal(generic, want, equal_nan=True)
        _same_or_close(bicubic_mode, fast, want, "fast vs oracle")
        assert np.array_equal(np.isinf(fast), np.isinf(want))


def test_windowed2d_bounds_error_in_late_chunk(pyi, orc_best):
    """Host buffers are processed in chunks; the reported index is global."""
    s = pyi.synthetic
    lon, lat = s.lon_axis_values(144), s.lat_axis_values(72)
    g = s.field(lon, lat)
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g)
    n = 20_000_000
    x = np.full(n, 10.0)
    y = np.full(n, 1.0)
    y[n - 5] = 95.0
    for method, half in (("bicubic", 3), ("bicubic", 2), ("bilinear", 3)):
        cfg = getattr(pyi.windowed.Bivariate, method)().with_half_window_size_x(half) \
            .with_half_window_size_y(half).with_bounds_error(True)
        with pytest.raises(IndexError) as ei:
            pyi.core.bivariate(grid, x, y, cfg)
        assert str(ei.value).startswith("95 is out of bounds for axis y"), str(ei.value)
    import torch
    with pytest.raises(IndexError):
        pyi.bivariate(grid, torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), "bicubic", bounds_error=True)


class _CudaArray:
    """A foreign device array: only ``__cuda_array_interface__`` (what CuPy /
    Numba / RAPIDS objects expose), backed by a torch allocation."""

    def __init__(self, t):
        self._t = t
        self.__cuda_array_interface__ = t.__cuda_array_interface__


def test_cuda_array_interface_inputs(pyi):
    """Zero-copy front end: coordinates and the grid given as
    foreign device arrays; the result stays on the device."""
    import torch
    s = pyi.synthetic
    lon, lat = s.lon_axis_values(144), s.lat_axis_values(72)
    g = s.field(lon, lat)
    x, y = s.queries(100_000, 7)
    host_grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g)
    want = pyi.bivariate(host_grid, x, y)
    dg = _CudaArray(torch.from_numpy(g).cuda())
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), dg)
    dx, dy = _CudaArray(torch.from_numpy(x).cuda()), _CudaArray(torch.from_numpy(y).cuda())
    got = pyi.bivariate(grid, dx, dy)
    assert got.is_cuda and np.array_equal(got.cpu().numpy(), want, equal_nan=True)
    got = pyi.bivariate(grid, dx, dy, "c_spline")
    assert np.array_equal(got.cpu().numpy(), pyi.bivariate(host_grid, x, y, "c_spline"), equal_nan=True)
    idx = grid.x.find_indexes(dx)
    assert idx.is_cuda and np.array_equal(idx.cpu().numpy(), host_grid.x.find_indexes(x))


@pytest.mark.parametrize("unit", ["ns", "s", "h", "D", "W"])
def test_temporal_bounds_error_message_is_the_reference_text(pyi, orc_best, unit):
    """The whole bounds-error message of a temporal z axis: coordinate and axis
    limits printed by TemporalAxis::coordinate_repr (math/temporal_axis.hpp:80-83,
    dateutils.hpp:729-772) in the axis resolution, compared with the reference
    build's own strings (the geometric driver, pybind/grid.hpp:362-371,376-381,
    and the windowed one)."""
    lon, lat = np.arange(36) * 10.0, -85.0 + np.arange(18) * 10.0
    t = (np.datetime64("2021-01-04", unit) + np.arange(8) * np.timedelta64(3, unit)).astype(f"datetime64[{unit}]")
    g = np.random.default_rng(2).normal(size=(36, 18, 8))
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), pyi.TemporalAxis(t), g)
    ti = t.astype(np.int64)
    x, y = np.array([5.0, 15.0, 25.0]), np.array([0.0, 10.0, -20.0])
    z = np.array([ti[2], ti[-1] + 5, ti[0] - 1], dtype=np.int64).view(f"datetime64[{unit}]")
    want = (f"{orc_best.temporal_repr(int(ti[-1]) + 5, unit)} is out of bounds for axis z "
            f"[{orc_best.temporal_repr(int(ti[0]), unit)}, ..., {orc_best.temporal_repr(int(ti[-1]), unit)}]")
    with pytest.raises(ValueError) as ei:
        pyi.trivariate(grid, x, y, z, bounds_error=True)
    assert str(ei.value) == want
    with pytest.raises(IndexError) as ei:
        pyi.trivariate(grid, x, y, z, "bicubic", bounds_error=True, half_window_size_x=2,
                       half_window_size_y=2, boundary_mode="shrink")
    assert str(ei.value) == want
    # numpy prints the same ISO text for these resolutions
    assert orc_best.temporal_repr(int(ti[0]), unit) == str(t[0])
    # a coarser-resolution query array is converted to the axis resolution first
    assert grid.z.coordinate_repr(int(ti[3])) == str(t[3])


@pytest.mark.parametrize("method", ["bilinear", "bicubic"])
def test_offset_axes_keep_the_bit_identical_path(pyi, orc_best, method):
    """Axes with a large |coordinate| / step ratio (start = 1e7, step = 0.01):
    the tolerance kernels assume nodes that are uniform to rounding, which such
    axes are not (ulp(1e7) / 0.01 = 1.9e-7 relative spacing error), so the
    library must route them through the bit-identical kernels in the DEFAULT
    value mode (bicubic weights kernel: max |coordinate| / |step| < 2^16;
    bilinear tolerance kernel: < 2^26).  A field crossing zero makes any
    derivative error visible."""
    xs = 1.0e7 + np.arange(400) * 0.01
    ys = -3.0e6 + np.arange(300) * 0.01
    rng = np.random.default_rng(12)
    g = np.sin(np.arange(400) * 0.3)[:, None] * np.cos(np.arange(300) * 0.2)[None, :] * 50.0 + rng.normal(size=(400, 300))
    grid = pyi.Grid(pyi.Axis(xs), pyi.Axis(ys), g)
    cax = [orc_best.axis(xs), orc_best.axis(ys)]
    n = 200_000
    x = rng.uniform(xs[0] - 0.05, xs[-1] + 0.05, n)
    y = rng.uniform(ys[0] - 0.05, ys[-1] + 0.05, n)
    x[::17] = xs[rng.integers(0, xs.size, x[::17].size)]
    if method == "bilinear":
        want, _ = orc_best.geometric(cax, g, x, y, method=0, num_threads=0)
    else:
        want, _ = orc_best.windowed2d(cax, g, x, y, num_threads=0, stateless=True)
    got = pyi.bivariate(grid, x, y, method)
    assert np.isfinite(want).mean() > 0.5
    assert np.array_equal(got, want, equal_nan=True)


@pytest.mark.parametrize("kind", ["step_0p1", "offset_1000", "lon_0p1_periodic", "lon_180_0p05"])
def test_bilinear_tolerance_kernel_on_inexact_nodes(pyi, orc_best, kind):
    """The general variant of the tolerance kernel (nodes start + k * step that
    are NOT exactly representable, spacings that differ from `step` by
    rounding): indexes identical, values within 1e-12 of the reference on a
    rough field with zero crossings, scaled by the cell's own magnitude."""
    rng = np.random.default_rng(5)
    per = None
    if kind == "step_0p1":
        xs, ys = np.arange(900) * 0.1 + 0.03, -40.0 + np.arange(700) * 0.1
    elif kind == "offset_1000":
        xs, ys = 1000.3 + np.arange(500) * 0.1, 5000.7 + np.arange(400) * 0.3
    elif kind == "lon_0p1_periodic":
        xs, ys, per = np.arange(3600) * 0.1, -89.95 + np.arange(1800) * 0.1, 360.0
    else:
        xs, ys, per = -180.0 + np.arange(7200) * 0.05, -60.0 + np.arange(500) * 0.05, 360.0
    g = rng.normal(size=(xs.size, ys.size)) * 50.0
    g[::31, ::7] = np.nan
    grid = pyi.Grid(pyi.Axis(xs, period=per), pyi.Axis(ys), g)
    cax = [orc_best.axis(xs, period=per), orc_best.axis(ys)]
    n = 500_000
    lo, hi = (xs[0] - 400.0, xs[-1] + 400.0) if per else (xs[0] - 1.0, xs[-1] + 1.0)
    x = rng.uniform(lo, hi, n)
    y = rng.uniform(ys[0] - 0.5, ys[-1] + 0.5, n)
    x[::19] = xs[rng.integers(0, xs.size, x[::19].size)]
    y[3::23] = ys[rng.integers(0, ys.size, y[3::23].size)]
    x[5::1009] = rng.uniform(-1e9, 1e9, x[5::1009].size)
    want, _ = orc_best.geometric(cax, g, x, y, method=0, num_threads=0)
    tol = pyi.core.bivariate(grid, x, y, pyi.geometric.Bivariate.bilinear().with_value_mode("tolerance"))
    exact = pyi.core.bivariate(grid, x, y, pyi.geometric.Bivariate.bilinear().with_value_mode("exact"))
    assert np.array_equal(exact, want, equal_nan=True)
    assert np.array_equal(np.isnan(tol), np.isnan(want))
    m = ~np.isnan(want)
    assert m.mean() > 0.3
    err = np.abs(tol[m] - want[m])
    assert np.all(err <= 1e-12 * np.maximum(np.abs(want[m]), 50.0)), err.max()
    # rounding only: a few ulp of the cell's values (|q| up to ~250)
    assert err.max() < 5e-13, err.max()


def test_bicubic_reference_order_vs_stateless_near_the_seam(pyi, orc_best):
    """the reference keeps a per-thread window cache whose
    contents depend on the previous point (cache_loader.hpp:110-133,194-203);
    the product is stateless.  Quantify the gap where it can exist -- queries
    hopping across the 0 / 360 seam -- between the reference-order run (one
    thread, points in order) and the stateless form, and hold the product to
    the stateless form."""
    lon, lat = np.arange(144) * 2.5, -88.75 + np.arange(72) * 2.5
    g = pyi.synthetic.field(lon, lat)
    cax = [orc_best.axis(lon, period=360.0), orc_best.axis(lat)]
    rng = np.random.default_rng(8)
    n = 60_000
    x = np.where(rng.random(n) < 0.5, rng.uniform(350.0, 360.0, n), rng.uniform(0.0, 10.0, n))
    x[::3] += 360.0 * rng.integers(-2, 3, x[::3].size)
    y = rng.uniform(-80.0, 80.0, n)
    stateless, _ = orc_best.windowed2d(cax, g, x, y, num_threads=1, stateless=True)
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g)
    got = pyi.core.bivariate(grid, x, y, pyi.windowed.Bivariate.bicubic().with_value_mode("exact"))
    assert np.array_equal(got, stateless, equal_nan=True)
    if orc_best.kind == "reference":
        stateful, _ = orc_best.windowed2d(cax, g, x, y, num_threads=1, stateless=False)
        assert np.array_equal(np.isnan(stateful), np.isnan(stateless))
        gap = np.nanmax(np.abs(stateful - stateless))
        print(f"max |reference-order - stateless| near the seam: {gap:.3e}")
        assert gap <= 1e-12


def test_divisors_with_all_ones_significand(pyi, orc_best, monkeypatch):
    """Cell spacings one ulp below a power of two (significand all ones): the
    Newton reciprocal behind the kernels' own IEEE division ends on an exact tie
    for such a divisor unless it is seeded as nvcc seeds it (low word 1,
    pyi_common.cuh rcp_seed).  Irregular axes whose EVERY spacing is
    2^k (1 - 2^-53): bilinear and bicubic must stay bit-identical in exact mode
    (found as 1-ulp differences on a 360 / 21 degree periodic axis, where the
    spline's 1 / 63.99999999999999 came out as 2^-6)."""
    one = 1.0 - 2.0 ** -53
    nodes = np.r_[0.0, one * 2.0 ** np.arange(0, 10)]
    h = np.diff(nodes)
    assert np.all((h.view(np.uint64) & np.uint64(0xFFFFFFFFFFFFF)) == np.uint64(0xFFFFFFFFFFFFF))
    rng = np.random.default_rng(77)
    g = rng.normal(size=(nodes.size, nodes.size)) * 10.0
    n = 200_000
    x = rng.uniform(0.0, nodes[-1], n)
    y = rng.uniform(0.0, nodes[-1], n)
    # more points in the small cells
    x[::2] = rng.uniform(0.0, 8.0, x[::2].size)
    y[::3] = rng.uniform(0.0, 8.0, y[::3].size)
    oaxes = [orc_best.axis(nodes), orc_best.axis(nodes)]
    assert oaxes[0].describe()["kind"] == 1
    grid = pyi.Grid(pyi.Axis(nodes), pyi.Axis(nodes), g)
    monkeypatch.setenv("PYINTERP_B200_BILINEAR", "exact")
    monkeypatch.setenv("PYINTERP_B200_BICUBIC", "exact")
    want, _ = orc_best.geometric(oaxes, g, x, y, method=0, num_threads=0)
    assert np.array_equal(pyi.bivariate(grid, x, y), want, equal_nan=True)
    want, _ = orc_best.windowed2d(oaxes, g, x, y, num_threads=0, stateless=True)
    got = pyi.bivariate(grid, x, y, "bicubic")
    assert np.isfinite(want).mean() > 0.2
    assert np.array_equal(got, want, equal_nan=True)
    # regular periodic longitude with the same kind of divisor inside the spline solve
    lon = 10.0 + np.arange(21) * (360.0 / 21)
    lat = np.arange(24) * 0.5
    g2 = rng.normal(size=(21, 24)) * 10.0
    x2 = rng.uniform(-400.0, 800.0, n)
    y2 = rng.uniform(1.2, 10.2, n)
    o2 = [orc_best.axis(lon, period=360.0), orc_best.axis(lat)]
    grid2 = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g2)
    want, _ = orc_best.windowed2d(o2, g2, x2, y2, num_threads=0, stateless=True)
    assert np.array_equal(pyi.bivariate(grid2, x2, y2, "bicubic"), want, equal_nan=True)
    want, _ = orc_best.geometric(o2, g2, x2, y2, method=0, num_threads=0)
    assert np.array_equal(pyi.bivariate(grid2, x2, y2), want, equal_nan=True)


def test_tracked_host_grid_in_place_and_refresh(pyi, synthetic, orc_best):
    """A grid over a borrowed host array (pyi_grid_track_host): small calls read the page-locked array
    in place, large calls refresh the device copies first; both see what the caller wrote into the
    array since the last call (pyinterp/tests/core/windowed/test_bivariate.py:234-250), and agree with
    a resident grid over a copy of the same values.  3-D as well; two handles over one array."""
    lon, lat = synthetic.lon_axis_values(1440), synthetic.lat_axis_values(720)
    g = synthetic.field(lon, lat)
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g)
    assert grid._tracked
    caxes = [orc_best.axis(lon, period=360.0), orc_best.axis(lat)]
    rng = np.random.default_rng(12)
    # in place; in place; plain array refreshed, generic kernels (400 000 x 128 B x 4 > 8.3 MB, below
    # 2^20 points); packed copy rebuilt, fast kernels
    for n in (50, 3_000, 400_000, 1_200_000):
        x, y = rng.uniform(-200.0, 500.0, n), rng.uniform(-91.0, 91.0, n)
        for round_ in range(2):
            g[rng.integers(0, 1440, 50), rng.integers(0, 720, 50)] = rng.normal(size=50)  # the caller writes
            g[5 + round_, 7] = np.nan
            resident = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g.copy(), resident=True)
            for method in ("bilinear", "nearest", "bicubic", "akima"):
                got = pyi.bivariate(grid, x, y, method)
                want = pyi.bivariate(resident, x, y, method)
                assert np.array_equal(np.isnan(got), np.isnan(want)), (n, method)
                close(got, want, 1.0, f"tracked {method} n={n}")
            want, _ = orc_best.geometric(caxes, g, x, y, num_threads=0)
            got = pyi.core.bivariate(grid, x, y, pyi.geometric.Bivariate.bilinear().with_value_mode("exact"))
            assert np.array_equal(got, want, equal_nan=True), n
    # bounds error from the in-place path
    with pytest.raises(ValueError, match="out of bounds"):
        pyi.bivariate(grid, np.array([10.0, 20.0]), np.array([0.0, 95.0]), "bilinear", bounds_error=True)
    # a second handle over the same array; the first one goes away
    other = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g)
    x, y = rng.uniform(0.0, 360.0, 100), rng.uniform(-80.0, 80.0, 100)
    want = pyi.bivariate(grid, x, y)
    del grid
    g[:, :] += 1.0
    assert np.allclose(pyi.bivariate(other, x, y), want + 1.0, rtol=0, atol=1e-12, equal_nan=True)
    # 3-D, temporal axis
    t = synthetic.time_axis_values(40)
    g3 = synthetic.field(lon[::4], lat[::4], nz=40, dtype="float32")
    grid3 = pyi.Grid(pyi.Axis(lon[::4], period=360.0), pyi.Axis(lat[::4]), pyi.TemporalAxis(t), g3)
    res3 = pyi.Grid(pyi.Axis(lon[::4], period=360.0), pyi.Axis(lat[::4]), pyi.TemporalAxis(t), g3.copy(), resident=True)
    for n in (200, 300_000):
        x, y, z = synthetic.queries(n, 3, ndim=3, time_values=t)
        for method in ("bilinear", "c_spline"):
            got = pyi.trivariate(grid3, x, y, z.view("datetime64[ns]"), method)
            want = pyi.trivariate(res3, x, y, z.view("datetime64[ns]"), method)
            close(got, want, 1.0, f"tracked 3-D {method} n={n}")
