"""GPU parity at the BASELINE.json sizes (configs 2, 3, 4): the full
1440 x 720 x 365 float32 grid with a daily datetime64[ns] axis, the full
1440 x 720 x 12 x 50 float32 grid with an irregular depth axis (cell-packed
copy of 12 / 40 GB, element offsets above 2^31), and 1e7 bicubic points on the
0.25 degree grid -- all against the CPU oracle (oracle/_ref when present) on
identical SURVEY.md 8d inputs: x ~ U(-180, 540), y ~ U(-90, 90), NaN / NaT,
exact nodes, first and last nodes.

Bars: geometric bilinear bit-identical (packed and plain layout); bicubic
bit-identical in exact mode, within 1e-12 in the default (weights) mode."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orc_best():
    import orc
    return orc.Oracle("reference") if orc.available("reference") else orc.Oracle("port")


def _adversarial(q, case, rng):
    """Exact nodes, first / last nodes and the seam on every axis (on top of
    the NaN / NaT coordinates `synthetic.queries` already plants)."""
    n = q[0].size
    lon, lat = case["lon"], case["lat"]
    q[0][1::41] = lon[rng.integers(0, lon.size, q[0][1::41].size)]
    q[0][2::4001] = lon[-1]
    q[0][3::4001] = rng.uniform(359.75, 360.0, q[0][3::4001].size)
    q[0][4::4001] = lon[0] + 360.0
    q[1][5::43] = lat[rng.integers(0, lat.size, q[1][5::43].size)]
    q[1][6::4003] = lat[-1]
    q[1][7::4003] = lat[0]
    if len(q) >= 3:
        tv = case["time"].astype("datetime64[ns]").astype(np.int64)
        q[2][8::47] = tv[rng.integers(0, tv.size, q[2][8::47].size)]
        q[2][9::4007] = tv[-1]
        q[2][10::4007] = tv[0]
        q[2][11::4007] = tv[-1] + 1
        q[2][12::4007] = tv[0] - 1
    if len(q) >= 4:
        dep = case["depth"]
        q[3][13::53] = dep[rng.integers(0, dep.size, q[3][13::53].size)]
        q[3][14::4009] = dep[-1]
        q[3][15::4009] = dep[0]
    assert n == q[0].size
    return q


def _case(synthetic, nd):
    lon, lat = synthetic.lon_axis_values(), synthetic.lat_axis_values()
    if nd == 3:
        t = synthetic.time_axis_values(365)
        return dict(nd=3, lon=lon, lat=lat, time=t,
                    grid=synthetic.field(lon, lat, nz=365, dtype="float32", noise=0.0))
    t = synthetic.time_axis_values(12, 30)
    depth = synthetic.depth_axis_values(50)
    return dict(nd=4, lon=lon, lat=lat, time=t, depth=depth,
                grid=synthetic.field(lon, lat, nz=12, nu=50, dtype="float32", noise=0.0))


@pytest.mark.parametrize("nd", [3, 4])
def test_geometric_bilinear_at_baseline_size(pyi, synthetic, orc_best, nd):
    """BASELINE configs 3 / 4 at size, bit-identical to the oracle through the
    cell-packed fast kernels AND the plain C-order layout, host and device
    entry, linear and nearest along z / u."""
    import torch
    case = _case(synthetic, nd)
    g = case["grid"]
    # a land mask: NaN cells must propagate exactly as in the reference
    g[::29, ::13] = np.nan
    n = 1_200_001
    rng = np.random.default_rng(20 + nd)
    q = _adversarial(synthetic.queries(n, 1000 + nd, ndim=nd, time_values=case["time"]), case, rng)
    oaxes = [orc_best.axis(case["lon"], period=360.0), orc_best.axis(case["lat"]),
             orc_best.axis(case["time"].astype("datetime64[ns]").astype(np.int64))]
    if nd == 4:
        oaxes.append(orc_best.axis(case["depth"]))
        assert oaxes[3].describe()["kind"] == 1  # irregular depth
    want = {zm: orc_best.geometric(oaxes, g, *q, method=0, z_method=zm, u_method=zm, num_threads=0)[0]
            for zm in (0, 1)}
    assert np.isfinite(want[0]).mean() > 0.9
    qq = list(q)
    qq[2] = qq[2].view("datetime64[ns]")
    cls = {3: pyi.geometric.Trivariate, 4: pyi.geometric.Quadrivariate}[nd]
    fn = {3: pyi.core.trivariate, 4: pyi.core.quadrivariate}[nd]
    for pack in (True, False):
        axes = [pyi.Axis(case["lon"], period=360.0), pyi.Axis(case["lat"]), pyi.TemporalAxis(case["time"])]
        if nd == 4:
            axes.append(pyi.Axis(case["depth"]))
        grid = pyi.Grid(*axes, g, resident=True, pack_cells=pack)
        for zm in (0, 1):
            cfg = cls.bilinear().with_third_axis("nearest" if zm else "linear")
            if nd == 4:
                cfg = cfg.with_fourth_axis("nearest" if zm else "linear")
            got = fn(grid, *qq, cfg)
            assert got.dtype == np.float32
            assert np.array_equal(got, want[zm], equal_nan=True), (nd, pack, zm)
        # device entry, default configuration
        dq = [torch.from_numpy(a).cuda() for a in q]
        got = fn(grid, *dq, cls.bilinear()).cpu().numpy()
        assert np.array_equal(got, want[0], equal_nan=True), (nd, pack, "device")
        del grid, dq
        torch.cuda.empty_cache()


def test_bicubic_at_baseline_size(pyi, synthetic, orc_best, bicubic_mode):
    """BASELINE config 2: 1e7 bicubic points (6 x 6 window, bounds_error=False)
    on the 0.25 degree grid with a NaN land mask."""
    lon, lat = synthetic.lon_axis_values(), synthetic.lat_axis_values()
    g = synthetic.field(lon, lat)
    g[::37, ::11] = np.nan
    case = dict(lon=lon, lat=lat)
    n = 10_000_000
    rng = np.random.default_rng(31)
    q = _adversarial(synthetic.queries(n, 1002), case, rng)
    oaxes = [orc_best.axis(lon, period=360.0), orc_best.axis(lat)]
    want, _ = orc_best.windowed2d(oaxes, g, q[0], q[1], num_threads=0, stateless=True)
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g, resident=True)
    got = pyi.bivariate(grid, q[0], q[1], "bicubic")
    assert np.array_equal(np.isnan(got), np.isnan(want))
    m = ~np.isnan(want)
    assert m.mean() > 0.8
    err = np.abs(got[m] - want[m])
    if bicubic_mode == "exact":
        assert err.max() == 0.0
    else:
        assert np.all(err <= 1e-12 * np.maximum(np.abs(want[m]), 1.0)), err.max()


@pytest.mark.parametrize("dtype,ny", [("float32", 720), ("float32", 719), ("float64", 719)])
def test_bicubic_separable_dtypes_and_odd_ny(pyi, synthetic, orc_best, bicubic_mode, dtype, ny):
    """The separable stencil on a FLOAT32 0.25 degree grid (8-byte pieces of two
    values) and on grids with an ODD ny (the rows of a window alternate parity;
    a window that wraps onto row 0 at wy0 = 0 once read past the array): results
    within 1e-5 / 1e-12 of the oracle, NaN positions identical (the land mask has
    a NaN at element 0); device and host entry."""
    import torch
    lon = synthetic.lon_axis_values()
    lat = synthetic.lat_axis_values()[:ny]
    g = synthetic.field(lon, lat).astype(dtype)
    g[::37, ::11] = np.nan
    n = 2_000_003
    rng = np.random.default_rng(43)
    q = _adversarial(synthetic.queries(n, 1002), dict(lon=lon, lat=lat), rng)
    oaxes = [orc_best.axis(lon, period=360.0), orc_best.axis(lat)]
    want, _ = orc_best.windowed2d(oaxes, g, q[0], q[1], num_threads=0, stateless=True)
    assert want.dtype == np.dtype(dtype)
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g, resident=True)
    got = pyi.bivariate(grid, q[0], q[1], "bicubic")
    dev = pyi.bivariate(grid, torch.from_numpy(q[0]).cuda(), torch.from_numpy(q[1]).cuda(), "bicubic").cpu().numpy()
    assert got.dtype == np.dtype(dtype) and np.array_equal(got, dev, equal_nan=True)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    m = ~np.isnan(want)
    assert m.mean() > 0.8
    err = np.abs(got[m].astype(np.float64) - want[m].astype(np.float64))
    if bicubic_mode == "exact":
        assert err.max() == 0.0
    else:
        tol = 1e-5 if dtype == "float32" else 1e-12
        assert np.all(err <= tol * np.maximum(np.abs(want[m]), 1.0)), err.max()
        if dtype == "float32":
            assert np.mean(got[m] == want[m]) > 0.99  # float64 arithmetic, one final rounding to float32


@pytest.mark.parametrize("lat_kind", ["regular", "irregular"])
def test_bilinear_at_baseline_size(pyi, synthetic, orc_best, lat_kind, bilinear_mode):
    """BASELINE configs 1 / 5 at 1e7 points (the CPU oracle's budget), both
    value modes: indexes (NaN positions) identical, values bit-identical in
    exact mode and within 1e-12 in tolerance mode."""
    lon = synthetic.lon_axis_values()
    lat = synthetic.lat_axis_values() if lat_kind == "regular" else synthetic.irregular_lat_values()
    g = synthetic.field(lon, lat)
    g[::37, ::11] = np.nan
    n = 10_000_003
    rng = np.random.default_rng(37)
    q = _adversarial(synthetic.queries(n, 1001), dict(lon=lon, lat=lat), rng)
    oaxes = [orc_best.axis(lon, period=360.0), orc_best.axis(lat)]
    want, _ = orc_best.geometric(oaxes, g, q[0], q[1], method=0, num_threads=0)
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g, resident=True)
    got = pyi.bivariate(grid, q[0], q[1])
    assert np.array_equal(np.isnan(got), np.isnan(want))
    m = ~np.isnan(want)
    err = np.abs(got[m] - want[m])
    if bilinear_mode == "exact":
        assert err.max() == 0.0
    else:
        assert np.all(err <= 1e-12 * np.maximum(np.abs(want[m]), 1.0)), err.max()
        # the tolerance kernel differs by rounding only: a few ulp of the cell values
        assert err.max() < 1e-14


def test_every_query_on_a_node_overflows_the_queues(pyi, synthetic, orc_best, monkeypatch):
    """Regridding onto the grid's own lattice: EVERY query sits exactly on a node,
    which the tolerance kernels hand to their redo queues.  4.5e6 points overflow
    the 2^20-entry queues: bilinear must then be recomputed by the bit-identical
    kernel launched behind the queue (device entry, one launch), bicubic by the
    scanning redo pass -- both bit-identical to the oracle."""
    import torch
    monkeypatch.setenv("PYINTERP_B200_BILINEAR", "tolerance")
    lon, lat = synthetic.lon_axis_values(), synthetic.lat_axis_values()
    g = synthetic.field(lon, lat)
    g[::37, ::11] = np.nan
    rng = np.random.default_rng(41)
    n = 4_500_000
    x = lon[rng.integers(0, lon.size, n)] + 360.0 * rng.integers(-1, 2, n)
    y = lat[rng.integers(0, lat.size, n)]
    oaxes = [orc_best.axis(lon, period=360.0), orc_best.axis(lat)]
    grid = pyi.Grid(pyi.Axis(lon, period=360.0), pyi.Axis(lat), g, resident=True)
    xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    want, _ = orc_best.geometric(oaxes, g, x, y, method=0, num_threads=0)
    got = pyi.bivariate(grid, xd, yd).cpu().numpy()
    assert np.array_equal(got, want, equal_nan=True)
    # a mix: half of the points on nodes (overflow), half random (the exact kernel recomputes them all)
    x[::2] = rng.uniform(-180.0, 540.0, x[::2].size)
    y[::2] = rng.uniform(-90.0, 90.0, y[::2].size)
    want, _ = orc_best.geometric(oaxes, g, x, y, method=0, num_threads=0)
    got = pyi.bivariate(grid, torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()).cpu().numpy()
    assert np.array_equal(got, want, equal_nan=True)
    # bicubic: every point tagged -> the redo pass scans
    m = 1_500_000
    xb = lon[rng.integers(0, lon.size, m)]
    yb = lat[rng.integers(3, lat.size - 3, m)]
    want, _ = orc_best.windowed2d(oaxes, g, xb, yb, num_threads=0, stateless=True)
    got = pyi.bivariate(grid, torch.from_numpy(xb).cuda(), torch.from_numpy(yb).cuda(), "bicubic").cpu().numpy()
    assert np.array_equal(got, want, equal_nan=True)
    assert np.isfinite(want).mean() > 0.5
