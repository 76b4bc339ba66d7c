"""The labelled-array front end (pangeo-pyinterp_b200/backends/xarray.py), mirror
of pyinterp/backends/xarray.py: dimension ordering by CF units / dtype, the
coordinate dict protocol and its errors (host logic, no GPU), and the
RegularGridInterpolator on host and device arrays (GPU)."""
import importlib

import numpy as np
import pytest

xb = importlib.import_module("pangeo-pyinterp_b200.backends.xarray")


def _array(dims, temporal=True):
    lon = np.arange(36) * 10.0
    lat = -85.0 + np.arange(18) * 10.0
    time = np.datetime64("2021-01-01", "ns") + np.arange(5) * np.timedelta64(1, "D")
    depth = np.array([0.0, 10.0, 50.0, 200.0])
    coords = {"lon": lon, "lat": lat, "time": time if temporal else np.arange(5.0), "depth": depth}
    shape = [coords[d].size for d in dims]
    data = np.random.default_rng(0).normal(size=shape)
    return xb.DataArray(data, dims, {d: coords[d] for d in dims},
                        units={"lon": "degrees_east", "lat": "degrees_north"})


def test_canonical_dimensions_follow_cf_units_and_dtype():
    """xarray.py:169-252: (longitude, latitude, time, others) whatever the storage order."""
    info = xb._get_canonical_dimensions(_array(("time", "lat", "lon")), ndims=3)
    assert info.dims == ("lon", "lat", "time") and info.has_longitude and info.has_temporal
    assert info.should_be_transposed and info.data_array.shape == (36, 18, 5)
    info = xb._get_canonical_dimensions(_array(("depth", "time", "lat", "lon")), ndims=4)
    assert info.dims == ("lon", "lat", "time", "depth")
    info = xb._get_canonical_dimensions(_array(("lon", "lat")), ndims=2)
    assert info.dims == ("lon", "lat") and not info.should_be_transposed and not info.has_temporal
    # no CF units: storage order is kept
    a = _array(("lat", "lon"))
    for c in a.coords.values():
        c.attrs.clear()
        if hasattr(c, "units"):
            del c.units
    info = xb._get_canonical_dimensions(a, ndims=2)
    assert info.dims == ("lat", "lon") and not info.has_longitude
    with pytest.raises(ValueError, match="number of dimensions of the variable is incorrect"):
        xb._get_canonical_dimensions(_array(("lon", "lat")), ndims=3)
    ax = xb._get_canonical_dimensions(_array(("lon", "lat")), ndims=2).axis(0)
    assert ax.is_periodic and ax.period == 360.0


def test_coords_protocol_and_errors():
    """xarray.py:256-295"""
    dims = ("lon", "lat")
    x, y = np.zeros(3), np.ones(3)
    assert xb._coords({"lat": y, "lon": x}, dims)[0] is x
    with pytest.raises(TypeError, match="coords must be an instance of dict"):
        xb._coords([x, y], dims)
    with pytest.raises(IndexError, match="Number of coordinates"):
        xb._coords({"lon": x}, dims)
    with pytest.raises(IndexError, match="axes not handled by this grid: depth"):
        xb._coords({"lon": x, "depth": y}, dims)
    pyi = importlib.import_module("pangeo-pyinterp_b200")
    t = np.datetime64("2021-01-01", "ns") + np.arange(5) * np.timedelta64(1, "D")
    axis = pyi.TemporalAxis(t)
    # same resolution as the axis: no conversion (a change of resolution runs on the device)
    z = np.array(["2021-01-02", "2021-01-03"], dtype="datetime64[ns]")
    out = xb._coords({"lon": x[:2], "lat": y[:2], "time": z}, ("lon", "lat", "time"), ("time", axis))
    assert out[2].dtype == np.dtype("datetime64[ns]") and out[2][0] == np.datetime64("2021-01-02", "ns")
    with pytest.raises(NotImplementedError, match="Only 2D, 3D, and 4D"):
        xb.RegularGridInterpolator(xb.DataArray(np.zeros(3), ("lon",), {"lon": np.arange(3.0)}))


@pytest.mark.gpu
@pytest.mark.parametrize("dims", [("lon", "lat"), ("lat", "lon"), ("time", "lat", "lon"),
                                  ("depth", "time", "lat", "lon")])
def test_regular_grid_interpolator_matches_core(dims):
    """The interpolator on a labelled array == the core call on the canonical
    (lon, lat, time, depth) grid; device arrays are forwarded without a copy."""
    import torch
    pyi = importlib.import_module("pangeo-pyinterp_b200")
    arr = _array(dims)
    interp = xb.RegularGridInterpolator(arr)
    assert interp.ndim == len(dims)
    canonical = [d for d in ("lon", "lat", "time", "depth") if d in dims]
    order = [dims.index(d) for d in canonical]
    g = np.ascontiguousarray(np.transpose(arr.data, order))
    axes = [pyi.Axis(arr.coords["lon"].values, period=360.0), pyi.Axis(arr.coords["lat"].values)]
    if "time" in dims:
        axes.append(pyi.TemporalAxis(arr.coords["time"].values))
    if "depth" in dims:
        axes.append(pyi.Axis(arr.coords["depth"].values))
    grid = pyi.Grid(*axes, g)
    rng = np.random.default_rng(1)
    n = 5000
    q = {"lon": rng.uniform(-200, 400, n), "lat": rng.uniform(-90, 90, n),
         "time": (np.datetime64("2021-01-01", "s") + rng.integers(0, 4 * 86400, n).astype("timedelta64[s]")),
         "depth": rng.uniform(0, 200, n)}
    q = {d: q[d] for d in dims}
    fn = {2: pyi.bivariate, 3: pyi.trivariate, 4: pyi.quadrivariate}[len(dims)]
    for method in ("bilinear", "nearest", "bicubic"):
        kw = dict(half_window_size_x=2, half_window_size_y=2, boundary_mode="shrink") if method == "bicubic" else {}
        want = fn(grid, *[q[d] for d in canonical], method, **kw)
        got = interp(q, method=method, **kw)
        assert np.array_equal(got, want, equal_nan=True), (dims, method)
    assert np.isfinite(want).mean() > 0.5
    # device-resident grid and coordinates: the result stays on the device
    dev = xb.RegularGridInterpolator(xb.DataArray(torch.from_numpy(arr.data).cuda(), dims, arr.coords))
    dq = {}
    for d in dims:
        v = q[d]
        if d == "time":
            v = v.astype("datetime64[ns]").astype(np.int64)
        dq[d] = torch.from_numpy(np.ascontiguousarray(v)).cuda()
    out = dev(dq, method="bilinear")
    assert out.is_cuda
    want = fn(grid, *[q[d] for d in canonical], "bilinear")
    assert np.array_equal(out.cpu().numpy(), want, equal_nan=True)
    with pytest.raises(IndexError, match="axes not handled by this grid"):
        interp({**{d: q[d] for d in dims[:-1]}, "foo": q[dims[-1]]})
