"""Labelled-array front end, mirroring ``pyinterp/backends/xarray.py`` of the
reference: ``Grid2D`` / ``Grid3D`` / ``Grid4D`` and ``RegularGridInterpolator``
built from a data array whose dimensions are named, with longitude / latitude
found through their CF ``units`` and a temporal axis through its dtype; the
interpolation is called with a dict of coordinates
(``_coords``, xarray.py:256-295; ``RegularGridInterpolator``, :582-809).

Nothing here needs the xarray package (absent from this image): any object
with ``dims``, ``coords[name].values`` (+ an optional ``units`` attribute or
``attrs["units"]``), ``values`` / ``data`` and ``transpose(*dims)`` works -- an
``xarray.DataArray`` does -- and ``DataArray`` below is a minimal stand-in.

Zero-copy: the grid values may live on the device (a torch
CUDA tensor or any ``__cuda_array_interface__`` / CUDA DLPack object) and so
may the query coordinates; both are forwarded untouched to the kernels and the
result stays on the device, so the path is not PCIe-bound.
"""
from __future__ import annotations

import dataclasses
from typing import Any, Hashable

import numpy as np

from .. import core
from ..regular_grid_interpolator import bivariate, quadrivariate, trivariate

__all__ = ["DataArray", "Grid2D", "Grid3D", "Grid4D", "RegularGridInterpolator"]

TWO_DIMENSIONS, THREE_DIMENSIONS, FOUR_DIMENSIONS = 2, 3, 4
LONGITUDE_AXIS_INDEX, TEMPORAL_AXIS_INDEX = 0, 2

#: pyinterp/cf.py:28-66
LATITUDE_UNITS = {"degrees_north", "degree_north", "degree_N", "degrees_N", "degreeN", "degreesN"}
LONGITUDE_UNITS = {"degrees_east", "degree_east", "degree_E", "degrees_E", "degreeE", "degreesE"}


class _Coord:
    def __init__(self, values, units=None):
        self.values = np.asarray(values)
        self.attrs = {} if units is None else {"units": units}
        if units is not None:
            self.units = units

    @property
    def dtype(self):
        return self.values.dtype


class DataArray:
    """Minimal labelled array: ``DataArray(data, dims, coords, units=None)``.
    ``data`` is a numpy array or a device array; ``coords`` maps each dimension
    name to its 1-D coordinate values; ``units`` maps names to CF units."""

    def __init__(self, data, dims, coords, units=None):
        self.data = data
        self.dims = tuple(dims)
        units = units or {}
        self.coords = {k: v if isinstance(v, _Coord) else _Coord(v, units.get(k)) for k, v in coords.items()}
        if len(self.dims) != len(self.shape):
            raise ValueError("dims does not match the number of array dimensions")

    @property
    def shape(self):
        return tuple(self.data.shape)

    @property
    def ndim(self):
        return len(self.shape)

    @property
    def values(self):
        return self.data

    def transpose(self, *dims):
        order = [self.dims.index(d) for d in dims]
        data = self.data
        if core._is_torch_cuda(core._device_view(data)):
            data = core._device_view(data).permute(*order).contiguous()
        else:
            data = np.transpose(np.asarray(data), order)
        return DataArray(data, dims, self.coords)


def _units(coord) -> str | None:
    u = getattr(coord, "units", None)
    if u is None:
        u = getattr(coord, "attrs", {}).get("units")
    return u


class AxisIdentifier:
    """xarray.py:52-98: the dimensions holding longitudes / latitudes (CF units)."""

    def __init__(self, data_array):
        self.data_array = data_array

    def _axis(self, units) -> str | None:
        for name, coord in self.data_array.coords.items():
            if _units(coord) in units:
                return str(name)
        return None

    def longitude(self):
        return self._axis(LONGITUDE_UNITS)

    def latitude(self):
        return self._axis(LATITUDE_UNITS)


def _identify_temporal_axis(data_array, dims):
    """xarray.py:101-127: the first dimension whose coordinate is datetime64 / timedelta64."""
    for dim in dims:
        if dim in data_array.coords:
            dt = np.asarray(data_array.coords[dim].values).dtype
            if np.issubdtype(dt, np.datetime64) or np.issubdtype(dt, np.timedelta64):
                return dim
    return None


@dataclasses.dataclass(frozen=True)
class _DimInfo:
    _data_array: Any
    dims: tuple
    has_longitude: bool = False
    has_temporal: bool = False
    should_be_transposed: bool = False

    def axis(self, index: int):
        values = np.asarray(self.data_array.coords[self.dims[index]].values)
        if index == LONGITUDE_AXIS_INDEX and self.has_longitude:
            return core.Axis(values, period=360.0)
        if index == TEMPORAL_AXIS_INDEX and self.has_temporal:
            return core.TemporalAxis(values)
        return core.Axis(values)

    @property
    def data_array(self):
        if self.should_be_transposed:
            return self._data_array.transpose(*self.dims)
        return self._data_array


def _get_canonical_dimensions(data_array, ndims: int = 2) -> _DimInfo:
    """xarray.py:169-252: dimensions in the order (longitude, latitude, time, others)."""
    if data_array.ndim != ndims:
        raise ValueError("The number of dimensions of the variable is incorrect. "
                         f"Expected {ndims}, found {data_array.ndim}.")
    current_dims = list(data_array.dims)
    ident = AxisIdentifier(data_array)
    lon_dim, lat_dim = ident.longitude(), ident.latitude()
    time_dim = _identify_temporal_axis(data_array, current_dims)
    has_longitude, has_temporal = lon_dim is not None, False
    special = {lon_dim, lat_dim, time_dim}
    remaining = [d for d in current_dims if d not in special]
    final: list = []
    if has_longitude:
        final.append(lon_dim)
    elif remaining:
        final.append(remaining.pop(0))
    if lat_dim is not None:
        final.append(lat_dim)
    elif remaining:
        final.append(remaining.pop(0))
    if ndims >= THREE_DIMENSIONS:
        if time_dim is not None:
            final.append(time_dim)
            has_temporal = True
        elif remaining:
            final.append(remaining.pop(0))
    final.extend(remaining)
    assert len(final) == ndims
    return _DimInfo(data_array, tuple(final), has_longitude, has_temporal, tuple(current_dims) != tuple(final))


def _coords(coords: dict, dims: tuple, datetime64=None) -> tuple:
    """xarray.py:256-295: the positional coordinate arrays of an interpolation call."""
    if not isinstance(coords, dict):
        raise TypeError("coords must be an instance of dict")
    if len(coords) != len(dims):
        raise IndexError(f"Number of coordinates ({len(coords)}) doesn't match "
                         f"number of dimensions ({len(dims)})")
    unknown = set(coords) - set(dims)
    if unknown:
        raise IndexError("axes not handled by this grid: " + ", ".join(str(item) for item in unknown))
    if datetime64 is not None:
        temporal_dim, temporal_axis = datetime64
        return tuple(coords[d] if d != temporal_dim else temporal_axis.cast_to_temporal_axis(coords[d])
                     for d in dims)
    return tuple(coords[d] for d in dims)


def _grid_values(data_array):
    v = getattr(data_array, "values", None)
    data = getattr(data_array, "data", None)
    # device-backed arrays keep their device buffer (``values`` would copy to the host)
    if data is not None and core._is_torch_cuda(core._device_view(data)):
        return data
    return v if v is not None else data


class _GridHolder:
    """xarray.py:298-338"""

    _NDIMS = 0

    def __init__(self, data_array, **grid_options):
        object.__setattr__(self, "_data_array", data_array)
        info = _get_canonical_dimensions(data_array, ndims=self._NDIMS)
        axes = [info.axis(k) for k in range(self._NDIMS)]
        self._instance = core.Grid(*axes, _grid_values(info.data_array), **grid_options)
        self._dims = info.dims
        self._datetime64 = (self._dims[2], self._instance.z) if self._instance.has_temporal_axis else None

    def __getattr__(self, name: str):
        return getattr(self.__dict__["_instance"], name)

    def __repr__(self) -> str:
        return repr(self._instance)

    def __reduce__(self):
        return (self.__class__, (self._data_array,))


class Grid2D(_GridHolder):
    """xarray.py:341-407"""
    _NDIMS = TWO_DIMENSIONS

    def bivariate(self, coords: dict, method="bilinear", **kwargs):
        x, y = _coords(coords, self._dims, self._datetime64)
        return bivariate(self._instance, x, y, method=method, **kwargs)


class Grid3D(_GridHolder):
    """xarray.py:410-482"""
    _NDIMS = THREE_DIMENSIONS

    def trivariate(self, coords: dict, method="bilinear", **kwargs):
        x, y, z = _coords(coords, self._dims, self._datetime64)
        return trivariate(self._instance, x, y, z, method=method, **kwargs)


class Grid4D(_GridHolder):
    """xarray.py:485-579"""
    _NDIMS = FOUR_DIMENSIONS

    def quadrivariate(self, coords: dict, method="bilinear", **kwargs):
        x, y, z, u = _coords(coords, self._dims, self._datetime64)
        return quadrivariate(self._instance, x, y, z, u, method=method, **kwargs)


class RegularGridInterpolator:
    """xarray.py:582-809: interpolation on a 2-D, 3-D or 4-D labelled grid,
    called with a dict of coordinates."""

    def __init__(self, array, **grid_options):
        ndim = len(array.shape)
        if ndim == TWO_DIMENSIONS:
            self._grid = Grid2D(array, **grid_options)
            self._interp = self._grid.bivariate
        elif ndim == THREE_DIMENSIONS:
            self._grid = Grid3D(array, **grid_options)
            self._interp = self._grid.trivariate
        elif ndim == FOUR_DIMENSIONS:
            self._grid = Grid4D(array, **grid_options)
            self._interp = self._grid.quadrivariate
        else:
            raise NotImplementedError("Only 2D, 3D, and 4D grids can be interpolated. "
                                      f"Got {ndim}D grid.")

    @property
    def ndim(self) -> int:
        return len(self._grid._dims)

    @property
    def grid(self):
        return self._grid

    def __call__(self, coords: dict, method="bilinear", **kwargs):
        return self._interp(coords, method=method, **kwargs)
