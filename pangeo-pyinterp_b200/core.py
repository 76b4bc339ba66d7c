"""Host-side mirror of the reference's ``pyinterp.core`` surface for the
regular-grid interpolation path: ``Axis``, ``TemporalAxis``, ``Grid`` and the
``bivariate`` / ``trivariate`` / ``quadrivariate`` entry points
(cxx/src/pybind/{axis,temporal_axis,grid}.cpp,
cxx/src/pybind/geometric/*_geometric.cpp, windowed/bivariate_windowed.cpp,
cxx/include/pyinterp/pybind/grid_dispatch.hpp).

Same names, argument meaning, result dtypes and error behaviour as the
reference; every index or value is computed by the sm_100a kernels behind the
C-ABI of ``include/pyinterp_b200.h``.  numpy inputs take the pipelined
host-buffer path; torch CUDA tensors stay on the device.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Any

import numpy as np

from . import _abi, config as _config

_P = C.c_void_p


def _lib():
    return _abi.load()


def _is_torch_cuda(a: Any) -> bool:
    return type(a).__module__.startswith("torch") and getattr(a, "is_cuda", False)


def _device_view(a: Any):
    """Zero-copy view of a device array as a CUDA tensor: torch
    CUDA tensors pass through; objects exposing ``__cuda_array_interface__``
    (CuPy, Numba, RAPIDS) or a CUDA ``__dlpack__`` are wrapped without a copy.
    Host objects are returned unchanged."""
    if _is_torch_cuda(a) or isinstance(a, np.ndarray):
        return a
    if hasattr(a, "__cuda_array_interface__"):
        import torch
        return torch.as_tensor(a, device="cuda")
    if hasattr(a, "__dlpack__") and hasattr(a, "__dlpack_device__"):
        try:
            if int(a.__dlpack_device__()[0]) == 2:  # kDLCUDA
                import torch
                return torch.from_dlpack(a)
        except Exception:
            pass
    return a


def _np_ptr(a: np.ndarray):
    return a.ctypes.data_as(_P)


def _host_result(n: int, dtype) -> np.ndarray:
    """Result buffer of a host call.  Large results are allocated page-locked
    (through torch's caching host allocator) so that the device-to-host copy is
    a direct DMA instead of a staged one; the numpy array keeps the allocation
    alive."""
    # (from 2^17 elements: a fresh pageable array of 1 MB or more costs more in first-touch
    # page faults than the whole device round trip; torch caches page-locked blocks by size class)
    if n >= (1 << 17):
        try:
            import torch
            t = torch.empty(n, dtype=torch.float64 if np.dtype(dtype) == np.float64 else torch.float32,
                            pin_memory=True)
            return t.numpy()
        except Exception:  # no torch / no pinned memory: pageable numpy
            pass
    return np.empty(n, dtype=dtype)


# ---------------------------------------------------------------------------
# Axis
# ---------------------------------------------------------------------------
class Axis:
    """``pyinterp.core.Axis(values, epsilon=1e-6, period=None)``
    (cxx/src/pybind/axis.cpp; math/axis.hpp:683-716)."""

    _is_temporal = False

    def __init__(self, values, epsilon: float = 1e-6, period: float | None = None, *, _handle=None):
        if _handle is not None:
            self._h = _handle
            return
        v = np.ascontiguousarray(values, dtype=np.float64)
        if v.ndim != 1:
            raise ValueError("values must be a one-dimensional array")
        st = C.c_int(0)
        self._h = _lib().pyi_axis_create_f64(_np_ptr(v), v.size, float(epsilon), period is not None,
                                             float(period or 0.0), C.byref(st))
        if not self._h:
            raise ValueError(_lib().pyi_last_error().decode())

    def __del__(self):
        h = getattr(self, "_h", None)
        lib = getattr(_abi, "_lib", None)  # module globals are None while the interpreter shuts down
        if h and lib is not None:
            lib.pyi_axis_destroy(h)
            self._h = None

    # -- descriptor ----------------------------------------------------------
    def _desc(self) -> _abi.AxisDesc:
        d = _abi.AxisDesc()
        _lib().pyi_axis_describe(self._h, C.byref(d))
        return d

    def _k(self):
        return "i" if self._is_temporal else "f"

    def _wrap(self, v):
        return float(v)

    def _values(self) -> np.ndarray:
        d = self._desc()
        if d.kind == _abi.PYI_AXIS_REGULAR:
            if self._is_temporal:
                return d.start.i + np.arange(d.size, dtype=np.int64) * d.step.i
            return d.start.f + np.arange(d.size, dtype=np.float64) * d.step.f
        ty = C.c_int64 if self._is_temporal else C.c_double
        return np.ctypeslib.as_array(C.cast(d.points, C.POINTER(ty)), (d.size,)).copy()

    def __len__(self) -> int:
        return int(self._desc().size)

    def __getitem__(self, index):
        vals = self._values()
        if isinstance(index, slice):
            return self._wrap_array(vals[index])
        n = len(vals)
        if index < -n or index >= n:
            raise IndexError("axis index out of range")
        return self._wrap(vals[index])

    def _wrap_array(self, a):
        return a

    def __iter__(self):
        return (self._wrap(v) for v in self._values())

    def front(self):
        return self._wrap(getattr(self._desc().start, self._k()))

    def back(self):
        return self._wrap(self._values()[-1])

    def min_value(self):
        return self._wrap(getattr(self._desc().min_value, self._k()))

    def max_value(self):
        return self._wrap(getattr(self._desc().max_value, self._k()))

    def is_regular(self) -> bool:
        return self._desc().kind == _abi.PYI_AXIS_REGULAR

    def is_ascending(self) -> bool:
        return bool(self._desc().is_ascending)

    def increment(self):
        d = self._desc()
        if d.kind != _abi.PYI_AXIS_REGULAR:
            raise RuntimeError("Increment cannot be retrieved because the axis is not regular.")
        return self._wrap_delta(getattr(d.step, self._k()))

    def _wrap_delta(self, v):
        return float(v)

    @property
    def is_periodic(self) -> bool:
        return bool(self._desc().is_periodic)

    @property
    def period(self):
        d = self._desc()
        return self._wrap_delta(getattr(d.period, self._k())) if d.is_periodic else None

    def flip(self, inplace: bool = False):
        target = self if inplace else self.__copy__()
        _abi.check(_lib().pyi_axis_flip(target._h), "Axis.flip")
        return target

    def __copy__(self):
        st = C.c_int(0)
        d = self._desc()
        h = _lib().pyi_axis_create_from_desc(C.byref(d), C.byref(st))
        if not h:
            raise ValueError(_lib().pyi_last_error().decode())
        new = object.__new__(type(self))
        new._h = h
        if self._is_temporal:
            new._dtype = self._dtype
        return new

    def __eq__(self, other):
        if not isinstance(other, Axis) or self._is_temporal != other._is_temporal:
            return NotImplemented
        a, b = self._desc(), other._desc()
        k = self._k()
        return (a.kind == b.kind and a.size == b.size and getattr(a.period, k) == getattr(b.period, k)
                and np.array_equal(self._values(), other._values()))

    def __ne__(self, other):
        r = self.__eq__(other)
        return r if r is NotImplemented else not r

    # -- lookups (GPU) -------------------------------------------------------
    def _coords(self, coordinates):
        coordinates = _device_view(coordinates)
        if _is_torch_cuda(coordinates):
            return coordinates
        return np.ascontiguousarray(coordinates, dtype=np.float64)

    def _lookup(self, coordinates, pairs: bool, bounded: bool = False):
        lib = _lib()
        c = self._coords(coordinates)
        if _is_torch_cuda(c):
            import torch
            want = torch.int64 if self._is_temporal else torch.float64
            if c.dtype != want:
                raise ValueError(f"device coordinates must be {want}")
            c = c.contiguous()
            out = torch.empty((c.numel(), 2) if pairs else (c.numel(),), dtype=torch.int64, device=c.device)
            stream = torch.cuda.current_stream(c.device).cuda_stream
            with torch.cuda.device(c.device):
                if pairs:
                    st = lib.pyi_axis_find_indexes(self._h, c.numel(), c.data_ptr(), out.data_ptr(),
                                                   _abi.PYI_MEM_DEVICE, stream)
                else:
                    st = lib.pyi_axis_find_index(self._h, c.numel(), c.data_ptr(), int(bounded),
                                                 out.data_ptr(), _abi.PYI_MEM_DEVICE, stream)
            _abi.check(st, "Axis.find_indexes")
            return out
        if c.ndim != 1:
            raise ValueError("coordinates must be a one-dimensional array")
        out = np.empty((c.size, 2) if pairs else (c.size,), dtype=np.int64)
        if pairs:
            st = lib.pyi_axis_find_indexes(self._h, c.size, _np_ptr(c), _np_ptr(out), _abi.PYI_MEM_HOST, None)
        else:
            st = lib.pyi_axis_find_index(self._h, c.size, _np_ptr(c), int(bounded), _np_ptr(out),
                                         _abi.PYI_MEM_HOST, None)
        _abi.check(st, "Axis.find_indexes")
        return out

    def find_indexes(self, coordinates):
        """(n, 2) int64 bracket pairs, (-1, -1) where the coordinate cannot be
        bracketed (cxx/include/pyinterp/pybind/axis.hpp:84-101)."""
        return self._lookup(coordinates, True)

    def find_index(self, coordinates, bounded: bool = False):
        return self._lookup(coordinates, False, bounded)

    def __repr__(self):
        d = self._desc()
        kind = "regular" if d.kind == 0 else "irregular"
        return (f"<pyinterp_b200.{type(self).__name__} {kind} size={d.size} "
                f"[{self.min_value()}, {self.max_value()}] periodic={bool(d.is_periodic)}>")


class TemporalAxis(Axis):
    """``pyinterp.core.TemporalAxis(points, epsilon=None, period=None)``:
    an int64 axis tagged with a numpy datetime64 / timedelta64 dtype
    (math/temporal_axis.hpp, cxx/src/pybind/temporal_axis.cpp:100-112)."""

    _is_temporal = True

    def __init__(self, points, epsilon=None, period=None):
        pts = np.asarray(points)
        if pts.dtype.kind not in "Mm":
            raise ValueError("points must be a numpy datetime64 or timedelta64 array")
        if pts.ndim != 1:
            raise ValueError("points must be a one-dimensional array")
        self._dtype = pts.dtype
        unit = np.datetime_data(pts.dtype)[0]
        tdelta = np.dtype(f"timedelta64[{unit}]")

        def as_int(name, value):
            if value is None:
                return None
            value = np.asarray(value)
            if value.dtype.kind != "m":
                raise ValueError(f"{name} must be a numpy.timedelta64")
            return int(value.astype(tdelta).astype(np.int64))

        eps = as_int("epsilon", epsilon) or 0
        per = as_int("period", period)
        v = np.ascontiguousarray(pts.view(np.int64))
        st = C.c_int(0)
        self._h = _lib().pyi_axis_create_i64(_np_ptr(v), v.size, eps, per is not None, per or 0,
                                             C.byref(st))
        if not self._h:
            raise ValueError(_lib().pyi_last_error().decode())

    @property
    def dtype(self) -> np.dtype:
        return self._dtype

    def _delta_dtype(self):
        return np.dtype(f"timedelta64[{np.datetime_data(self._dtype)[0]}]")

    def _wrap(self, v):
        return np.int64(v).view(self._dtype) if self._dtype.kind == "M" else np.timedelta64(int(v), np.datetime_data(self._dtype)[0])

    def _wrap_array(self, a):
        return np.asarray(a, dtype=np.int64).view(self._dtype)

    def _wrap_delta(self, v):
        return np.timedelta64(int(v), np.datetime_data(self._dtype)[0])

    def cast_to_int64(self, array) -> np.ndarray:
        """TemporalAxis::cast_to_int64 (src/pybind/temporal_axis.cpp:78-98,216-220):
        values converted to the axis resolution, as int64.  A resolution change
        runs dateutils::convert on the device (pyi_datetime_convert); like the
        reference it raises OverflowError when the checked arithmetic overflows
        (NaT included: the reference converts it as the integer it is)."""
        array = _device_view(array)
        if _is_torch_cuda(array):
            return array
        a = np.asarray(array)
        if a.dtype.kind in "iu" and a.dtype.itemsize == 8:
            return np.ascontiguousarray(a.astype(np.int64, copy=False))
        if a.dtype.kind != self._dtype.kind:
            # temporal_axis.cpp:84-89
            raise ValueError(f"array must be of type {'datetime64' if self._dtype.kind == 'M' else 'timedelta64'}, "
                             f"got {a.dtype}")
        src = np.datetime_data(a.dtype)[0]
        dst = np.datetime_data(self._dtype)[0]
        v = np.ascontiguousarray(a.view(np.int64))
        if src == dst:
            return v
        return convert_resolution(v, src, dst)

    def cast_to_temporal_axis(self, array):
        v = self.cast_to_int64(array)
        if _is_torch_cuda(v):  # device int64 values in the axis resolution stay where they are
            return v
        return v.view(self._dtype)

    def _coords(self, coordinates):
        return self.cast_to_int64(coordinates)

    def coordinate_repr(self, value: int) -> str:
        """math/temporal_axis.hpp:80-83: the reference's own text of a coordinate
        (ISO-8601 in the axis resolution, ``"<n> <unit>s"`` for timedelta64),
        formatted by the library (``pyi_format_temporal``)."""
        buf = C.create_string_buffer(128)
        n = _lib().pyi_format_temporal(int(self._dtype.kind == "m"),
                                       _abi.RESOLUTIONS[np.datetime_data(self._dtype)[0]], int(value), buf, 128)
        if n < 0:
            raise ValueError(f"unsupported temporal dtype {self._dtype}")
        return buf.value.decode()


def convert_resolution(values, src: str, dst: str):
    """``dateutils::convert`` (dateutils.hpp:821-935) on the device: int64
    datetime64 / timedelta64 values from numpy unit ``src`` to unit ``dst``.
    numpy int64 arrays or CUDA int64 tensors; returns the same kind."""
    lib = _lib()
    for u in (src, dst):
        if u not in _abi.RESOLUTIONS:
            raise ValueError(f"unknown datetime unit {u!r}")
    oi, op = C.c_int64(-1), C.c_int(0)
    if _is_torch_cuda(values):
        import torch
        if values.dtype != torch.int64:
            raise ValueError("values must be an int64 tensor")
        v = values.contiguous()
        out = torch.empty_like(v)
        with torch.cuda.device(v.device):
            st = lib.pyi_datetime_convert(_abi.RESOLUTIONS[src], _abi.RESOLUTIONS[dst], v.numel(), v.data_ptr(),
                                          out.data_ptr(), _abi.PYI_MEM_DEVICE, C.byref(oi), C.byref(op),
                                          torch.cuda.current_stream(v.device).cuda_stream)
    else:
        v = np.ascontiguousarray(values, dtype=np.int64)
        out = np.empty_like(v)
        st = lib.pyi_datetime_convert(_abi.RESOLUTIONS[src], _abi.RESOLUTIONS[dst], v.size, _np_ptr(v),
                                      _np_ptr(out), _abi.PYI_MEM_HOST, C.byref(oi), C.byref(op), None)
    _abi.check(st, "convert_resolution")
    return out


# ---------------------------------------------------------------------------
# Grid
# ---------------------------------------------------------------------------
class Grid:
    """``pyinterp.core.Grid(x, y[, z[, u]], array)`` for 2-D to 4-D grids
    (cxx/include/pyinterp/pybind/grid.hpp:214-226,443-481).

    The reference borrows the caller's numpy buffer, which may be mutated
    between calls (tests/core/windowed/test_bivariate.py:234-250); the device
    copy is therefore refreshed before every interpolation unless the grid was
    created with ``resident=True`` (benchmarks, multi-GPU shards) or from a
    device tensor.
    """

    def __init__(self, *args, resident: bool = False, pack_cells: bool = True):
        if len(args) < 2 or len(args) > 5:
            raise ValueError("Grid(x[, y[, z[, u]]], array): 1 to 4 axes and an array are required")
        *axes, array = args
        array = _device_view(array)
        for a in axes:
            if not isinstance(a, Axis):
                raise TypeError("axes must be Axis or TemporalAxis instances")
        self._axes = list(axes)
        self._device_tensor = None
        lib = _lib()
        if _is_torch_cuda(array):
            import torch
            t = array.contiguous()
            name = str(t.dtype).replace("torch.", "")
            self._device_tensor = t
            self._array = None
            shape, dtype_name, ptr, mem = tuple(t.shape), name, t.data_ptr(), _abi.PYI_MEM_DEVICE
            self._resident = True
            dev_ctx = torch.cuda.device(t.device)
        else:
            arr = np.asarray(array)
            if arr.dtype.name not in _abi.DTYPES:
                raise ValueError(f"unsupported grid dtype {arr.dtype}")
            self._array = arr
            src = np.ascontiguousarray(arr)
            shape, dtype_name, ptr, mem = src.shape, src.dtype.name, _np_ptr(src), _abi.PYI_MEM_HOST
            self._resident = bool(resident)
            dev_ctx = None
        if dtype_name not in _abi.DTYPES:
            raise ValueError(f"unsupported grid dtype {dtype_name}")
        if len(shape) != len(axes):
            raise ValueError(f"array has {len(shape)} dimensions but {len(axes)} axes were given")
        self._dtype_name = dtype_name
        self._shape = tuple(int(s) for s in shape)
        handles = (_P * len(axes))(*[a._h for a in axes])
        shp = (C.c_int64 * len(axes))(*self._shape)
        st = C.c_int(0)
        flags = _abi.PYI_GRID_PACK_CELLS if pack_cells else _abi.PYI_GRID_DEFAULT
        if dev_ctx is not None:
            with dev_ctx:
                self._h = lib.pyi_grid_create(len(axes), handles, _abi.DTYPES[dtype_name], ptr, shp, mem,
                                              flags, C.byref(st))
        else:
            self._h = lib.pyi_grid_create(len(axes), handles, _abi.DTYPES[dtype_name], ptr, shp, mem,
                                          flags, C.byref(st))
        if not self._h:
            _abi.check(st.value, "Grid")
            raise ValueError(lib.pyi_last_error().decode())
        # a borrowed C-contiguous host array is handed over once (page-locked in place): calls then
        # only say that it may have changed, and the library decides between reading it in place
        # and refreshing the device copies (pyi_grid_track_host)
        self._tracked = False
        if (self._array is not None and not self._resident and src is self._array and src.nbytes > 0
                and os.environ.get("PYINTERP_B200_TRACK_HOST", "1") != "0"):
            self._tracked = lib.pyi_grid_track_host(self._h, ptr) == _abi.PYI_OK

    def __del__(self):
        h = getattr(self, "_h", None)
        lib = getattr(_abi, "_lib", None)  # module globals are None while the interpreter shuts down
        if h and lib is not None:
            lib.pyi_grid_destroy(h)
            self._h = None

    # -- reference surface -----------------------------------------------------
    @property
    def ndim(self) -> int:
        return len(self._axes)

    @property
    def shape(self):
        return self._shape

    @property
    def dtype(self) -> np.dtype:
        return np.dtype(self._dtype_name)

    @property
    def array(self):
        return self._array if self._array is not None else self._device_tensor

    @property
    def x(self):
        return self._axes[0]

    @property
    def y(self):
        if len(self._axes) < 2:
            raise AttributeError("grid has no y axis")
        return self._axes[1]

    @property
    def z(self):
        if len(self._axes) < 3:
            raise AttributeError("grid has no z axis")
        return self._axes[2]

    @property
    def u(self):
        if len(self._axes) < 4:
            raise AttributeError("grid has no u axis")
        return self._axes[3]

    @property
    def has_temporal_axis(self) -> bool:
        return len(self._axes) >= 3 and self._axes[2]._is_temporal

    @property
    def result_dtype(self) -> np.dtype:
        return np.dtype(_abi.result_dtype(self._dtype_name))

    def refresh(self) -> None:
        """Re-upload the (possibly mutated) host array."""
        if self._array is not None:
            src = np.ascontiguousarray(self._array)
            _abi.check(_lib().pyi_grid_update(self._h, _np_ptr(src), _abi.PYI_MEM_HOST), "Grid.refresh")

    def _before_call(self) -> None:
        if self._resident:
            return
        if self._tracked:
            _abi.check(_lib().pyi_grid_mark_stale(self._h), "Grid")
        else:
            self.refresh()


# ---------------------------------------------------------------------------
# entry points
# ---------------------------------------------------------------------------
def _bounds_message(grid: Grid, axis: int, coord) -> str:
    ax = grid._axes[axis]
    labels = "xyzu"
    if ax._is_temporal:
        return (f"{ax.coordinate_repr(int(coord))} is out of bounds for axis {labels[axis]} "
                f"[{ax.coordinate_repr(int(ax._desc().min_value.i))}, ..., "
                f"{ax.coordinate_repr(int(ax._desc().max_value.i))}]")
    buf = C.create_string_buffer(512)
    c = C.c_double(float(coord))
    _lib().pyi_format_bounds_error(grid._h, axis, C.byref(c), buf, 512)
    return buf.value.decode()


def _check_shapes(names, arrays):
    """broadcast::check_eigen_shape (broadcast.hpp:29-37)."""
    n0 = arrays[0].shape[0] if hasattr(arrays[0], "shape") else len(arrays[0])
    for name, a in zip(names[1:], arrays[1:]):
        if a.ndim != 1 or a.shape[0] != n0:
            raise ValueError(
                f"{names[0]}, {name} could not be broadcast together with shape "
                f"({n0},) ({', '.join(str(s) for s in a.shape)},)")
    if arrays[0].ndim != 1:
        raise ValueError(f"{names[0]} must be a one-dimensional array")


def _run(grid: Grid, cfg, coords: list, windowed: bool):
    lib = _lib()
    nd = grid.ndim
    names = ["x", "y", "z", "u"][:nd]
    coords = [_device_view(c) for c in coords]
    on_device = any(_is_torch_cuda(c) for c in coords)
    oi, oa = C.c_int64(-1), C.c_int(-1)
    # 2-D bicubic / bilinear keep the dedicated entry point (and its fast kernel)
    two_d_entry = windowed and nd == 2 and cfg._method <= _abi.PYI_W_BILINEAR
    grid._before_call()
    if on_device:
        import torch
        if not all(_is_torch_cuda(c) for c in coords):
            raise ValueError("either every coordinate array is a CUDA tensor or none is")
        want = [torch.float64, torch.float64, torch.int64 if grid.has_temporal_axis else torch.float64,
                torch.float64][:nd]
        cs = []
        for name, c, w in zip(names, coords, want):
            if c.dtype != w:
                raise ValueError(f"{name} must be a {w} tensor")
            cs.append(c.contiguous())
        _check_shapes(names, cs)
        n = cs[0].numel()
        out = torch.empty(n, dtype=torch.float64 if grid.result_dtype == np.float64 else torch.float32,
                          device=cs[0].device)
        ptrs = [c.data_ptr() for c in cs] + [None] * (4 - nd)
        stream = torch.cuda.current_stream(cs[0].device).cuda_stream
        with torch.cuda.device(cs[0].device):
            if windowed and two_d_entry:
                pod = cfg._pod()
                st = lib.pyi_windowed2d(grid._h, C.byref(pod), n, ptrs[0], ptrs[1], out.data_ptr(),
                                        _abi.PYI_MEM_DEVICE, C.byref(oi), C.byref(oa), stream)
            elif windowed:
                pod = cfg._pod_nd()
                st = lib.pyi_windowed(grid._h, C.byref(pod), n, ptrs[0], ptrs[1], ptrs[2], ptrs[3],
                                      out.data_ptr(), _abi.PYI_MEM_DEVICE, C.byref(oi), C.byref(oa),
                                      stream)
            else:
                pod = cfg._pod()
                st = lib.pyi_geometric(grid._h, C.byref(pod), n, ptrs[0], ptrs[1], ptrs[2], ptrs[3],
                                       out.data_ptr(), _abi.PYI_MEM_DEVICE, C.byref(oi), C.byref(oa), stream)
        if st == _abi.PYI_OOB:
            bad = cs[oa.value][oi.value].item()
    else:
        cs = [np.ascontiguousarray(c, dtype=np.float64) for c in coords[:min(nd, 2)]]
        if nd >= 3:
            if grid.has_temporal_axis:
                cs.append(grid._axes[2].cast_to_int64(coords[2]))  # grid_dispatch.hpp:190-195
            else:
                cs.append(np.ascontiguousarray(coords[2], dtype=np.float64))
        if nd >= 4:
            cs.append(np.ascontiguousarray(coords[3], dtype=np.float64))
        _check_shapes(names, cs)
        n = cs[0].size
        out = _host_result(n, grid.result_dtype)
        ptrs = [_np_ptr(c) for c in cs] + [None] * (4 - nd)
        if windowed and two_d_entry:
            pod = cfg._pod()
            st = lib.pyi_windowed2d(grid._h, C.byref(pod), n, ptrs[0], ptrs[1], _np_ptr(out),
                                    _abi.PYI_MEM_HOST, C.byref(oi), C.byref(oa), None)
        elif windowed:
            pod = cfg._pod_nd()
            st = lib.pyi_windowed(grid._h, C.byref(pod), n, ptrs[0], ptrs[1], ptrs[2], ptrs[3],
                                  _np_ptr(out), _abi.PYI_MEM_HOST, C.byref(oi), C.byref(oa), None)
        else:
            pod = cfg._pod()
            st = lib.pyi_geometric(grid._h, C.byref(pod), n, ptrs[0], ptrs[1], ptrs[2], ptrs[3],
                                   _np_ptr(out), _abi.PYI_MEM_HOST, C.byref(oi), C.byref(oa), None)
        if st == _abi.PYI_OOB:
            bad = cs[oa.value][oi.value]
    if st == _abi.PYI_OOB:
        msg = _bounds_message(grid, oa.value, bad)
        # geometric: std::invalid_argument -> ValueError (pybind/grid.hpp:379-380);
        # windowed: std::out_of_range -> IndexError (pybind/windowed/bivariate.hpp:58-60)
        raise (IndexError if windowed else ValueError)(msg)
    _abi.check(st, "interpolation")
    return out


def _dispatch(grid: Grid, nd: int, name: str, coords: list, cfg):
    if not isinstance(grid, Grid):
        raise TypeError("grid must be a pyinterp_b200 Grid")
    if grid.ndim != nd:
        # grid_dispatch.hpp:138-141
        raise ValueError(f"{name} requires {nd}D grid, got {grid.ndim}D")
    if isinstance(cfg, _config._Geometric):
        return _run(grid, cfg, coords, windowed=False)
    if isinstance(cfg, _config._Windowed):
        if cfg._ndim != nd:
            raise TypeError(f"{name} takes a {nd}-D windowed configuration")
        return _run(grid, cfg, coords, windowed=True)
    raise TypeError(f"unsupported configuration object {type(cfg).__name__}")


def bivariate(grid: Grid, x, y, config=None):
    """``core.bivariate(grid, x, y, config)``
    (geometric/bivariate_geometric.cpp:46-55, windowed/bivariate_windowed.cpp:63-72)."""
    return _dispatch(grid, 2, "bivariate", [x, y], config or _config.geometric.Bivariate())


def trivariate(grid: Grid, x, y, z, config=None):
    """``core.trivariate(grid, x, y, z, config)`` (trivariate_geometric.cpp:49-58)."""
    return _dispatch(grid, 3, "trivariate", [x, y, z], config or _config.geometric.Trivariate())


def quadrivariate(grid: Grid, x, y, z, u, config=None):
    """``core.quadrivariate(grid, x, y, z, u, config)`` (quadrivariate_geometric.cpp:50-60)."""
    return _dispatch(grid, 4, "quadrivariate", [x, y, z, u], config or _config.geometric.Quadrivariate())


def univariate(grid: Grid, x, config=None):
    """``core.univariate(grid, x, config)`` (pybind/windowed/univariate.hpp:122-152)."""
    return _dispatch(grid, 1, "univariate", [x], config or _config.windowed.Univariate())


def univariate_derivative(grid: Grid, x, config=None):
    """``core.univariate_derivative(grid, x, config)`` (pybind/windowed/univariate.hpp:160-190)."""
    cfg = (config or _config.windowed.Univariate())._clone()
    cfg._derivative = 1
    return _dispatch(grid, 1, "univariate_derivative", [x], cfg)
