"""ctypes wrapper of the two CPU checkers (orc_abi.h).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs import this module.  ``Oracle("port")`` = oracle/liboracle.so
(plain-C restatement, always available after `make -C oracle`),
``Oracle("reference")`` = oracle/_ref/liboracle_ref.so (the reference's own
headers compiled in place; prebuilt where /root/reference exists).
"""
from __future__ import annotations

import ctypes as C
import importlib
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_abi = importlib.import_module("pangeo-pyinterp_b200._abi")
AxisDesc, GeometricConfig, WindowedConfig = _abi.AxisDesc, _abi.GeometricConfig, _abi.WindowedConfig
WindowedNdConfig = _abi.WindowedNdConfig

PATHS = {
    "port": os.path.join(_HERE, "liboracle.so"),
    "reference": os.path.join(_HERE, "_ref", "liboracle_ref.so"),
}
BOUNDARY = {"undef": 0, "shrink": 1, "expand": 2, "wrap": 3, "sym": 4}

_P = C.c_void_p
_SIG = {
    "orc_kind": (C.c_char_p, []),
    "orc_axis_new_f64": (_P, [_P, C.c_int64, C.c_double, C.c_int, C.c_double, C.c_char_p, C.c_size_t]),
    "orc_axis_new_i64": (_P, [_P, C.c_int64, C.c_int64, C.c_int, C.c_int64, C.c_char_p, C.c_size_t]),
    "orc_axis_free": (None, [_P]),
    "orc_axis_flip": (None, [_P]),
    "orc_axis_describe": (C.c_int, [_P, C.POINTER(AxisDesc)]),
    "orc_axis_find_indexes": (None, [_P, C.c_int64, _P, _P]),
    "orc_axis_find_index": (None, [_P, C.c_int64, _P, C.c_int, _P]),
    "orc_axis_find_window": (C.c_int, [_P, _P, C.c_int, C.c_int, _P, C.POINTER(C.c_int),
                                       C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "orc_geometric": (C.c_int, [C.c_int, C.POINTER(_P), C.c_int, _P, C.POINTER(GeometricConfig),
                                C.c_int64, _P, _P, _P, _P, _P, C.POINTER(C.c_int64),
                                C.POINTER(C.c_int), C.c_char_p, C.c_size_t]),
    "orc_windowed2d": (C.c_int, [C.POINTER(_P), C.c_int, _P, C.POINTER(WindowedConfig), C.c_int,
                                 C.c_int64, _P, _P, _P, C.POINTER(C.c_int64), C.POINTER(C.c_int),
                                 C.c_char_p, C.c_size_t]),
    "orc_windowed_nd": (C.c_int, [C.c_int, C.POINTER(_P), C.c_int, _P, C.POINTER(WindowedNdConfig), C.c_int,
                                  C.c_int64, _P, _P, _P, _P, _P, C.POINTER(C.c_int64), C.POINTER(C.c_int),
                                  C.c_char_p, C.c_size_t]),
    "orc_univariate": (C.c_double, [C.c_int, _P, _P, C.c_int, C.c_double, C.c_int]),
    "orc_datetime_convert": (C.c_int, [C.c_int, C.c_int, C.c_int64, _P, _P, C.c_char_p, C.c_size_t]),
    "orc_temporal_repr": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_char_p, C.c_size_t]),
    "orc_fill_gauss_seidel": (C.c_int, [_P, C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_uint64,
                                        C.c_double, C.c_double, C.c_int, C.POINTER(C.c_uint64),
                                        C.POINTER(C.c_double), C.c_char_p, C.c_size_t]),
    "orc_fill_multigrid": (C.c_int, [_P, C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_uint64,
                                     C.c_double, C.c_uint32, C.c_uint32, C.c_int, C.POINTER(C.c_uint64),
                                     C.POINTER(C.c_double), C.c_char_p, C.c_size_t]),
    "orc_fill_loess": (C.c_int, [_P, _P, C.c_int, C.c_int64, C.c_int64, C.c_uint32, C.c_uint32, C.c_int, C.c_int,
                                 C.c_int, C.c_uint64, C.c_double, C.c_int, C.c_char_p, C.c_size_t]),
    "orc_eval2d_f64": (C.c_double, [C.c_int, C.c_int, C.c_int] + [C.c_double] * 10),
    "orc_eval2d_f32": (C.c_float, [C.c_int, C.c_int, C.c_int] + [C.c_float] * 10),
    "orc_axis_interp_f64": (C.c_double, [C.c_int] + [C.c_double] * 5),
    "orc_axis_interp_i64": (C.c_double, [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_double]),
    "orc_axis_interp_i64_f32": (C.c_float, [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_float, C.c_float]),
    "orc_window_eval": (C.c_double, [C.c_int, _P, C.c_int, _P, C.c_int, _P, C.c_double, C.c_double]),
    "orc_cspline": (C.c_double, [_P, _P, C.c_int, C.c_double, C.c_int]),
}


def available(kind: str) -> bool:
    return os.path.exists(PATHS[kind])


def _ptr(a):
    return a.ctypes.data_as(_P) if a is not None else None


class OrcAxis:
    def __init__(self, orc, handle, is_i64):
        self.orc, self.h, self.is_i64 = orc, handle, is_i64

    def __del__(self):
        if getattr(self, "h", None):
            self.orc.lib.orc_axis_free(self.h)
            self.h = None

    def flip(self):
        self.orc.lib.orc_axis_flip(self.h)
        return self

    def describe(self) -> dict:
        d = AxisDesc()
        self.orc.lib.orc_axis_describe(self.h, C.byref(d))
        k = "i" if self.is_i64 else "f"
        out = {
            "kind": d.kind, "coord_type": d.coord_type, "size": d.size,
            "is_ascending": d.is_ascending, "is_periodic": d.is_periodic,
            "period": getattr(d.period, k), "start": getattr(d.start, k),
            "min_value": getattr(d.min_value, k), "max_value": getattr(d.max_value, k),
        }
        if d.kind == 0:
            out["step"] = getattr(d.step, k)
            if self.is_i64:
                out["step_2"] = d.step_2
            else:
                out["inv_step"] = d.inv_step
        else:
            ty = C.c_int64 if self.is_i64 else C.c_double
            out["points"] = np.ctypeslib.as_array(C.cast(d.points, C.POINTER(ty)), (d.size,)).copy()
        return out

    def raw_desc(self) -> AxisDesc:
        d = AxisDesc()
        self.orc.lib.orc_axis_describe(self.h, C.byref(d))
        return d

    def _coords(self, c):
        return np.ascontiguousarray(c, dtype=np.int64 if self.is_i64 else np.float64)

    def find_indexes(self, coords):
        c = self._coords(coords)
        out = np.empty((c.size, 2), dtype=np.int64)
        self.orc.lib.orc_axis_find_indexes(self.h, c.size, _ptr(c), _ptr(out))
        return out

    def find_index(self, coords, bounded=False):
        c = self._coords(coords)
        out = np.empty(c.size, dtype=np.int64)
        self.orc.lib.orc_axis_find_index(self.h, c.size, _ptr(c), int(bounded), _ptr(out))
        return out

    def find_window(self, coord, half, boundary="undef"):
        c = self._coords([coord])
        idx = np.zeros(2 * half, dtype=np.int64)
        cnt, c0, c1 = C.c_int(0), C.c_int(0), C.c_int(0)
        ok = self.orc.lib.orc_axis_find_window(self.h, _ptr(c), half, BOUNDARY[boundary], _ptr(idx),
                                               C.byref(cnt), C.byref(c0), C.byref(c1))
        if not ok:
            return None
        return idx[: cnt.value].tolist(), (c0.value, c1.value)


class Oracle:
    def __init__(self, kind: str = "port"):
        path = PATHS[kind]
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path}: run `make -C oracle`")
        self.kind = kind
        self.lib = _abi.bind(C.CDLL(path), _SIG)
        assert self.lib.orc_kind().decode() == kind

    def axis(self, values, epsilon=None, period=None) -> OrcAxis:
        v = np.asarray(values)
        err = C.create_string_buffer(512)
        if v.dtype.kind in "iuMm":
            v = np.ascontiguousarray(v.astype(np.int64))
            h = self.lib.orc_axis_new_i64(_ptr(v), v.size, int(epsilon or 0), period is not None,
                                          int(period or 0), err, 512)
            is_i64 = True
        else:
            v = np.ascontiguousarray(v, dtype=np.float64)
            h = self.lib.orc_axis_new_f64(_ptr(v), v.size, 1e-6 if epsilon is None else float(epsilon),
                                          period is not None, float(period or 0.0), err, 512)
            is_i64 = False
        if not h:
            raise ValueError(err.value.decode())
        return OrcAxis(self, h, is_i64)

    def _axes(self, axes):
        arr = (_P * len(axes))(*[a.h for a in axes])
        return arr

    def geometric(self, axes, data, x, y, z=None, u=None, method=0, idw_exp=2, z_method=0,
                  u_method=0, bounds_error=False, num_threads=1):
        data = np.ascontiguousarray(data)
        nd = data.ndim
        assert nd == len(axes)
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.float64)
        if z is not None:
            z = np.ascontiguousarray(z, dtype=np.int64 if axes[2].is_i64 else np.float64)
        if u is not None:
            u = np.ascontiguousarray(u, dtype=np.float64)
        out = np.empty(x.size, dtype=_abi.result_dtype(data.dtype.name))
        cfg = GeometricConfig(method, idw_exp, z_method, u_method, int(bounds_error), num_threads)
        oi, oa = C.c_int64(-1), C.c_int(-1)
        msg = C.create_string_buffer(512)
        st = self.lib.orc_geometric(nd, self._axes(axes), _abi.DTYPES[data.dtype.name], _ptr(data),
                                    C.byref(cfg), x.size, _ptr(x), _ptr(y), _ptr(z), _ptr(u), _ptr(out),
                                    C.byref(oi), C.byref(oa), msg, 512)
        if st == _abi.PYI_OOB:
            return out, (oi.value, oa.value, msg.value.decode())
        assert st == 0, st
        return out, None

    def windowed2d(self, axes, data, x, y, method=0, half_x=3, half_y=3, boundary="undef",
                   bounds_error=False, num_threads=1, stateless=True):
        data = np.ascontiguousarray(data)
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.float64)
        out = np.empty(x.size, dtype=_abi.result_dtype(data.dtype.name))
        cfg = WindowedConfig(method, half_x, half_y, BOUNDARY[boundary], int(bounds_error), num_threads)
        oi, oa = C.c_int64(-1), C.c_int(-1)
        msg = C.create_string_buffer(512)
        st = self.lib.orc_windowed2d(self._axes(axes), _abi.DTYPES[data.dtype.name], _ptr(data),
                                     C.byref(cfg), int(stateless), x.size, _ptr(x), _ptr(y), _ptr(out),
                                     C.byref(oi), C.byref(oa), msg, 512)
        if st == _abi.PYI_OOB:
            return out, (oi.value, oa.value, msg.value.decode())
        assert st == 0, st
        return out, None

    def windowed_nd(self, axes, data, x, y=None, z=None, u=None, method="bicubic", half_x=3, half_y=3,
                    boundary="undef", bounds_error=False, num_threads=1, stateless=True, z_method=0,
                    u_method=0, derivative=False):
        """windowed::pybind::{uni,bi,tri,quadri}variate with any fitting model."""
        data = np.ascontiguousarray(data)
        nd = data.ndim
        assert nd == len(axes)
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = None if y is None else np.ascontiguousarray(y, dtype=np.float64)
        if z is not None:
            z = np.ascontiguousarray(z, dtype=np.int64 if axes[2].is_i64 else np.float64)
        u = None if u is None else np.ascontiguousarray(u, dtype=np.float64)
        out = np.empty(x.size, dtype=_abi.result_dtype(data.dtype.name))
        m = _abi.SPATIAL_METHODS[method] if isinstance(method, str) else int(method)
        cfg = WindowedNdConfig(m, half_x, half_y, BOUNDARY[boundary], int(bounds_error), num_threads,
                               z_method, u_method, int(derivative))
        oi, oa = C.c_int64(-1), C.c_int(-1)
        msg = C.create_string_buffer(512)
        st = self.lib.orc_windowed_nd(nd, self._axes(axes), _abi.DTYPES[data.dtype.name], _ptr(data),
                                      C.byref(cfg), int(stateless), x.size, _ptr(x), _ptr(y), _ptr(z),
                                      _ptr(u), _ptr(out), C.byref(oi), C.byref(oa), msg, 512)
        if st == _abi.PYI_OOB:
            return out, (oi.value, oa.value, msg.value.decode())
        assert st == 0, st
        return out, None

    def univariate(self, method, xa, ya, x, derivative=False):
        xa = np.ascontiguousarray(xa, dtype=np.float64)
        ya = np.ascontiguousarray(ya, dtype=np.float64)
        m = _abi.SPATIAL_METHODS[method] if isinstance(method, str) else int(method)
        return self.lib.orc_univariate(m, _ptr(xa), _ptr(ya), xa.size, x, int(derivative))

    def datetime_convert(self, values, src: str, dst: str):
        """dateutils::convert between numpy units; raises OverflowError like the reference."""
        v = np.ascontiguousarray(values, dtype=np.int64)
        out = np.empty_like(v)
        msg = C.create_string_buffer(256)
        st = self.lib.orc_datetime_convert(_abi.RESOLUTIONS[src], _abi.RESOLUTIONS[dst], v.size, _ptr(v),
                                           _ptr(out), msg, 256)
        if st == _abi.PYI_OVERFLOW:
            raise OverflowError(msg.value.decode())
        assert st == 0, (st, msg.value)
        return out

    def fill_loess(self, grid, nx=3, ny=3, value_type="undefined", first_guess="zonal_average", is_periodic=True,
                   max_iterations=500, epsilon=1e-4, num_threads=1):
        """fill::loess: returns the processed copy of `grid`."""
        dt, rows, cols = self._fill_args(grid)
        out = np.empty_like(grid)
        msg = C.create_string_buffer(256)
        st = self.lib.orc_fill_loess(_ptr(grid), _ptr(out), dt, rows, cols, nx, ny,
                                     {"undefined": 0, "defined": 1, "all": 2}[value_type],
                                     _abi.FIRST_GUESS[first_guess], int(is_periodic), max_iterations, epsilon,
                                     num_threads, msg, 256)
        assert st == 0, msg.value
        return out

    def temporal_repr(self, value: int, unit: str, timedelta: bool = False) -> str:
        """TemporalAxis::coordinate_repr: the reference's text of an int64 coordinate."""
        buf = C.create_string_buffer(128)
        n = self.lib.orc_temporal_repr(int(timedelta), _abi.RESOLUTIONS[unit], int(value), buf, 128)
        assert n >= 0
        return buf.value.decode()

    def _fill_args(self, grid):
        assert isinstance(grid, np.ndarray) and grid.ndim == 2 and grid.flags.c_contiguous
        assert grid.dtype in (np.float32, np.float64)
        return _abi.DTYPES[grid.dtype.name], grid.shape[0], grid.shape[1]

    def fill_gauss_seidel(self, grid, first_guess="zonal_average", is_periodic=True, max_iterations=2000,
                          epsilon=1e-4, relaxation=1.0, num_threads=1):
        """fill::gauss_seidel in place; returns (iterations, max_residual)."""
        dt, nx, ny = self._fill_args(grid)
        it, res, msg = C.c_uint64(0), C.c_double(0), C.create_string_buffer(256)
        st = self.lib.orc_fill_gauss_seidel(_ptr(grid), dt, nx, ny, _abi.FIRST_GUESS[first_guess],
                                            int(is_periodic), max_iterations, epsilon, relaxation,
                                            num_threads, C.byref(it), C.byref(res), msg, 256)
        assert st == 0, (st, msg.value)
        return it.value, res.value

    def fill_multigrid(self, grid, first_guess="zonal_average", is_periodic=True, max_iterations=100,
                       epsilon=1e-4, pre_smooth=2, post_smooth=2, num_threads=1):
        """fill::multigrid in place; returns (iterations, max_residual)."""
        dt, nx, ny = self._fill_args(grid)
        it, res, msg = C.c_uint64(0), C.c_double(0), C.create_string_buffer(256)
        st = self.lib.orc_fill_multigrid(_ptr(grid), dt, nx, ny, _abi.FIRST_GUESS[first_guess],
                                         int(is_periodic), max_iterations, epsilon, pre_smooth,
                                         post_smooth, num_threads, C.byref(it), C.byref(res), msg, 256)
        assert st == 0, (st, msg.value)
        return it.value, res.value

    def window_eval(self, method, xa, ya, za, x, y):
        xa = np.ascontiguousarray(xa, dtype=np.float64)
        ya = np.ascontiguousarray(ya, dtype=np.float64)
        za = np.ascontiguousarray(za, dtype=np.float64)
        assert za.shape == (xa.size, ya.size)
        return self.lib.orc_window_eval(method, _ptr(xa), xa.size, _ptr(ya), ya.size, _ptr(za), x, y)

    def cspline(self, xa, ya, x, derivative=False):
        xa = np.ascontiguousarray(xa, dtype=np.float64)
        ya = np.ascontiguousarray(ya, dtype=np.float64)
        return self.lib.orc_cspline(_ptr(xa), _ptr(ya), xa.size, x, int(derivative))
