"""Configuration objects of the interpolation path, mirroring the reference's
``pyinterp.core.config.geometric`` / ``windowed`` builders
(cxx/include/pyinterp/config/{common,geometric,windowed}.hpp,
cxx/src/pybind/config.cpp): immutable, ``with_*`` methods return a copy.

Only what the hot path consumes is mirrored: the geometric methods
(bilinear / idw / nearest) for 2-4 D and the windowed bivariate methods this
build runs on the GPU (bicubic, windowed bilinear).
"""
from __future__ import annotations

import copy
import types

from . import _abi


class _Common:
    _bounds_error = False
    _num_threads = 0

    def _clone(self):
        return copy.copy(self)

    def with_bounds_error(self, value: bool):
        c = self._clone()
        c._bounds_error = bool(value)
        return c

    def with_num_threads(self, value: int):
        # accepted for API parity (config/common.hpp:66-73); the GPU ignores it
        if int(value) < 0:
            raise ValueError("num_threads must be non-negative")
        c = self._clone()
        c._num_threads = int(value)
        return c

    @property
    def bounds_error(self) -> bool:
        return self._bounds_error

    @property
    def num_threads(self) -> int:
        return self._num_threads


class _Geometric(_Common):
    """config/geometric.hpp: spatial method + idw exponent; the third / fourth
    axis method follows the spatial one (bilinear, idw -> linear; nearest ->
    nearest, config/geometric.hpp:127-139) unless overridden from C++."""

    _ndim = 2

    def __init__(self):
        self._method = _abi.PYI_BILINEAR
        self._exp = 2
        self._z_method = _abi.PYI_AXIS_LINEAR
        self._u_method = _abi.PYI_AXIS_LINEAR

    @classmethod
    def bilinear(cls):
        return cls()

    @classmethod
    def idw(cls, exp: int = 2):
        c = cls()
        c._method, c._exp = _abi.PYI_IDW, int(exp)
        return c

    @classmethod
    def nearest(cls):
        c = cls()
        c._method = _abi.PYI_NEAREST
        c._z_method = c._u_method = _abi.PYI_AXIS_NEAREST
        return c

    # C++-only builders of the reference (config/geometric.hpp:231-236,297-312)
    def with_third_axis(self, method: str):
        c = self._clone()
        c._z_method = {"linear": 0, "nearest": 1}[method]
        return c

    def with_fourth_axis(self, method: str):
        c = self._clone()
        c._u_method = {"linear": 0, "nearest": 1}[method]
        return c

    def _pod(self) -> _abi.GeometricConfig:
        return _abi.GeometricConfig(self._method, self._exp, self._z_method, self._u_method,
                                    int(self._bounds_error), self._num_threads)


class _GBivariate(_Geometric):
    _ndim = 2


class _GTrivariate(_Geometric):
    _ndim = 3


class _GQuadrivariate(_Geometric):
    _ndim = 4


class BoundaryConfig:
    """config/windowed.hpp:75-103"""

    def __init__(self, mode: int = _abi.PYI_BOUNDARY_UNDEF):
        self.mode = mode

    @classmethod
    def undef(cls):
        return cls(_abi.PYI_BOUNDARY_UNDEF)

    @classmethod
    def shrink(cls):
        return cls(_abi.PYI_BOUNDARY_SHRINK)


class _WBivariate(_Common):
    """config/windowed.hpp: method, half window sizes (default 3 / 3,
    :244-247), boundary mode."""

    _GPU_METHODS = {"bicubic": _abi.PYI_W_BICUBIC, "bilinear": _abi.PYI_W_BILINEAR}

    def __init__(self):
        self._method = _abi.PYI_W_BICUBIC
        self._half_x = 3
        self._half_y = 3
        self._boundary = _abi.PYI_BOUNDARY_UNDEF

    @classmethod
    def bicubic(cls):
        return cls()

    @classmethod
    def bilinear(cls):
        c = cls()
        c._method = _abi.PYI_W_BILINEAR
        return c

    def with_boundary_mode(self, config: BoundaryConfig):
        c = self._clone()
        c._boundary = config.mode
        return c

    def with_half_window_size_x(self, size: int):
        if int(size) < 1:
            raise ValueError("half_window_size_x must be at least 1")
        c = self._clone()
        c._half_x = int(size)
        return c

    def with_half_window_size_y(self, size: int):
        if int(size) < 1:
            raise ValueError("half_window_size_y must be at least 1")
        c = self._clone()
        c._half_y = int(size)
        return c

    def _pod(self) -> _abi.WindowedConfig:
        return _abi.WindowedConfig(self._method, self._half_x, self._half_y, self._boundary,
                                   int(self._bounds_error), self._num_threads)


def _unsupported(name):
    def factory(cls):
        raise NotImplementedError(
            f"windowed method '{name}' is outside this build's GPU path "
            "(SURVEY.md 8f: next); available: bicubic, bilinear")
    return classmethod(factory)


for _name in ("akima", "akima_periodic", "c_spline", "c_spline_not_a_knot", "c_spline_periodic",
              "linear", "polynomial", "steffen"):
    setattr(_WBivariate, _name, _unsupported(_name))

geometric = types.SimpleNamespace(Bivariate=_GBivariate, Trivariate=_GTrivariate,
                                  Quadrivariate=_GQuadrivariate)
windowed = types.SimpleNamespace(Bivariate=_WBivariate, BoundaryConfig=BoundaryConfig)
