"""Configuration objects of the interpolation path, mirroring the reference's
``pyinterp.core.config.geometric`` / ``windowed`` builders
(cxx/include/pyinterp/config/{common,geometric,windowed}.hpp,
cxx/src/pybind/config.cpp): immutable, ``with_*`` methods return a copy.

The geometric methods (bilinear / idw / nearest) for 2-4 D and every windowed
fitting model (bicubic, bilinear and the univariate family) for 1-4 D.
"""
from __future__ import annotations

import copy
import types

from . import _abi


class _Common:
    _bounds_error = False
    _num_threads = 0
    _value_mode = _abi.PYI_VALUES_DEFAULT

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

    def with_value_mode(self, mode: str):
        """Not in the reference: how VALUES are computed (``pyi_value_mode`` of
        include/pyinterp_b200.h).  ``"exact"``: bit-identical to the reference;
        ``"tolerance"``: FMA / reciprocal weights, within 1e-12 / 1e-5;
        ``"default"``: tolerance where the axes are well conditioned for it.
        Index selection is bit-exact in every mode."""
        c = self._clone()
        c._value_mode = _abi.VALUE_MODES[mode]
        return c

    @property
    def value_mode(self) -> str:
        return {v: k for k, v in _abi.VALUE_MODES.items()}[self._value_mode]

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
                                    int(self._bounds_error), self._num_threads, self._value_mode)


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


_AXIS_METHODS = {"linear": _abi.PYI_AXIS_LINEAR, "nearest": _abi.PYI_AXIS_NEAREST}
_UNIVARIATE_MODELS = ("akima", "akima_periodic", "c_spline", "c_spline_not_a_knot",
                      "c_spline_periodic", "linear", "polynomial", "steffen")
_SPATIAL_MODELS = _UNIVARIATE_MODELS + ("bilinear", "bicubic")


class AxisConfig:
    """config/common.hpp AxisConfig: linear / nearest along the third and
    fourth axes."""

    def __init__(self, method: int = _abi.PYI_AXIS_LINEAR):
        self.method = method

    @classmethod
    def linear(cls):
        return cls(_abi.PYI_AXIS_LINEAR)

    @classmethod
    def nearest(cls):
        return cls(_abi.PYI_AXIS_NEAREST)


class _Windowed(_Common):
    """config/windowed.hpp: fitting model (factories :404-436), half window
    sizes (default 3 / 3, :244-247), boundary mode, third / fourth axis
    methods (default linear, :392-401)."""

    _ndim = 2
    _models = _SPATIAL_MODELS

    def __init__(self, model: str = "bicubic"):
        if model not in self._models:
            raise ValueError(f"Unknown fitting model: {model}")  # config/windowed.hpp:54-55
        self._method = _abi.SPATIAL_METHODS[model]
        self._half_x = 3
        self._half_y = 3
        self._boundary = _abi.PYI_BOUNDARY_UNDEF
        self._z_method = _abi.PYI_AXIS_LINEAR
        self._u_method = _abi.PYI_AXIS_LINEAR
        self._derivative = 0

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
        """POD of pyi_windowed2d (bicubic / bilinear on 2-D grids)."""
        return _abi.WindowedConfig(self._method, self._half_x, self._half_y, self._boundary,
                                   int(self._bounds_error), self._num_threads, self._value_mode)

    def _pod_nd(self) -> _abi.WindowedNdConfig:
        return _abi.WindowedNdConfig(self._method, self._half_x, self._half_y, self._boundary,
                                     int(self._bounds_error), self._num_threads, self._z_method,
                                     self._u_method, self._derivative, self._value_mode)


def _model_factory(name):
    def factory(cls):
        return cls(name)
    factory.__name__ = name
    return classmethod(factory)


class _WUnivariate(_Windowed):
    """config/windowed.hpp Univariate / UnivariateMethod (:156-222,:313-352):
    univariate models only, default linear, one half window size."""

    _ndim = 1
    _models = _UNIVARIATE_MODELS

    def __init__(self, model: str = "linear"):
        super().__init__(model)
        self._half_y = 1

    def with_half_window_size(self, size: int):
        if int(size) < 1:
            raise ValueError("half_window_size must be at least 1")
        c = self._clone()
        c._half_x = int(size)
        return c


class _WBivariate(_Windowed):
    _ndim = 2


class _WTrivariate(_Windowed):
    _ndim = 3

    def with_third_axis(self, config: AxisConfig):
        c = self._clone()
        c._z_method = config.method
        return c


class _WQuadrivariate(_WTrivariate):
    _ndim = 4

    def with_fourth_axis(self, config: AxisConfig):
        c = self._clone()
        c._u_method = config.method
        return c


for _cls in (_WUnivariate, _WBivariate, _WTrivariate, _WQuadrivariate):
    for _name in _cls._models:
        setattr(_cls, _name, _model_factory(_name))

geometric = types.SimpleNamespace(Bivariate=_GBivariate, Trivariate=_GTrivariate,
                                  Quadrivariate=_GQuadrivariate)
windowed = types.SimpleNamespace(Univariate=_WUnivariate, Bivariate=_WBivariate,
                                 Trivariate=_WTrivariate, Quadrivariate=_WQuadrivariate,
                                 BoundaryConfig=BoundaryConfig, AxisConfig=AxisConfig)


# ---------------------------------------------------------------------------
# pyinterp.core.config.fill (config/fill.hpp, src/pybind/config.cpp:404-515)
# ---------------------------------------------------------------------------
import enum  # noqa: E402


class FirstGuess(enum.IntEnum):
    """config::fill::FirstGuess (config/fill.hpp:18-23)."""
    ZONAL_AVERAGE = _abi.FIRST_GUESS["zonal_average"]
    ZERO = _abi.FIRST_GUESS["zero"]


class _FillBase:
    """FillBase<Derived> (config/fill.hpp:67-160): immutable, ``with_*`` return
    a copy; ``num_threads`` is accepted and ignored (the result of the
    reference does not depend on it either)."""

    _first_guess = FirstGuess.ZONAL_AVERAGE
    _is_periodic = True
    _max_iterations = 500
    _epsilon = 1e-4
    _num_threads = 0

    def _with(self, name, value):
        c = copy.copy(self)
        setattr(c, name, value)
        return c

    def with_first_guess(self, value):
        if not isinstance(value, FirstGuess):
            raise TypeError("with_first_guess(): value must be a FirstGuess")
        return self._with("_first_guess", value)

    def with_is_periodic(self, value: bool):
        return self._with("_is_periodic", bool(value))

    def with_max_iterations(self, value: int):
        value = int(value)
        if not 0 <= value < 2 ** 32:  # uint32_t argument of the reference binding
            raise TypeError("with_max_iterations(): value must fit an uint32")
        return self._with("_max_iterations", value)

    def with_epsilon(self, value: float):
        return self._with("_epsilon", float(value))

    def with_num_threads(self, value: int):
        if int(value) < 0:
            raise TypeError("with_num_threads(): value must be non-negative")
        return self._with("_num_threads", int(value))

    def first_guess(self):
        return self._first_guess

    def is_periodic(self) -> bool:
        return self._is_periodic

    def max_iterations(self) -> int:
        return self._max_iterations

    def epsilon(self) -> float:
        return self._epsilon

    def num_threads(self) -> int:
        return self._num_threads


class GaussSeidel(_FillBase):
    """config::fill::GaussSeidel (config/fill.hpp:255-279)."""

    _max_iterations = 2000
    _relaxation = 1.0

    def with_relaxation(self, value: float):
        return self._with("_relaxation", float(value))

    def relaxation(self) -> float:
        return self._relaxation


class Multigrid(_FillBase):
    """config::fill::Multigrid (config/fill.hpp:282-324)."""

    _max_iterations = 100
    _pre_smooth = 2
    _post_smooth = 2

    def _smooth(self, name, value):
        value = int(value)
        if not 0 <= value < 2 ** 32:  # uint32_t argument of the reference binding
            raise TypeError(f"with{name}(): value must fit an uint32")
        return self._with(name, value)

    def with_pre_smooth(self, value: int):
        return self._smooth("_pre_smooth", value)

    def with_post_smooth(self, value: int):
        return self._smooth("_post_smooth", value)

    def pre_smooth(self) -> int:
        return self._pre_smooth

    def post_smooth(self) -> int:
        return self._post_smooth


class LoessValueType(enum.IntEnum):
    """config::fill::LoessValueType (config/fill.hpp:26-30)."""
    UNDEFINED = 0
    DEFINED = 1
    ALL = 2


class Loess(_FillBase):
    """config::fill::Loess (config/fill.hpp:164-219)."""

    _value_type = LoessValueType.UNDEFINED
    _nx = 3
    _ny = 3

    def _size(self, name, value):
        value = int(value)
        if not 0 <= value < 2 ** 32:  # uint32_t argument of the reference binding
            raise TypeError(f"with{name}(): value must fit an uint32")
        return self._with(name, value)

    def with_value_type(self, value):
        return self._with("_value_type", LoessValueType(value))

    def with_nx(self, value: int):
        return self._size("_nx", value)

    def with_ny(self, value: int):
        return self._size("_ny", value)

    def value_type(self):
        return self._value_type

    def nx(self) -> int:
        return self._nx

    def ny(self) -> int:
        return self._ny


fill = types.SimpleNamespace(FirstGuess=FirstGuess, GaussSeidel=GaussSeidel, Multigrid=Multigrid, Loess=Loess,
                             LoessValueType=LoessValueType)
