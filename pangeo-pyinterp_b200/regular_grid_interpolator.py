"""String-driven front end, mirroring ``pyinterp/regular_grid_interpolator.py``
of the reference (:424-502 bivariate, :545-624 trivariate, quadrivariate):
``"bilinear"`` is the geometric bilinear (the geometric list is tested first,
:474), ``"bicubic"`` the windowed one."""
from __future__ import annotations

from . import core
from .config import geometric, windowed

_GEOMETRIC_METHODS = ("bilinear", "idw", "nearest")
_WINDOWED_METHODS = ("akima", "akima_periodic", "bicubic", "bilinear", "c_spline", "c_spline_not_a_knot",
                     "c_spline_periodic", "linear", "polynomial", "steffen")
_BOUNDARY_MAP = {"undef": windowed.BoundaryConfig.undef, "shrink": windowed.BoundaryConfig.shrink}


def _make_geometric_config(method, class_type, *, bounds_error=False, num_threads=0):
    if method not in _GEOMETRIC_METHODS:
        raise ValueError(f"Unknown geometric method: '{method}'. "
                         f"Valid methods: {', '.join(_GEOMETRIC_METHODS)}")
    return getattr(class_type, method)().with_bounds_error(bounds_error).with_num_threads(num_threads)


def _validate_no_windowed_options(method, **options):
    """regular_grid_interpolator.py:225-253 of the reference."""
    given = [k for k, v in options.items() if v is not None]
    if given:
        raise TypeError(f"Options {given} are not valid for geometric method '{method}'. These options "
                        f"are only available for windowed methods: {', '.join(_WINDOWED_METHODS)}")


def _make_windowed_config(method, class_type, *, bounds_error=False, num_threads=0,
                          half_window_size=None, half_window_size_x=None, half_window_size_y=None,
                          boundary_mode=None, third_axis=None, fourth_axis=None):
    """regular_grid_interpolator.py:180-222 of the reference."""
    if method not in _WINDOWED_METHODS:
        raise ValueError(f"Unknown windowed method: '{method}'. "
                         f"Valid methods: {', '.join(_WINDOWED_METHODS)}")
    cfg = getattr(class_type, method)()
    if half_window_size is not None:
        cfg = cfg.with_half_window_size(half_window_size)
    if half_window_size_x is not None:
        cfg = cfg.with_half_window_size_x(half_window_size_x)
    if half_window_size_y is not None:
        cfg = cfg.with_half_window_size_y(half_window_size_y)
    if boundary_mode is not None:
        if boundary_mode not in _BOUNDARY_MAP:
            raise ValueError(f"Unknown boundary mode: '{boundary_mode}'")
        cfg = cfg.with_boundary_mode(_BOUNDARY_MAP[boundary_mode]())
    if third_axis is not None:
        cfg = cfg.with_third_axis(_axis_config(third_axis))
    if fourth_axis is not None:
        cfg = cfg.with_fourth_axis(_axis_config(fourth_axis))
    return cfg.with_bounds_error(bounds_error).with_num_threads(num_threads)


def _axis_config(name):
    if name not in ("linear", "nearest"):
        raise ValueError(f"Unknown axis method: '{name}'")
    return getattr(windowed.AxisConfig, name)()


def univariate(grid, x, method="linear", *, bounds_error=False, num_threads=0, half_window_size=None,
               boundary_mode=None):
    """regular_grid_interpolator.py:277-329: windowed models on a 1-D grid."""
    if not isinstance(method, str):
        return core.univariate(grid, x, method)
    cfg = _make_windowed_config(method, windowed.Univariate, bounds_error=bounds_error,
                                num_threads=num_threads, half_window_size=half_window_size,
                                boundary_mode=boundary_mode)
    return core.univariate(grid, x, cfg)


def univariate_derivative(grid, x, method="linear", *, bounds_error=False, num_threads=0,
                          half_window_size=None, boundary_mode=None):
    """regular_grid_interpolator.py:332-386."""
    if not isinstance(method, str):
        return core.univariate_derivative(grid, x, method)
    cfg = _make_windowed_config(method, windowed.Univariate, bounds_error=bounds_error,
                                num_threads=num_threads, half_window_size=half_window_size,
                                boundary_mode=boundary_mode)
    return core.univariate_derivative(grid, x, cfg)


def bivariate(grid, x, y, method="bilinear", *, bounds_error=False, num_threads=0,
              half_window_size_x=None, half_window_size_y=None, boundary_mode=None):
    if not isinstance(method, str):
        return core.bivariate(grid, x, y, method)
    if method in _GEOMETRIC_METHODS:
        _validate_no_windowed_options(method, half_window_size_x=half_window_size_x,
                                      half_window_size_y=half_window_size_y, boundary_mode=boundary_mode)
        cfg = _make_geometric_config(method, geometric.Bivariate, bounds_error=bounds_error,
                                     num_threads=num_threads)
    else:
        cfg = _make_windowed_config(method, windowed.Bivariate, bounds_error=bounds_error,
                                    num_threads=num_threads, half_window_size_x=half_window_size_x,
                                    half_window_size_y=half_window_size_y, boundary_mode=boundary_mode)
    return core.bivariate(grid, x, y, cfg)


def trivariate(grid, x, y, z, method="bilinear", *, bounds_error=False, num_threads=0,
               half_window_size_x=None, half_window_size_y=None, boundary_mode=None, third_axis=None):
    if not isinstance(method, str):
        return core.trivariate(grid, x, y, z, method)
    if method in _GEOMETRIC_METHODS:
        _validate_no_windowed_options(method, half_window_size_x=half_window_size_x,
                                      half_window_size_y=half_window_size_y, boundary_mode=boundary_mode,
                                      third_axis=third_axis)
        cfg = _make_geometric_config(method, geometric.Trivariate, bounds_error=bounds_error,
                                     num_threads=num_threads)
    else:
        cfg = _make_windowed_config(method, windowed.Trivariate, bounds_error=bounds_error,
                                    num_threads=num_threads, half_window_size_x=half_window_size_x,
                                    half_window_size_y=half_window_size_y, boundary_mode=boundary_mode,
                                    third_axis=third_axis)
    return core.trivariate(grid, x, y, z, cfg)


def quadrivariate(grid, x, y, z, u, method="bilinear", *, bounds_error=False, num_threads=0,
                  half_window_size_x=None, half_window_size_y=None, boundary_mode=None,
                  third_axis=None, fourth_axis=None):
    if not isinstance(method, str):
        return core.quadrivariate(grid, x, y, z, u, method)
    if method in _GEOMETRIC_METHODS:
        _validate_no_windowed_options(method, half_window_size_x=half_window_size_x,
                                      half_window_size_y=half_window_size_y, boundary_mode=boundary_mode,
                                      third_axis=third_axis, fourth_axis=fourth_axis)
        cfg = _make_geometric_config(method, geometric.Quadrivariate, bounds_error=bounds_error,
                                     num_threads=num_threads)
    else:
        cfg = _make_windowed_config(method, windowed.Quadrivariate, bounds_error=bounds_error,
                                    num_threads=num_threads, half_window_size_x=half_window_size_x,
                                    half_window_size_y=half_window_size_y, boundary_mode=boundary_mode,
                                    third_axis=third_axis, fourth_axis=fourth_axis)
    return core.quadrivariate(grid, x, y, z, u, cfg)
