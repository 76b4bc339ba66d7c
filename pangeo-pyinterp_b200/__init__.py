"""pyinterp_b200 -- B200-native (sm_100a) drop-in for ONE hot path of
CNES/pangeo-pyinterp: n-variate regular-grid interpolation
(``bivariate`` / ``trivariate`` / ``quadrivariate`` over ``Grid2D/3D/4D`` with
regular, irregular, circular-longitude and temporal axes).

Everything is computed by hand-written CUDA kernels behind the C-ABI of
``include/pyinterp_b200.h`` (``csrc/libpyinterp_b200.so``).  There is no CPU
fallback: importing works without a GPU (axes are host objects), every lookup
or interpolation raises without one.
"""
from . import _abi, backends, config, core, dist, fill, synthetic
from .config import geometric, windowed
from .core import Axis, Grid, TemporalAxis
from .regular_grid_interpolator import (bivariate, quadrivariate, trivariate, univariate,
                                        univariate_derivative)

Grid1D = Grid2D = Grid3D = Grid4D = Grid  # pyinterp/__init__.py:92-95

__all__ = ["Axis", "TemporalAxis", "Grid", "Grid1D", "Grid2D", "Grid3D", "Grid4D", "bivariate", "trivariate",
           "quadrivariate", "univariate", "univariate_derivative", "geometric", "windowed", "config", "core", "dist", "fill", "synthetic", "backends", "device_count",
           "launch_count"]


def device_count() -> int:
    return _abi.load().pyi_device_count()


def launch_count() -> int:
    """Number of CUDA kernels this library has launched in this process."""
    return int(_abi.load().pyi_launch_count())
