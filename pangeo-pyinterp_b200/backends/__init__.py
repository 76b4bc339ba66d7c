"""Front ends that build grids from labelled arrays (``pyinterp.backends``)."""
from . import xarray  # noqa: F401
