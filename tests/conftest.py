import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port():
    import orc
    return orc.Oracle("port")


@pytest.fixture(scope="session")
def ref():
    import orc
    if not orc.available("reference"):
        pytest.skip("oracle/_ref/liboracle_ref.so not built (needs /root/reference)")
    return orc.Oracle("reference")


@pytest.fixture(scope="session")
def synthetic():
    return importlib.import_module("pangeo-pyinterp_b200.synthetic")


@pytest.fixture(scope="session")
def pyi():
    """The product package (needs the CUDA library; GPU tests only)."""
    return importlib.import_module("pangeo-pyinterp_b200")


@pytest.fixture(params=["weights", "exact"])
def bicubic_mode(request, monkeypatch):
    """The two 6 x 6 bicubic fast paths: the default precomputed-weights kernel
    (pyi_bicubicw.cuh; values within the north-star tolerance) and the
    bit-identical one (PYINTERP_B200_BICUBIC=exact, pyi_fastwin*.cuh)."""
    if request.param == "exact":
        monkeypatch.setenv("PYINTERP_B200_BICUBIC", "exact")
    else:
        monkeypatch.delenv("PYINTERP_B200_BICUBIC", raising=False)
    return request.param


@pytest.fixture(params=["tolerance", "exact"])
def bilinear_mode(request, monkeypatch):
    """The two value modes of geometric bilinear on float64-result 2-D grids:
    the default tolerance kernel (pyi_fast2d_tol.cuh: verified brackets, FMA /
    reciprocal weights, values within the north-star 1e-12) and the
    bit-identical one (PYINTERP_B200_BILINEAR=exact, pyi_fast2d.cuh).  Every
    other kernel is bit-identical in both modes."""
    # "tolerance" in the environment also lifts the size threshold below which
    # the default mode picks the (launch-cheaper) bit-identical kernel
    monkeypatch.setenv("PYINTERP_B200_BILINEAR", request.param)
    return request.param
