"""GPU parity of the grid filling: the product's CUDA kernels
behind pyi_fill_gauss_seidel against the reference build (oracle/_ref) or the
pinned C restatement.  Everything is compared BIT FOR BIT: filled grid,
iteration count and final residual."""
import numpy as np
import pytest

import golden_util as gu
from test_gpu_parity import orc_best  # noqa: F401  (fixture)

pytestmark = pytest.mark.gpu

SHAPES = [(40, 30), (37, 23), (1, 9), (9, 1), (2, 2), (3, 3), (5, 64), (64, 5), (1, 1), (129, 67), (300, 2)]


def _same(a, b, what):
    assert a.dtype == b.dtype and a.shape == b.shape, what
    assert np.array_equal(a, b, equal_nan=True), (what, int((a != b).sum()))


@pytest.mark.parametrize("dtype", ["float64", "float32"])
@pytest.mark.parametrize("first_guess", ["zonal_average", "zero"])
@pytest.mark.parametrize("is_periodic", [True, False])
def test_gauss_seidel_bit_exact(pyi, orc_best, dtype, first_guess, is_periodic):
    for k, shape in enumerate(SHAPES):
        g0 = gu.fill_field(shape, dtype, 200 + k, blobs=2)
        for relax, its, eps in ((1.0, 80, 1e-4), (1.6, 9, 0.0)):
            want = g0.copy()
            r_want = orc_best.fill_gauss_seidel(want, first_guess, is_periodic, its, eps, relax, num_threads=2)
            got = g0.copy()
            r_got = pyi.fill.gauss_seidel(got, first_guess=first_guess, is_periodic=is_periodic,
                                          max_iterations=its, epsilon=eps, relaxation=relax)
            assert r_got == r_want, (shape, relax)
            _same(got, want, (shape, relax))


def test_gauss_seidel_full_size_converges(pyi, orc_best):
    """BASELINE grid size (1440 x 720, circular longitude): same iteration count,
    residual and grid as the reference; the filled field is finite everywhere
    and the defined cells are untouched."""
    g0 = gu.fill_field((1440, 720), "float64", 7, fraction=0.05, blobs=12)
    want = g0.copy()
    r_want = orc_best.fill_gauss_seidel(want, "zonal_average", True, 300, 1e-3, 1.0, num_threads=0)
    got = g0.copy()
    r_got = pyi.fill.gauss_seidel(got, max_iterations=300, epsilon=1e-3)
    assert r_got == r_want and r_got[0] > 5
    _same(got, want, "full size")
    assert np.isfinite(got).all()
    m = ~np.isnan(g0)
    assert np.array_equal(got[m], g0[m])


def test_gauss_seidel_config_object_and_device_tensor(pyi, orc_best):
    import torch
    g0 = gu.fill_field((90, 45), "float32", 11)
    cfg = (pyi.config.fill.GaussSeidel().with_max_iterations(50).with_epsilon(1e-5).with_relaxation(1.3)
           .with_first_guess(pyi.config.fill.FirstGuess.ZERO).with_is_periodic(False).with_num_threads(4))
    want = g0.copy()
    r_want = orc_best.fill_gauss_seidel(want, "zero", False, 50, 1e-5, 1.3)
    t = torch.from_numpy(g0.copy()).cuda()
    r_got = pyi.fill.gauss_seidel(t, cfg)
    assert r_got == r_want
    _same(t.cpu().numpy(), want, "device tensor")
    # non-contiguous host view: filled through a dense copy, written back
    big = np.full((90, 90), -1.0, dtype=np.float32)
    view = big[:, ::2]
    view[...] = g0
    assert pyi.fill.gauss_seidel(view, cfg) == r_want
    _same(np.ascontiguousarray(view), want, "strided view")
    assert (big[:, 1::2] == -1.0).all()


def test_gauss_seidel_trivial_and_errors(pyi):
    g = np.arange(12, dtype=np.float64).reshape(3, 4)
    assert pyi.fill.gauss_seidel(g) == (0, 0.0)
    g = np.full((6, 5), np.nan)
    assert pyi.fill.gauss_seidel(g, max_iterations=10) == (1, 0.0) and not g.any()
    g = np.full((6, 5), np.nan)
    assert pyi.fill.gauss_seidel(g, max_iterations=0) == (0, 0.0)
    assert pyi.fill.gauss_seidel(np.empty((0, 4))) == (0, 0.0)
    with pytest.raises(TypeError):
        pyi.fill.gauss_seidel(np.zeros((3, 3), dtype=np.int32))
    with pytest.raises(TypeError):
        pyi.fill.gauss_seidel(np.zeros(5))
    with pytest.raises(KeyError):
        pyi.fill.gauss_seidel(np.zeros((3, 3)), first_guess="mean")
    with pytest.raises(TypeError):
        pyi.fill.gauss_seidel(np.zeros((3, 3)), pyi.windowed.Bivariate.bicubic())


MG_SHAPES = [(40, 30), (37, 23), (77, 41), (9, 9), (5, 5), (4, 50), (64, 5), (6, 7), (1, 1), (129, 67), (300, 2),
             (200, 160)]


@pytest.mark.parametrize("dtype", ["float64", "float32"])
@pytest.mark.parametrize("is_periodic", [True, False])
def test_multigrid_bit_exact(pyi, orc_best, dtype, is_periodic):
    for k, shape in enumerate(MG_SHAPES):
        g0 = gu.fill_field(shape, dtype, 400 + k, blobs=2)
        for fg, pre, post, its, eps in (("zonal_average", 2, 2, 8, 1e-6), ("zero", 1, 3, 3, 0.0), ("zero", 0, 1, 2, 0.0)):
            want = g0.copy()
            r_want = orc_best.fill_multigrid(want, fg, is_periodic, its, eps, pre, post, num_threads=2)
            got = g0.copy()
            r_got = pyi.fill.multigrid(got, first_guess=fg, is_periodic=is_periodic, max_iterations=its,
                                       epsilon=eps, pre_smooth=pre, post_smooth=post)
            assert r_got == r_want, (shape, fg, pre, post)
            _same(got, want, (shape, fg, pre, post))


def test_multigrid_full_size(pyi, orc_best):
    """BASELINE grid size: multi-block sweeps on the fine levels, single-block
    smoothing below, nine levels."""
    g0 = gu.fill_field((1440, 720), "float64", 8, fraction=0.05, blobs=12)
    want = g0.copy()
    r_want = orc_best.fill_multigrid(want, "zonal_average", True, 5, 1e-9, 2, 2, num_threads=0)
    got = g0.copy()
    r_got = pyi.fill.multigrid(got, max_iterations=5, epsilon=1e-9)
    assert r_got == r_want
    _same(got, want, "full size")
    m = ~np.isnan(g0)
    assert np.isfinite(got).all() and np.array_equal(got[m], g0[m])


def test_multigrid_config_device_tensor_and_errors(pyi, orc_best):
    import torch
    g0 = gu.fill_field((90, 45), "float32", 12)
    cfg = (pyi.config.fill.Multigrid().with_max_iterations(4).with_epsilon(1e-7).with_pre_smooth(3)
           .with_post_smooth(1).with_first_guess(pyi.config.fill.FirstGuess.ZERO).with_is_periodic(False))
    want = g0.copy()
    r_want = orc_best.fill_multigrid(want, "zero", False, 4, 1e-7, 3, 1)
    t = torch.from_numpy(g0.copy()).cuda()
    assert pyi.fill.multigrid(t, cfg) == r_want
    _same(t.cpu().numpy(), want, "device tensor")
    assert pyi.fill.multigrid(np.arange(12.0).reshape(3, 4)) == (0, 0.0)
    g = np.full((9, 9), np.nan, dtype=np.float32)
    its, res = pyi.fill.multigrid(g, max_iterations=0)
    assert its == 0 and res == float(np.finfo(np.float32).max)  # multi_grid.hpp:311
    with pytest.raises(TypeError):
        pyi.fill.multigrid(np.zeros((3, 3)), pyi.config.fill.GaussSeidel())


@pytest.mark.parametrize("dtype", ["float64", "float32"])
def test_loess_bit_identical(pyi, dtype):
    """fill::loess on the device against the CPU oracle: single pass and
    iterated, the three value types, periodic or mirrored x, both first guesses,
    host arrays and device tensors, at toy sizes and on the 0.25 degree grid."""
    import orc
    import torch
    import golden_util as gu
    o = orc.Oracle("reference") if orc.available("reference") else orc.Oracle("port")
    for k, shape in enumerate([(40, 30), (37, 23), (1, 9), (9, 1), (2, 2), (64, 5), (1, 1), (130, 67)]):
        g0 = gu.fill_field(shape, dtype, 700 + k, blobs=2)
        for vt in ("undefined", "defined", "all"):
            for fg, per, its, eps, hx, hy in (("zonal_average", True, 1, 1e-4, 3, 3), ("zero", False, 4, 0.0, 2, 5),
                                              ("zonal_average", False, 30, 1e-3, 4, 1), ("zonal_average", True, 0, 1e-4, 3, 3)):
                want = o.fill_loess(g0, hx, hy, vt, fg, per, its, eps)
                kw = dict(nx=hx, ny=hy, value_type=vt, first_guess=fg, is_periodic=per, max_iterations=its, epsilon=eps)
                keep = g0.copy()
                got = pyi.fill.loess(g0, **kw)
                assert np.array_equal(g0, keep, equal_nan=True)  # the input is left untouched
                assert got.dtype == g0.dtype and np.array_equal(got, want, equal_nan=True), (shape, vt, fg, per, its)
        dev = pyi.fill.loess(torch.from_numpy(g0).cuda(), nx=3, ny=3, max_iterations=5)
        assert dev.is_cuda and np.array_equal(dev.cpu().numpy(), o.fill_loess(g0, 3, 3, max_iterations=5), equal_nan=True)
    s = pyi.synthetic
    g = s.field(s.lon_axis_values(), s.lat_axis_values()).astype(dtype)
    rng = np.random.default_rng(4)
    g[rng.random(g.shape) < 0.2] = np.nan
    g[200:260, 300:380] = np.nan
    for its in (1, 4):
        want = o.fill_loess(g, 5, 5, "undefined", "zonal_average", True, its, 1e-6, num_threads=0)
        got = pyi.fill.loess(g, nx=5, ny=5, max_iterations=its, epsilon=1e-6)
        assert np.array_equal(got, want, equal_nan=True)
    with pytest.raises(ValueError, match="at least 1"):
        pyi.fill.loess(g, nx=0)
    with pytest.raises(TypeError):
        pyi.fill.loess(g.astype(np.int32))
