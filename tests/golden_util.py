"""Shared helpers of the golden-vector tests (test infrastructure).

A "case" is a small, self-contained problem of the hot path: axis values,
the grid array, the query coordinates and the configuration.  The fixtures
tests/golden/ref_vectors_<case>.npz additionally hold the output the
reference's own headers (oracle/_ref) produced for it in the build container;
tests/golden/make_golden.py writes them."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
DAY = 86_400_000_000_000
T0 = int(np.datetime64("2021-01-01", "ns").astype(np.int64))
METHOD = {"bilinear": 0, "idw": 1, "nearest": 2}
WMETHOD = {"bicubic": 0, "bilinear": 1}


# ---- known-answer helpers ---------------------------------------------------
def arange(spec):
    start, step, n = spec
    return start + np.arange(n) * step


def axis_values(e):
    if "values" in e:
        return np.array(e["values"], dtype=np.float64)
    if "linspace" in e:
        a, b, n = e["linspace"]
        return np.linspace(a, b, n)
    return arange(e["arange"])


def axis_values_i64(e):
    return e["start"] + np.arange(e["size"], dtype=np.int64) * e["step"]


# ---- seeded cases -----------------------------------------------------------
def _field(shape, dtype, seed, nan_fraction=0.0):
    rng = np.random.Generator(np.random.PCG64(seed))
    idx = np.indices(shape).astype(np.float64)
    g = np.sin(idx[0] * 0.37)
    if len(shape) >= 2:
        g = g * np.cos(idx[1] * 0.23)
    for k in range(2, len(shape)):
        g = g * (1.0 + 0.1 * idx[k])
    g = g + 0.05 * rng.standard_normal(shape)
    dt = np.dtype(dtype)
    if dt.kind in "iu":
        info = np.iinfo(dt)
        scale = 100.0 if dt.itemsize == 1 else 1000.0
        if dt.kind == "u":
            g = g + 1.5
        g = np.clip(np.round(g * scale), info.min, info.max)
    g = g.astype(dt)
    if nan_fraction and dt.kind == "f":
        mask = rng.uniform(size=shape) < nan_fraction
        g[mask] = np.nan
    return np.ascontiguousarray(g)


def _coords(rng, lo, hi, n, nodes=None, integer=False):
    span = hi - lo
    if integer:
        c = rng.integers(int(lo - 0.1 * span), int(hi + 0.1 * span), n, dtype=np.int64)
        if nodes is not None:
            k = n // 8
            c[:k] = rng.choice(nodes, k)
            c[k] = np.iinfo(np.int64).min  # NaT
        return c
    c = rng.uniform(lo - 0.1 * span, hi + 0.1 * span, n)
    if nodes is not None:
        k = n // 8
        c[:k] = rng.choice(nodes, k)  # exact node hits
        c[k:2 * k] = np.nextafter(rng.choice(nodes, k), np.inf)
        c[2 * k] = np.nan
    return c


LON72 = np.arange(72) * 5.0
LAT36 = -87.5 + np.arange(36) * 5.0
LAT_IRR = np.degrees(np.arcsin(np.linspace(-0.999, 0.999, 36)))
TIME10 = T0 + np.arange(10, dtype=np.int64) * DAY
DEPTH8 = np.array([0.0, 5.0, 12.5, 30.0, 75.0, 200.0, 1000.0, 5500.0])

# name -> dict(kind, axes=[(values, period, flip)], dtype, n, seed, cfg...)
CASES = {
    "geo2d_f64_bilinear": dict(kind="geo", axes=[(LON72, 360.0, False), (LAT36, None, False)], dtype="float64",
                               method="bilinear", n=3000, seed=1),
    "geo2d_f32_bilinear": dict(kind="geo", axes=[(LON72, 360.0, False), (LAT36, None, False)], dtype="float32",
                               method="bilinear", n=3000, seed=2),
    "geo2d_f64_irregular_lat": dict(kind="geo", axes=[(LON72, 360.0, False), (LAT_IRR, None, False)],
                                    dtype="float64", method="bilinear", n=3000, seed=3),
    "geo2d_f64_descending": dict(kind="geo", axes=[(LON72, 360.0, True), (LAT36, None, True)], dtype="float64",
                                 method="bilinear", n=3000, seed=4),
    "geo2d_i16_bilinear": dict(kind="geo", axes=[(LON72 - 180.0, 360.0, False), (LAT36, None, False)],
                               dtype="int16", method="bilinear", n=2000, seed=5),
    "geo2d_f64_idw": dict(kind="geo", axes=[(LON72, 360.0, False), (LAT36, None, False)], dtype="float64",
                          method="idw", n=2000, seed=6),
    "geo2d_f32_nearest": dict(kind="geo", axes=[(LON72, 360.0, False), (LAT36, None, False)], dtype="float32",
                              method="nearest", n=2000, seed=7),
    "geo2d_f64_nan_grid": dict(kind="geo", axes=[(LON72, None, False), (LAT36, None, False)], dtype="float64",
                               method="bilinear", n=2000, seed=8, nan_fraction=0.05),
    "geo3d_f32_time": dict(kind="geo", axes=[(LON72, 360.0, False), (LAT36, None, False), (TIME10, None, False)],
                           dtype="float32", method="bilinear", n=3000, seed=9),
    "geo3d_f64_nearest_z": dict(kind="geo", axes=[(LON72, 360.0, False), (LAT36, None, False),
                                                  (np.arange(6) * 2.5, None, False)],
                                dtype="float64", method="nearest", n=2000, seed=10),
    "geo4d_f32_depth": dict(kind="geo", axes=[(LON72[::2], 360.0, False), (LAT36[::2], None, False),
                                              (TIME10[:6], None, False), (DEPTH8, None, False)],
                            dtype="float32", method="bilinear", n=3000, seed=11),
    "geo4d_f64_idw": dict(kind="geo", axes=[(LON72[::2], 360.0, False), (LAT36[::2], None, False),
                                            (np.arange(5) * 1.0, None, False), (DEPTH8, None, True)],
                          dtype="float64", method="idw", n=1500, seed=12),
    "bicubic_undef": dict(kind="win", axes=[(LON72, 360.0, False), (LAT36, None, False)], dtype="float64",
                          method="bicubic", half=(3, 3), boundary="undef", n=3000, seed=13),
    "bicubic_shrink_f32": dict(kind="win", axes=[(LON72, None, False), (LAT_IRR, None, False)], dtype="float32",
                               method="bicubic", half=(3, 2), boundary="shrink", n=3000, seed=14),
    "bicubic_nan_grid": dict(kind="win", axes=[(LON72, 360.0, False), (LAT36, None, False)], dtype="float64",
                             method="bicubic", half=(2, 4), boundary="undef", n=2000, seed=15, nan_fraction=0.01),
    "wbilinear_undef": dict(kind="win", axes=[(LON72, 360.0, False), (LAT36, None, True)], dtype="float64",
                            method="bilinear", half=(1, 1), boundary="undef", n=2000, seed=16),
    "lookup_lon_periodic": dict(kind="idx", axes=[(np.arange(1440) * 0.25, 360.0, False)], n=4000, seed=17),
    "lookup_lat_irregular_desc": dict(kind="idx", axes=[(LAT_IRR, None, True)], n=4000, seed=18),
    "lookup_time_daily": dict(kind="idx", axes=[(T0 + np.arange(365, dtype=np.int64) * DAY, None, False)],
                              n=4000, seed=19),
    "lookup_time_irregular": dict(kind="idx", axes=[(T0 + np.cumsum(np.arange(1, 41, dtype=np.int64)) * 3_600_000_000_000,
                                                     None, False)], n=3000, seed=20),
}

# windowed 1-D .. 4-D with every fitting model
WND_MODELS = ("akima", "akima_periodic", "c_spline", "c_spline_not_a_knot", "c_spline_periodic", "linear",
              "steffen", "polynomial")
for _k, _m in enumerate(WND_MODELS):
    CASES[f"wnd1d_{_m}"] = dict(kind="wnd", axes=[(LON72, 360.0, False)], dtype="float64",
                                method=_m, half=(3, 1), boundary="undef", n=1500, seed=30 + _k,
                                derivative=bool(_k % 2))
    CASES[f"wnd2d_{_m}"] = dict(kind="wnd", axes=[(LON72, 360.0, False), (LAT_IRR, None, False)],
                                dtype="float64" if _k % 2 == 0 else "float32", method=_m,
                                half=(3, 3) if _k % 3 else (4, 3), boundary="shrink" if _k % 2 else "undef",
                                n=1500, seed=40 + _k)
CASES["wnd3d_bicubic_time"] = dict(kind="wnd",
                                   axes=[(LON72, 360.0, False), (LAT36, None, False), (TIME10, None, False)],
                                   dtype="float32", method="bicubic", half=(3, 3), boundary="undef", n=1500,
                                   seed=50)
CASES["wnd3d_c_spline_nearest_z"] = dict(kind="wnd",
                                         axes=[(LON72, 360.0, False), (LAT36, None, False),
                                               (np.arange(6) * 2.5, None, False)],
                                         dtype="float64", method="c_spline", half=(3, 3), boundary="shrink",
                                         n=1500, seed=51, z_method=1)
CASES["wnd4d_bilinear_depth"] = dict(kind="wnd",
                                     axes=[(LON72[::2], 360.0, False), (LAT36[::2], None, False),
                                           (TIME10[:6], None, False), (DEPTH8, None, False)],
                                     dtype="float32", method="bilinear", half=(2, 2), boundary="undef", n=1500,
                                     seed=52)
CASES["wnd4d_akima_f64"] = dict(kind="wnd",
                                axes=[(LON72[::2], 360.0, False), (LAT36[::2], None, False),
                                      (np.arange(5) * 1.0, None, False), (DEPTH8, None, True)],
                                dtype="float64", method="akima", half=(3, 3), boundary="undef", n=1200, seed=53,
                                u_method=1)
# grid filling: the whole filled grid, the iteration count and
# the final residual of the reference build are the fixture
CASES["fill_gs_f64_periodic"] = dict(kind="fill", algo="gauss_seidel", shape=(72, 36), dtype="float64", seed=61,
                                     first_guess="zonal_average", is_periodic=True, max_iterations=500,
                                     epsilon=1e-4, relaxation=1.0)
CASES["fill_gs_f32_zero_reflect"] = dict(kind="fill", algo="gauss_seidel", shape=(45, 31), dtype="float32", seed=62,
                                         first_guess="zero", is_periodic=False, max_iterations=300,
                                         epsilon=1e-4, relaxation=1.4)
CASES["fill_gs_f64_odd_seam"] = dict(kind="fill", algo="gauss_seidel", shape=(37, 23), dtype="float64", seed=63,
                                     first_guess="zonal_average", is_periodic=True, max_iterations=40,
                                     epsilon=1e-12, relaxation=1.2)
CASES["fill_mg_f64_periodic"] = dict(kind="fill", algo="multigrid", shape=(72, 36), dtype="float64", seed=64,
                                     first_guess="zonal_average", is_periodic=True, max_iterations=30,
                                     epsilon=1e-6, pre_smooth=2, post_smooth=2)
CASES["fill_mg_f32_zero_reflect"] = dict(kind="fill", algo="multigrid", shape=(45, 31), dtype="float32", seed=65,
                                         first_guess="zero", is_periodic=False, max_iterations=12,
                                         epsilon=1e-4, pre_smooth=1, post_smooth=3)
CASES["fill_mg_f64_odd_sizes"] = dict(kind="fill", algo="multigrid", shape=(77, 41), dtype="float64", seed=66,
                                      first_guess="zonal_average", is_periodic=True, max_iterations=6,
                                      epsilon=0.0, pre_smooth=2, post_smooth=1)
# fill::loess: the whole processed grid of the reference build is the fixture
CASES["fill_loess_f64_single_pass"] = dict(kind="fill", algo="loess", shape=(72, 36), dtype="float64", seed=67,
                                           first_guess="zonal_average", is_periodic=True, max_iterations=1,
                                           epsilon=1e-4, nx=3, ny=3, value_type="undefined")
CASES["fill_loess_f32_iterative_reflect"] = dict(kind="fill", algo="loess", shape=(45, 31), dtype="float32", seed=68,
                                                 first_guess="zonal_average", is_periodic=False, max_iterations=25,
                                                 epsilon=1e-3, nx=2, ny=4, value_type="undefined")
CASES["fill_loess_f64_all_iterative"] = dict(kind="fill", algo="loess", shape=(37, 23), dtype="float64", seed=69,
                                             first_guess="zero", is_periodic=True, max_iterations=6,
                                             epsilon=0.0, nx=5, ny=2, value_type="all")
CASES["fill_loess_f64_defined"] = dict(kind="fill", algo="loess", shape=(40, 30), dtype="float64", seed=70,
                                       first_guess="zonal_average", is_periodic=True, max_iterations=1,
                                       epsilon=1e-4, nx=4, ny=4, value_type="defined")
VECTOR_CASES = sorted(CASES)
PORT_CASES = [k for k in VECTOR_CASES if not CASES[k].get("ref_only")]


def fill_field(shape, dtype, seed, fraction=0.15, blobs=4):
    """Smooth field + noise with scattered and clustered NaN cells (an empty
    latitude band included when the grid is large enough)."""
    rng = np.random.Generator(np.random.PCG64(5000 + seed))
    nx, ny = shape
    x = np.linspace(0.0, 2.0 * np.pi, nx, endpoint=False)[:, None]
    y = np.linspace(-1.0, 1.0, ny)[None, :]
    g = 10.0 * np.sin(x) * np.cos(2.0 * y) + 3.0 * y + 0.3 * rng.standard_normal((nx, ny))
    m = rng.random((nx, ny)) < fraction
    for _ in range(blobs):
        i, j = int(rng.integers(0, nx)), int(rng.integers(0, ny))
        w, h = max(nx // 8, 1), max(ny // 8, 1)
        m[max(i - w, 0): i + w, max(j - h, 0): j + h] = True
    if ny > 8:
        m[:, ny // 3] = True  # a band without any defined value
    g = g.astype(dtype)
    g[m] = np.nan
    return g


def build_case(name):
    """Inputs of a case (no outputs), deterministic."""
    c = CASES[name]
    if c["kind"] == "fill":
        return {"grid": fill_field(c["shape"], c["dtype"], c["seed"])}
    rng = np.random.Generator(np.random.PCG64(1000 + c["seed"]))
    d = {}
    for k, (values, _, _) in enumerate(c["axes"]):
        d[f"axis{k}"] = np.asarray(values)
    coords = []
    for k, (values, period, _) in enumerate(c["axes"]):
        v = np.asarray(values)
        integer = v.dtype.kind in "iu"
        lo, hi = v.min(), v.max()
        if period is not None and c["kind"] not in ("win", "wnd"):
            lo, hi = lo - float(period), hi + float(period)
        coords.append(_coords(rng, float(lo), float(hi), c["n"], nodes=v, integer=integer))
    for k, name_ in enumerate("xyzu"[: len(coords)]):
        d[name_] = coords[k]
    if c["kind"] != "idx":
        shape = tuple(np.asarray(a[0]).size for a in c["axes"])
        d["grid"] = _field(shape, c["dtype"], c["seed"], c.get("nan_fraction", 0.0))
    return d


def fixture_path(name):
    return os.path.join(GOLDEN, f"ref_vectors_{name}.npz")


def load_vectors(name):
    """Inputs rebuilt deterministically + the committed reference outputs; the
    fixture also stores a checksum of the inputs it was generated from."""
    d = build_case(name)
    with np.load(fixture_path(name)) as f:
        for k in f.files:
            d["want_" + k if not k.startswith("sum_") else k] = f[k]
    for k in ("x", "y", "z", "u", "grid"):
        if k in d:
            assert checksum(d[k]) == int(d["sum_" + k]), f"{name}: input {k} is not what the fixture was made from"
    return d


def checksum(a):
    b = np.ascontiguousarray(a).reshape(-1).view(np.uint8)
    w = (np.arange(b.size, dtype=np.uint64) % np.uint64(65521)) + np.uint64(1)
    return int((b.astype(np.uint64) * w).sum() % np.uint64(2 ** 61 - 1))


def _orc_axes(o, c):
    out = []
    for values, period, flip in c["axes"]:
        a = o.axis(values, period=period)
        out.append(a.flip() if flip else a)
    return out


def _fill_result(grid, its, res):
    return {"out": grid, "iterations": np.array([its], dtype=np.int64),
            "residual": np.array([res], dtype=np.float64)}


def run_case_on_oracle(o, name, d):
    c = CASES[name]
    if c["kind"] == "fill":
        g = d["grid"].copy()
        if c["algo"] == "loess":
            return {"out": o.fill_loess(g, c["nx"], c["ny"], c["value_type"], c["first_guess"], c["is_periodic"],
                                        c["max_iterations"], c["epsilon"])}
        if c["algo"] == "gauss_seidel":
            its, res = o.fill_gauss_seidel(g, c["first_guess"], c["is_periodic"], c["max_iterations"],
                                           c["epsilon"], c["relaxation"])
        else:
            its, res = o.fill_multigrid(g, c["first_guess"], c["is_periodic"], c["max_iterations"],
                                        c["epsilon"], c["pre_smooth"], c["post_smooth"])
        return _fill_result(g, its, res)
    axes = _orc_axes(o, c)
    if c["kind"] == "idx":
        return {"pairs": axes[0].find_indexes(d["x"]), "index": axes[0].find_index(d["x"], False),
                "index_bounded": axes[0].find_index(d["x"], True)}
    if c["kind"] == "geo":
        m = METHOD[c["method"]]
        am = 1 if c["method"] == "nearest" else 0
        out, _ = o.geometric(axes, d["grid"], d["x"], d["y"], d.get("z"), d.get("u"), method=m, z_method=am,
                             u_method=am)
        return {"out": out}
    if c["kind"] == "wnd":
        out, _ = o.windowed_nd(axes, d["grid"], d["x"], d.get("y"), d.get("z"), d.get("u"), method=c["method"],
                               half_x=c["half"][0], half_y=c["half"][1], boundary=c["boundary"], stateless=True,
                               z_method=c.get("z_method", 0), u_method=c.get("u_method", 0),
                               derivative=c.get("derivative", False))
        return {"out": out}
    out, _ = o.windowed2d(axes, d["grid"], d["x"], d["y"], method=WMETHOD[c["method"]], half_x=c["half"][0],
                          half_y=c["half"][1], boundary=c["boundary"], stateless=True)
    return {"out": out}


def _pyi_axes(pyi, c):
    out = []
    for values, period, flip in c["axes"]:
        v = np.asarray(values)
        if v.dtype.kind in "iu":
            a = pyi.TemporalAxis(v.astype("datetime64[ns]"))
        else:
            a = pyi.Axis(v, period=period)
        out.append(a.flip(inplace=False) if flip else a)
    return out


def run_case_on_gpu(pyi, name, d):
    """Same case through the product's public API (CUDA kernels behind the C-ABI)."""
    c = CASES[name]
    if c["kind"] == "fill":
        g = d["grid"].copy()
        kw = dict(first_guess=c["first_guess"], is_periodic=c["is_periodic"],
                  max_iterations=c["max_iterations"], epsilon=c["epsilon"])
        if c["algo"] == "loess":
            return {"out": pyi.fill.loess(g, nx=c["nx"], ny=c["ny"], value_type=c["value_type"], **kw)}
        if c["algo"] == "gauss_seidel":
            its, res = pyi.fill.gauss_seidel(g, relaxation=c["relaxation"], **kw)
        else:
            its, res = pyi.fill.multigrid(g, pre_smooth=c["pre_smooth"], post_smooth=c["post_smooth"], **kw)
        return _fill_result(g, its, res)
    axes = _pyi_axes(pyi, c)
    if c["kind"] == "idx":
        x = d["x"].astype("datetime64[ns]") if d["x"].dtype.kind == "i" else d["x"]
        return {"pairs": axes[0].find_indexes(x), "index": axes[0].find_index(x, bounded=False),
                "index_bounded": axes[0].find_index(x, bounded=True)}
    grid = pyi.Grid(*axes, d["grid"])
    coords = [d[k] for k in "xyzu"[: len(axes)]]
    if len(axes) >= 3 and coords[2].dtype.kind == "i":
        coords[2] = coords[2].view("datetime64[ns]")
    if c["kind"] == "geo":
        fn = {2: pyi.bivariate, 3: pyi.trivariate, 4: pyi.quadrivariate}[len(axes)]
        return {"out": fn(grid, *coords, c["method"])}
    if c["kind"] == "wnd":
        return {"out": _run_wnd(pyi, c, grid, coords)}
    return {"out": pyi.core.bivariate(grid, *coords, _windowed_cfg(pyi, c))}


def _run_wnd(pyi, c, grid, coords):
    nd = len(coords)
    cls = {1: pyi.windowed.Univariate, 2: pyi.windowed.Bivariate, 3: pyi.windowed.Trivariate,
           4: pyi.windowed.Quadrivariate}[nd]
    cfg = getattr(cls, c["method"])()
    bc = pyi.config.BoundaryConfig.shrink() if c["boundary"] == "shrink" else pyi.config.BoundaryConfig.undef()
    cfg = cfg.with_boundary_mode(bc)
    if nd == 1:
        cfg = cfg.with_half_window_size(c["half"][0])
        fn = pyi.core.univariate_derivative if c.get("derivative") else pyi.core.univariate
        return fn(grid, coords[0], cfg)
    cfg = cfg.with_half_window_size_x(c["half"][0]).with_half_window_size_y(c["half"][1])
    am = {0: pyi.windowed.AxisConfig.linear, 1: pyi.windowed.AxisConfig.nearest}
    if nd >= 3:
        cfg = cfg.with_third_axis(am[c.get("z_method", 0)]())
    if nd >= 4:
        cfg = cfg.with_fourth_axis(am[c.get("u_method", 0)]())
    fn = {2: pyi.core.bivariate, 3: pyi.core.trivariate, 4: pyi.core.quadrivariate}[nd]
    return fn(grid, *coords, cfg)


def _windowed_cfg(pyi, c):
    cfg = pyi.windowed.Bivariate.bilinear() if c["method"] == "bilinear" else pyi.windowed.Bivariate.bicubic()
    cfg = cfg.with_half_window_size_x(c["half"][0]).with_half_window_size_y(c["half"][1])
    bc = pyi.config.BoundaryConfig.shrink() if c["boundary"] == "shrink" else pyi.config.BoundaryConfig.undef()
    return cfg.with_boundary_mode(bc)


def compare(name, got, d, exact):
    """exact: bit-for-bit (CPU restatement vs reference build).  Otherwise the
    north-star tolerances: indexes bit-exact, float64 values 1e-12, float32
    1e-5 relative (scale guard: max(|want|, 1))."""
    exact = exact or CASES[name]["kind"] == "fill"  # the fill kernels are bit-identical by construction
    for k, g in got.items():
        want = d["want_" + k]
        g = np.asarray(g)
        assert g.shape == want.shape and g.dtype == want.dtype, (name, k, g.dtype, want.dtype)
        if g.dtype.kind in "iu":
            assert np.array_equal(g, want), (name, k)
            continue
        assert np.array_equal(np.isnan(g), np.isnan(want)), (name, k, "NaN positions differ")
        if exact or CASES[name].get("method") == "bilinear" and CASES[name]["kind"] == "geo":
            m = ~np.isnan(want)
            if exact:
                assert np.array_equal(g[m], want[m]), (name, k, np.abs(g[m] - want[m]).max())
                continue
        tol = 1e-12 if g.dtype == np.float64 else 1e-5
        m = ~np.isnan(want)
        err = np.abs(g[m].astype(np.float64) - want[m].astype(np.float64))
        lim = tol * np.maximum(np.abs(want[m].astype(np.float64)), 1.0)
        assert np.all(err <= lim), (name, k, float((err / lim).max()))


def dump_json(path, obj):
    with open(path, "w") as f:
        json.dump(obj, f, indent=1)
