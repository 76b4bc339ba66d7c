"""Synthetic axes, grids and query points of the BASELINE.json configurations
(SURVEY.md section 8d).  Pure numpy; shared by tests/ and bench.py."""
from __future__ import annotations

import numpy as np


def lon_axis_values(n: int = 1440) -> np.ndarray:
    return np.arange(n) * (360.0 / n)


def lat_axis_values(n: int = 720) -> np.ndarray:
    step = 180.0 / n
    return -90.0 + step / 2 + np.arange(n) * step


def irregular_lat_values(n: int = 720) -> np.ndarray:
    return np.degrees(np.arcsin(np.linspace(-0.99999, 0.99999, n)))


def time_axis_values(n: int = 365, step_days: int = 1) -> np.ndarray:
    t0 = np.datetime64("2021-01-01", "ns")
    return (t0 + np.arange(n) * np.timedelta64(step_days, "D")).astype("datetime64[ns]")


def depth_axis_values(n: int = 50) -> np.ndarray:
    return np.round(5500.0 * (np.arange(n) / (n - 1)) ** 2.2, 1)


def field(lon, lat, nz: int = 0, nu: int = 0, dtype="float64", seed: int = 42, noise: float = 1e-3):
    """sin(lon) cos(lat) (1 + 0.1 k_z + 0.01 k_u) + noise, C-order [lon][lat][z][u].
    Built slab by slab along lon (float64 arithmetic, then the cast) so that a
    2.5 GB float32 4-D grid does not need a 5 GB float64 temporary; values and
    the noise stream are those of the one-shot expression."""
    rng = np.random.Generator(np.random.PCG64(seed))
    dt = np.dtype(dtype)
    shape = [lon.size, lat.size] + ([nz] if nz else []) + ([nu] if nu else [])
    out = np.empty(shape, dtype=dt)
    sin_lon = np.sin(np.radians(lon))
    cos_lat = np.cos(np.radians(lat))
    rows = max(1, (1 << 24) // max(1, int(np.prod(shape[1:]))))
    for a in range(0, lon.size, rows):
        g = sin_lon[a:a + rows, None] * cos_lat[None, :]
        if nz:
            g = g[..., None] * (1.0 + 0.1 * np.arange(nz))
        if nu:
            g = g[..., None] * (1.0 + 0.01 * np.arange(nu)) if nz else g
        g = np.ascontiguousarray(np.broadcast_to(g, [g.shape[0]] + shape[1:])).astype(np.float64)
        if noise:
            g = g + noise * rng.standard_normal(g.shape)
        if dt.kind in "iu":
            info = np.iinfo(dt)
            g = np.clip(np.round(g * 1000.0), info.min, info.max)
        out[a:a + rows] = g.astype(dt)
    return out


def queries(n: int, seed: int, ndim: int = 2, time_values=None, depth_max: float = 5600.0,
            nan_fraction: float = 1e-4, x_range=(-180.0, 540.0), y_range=(-90.0, 90.0)):
    """x ~ U(-180, 540), y ~ U(-90, 90), z ~ U int64 over [t0 - 1 day, t_last + 1 day],
    u ~ U(-10, depth_max); a few NaN / NaT coordinates."""
    rng = np.random.Generator(np.random.PCG64(seed))
    x = rng.uniform(*x_range, n)
    y = rng.uniform(*y_range, n)
    out = [x, y]
    if nan_fraction and n >= 16:
        k = max(1, int(n * nan_fraction))
        x[rng.integers(0, n, k)] = np.nan
        y[rng.integers(0, n, k)] = np.nan
    if ndim >= 3:
        tv = np.asarray(time_values).astype("datetime64[ns]").astype(np.int64)
        day = 86_400_000_000_000
        z = rng.integers(tv.min() - day, tv.max() + day, n, dtype=np.int64)
        if nan_fraction and n >= 16:
            z[rng.integers(0, n, max(1, int(n * nan_fraction)))] = np.iinfo(np.int64).min
        out.append(z)
    if ndim >= 4:
        out.append(rng.uniform(-10.0, depth_max, n))
    return out
