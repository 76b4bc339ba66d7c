"""Axis definitions and adversarial coordinates shared by the parity tests."""
import numpy as np

DAY = 86_400_000_000_000
T0 = int(np.datetime64("2021-01-01", "ns").astype(np.int64))


def f64_axes():
    """name -> (values, kwargs, flip)"""
    lon = np.arange(1440) * 0.25
    lat = -89.875 + np.arange(720) * 0.25
    zoo = {
        "lon_periodic": (lon, dict(period=360.0), False),
        "lon_180": (lon - 180.0, dict(period=360.0), False),
        "lon_periodic_desc": (lon, dict(period=360.0), True),
        "lon_partial_period": (np.arange(101.0), dict(period=360.0), False),
        "lon_wrapped_values": (np.r_[np.arange(170.0, 180.0, 1.0), np.arange(-180.0, -170.0, 1.0)],
                               dict(period=360.0), False),
        "lon_0p1": (np.arange(3600) * 0.1, dict(period=360.0), False),
        "lat": (lat, {}, False),
        "lat_desc": (lat, {}, True),
        "lat_desc_inexact": (-89.8 + np.arange(600) * 0.3, {}, True),   # nodes are not fma(k, step, start)
        "levels_desc": (np.arange(20) * 250.0, {}, True),                # regular descending z / u
        "lat_irregular": (np.degrees(np.arcsin(np.linspace(-0.99999, 0.99999, 720))), {}, False),
        "lat_irregular_desc": (np.degrees(np.arcsin(np.linspace(-0.99999, 0.99999, 720))), {}, True),
        "small_irregular": (np.array([0.0, 1.0, 4.0, 8.0, 20.0, 50.0]), {}, False),
        "small_irregular_desc": (np.array([0.0, 1.0, 4.0, 8.0, 20.0, 50.0]), {}, True),
        "irregular_periodic": (np.r_[np.arange(0.0, 350.0, 10.0), [352.0, 357.0]], dict(period=360.0), False),
        "depth": (np.round(5500.0 * (np.arange(50) / 49) ** 2.2, 1), {}, False),
        "two_points": (np.array([1.0, 3.0]), {}, False),
        "single": (np.array([5.0]), {}, False),
        "tiny_step": (1e6 + np.arange(1000) * 1e-3, {}, False),
        "negative_step_input": (np.linspace(10.0, -10.0, 81), {}, False),
    }
    return zoo


def i64_axes():
    daily = T0 + np.arange(365, dtype=np.int64) * DAY
    zoo = {
        "daily_ns": (daily, {}, False),
        "daily_ns_desc": (daily, {}, True),
        "monthly30": (T0 + np.arange(12, dtype=np.int64) * 30 * DAY, {}, False),
        "irregular_time": (T0 + np.cumsum(np.r_[0, np.random.default_rng(7).integers(1, 10, 99)]).astype(np.int64) * 3_600_000_000_000, {}, False),
        "irregular_time_desc": (T0 + np.cumsum(np.r_[0, np.random.default_rng(7).integers(1, 10, 99)]).astype(np.int64) * 3_600_000_000_000, {}, True),
        "hours_small": (np.arange(0, 48, 1, dtype=np.int64), {}, False),
        "odd_step": (np.arange(10, dtype=np.int64) * 7 - 31, {}, False),
        "int_periodic": (np.arange(0, 360, 10, dtype=np.int64), dict(period=360), False),
    }
    return zoo


def coords_for(desc: dict, n: int, seed: int):
    """Adversarial coordinates for an axis described by Oracle.describe()."""
    rng = np.random.default_rng(seed)
    lo, hi = desc["min_value"], desc["max_value"]
    if desc["coord_type"] == 0:
        span = max(hi - lo, 1.0)
        base = rng.uniform(lo - 1.5 * span, hi + 1.5 * span, n)
        if desc["kind"] == 0:
            nodes = desc["start"] + np.arange(desc["size"]) * desc["step"]
        else:
            nodes = desc["points"]
        pick = nodes[rng.integers(0, len(nodes), n // 4 + 1)]
        mids = pick + (0.5 * desc.get("step", 1.0) if desc["kind"] == 0 else 0.0)
        extra = np.r_[
            pick, np.nextafter(pick, np.inf), np.nextafter(pick, -np.inf), mids,
            np.nextafter(mids, np.inf), np.nextafter(mids, -np.inf),
            pick + 360.0, pick - 360.0, pick + 720.0,
            [lo, hi, np.nextafter(lo, -np.inf), np.nextafter(hi, np.inf), np.nan, np.inf, -np.inf,
             1e300, -1e300, 1e19, -1e19, 0.0, -0.0, lo + 360.0, hi - 360.0],
        ]
        return np.r_[base, extra]
    span = max(int(hi - lo), 1)
    base = rng.integers(lo - span, hi + span, n, dtype=np.int64)
    if desc["kind"] == 0:
        nodes = desc["start"] + np.arange(desc["size"], dtype=np.int64) * desc["step"]
        half = abs(int(desc["step"])) // 2
    else:
        nodes = desc["points"]
        half = 1
    pick = nodes[rng.integers(0, len(nodes), n // 4 + 1)]
    extra = np.r_[pick, pick + 1, pick - 1, pick + half, pick - half, pick + half + 1, pick - half - 1,
                  pick + half - 1,
                  np.array([lo, hi, lo - 1, hi + 1, np.iinfo(np.int64).min, 0], dtype=np.int64)]
    return np.r_[base, extra].astype(np.int64)
