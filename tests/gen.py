"""Seeded synthetic inputs shared by the oracle tests and the GPU parity tests.

Integer data exercises wrap-around (full-range) unless `small`; divisors are never 0 and never -1
(x/0 and MIN/-1 are undefined in OpenCL C)."""
import numpy as np

INT_NAMES = ["u8", "i8", "u16", "i16", "u32", "i32", "u64", "i64"]
FLOAT_NAMES = ["f32", "f64"]
ALL_NAMES = INT_NAMES + FLOAT_NAMES


def values(name, n, seed, small=False):
    from oracle import NP
    dt = np.dtype(NP[name])
    rng = np.random.default_rng(seed)
    if dt.kind == "f":
        return rng.uniform(-1.0, 1.0, size=n).astype(dt)
    if small:
        lo, hi = (0, 100) if dt.kind == "u" else (-100, 100)
        return rng.integers(lo, hi, size=n, endpoint=True).astype(dt)
    info = np.iinfo(dt)
    return rng.integers(info.min, info.max, size=n, dtype=dt, endpoint=True)


def divisors(name, n, seed):
    from oracle import NP
    dt = np.dtype(NP[name])
    rng = np.random.default_rng(seed)
    if dt.kind == "f":
        return rng.uniform(1.0, 2.0, size=n).astype(dt)
    info = np.iinfo(dt)
    hi = min(info.max, 1000)
    d = rng.integers(1, hi, size=n, endpoint=True).astype(dt)
    if dt.kind == "i":
        sign = rng.integers(0, 1, size=n, endpoint=True).astype(bool)
        d = np.where(sign & (d != 1), -d, d).astype(dt)   # negative divisors, never -1
    return d
