"""Host mirror of la_synth (csrc/elementwise.cu): the same counter-based generator in numpy, so a
test or a bench can check any slice of a device-generated operand without downloading all of it."""
import numpy as np

from .traits import parameter

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M
    x = ((x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M
    x = ((x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M
    return x ^ (x >> np.uint64(31))


def values(dtype, seed, start, count):
    """Elements [start, start+count) of the sequence la_synth(dtype, seed) produces."""
    dt = parameter(dtype).np
    with np.errstate(over="ignore"):
        bits = splitmix64(np.uint64(seed) ^ np.arange(start, start + count, dtype=np.uint64))
    if dt == np.float32:
        return ((bits >> np.uint64(40)).astype(np.float32) * np.float32(2.0 / 16777216.0) - np.float32(1.0)).astype(np.float32)
    if dt == np.float64:
        return (bits >> np.uint64(11)).astype(np.float64) * (2.0 / 9007199254740992.0) - 1.0
    return (bits >> np.uint64(64 - 8 * dt.itemsize)).astype(np.dtype("u%d" % dt.itemsize)).view(dt) if dt.kind == "i" else (bits >> np.uint64(64 - 8 * dt.itemsize)).astype(dt)
