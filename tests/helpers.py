"""Shared test helpers: the SURVEY §8d counter-based initial-state generator."""
import numpy as np

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M
    return z ^ (z >> np.uint64(31))


def uniform(n, ncomp, seed=42, stream=0):
    """u(i,c) = (splitmix64(seed ^ (i*8+c)) >> 11) * 2^-53, c offset by `stream`*4."""
    with np.errstate(over="ignore"):
        i = np.arange(n, dtype=np.uint64)[:, None]
        c = (np.arange(ncomp, dtype=np.uint64) + np.uint64(stream * 4))[None, :]
        k = np.uint64(seed) ^ (i * np.uint64(8) + c)
        z = splitmix64(k)
    return (z >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)


def flock_init(n, extent, seed=42, dtype=np.float64):
    """pos = u*extent (random_position, continuous.jl:107-109); vel = 2u-1 (flock.jl:61)."""
    D = len(extent)
    pos = uniform(n, D, seed, 0) * np.asarray(extent, dtype=np.float64)[None, :]
    vel = uniform(n, D, seed, 1) * 2.0 - 1.0
    return pos.astype(dtype), vel.astype(dtype)


FLOCK = dict(speed=1.0, cohere_factor=0.1, separation=2.0, separate_factor=0.25, match_factor=0.04,
             visual_distance=5.0)  # examples/flock.jl:45-55


def csr_sets(offsets, ids):
    return [np.sort(ids[offsets[i]:offsets[i + 1]]) for i in range(len(offsets) - 1)]


def rel_err(a, b):
    """max |a - b| / max(1, |b|): RELATIVE for |b| > 1, ABSOLUTE below.  For velocities (|v| = 1) "1e-13" is an
    absolute 1e-13; for a coordinate near 2000 (F2D-1M) the same bound admits 2e-10 absolute — about one ulp
    (2.3e-13) times the few hundred roundings of a step, stated here so that nobody reads it as 1e-13 absolute."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b))))


def periodic_rel_err(a, b, extent):
    """max |a-b| / max(1, |b|) with the periodic image removed (a bird that wraps one ulp apart lands L away).
    Same scaling as rel_err: absolute below 1, relative above."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    L = np.asarray(extent, dtype=np.float64)[None, :]
    d = np.abs(a - b)
    d = np.minimum(d, L - d)
    return float(np.max(d / np.maximum(1.0, np.abs(b))))
