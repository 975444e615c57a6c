"""Deterministic synthetic fields and coefficients.

The reference fills its arrays with U(0,1) doubles from an *unseeded* mt19937_64
(src/multiarray.cpp:13-40) and, in the single-node drivers, 129 U(0,1) coefficients
(single/cpu.cpp:11-17), so no two reference runs see the same data.  Parity needs identical
inputs on both sides, so this module restates the distribution (uniform on [0,1), 53 random
bits) with a counter-based generator (splitmix64 of the element index) that gives the same
values on any machine, in any order, without storing them.
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
    return z ^ (z >> np.uint64(31))


def uniform(n, seed=0, start=0):
    """n doubles in [0,1): element i is a pure function of (seed, start + i)."""
    with np.errstate(over="ignore"):
        idx = np.arange(start, start + n, dtype=np.uint64)
        key = _splitmix64(np.uint64(seed) * np.uint64(0xD1342543DE82EF95) + np.uint64(1))
        bits = _splitmix64(idx ^ key)
    return (bits >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def coefficients(seed=42, n=129):
    """The 129 U(0,1) coefficients of single/cpu.cpp:11-17 (7pt.py reads coeff[0..6])."""
    return uniform(n, seed=seed ^ 0x5EED)
