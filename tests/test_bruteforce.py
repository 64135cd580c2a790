"""Two independent restatements of the floating-point rules must agree, and the device must agree with both.

`tests/bruteforce.py` is an O(N^2) numpy restatement of examples/flock.jl:79-108 and
examples/celllistmap.jl:143-155, 176-191 that shares no code and no data structure with
`oracle/abm_oracle.cpp` (no cell list, no offset table, sums in id order).  Neither can be checked against
Julia here (the reference has no vector for a rule trajectory: SURVEY §8c), so the pin is their agreement:
neighbour counts exactly, positions / velocities to 1e-13 relative (f64) after one step from the same state.
CPU tests compare oracle and brute force; the `gpu` tests compare the CUDA path with the brute force.
"""
import numpy as np
import pytest

import bruteforce as bf
from helpers import FLOCK, flock_init, periodic_rel_err, rel_err, uniform
from oracle.pyoracle import ContinuousOracle

TOL64, TOL32 = 1e-13, 1e-5

#            extent                 spacing periodic        n     dtype
FLOCK_CASES = [
    ((100.0, 100.0), 5.0, True, 3000, np.float64),
    ((100.0, 60.0), 5.0, False, 2000, np.float64),
    ((80.0, 100.0), 5.0, (True, False), 2000, np.float64),
    ((60.0, 60.0, 60.0), 5.0, True, 4000, np.float64),
    ((60.0, 60.0, 60.0), 5.0, (False, True, False), 3000, np.float64),
    ((100.0, 100.0), 5.0, True, 3000, np.float32),
]


def _clustered(n, extent, seed, dtype):
    """A state with crowded cells: half of the birds drawn into a few clumps (the F2D-1M flock clusters)."""
    pos, vel = flock_init(n, extent, seed, dtype=np.float64)
    u = uniform(n, len(extent), seed, 2)
    centres = uniform(8, len(extent), seed + 1, 0) * np.asarray(extent)[None, :]
    k = n // 2
    pos[:k] = np.mod(centres[np.arange(k) % 8] + (u[:k] - 0.5) * 6.0, np.asarray(extent)[None, :])
    pos = np.clip(pos, 0.0, np.nextafter(np.asarray(extent), 0.0)[None, :])
    return pos.astype(dtype), vel.astype(dtype)


@pytest.mark.parametrize("search", ["exact", "approximate"])
@pytest.mark.parametrize("extent,spacing,periodic,n,dtype", FLOCK_CASES)
def test_oracle_equals_bruteforce_flock(extent, spacing, periodic, n, dtype, search):
    tol = TOL64 if dtype == np.float64 else TOL32
    for maker in (lambda: flock_init(n, extent, 11, dtype), lambda: _clustered(n, extent, 12, dtype)):
        pos, vel = maker()
        o = ContinuousOracle(extent, spacing=spacing, periodic=periodic, dtype=dtype)
        o.add_agents(pos, vel)
        cnt = o.flock_step_sync(FLOCK, search=search, return_counts=True)
        bp, bv, bn = bf.flock_step(pos, vel, extent, spacing, periodic, FLOCK, search)
        assert np.array_equal(cnt, bn), "neighbour counts: oracle vs brute force"
        po, vo = o.state()
        assert periodic_rel_err(po, bp, extent) <= tol
        assert rel_err(vo, bv) <= tol


PF_CASES = [
    ((400.0, 400.0), 20.0, True, 3000),
    ((400.0, 300.0), 25.0, False, 2500),
    ((400.0, 400.0), 20.0, (True, False), 2500),
    ((120.0, 120.0, 100.0), 20.0, True, 2500),
]


def _particles(n, extent, seed=42, max_radius=10.0, vscale=100.0):
    D = len(extent)
    pos = uniform(n, D, seed, 0) * np.asarray(extent)[None, :]
    u1 = np.maximum(uniform(n, D, seed, 1), 1e-300)
    u2 = uniform(n, D, seed, 2)
    vel = vscale * np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)
    w = uniform(n, 3, seed, 3)
    return pos, vel, (0.5 + 0.9 * w[:, 0]) * max_radius, 10.0 + 20.0 * w[:, 1], 10.0 + 100.0 * w[:, 2]


@pytest.mark.parametrize("extent,spacing,periodic,n", PF_CASES)
def test_oracle_equals_bruteforce_pairforce(extent, spacing, periodic, n):
    pos, vel, rad, k, mass = _particles(n, extent)
    o = ContinuousOracle(extent, spacing=spacing, periodic=periodic)
    o.add_agents(pos, vel)
    o.pairforce_step(rad, k, mass, 0.001, 20.0)
    bp, bv = bf.pairforce_step(pos, vel, rad, k, mass, extent, periodic, 0.001, 20.0)
    po, vo = o.state()
    assert periodic_rel_err(po, bp, extent) <= 1e-12
    assert rel_err(vo, bv) <= 1e-12


# ---------------------------------------------------------------------------------------------------------
# device vs brute force
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["generic", "auto"])
@pytest.mark.parametrize("search", ["exact", "approximate"])
@pytest.mark.parametrize("extent,spacing,periodic,n,dtype", FLOCK_CASES)
def test_device_equals_bruteforce_flock(extent, spacing, periodic, n, dtype, search, kernel):
    import agents_jl_b200 as ab

    tol = TOL64 if dtype == np.float64 else TOL32
    for maker in (lambda: flock_init(n, extent, 11, dtype), lambda: _clustered(n, extent, 12, dtype)):
        pos, vel = maker()
        m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing, periodic=periodic, dtype=dtype),
                           model_step=ab.FlockRule(search=search, kernel=kernel, **FLOCK))
        m.add_agents(pos, vel)
        m.step(1)
        bp, bv, bn = bf.flock_step(pos, vel, extent, spacing, periodic, FLOCK, search)
        assert np.array_equal(m.last_neighbor_counts(), bn), "neighbour counts: device vs brute force"
        assert periodic_rel_err(m.pos, bp, extent) <= tol
        assert rel_err(m.vel, bv) <= tol
        m.close()


@pytest.mark.gpu
@pytest.mark.parametrize("extent,spacing,periodic,n", PF_CASES)
def test_device_equals_bruteforce_pairforce(extent, spacing, periodic, n):
    import agents_jl_b200 as ab

    pos, vel, rad, k, mass = _particles(n, extent)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing, periodic=periodic),
                       model_step=ab.PairForceRule(dt=0.001, cutoff=20.0))
    m.add_agents(pos, vel, r=rad, k=k, mass=mass)
    m.step(1)
    bp, bv = bf.pairforce_step(pos, vel, rad, k, mass, extent, periodic, 0.001, 20.0)
    assert periodic_rel_err(m.pos, bp, extent) <= 1e-12
    assert rel_err(m.vel, bv) <= 1e-12
    m.close()
