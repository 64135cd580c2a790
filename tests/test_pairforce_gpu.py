"""GPU parity tests for the fused pair-force step (csrc/ab_pairforce.cu; examples/celllistmap.jl:143-191)
against the CPU oracle.  Parity is UNPINNED by the reference (no reference test holds a vector for the
celllistmap example): oracle-vs-device only.  Tolerance: relative 1e-12 (f64) / 2e-5 (f32) after one
step from identical state — the summation order over pairs differs (the oracle visits each unique pair
once in discovery order, the device sums per agent in slot order)."""
import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import periodic_rel_err, rel_err, uniform
from oracle.pyoracle import ContinuousOracle

pytestmark = pytest.mark.gpu

TOL64, TOL32 = 1e-12, 2e-5


def particles(n, extent, seed=42, max_radius=10.0, vscale=100.0):
    """celllistmap.jl:118-123 on the SURVEY §8d counter stream (Box-Muller for the velocities)."""
    D = len(extent)
    pos = uniform(n, D, seed, 0) * np.asarray(extent)[None, :]
    u1 = np.maximum(uniform(n, D, seed, 1), 1e-300)
    u2 = uniform(n, D, seed, 2)
    vel = vscale * np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)
    w = uniform(n, 3, seed, 3)
    rad = (0.5 + 0.9 * w[:, 0]) * max_radius
    k = 10.0 + 20.0 * w[:, 1]
    mass = 10.0 + 100.0 * w[:, 2]
    return pos, vel, rad, k, mass


def make(extent, spacing, periodic, n, dtype=np.float64, dt=0.001, cutoff=20.0, **kw):
    pos, vel, rad, k, mass = particles(n, extent, **kw)
    pos, vel, rad, k, mass = (a.astype(dtype) for a in (pos, vel, rad, k, mass))
    o = ContinuousOracle(extent, spacing=spacing, periodic=periodic, dtype=dtype)
    o.add_agents(pos, vel)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing, periodic=periodic, dtype=dtype),
                       model_step=ab.PairForceRule(dt=dt, cutoff=cutoff))
    m.add_agents(pos, vel, r=rad, k=k, mass=mass)
    return o, m, (rad.astype(np.float64), k.astype(np.float64), mass.astype(np.float64))


CASES = [
    ((400.0, 400.0), 20.0, True, 6000),            # fine path: cutoff == spacing (PF2D-10M's geometry)
    ((400.0, 300.0), 25.0, True, 3000),            # fine path: cutoff < spacing
    ((400.0, 400.0), 20.0, False, 5000),           # fine path, walls
    ((400.0, 400.0), 20.0, (True, False), 5000),   # fine path, mixed periodicity
    ((200.0, 200.0), 10.0, True, 1500),            # generic path: cutoff = 2 * spacing
    ((60.0, 40.0), 20.0, True, 200),               # generic path: fewer than 3... exactly 3 x 2 cells
    ((120.0, 120.0, 100.0), 20.0, True, 4000),     # 3-D fine path
]


@pytest.mark.parametrize("extent,spacing,periodic,n", CASES)
def test_pairforce_one_step_parity(extent, spacing, periodic, n):
    o, m, (rad, k, mass) = make(extent, spacing, periodic, n)
    o.pairforce_step(rad, k, mass, 0.001, 20.0)
    m.step(1)
    po, vo = o.state()
    assert periodic_rel_err(m.pos, po, extent) <= TOL64
    assert rel_err(m.vel, vo) <= TOL64
    # the per-agent columns travel with the agents through the re-sort
    assert np.array_equal(m.column("r"), rad) and np.array_equal(m.column("mass"), mass)
    m.close()


def test_pairforce_trajectory_and_momentum():
    extent = (400.0, 400.0)
    o, m, (rad, k, mass) = make(extent, 20.0, True, 6000)
    p0 = (m.vel * mass[:, None]).sum(0)
    for step in range(10):
        # one-step parity from identical state, repeated along the trajectory
        po, vo = m.pos, m.vel
        oo = ContinuousOracle(extent, spacing=20.0)
        oo.add_agents(po, vo)
        oo.pairforce_step(rad, k, mass, 0.001, 20.0)
        m.step(1)
        assert periodic_rel_err(m.pos, oo.state()[0], extent) <= TOL64
        assert rel_err(m.vel, oo.state()[1]) <= TOL64
    p1 = (m.vel * mass[:, None]).sum(0)
    assert np.allclose(p0, p1, rtol=1e-9)          # f_i += fij, f_j -= fij (celllistmap.jl:151-152)
    # batched call == single steps, bit for bit (deterministic summation order)
    o2, m2, _ = make(extent, 20.0, True, 6000)
    m2.step(10)
    assert np.array_equal(m2.pos, m.pos) and np.array_equal(m2.vel, m.vel)
    m.close(); m2.close()


def test_pairforce_fast_agents_take_the_generic_resort():
    # velocities large enough to cross more than one cell per step: the device notices and
    # re-indexes with the general counting sort instead of the tiled re-sort
    extent = (400.0, 400.0)
    o, m, (rad, k, mass) = make(extent, 20.0, True, 3000, dt=0.5, vscale=100.0)
    o.pairforce_step(rad, k, mass, 0.5, 20.0)
    m.step(1)
    assert periodic_rel_err(m.pos, o.state()[0], extent) <= 1e-9
    ids = np.sort(m.nearby_ids(5, 20.0, search="exact"))
    assert np.array_equal(ids, np.sort(o.nearby_ids_agent(5, 20.0, "exact")))
    m.close()


def test_pairforce_float32():
    extent = (400.0, 400.0)
    o, m, (rad, k, mass) = make(extent, 20.0, True, 4000, dtype=np.float32)
    o.pairforce_step(rad, k, mass, 0.001, 20.0)
    m.step(1)
    assert periodic_rel_err(m.pos, o.state()[0], extent) <= TOL32
    assert rel_err(m.vel, o.state()[1]) <= 5e-4     # |v| ~ 100 and forces ~ 1e5 in f32
    m.close()


def test_pairforce_needs_its_columns():
    m = ab.StandardABM(ab.ContinuousSpace((100.0, 100.0), spacing=20.0), model_step=ab.PairForceRule())
    m.add_agents(np.array([[1.0, 1.0], [2.0, 2.0]]), np.zeros((2, 2)))
    with pytest.raises(KeyError):
        m.step(1)
    m.close()


@pytest.mark.parametrize("k", [2, 3])
def test_pairforce_known_answers(k):
    """The device against numbers worked out by hand from examples/celllistmap.jl:143-155, 176-191
    (tests/known_answers.py) — not against the oracle."""
    import known_answers as ka

    m = ab.StandardABM(ab.ContinuousSpace(ka.PF_EXTENT, spacing=ka.PF_SPACING),
                       model_step=ab.PairForceRule(dt=ka.PF_DT, cutoff=ka.PF_CUTOFF))
    m.add_agents(ka.PF_POS[:k], ka.PF_VEL[:k], r=ka.PF_R[:k], k=ka.PF_K[:k], mass=ka.PF_MASS[:k])
    m.step(1)
    pos1, vel1 = (ka.PF2_POS1, ka.PF2_VEL1) if k == 2 else (ka.PF3_POS1, ka.PF3_VEL1)
    assert np.max(np.abs(m.pos - pos1)) <= ka.PF_TOL and np.max(np.abs(m.vel - vel1)) <= ka.PF_TOL
    m.close()
