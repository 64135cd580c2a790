"""Parity at BASELINE.json's FULL sizes (F2D-1M, PF2D-10M, SCH-8192, F3D-12.5M = one GPU's share of F3D-100M).
The oracle cannot step 10^7 agents in seconds, so these tests use what the domain offers:

* F2D-1M: the oracle does step 10^6 boids (a few seconds per step): direct one-step parity from identical state,
  at the uniform seed state and at a clustered state the device produced, neighbour counts exact.
* Locality.  Flock and pair force are short-range: the new state of the agents inside a WINDOW depends only on the
  agents within the interaction range around it.  The oracle steps the window + margin (a few 10^4 agents cut out of
  the full state, shifted to the origin by a multiple of the spacing so that the cells coincide) and must agree with
  the device's full-size step on the window's agents — neighbour counts exactly, positions / velocities to 1e-11
  absolute (the shift costs one ulp of a coordinate ~10^3).
* Invariants over ALL agents: |vel| = 1 after a flock step, positions inside the extent, cells consistent with
  positions, total momentum of the pair-force step conserved, Schelling moods equal to a vectorised numpy
  restatement of the rule on the whole 8192^2 grid, relocation invariants on all 50 M agents.
"""
import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import FLOCK, flock_init, periodic_rel_err, rel_err, uniform
from oracle.pyoracle import ContinuousOracle
from test_pairforce_gpu import particles

pytestmark = pytest.mark.gpu


def _flock_invariants(m, extent, spacing):
    p, v = m.pos, m.vel
    assert np.all(np.abs(np.sqrt((v * v).sum(axis=1)) - 1.0) <= 1e-15)
    assert np.all(p >= 0.0) and np.all(p <= np.asarray(extent))
    want = np.maximum(1, np.ceil(p / spacing)).astype(np.int32)            # pos2cell, continuous.jl:112-113
    assert np.array_equal(m.cells(), want)


def test_f2d_1m_one_step_against_the_oracle():
    n, extent = 1_000_000, (2000.0, 2000.0)
    pos, vel = flock_init(n, extent, 42)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos, vel)
    o = ContinuousOracle(extent, spacing=5.0, periodic=True)
    o.add_agents(pos, vel)
    for presteps in (0, 60):                                   # uniform seed state, then the clustered state
        if presteps:
            m.step(presteps)
            p, v = m.pos, m.vel
            o = ContinuousOracle(extent, spacing=5.0, periodic=True)
            o.add_agents(p, v)
        cnt = o.flock_step_sync(FLOCK, nsteps=1, search="approximate", return_counts=True)
        m.step(1)
        assert np.array_equal(m.last_neighbor_counts(), cnt)
        po, vo = o.state()
        # a bird that wraps from ~2000 to ~0 keeps the absolute rounding of the large coordinate (ulp(2000) = 2.3e-13)
        assert periodic_rel_err(m.pos, po, extent) <= 1e-12
        # summation order differs (cell aggregates vs the oracle's neighbour walk); the normalisation amplifies it
        # where vel + steering nearly cancels: the worst of 2 x 10^6 components sits above the 1e-13 of the small cases
        dv = np.abs(m.vel - vo)
        assert dv.max() <= 1e-12 and dv.mean() <= 1e-15
        _flock_invariants(m, extent, 5.0)
    assert cnt.mean() > 100                                    # the flock has clustered (81 at the uniform state)
    m.close()


def _window(p, lo, hi):
    return np.all((p >= np.asarray(lo)) & (p < np.asarray(hi)), axis=1)


def test_f3d_12m_window_against_the_oracle():
    extent, spacing = (1400.0, 1400.0, 175.0), 5.0
    n = int(4.5554 * 280 * 280 * 35)
    pos, vel = flock_init(n, extent, 42)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos, vel)
    m.step(3)
    p0, v0 = m.pos, m.vel
    m.step(1)
    p1, v1, cnt1 = m.pos, m.vel, m.last_neighbor_counts()
    _flock_invariants(m, extent, spacing)
    m.close()
    margin = 2 * spacing                                       # grid_r = 2 cells for :approximate with spacing = r
    for lo, hi in (((500.0, 600.0, 50.0), (560.0, 660.0, 100.0)), ((1000.0, 95.0, 20.0), (1080.0, 150.0, 60.0))):
        lo_o, hi_o = np.asarray(lo) - margin, np.asarray(hi) + margin
        outer = np.flatnonzero(_window(p0, lo_o, hi_o))
        inner = _window(p0[outer], lo, hi)
        assert inner.sum() > 3000
        o = ContinuousOracle(tuple(hi_o - lo_o), spacing=spacing, periodic=False)
        o.add_agents(p0[outer] - lo_o, v0[outer])
        cnt = o.flock_step_sync(FLOCK, nsteps=1, search="approximate", return_counts=True)
        po, vo = o.state()
        sel = outer[inner]
        assert np.array_equal(cnt1[sel], cnt[inner])           # same candidate sets, bird by bird
        assert np.max(np.abs(p1[sel] - (po[inner] + lo_o))) <= 1e-11
        assert np.max(np.abs(v1[sel] - vo[inner])) <= 1e-11


def test_pf2d_10m_window_and_momentum():
    side = 791
    extent = (20.0 * side, 20.0 * side)
    n = int(0.04 * extent[0] * extent[1])
    pos, vel, rad, k, mass = particles(n, extent)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=20.0), model_step=ab.PairForceRule(dt=0.001, cutoff=20.0))
    m.add_agents(pos, vel, r=rad, k=k, mass=mass)
    m.step(2)
    p0, v0 = m.pos, m.vel
    m.step(1)
    p1, v1 = m.pos, m.vel
    m.close()
    assert np.all(p1 >= 0.0) and np.all(p1 <= np.asarray(extent))
    # f_ij = -f_ji (celllistmap.jl:150-152): the total momentum does not change
    mom0, mom1 = (mass[:, None] * v0).sum(axis=0), (mass[:, None] * v1).sum(axis=0)
    assert np.max(np.abs(mom1 - mom0)) <= 1e-13 * np.abs(mass[:, None] * v0).sum()
    assert not np.array_equal(v0, v1)
    margin = 20.0
    for lo, hi in (((3000.0, 3000.0), (3600.0, 3600.0)), ((15000.0, 200.0), (15700.0, 700.0))):
        lo_o, hi_o = np.asarray(lo) - margin, np.asarray(hi) + margin
        outer = np.flatnonzero(_window(p0, lo_o, hi_o))
        inner = _window(p0[outer], lo, hi)
        assert inner.sum() > 10000
        o = ContinuousOracle(tuple(hi_o - lo_o), spacing=20.0, periodic=False)
        o.add_agents(p0[outer] - lo_o, v0[outer])
        o.pairforce_step(rad[outer], k[outer], mass[outer], 0.001, 20.0)
        po, vo = o.state()
        sel = outer[inner]
        assert np.max(np.abs(v1[sel] - vo[inner])) <= 1e-10    # |v| ~ 100: 1e-12 relative
        assert np.max(np.abs(p1[sel] - (po[inner] + lo_o))) <= 1e-10
        assert np.max(np.abs(v1[sel] - v0[sel])) > 1.0        # forces did act on the window


def test_sch_8192_moods_and_relocation():
    side = 8192
    ncell = side * side
    n = int(ncell * 0.75)
    perm = np.argsort(uniform(ncell, 1, 42)[:, 0], kind="stable")[:n]
    pos = np.stack([perm % side + 1, perm // side + 1], axis=1).astype(np.int32)
    group = (np.arange(n) % 2 + 1).astype(np.int32)
    del perm

    def lin(p):
        return (p[:, 0].astype(np.int64) - 1) + (p[:, 1].astype(np.int64) - 1) * side

    def numpy_moods(p, min_to_be_happy=3):
        # docs/src/tutorial.jl:351-372: count the agents of the own group in the 8 surrounding cells
        # (GridSpaceSingle, chebyshev, non-periodic: cells outside the grid do not exist)
        G = np.zeros((side + 2, side + 2), dtype=np.int8)
        G[p[:, 1], p[:, 0]] = group
        C = G[1:-1, 1:-1]
        cnt = np.zeros((side, side), dtype=np.int8)
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                if dx or dy:
                    cnt += (G[1 + dy:side + 1 + dy, 1 + dx:side + 1 + dx] == C) & (C != 0)
        return (cnt[p[:, 1] - 1, p[:, 0] - 1] >= min_to_be_happy).astype(np.uint8)

    m = ab.StandardABM(ab.GridSpaceSingle((side, side), periodic=False),
                       model_step=ab.SchellingRule(min_to_be_happy=3, relocate=False))
    m.add_agents(pos, group=group, mood=np.zeros(n, dtype=np.uint8))
    m.step(1)
    want = numpy_moods(pos)
    assert np.array_equal(m.column("mood"), want) and np.array_equal(m.pos, pos)
    assert 0.05 < 1.0 - want.mean() < 0.6
    # full step with relocation: invariants of SURVEY §8a row 17 over all agents
    rule = ab.SchellingRule(min_to_be_happy=3, relocate=True, seed=42)
    m.model_step = rule
    m.step(1)
    new, mood = m.pos, m.column("mood")
    unhappy = want == 0
    moved = np.any(new != pos, axis=1)
    assert rule.last_moved == int(moved.sum())
    # only unhappy agents move; here they outnumber the empty cells (42 % unhappy at the random start against 25 %
    # empty), so exactly |E| of them do — move_agent_single! is a no-op once no empty cell is left (discrete.jl:283-284)
    assert not (moved & ~unhappy).any()
    assert int(moved.sum()) == min(int(unhappy.sum()), ncell - n)
    assert np.array_equal(mood, want)                          # moods are those of the snapshot
    occupied = np.zeros(ncell, dtype=bool)
    occupied[lin(pos)] = True
    assert not occupied[lin(new[moved])].any()                 # every destination was empty at the start of the step
    assert len(np.unique(lin(new[moved]))) == int(moved.sum())  # no two movers share a destination
    assert new.min() >= 1 and new.max() <= side
    assert np.array_equal(m.column("group"), group)
    # the next happiness pass sees the relocated population
    m.model_step = ab.SchellingRule(min_to_be_happy=3, relocate=False)
    m.step(1)
    assert np.array_equal(m.column("mood"), numpy_moods(new))
    m.close()
