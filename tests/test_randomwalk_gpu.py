"""Bulk randomwalk! against the oracle's restatement on the same counter stream.
Grids (src/spaces/walk.jl:154-214): positions exact, every call, including the sequential `ifempty` /
`force_motion` semantics; the scenario of the reference's own testset (test/grid_space_tests.jl:407-489) is
replayed with its assertions.  ContinuousSpace (walk.jl:389-408): new velocity and position within 1e-13
relative (f64) / 1e-5 (f32) — the normals go through log / cos / sin, whose device and libm results may
differ in the last place — plus the invariants |vel| == r and displacement == vel."""
import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import uniform
from oracle.pyoracle import (ContinuousOracle, GridOracle, continuous_randomwalk_directions, grid_offsets_at_radius,
                             grid_randomwalk)
from test_grid_gpu import make, random_positions

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("metric,r,expect", [("chebyshev", 1, 8), ("manhattan", 1, 4), ("chebyshev", 2, 16),
                                             ("manhattan", 2.7, 8), ("chebyshev", 0, 1)])
def test_offsets_at_radius_sizes(metric, r, expect):
    # "for Chebyshev metric and r=1 ... any of the 8 surrounding cells. For Manhattan ... 4" (walk.jl:141-143)
    assert len(grid_offsets_at_radius((20, 20), True, metric, r)) == expect


@pytest.mark.parametrize("single,ifempty,force", [(False, False, False), (False, True, False), (False, True, True),
                                                  (False, False, True), (True, True, False), (True, True, True)])
@pytest.mark.parametrize("metric,periodic,r", [("chebyshev", True, 1), ("manhattan", (True, False), 2),
                                               ("chebyshev", False, 1.5)])
def test_grid_randomwalk_equals_oracle(single, ifempty, force, metric, periodic, r):
    dims = (11, 8)
    pos = random_positions(dims, 60, 4, single)              # dense: most cells around an agent are taken
    m, o = make(dims, periodic, metric, single, pos)
    n = len(pos)
    rng = np.random.default_rng(9)
    moved = 0
    for tick in range(4):
        ids = None if tick == 0 else rng.permutation(n)[: n // 2] + 1
        before = m.pos.copy()
        m.randomwalk(r, seed=1000 + tick, ids=ids, ifempty=ifempty, force_motion=force)
        grid_randomwalk(o, range(1, n + 1) if ids is None else [int(i) for i in ids], r, 1000 + tick, ifempty, force)
        assert np.array_equal(m.pos, o.positions())
        moved += int(np.any(m.pos != before, axis=1).sum())
        if single:
            assert len({tuple(p) for p in m.pos}) == n         # still one agent per position
    assert moved > 0
    m.close()


@pytest.mark.parametrize("single", [False, True])
@pytest.mark.parametrize("metric,periodic", [("chebyshev", True), ("manhattan", (True, False)), ("chebyshev", False)])
def test_grid_randomwalk_force_motion_as_written(single, metric, periodic):
    """force_motion = 2: the reference's code as written — the empty POSITION found by random_empty_pos_in_offsets is
    passed to walk! as a direction (walk.jl:170-172, :193-195).  Both readings are offered; both are checked."""
    dims = (11, 8)
    pos = random_positions(dims, 60, 4, single)
    m, o = make(dims, periodic, metric, single, pos)
    m2, o2 = make(dims, periodic, metric, single, pos)
    n = len(pos)
    for tick in range(3):
        m.randomwalk(1, seed=77 + tick, ifempty=True, force_motion=2)
        grid_randomwalk(o, range(1, n + 1), 1, 77 + tick, True, True, literal=True)
        assert np.array_equal(m.pos, o.positions())
        m2.randomwalk(1, seed=77 + tick, ifempty=True, force_motion=1)
        grid_randomwalk(o2, range(1, n + 1), 1, 77 + tick, True, True)
        assert np.array_equal(m2.pos, o2.positions())
        if single:
            assert len({tuple(p) for p in m.pos}) == n
    assert not np.array_equal(m.pos, m2.pos)                  # the two readings are different rules
    m.close(); m2.close()


def test_grid_randomwalk_3d_and_errors():
    dims = (6, 5, 4)
    pos = random_positions(dims, 30, 2, True)
    m, o = make(dims, True, "chebyshev", True, pos)
    m.randomwalk(1, seed=5, force_motion=True)
    grid_randomwalk(o, range(1, len(pos) + 1), 1, 5, True, True)
    assert np.array_equal(m.pos, o.positions())
    m.close()
    m, _ = make((6, 6), True, "euclidean", False, [(1, 1)])
    with pytest.raises(ab.AgentsError, match="Euclidean metric are not defined"):
        m.randomwalk(1)
    m.close()


def test_reference_randomwalk_scenarios():
    # test/grid_space_tests.jl:407-431: a lone agent always moves, by exactly r under the metric
    for metric, r in (("chebyshev", 1), ("manhattan", 1), ("chebyshev", 3), ("manhattan", 2)):
        m, _ = make((100, 100), True, metric, False, [(50, 50)])
        for k in range(20):
            p0 = m.pos[0].astype(int)
            m.randomwalk(r, seed=k)
            d = np.abs(m.pos[0].astype(int) - p0)
            assert (d.max() if metric == "chebyshev" else d.sum()) == r
        m.close()
    # grid_space_tests.jl:462-489: agent 1 at (50,50) walled in by 8 neighbours cannot move, even with force_motion;
    # once a neighbour has been moved away, the freed cell is the only place agent 1 can go
    ring = [(49, 49), (50, 49), (51, 49), (49, 50), (51, 50), (49, 51), (50, 51), (51, 51)]
    for single in (True, False):
        m, _ = make((100, 100), True, "chebyshev", single, [(50, 50), (49, 50)] + [p for p in ring if p != (49, 50)])
        m.randomwalk(1, seed=1, ids=[1], force_motion=True)
        assert tuple(m.pos[0]) == (50, 50)
        m.randomwalk(1, seed=2, ids=[2], force_motion=True)
        assert tuple(m.pos[1]) != (49, 50) and np.abs(m.pos[1].astype(int) - (49, 50)).max() == 1
        m.randomwalk(1, seed=3, ids=[1], force_motion=True)     # the only empty target
        assert tuple(m.pos[0]) == (49, 50)
        m.close()


@pytest.mark.parametrize("D", [2, 3])
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-13), (np.float32, 1e-5)])
@pytest.mark.parametrize("r", [None, 0.7])
def test_continuous_randomwalk(D, dtype, tol, r):
    n = 500
    extent = (10.0,) * D
    pos = (uniform(n, D, 3) * 10.0).astype(dtype)
    vel = (uniform(n, D, 4) * 2.0 - 1.0).astype(dtype)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=1.0, periodic=True, dtype=dtype))
    m.add_agents(pos, vel)
    o = ContinuousOracle(extent, spacing=1.0, periodic=True, dtype=dtype)
    o.add_agents(pos, vel)
    for tick, ids in enumerate((None, np.arange(n, 0, -3))):
        idl = list(range(1, n + 1)) if ids is None else [int(i) for i in ids]
        p_before, v_before = m.pos.copy(), m.vel.copy()
        want_dir = continuous_randomwalk_directions(v_before, idl, r, 77 + tick, dtype)
        m.randomwalk(r, seed=77 + tick, ids=ids)
        got_v, got_p = m.vel, m.pos
        sel = np.asarray(idl) - 1
        assert np.max(np.abs(got_v[sel] - want_dir)) <= tol * max(1.0, np.abs(want_dir).max())
        speed = np.sqrt((got_v[sel].astype(np.float64) ** 2).sum(axis=1))
        want_speed = np.sqrt((v_before[sel].astype(np.float64) ** 2).sum(axis=1)) if r is None else r
        assert np.allclose(speed, want_speed, rtol=10 * tol)
        for i, d in zip(idl, want_dir):
            o.walk(i, d)
        want_p = o.state()[0]
        err = np.abs(got_p - want_p)
        err = np.minimum(err, 10.0 - err)                      # a wrap decided one ulp apart is the same point
        assert err.max() <= 2 * tol * 10.0
        rest = np.setdiff1d(np.arange(n), sel)
        assert np.array_equal(got_p[rest], p_before[rest]) and np.array_equal(got_v[rest], v_before[rest])
        o.set_vel(got_v)
        # keep the two states identical for the next round
        for i in idl:
            o.move_agent_to(i, got_p[i - 1])
    # the index follows the walk
    want = [np.sort(o.nearby_ids_agent(i, 1.0, "exact")) for i in (1, 7, 300)]
    off, ids2 = m.nearby_ids_batch(ids=np.array([1, 7, 300]), r=1.0, search="exact")
    for k in range(3):
        assert np.array_equal(np.sort(ids2[off[k]:off[k + 1]]), want[k])
    m.close()


def test_continuous_randomwalk_errors():
    m = ab.StandardABM(ab.ContinuousSpace((10.0, 10.0), spacing=1.0))
    m.add_agents(np.array([[1.0, 1.0], [2.0, 2.0]]), np.array([[0.0, 0.0], [1.0, 0.0]]))
    with pytest.raises(ab.AgentsError, match="displacement must be larger than 0"):
        m.randomwalk(0.0)
    with pytest.raises(ab.AgentsError, match="displacement must be larger than 0"):
        m.randomwalk()                                          # agent 1 is at rest: default r = 0 (walk.jl:394-396)
    m.close()


@pytest.mark.parametrize("D", [2, 3])
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-13), (np.float32, 1e-5)])
@pytest.mark.parametrize("r", [None, 1.3])
def test_continuous_randomwalk_with_angles(D, dtype, tol, r):
    # walk.jl:236-330: velocity rotated by the drawn angles, scaled to r, stored, walked
    from oracle.pyoracle import continuous_randomwalk_angle_directions
    n = 400
    extent = (10.0,) * D
    pos = (uniform(n, D, 3) * 10.0).astype(dtype)
    vel = (uniform(n, D, 4) * 2.0 - 1.0).astype(dtype)
    vel[5, 0] = 0.0                                            # first nonzero component is not the first one
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=1.0, dtype=dtype))
    m.add_agents(pos, vel)
    o = ContinuousOracle(extent, spacing=1.0, dtype=dtype)
    o.add_agents(pos, vel)
    ids = np.arange(1, n + 1, 2)
    th = (uniform(len(ids), 1, 8)[:, 0] * 2 - 1) * np.pi       # Uniform(-π, π)
    ph = np.arccos(uniform(len(ids), 1, 9)[:, 0] * 2 - 1)      # Arccos(-1, 1), walk.jl:359-376
    want = continuous_randomwalk_angle_directions(vel, [int(i) for i in ids], th, ph, r).astype(dtype)
    m.randomwalk_angles(th, ph if D == 3 else None, r=r, ids=ids)
    got_v, got_p = m.vel, m.pos
    sel = ids - 1
    assert np.max(np.abs(got_v[sel] - want)) <= tol * max(1.0, np.abs(want).max())
    speed = np.sqrt((got_v[sel].astype(np.float64) ** 2).sum(axis=1))
    want_speed = np.sqrt((vel[sel].astype(np.float64) ** 2).sum(axis=1)) if r is None else r
    assert np.allclose(speed, want_speed, rtol=20 * tol)
    # the polar angle is the angle between the old and the new velocity (walk.jl:347-357 docstring)
    cosang = (got_v[sel].astype(np.float64) * vel[sel]).sum(axis=1) / (speed * np.sqrt((vel[sel].astype(np.float64) ** 2).sum(axis=1)))
    assert np.allclose(cosang, np.cos(th), atol=50 * tol)
    for i, d in zip(ids, want):
        o.walk(int(i), d)
    err = np.abs(got_p - o.state()[0])
    err = np.minimum(err, 10.0 - err)
    assert err.max() <= 2 * tol * 10.0
    rest = np.setdiff1d(np.arange(n), sel)
    assert np.array_equal(got_p[rest], pos[rest]) and np.array_equal(got_v[rest], vel[rest])
    # callables as distributions, all agents
    m.randomwalk_angles(lambda k: np.zeros(k), (lambda k: np.zeros(k)) if D == 3 else None)
    assert np.max(np.abs(m.vel - got_v)) <= tol              # zero angles: the velocity is kept
    m.close()


def test_randomwalk_angles_errors():
    m = ab.StandardABM(ab.ContinuousSpace((10.0, 10.0, 10.0), spacing=1.0))
    m.add_agents(np.full((2, 3), 5.0), np.array([[0.0, 0.0, 0.0], [1.0, 0.0, 0.0]]))
    with pytest.raises(ab.AgentsError, match="azimuthal"):
        m.randomwalk_angles(np.zeros(2))
    with pytest.raises(ab.AgentsError, match="resting agent"):
        m.randomwalk_angles(np.zeros(2), np.zeros(2))
    m.close()
