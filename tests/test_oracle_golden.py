"""Pins the CPU oracle against every golden vector / known-answer test the
reference (Agents.jl v7.0.3) holds for the hot path.
Each test cites the reference test file:line it restates.  CPU only.
"""
import math

import numpy as np
import pytest

from oracle.pyoracle import ContinuousOracle, GridOracle, OracleError, grid_offsets, jl_mod


# --------------------------------------------------------------------------
# test/continuous_space_tests.jl
# --------------------------------------------------------------------------
def test_space_divisibility():
    # src/spaces/continuous.jl:99-101
    with pytest.raises(OracleError):
        ContinuousOracle((1.0, 1.0), spacing=0.3)
    m = ContinuousOracle((1, 1))  # default spacing = min(extent)/20
    assert m.dims == (20, 20)
    # flock.jl:56 — extent 100, spacing 5/1.5 → 30×30
    m = ContinuousOracle((100, 100), spacing=5.0 / 1.5)
    assert m.dims == (30, 30)
    m = ContinuousOracle((1, 1, 1), spacing=0.25, periodic=False)
    assert m.dims == (4, 4, 4)


def test_euler_move_and_outside_extent():
    # continuous_space_tests.jl:23-47
    m = ContinuousOracle((1, 1))
    a = m.add_agent((0.51, 0.51), (0.2, 0.1))
    m.move_agent_dt(a, 1)
    pos, _ = m.state()
    assert tuple(pos[0]) == (0.71, 0.61)
    m.move_agent_to(a, (0.51, 0.51))
    m.move_agent_to(a, (0.51, 0.51))
    assert tuple(m.state()[0][0]) == (0.51, 0.51)
    with pytest.raises(OracleError, match="outside space extent"):
        m.move_agent_to(a, (1.5, 1.5))


def test_periodic_wrap_edge_case_853():
    # continuous_space_tests.jl:52-58: result is exactly (0.5, 1.0)
    m = ContinuousOracle((1, 1), periodic=True)
    m.add_agent((0.4936805014512577, 0.007750092779602313), (0.6319498548742291, -0.7750092779602327))
    m.move_agent_dt(1, 0.01)
    assert tuple(m.state()[0][0]) == (0.5, 1.0)
    # and the agent sits in the last cell row (pos == extent is legal, A2)
    assert tuple(m.pos2cell(m.state()[0])[0]) == (10, 20)


def test_julia_mod_semantics():
    assert jl_mod(-1.3877787807814457e-17, 1.0) == 1.0
    assert jl_mod(14.0, 12.0) == 2.0
    assert jl_mod(10.0, 10.0) == 0.0
    assert jl_mod(-0.25, 1.0) == 0.75
    assert math.copysign(1.0, jl_mod(-0.0, 1.0)) == 1.0  # copysign(r, y)


def test_tuple_agent_motion():
    # continuous_space_tests.jl:65-80
    m = ContinuousOracle((1, 1))
    m.add_agent((0.0, 0.0), (0.1, 0.0))
    m.move_agent_dt(1, 1.0)
    assert tuple(m.state()[0][0]) == (0.1, 0.0)
    m.move_agent_to(1, (0.5, 0.2))
    m.walk(1, (0.2, 0.0))
    assert tuple(m.state()[0][0]) == (0.5 + 2 * 0.1, 0.2)


def test_nearby_ids_cell_centres():
    # continuous_space_tests.jl:184-211
    m = ContinuousOracle((1.0, 1.0), spacing=0.1)
    # Julia comprehension [f(i,j) for i in 0:9, j in 0:9] iterates i fastest
    centres = [(0.05 + 0.1 * i, 0.05 + 0.1 * j) for j in range(10) for i in range(10)]
    m.add_agents(centres, np.zeros((100, 2)))
    centre = (0.45, 0.45)
    centre_id = 25 + 20  # Julia id of the agent at (0.45, 0.45): i=4, j=4 → 4 + 10*4 + 1
    assert tuple(m.state()[0][centre_id - 1]) == pytest.approx(centre)
    rs = [x * 0.1 + 0.001 for x in (1, 2, 3.4)]
    ns = (4, 12, 36)
    for r, nn in zip(rs, ns):
        assert len(m.nearby_ids(centre, r, "exact")) == nn + 1
        assert len(m.nearby_ids_agent(centre_id, r, "exact")) == nn
        assert len(m.nearby_ids(centre, r)) >= nn + 1
        assert len(m.nearby_ids_agent(centre_id, r)) >= nn
        # the reference test itself uses `center_id = 25` (the agent at (0.45, 0.25))
        assert len(m.nearby_ids_agent(25, r, "exact")) == nn
        assert len(m.nearby_ids_agent(25, r)) >= nn


def test_nearby_ids_within_same_cell():
    # continuous_space_tests.jl:213-236 (collect order pinned too)
    m = ContinuousOracle((1.0, 1.0), spacing=0.1)
    r0, r1 = 0.01, 0.08
    a = m.add_agent((0.51, 0.51), (0.0, 0.0))
    b = m.add_agent((0.51 + r1, 0.51), (0.0, 0.0))
    cells = m.pos2cell(m.state()[0])
    assert tuple(cells[0]) == tuple(cells[1])
    assert list(m.nearby_ids_agent(a, r0)) == [2]
    assert list(m.nearby_ids_agent(a, r0, "exact")) == []
    assert list(m.nearby_ids_agent(a, r1)) == [2]
    assert list(m.nearby_ids_agent(a, r1, "exact")) == [2]
    pa = m.state()[0][0]
    assert list(m.nearby_ids(pa, r0)) == [1, 2]
    assert list(m.nearby_ids(pa, r0, "exact")) == [1]
    assert list(m.nearby_ids(pa, r1)) == [1, 2]
    assert list(m.nearby_ids(pa, r1, "exact")) == [1, 2]
    assert b == 2


PAIR_POS = [
    (7.074386436066224, 4.963014649338054),
    (5.831962448496828, 4.926297135685473),
    (5.122087781793935, 5.300031210394806),
    (3.9715633336430156, 4.8106570045816675),
]


def test_interacting_pairs_standard():
    # continuous_space_tests.jl:240-269
    m = ContinuousOracle((10, 10), spacing=0.2, periodic=False)
    m.add_agents(PAIR_POS, np.zeros((4, 2)))
    pairs = {tuple(p) for p in m.interacting_pairs_all(2.0)}
    assert len(pairs) == 5
    assert (3, 6) not in pairs
    m2 = ContinuousOracle((10, 10), spacing=0.1, periodic=False)
    m2.add_agents(PAIR_POS, np.zeros((4, 2)))
    pairs = {tuple(p) for p in m2.interacting_pairs_all(2.5)}
    assert len(pairs) == 5
    assert (1, 4) not in pairs


def test_interacting_pairs_fix_288():
    # continuous_space_tests.jl:303-320
    m = ContinuousOracle((1, 1), spacing=0.1, periodic=True)
    m.add_agents([(0.01, 0.01), (0.2, 0.2), (0.5, 0.5)], np.zeros((3, 2)))
    for r in (0.29, 0.3):
        pairs = [tuple(p) for p in m.interacting_pairs_all(r)]
        assert pairs == [(1, 2)]


def test_nearest_neighbor():
    # continuous_space_tests.jl:323-349
    m = ContinuousOracle((1, 1), spacing=0.1, periodic=True)
    m.add_agents([(0.01, 0.01), (0.2, 0.01), (0.2, 0.2), (0.5, 0.5)], np.zeros((4, 2)))
    nn = [m.nearest_neighbor(i, math.sqrt(2)) for i in (1, 2, 3, 4)]
    assert nn == [2, 1, 2, 3]
    m = ContinuousOracle((1, 1), spacing=0.02, periodic=False)
    m.add_agents([(0.30386961108678245, 0.17690859963285543), (0.33020500850053125, 0.178391311081449)],
                 np.zeros((2, 2)))
    assert m.nearest_neighbor(1, 1e-100) is None
    assert m.nearest_neighbor(2, 1e-100) is None


def test_interacting_pairs_nearest_and_collision():
    # continuous_space_tests.jl:244-265: 4 agents, r = 2.0, :nearest -> exactly one pair
    from oracle.pyoracle import elastic_collision

    pos = [(7.074386436066224, 4.963014649338054), (5.831962448496828, 4.926297135685473),
           (5.122087781793935, 5.300031210394806), (3.9715633336430156, 4.8106570045816675)]
    m = ContinuousOracle((10, 10), spacing=0.1, periodic=False)
    m.add_agents(pos, np.zeros((4, 2)))
    assert m.interacting_pairs_nearest(2.0) == [(2, 3)]
    # elastic_collision! (continuous.jl:333-366): head-on equal masses swap velocities; disks moving apart do
    # not collide; an immovable disk reflects; coincident disks do nothing; two immovable disks do nothing
    assert elastic_collision((0, 0), (1, 0), (1, 0), (-1, 0)) == (True, (-1.0, 0.0), (1.0, 0.0))
    assert elastic_collision((0, 0), (-1, 0), (1, 0), (1, 0))[0] is False
    h, v1, v2 = elastic_collision((0, 0), (0, 0), (1, 0), (-1, 0.5), math.inf, 1.33)
    assert h and v1 == (0.0, 0.0) and v2 == (1.0, 0.5)
    assert elastic_collision((2, 2), (1, 0), (2, 2), (-1, 0))[0] is False
    assert elastic_collision((0, 0), (0, 0), (1, 0), (0, 0), math.inf, math.inf)[0] is False
    # kinetic energy and momentum of a movable pair are conserved
    h, v1, v2 = elastic_collision((0.1, 0.2), (0.3, -0.1), (0.4, 0.35), (-0.2, 0.05), 2.0, 0.7)
    assert h
    k0 = 2.0 * (0.3 ** 2 + 0.1 ** 2) + 0.7 * (0.2 ** 2 + 0.05 ** 2)
    k1 = 2.0 * (v1[0] ** 2 + v1[1] ** 2) + 0.7 * (v2[0] ** 2 + v2[1] ** 2)
    assert abs(k0 - k1) < 1e-15
    assert abs(2.0 * v1[0] + 0.7 * v2[0] - (2.0 * 0.3 - 0.7 * 0.2)) < 1e-15


def test_walk_continuous():
    # continuous_space_tests.jl:351-378
    m = ContinuousOracle((12, 10), periodic=False)
    a = m.add_agent((0.0, 0.0), (0.0, 0.0))
    m.walk(a, (1.0, 1.0))
    assert tuple(m.state()[0][0]) == (1.0, 1.0)
    m.walk(a, (15.0, 1.0))
    assert tuple(m.state()[0][0]) == (np.nextafter(12.0, -np.inf), 2.0)
    m = ContinuousOracle((12, 10, 5), spacing=0.2)
    a = m.add_agent((0.0, 0.0, 0.0), (0.0, 0.0, 0.0))
    m.walk(a, (1.0, 1.0, 1.0))
    assert tuple(m.state()[0][0]) == (1.0, 1.0, 1.0)
    m.walk(a, (15.0, 1.2, 3.9))
    assert tuple(m.state()[0][0]) == (4.0, 2.2, 4.9)
    m = ContinuousOracle((12, 10), periodic=True)
    a = m.add_agent((11.0, 9.0), (3.0, 1.0))
    m.move_agent_dt(a, 1.0)
    p = m.state()[0][0]
    assert p[0] == 2 and p[1] == 0.0


def test_float32_agents():
    # continuous_space_tests.jl:606-637 — Float32 agents work end to end
    m = ContinuousOracle((1.0, 1.0), spacing=0.1, dtype=np.float32)
    m.add_agent((0.51, 0.51), (0.2, 0.1))
    m.move_agent_dt(1, 1.0)
    p = m.state()[0][0]
    assert p.dtype == np.float32
    assert tuple(p) == (np.float32(0.51) + np.float32(0.2), np.float32(0.51) + np.float32(0.1))
    assert len(m.nearby_ids(p, 0.1)) == 1


# --------------------------------------------------------------------------
# test/grid_space_tests.jl
# --------------------------------------------------------------------------
def test_offset_counts():
    # 2-D euclidean 5/13/29, 3-D 7/33/123, chebyshev (2r+1)^D
    for r0, n in zip((1, 2, 3), (5, 13, 29)):
        assert len(grid_offsets((100, 100), True, "euclidean", r0)) == n
    for r0, n in zip((1, 2, 3), (7, 33, 123)):
        assert len(grid_offsets((100, 100, 100), True, "euclidean", r0)) == n
    assert len(grid_offsets((100, 100), True, "chebyshev", 2)) == 25
    assert len(grid_offsets((100, 100), True, "manhattan", 2)) == 13
    # product order: first dimension fastest
    o = grid_offsets((100, 100), True, "chebyshev", 1)
    assert [tuple(x) for x in o[:4]] == [(-1, -1), (0, -1), (1, -1), (-1, 0)]
    # r cached by floor(r+1): r = 3.4 behaves as r0 = 3
    assert len(grid_offsets((100, 100), True, "euclidean", 3.4)) == 29


@pytest.mark.parametrize("single", [False, True])
def test_grid_distances(single):
    # grid_space_tests.jl:134-149 — distances on grids use the same formulas
    # as the continuous space (src/spaces/utilities.jl:84-116)
    m = ContinuousOracle((12, 10), spacing=1.0, periodic=True)
    assert m.euclidean_distance((1, 6), (11, 4)) == pytest.approx(2.82842712)
    m = ContinuousOracle((12, 10), spacing=1.0, periodic=False)
    assert m.euclidean_distance((1, 6), (11, 4)) == pytest.approx(10.198039)
    m = ContinuousOracle((10, 10), spacing=1.0, periodic=(False, True))
    assert m.euclidean_distance((1, 1), (9, 9)) == pytest.approx(8.24621125)


@pytest.mark.parametrize("single", [False, True])
@pytest.mark.parametrize("periodic", [False, True, (True, False)])
@pytest.mark.parametrize("metric", ["euclidean", "manhattan", "chebyshev"])
def test_grid_nearby_r1(single, periodic, metric):
    # grid_space_tests.jl:168-231
    g = GridOracle((5, 5), periodic=periodic, metric=metric, single=single)
    s = lambda a: sorted(tuple(int(v) for v in x) for x in a)
    if metric in ("euclidean", "manhattan"):
        assert s(g.nearby_positions((2, 2))) == sorted([(2, 1), (1, 2), (3, 2), (2, 3)])
        exp = {False: [(1, 2), (2, 1)], True: [(1, 2), (2, 1), (1, 5), (5, 1)],
               (True, False): [(1, 2), (2, 1), (5, 1)]}[periodic]
        assert s(g.nearby_positions((1, 1))) == sorted(exp)
    else:
        assert s(g.nearby_positions((2, 2))) == [(1, 1), (1, 2), (1, 3), (2, 1), (2, 3), (3, 1), (3, 2), (3, 3)]
        exp = {False: [(1, 2), (2, 1), (2, 2)],
               True: [(1, 2), (1, 5), (2, 1), (2, 2), (2, 5), (5, 1), (5, 2), (5, 5)],
               (True, False): [(1, 2), (2, 1), (2, 2), (5, 1), (5, 2)]}[periodic]
        assert s(g.nearby_positions((1, 1))) == exp
    g.add_agent((1, 1))
    a = g.add_agent((2, 1))
    g.add_agent((3, 2))
    g.add_agent((2, 5))
    near_agent = sorted(g.nearby_ids_agent(a))
    near_pos = sorted(g.nearby_ids((2, 1)))
    assert 2 in near_pos and 2 not in near_agent
    assert near_pos == sorted(near_agent + [2])
    if metric in ("euclidean", "manhattan"):
        exp = {False: [1], True: [1, 4], (True, False): [1]}[periodic]
    else:
        exp = {False: [1, 3], True: [1, 3, 4], (True, False): [1, 3]}[periodic]
    assert near_agent == exp


@pytest.mark.parametrize("single", [False, True])
@pytest.mark.parametrize("periodic", [False, True, (True, False)])
def test_grid_nearby_r34(single, periodic):
    # grid_space_tests.jl:233-245 (fill_space! puts one agent in every cell)
    res = {}
    for metric in ("euclidean", "manhattan", "chebyshev"):
        g = GridOracle((9, 9), periodic=periodic, metric=metric, single=single)
        g.add_agents([(x, y) for y in range(1, 10) for x in range(1, 10)])
        res[metric] = (len(g.nearby_positions((5, 5), 3.4)), len(g.nearby_ids((5, 5), 3.4)))
    assert res["euclidean"][0] == res["manhattan"][0] + 4
    assert res["chebyshev"][0] == 7 ** 2 - 1
    assert res["euclidean"][1] == res["manhattan"][1] + 4
    assert res["euclidean"][1] == res["euclidean"][0] + 1
    assert res["chebyshev"][1] == 7 ** 2


@pytest.mark.parametrize("periodic", [False, True, (True, False, True)])
def test_grid_radius_larger_than_extent(periodic):
    # grid_space_tests.jl:247-256
    g = GridOracle((3, 4, 15), periodic=periodic)
    allp = [tuple(int(v) for v in x) for x in g.nearby_positions((2, 2, 7), 5)]
    assert len(allp) == 131
    assert len(set(allp)) == 131
    arr = np.array(allp)
    assert tuple(arr.min(0)) == (1, 1, 2) and tuple(arr.max(0)) == (3, 4, 12)


def test_grid_single_refuses_second_agent():
    # grid_space_tests.jl:51-53, grid_single.jl:47-52
    g = GridOracle((5, 5), single=True)
    g.add_agent((2, 2))
    with pytest.raises(OracleError, match="occupied"):
        g.add_agent((2, 2))


def test_grid_walk():
    # grid_space_tests.jl:333-405
    g = GridOracle((3, 3), periodic=True)
    a = g.add_agent((1, 1))
    seq = [((0, 1), (1, 2)), ((1, 1), (2, 3)), ((1, 0), (3, 3)), ((2, 0), (2, 3)), ((1, -1), (3, 2)),
           ((0, -1), (3, 1)), ((-1, -1), (2, 3)), ((-1, 0), (1, 3)), ((0, -8), (1, 1))]
    for d, exp in seq:
        g.walk(a, d, ifempty=False)
        assert tuple(g.positions()[a - 1]) == exp
    a = g.add_agent((1, 1))
    g.add_agent((2, 2))
    g.walk(a, (1, 1), ifempty=True)
    assert tuple(g.positions()[a - 1]) == (1, 1)
    g.walk(a, (1, 0), ifempty=True)
    assert tuple(g.positions()[a - 1]) == (2, 1)
    g = GridOracle((3, 3), periodic=False)
    a = g.add_agent((1, 1))
    seq = [((0, 1), (1, 2)), ((1, 1), (2, 3)), ((1, 0), (3, 3)), ((1, 0), (3, 3)), ((-5, 0), (1, 3)),
           ((-1, -1), (1, 2))]
    for d, exp in seq:
        g.walk(a, d)
        assert tuple(g.positions()[a - 1]) == exp
    g = GridOracle((3, 3), periodic=(True, False))
    a = g.add_agent((1, 1))
    seq = [((0, 1), (1, 2)), ((1, 1), (2, 3)), ((1, 0), (3, 3)), ((1, 0), (1, 3)), ((0, 1), (1, 3)),
           ((-1, -3), (3, 1))]
    for d, exp in seq:
        g.walk(a, d)
        assert tuple(g.positions()[a - 1]) == exp
    g = GridOracle((5, 5), single=True)
    a = g.add_agent((3, 3))
    g.walk(a, (1, 1))
    assert tuple(g.positions()[0]) == (4, 4)
    g = GridOracle((3, 3, 2))
    a = g.add_agent((1, 1, 1))
    g.walk(a, (1, 1, 1))
    assert tuple(g.positions()[0]) == (2, 2, 2)


# --------------------------------------------------------------------------
# Rules: properties the reference implies (no reference test pins a trajectory)
# --------------------------------------------------------------------------
FLOCK = dict(speed=1.0, cohere_factor=0.1, separation=2.0, separate_factor=0.25, match_factor=0.04,
             visual_distance=5.0)  # examples/flock.jl:45-55


def _flock_model(n=300, seed=42):
    rng = np.random.default_rng(seed)
    m = ContinuousOracle((100, 100), spacing=5.0 / 1.5)
    pos = rng.random((n, 2)) * 100
    vel = rng.random((n, 2)) * 2 - 1
    m.add_agents(pos, vel)
    return m


def test_flock_runs_and_normalises():
    # test/jld2_tests.jl:97-148 only checks the flock model survives 100 steps
    m = _flock_model()
    m.flock_step_sync(FLOCK, nsteps=100)
    pos, vel = m.state()
    assert np.all(pos >= 0) and np.all(pos <= 100)
    assert np.allclose(np.linalg.norm(vel, axis=1), 1.0, atol=1e-14)
    m2 = _flock_model()
    m2.flock_step_async(FLOCK, nsteps=100)
    pos2, vel2 = m2.state()
    assert np.allclose(np.linalg.norm(vel2, axis=1), 1.0, atol=1e-14)
    # index stays consistent with positions after 100 steps of re-bucketing
    cells = m.pos2cell(pos)
    for i in (0, 17, 299):
        assert (i + 1) in m.cell_ids(cells[i])


def test_flock_sync_isolated_bird_moves_straight():
    m = ContinuousOracle((100, 100), spacing=5.0)
    m.add_agent((50.0, 50.0), (3.0, 4.0))
    cnt = m.flock_step_sync(FLOCK, return_counts=True)
    pos, vel = m.state()
    assert cnt[0] == 0
    assert tuple(vel[0]) == (3.0 / 5.0, 4.0 / 5.0)
    assert tuple(pos[0]) == (50.0 + 0.6, 50.0 + 0.8)


def test_flock_per_agent_params_equal_uniform():
    m1, m2 = _flock_model(100), _flock_model(100)
    keys = ("speed", "cohere_factor", "separation", "separate_factor", "match_factor", "visual_distance")
    pa = np.stack([np.full(100, FLOCK[k]) for k in keys])
    m1.flock_step_sync(FLOCK, nsteps=3)
    m2.flock_step_sync(FLOCK, nsteps=3, per_agent=pa)
    assert np.array_equal(m1.state()[0], m2.state()[0])


def test_pairforce_momentum_conserved():
    # celllistmap.jl:151-152: f_i += fij, f_j -= fij ⇒ Σ m v is conserved
    rng = np.random.default_rng(1)
    n = 200
    m = ContinuousOracle((200.0, 200.0), spacing=20.0)
    m.add_agents(rng.random((n, 2)) * 200, rng.normal(size=(n, 2)) * 100)
    rad = (0.5 + 0.9 * rng.random(n)) * 10
    k = 10 + 20 * rng.random(n)
    mass = 10 + 100 * rng.random(n)
    p0 = (m.state()[1] * mass[:, None]).sum(0)
    m.pairforce_step(rad, k, mass, 0.001, 20.0, nsteps=20)
    p1 = (m.state()[1] * mass[:, None]).sum(0)
    assert np.allclose(p0, p1, rtol=1e-9)


def test_schelling_happiness_small():
    # docs/src/tutorial.jl:351-372 on a hand-checkable 3×3 non-periodic grid
    g = GridOracle((3, 3), periodic=False, single=True)
    #  y=3:  . 2 .
    #  y=2:  1 1 2
    #  y=1:  1 . 2
    pts = [(1, 1), (1, 2), (2, 2), (3, 1), (3, 2), (2, 3)]
    grp = [1, 1, 1, 2, 2, 2]
    g.add_agents(pts)
    cnt, mood = g.schelling_happiness(grp, 2)
    assert list(cnt) == [2, 2, 2, 1, 2, 1]
    assert list(mood) == [1, 1, 1, 0, 1, 0]


def test_randomwalk_restatement_reference_scenarios():
    # test/grid_space_tests.jl:407-431, 462-489 on the oracle's restatement of randomwalk! (walk.jl:154-214)
    from oracle.pyoracle import grid_offsets_at_radius, grid_randomwalk
    assert len(grid_offsets_at_radius((20, 20), True, "chebyshev", 1)) == 8
    assert len(grid_offsets_at_radius((20, 20), True, "manhattan", 1)) == 4
    assert len(grid_offsets_at_radius((20, 20, 20), False, "chebyshev", 1)) == 26
    for metric, r in (("chebyshev", 1), ("manhattan", 1), ("chebyshev", 3), ("manhattan", 2)):
        g = GridOracle((100, 100), periodic=True, metric=metric)
        g.add_agents([(50, 50)])
        for k in range(20):
            p0 = g.positions()[0]
            grid_randomwalk(g, [1], r, k)
            d = np.abs(g.positions()[0] - p0)
            assert (d.max() if metric == "chebyshev" else d.sum()) == r
    ring = [(49, 49), (50, 49), (51, 49), (49, 50), (51, 50), (49, 51), (50, 51), (51, 51)]
    for single in (True, False):
        g = GridOracle((100, 100), periodic=True, metric="chebyshev", single=single)
        g.add_agents([(50, 50), (49, 50)] + [p for p in ring if p != (49, 50)])
        grid_randomwalk(g, [1], 1, 1, True, True)
        assert tuple(g.positions()[0]) == (50, 50)
        grid_randomwalk(g, [2], 1, 2, True, True)
        assert tuple(g.positions()[1]) != (49, 50)
        grid_randomwalk(g, [1], 1, 3, True, True)
        assert tuple(g.positions()[0]) == (49, 50)


def test_randomwalk_continuous_restatement():
    # walk.jl:389-408: |direction| == r, isotropic (mean of many directions near zero)
    from oracle.pyoracle import continuous_randomwalk_directions
    vel = np.ones((4000, 2))
    d = continuous_randomwalk_directions(vel, list(range(1, 4001)), 0.5, 11)
    assert np.allclose(np.sqrt((d ** 2).sum(axis=1)), 0.5, rtol=1e-14)
    assert np.abs(d.mean(axis=0)).max() < 0.03
    d2 = continuous_randomwalk_directions(vel, [1, 2], None, 11)
    assert np.allclose(np.sqrt((d2 ** 2).sum(axis=1)), np.sqrt(2.0), rtol=1e-14)


def test_rotate_restatement():
    # walk.jl:347-357 docstring: the rotated vector v satisfies (v·w)/(|v||w|) = cos(θ) for every ϕ, and |v| = |w|
    from oracle.pyoracle import rotate2, rotate3
    rng = np.random.default_rng(3)
    for _ in range(50):
        w = rng.normal(size=3)
        if rng.random() < 0.3:
            w[0] = 0.0
        th, ph = rng.uniform(-np.pi, np.pi), rng.uniform(0, np.pi)
        v = rotate3(w, th, ph)
        assert abs(np.dot(v, w) / (np.linalg.norm(v) * np.linalg.norm(w)) - np.cos(th)) < 1e-12
        assert abs(np.linalg.norm(v) - np.linalg.norm(w)) < 1e-12
    assert np.allclose(rotate2(np.array([1.0, 0.0]), np.pi / 2), [0.0, 1.0], atol=1e-15)
    assert np.allclose(rotate3(np.array([0.0, 0.0, 2.0]), 0.0, 1.3), [0.0, 0.0, 2.0], atol=1e-15)


# ---- rows without a vector in the reference's test-suite: known answers worked out by hand from the source ----------
def test_pairforce_known_answers():
    """examples/celllistmap.jl:143-155 (calc_forces!) + :176-191 (agent_step!): two- and three-particle systems, one
    pair across the periodic edge, one pair beyond r_i + r_j (tests/known_answers.py has the arithmetic)."""
    import known_answers as ka

    for k, pos1, vel1 in ((2, ka.PF2_POS1, ka.PF2_VEL1), (3, ka.PF3_POS1, ka.PF3_VEL1)):
        o = ContinuousOracle(ka.PF_EXTENT, spacing=ka.PF_SPACING)
        o.add_agents(ka.PF_POS[:k], ka.PF_VEL[:k])
        o.pairforce_step(ka.PF_R[:k], ka.PF_K[:k], ka.PF_MASS[:k], ka.PF_DT, ka.PF_CUTOFF, nsteps=1)
        pos, vel = o.state()[:2]
        assert np.max(np.abs(pos - pos1)) <= ka.PF_TOL and np.max(np.abs(vel - vel1)) <= ka.PF_TOL


def test_spatial_index_known_answers():
    """get_spatial_index, src/spaces/continuous.jl:566-573: positions on property-cell edges, a property of lower
    dimension than the space, and pos == extent (index size + 1: the reference's BoundsError)."""
    import known_answers as ka
    from oracle.pyoracle import spatial_index

    assert np.array_equal(spatial_index(ka.SI_POS, ka.SI_EXTENT, (5, 3)), ka.SI_IDX_5x3)
    assert np.array_equal(spatial_index(ka.SI_POS, ka.SI_EXTENT, (4,)), ka.SI_IDX_4)
    assert spatial_index(ka.SI_POS_AT_EXTENT, ka.SI_EXTENT, (5, 3))[0, 0] == 6      # one past the property's 5 rows
