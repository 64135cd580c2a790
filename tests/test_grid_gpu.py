"""GPU parity tests for GridSpace / GridSpaceSingle (csrc/ab_grid.cu) through the C ABI:
nearby_ids under all metrics / periodicities, walk!, and the Schelling rule — against the CPU
oracle and the reference's own golden vectors (test/grid_space_tests.jl)."""
import numpy as np
import pytest

import agents_jl_b200 as ab
from agents_jl_b200 import _lib as L
from helpers import csr_sets, uniform
from oracle.pyoracle import GridOracle

pytestmark = pytest.mark.gpu


def make(dims, periodic, metric, single, pos):
    sp = (ab.GridSpaceSingle if single else ab.GridSpace)(dims, periodic=periodic, metric=metric)
    m = ab.StandardABM(sp)
    m.add_agents(np.asarray(pos, dtype=np.int32))
    o = GridOracle(dims, periodic=periodic, metric=metric, single=single)
    o.add_agents(pos)
    return m, o


def random_positions(dims, n, seed, single):
    u = uniform(n * 3, len(dims), seed)
    p = (np.floor(u * np.asarray(dims)[None, :]).astype(np.int64) + 1)
    if single:
        _, idx = np.unique(p, axis=0, return_index=True)
        p = p[np.sort(idx)]
    return p[:n]


@pytest.mark.parametrize("single", [False, True])
@pytest.mark.parametrize("periodic", [False, True, (True, False)])
@pytest.mark.parametrize("metric", ["euclidean", "manhattan", "chebyshev"])
def test_golden_nearby_r1(single, periodic, metric):
    # grid_space_tests.jl:168-231
    m, o = make((5, 5), periodic, metric, single, [(1, 1), (2, 1), (3, 2), (2, 5)])
    near_agent = sorted(m.nearby_ids(2, 1).tolist())
    near_pos = sorted(m.nearby_ids((2, 1), 1).tolist())
    assert 2 in near_pos and 2 not in near_agent
    assert near_pos == sorted(near_agent + [2])
    if metric in ("euclidean", "manhattan"):
        exp = {False: [1], True: [1, 4], (True, False): [1]}[periodic]
    else:
        exp = {False: [1, 3], True: [1, 3, 4], (True, False): [1, 3]}[periodic]
    assert near_agent == exp
    m.close()


@pytest.mark.parametrize("single", [False, True])
@pytest.mark.parametrize("periodic", [False, True, (True, False)])
def test_golden_nearby_r34(single, periodic):
    # grid_space_tests.jl:233-245 (fill_space!: one agent in every cell)
    res = {}
    for metric in ("euclidean", "manhattan", "chebyshev"):
        m, _ = make((9, 9), periodic, metric, single, [(x, y) for y in range(1, 10) for x in range(1, 10)])
        res[metric] = len(m.nearby_ids((5, 5), 3.4))
        m.close()
    assert res["euclidean"] == res["manhattan"] + 4
    assert res["chebyshev"] == 7 ** 2


@pytest.mark.parametrize("periodic", [False, True, (True, False, True)])
def test_golden_radius_larger_than_extent_3d(periodic):
    # grid_space_tests.jl:247-256: 131 distinct neighbouring positions (+ the origin itself)
    pos = [(x, y, z) for z in range(1, 16) for y in range(1, 5) for x in range(1, 4)]
    m, o = make((3, 4, 15), periodic, "chebyshev", False, pos)
    ids = m.nearby_ids((2, 2, 7), 5)
    assert len(ids) == 132 and len(set(ids.tolist())) == 132
    assert sorted(ids.tolist()) == sorted(o.nearby_ids((2, 2, 7), 5).tolist())
    m.close()


@pytest.mark.parametrize("single", [False, True])
@pytest.mark.parametrize("dims,periodic", [((17, 13), True), ((17, 13), False), ((17, 13), (True, False)),
                                           ((4, 5), True), ((6, 7, 5), (True, False, True))])
@pytest.mark.parametrize("metric", ["euclidean", "manhattan", "chebyshev"])
def test_nearby_ids_sets_bit_exact(single, dims, periodic, metric):
    ncell = int(np.prod(dims))
    pos = random_positions(dims, ncell // 2 if single else ncell * 2, 7, single)
    m, o = make(dims, periodic, metric, single, pos)
    n = len(pos)
    qids = np.arange(1, n + 1, 7)
    qpos = random_positions(dims, 25, 11, False)
    for r in (0, 1, 1.5, 2, 3.4, 6):
        off, ids = m.nearby_ids_batch(ids=qids, r=r)
        for s, i in zip(csr_sets(off, ids), qids):
            assert np.array_equal(s, np.sort(o.nearby_ids_agent(int(i), r))), (r, i)
        off, ids = m.nearby_ids_batch(positions=qpos.astype(np.int32), r=r)
        for s, q in zip(csr_sets(off, ids), qpos):
            assert np.array_equal(s, np.sort(o.nearby_ids(q, r))), (r, q)
    m.close()


def test_single_refuses_second_agent_and_bounds():
    # grid_space_tests.jl:51-53, grid_single.jl:47-52
    m = ab.StandardABM(ab.GridSpaceSingle((5, 5)))
    with pytest.raises(ab.AgentsError) as ei:
        m.add_agents(np.array([(2, 2), (3, 3), (2, 2)], dtype=np.int32))
    assert ei.value.code == L.ERR_OCCUPIED
    with pytest.raises(ab.AgentsError) as ei:
        m.add_agents(np.array([(2, 2), (6, 3)], dtype=np.int32))
    assert ei.value.code == L.ERR_OUTSIDE
    m.close()


def test_golden_walk():
    # grid_space_tests.jl:333-405
    def run(dims, periodic, single, start, seq, ifempty=False, others=()):
        m, _ = make(dims, periodic, "chebyshev", single, [start, *others])
        for d, exp in seq:
            m.walk([d], ids=[1], ifempty=ifempty)
            assert tuple(m.pos[0]) == exp, (d, exp)
        m.close()

    run((3, 3), True, False, (1, 1), [((0, 1), (1, 2)), ((1, 1), (2, 3)), ((1, 0), (3, 3)), ((2, 0), (2, 3)),
                                      ((1, -1), (3, 2)), ((0, -1), (3, 1)), ((-1, -1), (2, 3)), ((-1, 0), (1, 3)),
                                      ((0, -8), (1, 1))])
    run((3, 3), True, False, (1, 1), [((1, 1), (1, 1)), ((1, 0), (2, 1))], ifempty=True, others=[(2, 2)])
    run((3, 3), False, False, (1, 1), [((0, 1), (1, 2)), ((1, 1), (2, 3)), ((1, 0), (3, 3)), ((1, 0), (3, 3)),
                                       ((-5, 0), (1, 3)), ((-1, -1), (1, 2))], ifempty=True)
    run((3, 3), (True, False), False, (1, 1), [((0, 1), (1, 2)), ((1, 1), (2, 3)), ((1, 0), (3, 3)), ((1, 0), (1, 3)),
                                               ((0, 1), (1, 3)), ((-1, -3), (3, 1))], ifempty=True)
    run((5, 5), True, True, (3, 3), [((1, 1), (4, 4))])
    run((3, 3, 2), True, False, (1, 1, 1), [((1, 1, 1), (2, 2, 2))])


@pytest.mark.parametrize("single,ifempty", [(False, False), (False, True), (True, True)])
@pytest.mark.parametrize("periodic", [True, (True, False)])
def test_walk_matches_sequential_reference_order(single, ifempty, periodic):
    dims = (12, 9)
    pos = random_positions(dims, 40, 3, single)
    m, o = make(dims, periodic, "chebyshev", single, pos)
    n = len(pos)
    rng = np.random.default_rng(5)
    for _ in range(3):
        ids = rng.permutation(n)[: n // 2] + 1
        dirs = rng.integers(-2, 3, size=(len(ids), 2))
        m.walk(dirs.astype(np.int32), ids=ids, ifempty=ifempty)
        for i, d in zip(ids, dirs):
            o.walk(int(i), d, ifempty=ifempty)
        assert np.array_equal(m.pos, o.positions())
        # the index follows the moves
        off, ids2 = m.nearby_ids_batch(ids=np.arange(1, n + 1), r=2)
        for s, i in zip(csr_sets(off, ids2), range(1, n + 1)):
            assert np.array_equal(s, np.sort(o.nearby_ids_agent(i, 2)))
    m.close()


def schelling_setup(dims, periodic, metric, occupancy=0.75, seed=42):
    ncell = int(np.prod(dims))
    n = int(ncell * occupancy)
    perm = np.argsort(uniform(ncell, 1, seed)[:, 0], kind="stable")[:n]
    pos = np.stack([perm % dims[0] + 1, perm // dims[0] + 1], axis=1)
    group = (np.arange(n) % 2 + 1).astype(np.int32)
    m = ab.StandardABM(ab.GridSpaceSingle(dims, periodic=periodic, metric=metric),
                       model_step=ab.SchellingRule(min_to_be_happy=3, relocate=False))
    m.add_agents(pos.astype(np.int32), group=group, mood=np.zeros(n, dtype=np.uint8))
    o = GridOracle(dims, periodic=periodic, metric=metric, single=True)
    o.add_agents(pos)
    return m, o, pos, group


@pytest.mark.parametrize("dims", [(64, 48), (33, 7), (300, 257), (256, 37), (128, 5), (1152, 20)])   # last three: whole-warp rows (k_sch_happy4)
@pytest.mark.parametrize("periodic", [False, True, (True, False), (False, True)])
@pytest.mark.parametrize("metric", ["chebyshev", "manhattan"])
def test_schelling_happiness_bit_exact(dims, periodic, metric):
    m, o, pos, group = schelling_setup(dims, periodic, metric)
    for mth in (1, 3, 5):
        m.model_step = ab.SchellingRule(min_to_be_happy=mth, relocate=False)
        m.step(1)
        _, mood = o.schelling_happiness(group, mth)
        assert np.array_equal(m.column("mood"), mood)
        assert np.array_equal(m.pos, pos)       # nobody moves in the happiness pass
    m.close()


@pytest.mark.parametrize("dims", [(96, 80), (128, 64)])      # (128, 64): whole-warp rows (k_sch_happy4 / k_sch_compact4)
@pytest.mark.parametrize("periodic", [False, True])
@pytest.mark.parametrize("occupancy", [0.75, 0.97])
def test_schelling_relocation_invariants(dims, periodic, occupancy):
    m, o, pos0, group = schelling_setup(dims, periodic, "chebyshev", occupancy=occupancy)
    n = len(pos0)
    ncell = dims[0] * dims[1]
    rule = ab.SchellingRule(min_to_be_happy=3, relocate=True, seed=7)
    m.model_step = rule
    pos = pos0.copy()
    for step in range(4):
        # expected unhappy set from the oracle on the current state
        o = GridOracle(dims, periodic=periodic, metric="chebyshev", single=True)
        o.add_agents(pos)
        _, mood_exp = o.schelling_happiness(group, 3)
        unhappy = np.flatnonzero(mood_exp == 0)
        occupied_before = set(map(tuple, pos))
        m.step(1)
        new = m.pos.astype(np.int64)
        mood = m.column("mood")
        moved = np.flatnonzero(np.any(new != pos, axis=1))
        # movers are unhappy agents; as many as there are empty cells, at most
        assert set(moved.tolist()) <= set(unhappy.tolist())
        assert len(moved) == min(len(unhappy), ncell - n) == rule.last_moved
        # every destination was empty at the start of the step; one agent per cell
        assert all(tuple(p) not in occupied_before for p in new[moved])
        assert len(set(map(tuple, new))) == n
        assert new.min() >= 1 and np.all(new.max(0) <= np.asarray(dims))
        # moods: happy agents true; unhappy ones (moved or not) false
        assert np.array_equal(mood, mood_exp)
        # the id grid follows: every agent finds itself through a position query
        off, ids = m.nearby_ids_batch(positions=new.astype(np.int32), r=0)
        assert np.array_equal(ids, np.arange(1, n + 1)) and np.array_equal(np.diff(off), np.ones(n, dtype=np.int64))
        assert np.array_equal(m.column("group"), group)
        pos = new
    # reproducible: same seed, same trajectory
    m2, _, _, _ = schelling_setup(dims, periodic, "chebyshev", occupancy=occupancy)
    m2.model_step = ab.SchellingRule(min_to_be_happy=3, relocate=True, seed=7)
    m2.step(4)
    assert np.array_equal(m2.pos, pos)
    # segregation progresses: fewer unhappy agents than at the start
    assert (m.column("mood") == 0).sum() <= (mood_exp == 0).sum()
    m.close(); m2.close()


@pytest.mark.parametrize("single", [False, True])
@pytest.mark.parametrize("dims,periodic", [((12, 9), True), ((12, 9), False), ((7, 5, 6), (True, False, True)), ((5, 5), True)])
@pytest.mark.parametrize("metric", ["chebyshev", "manhattan", "euclidean"])
def test_nearby_positions_and_empty_positions(single, dims, periodic, metric):
    """nearby_positions (grid_general.jl:185-232; golden vectors grid_space_tests.jl:168-257 pin the oracle) and
    empty_nearby_positions (discrete.jl:116-129) — position SETS exact."""
    ncell = int(np.prod(dims))
    n = ncell // 2
    pos = random_positions(dims, n, 5, single=True)            # distinct cells, so both space kinds accept them
    n = len(pos)
    Space = ab.GridSpaceSingle if single else ab.GridSpace
    m = ab.StandardABM(Space(dims, periodic=periodic, metric=metric))
    m.add_agents(pos.astype(np.int32))
    o = GridOracle(dims, periodic=periodic, metric=metric, single=single)
    o.add_agents(pos)
    occupied = {tuple(p) for p in pos}
    D = len(dims)
    qs = np.array([[1] * D, list(dims), [max(1, d // 2) for d in dims], [2] * D, [d - 1 for d in dims]], dtype=np.int32)
    for r in (1, 2, 3.4, 6):
        off, got = m.nearby_positions_batch(qs, r)
        offe, gote = m.nearby_positions_batch(qs, r, empty=True)
        for i, q in enumerate(qs):
            want = sorted(map(tuple, o.nearby_positions(q, r)))
            assert sorted(map(tuple, got[off[i]:off[i + 1]])) == want
            assert sorted(map(tuple, gote[offe[i]:offe[i + 1]])) == [p for p in want if p not in occupied]
    assert sorted(map(tuple, m.nearby_positions(qs[2], 1))) == sorted(map(tuple, o.nearby_positions(qs[2], 1)))
    with pytest.raises(ab.AgentsError):
        m.nearby_positions([0] * D, 1)
    m.close()
