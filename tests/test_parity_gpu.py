"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on
the same seeded inputs, and against the reference's golden vectors.

Bars: cells, neighbour-id sets, pair sets — bit-exact after
sorting; positions / velocities — one-step relative error <= 1e-13 (f64) from
an identical state; N-step trajectories — stated tolerance below, plus the
number of birds whose neighbour count ever differed.
"""
import math

import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import FLOCK, csr_sets, flock_init, periodic_rel_err, rel_err, uniform
from oracle.pyoracle import ContinuousOracle

pytestmark = pytest.mark.gpu

ONE_STEP_TOL_F64 = 1e-13
ONE_STEP_TOL_F32 = 1e-5


def make_pair(extent, spacing, periodic, n, seed=42, dtype=np.float64):
    pos, vel = flock_init(n, extent, seed, dtype)
    o = ContinuousOracle(extent, spacing=spacing, periodic=periodic, dtype=dtype)
    o.add_agents(pos, vel)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing, periodic=periodic, dtype=dtype))
    m.add_agents(pos, vel)
    return o, m, pos, vel


# ---------------------------------------------------------------------------
# boundary behaviour
# ---------------------------------------------------------------------------
def test_create_errors():
    with pytest.raises(ab.AgentsError) as ei:
        ab.StandardABM(ab.ContinuousSpace((1.0, 1.0), spacing=0.3))
    assert "divisible" in str(ei.value)          # continuous.jl:99-101
    m = ab.StandardABM(ab.ContinuousSpace((100, 100), spacing=5.0 / 1.5))
    assert m.space.dims == (30, 30)
    with pytest.raises(ab.AgentsError):
        ab.StandardABM(ab.ContinuousSpace((-1.0, 1.0), spacing=0.1))


def test_empty_model():
    m = ab.StandardABM(ab.ContinuousSpace((10.0, 10.0), spacing=1.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(np.zeros((0, 2)))
    assert m.nagents() == 0
    m.step(3)
    off, ids = m.nearby_ids_batch(positions=[(1.0, 1.0)], r=2.0)
    assert list(off) == [0, 0] and len(ids) == 0
    assert m.interacting_pairs(1.0).shape == (0, 2)


def test_upload_download_roundtrip_after_sort():
    o, m, pos, vel = make_pair((50.0, 30.0), 2.5, True, 5000, seed=7)
    assert np.array_equal(m.pos, pos) and np.array_equal(m.vel, vel)
    extra = uniform(5000, 3, 9)
    m.add_column("extra", extra)
    m.lib.ab_build_index(m.h)
    assert np.array_equal(m.column("extra"), extra)
    m.add_column("flag", (np.arange(5000) % 3).astype(np.uint8))
    m.move_agent(0.5)
    m.lib.ab_build_index(m.h)
    assert np.array_equal(m.column("flag"), (np.arange(5000) % 3).astype(np.uint8))
    assert np.array_equal(m.column("extra"), extra)


def test_outside_extent_errors():
    m = ab.StandardABM(ab.ContinuousSpace((1.0, 1.0), spacing=0.1, periodic=False))
    with pytest.raises(ab.AgentsError) as ei:
        m.add_agents([(0.5, 0.5), (1.5, 1.5)])
    assert "outside space extent" in str(ei.value)      # continuous.jl:160
    with pytest.raises(ab.AgentsError) as ei:
        m.add_agents([(0.5, 0.5), (float("nan"), 0.2)])
    assert ei.value.code == ab._lib.ERR_NUMERIC


# ---------------------------------------------------------------------------
# cells (bit-exact)
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("extent,spacing", [((100.0, 100.0), 5.0 / 1.5), ((2000.0, 2000.0), 5.0),
                                            ((12.0, 10.0, 5.0), 0.2), ((1.0, 1.0), 0.1)])
def test_cells_bit_exact(extent, spacing, dtype):
    n = 20000
    pos, vel = flock_init(n, extent, 11, dtype)
    # edge cases: zero, exactly on cell edges, exactly the extent (legal, A2/A3)
    pos[0] = 0
    pos[1] = np.asarray(extent, dtype=dtype)
    k = np.arange(2, 200)
    pos[k, 0] = (np.asarray(spacing, dtype=dtype) * (k % 7)).astype(dtype)
    pos = np.minimum(pos, np.asarray(extent, dtype=dtype))
    o = ContinuousOracle(extent, spacing=spacing, dtype=dtype)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing, dtype=dtype))
    m.add_agents(pos, vel)
    assert np.array_equal(m.cells(), o.pos2cell(pos))


# ---------------------------------------------------------------------------
# neighbour queries (bit-exact id sets)
# ---------------------------------------------------------------------------
def test_golden_cell_centres():
    # test/continuous_space_tests.jl:184-211
    m = ab.StandardABM(ab.ContinuousSpace((1.0, 1.0), spacing=0.1))
    centres = [(0.05 + 0.1 * i, 0.05 + 0.1 * j) for j in range(10) for i in range(10)]
    m.add_agents(centres)
    rs = [x * 0.1 + 0.001 for x in (1, 2, 3.4)]
    for r, nn in zip(rs, (4, 12, 36)):
        assert len(m.nearby_ids((0.45, 0.45), r, "exact")) == nn + 1
        assert len(m.nearby_ids(25, r, "exact")) == nn
        assert len(m.nearby_ids((0.45, 0.45), r)) >= nn + 1
        assert len(m.nearby_ids(25, r)) >= nn


def test_golden_within_same_cell():
    # test/continuous_space_tests.jl:213-236
    m = ab.StandardABM(ab.ContinuousSpace((1.0, 1.0), spacing=0.1))
    m.add_agents([(0.51, 0.51), (0.51 + 0.08, 0.51)])
    r0, r1 = 0.01, 0.08
    assert list(m.nearby_ids(1, r0)) == [2]
    assert list(m.nearby_ids(1, r0, "exact")) == []
    assert list(m.nearby_ids(1, r1)) == [2]
    assert list(m.nearby_ids(1, r1, "exact")) == [2]
    assert sorted(m.nearby_ids((0.51, 0.51), r0)) == [1, 2]
    assert sorted(m.nearby_ids((0.51, 0.51), r0, "exact")) == [1]
    assert sorted(m.nearby_ids((0.51, 0.51), r1, "exact")) == [1, 2]
    with pytest.raises(ab.AgentsError):
        m.nearby_ids((0.5, 0.5), 0.1, "fuzzy")       # continuous.jl:218


CASES = [
    # extent, spacing, periodic, n, radii
    ((100.0, 100.0), 5.0 / 1.5, True, 3000, (5.0, 1.0, 9.7)),
    ((100.0, 100.0), 5.0, False, 3000, (5.0, 3.3)),
    ((40.0, 60.0), 4.0, (True, False), 2000, (5.0, 11.0)),
    ((1.0, 1.0), 0.1, True, 400, (0.101, 0.341, 0.49, 0.75, 3.0)),      # grid_r >= dims/2: clipped tables
    ((1.0, 1.0), 0.25, False, 300, (0.3, 2.0)),
    ((30.0, 30.0, 30.0), 5.0, True, 4000, (5.0, 7.5)),
    ((12.0, 10.0, 6.0), 2.0, (True, False, True), 1500, (2.0, 3.1, 9.0)),
]


@pytest.mark.parametrize("extent,spacing,periodic,n,radii", CASES)
@pytest.mark.parametrize("search", ["approximate", "exact"])
def test_nearby_ids_sets_bit_exact(extent, spacing, periodic, n, radii, search):
    o, m, pos, _ = make_pair(extent, spacing, periodic, n, seed=3)
    qids = np.arange(1, n + 1, 7)
    qpos = flock_init(97, extent, 99)[0]
    for r in radii:
        off, ids = m.nearby_ids_batch(ids=qids, r=r, search=search)
        got = csr_sets(off, ids)
        for k, i in enumerate(qids):
            assert np.array_equal(got[k], np.sort(o.nearby_ids_agent(int(i), r, search))), (r, i)
        off, ids = m.nearby_ids_batch(positions=qpos, r=r, search=search)
        got = csr_sets(off, ids)
        for k in range(len(qpos)):
            assert np.array_equal(got[k], np.sort(o.nearby_ids(qpos[k], r, search))), (r, k)


def test_nearby_ids_float32():
    o, m, pos, _ = make_pair((100.0, 100.0), 5.0 / 1.5, True, 3000, seed=5, dtype=np.float32)
    qids = np.arange(1, 3001, 5)
    for search in ("approximate", "exact"):
        off, ids = m.nearby_ids_batch(ids=qids, r=5.0, search=search)
        got = csr_sets(off, ids)
        for k, i in enumerate(qids):
            assert np.array_equal(got[k], np.sort(o.nearby_ids_agent(int(i), 5.0, search)))


def test_golden_interacting_pairs():
    # test/continuous_space_tests.jl:240-269, :303-320
    P = [(7.074386436066224, 4.963014649338054), (5.831962448496828, 4.926297135685473),
         (5.122087781793935, 5.300031210394806), (3.9715633336430156, 4.8106570045816675)]
    m = ab.StandardABM(ab.ContinuousSpace((10, 10), spacing=0.2, periodic=False))
    m.add_agents(P)
    pairs = {tuple(p) for p in m.interacting_pairs(2.0)}
    assert len(pairs) == 5 and (3, 6) not in pairs
    m = ab.StandardABM(ab.ContinuousSpace((10, 10), spacing=0.1, periodic=False))
    m.add_agents(P)
    pairs = {tuple(p) for p in m.interacting_pairs(2.5)}
    assert len(pairs) == 5 and (1, 4) not in pairs
    m = ab.StandardABM(ab.ContinuousSpace((1, 1), spacing=0.1, periodic=True))
    m.add_agents([(0.01, 0.01), (0.2, 0.2), (0.5, 0.5)])
    for r in (0.29, 0.3):
        assert [tuple(p) for p in m.interacting_pairs(r)] == [(1, 2)]


@pytest.mark.parametrize("extent,spacing,periodic,n,radii", CASES[:5] + CASES[6:])
@pytest.mark.parametrize("search", ["approximate", "exact"])
def test_interacting_pairs_bit_exact(extent, spacing, periodic, n, radii, search):
    n = min(n, 1200)
    o, m, _, _ = make_pair(extent, spacing, periodic, n, seed=8)
    for r in radii[:2]:
        ref = o.interacting_pairs_all(r, search)
        got = m.interacting_pairs(r, "all", search)
        assert len({tuple(p) for p in got}) == len(got)            # unique
        ref = ref[np.lexsort((ref[:, 1], ref[:, 0]))]
        got = got[np.lexsort((got[:, 1], got[:, 0]))]
        assert np.array_equal(ref, got)


def test_golden_nearest_neighbor():
    # test/continuous_space_tests.jl:323-349
    m = ab.StandardABM(ab.ContinuousSpace((1, 1), spacing=0.1, periodic=True))
    m.add_agents([(0.01, 0.01), (0.2, 0.01), (0.2, 0.2), (0.5, 0.5)])
    assert list(m.nearest_neighbor(math.sqrt(2))) == [2, 1, 2, 3]
    m = ab.StandardABM(ab.ContinuousSpace((1, 1), spacing=0.02, periodic=False))
    m.add_agents([(0.30386961108678245, 0.17690859963285543), (0.33020500850053125, 0.178391311081449)])
    assert list(m.nearest_neighbor(1e-100)) == [0, 0]
    o, m, _, _ = make_pair((100.0, 100.0), 5.0, True, 2000, seed=21)
    ref = [o.nearest_neighbor(i, 6.0) or 0 for i in range(1, 2001)]
    assert list(m.nearest_neighbor(6.0)) == ref


# ---------------------------------------------------------------------------
# movement
# ---------------------------------------------------------------------------
def test_golden_moves():
    # test/continuous_space_tests.jl:28-36, :52-58, :353-378
    m = ab.StandardABM(ab.ContinuousSpace((1, 1)))
    m.add_agents([(0.51, 0.51)], [(0.2, 0.1)])
    m.move_agent(1)
    assert tuple(m.pos[0]) == (0.71, 0.61)
    m = ab.StandardABM(ab.ContinuousSpace((1, 1), periodic=True))
    m.add_agents([(0.4936805014512577, 0.007750092779602313)], [(0.6319498548742291, -0.7750092779602327)])
    m.move_agent(0.01)
    assert tuple(m.pos[0]) == (0.5, 1.0)
    assert tuple(m.cells()[0]) == (10, 20)
    m = ab.StandardABM(ab.ContinuousSpace((12, 10), periodic=False))
    m.add_agents([(0.0, 0.0)])
    m.walk([(1.0, 1.0)])
    assert tuple(m.pos[0]) == (1.0, 1.0)
    m.walk([(15.0, 1.0)])
    assert tuple(m.pos[0]) == (np.nextafter(12.0, -np.inf), 2.0)
    m = ab.StandardABM(ab.ContinuousSpace((12, 10, 5), spacing=0.2))
    m.add_agents([(0.0, 0.0, 0.0)])
    m.walk([(1.0, 1.0, 1.0)])
    m.walk([(15.0, 1.2, 3.9)])
    assert tuple(m.pos[0]) == (4.0, 2.2, 4.9)
    m = ab.StandardABM(ab.ContinuousSpace((12, 10), periodic=True))
    m.add_agents([(11.0, 9.0)], [(3.0, 1.0)])
    m.move_agent(1.0)
    assert tuple(m.pos[0]) == (2.0, 0.0)


@pytest.mark.parametrize("periodic", [True, False, (True, False)])
def test_move_bit_exact_and_masked(periodic):
    o, m, pos, vel = make_pair((20.0, 30.0), 1.0, periodic, 3000, seed=13)
    active = np.arange(1, 3001, 3)
    m.move_agent(7.3, ids=active)             # dt large enough to wrap / clamp several times
    for i in active:
        o.move_agent_dt(int(i), 7.3)
    assert np.array_equal(m.pos, o.state()[0])
    m.move_agent(-2.1)
    for i in range(1, 3001):
        o.move_agent_dt(i, -2.1)
    assert np.array_equal(m.pos, o.state()[0])
    assert np.array_equal(m.cells(), o.pos2cell(o.state()[0]))


# ---------------------------------------------------------------------------
# flock rule
# ---------------------------------------------------------------------------
FLOCK_CASES = [
    ((100.0, 100.0), 5.0 / 1.5, True, 300),       # BASELINE configs[0]: examples/flock.jl, 300 boids
    ((100.0, 100.0), 5.0, True, 2000),
    ((60.0, 40.0), 5.0, False, 1500),
    ((60.0, 40.0), 2.5, (False, True), 1500),
    ((20.0, 20.0), 5.0, True, 200),               # 4x4 cells: every offset table is clipped
    ((40.0, 40.0, 40.0), 5.0, True, 3000),
]


@pytest.mark.parametrize("extent,spacing,periodic,n", FLOCK_CASES)
@pytest.mark.parametrize("search", ["approximate", "exact"])
@pytest.mark.parametrize("kernel", ["generic", "auto"])
def test_flock_one_step_parity(extent, spacing, periodic, n, search, kernel):
    o, m, _, _ = make_pair(extent, spacing, periodic, n, seed=42)
    m.model_step = ab.FlockRule(search=search, kernel=kernel, **FLOCK)
    # one-step parity from identical state, repeated along a trajectory
    for chunk in (1, 1, 5, 1):
        if chunk > 1:       # advance both, then re-synchronise the device to the oracle's exact state
            o.flock_step_sync(FLOCK, nsteps=chunk, search=search)
            p, v = o.state()
            m.add_agents(p, v)
            continue
        cnt = o.flock_step_sync(FLOCK, nsteps=1, search=search, return_counts=True)
        m.step(1)
        assert np.array_equal(m.last_neighbor_counts(), cnt)        # neighbour sets: exact
        p, v = o.state()
        assert periodic_rel_err(m.pos, p, extent) <= ONE_STEP_TOL_F64
        assert rel_err(m.vel, v) <= ONE_STEP_TOL_F64


def test_flock_config0_trajectory_100_steps():
    """BASELINE configs[0]: examples/flock.jl, 300 boids, extent (100,100), visual_distance 5, 100 steps."""
    extent = (100.0, 100.0)
    o, m, _, _ = make_pair(extent, 5.0 / 1.5, True, 300, seed=42)
    m.model_step = ab.FlockRule(**FLOCK)
    flips = 0
    worst = 0.0
    for _ in range(100):
        cnt = o.flock_step_sync(FLOCK, return_counts=True)
        m.step(1)
        flips += int(np.sum(m.last_neighbor_counts() != cnt))
        worst = max(worst, periodic_rel_err(m.pos, o.state()[0], extent))
    # chaotic amplification of 1-ulp summation-order differences over 100 steps
    assert flips == 0
    assert worst <= 1e-9
    assert np.allclose(np.linalg.norm(m.vel, axis=1), 1.0, atol=1e-14)


def test_flock_multi_step_call_equals_single_steps():
    _, m1, pos, vel = make_pair((100.0, 100.0), 5.0, True, 2000, seed=4)
    m2 = ab.StandardABM(ab.ContinuousSpace((100.0, 100.0), spacing=5.0))
    m2.add_agents(pos, vel)
    m1.model_step = ab.FlockRule(**FLOCK)
    m2.model_step = ab.FlockRule(**FLOCK)
    m1.step(7)
    for _ in range(7):
        m2.step(1)
    assert np.array_equal(m1.pos, m2.pos) and np.array_equal(m1.vel, m2.vel)   # bitwise reproducible


def test_flock_per_agent_params():
    n = 1500
    o, m, _, _ = make_pair((100.0, 100.0), 5.0, True, n, seed=6)
    u = uniform(n, 6, 77)
    pa = np.stack([0.5 + u[:, 0], 0.05 + 0.1 * u[:, 1], 1.0 + 2.0 * u[:, 2], 0.1 + 0.3 * u[:, 3],
                   0.02 + 0.04 * u[:, 4], 3.0 + 4.0 * u[:, 5]], axis=1)
    m.model_step = ab.FlockRule(per_agent=pa, **FLOCK)
    for _ in range(3):
        cnt = o.flock_step_sync(FLOCK, per_agent=pa.T.copy(), return_counts=True)
        m.step(1)
        assert np.array_equal(m.last_neighbor_counts(), cnt)
        p, v = o.state()
        assert periodic_rel_err(m.pos, p, (100.0, 100.0)) <= ONE_STEP_TOL_F64
        m.add_agents(p, v)
        m.add_column("_flock_params", pa)


def test_flock_float32_one_step():
    extent = (100.0, 100.0)
    o, m, _, _ = make_pair(extent, 5.0, True, 2000, seed=42, dtype=np.float32)
    m.model_step = ab.FlockRule(**FLOCK)
    cnt = o.flock_step_sync(FLOCK, return_counts=True)
    m.step(1)
    assert np.array_equal(m.last_neighbor_counts(), cnt)
    assert periodic_rel_err(m.pos, o.state()[0], extent) <= ONE_STEP_TOL_F32
    assert rel_err(m.vel, o.state()[1]) <= ONE_STEP_TOL_F32


def test_flock_zero_velocity_is_an_error_not_a_fix():
    # flock.jl:105 divides by norm(vel): a lone bird with zero velocity gives NaN;
    # the reference then throws in pos2cell (InexactError).  Same here.
    m = ab.StandardABM(ab.ContinuousSpace((100.0, 100.0), spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents([(50.0, 50.0)], [(0.0, 0.0)])
    with pytest.raises(ab.AgentsError):
        m.step(1)


# ---------------------------------------------------------------------------
# tiled flock path (aggregated approximate sweep / block exact sweep + tiled re-sort)
# ---------------------------------------------------------------------------
TILED_CASES = [
    ((400.0, 400.0), 5.0, True, 40000),
    ((400.0, 300.0), 5.0, False, 30000),
    ((300.0, 400.0), 5.0, (True, False), 30000),
    ((60.0, 60.0, 60.0), 5.0, True, 8000),
    ((60.0, 45.0, 50.0), 5.0, (True, False, True), 6000),
    ((400.0, 400.0), 8.0, True, 20000),             # visual_distance 5 < spacing: grid_r 1 and 2 both occur
]


@pytest.mark.parametrize("extent,spacing,periodic,n", TILED_CASES)
@pytest.mark.parametrize("search", ["approximate", "exact"])
def test_flock_tiled_parity(extent, spacing, periodic, n, search):
    o, m, _, _ = make_pair(extent, spacing, periodic, n, seed=11)
    m.model_step = ab.FlockRule(search=search, kernel="tiled", **FLOCK)
    for _ in range(3):
        cnt = o.flock_step_sync(FLOCK, nsteps=1, search=search, return_counts=True)
        m.step(1)
        assert np.array_equal(m.last_neighbor_counts(), cnt)
        p, v = o.state()
        assert periodic_rel_err(m.pos, p, extent) <= ONE_STEP_TOL_F64
        assert rel_err(m.vel, v) <= ONE_STEP_TOL_F64
        assert np.array_equal(m.cells(), o.pos2cell(p))
        m.add_agents(p, v)


@pytest.mark.parametrize("search", ["approximate", "exact"])
def test_flock_tiled_trajectory_and_index_consistency(search):
    """20 chained tiled steps (no re-synchronisation): the index the tiled re-sort leaves behind must
    answer neighbour queries exactly like a fresh build, and the trajectory must follow the oracle."""
    extent = (300.0, 300.0)
    o, m, _, _ = make_pair(extent, 5.0, True, 25000, seed=3)
    m.model_step = ab.FlockRule(search=search, kernel="tiled", **FLOCK)
    flips = 0
    for _ in range(20):
        cnt = o.flock_step_sync(FLOCK, search=search, return_counts=True)
        m.step(1)
        flips += int(np.sum(m.last_neighbor_counts() != cnt))
    p, v = o.state()
    assert flips == 0
    assert periodic_rel_err(m.pos, p, extent) <= 1e-10
    q = np.arange(1, 2001)
    off, ids = m.nearby_ids_batch(ids=q, r=5.0, search="exact")
    got = csr_sets(off, ids)
    m2 = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0))
    m2.add_agents(m.pos, m.vel)
    off2, ids2 = m2.nearby_ids_batch(ids=q, r=5.0, search="exact")
    want = csr_sets(off2, ids2)
    assert all(np.array_equal(a, b) for a, b in zip(got, want))


def test_flock_tiled_equals_generic_and_is_reproducible():
    extent = (400.0, 400.0)
    pos, vel = flock_init(40000, extent, 5)
    out = []
    for kernel in ("tiled", "tiled", "generic"):
        m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(kernel=kernel, **FLOCK))
        m.add_agents(pos, vel)
        m.step(5)
        out.append((m.pos, m.vel, m.last_neighbor_counts()))
        m.close()
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])     # bitwise reproducible
    assert np.array_equal(out[0][2], out[2][2])
    assert periodic_rel_err(out[0][0], out[2][0], extent) <= 1e-12


@pytest.mark.parametrize("search", ["approximate", "exact"])
def test_flock_tiled_crowded_and_sparse_states(search):
    """States the tiles do not fit (crowded cells, nearly empty rows) take the in-kernel fallbacks."""
    extent = (200.0, 200.0)
    rng = np.random.default_rng(5)
    n = 12000
    pos = rng.uniform(0, 200.0, size=(n, 2))
    pos[:6000] = 100.0 + rng.normal(0, 3.0, size=(6000, 2))          # ~6000 birds in a few cells
    pos[6000:6050, 1] = 37.5                                          # and a thin line
    pos = np.mod(pos, 200.0)
    vel = rng.uniform(-1, 1, size=(n, 2))
    o = ContinuousOracle(extent, spacing=5.0)
    o.add_agents(pos, vel)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(search=search, kernel="tiled", **FLOCK))
    m.add_agents(pos, vel)
    for _ in range(2):
        cnt = o.flock_step_sync(FLOCK, search=search, return_counts=True)
        m.step(1)
        assert np.array_equal(m.last_neighbor_counts(), cnt)
        assert periodic_rel_err(m.pos, o.state()[0], extent) <= 1e-12
    # sparse: 300 birds in 160x160 cells
    extent = (800.0, 800.0)
    o2, m2, _, _ = make_pair(extent, 5.0, True, 300, seed=8)
    m2.model_step = ab.FlockRule(search=search, kernel="tiled", **FLOCK)
    cnt = o2.flock_step_sync(FLOCK, search=search, return_counts=True)
    m2.step(1)
    assert np.array_equal(m2.last_neighbor_counts(), cnt)
    assert periodic_rel_err(m2.pos, o2.state()[0], extent) <= ONE_STEP_TOL_F64


def test_refinement_does_not_change_queries():
    extent = (100.0, 100.0)
    o, m, pos, _ = make_pair(extent, 5.0, True, 3000, seed=12)
    m.set_refinement(2)
    q = np.arange(1, 501)
    for search in ("approximate", "exact"):
        off, ids = m.nearby_ids_batch(ids=q, r=6.5, search=search)
        for i, got in zip(q, csr_sets(off, ids)):
            assert np.array_equal(got, np.sort(o.nearby_ids_agent(int(i), 6.5, search)))
    got = m.interacting_pairs(3.0)
    want = o.interacting_pairs_all(3.0)
    assert np.array_equal(np.unique(got, axis=0), np.unique(want, axis=0))
    assert np.array_equal(m.cells(), o.pos2cell(pos))


@pytest.mark.parametrize("search", ["approximate", "exact"])
@pytest.mark.parametrize("kernel", ["generic", "auto"])
def test_flock_scheduled_subset(search, kernel):
    """Schedulers.Partially / an id list from schedule(model) (schedulers.jl:8, 68-120): only the scheduled agents
    update, the others keep their state and stay neighbours.  Under the synchronous update the expected state is the
    oracle's full step for the scheduled agents and the old state for the rest."""
    extent, n = (100.0, 100.0), 3000
    o, m, pos, vel = make_pair(extent, 5.0, True, n, seed=42)
    m.model_step = ab.FlockRule(search=search, kernel=kernel, **FLOCK)
    rng = np.random.default_rng(4)
    for tick in range(3):
        ids = np.sort(rng.permutation(n)[: n // 3]) + 1
        p0, v0 = o.state()
        cnt = o.flock_step_sync(FLOCK, nsteps=1, search=search, return_counts=True)
        p1, v1 = o.state()
        mask = np.zeros(n, dtype=bool)
        mask[ids - 1] = True
        want_p, want_v = np.where(mask[:, None], p1, p0), np.where(mask[:, None], v1, v0)
        m.step(1, scheduler=ids)
        got_p, got_v = m.pos, m.vel
        assert np.array_equal(got_p[~mask], p0[~mask]) and np.array_equal(got_v[~mask], v0[~mask])     # untouched, bit for bit
        assert periodic_rel_err(got_p, want_p, extent) <= ONE_STEP_TOL_F64
        assert rel_err(got_v, want_v) <= ONE_STEP_TOL_F64
        assert np.array_equal(m.last_neighbor_counts()[mask], cnt[mask])
        # the oracle continues from the masked state
        o = ContinuousOracle(extent, spacing=5.0, periodic=True)
        o.add_agents(got_p, got_v)
    assert m.time == 3
    # a callable scheduler is evaluated every tick; the index stays consistent
    m.step(2, scheduler=lambda mod: np.arange(1, n + 1, 2))
    off, idl = m.nearby_ids_batch(ids=np.array([1, 2, 3]), r=5.0, search="exact")
    o2 = ContinuousOracle(extent, spacing=5.0, periodic=True)
    o2.add_agents(m.pos, m.vel)
    for k, i in enumerate((1, 2, 3)):
        assert np.array_equal(np.sort(idl[off[k]:off[k + 1]]), np.sort(o2.nearby_ids_agent(i, 5.0, "exact")))
    with pytest.raises(ab.AgentsError):
        m.step(1, scheduler=[0])
    m.close()
