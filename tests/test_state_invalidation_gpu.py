"""Regression tests for host-side state tracking of the device handle:

* the per-cell aggregates of the approximate flock sweep (Σ(pos - centre), Σ vel per cell, csrc/ab_resort.cu)
  must be rebuilt when a caller rewrites `vel` alone (ab_upload of field 1, ab_elastic_collisions);
* entry points that are not documented as asynchronous stay synchronous after ab_set_async(1): results handed
  back through host pointers are complete on return, timing events are resolved, no spurious AB_ERR_CUDA.
Checked against the CPU oracle (examples/flock.jl:79-108, examples/celllistmap.jl:143-191,
docs/src/tutorial.jl:351-376, src/spaces/continuous.jl:212-239).
"""
import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import FLOCK, csr_sets, flock_init, periodic_rel_err, rel_err, uniform
from oracle.pyoracle import ContinuousOracle, GridOracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kernel", ["auto", "generic"])
def test_vel_upload_between_steps_refreshes_cell_sums(kernel):
    extent, spacing, n = (300.0, 300.0), 5.0, 40000
    pos, vel = flock_init(n, extent, 3)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=ab.FlockRule(kernel=kernel, **FLOCK))
    m.add_agents(pos, vel)
    m.step(2)                                   # leaves cell sums valid for the state after step 2
    p2 = m.pos
    _, v_new = flock_init(n, extent, 99)        # velocities replaced by the caller; positions untouched
    m.set_column("vel", v_new)
    o = ContinuousOracle(extent, spacing=spacing)
    o.add_agents(p2, v_new)
    cnt = o.flock_step_sync(FLOCK, return_counts=True)
    m.step(1)
    assert np.array_equal(m.last_neighbor_counts(), cnt)
    assert periodic_rel_err(m.pos, o.state()[0], extent) <= 1e-13
    assert rel_err(m.vel, o.state()[1]) <= 1e-13
    m.close()


def test_elastic_collisions_between_steps_refresh_cell_sums():
    extent, spacing, n = (200.0, 200.0), 5.0, 20000
    pos, vel = flock_init(n, extent, 5)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos, vel)
    m.step(2)
    pairs = m.interacting_pairs(0.5, "nearest")
    assert len(pairs) > 10
    m.elastic_collisions(pairs)                 # rewrites vel in place
    p, v = m.pos, m.vel
    o = ContinuousOracle(extent, spacing=spacing)
    o.add_agents(p, v)
    o.flock_step_sync(FLOCK)
    m.step(1)
    assert periodic_rel_err(m.pos, o.state()[0], extent) <= 1e-13
    assert rel_err(m.vel, o.state()[1]) <= 1e-13
    m.close()


def test_async_mode_keeps_synchronous_entry_points_synchronous():
    extent, spacing, n = (200.0, 200.0), 5.0, 8000
    pos, vel = flock_init(n, extent, 8)
    o = ContinuousOracle(extent, spacing=spacing)
    o.add_agents(pos, vel)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing))
    m.add_agents(pos, vel)
    m.set_async(True)
    # queries hand results back through host pointers: complete on return
    off, ids = m.nearby_ids_batch(ids=np.arange(1, 201), r=5.0, search="exact")
    for s, i in zip(csr_sets(off, ids), range(1, 201)):
        assert np.array_equal(s, np.sort(o.nearby_ids_agent(i, 5.0, "exact")))
    pr = m.interacting_pairs(2.0)
    assert set(map(tuple, pr)) == set(map(tuple, o.interacting_pairs_all(2.0)))
    # bulk move with caller-owned directions
    d = (uniform(n, 2, 77) - 0.5).astype(np.float64)
    m.walk(d)
    for i in range(1, 50):
        o.walk(i, d[i - 1])
    m.set_async(False)                          # (downloads are asynchronous in asynchronous mode)
    assert np.array_equal(m.pos[:49], o.state()[0][:49])
    m.close()


def test_async_mode_pairforce_and_schelling_return_valid_timing():
    extent = (400.0, 400.0)
    n = 4000
    pos = uniform(n, 2, 1, 0) * 400.0
    vel = (uniform(n, 2, 1, 1) - 0.5) * 50.0
    w = uniform(n, 3, 1, 3)
    rad, k, mass = (0.5 + 0.9 * w[:, 0]) * 10.0, 10.0 + 20.0 * w[:, 1], 10.0 + 100.0 * w[:, 2]
    o = ContinuousOracle(extent, spacing=20.0)
    o.add_agents(pos, vel)
    o.pairforce_step(rad, k, mass, 0.001, 20.0, nsteps=3)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=20.0), model_step=ab.PairForceRule(dt=0.001, cutoff=20.0))
    m.add_agents(pos, vel, r=rad, k=k, mass=mass)
    m.set_async(True)
    m.step(3)                                   # used to return a spurious AB_ERR_CUDA (event not ready)
    assert m.last_step_ms() > 0
    m.set_async(False)
    assert periodic_rel_err(m.pos, o.state()[0], extent) <= 1e-11
    m.close()

    dims = (128, 64)
    ncell = dims[0] * dims[1]
    ng = int(ncell * 0.8)
    perm = np.argsort(uniform(ncell, 1, 4)[:, 0], kind="stable")[:ng]
    gpos = np.stack([perm % dims[0] + 1, perm // dims[0] + 1], axis=1)
    group = (np.arange(ng) % 2 + 1).astype(np.int32)
    rule = ab.SchellingRule(min_to_be_happy=3, relocate=True, seed=5)
    g = ab.StandardABM(ab.GridSpaceSingle(dims, periodic=False), model_step=rule)
    g.add_agents(gpos.astype(np.int32), group=group, mood=np.zeros(ng, dtype=np.uint8))
    go = GridOracle(dims, periodic=False, metric="chebyshev", single=True)
    go.add_agents(gpos)
    _, mood = go.schelling_happiness(group, 3)
    g.set_async(True)
    g.step(1)
    assert rule.last_moved == min(int((mood == 0).sum()), ncell - ng)      # *h_moved was read after its copy landed
    assert g.last_step_ms() > 0
    g.set_async(False)
    g.close()


def test_index_canonical_flag():
    """ab_index_canonical: 1 while every cell of every index build was ordered by id; 0 once a cell held more than 4096
    agents at a build (it keeps arrival order: correct, but not bitwise reproducible — ADVICE round 1)."""
    extent = (100.0, 100.0)
    pos, vel = flock_init(20000, extent, 3)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos, vel)
    m.step(3)
    assert m.lib.ab_index_canonical(m.h) == 1
    m.close()
    crowd = pos.copy()
    crowd[:6000] = 52.0 + 0.4 * crowd[:6000] / 100.0            # 6000 agents inside one reference cell
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(crowd, vel)
    assert m.lib.ab_index_canonical(m.h) == 0
    assert len(m.nearby_ids(1, 1.0)) >= 5999                    # still a correct index
    m.close()


@pytest.mark.parametrize("search", ["approximate", "exact"])
def test_step_after_a_failed_step_starts_from_clean_counters(search):
    """A step that fails on the device (here: a NaN velocity makes a position NaN — the reference's InexactError in
    pos2cell, continuous.jl:112-113) leaves the re-sort's per-tile histograms of that step behind.  The next step on the
    repaired state must not see them: it equals, bit for bit, the same step on a fresh model."""
    extent, spacing, n = (300.0, 300.0), 5.0, 40000
    pos, vel = flock_init(n, extent, 17)
    rule = dict(search=search, **FLOCK)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=ab.FlockRule(**rule))
    m.add_agents(pos, vel)
    m.step(2)
    P, V = m.pos.copy(), m.vel.copy()
    bad = V.copy()
    bad[[7, 20003]] = np.nan
    m.set_column("vel", bad)
    with pytest.raises(ab.AgentsError) as ei:
        m.step(1)
    assert ei.value.code == ab._lib.ERR_NUMERIC
    m.set_column("pos", P)
    m.set_column("vel", V)
    m.step(3)
    f = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=ab.FlockRule(**rule))
    f.add_agents(P, V)
    f.step(3)
    assert np.array_equal(m.pos, f.pos) and np.array_equal(m.vel, f.vel)
    assert np.array_equal(m.last_neighbor_counts(), f.last_neighbor_counts())
    m.close()
    f.close()
