"""Row a18 of SURVEY §8: step!(model, f) (src/simulations/step.jl:7-27, :38) and scheduled subsets
(Schedulers.Partially / ByProperty / id lists, src/submodules/schedulers.jl:68-151) for the pair-force and
Schelling rules — the flock rule's subset test lives in tests/test_parity_gpu.py::test_flock_scheduled_subset.
Parity: the scheduled agents equal the oracle's full step from the same state (the synchronous update reads only
the state at the start of the tick), the unscheduled ones are untouched bit for bit."""
import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import FLOCK, flock_init, periodic_rel_err, rel_err
from oracle.pyoracle import ContinuousOracle, GridOracle
from test_pairforce_gpu import TOL64, make as pf_make
from test_grid_gpu import schelling_setup

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("extent,spacing", [((400.0, 400.0), 20.0), ((200.0, 200.0), 10.0)], ids=["fine", "generic"])
def test_pairforce_scheduled_subset(extent, spacing):
    o, m, (rad, k, mass) = pf_make(extent, spacing, True, 4000)
    n = 4000
    rng = np.random.default_rng(5)
    ids = m.schedule_partially(0.4, rng)
    assert 0 < len(ids) < n
    p0, v0 = m.pos, m.vel
    o.pairforce_step(rad, k, mass, 0.001, 20.0)
    m.step(1, scheduler=ids)
    po, vo = o.state()
    p1, v1 = m.pos, m.vel
    act = np.zeros(n, dtype=bool)
    act[ids - 1] = True
    assert periodic_rel_err(p1[act], po[act], extent) <= TOL64 and rel_err(v1[act], vo[act]) <= TOL64
    assert np.array_equal(p1[~act], p0[~act]) and np.array_equal(v1[~act], v0[~act])
    # a following full step still agrees with the oracle (index and per-agent columns stayed consistent)
    oo = ContinuousOracle(extent, spacing=spacing)
    oo.add_agents(p1, v1)
    oo.pairforce_step(rad, k, mass, 0.001, 20.0)
    m.step(1)
    assert periodic_rel_err(m.pos, oo.state()[0], extent) <= TOL64
    assert np.array_equal(m.column("mass"), mass)
    m.close()


@pytest.mark.parametrize("dims", [(96, 80), (128, 64)])          # (128, 64): whole-warp rows (k_sch_happy4 / k_sch_compact4)
@pytest.mark.parametrize("relocate", [False, True])
def test_schelling_scheduled_subset(dims, relocate):
    m, o, pos0, group = schelling_setup(dims, True, "chebyshev", occupancy=0.8)
    n = len(pos0)
    ncell = dims[0] * dims[1]
    # tick 1: everybody computes a mood (no relocation) so that "previous mood" is defined
    m.model_step = ab.SchellingRule(min_to_be_happy=3, relocate=False)
    m.step(1)
    mood_prev = m.column("mood").copy()
    # tick 2 with a stricter rule and a 50 % schedule
    rule = ab.SchellingRule(min_to_be_happy=5, relocate=relocate, seed=11)
    m.model_step = rule
    ids = m.schedule_partially(0.5, np.random.default_rng(3))
    act = np.zeros(n, dtype=bool)
    act[ids - 1] = True
    _, mood_exp = o.schelling_happiness(group, 5)
    m.step(1, scheduler=ids)
    new, mood = m.pos.astype(np.int64), m.column("mood")
    moved = np.flatnonzero(np.any(new != pos0, axis=1))
    # unscheduled agents: same position, same mood as before
    assert np.array_equal(new[~act], pos0[~act]) and np.array_equal(mood[~act], mood_prev[~act])
    # scheduled agents: the oracle's mood; movers are scheduled unhappy agents, min(|U_sched|, |E|) of them
    assert np.array_equal(mood[act], mood_exp[act])
    unhappy_sched = np.flatnonzero(act & (mood_exp == 0))
    if relocate:
        assert set(moved.tolist()) <= set(unhappy_sched.tolist())
        assert len(moved) == min(len(unhappy_sched), ncell - n) == rule.last_moved
        occupied_before = set(map(tuple, pos0))
        assert all(tuple(p) not in occupied_before for p in new[moved])
        assert len(set(map(tuple, new))) == n
    else:
        assert len(moved) == 0
    assert np.array_equal(m.column("group"), group)
    m.close()


def test_step_until_predicate():
    """step!(model, f): steps until f(model, t) is true; t counts the ticks of this call (step.jl:38)."""
    extent = (100.0, 100.0)
    pos, vel = flock_init(300, extent, 42)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos, vel)
    seen = []
    m.step(lambda model, t: (seen.append(t) or t >= 7))
    assert m.time == 7 and seen == list(range(8))
    ref = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    ref.add_agents(pos, vel)
    ref.step(7)
    assert np.array_equal(m.pos, ref.pos) and np.array_equal(m.vel, ref.vel)
    # a predicate on the model state, checked every 4 ticks; it is already true at t = 0 on the second call
    m.step(lambda model, t: model.time >= 15, check_every=4)
    assert m.time == 15
    m.step(lambda model, t: True)
    assert m.time == 15
    # with a per-tick scheduler
    m.step(lambda model, t: t >= 3, scheduler=lambda model: model.schedule_partially(0.5, np.random.default_rng(model.time)))
    assert m.time == 18
    m.close(); ref.close()


def test_schedule_by_property_is_sortperm():
    """Schedulers.ByProperty(:field) = ids in sortperm order of the column (schedulers.jl:130-151)."""
    extent = (100.0, 100.0)
    pos, vel = flock_init(500, extent, 7)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    w = np.round(np.random.default_rng(1).random(500) * 20.0)          # ties: the order must be stable
    m.add_agents(pos, vel, weight=w)
    m.step(3)                                                            # the device order is not the id order any more
    ids = m.schedule_by_property("weight")
    assert np.array_equal(ids, np.argsort(w, kind="stable") + 1)
    # every agent scheduled: equal to the unscheduled step under the synchronous update
    ref = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    ref.add_agents(pos, vel)
    ref.step(4)
    m.step(1, scheduler=ids)
    assert np.array_equal(m.pos, ref.pos) and np.array_equal(m.vel, ref.vel)
    m.close(); ref.close()


def test_ids_by_property_device_sort():
    """ab_ids_by_property: stable ascending sortperm on the device for float / integer columns, NaN last, descending
    = the reversed result; grid handles too (columns in id order there)."""
    rng = np.random.default_rng(11)
    n = 3001                                                             # not a power of two: the network is padded
    extent = (100.0, 100.0)
    pos, vel = flock_init(n, extent, 9)
    w = np.round(rng.standard_normal(n) * 3.0) + 0.0                     # ties and negative values (+ 0.0: no -0.0, see below)
    w[::97] = np.nan
    lvl = rng.integers(-5, 6, n).astype(np.int32)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos, vel, weight=w, level=lvl)
    m.step(2)
    assert np.array_equal(m.schedule_by_property("weight"), np.argsort(w, kind="stable") + 1)      # numpy sorts NaN last too
    assert np.array_equal(m.schedule_by_property("level"), np.argsort(lvl, kind="stable") + 1)
    assert np.array_equal(m.schedule_by_property("pos", comp=1), np.argsort(m.pos[:, 1], kind="stable") + 1)
    out = np.zeros(n, dtype=np.int64)
    m._ck(m.lib.ab_ids_by_property(m.h, m.field_id("level"), 0, 1, m._ptr(out), n))
    assert np.array_equal(out, (np.argsort(lvl, kind="stable") + 1)[::-1])
    m.close()
    # isless(-0.0, 0.0) is true in Julia (sortperm orders by isless): -0.0 sorts before 0.0 — numpy would call them equal
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos[:4], vel[:4], weight=np.array([0.0, -0.0, 1.0, -1.0]))
    assert m.schedule_by_property("weight").tolist() == [4, 2, 1, 3]
    m.close()
    g = ab.StandardABM(ab.GridSpace((20, 20)))
    gp = rng.integers(1, 21, (500, 2)).astype(np.int32)
    g.add_agents(gp, energy=rng.random(500))
    e = g.column("energy")
    assert np.array_equal(g.schedule_by_property("energy"), np.argsort(e, kind="stable") + 1)
    g.close()
