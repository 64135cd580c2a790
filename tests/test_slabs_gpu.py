"""Slab decomposition on ONE GPU: several slab handles of one process (ab_comm_link_local +
ab_step_flock_group) must reproduce the unsharded run BITWISE — same kernels, same window / halo /
migration code as the NCCL path (tests/test_multigpu.py), the boundary messages move by device copies."""
import numpy as np
import pytest

import agents_jl_b200 as ab
from agents_jl_b200.parallel import LocalSlabs
from helpers import FLOCK, flock_init

pytestmark = pytest.mark.gpu

CASES = [
    # extent, nslabs, n, search, dtype, steps[, periodic]
    ((400.0, 400.0), 2, 60000, "approximate", np.float64, 12),
    ((400.0, 400.0), 2, 60000, "approximate", np.float64, 12, (True, False)),     # non-periodic last axis: end slabs have one neighbour
    ((400.0, 400.0), 3, 60000, "exact", np.float64, 9, False),                    # walls everywhere
    ((60.0, 60.0, 120.0), 2, 40000, "approximate", np.float64, 6, (True, True, False)),
    ((400.0, 400.0), 4, 60000, "exact", np.float64, 9),
    ((300.0, 405.0), 3, 50000, "approximate", np.float64, 9),      # 81 layers over 3 slabs
    ((300.0, 415.0), 4, 50000, "approximate", np.float64, 8),      # 83 layers: uneven slabs (21,21,21,20)
    ((60.0, 60.0, 120.0), 2, 40000, "approximate", np.float64, 6),
    ((60.0, 60.0, 120.0), 3, 40000, "exact", np.float64, 5),
    ((400.0, 400.0), 2, 60000, "approximate", np.float32, 6),
]


@pytest.mark.parametrize("case", CASES)
def test_local_slabs_bitwise_equal_to_unsharded(case):
    extent, nslabs, n, search, dtype, steps = case[:6]
    periodic = case[6] if len(case) > 6 else True
    pos, vel = flock_init(n, extent, 42, dtype=dtype)
    rule = ab.FlockRule(search=search, kernel="fine", **FLOCK)
    mk = lambda: ab.ContinuousSpace(extent, spacing=5.0, dtype=dtype, periodic=periodic)
    ls = LocalSlabs(mk, rule, nslabs)
    owned0 = ls.scatter(pos, vel)
    assert sum(owned0) == n
    hist = [owned0]
    for k in (steps // 2, steps - steps // 2):          # two calls: the state carries over between them
        ls.step(k)
        hist.append(ls.n_owned())
        assert sum(hist[-1]) == n                        # migration conserves the population
    gp, gv = ls.gather()                                 # raises if an agent was lost or duplicated
    ref = ab.StandardABM(mk(), model_step=ab.FlockRule(search=search, kernel="fine", **FLOCK))
    ref.add_agents(pos, vel)
    ref.step(steps // 2)
    ref.step(steps - steps // 2)
    assert np.array_equal(gp, ref.pos) and np.array_equal(gv, ref.vel)
    assert any(h != hist[0] for h in hist[1:]), "the test state should make agents cross slab boundaries"
    ls.close(); ref.close()


def test_sharded_handle_refuses_id_addressed_calls_and_bad_layouts():
    extent = (400.0, 400.0)
    pos, vel = flock_init(20000, extent, 1)
    rule = ab.FlockRule(kernel="fine", **FLOCK)
    ls = LocalSlabs(lambda: ab.ContinuousSpace(extent, spacing=5.0), rule, 2)
    ls.scatter(pos, vel)
    m = ls.models[0]
    with pytest.raises(ab.AgentsError):
        m.nearby_ids(7, 5.0)                             # id-addressed query on a slab window
    with pytest.raises(ab.AgentsError):
        m.step(1)                                        # linked slabs step together
    ls.close()
    with pytest.raises(ab.AgentsError):                  # 20 layers over 8 slabs: fewer than 4 per slab
        LocalSlabs(lambda: ab.ContinuousSpace((100.0, 100.0), spacing=5.0), rule, 8)


def test_halo_capacity_overflow_is_an_error():
    extent = (400.0, 400.0)
    pos, vel = flock_init(60000, extent, 2)
    ls = LocalSlabs(lambda: ab.ContinuousSpace(extent, spacing=5.0), ab.FlockRule(kernel="fine", **FLOCK), 2)
    ls.scatter(pos, vel, halo_capacity=64)               # far fewer than the ~2800 agents of 3 layers
    with pytest.raises(ab.AgentsError) as ei:
        ls.step(1)
    assert ei.value.code == -8                           # AB_ERR_COMM
    ls.close()
