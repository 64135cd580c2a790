"""Checkpoint of a device-resident model (AgentsIO.save_checkpoint / load_checkpoint,
src/submodules/io/jld2_integration.jl:85-135, 196-223; the reference's test: test/jld2_tests.jl — the loaded model
equals the saved one field by field and keeps stepping).  Here: save at tick t, continue to t + k; load, step k;
both states must be equal BIT FOR BIT (the checkpoint carries the in-cell order the floating-point sums follow)."""
import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import FLOCK, flock_init, uniform
from test_grid_gpu import schelling_setup

pytestmark = pytest.mark.gpu


def _same_model(a, b):
    assert a.time == b.time and a.nagents() == b.nagents() and a.properties == b.properties
    assert set(a._columns) == set(b._columns)
    for name in a._columns:
        x, y = a.column(name), b.column(name)
        assert x.dtype == y.dtype and np.array_equal(x, y), name


@pytest.mark.parametrize("D,dtype,search,per_agent", [(2, np.float64, "approximate", False), (2, np.float64, "exact", True),
                                                      (3, np.float32, "approximate", False), (3, np.float64, "exact", False)])
def test_flock_resumes_bit_for_bit(tmp_path, D, dtype, search, per_agent):
    n = 30000 if D == 2 else 20000
    extent = (400.0, 300.0) if D == 2 else (100.0, 100.0, 60.0)
    pos, vel = flock_init(n, extent, 42, dtype)
    pa = None
    if per_agent:
        pa = np.tile(np.array([FLOCK[k] for k in ("speed", "cohere_factor", "separation", "separate_factor",
                                                  "match_factor", "visual_distance")]), (n, 1))
        pa[:, 0] *= 0.5 + uniform(n, 1, 9)[:, 0]
    rule = ab.FlockRule(search=search, per_agent=pa, **FLOCK)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0, dtype=dtype), model_step=rule,
                       properties={"note": "flock", "k": 3})
    m.add_agents(pos, vel, tag=np.arange(n, dtype=np.int32))
    m.step(7)
    path = tmp_path / "flock.npz"
    m.save_checkpoint(path)
    saved_pos = m.pos.copy()
    r = ab.load_checkpoint(path)
    _same_model(m, r)
    assert isinstance(r.model_step, ab.FlockRule) and r.model_step.search == search and r.model_step.p == rule.p
    assert np.array_equal(r.pos, saved_pos) and r.time == 7
    m.step(6)
    r.step(6)
    _same_model(m, r)
    assert not np.array_equal(m.pos, saved_pos)
    # a model rebuilt from the columns alone (cells in id order) is the same state, but its sums run in
    # another order: this is what the saved layout is for
    q = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0, dtype=dtype), model_step=ab.FlockRule(search=search, per_agent=pa, **FLOCK))
    with np.load(path) as z:
        q.add_agents(z["col_pos"], z["col_vel"])
    q.step(6)
    tol = 1e-9 if dtype == np.float64 else 1e-3
    d = np.abs(q.pos.astype(np.float64) - m.pos.astype(np.float64))
    d = np.minimum(d, np.asarray(extent) - d)
    assert np.quantile(d, 0.99) < tol
    for x in (m, r, q):
        x.close()


def test_pairforce_resumes_bit_for_bit(tmp_path):
    n, L = 20000, 700.0
    pos = uniform(n, 2, 1) * L
    vel = (uniform(n, 2, 2) - 0.5) * 100.0
    cols = dict(r=(0.5 + 0.9 * uniform(n, 1, 3)[:, 0]) * 10.0, k=10 + 20 * uniform(n, 1, 4)[:, 0],
                mass=10 + 100 * uniform(n, 1, 5)[:, 0])
    m = ab.StandardABM(ab.ContinuousSpace((L, L), spacing=20.0), model_step=ab.PairForceRule(dt=0.001, cutoff=20.0))
    m.add_agents(pos, vel, **cols)
    m.step(5)
    m.save_checkpoint(tmp_path / "pf.npz")
    r = ab.load_checkpoint(tmp_path / "pf.npz")
    _same_model(m, r)
    m.step(5)
    r.step(5)
    _same_model(m, r)
    m.close()
    r.close()


def test_schelling_resumes_bit_for_bit(tmp_path):
    m, _, _, _ = schelling_setup((96, 80), False, "chebyshev")
    m.model_step = ab.SchellingRule(min_to_be_happy=3, relocate=True, seed=7)
    m.step(3)
    m.save_checkpoint(tmp_path / "sch.npz")
    r = ab.load_checkpoint(tmp_path / "sch.npz")
    _same_model(m, r)
    assert isinstance(r.space, ab.GridSpaceSingle) and r.model_step.seed == 7
    m.step(4)
    r.step(4)
    _same_model(m, r)                      # same relocations: the step counter behind the permutations travelled
    assert r.model_step.last_moved == m.model_step.last_moved
    m.close()
    r.close()


def test_layout_errors(tmp_path):
    m = ab.StandardABM(ab.ContinuousSpace((10.0, 10.0), spacing=1.0))
    m.add_agents(uniform(50, 2, 1) * 10.0)
    bad = np.arange(1, 51, dtype=np.int64)
    bad[3] = 1
    with pytest.raises(ab.AgentsError, match="not a permutation"):
        m._ck(m.lib.ab_layout_set(m.h, m._ptr(bad), 50, 1, 0))
    np.savez(tmp_path / "x.npz", meta=np.frombuffer(b'{"format": "other"}', dtype=np.uint8))
    with pytest.raises(ValueError):
        ab.load_checkpoint(tmp_path / "x.npz")
    m.close()
