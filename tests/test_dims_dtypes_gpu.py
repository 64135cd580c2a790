"""1-D ContinuousSpace and Float16 agents (missing #5; test/continuous_space_tests.jl:606-637).

A 1-D space is embedded by the host mirror in a 2-D one (one non-periodic row of cells); Float16 columns are held as
Float32 on the device and rounded when read back.  The 1-D checks are brute-force numpy restatements of
continuous.jl:112-121, :212-239 (no oracle for D = 1); the Float16 checks replay the reference's own testset."""
import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import uniform

pytestmark = pytest.mark.gpu


def _near_1d(pos, q, r, L, spacing, periodic, search):
    """nearby_ids(q, model, r) on a 1-D ContinuousSpace, by brute force over all agents."""
    n = int(round(L / spacing))
    cell = lambda x: max(1, int(np.ceil(x / spacing)))
    c = cell(q)
    centre = spacing * (c - 0.5)
    d = abs(q - centre)
    delta = min(d, L - d) if periodic else d
    gr = int(np.ceil((r + delta) / spacing))
    assert gr < n // 2                                  # the test stays in the unclipped regime of grid_general.jl:47-89
    out = []
    for i, x in enumerate(pos):
        ci = cell(x)
        dc = ci - c
        if periodic:
            dc = (dc + n // 2) % n - n // 2 if n % 2 == 0 else ((dc + n // 2) % n) - n // 2
            ok = abs(dc) <= gr
        else:
            ok = abs(dc) <= gr
        if not ok:
            continue
        if search == "exact":
            dx = abs(x - q)
            if periodic:
                dx = min(dx, L - dx)
            if not dx <= r:
                continue
        out.append(i + 1)
    return np.asarray(out, dtype=np.int64)


@pytest.mark.parametrize("periodic", [True, False])
def test_one_dimensional_space(periodic):
    L, spacing, n = 100.0, 2.5, 1500
    pos = uniform(n, 1, 7, 0) * L
    vel = uniform(n, 1, 7, 1) * 2.0 - 1.0
    m = ab.StandardABM(ab.ContinuousSpace((L,), spacing=spacing, periodic=periodic))
    m.add_agents(pos, vel)
    assert m.pos.shape == (n, 1) and np.array_equal(m.pos, pos) and np.array_equal(m.vel, vel)
    cells = m.cells()
    assert cells.shape == (n, 1)
    assert np.array_equal(cells[:, 0], np.maximum(1, np.ceil(pos[:, 0] / spacing)).astype(np.int32))
    for search in ("approximate", "exact"):
        for q, r in ((12.3, 1.0), (50.0, 3.7), (97.9, 2.0), (1.1, 4.9)):
            got = np.sort(m.nearby_ids([q], r, search=search))
            want = _near_1d(pos[:, 0], q, r, L, spacing, periodic, search)
            assert np.array_equal(got, want), (search, q, r)
    # nearby_ids(agent): drops the agent itself only
    got = np.sort(m.nearby_ids(17, 2.0, search="exact"))
    want = _near_1d(pos[:, 0], pos[16, 0], 2.0, L, spacing, periodic, "exact")
    assert np.array_equal(got, want[want != 17])
    # move_agent!(agent, model, dt): Euler step + normalize_position (walk.jl:77-91)
    m.move_agent(1.5)
    x = pos[:, 0] + 1.5 * vel[:, 0]
    x = np.mod(x, L) if periodic else np.clip(x, 0.0, np.nextafter(L, -np.inf))
    assert np.array_equal(m.pos[:, 0], x)
    # walk!(agent, direction, model) for a subset
    ids = np.arange(1, n + 1, 3)
    d = np.full((len(ids), 1), -0.75)
    m.walk(d, ids=ids)
    x2 = x.copy()
    x2[ids - 1] = np.mod(x[ids - 1] - 0.75, L) if periodic else np.clip(x[ids - 1] - 0.75, 0.0, np.nextafter(L, -np.inf))
    assert np.array_equal(m.pos[:, 0], x2)
    with pytest.raises(ab.AgentsError):
        m.randomwalk(1.0)
    m.close()


def test_one_dimensional_space_spatial_fields():
    """get_spatial_index / get_spatial_property / the update_vel! field (continuous.jl:86-91, 545-573) on a 1-D space:
    rows are (n, 1) like every other 1-D call — the device's padded second coordinate never shows."""
    L, spacing, n, cells = 100.0, 2.5, 1501, 8                     # odd n: a (n, 1) array cannot be read as (n / 2, 2)
    pos = uniform(n, 1, 21, 0) * L
    vel = uniform(n, 1, 21, 1) - 0.5
    m = ab.StandardABM(ab.ContinuousSpace((L,), spacing=spacing))
    m.add_agents(pos, vel)
    assert m.column("pos").shape == (n, 1) and np.array_equal(m.column("pos"), pos)
    want = np.floor(pos[:, 0] / (L / cells)).astype(np.int64) + 1   # floor(Int, pos / eps) + 1, continuous.jl:566-573
    assert np.array_equal(m.get_spatial_index((cells,))[:, 0], want)
    q = uniform(333, 1, 22, 0) * L
    wq = np.floor(q[:, 0] / (L / cells)).astype(np.int64) + 1
    assert np.array_equal(m.get_spatial_index((cells,), positions=q)[:, 0], wq)
    prop = uniform(cells, 1, 23)[:, 0]
    assert np.array_equal(m.get_spatial_property(prop), prop[want - 1])
    assert np.array_equal(m.get_spatial_property(prop, positions=q), prop[wq - 1])
    field = (uniform(cells, 1, 24) - 0.5).reshape(cells, 1)
    m.update_vel_field(field, 0.5)
    assert np.array_equal(m.vel, vel + 0.5 * field[want - 1])
    m.set_column("vel", vel)
    assert np.array_equal(m.vel, vel)
    with pytest.raises(ab.AgentsError):
        m.get_spatial_index((cells, 2))                             # a 2-D property over a 1-D space
    with pytest.raises(ab.AgentsError):
        m.update_vel_field(np.zeros((cells, 2)), 1.0)               # two velocity components
    with pytest.raises(ab.AgentsError):
        m.randomwalk_angles(np.zeros(n))
    m.close()


def test_float16_agents_reference_testset():
    """test/continuous_space_tests.jl:606-637 ("different continuous agent types") for T = Float16."""
    T = np.float16
    m = ab.StandardABM(ab.ContinuousSpace((1.0, 1.0), dtype=T))
    x = np.array([0.5, 0.5], dtype=T)
    vel = (np.random.default_rng(3).standard_normal(2) / 20).astype(T)
    m.add_agents(x[None, :], vel[None, :])
    assert m.pos.dtype == T and np.array_equal(m.pos[0], x) and np.array_equal(m.vel[0], vel)
    m.move_agent(1.0)
    assert np.array_equal(m.pos[0], x + vel * T(1.0))                 # Float16 arithmetic, like `x + vel .* dt`
    y2 = np.array([0.6, 0.3], dtype=T)
    m.set_column("pos", y2[None, :].astype(np.float32))
    assert np.array_equal(m.pos[0], y2)
    m.walk(m.vel, ids=[1])
    assert np.array_equal(m.pos[0], y2 + vel)
    # a population: every column comes back as Float16 and equals the Float32 state rounded once
    n = 3000
    p = (uniform(n, 2, 5, 0)).astype(T)
    v = (uniform(n, 2, 5, 1) * 0.1 - 0.05).astype(T)
    m2 = ab.StandardABM(ab.ContinuousSpace((1.0, 1.0), dtype=T))
    m2.add_agents(p, v)
    m2.move_agent(1.0)
    want = np.mod(p.astype(np.float32) + v.astype(np.float32), np.float32(1.0)).astype(T)
    assert m2.pos.dtype == T and np.array_equal(m2.pos, want)
    m.close(); m2.close()
