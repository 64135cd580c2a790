"""Committed fixtures of tests/golden/ (made by tests/golden/make_golden.py from the CPU oracle — the reference is
Julia and cannot run here; the oracle itself is pinned to the reference's golden vectors in test_oracle_golden.py).
CPU: the oracle still reproduces them bit for bit.  GPU: the device matches them — integer results exactly,
positions / velocities within the one-step (1e-13) and ten-step (1e-11) tolerances."""
import os

import numpy as np
import pytest

from helpers import FLOCK, flock_init, periodic_rel_err, rel_err

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EXTENT, SPACING, N = (100.0, 100.0), 5.0 / 1.5, 300


def _csr(f, search):
    off, ids = f[f"nearby_{search}_offsets"], f[f"nearby_{search}_ids"]
    return [ids[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def test_oracle_reproduces_fixtures():
    from oracle.pyoracle import ContinuousOracle, GridOracle

    f = np.load(os.path.join(G, "flock300.npz"))
    pos, vel = flock_init(N, EXTENT, 42)
    o = ContinuousOracle(EXTENT, spacing=SPACING)
    o.add_agents(pos, vel)
    assert np.array_equal(o.pos2cell(pos), f["cells"])
    for search in ("approximate", "exact"):
        for i, want in zip(f["probe_ids"], _csr(f, search)):
            assert np.array_equal(np.sort(o.nearby_ids_agent(int(i), 5.0, search=search)), want)
    p = np.load(os.path.join(G, "pairs300.npz"))
    assert np.array_equal(np.array(sorted(map(tuple, o.interacting_pairs_all(3.0, search="exact")))), p["all_exact"])
    assert np.array_equal(np.array(o.interacting_pairs_nearest(3.0)), p["nearest"])
    assert np.array_equal(o.flock_step_sync(FLOCK, return_counts=True), f["counts_step1"])
    assert np.array_equal(o.state()[0], f["pos_1"]) and np.array_equal(o.state()[1], f["vel_1"])
    o.flock_step_sync(FLOCK, nsteps=9)
    assert np.array_equal(o.state()[0], f["pos_10"]) and np.array_equal(o.state()[1], f["vel_10"])
    s = np.load(os.path.join(G, "grid_sch.npz"))
    for periodic in (True, False):
        g = GridOracle((40, 30), periodic=periodic, metric="chebyshev", single=True)
        g.add_agents(s["pos"])
        cnt, mood = g.schelling_happiness(s["group"], 3)
        assert np.array_equal(cnt, s[f"count_periodic{int(periodic)}"]) and np.array_equal(mood, s[f"mood_periodic{int(periodic)}"])


@pytest.mark.gpu
def test_device_matches_fixtures():
    import agents_jl_b200 as ab

    f = np.load(os.path.join(G, "flock300.npz"))
    pos, vel = flock_init(N, EXTENT, 42)
    m = ab.StandardABM(ab.ContinuousSpace(EXTENT, spacing=SPACING), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos, vel)
    assert np.array_equal(m.cells(), f["cells"])
    for search in ("approximate", "exact"):
        for i, want in zip(f["probe_ids"], _csr(f, search)):
            assert np.array_equal(np.sort(m.nearby_ids(int(i), 5.0, search=search)), want)
    p = np.load(os.path.join(G, "pairs300.npz"))
    for search in ("exact", "approximate"):
        got = m.interacting_pairs(3.0, "all", search=search)
        assert np.array_equal(np.array(sorted(map(tuple, got))), p[f"all_{search}"])
    assert np.array_equal(m.interacting_pairs(3.0, "nearest"), p["nearest"])
    m.step(1)
    assert np.array_equal(m.last_neighbor_counts(), f["counts_step1"])
    assert periodic_rel_err(m.pos, f["pos_1"], EXTENT) <= 1e-13 and rel_err(m.vel, f["vel_1"]) <= 1e-13
    m.step(9)
    assert periodic_rel_err(m.pos, f["pos_10"], EXTENT) <= 1e-11 and rel_err(m.vel, f["vel_10"]) <= 1e-11
    m.close()
    s = np.load(os.path.join(G, "grid_sch.npz"))
    for periodic in (True, False):
        g = ab.StandardABM(ab.GridSpaceSingle((40, 30), periodic=periodic, metric="chebyshev"),
                           model_step=ab.SchellingRule(min_to_be_happy=3, relocate=False))
        g.add_agents(s["pos"].astype(np.int32), group=s["group"].astype(np.int32), mood=np.zeros(len(s["pos"]), dtype=np.uint8))
        g.step(1)
        assert np.array_equal(g.column("mood").ravel(), s[f"mood_periodic{int(periodic)}"].ravel())
        g.close()


def _rw_grid_cases():
    return [(single, periodic) for single in (True, False) for periodic in (True, False)]


def test_oracle_reproduces_randomwalk_fixture():
    from oracle.pyoracle import GridOracle, continuous_randomwalk_directions, grid_randomwalk, rw_u01

    f = np.load(os.path.join(G, "randomwalk.npz"))
    u = np.array([[rw_u01(s, i, c) for c in range(4)] for s in (0, 42, 2 ** 63 + 5) for i in (1, 2, 1000000)])
    assert np.array_equal(u, f["u01"]) and np.all((u >= 0) & (u < 1))
    pos = f["grid_pos"]
    ids = list(range(1, len(pos) + 1))
    for single, periodic in _rw_grid_cases():
        g = GridOracle((11, 8), periodic=periodic, metric="chebyshev", single=single)
        g.add_agents(pos)
        grid_randomwalk(g, ids, 1, 11, ifempty=False, force_motion=False)
        grid_randomwalk(g, ids[::2], 2, 12, ifempty=True, force_motion=False)
        grid_randomwalk(g, ids, 1, 13, ifempty=True, force_motion=True)
        assert np.array_equal(g.positions(), f[f"grid_single{int(single)}_periodic{int(periodic)}"])
    for D in (2, 3):
        vel = f[f"cont_vel{D}"]
        assert np.array_equal(continuous_randomwalk_directions(vel, list(range(1, 201)), 0.7, 21), f[f"cont_dir{D}_r"])
        assert np.array_equal(continuous_randomwalk_directions(vel, list(range(1, 201)), None, 22), f[f"cont_dir{D}_speed"])


@pytest.mark.gpu
def test_device_matches_randomwalk_fixture():
    import agents_jl_b200 as ab

    f = np.load(os.path.join(G, "randomwalk.npz"))
    pos = f["grid_pos"]
    ids = np.arange(1, len(pos) + 1)
    for single, periodic in _rw_grid_cases():
        sp = (ab.GridSpaceSingle if single else ab.GridSpace)((11, 8), periodic=periodic, metric="chebyshev")
        m = ab.StandardABM(sp)
        m.add_agents(pos.astype(np.int32))
        m.randomwalk(1, seed=11, ifempty=False, force_motion=False)
        m.randomwalk(2, seed=12, ids=ids[::2], ifempty=True, force_motion=False)
        m.randomwalk(1, seed=13, ifempty=True, force_motion=True)
        assert np.array_equal(m.pos, f[f"grid_single{int(single)}_periodic{int(periodic)}"])
        m.close()
    for D in (2, 3):
        vel = f[f"cont_vel{D}"]
        for r, seed, key in ((0.7, 21, "r"), (None, 22, "speed")):
            m = ab.StandardABM(ab.ContinuousSpace((10.0,) * D, spacing=1.0))
            m.add_agents(np.full((200, D), 5.0), vel)
            m.randomwalk(r, seed=seed)
            assert np.max(np.abs(m.vel - f[f"cont_dir{D}_{key}"])) <= 1e-13
            m.close()
