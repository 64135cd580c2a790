"""Generates the committed fixtures of tests/golden/ (run from the repo root: python tests/golden/make_golden.py).

The reference is Julia and cannot run in the build container (no `julia` binary), so these are NOT outputs of
Agents.jl: they are outputs of the CPU oracle (oracle/abm_oracle.cpp, itself pinned to the reference's own golden
vectors by tests/test_oracle_golden.py), frozen so that (a) a change of the oracle is noticed and (b) the GPU box,
where /root/reference does not exist either, compares the device against committed numbers as well as against
the live oracle.  Everything is derived from the counter-based generator of SURVEY §8d (tests/helpers.py).

  flock300.npz    BASELINE configs[0] (examples/flock.jl rule, 300 boids, extent (100,100), spacing 5/1.5,
                  seed 42, synchronous update): cells of the initial state, neighbour ids of 12 agents
                  (approximate + exact, r = 5), per-agent neighbour counts of step 1, pos/vel after 1 and 10 steps
  pairs300.npz    interacting_pairs(:all, exact / approximate, r = 3) and (:nearest, r = 3, ByID) of that state
  grid_sch.npz    Schelling happiness pass on a 40 x 30 GridSpaceSingle (periodic and not), 70 % occupancy
  randomwalk.npz  randomwalk! on the counter stream of include/agents_b200.h: first draws of the stream, positions of a
                  60-agent GridSpaceSingle / GridSpace (11 x 8) after 3 calls (plain, ifempty, force_motion), and the
                  directions of a ContinuousSpace uniform walk (2-D, 3-D)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import FLOCK, flock_init, uniform  # noqa: E402
from oracle.pyoracle import ContinuousOracle, GridOracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def flock300():
    extent, spacing, n = (100.0, 100.0), 5.0 / 1.5, 300
    pos, vel = flock_init(n, extent, 42)
    o = ContinuousOracle(extent, spacing=spacing)
    o.add_agents(pos, vel)
    out = {"cells": o.pos2cell(pos)}
    probe = np.arange(1, 300, 25)
    out["probe_ids"] = probe
    for search in ("approximate", "exact"):
        lists = [np.sort(o.nearby_ids_agent(int(i), 5.0, search=search)) for i in probe]
        out[f"nearby_{search}_offsets"] = np.cumsum([0] + [len(x) for x in lists])
        out[f"nearby_{search}_ids"] = np.concatenate(lists)
    pa = np.array(sorted(map(tuple, o.interacting_pairs_all(3.0, search="exact"))))
    pp = np.array(sorted(map(tuple, o.interacting_pairs_all(3.0, search="approximate"))))
    pn = np.array(o.interacting_pairs_nearest(3.0))
    np.savez_compressed(os.path.join(HERE, "pairs300.npz"), all_exact=pa, all_approximate=pp, nearest=pn)
    out["counts_step1"] = o.flock_step_sync(FLOCK, return_counts=True)
    out["pos_1"], out["vel_1"] = o.state()
    o.flock_step_sync(FLOCK, nsteps=9)
    out["pos_10"], out["vel_10"] = o.state()
    np.savez_compressed(os.path.join(HERE, "flock300.npz"), **out)


def grid_sch():
    dims = (40, 30)
    ncell = dims[0] * dims[1]
    n = int(0.7 * ncell)
    perm = np.argsort(uniform(ncell, 1, 7)[:, 0], kind="stable")[:n]
    pos = np.stack([perm % dims[0] + 1, perm // dims[0] + 1], axis=1).astype(np.int64)
    group = (np.arange(n) % 2 + 1).astype(np.int32)
    out = {"pos": pos, "group": group}
    for periodic in (True, False):
        g = GridOracle(dims, periodic=periodic, metric="chebyshev", single=True)
        g.add_agents(pos)
        cnt, mood = g.schelling_happiness(group, 3)
        out[f"count_periodic{int(periodic)}"] = np.asarray(cnt)
        out[f"mood_periodic{int(periodic)}"] = np.asarray(mood)
    np.savez_compressed(os.path.join(HERE, "grid_sch.npz"), **out)


def randomwalk():
    from oracle.pyoracle import continuous_randomwalk_directions, grid_randomwalk, rw_u01

    out = {"u01": np.array([[rw_u01(s, i, c) for c in range(4)] for s in (0, 42, 2 ** 63 + 5) for i in (1, 2, 1000000)])}
    dims = (11, 8)
    u = uniform(180, 2, 4)
    p = np.floor(u * np.asarray(dims)[None, :]).astype(np.int64) + 1
    _, idx = np.unique(p, axis=0, return_index=True)
    pos = p[np.sort(idx)][:60]
    out["grid_pos"] = pos
    for single in (True, False):
        for periodic in (True, False):
            g = GridOracle(dims, periodic=periodic, metric="chebyshev", single=single)
            g.add_agents(pos)
            ids = list(range(1, len(pos) + 1))
            grid_randomwalk(g, ids, 1, 11, ifempty=False, force_motion=False)
            grid_randomwalk(g, ids[::2], 2, 12, ifempty=True, force_motion=False)
            grid_randomwalk(g, ids, 1, 13, ifempty=True, force_motion=True)
            out[f"grid_single{int(single)}_periodic{int(periodic)}"] = g.positions()
    for D in (2, 3):
        vel = uniform(200, D, 4) * 2.0 - 1.0
        out[f"cont_vel{D}"] = vel
        out[f"cont_dir{D}_r"] = continuous_randomwalk_directions(vel, list(range(1, 201)), 0.7, 21)
        out[f"cont_dir{D}_speed"] = continuous_randomwalk_directions(vel, list(range(1, 201)), None, 22)
    np.savez_compressed(os.path.join(HERE, "randomwalk.npz"), **out)


if __name__ == "__main__":
    flock300()
    grid_sch()
    randomwalk()
    print("written:", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))
