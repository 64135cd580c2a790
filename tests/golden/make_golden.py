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


if __name__ == "__main__":
    flock300()
    grid_sch()
    print("written:", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))
