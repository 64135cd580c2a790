"""Small end-to-end workload: every kernel family of the library once (meant for compute-sanitizer memcheck / racecheck;
the tool is closed on this pool, so it only serves as a quick everything-once run there)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import agents_jl_b200 as ab
from agents_jl_b200.parallel import LocalSlabs
from helpers import FLOCK, flock_init, uniform
from bench import particles

# flock: fine path (approximate + exact), generic path, scheduled subset, 3-D
for extent, n, search, kernel in (((200.0, 200.0), 12000, "approximate", "auto"), ((200.0, 200.0), 12000, "exact", "auto"),
                                  ((100.0, 60.0), 3000, "approximate", "generic"), ((60.0, 60.0, 60.0), 8000, "approximate", "auto")):
    pos, vel = flock_init(n, extent, 42)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(search=search, kernel=kernel, **FLOCK))
    m.add_agents(pos, vel, weight=uniform(n, 1, 3)[:, 0])
    m.step(4)
    m.step(2, scheduler=np.arange(1, n + 1, 3))
    m.nearby_ids(7, 5.0, search="exact"); m.interacting_pairs(2.0); m.schedule_by_property("weight"); m.collect([(("vel", 0), "mean")])
    m.close()
# slabs on one GPU (fused delivery of the boundary messages, halo install with aggregates), walls
for periodic in (True, (True, False)):
    pos, vel = flock_init(30000, (300.0, 300.0), 1)
    ls = LocalSlabs(lambda: ab.ContinuousSpace((300.0, 300.0), spacing=5.0, periodic=periodic), ab.FlockRule(kernel="fine", **FLOCK), 3)
    ls.scatter(pos, vel); ls.step(4); ls.gather(); ls.close()
# pair force: fine and generic, scheduled subset
for extent, spacing in (((400.0, 400.0), 20.0), ((200.0, 200.0), 10.0)):
    n = 3000
    pos, vel, rad, k, mass = particles(n, extent)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=ab.PairForceRule(dt=0.001, cutoff=20.0))
    m.add_agents(pos, vel, r=rad, k=k, mass=mass)
    m.step(3); m.step(1, scheduler=np.arange(1, n + 1, 2)); m.close()
# Schelling: fast and slow rows, relocation, scheduled subset; grid queries and walks
for dims in ((128, 40), (50, 30)):
    ncell = dims[0] * dims[1]; n = int(0.8 * ncell)
    perm = np.argsort(uniform(ncell, 1, 42)[:, 0], kind="stable")[:n]
    gp = np.stack([perm % dims[0] + 1, perm // dims[0] + 1], axis=1).astype(np.int32)
    m = ab.StandardABM(ab.GridSpaceSingle(dims, periodic=True), model_step=ab.SchellingRule(min_to_be_happy=4, relocate=True, seed=1))
    m.add_agents(gp, group=(np.arange(n) % 2 + 1).astype(np.int32), mood=np.zeros(n, dtype=np.uint8))
    m.step(3); m.step(1, scheduler=np.arange(1, n + 1, 2)); m.randomwalk(1, seed=3, force_motion=2); m.nearby_ids(5, 2); m.close()
print("sanitize workload done")
