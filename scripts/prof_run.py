"""Workload for ncu captures: F2D-1M flock, `presteps` steps then 3 more (profile those with --launch-skip)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import agents_jl_b200 as ab
from helpers import FLOCK, flock_init

search = sys.argv[1] if len(sys.argv) > 1 else "approximate"
presteps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
n, extent, spacing = 1_000_000, (2000.0, 2000.0), 5.0
pos, vel = flock_init(n, extent, 42)
m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=ab.FlockRule(search=search, **FLOCK))
m.add_agents(pos, vel)
m.step(presteps)
m.step(3)
print("done", m.last_step_ms() / 3)
