"""Workload for ncu captures of the pair-force step: PF2D at `scale` x 10M particles (cutoff = spacing = 20), `presteps` + 2 steps."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import agents_jl_b200 as ab
from bench import particles

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.25
presteps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
side = int(round(791 * scale ** 0.5))
extent = (20.0 * side, 20.0 * side)
n = int(0.04 * extent[0] * extent[1])
pos, vel, rad, k, mass = particles(n, extent)
m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=20.0), model_step=ab.PairForceRule(dt=0.001, cutoff=20.0))
m.add_agents(pos, vel, r=rad, k=k, mass=mass)
m.step(presteps)
m.step(2)
print("done", n, m.last_step_ms() / 2)
