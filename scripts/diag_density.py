"""GPU diagnostic: how the F2D-1M flock state and the step time evolve with the step count."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import agents_jl_b200 as ab
from helpers import FLOCK, flock_init

search = sys.argv[1] if len(sys.argv) > 1 else "approximate"
kernel = sys.argv[2] if len(sys.argv) > 2 else "auto"
total = int(sys.argv[3]) if len(sys.argv) > 3 else 200
n, extent, spacing = 1_000_000, (2000.0, 2000.0), 5.0
pos, vel = flock_init(n, extent, 42)
m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=ab.FlockRule(search=search, kernel=kernel, **FLOCK))
m.add_agents(pos, vel)
done = 0
for chunk in [1, 4, 5, 10, 10, 20, 25, 25, 50, 50, 100, 100, 200, 400][:]:
    if done >= total: break
    m.step(chunk); done += chunk
    ms = m.last_step_ms() / chunk
    c = m.cells().astype(np.int64)
    lin = (c[:, 1] - 1) * 400 + (c[:, 0] - 1)
    cnt = np.bincount(lin, minlength=160000)
    p = m.pos
    sub = (np.floor(p[:, 1] / 2.5).astype(np.int64).clip(0, 799)) * 800 + np.floor(p[:, 0] / 2.5).astype(np.int64).clip(0, 799)
    sc = np.bincount(sub, minlength=640000)
    print(json.dumps(dict(step=done, ms_per_step=round(ms, 4), cell_max=int(cnt.max()), cell_p99=int(np.percentile(cnt, 99)),
                          empty_frac=float((cnt == 0).mean()), sum_n2_per_agent=float((cnt.astype(float) ** 2).sum() / n),
                          sub_max=int(sc.max()), sub_n2_per_agent=float((sc.astype(float) ** 2).sum() / n))), flush=True)
