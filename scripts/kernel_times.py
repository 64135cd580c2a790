"""GPU diagnostic: per-kernel CUDA-event times of the flock step at chosen points of the F2D-1M trajectory."""
import ctypes as C, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import agents_jl_b200 as ab
from helpers import FLOCK, flock_init

search = sys.argv[1] if len(sys.argv) > 1 else "approximate"
points = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "10,100,300").split(",")]
n, extent, spacing = 1_000_000, (2000.0, 2000.0), 5.0
pos, vel = flock_init(n, extent, 42)
m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=ab.FlockRule(search=search, **FLOCK))
m.add_agents(pos, vel)
lib = m.lib
done = 0
for pt in points:
    if pt > done:
        m.step(pt - done); done = pt
    m.step(10); done += 10
    warm = m.last_step_ms() / 10
    lib.ab_set_profiling(m.h, 1)
    m.step(10); done += 10
    names = (C.c_char_p * 32)(); kms = (C.c_double * 32)(); kl = (C.c_int64 * 32)()
    nk = lib.ab_kernel_times(m.h, names, kms, kl, 32)
    lib.ab_set_profiling(m.h, 0)
    print(json.dumps({"step": pt, "ms_per_step_back_to_back": round(warm, 4),
                      "kernels_us": {names[i].decode(): round(kms[i] / kl[i] * 1e3, 1) for i in range(nk)}}), flush=True)
