"""Per-kernel CUDA-event times of the sharded F3D step (12.5 M boids per GPU) on rank 0, next to the unsharded step.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/prof_f3d_mgpu.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch, torch.distributed as dist
import agents_jl_b200 as ab
from agents_jl_b200 import _lib as L
from agents_jl_b200.parallel import Slabs, default_halo_capacity
from bench import FLOCK, flock_chunks, kernel_table

world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank = dist.get_rank(); lib = L.lib()
per = 12_500_000; n = per * world; extent = (1400.0, 1400.0, 175.0 * world)
m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(kernel="fine", **FLOCK), device=local)
s = Slabs(m, dist)
s.scatter_chunks(flock_chunks(n, extent, 42, 4_000_000), n, halo_capacity=default_halo_capacity(n, s.n_layers, factor=1.25))
m.step(3); torch.cuda.synchronize(); dist.barrier()
lib.ab_comm_align(m.h); m.step(10); t_plain = m.last_step_ms() / 10
lib.ab_set_profiling(m.h, 1); lib.ab_comm_align(m.h); m.step(5); k = kernel_table(lib, m); lib.ab_set_profiling(m.h, 0)
if rank == 0:
    print(json.dumps({"sharded_ms_per_step": t_plain, "kernels_us": {a: round(b["ms_per_launch"] * 1e3, 1) for a, b in k.items()}}), flush=True)
m.close()
if rank == 0:
    parts = list(flock_chunks(per, (1400.0, 1400.0, 175.0), 42, 4_000_000))
    u = ab.StandardABM(ab.ContinuousSpace((1400.0, 1400.0, 175.0), spacing=5.0), model_step=ab.FlockRule(kernel="fine", **FLOCK), device=local)
    u.add_agents(np.concatenate([p for _, p, _ in parts]), np.concatenate([v for _, _, v in parts]))
    u.step(3); u.step(10); t1 = u.last_step_ms() / 10
    lib.ab_set_profiling(u.h, 1); u.step(5); k1 = kernel_table(lib, u)
    print(json.dumps({"unsharded_ms_per_step": t1, "kernels_us": {a: round(b["ms_per_launch"] * 1e3, 1) for a, b in k1.items()}}), flush=True)
    u.close()
dist.barrier(); dist.destroy_process_group()
