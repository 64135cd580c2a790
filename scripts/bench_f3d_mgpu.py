"""F3D-100M (BASELINE configs[4]): 3-D flock, 12.5 M boids per GPU, slabs along z, NCCL halo + migration.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
      scripts/bench_f3d_mgpu.py --steps 10 --warmup 3

Weak scaling: extent (1400, 1400, 175 x world), spacing = visual_distance = 5 (280 x 280 x 35 cells per
GPU); world = 8 is the 100 M-boid box of SURVEY §8d.  The initial state is the counter-based generator of
SURVEY §8d, produced in id chunks so that no rank ever holds the global arrays.  Rank 0 prints one JSON line
(device time of one ab_step_flock(K) call, max over ranks).  world = 1 runs the unsharded library on the same
per-GPU problem (periodic 175-deep box) as the weak-scaling base.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def chunks(n, extent, seed, step):
    from helpers import splitmix64

    ext = np.asarray(extent, dtype=np.float64)[None, :]
    D = len(extent)
    for first in range(0, n, step):
        m = min(step, n - first)
        with np.errstate(over="ignore"):
            i = np.arange(first, first + m, dtype=np.uint64)[:, None] * np.uint64(8)
            cp = np.arange(D, dtype=np.uint64)[None, :]
            up = (splitmix64(np.uint64(seed) ^ (i + cp)) >> np.uint64(11)).astype(np.float64) * 2.0 ** -53
            uv = (splitmix64(np.uint64(seed) ^ (i + cp + np.uint64(4))) >> np.uint64(11)).astype(np.float64) * 2.0 ** -53
        yield first, up * ext, uv * 2.0 - 1.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--per-gpu", type=int, default=12_500_000)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--search", default="approximate")
    ap.add_argument("--halo-factor", type=float, default=1.25)
    ap.add_argument("--profile", action="store_true", help="per-kernel CUDA-event times on rank 0 (perturbs the total)")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    import agents_jl_b200 as ab
    from agents_jl_b200.parallel import Slabs
    from helpers import FLOCK

    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank = dist.get_rank() if world > 1 else 0
    n = args.per_gpu * world
    extent = (1400.0, 1400.0, 175.0 * world)
    rule = ab.FlockRule(search=args.search, kernel="fine", **FLOCK)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0, periodic=True), model_step=rule, device=local)
    if world > 1:
        s = Slabs(m, dist)
        # message capacity: the boundary messages are fixed-size (capacity, not content, travels), so it is
        # set close to the content — 1.25 x the mean population of the 3 boundary layers (a layer of
        # 280 x 280 cells holds 357 k boids +- 0.2 %)
        from agents_jl_b200.parallel import default_halo_capacity

        cap = default_halo_capacity(n, s.n_layers, factor=args.halo_factor)
        owned = s.scatter_chunks(chunks(n, extent, 42, 4_000_000), n, halo_capacity=cap)
    else:
        parts = list(chunks(n, extent, 42, 4_000_000))
        m.add_agents(np.concatenate([p for _, p, _ in parts]), np.concatenate([v for _, _, v in parts]))
        owned = n

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    m.step(args.warmup)
    barrier()
    if args.profile:
        m.lib.ab_set_profiling(m.h, 1)
    m.step(args.steps)
    ms = m.last_step_ms()
    barrier()
    kernels = None
    if args.profile:
        import ctypes as C

        names, kms, kl = (C.c_char_p * 32)(), (C.c_double * 32)(), (C.c_int64 * 32)()
        nk = m.lib.ab_kernel_times(m.h, names, kms, kl, 32)
        kernels = {names[i].decode(): round(kms[i] / max(kl[i], 1), 4) for i in range(min(nk, 32))}
        m.lib.ab_set_profiling(m.h, 0)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    if rank == 0:
        print(json.dumps({"workload": "F3D-100M" if world == 8 else f"F3D-{n / 1e6:g}M", "n_gpus": world, "agents": n,
                          "extent": list(extent), "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
                          "value": n * args.steps / (ms * 1e-3), "unit": "agent-updates/s", "search": args.search,
                          "owned_rank0": int(owned), "kernel_ms_per_launch_rank0": kernels, "scaling": "weak", "dtype": "f64",
                          "hbm_frac_per_gpu": 160.0 * args.per_gpu * args.steps / (ms * 1e-3) / 1e9 / 6520.5}))
    m.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
