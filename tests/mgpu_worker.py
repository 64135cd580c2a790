"""Multi-GPU parity worker (launched by torchrun from tests/test_multigpu.py or by hand):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      tests/mgpu_worker.py --agents 60000 --extent 400 400 --steps 12

Every rank holds one slab of the same global flock; after `--steps` fused steps the gathered state must
be BITWISE equal to the single-GPU run of the same library on the same initial state (global ids keep
the in-cell order canonical), and the single-GPU run itself is checked against the CPU oracle by the
single-GPU parity tests.  Rank 0 prints one JSON line with the verdict.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--agents", dest="n", type=int, default=60000)
    ap.add_argument("--extent", type=float, nargs="+", default=[400.0, 400.0])
    ap.add_argument("--spacing", type=float, default=5.0)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--chunk", type=int, default=4, help="steps per ab_step_flock call")
    ap.add_argument("--search", default="approximate")
    ap.add_argument("--dtype", default="float64")
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--rule", default="flock", choices=["flock", "pairforce"])
    ap.add_argument("--walls", action="store_true", help="non-periodic last axis (the end slabs have one neighbour)")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    import agents_jl_b200 as ab
    from agents_jl_b200.parallel import Slabs
    from helpers import FLOCK, flock_init

    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    extent = tuple(args.extent)
    dt = np.dtype(args.dtype)
    cols = None
    if args.rule == "flock":
        pos, vel = flock_init(args.n, extent, args.seed, dtype=dt)
        make_rule = lambda: ab.FlockRule(search=args.search, kernel="fine", **FLOCK)
    else:
        # examples/celllistmap.jl-style particles; the velocity scale makes particles cross slab boundaries within a few steps
        from test_pairforce_gpu import particles

        pos, vel, rad, kk, mass = (a.astype(dt) for a in particles(args.n, extent, seed=args.seed, vscale=300.0))
        cols = {"r": rad, "k": kk, "mass": mass}
        make_rule = lambda: ab.PairForceRule(dt=0.005, cutoff=args.spacing)

    rule = make_rule()
    periodic = tuple([True] * (len(extent) - 1) + [not args.walls])
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=args.spacing, dtype=dt, periodic=periodic), model_step=rule, device=local)
    slabs = Slabs(m, dist)
    owned0 = slabs.scatter(pos, vel, columns=cols)
    done, owned_hist = 0, [owned0]
    while done < args.steps:
        k = min(args.chunk, args.steps - done)
        m.step(k)
        done += k
        owned_hist.append(slabs.n_owned())
    gp, gv = slabs.gather()                      # raises if an agent was lost or duplicated
    tot = torch.tensor([slabs.n_owned()], dtype=torch.int64, device="cuda")
    dist.all_reduce(tot)
    result = None
    if rank == 0:
        ref = ab.StandardABM(ab.ContinuousSpace(extent, spacing=args.spacing, dtype=dt, periodic=periodic), model_step=make_rule(), device=local)
        ref.add_agents(pos, vel, **(cols or {}))
        done = 0
        while done < args.steps:
            k = min(args.chunk, args.steps - done)
            ref.step(k)
            done += k
        rp, rv = ref.pos, ref.vel
        result = {
            "world": world, "n": args.n, "extent": list(extent), "steps": args.steps, "search": args.search,
            "dtype": args.dtype, "conserved": int(tot.item()) == args.n,
            "pos_bitwise_equal": bool(np.array_equal(gp, rp)), "vel_bitwise_equal": bool(np.array_equal(gv, rv)),
            "max_abs_pos_diff": float(np.max(np.abs(gp - rp))), "owned_rank0": owned_hist,
            "migrated": bool(len(set(owned_hist)) > 1), "launches": m.launch_count(),
            "direct": bool(slabs.direct), "rule": args.rule,
        }
        ref.close()
        print("MGPU_RESULT " + json.dumps(result), flush=True)
    dist.barrier()
    m.close()
    dist.destroy_process_group()
    if rank == 0 and not (result["conserved"] and result["pos_bitwise_equal"] and result["vel_bitwise_equal"]):
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
