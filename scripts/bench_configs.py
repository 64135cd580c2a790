#!/usr/bin/env python
"""Throughput of the BASELINE.json configurations other than the headline (bench.py covers F2D-1M):

  python scripts/bench_configs.py [--only f2d_exact,f2d_f32,pf2d,sch,f3d] [--scale 1.0] [--profile]

Each line printed is one JSON object: config, agents, steps, ms_per_step (CUDA events on the
library's stream, state resident), agent-updates/s, the algorithmic-bytes roofline fraction
 and the per-kernel times.  Sizes: PF2D-10M (10 M particles, cutoff 20, 791x791 cells),
SCH-8192 (8192x8192, 75 % occupancy), F3D-12.5M (one GPU's share of F3D-100M: 280x280x35 cells),
F2D-1M with search=:exact and with Float32 agents.  --scale shrinks the populations (and the
extents with them) for smoke runs.
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import agents_jl_b200 as ab  # noqa: E402
from helpers import FLOCK, flock_init, uniform  # noqa: E402

PEAK = 6530.6
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def kernel_times(m):
    lib = m.lib
    names = (C.c_char_p * 32)()
    kms = (C.c_double * 32)()
    kl = (C.c_int64 * 32)()
    nk = lib.ab_kernel_times(m.h, names, kms, kl, 32)
    return {names[i].decode(): {"ms_per_launch": kms[i] / max(kl[i], 1), "launches": int(kl[i])} for i in range(min(nk, 32))}


def run(m, name, n, steps, warmup, alg_bytes, extra=None, profile=False):
    import torch

    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")
    for _ in range(warmup):
        m.step(1)
    tot = 0.0
    for _ in range(steps):
        flush.add_(1.0)
        torch.cuda.synchronize()
        m.step(1)
        tot += m.last_step_ms()
    ms = tot / steps
    m.step(steps)
    warm = m.last_step_ms() / steps
    kern = None
    if profile:
        m.lib.ab_set_profiling(m.h, 1)
        for _ in range(min(steps, 5)):
            m.step(1)
        kern = kernel_times(m)
        m.lib.ab_set_profiling(m.h, 0)
    rate = n / (ms * 1e-3)
    line = {"config": name, "agents": n, "steps": steps, "warmup": warmup, "ms_per_step": ms,
            "agent_updates_per_s": rate, "ms_per_step_warm_l2": warm,
            "algorithmic_bytes_per_agent_update": alg_bytes,
            "hbm_roofline_frac_whole_step": rate * alg_bytes / 1e9 / PEAK, "hbm_peak_gbs": PEAK, "kernels": kern}
    if extra:
        line.update(extra)
    print(json.dumps(line), flush=True)


def particles(n, extent, seed=42, max_radius=10.0, vscale=100.0):
    D = len(extent)
    pos = uniform(n, D, seed, 0) * np.asarray(extent)[None, :]
    u1 = np.maximum(uniform(n, D, seed, 1), 1e-300)
    u2 = uniform(n, D, seed, 2)
    vel = vscale * np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)
    w = uniform(n, 3, seed, 3)
    return pos, vel, (0.5 + 0.9 * w[:, 0]) * max_radius, 10.0 + 20.0 * w[:, 1], 10.0 + 100.0 * w[:, 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="f2d_exact,f2d_f32,pf2d,pf2d_pairs,sch,f3d")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--profile", action="store_true")
    a = ap.parse_args()
    only = set(a.only.split(","))
    sc = a.scale

    if "f2d_exact" in only or "f2d_f32" in only:
        side = int(round(400 * sc ** 0.5))
        extent = (5.0 * side, 5.0 * side)
        n = int(6.25 * side * side)
        for name, dt, search, alg in (("F2D-1M exact f64", np.float64, "exact", 128.0),
                                      ("F2D-1M approximate f32", np.float32, "approximate", 72.0)):
            if ("f2d_exact" if search == "exact" else "f2d_f32") not in only:
                continue
            pos, vel = flock_init(n, extent, 42, dtype=dt)
            m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0, dtype=dt), model_step=ab.FlockRule(search=search, **FLOCK))
            m.add_agents(pos, vel)
            run(m, name, n, a.steps, a.warmup, alg, profile=a.profile)
            m.close()

    if "pf2d" in only:
        side = int(round(791 * sc ** 0.5))
        extent = (20.0 * side, 20.0 * side)
        n = int(0.04 * extent[0] * extent[1])
        pos, vel, rad, k, mass = particles(n, extent)
        m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=20.0), model_step=ab.PairForceRule(dt=0.001, cutoff=20.0))
        t0 = time.perf_counter()
        m.add_agents(pos, vel, r=rad, k=k, mass=mass)
        run(m, "PF2D-10M", n, a.steps, a.warmup, 104.0, {"setup_s": time.perf_counter() - t0}, profile=a.profile)
        if "pf2d_pairs" in only:
            # the separate row of SURVEY §8d: materialising interacting_pairs(model, cutoff, :all; search = :exact)
            # (continuous.jl:418-452) — 16 B per pair on the device, copied to two host Int64 vectors
            import ctypes as C

            lib = m.lib
            npairs = C.c_int64(0)
            t0 = time.perf_counter()
            m._ck(lib.ab_pairs_count(m.h, 20.0, 1, C.byref(npairs)))
            t_count = time.perf_counter() - t0
            P = npairs.value
            lo, hi = np.empty(P, dtype=np.int64), np.empty(P, dtype=np.int64)
            t0 = time.perf_counter()
            m._ck(lib.ab_pairs_fill(m.h, m._ptr(lo), m._ptr(hi), P))
            t_fill = time.perf_counter() - t0
            ok = bool(np.all(lo < hi))
            print(json.dumps({"config": "PF2D-10M interacting_pairs(:all, exact, r = 20) materialised", "agents": n, "pairs": P,
                              "pairs_per_agent": 2.0 * P / n, "count_s": t_count, "fill_and_copy_s": t_fill,
                              "pairs_per_s": P / (t_count + t_fill), "bytes": 16 * P, "sorted_pairs": ok}), flush=True)
        m.close()

    if "sch" in only:
        side = int(round(8192 * sc ** 0.5))
        dims = (side, side)
        ncell = side * side
        n = int(ncell * 0.75)
        perm = np.argsort(uniform(ncell, 1, 42)[:, 0], kind="stable")[:n]
        pos = np.stack([perm % side + 1, perm // side + 1], axis=1).astype(np.int32)
        group = (np.arange(n) % 2 + 1).astype(np.int32)
        for relocate, alg in ((False, 10.33), (True, 10.33)):
            m = ab.StandardABM(ab.GridSpaceSingle(dims, periodic=False),
                               model_step=ab.SchellingRule(min_to_be_happy=3, relocate=relocate, seed=42))
            m.add_agents(pos, group=group, mood=np.zeros(n, dtype=np.uint8))
            run(m, "SCH-8192 " + ("full step (relocation)" if relocate else "happiness pass"), n, a.steps, a.warmup, alg,
                profile=a.profile)
            m.close()

    if "f3d" in only:
        side = int(round(280 * sc ** (1.0 / 3.0)))
        layers = max(8, int(round(35 * sc ** (1.0 / 3.0))))
        extent = (5.0 * side, 5.0 * side, 5.0 * layers)
        n = int(4.5554 * side * side * layers)
        pos, vel = flock_init(n, extent, 42)
        m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
        m.add_agents(pos, vel)
        run(m, "F3D-12.5M (one GPU's share of F3D-100M, periodic box)", n, max(5, a.steps // 2), a.warmup, 160.0,
            profile=a.profile)
        m.close()


if __name__ == "__main__":
    main()
