#!/usr/bin/env python
"""bench.py — agent-updates/s of the fused flock step (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # our arm (CUDA, C ABI)
  python bench.py --impl reference --gpus N --steps K ...  # CPU arm (oracle port of Agents.jl)

Workload (config.workload = "F2D-1M", BASELINE configs[1]): 1,000,000 boids,
ContinuousSpace((2000,2000)) periodic, spacing = visual_distance = 5, f64,
examples/flock.jl rule constants, `search = :approximate` (the reference's
default), synthetic counter-based initial state, synchronous update.
A "step" = one flock step over the whole population.

value  : agents*steps / device time of ONE step!(model, K) call after W warm-up
         steps, state resident in HBM (CUDA events on the library's stream,
         max over ranks; the per-step working set exceeds L2, no flush).
e2e    : same metric through the public API with HOST (pinned) buffers: every
         model step uploads pos+vel, steps once, downloads pos+vel; an ensemble of
         6 independent models is kept in flight through the asynchronous C ABI
         (the one-synchronous-model figure is reported beside it).
roofline: dominant kernel (the sweep) — algorithmic 128 B/agent-update
          x agents per launch / its CUDA-event duration, against the
         measured HBM copy bandwidth in MEASURED_PEAKS.json.  The per-launch events
         are recorded in a twin pass over the same steps (they cost ~22 us per step,
         so the timed region carries none).
cpu_baseline: the CPU oracle's reference-faithful asynchronous step
         (vector-per-cell index, in-place sequential update, 1 thread — the
         reference's step! is single-threaded) on a bounded sample.
With --gpus N > 1 (torchrun): weak scaling, 1M boids per GPU, slab decomposition
along y, halo / migration exchange by direct peer-memory stores (AB_HALO=nccl:
ncclSend/ncclRecv).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

ALG_BYTES = {"flock2d_f64": 128.0, "flock3d_f64": 160.0, "flock2d_f32": 72.0}
N_PER_GPU = 1_000_000
EXTENT = (2000.0, 2000.0)
SPACING = 5.0
FLOCK = dict(speed=1.0, cohere_factor=0.1, separation=2.0, separate_factor=0.25, match_factor=0.04,
             visual_distance=5.0)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """SM clock + clock-event (throttle) reasons sampled DURING the timed region.

    The timed region of this workload is milliseconds long, far below nvidia-smi's
    polling period, so the samples come from NVML directly (pynvml, the library
    nvidia-smi itself reads) on a background thread every ~1 ms; `nvidia-smi
    --query-gpu` is the fallback when pynvml is missing."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.sm, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        self.source = None

    def _nvml_loop(self, nv, dev):
        bits = (("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown),
                ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown),
                ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap))
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(dev, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(dev)
                for nm, b in bits:
                    if r & b:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.001)

    def _smi_loop(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
                r = [x.strip() for x in out.strip().split(",")]
                self.sm.append(float(r[0]))
                self.max_mhz = float(r[1])
                for nm, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                pass

    def start(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it lists ordinals
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            idx = self.index
            if vis and all(t.strip().isdigit() for t in vis.split(",")):
                idx = int(vis.split(",")[self.index])
            dev = nv.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(dev, nv.NVML_CLOCK_SM))
            self.source = "nvml"
            self._thread = threading.Thread(target=self._nvml_loop, args=(nv, dev), daemon=True)
        except Exception:
            self.source = "nvidia-smi"
            self._thread = threading.Thread(target=self._smi_loop, daemon=True)
        self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=15)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": self.source}


def flock_init_seed(n, extent, seed):
    from helpers import flock_init

    return flock_init(n, extent, seed)


def flock_state(n, extent, seed=42):
    from helpers import flock_init

    return flock_init(n, extent, seed)


# ---------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's own (asynchronous, 1-thread) step
# ---------------------------------------------------------------------------
def cpu_flock_rate(n, extent, steps, warmup=0):
    from oracle.pyoracle import ContinuousOracle

    pos, vel = flock_state(n, extent)
    o = ContinuousOracle(extent, spacing=SPACING)
    o.add_agents(pos, vel)
    rng_order = np.random.default_rng(1)
    for _ in range(warmup):
        o.flock_step_async(FLOCK, order=rng_order.permutation(n) + 1)     # Schedulers.Randomly (flock.jl:59)
    t0 = time.perf_counter()
    for _ in range(steps):
        o.flock_step_async(FLOCK, order=rng_order.permutation(n) + 1)
    dt = time.perf_counter() - t0
    return n * steps / dt, dt


def cpu_flock_rate_sync_all_cores(n, extent, steps):
    """The synchronous CPU formulation (the parity oracle) with its gather phase on all host cores — NOT the
    reference (Agents.jl's step! is single-task); reported beside the single-thread baseline."""
    from oracle.pyoracle import ContinuousOracle, set_threads

    cores = set_threads(0)
    pos, vel = flock_state(n, extent)
    o = ContinuousOracle(extent, spacing=SPACING)
    o.add_agents(pos, vel)
    t0 = time.perf_counter()
    o.flock_step_sync(FLOCK, nsteps=steps)
    dt = time.perf_counter() - t0
    set_threads(1)
    return {"value": n * steps / dt, "unit": "agent-updates/s", "cores": cores, "kind": "port",
            "sample": f"{n} boids x {steps} synchronous steps, gather phase on {cores} threads, {dt:.1f} s",
            "note": "CPU restatement with the synchronous update, not Agents.jl"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n = N_PER_GPU
    # bounded sample: the full 1M-boid population, few steps (each ~10 s on one core)
    steps = max(1, min(args.steps, 3))
    warm = 1 if args.warmup > 0 else 0
    rate, dt = cpu_flock_rate(n, EXTENT, steps, warm)
    sample = f"{n} boids x {steps} asynchronous steps (Schedulers.Randomly order), 1 thread"
    line = {
        "impl": "reference", "metric": "agent-updates/sec", "value": rate, "unit": "agent-updates/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": dt / steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "F2D-1M", "agents": n, "extent": list(EXTENT), "spacing": SPACING,
                   "search": "approximate", "update": "asynchronous in place (reference semantics)"},
        "cpu_baseline": {"value": rate, "unit": "agent-updates/s", "cores": 1, "kind": "port", "sample": sample,
                         "note": "Julia is not installed; this is the CPU oracle port of Agents.jl's step! "
                                 "(single-threaded by design: src/simulations/step_standard.jl:9-23)"},
        "e2e": {"value": rate, "unit": "agent-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    line["cpu_baseline"]["sync_all_cores"] = cpu_flock_rate_sync_all_cores(n, EXTENT, 2)
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
def run_gpu(args):
    import torch

    import agents_jl_b200 as ab
    from agents_jl_b200 import _lib as L

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    n = N_PER_GPU
    search = args.search
    extent = EXTENT if world == 1 else (EXTENT[0], EXTENT[1] * world)     # weak scaling: slabs along y
    lib = L.lib()
    sampler = ClockSampler(local)

    rule = ab.FlockRule(search=search, kernel=args.kernel, **FLOCK)
    space = ab.ContinuousSpace(extent, spacing=SPACING, periodic=True)
    model = ab.StandardABM(space, model_step=rule, device=local)
    if world == 1:
        pos, vel = flock_state(n, extent)
        model.add_agents(pos, vel)
    else:
        from agents_jl_b200.parallel import init_slabs

        slabs = init_slabs(model, dist, n * world, extent, flock_state)
    nloc = n

    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def kernel_table(m):
        names = (C.c_char_p * 32)()
        kms = (C.c_double * 32)()
        kl = (C.c_int64 * 32)()
        nk = lib.ab_kernel_times(m.h, names, kms, kl, 32)
        return {names[i].decode(): {"ms_per_launch": kms[i] / max(kl[i], 1), "launches": int(kl[i]), "ms": kms[i]}
                for i in range(min(nk, 32))}

    # ---- device-resident throughput: the timed region is ONE step!(model, K) call ---------------
    # (K fused steps back to back on the library's stream, no host work in between; the state a step
    #  streams — 2 x (pos, vel, id, key) + step scratch = 130 MB per 1 M boids — exceeds the 126 MB L2, so
    #  there is no flush between steps; the flushed per-step figure is reported beside it.)
    # Per-launch CUDA events cost ~5 us per launch (22 us of a 260 us step), so the timed region runs
    # WITHOUT them and the per-kernel durations come from a twin pass over the SAME steps (W..W+K from the
    # same initial state; N = 1) with the events on; with several ranks the twin pass is the next K steps.
    model.step(args.warmup)
    barrier()
    sampler.start()
    launches0 = model.launch_count()
    model.step(args.steps)
    ms = model.last_step_ms()               # CUDA events around the K steps on the library's stream
    launches = model.launch_count() - launches0
    barrier()
    clocks = sampler.stop()
    if world == 1:
        twin = ab.StandardABM(ab.ContinuousSpace(extent, spacing=SPACING, periodic=True), model_step=rule, device=local)
        twin.add_agents(pos, vel)
        twin.step(args.warmup)
    else:
        twin = model
    barrier()
    lib.ab_set_profiling(twin.h, 1)         # events recorded around every launch, resolved after the call
    twin.step(args.steps)
    ms_profiled = twin.last_step_ms()
    barrier()
    kernels = kernel_table(twin)
    lib.ab_set_profiling(twin.h, 0)
    if twin is not model:
        twin.close()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    total_agents = n * world
    value = total_agents * args.steps / (ms * 1e-3)

    # per-step calls with an L2 flush (256 MB write) before each one; N = 1 only: with several ranks the
    # unsynchronised host-side flushes would be timed as waiting for the neighbour
    flushed = None
    if world == 1:
        kf = min(args.steps, 20)
        tot = 0.0
        for _ in range(kf):
            flush.add_(1.0)
            torch.cuda.synchronize()
            model.step(1)
            tot += model.last_step_ms()
        flushed = {"value": n * kf / (tot * 1e-3), "ms_per_step": tot / kf, "steps": kf,
                   "note": "the steps after the timed region (the flock is more clustered there), one call per step, "
                           "L2 flushed before each"}

    step_kernel_ms = sum(v["ms"] for v in kernels.values()) / args.steps
    dom = max(kernels, key=lambda k: kernels[k]["ms"]) if kernels else None
    peak, peak_src = measured_peaks()
    alg = ALG_BYTES["flock2d_f64"]
    roofline = None
    if dom:
        dur = kernels[dom]["ms_per_launch"] * 1e-3
        ach = alg * nloc / dur / 1e9
        roofline = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                    "frac": ach / peak, "traffic": None, "peak_source": peak_src,
                    "algorithmic_bytes_per_agent_update": alg, "kernel_ms": kernels[dom]["ms_per_launch"],
                    "kernel_share_of_step": kernels[dom]["ms"] / max(sum(v["ms"] for v in kernels.values()), 1e-12),
                    "whole_step": {"achieved": alg * nloc / (ms / args.steps * 1e-3) / 1e9,
                                   "frac": alg * nloc / (ms / args.steps * 1e-3) / 1e9 / peak},
                    "profiled_pass_ms_per_step": ms_profiled / args.steps,
                    "note": "launch durations from CUDA events around every launch in a twin pass over the same steps "
                            "(rank 0); the timed region itself carries no per-launch events (they cost ~22 us per step)"}
        tpath = os.path.join(ROOT, "profiles", "traffic.json")      # dram bytes per launch from an ncu --set full capture
        if os.path.exists(tpath):
            try:
                roofline["traffic"] = json.load(open(tpath)).get(dom)
            except Exception:
                pass

    # ---- end to end through the public API with host buffers --------------------
    # Every model step = upload pos+vel from pinned host rows, step!(model, 1), download pos+vel into
    # the same rows (the next step uploads what this one downloaded).  `single_model`: one model, one
    # ab_sync per step.  `value`: an ensemble of R independent models (ensemblerun!-style) queued
    # asynchronously on their own streams, so one model's download overlaps the other's upload and step.
    e2e = None
    if world == 1:
        def pinned_rows(m):
            hp = torch.empty((n, 2), dtype=torch.float64, pin_memory=True).numpy()
            hv = torch.empty((n, 2), dtype=torch.float64, pin_memory=True).numpy()
            hp[:] = m.pos
            hv[:] = m.vel
            return hp, hv

        def queue_step(m, hp, hv):
            m.upload_rows("pos", hp)
            m.upload_rows("vel", hv)
            m.step(1)
            m.download_rows("pos", hp)
            m.download_rows("vel", hv)

        R = max(2, args.ensemble)
        ens = [(model, *pinned_rows(model))]
        for r in range(1, R):
            mr = ab.StandardABM(ab.ContinuousSpace(extent, spacing=SPACING, periodic=True),
                                model_step=ab.FlockRule(search=search, kernel=args.kernel, **FLOCK), device=local)
            pr, vr = flock_init_seed(n, extent, 42 + r)
            mr.add_agents(pr, vr)
            mr.step(args.warmup)
            ens.append((mr, *pinned_rows(mr)))
        for m, _, _ in ens:
            m.set_async(True)
        m0, hp0, hv0 = ens[0]
        for _ in range(max(args.warmup, 3)):
            queue_step(m0, hp0, hv0)
            m0.sync()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            queue_step(m0, hp0, hv0)
            m0.sync()
        dt1 = time.perf_counter() - t0
        for _ in range(3):
            for m, hp, hv in ens:
                queue_step(m, hp, hv)
        for m, _, _ in ens:
            m.sync()
        t0 = time.perf_counter()
        for it in range(args.steps):
            for m, hp, hv in ens:
                queue_step(m, hp, hv)
            if it % 8 == 7:
                ens[0][0].sync()              # bound the queue depth
        for m, _, _ in ens:
            m.sync()
        dt2 = time.perf_counter() - t0
        for m, _, _ in ens:
            m.set_async(False)
        nb = int(hp0.nbytes + hv0.nbytes)
        e2e = {"value": R * n * args.steps / dt2, "unit": "agent-updates/s", "h2d_bytes_per_step": nb,
               "d2h_bytes_per_step": nb, "ms_per_step": dt2 / (R * args.steps) * 1e3,
               "mode": f"ensemble of {R} independent F2D-1M models, asynchronous C-ABI calls, one stream per model; every model "
                       "step uploads and downloads pos+vel (pinned host rows); ms_per_step is per model step",
               "single_model": {"value": n * args.steps / dt1, "ms_per_step": dt1 / args.steps * 1e3,
                                "note": "one model, upload -> step -> download -> ab_sync every step"}}
        for m, _, _ in ens[1:]:
            m.close()

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only) ------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        csteps = 2
        rate, dt = cpu_flock_rate(n, EXTENT, csteps)
        cpu = {"value": rate, "unit": "agent-updates/s", "cores": 1, "kind": "port",
               "sample": f"{n} boids x {csteps} asynchronous steps (reference semantics), {dt:.1f} s, 1 thread",
               "sync_all_cores": cpu_flock_rate_sync_all_cores(n, EXTENT, 2)}

    if rank == 0:
        line = {
            "metric": "agent-updates/sec", "value": value, "unit": "agent-updates/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "F2D-1M", "agents_per_gpu": n, "agents": total_agents, "extent": list(extent),
                       "spacing": SPACING, "periodic": True, "search": search, "update": "synchronous",
                       "l2": "no flush: one call of K fused steps; per-step working set (130 MB per 1 M boids) exceeds the 126 MB L2",
                       "kernel": args.kernel,
                       "parallelism": "single GPU" if world == 1 else f"{world} slabs along y, halo+migration by " + ("direct peer-memory stores" if slabs.direct else "NCCL send/recv")},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "l2_flushed_per_step": flushed,
            **({} if world == 1 else {"e2e_note": "the host-buffer leg is measured at N = 1 (one model per handle); a sharded model "
                                                  "stays resident across steps, its halo and migration travel over NVLink, not PCIe"}),
            "kernels": kernels, "kernel_ms_per_step": step_kernel_ms,
        }
        print(json.dumps(line))
    model.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--search", default="approximate", choices=["approximate", "exact"])
    ap.add_argument("--kernel", default="auto", choices=["auto", "generic", "fine"])
    ap.add_argument("--ensemble", type=int, default=6, help="independent models in flight in the e2e leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
