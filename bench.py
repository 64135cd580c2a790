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
         model step uploads pos+vel, steps once, downloads pos+vel.  `e2e.value`
         is ONE model, one synchronisation per step (like for like with the
         single-model reference arm); `e2e.ensemble` is the throughput of several
         independent models in flight through the asynchronous C ABI.
roofline: dominant kernel (the sweep) — algorithmic 128 B/agent-update
          x agents per launch / its CUDA-event duration, against the
         measured HBM copy bandwidth in MEASURED_PEAKS.json.  The per-launch events
         are recorded in a twin pass over the SAME steps (same initial state, same
         warm-up; also with several ranks), so the timed region carries none.
configs: (N = 1) driver-timed rows for the other BASELINE configurations on one
         GPU: F2D-1M :exact, F2D-1M Float32, PF2D-10M, SCH-8192 (happiness pass and
         full step), F3D-12.5M (one GPU's share of F3D-100M) — each with its own
         ms_per_step, roofline fraction and NVML clock sample.
cpu_baseline: the CPU oracle's reference-faithful asynchronous step
         (vector-per-cell index, in-place sequential update, 1 thread — the
         reference's step! is single-threaded) on a bounded sample.
With --gpus N > 1 (torchrun): weak scaling, 1M boids per GPU, slab decomposition
along y, halo / migration exchange by direct peer-memory stores (AB_HALO=nccl:
ncclSend/ncclRecv).  Before the timed region the ranks run reduced sharded cases
(flock 2-D, flock 3-D, pair force 2-D) and compare the gathered state with an
unsharded run on rank 0: `parity_check.bitwise_equal` (the process exits non-zero
on a mismatch).  The timed call starts behind a device-side all-reduce
(ab_comm_align) so rank launch skew is not charged to the K steps;
`halo_wait_us_per_step` says how long each rank's halo install waited for its
neighbours.  `f3d_weak` is the F3D row of BASELINE configs[4]: 12.5 M boids per
GPU, slabs along z (8 GPUs = F3D-100M); `pf2d_weak` the pair-force rule on slabs,
10 M particles per GPU; `e2e` the host-buffer leg of the sharded model (every rank
uploads / downloads the rows it owns, collective halo refresh in between).
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

ALG_BYTES = {"flock2d_f64": 128.0, "flock3d_f64": 160.0, "flock2d_f32": 72.0, "pairforce2d_f64": 104.0, "schelling": 10.33}
# what the fast path really moves per agent-update when the six rule parameters are uniform (they travel as kernel
# arguments, not as the 48 B of per-agent fields SURVEY 8(d) counts): read pos + vel 32 + key 4, write pos + vel 32 + key 4,
# id read + write 8
ALG_BYTES_MOVED = {"flock2d_f64_uniform": 80.0}
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
        return self

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


def flock_chunks(n, extent, seed, step):
    """The counter-based initial state in id chunks (no rank ever holds a 100 M-boid array)."""
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


def particles(n, extent, seed=42, max_radius=10.0, vscale=100.0):
    """examples/celllistmap.jl:73-128-style particles: uniform positions, normal velocities, radius / k / mass columns."""
    from helpers import uniform

    D = len(extent)
    pos = uniform(n, D, seed, 0) * np.asarray(extent)[None, :]
    u1 = np.maximum(uniform(n, D, seed, 1), 1e-300)
    u2 = uniform(n, D, seed, 2)
    vel = vscale * np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)
    w = uniform(n, 3, seed, 3)
    return pos, vel, (0.5 + 0.9 * w[:, 0]) * max_radius, 10.0 + 20.0 * w[:, 1], 10.0 + 100.0 * w[:, 2]


def particle_chunks(n, extent, seed, step, max_radius=10.0, vscale=100.0):
    """`particles` in id chunks: (first id, pos, vel, {r, k, mass}) — the same numbers for the same ids."""
    from helpers import splitmix64

    D = len(extent)
    ext = np.asarray(extent, dtype=np.float64)[None, :]

    def u(first, m, ncomp, stream):
        with np.errstate(over="ignore"):
            i = np.arange(first, first + m, dtype=np.uint64)[:, None]
            c = (np.arange(ncomp, dtype=np.uint64) + np.uint64(stream * 4))[None, :]
            z = splitmix64(np.uint64(seed) ^ (i * np.uint64(8) + c))
        return (z >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)

    for first in range(0, n, step):
        m = min(step, n - first)
        pos = u(first, m, D, 0) * ext
        vel = vscale * np.sqrt(-2.0 * np.log(np.maximum(u(first, m, D, 1), 1e-300))) * np.cos(2.0 * np.pi * u(first, m, D, 2))
        w = u(first, m, 3, 3)
        yield first, pos, vel, {"r": (0.5 + 0.9 * w[:, 0]) * max_radius, "k": 10.0 + 20.0 * w[:, 1], "mass": 10.0 + 100.0 * w[:, 2]}


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


def sync_all_cores_guarded(n, extent, steps):
    """The all-cores leg in a child process: it is optional context, so a crash or a hang in it (round 2: a data race in
    the oracle's offset-table cache aborted the N = 8 reference arm) must not take the result line with it."""
    code = ("import json, bench; print('SYNCLEG ' + json.dumps(bench.cpu_flock_rate_sync_all_cores(%d, %r, %d)))"
            % (n, tuple(extent), steps))
    try:
        r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=600)
        for ln in r.stdout.splitlines():
            if ln.startswith("SYNCLEG "):
                return json.loads(ln[8:])
        return {"unavailable": f"child exited with {r.returncode}: {r.stderr.strip()[-200:]}"}
    except Exception as e:                                     # timeout, spawn failure
        return {"unavailable": f"{type(e).__name__}: {e}"[:240]}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n = N_PER_GPU
    # bounded sample: the full 1M-boid population, 1 warm-up + 5 timed steps (each 2-4 s on one core)
    steps = max(1, min(args.steps, 5))
    warm = 1 if args.warmup > 0 else 0
    rate, dt = cpu_flock_rate(n, EXTENT, steps, warm)
    sample = f"{n} boids x {steps} asynchronous steps after {warm} warm-up step (Schedulers.Randomly order), 1 thread"
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
    line["cpu_baseline"]["sync_all_cores"] = sync_all_cores_guarded(n, EXTENT, 2)
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
def kernel_table(lib, m):
    names = (C.c_char_p * 32)()
    kms = (C.c_double * 32)()
    kl = (C.c_int64 * 32)()
    nk = lib.ab_kernel_times(m.h, names, kms, kl, 32)
    return {names[i].decode(): {"ms_per_launch": kms[i] / max(kl[i], 1), "launches": int(kl[i]), "ms": kms[i]}
            for i in range(min(nk, 32))}


FLOP_PER_CANDIDATE = 20.0          # min-image difference, d2, compare, masked accumulate
_FP64_PEAK = {}


def fp64_peak(lib, local):
    """Measured FP64 FMA peak of this GPU in TFLOP/s (ab_fp64_peak: independent fma chains, CUDA events), once per run."""
    if local not in _FP64_PEAK:
        t = C.c_double(0)
        rc = lib.ab_fp64_peak(int(local), C.byref(t))
        _FP64_PEAK[local] = t.value if rc == 0 and t.value > 0 else None
    return _FP64_PEAK[local]


def fp64_roofline(lib, local, candidates, updates_per_s, kernel_updates_per_s=None):
    """Secondary roofline: flop per agent-update = expected candidates per agent at the row's uniform
    density x 20 flop, against the measured FP64 peak."""
    pk = fp64_peak(lib, local)
    if not pk:
        return None
    f = candidates * FLOP_PER_CANDIDATE
    out = {"candidates_per_agent": candidates, "flops_per_agent_update": f, "peak_tflops": pk,
           "peak_source": "ab_fp64_peak, measured in this run",
           "whole_step": {"achieved_tflops": f * updates_per_s / 1e12, "frac": f * updates_per_s / 1e12 / pk}}
    if kernel_updates_per_s:
        out["achieved_tflops"] = f * kernel_updates_per_s / 1e12
        out["frac"] = out["achieved_tflops"] / pk
    return out


def timed_row(lib, m, name, n, steps, warmup, alg_bytes, local, peak, extra=None, candidates=None):
    """One driver-timed row of `configs`: W warm-up steps, ONE call of K fused steps timed by CUDA events on the
    library's stream (state resident), NVML clock samples inside the timed region; then the per-kernel durations of
    the NEXT min(K, 5) steps (per-launch events cost a few us each, so the timed call carries none)."""
    import torch

    m.step(warmup)
    torch.cuda.synchronize()
    sampler = ClockSampler(local).start()
    l0 = m.launch_count()
    m.step(steps)
    ms = m.last_step_ms() / steps
    launches = m.launch_count() - l0
    torch.cuda.synchronize()
    clocks = sampler.stop()
    lib.ab_set_profiling(m.h, 1)
    kp = min(steps, 5)
    m.step(kp)
    kern = kernel_table(lib, m)
    lib.ab_set_profiling(m.h, 0)
    dom = max(kern, key=lambda k: kern[k]["ms"]) if kern else None
    rate = n / (ms * 1e-3)
    row = {"config": name, "agents": int(n), "steps": steps, "warmup": warmup, "ms_per_step": ms, "value": rate,
           "unit": "agent-updates/s", "algorithmic_bytes_per_agent_update": alg_bytes,
           "roofline": {"bound": "hbm", "peak": peak, "unit": "GB/s",
                        "whole_step": {"achieved": rate * alg_bytes / 1e9, "frac": rate * alg_bytes / 1e9 / peak},
                        "kernel": dom,
                        "achieved": (alg_bytes * n / (kern[dom]["ms_per_launch"] * 1e-3) / 1e9) if dom else None,
                        "frac": (alg_bytes * n / (kern[dom]["ms_per_launch"] * 1e-3) / 1e9 / peak) if dom else None},
           "kernels_us_per_launch": {k: round(v["ms_per_launch"] * 1e3, 2) for k, v in kern.items()},
           "gpu_launches": int(launches), "clocks": clocks}
    if candidates:
        row["roofline"]["fp64"] = fp64_roofline(lib, local, candidates, rate,
                                                n / (kern[dom]["ms_per_launch"] * 1e-3) if dom else None)
    if extra:
        row.update(extra)
    return row


def config_rows(ab, lib, local, peak, scale=1.0):
    """BASELINE.json configs other than the headline, on one GPU (SURVEY §8d sizes)."""
    from helpers import flock_init, uniform

    rows = []
    sc = scale
    side = int(round(400 * sc ** 0.5))
    extent = (5.0 * side, 5.0 * side)
    n = int(6.25 * side * side)
    # expected candidates per agent at the row's uniform density: exact 2-D sweep 5 x 5 fine cells of 2.5 x 2.5 at
    # 0.25 / unit^2 = 39; pair force 3 x 3 cells of 20 x 20 at 0.04 = 144; 3-D sweep 27 fine cells of 2.5^3 at 0.0364 = 15.4
    for name, dt, search, alg, cand in (("F2D-1M exact f64", np.float64, "exact", ALG_BYTES["flock2d_f64"], 39.0),
                                        ("F2D-1M approximate f32", np.float32, "approximate", ALG_BYTES["flock2d_f32"], None)):
        pos, vel = flock_init(n, extent, 42, dtype=dt)
        m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0, dtype=dt), model_step=ab.FlockRule(search=search, **FLOCK), device=local)
        m.add_agents(pos, vel)
        rows.append(timed_row(lib, m, name, n, 20, 5, alg, local, peak, candidates=cand))
        m.close()

    side = int(round(791 * sc ** 0.5))
    extent = (20.0 * side, 20.0 * side)
    n = int(0.04 * extent[0] * extent[1])
    pos, vel, rad, k, mass = particles(n, extent)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=20.0), model_step=ab.PairForceRule(dt=0.001, cutoff=20.0), device=local)
    m.add_agents(pos, vel, r=rad, k=k, mass=mass)
    rows.append(timed_row(lib, m, "PF2D-10M (celllistmap.jl-style pair force, cutoff 20, interacting_pairs(:all) exact set)", n, 10, 3,
                          ALG_BYTES["pairforce2d_f64"], local, peak, candidates=144.0))
    # SURVEY 8(d)'s separate row: materialising interacting_pairs(model, cutoff, :all; search = :exact)
    # (continuous.jl:418-452) on the stepped state — count on the device, then fill and copy 16 B per pair to two host
    # Int64 vectors (the caller-owned result of the two-phase ABI)
    try:
        npairs = C.c_int64(0)
        t0 = time.perf_counter()
        m._ck(lib.ab_pairs_count(m.h, 20.0, 1, C.byref(npairs)))
        t_count = time.perf_counter() - t0
        P = npairs.value
        lo, hi = np.empty(P, dtype=np.int64), np.empty(P, dtype=np.int64)
        t0 = time.perf_counter()
        m._ck(lib.ab_pairs_fill(m.h, m._ptr(lo), m._ptr(hi), P))
        t_fill = time.perf_counter() - t0
        rows.append({"config": "PF2D-10M interacting_pairs(:all, exact, r = 20) materialised", "agents": int(n), "pairs": int(P),
                     "pairs_per_agent": 2.0 * P / n, "count_s": t_count, "fill_and_copy_s": t_fill,
                     "value": P / (t_count + t_fill), "unit": "pairs/s", "bytes_to_host": 16 * P,
                     "every_pair_lo_lt_hi": bool(np.all(lo < hi)), "timing": "host wall clock around the two synchronous calls"})
        del lo, hi
    except Exception as e:                                    # host memory for 2 x 8 B x 250 M pairs may be missing
        rows.append({"config": "PF2D-10M interacting_pairs(:all, exact, r = 20) materialised", "unavailable": str(e)[:200]})
    m.close()
    del pos, vel, rad, k, mass

    side = int(round(8192 * sc ** 0.5))
    ncell = side * side
    n = int(ncell * 0.75)
    perm = np.argsort(uniform(ncell, 1, 42)[:, 0], kind="stable")[:n]
    pos = np.stack([perm % side + 1, perm // side + 1], axis=1).astype(np.int32)
    group = (np.arange(n) % 2 + 1).astype(np.int32)
    for relocate in (False, True):
        m = ab.StandardABM(ab.GridSpaceSingle((side, side), periodic=False),
                           model_step=ab.SchellingRule(min_to_be_happy=3, relocate=relocate, seed=42), device=local)
        m.add_agents(pos, group=group, mood=np.zeros(n, dtype=np.uint8))
        rows.append(timed_row(lib, m, "SCH-8192 " + ("full step (happiness + relocation)" if relocate else "happiness pass (nearby_ids radius 1)"),
                              n, 20, 5, ALG_BYTES["schelling"], local, peak))
        m.close()
    del pos, group, perm

    side = int(round(280 * sc ** (1.0 / 3.0)))
    layers = max(8, int(round(35 * sc ** (1.0 / 3.0))))
    extent = (5.0 * side, 5.0 * side, 5.0 * layers)
    n = int(4.5554 * side * side * layers)
    pos, vel = flock_init(n, extent, 42)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK), device=local)
    m.add_agents(pos, vel)
    rows.append(timed_row(lib, m, "F3D-12.5M (one GPU's share of F3D-100M, periodic box)", n, 10, 3, ALG_BYTES["flock3d_f64"], local, peak, candidates=15.4))
    m.close()
    return rows


def parity_check(ab, dist, local, rank, world):
    """Reduced sharded cases (the cases of tests/test_multigpu.py, stretched so every rank owns >= 4 cell layers):
    after 8 fused steps the gathered state must be BITWISE equal to the unsharded run of the same library on rank 0."""
    from agents_jl_b200.parallel import Slabs
    from helpers import flock_init

    # the last case forces the path the timed F3D-100M row takes (messages too large for the re-sort to deliver:
    # k_halo_pack launch, cell sums split around the side-stream install) onto a state small enough to compare
    cases = [("flock 2d", (400.0, float(max(400, 100 * world))), 0.375, SPACING, {}),
             ("flock 3d", (60.0, 60.0, float(max(120, 40 * world))), 0.0926, SPACING, {}),
             ("pairforce 2d", (400.0, float(max(400, 100 * world))), 0.0375, 20.0, {}),
             ("flock 3d, large-message path (AB_FUSED_PACK_MAXMB=0)", (60.0, 60.0, float(max(120, 40 * world))), 0.0926, SPACING,
              {"AB_FUSED_PACK_MAXMB": "0"})]
    out = {"bitwise_equal": True, "migrated": 0, "direct": True, "cases": []}
    for name, extent, dens, spacing, env in cases:
        saved = {k: os.environ.get(k) for k in env}
        os.environ.update(env)                                # read by ab_comm_init (Slabs) of this case only
        n = int(dens * float(np.prod(extent)))
        cols = None
        if name.startswith("flock"):
            pos, vel = flock_init(n, extent, 42)
            make_rule = lambda: ab.FlockRule(kernel="fine", **FLOCK)
        else:
            pos, vel, rad, kk, mass = particles(n, extent, vscale=300.0)
            cols = {"r": rad, "k": kk, "mass": mass}
            make_rule = lambda: ab.PairForceRule(dt=0.005, cutoff=20.0)
        m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=make_rule(), device=local)
        s = Slabs(m, dist)
        for k, v in saved.items():
            os.environ.pop(k, None) if v is None else os.environ.__setitem__(k, v)
        own0 = s.scatter(pos, vel, columns=cols)
        m.step(4)
        own1 = s.n_owned()
        m.step(4)
        gp, gv = s.gather()                                   # raises if an agent was lost or duplicated
        direct = bool(s.direct)
        m.close()
        ok, migrated = True, abs(own1 - own0)
        if rank == 0:
            ref = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing), model_step=make_rule(), device=local)
            ref.add_agents(pos, vel, **(cols or {}))
            ref.step(4)
            ref.step(4)
            ok = bool(np.array_equal(gp, ref.pos) and np.array_equal(gv, ref.vel))
            ref.close()
        flags = [None] * world
        dist.all_gather_object(flags, (ok, migrated, direct))
        ok = all(f[0] for f in flags)
        out["cases"].append({"case": name, "agents": n, "extent": list(extent), "steps": 8, "bitwise_equal": ok,
                             "rank_population_change": [int(f[1]) for f in flags], "direct": all(f[2] for f in flags)})
        gates = not env                                       # the forced large-message case is reported, not gating
        out["cases"][-1]["gates_exit"] = gates
        if gates:
            out["bitwise_equal"] = out["bitwise_equal"] and ok
        out["all_cases_equal"] = out.get("all_cases_equal", True) and ok
        out["migrated"] += int(sum(f[1] for f in flags))
        out["direct"] = out["direct"] and all(f[2] for f in flags)
    return out


def f3d_weak(ab, lib, dist, local, rank, world, peak, steps=10, warmup=3, per_gpu=12_500_000):
    """BASELINE configs[4]: 3-D flock, 12.5 M boids per GPU, slabs along z (world = 8: F3D-100M)."""
    import torch
    from agents_jl_b200.parallel import Slabs, default_halo_capacity

    n = per_gpu * world
    extent = (1400.0, 1400.0, 175.0 * world)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0, periodic=True), model_step=ab.FlockRule(kernel="fine", **FLOCK), device=local)
    s = Slabs(m, dist)
    cap = default_halo_capacity(n, s.n_layers, factor=1.25)
    s.scatter_chunks(flock_chunks(n, extent, 42, 4_000_000), n, halo_capacity=cap)
    m.step(warmup)
    torch.cuda.synchronize()
    dist.barrier()
    sampler = ClockSampler(local).start()
    lib.ab_comm_wait_stats(m.h, None, None, 1)
    lib.ab_comm_align(m.h)
    m.step(steps)
    ms = m.last_step_ms()
    wait_us, nex = C.c_double(0), C.c_int64(0)
    lib.ab_comm_wait_stats(m.h, C.byref(wait_us), C.byref(nex), 1)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    waits = [None] * world
    dist.all_gather_object(waits, wait_us.value / max(steps, 1))
    m.close()
    rate = n * steps / (ms * 1e-3)
    return {"config": "F3D-100M" if world == 8 else f"F3D-{n / 1e6:g}M ({world} x 12.5M, weak scaling of F3D-100M)", "agents": n,
            "n_gpus": world, "extent": list(extent), "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "value": rate,
            "unit": "agent-updates/s", "direct": bool(s.direct), "halo_wait_us_per_step": [round(w, 1) for w in waits],
            "roofline": {"bound": "hbm", "peak": peak, "unit": "GB/s",
                         "whole_step_per_gpu": {"achieved": ALG_BYTES["flock3d_f64"] * rate / world / 1e9,
                                                "frac": ALG_BYTES["flock3d_f64"] * rate / world / 1e9 / peak}},
            "clocks_rank0": clocks}


def pf2d_weak(ab, lib, dist, local, rank, world, peak, steps=10, warmup=3):
    """Pair-force weak-scaling row (the pair-force rule shards too): ~10 M particles per GPU, slabs along y,
    the r / k / mass columns travel in the boundary messages."""
    import torch
    from agents_jl_b200.parallel import Slabs, default_halo_capacity

    side = 791
    extent = (20.0 * side, 20.0 * side * world)
    n = int(0.04 * extent[0] * extent[1])
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=20.0, periodic=True), model_step=ab.PairForceRule(dt=0.001, cutoff=20.0), device=local)
    s = Slabs(m, dist)
    cap = default_halo_capacity(n, s.n_layers, factor=1.5)
    s.scatter_chunks(particle_chunks(n, extent, 42, 4_000_000), n, halo_capacity=cap)
    m.step(warmup)
    torch.cuda.synchronize()
    dist.barrier()
    sampler = ClockSampler(local).start()
    lib.ab_comm_align(m.h)
    m.step(steps)
    ms = m.last_step_ms()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    m.close()
    rate = n * steps / (ms * 1e-3)
    return {"config": f"PF2D-{n / 1e6:.0f}M ({world} x 10M, pair-force rule on slabs)", "agents": n, "n_gpus": world, "extent": list(extent),
            "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "value": rate, "unit": "agent-updates/s", "direct": bool(s.direct),
            "roofline": {"bound": "hbm", "peak": peak, "unit": "GB/s",
                         "whole_step_per_gpu": {"achieved": ALG_BYTES["pairforce2d_f64"] * rate / world / 1e9,
                                                "frac": ALG_BYTES["pairforce2d_f64"] * rate / world / 1e9 / peak}},
            "clocks_rank0": clocks}


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
    peak, peak_src = measured_peaks()

    # ---- multi-GPU parity gate (before anything is timed) ------------------------
    parity = None
    if world > 1 and not args.no_parity:
        parity = parity_check(ab, dist, local, rank, world)
        if not parity["bitwise_equal"]:
            if rank == 0:
                print(json.dumps({"error": "sharded run differs from the unsharded run", "parity_check": parity}))
            dist.destroy_process_group()
            return 3

    rule = ab.FlockRule(search=search, kernel=args.kernel, **FLOCK)
    pos = vel = None
    if world == 1:
        pos, vel = flock_state(n, extent)
    else:
        gpos, gvel = flock_state(n * world, extent)

    def make_model():
        m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=SPACING, periodic=True), model_step=rule, device=local)
        if world == 1:
            m.add_agents(pos, vel)
            return m, None
        from agents_jl_b200.parallel import Slabs

        s = Slabs(m, dist)
        s.scatter(gpos, gvel)
        return m, s

    model, slabs = make_model()
    nloc = n
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput: the timed region is ONE step!(model, K) call ---------------
    # (K fused steps back to back on the library's stream, no host work in between; the state a step
    #  streams — 2 x (pos, vel, id, key) + step scratch = 130 MB per 1 M boids — exceeds the 126 MB L2, so
    #  there is no flush between steps; the flushed per-step figure is reported beside it.)
    # Per-launch CUDA events cost a few us per launch, so the timed region runs WITHOUT them and the per-kernel
    # durations come from a twin pass over the SAME steps (W..W+K from the same initial state) with the events on.
    # N > 1: the timed call is queued behind a device-side all-reduce (ab_comm_align): every rank's K steps start
    # when the slowest rank has arrived, so host launch skew after the barrier is not charged to them.
    model.step(args.warmup)
    barrier()
    sampler = ClockSampler(local).start()
    launches0 = model.launch_count()
    if world > 1:
        lib.ab_comm_wait_stats(model.h, None, None, 1)
        lib.ab_comm_align(model.h)
    model.step(args.steps)
    ms = model.last_step_ms()               # CUDA events around the K steps on the library's stream
    launches = model.launch_count() - launches0
    halo_wait = None
    if world > 1:
        wait_us, nex = C.c_double(0), C.c_int64(0)
        lib.ab_comm_wait_stats(model.h, C.byref(wait_us), C.byref(nex), 1)
        halo_wait = wait_us.value / max(args.steps, 1)
    barrier()
    clocks = sampler.stop()
    # ---- the timed configuration itself, compared bit for bit (N > 1): the state after W + K steps of the sharded run,
    # gathered by global id, against the unsharded run of the same steps on rank 0 (N x 1 M boids fit one GPU)
    if world > 1 and parity is not None and not args.no_fullsize_parity:
        gp, gv = slabs.gather()
        ok = True
        if rank == 0:
            ref = ab.StandardABM(ab.ContinuousSpace(extent, spacing=SPACING, periodic=True), model_step=rule, device=local)
            ref.add_agents(gpos, gvel)
            ref.step(args.warmup)
            ref.step(args.steps)
            ok = bool(np.array_equal(gp, ref.pos) and np.array_equal(gv, ref.vel))
            ref.close()
        del gp, gv
        flags = [None] * world
        dist.all_gather_object(flags, ok)
        ok = bool(flags[0])
        parity["cases"].append({"case": "flock 2d, 1M per GPU (the timed run itself)", "agents": int(n * world),
                                "extent": list(extent), "steps": int(args.warmup + args.steps), "bitwise_equal": ok,
                                "direct": bool(slabs.direct), "gates_exit": False})
        parity["all_cases_equal"] = parity.get("all_cases_equal", True) and ok   # reported in the line; the timing is kept
        barrier()
    twin, _tw_slabs = make_model()
    twin.step(args.warmup)
    barrier()
    lib.ab_set_profiling(twin.h, 1)         # events recorded around every launch, resolved after the call
    if world > 1:
        lib.ab_comm_align(twin.h)
    twin.step(args.steps)
    ms_profiled = twin.last_step_ms()
    barrier()
    kernels = kernel_table(lib, twin)
    lib.ab_set_profiling(twin.h, 0)
    twin.close()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_rank = ms
    ms = float(t[0])
    total_agents = n * world
    value = total_agents * args.steps / (ms * 1e-3)
    per_rank = None
    if dist is not None:
        per_rank = [None] * world
        dist.all_gather_object(per_rank, {"ms_per_step": ms_rank / args.steps, "halo_wait_us_per_step": round(halo_wait, 2)})

    # per-step calls with an L2 flush (256 MB write) before each one; N = 1 only: with several ranks the
    # unsynchronised host-side flushes would be timed as waiting for the neighbour
    flushed = None
    if world == 1:
        kf = args.steps
        fm, _ = make_model()                                  # the SAME steps as `value`: same initial state, same warm-up
        fm.step(args.warmup)
        tot = 0.0
        for _ in range(kf):
            flush.add_(1.0)
            torch.cuda.synchronize()
            fm.step(1)
            tot += fm.last_step_ms()
        fm.close()
        flushed = {"value": n * kf / (tot * 1e-3), "ms_per_step": tot / kf, "steps": kf,
                   "note": "steps W..W+K of a re-created model (the steps `value` is timed on), one call per step, "
                           "L2 flushed (256 MB write) before each"}

    step_kernel_ms = sum(v["ms"] for v in kernels.values()) / args.steps
    dom = max(kernels, key=lambda k: kernels[k]["ms"]) if kernels else None
    alg = ALG_BYTES["flock2d_f64"]
    roofline = None
    if dom:
        dur = kernels[dom]["ms_per_launch"] * 1e-3
        ach = alg * nloc / dur / 1e9
        roofline = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                    "frac": ach / peak, "traffic": None, "peak_source": peak_src,
                    "algorithmic_bytes_per_agent_update": alg,
                    "algorithmic_bytes_moved_per_agent_update": ALG_BYTES_MOVED["flock2d_f64_uniform"],
                    "frac_on_bytes_moved": ALG_BYTES_MOVED["flock2d_f64_uniform"] * nloc / dur / 1e9 / peak,
                    "kernel_ms": kernels[dom]["ms_per_launch"],
                    "kernel_share_of_step": kernels[dom]["ms"] / max(sum(v["ms"] for v in kernels.values()), 1e-12),
                    "whole_step": {"achieved": alg * nloc / (ms / args.steps * 1e-3) / 1e9,
                                   "frac": alg * nloc / (ms / args.steps * 1e-3) / 1e9 / peak,
                                   "frac_on_bytes_moved": ALG_BYTES_MOVED["flock2d_f64_uniform"] * nloc / (ms / args.steps * 1e-3) / 1e9 / peak},
                    "profiled_pass_ms_per_step": ms_profiled / args.steps,
                    "note": "launch durations from CUDA events around every launch in a twin pass over the SAME steps "
                            "(same initial state and warm-up, rank 0); the timed region itself carries no per-launch events"}
        # 3 x 3 fine cells of 2.5 x 2.5 at 0.25 / unit^2 = 14 candidates per bird in the separation sweep
        roofline["fp64"] = fp64_roofline(lib, local, 14.0, value / world, nloc / dur)
        tpath = os.path.join(ROOT, "profiles", "traffic.json")      # dram bytes per launch from an ncu --set full capture
        if os.path.exists(tpath):
            try:
                tj = json.load(open(tpath))
                roofline["traffic"] = tj.get(dom)
                roofline["traffic_source"] = ("profiles/traffic.json (committed; not measured in this run): "
                                              + str(tj.get("_note", "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full")))
            except Exception:
                pass

    # ---- end to end through the public API with host buffers --------------------
    # Every model step = upload pos+vel from pinned host rows, step!(model, 1), download pos+vel into
    # the same rows (the next step uploads what this one downloaded).  `value`: ONE model, one ab_sync per
    # step.  `ensemble`: R independent models (ensemblerun!-style) queued asynchronously on their own
    # streams, so one model's download overlaps another's upload and step.
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
        e2e = {"value": n * args.steps / dt1, "unit": "agent-updates/s", "h2d_bytes_per_step": nb,
               "d2h_bytes_per_step": nb, "ms_per_step": dt1 / args.steps * 1e3,
               "mode": "ONE F2D-1M model: upload pos+vel (pinned host rows) -> step!(model, 1) -> download pos+vel -> ab_sync, every step",
               "ensemble": {"value": R * n * args.steps / dt2, "models": R, "ms_per_model_step": dt2 / (R * args.steps) * 1e3,
                            "note": f"{R} independent F2D-1M models in flight (ensemblerun!-style), asynchronous C-ABI calls, one stream "
                                    "per model; every model step uploads and downloads pos+vel"}}
        for m, _, _ in ens[1:]:
            m.close()
    else:
        # sharded model: every rank keeps pinned host rows of the agents it owns; a step uploads them (in slot
        # order, global ids attached), refreshes the halo from the neighbours, steps once and downloads the owned rows
        e2e = e2e_sharded(torch, dist, lib, model, slabs, args, total_agents)

    # ---- the other BASELINE configurations, one GPU (N = 1 only) ----------------
    configs = None
    if world == 1 and not args.no_configs:
        model.close()
        model = None
        del flush
        torch.cuda.empty_cache()
        configs = config_rows(ab, lib, local, peak, args.config_scale)
    f3d = pf_weak = None
    if world > 1 and not args.no_f3d:
        model.close()
        model = None
        del flush
        torch.cuda.empty_cache()
        f3d = f3d_weak(ab, lib, dist, local, rank, world, peak)
        pf_weak = pf2d_weak(ab, lib, dist, local, rank, world, peak) if not args.no_pf else None

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only) ------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        csteps = 5
        rate, dt = cpu_flock_rate(n, EXTENT, csteps, 1)
        cpu = {"value": rate, "unit": "agent-updates/s", "cores": 1, "kind": "port",
               "sample": f"{n} boids x {csteps} asynchronous steps after 1 warm-up step (reference semantics), {dt:.1f} s, 1 thread",
               "sync_all_cores": sync_all_cores_guarded(n, EXTENT, 2)}

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
            "kernels": kernels, "kernel_ms_per_step": step_kernel_ms,
        }
        if world > 1:
            line["parity_check"] = parity
            line["per_rank"] = per_rank
            line["halo_wait_us_per_step"] = [p["halo_wait_us_per_step"] for p in per_rank]
            line["start_alignment"] = "device-side all-reduce on the library's stream before the first timed step (ab_comm_align)"
            line["f3d_weak"] = f3d
            line["pf2d_weak"] = pf_weak
        if configs is not None:
            line["configs"] = configs
        print(json.dumps(line))
    if model is not None:
        model.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def e2e_sharded(torch, dist, lib, model, slabs, args, total_agents):
    """Host-buffer leg of a sharded model (every rank, its own slab)."""
    if not hasattr(slabs, "host_step"):
        return None
    try:
        st = slabs.host_state()
        for _ in range(3):
            slabs.host_step(st, 1)
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            slabs.host_step(st, 1)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t[0])
        nb = torch.tensor([float(st["bytes_up"]), float(st["bytes_down"])], dtype=torch.float64, device="cuda")
        dist.all_reduce(nb)
        return {"value": total_agents * args.steps / dt, "unit": "agent-updates/s", "h2d_bytes_per_step": int(nb[0].item()),
                "d2h_bytes_per_step": int(nb[1].item()), "ms_per_step": dt / args.steps * 1e3,
                "mode": "every rank: upload the pos+vel rows it owns (pinned host memory) -> halo refresh from the neighbours -> "
                        "step!(model, 1) -> download the owned rows; wall clock, max over ranks; bytes summed over ranks"}
    except Exception as e:                                   # the device-resident figures above stay valid
        return {"value": None, "error": str(e)[:300]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--search", default="approximate", choices=["approximate", "exact"])
    ap.add_argument("--kernel", default="auto", choices=["auto", "generic", "fine"])
    ap.add_argument("--ensemble", type=int, default=6, help="independent models in flight in the e2e ensemble figure")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the rows of the other BASELINE configurations (N = 1)")
    ap.add_argument("--config-scale", type=float, default=1.0, help="shrink the populations of the config rows (smoke runs)")
    ap.add_argument("--no-fullsize-parity", action="store_true", help="skip the bitwise check of the timed sharded run (N > 1)")
    ap.add_argument("--no-parity", action="store_true", help="skip the sharded-vs-unsharded parity gate (N > 1)")
    ap.add_argument("--no-f3d", action="store_true", help="skip the F3D and PF2D weak-scaling rows (N > 1)")
    ap.add_argument("--no-pf", action="store_true", help="skip the PF2D weak-scaling row (N > 1)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
