"""Asynchronous mode of the C ABI (ab_set_async): two independent models queued on their own streams
(an ensemblerun!-style driver, src/simulations/ensemblerun.jl) give bitwise the results of the same
calls made synchronously; device-side failures surface at ab_sync with the reference's message
(src/spaces/continuous.jl:160)."""
import ctypes as C

import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import FLOCK, flock_init

pytestmark = pytest.mark.gpu


def _pinned(lib, shape):
    n = int(np.prod(shape)) * 8
    p = lib.ab_alloc_pinned(n)
    assert p
    buf = (C.c_double * (n // 8)).from_address(p)
    return p, np.frombuffer(buf, dtype=np.float64).reshape(shape)


def _model(extent, n, seed):
    pos, vel = flock_init(n, extent, seed)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos, vel)
    return m, pos, vel


def test_async_ensemble_equals_sync():
    extent, n, steps = (300.0, 300.0), 30000, 6
    want = []
    for seed in (5, 6):
        m, pos, vel = _model(extent, n, seed)
        p, v = pos.copy(), vel.copy()
        for _ in range(steps):                      # synchronous round trip through host rows every step
            m.upload_rows("pos", p); m.upload_rows("vel", v)
            m.step(1)
            m.download_rows("pos", p); m.download_rows("vel", v)
        want.append((p, v))
        m.close()
    ens = []
    for seed in (5, 6):
        m, pos, vel = _model(extent, n, seed)
        pp, hp = _pinned(m.lib, (n, 2))
        pv, hv = _pinned(m.lib, (n, 2))
        hp[:] = pos; hv[:] = vel
        m.set_async(True)
        ens.append((m, hp, hv, pp, pv))
    for _ in range(steps):
        for m, hp, hv, _, _ in ens:                 # nothing waits in here
            m.upload_rows("pos", hp); m.upload_rows("vel", hv)
            m.step(1)
            m.download_rows("pos", hp); m.download_rows("vel", hv)
    for (m, hp, hv, pp, pv), (p, v) in zip(ens, want):
        m.sync()
        assert np.array_equal(hp, p) and np.array_equal(hv, v)
        m.set_async(False)
        assert np.array_equal(m.pos, p)
        m.close()
        m.lib.ab_free_pinned(pp); m.lib.ab_free_pinned(pv)


def test_async_errors_surface_at_sync():
    extent, n = (100.0, 100.0), 500
    m, pos, vel = _model(extent, n, 3)
    m.set_async(True)
    bad = pos.copy()
    bad[17, 0] = 250.0                              # outside the extent
    m.upload_rows("pos", bad)
    m.step(1)                                       # queued: no error yet
    with pytest.raises(ab.AgentsError, match="outside space extent"):
        m.sync()
    m.set_async(False)
    m.close()
