"""Device-side aggregates for data collection (csrc/ab_collect.cu) against numpy on the downloaded
columns — the reference's collect_agent_data! with tuple-form adata (src/simulations/collect.jl:466-502).
Sums / means: relative 1e-12 (the reduction order differs from the reference's left fold); integer columns,
minimum / maximum / count: exact."""
import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import FLOCK, flock_init, uniform

pytestmark = pytest.mark.gpu


def test_flock_aggregates_follow_the_columns():
    extent = (400.0, 400.0)
    n = 70001                                     # not a multiple of the reduction chunk
    pos, vel = flock_init(n, extent, 4)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    energy = uniform(n, 1, 9)[:, 0] * 10.0
    kind = (np.arange(n) % 3).astype(np.int32)
    m.add_agents(pos, vel, energy=energy, kind=kind)
    m.step(3)                                     # columns are carried through the re-sorts
    P, V = m.pos, m.vel
    adata = [(("pos", 0), "mean"), (("vel", 1), "sum"), ("energy", "maximum"), ("energy", "minimum"),
             ("energy", "sum", ("kind", "==", 1)), ("kind", "sum"), ("kind", "count", ("energy", ">", 5.0)),
             (("pos", 1), "mean", (("pos", 0), "<", 200.0)), ("energy", "mean", ("kind", ">=", 7))]
    got = m.collect(adata)
    E, K = m.column("energy"), m.column("kind")
    assert np.array_equal(np.sort(E), np.sort(energy)) and np.array_equal(K, kind)
    want = [P[:, 0].mean(), V[:, 1].sum(), E.max(), E.min(), E[K == 1].sum(), K.sum(), (E > 5.0).sum(),
            P[P[:, 0] < 200.0, 1].mean()]
    for g, w in zip(got[:8], want):
        assert abs(g - w) <= 1e-12 * max(1.0, abs(w)), (g, w)
    assert got[2] == E.max() and got[3] == E.min() and got[5] == K.sum() and got[6] == (E > 5.0).sum()
    assert np.isnan(got[8])                       # mean over no agent
    with pytest.raises(ValueError):
        m.collect([("energy", "median")])
    with pytest.raises(KeyError):
        m.collect([("nope", "sum")])
    m.close()


def test_run_collects_every_when_ticks():
    extent = (200.0, 200.0)
    pos, vel = flock_init(8000, extent, 6)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos, vel)
    df = m.run(6, adata=[(("vel", 0), "mean"), (("pos", 0), "maximum")], when=2)
    assert df["time"] == [0, 2, 4, 6]
    assert len(df["mean_vel[1]"]) == 4 and len(df["maximum_pos[1]"]) == 4
    assert abs(df["mean_vel[1]"][0] - vel[:, 0].mean()) <= 1e-12
    assert abs(df["mean_vel[1]"][-1] - m.vel[:, 0].mean()) <= 1e-12
    m.close()


def test_schelling_segregation_series():
    """The tutorial's data collection: number of happy agents per step (docs/src/tutorial.jl:533-560)."""
    dims = (128, 96)
    ncell = dims[0] * dims[1]
    n = int(0.8 * ncell)
    perm = np.argsort(uniform(ncell, 1, 3)[:, 0], kind="stable")[:n]
    pos = np.stack([perm % dims[0] + 1, perm // dims[0] + 1], axis=1).astype(np.int32)
    group = (np.arange(n) % 2 + 1).astype(np.int32)
    m = ab.StandardABM(ab.GridSpaceSingle(dims, periodic=False), model_step=ab.SchellingRule(min_to_be_happy=3, seed=5))
    m.add_agents(pos, group=group, mood=np.zeros(n, dtype=np.uint8))
    df = m.run(5, adata=[("mood", "sum"), ("mood", "count", ("group", "==", 1)), (("pos", 0), "sum")], when=1)
    assert df["time"] == [0, 1, 2, 3, 4, 5]
    assert df["sum_mood"][0] == 0 and df["sum_mood"][-1] == int(m.column("mood").sum())
    assert df["sum_mood"][-1] > df["sum_mood"][1] * 0.9 and df["sum_mood"][-1] > 0.5 * n     # segregation proceeds
    assert all(c == (group == 1).sum() for c in df["count_mood_filtered"])
    assert df["sum_pos[1]"][-1] == int(m.pos[:, 0].astype(np.int64).sum())
    m.close()
