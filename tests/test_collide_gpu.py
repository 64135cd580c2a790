"""interacting_pairs(model, r, :nearest) and elastic_collision! on the device against the
oracle's literal restatement of true_pairs! / elastic_collision! (src/spaces/continuous.jl:333-366, 454-497).
Pair lists: exact, including list order.  Velocities after the collisions: relative error <= 1e-14 (same
operation order in f64; only the compiler's choice of sqrt/div sequences may differ).  The scenario of the
reference's own "collisions" testset (test/continuous_space_tests.jl:382-460) is replayed with its assertions."""
import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import uniform
from oracle.pyoracle import ContinuousOracle, elastic_collision

pytestmark = pytest.mark.gpu


def _pair_models(extent, spacing, periodic, pos, vel):
    o = ContinuousOracle(extent, spacing=spacing, periodic=periodic)
    o.add_agents(pos, vel)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=spacing, periodic=periodic))
    m.add_agents(pos, vel)
    return o, m


def test_reference_golden_nearest_pairs():
    # test/continuous_space_tests.jl:244-265
    pos = np.array([(7.074386436066224, 4.963014649338054), (5.831962448496828, 4.926297135685473),
                    (5.122087781793935, 5.300031210394806), (3.9715633336430156, 4.8106570045816675)])
    o, m = _pair_models((10.0, 10.0), 0.1, False, pos, np.zeros((4, 2)))
    got = m.interacting_pairs(2.0, "nearest")
    assert len(got) == 1 and [tuple(p) for p in got] == o.interacting_pairs_nearest(2.0) == [(2, 3)]
    m.close()


@pytest.mark.parametrize("periodic", [True, False])
@pytest.mark.parametrize("r", [0.3, 0.8, 2.5])
def test_nearest_pairs_equal_oracle(periodic, r):
    n, extent = 700, (20.0, 20.0)
    pos = uniform(n, 2, 11) * 20.0
    o, m = _pair_models(extent, 1.0, periodic, pos, np.zeros((n, 2)))
    for order in (None, np.random.default_rng(3).permutation(n) + 1, np.arange(n, 0, -2)):
        want = o.interacting_pairs_nearest(r, order)
        got = [tuple(int(x) for x in p) for p in m.interacting_pairs(r, "nearest", scheduler=order)]
        assert got == want
        ids = [i for p in got for i in p]
        assert len(got) > 0
    m.close()


def _apply_oracle(pairs, pos, vel, mass):
    vel = vel.copy()
    hap = []
    for a, b in pairs:
        h, v1, v2 = elastic_collision(pos[a - 1], vel[a - 1], pos[b - 1], vel[b - 1],
                                      1.0 if mass is None else mass[a - 1], 1.0 if mass is None else mass[b - 1])
        vel[a - 1], vel[b - 1] = v1, v2
        hap.append(h)
    return np.array(hap), vel


@pytest.mark.parametrize("with_mass", [False, True])
def test_elastic_collisions_equal_oracle(with_mass):
    n, extent = 600, (20.0, 20.0)
    pos = uniform(n, 2, 5) * 20.0
    vel = uniform(n, 2, 6) * 2.0 - 1.0
    mass = None
    if with_mass:
        mass = 0.5 + 3.0 * uniform(n, 1, 7)[:, 0]
        mass[::7] = np.inf
        vel[::7] = 0.0
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=1.0))
    m.add_agents(pos, vel, **({"mass": mass} if with_mass else {}))
    pairs = m.interacting_pairs(1.2, "nearest")
    # a chain (shared agents, sequential semantics), a repeated pair and a coincident pair on top
    extra = np.array([[1, 2], [2, 3], [3, 1], [1, 2]])
    pairs = np.concatenate([pairs, extra])
    want_h, want_v = _apply_oracle([tuple(p) for p in pairs], pos, vel, mass)
    got_h = m.elastic_collisions(pairs, mass="mass" if with_mass else None)
    assert np.array_equal(got_h, want_h) and want_h.any() and not want_h.all()
    err = np.max(np.abs(m.vel - want_v) / np.maximum(1.0, np.abs(want_v)))
    assert err <= 1e-14, err
    assert np.array_equal(m.pos, pos)
    m.close()


def test_infinite_mass_must_rest():
    m = ab.StandardABM(ab.ContinuousSpace((10.0, 10.0), spacing=1.0))
    m.add_agents(np.array([[1.0, 1.0], [1.5, 1.0]]), np.array([[0.1, 0.0], [-0.1, 0.0]]), mass=np.array([np.inf, 1.0]))
    with pytest.raises(ab.AgentsError, match="mass cannot have nonzero velocity"):
        m.elastic_collisions([[1, 2]], mass="mass")
    m.close()


def test_reference_collision_scenario():
    # test/continuous_space_tests.jl:382-460: 10 x 10 disks, the left half immovable (mass Inf, at rest)
    speed, dt, diameter = 0.002, 1.0, 0.1
    ij = [(i, j) for i in range(1, 11) for j in range(1, 11)]
    pos = np.array([(i / 10, j / 10) for i, j in ij])
    ang = 2 * np.pi * uniform(100, 1, 42)[:, 0]
    vel = np.stack([np.sin(ang), np.cos(ang)], axis=1) * speed
    mass = np.full(100, 1.33)
    fixed = np.array([i <= 5 for i, _ in ij])
    vel[fixed] = 0.0
    mass[fixed] = np.inf
    m = ab.StandardABM(ab.ContinuousSpace((10.0, 10.0), periodic=True))
    m.add_agents(pos, vel, mass=mass)
    init = m.vel.copy()
    p0 = init.sum(axis=0)
    c = 0

    def step(k):
        nonlocal c
        for _ in range(k):
            m.move_agent(dt)                                   # agent_step! (agents first)
            pairs = m.interacting_pairs(diameter, "nearest")   # model_step!
            c += int(m.elastic_collisions(pairs, mass="mass").sum())

    step(10)
    npairs = len(m.interacting_pairs(diameter, "nearest"))
    assert npairs not in (0, 100)
    step(10)
    v = m.vel
    x = int(np.sum(np.any(init != v, axis=1)))
    assert np.array_equal(v[fixed], init[fixed])               # y == 0: immovable agents keep their velocity
    assert 0 < x <= 50
    assert c > 0
    assert not np.array_equal(v.sum(axis=0), p0)               # momentum is not conserved next to walls
    m.close()
