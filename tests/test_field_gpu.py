"""Spatial fields: get_spatial_index / get_spatial_property (src/spaces/continuous.jl:536-573)
and the force-field form of the update_vel! hook (continuous.jl:86-91, :186-194) against the numpy restatement.
Indices: exact.  Velocities after the kick: bitwise (one multiply and one add per component, no contraction).
The reference holds no test for these functions (test/continuous_space_tests.jl:5): unpinned by the reference; the
last test holds the device to known answers worked out by hand from the source (tests/known_answers.py)."""
import numpy as np
import pytest

import agents_jl_b200 as ab
from helpers import flock_init, uniform
from oracle.pyoracle import spatial_index

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("extent,shape", [((100.0, 60.0), (10, 6)), ((100.0, 60.0), (7, 13)), ((100.0, 60.0), (25,)),
                                          ((30.0, 20.0, 10.0), (3, 4, 5)), ((30.0, 20.0, 10.0), (6, 2))])
def test_spatial_index_and_property(extent, shape, dtype):
    n = 5000
    pos, vel = flock_init(n, extent, 8, dtype=dtype)
    pos[:50] = np.floor(pos[:50])                                  # some agents exactly on property-cell edges
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0 if len(extent) == 2 else 2.5, dtype=dtype))
    m.add_agents(pos, vel)
    want = spatial_index(pos, extent, shape, dtype)
    assert np.array_equal(m.get_spatial_index(shape), want)
    prop = uniform(int(np.prod(shape)), 1, 3)[:, 0].reshape(shape, order="F")
    got = m.get_spatial_property(prop)
    assert np.array_equal(got, prop[tuple(want[:, d] - 1 for d in range(len(shape)))])
    q = uniform(333, len(extent), 9) * np.asarray(extent)[None, :]
    assert np.array_equal(m.get_spatial_index(shape, positions=q), spatial_index(q, extent, shape, dtype))
    assert np.array_equal(m.get_spatial_property(prop, positions=q),
                          prop[tuple(spatial_index(q, extent, shape, dtype)[:, d] - 1 for d in range(len(shape)))])
    m.close()


def test_index_outside_the_property_is_an_error():
    m = ab.StandardABM(ab.ContinuousSpace((10.0, 10.0), spacing=1.0, periodic=False))
    m.add_agents(np.array([[1.0, 1.0], [10.0, 3.0]]), np.zeros((2, 2)))      # pos == extent: BoundsError in the reference
    with pytest.raises(ab.AgentsError):
        m.get_spatial_index((5, 5))
    m.close()


def test_update_vel_field_then_move():
    extent, n, shape = (50.0, 40.0), 4000, (10, 8)
    pos, vel = flock_init(n, extent, 12)
    field = (uniform(80, 2, 4) - 0.5).reshape(shape + (2,), order="F")
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=2.5))
    m.add_agents(pos, vel)
    idx = spatial_index(pos, extent, shape) - 1
    ids = np.arange(1, n + 1, 3)
    want = vel.copy()
    want[ids - 1] = vel[ids - 1] + 0.25 * field[idx[ids - 1, 0], idx[ids - 1, 1]]
    m.update_vel_field(field, 0.25, ids=ids)
    assert np.array_equal(m.vel, want)
    m.update_vel_field(field, 0.25)                                 # all agents
    want = want + 0.25 * field[idx[:, 0], idx[:, 1]]
    assert np.array_equal(m.vel, want)
    m.move_agent(0.5)                                               # move_agent!(agent, model, dt) after update_vel!
    moved = np.mod(pos + 0.5 * want, np.asarray(extent)[None, :])
    assert np.max(np.abs(m.pos - moved)) <= 1e-12
    m.close()


def test_spatial_index_known_answers():
    """The device against indices worked out by hand from src/spaces/continuous.jl:566-573 (tests/known_answers.py)."""
    import known_answers as ka

    m = ab.StandardABM(ab.ContinuousSpace(ka.SI_EXTENT, spacing=1.0, periodic=False))
    m.add_agents(ka.SI_POS, np.zeros_like(ka.SI_POS))
    assert np.array_equal(m.get_spatial_index((5, 3)), ka.SI_IDX_5x3)
    assert np.array_equal(m.get_spatial_index((4,)), ka.SI_IDX_4)
    prop = np.arange(15, dtype=np.float64).reshape((5, 3), order="F")           # property[i, j] = (i - 1) + 5 (j - 1)
    want = (ka.SI_IDX_5x3[:, 0] - 1) + 5.0 * (ka.SI_IDX_5x3[:, 1] - 1)
    assert np.array_equal(m.get_spatial_property(prop), want)
    assert np.array_equal(m.get_spatial_index((5, 3), positions=ka.SI_POS[2:5]), ka.SI_IDX_5x3[2:5])
    with pytest.raises(ab.AgentsError):                                          # BoundsError in the reference
        m.get_spatial_index((5, 3), positions=ka.SI_POS_AT_EXTENT)
    m.close()
