"""Host logic of the slab decomposition on CPU: layout, window selection, and the world_size-2 gather
over the gloo backend (no GPU, no CUDA library calls)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import agents_jl_b200  # noqa: F401,E402  (registers the package alias)
from agents_jl_b200 import parallel as P  # noqa: E402
from helpers import flock_init  # noqa: E402


@pytest.mark.parametrize("n_layers,nranks", [(400, 2), (400, 8), (81, 2), (280, 8), (83, 4)])
def test_slab_layout_partitions_the_layers(n_layers, nranks):
    lay = P.slab_layout(n_layers, nranks)
    assert lay[0][0] == 0 and sum(c for _, c in lay) == n_layers
    for (f0, c0), (f1, _) in zip(lay, lay[1:]):
        assert f1 == f0 + c0
    assert max(c for _, c in lay) - min(c for _, c in lay) <= 1


def test_window_layers_order_matches_the_device_window():
    own, up, lo = P.window_layers(first=40, count=41, n_layers=81)
    assert own == list(range(40, 81)) and up == [0, 1, 2] and lo == [37, 38, 39]
    own, up, lo = P.window_layers(first=0, count=40, n_layers=81)
    assert up == [40, 41, 42] and lo == [78, 79, 80]


def test_cell_layer_is_pos2cell():
    # max(1, ceil(p/spacing)) - 1 with the edge cases of continuous.jl:112-113 (p = 0 and p = extent)
    p = np.array([0.0, 1e-300, 5.0, 5.000000000000001, 399.99999, 400.0])
    assert P.cell_layer(p, 5.0, 80).tolist() == [0, 0, 0, 1, 79, 79]
    p32 = np.array([0.0, 5.0, 5.0000005, 400.0], dtype=np.float32)
    assert P.cell_layer(p32, 5.0, 80).tolist() == [0, 0, 1, 79]


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_select_window_owned_sets_partition_and_halo_is_complete(nranks):
    extent, spacing = (200.0, 400.0), 5.0
    pos, _ = flock_init(20000, extent, 3)
    n_layers = 80
    layer = P.cell_layer(pos[:, 1], spacing, n_layers)
    owned_total = np.zeros(len(pos), dtype=int)
    for first, count in P.slab_layout(n_layers, nranks):
        idx, owned = P.select_window(layer, first, count, n_layers)
        owned_total[idx[owned]] += 1
        own, up, lo = P.window_layers(first, count, n_layers)
        assert set(np.unique(layer[idx[owned]])) <= set(own)
        want = np.isin(layer, own + up + lo)
        assert np.array_equal(np.sort(idx), np.nonzero(want)[0])
    assert np.all(owned_total == 1)


@pytest.mark.parametrize("nranks", [2, 3, 5])
def test_select_window_without_wrap_on_a_non_periodic_axis(nranks):
    """Non-periodic last axis: the end slabs have ONE halo (no layers beyond the space), owned sets still partition."""
    extent, spacing = (200.0, 400.0), 5.0
    pos, _ = flock_init(20000, extent, 5)
    n_layers = 80
    layer = P.cell_layer(pos[:, 1], spacing, n_layers)
    owned_total = np.zeros(len(pos), dtype=int)
    for r, (first, count) in enumerate(P.slab_layout(n_layers, nranks)):
        idx, owned = P.select_window(layer, first, count, n_layers, periodic=False)
        owned_total[idx[owned]] += 1
        own, up, lo = P.window_layers(first, count, n_layers, periodic=False)
        assert (r == 0) == all(l is None for l in lo) and (r == nranks - 1) == all(u is None for u in up)
        layers = own + [l for l in up + lo if l is not None]
        assert np.array_equal(np.sort(idx), np.nonzero(np.isin(layer, layers))[0])
    assert np.all(owned_total == 1)


def test_merge_parts_detects_loss_and_duplication():
    ids = np.arange(1, 11)
    p = np.zeros((10, 2))
    ok = P.merge_parts([(ids[:6], p[:6], p[:6]), (ids[6:], p[6:], p[6:])], 10)
    assert ok[0].shape == (10, 2)
    with pytest.raises(RuntimeError):
        P.merge_parts([(ids[:6], p[:6], p[:6]), (ids[7:], p[7:], p[7:])], 10)
    with pytest.raises(RuntimeError):
        P.merge_parts([(ids[:6], p[:6], p[:6]), (ids[5:], p[5:], p[5:])], 10)


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        extent, spacing, n_layers = (200.0, 400.0), 5.0, 80
        pos, vel = flock_init(5000, extent, 9)
        first, count = P.slab_layout(n_layers, world)[rank]
        idx, owned = P.select_window(P.cell_layer(pos[:, 1], spacing, n_layers), first, count, n_layers)
        mine = idx[owned]
        parts = [None] * world
        dist.all_gather_object(parts, (mine + 1, pos[mine], vel[mine]))       # what Slabs.gather does
        gp, gv = P.merge_parts(parts, len(pos))
        uid = [b"x" * 128 if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)                                 # how the NCCL id travels
        q.put((rank, bool(np.array_equal(gp, pos) and np.array_equal(gv, vel)), len(uid[0])))
    finally:
        dist.destroy_process_group()


def test_gather_over_gloo_world_size_2():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29541 + (os.getpid() % 200)
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] and r[2] == 128 for r in res)


def test_chunked_generators_equal_the_whole_state():
    """bench.py builds the 100 M-boid / 80 M-particle states in id chunks: the chunks must be the same numbers as the
    one-shot generators the single-GPU rows and the parity tests use."""
    import bench

    extent = (100.0, 200.0)
    a = bench.particles(1000, extent)
    parts = list(bench.particle_chunks(1000, extent, 42, 300))
    assert [p[0] for p in parts] == [0, 300, 600, 900]
    assert np.array_equal(np.concatenate([p[1] for p in parts]), a[0])
    assert np.array_equal(np.concatenate([p[2] for p in parts]), a[1])
    for name, col in zip(("r", "k", "mass"), a[2:]):
        assert np.array_equal(np.concatenate([p[3][name] for p in parts]), col)
    pos, vel = flock_init(1000, (50.0, 60.0, 70.0), 42)
    fparts = list(bench.flock_chunks(1000, (50.0, 60.0, 70.0), 42, 256))
    assert np.array_equal(np.concatenate([p[1] for p in fparts]), pos)
    assert np.array_equal(np.concatenate([p[2] for p in fparts]), vel)
