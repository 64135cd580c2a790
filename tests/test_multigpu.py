"""Multi-GPU slab decomposition: N-GPU runs must be bitwise equal to the 1-GPU run (needs >= 2 GPUs;
run with `gpurun --gpus 2 -- python -m pytest tests/test_multigpu.py -m gpu`)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    try:
        import torch

        return torch.cuda.device_count()
    except Exception:
        return 0


def _run(world, port, *wargs, halo="direct", overlap="0", **extra_env):
    env = dict(os.environ, AB_HALO=halo, AB_OVERLAP=overlap, **extra_env)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "mgpu_worker.py"), *wargs]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    lines = [l for l in out.stdout.splitlines() if l.startswith("MGPU_RESULT ")]
    assert out.returncode == 0 and lines, out.stdout[-2000:] + out.stderr[-4000:]
    return json.loads(lines[-1][len("MGPU_RESULT "):])


CASES = [
    ("2d-approx", ["--agents", "60000", "--extent", "400", "400", "--steps", "12"]),
    ("2d-exact", ["--agents", "60000", "--extent", "400", "400", "--steps", "12", "--search", "exact"]),
    ("2d-uneven-layers", ["--agents", "50000", "--extent", "300", "405", "--steps", "9", "--chunk", "3"]),
    ("3d-approx", ["--agents", "40000", "--extent", "60", "60", "120", "--steps", "8"]),
    ("2d-f32", ["--agents", "60000", "--extent", "400", "400", "--steps", "6", "--dtype", "float32"]),
    ("2d-walls", ["--agents", "60000", "--extent", "400", "400", "--steps", "12", "--walls"]),
    # large enough for the halo install to cover every SM: with a grid of spinning CTAs beside the re-sort this size
    # dead-locked two ranks (the install now waits in ONE thread, k_halo_wait)
    ("2d-1M", ["--agents", "1000000", "--extent", "2000", "2000", "--steps", "6", "--chunk", "3"]),
    ("3d-walls-exact", ["--agents", "40000", "--extent", "60", "60", "120", "--steps", "8", "--walls", "--search", "exact"]),
]


@pytest.mark.gpu
@pytest.mark.parametrize("name,wargs", CASES, ids=[c[0] for c in CASES])
def test_two_gpus_bitwise_equal_to_one(name, wargs):
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    r = _run(2, 29531, *wargs)
    assert r["conserved"] and r["pos_bitwise_equal"] and r["vel_bitwise_equal"], r
    assert r["migrated"], "the test state should make agents cross the slab boundary"
    assert r["direct"], "the boundary layers should travel by direct peer-memory stores"


@pytest.mark.gpu
def test_two_gpus_nccl_exchange():
    """The ncclSend/ncclRecv exchange (AB_HALO=nccl) stays available and gives the same bits."""
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    r = _run(2, 29535, "--agents", "60000", "--extent", "400", "400", "--steps", "12", halo="nccl")
    assert r["conserved"] and r["pos_bitwise_equal"] and r["vel_bitwise_equal"] and not r["direct"], r


@pytest.mark.gpu
@pytest.mark.parametrize("wargs", [["--agents", "6000", "--extent", "400", "400", "--spacing", "20", "--steps", "10"],
                                   ["--agents", "8000", "--extent", "120", "120", "240", "--spacing", "20", "--steps", "8"]],
                         ids=["2d", "3d"])
def test_two_gpus_pairforce_bitwise_equal_to_one(wargs):
    """The pair-force rule on slabs: the r / k / mass columns travel in the boundary messages."""
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    r = _run(2, 29539, *wargs, "--rule", "pairforce")
    assert r["conserved"] and r["pos_bitwise_equal"] and r["vel_bitwise_equal"], r
    assert r["migrated"] and r["direct"], r


@pytest.mark.gpu
@pytest.mark.parametrize("wargs", [["--agents", "60000", "--extent", "400", "400", "--steps", "12"],
                                   ["--agents", "40000", "--extent", "60", "60", "120", "--steps", "8"]], ids=["2d", "3d"])
def test_two_gpus_separate_pack_and_split_cell_sums(wargs):
    """Large boundary messages (F3D) are packed by k_halo_pack instead of the re-sort, and the halo install then hides
    behind the cell sums of the interior rows (k_cell_M split in two launches): forced here on small cases."""
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    r = _run(2, 29541, *wargs, AB_FUSED_PACK="0")
    assert r["conserved"] and r["pos_bitwise_equal"] and r["vel_bitwise_equal"] and r["direct"], r


@pytest.mark.gpu
def test_four_gpus_bitwise_equal_to_one():
    if _ngpu() < 4:
        pytest.skip("needs >= 4 GPUs")
    r = _run(4, 29533, "--agents", "120000", "--extent", "400", "800", "--steps", "10")
    assert r["conserved"] and r["pos_bitwise_equal"] and r["vel_bitwise_equal"], r


@pytest.mark.gpu
@pytest.mark.parametrize("wargs", [["--agents", "60000", "--extent", "400", "400", "--steps", "12"],
                                   ["--agents", "40000", "--extent", "60", "60", "120", "--steps", "8"],
                                   ["--agents", "60000", "--extent", "400", "400", "--steps", "9", "--search", "exact", "--chunk", "3"]],
                         ids=["2d", "3d", "2d-exact"])
def test_overlapped_step_order_gives_the_same_bits(wargs):
    """AB_OVERLAP=1: interior layers swept while the halo travels, halo install + boundary sweep on a side stream."""
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    r = _run(2, 29537, *wargs, overlap="1")
    assert r["conserved"] and r["pos_bitwise_equal"] and r["vel_bitwise_equal"] and r["direct"], r
