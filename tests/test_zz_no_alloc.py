"""ab_reduce_static: the data-collection reduction for hosts that cannot allocate device memory.

Some users of this library keep every byte of the GPU inside their own pool (CUDA.jl's, torch's caching
allocator), run many models at once and cannot take the device-wide stall of a cudaMalloc, or live inside a
fixed memory budget.  They cannot create a handle; they hand ab_reduce_static device pointers from their pool
and the library keeps its partial results in a fixed array that is part of the loaded library.  Here torch
plays that host: the column, the filter column and the two result words are torch tensors, the call goes through
the C ABI (ctypes), and the answer is compared with ab_reduce on a handle holding the same column — sums and
means to 1e-12 relative (the two walk the values in different orders: slot order there, array order here), minimum
/ maximum / count exactly, the tolerances of tests/test_collect_gpu.py.  Calls are made back to back, the
second shorter than the first, so partial results left in the shared array by one call would show in the next.

This file sorts last on purpose (pytest -x stops at the first failure; a failure here hides nothing else) and must
pass both alone in a fresh process and after every other test file."""
import ctypes as C

import numpy as np
import pytest

import agents_jl_b200 as ab
from agents_jl_b200 import _lib as L
from helpers import FLOCK, flock_init, uniform

pytestmark = pytest.mark.gpu

MAX_N = 1 << 26


def _static(col, op, flt=None, out=None, cnt=None, n=None):
    """One call through the C ABI: col / flt[0] / out / cnt are torch CUDA tensors of the caller."""
    import torch

    dt = {torch.float64: L.F64, torch.float32: L.F32, torch.int32: L.I32, torch.int64: L.I64, torch.uint8: L.U8}
    fcol, fdt, fcmp, fval = None, 0, 0, 0.0
    if flt is not None:
        fcol, fdt, fcmp, fval = flt[0].data_ptr(), dt[flt[0].dtype], L.CMP[flt[1]], float(flt[2])
    return ab.lib().ab_reduce_static(col.device.index, None, col.data_ptr(), dt[col.dtype],
                                     col.numel() if n is None else n, L.RED[op], fcol, fdt, fcmp, fval,
                                     out.data_ptr(), None if cnt is None else cnt.data_ptr())


def _close(got, want, exact):
    if isinstance(want, float) and np.isnan(want):
        return np.isnan(got)
    return got == want if exact else abs(got - want) <= 1e-12 * max(1.0, abs(want))


def test_static_reduction_equals_the_handle_reduction_twice_in_a_row():
    import torch

    extent = (400.0, 400.0)
    n = 300001                                    # 74 chunks, the last one ragged
    pos, vel = flock_init(n, extent, 11)
    energy = uniform(n, 1, 12)[:, 0] * 10.0 - 2.0
    kind = (np.arange(n) % 5).astype(np.int32)
    m = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m.add_agents(pos, vel, energy=energy, kind=kind)
    dev = torch.device("cuda", 0)
    # the caller's pool: everything the entry point touches is allocated by the caller, before the calls
    d_energy = torch.from_numpy(energy).to(dev)
    d_kind = torch.from_numpy(kind).to(dev)
    d_px = torch.from_numpy(np.ascontiguousarray(pos[:, 0])).to(dev)
    d_out = torch.zeros(8, dtype=torch.float64, device=dev)
    d_cnt = torch.zeros(8, dtype=torch.int64, device=dev)
    short = 5000                                  # second pass: 2 chunks — the 72 stale partials must not be read
    m2 = ab.StandardABM(ab.ContinuousSpace(extent, spacing=5.0), model_step=ab.FlockRule(**FLOCK))
    m2.add_agents(pos[:short], vel[:short], energy=energy[:short], kind=kind[:short])

    cases = [  # (model, column tensor, handle spec, static args, exact?)
        (m, d_energy, ("energy", "sum"), ("sum", None), False),
        (m2, d_energy[:short], ("energy", "sum"), ("sum", None), False),
        (m, d_energy, ("energy", "mean", ("kind", "==", 3)), ("mean", (d_kind, "==", 3)), False),
        (m2, d_energy[:short], ("energy", "maximum"), ("maximum", None), True),
        (m, d_energy, ("energy", "minimum", ("kind", "!=", 0)), ("minimum", (d_kind, "!=", 0)), True),
        (m2, d_kind[:short], ("kind", "sum"), ("sum", None), True),
        (m, d_kind, ("kind", "count", ("energy", ">", 5.0)), ("count", (d_energy, ">", 5.0)), True),
        (m, d_px, (("pos", 0), "mean", ("energy", "<", -100.0)), ("mean", (d_energy, "<", -100.0)), False),  # nobody passes
    ]
    free_before = None
    for rep in range(2):                          # the whole series twice: every call follows another one
        d_out.fill_(-7.0)
        d_cnt.fill_(-7)
        for k, (model, col, spec, (op, flt), exact) in enumerate(cases):
            if flt is not None and flt[0].numel() != col.numel():
                flt = (flt[0][:col.numel()], flt[1], flt[2])
            rc = _static(col, op, flt, d_out[k:], d_cnt[k:])
            assert rc == L.AB_OK, ab.lib().ab_last_error(None)
        torch.cuda.synchronize()
        got, cnt = d_out.cpu().numpy(), d_cnt.cpu().numpy()
        for k, (model, col, spec, (op, flt), exact) in enumerate(cases):
            want = model.collect([spec])[0]
            assert _close(got[k], want, exact), (rep, k, spec, got[k], want)
            h = col.cpu().numpy()
            if flt is None:
                passed = h.size
            else:
                f = flt[0][:col.numel()].cpu().numpy()
                passed = int({"==": f == flt[2], "!=": f != flt[2], ">": f > flt[2], "<": f < flt[2]}[flt[1]].sum())
            assert cnt[k] == passed, (rep, k, cnt[k], passed)
        # no device memory was taken by the calls of the second round (the first may load the kernels)
        free_now = torch.cuda.mem_get_info(dev)[0]
        if rep == 1:
            assert free_now == free_before
        free_before = free_now
    m.close()
    m2.close()


def test_static_reduction_limits_and_errors():
    import torch

    dev = torch.device("cuda", 0)
    col = torch.arange(10, dtype=torch.float64, device=dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)
    lib = ab.lib()
    # beyond the fixed scratch: refused before anything is read, with the limit in the message
    assert _static(col, "sum", None, out, cnt, n=MAX_N + 1) == L.ERR_INVALID
    msg = lib.ab_last_error(None).decode()
    assert "exceed" in msg and str(MAX_N) in msg
    assert _static(col, "sum", None, out, cnt, n=-1) == L.ERR_INVALID
    assert lib.ab_reduce_static(0, None, col.data_ptr(), 9, 10, 0, None, 0, 0, 0.0, out.data_ptr(), None) == L.ERR_INVALID
    assert lib.ab_reduce_static(0, None, col.data_ptr(), L.F64, 10, 17, None, 0, 0, 0.0, out.data_ptr(), None) == L.ERR_INVALID
    assert lib.ab_reduce_static(0, None, col.data_ptr(), L.F64, 10, 0, None, 0, 3, 0.0, out.data_ptr(), None) == L.ERR_INVALID
    assert lib.ab_reduce_static(0, None, col.data_ptr(), L.F64, 10, 0, None, 0, 0, 0.0, None, None) == L.ERR_INVALID
    # an empty column: identity results like ab_reduce (sum 0, mean NaN), count 0; NULL count pointer accepted
    assert _static(col, "sum", None, out, cnt, n=0) == L.AB_OK
    torch.cuda.synchronize()
    assert out.item() == 0.0 and cnt.item() == 0
    assert _static(col, "mean", None, out, None, n=0) == L.AB_OK
    torch.cuda.synchronize()
    assert np.isnan(out.item())
    # on a stream of the caller, and at the largest chunk-aligned tail: 10 values, then all 4096 of one chunk + 1
    big = torch.ones(4097, dtype=torch.float32, device=dev)
    st = torch.cuda.Stream(device=dev)
    st.wait_stream(torch.cuda.current_stream(dev))
    rc = lib.ab_reduce_static(0, C.c_void_p(st.cuda_stream), big.data_ptr(), L.F32, 4097, L.RED["sum"], None, 0, 0, 0.0,
                              out.data_ptr(), cnt.data_ptr())
    assert rc == L.AB_OK
    st.synchronize()
    assert out.item() == 4097.0 and cnt.item() == 4097
