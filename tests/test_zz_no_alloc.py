"""`-m gpu` test of the allocation-free entry point `asb_vertex_energy_static` (include/asb.h; no counterpart in the reference).
Collected last on purpose (file name).

The new entry point is called straight through the C-ABI (ctypes on libasb.so; torch only owns the input / output buffers,
which are allocated BEFORE the calls) and compared with the existing `asb_vertex_energy` on the same inputs: both run the
same kernels with the same frame split, so the results must be bit-equal; two calls in a row on different inputs show the
shared static scratch carries nothing over.  The torch caching allocator's counters bracket the calls as a cheap check
that nothing was allocated through torch meanwhile."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _inputs(torch, F, N, ld, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    R = torch.zeros((F, 3, ld), dtype=torch.float64)
    R[:, :, :N] = torch.randn((F, 3, N), generator=g, dtype=torch.float64)
    return R.cuda()


def test_vertex_energy_static_matches_workspace_form_twice():
    import torch
    from animsnapbases_b200 import _lib

    lib = _lib.load()
    torch.cuda.set_device(0)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    limit = int(lib.asb_static_scratch_doubles())
    assert limit == 131072

    # (F, N): frame splits of 25 and 64 slabs (slabs * ld = 75 200 and 73 728 doubles of scratch) and one without scratch
    cases = [(50, 3000, 11), (200, 1100, 12), (8, 80000, 13)]
    bufs = []
    for F, N, seed in cases:
        ld = (N + 31) // 32 * 32
        R = _inputs(torch, F, N, ld, seed)
        ws = torch.empty(int(lib.asb_workspace_bytes(F, 3 * ld)) + 256, dtype=torch.uint8, device="cuda")
        e_ws = torch.full((ld,), -1.0, dtype=torch.float64, device="cuda")
        e_st = torch.full((ld,), -2.0, dtype=torch.float64, device="cuda")
        bufs.append((F, N, ld, R, ws, e_ws, e_st))
    torch.cuda.synchronize()

    allocs_before = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
    for rep in range(2):                                    # two calls in a row per case, alternating inputs between calls
        for F, N, ld, R, ws, e_ws, e_st in bufs:
            e_st.fill_(-2.0 - rep)
            _lib.check(lib.asb_vertex_energy(_ptr(R), _ptr(e_ws), F, N, ld, _ptr(ws), stream), "asb_vertex_energy")
            _lib.check(lib.asb_vertex_energy_static(_ptr(R), _ptr(e_st), F, N, ld, stream), "asb_vertex_energy_static")
    allocs_after = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
    torch.cuda.synchronize()
    assert allocs_after == allocs_before

    for F, N, ld, R, ws, e_ws, e_st in bufs:
        a, b = e_ws.cpu().numpy(), e_st.cpu().numpy()
        assert np.array_equal(a[:N], b[:N]), (F, N, float(np.abs(a[:N] - b[:N]).max()))   # same kernels, same order: bit-equal
        Rh = R.cpu().numpy()[:, :, :N]
        ref = ((Rh[:, 0] ** 2 + Rh[:, 1] ** 2) + Rh[:, 2] ** 2).sum(axis=0)
        assert np.allclose(b[:N], ref, rtol=1e-12, atol=0.0)                               # and both are the right numbers


def test_vertex_energy_static_back_to_back_inputs_do_not_leak():
    """Call 1 fills the scratch with large partials, call 2 (fewer rows, other values) must not see them."""
    import torch
    from animsnapbases_b200 import _lib

    lib = _lib.load()
    torch.cuda.set_device(0)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    F1, N1, F2, N2 = 64, 2000, 40, 1500
    ld1, ld2 = (N1 + 31) // 32 * 32, (N2 + 31) // 32 * 32
    R1 = _inputs(torch, F1, N1, ld1, 21) * 1e6
    R2 = _inputs(torch, F2, N2, ld2, 22)
    e1 = torch.empty((ld1,), dtype=torch.float64, device="cuda")
    e2 = torch.empty((ld2,), dtype=torch.float64, device="cuda")
    e2_ws = torch.empty((ld2,), dtype=torch.float64, device="cuda")
    ws = torch.empty(int(lib.asb_workspace_bytes(F2, 3 * ld2)) + 256, dtype=torch.uint8, device="cuda")
    _lib.check(lib.asb_vertex_energy_static(_ptr(R1), _ptr(e1), F1, N1, ld1, stream), "asb_vertex_energy_static")
    _lib.check(lib.asb_vertex_energy_static(_ptr(R2), _ptr(e2), F2, N2, ld2, stream), "asb_vertex_energy_static")
    _lib.check(lib.asb_vertex_energy(_ptr(R2), _ptr(e2_ws), F2, N2, ld2, _ptr(ws), stream), "asb_vertex_energy")
    torch.cuda.synchronize()
    assert np.array_equal(e2.cpu().numpy()[:N2], e2_ws.cpu().numpy()[:N2])
    assert float(e1[:N1].min()) > 1e9


def test_vertex_energy_static_rejects_sizes_beyond_the_scratch():
    import torch
    from animsnapbases_b200 import _lib

    lib = _lib.load()
    torch.cuda.set_device(0)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    F, N = 4, 70000                                         # 2 slabs x ld = 140 032 doubles > 131 072 on a 148-SM device
    ld = (N + 31) // 32 * 32
    R = torch.zeros((F, 3, ld), dtype=torch.float64, device="cuda")
    e = torch.zeros((ld,), dtype=torch.float64, device="cuda")
    with pytest.raises(_lib.AsbError, match="static scratch"):
        _lib.check(lib.asb_vertex_energy_static(_ptr(R), _ptr(e), F, N, ld, stream), "asb_vertex_energy_static")
