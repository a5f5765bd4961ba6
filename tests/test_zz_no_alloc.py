"""relp_ratio_test_static: the allocation-free, context-free K4 ratio test (include/relp_b200.h; no counterpart in
the reference).  Its output is compared with the existing entry
point's (relp_ratio_test on a context) on the same generated column, on consecutive calls that share the static
scratch.  Collected last on purpose (zz)."""
import ctypes as C

import numpy as np
import pytest

from helpers import feasible_lp, pa_from_lp
from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp

MAX_ROWS = 1 << 20


def test_static_rejects_sizes_beyond_the_maximum():
    """refused before any CUDA call, so this runs without a GPU too"""
    import __graft_entry__ as entry
    entry.build_library()
    L = capi.load_library()
    dummy = np.zeros(4, np.int64)
    p = dummy.ctypes.data_as(C.c_void_p)
    st = L.relp_ratio_test_static(p, p, p, p, p, MAX_ROWS + 1, p, None)
    assert st == capi.ERR_ARGUMENT
    msg = L.relp_last_error().decode()
    assert "RELP_STATIC_MAX_ROWS" in msg and str(MAX_ROWS) in msg
    assert L.relp_ratio_test_static(None, p, p, p, p, 4, p, None) == capi.ERR_ARGUMENT


def _static_ratio(L, torch, b, d, basis):
    """(row, status) from the static entry point for host arrays b = (num, den), d = (num, den), basis"""
    dev = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (b[0], b[1], d[0], d[1])]
    bas = torch.from_numpy(np.ascontiguousarray(basis, dtype=np.int32)).cuda()
    res = torch.full((2,), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    st = L.relp_ratio_test_static(*[C.c_void_p(t.data_ptr()) for t in dev], C.c_void_p(bas.data_ptr()), len(basis),
                                  C.c_void_p(res.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert st == capi.OK, L.relp_last_error().decode()
    torch.cuda.synchronize()
    r, status = res.cpu().tolist()
    return r, status


def _compare_over_pivots(pa, rule, pivots, heap_words=1 << 22, phase2=False):
    import torch
    torch.cuda.set_device(0)
    c = capi.Carry(pa, heap_words=heap_words)
    L = c.L
    if phase2:          # an LP without artificial columns: phase 1 ends at once, compare on phase-2 pivots
        assert c.solve_to_handover() == capi.OK
    else:
        c.create_for_partially_artificial()
    seen = 0
    for _ in range(pivots):
        sel = c.select_primal_pivot_column(rule)
        if sel is None:
            break
        c.L.relp_generate_column(c.h, sel[0])
        expected = c.select_primal_pivot_row()              # the existing entry point: relp_ratio_test
        dn, dd, _ = c._read_raw(L.relp_read_column, c.mloc)
        bn, bd, _ = c._read_raw(L.relp_read_b, c.mloc)
        basis = c.basis_indices()
        assert (dd > 0).all() and (bd > 0).all(), "test LP left the int64 tier"
        for rep in range(2):                                # two calls in a row: the shared scratch is reused
            r, status = _static_ratio(L, torch, (bn.copy(), bd.copy()), (dn.copy(), dd.copy()), basis)
            if expected is None:
                assert (r, status) == (-1, capi.UNBOUNDED), (r, status, rep)
            else:
                assert (r, status) == (expected, capi.OK), (r, status, expected, rep)
        seen += 1
        if expected is None:
            break
        c.change_basis(expected)
    c.close()
    return seen


@pytest.mark.gpu
def test_static_matches_context_ratio_test_rationals():
    """fractional data (ties broken by basis column, reduced and unreduced quotients), one block"""
    lp = feasible_lp(7, 6, 2, 8, 4, 30, frac=True, density=0.5)
    assert _compare_over_pivots(pa_from_lp(lp), capi.FIRST_PROFITABLE_WITH_MEMORY, 12) >= 4


@pytest.mark.gpu
def test_static_matches_context_ratio_test_many_blocks():
    """m = 5000 rows = 5 blocks: the last-arriving block reduces the per-block winners held in the static scratch"""
    pa = synthetic_lp(5000, 20000, 20, "tu")
    assert _compare_over_pivots(pa, capi.STEEPEST_DESCENT, 6, heap_words=1 << 24, phase2=True) >= 4


@pytest.mark.gpu
def test_static_edge_cases():
    import torch
    torch.cuda.set_device(0)
    L = capi.load_library()
    one = np.ones(3000, np.int64)
    # no row with d_i > 0: unbounded; then (same scratch, next call) a plain argmin with a tie on the ratio
    d = -one.copy()
    assert _static_ratio(L, torch, (one, one), (d, one), np.arange(3000)) == (-1, capi.UNBOUNDED)
    d = one.copy(); b = 5 * one; b[2500] = 3; b[17] = 3
    basis = np.arange(3000)[::-1].copy()                    # row 2500 has the smaller basis column
    assert _static_ratio(L, torch, (b, one), (d, one), basis) == (2500, capi.OK)
    # a multi-limb handle (den == 0) is reported, not misread
    dd = one.copy(); dd[1234] = 0
    assert _static_ratio(L, torch, (b, one), (d, dd), basis) == (-1, capi.ERR_LIMB_CAPACITY)
    # and the flag does not leak into the next call
    assert _static_ratio(L, torch, (b, one), (d, one), basis) == (2500, capi.OK)
    # m = 0
    z = np.zeros(1, np.int64)
    dev = torch.zeros(2, dtype=torch.int64, device="cuda"); res = torch.zeros(2, dtype=torch.int32, device="cuda")
    p = C.c_void_p(dev.data_ptr())
    assert L.relp_ratio_test_static(p, p, p, p, p, 0, C.c_void_p(res.data_ptr()), None) == capi.OK
    torch.cuda.synchronize()
    assert res.cpu().tolist() == [-1, capi.UNBOUNDED]
    del z
