"""ORACLE (test infrastructure, not product): ctypes binding of oracle/_build/liboracle.so.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this.  `build()` compiles the C++ restatement with g++ (oracle/Makefile).
"""
import ctypes as C
import os
import subprocess
from fractions import Fraction as Fr

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None

RUNNING, OPTIMAL, UNBOUNDED, INFEASIBLE = 0, 1, 2, 3


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("simplex.cpp", "rational.hpp", "bigint.hpp")]
    if (not force and os.path.exists(_LIB_PATH)
            and all(os.path.getmtime(_LIB_PATH) >= os.path.getmtime(s) for s in srcs)):
        return _LIB_PATH
    subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 8 + [C.c_int, C.c_void_p, C.c_void_p]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_set_cleanup_any.argtypes = [C.c_void_p, C.c_int]
        for f in ("orc_set_cost", "orc_set_rhs"):
            getattr(L, f).argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        L.orc_set_value.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p]
        L.orc_begin.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_from_basis.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_advance.argtypes = [C.c_void_p, C.c_long]
        L.orc_advance.restype = C.c_int
        L.orc_advance_to_handover.argtypes = [C.c_void_p]
        L.orc_advance_to_handover.restype = C.c_int
        for f in ("orc_status", "orc_stage", "orc_m", "orc_nr_artificial", "orc_nr_rows_removed"):
            getattr(L, f).argtypes = [C.c_void_p]
            getattr(L, f).restype = C.c_int
        for f in ("orc_nr_pivots", "orc_max_bits", "orc_nnz_inverse"):
            getattr(L, f).argtypes = [C.c_void_p]
            getattr(L, f).restype = C.c_long
        L.orc_loop_seconds.argtypes = [C.c_void_p]
        L.orc_loop_seconds.restype = C.c_double
        L.orc_log_bulk.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_log_rc.argtypes = [C.c_void_p, C.c_long]
        L.orc_log_rc.restype = C.c_char_p
        for f in ("orc_objective", "orc_dump_carry"):
            getattr(L, f).argtypes = [C.c_void_p]
            getattr(L, f).restype = C.c_char_p
        for f in ("orc_b", "orc_minus_pi"):
            getattr(L, f).argtypes = [C.c_void_p, C.c_int]
            getattr(L, f).restype = C.c_char_p
        L.orc_basis_indices.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_rat_op.argtypes = [C.c_char_p, C.c_char_p, C.c_char]
        L.orc_rat_op.restype = C.c_char_p
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class COracle:
    """Resumable CPU solve of one provider (ProviderArrays-like object with the
    `relp_load_matrix` arrays)."""

    def __init__(self, pa):
        L = lib()
        self.pa = pa
        self.h = L.orc_create(pa.m, pa.n, _p(pa.col_ptr), _p(pa.row_idx), _p(pa.val_num), _p(pa.val_den),
                              _p(pa.cost_num), _p(pa.cost_den), _p(pa.rhs_num), _p(pa.rhs_den),
                              len(pa.piv_rows), _p(pa.piv_rows), _p(pa.piv_cols))
        for i, f in (getattr(pa, "rhs_exact", None) or {}).items():       # right-hand sides beyond int64/int64
            L.orc_set_rhs(self.h, int(i), f"{f.numerator}/{f.denominator}".encode())

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_destroy(self.h)
            self.h = None

    def set_cleanup_any(self, on=True):
        """clean-up pivots on any nonzero element of the artificial's row (default: the reference's zero-cost columns only)"""
        lib().orc_set_cleanup_any(self.h, 1 if on else 0)

    def begin(self, rule1=1, rule2=2):
        lib().orc_begin(self.h, rule1, rule2)

    def from_basis(self, basis, rule2=2):
        """carry/mod.rs:428-463: phase-two Carry of a given basis (basis[i] = column in row i)."""
        lib().orc_begin(self.h, 1, rule2)
        b = np.ascontiguousarray(basis, dtype=np.int32)
        return lib().orc_from_basis(self.h, b.ctypes.data_as(C.c_void_p))

    def advance(self, max_new_pivots=-1):
        return lib().orc_advance(self.h, max_new_pivots)

    def advance_to_handover(self):
        """Finish phase 1 incl. artificial clean-up and the Carry hand-over; no phase-2 pivot."""
        return lib().orc_advance_to_handover(self.h)

    def solve(self, rule1=1, rule2=2):
        self.begin(rule1, rule2)
        return self.advance(-1)

    @property
    def nr_pivots(self):
        return lib().orc_nr_pivots(self.h)

    @property
    def nr_artificial(self):
        return lib().orc_nr_artificial(self.h)

    @property
    def m(self):
        return lib().orc_m(self.h)

    @property
    def stage(self):
        return lib().orc_stage(self.h)

    @property
    def loop_seconds(self):
        return lib().orc_loop_seconds(self.h)

    def max_bits(self):
        return lib().orc_max_bits(self.h)

    def nnz_inverse(self):
        return lib().orc_nnz_inverse(self.h)

    def log(self):
        """int32 array [n_pivots, 6]: phase tag (ord), j, r, leaving, T, S."""
        n = self.nr_pivots
        out = np.zeros((n, 6), dtype=np.int32)
        if n:
            lib().orc_log_bulk(self.h, _p(out))
        return out

    def log_rc(self, k):
        return Fr(lib().orc_log_rc(self.h, k).decode())

    def objective(self):
        return Fr(lib().orc_objective(self.h).decode())

    def basis_indices(self):
        out = np.zeros(self.m, dtype=np.int32)
        lib().orc_basis_indices(self.h, _p(out))
        return out

    def carry(self):
        """(minus_obj, minus_pi[], b[], basis[], rows[] as dict col->Fraction)."""
        lines = lib().orc_dump_carry(self.h).decode().split("\n")
        mo = Fr(lines[0])
        mpi = [Fr(x) for x in lines[1].split()]
        b = [Fr(x) for x in lines[2].split()]
        basis = [int(x) for x in lines[3].split()]
        rows = []
        for i in range(len(b)):
            d = {}
            for tok in lines[4 + i].split():
                k, v = tok.split(":")
                d[int(k)] = Fr(v)
            rows.append(d)
        return mo, mpi, b, basis, rows

    def current_bfs(self):
        _, _, b, basis, _ = self.carry()
        t = [(basis[i], v) for i, v in enumerate(b) if v != 0]
        t.sort(key=lambda x: x[0])
        return t


def rat_op(a, b, op):
    return lib().orc_rat_op(str(a).encode(), str(b).encode(), op.encode()).decode()
