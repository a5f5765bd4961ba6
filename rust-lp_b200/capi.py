"""ctypes binding of librelp_b200.so (include/relp_b200.h) and a thin host mirror of the reference's
interface for the hot path: `Carry` (InverseMaintener, inverse_maintenance/mod.rs:29-267),
`PivotRule` constants (strategy/pivot_rule.rs), `solve_relaxation` (two_phase/mod.rs:30-75).

The library is CUDA-only.  Importing this module never falls back to a CPU implementation:
`Carry(...)` raises `RelpError` when the shared library is missing or no GPU is present.
"""
import ctypes as C
import os
from fractions import Fraction as Fr

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librelp_b200.so")

OK, OPTIMAL, UNBOUNDED, INFEASIBLE = 0, 1, 2, 3
ERR_CUDA, ERR_ZERO_PIVOT, ERR_LIMB_CAPACITY, ERR_HEAP_EXHAUSTED, ERR_ARGUMENT, ERR_PHASE1_UNBOUNDED, ERR_COMM, \
    ERR_ARTIFICIAL_BASIC = -1, -2, -3, -4, -5, -6, -7, -8
FIRST_PROFITABLE, FIRST_PROFITABLE_WITH_MEMORY, STEEPEST_DESCENT = 0, 1, 2

# every symbol include/relp_b200.h declares (tests check the library exports all of them)
SYMBOLS = [
    "relp_create", "relp_destroy", "relp_last_error", "relp_abi_version", "relp_set_option", "relp_stream", "relp_debug_trace", "relp_debug_k1_dense", "relp_debug_k1_verify", "relp_debug_rat_op", "relp_comm_unique_id",
    "relp_attach_comm", "relp_attach_comm_callbacks", "relp_peer_export", "relp_peer_attach", "relp_load_matrix", "relp_set_costs", "relp_set_rhs", "relp_carry_create_partially_artificial",
    "relp_carry_from_artificial", "relp_carry_from_artificial_remove_rows", "relp_carry_from_basis", "relp_price",
    "relp_generate_column", "relp_ratio_test", "relp_change_basis", "relp_generate_element",
    "relp_iterate", "relp_solve_relaxation", "relp_solve_continue", "relp_solve_to_handover",
    "relp_dims", "relp_read_basis", "relp_read_block", "relp_read_carry", "relp_read_b",
    "relp_read_minus_pi", "relp_read_minus_objective", "relp_read_column", "relp_log_size",
    "relp_read_log", "relp_stats_get", "relp_rows_removed", "relp_ratio_test_static",
]


class RelpError(RuntimeError):
    def __init__(self, status, msg=""):
        super().__init__(f"relp status {status}: {msg}")
        self.status = status


class IterLog(C.Structure):
    _fields_ = [("tag", C.c_int32), ("j", C.c_int32), ("r", C.c_int32), ("leaving", C.c_int32),
                ("T", C.c_int32), ("S", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("pivots", C.c_uint64), ("escalated", C.c_uint64), ("sum_T", C.c_uint64),
                ("sum_S", C.c_uint64), ("sum_TS", C.c_uint64), ("heap_peak_words", C.c_uint64),
                ("compactions", C.c_uint64), ("kernel_launches", C.c_uint64), ("max_limbs", C.c_int32),
                ("m", C.c_int32), ("n", C.c_int32), ("nr_artificial", C.c_int32), ("phase", C.c_int32),
                ("loop_ms", C.c_double), ("stage_ns", C.c_uint64 * 8), ("narrow_pivots", C.c_uint64), ("wide_handoffs", C.c_uint64)]


ALLGATHER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t)
ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_int64), C.c_size_t)

_libs = {}


def load_library(path=None):
    """dlopen the C-ABI library and declare its prototypes.  No fallback of any kind."""
    path = path or os.environ.get("RELP_B200_LIB") or LIB_PATH      # the env override is for A/B builds of the same library
    if path in _libs:
        return _libs[path]
    if not os.path.exists(path):
        raise RelpError(ERR_CUDA, f"{path} not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
    L = C.CDLL(path)
    vp, i32, i64, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t
    L.relp_create.restype = vp
    L.relp_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint64]
    L.relp_destroy.argtypes = [vp]
    L.relp_destroy.restype = None
    L.relp_last_error.restype = C.c_char_p
    L.relp_set_option.argtypes = [vp, C.c_char_p, i64]
    L.relp_stream.argtypes = [vp]
    L.relp_stream.restype = vp
    L.relp_debug_trace.argtypes = [vp, vp, i32]
    if hasattr(L, "relp_debug_k1_dense"):      # absent from A/B builds of older sources only
        L.relp_debug_k1_dense.argtypes = [vp, C.c_uint64, i32, vp, vp, vp, vp]
    if hasattr(L, "relp_debug_k1_verify"):
        L.relp_debug_k1_verify.argtypes = [vp, C.c_uint64, vp, vp]
    if hasattr(L, "relp_debug_rat_op"):
        L.relp_debug_rat_op.argtypes = [vp, C.c_int, i32, vp, vp, sz, C.c_int, vp, vp, vp, sz, vp]
    L.relp_comm_unique_id.argtypes = [vp, C.c_char_p]
    L.relp_attach_comm.argtypes = [vp, vp, C.c_char_p]
    L.relp_attach_comm_callbacks.argtypes = [vp, ALLGATHER_FN, ALLREDUCE_FN, vp]
    L.relp_peer_export.argtypes = [vp, i32, vp]
    L.relp_peer_attach.argtypes = [vp, vp]
    L.relp_load_matrix.argtypes = [vp, i32, i32] + [vp] * 8
    L.relp_set_costs.argtypes = [vp, i32, vp, vp]
    L.relp_set_rhs.argtypes = [vp, i32, vp, vp, vp, C.c_size_t]
    L.relp_carry_create_partially_artificial.argtypes = [vp, i32, vp, vp]
    L.relp_carry_from_artificial.argtypes = [vp]
    L.relp_carry_from_artificial_remove_rows.argtypes = [vp, i32, vp]
    L.relp_carry_from_basis.argtypes = [vp, vp]
    L.relp_price.argtypes = [vp, C.c_int, vp, vp, vp, vp, sz]
    L.relp_generate_column.argtypes = [vp, i32]
    L.relp_ratio_test.argtypes = [vp, vp]
    L.relp_change_basis.argtypes = [vp, i32, vp]
    L.relp_generate_element.argtypes = [vp, i32, i32, vp, vp, vp, sz]
    L.relp_iterate.argtypes = [vp, C.c_int, i64]
    L.relp_solve_relaxation.argtypes = [vp, C.c_int, C.c_int, i32, vp, vp, i64]
    L.relp_solve_continue.argtypes = [vp, i64]
    L.relp_solve_to_handover.argtypes = [vp, C.c_int, i32, vp, vp]
    L.relp_dims.argtypes = [vp, vp, vp, vp, vp]
    L.relp_read_basis.argtypes = [vp, vp]
    L.relp_read_block.argtypes = [vp, i32, i32, i32, i32, vp, vp, vp, sz, vp]
    for f in ("relp_read_carry", "relp_read_b", "relp_read_minus_pi", "relp_read_minus_objective", "relp_read_column"):
        getattr(L, f).argtypes = [vp, vp, vp, vp, sz, vp]
    L.relp_log_size.argtypes = [vp]
    L.relp_log_size.restype = i64
    L.relp_read_log.argtypes = [vp, i64, i64, vp]
    L.relp_stats_get.argtypes = [vp, vp]
    L.relp_rows_removed.argtypes = [vp, vp, i32]
    if hasattr(L, "relp_ratio_test_static"):
        L.relp_ratio_test_static.argtypes = [vp, vp, vp, vp, vp, i32, vp, vp]
    _libs[path] = L
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


_from_coprime = getattr(Fr, "_from_coprime_ints", None)     # Python >= 3.12: no gcd when the pair is known reduced


def decode_rats(num, den, limbs):
    """Rat array (see relp_b200.h) -> list of Fraction.  Asserts canonical form (lowest terms,
    positive denominator): the gcds of the int64 tier are checked in one numpy call."""
    num = np.asarray(num, np.int64); den = np.asarray(den, np.int64)
    small = den > 0
    if small.any():
        g = np.gcd(np.abs(num[small]), den[small])
        assert bool(np.all((g == 1) | ((num[small] == 0) & (den[small] == 1)))), "device returned a non-canonical rational"
    out = []
    mk = _from_coprime if _from_coprime is not None else Fr
    for n, d in zip(num.tolist(), den.tolist()):
        if d > 0:
            out.append(mk(n, d))
            continue
        assert d == 0, "absent value crossed the boundary"
        h = int(limbs[n]); nn = h & 0x3fff; nd = (h >> 14) & 0x3fff
        a = 0
        for i in range(nn):
            a |= int(limbs[n + 1 + i]) << (32 * i)
        b = 0
        for i in range(nd):
            b |= int(limbs[n + 1 + nn + i]) << (32 * i)
        f = Fr(a, b)
        assert f.numerator == a and f.denominator == b, "device returned a non-canonical rational"
        out.append(-f if h >> 31 else f)
    return out


class Carry:
    """Device-resident `Carry<RationalBig, BasisInverseRows>` (carry/mod.rs:45-65) + the tableau
    bookkeeping the fused loop needs.  Method names follow the reference's trait methods."""

    def __init__(self, provider, device=0, rank=0, world=1, heap_words=0, lib_path=None):
        self.L = load_library(lib_path)
        self.pa = provider
        self.h = self.L.relp_create(device, rank, world, heap_words)
        if not self.h:
            raise RelpError(ERR_CUDA, self.L.relp_last_error().decode())
        self.rank, self.world = rank, world
        self._limb_cap = 1 << 16
        self.load_matrix(provider)

    def set_option(self, name, value):
        self._chk(self.L.relp_set_option(self.h, name.encode(), int(value)))

    def stream(self):
        return self.L.relp_stream(self.h)

    # ---- multi-GPU plumbing -------------------------------------------------------------------
    def comm_unique_id(self, nccl_lib=None):
        buf = (C.c_char * 128)()
        self._chk(self.L.relp_comm_unique_id(buf, nccl_lib.encode() if nccl_lib else None))
        return bytes(buf)

    def attach_comm(self, unique_id, nccl_lib=None):
        buf = (C.c_char * 128).from_buffer_copy(unique_id)
        self._chk(self.L.relp_attach_comm(self.h, buf, nccl_lib.encode() if nccl_lib else None))

    def attach_comm_callbacks(self, allgather, allreduce):
        self._cb = (ALLGATHER_FN(allgather), ALLREDUCE_FN(allreduce))     # keep alive
        self._chk(self.L.relp_attach_comm_callbacks(self.h, self._cb[0], self._cb[1], None))

    def peer_export(self, max_m):
        h = (C.c_ubyte * 64)()
        self._chk(self.L.relp_peer_export(self.h, max_m, h))
        return bytes(h)

    def peer_attach(self, handles):
        buf = b"".join(handles)
        self._chk(self.L.relp_peer_attach(self.h, buf))

    def debug_trace(self, n_ctas=148):
        out = np.zeros((n_ctas, 16), np.uint64)
        self._chk(self.L.relp_debug_trace(self.h, _p(out), n_ctas))
        return out

    def debug_k1_dense(self, seed=1, reps=4):
        """K1 micro-benchmark (dense rank-1 update of the whole Carry block); the Carry is invalid afterwards.
        Returns (ms pivot row, ms update, T, S)."""
        a, b = C.c_float(), C.c_float()
        t, s_ = C.c_int64(), C.c_int64()
        self._chk(self.L.relp_debug_k1_dense(self.h, seed, reps, C.byref(a), C.byref(b), C.byref(t), C.byref(s_)))
        return a.value, b.value, t.value, s_.value

    def debug_rat_op(self, op, xs, as_, bs, s=1):
        """Device arithmetic test hook: op 0: x + s a b, 1: x / a, 2: sign(x - a), 3: x + s a b by whole warps; returns the
        list of results (Fractions; ints for op 2), each computed by one GPU thread through val_fma / val_div / val_cmp."""
        count = len(xs)
        ops = np.zeros((6, count), np.int64)
        limbs = []

        def enc(row, k, f):
            f = Fr(f)
            n, d = f.numerator, f.denominator
            if abs(n) <= 2 ** 63 - 1 and d <= 2 ** 63 - 1:
                ops[row, k], ops[row + 1, k] = n, d
                return
            a = abs(n); nl = []; dl = []
            while a:
                nl.append(a & 0xffffffff); a >>= 32
            while d:
                dl.append(d & 0xffffffff); d >>= 32
            ops[row, k], ops[row + 1, k] = len(limbs), 0
            limbs.extend([len(nl) | len(dl) << 14 | (0x80000000 if n < 0 else 0)] + nl + dl)
        for k in range(count):
            enc(0, k, xs[k]); enc(2, k, as_[k]); enc(4, k, bs[k] if bs is not None else 0)
        lin = np.array(limbs if limbs else [0], np.uint32)
        n = np.zeros(count, np.int64); d = np.zeros(count, np.int64)
        cap = max(1 << 16, 8 * len(limbs) + 1024)
        while True:
            lout = np.zeros(cap, np.uint32); used = C.c_size_t(0)
            st = self.L.relp_debug_rat_op(self.h, op, count, _p(ops), _p(lin), len(limbs), s, _p(n), _p(d), _p(lout), cap, C.byref(used))
            if st == ERR_ARGUMENT and used.value > cap:
                cap = int(used.value) * 2
                continue
            self._chk(st)
            break
        if op == 2:
            return [int(v) for v in n]
        return decode_rats(n, d, lout)

    def debug_k1_verify(self, seed=1):
        """Dense K1 self-check (staged path vs element-wise path on the same fabricated block): (mismatches, elements)."""
        bad, n = C.c_int64(-1), C.c_int64(0)
        self._chk(self.L.relp_debug_k1_verify(self.h, seed, C.byref(bad), C.byref(n)))
        return bad.value, n.value

    def close(self):
        if getattr(self, "h", None):
            self.L.relp_destroy(self.h)
            self.h = None

    __del__ = close

    def _chk(self, st):
        if st < 0:
            raise RelpError(st, self.L.relp_last_error().decode())
        return st

    # ---- MatrixProvider side -----------------------------------------------------------------
    def load_matrix(self, pa):
        self.pa = pa
        self._chk(self.L.relp_load_matrix(self.h, pa.m, pa.n, _p(pa.col_ptr), _p(pa.row_idx), _p(pa.val_num),
                                          _p(pa.val_den), _p(pa.cost_num), _p(pa.cost_den), _p(pa.rhs_num),
                                          _p(pa.rhs_den)))
        if getattr(pa, "rhs_exact", None):
            self.set_rhs([pa.rhs_exact.get(i, Fr(int(pa.rhs_num[i]), int(pa.rhs_den[i]))) for i in range(pa.m)])

    def set_rhs(self, rhs):
        """replace the right-hand side by exact Fractions of any size (relp_set_rhs; the reference's Rhs is a RationalBig)"""
        n = np.zeros(len(rhs), np.int64); d = np.ones(len(rhs), np.int64); limbs = []
        for i, f in enumerate(rhs):
            f = Fr(f)
            if f.numerator <= 2 ** 63 - 1 and f.denominator <= 2 ** 63 - 1:
                n[i], d[i] = f.numerator, f.denominator
                continue
            nl, dl, a, b = [], [], f.numerator, f.denominator
            while a:
                nl.append(a & 0xffffffff); a >>= 32
            while b:
                dl.append(b & 0xffffffff); b >>= 32
            n[i], d[i] = len(limbs), 0
            limbs.extend([len(nl) | len(dl) << 14] + nl + dl)
        lin = np.array(limbs if limbs else [0], np.uint32)
        self._chk(self.L.relp_set_rhs(self.h, len(rhs), _p(n), _p(d), _p(lin), len(limbs)))

    def set_costs(self, cost_num, cost_den):
        """replace cost_value(j) of the loaded provider (phase two reads them; refused while a phase-two tableau is live)"""
        cn = np.ascontiguousarray(cost_num, dtype=np.int64); cd = np.ascontiguousarray(cost_den, dtype=np.int64)
        self._chk(self.L.relp_set_costs(self.h, len(cn), _p(cn), _p(cd)))

    # ---- InverseMaintener constructors ---------------------------------------------------------
    def create_for_partially_artificial(self, piv_rows=None, piv_cols=None):
        pr = self.pa.piv_rows if piv_rows is None else np.ascontiguousarray(piv_rows, dtype=np.int32)
        pc = self.pa.piv_cols if piv_cols is None else np.ascontiguousarray(piv_cols, dtype=np.int32)
        self._chk(self.L.relp_carry_create_partially_artificial(self.h, len(pr), _p(pr), _p(pc)))

    def from_artificial(self):
        self._chk(self.L.relp_carry_from_artificial(self.h))

    def from_basis(self, basis):
        """`IM::from_basis(&basis, provider)` (inverse_maintenance/mod.rs:91): basis[i] = column in row i."""
        b = np.ascontiguousarray(basis, dtype=np.int32)
        assert b.size == self.pa.m
        self._chk(self.L.relp_carry_from_basis(self.h, _p(b)))

    def from_artificial_remove_rows(self, rows):
        r = np.ascontiguousarray(rows, dtype=np.int32)
        self._chk(self.L.relp_carry_from_artificial_remove_rows(self.h, len(r), _p(r)))

    # ---- per-iteration trait methods -------------------------------------------------------------
    def select_primal_pivot_column(self, rule):
        """PivotRule::select_primal_pivot_column -> (j, rc) or None."""
        j = C.c_int32(-1)
        n = np.zeros(1, np.int64); d = np.zeros(1, np.int64); limbs = np.zeros(self._limb_cap, np.uint32)
        st = self._chk(self.L.relp_price(self.h, rule, C.byref(j), _p(n), _p(d), _p(limbs), limbs.size))
        if st == OPTIMAL:
            return None
        return j.value, decode_rats(n, d, limbs)[0]

    def generate_column(self, j):
        self._chk(self.L.relp_generate_column(self.h, j))
        vals = self._read(self.L.relp_read_column, self.mloc)
        return [(self.row_lo + i, v) for i, v in enumerate(vals) if v != 0]

    def select_primal_pivot_row(self):
        r = C.c_int32(-1)
        st = self._chk(self.L.relp_ratio_test(self.h, C.byref(r)))
        return None if st == UNBOUNDED else r.value

    def change_basis(self, r):
        leaving = C.c_int32(-1)
        self._chk(self.L.relp_change_basis(self.h, r, C.byref(leaving)))
        return leaving.value

    def generate_element(self, i, j):
        n = np.zeros(1, np.int64); d = np.zeros(1, np.int64); limbs = np.zeros(self._limb_cap, np.uint32)
        self._chk(self.L.relp_generate_element(self.h, i, j, _p(n), _p(d), _p(limbs), limbs.size))
        v = decode_rats(n, d, limbs)[0]
        return None if v == 0 else v

    # ---- fused drivers ------------------------------------------------------------------------
    def iterate(self, rule, max_iters=-1):
        return self._chk(self.L.relp_iterate(self.h, rule, max_iters))

    def solve_relaxation(self, rule1=FIRST_PROFITABLE_WITH_MEMORY, rule2=STEEPEST_DESCENT, max_pivots=-1):
        pa = self.pa
        return self._chk(self.L.relp_solve_relaxation(self.h, rule1, rule2, len(pa.piv_rows), _p(pa.piv_rows),
                                                      _p(pa.piv_cols), max_pivots))

    def solve_continue(self, max_pivots=-1):
        return self._chk(self.L.relp_solve_continue(self.h, max_pivots))

    def solve_to_handover(self, rule1=FIRST_PROFITABLE_WITH_MEMORY):
        pa = self.pa
        return self._chk(self.L.relp_solve_to_handover(self.h, rule1, len(pa.piv_rows), _p(pa.piv_rows), _p(pa.piv_cols)))

    # ---- readers ------------------------------------------------------------------------------
    def dims(self):
        v = [C.c_int32() for _ in range(4)]
        self._chk(self.L.relp_dims(self.h, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)     # m, n, nr_artificial, phase

    @property
    def m(self):
        return self.dims()[0]

    @property
    def nr_artificial(self):
        return self.dims()[2]

    @property
    def mloc(self):
        m = self.m
        per = (m + self.world - 1) // self.world
        lo = min(m, self.rank * per)
        return max(0, min(m, lo + per) - lo)

    @property
    def row_lo(self):
        m = self.m
        per = (m + self.world - 1) // self.world
        return min(m, self.rank * per)

    def _read_raw(self, fn, count, *args):
        n = np.zeros(max(count, 1), np.int64); d = np.zeros(max(count, 1), np.int64)
        while True:
            limbs = np.zeros(self._limb_cap, np.uint32)
            used = C.c_size_t(0)
            st = fn(self.h, *args, _p(n), _p(d), _p(limbs), limbs.size, C.byref(used))
            if st == ERR_ARGUMENT and used.value > limbs.size:
                self._limb_cap = int(used.value) * 2
                continue
            self._chk(st)
            return n[:count], d[:count], limbs

    def _read(self, fn, count, *args):
        return decode_rats(*self._read_raw(fn, count, *args))

    def basis_indices(self):
        out = np.zeros(self.m, np.int32)
        self._chk(self.L.relp_read_basis(self.h, _p(out)))
        return out

    def b(self):
        return self._read(self.L.relp_read_b, self.mloc)

    def minus_pi(self):
        return self._read(self.L.relp_read_minus_pi, self.m)

    def minus_objective(self):
        return self._read(self.L.relp_read_minus_objective, 1)[0]

    def get_objective_function_value(self):
        return -self.minus_objective()

    def carry(self):
        """(minus_obj, minus_pi[], b[], basis[], rows[] as dict col->Fraction) -- the oracle's layout."""
        m, mloc = self.m, self.mloc
        flat = self._read(self.L.relp_read_carry, (mloc + 1) * (m + 1))
        rows = [flat[i * (m + 1):(i + 1) * (m + 1)] for i in range(mloc + 1)]
        z = rows[mloc]
        return (z[0], z[1:], [r[0] for r in rows[:mloc]], [int(x) for x in self.basis_indices()],
                [{k: v for k, v in enumerate(r[1:]) if v != 0} for r in rows[:mloc]])

    def current_bfs(self):
        """carry/mod.rs:616-625: (basis column, value) for b_i != 0, sorted by column."""
        n, d, limbs = self._read_raw(self.L.relp_read_b, self.mloc)
        basis = self.basis_indices()
        nz = np.nonzero((n != 0) | (d == 0))[0]    # d == 0: multi-limb value (exported handles start at 0), never zero
        vals = decode_rats(n[nz], d[nz], limbs)
        cols = basis[self.row_lo + nz]
        order = np.argsort(cols, kind="stable")
        return [(int(cols[i]), vals[i]) for i in order]

    def current_bfs_arrays(self):
        """The same answer as current_bfs() without building Python objects: (columns, numerators,
        denominators, limbs) as numpy arrays sorted by column; a denominator of 0 marks a multi-limb value
        whose numerator is its offset into `limbs` (decode_rats turns the three arrays into Fractions)."""
        n, d, limbs = self._read_raw(self.L.relp_read_b, self.mloc)
        basis = self.basis_indices()
        nz = np.nonzero((n != 0) | (d == 0))[0]
        cols = basis[self.row_lo + nz]
        order = np.argsort(cols, kind="stable")
        return cols[order], n[nz][order], d[nz][order], limbs

    def log(self):
        n = self.L.relp_log_size(self.h)
        arr = (IterLog * max(n, 1))()
        self._chk(self.L.relp_read_log(self.h, 0, n, arr))
        return [(chr(arr[k].tag), arr[k].j, arr[k].r, arr[k].leaving, arr[k].T, arr[k].S) for k in range(n)]

    def stats(self):
        s = Stats()
        self._chk(self.L.relp_stats_get(self.h, C.byref(s)))
        return s

    def rows_removed(self):
        out = np.zeros(max(self.pa.m, 1), np.int32)
        n = self.L.relp_rows_removed(self.h, _p(out), out.size)
        return [int(x) for x in out[:n]]


def solve_relaxation(provider, rule1=FIRST_PROFITABLE_WITH_MEMORY, rule2=STEEPEST_DESCENT, **kw):
    """`provider.solve_relaxation::<Carry<RationalBig, BasisInverseRows<_>>>()` (two_phase/mod.rs:30-75).
    Returns (status, carry)."""
    c = Carry(provider, **kw)
    return c.solve_relaxation(rule1, rule2), c
