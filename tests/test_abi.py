"""CPU: the CUDA library builds (nvcc cross-compiles sm_100a without a GPU), loads, exports every
symbol include/relp_b200.h declares, and refuses to run without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

import __graft_entry__ as entry
from rust_lp_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    entry.build_library()
    return C.CDLL(capi.LIB_PATH)


def test_header_symbols_match_binding_list():
    hdr = open(os.path.join(ROOT, "include", "relp_b200.h")).read()
    declared = set(re.findall(r"\b(relp_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(capi.SYMBOLS)


def test_rust_sys_crate_names_every_symbol():
    """integration/rust/relp-b200-sys (never compiled: no Rust toolchain here) declares exactly the header's symbols"""
    src = open(os.path.join(ROOT, "integration", "rust", "relp-b200-sys", "src", "lib.rs")).read()
    block = src[src.index('extern "C" {'):]
    declared = set(re.findall(r"pub fn (relp_[a-z0-9_]+)\s*\(", block[:block.index("\n}\n")]))
    assert declared == set(capi.SYMBOLS)


def test_rust_trait_shim_is_complete():
    """integration/rust/relp-b200 (source only): every method of the reference's InverseMaintener /
    ColumnComputationInfo / PivotRule traits has a body, nothing is elided, and only declared C symbols are called"""
    src = open(os.path.join(ROOT, "integration", "rust", "relp-b200", "src", "lib.rs")).read()
    used = set(re.findall(r"sys::(relp_[a-z0-9_]+)", src)) - {"relp_ctx"}
    assert used <= set(capi.SYMBOLS), used - set(capi.SYMBOLS)
    for method in ("create_for_fully_artificial", "create_for_partially_artificial", "from_basis", "from_basis_pivots",
                   "from_artificial", "from_artificial_remove_rows", "change_basis", "cost_difference", "generate_column",
                   "generate_element", "after_basis_change", "current_bfs", "basis_column_index_for_row", "b",
                   "get_objective_function_value", "get_constraint_value", "column", "into_column", "new",
                   "select_primal_pivot_column", "identity", "invert", "should_refactor", "basis_inverse_row", "m"):
        assert re.search(r"\bfn %s\b[^;]*?\{" % method, src, re.S), method
    assert "unimplemented!" not in src and "todo!" not in src and "..." not in src.replace("0..", "")


def test_library_exports_every_symbol(lib):
    for s in capi.SYMBOLS:
        assert hasattr(lib, s), s
    assert lib.relp_abi_version() == 1


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    L = capi.load_library()
    assert not L.relp_create(0, 0, 1, 1 << 20)
    assert b"no CUDA device" in L.relp_last_error()
    from rust_lp_b200.synth import synthetic_lp
    with pytest.raises(capi.RelpError):
        capi.Carry(synthetic_lp(30, 60, 5))


def test_sass_is_sm100(lib):
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
