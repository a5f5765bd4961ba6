"""Importable alias of the package directory `rust-lp_b200/` (a hyphen is not a valid Python
identifier).  `import rust_lp_b200.capi` resolves to `rust-lp_b200/capi.py`."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "rust-lp_b200")]
