"""ORACLE -- TEST INFRASTRUCTURE ONLY.

The same Python binding as oracle/pyoracle.py, bound to oracle/_ref/libgforb_ref.so (or libgforb_ref_native.so): the
REFERENCE'S OWN CODE -- /root/reference/src/ORBextractor.cc compiled whole and unmodified, and the Frame / ORBmatcher
member functions oracle/ref_extract.py lifts verbatim -- behind the entry points of oracle/gforb_oracle.h
(oracle/ref_capi.cpp).  tests/test_oracle_vs_ref.py holds the restatement to it bit for bit; bench.py's
`--impl reference` arm times it.  Entry points the reference has no code for raise AttributeError when called.
"""
import importlib.util
import sys
from pathlib import Path

_HERE = Path(__file__).resolve().parent
_LOADED = {}


def load(variant="ref"):
    """variant: "ref" (contraction off, x86-64-v3: travels to the GPU box) or "ref_native" (the reference's own
    -O3 -march=native flags; only meaningful on the machine that built it)."""
    assert variant in ("ref", "ref_native")
    if variant not in _LOADED:
        name = f"pyoracle_{variant}"
        spec = importlib.util.spec_from_file_location(name, _HERE / "pyoracle.py")
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
        mod._VARIANT = variant
        _LOADED[variant] = mod
    return _LOADED[variant]


def available(variant="ref"):
    return (_HERE / "_ref" / f"libgforb_{variant}.so").exists() or Path("/root/reference").exists()
