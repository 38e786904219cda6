"""The product library builds from a CLEAN tree with the committed recipe (nvcc cross-compiles sm_100a without a GPU) and
the result exports every symbol include/gforb.h declares -- what __graft_entry__.build() runs, minus the leftovers of
earlier builds in the working tree."""
import ctypes as C
import os
import re
import shutil
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_clean_tree_build(tmp_path):
    pkg = tmp_path / "gf-orb-slam2_b200"
    shutil.copytree(ROOT / "gf-orb-slam2_b200" / "csrc", pkg / "csrc", ignore=shutil.ignore_patterns("_build"))
    shutil.copytree(ROOT / "include", tmp_path / "include")
    assert not (pkg / "libgforb.so").exists() and not (pkg / "csrc" / "_build").exists()
    jobs = str(max(1, min(8, os.cpu_count() or 1)))
    subprocess.check_call(["make", "-C", str(pkg / "csrc"), "-j", jobs], stdout=subprocess.DEVNULL)
    so = pkg / "libgforb.so"
    assert so.exists()
    sass = subprocess.run(["cuobjdump", "-lelf", str(so)], capture_output=True, text=True).stdout
    assert "sm_100a" in sass                                               # built for the target, nothing else
    assert not re.search(r"sm_(?!100a)\d+", sass)
    declared = set(re.findall(r"GFORB_API[^;(]*?\b(gforb_\w+)\s*\(", (ROOT / "include/gforb.h").read_text()))
    lib = C.CDLL(str(so))
    for name in sorted(declared):
        assert hasattr(lib, name), name
