"""Randomised parity of the extraction + stereo path (tools/stress_extract.py): random image sizes with every w % 4, levels
narrower than a warp column, extractor parameters, through the single-image handles and the batch handles, bit-exact against
the oracle.  Plus the case that stress found: a key capacity (K around 2,000) that puts the stereo kernel's dynamic shared
memory just under 48 KB, where static + dynamic together exceed what a launch gets without the opt-in."""
import importlib
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

from conftest import B, BF
from synth import synth_stereo

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def test_random_sizes_and_parameters_bit_exact():
    r = subprocess.run([sys.executable, str(ROOT / "tools" / "stress_extract.py"), "24", "4242"], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert '"mismatches": 0' in r.stdout


def test_stereo_shared_memory_window_just_under_48k():
    from oracle import pyoracle as O
    gf = importlib.import_module("gf-orb-slam2_b200")
    w, h, args = 789, 389, (2000, 1.2, 2, 20, 5)
    left, right = synth_stereo(1005, w, h, nshapes=300)
    sb = gf.StereoBatch(w, h, 1, nfeatures=args[0], scaleFactor=args[1], nlevels=args[2], iniThFAST=args[3], minThFAST=args[4])
    out = sb.alloc_outputs(1)
    sb.process(left[None], right[None], out, BF, B)          # raised "stereo launch: invalid argument" before the fix
    oxL, oxR = O.Extractor(*args), O.Extractor(*args)
    okL, odL, _ = oxL.extract(left); okR, odR, _ = oxR.extract(right)
    o = O.stereo_match(oxL, oxR, okL, odL, okR, odR, BF, B)
    n = out["counts"][0]
    assert n == len(okL) and np.array_equal(out["descriptors"][0, :n], odL)
    assert np.array_equal(out["u_right"][0, :n], o["uRight"]) and np.array_equal(out["depth"][0, :n], o["depth"])
