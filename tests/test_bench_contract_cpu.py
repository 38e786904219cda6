"""bench.py contract on a machine without a GPU: the reference arm prints exactly ONE JSON line on stdout with the
keys the driver reads, and the gforb arm refuses to run without a CUDA device (no CPU fallback)."""
import json
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-500:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "stereo_frames/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0
    # oracle/_ref (the reference's own code) is the CPU arm wherever it was built; the port only stands in without it
    have_ref = (ROOT / "oracle" / "_ref" / "libgforb_ref.so").exists()
    assert d["cpu_baseline"]["kind"] == ("reference" if have_ref else "port")
    assert d["cpu_baseline"]["cores"] >= 1 and d["gpu_launches"] == 0 and d["cpu_baseline"]["cpu_model"]
    assert "752x480" in d["config"]["workload"]


def test_reference_arm_other_config():
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--config", "2tum", "--steps", "1",
                          "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-500:]
    d = json.loads(out.stdout.strip())
    assert d["data"] == "bundled" and "512x512" in d["metric"] and d["value"] > 0


def test_gforb_arm_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--steps", "1"], capture_output=True, text=True, timeout=600)
    assert out.returncode != 0 and "no CPU fallback" in (out.stderr + out.stdout)
