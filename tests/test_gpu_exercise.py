"""Runs scripts/sanitize_small.py (every public entry point once on small inputs: frames whole / sharded / odd sizes, batch,
termination rounds on an opaque scene, the peer frame with its fence kernels, stage probes, the wrapped-dense level mode,
density grid, post-process) in a fresh process.  compute-sanitizer is closed on this GPU pool, so the exercise is run plain:
a bad access shows up as a CUDA error (every call's status is checked) or a crash of the child."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_small_end_to_end_exercise(built_lib):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "sanitize_small.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "sanitize exercise done" in r.stdout, (r.stdout[-500:], r.stderr[-2000:])
