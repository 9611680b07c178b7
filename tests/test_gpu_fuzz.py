"""Randomised parity sweep (tools/fuzz_parity.py) as part of the GPU suite: random odd/even geometries,
up- and down-scales, every caps combination, sink format and path flag, byte for byte against the oracle."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_random_geometries_and_caps():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_parity.py"), "250", "20261106"],
                       capture_output=True, text=True, timeout=400)
    assert r.returncode == 0 and "0 mismatches" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]
