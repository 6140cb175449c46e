"""Randomised parity (tools/fuzz_parity.py) as part of the GPU suite: about twenty seconds of random model sizes,
lattices, frusta and PCD origins, every field / set / grouped gain / cluster against the oracle."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT


@pytest.mark.gpu
def test_fuzz_parity_short():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_parity.py"), "20", "11"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "TOTAL MISMATCHES 0" in out.stdout
