"""The edge-case and size sweeps of tools/ as part of the GPU suite: blank frames inside a batch, stereo with an empty
side, matcher problems with nothing on one side, disjoint vocabularies, tiny 2-NN problems, degenerate undistortion maps
(tools/edge_cases.py); 4K frames, 12 000 features, frames too small for a FAST cell (tools/big_cases.py)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("script", ["edge_cases.py", "big_cases.py"])
def test_sweep(script):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", script)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "failures: 0" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]
    assert "BAD" not in r.stdout
