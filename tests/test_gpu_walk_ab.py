"""The persistent walk kernels (csrc/conv4d_tc_walk16.cu, csrc/conv4d_tc_walk1.cu) against the per-plane tensor-core
kernel they replace inside a chained stack: same fp16 products, fp32 accumulation, only the summation order differs.
scripts/w16_check.py switches between them with NCNET_NO_W16 / NCNET_NO_W1 (read per call) and prints one JSON line per
case; it covers what the oracle-pinned tests cannot afford on the CPU (two chained 16->16 layers, the full PF-Pascal and
InLoc shapes)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_walk_kernels_match_per_plane_kernels(cuda):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "w16_check.py"), "quick"], cwd=ROOT, capture_output=True,
                       text=True, timeout=600)
    print(r.stdout[-4000:], r.stderr[-2000:])
    assert r.returncode == 0 and "W16 CHECK PASS" in r.stdout
