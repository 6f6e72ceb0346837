"""Two ranks on the CUDA head (torchrun-launched scripts/slab_check.py): the slab-sharded head with its in-place CB16
halo exchange, the cross-slab corr_to_matches merge and the batch-sharded training gradient against their
single-GPU results.  With two GPUs the ranks talk NCCL; on a one-GPU box both ranks share GPU 0 and talk gloo."""
import json
import os
import socket
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_two_rank_slab_head_matches_merge_and_sharded_gradient(cuda, tmp_path):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = tmp_path / "slab.json"
    env = dict(os.environ, NCNET_CONV_MODE="tc16")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "scripts", "slab_check.py"), "--json", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    rep = json.load(open(out))
    assert rep["ok"] and rep["world"] == 2
    for c in rep["cases"]:
        if "bit_identical" in c:
            assert c["bit_identical"] and c["corr_to_matches_merge_ok"], c
        else:
            assert c["max_nerr_over_ranks"] < 1e-3, c
