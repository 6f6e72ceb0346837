import os
import sys

import numpy as np
import pytest
import torch

# the parity tests use the fp32 CUDA-core Conv4d unless a test asks for the tensor-core mode explicitly
os.environ.setdefault("NCNET_CONV_MODE", "fp32")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name, device="cpu"):
    """tests/golden/<name>.npz -> dict of torch tensors (fixtures written by oracle/make_golden.py)."""
    out = {}
    with np.load(os.path.join(GOLD, name + ".npz"), allow_pickle=False) as z:
        for k in z.files:
            a = z[k]
            if a.dtype.kind in "US":
                out[k] = str(a)
            else:
                out[k] = torch.from_numpy(np.ascontiguousarray(a)).to(device)
    return out


def nerr(a, b):
    """max-abs error normalised by the reference's max-abs (the tolerance metric of DESIGN.md)."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.numel() == 0:
        return 0.0
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.fixture(scope="session")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def make_head(nb, kernel_sizes, channels, state_dict=None, reloc=0, half=False):
    """An ImMatchNet without its trunk (no torchvision download): the matching head the GPU tests drive through
    ``match_features``.  The real constructor is exercised by tests/test_gpu_model.py."""
    net = nb.ImMatchNet.__new__(nb.ImMatchNet)
    torch.nn.Module.__init__(net)
    net.relocalization_k_size, net.half_precision = reloc, half
    net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
    net.NeighConsensus = nb.NeighConsensus(True, list(kernel_sizes), list(channels))
    if state_dict is not None:
        net.NeighConsensus.load_state_dict(state_dict)
    if half:
        net.NeighConsensus.half()
    return net


def record(key, values):
    """When NCNET_PARITY_REPORT names a file, the parity tests also write what they measured into it (JSON, one entry
    per key) -- that is how profiles/r2_parity_tc16.json is produced on the B200."""
    path = os.environ.get("NCNET_PARITY_REPORT")
    if not path:
        return
    import json
    data = {}
    if os.path.exists(path):
        with open(path) as fh:
            data = json.load(fh)
    data[key] = values
    with open(path, "w") as fh:
        json.dump(data, fh, indent=1, sort_keys=True)
