"""CPU: the C-ABI library loads, exports every symbol include/ncnet_b200.h declares, validates
arguments without touching a GPU; the Python drop-in mirrors the reference's module structure."""
import ctypes
import os
import re

import pytest
import torch

from conftest import ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "ncnet_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ncnet_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from ncnet_b200 import _cabi
    lib = _cabi.lib()
    names = _declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ncnet_b200.h but not exported"
    assert set(names) == set(_cabi.SIGNATURES), set(names) ^ set(_cabi.SIGNATURES)
    assert lib.ncnet_version() >= 100


def test_argument_validation_without_gpu():
    from ncnet_b200 import _cabi
    lib = _cabi.lib()
    # null pointers / bad shapes are rejected before any CUDA call
    assert lib.ncnet_mutual_matching_fwd(None, None, None, 1, 4, 4, 1, 0, None) != 0
    assert b"null" in lib.ncnet_last_error()
    one = ctypes.c_void_p(16)
    rc = lib.ncnet_conv4d_fwd(one, one, None, one, one, 1, 1, 1, 4, 4, 4, 4, 4, 0, 0, 1, 0, 0, 0, 0, None)
    assert rc != 0 and b"even kernel" in lib.ncnet_last_error()
    rc = lib.ncnet_maxpool4d_fwd(one, one, one, one, one, one, 5, 4, 4, 4, 2, 8, 0, None)
    assert rc != 0 and b"divisible" in lib.ncnet_last_error()
    assert lib.ncnet_mutual_matching_keys_bytes(2, 10, 12) == 2 * 22 * 8
    assert lib.ncnet_conv4d_scratch_bytes(1, 16, 16, 25, 25, 25, 25, 5, 1, 0) >= 2 * 16 * 16 * 625 * 4


def test_cpu_tensors_raise_no_fallback():
    import ncnet_b200 as nb
    with pytest.raises(RuntimeError, match="no CPU path"):
        nb.MutualMatching(torch.rand(1, 1, 2, 2, 2, 2))
    with pytest.raises(RuntimeError, match="no CPU path"):
        nb.corr_to_matches(torch.rand(1, 1, 2, 2, 2, 2))
    with pytest.raises(RuntimeError, match="no CPU path"):
        nb.conv4d(torch.rand(1, 1, 3, 3, 3, 3), torch.rand(1, 1, 3, 3, 3, 3))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "ncnet_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src, f"{f} mentions the oracle"


def test_module_structure_matches_reference():
    """state_dict keys/shapes of lib/model.py:122-141 + lib/conv4d.py:76-77."""
    import ncnet_b200 as nb
    nc = nb.NeighConsensus(use_cuda=False, kernel_sizes=[5, 5, 5], channels=[16, 16, 1])
    sd = nc.state_dict()
    assert list(sd) == ["conv.0.weight", "conv.0.bias", "conv.2.weight", "conv.2.bias", "conv.4.weight", "conv.4.bias"]
    assert tuple(sd["conv.0.weight"].shape) == (5, 16, 1, 5, 5, 5)
    assert tuple(sd["conv.2.weight"].shape) == (5, 16, 16, 5, 5, 5)
    assert tuple(sd["conv.4.weight"].shape) == (5, 1, 16, 5, 5, 5)
    assert sum(p.numel() for p in nc.parameters()) == 180033
    assert all(isinstance(nc.conv[i], torch.nn.ReLU) for i in (1, 3, 5))
    assert all(hasattr(nc.conv[i], "use_half") for i in (0, 2, 4))
    c = nb.Conv4d(3, 4, 3, pre_permuted_filters=False)
    assert tuple(c.weight.shape) == (4, 3, 3, 3, 3, 3)
    bound = 1.0 / (3 * 81) ** 0.5
    assert c.weight.abs().max() <= bound and c.bias.abs().max() <= bound
    d = nb.NeighConsensus(use_cuda=False)
    assert d.kernel_sizes == [3, 3, 3] and d.channels == [10, 10, 1] and d.symmetric_mode


def test_axis_helpers():
    import ncnet_b200 as nb
    x = torch.tensor([1.0, 13.0, 25.0])
    n = nb.normalize_axis(x, 25)
    assert torch.allclose(n, torch.tensor([-1.0, 0.0, 1.0]))
    assert torch.allclose(nb.unnormalize_axis(n, 25), x)


def test_lib_shim_resolves_head_modules_to_this_library(tmp_path, monkeypatch):
    """ncnet_b200/dropin/lib: `from lib.model import ImMatchNet` (train.py:15) gives this library's class while the
    reference's other lib.* modules keep resolving to the reference checkout."""
    import importlib
    import sys
    fake = tmp_path / "ref" / "lib"
    fake.mkdir(parents=True)
    (fake / "torch_util.py").write_text("def str_to_bool(v):\n    return v == 'True'\n")
    (fake / "model.py").write_text("raise ImportError('the reference head must not be imported')\n")
    dropin = os.path.join(ROOT, "ncnet_b200", "dropin")
    monkeypatch.setenv("NCNET_REFERENCE_ROOT", str(tmp_path / "ref"))
    monkeypatch.syspath_prepend(dropin)
    for name in [m for m in sys.modules if m == "lib" or m.startswith("lib.")]:
        monkeypatch.delitem(sys.modules, name)
    lib = importlib.import_module("lib")
    assert os.path.dirname(lib.__file__) == os.path.join(dropin, "lib")
    import ncnet_b200
    from lib.conv4d import Conv4d, conv4d
    from lib.eval_util import pck_metric
    from lib.model import ImMatchNet, MutualMatching
    from lib.point_tnf import bilinearInterpPointTnf, corr_to_matches, normalize_axis
    from lib.torch_util import str_to_bool
    assert ImMatchNet is ncnet_b200.ImMatchNet and Conv4d is ncnet_b200.Conv4d and corr_to_matches is ncnet_b200.corr_to_matches
    assert str_to_bool("True") and callable(pck_metric) and callable(bilinearInterpPointTnf) and callable(MutualMatching)
    for name in [m for m in sys.modules if m == "lib" or m.startswith("lib.")]:
        monkeypatch.delitem(sys.modules, name)
