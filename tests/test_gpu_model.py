"""a8: ``ImMatchNet`` built and called the way the reference's scripts do it (lib/model.py:193-282) -- real
constructor, random-initialised ResNet-101 trunk (no network here: NCNET_TRUNK_WEIGHTS=random), checkpoint round trip
with ``args`` overriding the constructor's NeighConsensus shape, ``forward({'source_image','target_image'})`` at the
PF-Pascal image size, and the InLoc variant (half precision + relocalisation).  The head is compared with the CPU
oracle fed the SAME trunk features, so the trunk (stock torchvision/cuDNN) drops out of the comparison."""
import argparse

import pytest
import torch

from conftest import nerr
from oracle import head as O

pytestmark = pytest.mark.gpu


def _oracle_nc(model, ks, chs):
    onc = O.NeighConsensus(ks, chs)
    onc.load_state_dict({k: v.detach().float().cpu() for k, v in model.NeighConsensus.state_dict().items()})
    return onc


def test_immatchnet_constructor_forward_and_checkpoint(cuda, tmp_path, monkeypatch):
    import ncnet_b200 as nb
    monkeypatch.setenv("NCNET_TRUNK_WEIGHTS", "random")
    torch.manual_seed(0)
    ks, chs = [5, 5, 5], [16, 16, 1]
    model = nb.ImMatchNet(ncons_kernel_sizes=ks, ncons_channels=chs, use_cuda=True)
    assert isinstance(model.FeatureExtraction, nb.FeatureExtraction) and isinstance(model.NeighConsensus, nb.NeighConsensus)
    assert not model.FeatureExtraction.training
    assert all(not p.requires_grad for p in model.FeatureExtraction.parameters())           # frozen trunk (train_fe=False)
    assert [tuple(p.shape) for p in model.NeighConsensus.parameters()] == [(5, 16, 1, 5, 5, 5), (16,), (5, 16, 16, 5, 5, 5),
                                                                           (16,), (5, 1, 16, 5, 5, 5), (1,)]
    batch = {"source_image": torch.randn(1, 3, 400, 400, device=cuda), "target_image": torch.randn(1, 3, 400, 400, device=cuda)}
    with torch.no_grad():
        fa, fb = model.FeatureExtraction(batch["source_image"]), model.FeatureExtraction(batch["target_image"])
        assert tuple(fa.shape) == (1, 1024, 25, 25)
        ref = O.head_forward(fa.cpu(), fb.cpu(), _oracle_nc(model, ks, chs))
        outs = {}
        for mode, tol in (("fp32", 1e-5), ("tc16", 2e-3)):
            monkeypatch.setenv("NCNET_CONV_MODE", mode)
            out = model(batch)
            assert tuple(out.shape) == (1, 1, 25, 25, 25, 25) and out.is_contiguous()
            out.view(1, 625, 25, 25)                                                        # train.py:125-126 relies on it
            assert nerr(out, ref) < tol, (mode, nerr(out, ref))
            outs[mode] = out
    # training through the real module: gradients reach the NeighConsensus parameters only
    monkeypatch.setenv("NCNET_CONV_MODE", "fp32")
    from ncnet_b200.train_util import weak_loss
    two = {"source_image": torch.randn(2, 3, 224, 224, device=cuda), "target_image": torch.randn(2, 3, 224, 224, device=cuda)}
    loss = weak_loss(model, two)
    loss.backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.NeighConsensus.parameters())
    assert all(p.grad is None for p in model.FeatureExtraction.parameters())

    # checkpoint round trip: 'args' override the constructor's ncons_* (lib/model.py:211-219), weights are copied
    path = str(tmp_path / "ckpt.pth.tar")
    torch.save({"state_dict": model.state_dict(), "args": argparse.Namespace(ncons_channels=chs, ncons_kernel_sizes=ks)}, path)
    again = nb.ImMatchNet(checkpoint=path, ncons_kernel_sizes=[3, 3, 3], ncons_channels=[10, 10, 1], use_cuda=True)
    assert again.NeighConsensus.kernel_sizes == ks and again.NeighConsensus.channels == chs
    for (ka, va), (kb, vb) in zip(model.state_dict().items(), again.state_dict().items()):
        assert ka == kb and torch.equal(va, vb), ka
    with torch.no_grad():
        assert torch.equal(again(batch), outs["fp32"])


def test_immatchnet_half_precision_relocalisation(cuda, tmp_path, monkeypatch):
    """The InLoc construction (eval_inloc.py:60-64): half_precision=True, relocalization_k_size=2, NC 3-3 / 16-1."""
    import ncnet_b200 as nb
    monkeypatch.setenv("NCNET_TRUNK_WEIGHTS", "random")
    monkeypatch.setenv("NCNET_CONV_MODE", "tc16")
    torch.manual_seed(1)
    ks, chs = [3, 3], [16, 1]
    model = nb.ImMatchNet(ncons_kernel_sizes=ks, ncons_channels=chs, use_cuda=True, half_precision=True,
                          relocalization_k_size=2)
    assert all(p.dtype == torch.float16 for p in model.NeighConsensus.parameters())
    assert all(m.use_half for m in model.NeighConsensus.conv if isinstance(m, nb.Conv4d))
    batch = {"source_image": torch.randn(1, 3, 384, 512, device=cuda), "target_image": torch.randn(1, 3, 512, 384, device=cuda)}
    with torch.no_grad():
        corr4d, delta4d = model(batch)
        fa = model.FeatureExtraction(batch["source_image"]).half()
        fb = model.FeatureExtraction(batch["target_image"]).half()
    assert corr4d.dtype == torch.float16 and tuple(corr4d.shape) == (1, 1, 12, 16, 16, 12) and len(delta4d) == 4
    ref, rdelta = O.head_forward(fa.float().cpu(), fb.float().cpu(), _oracle_nc(model, ks, chs), relocalization_k_size=2)
    assert nerr(corr4d.float(), ref) < 5e-3
    m = nb.corr_to_matches(corr4d, delta4d=delta4d, k_size=2, do_softmax=True, scale="positive", return_indices=True)
    mo = O.corr_to_matches(ref, delta4d=rdelta, k_size=2, do_softmax=True, scale="positive", return_indices=True)
    # a random-initialised trunk gives nearly constant feature maps, i.e. volumes full of near-ties: compare the match
    # scores (the arg-max itself is compared on structured inputs in test_gpu_tc_pinned.py)
    assert m[0].shape == mo[0].shape and m[5].dtype == torch.int64
    assert nerr(m[4].float(), mo[4]) < 2e-2
