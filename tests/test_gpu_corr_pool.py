"""f2: FeatureCorrelation with the relocalisation max-pool fused into the GEMM epilogue (csrc/corr4d_tc.cu) against
(a) the unfused GPU path, bit for bit, (b) the CPU oracle on inputs whose correlations are exact integers, ties
included, (c) the fused featureL2Norm prologue."""
import pytest
import torch

from conftest import nerr
from oracle import head as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def F_(cuda):
    from ncnet_b200 import functional
    return functional


@pytest.mark.parametrize("shape,dt,idt,norm", [
    ((96, 10, 14, 12, 8), torch.float32, torch.int64, False),
    ((64, 6, 22, 18, 6), torch.float32, torch.uint8, True),
    ((128, 40, 30, 34, 28), torch.float16, torch.int64, False),      # several M and N tiles, ragged last tiles
    ((200, 26, 26, 24, 50), torch.float16, torch.uint8, True),
])
def test_fused_equals_unfused_bit_for_bit(F_, cuda, monkeypatch, shape, dt, idt, norm):
    monkeypatch.setenv("NCNET_CONV_MODE", "tc16")
    torch.manual_seed(71)
    c, ha, wa, hb, wb = shape
    fa = O.feature_l2norm(torch.relu(torch.randn(1, c, ha, wa))).to(cuda).to(dt)
    fb = O.feature_l2norm(torch.relu(torch.randn(1, c, hb, wb))).to(cuda).to(dt)
    fused = F_.feature_correlation_pooled(fa, fb, 2, normalization=norm, index_dtype=idt)
    corr = F_.feature_correlation(fa, fb, "4D", norm)                 # tensor-core GEMM, volume written
    plain = F_.maxpool4d(corr, 2, index_dtype=idt)
    assert fused[0].dtype == dt and tuple(fused[0].shape) == (1, 1, ha // 2, wa // 2, hb // 2, wb // 2)
    for a, b in zip(fused, plain):
        assert a.dtype == b.dtype and torch.equal(a, b)


@pytest.mark.parametrize("dt", [torch.float32, torch.float16])
def test_fused_vs_oracle_exact_integers_with_ties(F_, cuda, monkeypatch, dt):
    """Small-integer features: every dot product is an exact integer in fp16 operands / fp32 accumulation and in the
    fp32 oracle alike, and the pooling windows are full of ties -> values AND offsets must equal the oracle's."""
    monkeypatch.setenv("NCNET_CONV_MODE", "tc16")
    torch.manual_seed(72)
    fa = torch.randint(0, 3, (1, 24, 12, 10)).float()
    fb = torch.randint(0, 3, (1, 24, 8, 14)).float()
    ref = O.maxpool4d(O.feature_correlation(fa, fb, "4D", False), 2)
    got = F_.feature_correlation_pooled(fa.to(cuda).to(dt), fb.to(cuda).to(dt), 2)
    assert torch.equal(got[0].float().cpu(), ref[0])
    ties = 0
    for a, b in zip(got[1:], ref[1:]):
        assert torch.equal(a.cpu(), b.long())
    win = O.feature_correlation(fa, fb, "4D", False)[0, 0].reshape(6, 2, 5, 2, 4, 2, 7, 2).permute(0, 2, 4, 6, 1, 3, 5, 7).reshape(-1, 16)
    ties = int(((win == win.max(1, keepdim=True)[0]).sum(1) > 1).sum())
    assert ties > 100            # the test really exercises the tie-break


def test_fused_l2norm_prologue(F_, cuda, monkeypatch):
    """l2norm_inputs: un-normalised trunk outputs in, featureL2Norm (lib/model.py:14-17) applied while packing."""
    monkeypatch.setenv("NCNET_CONV_MODE", "tc16")
    torch.manual_seed(73)
    raw_a = torch.relu(torch.randn(1, 256, 16, 12)) * 3.0
    raw_b = torch.relu(torch.randn(1, 256, 14, 18)) * 0.5
    ref = O.maxpool4d(O.feature_correlation(O.feature_l2norm(raw_a), O.feature_l2norm(raw_b), "4D", False), 2)
    got = F_.feature_correlation_pooled(raw_a.to(cuda), raw_b.to(cuda), 2, l2norm_inputs=True)
    assert nerr(got[0], ref[0]) < 1e-3
    agree = torch.stack([(a.cpu() == b.long()) for a, b in zip(got[1:], ref[1:])]).all(0).float().mean().item()
    assert agree > 0.97, agree
    # and through the plain (unpooled) tensor-core entry point
    from ncnet_b200 import _cabi
    L = _cabi.lib()
    out = torch.empty(1, 16 * 12, 14 * 18, device=cuda)
    scratch = torch.empty(L.ncnet_corr4d_tc_scratch_bytes(1, 256, 192, 252), device=cuda, dtype=torch.uint8)
    a, b = raw_a.to(cuda).contiguous(), raw_b.to(cuda).contiguous()
    _cabi.check(L.ncnet_corr4d_fwd_tc(a.data_ptr(), b.data_ptr(), out.data_ptr(), scratch.data_ptr(), 1, 256, 192, 252, _cabi.F32,
                                      0, 1, torch.cuda.current_stream().cuda_stream), "corr4d_fwd_tc")
    full = O.feature_correlation(O.feature_l2norm(raw_a), O.feature_l2norm(raw_b), "4D", False).reshape(1, 192, 252)
    assert nerr(out, full) < 1e-3
