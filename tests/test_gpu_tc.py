"""GPU parity of the tensor-core Conv4d mode (NCNET_CONV_TC16: fp16 operands, fp32 accumulation in TMEM).

Tolerance: max-abs error / max-abs reference <= 1e-3 per layer against the fp32 oracle (the north star's bound for
the FP32-accumulate mode); <= 2e-3 through the three-layer stack whose hidden activations are stored in fp16.
ReLU makes gradients discontinuous in the pre-activations, so gradient parity is checked where no pre-activation
sits within rounding distance of zero (positive weights and inputs), plus the arg-max agreement the north star asks
to report.
"""
import pytest
import torch

from conftest import load_golden, nerr
from oracle import head as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def F_(cuda):
    from ncnet_b200 import functional
    return functional


def _tc():
    from ncnet_b200._cabi import CONV_TC16
    return CONV_TC16


@pytest.mark.parametrize("tag", ["k3", "k5", "k5c16"])
def test_tc_conv4d_golden(F_, cuda, tag):
    g = load_golden("conv4d_" + tag, cuda)
    x = g["x"].clone().requires_grad_(True)
    w = g["weight"].clone().requires_grad_(True)
    b = g["bias"].clone().requires_grad_(True)
    y = F_.conv4d(x, w, b, permute_filters=False, mode=_tc())
    assert nerr(y, g["y"]) < 1e-3
    gx, gw, gb = torch.autograd.grad(y, (x, w, b), g["grad_y"])
    assert nerr(gx, g["grad_x"]) < 1e-3 and nerr(gw, g["grad_weight"]) < 1e-3 and nerr(gb, g["grad_bias"]) < 1e-3


@pytest.mark.parametrize("cin,cout,k,dims,bsz", [
    (16, 16, 5, (6, 7, 25, 25), 2), (1, 16, 5, (7, 6, 12, 13), 2), (16, 1, 5, (6, 6, 11, 9), 1),
    (16, 16, 3, (5, 6, 20, 37), 1), (10, 10, 3, (5, 5, 8, 9), 2), (1, 16, 3, (5, 6, 50, 37), 1),
    (16, 1, 3, (5, 5, 9, 10), 2), (1, 12, 3, (6, 5, 10, 9), 2), (12, 1, 5, (5, 6, 25, 25), 1)])
def test_tc_conv4d_vs_oracle(F_, cuda, cin, cout, k, dims, bsz):
    torch.manual_seed(31)
    x = torch.rand(bsz, cin, *dims)
    layer = O.Conv4d(cin, cout, k)
    w, b = layer.weight.detach(), layer.bias.detach()
    xo, wo, bo = x.clone().requires_grad_(True), w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    yo = O.conv4d(xo, wo, bo, permute_filters=False)
    gy = torch.randn_like(yo)
    rx, rw, rb = torch.autograd.grad(yo, (xo, wo, bo), gy)
    xg, wg, bg = (t.to(cuda).requires_grad_(True) for t in (x, w, b))
    y = F_.conv4d(xg, wg, bg, permute_filters=False, mode=_tc())
    assert nerr(y, yo) < 1e-3
    gx, gw, gb = torch.autograd.grad(y, (xg, wg, bg), gy.to(cuda))
    assert nerr(gx, rx) < 1e-3 and nerr(gw, rw) < 1e-3 and nerr(gb, rb) < 1e-3


@pytest.mark.parametrize("cin,cout,k", [(1, 16, 5), (16, 1, 5), (16, 16, 5)])
def test_tc_conv4d_half_tensors_vs_oracle(F_, cuda, cin, cout, k):
    """fp16 activations (the InLoc storage type) through the tensor-core forward / dgrad / wgrad kernels."""
    torch.manual_seed(39)
    dims, bsz = (5, 6, 12, 11), 2
    x = torch.rand(bsz, cin, *dims).half().float()
    layer = O.Conv4d(cin, cout, k)
    w, b = layer.weight.detach(), layer.bias.detach()
    xo, wo, bo = x.clone().requires_grad_(True), w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    yo = O.conv4d(xo, wo, bo, permute_filters=False)
    gy = torch.randn_like(yo).half().float()
    rx, rw, rb = torch.autograd.grad(yo, (xo, wo, bo), gy)
    xg = x.half().to(cuda).requires_grad_(True)
    wg, bg = (t.to(cuda).requires_grad_(True) for t in (w, b))
    y = F_.conv4d(xg, wg, bg, permute_filters=False, mode=_tc())
    assert y.dtype == torch.float16 and nerr(y.float(), yo) < 3e-3
    gx, gw, gb = torch.autograd.grad(y, (xg, wg, bg), gy.half().to(cuda))
    assert nerr(gx.float(), rx) < 3e-3 and nerr(gw.float(), rw) < 3e-3 and nerr(gb.float(), rb) < 3e-3


def test_cb16_roundtrip_and_halo(F_, cuda):
    from ncnet_b200 import _cabi
    lib = _cabi.lib()
    torch.manual_seed(32)
    for k, ch in ((5, 16), (3, 10)):
        x = (torch.randint(-2048, 2048, (3, ch, 4, 5, 6, 7), device=cuda).float() / 64).contiguous()   # exact in fp16
        nbytes = lib.ncnet_cb16_bytes(3, 4, 5, 6, 7, k)
        buf = torch.full((nbytes,), 0x7F, device=cuda, dtype=torch.uint8)
        st = torch.cuda.current_stream().cuda_stream
        _cabi.check(lib.ncnet_cb16_pack(x.data_ptr(), buf.data_ptr(), 3, ch, 4, 5, 6, 7, k, _cabi.F32, st), "pack")
        back = torch.empty_like(x)
        _cabi.check(lib.ncnet_cb16_unpack(buf.data_ptr(), back.data_ptr(), 3, ch, 4, 5, 6, 7, k, _cabi.F32, st), "unpack")
        assert torch.equal(back, x)
        p = k // 2
        cb = buf.view(torch.float16).view(3, 2, 20, 6 + 2 * p, 7 + 2 * p, 8)
        assert cb[:, :, :, :p].abs().sum() == 0 and cb[:, :, :, -p:].abs().sum() == 0
        assert cb[:, :, :, :, :p].abs().sum() == 0 and cb[:, :, :, :, -p:].abs().sum() == 0
        if ch < 16:
            assert cb[:, 1, :, :, :, ch - 8:].abs().sum() == 0          # channel padding is zero


def _stack(nb, F_, cuda, ks, chs, dims, bsz, positive):
    torch.manual_seed(33)
    onc = O.NeighConsensus(ks, chs)
    if positive:
        with torch.no_grad():
            for p in onc.parameters():
                p.abs_()
    nc = nb.NeighConsensus(True, ks, chs)
    nc.load_state_dict(onc.state_dict())
    x = torch.rand(bsz, 1, *dims)
    xo = x.clone().requires_grad_(True)
    yo = onc(xo)
    go = torch.randn_like(yo)
    ro = torch.autograd.grad(yo, [xo] + list(onc.parameters()), go)
    xg = x.to(cuda).requires_grad_(True)
    ws, bs = nc._params()
    yb = F_.neigh_consensus_branches(xg, ws, bs, True, _tc())
    y = yb[:, :1] + yb[:, 1:]
    rg = torch.autograd.grad(y, [xg] + list(nc.parameters()), go.to(cuda))
    return y, yo, rg, ro


@pytest.mark.parametrize("ks,chs,dims,bsz", [([5, 5, 5], [16, 16, 1], (7, 6, 9, 10), 2), ([3, 3], [16, 1], (6, 5, 20, 13), 1),
                                              ([3, 3, 3], [10, 10, 1], (5, 6, 8, 9), 2)])
def test_tc_chained_stack_vs_oracle(F_, cuda, ks, chs, dims, bsz):
    import ncnet_b200 as nb
    y, yo, rg, ro = _stack(nb, F_, cuda, ks, chs, dims, bsz, positive=True)     # every ReLU active: a linear map
    assert nerr(y, yo) < 2e-3
    for a, b in zip(rg, ro):
        assert nerr(a, b) < 3e-3
    y, yo, _, _ = _stack(nb, F_, cuda, ks, chs, dims, bsz, positive=False)      # signed weights: forward only
    assert nerr(y, yo) < 2e-3


def test_tc_head_pf_shape_scores_and_argmax_agreement(F_, cuda, monkeypatch):
    """Whole head at the PF-Pascal shape in tensor-core mode against the fp32 oracle."""
    import ncnet_b200 as nb
    monkeypatch.setenv("NCNET_CONV_MODE", "tc16")
    torch.manual_seed(34)
    fa = O.feature_l2norm(torch.relu(torch.randn(1, 256, 25, 25)))
    fb = O.feature_l2norm(torch.relu(torch.randn(1, 256, 25, 25)))
    onc = O.NeighConsensus([5, 5, 5], [16, 16, 1])
    ref = O.head_forward(fa, fb, onc)
    net = nb.ImMatchNet.__new__(nb.ImMatchNet)
    torch.nn.Module.__init__(net)
    net.relocalization_k_size, net.half_precision = 0, False
    net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
    net.NeighConsensus = nb.NeighConsensus(True, [5, 5, 5], [16, 16, 1])
    net.NeighConsensus.load_state_dict(onc.state_dict())
    with torch.no_grad():
        out = net.match_features(fa.to(cuda), fb.to(cuda))
    e = nerr(out, ref)
    m = nb.corr_to_matches(out, do_softmax=True, return_indices=True)
    mo = O.corr_to_matches(ref.detach(), do_softmax=True, return_indices=True)
    agree = ((m[5].cpu() == mo[5]) & (m[6].cpu() == mo[6])).float().mean().item()
    print(f"tensor-core head: nerr {e:.2e}, arg-max agreement {agree:.4f}")
    assert e < 3e-3 and agree > 0.9


def test_tc_mode_routes_wide_layers_to_cuda_cores(F_, cuda):
    """Cin > 16 per group is outside the tensor-core envelope: TC16 mode must still give the fp32 answer."""
    torch.manual_seed(35)
    x = torch.randn(1, 20, 4, 4, 5, 6, device=cuda)
    w = torch.randn(3, 4, 20, 3, 3, 3, device=cuda) * 0.05
    a = F_.conv4d(x, w, None, permute_filters=False, mode=_tc())
    b = F_.conv4d(x, w, None, permute_filters=False, mode=0)
    assert torch.equal(a, b)


def test_tc_inloc_config_half_precision_relocalisation(F_, cuda, monkeypatch):
    """BASELINE.json configs[2] in miniature: fp16 features, maxpool4d k=2 relocalisation, NC 3,3 / 16,1 on the
    tensor-core kernels, corr_to_matches in both directions -- against the fp32 oracle."""
    import ncnet_b200 as nb
    monkeypatch.setenv("NCNET_CONV_MODE", "tc16")
    torch.manual_seed(36)
    fa = O.feature_l2norm(torch.relu(torch.randn(1, 128, 20, 16)))
    fb = O.feature_l2norm(torch.relu(torch.randn(1, 128, 16, 20)))
    onc = O.NeighConsensus([3, 3], [16, 1])
    ref, rdelta = O.head_forward(fa, fb, onc, relocalization_k_size=2)
    net = nb.ImMatchNet.__new__(nb.ImMatchNet)
    torch.nn.Module.__init__(net)
    net.relocalization_k_size, net.half_precision = 2, True
    net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
    net.NeighConsensus = nb.NeighConsensus(True, [3, 3], [16, 1])
    net.NeighConsensus.load_state_dict(onc.state_dict())
    net.NeighConsensus.half()
    with torch.no_grad():
        out, delta = net.match_features(fa.half().to(cuda), fb.half().to(cuda))
    assert out.dtype == torch.float16 and nerr(out.float(), ref) < 5e-3
    for inv in (False, True):
        m = nb.corr_to_matches(out, delta4d=delta, k_size=2, do_softmax=True, scale="positive", return_indices=True,
                               invert_matching_direction=inv)
        mo = O.corr_to_matches(ref, delta4d=rdelta, k_size=2, do_softmax=True, scale="positive", return_indices=True,
                               invert_matching_direction=inv)
        agree = ((m[5].cpu() == mo[5]) & (m[6].cpu() == mo[6]) & (m[7].cpu() == mo[7]) & (m[8].cpu() == mo[8])).float().mean()
        assert agree > 0.9, float(agree)


def test_slab_head_single_rank_equals_whole_volume(F_, cuda):
    """World size 1 of the slab-sharded head (ncnet_b200.dist) is the ordinary head; the 2-GPU equality run is
    scripts/slab_check.py under torchrun."""
    import ncnet_b200 as nb
    from ncnet_b200 import dist as D
    torch.manual_seed(37)
    fa = O.feature_l2norm(torch.relu(torch.randn(1, 32, 8, 7))).to(cuda)
    fb = O.feature_l2norm(torch.relu(torch.randn(1, 32, 6, 9))).to(cuda)
    net = nb.ImMatchNet.__new__(nb.ImMatchNet)
    torch.nn.Module.__init__(net)
    net.relocalization_k_size, net.half_precision = 0, False
    net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
    net.NeighConsensus = nb.NeighConsensus(True, [3, 3], [16, 1])
    with torch.no_grad():
        assert torch.equal(D.slab_match_features(net, fa, fb, 0, 1), net.match_features(fa, fb))


@pytest.mark.parametrize("shape,dt", [((2, 64, 25, 25, 25, 25), torch.float32), ((1, 96, 9, 13, 20, 7), torch.float32),
                                       ((1, 128, 40, 30, 33, 28), torch.float16)])
def test_tc_feature_correlation_vs_oracle(F_, cuda, monkeypatch, shape, dt):
    """tcgen05 correlation GEMM (operands fp16, fp32 accumulate) against the fp32 oracle, plain and normalised."""
    import ncnet_b200 as nb
    monkeypatch.setenv("NCNET_CONV_MODE", "tc16")
    torch.manual_seed(38)
    b, c, ha, wa, hb, wb = shape
    fa = O.feature_l2norm(torch.relu(torch.randn(b, c, ha, wa)))
    fb = O.feature_l2norm(torch.relu(torch.randn(b, c, hb, wb)))
    for norm in (False, True):
        ref = O.feature_correlation(fa, fb, "4D", norm)
        out = nb.FeatureCorrelation("4D", norm)(fa.to(cuda).to(dt), fb.to(cuda).to(dt))
        assert out.shape == ref.shape and out.dtype == dt
        assert nerr(out.float(), ref) < (2e-3 if dt == torch.float16 else 1e-3) * (3 if norm else 1)
