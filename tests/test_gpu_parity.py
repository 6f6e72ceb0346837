"""GPU parity: the CUDA path (through the C ABI) against the golden fixtures the reference produced
and against the CPU oracle on seeded inputs.

Tolerances (max-abs error / max-abs reference, `nerr`):
  fp32 CUDA-core mode ... 1e-5 forward, 1e-4 gradients (summation order differs from the CPU);
  fp16 storage .......... 2e-3;
  indices ............... bit-exact.
"""
import pytest
import torch

from conftest import load_golden, nerr
from oracle import head as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nb(cuda):
    import ncnet_b200
    return ncnet_b200


# --------------------------------------------------------------------------------------------- #
# golden fixtures (outputs of the reference itself)
# --------------------------------------------------------------------------------------------- #
def test_l2norm_golden(nb, cuda):
    g = load_golden("feature_l2norm", cuda)
    assert nerr(nb.featureL2Norm(g["x"]), g["y"]) < 1e-6


def test_feature_correlation_golden(nb, cuda):
    g = load_golden("feature_correlation", cuda)
    assert nerr(nb.FeatureCorrelation("4D", False)(g["fa"], g["fb"]), g["corr4d"]) < 1e-5
    assert nerr(nb.FeatureCorrelation("4D", True)(g["fa"], g["fb"]), g["corr4d_norm"]) < 1e-5
    assert nerr(nb.FeatureCorrelation("3D", True)(g["fa"], g["fb_sq"]), g["corr3d_norm"]) < 1e-5


def test_feature_correlation_backward(nb, cuda):
    torch.manual_seed(3)
    fa = torch.randn(2, 24, 5, 6, device=cuda, requires_grad=True)
    fb = torch.randn(2, 24, 4, 7, device=cuda, requires_grad=True)
    for norm in (False, True):
        out = nb.FeatureCorrelation("4D", norm)(fa, fb)
        go = torch.randn_like(out)
        ga, gb = torch.autograd.grad(out, (fa, fb), go)
        fa_c, fb_c = fa.detach().cpu().requires_grad_(True), fb.detach().cpu().requires_grad_(True)
        ref = O.feature_correlation(fa_c, fb_c, "4D", norm)
        ra, rb = torch.autograd.grad(ref, (fa_c, fb_c), go.cpu())
        # v/sqrt(v^2+1e-6) amplifies fp32 rounding of v near 0 by 1e3: looser bound for the normalised variant
        tol = 1e-3 if norm else 1e-5
        assert nerr(out, ref) < tol and nerr(ga, ra) < tol and nerr(gb, rb) < tol


def test_mutual_matching_golden(nb, cuda):
    g = load_golden("mutual_matching", cuda)
    x = g["x"].clone().requires_grad_(True)
    y = nb.MutualMatching(x)
    assert nerr(y, g["y"]) < 1e-6
    (gx,) = torch.autograd.grad(y, x, g["grad_y"])
    assert nerr(gx, g["grad_x"]) < 1e-5


@pytest.mark.parametrize("tag", ["k3", "k5", "k5c16"])
def test_conv4d_golden(nb, cuda, tag):
    g = load_golden("conv4d_" + tag, cuda)
    x = g["x"].clone().requires_grad_(True)
    w = g["weight"].clone().requires_grad_(True)
    b = g["bias"].clone().requires_grad_(True)
    y = nb.conv4d(x, w, b, permute_filters=False)
    assert nerr(y, g["y"]) < 1e-5
    gx, gw, gb = torch.autograd.grad(y, (x, w, b), g["grad_y"])
    assert nerr(gx, g["grad_x"]) < 1e-4
    assert nerr(gw, g["grad_weight"]) < 1e-4
    assert nerr(gb, g["grad_bias"]) < 1e-4
    # the un-permuted filter layout (permute_filters=True) gives the same result
    y2 = nb.conv4d(g["x"], O.from_prepermuted(g["weight"].cpu()).to(cuda), g["bias"], permute_filters=True)
    assert torch.equal(y2, y.detach())


def _make_nc(nb, cuda, g):
    ks, chs = g["kernel_sizes"].tolist(), g["channels"].tolist()
    nc = nb.NeighConsensus(use_cuda=True, kernel_sizes=ks, channels=chs)
    with torch.no_grad():
        for i in range(len(ks)):
            nc.conv[2 * i].weight.copy_(g[f"w{i}"]), nc.conv[2 * i].bias.copy_(g[f"b{i}"])
    return nc


@pytest.mark.parametrize("tag", ["k3", "k5", "k3_10"])
def test_neigh_consensus_golden(nb, cuda, tag):
    g = load_golden("neigh_consensus_" + tag, cuda)
    nc = _make_nc(nb, cuda, g)
    x = g["x"].clone().requires_grad_(True)
    y = nc(x)
    assert nerr(y, g["y_sym"]) < 1e-5
    grads = torch.autograd.grad(y, [x] + list(nc.parameters()), g["grad_y"])
    assert nerr(grads[0], g["grad_x"]) < 1e-4
    n = len(g["kernel_sizes"])
    for i in range(n):
        assert nerr(grads[1 + 2 * i], g[f"grad_w{i}"]) < 1e-4, f"w{i}"
        assert nerr(grads[2 + 2 * i], g[f"grad_b{i}"]) < 1e-4, f"b{i}"
    nc.symmetric_mode = False
    assert nerr(nc(g["x"]), g["y_one"]) < 1e-5
    # the layer-by-layer Sequential (Conv4d modules + ReLU) agrees with the fused stack
    assert nerr(nc.conv(g["x"].clone()), g["y_one"]) < 1e-5


def test_maxpool4d_golden(nb, cuda):
    g = load_golden("maxpool4d", cuda)
    out = nb.maxpool4d(g["x"], k_size=int(g["k_size"]))
    assert torch.equal(out[0], g["value"])
    for t, nm in zip(out[1:], ("max_i", "max_j", "max_k", "max_l")):
        assert t.dtype == torch.int64 and torch.equal(t, g[nm]), nm
    from ncnet_b200 import functional as F_
    out8 = F_.maxpool4d(g["x"], 2, index_dtype=torch.uint8)
    for t, nm in zip(out8[1:], ("max_i", "max_j", "max_k", "max_l")):
        assert torch.equal(t.long(), g[nm])


def test_corr_to_matches_golden(nb, cuda):
    g = load_golden("corr_to_matches", cuda)
    names = ("xA", "yA", "xB", "yB", "score", "iA", "jA", "iB", "jB")
    for inv in (0, 1):
        for sm in (0, 1):
            for sc in ("centered", "positive"):
                out = nb.corr_to_matches(g["vol"], do_softmax=bool(sm), scale=sc, return_indices=True,
                                         invert_matching_direction=bool(inv))
                for nm, t in zip(names, out):
                    ref = g[f"inv{inv}_sm{sm}_{sc}_{nm}"]
                    if nm == "score":
                        assert nerr(t, ref) < 2e-6, (inv, sm, sc)
                    else:
                        assert torch.equal(t, ref), (inv, sm, sc, nm)       # indices and LUT coordinates: exact
                assert len(nb.corr_to_matches(g["vol"], do_softmax=bool(sm), scale=sc)) == 5
        deltas = (g["d_i"], g["d_j"], g["d_k"], g["d_l"])
        for dl in (deltas, tuple(d.to(torch.uint8) for d in deltas)):
            out = nb.corr_to_matches(g["pooled"], delta4d=dl, k_size=2, do_softmax=True, scale="positive",
                                     return_indices=True, invert_matching_direction=bool(inv))
            for nm, t in zip(names, out):
                ref = g[f"reloc_inv{inv}_{nm}"]
                assert (nerr(t, ref) < 2e-6) if nm == "score" else torch.equal(t, ref), (inv, nm)


def test_head_and_weak_loss_golden(nb, cuda):
    from ncnet_b200.train_util import weak_loss_from_features
    g = load_golden("head_weak_loss", cuda)
    net = nb.ImMatchNet.__new__(nb.ImMatchNet)
    torch.nn.Module.__init__(net)
    net.relocalization_k_size, net.half_precision = 0, False
    net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
    net.NeighConsensus = nb.NeighConsensus(True, [5, 5, 5], [16, 16, 1])
    with torch.no_grad():
        for i in range(3):
            net.NeighConsensus.conv[2 * i].weight.copy_(g[f"w{i}"]), net.NeighConsensus.conv[2 * i].bias.copy_(g[f"b{i}"])
    assert nerr(net.match_features(g["fa"], g["fb"]), g["head_pos"]) < 1e-5
    assert nerr(net.match_features(torch.roll(g["fa"], -1, 0), g["fb"]), g["head_neg"]) < 1e-5
    loss = weak_loss_from_features(net, g["fa"], g["fb"])
    assert abs(float(loss) - float(g["loss64"])) < 1e-7
    loss.backward()
    for i in range(3):
        layer = net.NeighConsensus.conv[2 * i]
        assert nerr(layer.weight.grad, g[f"grad_w{i}"]) < 2e-3, f"w{i}"      # fp32 vs the fp64 gradient
        assert nerr(layer.bias.grad, g[f"grad_b{i}"]) < 2e-3, f"b{i}"


# --------------------------------------------------------------------------------------------- #
# seeded inputs against the oracle, shapes the fixtures do not cover
# --------------------------------------------------------------------------------------------- #
@pytest.mark.parametrize("shape", [(3, 1, 7, 5, 6, 9), (2, 1, 25, 25, 25, 25), (1, 1, 9, 40, 33, 11)])
def test_mutual_matching_vs_oracle(nb, cuda, shape):
    torch.manual_seed(5)
    x = torch.rand(*shape)
    xg = x.to(cuda).requires_grad_(True)
    y = nb.MutualMatching(xg)
    xo = x.clone().requires_grad_(True)
    yo = O.mutual_matching(xo)
    assert nerr(y, yo) < 1e-6
    go = torch.randn_like(yo)
    (gx,) = torch.autograd.grad(y, xg, go.to(cuda))
    (gr,) = torch.autograd.grad(yo, xo, go)
    assert nerr(gx, gr) < 2e-5


def test_mutual_matching_symmetry_and_negative(nb, cuda):
    """MM(x^T) == MM(x)^T (the 'important for symmetric output' remark, lib/model.py:173);
    negative entries go through the order-preserving key packing."""
    torch.manual_seed(6)
    x = torch.randn(2, 1, 6, 5, 4, 7, device=cuda)
    a = nb.MutualMatching(x)
    b = nb.MutualMatching(x.permute(0, 1, 4, 5, 2, 3).contiguous()).permute(0, 1, 4, 5, 2, 3)
    assert torch.equal(a, b)
    assert nerr(a, O.mutual_matching(x.cpu())) < 1e-6


@pytest.mark.parametrize("cin,cout,k,dims,bsz", [
    (1, 16, 5, (9, 8, 10, 11), 2), (16, 16, 5, (7, 9, 8, 25), 1), (16, 1, 5, (8, 7, 9, 10), 2),
    (1, 16, 3, (6, 7, 5, 37), 1), (16, 1, 3, (7, 5, 40, 6), 1), (10, 10, 3, (5, 6, 7, 8), 2), (3, 5, 5, (6, 5, 6, 33), 1)])
def test_conv4d_vs_oracle(nb, cuda, cin, cout, k, dims, bsz):
    torch.manual_seed(7)
    x = torch.randn(bsz, cin, *dims)
    layer = O.Conv4d(cin, cout, k)
    w, b = layer.weight.detach(), layer.bias.detach()
    xo, wo, bo = x.clone().requires_grad_(True), w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    yo = O.conv4d(xo, wo, bo, permute_filters=False)
    gy = torch.randn_like(yo)
    rx, rw, rb = torch.autograd.grad(yo, (xo, wo, bo), gy)
    xg, wg, bg = (t.to(cuda).requires_grad_(True) for t in (x, w, b))
    y = nb.conv4d(xg, wg, bg, permute_filters=False)
    assert nerr(y, yo) < 1e-5
    gx, gw, gb = torch.autograd.grad(y, (xg, wg, bg), gy.to(cuda))
    assert nerr(gx, rx) < 1e-4 and nerr(gw, rw) < 1e-4 and nerr(gb, rb) < 1e-4


def test_neigh_consensus_pf_shape_vs_oracle(nb, cuda):
    """PF-Pascal head shape: 25^4 volume, kernels 5,5,5, channels 16,16,1, symmetric."""
    torch.manual_seed(8)
    onc = O.NeighConsensus([5, 5, 5], [16, 16, 1])
    nc = nb.NeighConsensus(True, [5, 5, 5], [16, 16, 1])
    nc.load_state_dict(onc.state_dict())
    x = O.mutual_matching(torch.rand(1, 1, 25, 25, 25, 25))
    with torch.no_grad():
        ref = onc(x)
        out = nc(x.to(cuda))
    assert nerr(out, ref) < 1e-5
    # symmetric-mode property on the GPU path: NC(x^T) == NC(x)^T
    with torch.no_grad():
        outT = nc(x.to(cuda).permute(0, 1, 4, 5, 2, 3).contiguous()).permute(0, 1, 4, 5, 2, 3)
    assert nerr(outT, out) < 1e-5


def test_training_gradients_vs_oracle_batch(nb, cuda):
    from ncnet_b200.train_util import weak_loss_from_features
    torch.manual_seed(9)
    B, C, fs = 3, 64, 9
    fa = O.feature_l2norm(torch.relu(torch.randn(B, C, fs, fs)))
    fb = O.feature_l2norm(torch.relu(torch.randn(B, C, fs, fs)))
    onc = O.NeighConsensus([3, 3, 3], [10, 10, 1]).double()
    lo = O.weak_loss(fa.double(), fb.double(), onc)
    lo.backward()
    net = nb.ImMatchNet.__new__(nb.ImMatchNet)
    torch.nn.Module.__init__(net)
    net.relocalization_k_size, net.half_precision = 0, False
    net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
    net.NeighConsensus = nb.NeighConsensus(True, [3, 3, 3], [10, 10, 1])
    net.NeighConsensus.load_state_dict({k: v.float() for k, v in onc.state_dict().items()})
    lg = weak_loss_from_features(net, fa.to(cuda), fb.to(cuda))
    lg.backward()
    assert abs(float(lg) - float(lo)) < 1e-7
    for (n1, p), (n2, q) in zip(net.NeighConsensus.named_parameters(), onc.named_parameters()):
        assert nerr(p.grad, q.grad) < 2e-3, n1


def test_maxpool4d_vs_oracle_and_ties(nb, cuda):
    torch.manual_seed(10)
    for k, shp in ((2, (8, 6, 10, 14)), (3, (6, 9, 3, 12)), (4, (4, 8, 4, 8))):
        x = torch.randint(0, 5, (1, 1) + shp).float()          # many ties: first window element must win
        ref = O.maxpool4d(x, k)
        out = nb.maxpool4d(x.to(cuda), k)
        assert torch.equal(out[0].cpu(), ref[0])
        for a, b in zip(out[1:], ref[1:]):
            assert torch.equal(a.cpu(), b)
    xh = torch.rand(1, 1, 4, 6, 8, 10).half()
    ref = O.maxpool4d(xh.float(), 2)
    out = nb.maxpool4d(xh.to(cuda), 2)
    assert out[0].dtype == torch.float16 and torch.equal(out[0].cpu().float(), ref[0])
    assert all(torch.equal(a.cpu(), b) for a, b in zip(out[1:], ref[1:]))


def test_corr_to_matches_batch_ties_and_planted(nb, cuda):
    torch.manual_seed(11)
    # batch > 1 (the oracle extends the reference's batch-1 behaviour per item), non-square grids
    vol = torch.rand(3, 1, 5, 7, 6, 4)
    vol[1, 0, :, :, 3, 2] = 0.9       # tied column -> first source cell
    vol[2, 0, 4, 6, :, :] = 0.8       # tied row -> first target cell
    for inv in (False, True):
        for sm in (False, True):
            ref = O.corr_to_matches(vol, do_softmax=sm, return_indices=True, invert_matching_direction=inv)
            out = nb.corr_to_matches(vol.to(cuda), do_softmax=sm, return_indices=True, invert_matching_direction=inv)
            for i, (a, b) in enumerate(zip(out, ref)):
                assert (nerr(a, b) < 2e-6) if i == 4 else torch.equal(a.cpu(), b), (inv, sm, i)
    # planted correspondence at the PF-Pascal size: target cell t matches source cell perm[t]
    n = 625
    perm = torch.randperm(n)
    m = torch.rand(n, n) * 0.1
    m[perm, torch.arange(n)] += 1.0
    out = nb.corr_to_matches(m.reshape(1, 1, 25, 25, 25, 25).to(cuda), do_softmax=True, return_indices=True)
    flat = out[5] * 25 + out[6]
    assert torch.equal(flat.cpu().reshape(-1), perm)


# --------------------------------------------------------------------------------------------- #
# fp16 storage (the InLoc configuration), fp32 accumulation, against the fp32 oracle
# --------------------------------------------------------------------------------------------- #
def test_half_precision_head_vs_fp32_oracle(nb, cuda):
    torch.manual_seed(12)
    fa = O.feature_l2norm(torch.relu(torch.randn(1, 128, 12, 8)))
    fb = O.feature_l2norm(torch.relu(torch.randn(1, 128, 8, 12)))
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
    assert out.dtype == torch.float16
    assert nerr(out.float(), ref) < 5e-3
    # offsets can differ only where fp16 rounding created/removed a tie; demand near-total agreement
    agree = sum((a.cpu() == b).float().mean().item() for a, b in zip(delta, rdelta)) / 4
    assert agree > 0.97
    m = nb.corr_to_matches(out, delta4d=delta, k_size=2, do_softmax=True, scale="positive", return_indices=True)
    assert m[4].dtype == torch.float16 and all(t.shape == (1, 24) for t in m)


# --------------------------------------------------------------------------------------------- #
# error behaviour
# --------------------------------------------------------------------------------------------- #
def test_error_paths(nb, cuda):
    with pytest.raises(RuntimeError, match="even kernel"):
        nb.conv4d(torch.rand(1, 1, 4, 4, 4, 4, device=cuda), torch.rand(1, 1, 4, 4, 4, 4, device=cuda))
    with pytest.raises(RuntimeError, match="divisible"):
        nb.maxpool4d(torch.rand(1, 1, 5, 4, 4, 4, device=cuda), 2)
    with pytest.raises(RuntimeError, match="batch"):
        nb.maxpool4d(torch.rand(2, 1, 4, 4, 4, 4, device=cuda), 2)
    with pytest.raises(RuntimeError):
        nb.MutualMatching(torch.rand(1, 3, 4, 4, 4, 4, device=cuda))
    with pytest.raises(TypeError):
        nb.MutualMatching(torch.rand(1, 1, 4, 4, 4, 4, device=cuda).double())


@pytest.mark.parametrize("shape", [(2, 1, 6, 5, 6, 5), (3, 1, 25, 25, 25, 25), (1, 1, 7, 9, 8, 4)])
def test_fused_match_score_vs_oracle(nb, cuda, shape):
    """ncnet_match_score_fwd/bwd against the oracle's soft-max score (train.py:123-134) and its autograd."""
    from ncnet_b200.functional import match_score_softmax
    torch.manual_seed(13)
    x = torch.rand(*shape) * 3
    xo = x.double().requires_grad_(True)
    so = O.match_score(xo, "softmax")
    (go,) = torch.autograd.grad(so, xo)
    xg = x.to(cuda).requires_grad_(True)
    sg = match_score_softmax(xg)
    (gg,) = torch.autograd.grad(sg * 2.5, xg)
    assert abs(float(sg) - float(so)) < 1e-6 * max(1.0, abs(float(so)))
    assert nerr(gg, go * 2.5) < 1e-4


def test_stacked_weak_loss_equals_two_pass_loss(nb, cuda):
    """One pass of the head over [positives; rolled negatives] with per-item score weights (what bench.py times) gives
    the loss and the gradients of the reference's two passes (train.py:135-139)."""
    from ncnet_b200.train_util import weak_loss_from_features, weak_loss_stacked
    from conftest import make_head
    torch.manual_seed(61)
    fa = O.feature_l2norm(torch.relu(torch.randn(3, 32, 7, 6))).to(cuda)
    fb = O.feature_l2norm(torch.relu(torch.randn(3, 32, 6, 7))).to(cuda)
    net = make_head(nb, [3, 3], [16, 1])
    l2 = weak_loss_from_features(net, fa, fb)
    g2 = torch.autograd.grad(l2, list(net.NeighConsensus.parameters()))
    l1 = weak_loss_stacked(net, fa, torch.roll(fa, -1, 0), fb)
    g1 = torch.autograd.grad(l1, list(net.NeighConsensus.parameters()))
    assert abs(float(l1) - float(l2)) < 1e-7 + 1e-5 * abs(float(l2))
    for a, b in zip(g1, g2):
        assert nerr(a, b) < 1e-4


@pytest.mark.parametrize("dt,softmax", [(torch.float32, True), (torch.float32, False), (torch.float16, True)])
def test_corr_to_matches_long_reduction_split_rows(nb, cuda, dt, softmax):
    """nA >= 1024 source cells with a single batch item: the B->A reduction splits its rows over CTAs and merges them
    through scratch (csrc/pool_matches.cu); indices and coordinates still equal the oracle's, both directions."""
    torch.manual_seed(62)
    shape = (1, 1, 40, 30, 33, 28)
    x = torch.rand(shape).to(dt)
    for t in range(33 * 28):                          # a clear winner per target cell, plus exact ties below it
        x[0, 0, (7 * t) % 40, (11 * t) % 30, t // 28, t % 28] = 2.0 + (t % 5)
    x[0, 0, 3, 4, :, :] = x[0, 0, 5, 6, :, :]
    from ncnet_b200 import _cabi
    assert _cabi.lib().ncnet_corr_to_matches_scratch_bytes(1, 40, 30, 33, 28, 0, _cabi.F16 if dt == torch.float16 else _cabi.F32) > 0
    for inv in (False, True):
        m = nb.corr_to_matches(x.to(cuda), do_softmax=softmax, return_indices=True, invert_matching_direction=inv)
        # fp16 volumes: the reference's soft-max of a half tensor is rounded to half before the arg-max, which merges
        # near-equal candidates into ties (first index wins) -- the oracle gets the half tensor, as the reference would
        mo = O.corr_to_matches(x, do_softmax=softmax, return_indices=True, invert_matching_direction=inv)
        if dt == torch.float16 and softmax:
            # half soft-max values one ulp apart (summation order of the denominator) can split or merge a tie
            same = torch.stack([m[k].cpu() == mo[k] for k in (5, 6, 7, 8)]).all(0).float().mean().item()
            assert same > 0.995, (inv, same)
        else:
            for k in (5, 6, 7, 8):
                assert torch.equal(m[k].cpu(), mo[k]), (inv, k)
            for k in range(4):
                assert torch.equal(m[k].cpu(), mo[k])
        assert nerr(m[4].float(), mo[4].float()) < (2e-3 if dt == torch.float16 else 1e-5)
