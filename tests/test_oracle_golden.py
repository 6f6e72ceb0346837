"""CPU: the oracle restatement reproduces the fixtures the reference itself produced
(oracle/make_golden.py).  These fixtures are the pin for every parity claim."""
import torch

from conftest import load_golden, nerr
from oracle import head as O


def test_l2norm():
    g = load_golden("feature_l2norm")
    assert torch.equal(O.feature_l2norm(g["x"]), g["y"])


def test_feature_correlation():
    g = load_golden("feature_correlation")
    assert nerr(O.feature_correlation(g["fa"], g["fb"], "4D", False), g["corr4d"]) < 1e-6
    assert nerr(O.feature_correlation(g["fa"], g["fb"], "4D", True), g["corr4d_norm"]) < 1e-6
    assert nerr(O.feature_correlation(g["fa"], g["fb_sq"], "3D", True), g["corr3d_norm"]) < 1e-6


def test_mutual_matching_and_grad():
    g = load_golden("mutual_matching")
    x = g["x"].clone().requires_grad_(True)
    y = O.mutual_matching(x)
    assert nerr(y, g["y"]) < 1e-7
    (gx,) = torch.autograd.grad(y, x, g["grad_y"])
    assert nerr(gx, g["grad_x"]) < 1e-6


def test_conv4d_cases():
    for tag in ("k3", "k5", "k5c16"):
        g = load_golden("conv4d_" + tag)
        x = g["x"].clone().requires_grad_(True)
        w = g["weight"].clone().requires_grad_(True)
        b = g["bias"].clone().requires_grad_(True)
        y = O.conv4d(x, w, b, permute_filters=False)
        assert nerr(y, g["y"]) < 2e-6, tag
        gx, gw, gb = torch.autograd.grad(y, (x, w, b), g["grad_y"])
        assert nerr(gx, g["grad_x"]) < 1e-5 and nerr(gw, g["grad_weight"]) < 1e-5 and nerr(gb, g["grad_bias"]) < 1e-5, tag
        # definition-level check on the smallest case
        if tag == "k3":
            yd = O.conv4d_direct(g["x"], O.from_prepermuted(g["weight"]), g["bias"])
            assert nerr(yd, g["y"]) < 2e-6


def test_neigh_consensus():
    for tag in ("k3", "k5", "k3_10"):
        g = load_golden("neigh_consensus_" + tag)
        n = len(g["kernel_sizes"])
        ws, bs = [g[f"w{i}"] for i in range(n)], [g[f"b{i}"] for i in range(n)]
        assert nerr(O.neigh_consensus(g["x"], ws, bs, True), g["y_sym"]) < 5e-6, tag
        assert nerr(O.neigh_consensus(g["x"], ws, bs, False), g["y_one"]) < 5e-6, tag


def test_symmetric_branch_is_transposed_filter_conv():
    """conv_W(x^T)^T == conv_{W^T}(x) with W^T[o,c,a,b,c',d'] = W[o,c,c',d',a,b]."""
    torch.manual_seed(0)
    x = torch.randn(1, 2, 4, 5, 3, 6, dtype=torch.float64)
    w = torch.randn(3, 2, 3, 3, 3, 3, dtype=torch.float64)
    lhs = O.conv4d(x.permute(0, 1, 4, 5, 2, 3), w, None).permute(0, 1, 4, 5, 2, 3)
    rhs = O.conv4d(x, w.permute(0, 1, 4, 5, 2, 3).contiguous(), None)
    assert nerr(lhs, rhs) < 1e-12


def test_maxpool4d():
    g = load_golden("maxpool4d")
    v, mi, mj, mk, ml = O.maxpool4d(g["x"], int(g["k_size"]))
    assert torch.equal(v, g["value"])
    for a, b in ((mi, "max_i"), (mj, "max_j"), (mk, "max_k"), (ml, "max_l")):
        assert torch.equal(a, g[b])


def test_corr_to_matches():
    g = load_golden("corr_to_matches")
    names = ("xA", "yA", "xB", "yB", "score", "iA", "jA", "iB", "jB")
    for inv in (0, 1):
        for sm in (0, 1):
            for sc in ("centered", "positive"):
                out = O.corr_to_matches(g["vol"], do_softmax=bool(sm), scale=sc, return_indices=True,
                                        invert_matching_direction=bool(inv))
                for nm, t in zip(names, out):
                    ref = g[f"inv{inv}_sm{sm}_{sc}_{nm}"]
                    if t.dtype == torch.int64:
                        assert torch.equal(t, ref), (inv, sm, sc, nm)
                    else:
                        assert nerr(t, ref) < 1e-6, (inv, sm, sc, nm)
        deltas = (g["d_i"], g["d_j"], g["d_k"], g["d_l"])
        out = O.corr_to_matches(g["pooled"], delta4d=deltas, k_size=2, do_softmax=True, scale="positive",
                                return_indices=True, invert_matching_direction=bool(inv))
        for nm, t in zip(names, out):
            ref = g[f"reloc_inv{inv}_{nm}"]
            assert (torch.equal(t, ref) if t.dtype == torch.int64 else nerr(t, ref) < 1e-6), (inv, nm)


def test_head_and_weak_loss():
    g = load_golden("head_weak_loss")
    nc = O.NeighConsensus([5, 5, 5], [16, 16, 1])
    with torch.no_grad():
        for i in range(3):
            nc.conv[2 * i].weight.copy_(g[f"w{i}"]), nc.conv[2 * i].bias.copy_(g[f"b{i}"])
    assert nerr(O.head_forward(g["fa"], g["fb"], nc), g["head_pos"]) < 1e-5
    loss = O.weak_loss(g["fa"], g["fb"], nc)
    assert abs(float(loss) - float(g["loss"])) < 5e-8


def test_post_oracle_reproduces_reference_point_transfer_and_pck():
    """oracle/post.py against the reference's own bilinearInterpPointTnf / PointsTo*Coords / pck outputs."""
    from oracle import post as P
    g = load_golden("point_tnf")
    tpn = P.points_to_unit_coords(g["target_points"], g["target_im_size"])
    assert torch.equal(tpn, g["target_points_norm"])
    wpn = P.bilinear_interp_point_tnf((g["xA"], g["yA"], g["xB"], g["yB"]), tpn)
    assert torch.equal(wpn, g["warped_points_norm"])
    assert torch.equal(P.points_to_pixel_coords(wpn, g["source_im_size"]), g["warped_points"])
    assert torch.equal(P.pck(g["source_points"], g["warped_points"], g["L_pck"]), g["pck"])
    assert torch.equal(P.pck(g["source_points"], g["warped_points"], g["L_pck"], 0.05), g["pck05"])


def test_inloc_postprocess_oracle_removes_duplicates_keeping_best_score():
    from oracle import post as P
    torch.manual_seed(5)
    corr = torch.rand(1, 1, 4, 5, 6, 3)
    corr[0, 0, 1, 2, 3, 1] = 5.0                       # a mutual best match: found in both directions
    a = O.corr_to_matches(corr, do_softmax=True, scale="positive")
    b = O.corr_to_matches(corr, do_softmax=True, scale="positive", invert_matching_direction=True)
    rows = P.inloc_postprocess(a, b, (4, 5, 6, 3), 1)
    assert rows.shape[1] == 5 and rows.shape[0] < 18 + 20
    assert len({tuple(r[:4]) for r in rows}) == rows.shape[0]
    coords = [tuple(r[:4]) for r in rows]
    assert coords == sorted(coords)                    # np.unique order: lexicographic in (xA, yA, xB, yB)
