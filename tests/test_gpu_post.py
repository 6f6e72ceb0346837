"""GPU parity of the match post-processing kernels (csrc/match_post.cu): key-point transfer and PCK against the
fixture the reference's own functions produced (bit-exact), InLoc duplicate removal against oracle/post.py
(bit-exact rows, same order)."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import head as O
from oracle import post as P

pytestmark = pytest.mark.gpu


def test_point_transfer_and_pck_golden(cuda):
    import ncnet_b200.point_tnf as pt
    from ncnet_b200 import eval_util as eu
    g = load_golden("point_tnf", cuda)
    tpn = pt.PointsToUnitCoords(g["target_points"], g["target_im_size"])
    assert torch.equal(tpn, g["target_points_norm"])
    wpn = pt.bilinearInterpPointTnf((g["xA"], g["yA"], g["xB"], g["yB"]), tpn)
    assert torch.equal(wpn, g["warped_points_norm"])
    wp = pt.PointsToPixelCoords(wpn, g["source_im_size"])
    assert torch.equal(wp, g["warped_points"])
    assert torch.equal(eu.pck(g["source_points"], wp, g["L_pck"]), g["pck"])
    assert torch.equal(eu.pck(g["source_points"], wp, g["L_pck"], alpha=0.05), g["pck05"])
    # pck_metric: the reference's glue (lib/eval_util.py:28-51) writes into the stats array
    stats = {"point_tnf": {"pck": np.zeros((3, 1))}}
    batch = {"source_im_size": g["source_im_size"], "target_im_size": g["target_im_size"], "source_points": g["source_points"],
             "target_points": g["target_points"], "L_pck": g["L_pck"]}
    eu.pck_metric(batch, 1, (g["xA"], g["yA"], g["xB"], g["yB"]), stats, None)
    assert stats["point_tnf"]["pck"][1, 0] == g["pck"].item() and stats["point_tnf"]["pck"][0, 0] == 0


def test_point_transfer_many_points_vs_oracle(cuda):
    import ncnet_b200.point_tnf as pt
    torch.manual_seed(51)
    fs = 18
    corr = torch.rand(1, 1, fs, fs, fs, fs)
    m = O.corr_to_matches(corr, do_softmax=False)
    tp = torch.rand(1, 2, 777) * 2.2 - 1.1              # some points outside [-1,1] on the low side and slightly above
    tp = tp.clamp(max=1.0)                               # above +1 the reference indexes out of range
    ref = P.bilinear_interp_point_tnf(m[:4], tp)
    got = pt.bilinearInterpPointTnf(tuple(t.to(cuda) for t in m[:4]), tp.to(cuda))
    assert torch.equal(got.cpu(), ref)


@pytest.mark.parametrize("shape,k,dt", [((1, 1, 6, 7, 5, 8), 1, torch.float32), ((1, 1, 12, 9, 10, 11), 2, torch.float16),
                                         ((1, 1, 25, 25, 25, 25), 1, torch.float32)])
def test_inloc_matches_vs_oracle(cuda, shape, k, dt):
    """Both directions + duplicate removal + recentring on the device == eval_inloc.py:151-203 restated on the host."""
    import ncnet_b200 as nb
    from ncnet_b200 import eval_util as eu
    torch.manual_seed(52)
    corr = torch.rand(shape).to(dt)
    fs = shape[2:]
    for t in range(min(fs[2] * fs[3], fs[0] * fs[1]) // 2):      # plant mutual best matches: duplicates across directions
        corr[0, 0, (3 * t) % fs[0], (5 * t) % fs[1], t // fs[3] % fs[2], t % fs[3]] = 3.0 + t
    delta = None
    cg = corr.to(cuda)
    if k > 1:
        hi = torch.rand(1, 1, *(f * k for f in fs)).to(dt)
        cg, *dl = nb.maxpool4d(hi.to(cuda), k)
        delta = tuple(dl)
        corr = cg.cpu()
    dcpu = None if delta is None else tuple(d.cpu() for d in delta)
    a = O.corr_to_matches(corr.float(), delta4d=dcpu, k_size=k, do_softmax=True, scale="positive")
    b = O.corr_to_matches(corr.float(), delta4d=dcpu, k_size=k, do_softmax=True, scale="positive", invert_matching_direction=True)
    # the device path scores in the volume's dtype (fp16 heads give fp16 scores, like the reference): feed the oracle
    # the device's own corr_to_matches columns so this test isolates the post-processing
    ga = nb.corr_to_matches(cg, delta4d=delta, k_size=k, do_softmax=True, scale="positive")
    gb = nb.corr_to_matches(cg, delta4d=delta, k_size=k, do_softmax=True, scale="positive", invert_matching_direction=True)
    ref = P.inloc_postprocess(tuple(t.cpu().float() for t in ga), tuple(t.cpu().float() for t in gb), fs, k)
    rows, count = eu.inloc_matches(cg, delta4d=delta, k_size=k, do_softmax=True, matching_both_directions=True)
    n = int(count.item())
    assert n == ref.shape[0] and n < rows.shape[0]           # the planted mutual matches were de-duplicated
    assert np.array_equal(rows[:n].cpu().numpy(), ref)
    assert rows[n:].abs().sum().item() == 0
    # coordinates agree with the host oracle's own matches too (fp32 volumes: exact)
    if dt == torch.float32:
        ref2 = P.inloc_postprocess(a, b, fs, k)
        assert np.array_equal(rows[:n, :4].cpu().numpy(), ref2[:, :4])
    for flip in (False, True):
        rows1, c1 = eu.inloc_matches(cg, delta4d=delta, k_size=k, do_softmax=True, matching_both_directions=False,
                                     flip_matching_direction=flip)
        one = gb if flip else ga
        ref1 = P.inloc_postprocess(tuple(t.cpu().float() for t in one), None, fs, k)
        assert int(c1.item()) == ref1.shape[0] and np.array_equal(rows1.cpu().numpy(), ref1)
