"""Drop-in for the evaluation glue around the matching head: ``pck`` / ``pck_metric`` of the reference's
``lib/eval_util.py:12-51`` and the match post-processing that ``eval_inloc.py:151-203`` does inline (on the host,
with numpy) -- here on the device (``csrc/match_post.cu``)."""
from __future__ import annotations

import torch

from . import _cabi
from .functional import _count, _need_cuda, _ptr, _stream, corr_to_matches
from .point_tnf import PointsToPixelCoords, PointsToUnitCoords, bilinearInterpPointTnf


def pck(source_points, warped_points, L_pck, alpha=0.1):
    """Percentage of correct key points per batch item (lib/eval_util.py:12-25)."""
    _need_cuda(source_points, warped_points, L_pck)
    s = source_points.detach().float().contiguous()
    w = warped_points.detach().float().contiguous()
    l = L_pck.detach().float().reshape(-1).contiguous()
    b, _, n = s.shape
    out = torch.empty(b, device=s.device, dtype=torch.float32)
    with torch.cuda.device(s.device):
        _cabi.check(_cabi.lib().ncnet_pck(_ptr(s), _ptr(w), _ptr(l), float(alpha), _ptr(out), b, n, _stream()), "pck")
    _count(1)
    return out


def pck_metric(batch, batch_start_idx, matches, stats, args, use_cuda=True):
    """lib/eval_util.py:28-51: key-point transfer through the matches, then PCK into ``stats``."""
    target_points_norm = PointsToUnitCoords(batch['target_points'], batch['target_im_size'])
    warped_points_norm = bilinearInterpPointTnf(matches, target_points_norm)
    warped_points = PointsToPixelCoords(warped_points_norm, batch['source_im_size'])
    current_batch_size = batch['source_im_size'].size(0)
    indices = range(batch_start_idx, batch_start_idx + current_batch_size)
    pck_batch = pck(batch['source_points'].data, warped_points.data, batch['L_pck'].data)
    stats['point_tnf']['pck'][indices] = pck_batch.unsqueeze(1).cpu().numpy()
    return stats


def inloc_matches(corr4d, delta4d=None, k_size=1, do_softmax=True, matching_both_directions=True,
                  flip_matching_direction=False):
    """What eval_inloc.py:151-203 does with one image pair's volume: matches in one or both directions
    (``scale='positive'``), duplicates removed keeping the highest score, coordinates recentred to cell centres.
    Returns ``(rows, count)``: ``rows`` [N,5] fp32 = (xA, yA, xB, yB, score) with the first ``count`` (a device int32
    scalar) rows valid and the rest zero -- the layout the reference copies into ``matches[0, idx, :Npts, :]``.
    In the single-direction modes the reference removes nothing and keeps corr_to_matches' order; so does this."""
    _need_cuda(corr4d)
    fs1, fs2, fs3, fs4 = corr4d.shape[2:]
    kk = max(int(k_size), 1)
    kw = dict(scale='positive', do_softmax=do_softmax, delta4d=delta4d, k_size=kk)
    if matching_both_directions:
        first = corr_to_matches(corr4d, **kw)
        second = corr_to_matches(corr4d, invert_matching_direction=True, **kw)
    else:
        first = corr_to_matches(corr4d, invert_matching_direction=bool(flip_matching_direction), **kw)
        second = None
    if corr4d.shape[0] != 1:
        raise RuntimeError("inloc_matches: batch size must be 1 (as in eval_inloc.py)")
    a = [t.reshape(-1).float().contiguous() for t in first]
    n1 = a[0].numel()
    dev = corr4d.device
    n = (fs1 * kk, fs2 * kk, fs3 * kk, fs4 * kk)            # recentring divisors for yA, xA, yB, xB
    # single direction (n2 = 0): the reference removes nothing and keeps corr_to_matches' order
    b = [t.reshape(-1).float().contiguous() for t in second] if second is not None else [None] * 5
    n2 = b[0].numel() if second is not None else 0
    lib = _cabi.lib()
    rows = torch.empty(n1 + n2, 5, device=dev, dtype=torch.float32)
    count = torch.empty((), device=dev, dtype=torch.int32)
    scratch = torch.empty(lib.ncnet_matches_unique_scratch_bytes(n1 + n2), device=dev, dtype=torch.uint8)
    with torch.cuda.device(dev):
        _cabi.check(lib.ncnet_matches_sort_unique(*[_ptr(t) for t in a], n1, *[_ptr(t) for t in b], n2, n[0], n[1], n[2], n[3],
                                                  _ptr(rows), _ptr(count), _ptr(scratch), _stream()), "matches_sort_unique")
    _count(4 if n2 else 1)
    return rows, count
