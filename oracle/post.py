"""CPU restatement of what the reference does with matches after ``corr_to_matches``.  TEST INFRASTRUCTURE ONLY.

  * ``bilinear_interp_point_tnf``, ``points_to_unit_coords``, ``points_to_pixel_coords``  -- lib/point_tnf.py:96-167
  * ``pck``                                                                         -- lib/eval_util.py:12-25
  * ``inloc_postprocess``                                                           -- eval_inloc.py:151-203

The first four are pinned against the reference's own functions by ``oracle/make_golden_post.py`` (fixtures
``tests/golden/point_tnf.npz``).  ``inloc_postprocess`` restates script-level code of ``eval_inloc.py`` that cannot be
imported (it is the body of the evaluation loop and indexes with ``torch.cuda.LongTensor``): parity for it is
UNPINNED by a reference run; it uses the same ``torch.sort`` / ``np.unique(axis=1, return_index=True)`` calls.
"""
from __future__ import annotations

import numpy as np
import torch


def normalize_axis(x, L):          # lib/point_tnf.py:6-7
    return (x - 1 - (L - 1) / 2) * 2 / (L - 1)


def unnormalize_axis(x, L):        # lib/point_tnf.py:9-10
    return x * (L - 1) / 2 + 1 + (L - 1) / 2


def points_to_unit_coords(P, im_size):       # lib/point_tnf.py:151-158
    h, w = im_size[:, 0], im_size[:, 1]
    out = P.clone()
    out[:, 0, :] = normalize_axis(P[:, 0, :], w.unsqueeze(1).expand_as(P[:, 0, :]))
    out[:, 1, :] = normalize_axis(P[:, 1, :], h.unsqueeze(1).expand_as(P[:, 1, :]))
    return out


def points_to_pixel_coords(P, im_size):      # lib/point_tnf.py:160-167
    h, w = im_size[:, 0], im_size[:, 1]
    out = P.clone()
    out[:, 0, :] = unnormalize_axis(P[:, 0, :], w.unsqueeze(1).expand_as(P[:, 0, :]))
    out[:, 1, :] = unnormalize_axis(P[:, 1, :], h.unsqueeze(1).expand_as(P[:, 1, :]))
    return out


def bilinear_interp_point_tnf(matches, target_points_norm):     # lib/point_tnf.py:96-148
    xA, yA, xB, yB = matches
    fs = int(np.sqrt(xB.shape[-1]))
    b, _, n = target_points_norm.shape
    grid = torch.from_numpy(np.linspace(-1, 1, fs)).float().view(1, fs, 1)
    tx, ty = target_points_norm[:, 0:1, :], target_points_norm[:, 1:2, :]
    xm = ((tx - grid) > 0).long().sum(1, keepdim=True) - 1
    ym = ((ty - grid) > 0).long().sum(1, keepdim=True) - 1
    xm, ym = xm.clamp_min(0), ym.clamp_min(0)
    xp, yp = xm + 1, ym + 1
    flat = lambda x, y: (y * fs + x).view(-1)
    mm, pp, pm, mp = flat(xm, ym), flat(xp, yp), flat(xp, ym), flat(xm, yp)
    # indices address the flattened [B*fs*fs] arrays without a batch offset (reference quirk, batch 1 in practice)
    pick = lambda idx, X, Y: torch.cat((X.reshape(-1)[idx].view(b, 1, n), Y.reshape(-1)[idx].view(b, 1, n)), 1)
    area = lambda P: (target_points_norm - P).abs()[:, 0, :] * (target_points_norm - P).abs()[:, 1, :]
    f_pp, f_mm = area(pick(mm, xB, yB)), area(pick(pp, xB, yB))
    f_mp, f_pm = area(pick(pm, xB, yB)), area(pick(mp, xB, yB))
    f_pp, f_mm, f_mp, f_pm = (f.unsqueeze(1) for f in (f_pp, f_mm, f_mp, f_pm))
    Q = lambda idx: pick(idx, xA, yA)
    return (Q(mm) * f_mm + Q(pp) * f_pp + Q(mp) * f_mp + Q(pm) * f_pm) / (f_pp + f_mm + f_mp + f_pm)


def pck(source_points, warped_points, L_pck, alpha=0.1):        # lib/eval_util.py:12-25
    b = source_points.size(0)
    out = torch.zeros(b)
    for i in range(b):
        ps, pw = source_points[i], warped_points[i]
        npts = int(((ps[0] != -1) & (ps[1] != -1)).sum())
        d = ((ps[:, :npts] - pw[:, :npts]) ** 2).sum(0) ** 0.5
        out[i] = (d <= L_pck[i].expand_as(d) * alpha).float().mean()
    return out


def inloc_postprocess(first, second, fs, k_size):               # eval_inloc.py:151-203
    """first / second: (xA,yA,xB,yB,score) tuples of [1,N] tensors from corr_to_matches (second may be None);
    fs = (fs1,fs2,fs3,fs4) of the volume.  Returns an [N,5] float32 array of (xA,yA,xB,yB,score) rows."""
    if second is not None:
        cols = [torch.cat((a, b), 1) for a, b in zip(first, second)]
        order = torch.sort(-cols[4])[1].squeeze()
        cols = [c.squeeze()[order].unsqueeze(0) for c in cols]
        coords = np.concatenate([c.numpy() for c in cols[:4]], 0)
        _, uniq = np.unique(coords, axis=1, return_index=True)
        cols = [c.squeeze()[torch.from_numpy(uniq)].unsqueeze(0) for c in cols]
    else:
        cols = list(first)
    xA, yA, xB, yB, score = cols
    kk = max(k_size, 1)
    f1, f2, f3, f4 = (f * kk for f in fs)
    yA = yA * (f1 - 1) / f1 + 0.5 / f1
    xA = xA * (f2 - 1) / f2 + 0.5 / f2
    yB = yB * (f3 - 1) / f3 + 0.5 / f3
    xB = xB * (f4 - 1) / f4 + 0.5 / f4
    return np.stack([t.reshape(-1).float().numpy() for t in (xA, yA, xB, yB, score)], 1)
