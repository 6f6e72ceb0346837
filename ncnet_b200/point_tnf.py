"""Drop-in for the hot-path part of the reference's ``lib/point_tnf.py``: same names and arguments, CUDA tensors in,
CUDA tensors out, every arithmetic step on this library's kernels (``csrc/pool_matches.cu``, ``csrc/match_post.cu``)."""
from __future__ import annotations

import numpy as np
import torch

from . import _cabi
from .functional import _count, _need_cuda, _ptr, _stream, corr_to_matches  # noqa: F401   (lib/point_tnf.py:12-80)


def normalize_axis(x, L):
    """pixel coordinate (1-based) -> [-1, 1]   (lib/point_tnf.py:6-7)"""
    return (x - 1 - (L - 1) / 2) * 2 / (L - 1)


def unnormalize_axis(x, L):
    """[-1, 1] -> pixel coordinate (1-based)   (lib/point_tnf.py:9-10)"""
    return x * (L - 1) / 2 + 1 + (L - 1) / 2


def _axis_transform(P: torch.Tensor, im_size: torch.Tensor, to_unit: bool) -> torch.Tensor:
    _need_cuda(P, im_size)
    if P.dim() != 3 or P.shape[1] != 2:
        raise RuntimeError(f"points: expected [B,2,N], got {tuple(P.shape)}")
    p = P.detach().float().contiguous()
    s = im_size.detach().float().contiguous()
    if s.dim() != 2 or s.shape[0] != p.shape[0] or s.shape[1] < 2:
        raise RuntimeError(f"im_size: expected [B,>=2] rows (h, w, ...), got {tuple(im_size.shape)}")
    out = torch.empty_like(p)
    with torch.cuda.device(p.device):
        _cabi.check(_cabi.lib().ncnet_points_axis_transform(_ptr(p), _ptr(s), s.shape[1], _ptr(out), p.shape[0], p.shape[2],
                                                            int(to_unit), _stream()), "points_axis_transform")
    _count(1)
    return out


def PointsToUnitCoords(P, im_size):
    """pixel -> unit coordinates, x by the width and y by the height (lib/point_tnf.py:151-158)"""
    return _axis_transform(P, im_size, True)


def PointsToPixelCoords(P, im_size):
    """unit -> pixel coordinates (lib/point_tnf.py:160-167)"""
    return _axis_transform(P, im_size, False)


_GRID_CACHE = {}


def bilinearInterpPointTnf(matches, target_points_norm):
    """Transfers target key points to the source image by bilinear interpolation of the match grid
    (lib/point_tnf.py:96-148).  ``matches`` = (xA, yA, xB, yB) from ``corr_to_matches`` (B->A direction)."""
    xA, yA, xB, yB = matches
    _need_cuda(xA, yA, xB, yB, target_points_norm)
    fs = int(np.sqrt(xB.shape[-1]))
    b, _, n = target_points_norm.shape
    key = (fs, str(xB.device))
    grid = _GRID_CACHE.get(key)
    if grid is None:
        grid = _GRID_CACHE[key] = torch.from_numpy(np.linspace(-1, 1, fs)).float().to(xB.device)
    cols = [t.detach().float().contiguous() for t in (xA, yA, xB, yB)]
    tp = target_points_norm.detach().float().contiguous()
    out = torch.empty_like(tp)
    with torch.cuda.device(tp.device):
        _cabi.check(_cabi.lib().ncnet_bilinear_point_tnf(*[_ptr(c) for c in cols], _ptr(grid), fs, _ptr(tp), _ptr(out), b, n,
                                                         _stream()), "bilinear_point_tnf")
    _count(1)
    return out
