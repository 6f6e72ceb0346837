"""Pin ``oracle/post.py`` to the reference and freeze ``tests/golden/point_tnf.npz``.  TEST INFRASTRUCTURE ONLY.

    python oracle/make_golden_post.py       # build container only (needs /root/reference)

``lib.point_tnf`` imports as is.  ``lib/eval_util.py`` pulls in scikit-image (absent here) at import time, so its
``pck`` function is taken out of the file with ``ast`` and executed unchanged -- still the reference's own code.
"""
from __future__ import annotations

import ast
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("NCNET_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)

from oracle import head as O          # noqa: E402
from oracle import post as P          # noqa: E402
from oracle.make_golden import npz    # noqa: E402


def reference_pck():
    src = open(os.path.join(REF, "lib", "eval_util.py")).read()
    fn = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "pck")
    ns = {"torch": torch, "np": np}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "lib/eval_util.py", "exec"), ns)
    return ns["pck"]


def main():
    sys.path.insert(0, REF)
    import lib.point_tnf as rpt
    sys.path.pop(0)
    rpck = reference_pck()
    torch.manual_seed(4321)
    fs, n = 25, 20
    # matches of a random volume (batch 1, like eval_pf_pascal.py:52-53), B->A direction
    corr = torch.rand(1, 1, fs, fs, fs, fs)
    xA, yA, xB, yB, score = rpt.corr_to_matches(corr, do_softmax=True)
    im_t = torch.tensor([[333.0, 500.0, 3.0]])
    im_s = torch.tensor([[375.0, 480.0, 3.0]])
    tp = torch.stack([torch.rand(1, n) * 499 + 1, torch.rand(1, n) * 332 + 1], 1)      # pixel coords (x, y), 1-based
    tp[0, 0, 0], tp[0, 1, 0] = 1.0, 1.0                      # the corners: edge handling of the cell search
    tp[0, 0, 1], tp[0, 1, 1] = 500.0, 333.0
    tpn = rpt.PointsToUnitCoords(tp, im_t)
    wpn = rpt.bilinearInterpPointTnf((xA, yA, xB, yB), tpn)
    wp = rpt.PointsToPixelCoords(wpn, im_s)
    sp = wp + torch.randn_like(wp) * 25
    sp[:, :, 15:] = -1                                        # trailing invalid key points
    L = torch.tensor([400.0])
    pk = rpck(sp, wp, L)
    pk05 = rpck(sp, wp, L, alpha=0.05)
    for got, ref, what in ((P.points_to_unit_coords(tp, im_t), tpn, "unit"), (P.bilinear_interp_point_tnf((xA, yA, xB, yB), tpn), wpn, "bilinear"),
                           (P.points_to_pixel_coords(wpn, im_s), wp, "pixel"), (P.pck(sp, wp, L), pk, "pck"),
                           (P.pck(sp, wp, L, 0.05), pk05, "pck05")):
        assert torch.equal(got, ref), what
    npz("point_tnf", xA=xA, yA=yA, xB=xB, yB=yB, target_points=tp, target_im_size=im_t, source_im_size=im_s,
        target_points_norm=tpn, warped_points_norm=wpn, warped_points=wp, source_points=sp, L_pck=L, pck=pk, pck05=pk05,
        source="reference lib/point_tnf.py:96-167 and lib/eval_util.py:12-25 (pck extracted with ast, run unchanged)")
    print("oracle/post.py reproduces the reference bit for bit; fixture written")


if __name__ == "__main__":
    main()
