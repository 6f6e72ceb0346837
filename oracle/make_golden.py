"""Pin the oracle to the reference and freeze golden vectors.

TEST INFRASTRUCTURE ONLY.  Run in the build container, where the reference
checkout exists:

    python oracle/make_golden.py            # check + (re)write tests/golden/*.npz

It imports the reference's own modules from /root/reference (read-only) with
one constructor shim (``_ConvNd.__init__`` gained a ``padding_mode`` argument
after torch 0.4), runs them on seeded inputs, asserts that ``oracle.head``
reproduces every output, and stores the REFERENCE outputs (plus the inputs) as
small ``.npz`` fixtures.  The GPU box has no /root/reference; tests there read
only the fixtures.

Where the reference cannot run on this torch the fixture says
so in its ``source`` field and is produced from the documented minimal
deviation instead:
  * maxpool4d offsets: reference code run as is, offsets cast ``.long()``
    (they come back as integral floats on torch >= 1.6);
  * conv4d gradients: reference autograd raises, so gradients come from the
    definition-level ``conv4d_direct`` in float64.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("NCNET_REFERENCE", "/root/reference")
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)

from oracle import head as O  # noqa: E402


def import_reference():
    from torch.nn.modules import conv as _conv
    orig = _conv._ConvNd.__init__

    def shim(self, *a, **kw):
        if len(a) == 10 and "padding_mode" not in kw:
            a = a + ("zeros",)
        return orig(self, *a, **kw)

    _conv._ConvNd.__init__ = shim
    sys.path.insert(0, REF)
    import lib.conv4d as rconv
    import lib.model as rmodel
    import lib.point_tnf as rpt
    sys.path.pop(0)
    return rmodel, rconv, rpt


def npz(name, **arrays):
    os.makedirs(GOLD, exist_ok=True)
    out = {}
    for k, v in arrays.items():
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
        out[k] = np.asarray(v)
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    print("wrote", name, {k: tuple(v.shape) for k, v in out.items() if hasattr(v, "shape") and v.ndim})


def close(a, b, tol=0.0, what=""):
    a, b = torch.as_tensor(a), torch.as_tensor(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    if a.dtype in (torch.int64, torch.int32, torch.bool):
        assert torch.equal(a, b.to(a.dtype)), what
        return
    err = (a.double() - b.double()).abs().max().item() if a.numel() else 0.0
    ref = max(b.double().abs().max().item(), 1e-30) if b.numel() else 1.0
    assert err <= tol * ref + 0.0 if tol == 0 else err <= tol * ref, f"{what}: err {err:.3e} ref {ref:.3e}"
    return err / ref


def main():
    rmodel, rconv, rpt = import_reference()
    torch.manual_seed(1234)
    torch.set_num_threads(max(1, os.cpu_count() or 1))

    # ---- featureL2Norm ----------------------------------------------------
    x = torch.randn(2, 8, 3, 4)
    y = rmodel.featureL2Norm(x)
    close(O.feature_l2norm(x), y, 0.0, "l2norm")
    npz("feature_l2norm", x=x, y=y, source="reference lib/model.py:14-17")

    # ---- FeatureCorrelation -------------------------------------------------
    fa = torch.relu(torch.randn(2, 16, 3, 4))
    fb = torch.relu(torch.randn(2, 16, 5, 2))
    c4 = rmodel.FeatureCorrelation(shape="4D", normalization=False)(fa, fb)
    c4n = rmodel.FeatureCorrelation(shape="4D", normalization=True)(fa, fb)
    fb_sq = torch.relu(torch.randn(2, 16, 3, 4))
    c3 = rmodel.FeatureCorrelation(shape="3D", normalization=True)(fa, fb_sq)
    close(O.feature_correlation(fa, fb, "4D", False), c4, 0.0, "corr4d")
    close(O.feature_correlation(fa, fb, "4D", True), c4n, 0.0, "corr4d norm")
    close(O.feature_correlation(fa, fb_sq, "3D", True), c3, 1e-6, "corr3d norm")
    npz("feature_correlation", fa=fa, fb=fb, fb_sq=fb_sq, corr4d=c4, corr4d_norm=c4n, corr3d_norm=c3,
        source="reference lib/model.py:89-120")

    # ---- MutualMatching (+ reference autograd) ------------------------------
    v = torch.rand(2, 1, 3, 4, 5, 2)
    v[0, 0, 1, 2, 3, 1] = v.max() + 0.5           # a clear mutual maximum
    v[1, 0, 0, 0, :, :] = 0.25                     # a fully tied row: first index wins
    vr = v.clone().requires_grad_(True)
    mm = rmodel.MutualMatching(vr)
    g = torch.randn_like(mm)
    (gv,) = torch.autograd.grad(mm, vr, g)
    close(O.mutual_matching(v), mm, 0.0, "mutual matching")
    vo = v.clone().requires_grad_(True)
    (gvo,) = torch.autograd.grad(O.mutual_matching(vo), vo, g)
    close(gvo, gv, 1e-6, "mutual matching grad")
    npz("mutual_matching", x=v, y=mm, grad_y=g, grad_x=gv, source="reference lib/model.py:155-175 + autograd")

    # ---- conv4d / Conv4d ----------------------------------------------------
    cases = [("k3", (2, 3, 5, 4, 6, 3), 4, 3), ("k5", (1, 1, 6, 5, 7, 6), 2, 5), ("k5c16", (1, 16, 6, 6, 5, 7), 16, 5)]
    for tag, shp, cout, k in cases:
        xin = torch.randn(*shp)
        m = rconv.Conv4d(shp[1], cout, k)
        wt, bs = m.weight.detach().clone(), m.bias.detach().clone()
        assert tuple(wt.shape) == (k, cout, shp[1], k, k, k)
        with torch.no_grad():
            yref = m(xin)
            yref_nop = rconv.conv4d(xin, O.from_prepermuted(wt), bias=bs, permute_filters=True)
        yo = O.conv4d(xin, wt, bs, permute_filters=False)
        e = close(yo, yref, 2e-6, "conv4d " + tag)
        close(yref_nop, yref, 0.0, "conv4d permute_filters " + tag)
        # definition-level cross check + gradients in float64
        xd = xin.double().requires_grad_(True)
        wd = O.from_prepermuted(wt).double().requires_grad_(True)
        bd = bs.double().requires_grad_(True)
        yd = O.conv4d_direct(xd, wd, bd)
        close(yd.float(), yref, 2e-6, "conv4d direct " + tag)
        gy = torch.randn_like(yref)
        gx, gw, gb = torch.autograd.grad(yd, (xd, wd, bd), gy.double())
        # oracle autograd agrees with the definition
        xo = xin.double().requires_grad_(True)
        wo = wt.double().requires_grad_(True)
        bo = bs.double().requires_grad_(True)
        gxo, gwo, gbo = torch.autograd.grad(O.conv4d(xo, wo, bo, permute_filters=False), (xo, wo, bo), gy.double())
        close(gxo, gx, 1e-10, "conv4d gx"), close(O.from_prepermuted(gwo), gw, 1e-10, "conv4d gw"), close(gbo, gb, 1e-10, "gb")
        npz("conv4d_" + tag, x=xin, weight=wt, bias=bs, y=yref, grad_y=gy, grad_x=gx.float(),
            grad_weight=O.to_prepermuted(gw).float(), grad_bias=gb.float(),
            source="y: reference lib/conv4d.py Conv4d.forward; grads: float64 definition (reference autograd raises on torch>=1.x)")
        print("  conv4d", tag, "oracle-vs-reference rel err", e)

    # ---- NeighConsensus (symmetric and not) ---------------------------------
    for tag, shp, ks, chs in [("k3", (2, 1, 5, 4, 6, 3), [3, 3], [4, 1]), ("k5", (1, 1, 6, 6, 6, 6), [5, 5, 5], [16, 16, 1]),
                              ("k3_10", (1, 1, 4, 5, 4, 5), [3, 3, 3], [10, 10, 1])]:
        nc = rmodel.NeighConsensus(use_cuda=False, kernel_sizes=ks, channels=chs)
        ws = [m.weight.detach().clone() for m in nc.conv if isinstance(m, rconv.Conv4d)]
        bs = [m.bias.detach().clone() for m in nc.conv if isinstance(m, rconv.Conv4d)]
        xin = torch.rand(*shp)
        with torch.no_grad():
            ysym = nc(xin.clone())
            nc.symmetric_mode = False
            yone = nc(xin.clone())
        close(O.neigh_consensus(xin, ws, bs, True), ysym, 3e-6, "nc sym " + tag)
        close(O.neigh_consensus(xin, ws, bs, False), yone, 3e-6, "nc one " + tag)
        # gradients of the symmetric stack from the oracle in float64 (pinned above layer by layer)
        xd = xin.double().requires_grad_(True)
        wd = [w.double().requires_grad_(True) for w in ws]
        bd = [b.double().requires_grad_(True) for b in bs]
        yd = O.neigh_consensus(xd, wd, bd, True)
        gy = torch.randn_like(ysym)
        grads = torch.autograd.grad(yd, [xd] + wd + bd, gy.double())
        arrs = dict(x=xin, y_sym=ysym, y_one=yone, grad_y=gy, grad_x=grads[0].float())
        for i, (w, b) in enumerate(zip(ws, bs)):
            arrs[f"w{i}"], arrs[f"b{i}"] = w, b
            arrs[f"grad_w{i}"], arrs[f"grad_b{i}"] = grads[1 + i].float(), grads[1 + len(ws) + i].float()
        npz("neigh_consensus_" + tag, kernel_sizes=np.array(ks), channels=np.array(chs),
            source="y: reference lib/model.py NeighConsensus.forward; grads: oracle float64", **arrs)

    # ---- maxpool4d ----------------------------------------------------------
    hres = torch.rand(1, 1, 4, 6, 4, 6)
    hres[0, 0, 0:2, 0:2, 0:2, 0:2] = 0.5           # a fully tied window -> offsets (0,0,0,0)
    hres[0, 0, 2, 4, 2, 5] = 2.0
    hres[0, 0, 3, 4, 2, 5] = 2.0                   # tie inside a window: first in (i,j,k,l) order
    pv, pi, pj, pk, pl = rmodel.maxpool4d(hres, k_size=2)
    ov = O.maxpool4d(hres, 2)
    close(ov[0], pv, 0.0, "maxpool value")
    for a, bref, nm in zip(ov[1:], (pi, pj, pk, pl), "ijkl"):
        close(a, bref.long(), 0.0, "maxpool idx " + nm)
    npz("maxpool4d", x=hres, value=pv, max_i=pi.long(), max_j=pj.long(), max_k=pk.long(), max_l=pl.long(),
        k_size=2, source="reference lib/model.py:177-191 (offsets cast .long(); integral floats on torch>=1.6)")

    # ---- corr_to_matches ----------------------------------------------------
    # batch 1: the reference's `.view(-1)` on an expanded index raises for B>1
    vol = torch.rand(1, 1, 3, 4, 5, 2)
    vol[0, 0, :, :, 2, 1] = 0.7                    # tied column: first source index wins
    vol[0, 0, 1, 3, :, :] = 0.6                    # tied row: first target index wins
    out = {}
    for inv in (False, True):
        for sm in (False, True):
            for sc in ("centered", "positive"):
                r = rpt.corr_to_matches(vol, do_softmax=sm, scale=sc, return_indices=True, invert_matching_direction=inv)
                o = O.corr_to_matches(vol, do_softmax=sm, scale=sc, return_indices=True, invert_matching_direction=inv)
                for i, (a, bref) in enumerate(zip(o, r)):
                    close(a, bref, 0.0 if i >= 5 or not sm else 1e-6, f"c2m inv={inv} sm={sm} {sc} #{i}")
                key = f"inv{int(inv)}_sm{int(sm)}_{sc}"
                for nm, t in zip(("xA", "yA", "xB", "yB", "score", "iA", "jA", "iB", "jB"), r):
                    out[f"{key}_{nm}"] = t
    # relocalisation path, batch 1
    pooled = pv
    deltas = (pi.long(), pj.long(), pk.long(), pl.long())
    for inv in (False, True):
        r = rpt.corr_to_matches(pooled, delta4d=deltas, k_size=2, do_softmax=True, scale="positive",
                                return_indices=True, invert_matching_direction=inv)
        o = O.corr_to_matches(pooled, delta4d=deltas, k_size=2, do_softmax=True, scale="positive",
                              return_indices=True, invert_matching_direction=inv)
        for i, (a, bref) in enumerate(zip(o, r)):
            close(a, bref, 0.0 if i >= 5 else 1e-6, f"c2m reloc inv={inv} #{i}")
        for nm, t in zip(("xA", "yA", "xB", "yB", "score", "iA", "jA", "iB", "jB"), r):
            out[f"reloc_inv{int(inv)}_{nm}"] = t
    npz("corr_to_matches", vol=vol, pooled=pooled, d_i=deltas[0], d_j=deltas[1], d_k=deltas[2], d_l=deltas[3],
        source="reference lib/point_tnf.py:12-80", **out)

    # ---- whole head + weak loss ---------------------------------------------
    torch.manual_seed(7)
    B, C, fs = 2, 32, 6
    fa = rmodel.featureL2Norm(torch.relu(torch.randn(B, C, fs, fs)))
    fb = rmodel.featureL2Norm(torch.relu(torch.randn(B, C, fs, fs)))
    nc = rmodel.NeighConsensus(use_cuda=False, kernel_sizes=[5, 5, 5], channels=[16, 16, 1])
    ws = [m.weight.detach().clone() for m in nc.conv if isinstance(m, rconv.Conv4d)]
    bs = [m.bias.detach().clone() for m in nc.conv if isinstance(m, rconv.Conv4d)]

    def ref_head(a, b):
        c = rmodel.FeatureCorrelation(shape="4D", normalization=False)(a, b)
        c = rmodel.MutualMatching(c)
        c = nc(c)
        return rmodel.MutualMatching(c)

    with torch.no_grad():
        hp = ref_head(fa, fb)
        hn = ref_head(torch.roll(fa, -1, 0), fb)

    def ref_score(corr4d):        # train.py:123-134 verbatim semantics
        bsz, f = corr4d.size(0), corr4d.size(2)
        nb = torch.nn.functional.softmax(corr4d.view(bsz, f * f, f, f), 1)
        na = torch.nn.functional.softmax(corr4d.view(bsz, f, f, f * f).permute(0, 3, 1, 2), 1)
        return torch.mean(torch.max(na, dim=1)[0] + torch.max(nb, dim=1)[0]) / 2

    loss_ref = ref_score(hn) - ref_score(hp)
    onc = O.NeighConsensus([5, 5, 5], [16, 16, 1])
    with torch.no_grad():
        for i, (w, b) in enumerate(zip(ws, bs)):
            onc.conv[2 * i].weight.copy_(w), onc.conv[2 * i].bias.copy_(b)
    close(O.head_forward(fa, fb, onc), hp, 5e-6, "head")
    loss_o = O.weak_loss(fa, fb, onc)
    assert abs(float(loss_o) - float(loss_ref)) < 2e-8, ("weak loss", float(loss_o), float(loss_ref))  # |loss| ~ 3e-6: fp32 cancellation
    # gradient of the loss wrt the NC parameters: oracle in float64
    onc64 = O.NeighConsensus([5, 5, 5], [16, 16, 1]).double()
    with torch.no_grad():
        for i, (w, b) in enumerate(zip(ws, bs)):
            onc64.conv[2 * i].weight.copy_(w.double()), onc64.conv[2 * i].bias.copy_(b.double())
    l64 = O.weak_loss(fa.double(), fb.double(), onc64)
    l64.backward()
    arrs = dict(fa=fa, fb=fb, head_pos=hp, head_neg=hn, loss=loss_ref, loss64=l64.detach())
    for i in range(3):
        arrs[f"w{i}"], arrs[f"b{i}"] = ws[i], bs[i]
        arrs[f"grad_w{i}"] = onc64.conv[2 * i].weight.grad.float()
        arrs[f"grad_b{i}"] = onc64.conv[2 * i].bias.grad.float()
    npz("head_weak_loss", source="head/loss: reference modules + train.py:110-156 semantics; grads: oracle float64", **arrs)
    print("all reference checks passed; fixtures in", GOLD)


if __name__ == "__main__":
    main()
