"""CPU restatement of the NCNet matching head (the oracle).

TEST INFRASTRUCTURE ONLY -- never imported by ``ncnet_b200``.  Plain PyTorch on
CPU tensors, fp32 or fp64, written from the algorithm's definition and pinned
against the reference's own modules by ``oracle/make_golden.py``
(run in the build container where ``/root/reference`` exists) and by the golden
fixtures under ``tests/golden`` that script family generates.

The reference ships no tests and no golden vectors of its own,
so parity is pinned by *outputs of the reference itself run in the build
container* on seeded inputs; those outputs are the committed fixtures.

Each function cites the reference lines it restates (paths relative to the
reference checkout).  Three deliberate deviations from the reference, all
needed for it to run on torch >= 1.6 at all:
  * conv4d writes into a buffer that does not require grad (the reference's
    ``Variable(..., requires_grad=...)`` leaf breaks autograd),
  * maxpool4d decodes the window offsets with integer floor division (the
    reference's ``.div`` became true division and yields floats),
  * Conv4d parameters are created directly instead of through ``_ConvNd``.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------- #
# a1  featureL2Norm                                  reference lib/model.py:14-17
# --------------------------------------------------------------------------- #
def feature_l2norm(x: torch.Tensor) -> torch.Tensor:
    """x / sqrt(sum_c x^2 + 1e-6), reduction over dim 1, epsilon inside the root."""
    sq = (x * x).sum(dim=1, keepdim=True)
    return x / (sq + 1e-6).pow(0.5)


# --------------------------------------------------------------------------- #
# a2  FeatureCorrelation                             reference lib/model.py:89-120
# --------------------------------------------------------------------------- #
def feature_correlation(fa: torch.Tensor, fb: torch.Tensor, shape: str = "4D",
                        normalization: bool = True) -> torch.Tensor:
    """All-pairs dot products over channels.

    4D: out[b,0,i,j,k,l] = sum_c fa[b,c,i,j] * fb[b,c,k,l]       (model.py:106-115)
    3D: out[b, j*h+i, k, l] (source index column-major)          (model.py:97-105)
    normalization: featureL2Norm(relu(out)) over dim 1           (model.py:117-118)
    (for 4D that dim is a singleton -- a quirk the drop-in reproduces).
    """
    b, c, ha, wa = fa.shape
    _, _, hb, wb = fb.shape
    if shape == "4D":
        lhs = fa.reshape(b, c, ha * wa).transpose(1, 2)
        rhs = fb.reshape(b, c, hb * wb)
        out = torch.bmm(lhs, rhs).reshape(b, ha, wa, hb, wb).unsqueeze(1)
    elif shape == "3D":
        lhs = fb.reshape(b, c, hb * wb).transpose(1, 2)
        rhs = fa.transpose(2, 3).reshape(b, c, ha * wa)
        prod = torch.bmm(lhs, rhs)                       # [b, posB, posA(col-major)]
        out = prod.reshape(b, hb, wb, ha * wa).permute(0, 3, 1, 2)
    else:
        raise ValueError(shape)
    if normalization:
        out = feature_l2norm(torch.relu(out))
    return out


# --------------------------------------------------------------------------- #
# a3  MutualMatching                                reference lib/model.py:155-175
# --------------------------------------------------------------------------- #
def mutual_matching(x: torch.Tensor) -> torch.Tensor:
    """y = x * ((x / (max_B x + 1e-5)) * (x / (max_A x + 1e-5))); ch must be 1."""
    b, ch, f1, f2, f3, f4 = x.shape
    assert ch == 1
    m = x.reshape(b, f1 * f2, f3 * f4)
    max_over_a = m.max(dim=1, keepdim=True)[0]          # per target cell
    max_over_b = m.max(dim=2, keepdim=True)[0]          # per source cell
    ratio_b = m / (max_over_a + 1e-5)
    ratio_a = m / (max_over_b + 1e-5)
    return (m * (ratio_a * ratio_b)).reshape(x.shape)


# --------------------------------------------------------------------------- #
# a4  conv4d                                          reference lib/conv4d.py:11-51
# --------------------------------------------------------------------------- #
def to_prepermuted(w: torch.Tensor) -> torch.Tensor:
    """[Cout,Cin,k,k,k,k] -> the reference's stored layout [k,Cout,Cin,k,k,k]
    (lib/conv4d.py:76-77)."""
    return w.permute(2, 0, 1, 3, 4, 5).contiguous()


def from_prepermuted(w: torch.Tensor) -> torch.Tensor:
    return w.permute(1, 2, 0, 3, 4, 5).contiguous()


def conv4d(x: torch.Tensor, filters: torch.Tensor, bias: Optional[torch.Tensor] = None,
           permute_filters: bool = True) -> torch.Tensor:
    """4D cross-correlation, stride 1, zero 'same' padding, odd kernel.

    out[b,o,p] = bias[o] + sum_{c,d} x[b,c,p+d-k//2] * W[o,c,d]
    ``filters`` is [Cout,Cin,k,k,k,k] when permute_filters else the stored
    [k,Cout,Cin,k,k,k].  The first-dim taps are accumulated in the reference's
    order (centre, then -1,+1, -2,+2, ...; conv4d.py:41-48) so fp32 results
    round the same way; every tap is one conv3d over the merged (b*h) batch.
    """
    b, cin, h, w, d, t = x.shape
    wk = filters if not permute_filters else to_prepermuted(filters)
    k1, cout = wk.shape[0], wk.shape[1]
    pad = k1 // 2
    xp = F.pad(x, (0, 0, 0, 0, 0, 0, pad, pad))          # zero rows on the first 4D axis

    def tap(p: int, with_bias: bool) -> torch.Tensor:
        rows = xp[:, :, p:p + h].permute(0, 2, 1, 3, 4, 5).reshape(b * h, cin, w, d, t)
        y = F.conv3d(rows, wk[p], bias=bias if with_bias else None, stride=1,
                     padding=(wk.shape[3] // 2, wk.shape[4] // 2, wk.shape[5] // 2))
        return y

    acc = tap(pad, True)
    for q in range(1, pad + 1):
        acc = acc + tap(pad - q, False)
        acc = acc + tap(pad + q, False)
    return acc.reshape(b, h, cout, w, d, t).permute(0, 2, 1, 3, 4, 5).contiguous()


def conv4d_direct(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Independent definition-level check (pad + explicit tap loop), for tiny
    shapes only.  ``w`` is [Cout,Cin,k1,k2,k3,k4]."""
    b, cin, h, ww, d, t = x.shape
    cout, _, k1, k2, k3, k4 = w.shape
    xp = F.pad(x, (k4 // 2, k4 // 2, k3 // 2, k3 // 2, k2 // 2, k2 // 2, k1 // 2, k1 // 2))
    out = x.new_zeros(b, cout, h, ww, d, t)
    for a in range(k1):
        for bb in range(k2):
            for c3 in range(k3):
                for c4 in range(k4):
                    patch = xp[:, :, a:a + h, bb:bb + ww, c3:c3 + d, c4:c4 + t]
                    out = out + torch.einsum("bchwdt,oc->bohwdt", patch, w[:, :, a, bb, c3, c4])
    if bias is not None:
        out = out + bias.view(1, -1, 1, 1, 1, 1)
    return out


# --------------------------------------------------------------------------- #
# a5  Conv4d parameter holder                        reference lib/conv4d.py:53-82
# --------------------------------------------------------------------------- #
class Conv4d(torch.nn.Module):
    """weight stored as [k,Cout,Cin,k,k,k] (pre-permuted), bias [Cout]; default
    _ConvNd init = kaiming_uniform(a=sqrt(5)) == U(+-1/sqrt(fan_in)) for both."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size: int, bias: bool = True,
                 pre_permuted_filters: bool = True):
        super().__init__()
        k = kernel_size
        bound = 1.0 / math.sqrt(in_channels * k ** 4)
        w = torch.empty(out_channels, in_channels, k, k, k, k).uniform_(-bound, bound)
        self.pre_permuted_filters = pre_permuted_filters
        if pre_permuted_filters:
            w = to_prepermuted(w)
        self.weight = torch.nn.Parameter(w)
        self.bias = torch.nn.Parameter(torch.empty(out_channels).uniform_(-bound, bound)) if bias else None
        self.use_half = False

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return conv4d(x, self.weight, self.bias, permute_filters=not self.pre_permuted_filters)


# --------------------------------------------------------------------------- #
# a6  NeighConsensus                                reference lib/model.py:122-153
# --------------------------------------------------------------------------- #
class NeighConsensus(torch.nn.Module):
    def __init__(self, kernel_sizes: Sequence[int] = (3, 3, 3), channels: Sequence[int] = (10, 10, 1),
                 symmetric_mode: bool = True):
        super().__init__()
        self.symmetric_mode = symmetric_mode
        mods: List[torch.nn.Module] = []
        cin = 1
        for k, cout in zip(kernel_sizes, channels):
            mods.append(Conv4d(cin, cout, k))
            mods.append(torch.nn.ReLU())
            cin = cout
        self.conv = torch.nn.Sequential(*mods)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not self.symmetric_mode:
            return self.conv(x)
        swapped = x.permute(0, 1, 4, 5, 2, 3)
        return self.conv(x) + self.conv(swapped).permute(0, 1, 4, 5, 2, 3)   # model.py:147


def neigh_consensus(x: torch.Tensor, weights: Sequence[torch.Tensor], biases: Sequence[torch.Tensor],
                    symmetric: bool = True) -> torch.Tensor:
    """Functional form; ``weights`` in the stored [k,Cout,Cin,k,k,k] layout."""
    def stack(v: torch.Tensor) -> torch.Tensor:
        for w, bvec in zip(weights, biases):
            v = torch.relu(conv4d(v, w, bvec, permute_filters=False))
        return v
    if not symmetric:
        return stack(x)
    return stack(x) + stack(x.permute(0, 1, 4, 5, 2, 3)).permute(0, 1, 4, 5, 2, 3)


# --------------------------------------------------------------------------- #
# a7  maxpool4d                                     reference lib/model.py:177-191
# --------------------------------------------------------------------------- #
def maxpool4d(x: torch.Tensor, k_size: int = 4):
    """Non-overlapping k^4 max-pool of channel 0 with window-relative argmax
    offsets; window scanned (i,j,k,l) row-major, ties -> first.  Batch-1 only
    like the reference (for B>1 the reference mixes batch items; not restated)."""
    assert x.shape[0] == 1
    k = k_size
    v = x[0, 0]
    h1, h2, h3, h4 = (s // k for s in v.shape)
    win = v.reshape(h1, k, h2, k, h3, k, h4, k).permute(1, 3, 5, 7, 0, 2, 4, 6).reshape(k ** 4, h1, h2, h3, h4)
    val, idx = win.max(dim=0)
    max_l = idx % k
    max_k = (idx // k) % k
    max_j = (idx // (k * k)) % k
    max_i = idx // (k * k * k)
    shp = (1, 1, h1, h2, h3, h4)
    return (val.reshape(shp), max_i.reshape(shp), max_j.reshape(shp), max_k.reshape(shp), max_l.reshape(shp))


# --------------------------------------------------------------------------- #
# a9  corr_to_matches                               reference lib/point_tnf.py:12-80
# --------------------------------------------------------------------------- #
def corr_to_matches(corr4d: torch.Tensor, delta4d=None, k_size: int = 1, do_softmax: bool = False,
                    scale: str = "centered", return_indices: bool = False,
                    invert_matching_direction: bool = False):
    b, _, f1, f2, f3, f4 = corr4d.shape
    lo = -1.0 if scale == "centered" else 0.0
    # coordinate tables: float64 linspace cast to fp32 (point_tnf.py:16-27)
    ax = lambda n: torch.from_numpy(np.linspace(lo, 1, n * k_size)).float()
    xa_t, ya_t, xb_t, yb_t = ax(f2), ax(f1), ax(f4), ax(f3)

    if invert_matching_direction:           # one match per source cell (point_tnf.py:32-44)
        m = corr4d.reshape(b, f1 * f2, f3 * f4)
        if do_softmax:
            m = torch.softmax(m, dim=2)
        score, flat = m.max(dim=2)
        i_b, j_b = flat // f4, flat % f4
        own = torch.arange(f1 * f2).unsqueeze(0).expand(b, -1)
        i_a, j_a = own // f2, own % f2
    else:                                   # one match per target cell (point_tnf.py:46-57)
        m = corr4d.reshape(b, f1 * f2, f3 * f4)
        if do_softmax:
            m = torch.softmax(m, dim=1)
        score, flat = m.max(dim=1)
        i_a, j_a = flat // f2, flat % f2
        own = torch.arange(f3 * f4).unsqueeze(0).expand(b, -1)
        i_b, j_b = own // f4, own % f4

    if delta4d is not None:                 # relocalisation, batch-1 only (point_tnf.py:59-70)
        d_ia, d_ja, d_ib, d_jb = (d[0, 0] for d in delta4d)
        sel = (i_a.reshape(-1), j_a.reshape(-1), i_b.reshape(-1), j_b.reshape(-1))
        i_a2 = i_a * k_size + d_ia[sel].long().reshape(i_a.shape)
        j_a2 = j_a * k_size + d_ja[sel].long().reshape(i_a.shape)
        i_b2 = i_b * k_size + d_ib[sel].long().reshape(i_a.shape)
        j_b2 = j_b * k_size + d_jb[sel].long().reshape(i_a.shape)
        i_a, j_a, i_b, j_b = i_a2, j_a2, i_b2, j_b2

    x_a, y_a, x_b, y_b = xa_t[j_a], ya_t[i_a], xb_t[j_b], yb_t[i_b]
    if return_indices:
        return (x_a, y_a, x_b, y_b, score, i_a, j_a, i_b, j_b)
    return (x_a, y_a, x_b, y_b, score)


# --------------------------------------------------------------------------- #
# head assembly + weak loss          reference lib/model.py:269-282, train.py:110-156
# --------------------------------------------------------------------------- #
def head_forward(fa: torch.Tensor, fb: torch.Tensor, nc: NeighConsensus, relocalization_k_size: int = 0):
    """correlation -> [maxpool4d] -> MutualMatching -> NeighConsensus -> MutualMatching."""
    corr = feature_correlation(fa, fb, shape="4D", normalization=False)
    delta = None
    if relocalization_k_size > 1:
        corr, mi, mj, mk, ml = maxpool4d(corr, relocalization_k_size)
        delta = (mi, mj, mk, ml)
    corr = mutual_matching(corr)
    corr = nc(corr)
    corr = mutual_matching(corr)
    return corr if delta is None else (corr, delta)


def match_score(corr4d: torch.Tensor, normalization: Optional[str] = "softmax") -> torch.Tensor:
    """mean over cells of (max_A norm(x) + max_B norm(x)) / 2      (train.py:123-134)"""
    b, _, f1, f2, f3, f4 = corr4d.shape
    m = corr4d.reshape(b, f1 * f2, f3 * f4)
    if normalization == "softmax":
        over_a, over_b = torch.softmax(m, dim=1), torch.softmax(m, dim=2)
    elif normalization == "l1":
        over_a = m / (m.sum(dim=1, keepdim=True) + 0.0001)
        over_b = m / (m.sum(dim=2, keepdim=True) + 0.0001)
    else:
        over_a = over_b = m
    s_b = over_a.max(dim=1)[0]          # per target cell  [b, f3*f4]
    s_a = over_b.max(dim=2)[0]          # per source cell  [b, f1*f2]
    if f1 * f2 == f3 * f4:
        return torch.mean(s_a + s_b) / 2
    return (s_a.mean() + s_b.mean()) / 2


def weak_loss(fa: torch.Tensor, fb: torch.Tensor, nc: NeighConsensus,
              normalization: Optional[str] = "softmax") -> torch.Tensor:
    """score(negatives) - score(positives); negatives pair target b with source
    (b+1) mod B (train.py:137: roll of the source images by -1)."""
    pos = match_score(head_forward(fa, fb, nc), normalization)
    neg = match_score(head_forward(torch.roll(fa, shifts=-1, dims=0), fb, nc), normalization)
    return neg - pos
