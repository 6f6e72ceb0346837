"""CPU model of the tensor-core ("TC16") mode of the matching head.  TEST INFRASTRUCTURE ONLY.

``oracle/head.py`` restates the reference in full precision.  The TC16 kernels deviate from it in exactly one
way: some tensors are rounded to fp16 on their way into a tensor-core instruction (products are then exact and
sums are fp32).  This module applies the *same roundings at the same places* on top of the oracle's arithmetic
(run it in float64), so that what is left between it and the CUDA path is summation order only.  It lets the
tests separate two questions the fp32 oracle alone cannot:

  * is every index / mask / layout of the TC16 path right?  -> compare with this model, tight tolerance;
  * how far is TC16 from the reference's fp32 result?        -> compare with ``oracle/head.py``, stated tolerance.

Where the kernels round (ncnet_b200/csrc):
  correlation (corr4d_tc.cu)       both feature operands -> fp16 (inference / frozen trunk only)
  Conv4d forward (conv4d_tc.cu)    input activations and weights -> fp16; bias fp32; hidden activations of a chained
                                   stack are stored as fp16 after bias + ReLU; the stack's last output stays fp32
  Conv4d dgrad                     incoming gradient and weights -> fp16; result (after the ReLU mask) stored fp16
  Conv4d wgrad                     activations (as stored) and incoming gradient -> fp16; result fp32
  backward of a stack              the gradient arriving at the stack's output is multiplied by S = 2^e with
                                   max|g|*S in [8,16) (ncnet_grad_scale) and the weight / bias / input gradients by
                                   1/S at the end -- exact, and it keeps the fp16 gradient roundings out of fp16's
                                   subnormal range
Everything else (MutualMatching, match score, correlation backward) is fp32 in both modes.

Follows reference lib/model.py:143-153 (NeighConsensus.forward), lib/conv4d.py:11-51 and train.py:110-156 through
``oracle/head.py``.
"""
from __future__ import annotations

import math
from typing import Sequence

import torch

from . import head as O


class _Round16(torch.autograd.Function):
    """fp16 round-to-nearest-even of the value (``fwd``) and / or of the gradient flowing back (``bwd``)."""

    @staticmethod
    def forward(ctx, t, fwd, bwd):
        ctx.bwd = bwd
        return t.half().to(t.dtype) if fwd else t.clone()

    @staticmethod
    def backward(ctx, g):
        return (g.half().to(g.dtype) if ctx.bwd else g), None, None


def r16(t: torch.Tensor, fwd: bool = True, bwd: bool = False) -> torch.Tensor:
    return _Round16.apply(t, fwd, bwd)


class _ScaleOut(torch.autograd.Function):
    """identity; backward: g -> S*g with the library's power-of-two scale, S left in ``state``"""

    @staticmethod
    def forward(ctx, t, state):
        ctx.state = state
        return t.clone()

    @staticmethod
    def backward(ctx, g):
        m = float(g.detach().float().abs().max())       # the kernel takes the maximum of the fp32 gradient
        s = 1.0
        if m > 0 and math.isfinite(m):
            _, e = math.frexp(m)                        # m = f * 2^e, f in [0.5, 1)
            s = math.ldexp(1.0, max(-100, min(100, 4 - e)))
        ctx.state["S"] = s
        return g * s, None


class _Unscale(torch.autograd.Function):
    """identity; backward: g -> g / S"""

    @staticmethod
    def forward(ctx, t, state):
        ctx.state = state
        return t.clone()

    @staticmethod
    def backward(ctx, g):
        return g / ctx.state["S"], None


def neigh_consensus(x: torch.Tensor, weights: Sequence[torch.Tensor], biases: Sequence[torch.Tensor],
                    symmetric: bool = True) -> torch.Tensor:
    """The [Conv4d + ReLU] x L stack (lib/model.py:130-153) with the TC16 roundings.  ``weights`` in the stored
    [k,Cout,Cin,k,k,k] layout.  Gradients: d/dweights are what the wgrad kernels compute (fp16 activations x fp16
    incoming gradients, wide accumulation); d/dx uses the fp16 weights here whereas the library's first-layer input
    gradient runs on the fp32 CUDA-core kernel with unrounded weights -- compare that one loosely."""
    n = len(weights)
    state = {"S": 1.0}
    x = _Unscale.apply(x, state)
    weights = [_Unscale.apply(w, state) for w in weights]
    biases = [_Unscale.apply(b, state) for b in biases]

    def branch(v):
        v = r16(v)                                         # planar input -> fp16 operand
        for l, (w, b) in enumerate(zip(weights, biases)):
            z = O.conv4d(v, r16(w), b, permute_filters=False)
            # hidden layers: value stored fp16, gradient stored fp16; last layer: fp32 value, gradient rounded when
            # it is packed for the dgrad / wgrad launches (after the ReLU mask: rounding commutes with masking)
            v = torch.relu(r16(z, fwd=l < n - 1, bwd=True))
        return v

    y = branch(x)
    if symmetric:
        y = y + branch(x.permute(0, 1, 4, 5, 2, 3)).permute(0, 1, 4, 5, 2, 3)
    return _ScaleOut.apply(y, state)


def _params(nc):
    ws = [m.weight for m in nc.conv if isinstance(m, O.Conv4d)]
    bs = [m.bias for m in nc.conv if isinstance(m, O.Conv4d)]
    return ws, bs


def head_forward(fa: torch.Tensor, fb: torch.Tensor, nc: "O.NeighConsensus") -> torch.Tensor:
    """oracle.head.head_forward (lib/model.py:269-282, no relocalisation) as TC16 mode computes it for features
    that need no gradient: fp16 correlation operands, fp16 Conv4d operands, everything else unrounded."""
    corr = O.feature_correlation(r16(fa), r16(fb), "4D", False)
    x = O.mutual_matching(corr)
    ws, bs = _params(nc)
    return O.mutual_matching(neigh_consensus(x, ws, bs, nc.symmetric_mode))


def weak_loss(fa: torch.Tensor, fb: torch.Tensor, nc: "O.NeighConsensus") -> torch.Tensor:
    """train.py:110-156 on features: score(negatives: source rolled by -1) - score(positives)."""
    pos = O.match_score(head_forward(fa, fb, nc))
    neg = O.match_score(head_forward(torch.roll(fa, -1, 0), fb, nc))
    return neg - pos


def count_relu_flips(nc64: "O.NeighConsensus", x64: torch.Tensor):
    """(flipped, total): hidden units of the first branch whose sign differs between the exact stack and the
    fp16-operand stack -- the units whose whole gradient path switches on or off in TC16 mode."""
    r = lambda t: t.half().to(t.dtype)
    ws, bs = _params(nc64)
    v, ve, flips, total = x64, r(x64), 0, 0
    with torch.no_grad():
        for l, (w, b) in enumerate(zip(ws, bs)):
            z = O.conv4d(v, w, b, permute_filters=False)
            ze = O.conv4d(ve, r(w), b, permute_filters=False)
            if l < len(ws) - 1:
                ze = r(ze)
            flips += int(((z > 0) != (ze > 0)).sum())
            total += z.numel()
            v, ve = torch.relu(z), torch.relu(ze)
    return flips, total
