"""torch.autograd.Functions over the C ABI (include/ncnet_b200.h).

PyTorch is plumbing here: it owns device memory and streams; every arithmetic step of the
matching head is one of this library's kernels.  CUDA tensors only -- CPU tensors raise.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _cabi
from ._cabi import CONV_FP32, CONV_TC16, F16, F32, W_KOIKKK, W_OIKKKK

# number of kernel launches issued through the C ABI (bench.py reports it as gpu_launches)
LAUNCHES = {"n": 0}
# when a list: (tag, start_event, end_event) per 16->16 Conv4d forward launch (bench.py's roofline)
EVENT_LOG = None


def default_conv_mode() -> int:
    """NCNET_CONV_MODE selects the Conv4d back end: "tc16" (default; tcgen05 tensor cores, fp16 operands,
    fp32 accumulation, <= 1e-3 normalised error per layer) or "fp32" (CUDA-core fp32 FMA, parity mode)."""
    return CONV_FP32 if os.environ.get("NCNET_CONV_MODE", "tc16").lower() in ("fp32", "f32", "0") else CONV_TC16


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.float16:
        return F16
    raise TypeError(f"ncnet_b200: unsupported dtype {t.dtype} (float32 or float16)")


def _need_cuda(*ts: torch.Tensor) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("ncnet_b200: CUDA tensors required -- this library has no CPU path")


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _count(n: int) -> None:
    LAUNCHES["n"] += n


# ------------------------------------------------------------------------------------------- #
# featureL2Norm / FeatureCorrelation
# ------------------------------------------------------------------------------------------- #
def feature_l2norm(x: torch.Tensor) -> torch.Tensor:
    _need_cuda(x)
    xc = x.contiguous()
    b, c = xc.shape[0], xc.shape[1]
    hw = xc.numel() // (b * c)
    y = torch.empty_like(xc)
    with torch.cuda.device(xc.device):
        _cabi.check(_cabi.lib().ncnet_feature_l2norm(_ptr(xc), _ptr(y), b, c, hw, _dt(xc), _stream()), "feature_l2norm")
    _count(1)
    return y


def _corr_fwd_raw(fa: torch.Tensor, fb: torch.Tensor, norm: bool, tensor_cores: bool = False) -> torch.Tensor:
    b, c = fa.shape[0], fa.shape[1]
    na, nb = fa[0, 0].numel(), fb[0, 0].numel()
    out = torch.empty(b, na, nb, device=fa.device, dtype=fa.dtype)
    if tensor_cores:
        L = _cabi.lib()
        scratch = torch.empty(L.ncnet_corr4d_tc_scratch_bytes(b, c, na, nb), device=fa.device, dtype=torch.uint8)
        with torch.cuda.device(fa.device):
            _cabi.check(L.ncnet_corr4d_fwd_tc(_ptr(fa), _ptr(fb), _ptr(out), _ptr(scratch), b, c, na, nb, _dt(fa), int(norm), 0,
                                              _stream()), "corr4d_fwd_tc")
        _count(3)
        return out
    with torch.cuda.device(fa.device):
        _cabi.check(_cabi.lib().ncnet_corr4d_fwd(_ptr(fa), _ptr(fb), _ptr(out), b, c, na, nb, _dt(fa), int(norm),
                                                 _stream()), "corr4d_fwd")
    _count(1)
    return out


class _CorrFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, fa, fb):
        fa, fb = fa.contiguous(), fb.contiguous()
        ctx.save_for_backward(fa, fb)
        return _corr_fwd_raw(fa, fb, False)

    @staticmethod
    def backward(ctx, g):
        fa, fb = ctx.saved_tensors
        g = g.contiguous()
        b, c = fa.shape[0], fa.shape[1]
        na, nb = fa[0, 0].numel(), fb[0, 0].numel()
        ga = torch.empty_like(fa) if ctx.needs_input_grad[0] else None
        gb = torch.empty_like(fb) if ctx.needs_input_grad[1] else None
        with torch.cuda.device(fa.device):
            _cabi.check(_cabi.lib().ncnet_corr4d_bwd(_ptr(g), _ptr(fa), _ptr(fb), _ptr(ga), _ptr(gb), b, c, na, nb,
                                                     _dt(fa), _stream()), "corr4d_bwd")
        _count((ga is not None) + (gb is not None))
        return ga, gb


def feature_correlation(fa: torch.Tensor, fb: torch.Tensor, shape: str = "4D", normalization: bool = True) -> torch.Tensor:
    """FeatureCorrelation.forward (reference lib/model.py:96-120)."""
    _need_cuda(fa, fb)
    if fa.dtype != fb.dtype:
        raise TypeError("feature_correlation: dtype mismatch")
    b, c, ha, wa = fa.shape
    _, _, hb, wb = fb.shape
    needs_grad = torch.is_grad_enabled() and (fa.requires_grad or fb.requires_grad)
    if shape == "4D":
        lhs, rhs = fa, fb
    elif shape == "3D":       # out[b, j*h+i, k, l]: source index column-major, target cells as the plane
        lhs, rhs = fa.transpose(2, 3), fb
    else:
        raise ValueError(shape)
    if needs_grad:
        m = _CorrFn.apply(lhs, rhs)
        if normalization:     # ReLU + L2 norm over the channel dim, elementwise tail kept differentiable
            m = m.reshape(b, 1, ha, wa, hb, wb) if shape == "4D" else m.reshape(b, ha * wa, hb, wb)
            m = torch.relu(m)
            m = m / (m.pow(2).sum(1, keepdim=True) + 1e-6).pow(0.5)
            return m
    else:
        fuse = normalization and shape == "4D"
        m = _corr_fwd_raw(lhs.contiguous(), rhs.contiguous(), fuse, tensor_cores=default_conv_mode() == CONV_TC16)
        if normalization and shape == "3D":
            m = feature_l2norm(torch.relu(m.reshape(b, ha * wa, hb, wb)))
            return m
    if shape == "4D":
        return m.reshape(b, 1, ha, wa, hb, wb)
    return m.reshape(b, ha * wa, hb, wb)


def feature_correlation_pooled(fa: torch.Tensor, fb: torch.Tensor, k_size: int = 2, normalization: bool = False,
                               l2norm_inputs: bool = False, index_dtype: torch.dtype = torch.int64):
    """``maxpool4d(FeatureCorrelation('4D', normalization)(fa, fb), k_size)`` (reference lib/model.py:269-272) in one
    tensor-core kernel that never writes the un-pooled volume.  Inference only (no gradient), batch 1, k_size 2, even
    feature-map sides.  ``l2norm_inputs``: fa / fb are un-normalised trunk outputs and featureL2Norm
    (lib/model.py:85-86) is applied while the operands are packed.  Returns (value, max_i, max_j, max_k, max_l)."""
    _need_cuda(fa, fb)
    if fa.dtype != fb.dtype:
        raise TypeError("feature_correlation_pooled: dtype mismatch")
    b, c, ha, wa = fa.shape
    _, _, hb, wb = fb.shape
    if b != 1:
        raise RuntimeError("feature_correlation_pooled: batch size must be 1 (maxpool4d is batch-1 in the reference)")
    fa, fb = fa.detach().contiguous(), fb.detach().contiguous()
    shp = (1, 1, ha // 2, wa // 2, hb // 2, wb // 2)
    val = torch.empty(shp, device=fa.device, dtype=fa.dtype)
    idx = [torch.empty(shp, device=fa.device, dtype=index_dtype) for _ in range(4)]
    L = _cabi.lib()
    scratch = torch.empty(L.ncnet_corr4d_tc_scratch_bytes(1, c, ha * wa, hb * wb), device=fa.device, dtype=torch.uint8)
    with torch.cuda.device(fa.device):
        _cabi.check(L.ncnet_corr4d_pool_fwd_tc(_ptr(fa), _ptr(fb), _ptr(val), *[_ptr(t) for t in idx], _ptr(scratch), c, ha, wa,
                                               hb, wb, int(k_size), {torch.int64: 8, torch.uint8: 1}[index_dtype], _dt(fa),
                                               int(normalization), int(l2norm_inputs), _stream()), "corr4d_pool_fwd_tc")
    _count(3)
    return (val, idx[0], idx[1], idx[2], idx[3])


def can_fuse_correlation_pool(fa: torch.Tensor, fb: torch.Tensor, k_size: int) -> bool:
    return (k_size == 2 and fa.shape[0] == 1 and default_conv_mode() == CONV_TC16 and all(s % 2 == 0 for s in (*fa.shape[2:], *fb.shape[2:]))
            and not (torch.is_grad_enabled() and (fa.requires_grad or fb.requires_grad)))


# ------------------------------------------------------------------------------------------- #
# MutualMatching
# ------------------------------------------------------------------------------------------- #
class _MutualMatchingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        b, g, f1, f2, f3, f4 = x.shape
        if g not in (1, 2):
            raise RuntimeError("MutualMatching: channel dim must be 1 (or 2 = the two NeighConsensus branches)")
        xc = x.contiguous()
        na, nb = f1 * f2, f3 * f4
        y = torch.empty(b, 1, f1, f2, f3, f4, device=x.device, dtype=x.dtype)
        keys = torch.empty(b * (na + nb), device=x.device, dtype=torch.int64)
        # no gradient will be asked for: pass 1 need not track the arg-max indices the backward pass consumes
        infer = not ctx.needs_input_grad[0]
        fn = _cabi.lib().ncnet_mutual_matching_fwd_infer if infer else _cabi.lib().ncnet_mutual_matching_fwd
        with torch.cuda.device(x.device):
            _cabi.check(fn(_ptr(xc), _ptr(y), _ptr(keys), b, na, nb, g, _dt(xc), _stream()), "mutual_matching_fwd")
        _count(2)
        if not infer:
            ctx.save_for_backward(xc, keys)
        return y

    @staticmethod
    def backward(ctx, gy):
        xc, keys = ctx.saved_tensors
        b, g, f1, f2, f3, f4 = xc.shape
        na, nb = f1 * f2, f3 * f4
        gy = gy.contiguous()
        gx = torch.empty(b, 1, f1, f2, f3, f4, device=xc.device, dtype=xc.dtype)
        sums = torch.empty(_cabi.lib().ncnet_mutual_matching_sums_bytes(b, na, nb) // 4, device=xc.device, dtype=torch.float32)
        with torch.cuda.device(xc.device):
            _cabi.check(_cabi.lib().ncnet_mutual_matching_bwd(_ptr(xc), _ptr(gy), _ptr(keys), _ptr(sums), _ptr(gx), b,
                                                              na, nb, g, _dt(xc), _stream()), "mutual_matching_bwd")
        _count(2)
        return gx.expand(b, g, f1, f2, f3, f4) if g == 2 else gx


def mutual_matching(corr4d: torch.Tensor, sum_branches: bool = False) -> torch.Tensor:
    """MutualMatching (reference lib/model.py:155-175).  With ``sum_branches`` a 2-channel input is
    read as the two symmetric NeighConsensus branch outputs and summed on the fly."""
    _need_cuda(corr4d)
    if corr4d.dim() != 6 or corr4d.shape[1] != (2 if sum_branches else 1):
        raise RuntimeError(f"MutualMatching: expected [B,{2 if sum_branches else 1},fs1,fs2,fs3,fs4], got {tuple(corr4d.shape)}")
    return _MutualMatchingFn.apply(corr4d)


class _MatchScoreFn(torch.autograd.Function):
    """Soft-max matching score of the weak loss (reference train.py:123-134), fused: maxima, sums of exponentials,
    one reduction; the backward pass is one elementwise kernel."""

    @staticmethod
    def forward(ctx, corr4d, item_weight=None):
        b, _, f1, f2, f3, f4 = corr4d.shape
        x = corr4d.contiguous()
        wt = None if item_weight is None else item_weight.detach().to(device=x.device, dtype=torch.float32).contiguous()
        if wt is not None and wt.numel() != b:
            raise RuntimeError(f"match_score: {wt.numel()} item weights for a batch of {b}")
        na, nb = f1 * f2, f3 * f4
        keys = torch.empty(b * (na + nb), device=x.device, dtype=torch.int64)
        sums = torch.empty(_cabi.lib().ncnet_mutual_matching_sums_bytes(b, na, nb) // 4, device=x.device, dtype=torch.float32)
        score = torch.empty((), device=x.device, dtype=torch.float32)
        with torch.cuda.device(x.device):
            _cabi.check(_cabi.lib().ncnet_match_score_fwd(_ptr(x), _ptr(keys), _ptr(sums), _ptr(wt), _ptr(score), b, na, nb,
                                                          _dt(x), _stream()), "match_score_fwd")
        _count(4)
        ctx.has_weight = wt is not None
        ctx.save_for_backward(x, keys, sums, *([wt] if wt is not None else []))
        return score.to(x.dtype)

    @staticmethod
    def backward(ctx, g):
        x, keys, sums = ctx.saved_tensors[:3]
        wt = ctx.saved_tensors[3] if ctx.has_weight else None
        b, _, f1, f2, f3, f4 = x.shape
        gs = g.detach().reshape(1).float().contiguous()
        gx = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _cabi.check(_cabi.lib().ncnet_match_score_bwd(_ptr(x), _ptr(keys), _ptr(sums), _ptr(wt), _ptr(gs), _ptr(gx), b,
                                                          f1 * f2, f3 * f4, _dt(x), _stream()), "match_score_bwd")
        _count(1)
        return gx, None


def match_score_softmax(corr4d: torch.Tensor, item_weight: Optional[torch.Tensor] = None) -> torch.Tensor:
    """mean(max softmax over A + max softmax over B) / 2 of a [B,1,fs1,fs2,fs3,fs4] volume (train.py:123-134).
    ``item_weight`` [B]: weighted sum over the batch items instead of their mean (see include/ncnet_b200.h)."""
    _need_cuda(corr4d)
    if corr4d.dim() != 6 or corr4d.shape[1] != 1:
        raise RuntimeError(f"match_score: expected [B,1,fs1,fs2,fs3,fs4], got {tuple(corr4d.shape)}")
    return _MatchScoreFn.apply(corr4d, item_weight)


# ------------------------------------------------------------------------------------------- #
# Conv4d / NeighConsensus
# ------------------------------------------------------------------------------------------- #
def _wdims(w: torch.Tensor, layout: int) -> Tuple[int, int, int]:
    if layout == W_KOIKKK:
        return w.shape[2], w.shape[1], w.shape[0]       # Cin, Cout, k
    return w.shape[1], w.shape[0], w.shape[2]


class _CB16:
    """Handle of a channel-blocked, zero-haloed fp16 activation tensor (include/ncnet_b200.h: NCNET_CB16).
    ``buf`` is a flat uint8 torch tensor owned by PyTorch; ``shape`` is the planar shape it stands for.  ``pm``: the
    buffer also holds the position-major twin behind the CB16 part (NCNET_CB16PM), which the weight-gradient kernels
    read without a conversion pass."""
    __slots__ = ("buf", "shape", "k", "dtype", "pm")

    def __init__(self, buf, shape, k, dtype, pm=False):
        self.buf, self.shape, self.k, self.dtype, self.pm = buf, tuple(shape), k, dtype, pm

    @staticmethod
    def empty(shape, k, device, dtype, per_group, pm=False):
        bsz, ch, d1, d2, d3, d4 = shape
        nblk = bsz * (ch // per_group)
        nbytes = _cabi.lib().ncnet_cb16_bytes(nblk, d1, d2, d3, d4, k)
        if pm:
            nbytes = 2 * ((nbytes + 255) & ~255)
        return _CB16(torch.empty(nbytes, device=device, dtype=torch.uint8), shape, k, dtype, pm)

    @property
    def device(self):
        return self.buf.device

    def data_ptr(self):
        return self.buf.data_ptr()


def _adt(t) -> int:
    if isinstance(t, _CB16):
        return _cabi.CB16PM if t.pm else _cabi.CB16
    return _dt(t)


def _conv_fwd(x, w32, b32, layout, symmetric, shared_input, relu, mode, out_cb16=False, out_pm=False):
    cin, cout, k = _wdims(w32, layout)
    bsz, _, d1, d2, d3, d4 = x.shape
    g = 2 if symmetric else 1
    oshape = (bsz, g * cout, d1, d2, d3, d4)
    if out_cb16:
        y = _CB16.empty(oshape, k, x.device, x.dtype, cout, pm=out_pm)
    else:
        y = torch.empty(oshape, device=x.device, dtype=x.dtype)
    L = _cabi.lib()
    scratch = torch.empty(L.ncnet_conv4d_scratch_bytes(bsz, cin, cout, d1, d2, d3, d4, k, int(symmetric), mode),
                          device=x.device, dtype=torch.uint8)
    log = EVENT_LOG is not None and cin == 16 and cout == 16
    with torch.cuda.device(x.device):
        if log:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
        _cabi.check(L.ncnet_conv4d_fwd(x.data_ptr(), _ptr(w32), _ptr(b32), y.data_ptr(), _ptr(scratch), bsz, cin, cout, d1,
                                       d2, d3, d4, k, layout, int(symmetric), int(shared_input), int(relu), _adt(x),
                                       _adt(y), mode, _stream()), "conv4d_fwd")
        if log:
            ev1.record()
            EVENT_LOG.append(("conv4d_fwd_16_16", ev0, ev1))
    _count(2 if isinstance(x, _CB16) else 3 if mode == CONV_TC16 else 2)
    return y


def _conv_dgrad(gy, w32, mask, layout, symmetric, shared_input, mode, out_cb16=False, out_pm=False):
    cin, cout, k = _wdims(w32, layout)
    bsz, _, d1, d2, d3, d4 = gy.shape
    gin = 1 if (shared_input or not symmetric) else 2
    oshape = (bsz, gin * cin, d1, d2, d3, d4)
    if out_cb16:
        gx = _CB16.empty(oshape, k, gy.device, gy.dtype, cin, pm=out_pm)
    else:
        gx = torch.empty(oshape, device=gy.device, dtype=gy.dtype)
    L = _cabi.lib()
    scratch = torch.empty(L.ncnet_conv4d_scratch_bytes(bsz, cin, cout, d1, d2, d3, d4, k, int(symmetric), mode),
                          device=gy.device, dtype=torch.uint8)
    with torch.cuda.device(gy.device):
        _cabi.check(L.ncnet_conv4d_dgrad(gy.data_ptr(), _ptr(w32), None if mask is None else mask.data_ptr(),
                                         gx.data_ptr(), _ptr(scratch), bsz, cin, cout, d1, d2, d3, d4, k, layout,
                                         int(symmetric), int(shared_input), _adt(gy), 0 if mask is None else _adt(mask),
                                         _adt(gx), mode, _stream()), "conv4d_dgrad")
    _count(2 if isinstance(gy, _CB16) else 3 if mode == CONV_TC16 else 2)
    return gx


def _conv_wgrad(x, gy, w_like, want_bias, layout, symmetric, shared_input, mode, out_scale=None):
    cin, cout, k = _wdims(w_like, layout)
    bsz, _, d1, d2, d3, d4 = x.shape
    dev = x.device
    gw = torch.empty(w_like.shape, device=dev, dtype=torch.float32)
    gb = torch.empty(cout, device=dev, dtype=torch.float32) if want_bias else None
    L = _cabi.lib()
    nbytes = L.ncnet_conv4d_wgrad_scratch_bytes(bsz, cin, cout, d1, d2, d3, d4, k, int(symmetric), mode)
    scratch = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    log = EVENT_LOG is not None and cin == 16 and cout == 16
    with torch.cuda.device(dev):
        if log:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
        _cabi.check(L.ncnet_conv4d_wgrad(x.data_ptr(), gy.data_ptr(), _ptr(gw), _ptr(gb), _ptr(scratch), bsz, cin, cout, d1,
                                         d2, d3, d4, k, layout, int(symmetric), int(shared_input), _adt(x), _adt(gy), mode,
                                         _ptr(out_scale), _stream()), "conv4d_wgrad")
        if log:
            ev1.record()
            EVENT_LOG.append(("conv4d_wgrad_16_16", ev0, ev1))
    _count(4 if want_bias else 3)
    return gw, gb


def cb16_unpack(t: "_CB16", per_group: int) -> torch.Tensor:
    """CB16 handle -> planar tensor of its shape (tests, and the rare input-gradient path)."""
    bsz, ch, d1, d2, d3, d4 = t.shape
    out = torch.empty(t.shape, device=t.device, dtype=t.dtype)
    with torch.cuda.device(t.device):
        _cabi.check(_cabi.lib().ncnet_cb16_unpack(t.data_ptr(), _ptr(out), bsz * (ch // per_group), per_group, d1, d2, d3, d4,
                                                  t.k, _dt(out), _stream()), "cb16_unpack")
    _count(1)
    return out


class _Conv4dFn(torch.autograd.Function):
    """One Conv4d layer, no activation (reference lib/conv4d.py:11-51)."""

    @staticmethod
    def forward(ctx, x, weight, bias, layout, mode):
        xc = x.contiguous()
        w32 = weight.detach().float().contiguous()
        b32 = None if bias is None else bias.detach().float().contiguous()
        ctx.layout, ctx.mode = layout, mode
        ctx.has_bias = bias is not None
        ctx.save_for_backward(xc, w32)
        ctx.wdtype = weight.dtype
        return _conv_fwd(xc, w32, b32, layout, False, True, False, mode)

    @staticmethod
    def backward(ctx, gy):
        xc, w32 = ctx.saved_tensors
        gy = gy.contiguous()
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            gx = _conv_dgrad(gy, w32, None, ctx.layout, False, True, ctx.mode)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            gw, gb = _conv_wgrad(xc, gy, w32, ctx.has_bias, ctx.layout, False, True, ctx.mode)
            gw = gw.to(ctx.wdtype)
            gb = None if gb is None else gb.to(ctx.wdtype)
        return gx, gw, gb, None, None


def conv4d(data: torch.Tensor, filters: torch.Tensor, bias: Optional[torch.Tensor] = None,
           permute_filters: bool = True, use_half: bool = False, mode: Optional[int] = None) -> torch.Tensor:
    """Drop-in for the reference ``conv4d`` (lib/conv4d.py:11).  ``permute_filters=True`` means
    ``filters`` is [Cout,Cin,k,k,k,k]; False means the stored pre-permuted [k,Cout,Cin,k,k,k]."""
    _need_cuda(data, filters, bias)
    layout = W_OIKKKK if permute_filters else W_KOIKKK
    return _Conv4dFn.apply(data, filters, bias, layout, default_conv_mode() if mode is None else mode)


def _stack_can_chain(kernel_sizes, chans, d3, d4, mode, need_backward=True) -> bool:
    """True when every layer of the stack can hand its activations to the next one in CB16
    (tensor-core mode, one kernel size, <= 16 channels per group).  Inference only needs the forward envelope."""
    if mode != CONV_TC16 or len(set(kernel_sizes)) != 1 or len(kernel_sizes) < 2:
        return False
    L = _cabi.lib()
    ok = L.ncnet_conv4d_tc_supported if need_backward else L.ncnet_conv4d_tc_fwd_supported
    cin = 1
    for cout in chans:
        if not ok(cin, cout, kernel_sizes[0], d3, d4):
            return False
        cin = cout
    return True


class _NeighConsensusFn(torch.autograd.Function):
    """The whole [Conv4d + ReLU] x L stack, both symmetric branches per launch, ReLU fused into the
    conv epilogue and ReLU-backward fused into the dgrad epilogue (reference lib/model.py:143-153).
    In tensor-core mode the hidden activations (and their gradients) stay in the CB16 layout between
    launches.  Returns [B, G*Clast, ...] (G=2 branches when symmetric; the caller sums them)."""

    @staticmethod
    def forward(ctx, x, symmetric, mode, train, *wb):
        n_layers = len(wb) // 2
        xc = x.contiguous()
        w32s = [wb[2 * l].detach().float().contiguous() for l in range(n_layers)]
        b32s = [wb[2 * l + 1].detach().float().contiguous() for l in range(n_layers)]
        chain = _stack_can_chain([w.shape[0] for w in w32s], [w.shape[1] for w in w32s], xc.shape[4], xc.shape[5], mode,
                                 need_backward=train)
        acts = [xc]
        cur = xc
        for l in range(n_layers):
            cur = _conv_fwd(cur, w32s[l], b32s[l], W_KOIKKK, symmetric, l == 0, True, mode,
                            out_cb16=chain and l < n_layers - 1, out_pm=train)   # training: + position-major twin for wgrad
            acts.append(cur)
        ctx.symmetric, ctx.mode, ctx.n_layers, ctx.chain = symmetric, mode, n_layers, chain
        ctx.wdtype = wb[0].dtype
        ctx.meta = [(a.shape, a.k, a.dtype, a.pm) if isinstance(a, _CB16) else None for a in acts]
        ctx.save_for_backward(*[a.buf if isinstance(a, _CB16) else a for a in acts], *w32s)
        return cur

    @staticmethod
    def backward(ctx, g):
        n = ctx.n_layers
        saved = ctx.saved_tensors
        acts = [(_CB16(t, *m) if m is not None else t) for t, m in zip(saved[: n + 1], ctx.meta)]
        w32s = saved[n + 1:]
        sym, mode, chain = ctx.symmetric, ctx.mode, ctx.chain
        out = acts[n]
        bsz = out.shape[0]
        g = g.contiguous()
        # TC16 mode stores every gradient of the backward pass in fp16: scale the incoming one by a power of two so
        # that nothing falls into fp16's subnormal range, un-scale the (fp32) weight gradients at the end
        scale3 = inv_scale = None
        if mode == CONV_TC16:
            scale3 = torch.empty(3, device=out.device, dtype=torch.float32)
            inv_scale = scale3[1:]
            with torch.cuda.device(out.device):
                _cabi.check(_cabi.lib().ncnet_grad_scale(_ptr(g), g.numel(), _dt(g), _ptr(scale3), _stream()), "grad_scale")
            _count(2)
        # ReLU of the last layer: gz = g where the branch output is positive
        gz = torch.empty_like(out)
        per = out[0].numel()
        with torch.cuda.device(out.device):
            _cabi.check(_cabi.lib().ncnet_relu_mask_bcast(_ptr(out), _ptr(g), _ptr(gz), bsz, 1, per, _dt(out), _ptr(scale3),
                                                          _stream()), "relu_mask_bcast")
        _count(1)
        grads: List[Optional[torch.Tensor]] = [None] * (2 * n)
        gx = None
        for l in range(n - 1, -1, -1):
            gw, gb = _conv_wgrad(acts[l], gz, w32s[l], True, W_KOIKKK, sym, l == 0, mode, out_scale=inv_scale)
            grads[2 * l], grads[2 * l + 1] = gw.to(ctx.wdtype), gb.to(ctx.wdtype)
            if l > 0:
                gz = _conv_dgrad(gz, w32s[l], acts[l], W_KOIKKK, sym, False, mode, out_cb16=chain, out_pm=chain)
            elif ctx.needs_input_grad[0]:
                if isinstance(gz, _CB16):       # merged 2*Cout -> Cin gradient runs on the CUDA-core kernel (planar)
                    gz = cb16_unpack(gz, w32s[0].shape[1])
                gx = _conv_dgrad(gz, w32s[0], None, W_KOIKKK, sym, True, mode)
                if inv_scale is not None:
                    gx = gx * inv_scale[0]
        return (gx, None, None, None, *grads)


def neigh_consensus_branches(x: torch.Tensor, weights: Sequence[torch.Tensor], biases: Sequence[torch.Tensor],
                             symmetric: bool = True, mode: Optional[int] = None) -> torch.Tensor:
    """Runs the NeighConsensus stack; returns the branch outputs [B, G*Clast, fs1..fs4] un-summed."""
    _need_cuda(x, *weights, *biases)
    wb = []
    for w, b in zip(weights, biases):
        wb += [w, b]
    # inference (no gradient will be asked for) only needs the forward envelope to keep the activations in CB16
    train = torch.is_grad_enabled() and any(t.requires_grad for t in (x, *wb))
    return _NeighConsensusFn.apply(x, bool(symmetric), default_conv_mode() if mode is None else mode, train, *wb)


# ------------------------------------------------------------------------------------------- #
# maxpool4d / corr_to_matches
# ------------------------------------------------------------------------------------------- #
def maxpool4d(corr4d_hres: torch.Tensor, k_size: int = 4, index_dtype: torch.dtype = torch.int64):
    """maxpool4d (reference lib/model.py:177-191): batch-1 4D max-pool + window-relative arg-max
    offsets.  index_dtype=torch.uint8 writes 1-byte offsets (8x less index traffic)."""
    _need_cuda(corr4d_hres)
    if corr4d_hres.shape[0] != 1:
        raise RuntimeError("maxpool4d: batch size must be 1 (the reference mixes batch items otherwise)")
    x = corr4d_hres[0, 0].contiguous()
    h1, h2, h3, h4 = x.shape
    k = int(k_size)
    if any(s % k for s in x.shape):
        raise RuntimeError(f"maxpool4d: dims {tuple(x.shape)} not divisible by k_size {k}")
    shp = (1, 1, h1 // k, h2 // k, h3 // k, h4 // k)
    val = torch.empty(shp, device=x.device, dtype=x.dtype)
    idx = [torch.empty(shp, device=x.device, dtype=index_dtype) for _ in range(4)]
    nbytes = {torch.int64: 8, torch.uint8: 1}[index_dtype]
    with torch.cuda.device(x.device):
        _cabi.check(_cabi.lib().ncnet_maxpool4d_fwd(_ptr(x), _ptr(val), *[_ptr(t) for t in idx], h1, h2, h3, h4, k, nbytes,
                                                    _dt(x), _stream()), "maxpool4d")
    _count(1)
    return (val, idx[0], idx[1], idx[2], idx[3])


_LUT_CACHE = {}


def _axis_lut(lo: float, n: int, device) -> torch.Tensor:
    key = (lo, n, str(device))
    t = _LUT_CACHE.get(key)
    if t is None:     # float64 linspace cast to fp32, as the reference builds it (lib/point_tnf.py:16-27)
        t = torch.from_numpy(np.linspace(lo, 1, n)).float().to(device)
        _LUT_CACHE[key] = t
    return t


def corr_to_matches(corr4d, delta4d=None, k_size=1, do_softmax=False, scale="centered", return_indices=False,
                    invert_matching_direction=False):
    """corr_to_matches (reference lib/point_tnf.py:12-80)."""
    _need_cuda(corr4d)
    b, ch, f1, f2, f3, f4 = corr4d.shape
    x = corr4d.detach().contiguous()
    if scale == "centered":
        lo = -1.0
    elif scale == "positive":
        lo = 0.0
    else:
        raise ValueError(scale)
    dev = x.device
    luts = [_axis_lut(lo, f2 * k_size, dev), _axis_lut(lo, f1 * k_size, dev), _axis_lut(lo, f4 * k_size, dev),
            _axis_lut(lo, f3 * k_size, dev)]
    n = f1 * f2 if invert_matching_direction else f3 * f4
    fo = [torch.empty(b, n, device=dev, dtype=torch.float32) for _ in range(5)]
    io = [torch.empty(b, n, device=dev, dtype=torch.int64) for _ in range(4)]
    deltas = [None] * 4
    idx_bytes = 8
    if delta4d is not None:
        deltas = []
        for d in delta4d:
            if d.dtype not in (torch.int64, torch.uint8):
                d = d.long()
            deltas.append(d.contiguous())
        idx_bytes = 8 if deltas[0].dtype == torch.int64 else 1
        if any(d.dtype != deltas[0].dtype for d in deltas):
            raise TypeError("corr_to_matches: delta4d tensors must share one integer dtype")
    nscratch = _cabi.lib().ncnet_corr_to_matches_scratch_bytes(b, f1, f2, f3, f4, int(bool(invert_matching_direction)), _dt(x))
    scratch = torch.empty(nscratch, device=dev, dtype=torch.uint8) if nscratch else None
    with torch.cuda.device(dev):
        _cabi.check(_cabi.lib().ncnet_corr_to_matches(
            _ptr(x), b, f1, f2, f3, f4, int(bool(do_softmax)), int(bool(invert_matching_direction)), int(k_size),
            *[_ptr(d) for d in deltas], idx_bytes, *[_ptr(t) for t in luts], *[_ptr(t) for t in fo],
            *[_ptr(t) for t in io], _ptr(scratch), _dt(x), _stream()), "corr_to_matches")
    _count(4 if nscratch else 1)
    xa, ya, xb, yb, score = fo
    if x.dtype != torch.float32:
        score = score.to(x.dtype)
    if return_indices:
        return (xa, ya, xb, yb, score, io[0], io[1], io[2], io[3])
    return (xa, ya, xb, yb, score)
