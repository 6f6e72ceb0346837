"""Multi-GPU plumbing (one process per GPU, torch.distributed; NCCL on GPUs, gloo in CPU tests).

The reference has no parallelism; BASELINE.json defines two sharded modes:
  * training: image pairs sharded over the batch, one all-reduce of the 180 033-float head
    gradient per step (latency-bound: a single flat bucket);
  * large volumes: slabs along the first source axis with a k//2 halo per Conv4d layer
    (see ``slab_bounds`` / ``exchange_halo``).
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

import torch
import torch.distributed as dist


# ------------------------------------------------------------------------------------------- #
# batch sharding + gradient bucket
# ------------------------------------------------------------------------------------------- #
def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced [begin, end) of ``n`` items for ``rank``."""
    base, rem = divmod(n, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard_pairs(feature_A: torch.Tensor, feature_B: torch.Tensor, rank: int, world: int):
    """Split a global batch of pairs for data-parallel training.  Returns the rank's
    (source, target, negative source): the weak loss pairs target b with source (b+1) mod B of the
    GLOBAL batch (train.py:137 rolls the whole batch), so the negative sources are sliced from the
    globally rolled tensor and the sharded step reproduces the single-GPU gradients."""
    b = feature_A.shape[0]
    lo, hi = shard_range(b, rank, world)
    rolled = torch.roll(feature_A, shifts=-1, dims=0)
    return feature_A[lo:hi], feature_B[lo:hi], rolled[lo:hi]


class GradBucket:
    """All head gradients flattened into one fp32 buffer: one collective per step."""

    def __init__(self, params: Iterable[torch.nn.Parameter]):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        self.numel = sum(p.numel() for p in self.params)
        p0 = self.params[0]
        self.flat = torch.zeros(self.numel, device=p0.device, dtype=torch.float32)

    def _views(self):
        off, out = 0, []
        for p in self.params:
            n = p.numel()
            out.append(self.flat[off:off + n].view_as(p))
            off += n
        return out

    def pack(self) -> torch.Tensor:
        """One multi-tensor copy of all gradients into the flat buffer (a missing gradient counts as zero)."""
        views = self._views()
        have = [(v, p.grad) for v, p in zip(views, self.params) if p.grad is not None]
        for v, p in zip(views, self.params):
            if p.grad is None:
                v.zero_()
        if have:
            torch._foreach_copy_([v for v, _ in have], [g for _, g in have])
        return self.flat

    def unpack(self) -> None:
        views = self._views()
        for v, p in zip(views, self.params):
            if p.grad is None:
                p.grad = torch.empty_like(p)
        torch._foreach_copy_([p.grad for p in self.params], views)

    def all_reduce_mean(self, group=None, local_batch=None, global_batch=None) -> None:
        """grad <- gradient of the mean loss over the GLOBAL batch.  Each rank's loss is the mean over its own shard,
        so with equal shards this is the plain mean over ranks (the default); with unequal shards (batch not divisible
        by the world size) pass ``local_batch`` and ``global_batch``: the rank's gradient is weighted by its share."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return
        if (local_batch is None) != (global_batch is None):
            raise ValueError("all_reduce_mean: give both local_batch and global_batch, or neither")
        self.pack_weighted(group, local_batch, global_batch)
        self.all_reduce_flat(group)
        self.unpack()

    # the three phases on their own, for callers that replay the compute phases as CUDA graphs and keep the
    # collective outside the captures (bench.py)
    def pack_weighted(self, group=None, local_batch=None, global_batch=None) -> None:
        weight = 1.0 / dist.get_world_size(group) if local_batch is None else float(local_batch) / float(global_batch)
        self.pack()
        self.flat.mul_(weight)

    def all_reduce_flat(self, group=None) -> None:
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)


# ------------------------------------------------------------------------------------------- #
# slab sharding of a 4D volume along the first source axis (iA)
# ------------------------------------------------------------------------------------------- #
def slab_bounds(h: int, rank: int, world: int) -> Tuple[int, int]:
    return shard_range(h, rank, world)


def exchange_halo(slab: torch.Tensor, halo: int, rank: int, world: int, group=None) -> torch.Tensor:
    """``slab``: [B, C, rows, w, d, t] = this rank's rows of the volume.  Returns
    [B, C, halo_lo + rows + halo_hi, ...] with ``halo`` rows fetched from each existing neighbour
    (nothing is added at the outer edges: Conv4d zero-pads there).  Point-to-point only."""
    if world == 1 or halo == 0:
        return slab
    rows = slab.shape[2]
    if rows < halo:
        raise RuntimeError(f"exchange_halo: slab of {rows} rows is thinner than the halo {halo}")
    sends, recvs, recv_lo, recv_hi = [], [], None, None
    send_lo = slab[:, :, :halo].contiguous()
    send_hi = slab[:, :, rows - halo:].contiguous()
    if rank > 0:
        recv_lo = torch.empty_like(send_lo)
        sends.append((send_lo, rank - 1)), recvs.append((recv_lo, rank - 1))
    if rank < world - 1:
        recv_hi = torch.empty_like(send_hi)
        sends.append((send_hi, rank + 1)), recvs.append((recv_hi, rank + 1))
    _p2p(sends, recvs, group)
    parts = [t for t in (recv_lo, slab, recv_hi) if t is not None]
    return torch.cat(parts, dim=2)


def crop_halo(y: torch.Tensor, halo: int, rank: int, world: int) -> torch.Tensor:
    """Drop the rows that belong to the neighbours after a layer ran on a halo-extended slab."""
    lo = halo if rank > 0 else 0
    hi = y.shape[2] - (halo if rank < world - 1 else 0)
    return y[:, :, lo:hi].contiguous()


def all_reduce_max_(t: torch.Tensor, group=None) -> torch.Tensor:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return t


# ------------------------------------------------------------------------------------------- #
# slab-sharded head on the GPU kernels (BASELINE.json config 4: volumes too large for one GPU)
# ------------------------------------------------------------------------------------------- #
_SIGN64 = -(1 << 63)


def slab_mutual_matching(x_slab: torch.Tensor, group=None, sum_branches: bool = False) -> torch.Tensor:
    """MutualMatching (lib/model.py:155-175) of a volume sharded along iA.  ``x_slab``: this rank's rows
    [B,1|2,rows,fs2,fs3,fs4].  A slab holds whole source rows, so the max over B is local; the max over A of
    every target cell is an all-reduce(max) over ranks of the per-target key table between the two kernel passes."""
    from . import _cabi
    from .functional import _dt, _ptr, _stream, _count
    x = x_slab.contiguous()
    b, g, r, f2, f3, f4 = x.shape
    na, nb = r * f2, f3 * f4
    lib = _cabi.lib()
    keys = torch.empty(b, na + nb, device=x.device, dtype=torch.int64)
    with torch.cuda.device(x.device):
        _cabi.check(lib.ncnet_mutual_matching_max(_ptr(x), _ptr(keys), b, na, nb, g, _dt(x), _stream()), "mutual_matching_max")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        # keys order as UNSIGNED 64-bit; flipping the top bit makes the signed max agree
        colk = (keys[:, na:] ^ _SIGN64).contiguous()
        dist.all_reduce(colk, op=dist.ReduceOp.MAX, group=group)
        keys[:, na:] = colk ^ _SIGN64
    y = torch.empty(b, 1, r, f2, f3, f4, device=x.device, dtype=x.dtype)
    with torch.cuda.device(x.device):
        _cabi.check(lib.ncnet_mutual_matching_apply(_ptr(x), _ptr(keys), _ptr(y), b, na, nb, g, _dt(x), _stream()),
                    "mutual_matching_apply")
    _count(2)
    return y


def _p2p(sends, recvs, group=None):
    """One batch of point-to-point transfers: sends / recvs are lists of (tensor, peer).  NCCL moves device memory
    directly (NVLink); under gloo (the CPU-side tests, or several ranks sharing one GPU) CUDA tensors are staged through
    host memory because gloo has no device-to-device send."""
    if not sends and not recvs:
        return
    staged = dist.get_backend(group) == "gloo"
    ops, landing = [], []
    for t, peer in sends:
        ops.append(dist.P2POp(dist.isend, t.cpu() if (staged and t.is_cuda) else t, peer, group))
    for t, peer in recvs:
        if staged and t.is_cuda:
            h = torch.empty(t.shape, dtype=t.dtype)
            landing.append((t, h))
            ops.append(dist.P2POp(dist.irecv, h, peer, group))
        else:
            ops.append(dist.P2POp(dist.irecv, t, peer, group))
    for req in dist.batch_isend_irecv(ops):
        req.wait()
    for t, h in landing:
        t.copy_(h)


def _all_gather(outs, t, group=None):
    """dist.all_gather; under gloo CUDA tensors go through the host (gloo gathers CPU tensors only)."""
    if dist.get_backend(group) == "gloo" and t.is_cuda:
        houts = [torch.empty(o.shape, dtype=o.dtype) for o in outs]
        dist.all_gather(houts, t.cpu(), group=group)
        for o, h in zip(outs, houts):
            o.copy_(h)
    else:
        dist.all_gather(outs, t, group=group)


def exchange_halo_cb16_(act, halo: int, rank: int, world: int, group=None) -> int:
    """In-place halo exchange on a CB16 activation tensor (include/ncnet_b200.h: NCNET_CB16) that already has room for
    the neighbours' rows: its first-axis extent is halo_lo + rows + halo_hi.  A (block, channel-half) of CB16 is
    [d1][d2][haloed plane][8], so ``halo`` rows are ONE contiguous run per (block, half): the boundary rows go out of
    the tensor and the neighbours' rows land in it without any staging copy, concatenation or crop.
    Returns the bytes this rank received."""
    if world == 1 or halo == 0:
        return 0
    bsz, ch, d1e, d2, d3, d4 = act.shape
    pad = act.k // 2
    row_bytes = d2 * (d3 + 2 * pad) * (d4 + 2 * pad) * 16
    nrun = act.buf.numel() // (d1e * row_bytes) if not act.pm else None
    if act.pm:
        raise RuntimeError("exchange_halo_cb16_: CB16 tensors with a position-major twin are not exchanged")
    view = act.buf[: nrun * d1e * row_bytes].view(nrun, d1e, row_bytes)
    h_lo = halo if rank > 0 else 0
    h_hi = halo if rank < world - 1 else 0
    sends, recvs = [], []
    for i in range(nrun):
        if rank > 0:
            sends.append((view[i, h_lo:h_lo + halo], rank - 1))
            recvs.append((view[i, 0:h_lo], rank - 1))
        if rank < world - 1:
            sends.append((view[i, d1e - h_hi - halo:d1e - h_hi], rank + 1))
            recvs.append((view[i, d1e - h_hi:d1e], rank + 1))
    _p2p(sends, recvs, group)
    return sum(t.numel() for t, _ in recvs)


def slab_neigh_consensus(x_slab: torch.Tensor, nc, rank: int, world: int, group=None, stats=None) -> torch.Tensor:
    """NeighConsensus (lib/model.py:143-153) on a slab of rows.  The slab is extended ONCE by the halo of the whole
    stack's first layer (k//2 rows per existing neighbour; the 1-channel input rows are cheap to fetch); every layer
    then runs on that fixed extent.  A layer's output on the borrowed rows is computed from incomplete input and is
    wrong there -- before the next layer reads it, those rows are overwritten in place with the neighbour's (correct)
    rows (``exchange_halo_cb16_``).  In tensor-core mode the hidden activations stay in CB16 from launch to launch.
    Both symmetric branches stay slab-local (transposed-filter identity).
    Returns the un-summed branch outputs [B, G*C_last, rows, ...] like ``forward_branches``.
    ``stats`` (dict, optional) receives the halo bytes received per layer."""
    from .functional import _CB16, _conv_fwd, _stack_can_chain, default_conv_mode, W_KOIKKK
    ws, bs = nc._params()
    mode = default_conv_mode()
    ks = [w.shape[0] for w in ws]
    if len(set(ks)) != 1:
        raise RuntimeError("slab_neigh_consensus: one kernel size per stack")
    halo = ks[0] // 2
    rows = x_slab.shape[2]
    if world > 1 and rows < halo:
        raise RuntimeError(f"slab of {rows} rows is thinner than the halo {halo}")
    chain = _stack_can_chain(ks, [w.shape[1] for w in ws], x_slab.shape[4], x_slab.shape[5], mode, need_backward=False)
    cur = exchange_halo(x_slab.contiguous(), halo, rank, world, group)       # 1-channel input: rows fetched once
    received = [cur.numel() * cur.element_size() - x_slab.numel() * x_slab.element_size()]
    for l, (w, bvec) in enumerate(zip(ws, bs)):
        last = l == len(ws) - 1
        if l > 0:
            if isinstance(cur, _CB16):
                received.append(exchange_halo_cb16_(cur, halo, rank, world, group))
            else:                       # planar hidden activations (fp32 mode): crop + re-fetch
                own = crop_halo(cur, halo, rank, world)
                cur = exchange_halo(own, halo, rank, world, group)
                received.append((cur.numel() - own.numel()) * cur.element_size())
        cur = _conv_fwd(cur, w.detach().float().contiguous(), bvec.detach().float().contiguous(), W_KOIKKK,
                        nc.symmetric_mode, l == 0, True, mode, out_cb16=chain and not last)
    if stats is not None:
        stats["halo_bytes_received_per_layer"] = received
    return crop_halo(cur, halo, rank, world)


def slab_match_features(model, feature_A: torch.Tensor, feature_B: torch.Tensor, rank: int, world: int, group=None,
                        stats=None):
    """The head of ``ImMatchNet.match_features`` for this rank's rows of the volume: the correlation rows are
    computed locally from the (replicated) features, so only the halo rows and the target maxima cross NVLink."""
    lo, hi = slab_bounds(feature_A.shape[2], rank, world)
    corr = model.FeatureCorrelation(feature_A[:, :, lo:hi].contiguous(), feature_B)
    corr = slab_mutual_matching(corr, group)
    br = slab_neigh_consensus(corr, model.NeighConsensus, rank, world, group, stats)
    if model.NeighConsensus.symmetric_mode and br.shape[1] == 2:
        return slab_mutual_matching(br, group, sum_branches=True)
    c = br.shape[1] // 2 if model.NeighConsensus.symmetric_mode else br.shape[1]
    summed = br[:, :c] + br[:, c:] if model.NeighConsensus.symmetric_mode else br
    return slab_mutual_matching(summed, group)


def slab_corr_to_matches(corr_slab: torch.Tensor, fs1: int, rank: int, world: int, group=None, do_softmax: bool = False,
                         scale: str = "centered", invert_matching_direction: bool = False):
    """``corr_to_matches(..., return_indices=True)`` (lib/point_tnf.py:12-80, no relocalisation) for a volume sharded into
    slabs of source rows; every rank returns the matches of the WHOLE volume.
      * A->B (``invert_matching_direction``): each source cell's best target is local to its slab -- the per-rank
        results are gathered in rank order.
      * B->A: per target cell each rank contributes its local (max, first arg-max, sum of exp(x - max)) -- the column
        halves of ``ncnet_match_score_fwd``'s keys / sums -- and the ranks' candidates are merged: the winner is the
        largest maximum (lowest rank, i.e. lowest source row, on ties), the soft-max score is 1 / sum_r s_r exp(m_r - M).
    Equal to the single-GPU result except where two source cells of different slabs tie after the soft-max's rounding."""
    from . import _cabi
    from .functional import _dt, _ptr, _stream, _count, corr_to_matches, _axis_lut
    x = corr_slab.detach().contiguous()
    b, _, rows, f2, f3, f4 = x.shape
    lo, _hi = slab_bounds(fs1, rank, world)
    dev = x.device
    lo_v = -1.0 if scale == "centered" else 0.0
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1

    def gather(t):
        if not multi:
            return [t]
        counts = [(h - l) * f2 for l, h in (slab_bounds(fs1, r, world) for r in range(world))]
        padded = torch.zeros((b, max(counts)), device=dev, dtype=t.dtype)      # equal sizes for the collective
        padded[:, : t.shape[1]] = t
        outs = [torch.empty_like(padded) for _ in range(world)]
        _all_gather(outs, padded, group)
        return [o[:, :c] for o, c in zip(outs, counts)]

    if invert_matching_direction:
        m = corr_to_matches(x, do_softmax=do_softmax, scale=scale, return_indices=True, invert_matching_direction=True)
        ia = m[5] + lo                                        # local source row -> global
        ya = _axis_lut(lo_v, fs1, dev)[ia]
        parts = [gather(t) for t in (m[0], ya, m[2], m[3], m[4], ia, m[6], m[7], m[8])]
        return tuple(torch.cat(p, 1) for p in parts)
    na, nb = rows * f2, f3 * f4
    lib = _cabi.lib()
    keys = torch.empty(b * (na + nb), device=dev, dtype=torch.int64)
    sums = torch.empty(lib.ncnet_mutual_matching_sums_bytes(b, na, nb) // 4, device=dev, dtype=torch.float32)
    score = torch.empty((), device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        _cabi.check(lib.ncnet_match_score_fwd(_ptr(x), _ptr(keys), _ptr(sums), None, _ptr(score), b, na, nb, _dt(x), _stream()),
                    "match_score_fwd")
    _count(4)
    ck = keys.view(b, na + nb)[:, na:]                        # per target cell: (ordered max bits << 32) | ~first row
    hi32 = (ck >> 32) & 0xFFFFFFFF
    bits = torch.where(hi32 >= 0x80000000, hi32 - 0x80000000, 0xFFFFFFFF - hi32)       # inverse of common.cuh f2ord
    m_loc = (bits & 0xFFFFFFFF).to(torch.int32).view(torch.float32)
    a_loc = 0xFFFFFFFF - (ck & 0xFFFFFFFF)                    # local flat source index
    s_loc = sums.view(-1)[: b * (na + nb)].view(b, na + nb)[:, na:]
    ia_g = a_loc // f2 + lo
    ja_g = a_loc % f2
    cand = torch.stack([m_loc.double(), s_loc.double(), ia_g.double(), ja_g.double()], 0).contiguous()   # [4,b,nb]
    if multi:
        allc = [torch.empty_like(cand) for _ in range(world)]
        _all_gather(allc, cand, group)
        allc = torch.stack(allc, 0)                           # [world,4,b,nb]
    else:
        allc = cand[None]
    mr, sr = allc[:, 0], allc[:, 1]
    M, win = mr.max(0)                                        # first (lowest-rank) maximum wins
    S = (sr * torch.exp((mr - M[None]).float()).double()).float().sum(0) if do_softmax else None
    pick = lambda q: torch.gather(allc[:, q], 0, win[None])[0]
    ia, ja = pick(2).long(), pick(3).long()
    own = torch.arange(nb, device=dev)[None].expand(b, -1)
    ib, jb = own // f4, own % f4
    sc = (1.0 / S) if do_softmax else M.float()
    if x.dtype != torch.float32:
        sc = sc.to(x.dtype)
    luts = [_axis_lut(lo_v, n, dev) for n in (f2, fs1, f4, f3)]
    return (luts[0][ja], luts[1][ia], luts[2][jb], luts[3][ib], sc, ia, ja, ib, jb)
