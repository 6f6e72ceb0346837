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

    def pack(self) -> torch.Tensor:
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is None:
                self.flat[off:off + n].zero_()
            else:
                self.flat[off:off + n].copy_(p.grad.reshape(-1))
            off += n
        return self.flat

    def unpack(self) -> None:
        off = 0
        for p in self.params:
            n = p.numel()
            g = self.flat[off:off + n].view_as(p).to(p.dtype)
            if p.grad is None:
                p.grad = g.clone()
            else:
                p.grad.copy_(g)
            off += n

    def all_reduce_mean(self, group=None) -> None:
        """grad <- mean over ranks (each rank's loss is the mean over its equal-sized shard)."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return
        self.pack()
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        self.flat.div_(dist.get_world_size(group))
        self.unpack()


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
    ops, recv_lo, recv_hi = [], None, None
    send_lo = slab[:, :, :halo].contiguous()
    send_hi = slab[:, :, rows - halo:].contiguous()
    if rank > 0:
        recv_lo = torch.empty_like(send_lo)
        ops += [dist.P2POp(dist.isend, send_lo, rank - 1, group), dist.P2POp(dist.irecv, recv_lo, rank - 1, group)]
    if rank < world - 1:
        recv_hi = torch.empty_like(send_hi)
        ops += [dist.P2POp(dist.isend, send_hi, rank + 1, group), dist.P2POp(dist.irecv, recv_hi, rank + 1, group)]
    for req in dist.batch_isend_irecv(ops):
        req.wait()
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


def slab_neigh_consensus(x_slab: torch.Tensor, nc, rank: int, world: int, group=None) -> torch.Tensor:
    """NeighConsensus (lib/model.py:143-153) on a slab: before every Conv4d layer the k//2 boundary rows are
    exchanged with the neighbouring ranks (point-to-point), the layer runs on the extended slab and the borrowed
    rows are cropped.  Both symmetric branches stay slab-local (transposed-filter identity).
    Returns the un-summed branch outputs [B, G*C_last, rows, ...] like ``forward_branches``."""
    from .functional import _conv_fwd, default_conv_mode, W_KOIKKK
    ws, bs = nc._params()
    mode = default_conv_mode()
    cur = x_slab.contiguous()
    for l, (w, bvec) in enumerate(zip(ws, bs)):
        halo = w.shape[0] // 2
        ext = exchange_halo(cur, halo, rank, world, group)
        y = _conv_fwd(ext, w.detach().float().contiguous(), bvec.detach().float().contiguous(), W_KOIKKK,
                      nc.symmetric_mode, l == 0, True, mode)
        cur = crop_halo(y, halo, rank, world)
    return cur


def slab_match_features(model, feature_A: torch.Tensor, feature_B: torch.Tensor, rank: int, world: int, group=None):
    """The head of ``ImMatchNet.match_features`` for this rank's rows of the volume: the correlation rows are
    computed locally from the (replicated) features, so only the halo rows and the target maxima cross NVLink."""
    lo, hi = slab_bounds(feature_A.shape[2], rank, world)
    corr = model.FeatureCorrelation(feature_A[:, :, lo:hi].contiguous(), feature_B)
    corr = slab_mutual_matching(corr, group)
    br = slab_neigh_consensus(corr, model.NeighConsensus, rank, world, group)
    if model.NeighConsensus.symmetric_mode and br.shape[1] == 2:
        return slab_mutual_matching(br, group, sum_branches=True)
    c = br.shape[1] // 2 if model.NeighConsensus.symmetric_mode else br.shape[1]
    summed = br[:, :c] + br[:, c:] if model.NeighConsensus.symmetric_mode else br
    return slab_mutual_matching(summed, group)
