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
