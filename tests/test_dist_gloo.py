"""CPU, world_size 2, gloo: the host-side multi-GPU logic of ncnet_b200/dist.py.

  * batch sharding reproduces the single-process weak-loss gradient after one flat all-reduce
    (checked with the CPU oracle head as the model; the CUDA head is exercised by `bench.py --gpus N`);
  * slab sharding with a k//2 halo exchange reproduces a Conv4d layer computed on the whole volume;
  * all-reduce(max) of the per-target column maximum reproduces MutualMatching on slabs.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import head as O


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, case, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ncnet_b200 import dist as D
        torch.manual_seed(0)
        torch.set_num_threads(1)
        if case == "grad":
            B, C, fs = 4, 16, 5
            fa = O.feature_l2norm(torch.relu(torch.randn(B, C, fs, fs)))
            fb = O.feature_l2norm(torch.relu(torch.randn(B, C, fs, fs)))
            nc = O.NeighConsensus([3, 3], [4, 1]).double()
            ref = O.NeighConsensus([3, 3], [4, 1]).double()
            ref.load_state_dict(nc.state_dict())
            O.weak_loss(fa.double(), fb.double(), ref).backward()          # single process, whole batch
            src, tgt, neg = D.shard_pairs(fa.double(), fb.double(), rank, world)
            pos = O.match_score(O.head_forward(src, tgt, nc))
            ngs = O.match_score(O.head_forward(neg, tgt, nc))
            (ngs - pos).backward()
            bucket = D.GradBucket(nc.parameters())
            assert bucket.numel == sum(p.numel() for p in nc.parameters())
            bucket.all_reduce_mean()
            err = max((p.grad - q.grad).abs().max().item() / q.grad.abs().max().item()
                      for p, q in zip(nc.parameters(), ref.parameters()))
            ret[rank] = err
        elif case == "slab":
            x = torch.randn(1, 3, 8, 4, 5, 6, dtype=torch.float64)
            layer = O.Conv4d(3, 2, 3).double()
            full = layer(x)
            lo, hi = D.slab_bounds(8, rank, world)
            ext = D.exchange_halo(x[:, :, lo:hi].contiguous(), 1, rank, world)
            # rows added by the exchange are real neighbour rows; the outer edges stay zero-padded by conv4d itself
            y = D.crop_halo(layer(ext), 1, rank, world)
            ret[rank] = (y - full[:, :, lo:hi]).abs().max().item()
        elif case == "mm":
            v = torch.rand(2, 1, 6, 3, 4, 5, dtype=torch.float64)
            full = O.mutual_matching(v)
            lo, hi = D.slab_bounds(6, rank, world)
            s = v[:, :, lo:hi]
            b, _, f1, f2, f3, f4 = s.shape
            m = s.reshape(b, f1 * f2, f3 * f4)
            col = D.all_reduce_max_(m.max(dim=1, keepdim=True)[0].contiguous())     # max over ALL source rows
            row = m.max(dim=2, keepdim=True)[0]                                     # local: a slab holds whole rows
            y = (m * ((m / (row + 1e-5)) * (m / (col + 1e-5)))).reshape(s.shape)
            ret[rank] = (y - full[:, :, lo:hi]).abs().max().item()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case,tol", [("grad", 1e-6), ("slab", 1e-12), ("mm", 1e-12)])
def test_world2_gloo(case, tol):
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), case, ret), nprocs=world, join=True)
    assert len(ret) == world
    assert all(v <= tol for v in ret.values()), dict(ret)


def test_shard_range_covers_everything():
    from ncnet_b200.dist import shard_range
    for n in (1, 7, 16, 25):
        for w in (1, 2, 3, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(e - s for s, e in spans) - min(e - s for s, e in spans) <= 1
