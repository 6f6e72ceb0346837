"""torchrun, N ranks: multi-GPU paths of ncnet_b200/dist.py on the CUDA head against their single-GPU results.

  1. slab-sharded head (halo exchange in CB16, cross-slab maxima) == whole-volume head, bit for bit;
  2. cross-slab corr_to_matches merge == corr_to_matches of the whole volume (both directions, soft-max on/off);
  3. batch-sharded weak-loss gradient after the flat all-reduce == the unsharded gradient (fp32 reduction noise);
  4. (--time) the slab head at a large volume: ms per pair, halo bytes per layer, achieved link bandwidth.

Backend: NCCL when every rank has its own GPU, else gloo with all ranks on GPU 0 (point-to-point transfers staged
through the host) -- that is how the 2-rank case runs inside `pytest -m gpu` on a one-GPU box.
usage: torchrun --nproc-per-node N scripts/slab_check.py [--json out.json] [--time h,w,C]"""
import argparse, json, os, sys, time
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ncnet_b200 as nb
from ncnet_b200 import dist as D
from ncnet_b200.train_util import weak_loss_stacked

ap = argparse.ArgumentParser()
ap.add_argument("--json", default="")
ap.add_argument("--time", default="")
args = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
own_gpu = torch.cuda.device_count() >= world
dev = torch.device("cuda", local if own_gpu else 0)
torch.cuda.set_device(dev)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
if own_gpu:
    dist.init_process_group("nccl", device_id=dev)
else:
    dist.init_process_group("gloo")
report = {"world": world, "backend": dist.get_backend(), "cases": []}


def head(ks, chs, half):
    net = nb.ImMatchNet.__new__(nb.ImMatchNet); torch.nn.Module.__init__(net)
    net.relocalization_k_size, net.half_precision = 0, half
    net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
    net.NeighConsensus = nb.NeighConsensus(True, ks, chs)
    for p in net.NeighConsensus.parameters():
        dist.broadcast(p.data, 0)
    if half:
        net.NeighConsensus.half()
    return net


def feats(c, h, w, seed, half):
    g = torch.Generator().manual_seed(seed)
    f = torch.relu(torch.randn(1, c, h, w, generator=g))
    f = (f / (f.pow(2).sum(1, keepdim=True) + 1e-6).sqrt()).to(dev)
    return f.half() if half else f


ok = True
for half, ks, chs, shape in ((False, [3, 3], [16, 1], (16, 10, 14, 12)), (True, [3, 3], [16, 1], (24, 12, 20, 18)),
                             (False, [5, 5, 5], [16, 16, 1], (16, 9, 11, 10)), (True, [3, 3], [16, 1], (13, 12, 20, 18))):
    net = head(ks, chs, half)
    fa, fb = feats(64, shape[0], shape[1], 3, half), feats(64, shape[2], shape[3], 4, half)
    with torch.no_grad():
        full = net.match_features(fa, fb)
        slab = D.slab_match_features(net, fa, fb, rank, world)
        lo, hi = D.slab_bounds(shape[0], rank, world)
        same = torch.equal(slab, full[:, :, lo:hi])
        err = ((slab.float() - full[:, :, lo:hi].float()).abs().max() / full.float().abs().max()).item()
        match_ok, detail = True, []
        for inv in (False, True):
            for sm in (False, True):
                ref = nb.corr_to_matches(full, do_softmax=sm, return_indices=True, invert_matching_direction=inv)
                got = D.slab_corr_to_matches(slab, shape[0], rank, world, do_softmax=sm, invert_matching_direction=inv)
                idx_same = torch.stack([a == b for a, b in zip(got[5:], ref[5:])]).all(0)
                xy_same = torch.stack([a == b for a, b in zip(got[:4], ref[:4])]).all(0)
                sc_err = ((got[4].float() - ref[4].float()).abs().max() / ref[4].float().abs().max()).item()
                # fp32: exact.  fp16 with soft-max: the one-GPU kernel takes the first source cell whose ROUNDED half
                # soft-max equals the maximum, the merge takes the first raw maximum -- they differ on rounding ties
                need = 0.9 if (half and sm) else 1.0
                agree = lambda m: bool(m.all()) if need == 1.0 else m.float().mean().item() >= need
                ok_here = agree(idx_same) and agree(xy_same) and sc_err < (2e-3 if half else 1e-5)
                if not ok_here:
                    detail.append({"invert": inv, "softmax": sm, "idx_agree": idx_same.float().mean().item(),
                                   "xy_agree": xy_same.float().mean().item(), "score_nerr": sc_err})
                match_ok &= ok_here
    t = torch.tensor([err, 0.0 if same else 1.0, 0.0 if match_ok else 1.0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    case = {"case": f"slab head, half={half} ks={ks} volume={shape}", "max_nerr_over_ranks": t[0].item(),
            "bit_identical": t[1].item() == 0, "corr_to_matches_merge_ok": t[2].item() == 0, "mismatch_detail_rank%d" % rank: detail}
    ok &= case["bit_identical"] and case["corr_to_matches_merge_ok"]
    report["cases"].append(case)
    if rank == 0:
        print(case, flush=True)

# batch-sharded training gradient vs the unsharded one (PF-shaped volume, C = 128, TC16 mode)
B = 2 * world
g = torch.Generator().manual_seed(7)
fa = torch.relu(torch.randn(B, 128, 25, 25, generator=g)); fa = (fa / (fa.pow(2).sum(1, keepdim=True) + 1e-6).sqrt()).to(dev)
fb = torch.relu(torch.roll(fa.cpu(), (1, 2), (2, 3)) + 0.5 * torch.randn(B, 128, 25, 25, generator=g))
fb = (fb / (fb.pow(2).sum(1, keepdim=True) + 1e-6).sqrt()).to(dev)
net = head([5, 5, 5], [16, 16, 1], False)
params = list(net.NeighConsensus.parameters())
src, tgt, neg = D.shard_pairs(fa, fb, rank, world)
weak_loss_stacked(net, src, neg, tgt).backward()
D.GradBucket(params).all_reduce_mean()
sharded = torch.cat([p.grad.reshape(-1) for p in params]).clone()
net.zero_grad(set_to_none=True)
weak_loss_stacked(net, fa, torch.roll(fa, -1, 0), fb).backward()
whole = torch.cat([p.grad.reshape(-1) for p in params])
gerr = torch.tensor([((sharded - whole).abs().max() / whole.abs().max()).item()], device=dev)
dist.all_reduce(gerr, op=dist.ReduceOp.MAX)
case = {"case": f"weak-loss gradient, batch {B} sharded over {world} ranks vs unsharded (CUDA head, {os.environ.get('NCNET_CONV_MODE', 'tc16')})",
        "max_nerr_over_ranks": gerr.item()}
ok &= gerr.item() < 1e-3
report["cases"].append(case)
if rank == 0:
    print(case, flush=True)

if args.time:
    h, w, c = (int(v) for v in args.time.split(","))
    net = head([3, 3], [16, 1], True)
    fa, fb = feats(c, h, w, 5, True), feats(c, h, w, 6, True)
    best, stats = None, {}
    with torch.no_grad():
        for it in range(4):
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = D.slab_match_features(net, fa, fb, rank, world, stats=stats)
            m = D.slab_corr_to_matches(out, h, rank, world, do_softmax=True, scale="positive")
            e1.record(); torch.cuda.synchronize()
            tt = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            if it > 0:
                best = tt.item() if best is None else min(best, tt.item())
        # the halo exchange of the 16-channel layer alone: bytes / time = achieved link bandwidth
        from ncnet_b200.functional import _CB16
        lo, hi = D.slab_bounds(h, rank, world)
        ext = (hi - lo) + (1 if rank > 0 else 0) + (1 if rank < world - 1 else 0)
        act = _CB16.empty((1, 32, ext, w, h, w), 3, dev, torch.float16, 16)
        act.buf.zero_()
        nbytes, tx = 0, None
        for it in range(4):
            dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
            nbytes = D.exchange_halo_cb16_(act, 1, rank, world)
            torch.cuda.synchronize(); dt = time.perf_counter() - t0
            tt = torch.tensor([dt], device=dev); dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            if it > 0:
                tx = tt.item() if tx is None else min(tx, tt.item())
    nb_t = torch.tensor([float(nbytes)], device=dev); dist.all_reduce(nb_t, op=dist.ReduceOp.MAX)
    report["timing"] = {"volume": [h, w, h, w], "channels": c, "dtype": "fp16", "stack": "NC 3-3 / 16-1, no pooling",
                        "ms_per_pair_max_over_ranks": best, "halo_bytes_received_per_layer_rank0": stats.get("halo_bytes_received_per_layer"),
                        "halo_exchange_16ch_ms": tx * 1e3, "halo_exchange_16ch_bytes_received_per_rank_max": nb_t.item(),
                        "halo_exchange_GBs_per_rank_receive": nb_t.item() / tx / 1e9,
                        "rows_per_rank": hi - lo, "limiting_collective": "none: the two halo exchanges move ~%.0f MB per rank per pair" % (nb_t.item() / 1e6)}
    if rank == 0:
        print(report["timing"], flush=True)
report["ok"] = bool(ok)
if rank == 0:
    print("SLAB_OK" if ok else "SLAB_MISMATCH", flush=True)
    if args.json:
        with open(args.json, "w") as fh:
            json.dump(report, fh, indent=1)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
