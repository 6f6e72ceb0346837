"""torchrun, N GPUs: the slab-sharded head (halo exchange + cross-slab maxima) against the whole-volume head."""
import os, sys, time
import torch, torch.distributed as dist
sys.path.insert(0, ".")
import ncnet_b200 as nb
from ncnet_b200 import dist as D

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
worst = 0.0
for half, ks, chs, shape in ((False, [3, 3], [16, 1], (16, 10, 14, 12)), (True, [3, 3], [16, 1], (24, 12, 20, 18)),
                             (False, [5, 5, 5], [16, 16, 1], (16, 9, 11, 10))):
    torch.manual_seed(3)
    fa = torch.relu(torch.randn(1, 64, shape[0], shape[1], device=dev)); fa = fa / (fa.pow(2).sum(1, keepdim=True) + 1e-6).sqrt()
    fb = torch.relu(torch.randn(1, 64, shape[2], shape[3], device=dev)); fb = fb / (fb.pow(2).sum(1, keepdim=True) + 1e-6).sqrt()
    net = nb.ImMatchNet.__new__(nb.ImMatchNet); torch.nn.Module.__init__(net)
    net.relocalization_k_size, net.half_precision = 0, half
    net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
    net.NeighConsensus = nb.NeighConsensus(True, ks, chs)
    for p in net.NeighConsensus.parameters():
        dist.broadcast(p.data, 0)
    if half:
        fa, fb = fa.half(), fb.half(); net.NeighConsensus.half()
    with torch.no_grad():
        full = net.match_features(fa, fb)
        slab = D.slab_match_features(net, fa, fb, rank, world)
    lo, hi = D.slab_bounds(shape[0], rank, world)
    err = ((slab.float() - full[:, :, lo:hi].float()).abs().max() / full.float().abs().max()).item()
    t = torch.tensor([err], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    worst = max(worst, t.item())
    if rank == 0:
        print(f"slab vs whole volume, half={half} ks={ks}: max normalised error over ranks {t.item():.2e}", flush=True)
# timing at a larger InLoc-like volume (pooled 3200px: 100x75x100x75), fp16
torch.manual_seed(4)
fa = torch.rand(1, 256, 100, 75, device=dev).half(); fb = torch.rand(1, 256, 100, 75, device=dev).half()
net = nb.ImMatchNet.__new__(nb.ImMatchNet); torch.nn.Module.__init__(net)
net.relocalization_k_size, net.half_precision = 0, True
net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
net.NeighConsensus = nb.NeighConsensus(True, [3, 3], [16, 1]).half()
with torch.no_grad():
    for it in range(3):
        dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
        out = D.slab_match_features(net, fa, fb, rank, world)
        torch.cuda.synchronize(); dist.barrier(); dt = time.perf_counter() - t0
if rank == 0:
    print(f"IL32-pooled volume 100x75x100x75 fp16 over {world} slab(s): {dt*1e3:.1f} ms per pair", flush=True)
    print("SLAB_OK" if worst < 5e-3 else "SLAB_MISMATCH", flush=True)
dist.destroy_process_group()
