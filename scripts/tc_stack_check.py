"""Development check: the chained (CB16) tensor-core NeighConsensus stack against the fp32 CUDA-core stack."""
import sys, time
import torch
sys.path.insert(0, ".")
import ncnet_b200 as nb
from ncnet_b200 import functional as F_
from ncnet_b200._cabi import CONV_FP32, CONV_TC16

def nerr(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30)).item()

def run(ks, chs, dims, bsz):
    torch.manual_seed(1)
    nc = nb.NeighConsensus(True, ks, chs)
    ws, bs = nc._params()
    x = torch.rand(bsz, 1, *dims, device="cuda").requires_grad_(True)
    outs = {}
    for mode in (CONV_FP32, CONV_TC16):
        for p in nc.parameters():
            p.grad = None
        y = F_.neigh_consensus_branches(x, ws, bs, True, mode)
        go = torch.rand_like(y)
        torch.manual_seed(2)
        go = torch.randn(y.shape, device="cuda")
        grads = torch.autograd.grad(y, [x] + list(nc.parameters()), go)
        outs[mode] = (y.detach(), grads)
    torch.cuda.synchronize()
    e = nerr(outs[1][0], outs[0][0])
    eg = [nerr(a, b) for a, b in zip(outs[1][1], outs[0][1])]
    print(f"ks={ks} ch={chs} dims={dims} B={bsz}: fwd {e:.2e}  gx {eg[0]:.2e}  params " + " ".join(f"{v:.1e}" for v in eg[1:]), flush=True)

run([5, 5, 5], [16, 16, 1], (7, 6, 9, 10), 2)
run([3, 3], [16, 1], (6, 5, 20, 13), 1)
run([3, 3, 3], [10, 10, 1], (5, 6, 8, 9), 2)
run([5, 5, 5], [16, 16, 1], (25, 25, 25, 25), 1)

# training-step timing at PF shape, B=16
nc = nb.NeighConsensus(True, [5, 5, 5], [16, 16, 1])
ws, bs = nc._params()
x = torch.rand(16, 1, 25, 25, 25, 25, device="cuda")
for mode in (CONV_TC16,):
    for it in range(3):
        if it == 1:
            torch.cuda.synchronize(); t0 = time.perf_counter()
        y = F_.neigh_consensus_branches(x, ws, bs, True, mode)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        y.sum().backward()
        torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 2
    print(f"mode {mode}: NC stack fwd+bwd B=16: {dt*1e3:.1f} ms per evaluation (last fwd {1e3*(t1-t0)/2:.1f})", flush=True)
