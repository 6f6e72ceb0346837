"""ncu driver: NC stack forward+backward in tensor-core mode at the PF shape (B from argv)."""
import sys, time
import torch
sys.path.insert(0, ".")
import ncnet_b200 as nb
from ncnet_b200 import functional as F_
from ncnet_b200._cabi import CONV_TC16
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
torch.manual_seed(0)
nc = nb.NeighConsensus(True, [5, 5, 5], [16, 16, 1])
ws, bs = nc._params()
x = torch.rand(B, 1, 25, 25, 25, 25, device="cuda")
for it in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    y = F_.neigh_consensus_branches(x, ws, bs, True, CONV_TC16)
    y.sum().backward()
    torch.cuda.synchronize()
    print("fwd+bwd ms", (time.perf_counter() - t0) * 1e3, flush=True)
