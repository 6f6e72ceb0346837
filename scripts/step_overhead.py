"""Development: host enqueue time vs device time of one training step of the head at a small per-rank batch
(the 8-GPU run gives each rank 2 pairs).  usage: python scripts/step_overhead.py [batch]"""
import sys, time
import torch
sys.path.insert(0, ".")
import ncnet_b200 as nb
from ncnet_b200.train_util import match_score
B = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda")
torch.manual_seed(0)
net = nb.ImMatchNet.__new__(nb.ImMatchNet)
torch.nn.Module.__init__(net)
net.relocalization_k_size, net.half_precision = 0, False
net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
net.NeighConsensus = nb.NeighConsensus(True, [5, 5, 5], [16, 16, 1]).to(dev)
opt = torch.optim.Adam(net.NeighConsensus.parameters(), lr=5e-4)
fa = torch.rand(B, 1024, 25, 25, device=dev); fb = torch.rand(B, 1024, 25, 25, device=dev); fn = torch.roll(fa, -1, 0)

def step():
    opt.zero_grad(set_to_none=True)
    loss = match_score(net.match_features(fn, fb)) - match_score(net.match_features(fa, fb))
    loss.backward()
    opt.step()

for _ in range(5):
    step()
torch.cuda.synchronize()
n = 20
t0 = time.perf_counter()
for _ in range(n):
    step()
t_enq = time.perf_counter() - t0
torch.cuda.synchronize()
t_all = time.perf_counter() - t0
print(f"batch {B}: host enqueue {t_enq / n * 1e3:.2f} ms/step, wall {t_all / n * 1e3:.2f} ms/step")
