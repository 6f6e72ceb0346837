import sys, torch
sys.path.insert(0, ".")
import ncnet_b200 as nb
x = torch.rand(1, 1, 100, 75, 100, 75, device="cuda")
xh = x.half()
with torch.no_grad():
    for _ in range(2):
        nb.MutualMatching(x); nb.MutualMatching(xh)
        nb.corr_to_matches(x, do_softmax=True); nb.corr_to_matches(x, do_softmax=True, invert_matching_direction=True)
torch.cuda.synchronize()
print("ok")
