"""Small driver for ncu: one tensor-core forward and one tensor-core wgrad of the 16->16, k=5 layer at the PF shape."""
import sys, time
import torch
sys.path.insert(0, ".")
from ncnet_b200 import functional as F_
from ncnet_b200._cabi import CONV_TC16, W_KOIKKK
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
torch.manual_seed(0)
x = torch.rand(B, 32, 25, 25, 25, 25, device="cuda")
gy = torch.randn(B, 32, 25, 25, 25, 25, device="cuda")
w = (torch.rand(5, 16, 16, 5, 5, 5, device="cuda") * 2 - 1) * 0.01
b = torch.zeros(16, device="cuda")
for name, fn in (("fwd", lambda: F_._conv_fwd(x, w, b, W_KOIKKK, True, False, True, CONV_TC16)),
                 ("wgrad", lambda: F_._conv_wgrad(x, gy, w, True, W_KOIKKK, True, False, CONV_TC16))):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    print(name, (time.perf_counter() - t0) / 3 * 1e3, "ms", flush=True)
