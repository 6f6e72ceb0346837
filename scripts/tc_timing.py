"""Development: per-CTA phase timing of conv4d_tc_kernel (clock64 stamps)."""
import sys, ctypes
import torch
sys.path.insert(0, ".")
from ncnet_b200 import functional as F_, _cabi
from ncnet_b200._cabi import CONV_TC16, W_KOIKKK
B = 4
x = torch.rand(B, 32, 25, 25, 25, 25, device="cuda")
w = (torch.rand(5, 16, 16, 5, 5, 5, device="cuda") * 2 - 1) * 0.01
b = torch.zeros(16, device="cuda")
F_._conv_fwd(x, w, b, W_KOIKKK, True, False, True, CONV_TC16)
buf = torch.zeros(4096 * 8, dtype=torch.int64, device="cuda")
lib = _cabi.lib()
lib.ncnet_debug_set_tc_timing.argtypes = [ctypes.c_void_p]
lib.ncnet_debug_set_tc_timing(buf.data_ptr())
F_._conv_fwd(x, w, b, W_KOIKKK, True, False, True, CONV_TC16)
torch.cuda.synchronize()
lib.ncnet_debug_set_tc_timing(None)
t = buf.view(4096, 8).cpu().double()
names = ["first_full", "mma_issued_all", "accum_ready", "shift_sums_done", "store_done"]
for sel, lab in ((slice(0, 148), "first wave"), (slice(1000, 3000), "steady state")):
    print(lab, {n: int(t[sel, i].median()) for i, n in enumerate(names)})
