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
x = F_._conv_fwd(x, w, b, W_KOIKKK, True, False, True, CONV_TC16, out_cb16=True) if ("cb16" in sys.argv and "thin1" not in sys.argv) else x
CB = "cb16" in sys.argv
if CB:
    w = (torch.rand(5, 16, 16, 5, 5, 5, device="cuda") * 2 - 1) * 0.01
THIN = "thin" in sys.argv          # the 16 -> 1 layer (planar output) on the CB16 activations
if THIN:
    w = (torch.rand(5, 1, 16, 5, 5, 5, device="cuda") * 2 - 1) * 0.01
    b = torch.zeros(1, device="cuda")
    CB = False
THIN1 = "thin1" in sys.argv        # the 1 -> 16 layer (planar single-channel input, CB16 output)
if THIN1:
    x = torch.rand(B, 1, 25, 25, 25, 25, device="cuda")
    w = (torch.rand(5, 16, 1, 5, 5, 5, device="cuda") * 2 - 1) * 0.01
    b = torch.zeros(16, device="cuda")
    CB = True
F_._conv_fwd(x, w, b, W_KOIKKK, True, THIN1, True, CONV_TC16, out_cb16=CB)
buf = torch.zeros(4096 * 8, dtype=torch.int64, device="cuda")
lib = _cabi.lib()
lib.ncnet_debug_set_tc_timing.argtypes = [ctypes.c_void_p]
MODE = next((int(a[5:]) for a in sys.argv if a.startswith("mode=")), 0)
lib.ncnet_debug_set_tc_mode.argtypes = [ctypes.c_int]
lib.ncnet_debug_set_tc_mode(MODE)
lib.ncnet_debug_set_tc_timing(buf.data_ptr())
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
F_._conv_fwd(x, w, b, W_KOIKKK, True, THIN1, True, CONV_TC16, out_cb16=CB)
e1.record()
torch.cuda.synchronize()
print("mode", MODE, "call ms", e0.elapsed_time(e1))
lib.ncnet_debug_set_tc_mode(0)
lib.ncnet_debug_set_tc_timing(None)
t = buf.view(4096, 8).cpu().double()
names = ["first_full", "mma_issued_all", "accum_ready", "shift_sums_done", "store_done"]
for sel, lab in ((slice(0, 148), "first wave"), (slice(1000, 3000), "steady state")):
    print(lab, {n: int(t[sel, i].median()) for i, n in enumerate(names)})

# per-SM timeline: fraction of the SM's span during which at least one / both resident CTAs were in their MMA phase
import collections
by_sm = collections.defaultdict(list)
for r in t.tolist()[:3990]:
    if r[5] == 0: continue
    by_sm[int(r[6])].append((r[5], r[5] + r[0], r[5] + r[1], r[5] + r[2], r[5] + r[4]))
tot_span = tot_any = tot_mma = tot_life = 0.0
for sm, L in by_sm.items():
    L.sort()
    L = L[2:-2] if len(L) > 8 else L
    span = max(e[4] for e in L) - min(e[0] for e in L)
    ev = []
    for s0, f, m1, a, e in L:
        ev += [(f, 1), (a, -1)]
        tot_mma += a - f
        tot_life += e - s0
    ev.sort()
    depth, last, anyt = 0, None, 0.0
    for tt, d in ev:
        if depth > 0: anyt += tt - last
        depth += d; last = tt
    tot_span += span; tot_any += anyt
print("SMs", len(by_sm), "pipe-occupied fraction (>=1 CTA between first_full and accum_ready)", round(tot_any / tot_span, 3),
      " sum(MMA phase)/span", round(tot_mma / tot_span, 3), " sum(CTA life)/span", round(tot_life / tot_span, 3))

flat = buf.cpu().tolist()
print("CTA 2000 stamps :", [int(v) for v in flat[32000:32025]])
print("CTA 2000 full arrival   :", [int(v) for v in flat[32100:32125]])
print("CTA 2000 mma issued     :", [int(v) for v in flat[32200:32225]])
