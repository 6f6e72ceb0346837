"""A/B check of the 16->16 weight-gradient kernels: the walk along jA (grad-plane ring in shared memory, X tile shared by the
k taps through the A collector) against the per-plane-triple kernel (NCNET_NO_PMW=1), same inputs, planar operands.
Same fp16 products, fp32 accumulation; only the summation order differs: with a positive output gradient the sums are
coherent and the two agree to ~1e-5 on short sums; the tensor core adds into its accumulators with truncation, so on
long sums both drift low in proportion to their MMA chain lengths (equal by design, see kPwFlushRows) and differ by a
fraction of that drift (~2e-4 at 32 PF-Pascal volumes); with a random-sign gradient each weight gradient is a heavily cancelling sum over
up to 10^7 positions (|sum| ~ sqrt(n) of the summed magnitudes) and the fp32 order shows at ~1e-4 .. 1e-3.  usage: python scripts/wgrad_check.py [quick]"""
import os, sys, json
import torch
sys.path.insert(0, ".")
from ncnet_b200 import functional as F_
from ncnet_b200._cabi import CONV_TC16, W_KOIKKK

dev = torch.device("cuda")


def timed(fn, iters=3):
    fn()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


ok = True
torch.manual_seed(0)
cases = [(2, 16, 16, 5, (25, 25, 25, 25)), (1, 16, 16, 5, (6, 7, 9, 11)), (3, 12, 16, 5, (5, 6, 13, 7)), (2, 16, 9, 3, (8, 5, 12, 10)),
         (1, 16, 16, 5, (7, 31, 8, 9)), (2, 16, 16, 3, (4, 4, 20, 17))]
if "quick" not in sys.argv:
    cases.append((32, 16, 16, 5, (25, 25, 25, 25)))
for bsz, cin, cout, k, dims in cases:
    x = torch.rand(bsz, 2 * cin, *dims, device=dev)
    w = torch.empty(k, cout, cin, k, k, k, device=dev)
    for sign, tol in (("positive", 1e-3), ("random-sign", 5e-3)):
        gy = torch.rand(bsz, 2 * cout, *dims, device=dev) if sign == "positive" else torch.randn(bsz, 2 * cout, *dims, device=dev)
        fn = lambda: F_._conv_wgrad(x, gy, w, True, W_KOIKKK, True, False, CONV_TC16)
        os.environ.pop("NCNET_NO_PMW", None)
        gw_n, gb_n = fn(); tn = timed(fn)
        os.environ["NCNET_NO_PMW"] = "1"
        gw_o, gb_o = fn(); to = timed(fn)
        os.environ.pop("NCNET_NO_PMW", None)
        e = ((gw_n - gw_o).abs().max() / gw_o.abs().max()).item()
        eb = ((gb_n - gb_o).abs().max() / gb_o.abs().max()).item()
        good = e < tol and eb < 2e-5
        ok &= good
        print(json.dumps({"case": f"wgrad B={bsz} {cin}->{cout} k={k} {dims} {sign} g", "rel_diff_w": e, "rel_diff_b": eb, "ms_walk": round(tn, 4),
                          "ms_per_plane": round(to, 4), "ok": good}), flush=True)
        del gy
    del x
print("WGRAD CHECK", "PASS" if ok else "FAIL")
sys.exit(0 if ok else 1)
