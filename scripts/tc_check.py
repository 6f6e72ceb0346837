"""Development check: tensor-core Conv4d (NCNET_CONV_TC16) against the fp32 CUDA-core kernel."""
import sys, time
import torch
sys.path.insert(0, ".")
from ncnet_b200 import functional as F_
from ncnet_b200._cabi import CONV_FP32, CONV_TC16, W_KOIKKK

def nerr(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30)).item()

def run(cin, cout, k, dims, bsz, sym, shared, relu=True):
    torch.manual_seed(0)
    g = 2 if sym else 1
    gin = 1 if (shared or not sym) else 2
    x = torch.rand(bsz, gin * cin, *dims, device="cuda")
    bound = 1.0 / (cin * k ** 4) ** 0.5
    w = (torch.rand(k, cout, cin, k, k, k, device="cuda") * 2 - 1) * bound
    b = (torch.rand(cout, device="cuda") * 2 - 1) * bound
    ref = F_._conv_fwd(x, w, b, W_KOIKKK, sym, shared, relu, CONV_FP32)
    out = F_._conv_fwd(x, w, b, W_KOIKKK, sym, shared, relu, CONV_TC16)
    torch.cuda.synchronize()
    e = nerr(out, ref)
    # dgrad
    gy = torch.randn(bsz, g * cout, *dims, device="cuda")
    mask = torch.randn(bsz, gin * cin, *dims, device="cuda") if not shared else None
    rgx = F_._conv_dgrad(gy, w, mask, W_KOIKKK, sym, shared, CONV_FP32)
    ogx = F_._conv_dgrad(gy, w, mask, W_KOIKKK, sym, shared, CONV_TC16)
    torch.cuda.synchronize()
    e2 = nerr(ogx, rgx)
    # wgrad
    gyp = gy * (torch.rand_like(gy) > 0.3)
    rgw, rgb = F_._conv_wgrad(x, gyp, w, True, W_KOIKKK, sym, shared, CONV_FP32)
    ogw, ogb = F_._conv_wgrad(x, gyp, w, True, W_KOIKKK, sym, shared, CONV_TC16)
    torch.cuda.synchronize()
    e3 = max(nerr(ogw, rgw), nerr(ogb, rgb))
    print(f"cin={cin} cout={cout} k={k} dims={dims} B={bsz} sym={sym} shared={shared}: fwd nerr {e:.2e}  dgrad nerr {e2:.2e}  wgrad nerr {e3:.2e}", flush=True)
    return e, max(e2, e3)

cases = [
    (16, 16, 5, (6, 7, 9, 10), 1, False, True),
    (16, 16, 5, (7, 6, 25, 25), 2, True, False),
    (16, 1, 5, (6, 6, 11, 9), 1, True, False),
    (1, 16, 5, (6, 5, 12, 13), 2, True, True),
    (16, 16, 3, (5, 6, 20, 37), 1, True, False),
    (1, 16, 3, (6, 5, 12, 13), 1, True, True),
    (16, 1, 3, (5, 5, 50, 37), 1, True, False),
    (10, 10, 3, (5, 5, 8, 9), 1, True, False),
]
worst = 0.0
for c in cases:
    e, e2 = run(*c)
    worst = max(worst, e, e2)
print("worst", worst)
# timing at the PF shape
x = torch.rand(16, 32, 25, 25, 25, 25, device="cuda")
w = (torch.rand(5, 16, 16, 5, 5, 5, device="cuda") * 2 - 1) * 0.01
b = torch.zeros(16, device="cuda")
for mode in (CONV_TC16, CONV_FP32):
    for _ in range(2):
        F_._conv_fwd(x, w, b, W_KOIKKK, True, False, True, mode)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 5
    for _ in range(n):
        F_._conv_fwd(x, w, b, W_KOIKKK, True, False, True, mode)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / n
    print(f"mode {mode}: fwd 16->16 k=5 B=16 sym: {dt*1e3:.2f} ms  {2*16*25**4*16*16*625*2/dt/1e12:.1f} TFLOP/s dense", flush=True)
    gy = torch.randn(16, 32, 25, 25, 25, 25, device="cuda")
    for _ in range(2):
        F_._conv_wgrad(x, gy, w, True, W_KOIKKK, True, False, mode)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        F_._conv_wgrad(x, gy, w, True, W_KOIKKK, True, False, mode)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 3
    print(f"mode {mode}: wgrad 16->16 k=5 B=16 sym: {dt*1e3:.2f} ms  {2*16*25**4*16*16*625*2/dt/1e12:.1f} TFLOP/s dense", flush=True)
    del gy
