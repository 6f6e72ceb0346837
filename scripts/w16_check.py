"""A/B check of the persistent walk kernels inside a chained NeighConsensus stack -- csrc/conv4d_tc_walk16.cu (16->16, k = 5:
CTA pairs with resident weights) and csrc/conv4d_tc_walk1.cu (single input channel: walk along iA) -- against the
per-plane kernel (NCNET_NO_W16=1 NCNET_NO_W1=1), same inputs.  Both run the same fp16 x fp16 products with fp32
accumulation; only the summation order differs.
Prints one line per case: max abs difference / max abs value, and the time of both.
usage: python scripts/w16_check.py [quick]"""
import os, sys, json
import torch
sys.path.insert(0, ".")
import ncnet_b200 as nb
from ncnet_b200 import functional as F_
from ncnet_b200._cabi import CONV_TC16, W_KOIKKK

dev = torch.device("cuda")


def timed(fn, iters=5):
    fn(); fn()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def both(fn):
    for v in ("NCNET_NO_W16", "NCNET_NO_W1"):
        os.environ.pop(v, None)
    new = fn(); t_new = timed(fn)
    for v in ("NCNET_NO_W16", "NCNET_NO_W1"):
        os.environ[v] = "1"
    old = fn(); t_old = timed(fn)
    for v in ("NCNET_NO_W16", "NCNET_NO_W1"):
        os.environ.pop(v, None)
    return new, old, t_new, t_old


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


ok = True
quick = "quick" in sys.argv
torch.manual_seed(0)
# the chained stack (CB16 / CB16PM hidden activations, CB16 ReLU masks in dgrad), forward + backward
for bsz, ks, chans, dims in ((2, [5, 5, 5], [16, 16, 1], (25, 25, 25, 25)), (2, [3, 3, 3], [16, 16, 1], (9, 8, 10, 12)), (2, [5, 5], [7, 1], (3, 9, 30, 41)), (1, [3, 3], [16, 1], (6, 5, 70, 9)),
                             (1, [5, 5, 5, 5], [16, 16, 16, 1], (7, 6, 8, 9))) + (() if quick else ((32, [5, 5, 5], [16, 16, 1], (25, 25, 25, 25)),)):
    nc = nb.NeighConsensus(True, ks, chans).to(dev)
    x = torch.rand(bsz, 1, *dims, device=dev)
    gout = torch.rand(bsz, 1, *dims, device=dev)

    def run():
        for p_ in nc.parameters():
            p_.grad = None
        y = nc(x)
        y.backward(gout)
        return torch.cat([y.detach().flatten()] + [p_.grad.flatten() / p_.grad.abs().max().clamp_min(1e-30) for p_ in nc.parameters()])
    new, old, tn, to = both(run)
    ny = x.numel()
    e, eg = rel(new[:ny], old[:ny]), rel(new[ny:], old[ny:])
    # forward: only the summation order differs, but a hidden fp16 activation may then round the other way (one fp16 ulp,
    # 4.9e-4 of that activation); gradients: a pre-activation within rounding of zero may switch its ReLU mask, one term of
    # a sum over the volume (<~ 1/cells).  tests/test_gpu_tc_pinned.py holds both kernels to the fp16 model.
    good = e < 5e-4 and eg < 5e-3
    ok &= good
    print(json.dumps({"case": f"stack B={bsz} k={ks} ch={chans} {dims} fwd+bwd", "rel_diff_fwd": e, "rel_diff_grads": eg, "ms_walk16": round(tn, 4),
                      "ms_per_plane": round(to, 4), "ok": good}), flush=True)
# inference (no gradients): the PF-Pascal stack at batch 1 and 16
for bsz in (1, 16):
    nc = nb.NeighConsensus(True, [5, 5, 5], [16, 16, 1]).to(dev)
    x = torch.rand(bsz, 1, 25, 25, 25, 25, device=dev)
    with torch.no_grad():
        new, old, tn, to = both(lambda: nc(x))
    e = rel(new, old)
    good = e < 1e-4
    ok &= good
    print(json.dumps({"case": f"stack B={bsz} PF-Pascal forward only", "rel_diff": e, "ms_walk16": round(tn, 4), "ms_per_plane": round(to, 4), "ok": good}),
          flush=True)
# the InLoc head's stack (fp16 storage, k = 3, 16 -> 1), inference
for dims in ((50, 37, 50, 37),) + (() if quick else ((100, 75, 100, 75),)):
    nc = nb.NeighConsensus(True, [3, 3], [16, 1]).to(dev).half()
    for l in nc.conv:
        if isinstance(l, nb.Conv4d):
            l.use_half = True
    x = torch.rand(1, 1, *dims, device=dev).half()
    with torch.no_grad():
        new, old, tn, to = both(lambda: nc(x))
    e = rel(new, old)
    good = e < 2e-3          # fp16 output: one ulp
    ok &= good
    print(json.dumps({"case": f"InLoc stack fp16 {dims} forward only", "rel_diff": e, "ms_walk": round(tn, 4), "ms_per_plane": round(to, 4), "ok": good}),
          flush=True)
    del x, new, old
print("W16 CHECK", "PASS" if ok else "FAIL")
sys.exit(0 if ok else 1)
