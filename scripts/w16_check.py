"""A/B check of the many-channel Conv4d layers: the persistent CTA-pair walk with resident weights
(csrc/conv4d_tc_walk16.cu; taken for k = 5 layers writing CB16, i.e. inside a chained stack) against the per-plane kernel
(NCNET_NO_W16=1), same inputs.  Both run the same fp16 x fp16
products with fp32 accumulation; only the summation order differs, so results must agree to ~1e-6 relative.
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
    os.environ.pop("NCNET_NO_W16", None)
    new = fn(); t_new = timed(fn)
    os.environ["NCNET_NO_W16"] = "1"
    old = fn(); t_old = timed(fn)
    os.environ.pop("NCNET_NO_W16", None)
    return new, old, t_new, t_old


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


ok = True
quick = "quick" in sys.argv
torch.manual_seed(0)
# the chained stack (CB16 / CB16PM hidden activations, CB16 ReLU masks in dgrad), forward + backward
for bsz, ks, chans, dims in ((2, [5, 5, 5], [16, 16, 1], (25, 25, 25, 25)), (2, [3, 3, 3], [16, 16, 1], (9, 8, 10, 12)),
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
    # forward: only the summation order differs (a hidden fp16 activation may round the other way: ~1e-5); gradients: a
    # pre-activation within rounding of zero may switch its ReLU mask, one term of a sum over the volume (<~ 1/cells)
    good = e < 1e-4 and eg < 5e-3
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
print("W16 CHECK", "PASS" if ok else "FAIL")
sys.exit(0 if ok else 1)
