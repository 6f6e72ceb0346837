"""BASELINE.json configs 1 and 3 as inference timings of the head (features already extracted, resident in HBM):

  PF   : ImMatchNet head at the PF-Pascal shape (25^4 volume, NC 5-5-5 / 16-16-1), batch 1 and 16, fp32 features
  InLoc: eval_inloc.py's head -- fp16 features, correlation at full resolution, maxpool4d k=2 relocalisation,
         NC 3-3 / 16-1, corr_to_matches in both directions with the relocalisation offsets (lib/point_tnf.py:12-80)
         at the 1600-px (100x75 features) and 3200-px (200x150 features) image sizes of the reference's InLoc runs.

  full : BASELINE.json configs[0] -- ImMatchNet.forward on a 400x400 image pair through the real constructor
         (ResNet-101 trunk on stock torchvision/cuDNN, random init: NCNET_TRUNK_WEIGHTS=random), batch 1.

Prints one JSON line per case: ms per pair (CUDA events, median of 5 after 2 warm-ups) and the split by stage.
Synthetic features, random-init weights.  usage: python scripts/inloc_bench.py [pf] [il16] [il32] [full]"""
import json, sys
import torch
sys.path.insert(0, ".")
import ncnet_b200 as nb
from ncnet_b200 import functional as F_

dev = torch.device("cuda")


def head(kernels, channels, k_size, half):
    net = nb.ImMatchNet.__new__(nb.ImMatchNet)          # head only: no trunk
    torch.nn.Module.__init__(net)
    net.relocalization_k_size, net.half_precision = k_size, half
    net.FeatureCorrelation = nb.FeatureCorrelation("4D", False)
    net.NeighConsensus = nb.NeighConsensus(True, kernels, channels).to(dev)
    if half:
        net.NeighConsensus.half()
        for l in net.NeighConsensus.conv:
            if isinstance(l, nb.Conv4d):
                l.use_half = True
    return net


def timed(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def run(name, net, fa, fb, k_size):
    state = {}

    fused = k_size > 1 and F_.can_fuse_correlation_pool(fa, fb, k_size)

    def corr():
        if fused:        # correlation + relocalisation pooling in one kernel (what ImMatchNet.match_features runs)
            state["p"] = F_.feature_correlation_pooled(fa, fb, k_size)
            state["c"] = state["p"][0]
        else:
            state["c"] = net.FeatureCorrelation(fa, fb)

    def pool():
        if not fused:
            state["p"] = nb.maxpool4d(state["c"], k_size=k_size) if k_size > 1 else (state["c"],)

    def mm1():
        state["m"] = nb.MutualMatching(state["p"][0])

    def nc():
        ncm = net.NeighConsensus
        state["o"] = F_.mutual_matching(ncm.forward_branches(state["m"]), sum_branches=True)

    def matches():
        d = tuple(state["p"][1:]) if k_size > 1 else None
        if fa.shape[0] == 1:     # eval_inloc.py:151-203: both directions, duplicates removed, recentred rows [N,5]
            from ncnet_b200.eval_util import inloc_matches
            state["rows"] = inloc_matches(state["o"], delta4d=d, k_size=max(1, k_size), do_softmax=True)
            return
        for inv in (False, True):
            nb.corr_to_matches(state["o"], delta4d=d, k_size=max(1, k_size), do_softmax=True, invert_matching_direction=inv)

    def whole():
        corr(); pool(); mm1(); nc(); matches()

    with torch.no_grad():
        total = timed(whole)
        split = {}
        for lab, fn in (("correlation+maxpool4d (fused)" if fused else "correlation", corr), ("maxpool4d", pool), ("mutual_matching", mm1),
                        ("neigh_consensus+mutual_matching", nc), ("matches both directions + dedup", matches)):
            split[lab] = round(timed(fn), 4)
    b = fa.shape[0]
    print(json.dumps({"case": name, "batch": b, "ms_per_call": round(total, 4), "pairs_per_s": round(b / total * 1e3, 2),
                      "stage_ms": split, "volume": list(state["c"].shape), "dtype": str(fa.dtype)}), flush=True)


if __name__ == "__main__":
    want = set(sys.argv[1:]) or {"pf", "il16", "il32"}
    torch.manual_seed(0)
    if "pf" in want:
        net = head([5, 5, 5], [16, 16, 1], 0, False)
        for b in (1, 16):
            fa = torch.rand(b, 1024, 25, 25, device=dev); fb = torch.rand(b, 1024, 25, 25, device=dev)
            run(f"PF-Pascal head forward + matches, batch {b}", net, fa, fb, 0)
    if "full" in want:
        import os
        os.environ["NCNET_TRUNK_WEIGHTS"] = "random"
        model = nb.ImMatchNet(ncons_kernel_sizes=[5, 5, 5], ncons_channels=[16, 16, 1], use_cuda=True)
        batch = {"source_image": torch.randn(1, 3, 400, 400, device=dev), "target_image": torch.randn(1, 3, 400, 400, device=dev)}
        with torch.no_grad():
            t_full = timed(lambda: model(batch))
            fa = model.FeatureExtraction(batch["source_image"]); fb = model.FeatureExtraction(batch["target_image"])
            t_head = timed(lambda: model.match_features(fa, fb))
            t_trunk = timed(lambda: (model.FeatureExtraction(batch["source_image"]), model.FeatureExtraction(batch["target_image"])))
        print(json.dumps({"case": "config 1: ImMatchNet.forward, 400x400 pair, batch 1 (ResNet-101[:layer3] trunk on torchvision/cuDNN + head)",
                          "ms_full_forward": round(t_full, 4), "ms_head_only": round(t_head, 4), "ms_trunk_x2": round(t_trunk, 4),
                          "pairs_per_s_full": round(1e3 / t_full, 2), "pairs_per_s_head": round(1e3 / t_head, 2)}), flush=True)
    for tag, (h, w) in (("il16", (100, 74)), ("il32", (200, 150))):
        if tag not in want:
            continue
        net = head([3, 3], [16, 1], 2, True)
        fa = torch.rand(1, 1024, h, w, device=dev).half(); fb = torch.rand(1, 1024, h, w, device=dev).half()
        run(f"InLoc head ({tag}: {h}x{w} features, k_size 2, NC 3-3/16-1, fp16)", net, fa, fb, 2)
