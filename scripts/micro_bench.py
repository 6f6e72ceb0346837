"""Config 5 (BASELINE.json): micro-benchmarks of the memory-bound kernels and the Conv4d layers against their rooflines.
Prints one JSON line per measurement (CUDA events, L2 flushed between iterations for the bandwidth kernels)."""
import json, os, sys
import torch
sys.path.insert(0, ".")
import ncnet_b200 as nb
from ncnet_b200 import functional as F_
from ncnet_b200._cabi import CONV_FP32, CONV_TC16, W_KOIKKK

PEAKS = json.load(open("MEASURED_PEAKS.json")) if os.path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, iters=10, do_flush=True, reps=1):
    """median over `iters` of (one event pair around `reps` back-to-back calls) / reps.  reps > 1 hides the host's
    launch overhead for kernels of a few microseconds (use it only with inputs larger than L2 or do_flush=False)."""
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        if do_flush:
            flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / reps)
    ts.sort()
    return ts[len(ts) // 2]


def emit(**kw):
    print(json.dumps(kw), flush=True)


with torch.no_grad():
    # MutualMatching forward: algorithmic bytes = 1 read + 1 write
    for name, shape, dt in (("PF B=16", (16, 1, 25, 25, 25, 25), torch.float32), ("PF B=64", (64, 1, 25, 25, 25, 25), torch.float32),
                            ("50^4 B=8", (8, 1, 50, 50, 50, 50), torch.float32), ("IL32 pooled fp16", (1, 1, 100, 75, 100, 75), torch.float16),
                            ("IL32 pooled fp32", (1, 1, 100, 75, 100, 75), torch.float32),
                            ("PF B=1", (1, 1, 25, 25, 25, 25), torch.float32)):
        x = torch.rand(*shape, device=dev).to(dt)
        big = 2 * x.numel() * x.element_size() > (200 << 20)
        ms = timeit(lambda: nb.MutualMatching(x), reps=8 if big else 1)
        gbs = 2 * x.numel() * x.element_size() / ms / 1e6
        emit(kernel="MutualMatching fwd", case=name + (" (8 back-to-back calls, > L2)" if big else " (single call incl. host launch gaps)"), ms=ms, achieved_GBs=gbs, frac_of_hbm=gbs / PEAKS["hbm_gbs"], bytes="2*numel*sizeof")
    # maxpool4d k=2 (InLoc): bytes = input + pooled value + 4 offset bytes
    for name, shape in (("IL16 fp16", (1, 1, 100, 74, 100, 74)), ("IL32 fp16", (1, 1, 200, 150, 200, 150))):
        x = torch.rand(*shape, device=dev).half()
        ms = timeit(lambda: F_.maxpool4d(x, 2, index_dtype=torch.uint8), iters=5)
        nout = x.numel() // 16
        gbs = (x.numel() * 2 + nout * (2 + 4)) / ms / 1e6
        emit(kernel="maxpool4d k=2 (uint8 offsets)", case=name, ms=ms, achieved_GBs=gbs, frac_of_hbm=gbs / PEAKS["hbm_gbs"])
        ms = timeit(lambda: nb.maxpool4d(x, 2), iters=5)
        gbs = (x.numel() * 2 + nout * (2 + 32)) / ms / 1e6
        emit(kernel="maxpool4d k=2 (int64 offsets, reference dtype)", case=name, ms=ms, achieved_GBs=gbs, frac_of_hbm=gbs / PEAKS["hbm_gbs"])
        del x
    # corr_to_matches: one read of the volume
    for name, shape, dt in (("PF B=16", (16, 1, 25, 25, 25, 25), torch.float32), ("IL32 pooled fp16", (1, 1, 100, 75, 100, 75), torch.float16)):
        x = torch.rand(*shape, device=dev).to(dt)
        for inv in (False, True):
            ms = timeit(lambda: nb.corr_to_matches(x, do_softmax=True, invert_matching_direction=inv))
            gbs = x.numel() * x.element_size() / ms / 1e6
            emit(kernel="corr_to_matches softmax", case=f"{name} invert={inv}", ms=ms, achieved_GBs=gbs, frac_of_hbm=gbs / PEAKS["hbm_gbs"])
    # FeatureCorrelation: tcgen05 GEMM vs the fp32 CUDA-core GEMM
    for name, b, c, ha, wa, dt in (("PF B=16", 16, 1024, 25, 25, torch.float32), ("IL16 fp16", 1, 1024, 100, 74, torch.float16),
                                   ("IL32 fp16", 1, 1024, 200, 150, torch.float16)):
        fa = torch.rand(b, c, ha, wa, device=dev).to(dt); fb = torch.rand(b, c, ha, wa, device=dev).to(dt)
        for tcm, mname in ((True, "tc16"), (False, "fp32")):
            if not tcm and ha * wa > 10000:
                continue
            ms = timeit(lambda: F_._corr_fwd_raw(fa, fb, False, tensor_cores=tcm), iters=3, do_flush=False)
            tf = 2.0 * b * (ha * wa) ** 2 * c / ms / 1e9
            emit(kernel="FeatureCorrelation fwd (incl. operand packing)", case=name, mode=mname, ms=ms, achieved_TFLOPs=tf,
                 frac_of_bf16_peak=tf / PEAKS["bf16_tflops"])
        del fa, fb
    # Conv4d layers, forward, both branches, per mode
    for n in (16, 25, 32):
        for cin, cout in ((1, 16), (16, 16), (16, 1)):
            for k in (3, 5):
                bsz = max(1, (25 ** 4 * 4) // n ** 4)
                gin = 1 if cin == 1 else 2
                x = torch.rand(bsz, gin * cin, n, n, n, n, device=dev)
                w = (torch.rand(k, cout, cin, k, k, k, device=dev) - 0.5) * 0.1
                b = torch.zeros(cout, device=dev)
                for mode, mname in ((CONV_TC16, "tc16"), (CONV_FP32, "fp32")):
                    ms = timeit(lambda: F_._conv_fwd(x, w, b, W_KOIKKK, True, cin == 1, True, mode), iters=3, do_flush=False)
                    tf = 2.0 * bsz * n ** 4 * cin * cout * k ** 4 * 2 / ms / 1e9
                    emit(kernel="conv4d fwd (planar in/out, both branches)", case=f"N={n} {cin}->{cout} k={k} B={bsz}", mode=mname, ms=ms,
                         achieved_TFLOPs=tf, frac_of_bf16_peak=tf / PEAKS["bf16_tflops"])
