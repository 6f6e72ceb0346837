"""Config 5 (BASELINE.json): micro-benchmarks of the memory-bound kernels and the Conv4d layers against their rooflines,
volume N^4 for N in 16..50, (Cin,Cout) in {1,16}^2 \\ {(1,1)}, k in {3,5}, forward / dgrad / wgrad, with the CPU oracle
(the reference's PyTorch path restated, oracle/head.py) timed beside the rows where it finishes in seconds.
One JSON line per measurement.  Bandwidth kernels: called straight through the C ABI on preallocated buffers, `reps`
back-to-back launches per event pair on volumes larger than the 126 MB L2 (so launch gaps and L2 residency do not
flatter or hurt the number).  usage: python scripts/micro_bench.py [mem] [conv] [cpu]"""
import json, os, sys, time
import torch
sys.path.insert(0, ".")
import ncnet_b200 as nb
from ncnet_b200 import _cabi, functional as F_
from ncnet_b200._cabi import CONV_FP32, CONV_TC16, W_KOIKKK

PEAKS = json.load(open("MEASURED_PEAKS.json")) if os.path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
dev = torch.device("cuda")
want = set(sys.argv[1:]) or {"mem", "conv", "cpu"}
L = _cabi.lib()


def timeit(fn, iters=7, reps=1):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(reps):
            fn()
        b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / reps)
    ts.sort()
    return ts[len(ts) // 2]


def emit(**kw):
    print(json.dumps(kw), flush=True)


def cpu_ms(fn, budget=6.0):
    t0 = time.perf_counter(); fn(); dt = time.perf_counter() - t0
    if dt < budget / 3:
        t0 = time.perf_counter(); fn(); dt = min(dt, time.perf_counter() - t0)
    return dt * 1e3


st = lambda: torch.cuda.current_stream().cuda_stream
with torch.no_grad():
    if "mem" in want:
        from oracle import head as O
        cases = [(f"N={n} fp32", (max(1, (300 << 20) // (n ** 4 * 4)), n, n, n, n), torch.float32) for n in (16, 25, 32, 40, 50)]
        cases += [("IL32 pooled fp16 (100x75x100x75)", (3, 100, 75, 100, 75), torch.float16), ("PF B=32 fp32 (L2-resident)", (32, 25, 25, 25, 25), torch.float32)]
        for name, (b, f1, f2, f3, f4), dt in cases:
            x = torch.rand(b, 1, f1, f2, f3, f4, device=dev).to(dt)
            y = torch.empty_like(x)
            na, nbb = f1 * f2, f3 * f4
            keys = torch.empty(b * (na + nbb), device=dev, dtype=torch.int64)
            d = _cabi.F16 if dt == torch.float16 else _cabi.F32
            byts = 2 * x.numel() * x.element_size()
            fn = lambda: _cabi.check(L.ncnet_mutual_matching_fwd(x.data_ptr(), y.data_ptr(), keys.data_ptr(), b, na, nbb, 1, d, st()), "mm")
            ms_train = timeit(fn, reps=4)
            fn = lambda: _cabi.check(L.ncnet_mutual_matching_fwd_infer(x.data_ptr(), y.data_ptr(), keys.data_ptr(), b, na, nbb, 1, d, st()), "mm")
            ms = timeit(fn, reps=4)
            row = dict(kernel="MutualMatching fwd (memset + max + apply), inference entry point", case=f"{name} B={b}", volume_MB=round(byts / 2e6, 1),
                       ms=ms, achieved_GBs=byts / ms / 1e6, frac_of_hbm=byts / ms / 1e6 / PEAKS["hbm_gbs"], bytes="2*numel*sizeof (1 read + 1 write)",
                       ms_with_argmax_keys=ms_train, frac_of_hbm_with_argmax_keys=byts / ms_train / 1e6 / PEAKS["hbm_gbs"])
            if "cpu" in want and x.numel() * b <= 64 << 20:
                xc = x[:1].float().cpu()
                row["cpu_oracle_ms_per_item"] = cpu_ms(lambda: O.mutual_matching(xc)); row["gpu_ms_per_item"] = ms / b
            emit(**row)
            del x, y, keys
        # maxpool4d k=2 (batch 1 like the reference): the InLoc pre-pooling volumes, and 50^4
        for name, shape in (("N=50 fp16", (1, 1, 50, 50, 50, 50)), ("IL16 fp16 (100x74x100x74)", (1, 1, 100, 74, 100, 74)),
                            ("IL32 fp16 (200x150x200x150)", (1, 1, 200, 150, 200, 150))):
            x = torch.rand(*shape, device=dev).half()
            nout = x.numel() // 16
            for idt, ib in ((torch.int64, 8), (torch.uint8, 1)):
                ms = timeit(lambda: F_.maxpool4d(x, 2, index_dtype=idt), iters=5, reps=4 if x.numel() < (1 << 27) else 1)
                byts = x.numel() * 2 + nout * (2 + 4 * ib)
                emit(kernel=f"maxpool4d k=2 ({'int64' if ib == 8 else 'uint8'} offsets)", case=name, ms=ms,
                     achieved_GBs=byts / ms / 1e6, frac_of_hbm=byts / ms / 1e6 / PEAKS["hbm_gbs"])
            del x
        for name, shape, dt in (("PF B=16 fp32", (16, 1, 25, 25, 25, 25), torch.float32), ("N=50 B=8 fp32", (8, 1, 50, 50, 50, 50), torch.float32),
                                ("IL32 pooled fp16", (1, 1, 100, 75, 100, 75), torch.float16)):
            x = torch.rand(*shape, device=dev).to(dt)
            for inv in (False, True):
                for sm in (True, False):
                    ms = timeit(lambda: nb.corr_to_matches(x, do_softmax=sm, invert_matching_direction=inv))
                    gbs = x.numel() * x.element_size() / ms / 1e6
                    emit(kernel="corr_to_matches", case=f"{name} invert={inv} softmax={sm}", ms=ms, achieved_GBs=gbs, frac_of_hbm=gbs / PEAKS["hbm_gbs"],
                         bytes="numel*sizeof (one read; the soft-max variant reads twice)")
            del x
        for name, c, ha, wa, dt in (("IL16 fp16", 1024, 100, 74, torch.float16), ("IL32 fp16", 1024, 200, 150, torch.float16)):
            fa = torch.rand(1, c, ha, wa, device=dev).to(dt); fb = torch.rand(1, c, ha, wa, device=dev).to(dt)
            flop = 2.0 * (ha * wa) ** 2 * c
            ms1 = timeit(lambda: F_.maxpool4d(F_._corr_fwd_raw(fa, fb, False, tensor_cores=True).reshape(1, 1, ha, wa, ha, wa), 2), iters=5)
            ms2 = timeit(lambda: F_.feature_correlation_pooled(fa, fb, 2), iters=5)
            ms3 = timeit(lambda: F_.feature_correlation_pooled(fa, fb, 2, index_dtype=torch.uint8), iters=5)
            emit(kernel="correlation + maxpool4d k=2", case=name, ms_unfused=ms1, ms_fused_int64_offsets=ms2, ms_fused_uint8_offsets=ms3,
                 fused_TFLOPs=flop / ms2 / 1e9, frac_of_bf16_burst_peak=flop / ms2 / 1e9 / PEAKS["bf16_tflops"])
            del fa, fb
    if "conv" in want:
        from oracle import head as O
        for n in (16, 25, 32, 40, 50):
            for cin, cout in ((1, 16), (16, 16), (16, 1)):
                for k in (3, 5):
                    bsz = max(1, (25 ** 4 * 4) // n ** 4)
                    gin = 1 if cin == 1 else 2
                    x = torch.rand(bsz, gin * cin, n, n, n, n, device=dev)
                    w = (torch.rand(k, cout, cin, k, k, k, device=dev) - 0.5) * 0.1
                    bia = torch.zeros(cout, device=dev)
                    gy = torch.randn(bsz, 2 * cout, n, n, n, n, device=dev)
                    flop = 2.0 * bsz * n ** 4 * cin * cout * k ** 4 * 2
                    for mode, mname in ((CONV_TC16, "tc16"), (CONV_FP32, "fp32")):
                        if mode == CONV_FP32 and n > 25:
                            continue
                        row = dict(kernel="conv4d (planar in/out, both branches)", case=f"N={n} {cin}->{cout} k={k} B={bsz}", mode=mname)
                        for what, fn in (("fwd", lambda: F_._conv_fwd(x, w, bia, W_KOIKKK, True, cin == 1, True, mode)),
                                         ("dgrad", lambda: F_._conv_dgrad(gy, w, None, W_KOIKKK, True, cin == 1, mode)),
                                         ("wgrad", lambda: F_._conv_wgrad(x, gy, w, True, W_KOIKKK, True, cin == 1, mode))):
                            ms = timeit(fn, iters=3)
                            row[what + "_ms"] = ms
                            row[what + "_TFLOPs"] = flop / ms / 1e9
                            row[what + "_frac_of_bf16_burst_peak"] = flop / ms / 1e9 / PEAKS["bf16_tflops"]
                        if "cpu" in want and mode == CONV_TC16 and n == 16:
                            layer = O.Conv4d(cin, cout, k)
                            xc = torch.rand(1, cin, n, n, n, n)
                            row["cpu_oracle_fwd_ms_per_item_per_branch"] = cpu_ms(lambda: layer(xc))
                            row["gpu_fwd_ms_per_item_per_branch"] = row["fwd_ms"] / bsz / 2
                        emit(**row)
                    del x, gy
