"""A handful of launches of the bandwidth-bound kernels at sizes beyond L2, for `ncu` (launch list or --set full):
MutualMatching forward at 50^4 x 12 (fp32) and at the InLoc-3200 pooled size (fp16), corr_to_matches in both
directions, maxpool4d, fused correlation + pooling.  usage: ncu ... python scripts/mm_prof.py"""
import sys
import torch
sys.path.insert(0, ".")
import ncnet_b200 as nb
from ncnet_b200 import _cabi, functional as F_

dev = torch.device("cuda")
L = _cabi.lib()
st = lambda: torch.cuda.current_stream().cuda_stream
with torch.no_grad():
    for (b, f1, f2, f3, f4), dt in (((12, 50, 50, 50, 50), torch.float32), ((3, 100, 75, 100, 75), torch.float16)):
        x = torch.rand(b, 1, f1, f2, f3, f4, device=dev).to(dt)
        y = torch.empty_like(x)
        keys = torch.empty(b * (f1 * f2 + f3 * f4), device=dev, dtype=torch.int64)
        d = _cabi.F16 if dt == torch.float16 else _cabi.F32
        for _ in range(2):
            _cabi.check(L.ncnet_mutual_matching_fwd(x.data_ptr(), y.data_ptr(), keys.data_ptr(), b, f1 * f2, f3 * f4, 1, d, st()), "mm")
        for _ in range(2):      # inference entry point: pass 1 without the arg-max indices
            _cabi.check(L.ncnet_mutual_matching_fwd_infer(x.data_ptr(), y.data_ptr(), keys.data_ptr(), b, f1 * f2, f3 * f4, 1, d, st()), "mm")
        torch.cuda.synchronize()
        for inv in (False, True):
            for _ in range(2):
                nb.corr_to_matches(x[:1], do_softmax=True, invert_matching_direction=inv)
        del x, y, keys
    x = torch.rand(1, 1, 200, 150, 200, 150, device=dev).half()
    for _ in range(2):
        F_.maxpool4d(x, 2, index_dtype=torch.uint8)
        nb.maxpool4d(x, 2)
    del x
    fa = torch.rand(1, 1024, 200, 150, device=dev).half(); fb = torch.rand(1, 1024, 200, 150, device=dev).half()
    for _ in range(2):
        F_.feature_correlation_pooled(fa, fb, 2)
    torch.cuda.synchronize()
