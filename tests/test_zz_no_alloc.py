"""GPU: the allocation-free MutualMatching entry point, called straight through the C ABI.

Collected last on purpose (file name): a stop-at-first-failure run has already gone through every
other test when it gets here.  The entry point keeps its (max, arg-max) table in a device array that
is built into the library, for hosts that may not call cudaMalloc where the head runs (plugin enqueue
hooks, real-time loops).  Two calls in a row must each equal the ordinary entry point -- the second,
smaller problem runs over whatever the first left in the shared table.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(lib, x, groups):
    """(no-alloc result, ordinary result) for a [B,groups,f1,f2,f3,f4] volume."""
    from ncnet_b200 import _cabi
    b, g, f1, f2, f3, f4 = x.shape
    na, nb = f1 * f2, f3 * f4
    dt = _cabi.F32 if x.dtype == torch.float32 else _cabi.F16
    st = torch.cuda.current_stream().cuda_stream
    y_static = torch.full((b, 1, f1, f2, f3, f4), float("nan"), device=x.device, dtype=x.dtype)
    y_plain = torch.full_like(y_static, float("nan"))
    keys = torch.empty(b * (na + nb), device=x.device, dtype=torch.int64)
    _cabi.check(lib.ncnet_mutual_matching_fwd_noalloc(x.data_ptr(), y_static.data_ptr(), b, na, nb, groups, dt, st),
                "noalloc")
    _cabi.check(lib.ncnet_mutual_matching_fwd(x.data_ptr(), y_plain.data_ptr(), keys.data_ptr(), b, na, nb, groups, dt, st),
                "plain")
    torch.cuda.synchronize()
    return y_static, y_plain


def test_noalloc_matches_plain_entry_point_twice(cuda):
    from ncnet_b200 import _cabi
    lib = _cabi.lib()
    torch.manual_seed(21)
    # first call: PF-Pascal sized batch with large values, leaves big maxima in the shared table
    x1 = (torch.rand(4, 1, 25, 25, 25, 25, device=cuda) * 50.0).contiguous()
    a1, b1 = _both(lib, x1, 1)
    assert torch.isfinite(a1).all() and torch.equal(a1, b1)
    # second call, straight after: smaller problem, smaller values, two summed branches, fp16 storage
    x2 = (torch.rand(2, 2, 7, 5, 6, 9, device=cuda) * 0.1).half().contiguous()
    a2, b2 = _both(lib, x2, 2)
    assert torch.isfinite(a2.float()).all() and torch.equal(a2, b2)
    # the property itself: a warm call takes no device memory (neither from the driver nor from torch's pool)
    y = torch.empty_like(x2[:, :1])
    st = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    held0, reserved0 = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    _cabi.check(lib.ncnet_mutual_matching_fwd_noalloc(x2.data_ptr(), y.data_ptr(), 2, 35, 54, 2, _cabi.F16, st), "noalloc")
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0 and torch.cuda.memory_allocated() == held0 and torch.cuda.memory_reserved() == reserved0
    assert torch.equal(y, b2)
    # and against the CPU oracle, within the fp32 tolerance of the parity tests
    from oracle import head as O
    ref = O.mutual_matching(x1[:1].cpu())
    err = ((a1[:1].cpu() - ref).abs().max() / ref.abs().max()).item()
    assert err < 1e-6


def test_noalloc_rejects_oversize_and_bad_arguments(cuda):
    from ncnet_b200 import _cabi
    lib = _cabi.lib()
    one = ctypes.c_void_p(256)
    # 1 * (300000 + 10) keys > 262144: refused before anything is launched
    assert lib.ncnet_mutual_matching_fwd_noalloc(one, one, 1, 300000, 10, 1, 0, None) != 0
    assert b"exceeds the built-in table" in lib.ncnet_last_error()
    assert lib.ncnet_mutual_matching_fwd_noalloc(None, one, 1, 4, 4, 1, 0, None) != 0
    assert lib.ncnet_mutual_matching_fwd_noalloc(one, one, 1, 4, 4, 3, 0, None) != 0
