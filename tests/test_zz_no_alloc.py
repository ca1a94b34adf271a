"""ramp_waypoint_variance_static (include/ramp_b200.h): the no-allocation variant of the waypoint variance — its scratch is a
file-scope `static __device__` array of 1 MiB inside the library instead of caller memory
(no counterpart in the reference).  Called through the C ABI, compared with the existing entry point
(ramp_waypoint_variance, caller scratch) on two calls in a row; sizes beyond the static array must be rejected with a clear
error.  This file is collected last on purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(L, x):
    from ramp_b200._lib import check, ptr, stream_ptr
    n, H, D = x.shape
    L.ramp_waypoint_variance_workspace_bytes.restype = ctypes.c_size_t
    ws = torch.empty(L.ramp_waypoint_variance_workspace_bytes(n, H) // 8, dtype=torch.float64, device=x.device)
    ref = torch.zeros(1, device=x.device)
    check(L.ramp_waypoint_variance(ptr(x), n, H, D, ptr(ws), ptr(ref), stream_ptr()), "ramp_waypoint_variance")
    out = torch.full((1,), float("nan"), device=x.device)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(x.device)
    check(L.ramp_waypoint_variance_static(ptr(x), n, H, D, ptr(out), stream_ptr()), "ramp_waypoint_variance_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(x.device) == before          # nothing went through torch's allocator either
    return float(ref[0]), float(out[0])


def test_static_scratch_variant_matches_caller_scratch_variant(cuda):
    from ramp_b200._lib import RampError, check, lib, ptr, stream_ptr
    L = lib()
    L.ramp_waypoint_variance_static_max_bytes.restype = ctypes.c_size_t
    assert L.ramp_waypoint_variance_static_max_bytes() == 1 << 20
    g = torch.Generator("cpu").manual_seed(17)
    # two calls in a row on DIFFERENT sizes: the second must not see anything of the first in the shared array
    for n, H, D in [(1024, 64, 4), (300, 48, 4), (1024, 64, 4), (2, 8, 6), (4096, 64, 4)]:
        x = (torch.rand(n, H, D, generator=g) * 2 - 1).to(cuda).contiguous()
        ref, got = _both(L, x)
        assert ref == ref and got == got
        # same kernels, same summation order: the two entry points agree to the bit (tests/test_metrics_gpu.py holds the
        # kernels themselves to 1e-7 relative against the fp64 oracle)
        assert abs(got - ref) <= 1e-7 * max(1.0, abs(ref)), (n, H, got, ref)
        assert got == ref
    # beyond the static array: rejected with a clear error, output untouched
    n, H = 8192, 64
    x = torch.zeros(n, H, 4, device=cuda)
    out = torch.full((1,), 7.0, device=cuda)
    rc = L.ramp_waypoint_variance_static(ptr(x), n, H, 4, ptr(out), stream_ptr())
    assert rc != 0
    msg = L.ramp_last_error().decode()
    assert "static array" in msg and "8192" in msg, msg
    with pytest.raises(RampError):
        check(rc, "ramp_waypoint_variance_static")
    torch.cuda.synchronize()
    assert float(out[0]) == 7.0
