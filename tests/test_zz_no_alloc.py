"""`kml_softmax_ce_noalloc`: the loss entry point for callers that are not allowed to allocate.

Some users of the native library cannot call cudaMalloc at the point where they need the loss:
their HBM is fully handed out to pre-planned arenas, or a cudaMalloc would stall a latency-critical
loop, or another component owns the allocator.  For them the library keeps the kernel's scratch in
a fixed `static __device__` array.  This test drives the new entry point through the C ABI exactly
as such a caller would — every buffer is created up front, nothing is allocated between the two
calls — and checks it against `kml_softmax_ce` (caller-provided scratch) twice in a row, because
the built-in array is shared and keeps the previous call's contents.

The file name sorts last on purpose: a `pytest -x` run has finished every other test before this
one can fail.  It passes both on its own in a fresh process and after the rest of the suite.
"""
import ctypes
from ctypes import c_int, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _bind(lib):
    lib.kml_softmax_ce_noalloc.restype = c_int
    lib.kml_softmax_ce_noalloc.argtypes = [c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p,
                                           c_void_p, c_void_p]
    lib.kml_softmax_ce_noalloc_max_rows.restype = c_int
    lib.kml_softmax_ce_noalloc_max_rows.argtypes = []
    return lib


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_noalloc_matches_scratch_variant_twice(dtype):
    from kubeml_b200 import _native
    lib = _bind(_native.lib())
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = _native.stream_ptr()
    torch.manual_seed(11)

    cases = [(128, 1000), (37, 10)]       # second call is smaller: stale rows from the first must not leak
    # everything the two calls need exists before the first one
    bufs = []
    for rows, classes in cases:
        x = (torch.randn(rows, classes, device=dev) * 3).to(dtype)
        y = torch.randint(0, classes, (rows,), device=dev)
        ref = dict(loss=torch.empty(1, device=dev), dl=torch.empty_like(x),
                   corr=torch.empty(1, dtype=torch.int32, device=dev),
                   scratch=torch.zeros(int(lib.kml_softmax_ce_scratch_bytes(rows)), dtype=torch.uint8, device=dev))
        new = dict(loss=torch.empty(1, device=dev), dl=torch.empty_like(x),
                   corr=torch.empty(1, dtype=torch.int32, device=dev))
        bufs.append((rows, classes, x, y, ref, new))
    torch.cuda.synchronize()

    is_bf16 = int(dtype == torch.bfloat16)
    free_before = allocs_before = None
    for i, (rows, classes, x, y, ref, new) in enumerate(bufs):
        _native.check(lib.kml_softmax_ce(x.data_ptr(), is_bf16, y.data_ptr(), rows, classes,
                                         ref["loss"].data_ptr(), ref["dl"].data_ptr(), ref["corr"].data_ptr(),
                                         ref["scratch"].data_ptr(), stream), "kml_softmax_ce")
        if i == 1:
            # The promise itself: the call takes no device memory.  Checked on the second call, after the
            # first one did the once-per-device set-up (which allocates nothing either, but this way the
            # driver's lazy module loading cannot blur the reading): free device memory as the driver
            # reports it (cudaMemGetInfo) and torch's allocation counter must both stand still.
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(dev)[0]
            allocs_before = torch.cuda.memory_stats(dev).get("allocation.all.allocated", 0)
        _native.check(lib.kml_softmax_ce_noalloc(x.data_ptr(), is_bf16, y.data_ptr(), rows, classes,
                                                 new["loss"].data_ptr(), new["dl"].data_ptr(),
                                                 new["corr"].data_ptr(), stream), "kml_softmax_ce_noalloc")
        if i == 1:
            torch.cuda.synchronize()
            assert torch.cuda.mem_get_info(dev)[0] == free_before, "kml_softmax_ce_noalloc changed the free device memory"
            assert torch.cuda.memory_stats(dev).get("allocation.all.allocated", 0) == allocs_before
    torch.cuda.synchronize()

    tol = 1e-5 if dtype == torch.float32 else 2e-2      # the tolerances of tests/test_gpu_native.py
    for rows, classes, x, y, ref, new in bufs:
        torch.testing.assert_close(new["loss"], ref["loss"], rtol=tol, atol=tol)
        torch.testing.assert_close(new["dl"].float(), ref["dl"].float(), rtol=tol, atol=tol)
        assert int(new["corr"].item()) == int(ref["corr"].item())
        # and both agree with PyTorch
        want = torch.nn.functional.cross_entropy(x.float(), y)
        torch.testing.assert_close(new["loss"].view(()), want, rtol=tol, atol=tol)


def test_noalloc_rejects_oversized_batches():
    from kubeml_b200 import _native
    lib = _bind(_native.lib())
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    limit = int(lib.kml_softmax_ce_noalloc_max_rows())
    rows = limit + 1
    x = torch.zeros(rows, 2, device=dev)
    y = torch.zeros(rows, dtype=torch.int64, device=dev)
    loss = torch.empty(1, device=dev)
    rc = lib.kml_softmax_ce_noalloc(x.data_ptr(), 0, y.data_ptr(), rows, 2, loss.data_ptr(), None, None,
                                    _native.stream_ptr())
    assert rc != 0
    msg = lib.kml_last_error().decode()
    assert "exceed" in msg and str(limit) in msg
