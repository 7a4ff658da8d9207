"""GPU: vlgb_index_sa_extract_static -- the allocation-free variant of vlgb_index_sa_extract (scratch in a file-scope
device array, include/vlgb200.h) -- against the existing entry point, through the C ABI, on two calls in a row (the
second one finds the first one's values in the shared scratch).  Bit-exact, like every other test here.
Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_sa_extract_static_equals_allocating_entry_point():
    import vlg_matching_b200 as V
    from vlg_matching_b200 import capi, synth
    capi.init(0)
    lib = capi.lib
    text = synth.make_text(2, 300_000)
    idx = V.Index.build(text)
    idx.attach_rank()
    n = text.size
    assert n > capi.SA_EXTRACT_STATIC_MAX

    def static(first, count):
        out = np.full(max(count, 1), 0xDEADBEEFDEADBEEF, dtype=np.uint64)
        capi.check(lib.vlgb_index_sa_extract_static(idx.h, first, count, out.ctypes.data))
        return out[:count]

    # two calls in a row: a long range, then a SHORTER one somewhere else (stale values of the first call lie behind it)
    calls = [(1000, capi.SA_EXTRACT_STATIC_MAX), (n - 70_001, 70_001), (0, 1), (n // 2, 4097), (n, 0)]
    for first, count in calls:
        want = idx.sa_extract(first, count)
        got = static(first, count)
        assert got.shape == want.shape and np.array_equal(got, want), (first, count)
        assert np.array_equal(static(first, count), want), (first, count)     # and again: same answer
    # the whole suffix array in pieces of the maximum size is a permutation of the text positions
    pieces = [static(f, min(capi.SA_EXTRACT_STATIC_MAX, n - f)) for f in range(0, n, capi.SA_EXTRACT_STATIC_MAX)]
    assert np.array_equal(np.sort(np.concatenate(pieces)), np.arange(n, dtype=np.uint64))

    # a device-memory destination works too (the copy direction is taken from the pointer)
    import torch
    dev = torch.zeros(5000, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    capi.check(lib.vlgb_index_sa_extract_static(idx.h, 77, 5000, C.c_void_p(dev.data_ptr())))
    assert np.array_equal(dev.cpu().numpy().astype(np.uint64), idx.sa_extract(77, 5000))

    # no device memory is taken or returned by the call
    free0 = torch.cuda.mem_get_info(0)[0]
    for first, count in calls:
        static(first, count)
    assert torch.cuda.mem_get_info(0)[0] == free0

    # beyond the static scratch: a clear argument error, nothing written
    out = np.zeros(4, dtype=np.uint64)
    rc = lib.vlgb_index_sa_extract_static(idx.h, 0, capi.SA_EXTRACT_STATIC_MAX + 1, out.ctypes.data)
    assert rc == capi.ERR_ARG and b"VLGB_SA_EXTRACT_STATIC_MAX" in lib.vlgb_last_error()
    assert lib.vlgb_index_sa_extract_static(idx.h, n - 5, 6, out.ctypes.data) == capi.ERR_ARG   # range outside the SA
    plain = V.Index.build(text[:4096])
    assert lib.vlgb_index_sa_extract_static(plain.h, 0, 4, out.ctypes.data) == capi.ERR_ARG     # no rank structures
    plain.free()
    idx.free()
