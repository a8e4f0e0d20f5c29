"""bpc_sort_u64_noalloc: the key sort for hosts that may not allocate device memory.

Some users of this library receive their device buffers from somewhere else (an engine's fixed
arena, a preallocated pool) and are forbidden to cause a cudaMalloc / cudaFree while running, since
an allocation can stall every stream of the device.  For them the library keeps the sort's scratch
in a fixed array of the loaded module instead of growing buffers on demand.  This test drives the
entry point through the C ABI with caller-owned device buffers and checks it against the regular
bpc_sort_u64 on two calls in a row (the scratch is shared and keeps the previous call's contents,
so the second, smaller call must not be disturbed by what the first one left behind), and checks
that a request beyond the fixed limit is refused with a readable error.

Collected last on purpose (file name): a run that stops at the first failure has already run every
other test.
"""
import ctypes

import numpy as np
import pytest

from bradot_b200 import server

pytestmark = pytest.mark.gpu

MAX_KEYS = 65536  # BPC_SORT_NOALLOC_MAX_KEYS in include/bradot_physics_cuda.h


def _sort_noalloc(lib, torch, keys):
    """keys: host uint64 array -> sorted host uint64 array, through caller-owned device buffers."""
    d_in = torch.from_numpy(keys.view(np.int64).copy()).cuda()
    d_out = torch.empty_like(d_in)
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.bpc_sort_u64_noalloc(ctypes.c_void_p(d_in.data_ptr()), ctypes.c_void_p(d_out.data_ptr()), len(keys), ctypes.c_void_p(stream))
    assert rc == 0, lib.bpc_noalloc_last_error().decode()
    torch.cuda.synchronize()
    assert np.array_equal(d_in.cpu().numpy().view(np.uint64), keys), "the caller's input buffer was modified"
    return d_out.cpu().numpy().view(np.uint64)


def test_noalloc_sort_matches_regular_sort_twice_in_a_row():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.cuda.set_device(0)
    lib = server.load_library()
    ps = server.PhysicsServerCUDA(0)
    rng = np.random.default_rng(1234)
    # first call: the largest size allowed, full 64-bit keys; second call: small, with many duplicates
    for n, hi in ((MAX_KEYS, 2 ** 63), (5000, 2 ** 12)):
        keys = rng.integers(0, hi, n, dtype=np.uint64)
        want = ps.sort_u64(keys)                 # the existing entry point
        got = _sort_noalloc(lib, torch, keys)    # the allocation-free one
        assert np.array_equal(want, np.sort(keys))
        assert np.array_equal(got, want), "allocation-free sort differs from bpc_sort_u64 at n=%d" % n
    ps.finish()


def test_noalloc_sort_rejects_sizes_beyond_its_scratch():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.cuda.set_device(0)
    lib = server.load_library()
    d = torch.zeros(MAX_KEYS + 1, dtype=torch.int64, device="cuda")
    o = torch.empty_like(d)
    rc = lib.bpc_sort_u64_noalloc(ctypes.c_void_p(d.data_ptr()), ctypes.c_void_p(o.data_ptr()), MAX_KEYS + 1, ctypes.c_void_p(0))
    assert rc == -1
    msg = lib.bpc_noalloc_last_error().decode()
    assert str(MAX_KEYS) in msg and "exceed" in msg, msg
    # empty input is fine and touches nothing
    assert lib.bpc_sort_u64_noalloc(None, None, 0, None) == 0
