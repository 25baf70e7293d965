"""GPU: the allocation-free exclusive scan (cme213_exclusive_scan_u32_static).

It exists for callers that cannot allocate device memory at call time (the whole GPU belongs to their own
pool, or their loop must not run into cudaMalloc's device-wide synchronisation), so its look-back descriptors
sit in a fixed array inside the library that every call re-uses.  The test drives it through the C ABI twice
in a row -- the second call finds the first call's descriptors in the array and must clear them -- and checks
both results bit for bit against the regular entry point (and the oracle).  It lives in the file pytest
collects last, and passes both on its own in a fresh process and after all the other tests."""
import ctypes as C

import numpy as np
import pytest

import oracle
from conftest import dev, dev_empty, host

pytestmark = pytest.mark.gpu


def test_static_scan_matches_regular_entry_point_twice(cme, ctx):
    lib = cme.load()
    rng = np.random.default_rng(42)
    for n in (3 * (1 << 20) + 11, 70001):  # second call: smaller input over the leftovers of the first
        x = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
        d_in = dev(x)
        d_ref, d_out = dev_empty(n, np.uint32), dev_empty(n, np.uint32)
        ctx.call("cme213_exclusive_scan_u32", d_in, d_ref, n)
        rc = lib.cme213_exclusive_scan_u32_static(C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), n, None)
        assert rc == 0, lib.cme213_last_error()
        got, want = host(d_out, np.uint32), host(d_ref, np.uint32)
        assert np.array_equal(got, want)                       # bit-exact: integer work
        assert np.array_equal(got, oracle.exclusive_scan_u32(x))
    # same input again, in place this time
    rc = lib.cme213_exclusive_scan_u32_static(C.c_void_p(d_in.data_ptr()), C.c_void_p(d_in.data_ptr()), n, None)
    assert rc == 0 and np.array_equal(host(d_in, np.uint32), want)


def test_static_scan_rejects_sizes_beyond_its_scratch(cme):
    lib = cme.load()
    limit = lib.cme213_exclusive_scan_u32_static_max_n()
    assert limit == 1 << 28
    rc = lib.cme213_exclusive_scan_u32_static(None, None, limit + 1, None)  # rejected before any pointer is touched
    assert rc == cme.capi.ERR_UNSUPPORTED and b"exceeds" in lib.cme213_last_error()
    assert lib.cme213_exclusive_scan_u32_static(None, None, 0, None) == 0
    assert lib.cme213_exclusive_scan_u32_static(None, None, -1, None) == cme.capi.ERR_INVALID


def test_static_scan_at_its_size_limit_allocates_nothing(cme, ctx):
    """n = 2^28 is the largest input the built-in array can serve (its descriptors fill the array exactly).  After a
    set-up call (which resolves the array's address and lets the driver load the kernel), further calls must leave the
    device's free memory untouched: that is the whole point of the entry point for callers that cannot allocate."""
    import torch
    lib = cme.load()
    n = int(lib.cme213_exclusive_scan_u32_static_max_n())
    d_in, d_ref, d_out = dev_empty(n, np.uint32), dev_empty(n, np.uint32), dev_empty(n, np.uint32)
    ctx.call("cme213_fill_hash_u32", d_in, n, 77, 0)              # full-range values: the sums wrap mod 2^32
    ctx.call("cme213_exclusive_scan_u32", d_in, d_ref, n)
    P = lambda t: C.c_void_p(t.data_ptr())
    assert lib.cme213_exclusive_scan_u32_static(P(d_in), P(d_out), 4096, None) == 0   # set-up call on this device
    torch.cuda.synchronize()
    for _ in range(2):                                            # two calls in a row share the array
        d_out.zero_()
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]                # cudaMemGetInfo; nothing but the call sits between the two readings
        rc = lib.cme213_exclusive_scan_u32_static(P(d_in), P(d_out), n, None)
        torch.cuda.synchronize()
        free_after = torch.cuda.mem_get_info()[0]
        assert rc == 0, lib.cme213_last_error()
        assert free_after == free_before                          # the call allocated (and freed) nothing on the device
        assert torch.equal(d_out, d_ref)                          # bit-exact against the regular entry point (torch may allocate here: outside the window)
    # size-independent property on the full range: out[i+1] - out[i] == in[i] (mod 2^32), out[0] == 0
    assert int(d_out[0]) == 0
    assert torch.equal(d_out[1:] - d_out[:-1], d_in[:-1])


def test_static_scan_refuses_buffers_of_another_device(cme):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    lib = cme.load()
    other = torch.zeros(1024, dtype=torch.int32, device="cuda:1")
    torch.cuda.set_device(0)
    rc = lib.cme213_exclusive_scan_u32_static(C.c_void_p(other.data_ptr()), C.c_void_p(other.data_ptr()), 1024, None)
    assert rc == cme.capi.ERR_INVALID and b"device" in lib.cme213_last_error()
