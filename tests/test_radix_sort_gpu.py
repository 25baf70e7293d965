"""GPU parity: LSD radix sort on the scan (hw/hw4/programming/radixsort.cpp) through the C ABI vs the oracle,
the reference's golden pass outputs and sortedness properties."""
import ctypes as C

import numpy as np
import pytest

import oracle
from conftest import dev, dev_empty, host, golden

pytestmark = pytest.mark.gpu


def gpu_pass(ctx, keys, start_bit, num_bits):
    d_in = dev(keys)
    d_out = dev_empty(len(keys), np.uint32)
    ctx.call("cme213_radix_sort_pass_u32", d_in, d_out, len(keys), start_bit, num_bits)
    return host(d_out, np.uint32)


def gpu_sort(ctx, keys, num_bits=8):
    d_keys = dev(keys)
    d_tmp = dev_empty(len(keys), np.uint32)
    ctx.call("cme213_radix_sort_u32", d_keys, d_tmp, len(keys), num_bits)
    return host(d_keys, np.uint32)


def rand_keys(n, seed, hi=2**32):
    return np.random.default_rng(seed).integers(0, hi, n, dtype=np.uint64).astype(np.uint32)


def test_reference_golden_passes(ctx):
    gd = golden("radixsort_ref_n12000.npz")
    keys = gd["keys"]
    p0 = gpu_pass(ctx, keys, 0, 8)
    assert np.array_equal(p0, gd["pass_0_8"])
    assert np.array_equal(gpu_pass(ctx, p0, 8, 4), gd["pass_8_4_of_pass_0_8"])
    assert np.array_equal(gpu_pass(ctx, keys, 13, 5), gd["serial_pass_13_5"])
    assert np.array_equal(gpu_sort(ctx, keys, 8), gd["sorted_bits8"])
    assert np.array_equal(gpu_sort(ctx, keys, 4), gd["sorted_bits4"])


@pytest.mark.parametrize("n", [1, 2, 31, 33, 511, 512, 513, 8191, 8192, 8193, 65536, 65537, 8 * 8192 + 5, 1000003])
def test_pass_is_the_stable_counting_pass(ctx, n):
    keys = rand_keys(n, n)
    for start, bits in ((0, 8), (8, 8), (24, 8), (5, 3), (31, 1), (28, 8)):
        assert np.array_equal(gpu_pass(ctx, keys, start, bits), oracle.radix_pass_blocked_u32(keys, start, bits)), (start, bits)


@pytest.mark.parametrize("n", [1, 2, 100, 8192, 8193, 70001, 3000017])
@pytest.mark.parametrize("bits", [8, 4, 16])
def test_sort_equals_oracle_and_numpy(ctx, n, bits):
    keys = rand_keys(n, 3 * n + bits)
    got = gpu_sort(ctx, keys, bits)
    assert np.array_equal(got, oracle.radix_sort_u32(keys, bits))
    assert np.array_equal(got, np.sort(keys))


@pytest.mark.parametrize("bits", [1, 2, 3, 5, 6, 7, 11, 13])
def test_sort_any_digit_width(ctx, bits):
    keys = rand_keys(50021, bits)
    assert np.array_equal(gpu_sort(ctx, keys, bits), np.sort(keys))


def test_sort_degenerate_inputs(ctx):
    n = 40000
    for keys in (np.zeros(n, np.uint32), np.full(n, 0xFFFFFFFF, np.uint32), np.arange(n, dtype=np.uint32),
                 np.arange(n, dtype=np.uint32)[::-1].copy(), rand_keys(n, 1, hi=4), rand_keys(n, 2, hi=2**31) & np.uint32(0xFF00FF00)):
        assert np.array_equal(gpu_sort(ctx, keys), np.sort(keys))
    # rand()-style 31-bit keys as in the reference's main() (radixsort.cpp:173-175)
    keys = rand_keys(123457, 9, hi=2**31)
    assert np.array_equal(gpu_sort(ctx, keys), np.sort(keys))


def test_sort_unaligned_pointers_and_repeat(ctx):
    keys = rand_keys(30001, 12)
    d = dev(np.concatenate([[0], keys, [0, 0]]).astype(np.uint32))
    d_tmp = dev_empty(len(keys) + 3, np.uint32)
    for _ in range(2):   # second call sorts already sorted keys, workspace reused
        ctx.call("cme213_radix_sort_u32", d.data_ptr() + 4, d_tmp.data_ptr() + 8, len(keys), 8)
        assert np.array_equal(host(d, np.uint32)[1:-2], np.sort(keys))


def test_sort_errors_and_empty(ctx, cme):
    ctx.call("cme213_radix_sort_u32", None, None, 0, 8)
    d = dev(rand_keys(16, 1))
    t = dev_empty(16, np.uint32)
    for args in ((d, t, 16, 0), (d, t, 16, 17), (d, d, 16, 8), (d, t, -1, 8), (d, None, 16, 8)):
        with pytest.raises(cme.Cme213Error):
            ctx.call("cme213_radix_sort_u32", *args)
    for args in ((d, t, 16, 0, 9), (d, t, 16, 32, 8), (d, t, 16, -1, 8)):
        with pytest.raises(cme.Cme213Error):
            ctx.call("cme213_radix_sort_pass_u32", *args)


def test_sort_host_twin(cme):
    keys = rand_keys(200003, 21, hi=2**31)
    work = keys.copy()
    ms = C.c_float()
    rc = cme.load().cme213_radix_sort_u32_host(work.ctypes.data_as(C.c_void_p), C.c_int64(len(work)), 8, C.byref(ms))
    assert rc == 0 and ms.value > 0
    assert np.array_equal(work, np.sort(keys))


def test_pass_host_twin_reference_signature(cme):
    # radixSortParallelPass(keys, sorted, numBits, startBit, blockSize): also digits wider than 8 bits
    gd = golden("radixsort_ref_n12000.npz")
    keys = np.ascontiguousarray(gd["keys"])
    out = np.empty_like(keys)
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    fn = cme.load().cme213_radix_sort_pass_u32_host
    assert fn(P(keys), P(out), len(keys), 8, 0, 8192) == 0
    assert np.array_equal(out, gd["pass_0_8"])
    for bits, start in ((16, 0), (11, 5), (13, 16), (12, 24), (5, 13)):
        assert fn(P(keys), P(out), len(keys), bits, start, 1000) == 0
        assert np.array_equal(out, oracle.radix_pass_u32(keys, start, bits)), (bits, start)
    assert fn(P(keys), P(out), len(keys), 17, 0, 8192) == cme.capi.ERR_INVALID
    assert fn(None, P(out), len(keys), 8, 0, 8192) == cme.capi.ERR_INVALID


def test_sort_full_size_properties(ctx):
    # the reference's default size is 40 M keys (radixsort.cpp:165); checked at 2^28 on the device:
    # sortedness, and the multiset is kept (sum of the keys and of a hash of the keys).
    import torch
    n = 1 << 28
    x = torch.empty(n, dtype=torch.int32, device="cuda")
    ctx.call("cme213_fill_hash_u32", x, n, 11, 0)
    before = x.clone()
    tmp = torch.empty_like(x)
    ctx.call("cme213_radix_sort_u32", x, tmp, n, 8)
    torch.cuda.synchronize()
    u = x.to(torch.int64) & 0xFFFFFFFF
    assert bool((u[1:] >= u[:-1]).all())
    b = before.to(torch.int64) & 0xFFFFFFFF
    assert int(u.sum()) == int(b.sum())
    assert int((u * 2654435761 % 4294967291).sum()) == int((b * 2654435761 % 4294967291).sum())
    del u, b
    # a slice against the CPU: the 2^20 smallest keys
    want = np.sort(host(before, np.uint32))[: 1 << 20]
    assert np.array_equal(host(x[: 1 << 20], np.uint32), want)
