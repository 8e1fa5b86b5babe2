"""h2b_batch_invert_device_noalloc -- the one entry point of libh2b200.so that never allocates.

It exists for hosts that cannot let the library call cudaMalloc: a prover that already holds all of HBM in its own arena, a
pipeline that must not take the device-wide stall of an allocation while other streams run, a component with a buffer and a
stream but no h2b_ctx.  Its scratch is an array inside the library image (2 MiB per GPU, n <= 2^26), shared by all calls.

What is checked here, through the C ABI only: the result equals h2b_batch_invert_device's bit for bit (both run the same three
kernels; integer work, so the tolerance is zero) on TWO calls in a row -- the second call finds the first call's leftovers in the
shared scratch and must not care, whatever the order of sizes -- and sizes beyond the limit are refused with a message.

The file sorts last on purpose (pytest -x stops at the first failure; one here must not hide the suite), and it depends on no
state left by the other tests: it passes alone in a fresh process and after everything else.
"""
import ctypes

import numpy as np
import pytest

import h2b200

R_MOD = 0x30644E72E131A029B85045B68181585D2833E84879B9709143E1F593F0000001


def _random_fr(n, seed):
    """n field elements as (n, 4) uint64 limbs, every one < r (top limb kept below r's), with a few zeros (they stay zero)."""
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 1 << 63, size=(n, 4), dtype=np.uint64)
    a[:, 3] %= np.uint64(R_MOD >> 192)
    a[3::7] = 0
    return a


def _noalloc(lib, d, n):
    msg = ctypes.c_char_p()
    rc = lib.h2b_batch_invert_device_noalloc(ctypes.c_void_p(d.data_ptr()), n, None, ctypes.byref(msg))
    return rc, (msg.value or b"").decode()


@pytest.mark.gpu
def test_noalloc_batch_invert_matches_context_call_twice(ctx):
    import torch
    lib = h2b200.load()
    # a large call first, then a small one: the second sees a scratch full of the first one's tile products
    for i, (n, seed) in enumerate((((1 << 20) + 4097, 1), (5000, 2), (1, 3), (2048, 4))):
        a = _random_fr(n, seed)
        want = torch.from_numpy(a.view(np.int64)).cuda()
        ctx.check(lib.h2b_batch_invert_device(ctx.handle, ctypes.c_void_p(want.data_ptr()), n))
        got = torch.from_numpy(a.view(np.int64)).cuda()
        torch.cuda.synchronize()                         # the input copy is done before our kernels start on the default stream
        mem_before = torch.cuda.mem_get_info()[0]
        rc, msg = _noalloc(lib, got, n)
        assert rc == 0, msg
        torch.cuda.synchronize()                         # the call only enqueues
        if i:          # (the very first launch may make the driver load the module that holds the scratch array)
            assert torch.cuda.mem_get_info()[0] == mem_before, "the no-alloc entry point changed the free device memory"
        assert torch.equal(got, want), n
        zeros = np.flatnonzero(~a.any(axis=1))
        assert not got.cpu().numpy()[zeros].any()        # zeros untouched, as ff::BatchInvert leaves them
    # two calls back to back on one stream with no synchronisation in between: inverting twice gives the input back
    a = _random_fr(70001, 9)
    d = torch.from_numpy(a.view(np.int64)).cuda()
    torch.cuda.synchronize()
    assert _noalloc(lib, d, 70001)[0] == 0 and _noalloc(lib, d, 70001)[0] == 0
    torch.cuda.synchronize()
    assert np.array_equal(d.cpu().numpy().view(np.uint64), a)


def test_noalloc_rejects_oversize_and_null():
    """Needs no GPU: the argument checks come first."""
    lib = h2b200.load()
    msg = ctypes.c_char_p()
    # the size check comes before anything touches the pointer or the device
    rc = lib.h2b_batch_invert_device_noalloc(ctypes.c_void_p(256), (1 << 26) + 1, None, ctypes.byref(msg))
    assert rc == -1 and b"H2B_NOALLOC_MAX_ELEMS" in msg.value
    rc = lib.h2b_batch_invert_device_noalloc(None, 8, None, ctypes.byref(msg))
    assert rc == -1 and b"null" in msg.value
    assert lib.h2b_batch_invert_device_noalloc(None, 0, None, None) == 0          # empty input: nothing to do
