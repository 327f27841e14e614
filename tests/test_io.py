"""Bulk state I/O: setAmps / read-back from pageable and pinned host memory through the C ABI.
Large pageable transfers take the staged multi-thread path (qb_io.cu); the bytes that arrive must be the bytes sent."""
import ctypes

import numpy as np
import pytest

from quest_b200 import capi

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def product():
    lib = capi.load_product()
    lib.lib.qb200_getPipelinedTransferCount.restype = ctypes.c_ulonglong
    return lib


@pytest.fixture(scope="module")
def env(product):
    return product.createQuESTEnv()


def _ptr(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def test_pageable_round_trip_is_exact_and_takes_the_staged_path(product, env):
    nq = 23  # 64 MiB per array: 16 staging chunks, split over the worker threads
    n = 1 << nq
    rng = np.random.default_rng(5)
    re, im = rng.standard_normal(n), rng.standard_normal(n)
    q = product.createQureg(nq, env)
    before = product.lib.qb200_getPipelinedTransferCount()
    product.setAmps(q, 0, _ptr(re), _ptr(im), n)
    out_re, out_im = np.full(n, np.nan), np.full(n, np.nan)
    product.qb200_copyAmpsToHost(q, 0, n, _ptr(out_re), _ptr(out_im))
    after = product.lib.qb200_getPipelinedTransferCount()
    assert after - before == 4  # two uploads, two downloads
    assert np.array_equal(out_re, re) and np.array_equal(out_im, im)
    # the uploaded state is what the kernels see
    assert abs(product.calcTotalProb(q) - float(np.sum(re * re + im * im))) < 1e-6 * n
    product.destroyQureg(q, env)


@pytest.mark.parametrize("start,count", [(1, (1 << 21) + 3), (12345, (1 << 22) - 7777), ((1 << 23) - (1 << 20) - 5, (1 << 20) + 5)])
def test_partial_setamps_at_odd_offsets(product, env, start, count):
    """Slices that are neither chunk-aligned nor a multiple of the chunk size, around the staged-path threshold."""
    nq = 23
    n = 1 << nq
    rng = np.random.default_rng(start)
    q = product.createQureg(nq, env)
    product.initZeroState(q)
    re, im = rng.standard_normal(count), rng.standard_normal(count)
    product.setAmps(q, start, _ptr(re), _ptr(im), count)
    got = product.get_state(q)
    want = np.zeros(n, dtype=np.complex128)
    want[0] = 1
    want[start:start + count] = re + 1j * im
    assert np.array_equal(got, want)
    product.destroyQureg(q, env)


def test_pinned_memory_goes_straight_through(product, env):
    torch = pytest.importorskip("torch")
    nq = 22
    n = 1 << nq
    re = torch.randn(n, dtype=torch.float64).pin_memory()
    im = torch.randn(n, dtype=torch.float64).pin_memory()
    q = product.createQureg(nq, env)
    before = product.lib.qb200_getPipelinedTransferCount()
    product.setAmps(q, 0, ctypes.cast(re.data_ptr(), ctypes.POINTER(ctypes.c_double)),
                    ctypes.cast(im.data_ptr(), ctypes.POINTER(ctypes.c_double)), n)
    assert product.lib.qb200_getPipelinedTransferCount() == before
    got = product.get_state(q)
    assert np.array_equal(got.real, re.numpy()) and np.array_equal(got.imag, im.numpy())
    product.destroyQureg(q, env)


def test_deferred_upload_overlaps_passes_and_gives_the_same_state(product, env):
    """qb200_setDeferredUpload(1): a whole-register setAmps from pinned memory returns at once, the leading passes of
    the gates recorded afterwards run underneath the transfer slab by slab, and the result equals the blocking path
    bit for bit - also when a value-returning call, a second upload or a non-fusable call comes straight after."""
    import torch

    from quest_b200 import circuits

    lib = product.lib
    lib.qb200_setDeferredUpload.argtypes = [ctypes.c_int]
    lib.qb200_getDeferredUploadCount.restype = ctypes.c_ulonglong
    lib.qb200_getPassesUnderUploadCount.restype = ctypes.c_ulonglong
    nq = 25
    n = 1 << nq
    gen = torch.Generator().manual_seed(9)
    host_re = torch.randn(n, dtype=torch.float64, generator=gen).pin_memory()
    host_im = torch.randn(n, dtype=torch.float64, generator=gen).pin_memory()
    pre = ctypes.cast(host_re.data_ptr(), ctypes.POINTER(ctypes.c_double))
    pim = ctypes.cast(host_im.data_ptr(), ctypes.POINTER(ctypes.c_double))
    ops = circuits.config2_sweep(nq)  # low targets first: three of its passes can follow the upload
    q = product.createQureg(nq, env)

    def run(deferred, variant):
        lib.qb200_setDeferredUpload(deferred)
        product.setAmps(q, 0, pre, pim, n)
        if variant == "probe_first":      # a value-returning call straight after the upload
            first = product.getRealAmp(q, n - 1)
            assert first == host_re[n - 1].item()
        elif variant == "twice":          # a second whole-register upload on top of the first
            product.setAmps(q, 0, pre, pim, n)
        elif variant == "unfused_first":  # a call that is executed at once (not recorded)
            product.collapseToOutcome(q, nq - 1, 0)
        for op in ops:
            circuits.apply_op(product, q, op)
        total = product.calcTotalProb(q)
        out_re, out_im = np.empty(n), np.empty(n)
        product.qb200_copyAmpsToHost(q, 0, n, _ptr(out_re), _ptr(out_im))
        lib.qb200_setDeferredUpload(0)
        return total, out_re, out_im

    for variant in ("plain", "probe_first", "twice", "unfused_first"):
        want = run(0, variant)
        d0, u0 = lib.qb200_getDeferredUploadCount(), lib.qb200_getPassesUnderUploadCount()
        got = run(1, variant)
        assert lib.qb200_getDeferredUploadCount() - d0 == (2 if variant == "twice" else 1)
        if variant in ("plain", "twice"):
            assert lib.qb200_getPassesUnderUploadCount() - u0 >= 2  # passes really ran underneath the transfer
        assert got[0] == want[0]
        assert np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
    # pageable memory never takes the deferred path
    lib.qb200_setDeferredUpload(1)
    d0 = lib.qb200_getDeferredUploadCount()
    re2 = np.zeros(n); re2[0] = 1.0
    product.setAmps(q, 0, _ptr(re2), _ptr(np.zeros(n)), n)
    assert lib.qb200_getDeferredUploadCount() == d0 and product.calcTotalProb(q) == 1.0
    lib.qb200_setDeferredUpload(0)
    product.destroyQureg(q, env)
