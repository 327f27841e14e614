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
