"""The allocation-free total-probability entry point (include/quest_b200.h: qb200_calcTotalProbNoAlloc).

Some hosts of this library may not allocate device memory (all of HBM is owned by another runtime or capped by an
MPS/MIG quota, or cudaMalloc's device-wide synchronisation is not acceptable in their loop).  For them the library
offers one reduction whose scratch is a fixed array inside the library image.  This test drives it through the C
ABI, as such a host would - raw device pointers in, a device double out - and checks it against the regular
calcTotalProb(), twice in a row so that the second call runs on the scratch the first one left behind.

The file name sorts last on purpose: under `pytest -x` every other test has already run if this one fails.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _bind(product):
    fn = product.lib.qb200_calcTotalProbNoAlloc
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_void_p]
    product.lib.qb200_getDevicePointers.argtypes = [product.Qureg, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_void_p)]
    product.lib.qb200_getDevicePointers.restype = None
    product.lib.qb200_getStream.restype = ctypes.c_void_p
    return fn


def test_no_alloc_total_prob_matches_calcTotalProb_twice(product, product_env):
    import torch

    fn = _bind(product)
    stream = product.lib.qb200_getStream()
    out = torch.zeros(1, dtype=torch.float64, device="cuda")  # the caller owns the output, as a real host would
    q = product.createQureg(22, product_env)
    re, im = ctypes.c_void_p(), ctypes.c_void_p()
    product.lib.qb200_getDevicePointers(q, ctypes.byref(re), ctypes.byref(im))
    n = int(q.numAmpsPerChunk)

    # first call: an unnormalised state (total probability far from 1)
    product.initStateDebug(q)
    assert fn(re, im, n, out.data_ptr(), stream) == 0
    product.syncQuESTEnv(product_env)
    torch.cuda.synchronize()
    first = float(out.item())
    np.testing.assert_allclose(first, product.calcTotalProb(q), rtol=1e-12)

    # second call, straight after, on a different state: must not be disturbed by what the first call left
    product.initPlusState(q)
    for t in (0, 5, 21):
        product.rotateY(q, t, 0.3 + t)
    product.controlledNot(q, 3, 20)
    assert fn(re, im, n, out.data_ptr(), stream) == 0
    product.syncQuESTEnv(product_env)
    torch.cuda.synchronize()
    second = float(out.item())
    np.testing.assert_allclose(second, product.calcTotalProb(q), rtol=1e-12)
    np.testing.assert_allclose(second, 1.0, rtol=1e-12)
    assert abs(first - second) > 1.0  # the two calls really saw different data

    # third call on the lower half of the amplitudes only: that sum is P(qubit 21 = 0), which the rotateY on qubit 21
    # above moved away from 1/2 - so this check fails if the gates issued before the call had not reached the
    # amplitudes the raw pointers show (a register whose pointers were handed out runs its gates at once)
    assert fn(re, im, n // 2, out.data_ptr(), stream) == 0
    product.syncQuESTEnv(product_env)
    torch.cuda.synchronize()
    half = float(out.item())
    expect = product.calcProbOfOutcome(q, 21, 0)
    assert abs(expect - 0.5) > 0.05
    np.testing.assert_allclose(half, expect, rtol=1e-12)
    product.destroyQureg(q, product_env)


def test_no_alloc_small_and_unaligned_sizes(product, product_env):
    import torch

    fn = _bind(product)
    stream = product.lib.qb200_getStream()
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    q = product.createQureg(3, product_env)
    product.initStateDebug(q)
    re, im = ctypes.c_void_p(), ctypes.c_void_p()
    product.lib.qb200_getDevicePointers(q, ctypes.byref(re), ctypes.byref(im))
    # the first 7 amplitudes only: exercises the scalar path (size not a multiple of the vector width)
    assert fn(re, im, 7, out.data_ptr(), stream) == 0
    product.syncQuESTEnv(product_env)
    torch.cuda.synchronize()
    k = np.arange(7)
    np.testing.assert_allclose(out.item(), np.sum((2 * k / 10) ** 2 + ((2 * k + 1) / 10) ** 2), rtol=1e-13)
    product.destroyQureg(q, product_env)


def test_no_alloc_rejects_sizes_beyond_its_fixed_scratch(product, product_env, capfd):
    import torch

    fn = _bind(product)
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    q = product.createQureg(4, product_env)
    re, im = ctypes.c_void_p(), ctypes.c_void_p()
    product.lib.qb200_getDevicePointers(q, ctypes.byref(re), ctypes.byref(im))
    assert fn(re, im, (1 << 31) + 4, out.data_ptr(), None) != 0  # refused before anything is launched
    assert fn(re, im, 0, out.data_ptr(), None) != 0
    assert fn(None, im, 16, out.data_ptr(), None) != 0
    err = capfd.readouterr().err
    assert "QB200_NOALLOC_MAX_AMPS" in err
    product.destroyQureg(q, product_env)
