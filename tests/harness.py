"""Shared test plumbing: replay one converted reference test case through any QuEST-compatible library.

Mirrors what the reference's runner does per case (utilities/QuESTTest/QuESTCore.py:380-496): build the register
named by the case's type letter (z zero, p plus, d debug, c custom amplitudes, b bit-string; upper case = density
matrix), call the function under test, then read back total probability (P), per-qubit outcome probabilities
(M) and the full state (S), or the return value.
"""
import ctypes
import gzip
import json
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden_cases():
    with gzip.open(os.path.join(GOLDEN_DIR, "ref_unit_tests.json.gz"), "rt") as f:
        return json.load(f)


def make_qureg(lib, env, case):
    nq, qtype = case["nq"], case["qtype"]
    q = lib.createDensityQureg(nq, env) if case["density"] else lib.createQureg(nq, env)
    kind = qtype.upper()
    if kind == "Z":
        lib.initZeroState(q)
    elif kind == "P":
        lib.initPlusState(q)
    elif kind == "D":
        lib.initStateDebug(q)
    elif kind == "B":
        lib.initClassicalState(q, int(case["init"], 2))
    elif kind == "C":
        vals = np.array(case["init"], dtype=np.float64)
        lib.set_state(q, vals[0::2] + 1j * vals[1::2])
    else:
        raise ValueError(qtype)
    return q


def marshal(lib, arg):
    if isinstance(arg, dict):
        if "complex" in arg:
            return [lib.Complex(*arg["complex"])]
        if "matrix" in arg:
            m = arg["matrix"]
            return [lib.ComplexMatrix2(*[lib.Complex(m[i], m[i + 1]) for i in range(0, 8, 2)])]
        if "vector" in arg:
            return [lib.Vector(*arg["vector"])]
        if "ints" in arg:
            return [lib.int_array(arg["ints"])]
    return [arg]


def run_golden_case(lib, env, case):
    """Returns {'P': float, 'M': [[p0,p1]..], 'S': complex array} and/or {'ret': value} as the case asks."""
    q = make_qureg(lib, env, case)
    args = []
    for a in case["args"]:
        args += marshal(lib, a)
    ret = getattr(lib, case["func"])(q, *args)
    out = {}
    if "ret" in case:
        out["ret"] = [ret.real, ret.imag] if isinstance(ret, lib.Complex) else ret
    if "P" in case:
        out["P"] = lib.calcTotalProb(q)
    if "M" in case:
        out["M"] = [[lib.calcProbOfOutcome(q, t, 0), lib.calcProbOfOutcome(q, t, 1)] for t in range(case["nq"])]
    if "S" in case:
        out["S"] = lib.get_state(q)
    lib.destroyQureg(q, env)
    return out


def check_golden_case(case, got, tol=1e-10, rtol=0):
    """Harness tolerance 1e-10 absolute (reference QuESTTest/__main__.py:50); golden values carry 13 decimals."""
    for key in ("P", "ret"):
        if key in case:
            np.testing.assert_allclose(got[key], case[key], atol=tol, rtol=rtol, err_msg=f"{case['file']} {key}")
    if "M" in case:
        np.testing.assert_allclose(got["M"], case["M"], atol=tol, rtol=rtol, err_msg=f"{case['file']} M")
    if "S" in case:
        exp = np.array(case["S"][0::2]) + 1j * np.array(case["S"][1::2])
        np.testing.assert_allclose(got["S"], exp, atol=tol, rtol=rtol, err_msg=f"{case['file']} S")


def seed(lib, keys):
    arr = (ctypes.c_ulong * len(keys))(*keys)
    lib.seedQuEST(arr, len(keys))


def compare_with_cpu_register(product, q, cpu_q, chunk=1 << 24):
    """Amplitude-by-amplitude comparison of a product register with a register of a CPU library that keeps its
    state in host arrays (the reference build, the oracle), without materialising either as one complex array.
    Returns (max |diff| / max |amp|, every double bit-identical)."""
    n = int(q.numAmpsPerChunk)
    re = np.empty(min(chunk, n), dtype=product.np_real)
    im = np.empty_like(re)
    want_re = np.ctypeslib.as_array(cpu_q.stateVec.real, shape=(n,))
    want_im = np.ctypeslib.as_array(cpu_q.stateVec.imag, shape=(n,))
    worst, biggest, same = 0.0, 0.0, True
    for at in range(0, n, chunk):
        m = min(chunk, n - at)
        product.qb200_copyAmpsToHost(q, at, m, product.real_ptr(re), product.real_ptr(im))
        for got, want in ((re[:m], want_re[at:at + m]), (im[:m], want_im[at:at + m])):
            same = same and np.array_equal(got, want)
            worst = max(worst, float(np.abs(got - want).max()))
            biggest = max(biggest, float(np.abs(want).max()))
    return (worst / biggest if biggest else worst), same


def sum_squares_chunked(product, q, chunk=1 << 24):
    """sum |amp|^2 over a product register, from its amplitudes read back chunk by chunk and added pairwise by numpy
    (accurate to ~1e-16 relative): an independent check of the device reductions at sizes where the reference's own
    sequential sums are not accurate enough to serve as one."""
    n = int(q.numAmpsPerChunk)
    re = np.empty(min(chunk, n), dtype=product.np_real)
    im = np.empty_like(re)
    parts = []
    for at in range(0, n, chunk):
        m = min(chunk, n - at)
        product.qb200_copyAmpsToHost(q, at, m, product.real_ptr(re), product.real_ptr(im))
        parts.append(float(np.sum(re[:m].astype(np.float64) ** 2) + np.sum(im[:m].astype(np.float64) ** 2)))
    return float(np.sum(parts))
