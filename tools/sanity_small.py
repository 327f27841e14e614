#!/usr/bin/env python3
"""Small all-kernel exercise for compute-sanitizer (memcheck): every gate family, reductions, collapse, noise, on
tiny and odd-sized registers, fused and gate-at-a-time, all vector widths of the single-gate kernels."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from quest_b200 import capi, circuits  # noqa: E402

lib = capi.load_product()
env = lib.createQuESTEnv()
lib.lib.qb200_setFusion.argtypes = [ctypes.c_int]
for fusion in (0, 1):
    lib.lib.qb200_setFusion(fusion)
    for nq, dens, seed in [(1, False, 1), (2, False, 2), (4, False, 3), (5, False, 4), (11, False, 5), (1, True, 6), (2, True, 7), (5, True, 8)]:
        ops = circuits.random_circuit(nq, 150, seed, density=dens, measure=True)
        q = lib.createDensityQureg(nq, env) if dens else lib.createQureg(nq, env)
        circuits.prepare(lib, q, seed)
        circuits.run_circuit(lib, q, ops, seed=seed)
        st = lib.get_state(q)
        assert np.isfinite(st.view(np.float64)).all()
        lib.destroyQureg(q, env)
lib.destroyQuESTEnv(env)
print("sanity_small ok")
