#!/usr/bin/env python3
"""Wall-clock time of the two-qubit noise channels on a 15-qubit density matrix (2^30 elements): the one-pass
k_depol2 and the vectorised k_dephase2, for qubit pairs whose bits sit high, at both ends, and at the bottom."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from quest_b200 import capi  # noqa: E402

lib = capi.load_product()
env = lib.createQuESTEnv()
nq = int(os.environ.get("QB_DENSITY_QUBITS", "15"))
q = lib.createDensityQureg(nq, env)
lib.initPlusState(q)
for t in range(nq):
    lib.rotateY(q, t, 0.1 * (t + 1))
lib.calcPurity(q)
N = float(1 << (2 * nq))
for a, b in ((3, 9), (0, nq - 1), (0, 1)):
    for name, fn in (("applyTwoQubitDepolariseError", lib.applyTwoQubitDepolariseError), ("applyTwoQubitDephaseError", lib.applyTwoQubitDephaseError)):
        fn(q, a, b, 0.1)
        lib.syncQuESTEnv(env)
        t0 = time.perf_counter()
        fn(q, a, b, 0.1)
        lib.syncQuESTEnv(env)
        ms = 1e3 * (time.perf_counter() - t0)
        print(f"{name}({a},{b}): {ms:7.2f} ms  ({32 * N / ms / 1e6:7.0f} GB/s of one read + one write of the matrix)", flush=True)
print("purity", lib.calcPurity(q))
lib.destroyQureg(q, env)
lib.destroyQuESTEnv(env)
