#!/usr/bin/env python3
"""One launch of every kernel class on a 2^30-amplitude register, in a fixed order, for `ncu -k regex:<kernel>`:

    python tools/profile_classes.py            # prints the CUDA-event time of every call as well

State-vector (30 qubits): k_diag (phaseShift on qubit 16, controlledPhaseShift, 3-qubit phase), k_pairs (hadamard,
compactUnitary, controlledNot), k_probZero (qubit 0 and qubit 16), k_probAllZero (second query), k_sumSquares,
k_collapse.  Density matrix (15 qubits): k_noise1 for the three channels, then the noise tier of the fused pass on a
layer of the configs[3] circuit."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from quest_b200 import capi, circuits  # noqa: E402

lib = capi.load_product()
env = lib.createQuESTEnv()
lib.lib.qb200_getStream.restype = ctypes.c_void_p
lib.lib.qb200_setFusion.argtypes = [ctypes.c_int]
stream = torch.cuda.ExternalStream(lib.lib.qb200_getStream())
nq = int(os.environ.get("QB_PROFILE_QUBITS", "30"))
N = float(1 << nq)


def timed(label, algorithmic_bytes, fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    out = fn()
    lib.lib.qb200_flush()
    e1.record(stream)
    e1.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"{label:46s} {ms:8.3f} ms  {algorithmic_bytes / ms / 1e6:8.1f} GB/s of {algorithmic_bytes / 1e9:6.2f} GB algorithmic", flush=True)
    return out


lib.lib.qb200_setFusion(0)  # one kernel per call
q = lib.createQureg(nq, env)
lib.initPlusState(q)
for t in range(nq):
    lib.rotateY(q, t, 0.1 * (t + 1))
lib.calcTotalProb(q)
a, b = lib.complex(circuits.SWEEP_ALPHA), lib.complex(circuits.SWEEP_BETA)
timed("k_pairs hadamard(16)", 32 * N, lambda: lib.hadamard(q, 16))
timed("k_pairs compactUnitary(0)", 32 * N, lambda: lib.compactUnitary(q, 0, a, b))
timed("k_pairs controlledNot(17 -> 16)", 16 * N, lambda: lib.controlledNot(q, 17, 16))
timed("k_diag phaseShift(16)", 16 * N, lambda: lib.phaseShift(q, 16, 0.3))
timed("k_diag controlledPhaseShift(20, 16)", 8 * N, lambda: lib.controlledPhaseShift(q, 20, 16, 0.3))
timed("k_diag multiControlledPhaseShift(16,17,18)", 4 * N, lambda: lib.multiControlledPhaseShift(q, lib.int_array([16, 17, 18]), 3, 0.7))
timed("k_probZero calcProbOfOutcome(0)", 8 * N, lambda: lib.calcProbOfOutcome(q, 0, 0))
lib.hadamard(q, 3)  # state changed: the next query is a first query again
timed("k_probZero calcProbOfOutcome(16)", 8 * N, lambda: lib.calcProbOfOutcome(q, 16, 0))
timed("k_probAllZero calcProbOfOutcome(5), 2nd query", 16 * N, lambda: lib.calcProbOfOutcome(q, 5, 0))
timed("k_sumSquares calcTotalProb", 16 * N, lambda: lib.calcTotalProb(q))
timed("k_probZero + k_collapse collapseToOutcome(16,0)", 8 * N + 24 * N, lambda: lib.collapseToOutcome(q, 16, 0))
lib.destroyQureg(q, env)

nd = nq // 2
rho = lib.createDensityQureg(nd, env)
lib.initPlusState(rho)
for t in range(nd):
    lib.rotateY(rho, t, 0.1 * (t + 1))
lib.calcPurity(rho)
timed("k_noise1 dephase(7)", 16 * N, lambda: lib.applyOneQubitDephaseError(rho, 7, 0.02))
timed("k_noise1 depolarise(7)", 32 * N, lambda: lib.applyOneQubitDepolariseError(rho, 7, 0.03))
timed("k_noise1 damping(7)", 32 * N, lambda: lib.applyOneQubitDampingError(rho, 7, 0.05))
timed("k_noise1 depolarise(0) (row bit in the vector)", 32 * N, lambda: lib.applyOneQubitDepolariseError(rho, 0, 0.03))
lib.lib.qb200_setFusion(1)
layer = [op for op in circuits.config4_noisy(nd, layers=1) if op[0] != "calcPurity"]
lib.lib.qb200_getLaunchCount.restype = ctypes.c_ulonglong
n0 = lib.lib.qb200_getLaunchCount()
timed(f"k_tile noise tier: one configs[3] layer ({len(layer)} calls)", 0, lambda: [circuits.apply_op(lib, rho, op) for op in layer])
print(f"   ... in {lib.lib.qb200_getLaunchCount() - n0} launches")
print("purity", lib.calcPurity(rho))
lib.destroyQureg(rho, env)
lib.destroyQuESTEnv(env)
