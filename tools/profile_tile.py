#!/usr/bin/env python3
"""One configs[1] sweep step at 30 qubits through the fused pass, for ncu (`-k regex:k_tile`) and quick A/B timing:
prints the device time of every fused pass of `--steps` steps.  QUEST_B200_LIB selects another build of the product."""
import argparse
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from quest_b200 import capi, circuits  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--qubits", type=int, default=30)
ap.add_argument("--steps", type=int, default=1)
ap.add_argument("--fma", type=int, default=0)
ap.add_argument("--circuit", default="sweep", choices=["sweep", "layers", "density"])
ap.add_argument("--skip", default="", help="comma-separated API names to leave out of the circuit (ablation)")
args = ap.parse_args()
lib = capi.load_product()
env = lib.createQuESTEnv()
lib.lib.qb200_setKernelTiming.argtypes = [ctypes.c_int]
lib.lib.qb200_setArithmetic.argtypes = [ctypes.c_int]
nq = args.qubits
if args.circuit == "density":
    q = lib.createDensityQureg(nq // 2, env)
    ops = [op for op in circuits.config4_noisy(nq // 2, layers=1) if op[0] != "calcPurity"]
else:
    q = lib.createQureg(nq, env)
    ops = circuits.config2_sweep(nq) if args.circuit == "sweep" else circuits.config1_layers(nq, layers=2)
ops = [op for op in ops if op[0] not in args.skip.split(",")]
lib.initPlusState(q)
for t in range(q.numQubitsRepresented):
    lib.rotateY(q, t, 0.1 * (t + 1))
lib.lib.qb200_flush()
lib.lib.qb200_setArithmetic(args.fma)
for rep in range(args.steps + 1):
    lib.lib.qb200_setKernelTiming(1)
    for op in ops:
        circuits.apply_op(lib, q, op)
    lib.lib.qb200_flush()
    ms, n = ctypes.c_double(), ctypes.c_ulonglong()
    lib.lib.qb200_getFusedKernelTime(ctypes.byref(ms), ctypes.byref(n))
    if rep:
        print(f"{os.environ.get('QUEST_B200_LIB', 'product')}: {args.circuit} {nq}q fma={args.fma} skip={args.skip or '-'} ({len(ops)} gates): {n.value} fused passes, "
              f"{ms.value:.2f} ms in total, {ms.value / max(n.value, 1):.2f} ms per pass", flush=True)
print("total prob", lib.calcTotalProb(q))
lib.destroyQureg(q, env)
lib.destroyQuESTEnv(env)
