#!/usr/bin/env python3
"""Per-gate-class device timing sweep (CUDA events on the engine's stream): every target for the config-2 gate
classes plus reductions and noise.  Prints a table and writes JSON; used to find which kernel/target to tune."""
import argparse
import ctypes
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from quest_b200 import capi, circuits  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qubits", type=int, default=30)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--targets", type=str, default="all")
    ap.add_argument("--out", type=str, default="gpurun_out/gate_sweep.json")
    ap.add_argument("--density", action="store_true")
    args = ap.parse_args()
    lib = capi.load_product()
    env = lib.createQuESTEnv()
    lib.lib.qb200_getStream.restype = ctypes.c_void_p
    lib.lib.qb200_setFusion.argtypes = [ctypes.c_int]
    lib.lib.qb200_setFusion(0)  # this tool times the single-gate kernels: one launch per call
    stream = torch.cuda.ExternalStream(lib.lib.qb200_getStream())
    nq = args.qubits
    peaks = json.load(open("MEASURED_PEAKS.json")) if os.path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6650.0}
    peak = peaks["hbm_gbs"]
    if args.density:
        q = lib.createDensityQureg(nq // 2, env)
        nrep = nq // 2
    else:
        q = lib.createQureg(nq, env)
        nrep = nq
    lib.initPlusState(q)
    N = float(q.numAmpsTotal)
    targets = list(range(nrep)) if args.targets == "all" else [int(t) for t in args.targets.split(",")]
    a, b = lib.complex(circuits.SWEEP_ALPHA), lib.complex(circuits.SWEEP_BETA)

    def timed(fn, reps=args.reps):
        fn()  # warm
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn()
            e1.record(stream)
            e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        return ts[len(ts) // 2], ts[0]

    rows = []
    if args.density:
        classes = {
            "hadamard(2 passes)": (lambda t: lib.hadamard(q, t), 64),
            "dephase": (lambda t: lib.applyOneQubitDephaseError(q, t, 0.02), 16),
            "depolarise": (lambda t: lib.applyOneQubitDepolariseError(q, t, 0.03), 32),
            "damping": (lambda t: lib.applyOneQubitDampingError(q, t, 0.05), 32),
        }
    else:
        classes = {
            "hadamard": (lambda t: lib.hadamard(q, t), 32),
            "compactUnitary": (lambda t: lib.compactUnitary(q, t, a, b), 32),
            "controlledNot": (lambda t: lib.controlledNot(q, (t + 1) % nq, t), 16),
            "mcPhaseShift3": (lambda t: lib.multiControlledPhaseShift(q, lib.int_array([t, (t + 1) % nq, (t + 2) % nq]), 3, 0.7), 4),
            "phaseShift": (lambda t: lib.phaseShift(q, t, 0.3), 16),
            "calcProbOfOutcome": (lambda t: lib.calcProbOfOutcome(q, t, 0), 8),
        }
    for name, (fn, bytes_per_amp) in classes.items():
        for t in targets:
            med, best = timed(lambda: fn(t))
            gbs = bytes_per_amp * N / (med * 1e-3) / 1e9
            rows.append({"gate": name, "target": t, "ms_median": med, "ms_best": best, "GBps": gbs, "frac_measured": gbs / peak})
            print(f"{name:20s} t={t:2d}  {med:8.3f} ms (best {best:8.3f})  {gbs:8.1f} GB/s  {gbs / peak:5.2f} of measured", flush=True)
    # (the trace of a density matrix reads only 2^q diagonal elements: latency-bound, not listed)
    extra = {"calcPurity": (lambda: lib.calcPurity(q), 16)} if args.density else {"calcTotalProb": (lambda: lib.calcTotalProb(q), 16)}
    for name, (fn, bpa) in extra.items():
        med, best = timed(fn)
        gbs = bpa * N / (med * 1e-3) / 1e9
        rows.append({"gate": name, "target": -1, "ms_median": med, "ms_best": best, "GBps": gbs, "frac_measured": gbs / peak})
        print(f"{name:20s}       {med:8.3f} ms (best {best:8.3f})  {gbs:8.1f} GB/s  {gbs / peak:5.2f} of measured", flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump({"qubits": nq, "density": args.density, "peak_gbs": peak, "rows": rows}, open(args.out, "w"), indent=1)
    lib.destroyQureg(q, env)


if __name__ == "__main__":
    main()
