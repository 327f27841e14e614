#!/usr/bin/env python3
"""Device-timed runs of the other BASELINE.json configs on one GPU (parity cases that are also worth a number):
configs[2] 30-qubit QFT + P(1) of every qubit, configs[3] 15-qubit noisy density matrix with purity per layer.
Each is run gate-at-a-time and in the engine's default (fused) mode; results of the two modes must agree exactly."""
import argparse
import ctypes
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from quest_b200 import capi, circuits  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qft-qubits", type=int, default=30)
    ap.add_argument("--dm-qubits", type=int, default=15)
    ap.add_argument("--layers", type=int, default=10)
    ap.add_argument("--out", default="gpurun_out/config_bench.json")
    args = ap.parse_args()
    lib = capi.load_product()
    env = lib.createQuESTEnv()
    lib.lib.qb200_setFusion.argtypes = [ctypes.c_int]
    lib.lib.qb200_getLaunchCount.restype = ctypes.c_ulonglong
    out = {}

    def timed(make_qureg, prep, ops, label):
        res = {}
        for mode, name in ((0, "gate_at_a_time"), (1, "fused")):
            lib.lib.qb200_setFusion(mode)
            q = make_qureg()
            prep(q)
            circuits.run_circuit(lib, q, ops[: min(len(ops), 40)])  # warm-up
            prep(q)
            lib.syncQuESTEnv(env)
            l0 = lib.lib.qb200_getLaunchCount()
            t0 = time.perf_counter()
            scal = circuits.run_circuit(lib, q, ops)
            lib.syncQuESTEnv(env)
            dt = time.perf_counter() - t0
            res[name] = {"seconds": dt, "api_calls": len(ops), "calls_per_s": len(ops) / dt,
                         "kernel_launches": int(lib.lib.qb200_getLaunchCount() - l0), "scalars": scal}
            lib.destroyQureg(q, env)
        a, b = np.array(res["gate_at_a_time"]["scalars"]), np.array(res["fused"]["scalars"])
        res["modes_agree_exactly"] = bool(np.array_equal(a, b))
        res["speedup_fused"] = res["gate_at_a_time"]["seconds"] / res["fused"]["seconds"]
        out[label] = res
        print(label, {k: (v if not isinstance(v, dict) else {kk: vv for kk, vv in v.items() if kk != "scalars"}) for k, v in res.items()}, flush=True)
        return res

    nq = args.qft_qubits
    r = timed(lambda: lib.createQureg(nq, env), lambda q: lib.initClassicalState(q, 0x2AAAAAAA & ((1 << nq) - 1)),
              circuits.config3_qft(nq), f"config3_qft_{nq}q")
    probs = np.array(r["fused"]["scalars"])
    out[f"config3_qft_{nq}q"]["max_abs_dev_from_half"] = float(np.abs(probs - 0.5).max())
    assert np.abs(probs - 0.5).max() < 1e-12

    nd = args.dm_qubits
    r = timed(lambda: lib.createDensityQureg(nd, env), lambda q: lib.initPlusState(q),
              circuits.config4_noisy(nd, args.layers), f"config4_noisy_density_{nd}q")
    out[f"config4_noisy_density_{nd}q"]["purity_per_layer"] = r["fused"]["scalars"]
    for v in out.values():
        for m in ("gate_at_a_time", "fused"):
            v[m].pop("scalars", None)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
