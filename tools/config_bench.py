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
    ap.add_argument("--config1-qubits", type=int, default=24)
    ap.add_argument("--skip-config1", action="store_true")
    ap.add_argument("--out", default="gpurun_out/config_bench.json")
    args = ap.parse_args()
    lib = capi.load_product()
    env = lib.createQuESTEnv()
    lib.lib.qb200_setFusion.argtypes = [ctypes.c_int]
    lib.lib.qb200_setReordering.argtypes = [ctypes.c_int]
    lib.lib.qb200_getLaunchCount.restype = ctypes.c_ulonglong
    out = {}

    def timed(make_qureg, prep, ops, label):
        res = {}
        for mode, name in ((0, "gate_at_a_time"), (1, "fused"), (2, "fused_reordering")):
            lib.lib.qb200_setFusion(1 if mode else 0)
            lib.lib.qb200_setReordering(1 if mode == 2 else 0)
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
        c = np.array(res["fused_reordering"]["scalars"])
        res["reordering_max_rel_diff_of_scalars"] = float(np.max(np.abs(c - b) / np.maximum(np.abs(b), 1e-300))) if b.size else 0.0
        res["speedup_reordering"] = res["gate_at_a_time"]["seconds"] / res["fused_reordering"]["seconds"]
        lib.lib.qb200_setReordering(0)
        out[label] = res
        print(label, {k: (v if not isinstance(v, dict) else {kk: vv for kk, vv in v.items() if kk != "scalars"}) for k, v in res.items()}, flush=True)
        return res

    # configs[0]: 24-qubit layered circuit (hadamard / rotateY / controlledNot on every qubit, 20 layers = 1440 gates):
    # the reference's CPU/OpenMP build on this box's cores next to the GPU library, then every amplitude compared
    if os.path.isfile(capi.REFERENCE_LIB) and not args.skip_config1:
        n1 = args.config1_qubits
        ops1 = circuits.config1_layers(n1, 20)
        cpu = capi.load(capi.REFERENCE_LIB)
        cenv = cpu.createQuESTEnv()
        cq = cpu.createQureg(n1, cenv)
        t0 = time.perf_counter()
        circuits.run_circuit(cpu, cq, ops1)
        cpu_s = time.perf_counter() - t0
        want = cpu.get_state(cq)
        cpu.destroyQureg(cq, cenv)
        lib.lib.qb200_setFusion(1)
        gq = lib.createQureg(n1, env)
        circuits.run_circuit(lib, gq, ops1[:72])
        lib.initZeroState(gq)
        lib.syncQuESTEnv(env)
        t0 = time.perf_counter()
        circuits.run_circuit(lib, gq, ops1)
        lib.syncQuESTEnv(env)
        gpu_s = time.perf_counter() - t0
        got = lib.get_state(gq)
        # the same with opt-in reordering (qb200_setReordering): layers run qubit block by qubit block
        lib.lib.qb200_setReordering(1)
        lib.initZeroState(gq)
        lib.syncQuESTEnv(env)
        t0 = time.perf_counter()
        circuits.run_circuit(lib, gq, ops1)
        lib.syncQuESTEnv(env)
        gpu_reordered_s = time.perf_counter() - t0
        got_reordered = lib.get_state(gq)
        lib.lib.qb200_setReordering(0)
        lib.destroyQureg(gq, env)
        out[f"config1_layers_{n1}q"] = {
            "gpu_seconds_with_reordering": gpu_reordered_s, "gpu_gates_per_s_with_reordering": len(ops1) / gpu_reordered_s,
            "reordering_max_abs_diff": float(np.abs(got_reordered - want).max()),
            "reordering_max_rel_diff": float(np.abs(got_reordered - want).max() / np.abs(want).max()),
            "gates": len(ops1), "cpu_reference_seconds": cpu_s, "cpu_gates_per_s": len(ops1) / cpu_s,
            "cpu_threads": int(os.environ.get("OMP_NUM_THREADS", os.cpu_count())), "gpu_seconds": gpu_s,
            "gpu_gates_per_s": len(ops1) / gpu_s, "speedup": cpu_s / gpu_s,
            "all_amplitudes_bit_identical": bool(np.array_equal(got.view(np.float64), want.view(np.float64))),
            "max_abs_diff": float(np.abs(got - want).max())}
        print(f"config1_layers_{n1}q", out[f"config1_layers_{n1}q"], flush=True)

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
        for m in ("gate_at_a_time", "fused", "fused_reordering"):
            if m in v:
                v[m].pop("scalars", None)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
