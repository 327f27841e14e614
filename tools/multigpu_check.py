#!/usr/bin/env python3
"""Multi-GPU parity check, run under torchrun on N GPUs of one box:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/multigpu_check.py [--qubits 22]

Every rank drives its shard through the CUDA library; rank 0 gathers the shards and compares with the CPU oracle
running the same circuit on one unsharded register: measurement-free circuits must match bit for bit, circuits with
measurements must give the same outcomes and amplitudes to 1e-12 relative (north_star)."""
import argparse
import ctypes
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from quest_b200 import capi, circuits  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qubits", type=int, default=22)
    ap.add_argument("--depth", type=int, default=300)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    lib = capi.load_product()
    env = lib.createQuESTEnv()
    assert (env.rank, env.numRanks) == (rank, world)
    torch.cuda.set_device(lib.lib.qb200_getDevice())
    dist.init_process_group("gloo")
    cpu = capi.load(capi.ORACLE_LIB) if rank == 0 else None
    cenv = cpu.createQuESTEnv() if rank == 0 else None
    ok = True
    for case, (nq, measure, seed) in enumerate([(args.qubits, False, 301), (args.qubits, True, 302), (args.qubits - 5, False, 303),
                                                 (args.qubits - 3, True, 304)]):
        ops = circuits.random_circuit(nq, args.depth, seed, measure=measure)
        # bias: add gates on the shard-selecting qubits so the exchange path is hit hard
        g = int(np.log2(world))
        extra = []
        for t in range(nq - g, nq):
            extra += [("hadamard", t), ("compactUnitary", t, circuits.SWEEP_ALPHA, circuits.SWEEP_BETA), ("controlledNot", 0, t),
                      ("controlledNot", t, 1), ("pauliY", t), ("multiControlledUnitary", [2, (t - 1)], t, circuits._unitary(__import__("random").Random(t))),
                      ("calcProbOfOutcome", t, 1), ("phaseShift", t, 0.3), ("controlledPhaseShift", t, 3, 0.2)]
        # every qubit's probability, asked twice: the second query on an unchanged state fills the all-qubit cache
        # (one pass + shard-selecting bits + allreduce), the rest are answered from it - in the remapped layout
        probes = [("calcProbOfOutcome", t, o) for o in (0, 1) for t in range(nq)]
        ops = extra + ops[:len(ops) // 2] + probes + ops[len(ops) // 2:] + extra + probes
        if measure:  # a measurement whose qubit selects the shard: whole shards are kept or cleared
            ops += [("measure", nq - 1), ("calcTotalProb",), ("measureWithStats", nq - g), ("calcTotalProb",)]
        q = lib.createQureg(nq, env)
        assert q.numChunks == world and q.chunkId == rank and q.numAmpsPerChunk * world == q.numAmpsTotal
        full = circuits.random_state(1 << nq, seed)
        lib.set_state(q, full)  # every rank passes the full vector; each keeps its slice (reference setAmps semantics)
        scal = circuits.run_circuit(lib, q, ops, seed=seed)
        amp = lib.getAmp(q, (1 << nq) - 3)
        shard = lib.get_state(q)
        lib.destroyQureg(q, env)
        gathered = [torch.empty(2 * shard.size, dtype=torch.float64) for _ in range(world)] if rank == 0 else None
        dist.gather(torch.from_numpy(shard.view(np.float64).copy()), gathered, dst=0)
        if rank == 0:
            got = np.concatenate([t.numpy().view(np.complex128) for t in gathered])
            cq = cpu.createQureg(nq, cenv)
            cpu.set_state(cq, full)
            cscal = circuits.run_circuit(cpu, cq, ops, seed=seed)
            want = cpu.get_state(cq)
            camp = cpu.getAmp(cq, (1 << nq) - 3)
            cpu.destroyQureg(cq, cenv)
            same_bits = np.array_equal(got.view(np.float64), want.view(np.float64))
            err = np.abs(got - want).max() / np.abs(want).max()
            scal_ok = np.allclose(scal, cscal, rtol=1e-12, atol=1e-13)
            amp_ok = abs(complex(amp.real, amp.imag) - complex(camp.real, camp.imag)) <= 1e-12
            good = scal_ok and amp_ok and (same_bits if not measure else err < 1e-12)
            ok = ok and good
            print(f"case {case}: {nq} qubits over {world} GPUs, {len(ops)} ops, measure={measure}: bit-exact={same_bits} "
                  f"rel.err={err:.2e} scalars_ok={scal_ok} getAmp_ok={amp_ok} -> {'PASS' if good else 'FAIL'}", flush=True)
    # sharded density matrices: unitaries on row and column qubits, one-qubit noise channels (also on qubits whose
    # column bit selects the shard), purity and outcome probabilities
    for case, (nd, seed, meas) in enumerate([(11, 311, False), (9, 312, True)]):
        ops = circuits.random_circuit(nd, args.depth, seed, density=True, measure=meas, two_qubit_noise=True)
        ops += [(name, t, 0.05 + 0.01 * t) for t in range(nd) for name in circuits.NOISE] + [("calcPurity",), ("calcTotalProb",)]
        q = lib.createDensityQureg(nd, env)
        circuits.prepare(lib, q, seed)
        scal = circuits.run_circuit(lib, q, ops, seed=seed)
        shard = lib.get_state(q)
        lib.destroyQureg(q, env)
        gathered = [torch.empty(2 * shard.size, dtype=torch.float64) for _ in range(world)] if rank == 0 else None
        dist.gather(torch.from_numpy(shard.view(np.float64).copy()), gathered, dst=0)
        if rank == 0:
            got = np.concatenate([t.numpy().view(np.complex128) for t in gathered])
            cq = cpu.createDensityQureg(nd, cenv)
            circuits.prepare(cpu, cq, seed)
            cscal = circuits.run_circuit(cpu, cq, ops, seed=seed)
            want = cpu.get_state(cq)
            cpu.destroyQureg(cq, cenv)
            same_bits = np.array_equal(got.view(np.float64), want.view(np.float64))
            scal_ok = np.allclose(scal, cscal, rtol=1e-12, atol=1e-13)
            err = np.abs(got - want).max() / np.abs(want).max()
            good = scal_ok and (same_bits if not meas else err < 1e-12)
            ok = ok and good
            print(f"density case {case}: {nd}-qubit density matrix over {world} GPUs, {len(ops)} ops: bit-exact={same_bits} "
                  f"scalars_ok={scal_ok} -> {'PASS' if good else 'FAIL'}", flush=True)
    # multi-register API fuzz: copies and observers interleaved with gates on sharded registers
    for case, (nq, seed) in enumerate([(9, 321), (12, 322)]):
        calls = circuits.random_api_sequence(nq, 300, seed, extended=True)
        scal, states = circuits.run_api_sequence(lib, env, nq, calls, seed)
        parts = {}
        for name in sorted(states):
            shard = states[name]
            gathered = [torch.empty(2 * shard.size, dtype=torch.float64) for _ in range(world)] if rank == 0 else None
            dist.gather(torch.from_numpy(shard.view(np.float64).copy()), gathered, dst=0)
            if rank == 0:
                parts[name] = np.concatenate([t.numpy().view(np.complex128) for t in gathered])
        if rank == 0:
            cscal, cstates = circuits.run_api_sequence(cpu, cenv, nq, calls, seed)
            # the CPU checker sums up to 4^nq terms sequentially (one OpenMP thread under torchrun): its own rounding
            # error bound is ~N*eps, far above the GPU's tree reduction; compare scalars at that bound
            tol = max(1e-12, 4.0 ** nq * 4.5e-16)
            scal_ok = np.allclose(scal, cscal, rtol=tol, atol=1e-13)
            st_ok = all(np.allclose(parts[k], cstates[k], rtol=0, atol=1e-13 * max(1.0, np.abs(cstates[k]).max())) for k in parts)
            good = scal_ok and st_ok
            ok = ok and good
            if not scal_ok:
                bad = np.nonzero(~np.isclose(scal, cscal, rtol=tol, atol=1e-13))[0]
                for i in bad[:6]:
                    print("   scalar", i, circuits.run_api_sequence.last_labels[i], "gpu", scal[i], "cpu", cscal[i], flush=True)
            print(f"api fuzz case {case}: {nq} qubits, {len(calls)} calls over 4 registers on {world} GPUs: scalars_ok={scal_ok} "
                  f"states_ok={st_ok} -> {'PASS' if good else 'FAIL'}", flush=True)
    lib.lib.qb200_getOverlappedExchangeCount.restype = ctypes.c_ulonglong
    lib.lib.qb200_getSlotExchangeCount.restype = ctypes.c_ulonglong
    if rank == 0:
        print(f"slot exchanges: {lib.lib.qb200_getSlotExchangeCount()}, of which pipelined with the preceding fused pass: "
              f"{lib.lib.qb200_getOverlappedExchangeCount()}", flush=True)
    flag = torch.tensor([1 if ok else 0])
    dist.broadcast(flag, 0)
    lib.destroyQuESTEnv(env)
    dist.destroy_process_group()
    if rank == 0:
        print("MULTIGPU PARITY", "PASS" if ok else "FAIL", flush=True)
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
