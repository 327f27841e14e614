#!/usr/bin/env python3
"""bench.py — gate throughput of the quest-b200 engine on BASELINE.json's headline workload.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N = 1 : configs[1] — 30-qubit double-precision state-vector, one step = the single-gate sweep (every target
        0..29 for hadamard, compactUnitary, controlledNot, multiControlledPhaseShift = 120 gate calls).
N > 1 : one process per GPU (torchrun), the state sharded by its top log2(N) qubits; weak scaling of the same
        sweep: each GPU keeps a 2^30-amplitude shard, the register has 30+log2 N qubits and one step sweeps every
        target again, so log2 N of the targets select the GPU and run as fused NVLink exchange+combine gates.
        value = shard-gates/s summed over GPUs (gates/s x N).  For N >= 2 the line also carries
        "circuit_34q": seconds for the configs[4] 34-qubit random circuit with measurements.

The JSON line carries: value (inputs resident in HBM, CUDA events on the engine's stream, max over ranks), e2e
(the same through the public C API with the state uploaded from pinned host memory and a scalar read back every
step), roofline (dominant kernel: algorithmic bytes / measured launch time against MEASURED_PEAKS.json), and
cpu_baseline (the reference's own CPU/OpenMP library, oracle/_ref, timed on this box's host cores on a bounded
sample of the same sweep).  ``--impl reference`` runs only that CPU library and prints the same line shape.
Nothing here reads /root/reference.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from quest_b200 import capi, circuits  # noqa: E402

def target_of(op):
    """Target qubit of a circuit op (controlledNot carries its control first)."""
    name = op[0]
    if name in ("controlledNot", "controlledPauliY") or name.startswith("controlledRotate") or name in ("controlledCompactUnitary", "controlledUnitary"):
        return op[2]
    if name.startswith("multiControlled"):
        return op[2] if name == "multiControlledUnitary" else -1
    return op[1] if len(op) > 1 and isinstance(op[1], int) else -1


def read_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons every 200 ms while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.proc, self.lines = device, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.device)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(names, parts[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------------------
# the reference's own CPU library: reference arm, cpu_baseline and the parity leg
# ------------------------------------------------------------------------------------------------------------

METRIC = "gates/sec (30q double state-vector single-gate sweep)"
CLASSES = ("hadamard", "compactUnitary", "controlledNot", "multiControlledPhaseShift")


def sweep_config(nq, world, shard_bits, n_ops):
    """`config` of the JSON line: the same dict on both arms (the driver compares them)."""
    if world == 1:
        workload = (f"configs[1]: {nq}-qubit state-vector single-gate sweep, every target, 4 gate classes "
                    f"({n_ops} gates/step), state = initPlusState + rotateY(t, 0.1(t+1)) on every qubit")
    else:
        workload = (f"configs[1] sweep weak-scaled: {nq}-qubit state-vector sharded over {world} GPUs (2^{shard_bits} "
                    f"amplitudes per GPU), every target, 4 gate classes ({n_ops} gates/step, "
                    f"{4 * int(math.log2(world))} of them on shard-selecting qubits); value counts one gate call on the "
                    f"sharded register as {world} shard-gates (gates/s x {world})")
    return {"workload": workload, "qubits": nq, "amps_per_gpu": 1 << shard_bits, "gates_per_step": n_ops,
            "l2": "state (16 GiB per GPU) is far larger than L2: every pass streams from HBM"}


def sweep_prepare(lib, q, nq):
    """The dense, non-trivial input state of the sweep - the same calls on every arm."""
    lib.initPlusState(q)
    for t in range(nq):
        lib.rotateY(q, t, 0.1 * (t + 1))


def use_all_host_cores():
    # torchrun pins OMP_NUM_THREADS=1 for its workers; the CPU arm is meant to use every host core
    os.environ["OMP_NUM_THREADS"] = os.environ.get("QUEST_BENCH_THREADS", str(os.cpu_count() or 1))
    return int(os.environ["OMP_NUM_THREADS"])


def load_cpu_library():
    if os.path.isfile(capi.REFERENCE_LIB):
        return capi.load(capi.REFERENCE_LIB), "reference"
    if not os.path.isfile(capi.ORACLE_LIB):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, stdout=subprocess.DEVNULL)
    return capi.load(capi.ORACLE_LIB), "port"


def sampled_step_ops(nq, step):
    """The reference arm's bounded step: every target once, the gate class rotating with target and step, so that four
    consecutive steps are exactly one full sweep (120 gates at 30 qubits) and every step holds all four classes."""
    full = circuits.config2_sweep(nq)
    return [full[4 * t + (t + step) % 4] for t in range(nq)]


def run_reference_arm(args, rank, world, emit):
    if rank != 0:
        return
    cores = use_all_host_cores()
    nq = args.qubits  # at N > 1 too: a gate on the (nq + log2 N)-qubit register is N shard-gates of 2^nq amplitudes
    lib, kind = load_cpu_library()
    env = lib.createQuESTEnv()
    q = lib.createQureg(nq, env)
    sweep_prepare(lib, q, nq)
    warmup = max(args.warmup, 0)
    for w in range(warmup):
        circuits.run_circuit(lib, q, sampled_step_ops(nq, w))
    n_gates = 0
    t0 = time.perf_counter()
    for s in range(args.steps):
        ops = sampled_step_ops(nq, warmup + s)
        circuits.run_circuit(lib, q, ops)
        n_gates += len(ops)
    dt = time.perf_counter() - t0
    prob = lib.calcTotalProb(q)
    lib.destroyQureg(q, env)
    assert abs(prob - 1.0) < 1e-9, prob
    value = n_gates / dt
    n_full = len(circuits.config2_sweep(nq + int(math.log2(world))))
    sample = (f"{nq} gates per step ({args.steps} steps, {warmup} warm-up): every target of the configs[1] sweep once, the "
              f"gate class (hadamard / compactUnitary / controlledNot / multiControlledPhaseShift) rotating with target and "
              f"step so that 4 steps = one full 120-gate sweep; same state preparation as the quest_b200 arm; wall clock "
              f"{dt:.1f} s on {cores} threads"
              + ("" if world == 1 else f"; timed at the shard size ({nq} qubits), where one gate is one shard-gate"))
    line = {
        "impl": "reference", "metric": METRIC, "value": value,
        "unit": "gates/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": sweep_config(nq + int(math.log2(world)), world, nq, n_full),
        "cpu_baseline": {"value": value, "unit": "gates/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "gates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def parity_leg(lib, env, rank, world, nq, torch, dist):
    """Outside every timed region: ONE configs[1] sweep step (state preparation + every gate of the sweep) on a fresh
    nq-qubit register through the product, and the same calls through the reference's own CPU library (oracle/_ref) on
    rank 0; every amplitude is compared, then three measurements and a collapseToOutcome under the same seed, then
    every amplitude again.  Returns (parity dict, cpu_baseline dict) on rank 0, (None, None) elsewhere."""
    import numpy as np

    ops = circuits.config2_sweep(nq)
    tail = [("measure", 0), ("measure", nq // 2), ("measure", nq - 1), ("collapseToOutcome", 1, 0), ("calcTotalProb",)]
    q = lib.createQureg(nq, env)
    n_local = int(q.numAmpsPerChunk)
    sweep_prepare(lib, q, nq)
    circuits.run_circuit(lib, q, ops)

    cpu = cq = cenv = None
    cpu_info = None
    if rank == 0:
        cores = use_all_host_cores()
        cpu, kind = load_cpu_library()
        cenv = cpu.createQuESTEnv()
        cq = cpu.createQureg(nq, cenv)
        sweep_prepare(cpu, cq, nq)
        t0 = time.perf_counter()
        circuits.run_circuit(cpu, cq, ops)
        cpu_dt = time.perf_counter() - t0
        cpu_info = {"value": len(ops) / cpu_dt, "unit": "gates/s", "cores": cores, "kind": kind,
                    "sample": f"one full configs[1] sweep step ({len(ops)} gates) at {nq} qubits on the state the parity "
                              f"check compares, wall clock {cpu_dt:.1f} s"}

    chunk = 1 << 24

    def compare(scale=1.0):
        """max |got - scale * want| / max |want| over the whole register and whether every double is bit-identical."""
        worst, biggest, same = 0.0, 0.0, True
        if world == 1:
            re = np.empty(min(chunk, n_local), dtype=np.float64)
            im = np.empty_like(re)
            for at in range(0, n_local, chunk):
                n = min(chunk, n_local - at)
                lib.qb200_copyAmpsToHost(q, at, n, lib.real_ptr(re), lib.real_ptr(im))
                for got, want_ptr in ((re, cq.stateVec.real), (im, cq.stateVec.imag)):
                    want = np.ctypeslib.as_array(want_ptr, shape=(n_local,))[at:at + n]
                    same = same and np.array_equal(got[:n], want)
                    worst = max(worst, float(np.abs(got[:n] - scale * want).max()))
                    biggest = max(biggest, float(np.abs(want).max()))
            return worst / biggest, same
        shard = lib.get_state(q)  # this rank's shard, identity layout restored
        mine = torch.from_numpy(shard.view(np.float64).copy()).cuda()
        gathered = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
        dist.gather(mine, gathered, dst=0)
        if rank != 0:
            return None, None
        want_re = np.ctypeslib.as_array(cq.stateVec.real, shape=(n_local * world,))
        want_im = np.ctypeslib.as_array(cq.stateVec.imag, shape=(n_local * world,))
        for r, t in enumerate(gathered):
            got = t.cpu().numpy().view(np.complex128)
            sl = slice(r * n_local, (r + 1) * n_local)
            same = same and np.array_equal(got.real, want_re[sl]) and np.array_equal(got.imag, want_im[sl])
            worst = max(worst, float(np.abs(got.real - scale * want_re[sl]).max()), float(np.abs(got.imag - scale * want_im[sl]).max()))
            biggest = max(biggest, float(np.abs(want_re[sl]).max()), float(np.abs(want_im[sl]).max()))
        return worst / biggest, same

    rel1, same1 = compare()
    got_tail = circuits.run_circuit(lib, q, tail, seed=77)
    want_tail = circuits.run_circuit(cpu, cq, tail, seed=77) if rank == 0 else None
    # After a collapse each library has renormalised with ITS OWN outcome probability.  The reference adds |amp|^2
    # sequentially per OpenMP thread (QuEST_cpu.c:2880-2887), which at 2^28+ amplitudes leaves it up to ~1e-10 off, and
    # four collapses compound that.  Both final states are therefore the same direction with slightly different norms,
    # and the norms are known accurately (the product's device tree sum; the reference's Kahan sum,
    # QuEST_cpu_local.c:123-151).  The bar after the measurements: identical outcomes; the product's own norm 1 to 1e-12;
    # amplitudes within 1e-12 once the reference state is rescaled by sqrt(total_quest_b200 / total_cpu); and that
    # factor, like the outcome probability, within the 2e-9 bound on the reference's sequential sums.
    scale = 1.0
    if world > 1:
        t = torch.tensor([want_tail[-1] if rank == 0 else 0.0], dtype=torch.float64, device="cuda")
        dist.broadcast(t, 0)
        want_total = float(t.item())
    else:
        want_total = want_tail[-1]
    scale = math.sqrt(got_tail[-1] / want_total)
    rel2, same2 = compare(scale)
    lib.destroyQureg(q, env)
    if rank != 0:
        return None, None
    cpu.destroyQureg(cq, cenv)
    outcomes_equal = [int(x) for x in got_tail[:3]] == [int(x) for x in want_tail[:3]]
    scalars_rel = max(abs(a - b) / max(abs(b), 1e-300) for a, b in zip(got_tail[3:], want_tail[3:]))
    ref_bound = 2e-9
    ok = (rel1 <= 1e-12 and rel2 <= 1e-12 and outcomes_equal and abs(got_tail[-1] - 1.0) <= 1e-12
          and abs(scale - 1.0) <= ref_bound and scalars_rel <= ref_bound)
    parity = {"qubits": nq, "n_gpus": world, "against": "oracle/_ref (the reference's own CPU/OpenMP build)" if kind == "reference"
              else "oracle/ (CPU restatement)", "gates": nq + len(ops), "amplitudes_compared": n_local * world,
              "max_rel_diff": rel1, "bit_identical": bool(same1),
              "measurement_outcomes": [int(x) for x in got_tail[:3]], "outcomes_equal": outcomes_equal,
              "collapse_prob_and_total_prob_rel_diff": scalars_rel,
              "max_rel_diff_after_measurements": rel2, "bit_identical_after_measurements": bool(same2),
              "tolerance": 1e-12, "reference_rescaled_by": scale, "bound_on_reference_sequential_sums": ref_bound,
              "total_prob_after": {"quest_b200": got_tail[-1], "cpu": want_tail[-1]}, "ok": bool(ok),
              "what": "initPlusState + rotateY on every qubit + one full configs[1] sweep step, every amplitude compared "
                      "(max |diff| / max |amp|); then measure x3 + collapseToOutcome under seedQuEST({77,155}) and every "
                      "amplitude again, the reference state rescaled by sqrt(total_prob quest_b200 / cpu): each library "
                      "renormalises with its own outcome probabilities and the reference's sequential sums are only good "
                      "to ~1e-10 at this size, the product's norm stays 1 to 1e-15"}
    return parity, cpu_info


# ------------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------------

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="quest_b200", choices=["quest_b200", "reference"])
    ap.add_argument("--qubits", type=int, default=30, help="qubits per 1-GPU shard (default 30 = BASELINE config)")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity leg (and with it cpu_baseline)")
    ap.add_argument("--parity-qubits", type=int, default=0, help="register size of the parity leg (default: 30 at N=1, 28 sharded at N>1)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-34q", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # keep stdout to the one JSON line: NCCL writes its version banner / debug lines to stdout unless told otherwise
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    # stdout carries the ONE JSON line and nothing else: whatever libraries print (NCCL banners, C-level printf) goes to
    # stderr for the duration of the run
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())

    if args.impl == "reference":
        run_reference_arm(args, rank, world, emit)
        return

    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the quest-b200 arm has no CPU fallback")
    warmup = max(args.warmup, 3)
    lib = capi.load_product()
    lib.lib.qb200_getStream.restype = ctypes.c_void_p
    lib.lib.qb200_getLaunchCount.restype = ctypes.c_ulonglong
    env = lib.createQuESTEnv()
    device = lib.lib.qb200_getDevice()
    torch.cuda.set_device(device)
    stream = torch.cuda.ExternalStream(lib.lib.qb200_getStream(), device=device)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", device))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    shard_bits = args.qubits
    nq = shard_bits + int(math.log2(world))
    ops = circuits.config2_sweep(nq)
    q = lib.createQureg(nq, env)
    n_local = int(q.numAmpsPerChunk)
    sweep_prepare(lib, q, nq)  # a dense, non-trivial state: one rotation per qubit (same calls on the CPU arm)

    lib.lib.qb200_setFusion.argtypes = [ctypes.c_int]
    lib.lib.qb200_setKernelTiming.argtypes = [ctypes.c_int]

    def step():
        for op in ops:
            circuits.apply_op(lib, q, op)
        lib.lib.qb200_flush()  # recorded gates are launched here (no host wait)

    peak, peak_src = read_peaks()

    # ---- pass 1: one kernel per gate call (fusion off) -> per-gate-class device times, single-gate roofline -------
    lib.lib.qb200_setFusion(0)
    step()
    barrier()
    marks = []
    per_gate_steps = 1 if args.steps < 4 else 2
    for _ in range(per_gate_steps):
        for op in ops:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(stream)
            marks.append((op[0], target_of(op), ev))
            circuits.apply_op(lib, q, op)
    end1 = torch.cuda.Event(enable_timing=True)
    end1.record(stream)
    barrier()
    unfused_ms = marks[0][2].elapsed_time(end1) / per_gate_steps
    evs = [m[2] for m in marks] + [end1]
    durs = {}
    cross = {}  # gates whose target qubit selects the GPU: exchange fused into the gate kernel over NVLink
    for i, (name, tgt, _) in enumerate(marks):
        d = evs[i].elapsed_time(evs[i + 1])
        if world > 1 and tgt >= shard_bits:
            cross.setdefault(name, []).append(d)
        else:
            durs.setdefault(name, []).append(d)
    dense = [d for name in ("hadamard", "compactUnitary", "rotateY") for d in durs.get(name, [])]
    dense_ms = sum(dense) / max(len(dense), 1)
    achieved1 = 32.0 * n_local / (dense_ms * 1e-3) / 1e9 if dense else None
    traffic = None
    prof = os.path.join(ROOT, "profiles", "dominant_kernel.json")
    if os.path.isfile(prof):
        traffic = json.load(open(prof)).get("dram_bytes_per_launch")
    single_gate = {"kernel": "qb::k_pairs (one dense one-target gate per launch: hadamard/compactUnitary)",
                   "achieved": achieved1, "peak": peak, "unit": "GB/s", "frac": achieved1 / peak if achieved1 else None,
                   "traffic": traffic, "algorithmic_bytes_per_launch": 32.0 * n_local, "launches_timed": len(dense),
                   "ms_per_launch": dense_ms, "per_class_ms": {k: sum(v) / len(v) for k, v in durs.items()},
                   "gates_per_s_gate_at_a_time": world * len(ops) / (unfused_ms * 1e-3)}

    # ---- pass 2: the engine's default mode (gate fusion on) -> value --------------------------------------------
    lib.lib.qb200_setFusion(1)
    for _ in range(warmup):
        step()
    barrier()
    lib.lib.qb200_getSlotExchangeCount.restype = ctypes.c_ulonglong
    lib.lib.qb200_getOverlappedExchangeCount.restype = ctypes.c_ulonglong
    x0, xo0 = lib.lib.qb200_getSlotExchangeCount(), lib.lib.qb200_getOverlappedExchangeCount()
    sampler = ClockSampler(device)
    sampler.start()
    lib.lib.qb200_setKernelTiming(1)
    launches0 = lib.lib.qb200_getLaunchCount()
    fl0, fg0 = ctypes.c_ulonglong(), ctypes.c_ulonglong()
    lib.lib.qb200_getFusionStats(ctypes.byref(fl0), ctypes.byref(fg0))
    barrier()
    start = torch.cuda.Event(enable_timing=True)
    end = torch.cuda.Event(enable_timing=True)
    start.record(stream)
    for _ in range(args.steps):
        step()
    end.record(stream)
    barrier()
    total_ms = start.elapsed_time(end)
    launches = lib.lib.qb200_getLaunchCount() - launches0
    fused_ms, fused_n = ctypes.c_double(), ctypes.c_ulonglong()
    lib.lib.qb200_getFusedKernelTime(ctypes.byref(fused_ms), ctypes.byref(fused_n))
    lib.lib.qb200_setKernelTiming(0)
    fl1, fg1 = ctypes.c_ulonglong(), ctypes.c_ulonglong()
    lib.lib.qb200_getFusionStats(ctypes.byref(fl1), ctypes.byref(fg1))
    clocks = sampler.stop()
    exchanges = {"exchanges_per_step": (lib.lib.qb200_getSlotExchangeCount() - x0) / args.steps,
                 "pipelined_per_step": (lib.lib.qb200_getOverlappedExchangeCount() - xo0) / args.steps}
    if world > 1:
        t = torch.tensor([total_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    gates = args.steps * len(ops)
    value = world * gates / (total_ms * 1e-3)

    # roofline of the dominant kernel of the timed region: the fused pass (qb::k_tile).  One launch reads and
    # writes the whole shard once (32*N algorithmic bytes) whatever the number of gates it carries.
    n_fused = int(fused_n.value)
    fused_avg = fused_ms.value / n_fused if n_fused else None
    achieved = 32.0 * n_local / (fused_avg * 1e-3) / 1e9 if fused_avg else None
    passes_per_step = n_fused / args.steps if args.steps else None
    # fp64 work of one step, exact arithmetic (no FMA): per amplitude, hadamard 4 + compactUnitary 14 + 3-qubit phase 6/8
    # instructions for each of the nq targets; the pipe issues 64 fp64 lanes per SM per clock
    fp64_per_amp = (4 + 14 + 0.75) * nq
    sm_mhz = None
    roofline = {"bound": "hbm", "kernel": "qb::k_tile (one HBM pass applying a run of recorded gates on a shared-memory tile)",
                "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                "frac": achieved / peak if achieved else None, "traffic": None,
                "algorithmic_bytes_per_launch": 32.0 * n_local, "launches_timed": n_fused, "ms_per_launch": fused_avg,
                "passes_per_step": passes_per_step,
                "gates_per_launch": (fg1.value - fg0.value) / max(fl1.value - fl0.value, 1),
                "share_of_step": fused_ms.value / total_ms if total_ms else None,
                "note": "a pass carries ~30-44 gates (7-11 target qubits), so it is bound by the fp64 pipe, not by HBM: "
                        "frac < 1 against the HBM peak is by construction (fewer passes = less HBM time per gate); see fp64 "
                        "below for the binding pipe and single_gate_kernel for the pure-HBM case",
                "fp64": {"instr_per_amplitude_per_step": fp64_per_amp,
                         "achieved_Tinstr_per_s": fp64_per_amp * n_local / (total_ms / args.steps * 1e-3) / 1e12,
                         "peak_Tinstr_per_s_at_max_clock": 64 * 148 * 1.965e9 / 1e12,
                         "what": "DADD/DMUL lane-operations per second over the whole step (exact arithmetic, no FMA) "
                                 "against 64 lanes x 148 SMs x 1.965 GHz"},
                "single_gate_kernel": single_gate}
    roofline["fp64"]["frac_of_max_clock_peak"] = roofline["fp64"]["achieved_Tinstr_per_s"] / roofline["fp64"]["peak_Tinstr_per_s_at_max_clock"]
    roofline["fp64"]["peak_Tinstr_per_s_measured"] = 18.2  # tools/fp64_peak.cu on this pool: 62.6 lanes/clk/SM, data-independent
    roofline["fp64"]["frac_of_measured_peak"] = roofline["fp64"]["achieved_Tinstr_per_s"] / 18.2
    # the same step by SURVEY 8(d)'s per-gate algorithmic bytes (what one kernel per gate has to move): per target
    # hadamard 32N + compactUnitary 32N + controlledNot 16N + 3-qubit phase 4N
    gate_bytes = (32 + 32 + 16 + 4) * float(n_local) * nq
    roofline["gate_equivalent"] = {"algorithmic_bytes_per_step_one_kernel_per_gate": gate_bytes,
                                   "equivalent_GBps": gate_bytes / (total_ms / args.steps * 1e-3) / 1e9,
                                   "x_hbm_peak": gate_bytes / (total_ms / args.steps * 1e-3) / 1e9 / peak,
                                   "what": "bytes the gates of a step move when each is its own HBM pass, divided by the "
                                           "measured step time: how far fusion has taken the step past the per-gate HBM roofline"}
    prof2 = os.path.join(ROOT, "profiles", "fused_kernel.json")
    if os.path.isfile(prof2):
        roofline["traffic"] = json.load(open(prof2)).get("dram_bytes_per_launch")

    # opt-in contracted arithmetic (fma) of the fused pass: same step, reported beside the default, never as `value`
    contracted = None
    if world == 1 and hasattr(lib.lib, "qb200_setArithmetic"):
        lib.lib.qb200_setArithmetic.argtypes = [ctypes.c_int]
        lib.lib.qb200_setArithmetic(1)
        step()
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(stream)
        for _ in range(3):
            step()
        c1.record(stream)
        barrier()
        lib.lib.qb200_setArithmetic(0)
        cms = c0.elapsed_time(c1) / 3
        contracted = {"value": len(ops) / (cms * 1e-3), "unit": "gates/s", "ms_per_step": cms,
                      "what": "qb200_setArithmetic(1): fused pass with a*b+c contracted into fma; within 1e-12 of the "
                              "default mode but not bit-identical to the reference CPU arithmetic, hence opt-in"}

    # NVLink: the two kernels that move amplitudes between GPUs, timed on the top qubit with CUDA events
    nvlink = None
    if world > 1:
        lib.lib.qb200_setRemap.argtypes = [ctypes.c_int]
        lib.lib.qb200_getSlotExchangeCount.restype = ctypes.c_ulonglong
        top = nq - 1

        def timed_ms(fn, before=None, reps=3):
            out = []
            for _ in range(reps):
                if before:
                    before()
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                fn()
                lib.lib.qb200_flush()
                e1.record(stream)
                barrier()
                out.append(e0.elapsed_time(e1))
            t = torch.tensor([sorted(out)[len(out) // 2]], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())

        lib.lib.qb200_setFusion(0)
        lib.lib.qb200_setRemap(0)
        lib.hadamard(q, top)  # warm: NCCL pair connection
        cross_ms = timed_ms(lambda: lib.hadamard(q, top))
        # with remapping on, a hadamard on the (canonically placed) top qubit = one slot exchange + one local gate;
        # qb200_setRemap(1) puts every qubit back in its own slot before each repetition
        n0 = lib.lib.qb200_getSlotExchangeCount()
        swap_ms = timed_ms(lambda: lib.hadamard(q, top), before=lambda: lib.lib.qb200_setRemap(1))
        exchange_ms = swap_ms - dense_ms
        lib.lib.qb200_setFusion(1)
        per_dir_cross = 16.0 * n_local / (cross_ms * 1e-3) / 1e9
        nvlink = {"cross_shard_gate": {"kernel": "qb::k_pairsPeer (dense gate on a shard-selecting qubit, exchange fused "
                                                 "into the gate over peer memory; remapping off)",
                                       "ms": cross_ms, "bytes_each_way_per_gpu": 16.0 * n_local,
                                       "achieved_GBps_per_direction": per_dir_cross,
                                       "frac_of_measured_peer_copy_770": per_dir_cross / 770.0,
                                       "frac_of_nominal_900": per_dir_cross / 900.0,
                                       "includes": "two stream-ordered pair handshakes (sequence counter in the partner's memory + cuStreamWaitValue32)"},
                  "slot_exchange": {"kernel": "qb::k_swapPeer (makes a shard-selecting qubit local: half a shard each way)",
                                    "bytes_each_way_per_gpu": 8.0 * n_local, "ms": exchange_ms,
                                    "achieved_GBps_per_direction": 8.0 * n_local / (exchange_ms * 1e-3) / 1e9,
                                    "frac_of_measured_peer_copy_770": 8.0 * n_local / (exchange_ms * 1e-3) / 1e9 / 770.0,
                                    "how": "hadamard on the top qubit with remapping on (one exchange + one local dense "
                                           "gate) minus the local dense gate time",
                                    "exchanges_counted": int(lib.lib.qb200_getSlotExchangeCount() - n0)}}

    # e2e: every step uploads the state from pinned host memory through initStateFromAmps, runs the same gates and
    # reads a scalar back through calcTotalProb (host wall clock, max over ranks)
    e2e = None
    if not args.no_e2e:
        re = torch.empty(n_local, dtype=torch.float64, pin_memory=True)
        im = torch.empty(n_local, dtype=torch.float64, pin_memory=True)
        re.fill_(1.0 / math.sqrt(float(q.numAmpsTotal)))
        im.zero_()
        pre = ctypes.cast(re.data_ptr(), ctypes.POINTER(ctypes.c_double))
        pim = ctypes.cast(im.data_ptr(), ctypes.POINTER(ctypes.c_double))

        def e2e_step():
            # each rank uploads its own shard (setAmps keeps the slice that falls into the caller's shard)
            lib.setAmps(q, rank * n_local, pre, pim, n_local)
            step()
            return lib.calcTotalProb(q)

        def timed_e2e():
            e2e_step()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                prob = e2e_step()
            barrier()
            dt = time.perf_counter() - t0
            if world > 1:
                tt = torch.tensor([dt], device="cuda", dtype=torch.float64)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                dt = float(tt.item())
            assert abs(prob - 1.0) < 1e-9, prob
            return dt

        # reference semantics first: setAmps complete on return (the GPU idles during the PCIe transfer) ...
        dt_blocking = timed_e2e()
        # ... then the library's deferred upload (quest_b200.h: qb200_setDeferredUpload): the same setAmps call returns once
        # the copies are enqueued and the first passes follow the upload slab by slab; one GPU only
        lib.lib.qb200_setDeferredUpload.argtypes = [ctypes.c_int]
        lib.lib.qb200_getPassesUnderUploadCount.restype = ctypes.c_ulonglong
        deferred = world == 1
        dt = dt_blocking
        under = 0
        if deferred:
            lib.lib.qb200_setDeferredUpload(1)
            u0 = lib.lib.qb200_getPassesUnderUploadCount()
            dt = timed_e2e()
            under = int(lib.lib.qb200_getPassesUnderUploadCount() - u0)
            lib.lib.qb200_setDeferredUpload(0)
        e2e = {"value": world * gates / dt, "unit": "gates/s", "h2d_bytes_per_step": 16 * n_local * world,
               "d2h_bytes_per_step": 8 * world, "ms_per_step": 1e3 * dt / args.steps,
               "what": "setAmps(shard from pinned host) + the step's gate calls + calcTotalProb per step, host wall "
                       "clock, max over ranks" + ("; qb200_setDeferredUpload(1): the upload is cut into 8 slabs and the "
                       "leading passes follow it slab by slab" if deferred else ""),
               "passes_under_upload_per_step": under / (args.steps + 1) if deferred else 0,
               "blocking_setAmps": {"value": world * gates / dt_blocking, "ms_per_step": 1e3 * dt_blocking / args.steps,
                                    "what": "the same with setAmps complete on return (reference semantics)"}}
        del re, im
    lib.destroyQureg(q, env)

    # configs[4]: the 34-qubit random circuit with measurements, sharded over all GPUs (does not fit one GPU)
    circuit34 = None
    if world > 1 and not args.no_34q and (16 << 34) / world < 150e9:
        q34 = lib.createQureg(34, env)
        lib.initPlusState(q34)
        ops34 = circuits.config5_sharded(34, layers=3, seed=34)
        circuits.run_circuit(lib, q34, ops34[:8])  # warm-up
        barrier()
        t0 = time.perf_counter()
        out34 = circuits.run_circuit(lib, q34, ops34, seed=34)
        barrier()
        dt34 = time.perf_counter() - t0
        tt = torch.tensor([dt34], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        n_global = sum(1 for op in ops34 if target_of(op) >= 34 - int(math.log2(world)))
        circuit34 = {"seconds": float(tt.item()), "gates": len(ops34), "gates_on_shard_selecting_qubits": n_global,
                     "measurement_outcomes": [int(x) for x in out34[:3]], "total_prob_after": out34[-1],
                     "what": "configs[4]: 34-qubit random circuit (hadamard/rotateY/controlledNot/compactUnitary, 30% of "
                             "targets on the top 3 qubits) + 3 measurements, wall clock incl. scalar read-backs"}
        lib.destroyQureg(q34, env)

    # 34 qubits (256 GiB) do not fit one GPU; the one-GPU point of configs[4] is the same circuit recipe at
    # 33 qubits (128 GiB of the 180 GB HBM)
    circuit33 = None
    if world == 1 and not args.no_34q and args.qubits == 30:
        torch.cuda.empty_cache()
        free_b, _ = torch.cuda.mem_get_info()
        if free_b > (16 << 33) + (4 << 30):
            q33 = lib.createQureg(33, env)
            lib.initPlusState(q33)
            ops33 = circuits.config5_sharded(33, layers=3, seed=34)
            circuits.run_circuit(lib, q33, ops33[:8])  # warm-up
            barrier()
            t0 = time.perf_counter()
            out33 = circuits.run_circuit(lib, q33, ops33, seed=34)
            barrier()
            circuit33 = {"seconds": time.perf_counter() - t0, "gates": len(ops33), "qubits": 33,
                         "measurement_outcomes": [int(x) for x in out33[:3]], "total_prob_after": out33[-1],
                         "what": "configs[4] recipe at 33 qubits on one GPU (128 GiB state): random circuit + 3 "
                                 "measurements, wall clock incl. scalar read-backs"}
            lib.destroyQureg(q33, env)
        else:
            circuit33 = {"skipped": f"only {free_b >> 30} GiB of HBM free"}

    # parity at the size the numbers are quoted on (N = 1: 30 qubits; N > 1: 28 qubits sharded over all GPUs so that the
    # CPU can follow), against the reference's own CPU library; its timed sweep step doubles as cpu_baseline
    parity = cpu = None
    if not args.no_parity:
        parity_nq = args.parity_qubits or (nq if world == 1 else 28)
        parity, cpu = parity_leg(lib, env, rank, world, parity_nq, torch, dist if world > 1 else None)
        if world > 1:
            cpu = None  # contract: cpu_baseline on rank 0 at N = 1 only

    if rank == 0:
        line = {
            "metric": METRIC,
            "value": value, "unit": "gates/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": sweep_config(nq, world, shard_bits, len(ops)),
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
            "parity": parity,
        }
        if contracted:
            line["contracted_arithmetic"] = contracted
        if world > 1:
            # time of a step that is not spent inside a fused pass on rank 0: exchange time the passes did not hide (plus
            # single-gate launches and gaps); one qubit-slot exchange alone takes nvlink.slot_exchange.ms
            exchanges["exchange_ms_exposed"] = total_ms / args.steps - fused_ms.value / args.steps
            line["exchanges"] = exchanges
        if nvlink:
            line["nvlink"] = nvlink
        if circuit34:
            line["circuit_34q"] = circuit34
        if circuit33:
            line["circuit_33q"] = circuit33
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    lib.destroyQuESTEnv(env)
    if rank == 0 and parity is not None and not parity["ok"]:
        raise SystemExit(f"bench.py: PARITY FAILED against the reference CPU library: {parity}")


if __name__ == "__main__":
    main()
