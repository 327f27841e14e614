"""Synthetic circuits and workload recipes shared by the tests, the golden-vector generator and bench.py.

Each circuit is a plain list of ``(api_name, *python_args)`` tuples.  :func:`run_circuit` replays it through any
:class:`quest_b200.capi.QuESTLibrary` — so the same list drives the CUDA library, the CPU oracle and the
reference's own CPU build, which is how parity is checked.  Generation uses ``random.Random(seed)`` only, so a
seed names the same circuit everywhere.

The named workloads follow SURVEY.md §8(d) / BASELINE.json ``configs``.
"""
from __future__ import annotations

import hashlib
import math
import random
from typing import List, Sequence, Tuple

import numpy as np

Op = Tuple

DENSE_1Q = ("hadamard", "pauliX", "pauliY", "rotateX", "rotateY", "rotateZ", "compactUnitary", "unitary", "rotateAroundAxis")
DIAG_1Q = ("pauliZ", "sGate", "tGate", "phaseShift")
CTRL_DENSE = ("controlledNot", "controlledPauliY", "controlledRotateX", "controlledRotateY", "controlledRotateZ",
              "controlledCompactUnitary", "controlledUnitary", "controlledRotateAroundAxis")
CTRL_DIAG = ("controlledPhaseShift", "controlledPhaseFlip")
MULTI = ("multiControlledUnitary", "multiControlledPhaseShift", "multiControlledPhaseFlip")
NOISE = ("applyOneQubitDephaseError", "applyOneQubitDepolariseError", "applyOneQubitDampingError")
NOISE2 = ("applyTwoQubitDephaseError", "applyTwoQubitDepolariseError")
SCALARS = ("calcProbOfOutcome", "calcTotalProb")


def _unit_pair(rng):
    """(alpha, beta) with |alpha|^2+|beta|^2 = 1 to well below REAL_EPS."""
    th, a, b = rng.uniform(0, math.pi), rng.uniform(0, 2 * math.pi), rng.uniform(0, 2 * math.pi)
    return complex(math.cos(th) * math.cos(a), math.cos(th) * math.sin(a)), complex(math.sin(th) * math.cos(b), math.sin(th) * math.sin(b))


def _unitary(rng):
    al, be = _unit_pair(rng)
    ph = rng.uniform(0, 2 * math.pi)
    g = complex(math.cos(ph), math.sin(ph))
    return [g * al, g * -be.conjugate(), g * be, g * al.conjugate()]


def random_circuit(nq: int, depth: int, seed: int, density: bool = False, measure: bool = True,
                   two_qubit_noise: bool = True, families: Sequence[str] | None = None) -> List[Op]:
    """`depth` operations drawn from every gate family on the hot path; targets and controls uniform over the
    register, so low (intra-vector), middle and top (shard-crossing) qubits are all hit."""
    rng = random.Random(seed)
    ops: List[Op] = []
    pools = [DENSE_1Q, DIAG_1Q, CTRL_DENSE, CTRL_DIAG, MULTI, SCALARS]
    if density:
        pools = pools + [NOISE] + ([NOISE2] if two_qubit_noise and nq >= 2 else [])
    for _ in range(depth):
        pool = rng.choice(pools)
        name = rng.choice(pool)
        if families is not None and name not in families:
            continue
        t = rng.randrange(nq)
        others = [q for q in range(nq) if q != t]
        c = rng.choice(others) if others else None
        ang = rng.uniform(-math.pi, math.pi)
        if name in ("hadamard", "pauliX", "pauliY", "pauliZ", "sGate", "tGate"):
            ops.append((name, t))
        elif name in ("rotateX", "rotateY", "rotateZ", "phaseShift"):
            ops.append((name, t, ang))
        elif name == "compactUnitary":
            ops.append((name, t, *_unit_pair(rng)))
        elif name == "unitary":
            ops.append((name, t, _unitary(rng)))
        elif name == "rotateAroundAxis":
            ops.append((name, t, ang, [rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(0.1, 1)]))
        elif c is None:
            continue
        elif name in ("controlledNot", "controlledPauliY", "controlledPhaseFlip"):
            ops.append((name, c, t))
        elif name in ("controlledRotateX", "controlledRotateY", "controlledRotateZ", "controlledPhaseShift"):
            ops.append((name, c, t, ang))
        elif name == "controlledCompactUnitary":
            ops.append((name, c, t, *_unit_pair(rng)))
        elif name == "controlledUnitary":
            ops.append((name, c, t, _unitary(rng)))
        elif name == "controlledRotateAroundAxis":
            ops.append((name, c, t, ang, [rng.uniform(-1, 1), rng.uniform(0.1, 1), rng.uniform(-1, 1)]))
        elif name == "multiControlledUnitary":
            k = rng.randint(1, min(3, len(others)))
            ops.append((name, rng.sample(others, k), t, _unitary(rng)))
        elif name in ("multiControlledPhaseShift", "multiControlledPhaseFlip"):
            k = rng.randint(1, min(4, nq))
            qs = rng.sample(range(nq), k)
            ops.append((name, qs, ang) if name.endswith("Shift") else (name, qs))
        elif name in NOISE:
            ops.append((name, t, rng.uniform(0.0, 0.3)))
        elif name in NOISE2:
            ops.append((name, t, c, rng.uniform(0.0, 0.3)))
        elif name == "calcProbOfOutcome":
            ops.append((name, t, rng.randrange(2)))
        elif name == "calcTotalProb":
            ops.append((name,))
    if measure:
        ops.append(("measure", rng.randrange(nq)))
        ops.append(("calcTotalProb",))
        ops.append(("measureWithStats", rng.randrange(nq)))
    if density:
        ops.append(("calcPurity",))
    ops.append(("calcTotalProb",))
    return ops


def apply_op(lib, qureg, op: Op):
    """Apply one op; returns a float for value-returning calls, else None."""
    name, *a = op
    fn = getattr(lib, name)
    if name in ("compactUnitary",):
        return fn(qureg, a[0], lib.complex(a[1]), lib.complex(a[2]))
    if name == "controlledCompactUnitary":
        return fn(qureg, a[0], a[1], lib.complex(a[2]), lib.complex(a[3]))
    if name == "unitary":
        return fn(qureg, a[0], lib.matrix(a[1]))
    if name == "controlledUnitary":
        return fn(qureg, a[0], a[1], lib.matrix(a[2]))
    if name == "rotateAroundAxis":
        return fn(qureg, a[0], a[1], lib.vector(a[2]))
    if name == "controlledRotateAroundAxis":
        return fn(qureg, a[0], a[1], a[2], lib.vector(a[3]))
    if name == "multiControlledUnitary":
        return fn(qureg, lib.int_array(a[0]), len(a[0]), a[1], lib.matrix(a[2]))
    if name == "multiControlledPhaseShift":
        return fn(qureg, lib.int_array(a[0]), len(a[0]), a[1])
    if name == "multiControlledPhaseFlip":
        return fn(qureg, lib.int_array(a[0]), len(a[0]))
    if name == "measureWithStats":
        p = lib.qreal(0)
        import ctypes
        out = fn(qureg, a[0], ctypes.byref(p))
        return float(out) + 0.25 * float(p.value)  # outcome in the integer part, probability/4 in the fraction
    out = fn(qureg, *a)
    return None if out is None else float(out)


def run_circuit(lib, qureg, ops: Sequence[Op], seed: int | None = None) -> List[float]:
    """Replay `ops`; returns every scalar the circuit produced (probabilities, outcomes, purity), in order."""
    if seed is not None:
        import ctypes
        key = (ctypes.c_ulong * 2)(seed, 2 * seed + 1)
        lib.seedQuEST(key, 2)
    scalars = []
    for op in ops:
        v = apply_op(lib, qureg, op)
        if v is not None:
            scalars.append(v)
    return scalars


def state_digest(state: np.ndarray) -> np.ndarray:
    """Size-independent summary of a state: norm, plain sums and three fixed pseudo-random projections.
    Two states that agree to 1e-12 relative agree in every entry to ~1e-12*norm."""
    n = state.size
    idx = np.arange(n, dtype=np.float64)
    out = [float(np.vdot(state, state).real), float(state.real.sum()), float(state.imag.sum())]
    for k in (0.7548776662466927, 0.5698402909980532, 0.3819660112501051):  # irrational strides -> fixed weights
        w = np.cos(2 * math.pi * ((idx * k) % 1.0))
        out += [float(np.dot(w, state.real)), float(np.dot(w, state.imag))]
    return np.array(out, dtype=np.float64)


def state_sha(state: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(state).tobytes()).hexdigest()


def random_state(n_amps: int, seed: int, dtype=np.complex128) -> np.ndarray:
    """Fixed pseudo-random normalised vector (the config-2 input state)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    v = rng.standard_normal(n_amps) + 1j * rng.standard_normal(n_amps)
    v /= np.linalg.norm(v)
    return v.astype(dtype)


def prepare(lib, qureg, seed: int):
    """Starting point for the seeded circuits: a dense random normalised state-vector, or |+><+| for a
    density matrix (kept physical so noise channels and purity behave)."""
    if qureg.isDensityMatrix:
        lib.initPlusState(qureg)
    else:
        lib.set_state(qureg, random_state(int(qureg.numAmpsTotal), seed, lib.np_complex))


# ---- BASELINE.json workloads -----------------------------------------------------------------------------

def config1_layers(nq: int = 24, layers: int = 20) -> List[Op]:
    """configs[0]: layers of hadamard / rotateY / controlledNot on every qubit."""
    ops: List[Op] = []
    for l in range(layers):
        ops += [("hadamard", i) for i in range(nq)]
        ops += [("rotateY", i, 0.1 * (i + 1) + 0.01 * l) for i in range(nq)]
        ops += [("controlledNot", i, (i + 1) % nq) for i in range(nq)]
    return ops


SWEEP_ALPHA = complex(math.cos(1.232) * math.cos(0.423), math.cos(1.232) * math.sin(0.423))
SWEEP_BETA = complex(math.sin(1.232) * math.cos(-0.6523), math.sin(1.232) * math.sin(-0.6523))


def config2_sweep(nq: int = 30) -> List[Op]:
    """configs[1]: every target for hadamard, compactUnitary, controlledNot, multiControlledPhaseShift."""
    ops: List[Op] = []
    for t in range(nq):
        ops.append(("hadamard", t))
        ops.append(("compactUnitary", t, SWEEP_ALPHA, SWEEP_BETA))
        ops.append(("controlledNot", (t + 1) % nq, t))
        ops.append(("multiControlledPhaseShift", [t, (t + 1) % nq, (t + 2) % nq], 0.7))
    return ops


def config3_qft(nq: int = 30) -> List[Op]:
    """configs[2]: QFT as in the reference's tests/algor/QFT.test:40-45, then P(1) of every qubit."""
    ops: List[Op] = []
    for q in range(nq):
        ops.append(("hadamard", q))
        for a in range(q + 1, nq):
            ops.append(("controlledPhaseShift", a, q, math.pi / (2 ** (a - q))))
    ops += [("calcProbOfOutcome", q, 1) for q in range(nq)]
    return ops


def config4_noisy(nq: int = 15, layers: int = 10) -> List[Op]:
    """configs[3]: noisy density-matrix layers with a purity read-out per layer."""
    ops: List[Op] = []
    for l in range(layers):
        for i in range(nq):
            ops.append(("hadamard", i))
            ops.append(("rotateY", i, 0.1 * (i + 1) + 0.01 * l))
            if i + 1 < nq:
                ops.append(("controlledNot", i, i + 1))
        for i in range(nq):
            ops.append(("applyOneQubitDephaseError", i, 0.02))
            ops.append(("applyOneQubitDepolariseError", i, 0.03))
            ops.append(("applyOneQubitDampingError", i, 0.05))
        ops.append(("calcPurity",))
    return ops


def config5_sharded(nq: int = 34, layers: int = 3, seed: int = 34) -> List[Op]:
    """configs[4]: random circuit with >=25% of targets on the top three (shard-crossing) qubits, then three
    measurements (one on a global qubit) and a collapseToOutcome on another global qubit.  Same seed => same circuit for every GPU count."""
    rng = random.Random(seed)
    ops: List[Op] = []
    for _ in range(layers):
        for _ in range(nq):
            t = rng.randrange(nq - 3, nq) if rng.random() < 0.3 else rng.randrange(nq)
            kind = rng.choice(("hadamard", "rotateY", "controlledNot", "compactUnitary"))
            if kind == "hadamard":
                ops.append((kind, t))
            elif kind == "rotateY":
                ops.append((kind, t, rng.uniform(-math.pi, math.pi)))
            elif kind == "controlledNot":
                c = rng.choice([q for q in range(nq) if q != t])
                ops.append((kind, c, t))
            else:
                ops.append((kind, t, *_unit_pair(rng)))
    ops.append(("measure", nq - 1))
    ops.append(("measure", 3))
    ops.append(("measure", nq // 2))
    # ... and a collapse onto a GIVEN outcome of a shard-selecting qubit (reference QuEST.c:546): every shard is either
    # renormalised as a whole or cleared (QuEST_cpu_distributed.c:1274-1292)
    ops.append(("collapseToOutcome", nq - 2, 0))
    ops.append(("calcTotalProb",))
    return ops


# ---- API fuzz: several registers, gates interleaved with every kind of observer and copy ----------------------

def random_api_sequence(nq: int, depth: int, seed: int, extended: bool = False):
    """Random calls over two state-vectors (a, b) and two density matrices (r, s) of nq qubits: gates, noise,
    initialisers, partial setAmps, clone, inner product, fidelity, compare, mixing, probabilities, measurement.
    Exercises every path that must drain a register's recorded gates (and restore its qubit layout) first.
    extended=True adds initPureState (whose reference build has a sign slip, so only product-vs-oracle runs use it)
    and the two-qubit noise channels."""
    rng = random.Random(seed)
    calls = []
    gate_pool = DENSE_1Q + DIAG_1Q + CTRL_DENSE + CTRL_DIAG + MULTI
    for _ in range(depth):
        kind = rng.random()
        if kind < 0.62:
            reg = rng.choice("abrs")
            ops = random_circuit(nq, 3, rng.randrange(1 << 30), density=False, measure=False, families=gate_pool)
            calls += [("gate", reg, op) for op in ops[:2] if op[0] in gate_pool]
        elif kind < 0.70:
            reg = rng.choice("rs")
            calls.append(("gate", reg, (rng.choice(NOISE), rng.randrange(nq), rng.uniform(0, 0.3))))
        elif kind < 0.74:
            calls.append(("clone", *rng.choice([("a", "b"), ("b", "a"), ("r", "s"), ("s", "r")])))
        elif kind < 0.78:
            calls.append(("inner", "a", "b"))
        elif kind < 0.81:
            calls.append(("fidelity", rng.choice("ar"), "b"))
        elif kind < 0.84:
            calls.append(("compare", *rng.choice([("a", "b"), ("r", "s")])))
        elif kind < 0.87:
            calls.append(("mix", "r", rng.uniform(0, 1), "s"))
        elif kind < 0.90:
            start = rng.randrange(1 << nq)
            n = rng.randint(1, (1 << nq) - start)
            vals = [complex(rng.uniform(-0.1, 0.1), rng.uniform(-0.1, 0.1)) for _ in range(min(n, 7))]
            calls.append(("setamps", rng.choice("ab"), start, vals))
        elif kind < 0.93:
            calls.append(("init", rng.choice("abrs"), rng.choice(["zero", "plus", "debug", "classical"]), rng.randrange(1 << nq)))
        elif kind < 0.96:
            calls.append(("prob", rng.choice("abrs"), rng.randrange(nq), rng.randrange(2)))
        elif kind < 0.98:
            calls.append(("getamp", rng.choice("ab"), rng.randrange(1 << nq)))
        else:
            calls.append(("purity", rng.choice("rs")))
        if extended and rng.random() < 0.06:
            if rng.random() < 0.4:
                calls.append(("initpure", rng.choice("rs"), rng.choice("ab")))
            else:
                t1, t2 = rng.sample(range(nq), 2)
                calls.append(("gate", rng.choice("rs"), (rng.choice(NOISE2), t1, t2, rng.uniform(0, 0.3))))
    calls += [("totalprob", r) for r in "abrs"]
    return calls


def run_api_sequence(lib, env, nq: int, calls, seed: int):
    """Replay `calls`; returns (scalars, {register: final state})."""
    import ctypes
    regs = {"a": lib.createQureg(nq, env), "b": lib.createQureg(nq, env),
            "r": lib.createDensityQureg(nq, env), "s": lib.createDensityQureg(nq, env)}
    for name, q in regs.items():
        prepare(lib, q, seed + ord(name))
        if q.isDensityMatrix:
            lib.rotateY(q, 0, 0.3 + 0.1 * ord(name))
    key = (ctypes.c_ulong * 1)(seed)
    lib.seedQuEST(key, 1)
    out = []
    labels = []  # which call produced each scalar (for diagnostics)
    for ci, c in enumerate(calls):
        k = c[0]
        n_before = len(out)
        if k == "gate":
            v = apply_op(lib, regs[c[1]], c[2])
            if v is not None:
                out.append(v)
        elif k == "clone":
            lib.cloneQureg(regs[c[1]], regs[c[2]])
        elif k == "inner":
            z = lib.calcInnerProduct(regs[c[1]], regs[c[2]])
            out += [z.real, z.imag]
        elif k == "fidelity":
            out.append(lib.calcFidelity(regs[c[1]], regs[c[2]]))
        elif k == "compare":
            out.append(float(lib.compareStates(regs[c[1]], regs[c[2]], 1e-9)))
        elif k == "mix":
            lib.addDensityMatrix(regs[c[1]], c[2], regs[c[3]])
        elif k == "setamps":
            lib.set_state(regs[c[1]], np.array(c[3]), start=min(c[2], (1 << nq) - len(c[3])))
        elif k == "init":
            q = regs[c[1]]
            {"zero": lib.initZeroState, "plus": lib.initPlusState, "debug": lib.initStateDebug}.get(c[2], lambda r: lib.initClassicalState(r, c[3]))(q)
        elif k == "prob":
            out.append(lib.calcProbOfOutcome(regs[c[1]], c[2], c[3]))
        elif k == "getamp":
            z = lib.getAmp(regs[c[1]], c[2])
            out += [z.real, z.imag]
        elif k == "purity":
            out.append(lib.calcPurity(regs[c[1]]))
        elif k == "initpure":
            lib.initPureState(regs[c[1]], regs[c[2]])
        elif k == "totalprob":
            out.append(lib.calcTotalProb(regs[c[1]]))
        labels += [(ci, c[:3])] * (len(out) - n_before)
    run_api_sequence.last_labels = labels
    states = {name: lib.get_state(q) for name, q in regs.items()}
    for q in regs.values():
        lib.destroyQureg(q, env)
    return np.array(out, dtype=np.float64), states
