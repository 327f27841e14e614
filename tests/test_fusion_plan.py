"""CPU-only check of the gate-fusion planner (qb_fuse.h through the test hook qb200_debugTilePlan): passes preserve
program order, a pass's dense targets plus the four lowest index bits never need more than eleven tile bits, a
sub-group never more than three register bits, and the per-gate descriptors (opcode, register mask, tile mask, outer
mask) select exactly the amplitudes the unfused gate selects.  The tile kernel's indexing (CTA index -> tile base by bit
insertion, tile index -> amplitude index by deposit, thread block -> register elements) is replayed in numpy and
compared with applying the gates one by one."""
import ctypes
import os
import random

import numpy as np
import pytest

from quest_b200 import capi

K_MATRIX, K_HAD, K_SWAP, K_Y, K_PHASE, K_FLIP, K_DEPHASE, K_DEPOL, K_DAMP = range(9)
TILE_BITS, TILE_LOW, SUB_BITS, MAX_GATES, MAX_SUB_GATES = 11, 4, 3, 192, 32


def _lib():
    if not os.path.isfile(capi.PRODUCT_LIB):
        from quest_b200 import build
        build.build_product(2)
    lib = ctypes.CDLL(capi.PRODUCT_LIB, mode=ctypes.RTLD_LOCAL)
    lib.qb200_debugTilePlan.restype = ctypes.c_int
    return lib


def plan(lib, local_bits, ops, reorder=1):
    n = len(ops)
    kinds = (ctypes.c_int * n)(*[o[0] for o in ops])
    targets = (ctypes.c_int * n)(*[o[1] for o in ops])
    targets2 = (ctypes.c_int * n)(*[o[3] if len(o) > 3 else -1 for o in ops])
    masks = (ctypes.c_ulonglong * n)(*[o[2] for o in ops])
    passes = (ctypes.c_int * ((4 + TILE_BITS) * n))()
    order = (ctypes.c_int * n)()
    subs = (ctypes.c_int * (7 * n))()
    desc = (ctypes.c_ulonglong * (4 * n))()
    n_subs = ctypes.c_int()
    np_ = lib.qb200_debugTilePlan(local_bits, n, kinds, targets, targets2, masks, reorder, passes, order, subs, desc, ctypes.byref(n_subs))
    W = 4 + TILE_BITS
    pl = [dict(first=passes[W * i], count=passes[W * i + 1], ntile=passes[W * i + 2], nsub=passes[W * i + 3],
               bits=[b for b in passes[W * i + 4:W * i + W] if b >= 0], subs=[]) for i in range(np_)]
    for p in pl:
        p["order"] = list(order[p["first"]:p["first"] + p["count"]])
    for j in range(n_subs.value):
        pl[subs[7 * j]]["subs"].append(dict(first=subs[7 * j + 1], count=subs[7 * j + 2], rb=list(subs[7 * j + 3:7 * j + 6]),
                                            spect=subs[7 * j + 6]))
    dl = [dict(opcode=desc[4 * i], reg=desc[4 * i + 1], tile=desc[4 * i + 2], outer=desc[4 * i + 3]) for i in range(n)]
    return pl, dl, list(order)


def decode_opcode(opcode):
    """What k_tile's single table jump selects (qb_fuse.h TileOpcode): (kind, register position, second position)."""
    if opcode < 12:
        return (K_MATRIX, K_HAD, K_SWAP, K_Y)[opcode // 3], opcode % 3, None
    if opcode == 12:
        return K_FLIP, None, None
    assert opcode >= 16  # 13..15: hadamard + matrix fused by the flush-time peephole, never produced by the planner itself
    if opcode < 22:      # controlledNot with its control among the register bits: position pair -> (target, control)
        t, which = divmod(opcode - 16, 2)
        return K_SWAP, t, None
    if opcode < 30:      # phase: the register mask is part of the opcode
        return K_PHASE, None, None
    kind, slot = divmod(opcode - 30, 3)
    return K_DEPHASE + kind, (0, 0, 1)[slot], (1, 2, 2)[slot]


def unitary(rng):
    m = rng.standard_normal((2, 2)) + 1j * rng.standard_normal((2, 2))
    q, _ = np.linalg.qr(m)
    return q


def apply_unfused(state, op, u):
    kind, target, mask = op[:3]
    idx = np.arange(state.size)
    if kind in (K_PHASE, K_FLIP):
        sel = (idx & mask) == mask
        state[sel] *= (np.exp(0.3j) if kind == K_PHASE else -1)
        return
    if kind >= K_DEPHASE:  # stand-in for a noise channel: scale the two coherence blocks of the (target, target2) structure
        t2 = op[3]
        a, b = (idx >> target) & 1, (idx >> t2) & 1
        state[a != b] *= 0.9
        return
    sel = ((idx & mask) == mask) & ((idx >> target) & 1 == 0)
    up = idx[sel]
    lo = up | (1 << target)
    a, b = state[up].copy(), state[lo].copy()
    state[up], state[lo] = u[0, 0] * a + u[0, 1] * b, u[1, 0] * a + u[1, 1] * b


def insert_zeros(k, positions):
    for b in sorted(positions):
        low = k & ((1 << b) - 1)
        k = ((k >> b) << (b + 1)) | low
    return k


def apply_pass(state, ops, p, descs, mats, order):
    """numpy twin of k_tile: CTA = tile base (tile bits at 0), tile = 2^ntile amplitudes, thread = block of 8 elements."""
    bits, T = p["bits"], p["ntile"]
    bases = insert_zeros(np.arange(state.size >> T, dtype=np.int64), bits)
    dep = np.array([sum(((i >> j) & 1) << bits[j] for j in range(T)) for i in range(1 << T)], dtype=np.int64)
    assert all(bits[j] == j for j in range(min(TILE_LOW, T)))  # rows of the tile are contiguous in memory
    tile = state[bases[:, None] + dep[None, :]]  # [CTAs, 2^T]
    for sg in p["subs"]:
        rb, spect = sg["rb"], sg["spect"]
        assert rb == sorted(rb) and all(0 <= r < T for r in rb) and 0 <= spect < T and spect not in rb
        assert sg["count"] <= MAX_SUB_GATES
        # a thread holds the 16 elements spanned by the register bits and the spectator bit; no gate of the sub-group may
        # mention the spectator bit (both 8-element halves run exactly the same gates)
        for k in order[sg["first"]:sg["first"] + sg["count"]]:
            assert not (descs[k]["tile"] >> spect) & 1
        blocks = insert_zeros(np.arange((1 << T) >> (SUB_BITS + 1), dtype=np.int64), rb + [spect])
        blocks = np.concatenate([blocks, blocks | (1 << spect)])  # the two halves, as two 8-element blocks
        off = np.array([sum(((e >> j) & 1) << rb[j] for j in range(SUB_BITS)) for e in range(1 << SUB_BITS)], dtype=np.int64)
        for k in order[sg["first"]:sg["first"] + sg["count"]]:
            d, u = descs[k], mats[k]
            kind, pos, pos2 = decode_opcode(d["opcode"])
            assert kind == ops[k][0]
            if kind == K_PHASE:
                assert d["opcode"] - 22 == d["reg"]
            cta_on = (bases & d["outer"]) == d["outer"]
            thr_on = (blocks & d["tile"]) == d["tile"]
            assert d["tile"] & sum(1 << r for r in rb) == 0
            cc, tt = np.nonzero(cta_on)[0][:, None], blocks[thr_on][None, :]
            for e in range(1 << SUB_BITS):
                if (e & d["reg"]) != d["reg"]:
                    continue
                if kind in (K_PHASE, K_FLIP):
                    tile[cc, tt + off[e]] *= (np.exp(0.3j) if kind == K_PHASE else -1)
                elif kind >= K_DEPHASE:
                    if ((e >> pos) & 1) != ((e >> pos2) & 1):
                        tile[cc, tt + off[e]] *= 0.9
                else:
                    if (e >> pos) & 1:
                        continue
                    f = e | (1 << pos)
                    a, b = tile[cc, tt + off[e]].copy(), tile[cc, tt + off[f]].copy()
                    tile[cc, tt + off[e]], tile[cc, tt + off[f]] = u[0, 0] * a + u[0, 1] * b, u[1, 0] * a + u[1, 1] * b
    state[bases[:, None] + dep[None, :]] = tile


def random_ops(rng, local_bits, n_ops, noise=False):
    ops = []
    for _ in range(n_ops):
        kind = rng.choice([K_MATRIX, K_MATRIX, K_HAD, K_SWAP, K_Y, K_PHASE, K_PHASE, K_FLIP] + ([K_DEPHASE, K_DEPOL, K_DAMP] if noise else []))
        t = rng.randrange(local_bits)
        others = [q for q in range(local_bits) if q != t]
        if kind in (K_PHASE, K_FLIP):
            qs = rng.sample(range(local_bits), rng.randint(1, min(4, local_bits)))
            ops.append((kind, -1, sum(1 << q for q in qs)))
        elif kind >= K_DEPHASE:
            ops.append((kind, t, 0, rng.choice(others)))
        else:
            cs = rng.sample(others, rng.choice([0, 0, 1, 2])) if kind != K_HAD else []  # no controlled hadamard in the API
            ops.append((kind, t, sum(1 << c for c in cs)))
    return ops


def dense_bits(op):
    if op[0] in (K_PHASE, K_FLIP):
        return set()
    return {op[1]} | ({op[3]} if len(op) > 3 else set())


@pytest.mark.parametrize("reorder", [1, 0], ids=["reordering", "consecutive"])
@pytest.mark.parametrize("local_bits,seed,noise", [(4, 1, False), (4, 2, False), (5, 3, True), (7, 4, False), (10, 5, True),
                                                   (12, 6, False), (14, 7, True), (16, 8, False)])
def test_planned_passes_reproduce_gate_by_gate(local_bits, seed, noise, reorder):
    lib = _lib()
    rng = random.Random(seed)
    nrng = np.random.default_rng(seed)
    n_ops = 150
    ops = random_ops(rng, local_bits, n_ops, noise)
    mats = [unitary(nrng) for _ in ops]
    passes, descs, order = plan(lib, local_bits, ops, reorder)
    T = min(TILE_BITS, local_bits)
    low = set(range(min(TILE_LOW, T)))
    # structure: every gate in exactly one pass; inside a pass program order is kept; dense targets + low run fit the
    # tile; sub-groups cover the pass in order
    assert sorted(order) == list(range(n_ops)) and sum(p["count"] for p in passes) == n_ops
    for p in passes:
        assert p["order"] == sorted(p["order"])
        dense = set().union(*[dense_bits(ops[i]) for i in p["order"]])
        assert p["ntile"] == T and len(p["bits"]) == T and p["bits"] == sorted(set(p["bits"]))
        assert (dense | low) <= set(p["bits"]) and all(0 <= b < local_bits for b in p["bits"])
        assert p["subs"][0]["first"] == p["first"] and sum(s["count"] for s in p["subs"]) == p["count"]
        for s in p["subs"]:
            sd = set().union(*[dense_bits(ops[i]) for i in order[s["first"]:s["first"] + s["count"]]])
            assert len(sd) <= SUB_BITS and {p["bits"].index(b) for b in sd} <= set(s["rb"])
    if not reorder:  # consecutive passes, greedy: a pass ends only because the next gate does not fit, or it is full
        assert order == list(range(n_ops))
        loner = lambda op: len(dense_bits(op)) + bin(op[2]).count("1") > min(8, T - 1)
        for p, nxt in zip(passes, passes[1:]):
            if p["count"] == MAX_GATES or len(p["subs"]) == 64:
                continue
            if loner(ops[nxt["order"][0]]) or (p["count"] == 1 and loner(ops[p["order"][0]])):
                continue
            dense = set().union(*[dense_bits(ops[i]) for i in p["order"]])
            assert len(dense | low | dense_bits(ops[nxt["order"][0]])) > T
    # whatever was reordered must have been legal: random unitaries do not commute by accident
    state = nrng.standard_normal(1 << local_bits) + 1j * nrng.standard_normal(1 << local_bits)
    want = state.copy()
    for op, u in zip(ops, mats):
        apply_unfused(want, op, u)
    got = state.copy()
    for p in passes:
        if p["count"] == 1:  # flush() runs one-gate passes through the single-gate kernels
            apply_unfused(got, ops[p["order"][0]], mats[p["order"][0]])
        else:
            apply_pass(got, ops, p, descs, mats, order)
    np.testing.assert_allclose(got, want, atol=1e-12)


def test_layered_circuit_runs_qubit_block_by_qubit_block():
    """BASELINE configs[0]: layers of hadamard on every qubit, a rotation on every qubit, a ring of controlledNots.  In
    program order every layer of 24 one-qubit gates is 3 passes; with gates allowed to move forward past gates they
    commute with, a whole layer (72 gates) is 3 passes."""
    lib = _lib()
    nq, layers = 24, 4
    ops = []
    for _ in range(layers):
        ops += [(K_HAD, i, 0) for i in range(nq)] + [(K_MATRIX, i, 0) for i in range(nq)]
        ops += [(K_SWAP, (i + 1) % nq, 1 << i) for i in range(nq)]
    in_order, _, _ = plan(lib, nq, ops, 0)
    moved, descs, order = plan(lib, nq, ops, 1)
    assert len(in_order) >= 8 * layers and len(moved) <= 3 * layers + 2, (len(in_order), len(moved))
    rng = np.random.default_rng(3)
    mats = [unitary(rng) for _ in ops]
    state = rng.standard_normal(1 << 12) + 1j * rng.standard_normal(1 << 12)  # the same structure on 12 qubits, executed
    ops12 = []
    for _ in range(3):
        ops12 += [(K_HAD, i, 0) for i in range(12)] + [(K_MATRIX, i, 0) for i in range(12)] + [(K_SWAP, (i + 1) % 12, 1 << i) for i in range(12)]
    mats12 = [unitary(rng) for _ in ops12]
    passes, descs, order = plan(lib, 12, ops12, 1)
    want, got = state.copy(), state.copy()
    for op, u in zip(ops12, mats12):
        apply_unfused(want, op, u)
    for p in passes:
        if p["count"] == 1:
            apply_unfused(got, ops12[p["order"][0]], mats12[p["order"][0]])
        else:
            apply_pass(got, ops12, p, descs, mats12, order)
    np.testing.assert_allclose(got, want, atol=1e-12)


def test_sweep_step_needs_four_passes_at_30_qubits():
    """The benchmark's step (configs[1]: 120 gates, every target 0..29) is four passes over HBM."""
    lib = _lib()
    ops = []
    for t in range(30):
        ops += [(K_HAD, t, 0), (K_MATRIX, t, 0), (K_SWAP, t, 1 << ((t + 1) % 30)),
                (K_PHASE, -1, (1 << t) | (1 << ((t + 1) % 30)) | (1 << ((t + 2) % 30)))]
    passes, _, _ = plan(lib, 30, ops)
    assert len(passes) == 4, [p["count"] for p in passes]
    assert [p["count"] for p in passes][0] >= 44  # targets 0..10 (plus phase gates that could move forward)


def test_shared_memory_placement_is_a_conflict_light_bijection():
    """qb::tilePlace (mirrored here): a bijection of the tile, linear over GF(2), and for every choice of three register
    bits the 16 threads of a half-warp (which differ in the four lowest non-register bits) hit each bank at most twice."""
    import itertools

    def place(i):
        h = i >> 4
        return i ^ ((h ^ (h << 1) ^ (h << 3)) & 15)

    n = 1 << TILE_BITS
    assert sorted(place(i) for i in range(n)) == list(range(n))
    rng = random.Random(0)
    for _ in range(200):
        a, b = rng.randrange(n), rng.randrange(n)
        assert place(a ^ b) == place(a) ^ place(b)
    worst = 0
    for reg in itertools.combinations(range(TILE_BITS), SUB_BITS):
        free = [b for b in range(TILE_BITS) if b not in reg]
        for half in (0, 1):
            for e in range(1 << SUB_BITS):
                hits = {}
                for lane in range(16):
                    idx = sum(((lane >> j) & 1) << free[j] for j in range(4)) | (half << free[4])
                    idx |= sum(((e >> j) & 1) << reg[j] for j in range(SUB_BITS))
                    bank = place(idx) & 15
                    hits[bank] = hits.get(bank, 0) + 1
                worst = max(worst, max(hits.values()))
    assert worst <= 2


def test_peephole_fuses_only_adjacent_uncontrolled_hadamard_then_matrix_on_one_qubit():
    """qb::fuseAdjacentPairs: hadamard(t) directly followed by an uncontrolled matrix gate on t becomes one op (kind 10,
    carrying the matrix and the hadamard factor); anything else - another qubit, a control, the opposite order, a gate in
    between - stays as it is, in order."""
    lib = _lib()
    lib.qb200_debugPeephole.restype = ctypes.c_int
    ops = [(K_HAD, 3, 0), (K_MATRIX, 3, 0),            # fused
           (K_HAD, 3, 0), (K_MATRIX, 4, 0),            # different qubit
           (K_HAD, 5, 0), (K_MATRIX, 5, 1 << 2),       # controlled matrix
           (K_MATRIX, 6, 0), (K_HAD, 6, 0),            # opposite order
           (K_HAD, 7, 0), (K_PHASE, -1, 1 << 7), (K_MATRIX, 7, 0),   # a gate in between
           (K_HAD, 8, 0), (K_HAD, 8, 0), (K_MATRIX, 8, 0),           # only the adjacent pair
           (K_HAD, 9, 0)]                              # trailing hadamard
    n = len(ops)
    kinds_out, first_out = (ctypes.c_int * n)(), (ctypes.c_int * n)()
    m = lib.qb200_debugPeephole(n, (ctypes.c_int * n)(*[o[0] for o in ops]), (ctypes.c_int * n)(*[o[1] for o in ops]),
                                (ctypes.c_ulonglong * n)(*[o[2] for o in ops]), kinds_out, first_out)
    K_HM = 10
    got = [(kinds_out[k], first_out[k]) for k in range(m)]
    want = [(K_HM, 0), (K_HAD, 2), (K_MATRIX, 3), (K_HAD, 4), (K_MATRIX, 5), (K_MATRIX, 6), (K_HAD, 7), (K_HAD, 8), (K_PHASE, 9),
            (K_MATRIX, 10), (K_HAD, 11), (K_HM, 12), (K_HAD, 14)]
    assert got == want
