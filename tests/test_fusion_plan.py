"""CPU-only check of the gate-fusion scheduler (qb_fuse.h through qb200_debugFuse): groups preserve program order,
never need more than four dense target qubits, and the per-gate descriptors (target register bit, register mask,
outer mask) select exactly the amplitudes the unfused gate selects.  The fused kernel's indexing is replayed in
numpy and compared with applying the gates one by one."""
import ctypes
import os
import random

import numpy as np
import pytest

from quest_b200 import capi

K_MATRIX, K_HAD, K_SWAP, K_Y, K_PHASE, K_FLIP = range(6)


def _lib():
    if not os.path.isfile(capi.PRODUCT_LIB):
        from quest_b200 import build
        build.build_product(2)
    lib = ctypes.CDLL(capi.PRODUCT_LIB, mode=ctypes.RTLD_LOCAL)
    lib.qb200_debugFuse.restype = ctypes.c_int
    return lib


def fuse(lib, local_bits, ops):
    n = len(ops)
    kinds = (ctypes.c_int * n)(*[o[0] for o in ops])
    targets = (ctypes.c_int * n)(*[o[1] for o in ops])
    masks = (ctypes.c_ulonglong * n)(*[o[2] for o in ops])
    groups = (ctypes.c_int * (7 * n))()
    desc = (ctypes.c_ulonglong * (4 * n))()
    ng = lib.qb200_debugFuse(local_bits, n, kinds, targets, masks, groups, desc)
    gl = [dict(first=groups[7 * i], count=groups[7 * i + 1], bits=list(groups[7 * i + 2:7 * i + 6]), ndense=groups[7 * i + 6])
          for i in range(ng)]
    dl = [dict(kind=desc[4 * i] & 0xFF, opcode=(desc[4 * i] >> 8) & 0xFF, tier=desc[4 * i] >> 16, pos=desc[4 * i + 1], reg=desc[4 * i + 2],
               outer=desc[4 * i + 3]) for i in range(n)]
    return gl, dl


def decode_opcode(opcode):
    """What k_fused's single table jump selects (qb_fuse.h FusedOpcode): (kind, target position, register mask)."""
    if opcode < 4:
        return K_MATRIX, opcode, 0            # uncontrolled among the fused qubits
    if opcode < 8:
        return K_MATRIX, opcode - 4, None     # mask taken from the descriptor at run time
    if opcode < 12:
        return K_HAD, opcode - 8, 0
    if opcode < 16:
        return K_SWAP, opcode - 12, None
    if opcode < 20:
        return K_Y, opcode - 16, None
    assert opcode >= 32
    if opcode < 48:
        return K_PHASE, None, opcode - 32
    assert opcode < 64
    return K_FLIP, None, opcode - 48


def unitary(rng):
    m = rng.standard_normal((2, 2)) + 1j * rng.standard_normal((2, 2))
    q, _ = np.linalg.qr(m)
    return q


def apply_unfused(state, op, u):
    kind, target, mask = op
    idx = np.arange(state.size)
    if kind in (K_PHASE, K_FLIP):
        sel = (idx & mask) == mask
        state[sel] *= (np.exp(0.3j) if kind == K_PHASE else -1)
        return
    sel = ((idx & mask) == mask) & ((idx >> target) & 1 == 0)
    up = idx[sel]
    lo = up | (1 << target)
    a, b = state[up].copy(), state[lo].copy()
    state[up], state[lo] = u[0, 0] * a + u[0, 1] * b, u[1, 0] * a + u[1, 1] * b


def apply_fused_group(state, local_bits, grp, descs, mats):
    """numpy twin of k_fused: thread = outer index (fused qubits at 0), registers = 16 amplitudes."""
    bits = grp["bits"]
    outer = np.arange(state.size >> 4, dtype=np.int64)
    for b in bits:  # ascending insertion of zeros
        low = outer & ((1 << b) - 1)
        outer = ((outer >> b) << (b + 1)) | low
    off = np.array([sum(((e >> j) & 1) << bits[j] for j in range(4)) for e in range(16)], dtype=np.int64)
    reg = state[outer[:, None] + off[None, :]]  # [threads, 16]
    for d, u in zip(descs, mats):
        active = (outer & d["outer"]) == d["outer"]
        kind, p, mask = decode_opcode(d["opcode"])  # the kernel sees only the opcode and the run-time register mask
        assert kind == d["kind"]
        # core tier: matrix without a control among the fused qubits, hadamard, swap, phase; everything else needs tier 1
        assert d["tier"] == (0 if kind in (K_HAD, K_SWAP, K_PHASE) or (kind == K_MATRIX and d["reg"] == 0) else 1)
        if mask is None:
            mask = d["reg"]
        assert mask == d["reg"] and (p is None or p == d["pos"])
        for e in range(16):
            if (e & mask) != mask:
                continue
            if kind in (K_PHASE, K_FLIP):
                reg[active, e] *= (np.exp(0.3j) if kind == K_PHASE else -1)
            else:
                if (e >> p) & 1:
                    continue
                f = e | (1 << p)
                a, b = reg[active, e].copy(), reg[active, f].copy()
                reg[active, e], reg[active, f] = u[0, 0] * a + u[0, 1] * b, u[1, 0] * a + u[1, 1] * b
    state[outer[:, None] + off[None, :]] = reg


@pytest.mark.parametrize("local_bits,seed", [(4, 1), (5, 2), (7, 3), (10, 4), (12, 5)])
def test_fused_groups_reproduce_gate_by_gate(local_bits, seed):
    lib = _lib()
    rng = random.Random(seed)
    nrng = np.random.default_rng(seed)
    n_ops = 120
    ops, mats = [], []
    for _ in range(n_ops):
        kind = rng.choice([K_MATRIX, K_MATRIX, K_HAD, K_SWAP, K_Y, K_PHASE, K_PHASE, K_FLIP])
        t = rng.randrange(local_bits)
        others = [q for q in range(local_bits) if q != t]
        if kind in (K_PHASE, K_FLIP):
            qs = rng.sample(range(local_bits), rng.randint(1, min(4, local_bits)))
            ops.append((kind, -1, sum(1 << q for q in qs)))
        else:
            cs = rng.sample(others, rng.choice([0, 0, 1, 2])) if others and kind != K_HAD else []  # no controlled hadamard in the API
            ops.append((kind, t, sum(1 << c for c in cs)))
        mats.append(unitary(nrng))
    groups, descs = fuse(lib, local_bits, ops)
    # structure: contiguous cover in order; at most four dense targets per group, all of them among the fused bits
    assert groups[0]["first"] == 0 and sum(g["count"] for g in groups) == n_ops
    for a, b in zip(groups, groups[1:]):
        assert b["first"] == a["first"] + a["count"]
    for g in groups:
        dense_targets = {ops[i][1] for i in range(g["first"], g["first"] + g["count"]) if ops[i][0] < K_PHASE}
        assert len(dense_targets) <= 4 and dense_targets <= set(g["bits"])
        assert g["bits"] == sorted(set(g["bits"])) and all(0 <= b < local_bits for b in g["bits"])
    # greedy maximality: a group ends only because the next dense gate needs a fifth qubit, or it is full (96 gates)
    for g, nxt in zip(groups, groups[1:]):
        if g["count"] == 96:
            continue
        dense_targets = {ops[i][1] for i in range(g["first"], g["first"] + g["count"]) if ops[i][0] < K_PHASE}
        assert len(dense_targets) == 4 and ops[nxt["first"]][0] < K_PHASE and ops[nxt["first"]][1] not in dense_targets
    state = nrng.standard_normal(1 << local_bits) + 1j * nrng.standard_normal(1 << local_bits)
    want = state.copy()
    for op, u in zip(ops, mats):
        apply_unfused(want, op, u)
    got = state.copy()
    for g in groups:
        sl = slice(g["first"], g["first"] + g["count"])
        apply_fused_group(got, local_bits, g, descs[sl], mats[sl])
    np.testing.assert_allclose(got, want, atol=1e-12)
