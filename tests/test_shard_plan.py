"""CPU-only coverage of the multi-GPU path's host logic.

The product decides per gate and per rank whether to skip, run a local pass, or process its half of the pairs of a
shard-crossing gate (qb_plan.h: planShardedGate, exported for tests as qb200_debugPlanGate; no CUDA call is made).
Here that REAL plan is executed by a small numpy stand-in for the kernels:

* in one process, emulating every rank of a 2/4/8-way sharded register, against a dense numpy reference — checks that
  every amplitude pair is owned by exactly one task of exactly one rank;
* in world_size-2 and -4 gloo process groups, one shard per process, with the partner's amplitudes moved by
  torch.distributed send/recv where the GPUs would use NVLink peer access — checks the SPMD protocol (partner choice,
  upper/lower roles, half split, control localisation) end to end against the CPU oracle.
"""
import ctypes
import os
import random

import numpy as np
import pytest

from quest_b200 import capi, circuits

ACT_SKIP, ACT_LOCAL, ACT_CROSS = 0, 1, 2


def _plan_fn():
    if not os.path.isfile(capi.PRODUCT_LIB):
        from quest_b200 import build
        build.build_product(2)
    lib = ctypes.CDLL(capi.PRODUCT_LIB, mode=ctypes.RTLD_LOCAL)
    fn = lib.qb200_debugPlanGate
    fn.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_ulonglong, ctypes.c_int, ctypes.POINTER(ctypes.c_ulonglong)]
    fn.restype = None
    return fn


def plan_gate(fn, local_bits, chunk, target, ctrl_mask, vec):
    out = (ctypes.c_ulonglong * 18)()
    fn(local_bits, chunk, target, ctrl_mask, vec, out)
    o = list(out)
    return dict(action=o[0], partner=o[1], upper=bool(o[2]), launches=o[3], n_tasks=o[4], pos=o[8:8 + o[5]],
                set_mask=o[6], lo_mask=o[7], skip_mask=o[16], ctrl=o[17])


def selected_indices(p, vec, target_lane=None):
    """Amplitude indices of the 'up' element of every pair a launch with plan `p` visits (numpy twin of
    qb::expand + the lane predicate of k_pairs / k_pairsPeer)."""
    k = np.arange(p["n_tasks"], dtype=np.uint64)
    for b in p["pos"]:
        b = np.uint64(b)
        low = k & ((np.uint64(1) << b) - np.uint64(1))
        k = ((k >> b) << (b + np.uint64(1))) | low
    k |= np.uint64(p["set_mask"])
    k = k[(k & np.uint64(p["skip_mask"])) == np.uint64(p["skip_mask"])]
    lanes = [e for e in range(vec) if (e & p["lo_mask"]) == p["lo_mask"] and (target_lane is None or not (e >> target_lane) & 1)]
    return (k[:, None] * np.uint64(vec) + np.array(lanes, dtype=np.uint64)[None, :]).ravel().astype(np.int64)


def apply_2x2(u, up, lo):
    return u[0, 0] * up + u[0, 1] * lo, u[1, 0] * up + u[1, 1] * lo


def dense_reference(state, nq, target, ctrl_mask, u):
    idx = np.arange(state.size)
    sel = ((idx & ctrl_mask) == ctrl_mask) & ((idx >> target) & 1 == 0)
    up = idx[sel]
    lo = up | (1 << target)
    out = state.copy()
    out[up], out[lo] = apply_2x2(u, state[up], state[lo])
    return out


def rand_u(rng):
    m = np.array(circuits._unitary(rng)).reshape(2, 2)
    return m


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("vec", [1, 2, 4])
def test_plan_covers_every_pair_exactly_once(world, vec):
    fn = _plan_fn()
    rng = random.Random(100 * world + vec)
    nq = 9
    gbits = world.bit_length() - 1
    L = nq - gbits
    n = 1 << L
    v = vec.bit_length() - 1
    for trial in range(60):
        target = rng.randrange(nq)
        others = [q for q in range(nq) if q != target]
        ctrls = rng.sample(others, rng.choice([0, 0, 1, 2, 3, 8]))
        cm = sum(1 << c for c in ctrls)
        u = rand_u(rng)
        state = np.random.default_rng(trial).standard_normal(1 << nq) + 1j * np.random.default_rng(trial + 1).standard_normal(1 << nq)
        want = dense_reference(state, nq, target, cm, u)
        shards = [state[r * n:(r + 1) * n].copy() for r in range(world)]
        new = [s.copy() for s in shards]
        touched = [np.zeros(n, dtype=int) for _ in range(world)]
        for r in range(world):
            p = plan_gate(fn, L, r, target, cm, vec)
            if p["action"] == ACT_SKIP:
                continue
            if p["action"] == ACT_LOCAL:
                up = selected_indices(p, vec, target if target < v else None)
                lo = up | (1 << target)
                new[r][up], new[r][lo] = apply_2x2(u, shards[r][up], shards[r][lo])
                touched[r][up] += 1
                touched[r][lo] += 1
            else:
                assert p["partner"] == r ^ (1 << (target - L)) and p["upper"] == (((r >> (target - L)) & 1) == 0)
                if not p["launches"]:
                    continue
                idx = selected_indices(p, vec)
                ur, lr = (r, p["partner"]) if p["upper"] else (p["partner"], r)
                new[ur][idx], new[lr][idx] = apply_2x2(u, shards[ur][idx], shards[lr][idx])
                touched[ur][idx] += 1
                touched[lr][idx] += 1
        got = np.concatenate(new)
        np.testing.assert_allclose(got, want, atol=1e-13)
        # exactly the selected amplitudes were written, each once
        idx = np.arange(1 << nq)
        expect_touch = ((idx & cm) == cm).astype(int)
        assert np.array_equal(np.concatenate(touched), expect_touch), (world, vec, target, ctrls)


# ---- gloo process groups: one shard per process ------------------------------------------------------------------

def _worker(rank, world, port, nq, seed, result_path):
    import torch
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    fn = _plan_fn()
    gbits = world.bit_length() - 1
    L = nq - gbits
    n = 1 << L
    vec = 4
    v = 2
    full = circuits.random_state(1 << nq, seed)
    shard = full[rank * n:(rank + 1) * n].copy()
    rng = random.Random(seed)
    ops = []
    for _ in range(40):
        target = rng.randrange(nq) if rng.random() < 0.5 else rng.randrange(L, nq)  # half the gates cross shards
        others = [q for q in range(nq) if q != target]
        ctrls = rng.sample(others, rng.choice([0, 0, 1, 2]))
        ops.append((target, ctrls, circuits._unitary(rng)))
    for target, ctrls, mat in ops:
        u = np.array(mat).reshape(2, 2)
        cm = sum(1 << c for c in ctrls)
        p = plan_gate(fn, L, rank, target, cm, vec)
        if p["action"] == ACT_SKIP:
            continue
        if p["action"] == ACT_LOCAL:
            up = selected_indices(p, vec, target if target < v else None)
            lo = up | (1 << target)
            shard[up], shard[lo] = apply_2x2(u, shard[up], shard[lo])
            continue
        # cross-shard: what the GPUs do with peer loads/stores is done here with send/recv of the touched slices
        partner = p["partner"]
        mine = selected_indices(p, vec) if p["launches"] else np.zeros(0, dtype=np.int64)
        # the partner processes the complementary half: it needs my amplitudes at ITS indices
        theirs = selected_indices(plan_gate(fn, L, partner, target, cm, vec), vec) \
            if plan_gate(fn, L, partner, target, cm, vec)["launches"] else np.zeros(0, dtype=np.int64)
        send_buf = torch.from_numpy(np.ascontiguousarray(shard[theirs]).view(np.float64).copy())
        recv_buf = torch.empty(2 * mine.size, dtype=torch.float64)
        reqs = []
        if send_buf.numel():
            reqs.append(dist.isend(send_buf, partner))
        if recv_buf.numel():
            reqs.append(dist.irecv(recv_buf, partner))
        for r in reqs:
            r.wait()
        remote = recv_buf.numpy().view(np.complex128)
        if p["upper"]:
            new_mine, new_remote = apply_2x2(u, shard[mine], remote)
        else:
            new_remote, new_mine = apply_2x2(u, remote, shard[mine])
        shard[mine] = new_mine
        # "remote write": return the partner's updated amplitudes
        back = torch.from_numpy(np.ascontiguousarray(new_remote).view(np.float64).copy())
        got_back = torch.empty(2 * theirs.size, dtype=torch.float64)
        reqs = []
        if back.numel():
            reqs.append(dist.isend(back, partner))
        if got_back.numel():
            reqs.append(dist.irecv(got_back, partner))
        for r in reqs:
            r.wait()
        shard[theirs] = got_back.numpy().view(np.complex128)
    gathered = [torch.empty(2 * n, dtype=torch.float64) for _ in range(world)] if rank == 0 else None
    dist.gather(torch.from_numpy(shard.view(np.float64).copy()), gathered, dst=0)
    if rank == 0:
        np.save(result_path, np.concatenate([g.numpy().view(np.complex128) for g in gathered]))
        np.save(result_path + ".ops.npy", np.array([(t, sum(1 << c for c in cs)) for t, cs, _ in ops]))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_protocol_over_gloo_matches_oracle(world, oracle, oracle_env, tmp_path):
    import torch.multiprocessing as mp

    nq, seed = 10, 77 + world
    port = 29600 + world + (os.getpid() % 200)
    result = str(tmp_path / "state.npy")
    mp.spawn(_worker, args=(world, port, nq, seed, result), nprocs=world, join=True)
    got = np.load(result)
    # the same gate sequence through the CPU oracle on one unsharded register
    rng = random.Random(seed)
    L = nq - (world.bit_length() - 1)
    q = oracle.createQureg(nq, oracle_env)
    oracle.set_state(q, circuits.random_state(1 << nq, seed))
    for _ in range(40):
        target = rng.randrange(nq) if rng.random() < 0.5 else rng.randrange(L, nq)
        others = [x for x in range(nq) if x != target]
        ctrls = rng.sample(others, rng.choice([0, 0, 1, 2]))
        mat = circuits._unitary(rng)
        if ctrls:
            oracle.multiControlledUnitary(q, oracle.int_array(ctrls), len(ctrls), target, oracle.matrix(mat))
        else:
            oracle.unitary(q, target, oracle.matrix(mat))
    want = oracle.get_state(q)
    oracle.destroyQureg(q, oracle_env)
    np.testing.assert_allclose(got, want, atol=1e-13)


# ---- qubit remapping (slot exchange instead of a cross-shard pass per gate) -----------------------------------------

def _swap_bits(state, a, b):
    """Permute a full state so that index bits a and b trade places (what k_swapPeer does to the shards)."""
    idx = np.arange(state.size)
    ba, bb = (idx >> a) & 1, (idx >> b) & 1
    src = idx ^ ((ba ^ bb) << a) ^ ((ba ^ bb) << b)
    return state[src]


@pytest.mark.parametrize("world,seed", [(2, 1), (4, 2), (8, 3)])
def test_qubit_remap_keeps_the_state_and_ends_canonical(world, seed):
    if not os.path.isfile(capi.PRODUCT_LIB):
        from quest_b200 import build
        build.build_product(2)
    lib = ctypes.CDLL(capi.PRODUCT_LIB, mode=ctypes.RTLD_LOCAL)
    lib.qb200_debugRemap.restype = ctypes.c_int
    rng = random.Random(seed)
    nq = 10
    L = nq - (world.bit_length() - 1)
    n_ops = 80
    targets, masks, mats = [], [], []
    for _ in range(n_ops):
        t = rng.randrange(nq) if rng.random() < 0.5 else rng.randrange(L, nq)
        cs = rng.sample([q for q in range(nq) if q != t], rng.choice([0, 0, 1, 2]))
        targets.append(t); masks.append(sum(1 << c for c in cs)); mats.append(rand_u(rng))
    ev = (ctypes.c_int * (3 * n_ops))()
    phys = (ctypes.c_int * nq)()
    canon = (ctypes.c_int * 128)()
    n_back = lib.qb200_debugRemap(nq, L, n_ops, (ctypes.c_int * n_ops)(*targets), (ctypes.c_ulonglong * n_ops)(*masks), ev, phys, canon)
    state = circuits.random_state(1 << nq, seed)
    want = state.copy()
    layout = list(range(nq))  # logical -> physical, tracked independently of the library
    cur = state.copy()
    n_swaps = 0
    for i in range(n_ops):
        want = dense_reference(want, nq, targets[i], masks[i], mats[i])
        g, l, pt = ev[3 * i], ev[3 * i + 1], ev[3 * i + 2]
        if g >= 0:
            assert g >= L > l >= 2
            cur = _swap_bits(cur, g, l)
            a, b = layout.index(g), layout.index(l)
            layout[a], layout[b] = l, g
            n_swaps += 1
        assert layout[targets[i]] == pt < L          # every dense gate ends up on a local slot
        pm = sum(1 << layout[c] for c in range(nq) if (masks[i] >> c) & 1)
        cur = dense_reference(cur, nq, pt, pm, mats[i])
    assert list(phys) == layout
    # far fewer exchanges than shard-crossing gates
    n_cross = sum(1 for t in targets if t >= L)
    assert 0 < n_swaps < n_cross
    for k in range(n_back):
        cur = _swap_bits(cur, canon[2 * k], canon[2 * k + 1])
    np.testing.assert_allclose(cur, want, atol=1e-12)


# ---- slab pipeline: fused pass || slot exchange (qb::planSlabs, qb::planSlotSwap) ------------------------------

def _expand(k, pos, set_mask):
    for b in pos:
        low = k & ((1 << b) - 1)
        k = ((k >> b) << (b + 1)) | low
    return k | set_mask


@pytest.mark.parametrize("local_bits,vec,local_slot,busy_extra", [(10, 1, 4, 0b1100000011), (10, 2, 6, 0b0000001111),
                                                                   (12, 4, 10, 0b110000000011), (9, 4, 2, 0b111100000),
                                                                   (11, 2, 10, 0b01111000000)])
def test_slab_exchanges_cover_the_exchange_exactly_once(local_bits, vec, local_slot, busy_extra):
    """Every pair a[j | 2^l] <-> b[j] of a slot exchange is moved by exactly one (slab, rank) launch, stays inside its
    slab, and the slab bits avoid the busy bits; the union over slabs equals the unsliced exchange."""
    lib = ctypes.CDLL(capi.PRODUCT_LIB, mode=ctypes.RTLD_LOCAL)
    lib.qb200_debugSlabs.restype = ctypes.c_int
    busy = busy_extra | (1 << local_slot)
    v = vec.bit_length() - 1
    for want in (0, 1, 2, 3):
        bits = (ctypes.c_int * 3)()
        out = (ctypes.c_ulonglong * (11 * 2 * 8))()
        nb = lib.qb200_debugSlabs(local_bits, ctypes.c_ulonglong(busy), want, v + 1, vec, local_slot, bits, out)
        slab_bits = list(bits[:nb])
        assert nb <= want and slab_bits == sorted(slab_bits)
        assert all(not (busy >> b) & 1 and v < b < local_bits for b in slab_bits)
        # the highest free bits are taken
        free = [b for b in range(local_bits - 1, v, -1) if not (busy >> b) & 1]
        assert sorted(free[:min(want, len(free))]) == slab_bits
        seen = np.zeros(1 << local_bits, dtype=np.int32)  # per b-side index j: how many launches moved it
        for s in range(1 << nb):
            slab_val = sum(((s >> j) & 1) << slab_bits[j] for j in range(nb))
            for r in range(2):
                o = out[11 * (2 * s + r):11 * (2 * s + r + 1)]
                n_tasks, n_ins, set_mask = int(o[0]), int(o[1]), int(o[2])
                pos = [int(x) for x in o[3:3 + n_ins]]
                for k in range(n_tasks):
                    vi = _expand(k, pos, set_mask)
                    for e in range(vec):
                        j = vi * vec + e
                        assert not (j >> local_slot) & 1
                        assert all(((j >> slab_bits[b]) & 1) == ((s >> b) & 1) for b in range(nb)), (j, s)
                        assert (j & sum(1 << b for b in slab_bits)) == slab_val
                        seen[j] += 1
        want_seen = np.array([0 if (j >> local_slot) & 1 else 1 for j in range(1 << local_bits)], dtype=np.int32)
        assert np.array_equal(seen, want_seen)


def test_swap_in_prefers_high_slots_and_falls_back_to_low_ones():
    """A shard-selecting qubit is exchanged with a slot >= 10 of a large shard (long contiguous runs over NVLink); only
    when every such slot is pinned by the gate itself (or hosts a guest) does the search drop to the low slots."""
    lib = ctypes.CDLL(capi.PRODUCT_LIB, mode=ctypes.RTLD_LOCAL)
    lib.qb200_debugRemap.restype = ctypes.c_int
    nq, L = 15, 13  # 4 shards of 2^13 amplitudes: preferred victims are slots 10, 11, 12

    def run(targets, masks):
        n = len(targets)
        ev, phys, canon = (ctypes.c_int * (3 * n))(), (ctypes.c_int * nq)(), (ctypes.c_int * 128)()
        lib.qb200_debugRemap(nq, L, n, (ctypes.c_int * n)(*targets), (ctypes.c_ulonglong * n)(*masks), ev, phys, canon)
        return [(ev[3 * i], ev[3 * i + 1]) for i in range(n)], list(phys)

    events, _ = run([13], [0])
    assert events[0][0] == 13 and events[0][1] >= 10
    pinned = (1 << 10) | (1 << 11) | (1 << 12)
    events, _ = run([13], [pinned])           # all preferred slots are controls of the gate
    assert events[0][0] == 13 and 0 <= events[0][1] < 10
    # one preferred slot already hosts a guest, the other two are pinned
    events, phys = run([13, 14], [0, (1 << 13) | sum(1 << s for s in (10, 11, 12) if s != 0)])
    first_victim = events[0][1]
    assert first_victim >= 10 and events[1][0] == 14 and events[1][1] not in (10, 11, 12)
    assert sorted(phys) == list(range(nq))    # still a permutation



# ---- walking a recorded queue of logical gates (qb::planWalk: sharded registers, default mode) -----------------------

def _walk(lib, nq, local_bits, ops, state):
    """ops: (target, target2, mask); state: the 193-word layout/history block carried between flushes."""
    n = len(ops)
    steps = (ctypes.c_int * (3 * (2 * n + 8)))()
    ns = lib.qb200_debugWalk(nq, local_bits, n, (ctypes.c_int * n)(*[o[0] for o in ops]), (ctypes.c_int * n)(*[o[1] for o in ops]),
                             (ctypes.c_ulonglong * n)(*[o[2] for o in ops]), state, steps)
    return [(steps[3 * k], steps[3 * k + 1], steps[3 * k + 2]) for k in range(ns)]


@pytest.mark.parametrize("world,seed", [(2, 21), (4, 22), (8, 23)])
def test_walk_of_a_logical_queue_reproduces_the_circuit(world, seed):
    """Several flushes in a row: segments cover the queue in order, every dense target is local when its gate runs,
    exchanges swap one shard-selecting slot with one local slot, and replaying exchanges + translated gates on a full
    state - followed by the canonicalising exchanges for whatever layout resulted - gives the circuit's result."""
    lib = ctypes.CDLL(capi.PRODUCT_LIB, mode=ctypes.RTLD_LOCAL)
    lib.qb200_debugWalk.restype = ctypes.c_int
    lib.qb200_debugCanon.restype = ctypes.c_int
    rng = random.Random(seed)
    nq = 11
    L = nq - (world.bit_length() - 1)
    state = (ctypes.c_ulonglong * 193)()
    psi = circuits.random_state(1 << nq, seed)
    want, cur = psi.copy(), psi.copy()
    layout = list(range(nq))  # logical -> bit position
    n_exchanges = n_cross = 0
    for flush in range(6):
        ops, mats = [], []
        for _ in range(rng.randrange(1, 60)):
            t = rng.randrange(nq) if rng.random() < 0.6 else rng.randrange(L, nq)
            cs = rng.sample([q for q in range(nq) if q != t], rng.choice([0, 0, 1, 2]))
            if rng.random() < 0.2:   # diagonal gate: no dense target
                ops.append((-1, -1, sum(1 << c for c in cs) | (1 << t)))
            else:
                ops.append((t, -1, sum(1 << c for c in cs)))
                n_cross += t >= L
            mats.append(rand_u(rng))
        steps = _walk(lib, nq, L, ops, state)
        covered = 0
        for is_x, a, b in steps:
            if is_x:
                assert a >= L > b >= 0
                cur = _swap_bits(cur, a, b)
                ia, ib = layout.index(a), layout.index(b)
                layout[ia], layout[ib] = b, a
                n_exchanges += 1
                continue
            assert a == covered
            for i in range(a, a + b):
                t, _, m = ops[i]
                pm = sum(1 << layout[c] for c in range(nq) if (m >> c) & 1)
                if t < 0:
                    idx = np.arange(1 << nq)
                    sel = (idx & pm) == pm
                    cur[sel] *= np.exp(0.3j)
                    lm = m
                    want[(idx & lm) == lm] *= np.exp(0.3j)
                else:
                    assert layout[t] < L
                    cur = dense_reference(cur, nq, layout[t], pm, mats[i])
                    want = dense_reference(want, nq, t, m, mats[i])
            covered += b
        assert covered == len(ops)
        assert [int(state[q]) for q in range(nq)] == layout
    assert 0 < n_exchanges < n_cross
    back = (ctypes.c_int * 384)()
    nb = lib.qb200_debugCanon(nq, L, (ctypes.c_int * nq)(*layout), back)
    for k in range(nb):
        assert back[2 * k] >= L > back[2 * k + 1]
        cur = _swap_bits(cur, back[2 * k], back[2 * k + 1])
        ia, ib = layout.index(back[2 * k]), layout.index(back[2 * k + 1])
        layout[ia], layout[ib] = back[2 * k + 1], back[2 * k]
    assert layout == list(range(nq))
    np.testing.assert_allclose(cur, want, atol=1e-12)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_cyclic_sweep_costs_one_exchange_per_shard_selecting_qubit_per_step(world):
    """The benchmark's step sweeps every target in turn.  Evicting the local qubit whose next use is furthest away - the
    one just swept - costs log2(G) exchanges per step; least-recently-used eviction (round 1) cost twice that."""
    lib = ctypes.CDLL(capi.PRODUCT_LIB, mode=ctypes.RTLD_LOCAL)
    lib.qb200_debugWalk.restype = ctypes.c_int
    g = world.bit_length() - 1
    L, nq = 30, 30 + g
    ops = []
    for t in range(nq):
        ops += [(t, -1, 0), (t, -1, 0), (t, -1, 1 << ((t + 1) % nq)), (-1, -1, (1 << t) | (1 << ((t + 1) % nq)) | (1 << ((t + 2) % nq)))]
    state = (ctypes.c_ulonglong * 193)()
    counts, spacing = [], []
    for step in range(12):
        steps = _walk(lib, nq, L, ops, state)
        counts.append(sum(1 for s in steps if s[0]))
        spacing.append([s[2] for s in steps if not s[0]])
    assert all(g <= c <= g + 1 for c in counts) and sum(counts) <= 12 * g + 2, counts
    # after the first step the exchanges are spread over the step: at least seven targets' worth of gates between them
    if g > 1:
        assert all(min(sp[1:-1]) >= 28 for sp in spacing[1:6]), spacing


# ---- the default sharded mode over gloo: planned walk, per-rank gate dropping, pairwise slot exchanges ---------------

def _walk_worker(rank, world, port, nq, seed, result_path):
    """One process per shard, as on the GPUs: every rank plans the same walk (qb::planWalk through qb200_debugWalk), runs
    the segments on ITS shard - dropping the gates whose controls on shard-selecting slots it does not satisfy
    (qb::localiseControls) - and performs every slot exchange with its partner rank by send/recv of the half shard that
    k_swapPeer moves.  At the end the layout is canonicalised (qb200_debugCanon) and the shards are gathered."""
    import torch
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lib = ctypes.CDLL(capi.PRODUCT_LIB, mode=ctypes.RTLD_LOCAL)
    lib.qb200_debugWalk.restype = ctypes.c_int
    lib.qb200_debugCanon.restype = ctypes.c_int
    gbits = world.bit_length() - 1
    L = nq - gbits
    n = 1 << L
    full = circuits.random_state(1 << nq, seed)
    shard = full[rank * n:(rank + 1) * n].copy()
    layout = list(range(nq))
    state = (ctypes.c_ulonglong * 193)()
    idx = np.arange(n)

    def exchange(g, l):
        gb = g - L
        partner = rank ^ (1 << gb)
        upper = ((rank >> gb) & 1) == 0
        sel = ((idx >> l) & 1) == (1 if upper else 0)  # the upper shard trades its bit-l = 1 half, the lower its bit-l = 0 half
        out = torch.from_numpy(np.ascontiguousarray(shard[sel]).view(np.float64).copy())
        buf = torch.empty_like(out)
        reqs = [dist.isend(out, partner), dist.irecv(buf, partner)]
        for r in reqs:
            r.wait()
        shard[sel] = buf.numpy().view(np.complex128)
        a, b = layout.index(g), layout.index(l)
        layout[a], layout[b] = l, g

    rng = random.Random(seed)
    all_ops = []
    for flush in range(5):
        ops, mats = [], []
        for _ in range(rng.randrange(5, 50)):
            t = rng.randrange(nq) if rng.random() < 0.55 else rng.randrange(L, nq)
            cs = rng.sample([q for q in range(nq) if q != t], rng.choice([0, 0, 1, 2]))
            if rng.random() < 0.2:
                ops.append((-1, -1, sum(1 << c for c in cs) | (1 << t)))
            else:
                ops.append((t, -1, sum(1 << c for c in cs)))
            mats.append(rand_u(rng))
        all_ops += list(zip(ops, mats))
        for is_x, a, b in _walk(lib, nq, L, ops, state):
            if is_x:
                exchange(a, b)
                continue
            for i in range(a, a + b):
                t, _, m = ops[i]
                pm = sum(1 << layout[c] for c in range(nq) if (m >> c) & 1)
                high = pm >> L
                if (rank & high) != high:
                    continue  # a control on a shard-selecting slot this rank does not satisfy: nothing to do here
                lm = pm & (n - 1)
                if t < 0:
                    shard[(idx & lm) == lm] *= np.exp(0.3j)
                else:
                    assert layout[t] < L
                    shard[:] = dense_reference(shard, L, layout[t], lm, mats[i])
        assert [int(state[q]) for q in range(nq)] == layout
    back = (ctypes.c_int * 384)()
    for k in range(lib.qb200_debugCanon(nq, L, (ctypes.c_int * nq)(*layout), back)):
        exchange(back[2 * k], back[2 * k + 1])
    assert layout == list(range(nq))
    gathered = [torch.empty(2 * n, dtype=torch.float64) for _ in range(world)] if rank == 0 else None
    dist.gather(torch.from_numpy(shard.view(np.float64).copy()), gathered, dst=0)
    if rank == 0:
        want = full.copy()
        gi = np.arange(1 << nq)
        for (t, _, m), u in all_ops:
            if t < 0:
                want[(gi & m) == m] *= np.exp(0.3j)
            else:
                want = dense_reference(want, nq, t, m, u)
        got = np.concatenate([g.numpy().view(np.complex128) for g in gathered])
        np.save(result_path, np.array([np.abs(got - want).max()]))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_planned_walk_over_gloo_reproduces_the_circuit(world, tmp_path):
    import torch.multiprocessing as mp

    nq, seed = 10, 177 + world
    port = 29700 + world + (os.getpid() % 200)
    result = str(tmp_path / "err.npy")
    mp.spawn(_walk_worker, args=(world, port, nq, seed, result), nprocs=world, join=True)
    assert np.load(result)[0] < 1e-12
