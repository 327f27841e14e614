// quest-b200 host planning logic, free of any CUDA dependency so it can be unit-tested on a CPU-only machine
// (tests/test_shard_plan.py drives it through qb200_debugPlanGate with world_size-2/4 gloo processes).
//
//   BitPlan      - which vectors a launch visits (dense task index -> vector index by bit insertion)
//   ShardedGate  - for a register sharded over G GPUs by its top log2(G) qubits (reference model:
//                  QuEST/src/CPU/QuEST_cpu_distributed.c), what THIS rank does for a one-target gate:
//                  nothing, a local pass, or its half of a pass shared with the partner rank.
#pragma once

#include <algorithm>
#include <cstring>
#include <initializer_list>
#include <utility>
#include <vector>

namespace qb {

constexpr int kMaxInsert = 8;  // target + up to 7 controls by insertion; further controls by predicate

// Insert a 0 bit at each pos[i] (ascending, vector-index space), then OR setMask.
struct BitPlan {
    int nIns;
    int pos[kMaxInsert];
    unsigned long long setMask;   // control bits forced to 1 (vector-index space)
    unsigned long long skipMask;  // overflow controls: task is skipped unless all are 1 (vector-index space)
    unsigned int loMask;          // control bits that fall inside a vector (lane space)
};

struct Planned {
    BitPlan plan;
    unsigned long long nTasks;
};

inline int log2u(unsigned long long x) {
    int l = 0;
    while ((1ull << l) < x) ++l;
    return l;
}

// zeroBits: amplitude-index bits the enumeration holds at 0 (the kernel addresses their 1-partners itself);
// oneMask : amplitude-index bits that must be 1 (controls).  Bits below log2(vec) never enter the plan: a low
// zero-bit is handled inside the vector by the kernel (TLO), low one-bits become the lane predicate loMask.
inline Planned makePlan(long long n, int vec, const std::vector<int>& zeroBits, unsigned long long oneMask) {
    const int v = log2u((unsigned long long)vec);
    Planned out;
    std::memset(&out.plan, 0, sizeof out.plan);
    std::vector<int> pos;
    for (int b : zeroBits)
        if (b >= v) pos.push_back(b - v);
    out.plan.loMask = (unsigned int)(oneMask & (unsigned long long)(vec - 1));
    for (int b = v; b < 63; ++b) {
        if (!((oneMask >> b) & 1)) continue;
        if ((int)pos.size() < kMaxInsert) {
            pos.push_back(b - v);
            out.plan.setMask |= 1ull << (b - v);
        } else {
            out.plan.skipMask |= 1ull << (b - v);
        }
    }
    std::sort(pos.begin(), pos.end());
    out.plan.nIns = (int)pos.size();
    for (int i = 0; i < out.plan.nIns; ++i) out.plan.pos[i] = pos[i];
    out.nTasks = ((unsigned long long)n / vec) >> out.plan.nIns;
    return out;
}

inline Planned makePlan(long long n, int vec, std::initializer_list<int> zeroBits, unsigned long long oneMask) {
    return makePlan(n, vec, std::vector<int>(zeroBits), oneMask);
}

// host twin of the device-side qb::expand (qb_device.cuh)
inline unsigned long long expandHost(unsigned long long k, const BitPlan& p) {
    for (int i = 0; i < p.nIns; ++i) {
        const int b = p.pos[i];
        const unsigned long long low = k & ((1ull << b) - 1ull);
        k = ((k >> b) << (b + 1)) | low;
    }
    return k | p.setMask;
}

// Controls on qubits that select the shard cost nothing: this rank either satisfies them (the bits are dropped)
// or holds no selected amplitude (nothing to do).  Mirrors the global-index tests of QuEST_cpu.c:1841,2694.
inline bool localiseControls(int localBits, int chunkId, unsigned long long& mask) {
    const unsigned long long globalBits = mask >> localBits;
    if (globalBits) {
        if (((unsigned long long)chunkId & globalBits) != globalBits) return false;
        mask &= (1ull << localBits) - 1ull;
    }
    return true;
}

enum ShardAction { kShardSkip = 0, kShardLocal = 1, kShardCross = 2 };

struct ShardedGate {
    int action;           // ShardAction
    int partner;          // kShardCross: rank holding the other half of every pair (chunkId XOR 2^(t-L))
    int amUpper;          // kShardCross: this rank's amplitudes are the target-bit-0 halves
    int launches;         // kShardCross: 0 if this rank has no pairs to process (all local bits are controls)
    unsigned long long ctrlMask;  // local control bits that remain
    Planned planned;      // kShardLocal: plan over {target}; kShardCross: plan over this rank's half of the pairs
};

// local iff 2^target < shard size (halfMatrixBlockFitsInChunk, QuEST_cpu_distributed.c:360-365); otherwise the
// pairs are split between the two ranks on the highest local bit that is not a control, the upper rank
// (chunkIsUpper, :231-237) taking the half where that bit is 0.
inline ShardedGate planShardedGate(int localBits, int chunkId, int target, unsigned long long ctrlMask, int vec) {
    ShardedGate g;
    std::memset(&g, 0, sizeof g);
    const long long n = 1LL << localBits;
    if (!localiseControls(localBits, chunkId, ctrlMask)) {
        g.action = kShardSkip;
        return g;
    }
    g.ctrlMask = ctrlMask;
    if (target < localBits) {
        g.action = kShardLocal;
        g.planned = makePlan(n, vec, {target}, ctrlMask);
        return g;
    }
    g.action = kShardCross;
    const int gbit = target - localBits;
    g.partner = chunkId ^ (1 << gbit);
    g.amUpper = ((chunkId >> gbit) & 1) == 0;
    const int v = log2u((unsigned long long)vec);
    int split = -1;
    for (int b = localBits - 1; b >= v; --b)
        if (!((ctrlMask >> b) & 1)) {
            split = b;
            break;
        }
    if (split >= 0) {
        g.planned = makePlan(n, vec, {split}, ctrlMask);
        if (!g.amUpper) g.planned.plan.setMask |= 1ull << (split - v);
        g.launches = g.planned.nTasks ? 1 : 0;
    } else {
        g.planned = makePlan(n, vec, {}, ctrlMask);
        g.launches = (g.amUpper && g.planned.nTasks) ? 1 : 0;
    }
    return g;
}

}  // namespace qb

// ---------------------------------------------------------------------------------------------------------
// Qubit remapping for sharded registers.
//
// A dense gate on a qubit that selects the GPU needs the partner's amplitudes: 16 bytes per amplitude pair over
// NVLink for EVERY such gate.  Instead, the first gate on such a qubit swaps it with a local qubit - an exchange of
// half a shard, 8 bytes per amplitude - after which that qubit is local: its gates join the fusion queue and cost
// no communication until a gate needs the evicted qubit.  QubitMap records where each logical qubit currently
// lives.  Invariant: the map is a product of disjoint transpositions (global slot g <-> local slot l), so undoing
// it is replaying the same swaps.  All ranks run this logic on identical inputs and stay consistent.
// ---------------------------------------------------------------------------------------------------------
namespace qb {

struct QubitMap {
    int nQubits = 0, localBits = 0;
    int phys[64];                   // logical qubit -> bit position in the amplitude index
    int logicalAt[64];              // bit position -> logical qubit
    unsigned long long lastUse[64]; // logical qubit -> tick of its last use as a dense target
    unsigned long long prevUse[64]; // ... and of the use before that (0 = none): their distance predicts the next use
    unsigned long long tick = 0;

    void reset(int n, int local) {
        nQubits = n;
        localBits = local;
        tick = 0;
        for (int q = 0; q < 64; ++q) {
            phys[q] = logicalAt[q] = q;
            lastUse[q] = prevUse[q] = 0;
        }
    }
    bool identity() const {
        for (int q = 0; q < nQubits; ++q)
            if (phys[q] != q) return false;
        return true;
    }
    unsigned long long physMask(unsigned long long logicalMask) const {
        unsigned long long m = 0;
        while (logicalMask) {
            const int q = __builtin_ctzll(logicalMask);
            logicalMask &= logicalMask - 1;
            m |= 1ull << phys[q];
        }
        return m;
    }
    unsigned long long physIndex(unsigned long long logicalIndex) const {
        unsigned long long p = 0;
        for (int q = 0; q < nQubits; ++q)
            if ((logicalIndex >> q) & 1) p |= 1ull << phys[q];
        return p;
    }
    void touch(int logicalQubit) {
        prevUse[logicalQubit] = lastUse[logicalQubit];
        lastUse[logicalQubit] = ++tick;
    }
    // When will this qubit be a dense target again?  Its last two uses predict the next one (a qubit swept every N gates
    // comes back N gates after its last use; a qubit used once has no prediction).
    unsigned long long predictedNextUse(int logicalQubit) const {
        if (!prevUse[logicalQubit]) return ~0ull;
        return lastUse[logicalQubit] + (lastUse[logicalQubit] - prevUse[logicalQubit]);
    }
    void applySwap(int g, int l) {  // exchange the occupants of bit positions g and l
        const int a = logicalAt[g], b = logicalAt[l];
        logicalAt[g] = b;
        logicalAt[l] = a;
        phys[a] = l;
        phys[b] = g;
    }
};

struct SlotSwap {
    int globalSlot, localSlot;  // bit positions to exchange; globalSlot < 0: nothing to do
};

// Which exchange makes logical qubit q local?  `avoidMask` = logical qubits that should not be evicted (the
// gate's own controls).  (a) q sits in a global slot because it was evicted earlier: undo that transposition.
// (b) q is global-native in its own slot: evict the least recently used local-native qubit that is at home.
inline SlotSwap planSwapIn(const QubitMap& m, int q, unsigned long long avoidMask) {
    SlotSwap s{-1, -1};
    const int p = m.phys[q];
    if (p < m.localBits) return s;
    if (q < m.localBits) {  // (a) q's home slot is local and currently holds the qubit native to slot p
        s.globalSlot = p;
        s.localSlot = q;
        return s;
    }
    // (b) Low slots are never evicted: exchanging slot l moves runs of 2^l contiguous amplitudes, and NVLink needs
    // runs of a few KB to reach its rate (measured: slot 2 -> 404 GB/s per direction, slot >= 10 -> ~700 GB/s).
    // Only when every preferred slot is pinned or already hosts a guest does the search drop to the low slots.
    int best = -1;
    const int preferred = m.localBits > 12 ? 10 : 2;
    for (int lowest = preferred; best < 0 && lowest >= 0; lowest = lowest ? 0 : -1)
        for (int x = lowest; x < m.localBits; ++x) {
            if (m.phys[x] != x || ((avoidMask >> x) & 1)) continue;
            if (best < 0 || m.lastUse[x] < m.lastUse[best] || (m.lastUse[x] == m.lastUse[best] && x > best)) best = x;
        }
    if (best < 0) return s;  // every local qubit is pinned: caller falls back to the cross-shard gate
    s.globalSlot = p;
    s.localSlot = best;
    return s;
}

// Slabs: a pass and the slot exchange that follows it can be pipelined if both are cut along the same index bits -
// slab s = the amplitudes whose slab bits hold the value s.  The exchange of slab s needs only the pass on slab s
// (on both GPUs of the pair), so it runs over NVLink while the pass computes slab s+1.  Slab bits are the highest
// local bits that the pass does not hold in registers (`busyMask`) and that are not the exchange's local slot.
struct SlabPlan {
    int nBits;
    int bits[3];  // ascending
    unsigned long long valueMask(int slab) const {
        unsigned long long m = 0;
        for (int j = 0; j < nBits; ++j)
            if ((slab >> j) & 1) m |= 1ull << bits[j];
        return m;
    }
};

inline SlabPlan planSlabs(int localBits, unsigned long long busyMask, int want, int lowestAllowed) {
    SlabPlan p{0, {0, 0, 0}};
    int found[3];
    for (int b = localBits - 1; b >= lowestAllowed && p.nBits < want && p.nBits < 3; --b)
        if (!((busyMask >> b) & 1)) found[p.nBits++] = b;
    for (int j = 0; j < p.nBits; ++j) p.bits[j] = found[p.nBits - 1 - j];
    return p;
}

// This rank's k_swapPeer launch for exchanging local slot l with a shard-selecting slot: the pairs
// a[j | 2^l] <-> b[j] (j with bit l clear) are divided between the two GPUs on `split`, the highest local bit that
// is neither l nor a slab bit; the upper rank (shard-selecting bit 0) takes split = 0.  With a slab plan only the
// pairs inside slab `slab` are enumerated.
struct SwapLaunch {
    bool active;
    Planned planned;
};

inline SwapLaunch planSlotSwap(int localBits, int vec, int l, bool amUpper, const SlabPlan* sp, int slab) {
    SwapLaunch out;
    out.active = false;
    std::memset(&out.planned, 0, sizeof out.planned);
    const int v = log2u((unsigned long long)vec);
    unsigned long long slabBits = 0;
    if (sp)
        for (int j = 0; j < sp->nBits; ++j) slabBits |= 1ull << sp->bits[j];
    int split = -1;
    for (int b = localBits - 1; b >= v; --b)
        if (b != l && !((slabBits >> b) & 1)) {
            split = b;
            break;
        }
    if (split < 0 && !amUpper) return out;
    std::vector<int> zero{l};
    if (split >= 0) zero.push_back(split);
    if (sp)
        for (int j = 0; j < sp->nBits; ++j) zero.push_back(sp->bits[j]);
    out.planned = makePlan(1LL << localBits, vec, zero, 0);
    if (split >= 0 && !amUpper) out.planned.plan.setMask |= 1ull << (split - v);
    if (sp) out.planned.plan.setMask |= sp->valueMask(slab) >> v;
    out.active = out.planned.nTasks != 0;
    return out;
}

// Global<->local slot exchanges that restore the identity layout, for ANY placement the exchanges have produced.
// Phase 1 sends every shard-selecting qubit home (one exchange if it sits in a local slot, two if it sits in another
// shard-selecting slot); phase 2 resolves what is left - a permutation of the local qubits among the local slots - cycle
// by cycle through the first shard-selecting slot as a buffer (k + 1 exchanges for a cycle of k slots).  In the common
// case, a few disjoint global<->local transpositions, this is one exchange per transposition.
inline int planCanonicalize(const QubitMap& m0, SlotSwap* out /* >= 192 */) {
    QubitMap m = m0;
    int n = 0;
    auto swap = [&](int g, int l) {
        out[n++] = SlotSwap{g, l};
        m.applySwap(g, l);
    };
    for (int h = m.localBits; h < m.nQubits; ++h) {
        if (m.logicalAt[h] == h) continue;
        int at = m.phys[h];  // where the native of slot h sits
        if (at >= m.localBits) {  // in another shard-selecting slot: bring it into a local slot first
            swap(at, m.localBits - 1);
            at = m.localBits - 1;
        }
        swap(h, at);
    }
    if (m.nQubits > m.localBits) {
        const int g0 = m.localBits;
        for (int l = 0; l < m.localBits; ++l) {
            if (m.logicalAt[l] == l) continue;
            swap(g0, l);  // slot l now holds the native of g0, the buffer holds a displaced local qubit
            while (m.logicalAt[g0] != g0) swap(g0, m.logicalAt[g0]);  // ... which goes home, handing the buffer the next one
        }
    }
    return n;
}

// ---- walking a recorded queue of LOGICAL gates (sharded registers, default mode) ---------------------------------------
// The queue (the same on every rank) is cut into SEGMENTS of gates whose dense targets are all local under the layout
// of that moment, separated by slot exchanges that bring the next needed shard-selecting qubit in.  The victim of an
// exchange is the local qubit whose next use as a dense target lies furthest away: known from the rest of the queue
// where it is used again, predicted from its last two uses where it is not (a qubit swept every N gates comes back N
// gates after its last use), least recently used where there is no history.  Pure bookkeeping: the executor replays the
// steps on the GPUs.
struct WalkStep {
    int isExchange;
    int first, count;  // segment: gates [first, first + count) of the queue
    SlotSwap sw;       // exchange
};

// `skip`: take the (skip+1)-th best candidate of the best class instead of the best one.  The walk asks for that when
// exchanges follow each other within a few gates (a run of gates on several shard-selecting qubits): spreading the victims
// over qubits that were last used at different times spreads the NEXT round of exchanges over the program, where each
// one can hide behind the passes on both sides of it instead of queueing up behind the previous exchange.
inline int chooseVictimSlot(const QubitMap& m, const unsigned long long* nextUse /* per logical qubit, ~0 = none */,
                            unsigned long long keepLogical, int skip = 0) {
    const int preferred = m.localBits > 12 ? 10 : 2;  // low slots move short runs over NVLink (see planSwapIn)
    for (int lowest = preferred; lowest >= 0; lowest = lowest ? 0 : -1) {
        int bestClass = -1;
        std::vector<std::pair<unsigned long long, int>> ranked;  // (key, slot) of the best class
        for (int x = lowest; x < m.localBits; ++x) {
            const int lq = m.logicalAt[x];
            if ((keepLogical >> lq) & 1) continue;
            int cls;
            unsigned long long key;
            if (nextUse[lq] != ~0ull) {                 // used again in this queue: the later the better
                cls = 0;
                key = nextUse[lq];
            } else if (m.lastUse[lq] == 0) {            // never a dense target so far
                cls = 3;
                key = (unsigned long long)x;
            } else if (m.prevUse[lq] == 0) {            // used once: no prediction, least recently used first
                cls = 1;
                key = ~m.lastUse[lq];
            } else {                                    // predicted from its last two uses: the later the better
                cls = 2;
                key = m.predictedNextUse(lq);
            }
            if (cls > bestClass) {
                bestClass = cls;
                ranked.clear();
            }
            if (cls == bestClass) ranked.push_back({key, x});
        }
        if (ranked.empty()) continue;
        std::sort(ranked.begin(), ranked.end(), [](const std::pair<unsigned long long, int>& a, const std::pair<unsigned long long, int>& b) {
            return a.first != b.first ? a.first > b.first : a.second > b.second;
        });
        const int pick = bestClass == 0 ? 0 : std::min(skip, (int)ranked.size() - 1);  // never give up a known-later use
        return ranked[pick].second;
    }
    return -1;
}

struct WalkOp {  // what the walk needs to know about a recorded gate (logical qubits)
    int target, target2;          // dense targets (-1 = none; both -1 for a diagonal gate)
    unsigned long long mask;
};

// `map` is advanced to the layout (and use history) after the whole queue.  A gate whose dense target cannot be made
// local (every local qubit is pinned by the gate itself) ends up in a segment of its own with that target still in a
// shard-selecting slot: the executor runs it as a cross-shard pass.
inline std::vector<WalkStep> planWalk(QubitMap& map, const std::vector<WalkOp>& ops) {
    std::vector<WalkStep> steps;
    const int n = (int)ops.size();
    // next dense use of every logical qubit at or after position i, filled backwards
    std::vector<unsigned long long> next((size_t)(n + 1) * 64, ~0ull);
    for (int i = n - 1; i >= 0; --i) {
        for (int q = 0; q < 64; ++q) next[(size_t)i * 64 + q] = next[(size_t)(i + 1) * 64 + q];
        if (ops[i].target >= 0) next[(size_t)i * 64 + ops[i].target] = (unsigned long long)i;
        if (ops[i].target2 >= 0) next[(size_t)i * 64 + ops[i].target2] = (unsigned long long)i;
    }
    int segStart = 0;
    int lastExchangeAt = -1000000, closeRun = 0;  // exchanges that follow each other within kCloseGates gates
    constexpr int kCloseGates = 16, kSpread = 7;
    for (int i = 0; i < n; ++i) {
        const int t[2] = {ops[i].target, ops[i].target2};
        unsigned long long keep = ops[i].mask;
        for (int k = 0; k < 2; ++k)
            if (t[k] >= 0) keep |= 1ull << t[k];
        bool cross = false;
        for (int k = 0; k < 2; ++k) {
            if (t[k] < 0 || map.phys[t[k]] < map.localBits) continue;
            closeRun = (i - lastExchangeAt <= kCloseGates) ? closeRun + 1 : 0;
            const int slot = map.localBits >= 3 ? chooseVictimSlot(map, &next[(size_t)(i + 1) * 64], keep, closeRun * kSpread) : -1;
            if (slot < 0) {
                cross = true;
                continue;
            }
            if (i > segStart) steps.push_back(WalkStep{0, segStart, i - segStart, SlotSwap{-1, -1}});
            segStart = i;
            lastExchangeAt = i;
            const SlotSwap sw{map.phys[t[k]], slot};
            steps.push_back(WalkStep{1, 0, 0, sw});
            map.applySwap(sw.globalSlot, sw.localSlot);
        }
        if (cross) {  // isolate the gate
            if (i > segStart) steps.push_back(WalkStep{0, segStart, i - segStart, SlotSwap{-1, -1}});
            steps.push_back(WalkStep{0, i, 1, SlotSwap{-1, -1}});
            segStart = i + 1;
        }
        for (int k = 0; k < 2; ++k)
            if (t[k] >= 0) map.touch(t[k]);
    }
    if (n > segStart) steps.push_back(WalkStep{0, segStart, n - segStart, SlotSwap{-1, -1}});
    return steps;
}

}  // namespace qb
