// quest-b200 gate fusion: host-side queue and grouping (CUDA-free, unit-tested on CPU through qb200_debugFuse).
//
// Every gate of the QuEST API is one pass over HBM (16 GiB read + 16 GiB written at 30 qubits) however little
// arithmetic it does.  The API lets a gate call return before the GPU has run it (the reference GPU backend
// launches asynchronously, QuEST_gpu.cu:591), so gate calls are recorded in a per-register queue and executed at
// the next call that needs the amplitudes (any value-returning call, a copy, syncQuESTEnv, a shard-crossing gate).
// At that point consecutive gates are grouped: a group is a run of gates, in program order, whose dense (non-
// diagonal) targets fall on at most kFuseBits distinct qubits; diagonal gates and control qubits are unrestricted.
// One launch of k_fused then loads 2^kFuseBits amplitudes per thread - the sub-block spanned by those qubits -
// applies the whole run in registers, gate after gate in the original order with the original per-gate
// arithmetic, and writes the block back: one HBM pass for the whole group, bit-identical results.
#pragma once

#include <algorithm>
#include <cstring>
#include <vector>

#include "QuEST_precision.h"
#include "qb_plan.h"

namespace qb {

#ifndef QB_FUSE_BITS
#define QB_FUSE_BITS 4
#endif
constexpr int kFuseBits = QB_FUSE_BITS;  // qubits resolved in registers per launch (2^kFuseBits amplitudes per thread)
constexpr int kMaxFusedGates = 96;  // gates per launch: descriptors travel as kernel parameters (constant bank)

enum FusedKind : int { kFMatrix = 0, kFHadamard = 1, kFSwap = 2, kFPauliY = 3, kFPhase = 4, kFFlip = 5,
                       // one-qubit noise channels of a density matrix: act on the 2x2 block structure spanned by the
                       // qubit's row bit and column bit (two dense targets); symmetric under exchanging the two
                       kFDephase = 6, kFDepolarise = 7, kFDamping = 8,
                       // density-matrix collapse: block (o,o) of the (row,column) structure /= c[0], other blocks = 0; c[1] = o
                       kFDensCollapse = 9 };

inline bool isDiagonal(int kind) { return kind == kFPhase || kind == kFFlip; }
inline bool isNoise(int kind) { return kind >= kFDephase; }

// A recorded gate: target and control mask are LOCAL bit positions of this rank's shard (controls on
// shard-selecting qubits were already resolved when the gate was recorded).
struct PendingOp {
    int kind;
    int target;                    // dense kinds only
    int target2;                   // noise kinds: the second bit of the (row, column) pair; otherwise -1
    unsigned long long mask;       // dense: control mask; diagonal: all selecting bits
    qreal c[8];                    // Matrix: r00,i00,r01,i01,r10,i10,r11,i11; Hadamard: c[0]=1/sqrt2;
                                   // PauliY: c[0]=+-1; Phase: c[0]=cos, c[1]=sin; noise: c[0]=factor on
                                   // coherences, c[1]=level, c[2]=1-level (see k_noise1)
};

// What the kernel reads per gate.
struct FusedGate {
    int kind;
    int pos;                       // dense: which of the kFuseBits register bits is the target (noise: the lower one)
    unsigned int regMask;          // control/selecting bits that are register bits (bit j = j-th fused qubit)
    int pos2;                      // noise: register bit of the higher of the two targets
    unsigned long long outerMask;  // control/selecting bits outside the fused qubits (amplitude-index space)
    qreal c[8];
};

struct FusedGroup {
    int first, count;              // ops [first, first+count) of the queue
    int bits[kFuseBits];           // fused qubits, ascending; padded with unused qubits when the group needs fewer
    int nDenseBits;                // how many of them are targets of dense gates
};

// Greedy, order-preserving grouping.
// `avoidMask`: bits that must not be used to pad a group (they are held out for slab pipelining, qb_plan.h); they
// never hold a dense target of these ops.
inline std::vector<FusedGroup> groupOps(const std::vector<PendingOp>& ops, int localBits, unsigned long long avoidMask = 0) {
    std::vector<FusedGroup> out;
    const int n = (int)ops.size();
    int i = 0;
    while (i < n) {
        FusedGroup g;
        g.first = i;
        g.count = 0;
        g.nDenseBits = 0;
        unsigned long long chosen = 0;
        while (i < n && g.count < kMaxFusedGates) {
            const PendingOp& op = ops[i];
            if (!isDiagonal(op.kind)) {
                unsigned long long need = 1ull << op.target;
                if (op.target2 >= 0) need |= 1ull << op.target2;
                need &= ~chosen;
                const int extra = __builtin_popcountll(need);
                if (g.nDenseBits + extra > kFuseBits) break;
                chosen |= need;
                g.nDenseBits += extra;
            }
            ++g.count;
            ++i;
        }
        // pad with qubits the group's gates mention (keeps their predicates in registers), then with the
        // highest unused qubits (keeps the remaining index bits low => consecutive threads, consecutive amplitudes)
        int have = g.nDenseBits;
        for (int k = g.first; k < g.first + g.count && have < kFuseBits; ++k) {
            unsigned long long m = ops[k].mask & ~chosen;
            while (m && have < kFuseBits) {
                const int b = __builtin_ctzll(m);
                m &= m - 1;
                if (b < localBits && !((avoidMask >> b) & 1)) {
                    chosen |= 1ull << b;
                    ++have;
                }
            }
        }
        for (int b = localBits - 1; b >= 0 && have < kFuseBits; --b)
            if (!((chosen >> b) & 1) && !((avoidMask >> b) & 1)) {
                chosen |= 1ull << b;
                ++have;
            }
        int j = 0;
        for (int b = 0; b < 64 && j < kFuseBits; ++b)
            if ((chosen >> b) & 1) g.bits[j++] = b;
        out.push_back(g);
    }
    return out;
}

// Translate one queued op into the kernel's descriptor for a group with fused qubits `bits`.
inline FusedGate describe(const PendingOp& op, const int* bits) {
    FusedGate fg;
    std::memset(&fg, 0, sizeof fg);
    fg.kind = op.kind;
    fg.pos = 0;
    fg.pos2 = 0;
    unsigned long long outer = op.mask;
    const int lowT = op.target2 >= 0 && op.target2 < op.target ? op.target2 : op.target;
    const int highT = op.target2 >= 0 ? (op.target2 < op.target ? op.target : op.target2) : -1;
    for (int j = 0; j < kFuseBits; ++j) {
        if (!isDiagonal(op.kind) && bits[j] == lowT) fg.pos = j;
        if (bits[j] == highT) fg.pos2 = j;
        if ((op.mask >> bits[j]) & 1) {
            fg.regMask |= 1u << j;
            outer &= ~(1ull << bits[j]);
        }
    }
    fg.outerMask = outer;
    for (int k = 0; k < 8; ++k) fg.c[k] = op.c[k];
    return fg;
}

// ---- walking a queue of LOGICAL gates (distributed mode of sharded registers) ------------------------------------
// The queue holds gates in logical qubits with full control masks, the same on every rank.  It is cut, in program
// order, into groups (runs of gates on at most kFuseBits dense qubits); before each group the dense qubits that sit in
// shard-selecting slots are either left there (the group runs as a distributed pass) or exchanged with a local slot
// first.  With `lookahead` a qubit is made local iff a later gate of the queue uses it as a dense target again; without,
// never.  Pure bookkeeping on a copy of the qubit map: the executor replays the same exchanges on the real one.
struct WalkStep {
    int first, count;   // gates [first, first + count) of the queue form the group ...
    int nSwaps;
    SlotSwap swaps[kFuseBits];  // ... which runs after these slot exchanges
    unsigned long long dense;   // its dense logical qubits
};

inline std::vector<WalkStep> planWalk(QubitMap map, const std::vector<PendingOp>& ops, bool lookahead) {
    std::vector<WalkStep> steps;
    const int n = (int)ops.size();
    int i = 0;
    while (i < n) {
        WalkStep st;
        st.first = i;
        st.nSwaps = 0;
        st.dense = 0;
        int j = i;
        while (j < n && j - i < kMaxFusedGates) {
            const PendingOp& op = ops[j];
            if (!isDiagonal(op.kind)) {
                unsigned long long need = 1ull << op.target;
                if (op.target2 >= 0) need |= 1ull << op.target2;
                if (__builtin_popcountll(st.dense | need) > kFuseBits) break;
                st.dense |= need;
            }
            ++j;
        }
        st.count = j - i;
        for (unsigned long long m = st.dense; m; m &= m - 1) {
            const int lq = __builtin_ctzll(m);
            if (lookahead && map.phys[lq] >= map.localBits) {
                bool usedLater = false;
                for (int k = j; k < n && !usedLater; ++k)
                    usedLater = !isDiagonal(ops[k].kind) && (ops[k].target == lq || ops[k].target2 == lq);
                if (usedLater) {
                    const SlotSwap sw = planSwapIn(map, lq, st.dense);
                    if (sw.globalSlot >= 0) {
                        map.applySwap(sw.globalSlot, sw.localSlot);
                        st.swaps[st.nSwaps++] = sw;
                    }
                }
            }
            map.touch(lq);
        }
        steps.push_back(st);
        i = j;
    }
    return steps;
}

// ---- device dispatch ---------------------------------------------------------------------------------------
// The kernel resolves a gate with ONE table jump: kind, target position (dense gates), register mask (diagonal
// gates) or position pair (noise) are folded into a dense opcode on the host.
enum FusedOpcode : int {
    kOpMatrix = 0,      // +pos           uncontrolled among the fused qubits: straight-line arithmetic on all pairs
    kOpMatrixCtl = 4,   // +pos           some control is a fused qubit: per-pair predicate on regMask
    kOpHadamard = 8,    // +pos
    kOpSwap = 12,       // +pos
    kOpPauliY = 16,     // +pos           (20..31 unused: every family starts on a multiple of 4, diagonal ones of 16)
    kOpPhase = 32,      // +regMask (16)
    kOpFlip = 48,       // +regMask (16)
    kOpNoise = 64,      // +(kind - kFDephase) * 6 + pairSlot(pos, pos2)
    kOpCount = 88
};

// (p, q), p < q < 4  ->  0..5
inline int pairSlot(int p, int q) { return p == 0 ? q - 1 : (p == 1 ? q + 1 : 5); }

inline int opcodeOf(const FusedGate& g) {
    switch (g.kind) {
        case kFMatrix: return (g.regMask ? kOpMatrixCtl : kOpMatrix) + g.pos;
        case kFHadamard: return kOpHadamard + g.pos;
        case kFSwap: return kOpSwap + g.pos;
        case kFPauliY: return kOpPauliY + g.pos;
        case kFPhase: return kOpPhase + (int)(g.regMask & 15u);
        case kFFlip: return kOpFlip + (int)(g.regMask & 15u);
        default: return kOpNoise + (g.kind - kFDephase) * 6 + pairSlot(g.pos, g.pos2);
    }
}

// Kernel tier bits (qb_kernels.cuh: FusedTier) this opcode needs: bit 0 = the less common state-vector gates,
// bit 1 = density-matrix noise.  A group runs on the OR over its gates.
inline int tierOf(int opcode) {
    if (opcode >= kOpNoise) return 2;
    if ((opcode >= kOpMatrixCtl && opcode < kOpHadamard) || (opcode >= kOpPauliY && opcode < kOpPhase) || opcode >= kOpFlip) return 1;
    return 0;
}

// What the kernel reads per gate: 8 bytes of control, the outer-bit condition and 8 coefficients, in separate arrays:
// the conditions of 32 gates are folded into one bit mask before the gate loop, and the control word of the NEXT
// gate is fetched while the current one computes.
struct GateCtl {
    int opcode;
    unsigned int regMask;
};

}  // namespace qb
