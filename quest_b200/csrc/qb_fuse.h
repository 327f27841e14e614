// quest-b200 gate fusion: host-side queue and pass planning (CUDA-free, unit-tested on CPU through qb200_debugTilePlan).
//
// Every gate of the QuEST API is one pass over HBM (16 GiB read + 16 GiB written at 30 qubits) however little
// arithmetic it does.  The API lets a gate call return before the GPU has run it (the reference GPU backend
// launches asynchronously, QuEST_gpu.cu:591), so gate calls are recorded in a per-register queue and executed at
// the next call that needs the amplitudes (any value-returning call, a copy, syncQuESTEnv, a shard-crossing gate).
// At that point the queue is cut, in program order, into PASSES and each pass into SUB-GROUPS:
//
//   pass       a run of gates whose dense (non-diagonal) targets, together with the lowest kTileLow index bits, fall
//              on at most kTileBits distinct index bits - the TILE BITS.  One launch of k_tile (qb_tile.cuh) serves
//              the pass: every CTA stages the 2^kTileBits amplitudes that differ only in the tile bits in shared
//              memory (rows of 2^kTileLow contiguous amplitudes, so HBM sees whole 128-byte lines), runs all gates
//              of the pass on them, and writes them back - ONE read and ONE write of the state for the whole pass.
//   sub-group  a run of gates inside a pass whose dense targets fall on at most kSubBits tile bits.  Each thread
//              loads the 2^kSubBits amplitudes spanned by those bits from shared memory into registers, applies the
//              sub-group's gates in program order with the per-gate arithmetic of the single-gate kernels, and
//              stores them back; the next sub-group re-blocks the tile along other bits.
//
// Diagonal gates and control qubits never restrict a pass: a selecting bit outside the tile is a per-CTA condition,
// one inside the tile a per-thread condition, one among the sub-group's register bits a per-element condition.
// Results are bit-identical to gate-at-a-time execution.
#pragma once

#include <algorithm>
#include <cstring>
#include <vector>

#include "QuEST_precision.h"
#include "qb_plan.h"

namespace qb {

#ifndef QB_TILE_BITS
#define QB_TILE_BITS 11
#endif
constexpr int kTileBits = QB_TILE_BITS;  // amplitudes staged per CTA: 2^kTileBits (2 x 16 KiB of shared memory in double)
constexpr int kTileLow = 4;              // the lowest index bits are always tile bits: 2^4 doubles = one 128-byte line
constexpr int kSubBits = 3;              // dense-target qubits resolved in registers per sub-group ...
constexpr int kSpectBits = 1;            // ... plus one SPECTATOR tile bit no gate of the sub-group touches: a thread holds
                                         // 2^(3+1) = 16 amplitudes and every dispatched gate body serves two 8-blocks
constexpr int kMaxTouched = 8;           // a gate touching more index bits than this (target + controls) runs on its own
constexpr int kMaxTileGates = 192;       // gates per pass: descriptors travel as kernel parameters (constant bank)
constexpr int kMaxTileSubs = 64;         // sub-groups per pass
constexpr int kMaxSubGates = 32;         // gates per sub-group: their per-CTA activity is one 32-bit word
constexpr int kMaxSlabBits = 3;          // a launch may be confined to the amplitudes with given values on up to 3 bits

enum FusedKind : int { kFMatrix = 0, kFHadamard = 1, kFSwap = 2, kFPauliY = 3, kFPhase = 4, kFFlip = 5,
                       // one-qubit noise channels of a density matrix: act on the 2x2 block structure spanned by the
                       // qubit's row bit and column bit (two dense targets); symmetric under exchanging the two
                       kFDephase = 6, kFDepolarise = 7, kFDamping = 8,
                       // density-matrix collapse: block (o,o) of the (row,column) structure /= c[0], other blocks = 0; c[1] = o
                       kFDensCollapse = 9,
                       // made by the planner, never recorded: an uncontrolled hadamard immediately followed by an uncontrolled
                       // matrix gate on the same qubit, executed pair by pair with one dispatch (c = the matrix)
                       kFHadMatrix = 10 };

inline bool isDiagonal(int kind) { return kind == kFPhase || kind == kFFlip; }
inline bool isNoise(int kind) { return kind >= kFDephase && kind <= kFDensCollapse; }

struct PendingOp;
inline std::vector<PendingOp> fuseAdjacentPairs(const std::vector<PendingOp>& ops);

// A recorded gate: target and control mask are LOCAL bit positions of this rank's shard (controls on
// shard-selecting qubits were already resolved when the gate was recorded).
struct PendingOp {
    int kind;
    int target;                    // dense kinds only
    int target2;                   // noise kinds: the second bit of the (row, column) pair; otherwise -1
    unsigned long long mask;       // dense: control mask; diagonal: all selecting bits
    qreal c[9];                    // Matrix: r00,i00,r01,i01,r10,i10,r11,i11; Hadamard: c[0]=1/sqrt2; HadMatrix: matrix + c[8]=1/sqrt2;
                                   // PauliY: c[0]=+-1; Phase: c[0]=cos, c[1]=sin; noise: c[0]=factor on
                                   // coherences, c[1]=level, c[2]=1-level (see k_noise1)
};

inline unsigned long long denseBitsOf(const PendingOp& op) {
    if (isDiagonal(op.kind)) return 0;
    unsigned long long m = 1ull << op.target;
    if (op.target2 >= 0) m |= 1ull << op.target2;
    return m;
}

struct TileSub {
    int first, count;              // positions [first, first+count) of the pass's `order`
    int rb[kSubBits];              // register bits, ascending, as positions in TILE-INDEX space (bit j of a tile index
                                   // is amplitude-index bit tileBits[j])
    int spect;                     // spectator bit (tile-index space): no gate of the sub-group has it as target or in its mask
};

struct TilePass {
    std::vector<int> order;        // the pass's gates: indices into the queue, in execution order
    int count;                     // = order.size()
    int nTile;                     // number of tile bits = min(kTileBits, localBits)
    int tileBits[kTileBits];       // amplitude-index bit positions, ascending
    unsigned long long denseMask;  // amplitude-index bits that are dense targets of the pass
    std::vector<TileSub> subs;
};

// Two recorded gates commute when, on every index bit they share, both act diagonally (a control or phase-mask bit; a
// dense target never does): then both are block-diagonal in that bit and their blocks act on disjoint qubits.  For a
// candidate gate against a SET of gates this is two mask tests against the set's united dense bits and united bits.
struct Footprint {
    unsigned long long dense, all;  // dense targets; dense targets + mask
};
inline Footprint footprintOf(const PendingOp& op) {
    const unsigned long long d = denseBitsOf(op);
    return Footprint{d, d | op.mask};
}
inline bool commutesWithAll(const Footprint& f, unsigned long long setDense, unsigned long long setAll) {
    return !(f.dense & setAll) && !(f.all & setDense);
}

// Pass planning.  Program order is kept wherever it matters: a pass takes the gates of the queue in order, skipping a gate
// that does not fit (its dense targets would overflow the tile) - and a later gate may still join the pass if it
// commutes with every gate skipped before it.  Layered circuits (a hadamard on every qubit, then a rotation on every
// qubit, then a chain of controlledNots ...) thus run qubit block by qubit block instead of layer by layer: the passes of
// BASELINE configs[0] drop from 12 to 4 per layer, those of the noisy density-matrix configs[3] from 9 to 5.
// `avoidMask`: amplitude-index bits that must not become tile bits by padding (they are held out for slab pipelining,
// qb_plan.h); the caller guarantees they hold no dense target of these ops and none of the kTileLow lowest bits.
// `reorder` = false restores strictly consecutive passes.
constexpr int kPlanWindow = 2048;  // how far past the first skipped gate a pass looks for gates that may move forward

inline std::vector<TilePass> planPasses(const std::vector<PendingOp>& ops, int localBits, unsigned long long avoidMask = 0,
                                        bool reorder = true) {
    std::vector<TilePass> out;
    const int n = (int)ops.size();
    const int T = std::min(kTileBits, localBits);
    const int R = std::min(kSubBits, T);
    const unsigned long long lowMask = (1ull << std::min(kTileLow, T)) - 1ull;
    std::vector<int> remaining(n);
    for (int i = 0; i < n; ++i) remaining[i] = i;
    std::vector<char> taken(n, 0);
    size_t head = 0;  // first position of `remaining` not yet taken
    while (true) {
        while (head < remaining.size() && taken[remaining[head]]) ++head;
        if (head == remaining.size()) break;
        TilePass p;
        p.nTile = T;
        p.denseMask = 0;
        unsigned long long skippedDense = 0, skippedAll = 0;
        int firstSkipAt = -1;
        for (size_t pos = head; pos < remaining.size() && (int)p.order.size() < kMaxTileGates; ++pos) {
            const int k = remaining[pos];
            if (taken[k]) continue;
            if (firstSkipAt >= 0 && (!reorder || (int)pos - firstSkipAt > kPlanWindow)) break;
            const Footprint f = footprintOf(ops[k]);
            const bool manyBits = __builtin_popcountll(f.all) > std::min(kMaxTouched, T - 1);
            bool fits = !manyBits && __builtin_popcountll(p.denseMask | f.dense | lowMask) <= T;
            if (manyBits && p.order.empty() && firstSkipAt < 0) {  // many-control gate: a pass of its own (single-gate kernel)
                p.order.push_back(k);
                taken[k] = 1;
                break;
            }
            if (fits && firstSkipAt >= 0 && !commutesWithAll(f, skippedDense, skippedAll)) fits = false;
            if (!fits) {
                if (firstSkipAt < 0) firstSkipAt = (int)pos;
                skippedDense |= f.dense;
                skippedAll |= f.all;
                continue;
            }
            p.order.push_back(k);
            taken[k] = 1;
            p.denseMask |= f.dense;
        }
        p.count = (int)p.order.size();
        // sub-groups: consecutive gates of the pass whose dense targets fit the register bits, at most kMaxSubGates of
        // them, touching few enough index bits that a tile bit is left over as the spectator
        std::vector<unsigned long long> subMasks;
        {
            TileSub cur{0, 0, {0, 0, 0}, 0};
            unsigned long long subDense = 0, subTouched = 0;
            for (int j = 0; j < p.count; ++j) {
                const Footprint f = footprintOf(ops[p.order[j]]);
                if (cur.count && (__builtin_popcountll(subDense | f.dense) > R || cur.count == kMaxSubGates ||
                                  __builtin_popcountll(subTouched | f.all) > T - 1)) {
                    p.subs.push_back(cur);
                    subMasks.push_back(subDense);
                    cur = TileSub{j, 0, {0, 0, 0}, 0};
                    subDense = 0;
                    subTouched = 0;
                }
                subDense |= f.dense;
                subTouched |= f.all;
                ++cur.count;
            }
            p.subs.push_back(cur);
            subMasks.push_back(subDense);
        }
        // a pass may hold at most kMaxTileSubs sub-groups: give the surplus gates back to the queue
        if ((int)p.subs.size() > kMaxTileSubs) {
            const int keep = p.subs[kMaxTileSubs].first;
            for (int j = keep; j < p.count; ++j) taken[p.order[j]] = 0;
            p.order.resize(keep);
            p.count = keep;
            p.subs.resize(kMaxTileSubs);
            subMasks.resize(kMaxTileSubs);
            p.denseMask = 0;
            for (int k : p.order) p.denseMask |= denseBitsOf(ops[k]);
            head = 0;  // (returned gates may lie before `head`; rare, so a rescan is fine)
            while (head < remaining.size() && taken[remaining[head]]) ++head;
        }
        // tile bits: dense targets + the low run, padded with the lowest free bits (longer contiguous rows in HBM)
        unsigned long long chosen = p.denseMask | lowMask;
        int have = __builtin_popcountll(chosen);
        for (int round = 0; round < 2 && have < T; ++round)
            for (int b = 0; b < localBits && have < T; ++b)
                if (!((chosen >> b) & 1) && (round == 1 || !((avoidMask >> b) & 1))) {
                    chosen |= 1ull << b;
                    ++have;
                }
        int j = 0;
        int tilePosOf[64];
        for (int b = 0; b < 64; ++b) {
            tilePosOf[b] = -1;
            if (j < T && ((chosen >> b) & 1)) {
                tilePosOf[b] = j;
                p.tileBits[j++] = b;
            }
        }
        // register bits of every sub-group: its dense targets, then tile bits its gates use as controls / selecting bits
        // (a per-element condition instead of a per-thread one), then the highest free tile bits (keeps the threads of
        // a warp on the lowest tile-index bits: neighbouring shared-memory words)
        for (size_t s = 0; s < p.subs.size(); ++s) {
            TileSub& sg = p.subs[s];
            unsigned int reg = 0;
            for (unsigned long long m = subMasks[s]; m; m &= m - 1) {
                const int tp = tilePosOf[__builtin_ctzll(m)];
                if (tp >= 0) reg |= 1u << tp;
            }
            int nreg = __builtin_popcount(reg);
            for (int k = sg.first; k < sg.first + sg.count && nreg < R; ++k)
                for (unsigned long long m = ops[p.order[k]].mask; m && nreg < R; m &= m - 1) {
                    const int tp = tilePosOf[__builtin_ctzll(m)];
                    if (tp >= 0 && !((reg >> tp) & 1)) {
                        reg |= 1u << tp;
                        ++nreg;
                    }
                }
            for (int tp = T - 1; tp >= 0 && nreg < R; --tp)
                if (!((reg >> tp) & 1)) {
                    reg |= 1u << tp;
                    ++nreg;
                }
            int q = 0;
            for (int tp = 0; tp < T; ++tp)
                if ((reg >> tp) & 1) sg.rb[q++] = tp;
            for (; q < kSubBits; ++q) sg.rb[q] = -1;
            // spectator: the highest tile bit that is neither a register bit nor mentioned by any gate of the sub-group
            unsigned int busy = reg;
            for (int k = sg.first; k < sg.first + sg.count; ++k)
                for (unsigned long long m = ops[p.order[k]].mask; m; m &= m - 1) {
                    const int tp = tilePosOf[__builtin_ctzll(m)];
                    if (tp >= 0) busy |= 1u << tp;
                }
            sg.spect = -1;
            for (int tp = T - 1; tp >= 0 && sg.spect < 0; --tp)
                if (!((busy >> tp) & 1)) sg.spect = tp;
        }
        out.push_back(p);
    }
    return out;
}

// ---- device dispatch ---------------------------------------------------------------------------------------
// The kernel resolves a gate with ONE table jump: kind and target position among the sub-group's register bits
// (dense gates) or position pair (noise) are folded into a dense opcode on the host.
enum TileOpcode : int {   // dense numbering: the kernel's switch compiles to one jump table
    kOpMatrix = 0,      // +pos (0..2)
    kOpHadamard = 3,    // +pos
    kOpSwap = 6,        // +pos
    kOpPauliY = 9,      // +pos
    kOpFlip = 12,
    kOpHadMatrix = 13,  // +pos
    kOpSwapCtl = 16,    // +pos * 2 + (0: the control is the lower of the other two register bits, 1: the higher)
    kOpPhase = 22,      // +regMask (0..7): which register elements the phase multiplies is part of the opcode - one jump
                        // instead of a jump on the kind followed by a jump on the mask
    kOpNoise = 30,      // +(kind - kFDephase) * 3 + pairSlot(pos, pos2)
    kOpCount = 42
};
// (p, q), p < q < 3  ->  0..2
inline int pairSlot(int p, int q) { return p + q - 1; }

// What the kernel reads per gate.
struct TileGate {
    int opcode;
    unsigned int regMask;          // control/selecting bits that are register bits (bit j = rb[j])
    unsigned int tileMask;         // ... that are other tile bits (tile-index space): a per-thread condition
    unsigned long long outerMask;  // ... outside the tile (amplitude-index space): a per-CTA condition
    int pos, pos2;                 // register bit of the target (noise: of the lower and the higher of the two)
};

// Translate one queued op into the kernel's descriptor for a sub-group with register bits `rb` of a pass with tile
// bits `tileBits`.
inline TileGate describe(const PendingOp& op, const TilePass& p, const TileSub& sg) {
    TileGate tg;
    std::memset(&tg, 0, sizeof tg);
    int tilePosOf[64];
    for (int b = 0; b < 64; ++b) tilePosOf[b] = -1;
    for (int j = 0; j < p.nTile; ++j) tilePosOf[p.tileBits[j]] = j;
    int regOf[kTileBits];
    for (int j = 0; j < kTileBits; ++j) regOf[j] = -1;
    for (int j = 0; j < kSubBits; ++j)
        if (sg.rb[j] >= 0) regOf[sg.rb[j]] = j;
    for (unsigned long long m = op.mask; m; m &= m - 1) {
        const int b = __builtin_ctzll(m);
        const int tp = tilePosOf[b];
        if (tp < 0) tg.outerMask |= 1ull << b;
        else if (regOf[tp] >= 0) tg.regMask |= 1u << regOf[tp];
        else tg.tileMask |= 1u << tp;
    }
    if (!isDiagonal(op.kind)) {
        const int a = regOf[tilePosOf[op.target]];
        const int b = op.target2 >= 0 ? regOf[tilePosOf[op.target2]] : -1;
        tg.pos = (b >= 0 && b < a) ? b : a;
        tg.pos2 = b >= 0 ? (b < a ? a : b) : 0;
    }
    switch (op.kind) {
        case kFMatrix: tg.opcode = kOpMatrix + tg.pos; break;
        case kFHadamard: tg.opcode = kOpHadamard + tg.pos; break;
        case kFSwap:
            tg.opcode = kOpSwap + tg.pos;
            if (__builtin_popcount(tg.regMask) == 1) {  // controlledNot with its control among the register bits: which
                const int q = __builtin_ctz(tg.regMask);  // half of the pairs moves is known per (target, control) position
                tg.opcode = kOpSwapCtl + tg.pos * 2 + ((q > (tg.pos == 0 ? 1 : 0)) ? 1 : 0);
            }
            break;
        case kFPauliY: tg.opcode = kOpPauliY + tg.pos; break;
        case kFPhase: tg.opcode = kOpPhase + (int)(tg.regMask & 7u); break;
        case kFFlip: tg.opcode = kOpFlip; break;
        case kFHadMatrix: tg.opcode = kOpHadMatrix + tg.pos; break;
        default: tg.opcode = kOpNoise + (op.kind - kFDephase) * 3 + pairSlot(tg.pos, tg.pos2); break;
    }
    return tg;
}

// Shared-memory placement of tile index i (qb_tile.cuh uses the same function on the device): the low four bits are
// XOR-ed with a linear function of the higher ones, so that the 16 threads of a half-warp - which differ in the lowest
// tile-index bits that are not register bits, whatever those are - fall into different banks (counted over all 165
// choices of 3 register bits out of 11: at most 2-way conflicts, 1.22 on average; the identity placement reaches 8-way).
// Linear over GF(2): place(a ^ b) = place(a) ^ place(b), which lets the host precompute element offsets.
#if defined(__CUDACC__)
__host__ __device__
#endif
inline unsigned int tilePlace(unsigned int i) {
    const unsigned int h = i >> 4;
    return i ^ ((h ^ (h << 1) ^ (h << 3)) & 15u);
}

// Peephole on a recorded queue: hadamard(t) directly followed by an uncontrolled matrix gate on t becomes ONE op.  The
// kernel applies the same two per-pair expressions one after the other (every product and sum still rounded
// separately: bit-identical), but pays the gate dispatch - a dependent chain of constant loads and branches that costs
// as much as the hadamard's arithmetic - once instead of twice.
inline std::vector<PendingOp> fuseAdjacentPairs(const std::vector<PendingOp>& ops) {
    std::vector<PendingOp> out;
    out.reserve(ops.size());
    for (size_t i = 0; i < ops.size(); ++i) {
        if (i + 1 < ops.size() && ops[i].kind == kFHadamard && ops[i].mask == 0 && ops[i + 1].kind == kFMatrix &&
            ops[i + 1].mask == 0 && ops[i + 1].target == ops[i].target) {
            PendingOp f = ops[i + 1];
            f.kind = kFHadMatrix;
            f.c[8] = ops[i].c[0];  // the hadamard's own factor, (qreal)(1.0 / sqrt(2)) as the front end formed it
            out.push_back(f);
            ++i;
        } else {
            out.push_back(ops[i]);
        }
    }
    return out;
}

// Kernel tier this opcode needs: 0 = the common state-vector gates (matrix, hadamard, swap / controlledNot, phase), 1 = every
// state-vector gate (+ pauliY, phase flip), 2 = + density-matrix noise and collapse.  A pass runs on the maximum over its
// gates: the dispatch code a pass does not need is code its instruction cache does not have to hold.
inline int tierOf(int opcode) {
    if (opcode >= kOpNoise) return 2;
    if ((opcode >= kOpPauliY && opcode < kOpFlip) || opcode == kOpFlip) return 1;  // (kOpSwapCtl, kOpHadMatrix, kOpPhase: tier 0)
    return 0;
}

}  // namespace qb
