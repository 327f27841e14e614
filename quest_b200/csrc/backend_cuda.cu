// quest-b200 product backend: the statevec_* / densmatr_* seam (quest_internal.h) on B200 HBM.
//
// Replaces the reference's QuEST/src/GPU/QuEST_gpu.cu and QuEST/src/CPU/QuEST_cpu*.c for the gate path.
// One process drives one GPU (its shard of the state).  There is NO CPU fallback: if no CUDA device can be
// initialised the library prints the CUDA error and exits, following the reference's print-and-exit convention
// (QuEST_validation.c:82-88, QuEST_cpu.c:1216-1217).
//
// Each entry point below decides, on the host, which bits of the amplitude index are special for the gate
// (target, controls, density-matrix column partner), builds a qb::BitPlan, picks the kernel instantiation
// (vector width VEC, and where the target falls relative to the vector) and enqueues ONE kernel on the engine's
// stream.  Gate calls do not synchronise; value-returning calls do.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include <vector>

#include "qb_comm.h"
#include "qb_io.h"
#include "qb_fuse.h"
#include "qb_kernels.cuh"
#include "qb_tile.cuh"
#include "qb_debug.h"
#include "quest_b200.h"
#include "quest_internal.h"

// NOTE on host maths: nvcc's headers put float overloads of sqrt/cos/sin in the global namespace, while the
// reference (C99) always calls the double versions and rounds on assignment.  Every libm call below therefore
// casts its argument to double explicitly, so the single-precision build derives the same coefficients.
using namespace qb;
typedef long long int idx_t;
typedef unsigned long long u64;

// ---------------------------------------------------------------------------------------------------------
// engine state
// ---------------------------------------------------------------------------------------------------------

namespace {

#define CUDA_OK(call)                                                                                   \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) {                                                                        \
            std::printf("!!!\nquest-b200 CUDA failure: %s (%s) at %s:%d\n!!!\nexiting..\n",              \
                        cudaGetErrorString(e_), #call, __FILE__, __LINE__);                             \
            std::fflush(stdout);                                                                        \
            std::exit(EXIT_FAILURE);                                                                    \
        }                                                                                               \
    } while (0)

constexpr int kMaxReduceBlocks = 4096;  // per-block partial sums kept in scratch: 2 * kMaxReduceBlocks doubles

struct Engine {
    bool up = false;
    int device = 0;
    int numSMs = 148;
    int rank = 0, world = 1;
    cudaStream_t stream = nullptr;
    bool ownsStream = false;
    double* dPartials = nullptr;    // [2 * kMaxReduceBlocks]
    unsigned int* dTicket = nullptr;
    double* dResult = nullptr;      // [4]
    double* hResult = nullptr;      // pinned [4]
    double* dWidePartials = nullptr; // [kProbAll * wideBlocks]: scratch of the all-qubit probability pass
    unsigned int wideBlocks = 0;
    double* dWide = nullptr;        // [kProbAll]
    double* hWide = nullptr;        // pinned [kProbAll]
    bool probCache = true;          // QUEST_B200_PROB_CACHE
    int vecCap = 4;                 // QUEST_B200_VEC: upper bound on amplitudes per vector (1, 2 or 4)
    unsigned long long launches = 0;
    bool fuse = true;               // QUEST_B200_FUSE / qb200_setFusion: record gates and run them in fused passes
    unsigned long long fusedLaunches = 0, fusedGates = 0;
    bool reorder = false;           // QUEST_B200_REORDER / qb200_setReordering: gates may move forward past gates they commute
                                    // with (qb_fuse.h); off by default - it changes the order of roundings, not the mathematics
    bool peephole = true;           // QUEST_B200_PEEPHOLE: hadamard + matrix on the same target share one dispatch
    bool contract = false;          // QUEST_B200_FMA / qb200_setArithmetic: fused pass may contract a*b+c into fma
    bool remap = true;              // QUEST_B200_REMAP: make shard-selecting qubits local by slot exchange
    unsigned long long slotSwaps = 0;
    bool overlap = true;            // QUEST_B200_OVERLAP: pipeline a slot exchange with the fused pass before it
    int swapCtasPerSm = 1;          // QUEST_B200_SWAP_CTAS: footprint of a pipelined exchange
    bool swapSlim = true;           // QUEST_B200_SWAP_SLIM=0: pipelined exchanges use k_swapPeer instead of the 32-register k_swapPeerSlim
                                    // (2 GPUs: 68.4 -> 67.2 ms per step: the slim CTA no longer pushes a pass CTA off the SM)
    int swapThreads = 128;          // QUEST_B200_SWAP_THREADS: ... and the size of its CTAs (measured on 2 GPUs: 64 threads leave NVLink idle,
                                    // 256 push a pass CTA off every SM; 128 hide the exchange completely behind the passes)
    int overlapMinBits = 24;        // ... for shards of at least 2^24 amplitudes (QUEST_B200_OVERLAP_MIN_BITS: tests)
    cudaStream_t comm = nullptr;    // high-priority stream of the pipelined exchanges (sharded runs only)
    cudaEvent_t slabEv[8] = {};     // engine stream: the pass before an exchange has finished slab s
    cudaEvent_t xDone[8] = {};      // communication stream: slab s of the exchange is complete on both GPUs
    cudaEvent_t commDone = nullptr; // communication stream: the whole exchange is
    void* pendingReg = nullptr;     // register whose last exchange the engine stream has not waited for yet ...
    SlabPlan pendingPlan{0, {0, 0, 0}};  // ... and how it was cut
    unsigned long long overlapped = 0;
    // deferred uploads (qb200_setDeferredUpload): a whole-register setAmps from pinned memory is cut into slabs on the
    // copy stream and returns at once; the first passes recorded afterwards follow it slab by slab
    bool deferUpload = false;
    cudaStream_t copy = nullptr;
    cudaEvent_t upEv[8] = {};
    unsigned long long deferredUploads = 0, passesUnderUpload = 0;
    bool timeFused = false;         // qb200_setKernelTiming: CUDA events around every k_tile launch
    struct TimedLaunch {
        cudaEvent_t e0, e1;
        bool opensPass;  // a pass cut into slabs is several launches, each timed after its own wait
    };
    std::vector<TimedLaunch> fusedEvents;
} g;

struct QuregRec {
    qreal* re;
    qreal* im;
    idx_t n;       // amplitudes in this shard
    int localBits; // log2 n
    std::vector<PendingOp> queue;  // gates recorded but not yet executed (qb_fuse.h)
    // The same on every rank (ranks drop different gates, so their queues differ): how many gate calls arrived since
    // the last flush and which bit positions they used as dense targets.  Decisions two GPUs must agree on use these.
    size_t callsSinceFlush = 0;
    u64 denseSinceFlush = 0;
    // calcProbOfOutcome: the second query on an unchanged state computes P(qubit = 0) for every qubit in one pass;
    // later queries are answered from here until the state changes (dirty()).
    bool logical = false;  // sharded registers, default mode: the queue holds LOGICAL qubits and full control masks
    bool probValid = false;
    bool exposed = false;  // device pointers were handed out (qb200_getDevicePointers): the state can change unseen
    int probQueries = 0;
    double probZero[64];
    QubitMap map;                  // where each logical qubit lives (qb_plan.h); identity unless sharded
    bool uploading = false;        // a deferred upload of this register is in flight on the copy stream, cut along ...
    SlabPlan uploadPlan{0, {0, 0, 0}};  // ... these bits; slab s is in place once g.upEv[s] has happened
};

std::vector<QuregRec*> liveRegisters;  // for syncQuESTEnv / qb200_setFusion: every queue must be drainable
constexpr size_t kQueueCap = 8192;

inline QuregRec* rec(const Qureg& q) { return reinterpret_cast<QuregRec*>(q.pairStateVec.real); }

inline int log2i(u64 x) { return qb::log2u(x); }

// the amplitudes of this register are about to change (or may be changed behind our back)
inline void dirty(QuregRec* r) {
    r->probValid = false;
    r->probQueries = 0;
}

void requireEngine() {
    if (!g.up) {
        std::printf("!!!\nquest-b200: createQuESTEnv() must be called before any other QuEST function\n!!!\nexiting..\n");
        std::exit(EXIT_FAILURE);
    }
}

inline void launched() {
    ++g.launches;
    CUDA_OK(cudaGetLastError());
}

inline unsigned int blocksFor(u64 tasks) { return (unsigned int)((tasks + kBlock - 1) / kBlock); }

// reductions: enough blocks to fill the chip several times over, never more than the scratch holds
inline unsigned int reduceBlocks(u64 tasks) {
    u64 want = (tasks + kBlock - 1) / kBlock;
    u64 cap = std::min<u64>(kMaxReduceBlocks, (u64)g.numSMs * 16);
    return (unsigned int)std::max<u64>(1, std::min(want, cap));
}

// widest vector (in amplitudes) allowed for a shard of n amplitudes; `limit` adds a per-call bound
inline int pickVec(idx_t n, idx_t limit = 4) {
    int v = g.vecCap;
    while (v > 1 && (v > n || v > limit)) v >>= 1;
    return v;
}

// compile-time dispatch on (VEC, TLO): TLO = target lane bit when the target is inside the vector, else -1
template <class F>
void dispatchVT(int vec, int target, F&& f) {
    const int v = log2i(vec);
    const int tlo = target < v ? target : -1;
    if (vec == 4) {
        if (tlo < 0) f(std::integral_constant<int, 4>(), std::integral_constant<int, -1>());
        else if (tlo == 0) f(std::integral_constant<int, 4>(), std::integral_constant<int, 0>());
        else f(std::integral_constant<int, 4>(), std::integral_constant<int, 1>());
    } else if (vec == 2) {
        if (tlo < 0) f(std::integral_constant<int, 2>(), std::integral_constant<int, -1>());
        else f(std::integral_constant<int, 2>(), std::integral_constant<int, 0>());
    } else {
        f(std::integral_constant<int, 1>(), std::integral_constant<int, -1>());
    }
}

template <class F>
void dispatchV(int vec, F&& f) {
    if (vec == 4) f(std::integral_constant<int, 4>());
    else if (vec == 2) f(std::integral_constant<int, 2>());
    else f(std::integral_constant<int, 1>());
}

u64 maskOf(const int* qubits, int n) {
    u64 m = 0;
    for (int i = 0; i < n; ++i) m |= 1ull << qubits[i];
    return m;
}

// ---- gate execution ----------------------------------------------------------------------------------------
// Every gate becomes a qb::PendingOp.  It is either executed at once (runOpNow: one single-gate kernel, the
// shard-crossing path) or recorded in the register's queue and executed later inside a fused pass (flush).

// One dense gate on THIS shard only (target < localBits, controls already local).
template <class Op>
void runPairsNow(const Qureg& q, int target, u64 ctrlMask, const Op& op) {
    QuregRec* r = rec(q);
    const int vec = pickVec(r->n);
    const Planned p = makePlan(r->n, vec, {target}, ctrlMask);
    if (p.nTasks == 0) return;
    const int tvb = target - log2i(vec);
    dispatchVT(vec, target, [&](auto V, auto T) {
        k_pairs<Op, decltype(V)::value, decltype(T)::value>
            <<<blocksFor(p.nTasks), kBlock, 0, g.stream>>>(r->re, r->im, p.nTasks, p.plan, tvb, op);
    });
    launched();
}

// Dense gate whose target qubit selects the shard: fused exchange+combine over the partner's peer mapping
// (k_pairsPeer).  qb::planShardedGate (qb_plan.h) gives this rank's role and its half of the pairs; two
// stream-ordered handshakes bracket the kernel.
template <class Op>
void runPairsAcrossShards(const Qureg& q, int target, u64 ctrlMask, const Op& op) {
    QuregRec* r = rec(q);
    const int vec = pickVec(r->n);
    const ShardedGate sg = planShardedGate(r->localBits, q.chunkId, target, ctrlMask, vec);
    const Planned& p = sg.planned;
    const comm::Partner peer = comm::partnerOf(q, target - r->localBits);
    comm::pairBarrier(peer.rank, g.stream);  // partner has finished everything enqueued before this gate
    if (sg.launches) {
        qreal* upRe = sg.amUpper ? r->re : peer.re;
        qreal* upIm = sg.amUpper ? r->im : peer.im;
        qreal* loRe = sg.amUpper ? peer.re : r->re;
        qreal* loIm = sg.amUpper ? peer.im : r->im;
        dispatchV(vec, [&](auto V) {
            k_pairsPeer<Op, decltype(V)::value>
                <<<blocksFor(p.nTasks), kBlock, 0, g.stream>>>(upRe, upIm, loRe, loIm, p.nTasks, p.plan, op);
        });
        launched();
    }
    comm::pairBarrier(peer.rank, g.stream);  // partner's remote writes into this shard have landed
}

template <class Op>
void runDiagNow(const Qureg& q, u64 mask, const Op& op) {
    QuregRec* r = rec(q);
    const int vec = pickVec(r->n);
    const Planned p = makePlan(r->n, vec, {}, mask);
    if (p.nTasks == 0) return;
    dispatchV(vec, [&](auto V) {
        k_diag<Op, decltype(V)::value><<<blocksFor(p.nTasks), kBlock, 0, g.stream>>>(r->re, r->im, p.nTasks, p.plan, op);
    });
    launched();
}

// One noise channel in its own pass (k_noise1).  The channels are symmetric in the qubit's row and column bits, so
// the lower of the two bit positions plays the row and the higher - always above the vector - the column.
template <int KIND>
void runNoiseNow(const Qureg& q, int bitA, int bitB, qreal p0, qreal p1, qreal p2) {
    QuregRec* r = rec(q);
    const int lo = bitA < bitB ? bitA : bitB, hi = bitA < bitB ? bitB : bitA;
    const int vec = pickVec(r->n, 1LL << hi);
    const int v = log2i(vec);
    const Planned p = makePlan(r->n, vec, {lo, hi}, 0);
    dispatchVT(vec, lo, [&](auto V, auto T) {
        k_noise1<KIND, decltype(V)::value, decltype(T)::value><<<blocksFor(p.nTasks), kBlock, 0, g.stream>>>(
            r->re, r->im, p.nTasks, p.plan, lo - v, hi - v, p0, p1, p2);
    });
    launched();
}

// Density-matrix collapse in its own pass (k_densCollapse); symmetric in the two bits like the noise channels.
void runDensCollapseNow(const Qureg& q, int bitA, int bitB, int outcome, qreal prob) {
    QuregRec* r = rec(q);
    const int lo = bitA < bitB ? bitA : bitB, hi = bitA < bitB ? bitB : bitA;
    const int vec = pickVec(r->n, 1LL << hi);
    const int v = log2i(vec);
    const Planned p = makePlan(r->n, vec, {lo, hi}, 0);
    dispatchVT(vec, lo, [&](auto V, auto T) {
        k_densCollapse<decltype(V)::value, decltype(T)::value><<<blocksFor(p.nTasks), kBlock, 0, g.stream>>>(
            r->re, r->im, p.nTasks, p.plan, lo - v, hi - v, outcome, prob);
    });
    launched();
}

OpMatrix matrixOf(const PendingOp& op) {
    return OpMatrix{op.c[0], op.c[1], op.c[2], op.c[3], op.c[4], op.c[5], op.c[6], op.c[7]};
}

void runOpNow(const Qureg& q, const PendingOp& op) {
    const bool cross = !isDiagonal(op.kind) && op.target >= rec(q)->localBits;
    switch (op.kind) {
        case kFMatrix:
            cross ? runPairsAcrossShards(q, op.target, op.mask, matrixOf(op)) : runPairsNow(q, op.target, op.mask, matrixOf(op));
            break;
        case kFHadMatrix: {  // a fused pair that ended up alone in a pass: its two gates, one kernel each
            PendingOp h = op, m = op;
            h.kind = kFHadamard;
            h.c[0] = op.c[8];
            m.kind = kFMatrix;
            runOpNow(q, h);
            runOpNow(q, m);
            break;
        }
        case kFHadamard:
            cross ? runPairsAcrossShards(q, op.target, op.mask, OpHadamard{op.c[0]}) : runPairsNow(q, op.target, op.mask, OpHadamard{op.c[0]});
            break;
        case kFSwap:
            cross ? runPairsAcrossShards(q, op.target, op.mask, OpSwap{}) : runPairsNow(q, op.target, op.mask, OpSwap{});
            break;
        case kFPauliY:
            cross ? runPairsAcrossShards(q, op.target, op.mask, OpPauliY{op.c[0]}) : runPairsNow(q, op.target, op.mask, OpPauliY{op.c[0]});
            break;
        case kFPhase:
            runDiagNow(q, op.mask, DiagPhase{op.c[0], op.c[1]});
            break;
        case kFFlip:
            runDiagNow(q, op.mask, DiagFlip{});
            break;
        case kFDephase:
            runNoiseNow<kDephase>(q, op.target, op.target2, op.c[0], op.c[1], op.c[2]);
            break;
        case kFDepolarise:
            runNoiseNow<kDepolarise>(q, op.target, op.target2, op.c[0], op.c[1], op.c[2]);
            break;
        case kFDamping:
            runNoiseNow<kDamping>(q, op.target, op.target2, op.c[0], op.c[1], op.c[2]);
            break;
        default:
            runDensCollapseNow(q, op.target, op.target2, op.c[1] != 0 ? 1 : 0, op.c[0]);
            break;
    }
}

TileLaunch tl;  // ~21 KB: passed by value as the kernel parameter block

// Fill the parameter block for one planned pass (qb_fuse.h); returns the kernel tier it needs.  `alive` (sharded
// registers): gates this rank takes no part in (a control on a shard-selecting qubit is not satisfied here) are planned
// like the others - every rank cuts the queue the same way - but do not reach the kernel.
int preparePass(const std::vector<PendingOp>& ops, const TilePass& p, const std::vector<char>* alive = nullptr) {
    int tier = 0;
    tl.nTile = p.nTile;
    tl.nLow = std::min(kTileLow, p.nTile);
    tl.nSub = (int)p.subs.size();
    int g = 0;
    for (int s = 0; s < tl.nSub; ++s) {
        const TileSub& sg = p.subs[s];
        TileSubDesc& d = tl.sub[s];
        d.first = g;
        int pos[kRegBits];  // element bit j -> tile-index position: the register bits, then the spectator bit
        for (int j = 0; j < kSubBits; ++j) pos[j] = sg.rb[j];
        pos[kSubBits] = sg.spect;
        for (int j = 0; j < kRegBits; ++j) d.ins[j] = pos[j];
        std::sort(d.ins, d.ins + kRegBits);
        for (int j = 0; j < kRegBits; ++j) d.placeBit[j] = tilePlace(1u << pos[j]) * (unsigned int)sizeof(qreal);
        for (int j = sg.first; j < sg.first + sg.count; ++j) {
            const int k = p.order[j];
            if (alive && !(*alive)[k]) continue;
            const TileGate tg = describe(ops[k], p, sg);
            tl.ctl[g] = TileGateCtl{(unsigned short)tg.tileMask, (unsigned char)tg.regMask, (unsigned char)tg.opcode};
            tl.outerMask[g] = tg.outerMask;
            std::memcpy(tl.c[g], ops[k].c, sizeof ops[k].c);
            tier = std::max(tier, tierOf(tg.opcode));  // the tiers are nested
            ++g;
        }
        d.count = g - d.first;
    }
    tl.nGates = g;
    // amplitude offset of every row of the tile: tile-index bits [nLow, nTile) deposited at their index positions
    for (unsigned int r = 0; r < (1u << (p.nTile - tl.nLow)); ++r) {
        u64 o = 0;
        for (int j = tl.nLow; j < p.nTile; ++j)
            if ((r >> (j - tl.nLow)) & 1) o |= 1ull << p.tileBits[j];
        tl.rowOff[r] = o;
    }
    for (int j = 0; j < kTileCopies; ++j) {
        const unsigned int i = (unsigned int)j * kTileThreads;
        tl.stepOff[j] = i < (1u << p.nTile) ? tl.rowOff[i >> tl.nLow] + (i & ((1u << tl.nLow) - 1u)) : 0;
    }
    return tier;
}

// One launch of k_tile over the whole shard (sp == nullptr) or over slab `slab` of it.
void launchPrepared(QuregRec* r, const TilePass& p, int tier, const SlabPlan* sp, int slab) {
    // the index bits that are neither tile bits nor slab bits, as contiguous runs fed from the CTA index
    u64 held = 0;
    for (int j = 0; j < p.nTile; ++j) held |= 1ull << p.tileBits[j];
    if (sp)
        for (int j = 0; j < sp->nBits; ++j) held |= 1ull << sp->bits[j];
    int nRuns = 0, src = 0;
    for (int b = 0; b < r->localBits;) {
        if ((held >> b) & 1) {
            ++b;
            continue;
        }
        int e = b;
        while (e < r->localBits && !((held >> e) & 1)) ++e;
        tl.runSrc[nRuns] = (unsigned char)src;
        tl.runDst[nRuns] = (unsigned char)b;
        tl.runLen[nRuns] = (unsigned char)(e - b);
        src += e - b;
        ++nRuns;
        b = e;
    }
    if (tl.nGates == 0) return;  // nothing for this rank in this pass
    tl.nRuns = nRuns;
    tl.slabMask = sp ? sp->valueMask(slab) : 0;
    const unsigned int nb = (unsigned int)((u64)r->n >> __builtin_popcountll(held));
    auto go = [&](auto kernel) { kernel<<<nb, kTileThreads, 0, g.stream>>>(r->re, r->im, tl); };
    if (tier >= kTierNoise) g.contract ? go(k_tile<kTierNoise, true>) : go(k_tile<kTierNoise, false>);
    else if (tier == kTierGates) g.contract ? go(k_tile<kTierGates, true>) : go(k_tile<kTierGates, false>);
    else g.contract ? go(k_tile<kTierCore, true>) : go(k_tile<kTierCore, false>);
    launched();
}

struct FusedTimer {  // qb200_setKernelTiming: an event pair per fused launch (a pass cut into slabs counts as one pass)
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    void start() {
        if (!g.timeFused) return;
        CUDA_OK(cudaEventCreate(&e0));
        CUDA_OK(cudaEventCreate(&e1));
        CUDA_OK(cudaEventRecord(e0, g.stream));
    }
    void stop(bool opensPass = true) {
        if (!g.timeFused) return;
        CUDA_OK(cudaEventRecord(e1, g.stream));
        g.fusedEvents.push_back(Engine::TimedLaunch{e0, e1, opensPass});
    }
};

void runPass(const Qureg& q, const std::vector<PendingOp>& ops, const TilePass& p, const std::vector<char>* alive = nullptr) {
    if (p.count == 1) {
        if (!alive || (*alive)[p.order[0]]) runOpNow(q, ops[p.order[0]]);
        return;
    }
    const int tier = preparePass(ops, p, alive);
    FusedTimer t;
    t.start();
    launchPrepared(rec(q), p, tier, nullptr, 0);
    t.stop();
    ++g.fusedLaunches;
    g.fusedGates += p.count;
}

// A pipelined exchange (flushThenExchange) returns while its last slabs are still moving; the engine stream catches
// up here before anything else touches that register.
void settle() {
    if (!g.pendingReg) return;
    CUDA_OK(cudaStreamWaitEvent(g.stream, g.commDone, 0));
    g.pendingReg = nullptr;
}

inline u64 bitsOf(const SlabPlan& sp) {
    u64 m = 0;
    for (int j = 0; j < sp.nBits; ++j) m |= 1ull << sp.bits[j];
    return m;
}
inline u64 bitsOf(const TilePass& p) {
    u64 m = 0;
    for (int j = 0; j < p.nTile; ++j) m |= 1ull << p.tileBits[j];
    return m;
}

// One fused pass cut along `sp`: slab s starts once `waitFor[s]` has happened (the exchange before it delivered that
// slab) and `record[s]` marks its end (the exchange after it may take that slab).
void runPassInSlabs(const Qureg& q, const std::vector<PendingOp>& ops, const TilePass& p, const SlabPlan& sp,
                    const cudaEvent_t* waitFor, const cudaEvent_t* record, const std::vector<char>* alive = nullptr) {
    const int tier = preparePass(ops, p, alive);
    for (int s = 0; s < (1 << sp.nBits); ++s) {
        if (waitFor) CUDA_OK(cudaStreamWaitEvent(g.stream, waitFor[s], 0));
        FusedTimer t;  // after the wait, so that the measured time is the pass and not the exchange before it
        t.start();
        launchPrepared(rec(q), p, tier, &sp, s);
        t.stop(s == 0);
        if (record) CUDA_OK(cudaEventRecord(record[s], g.stream));
    }
    ++g.fusedLaunches;
    g.fusedGates += p.count;
}

void exchangeSlots(const Qureg& q, SlotSwap sw);
void flushSharded(const Qureg& q);

// Execute everything recorded for this register: single gates through their own kernel, runs of gates through
// k_tile (one HBM pass per planned pass, qb_fuse.h).
// The engine stream catches up with a deferred upload of this register (everything enqueued afterwards sees the data).
void settleUpload(QuregRec* r) {
    if (!r->uploading) return;
    for (int s = 0; s < (1 << r->uploadPlan.nBits); ++s) CUDA_OK(cudaStreamWaitEvent(g.stream, g.upEv[s], 0));
    r->uploading = false;
}

// Recorded gates behind a deferred upload: the leading passes that do not touch the slab bits run slab by slab, each
// slab as soon as its part of the state has arrived over PCIe - the upload and those passes overlap.
void flushBehindUpload(const Qureg& q) {
    QuregRec* r = rec(q);
    std::vector<PendingOp> ops;
    ops.swap(r->queue);
    if (g.peephole) ops = fuseAdjacentPairs(ops);
    const SlabPlan sp = r->uploadPlan;
    const std::vector<TilePass> passes = planPasses(ops, r->localBits, bitsOf(sp), g.reorder);
    size_t nFollow = 0;
    while (nFollow < passes.size() && passes[nFollow].count > 1 && !(bitsOf(passes[nFollow]) & bitsOf(sp))) ++nFollow;
    for (int s = 0; s < (1 << sp.nBits); ++s) {
        CUDA_OK(cudaStreamWaitEvent(g.stream, g.upEv[s], 0));
        for (size_t i = 0; i < nFollow; ++i) {
            const int tier = preparePass(ops, passes[i]);
            launchPrepared(r, passes[i], tier, &sp, s);
        }
    }
    r->uploading = false;
    for (size_t i = 0; i < nFollow; ++i) {
        ++g.fusedLaunches;
        ++g.passesUnderUpload;
        g.fusedGates += passes[i].count;
    }
    for (size_t i = nFollow; i < passes.size(); ++i) runPass(q, ops, passes[i]);
}

void flush(const Qureg& q) {
    QuregRec* r = rec(q);
    r->callsSinceFlush = 0;
    r->denseSinceFlush = 0;
    if (r->uploading) {
        if (r->queue.empty()) settleUpload(r);
        else flushBehindUpload(q);
        return;
    }
    if (r->logical) {
        flushSharded(q);
        return;
    }
    const bool pending = g.pendingReg == r;
    if (r->queue.empty()) {
        if (pending) settle();
        return;
    }
    std::vector<PendingOp> ops;
    ops.swap(r->queue);
    if (g.peephole) ops = fuseAdjacentPairs(ops);
    const std::vector<TilePass> passes = planPasses(ops, r->localBits, pending ? bitsOf(g.pendingPlan) : 0, g.reorder);
    size_t first = 0;
    if (pending) {
        // the first pass follows the exchange slab by slab if it can be cut the same way
        if (passes[0].count > 1 && !(bitsOf(passes[0]) & bitsOf(g.pendingPlan))) {
            runPassInSlabs(q, ops, passes[0], g.pendingPlan, g.xDone, nullptr);
            g.pendingReg = nullptr;  // the last slab waited for the last piece of the exchange
            first = 1;
        } else {
            settle();
        }
    }
    for (size_t i = first; i < passes.size(); ++i) runPass(q, ops, passes[i]);
}

void flushAll() {
    for (QuregRec* r : liveRegisters) {
        if (r->queue.empty() && !r->uploading) continue;
        Qureg q;
        std::memset(&q, 0, sizeof q);
        q.pairStateVec.real = reinterpret_cast<qreal*>(r);
        q.numAmpsPerChunk = r->n;
        q.chunkId = g.rank;
        q.numChunks = g.world;
        flush(q);
    }
    settle();
}

// The k_swapPeer launch of a slot exchange, whole shard (sp == nullptr) or one slab, bracketed by nothing: the
// callers place the handshakes.  Each GPU of the pair moves half of the (local slot = 1 here, = 0 there) pairs.
void launchSwap(const Qureg& q, SlotSwap sw, const SlabPlan* sp, int slab, cudaStream_t stream) {
    QuregRec* r = rec(q);
    const int gbit = sw.globalSlot - r->localBits, l = sw.localSlot;
    const comm::Partner peer = comm::partnerOf(q, gbit);
    const bool amUpper = ((q.chunkId >> gbit) & 1) == 0;
    const bool slim = sp != nullptr && g.swapSlim;  // next to a fused pass: the 32-register variant with 16-byte vectors
    const int vec = pickVec(r->n, slim ? std::min<idx_t>(2, 1LL << l) : (1LL << l));  // the local slot stays above the vector
    const int v = log2i(vec);
    const SwapLaunch plan = planSlotSwap(r->localBits, vec, l, amUpper, sp, slab);
    if (!plan.active) return;
    const Planned& p = plan.planned;
    qreal* aRe = amUpper ? r->re : peer.re;
    qreal* aIm = amUpper ? r->im : peer.im;
    qreal* bRe = amUpper ? peer.re : r->re;
    qreal* bIm = amUpper ? peer.im : r->im;
    // alone: up to 8 CTAs of 256 threads on every SM; next to a fused pass: g.swapCtasPerSm small CTAs (g.swapThreads
    // threads: they must fit into the registers four resident pass CTAs leave over, see qb_tile.cuh)
    const unsigned int threads = sp ? (unsigned int)g.swapThreads : (unsigned int)kBlock;
    const unsigned int cap = (unsigned int)g.numSMs * (sp ? (unsigned int)g.swapCtasPerSm : 8u);
    const unsigned int grid = std::min((unsigned int)((p.nTasks + threads - 1) / threads), cap);
    if (slim) {
        if (vec == 2) k_swapPeerSlim<2><<<grid, threads, 0, stream>>>(aRe, aIm, bRe, bIm, p.nTasks, p.plan, l - v);
        else k_swapPeerSlim<1><<<grid, threads, 0, stream>>>(aRe, aIm, bRe, bIm, p.nTasks, p.plan, l - v);
    } else {
        dispatchV(vec, [&](auto V) {
            k_swapPeer<decltype(V)::value><<<grid, threads, 0, stream>>>(aRe, aIm, bRe, bIm, p.nTasks, p.plan, l - v);
        });
    }
    launched();
}

// Exchange a shard-selecting bit position with a local one; caller has flushed the queue.
void exchangeSlots(const Qureg& q, SlotSwap sw) {
    QuregRec* r = rec(q);
    settle();
    const comm::Partner peer = comm::partnerOf(q, sw.globalSlot - r->localBits);
    comm::pairBarrier(peer.rank, g.stream);  // partner has finished everything enqueued before the exchange
    launchSwap(q, sw, nullptr, 0, g.stream);
    comm::pairBarrier(peer.rank, g.stream);  // partner's remote writes into this shard have landed
    r->map.applySwap(sw.globalSlot, sw.localSlot);
    ++g.slotSwaps;
}

// ---- sharded registers, default mode: plan the whole recorded queue, then execute ---------------------------------
// The queue holds logical gates - the same list on every rank.  qb::planWalk cuts it into segments whose dense targets
// are local and the slot exchanges between them (victims by furthest next use).  Every segment is planned into passes on
// the FULL gate list (so that all ranks cut alike); gates this rank takes no part in are dropped only when a pass is
// handed to the kernel.  An exchange is cut into slabs along index bits that neither it nor the passes next to it hold,
// and pipelined on the communication stream: slab s moves over NVLink once the pass before it has finished slab s on
// both GPUs of the pair, while that pass computes slab s+1; the pass after it follows slab by slab in the same way.
struct ShardSegment {
    std::vector<PendingOp> ops;    // bit positions under the layout of the segment, controls localised for this rank
    std::vector<char> alive;
    std::vector<TilePass> passes;
    u64 firstBits = 0, lastBits = 0;  // tile bits of the first / last pass
};

void planSegment(ShardSegment& seg, const std::vector<PendingOp>& full, int localBits, u64 avoid) {
    seg.passes = planPasses(full, localBits, avoid, g.reorder);
    seg.firstBits = bitsOf(seg.passes.front());
    seg.lastBits = bitsOf(seg.passes.back());
}

void flushSharded(const Qureg& q) {
    QuregRec* r = rec(q);
    r->callsSinceFlush = 0;
    r->denseSinceFlush = 0;
    r->logical = false;
    if (r->queue.empty()) {
        if (g.pendingReg == r) settle();
        return;
    }
    std::vector<PendingOp> logicalOps;
    logicalOps.swap(r->queue);
    if (g.peephole) logicalOps = fuseAdjacentPairs(logicalOps);
    std::vector<WalkOp> walk(logicalOps.size());
    for (size_t i = 0; i < logicalOps.size(); ++i) {
        const PendingOp& op = logicalOps[i];
        walk[i] = WalkOp{isDiagonal(op.kind) ? -1 : op.target, isDiagonal(op.kind) ? -1 : op.target2, op.mask};
    }
    QubitMap after = r->map;
    const std::vector<WalkStep> steps = planWalk(after, walk);

    // 1. every segment in bit positions under its own layout (rank-independent: full control masks), dry-planned
    QubitMap cur = r->map;
    std::vector<ShardSegment> segs(steps.size());
    std::vector<std::vector<PendingOp>> full(steps.size());
    for (size_t k = 0; k < steps.size(); ++k) {
        if (steps[k].isExchange) {
            cur.applySwap(steps[k].sw.globalSlot, steps[k].sw.localSlot);
            continue;
        }
        for (int i = steps[k].first; i < steps[k].first + steps[k].count; ++i) {
            PendingOp op = logicalOps[i];
            if (!isDiagonal(op.kind)) {
                op.target = cur.phys[op.target];
                if (op.target2 >= 0) op.target2 = cur.phys[op.target2];
            }
            op.mask = cur.physMask(op.mask);
            full[k].push_back(op);
        }
        planSegment(segs[k], full[k], r->localBits, 0);
    }
    // 2. slab plan of every exchange: the highest local bits held neither by the exchange nor by the passes around it
    const bool pipeline = g.overlap && g.comm && r->localBits >= g.overlapMinBits;
    std::vector<SlabPlan> plan(steps.size(), SlabPlan{0, {0, 0, 0}});
    std::vector<u64> avoid(steps.size(), 0);
    for (size_t k = 0; k < steps.size(); ++k) {
        if (!steps[k].isExchange || !pipeline) continue;
        const bool hasBefore = k > 0 && !steps[k - 1].isExchange, hasAfter = k + 1 < steps.size() && !steps[k + 1].isExchange;
        u64 busy = (1ull << steps[k].sw.localSlot) | ((1ull << kTileLow) - 1ull);
        if (hasBefore) busy |= segs[k - 1].lastBits;
        if (hasAfter) busy |= segs[k + 1].firstBits;
        if (!hasBefore && !hasAfter) continue;
        // a pass that sits alone between two exchanges is cut once: reuse the previous exchange's plan if it still fits
        SlabPlan sp{0, {0, 0, 0}};
        if (k >= 2 && steps[k - 2].isExchange && plan[k - 2].nBits && segs[k - 1].passes.size() == 1 && !(bitsOf(plan[k - 2]) & busy))
            sp = plan[k - 2];
        else
            sp = planSlabs(r->localBits, busy, kMaxSlabBits, std::min(16, g.overlapMinBits / 2 + 2));
        plan[k] = sp;
        if (hasBefore) avoid[k - 1] |= bitsOf(sp);
        if (hasAfter) avoid[k + 1] |= bitsOf(sp);
    }
    // 3. final plans (padding keeps clear of the slab bits; dense targets are unchanged), this rank's view of the gates
    for (size_t k = 0; k < steps.size(); ++k) {
        if (steps[k].isExchange) continue;
        if (avoid[k]) planSegment(segs[k], full[k], r->localBits, avoid[k]);
        segs[k].ops = full[k];
        segs[k].alive.assign(full[k].size(), 1);
        for (size_t i = 0; i < full[k].size(); ++i)
            if (!qb::localiseControls(r->localBits, q.chunkId, segs[k].ops[i].mask)) segs[k].alive[i] = 0;
    }
    // 4. execute
    bool signalled = false;             // the pass just run has recorded slabEv[] for the exchange that follows it
    bool inFlight = g.pendingReg == r;  // an exchange whose slabs the engine stream has not waited for yet ...
    SlabPlan inFlightPlan = g.pendingPlan;  // ... cut this way
    if (inFlight) {  // left over from an earlier call: not planned here
        settle();
        inFlight = false;
    }
    for (size_t k = 0; k < steps.size(); ++k) {
        if (steps[k].isExchange) {
            const SlotSwap sw = steps[k].sw;
            const SlabPlan& sp = plan[k];
            if (sp.nBits == 0) {
                if (inFlight) {
                    settle();
                    inFlight = false;
                }
                exchangeSlots(q, sw);
                continue;
            }
            const int nSlabs = 1 << sp.nBits;
            // Unless the pass before it recorded slabEv[] slab by slab, everything the engine stream holds so far has to
            // finish first.  (An exchange straight after another one simply queues up behind it on the communication
            // stream; an exchange after a segment finds no exchange in flight: the segment's first pass has waited.)
            if (!signalled)
                for (int sl = 0; sl < nSlabs; ++sl) CUDA_OK(cudaEventRecord(g.slabEv[sl], g.stream));
            signalled = false;
            const comm::Partner peer = comm::partnerOf(q, sw.globalSlot - r->localBits);
            for (int sl = 0; sl < nSlabs; ++sl) {
                CUDA_OK(cudaStreamWaitEvent(g.comm, g.slabEv[sl], 0));
                comm::pairBarrier(peer.rank, g.comm, 1);  // the partner has finished slab sl too
                launchSwap(q, sw, &sp, sl, g.comm);
                comm::pairBarrier(peer.rank, g.comm, 1);  // the partner's remote writes into this slab have landed
                CUDA_OK(cudaEventRecord(g.xDone[sl], g.comm));
            }
            CUDA_OK(cudaEventRecord(g.commDone, g.comm));
            g.pendingReg = r;
            g.pendingPlan = sp;
            inFlight = true;
            inFlightPlan = sp;
            r->map.applySwap(sw.globalSlot, sw.localSlot);
            ++g.slotSwaps;
            ++g.overlapped;
            continue;
        }
        ShardSegment& seg = segs[k];
        const bool exchangeNext = k + 1 < steps.size() && steps[k + 1].isExchange && plan[k + 1].nBits > 0;
        for (size_t i = 0; i < seg.passes.size(); ++i) {
            const TilePass& p = seg.passes[i];
            const bool isFirst = i == 0, isLast = i + 1 == seg.passes.size();
            const bool fused = p.count > 1;
            const bool follow = inFlight && isFirst && fused && !(bitsOf(p) & bitsOf(inFlightPlan));
            bool signal = exchangeNext && isLast && fused && !(bitsOf(p) & bitsOf(plan[k + 1]));
            if (follow && signal && bitsOf(inFlightPlan) != bitsOf(plan[k + 1])) signal = false;  // one cut per pass
            if (inFlight && isFirst && !follow) {
                settle();
                inFlight = false;
            }
            if (follow || signal) {
                runPassInSlabs(q, seg.ops, p, follow ? inFlightPlan : plan[k + 1], follow ? g.xDone : nullptr,
                               signal ? g.slabEv : nullptr, &seg.alive);
                if (follow) {  // the last slab waited for the last piece of the exchange
                    g.pendingReg = nullptr;
                    inFlight = false;
                }
            } else {
                runPass(q, seg.ops, p, &seg.alive);
            }
            if (isLast) signalled = signal;
        }
    }
    if (inFlight) settle();
    // the use history the walk accumulated (the layout itself was replayed exchange by exchange above)
    for (int lq = 0; lq < 64; ++lq) {
        r->map.lastUse[lq] = after.lastUse[lq];
        r->map.prevUse[lq] = after.prevUse[lq];
    }
    r->map.tick = after.tick;
}

// the register's record with no gate pending: every entry point that reads or overwrites amplitudes uses this
inline QuregRec* live(const Qureg& q) {
    flush(q);
    return rec(q);
}

// ... and with every qubit back in its own slot: entry points that address amplitudes by index range
inline QuregRec* canon(const Qureg& q) {
    flush(q);
    QuregRec* r = rec(q);
    if (!r->map.identity()) {
        SlotSwap sw[192];
        const int n = planCanonicalize(r->map, sw);
        for (int i = 0; i < n; ++i) exchangeSlots(q, sw[i]);
    }
    return r;
}

// the state is about to be overwritten as a whole: drop pending gates and forget the layout
inline QuregRec* fresh(const Qureg& q) {
    QuregRec* r = rec(q);
    if (g.pendingReg == r) settle();
    settleUpload(r);
    dirty(r);
    r->queue.clear();
    r->logical = false;
    r->callsSinceFlush = 0;
    r->denseSinceFlush = 0;
    r->map.reset(r->map.nQubits, r->localBits);
    return r;
}

// Entry for every gate of the API.  Controls on shard-selecting qubits are resolved here (QuEST_cpu.c:1841,2694):
// this rank either drops them or has nothing to do.
void submit(const Qureg& q, PendingOp op) {
    QuregRec* r = rec(q);
    dirty(r);
    if (g.world > 1 && g.fuse && g.remap && !r->exposed && r->localBits >= kRegBits) {
        // sharded register, default mode: the gate is recorded as it came - logical qubits, full control mask - so that
        // every rank holds the same queue; layout decisions are taken when the queue is executed (flushSharded)
        if (r->callsSinceFlush >= kQueueCap) flush(q);
        ++r->callsSinceFlush;
        r->logical = true;
        r->queue.push_back(op);
        return;
    }
    // logical qubits -> current bit positions; a dense target living in a shard-selecting slot is first made local
    if (g.world > 1) {
        if (!isDiagonal(op.kind)) {
            const int logical[2] = {op.target, op.target2};
            for (int k = 0; k < 2; ++k) {
                const int t = logical[k];
                if (t < 0) continue;
                if (g.remap && r->localBits >= 3 && r->map.phys[t] >= r->localBits) {
                    u64 keep = op.mask | (1ull << logical[0]);
                    if (logical[1] >= 0) keep |= 1ull << logical[1];
                    const SlotSwap sw = planSwapIn(r->map, t, keep);
                    if (sw.globalSlot >= 0) {
                        flush(q);
                        exchangeSlots(q, sw);
                    }
                }
                r->map.touch(t);
            }
            op.target = r->map.phys[logical[0]];
            if (logical[1] >= 0) op.target2 = r->map.phys[logical[1]];
        }
        op.mask = r->map.physMask(op.mask);
    }
    // registers whose device pointers were handed out run every gate at once: whoever holds the raw pointers sees the
    // amplitudes the API calls so far produce (stream-ordered), and nothing queued can overwrite what they write
    const bool recording = g.fuse && !r->exposed && r->localBits >= kRegBits;
    if (recording) {  // before any rank-dependent decision: every rank counts every gate call
        if (r->callsSinceFlush >= kQueueCap) flush(q);
        ++r->callsSinceFlush;
        if (!isDiagonal(op.kind)) {
            if (op.target < r->localBits) r->denseSinceFlush |= 1ull << op.target;
            if (op.target2 >= 0 && op.target2 < r->localBits) r->denseSinceFlush |= 1ull << op.target2;
        }
    }
    if (!recording) settleUpload(r);
    if (!qb::localiseControls(r->localBits, q.chunkId, op.mask)) return;
    if (isNoise(op.kind) && (op.target >= r->localBits || op.target2 >= r->localBits))
        comm::unsupported("one-qubit noise on a qubit that selects the shard (with QUEST_B200_REMAP=0)");
    const bool cross = !isDiagonal(op.kind) && op.target >= r->localBits;
    if (cross) {
        flush(q);
        runOpNow(q, op);
        return;
    }
    if (recording) {
        r->queue.push_back(op);
        return;
    }
    runOpNow(q, op);
}

PendingOp denseOp(int kind, int target, u64 ctrlMask) {
    PendingOp op;
    std::memset(&op, 0, sizeof op);
    op.kind = kind;
    op.target = target;
    op.target2 = -1;
    op.mask = ctrlMask;
    return op;
}

PendingOp matrixOp(int target, u64 ctrlMask, ComplexMatrix2 u) {
    PendingOp op = denseOp(kFMatrix, target, ctrlMask);
    const qreal c[8] = {u.r0c0.real, u.r0c0.imag, u.r0c1.real, u.r0c1.imag, u.r1c0.real, u.r1c0.imag, u.r1c1.real, u.r1c1.imag};
    std::memcpy(op.c, c, sizeof c);
    return op;
}

// [[alpha, -conj(beta)], [beta, conj(alpha)]] as a general matrix.  Evaluated by OpMatrix this gives the same IEEE
// bits as the reference's compact expressions (QuEST_cpu.c:1614-1623): x + (-b)*y == x - b*y and x - (-a)*y == x + a*y.
PendingOp compactOp(int target, u64 ctrlMask, Complex a, Complex b) {
    PendingOp op = denseOp(kFMatrix, target, ctrlMask);
    const qreal c[8] = {a.real, a.imag, -b.real, b.imag, b.real, b.imag, a.real, -a.imag};
    std::memcpy(op.c, c, sizeof c);
    return op;
}

PendingOp phaseOp(u64 mask, qreal c, qreal s) {
    PendingOp op = denseOp(kFPhase, -1, mask);
    op.c[0] = c;
    op.c[1] = s;
    return op;
}

// ---- reductions --------------------------------------------------------------------------------------------

// device result[0..nv) -> (sum over shards) -> host.  The only synchronisation point of a reduction.
void fetchResult(int nv, double* out) {
    if (g.world > 1) comm::allReduceSum(g.dResult, nv, g.stream);
    CUDA_OK(cudaMemcpyAsync(g.hResult, g.dResult, nv * sizeof(double), cudaMemcpyDeviceToHost, g.stream));
    CUDA_OK(cudaStreamSynchronize(g.stream));
    for (int i = 0; i < nv; ++i) out[i] = g.hResult[i];
}

void zeroResult(int nv) { CUDA_OK(cudaMemsetAsync(g.dResult, 0, nv * sizeof(double), g.stream)); }

void fillShard(const Qureg& q, qreal vr, qreal vi) {
    QuregRec* r = rec(q);
    const int vec = pickVec(r->n);
    const u64 nVec = (u64)r->n / vec;
    dispatchV(vec, [&](auto V) {
        k_fill<decltype(V)::value><<<blocksFor(nVec), kBlock, 0, g.stream>>>(r->re, r->im, nVec, vr, vi);
    });
    launched();
}

// set the amplitude with GLOBAL index gi, if this shard owns it
void setOneGlobal(const Qureg& q, idx_t gi, qreal vr, qreal vi) {
    QuregRec* r = rec(q);
    const idx_t off = (idx_t)q.chunkId * r->n;
    if (gi < off || gi >= off + r->n) return;
    k_setOne<<<1, 1, 0, g.stream>>>(r->re, r->im, (u64)(gi - off), vr, vi);
    launched();
}

// ---- helpers shared by several entry points ------------------------------------------------------------

qreal readOne(const Qureg& q, idx_t index, bool imag) {
    QuregRec* r = live(q);
    index = (idx_t)r->map.physIndex((u64)index);  // where that logical amplitude currently lives
    const int owner = (int)(index / r->n);
    qreal v = 0;
    if (owner == q.chunkId) {
        const qreal* src = (imag ? r->im : r->re) + (index - (idx_t)owner * r->n);
        CUDA_OK(cudaMemcpyAsync(&v, src, sizeof(qreal), cudaMemcpyDeviceToHost, g.stream));
        CUDA_OK(cudaStreamSynchronize(g.stream));
    }
    if (g.world > 1) comm::broadcastBytes(&v, sizeof v, owner);  // replaces MPI_Bcast, QuEST_cpu_distributed.c:208
    return v;
}

double sumSquares(const Qureg& q) {
    QuregRec* r = live(q);
    const int vec = pickVec(r->n);
    const u64 nVec = (u64)r->n / vec;
    dispatchV(vec, [&](auto V) {
        k_sumSquares<decltype(V)::value><<<reduceBlocks(nVec), kBlock, 0, g.stream>>>(r->re, r->im, nVec, g.dPartials,
                                                                                     g.dTicket, g.dResult);
    });
    launched();
    double s;
    fetchResult(1, &s);
    return s;
}

// All index bits of the shard in one pass (k_probAllZero), reduced over the shards, stored per LOGICAL qubit.
void fillProbCache(const Qureg& q, QuregRec* r) {
    const int vec = pickVec(r->n);
    const u64 nVec = (u64)r->n / vec;
    const unsigned int blocks = (unsigned int)std::min<u64>(g.wideBlocks, (nVec + kBlock - 1) / kBlock);
    dispatchV(vec, [&](auto V) {
        k_probAllZero<decltype(V)::value><<<blocks, kBlock, 0, g.stream>>>(r->re, r->im, nVec, r->localBits,
                                                                           g.dWidePartials, g.dTicket, g.dWide);
    });
    launched();
    if (g.world > 1) {
        k_probAllShardBits<<<1, 1, 0, g.stream>>>(g.dWide, r->localBits, log2i((u64)g.world), q.chunkId);
        launched();
        comm::allReduceSum(g.dWide, kProbAll, g.stream);
    }
    CUDA_OK(cudaMemcpyAsync(g.hWide, g.dWide, kProbAll * sizeof(double), cudaMemcpyDeviceToHost, g.stream));
    CUDA_OK(cudaStreamSynchronize(g.stream));
    for (int lq = 0; lq < r->map.nQubits; ++lq) r->probZero[lq] = g.hWide[r->map.phys[lq]];
    r->probValid = true;
}

double probOfZero(const Qureg& q, int tLogical) {
    QuregRec* r = live(q);
    if (g.probCache && !r->exposed && r->map.nQubits < kProbAll - 1) {
        if (!r->probValid && ++r->probQueries >= 2) fillProbCache(q, r);
        if (r->probValid) return r->probZero[tLogical];
    }
    const int t = r->map.phys[tLogical];
    if (t >= r->localBits) {
        // the measured qubit selects the shard: shards with the bit set contribute 0, the others everything
        // (QuEST_cpu_distributed.c:1227-1247)
        if ((q.chunkId >> (t - r->localBits)) & 1) {
            zeroResult(1);
            double s;
            fetchResult(1, &s);
            return s;
        }
        return sumSquares(q);
    }
    const int vec = pickVec(r->n);
    const Planned p = makePlan(r->n, vec, {t}, 0);
    dispatchVT(vec, t, [&](auto V, auto T) {
        k_probZero<decltype(V)::value, decltype(T)::value><<<reduceBlocks(p.nTasks), kBlock, 0, g.stream>>>(
            r->re, r->im, p.nTasks, p.plan, g.dPartials, g.dTicket, g.dResult);
    });
    launched();
    double s;
    fetchResult(1, &s);
    return s;
}

bool isPinnedHost(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

// Deferred upload (opt-in, qb200_setDeferredUpload): the whole shard, from pinned host memory, goes out as 8 slabs (cut
// along the three highest index bits) on the copy stream and the call returns at once; slab s is announced by g.upEv[s].
// The caller must leave the host arrays alone until the next call that synchronises (any value-returning call,
// syncQuESTEnv).  Returns false when the transfer does not qualify (the caller then copies the usual, blocking way).
bool uploadDeferred(QuregRec* r, idx_t lo, idx_t hi, const qreal* reals, const qreal* imags) {
    if (!g.deferUpload || g.world > 1 || lo != 0 || hi != r->n || r->localBits < 24) return false;
    if (!isPinnedHost(reals) || !isPinnedHost(imags)) return false;
    if (!g.copy) {
        CUDA_OK(cudaStreamCreateWithFlags(&g.copy, cudaStreamNonBlocking));
        for (cudaEvent_t& e : g.upEv) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    // the copy stream must not overtake what the engine stream still does to this register
    cudaEvent_t now = g.upEv[7];
    CUDA_OK(cudaEventRecord(now, g.stream));
    CUDA_OK(cudaStreamWaitEvent(g.copy, now, 0));
    r->uploadPlan = SlabPlan{3, {r->localBits - 3, r->localBits - 2, r->localBits - 1}};
    const idx_t slab = r->n >> 3;
    for (int s = 0; s < 8; ++s) {
        CUDA_OK(cudaMemcpyAsync(r->re + s * slab, reals + s * slab, slab * sizeof(qreal), cudaMemcpyHostToDevice, g.copy));
        CUDA_OK(cudaMemcpyAsync(r->im + s * slab, imags + s * slab, slab * sizeof(qreal), cudaMemcpyHostToDevice, g.copy));
        CUDA_OK(cudaEventRecord(g.upEv[s], g.copy));
    }
    r->uploading = true;
    ++g.deferredUploads;
    return true;
}

double densDiagSum(const Qureg& q, int bit) {
    QuregRec* r = canon(q);
    const u64 dim = 1ull << q.numQubitsRepresented;
    k_densDiagSum<<<reduceBlocks(dim), kBlock, 0, g.stream>>>(r->re, dim, bit, (u64)q.chunkId * r->n, (u64)r->n,
                                                             g.dPartials, g.dTicket, g.dResult);
    launched();
    double s;
    fetchResult(1, &s);
    return s;
}


}  // namespace

// ---------------------------------------------------------------------------------------------------------
// environment                               (replaces QuEST_gpu.cu:234-290 / QuEST_cpu_distributed.c:135-197)
// ---------------------------------------------------------------------------------------------------------

extern "C" {

QuESTEnv createQuESTEnv(void) {
    QuESTEnv env;
    env.rank = 0;
    env.numRanks = 1;
    if (g.up) {
        std::printf("ERROR: Trying to initialize QuESTEnv multiple times. Ignoring...\n");
        env.rank = g.rank;
        env.numRanks = g.world;
        return env;
    }
    // one process per GPU: rank / world size / local device come from the launcher's environment
    // (torchrun: RANK, WORLD_SIZE, LOCAL_RANK; overridable with QUEST_B200_RANK / _WORLD / _DEVICE)
    comm::Launch ln = comm::readLaunchEnv();
    int nDev = 0;
    cudaError_t e = cudaGetDeviceCount(&nDev);
    if (e != cudaSuccess || nDev == 0) {
        std::printf("!!!\nquest-b200: no CUDA device available (%s). This library has no CPU fallback.\n!!!\nexiting..\n",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        std::fflush(stdout);
        std::exit(EXIT_FAILURE);
    }
    g.device = ln.device % nDev;
    CUDA_OK(cudaSetDevice(g.device));
    cudaDeviceProp prop;
    CUDA_OK(cudaGetDeviceProperties(&prop, g.device));
    g.numSMs = prop.multiProcessorCount;
    CUDA_OK(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
    g.ownsStream = true;
    CUDA_OK(cudaMalloc(&g.dPartials, 2 * kMaxReduceBlocks * sizeof(double)));
    CUDA_OK(cudaMalloc(&g.dTicket, sizeof(unsigned int)));
    CUDA_OK(cudaMemset(g.dTicket, 0, sizeof(unsigned int)));
    CUDA_OK(cudaMalloc(&g.dResult, 4 * sizeof(double)));
    CUDA_OK(cudaHostAlloc(&g.hResult, 4 * sizeof(double), cudaHostAllocDefault));
    g.wideBlocks = (unsigned int)g.numSMs * 4;
    CUDA_OK(cudaMalloc(&g.dWidePartials, (size_t)kProbAll * g.wideBlocks * sizeof(double)));
    CUDA_OK(cudaMalloc(&g.dWide, kProbAll * sizeof(double)));
    CUDA_OK(cudaHostAlloc(&g.hWide, kProbAll * sizeof(double), cudaHostAllocDefault));
    // the fused pass keeps 32 KiB of shared memory per CTA and wants 3 CTAs per SM: ask for the large carve-out
    for (auto kernel : {k_tile<kTierCore, false>, k_tile<kTierCore, true>, k_tile<kTierGates, false>, k_tile<kTierGates, true>,
                        k_tile<kTierNoise, false>, k_tile<kTierNoise, true>})
        CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    if (const char* s = std::getenv("QUEST_B200_PROB_CACHE")) g.probCache = std::atoi(s) != 0;
    if (const char* s = std::getenv("QUEST_B200_FUSE")) g.fuse = std::atoi(s) != 0;
    if (const char* s = std::getenv("QUEST_B200_FMA")) g.contract = std::atoi(s) != 0;
    if (const char* s = std::getenv("QUEST_B200_PEEPHOLE")) g.peephole = std::atoi(s) != 0;
    if (const char* s = std::getenv("QUEST_B200_REORDER")) g.reorder = std::atoi(s) != 0;
    if (const char* s = std::getenv("QUEST_B200_REMAP")) g.remap = std::atoi(s) != 0;
    if (const char* s = std::getenv("QUEST_B200_VEC")) {
        int v = std::atoi(s);
        if (v == 1 || v == 2 || v == 4) g.vecCap = v;
    }
    g.rank = ln.rank;
    g.world = ln.world;
    g.up = true;
    if (g.world > 1) {
        comm::init(ln, g.device, g.stream);
        if (const char* s = std::getenv("QUEST_B200_OVERLAP")) g.overlap = std::atoi(s) != 0;
        if (const char* s = std::getenv("QUEST_B200_SWAP_CTAS")) g.swapCtasPerSm = std::max(1, std::min(8, std::atoi(s)));
        if (const char* s = std::getenv("QUEST_B200_SWAP_SLIM")) g.swapSlim = std::atoi(s) != 0;
        if (const char* s = std::getenv("QUEST_B200_SWAP_THREADS")) g.swapThreads = std::max(32, std::min(256, std::atoi(s) & ~31));
        // the exchange kernel shares SMs with the fused pass: both must ask for the same shared-memory carve-out
        CUDA_OK(cudaFuncSetAttribute(k_swapPeer<4>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CUDA_OK(cudaFuncSetAttribute(k_swapPeer<2>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CUDA_OK(cudaFuncSetAttribute(k_swapPeer<1>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CUDA_OK(cudaFuncSetAttribute(k_swapPeerSlim<2>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CUDA_OK(cudaFuncSetAttribute(k_swapPeerSlim<1>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        if (const char* s = std::getenv("QUEST_B200_OVERLAP_MIN_BITS")) g.overlapMinBits = std::max(8, std::atoi(s));
        int lowPrio = 0, highPrio = 0;
        CUDA_OK(cudaDeviceGetStreamPriorityRange(&lowPrio, &highPrio));
        CUDA_OK(cudaStreamCreateWithPriority(&g.comm, cudaStreamNonBlocking, highPrio));
        for (cudaEvent_t& e : g.slabEv) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (cudaEvent_t& e : g.xDone) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        CUDA_OK(cudaEventCreateWithFlags(&g.commDone, cudaEventDisableTiming));
    }
    env.rank = g.rank;
    env.numRanks = g.world;
    seedQuESTDefault();
    return env;
}

void destroyQuESTEnv(QuESTEnv env) {
    (void)env;
    if (!g.up) {
        std::printf("ERROR: Trying to close QuESTEnv multiple times. Ignoring\n");
        return;
    }
    flushAll();
    CUDA_OK(cudaStreamSynchronize(g.stream));
    if (g.comm) {
        CUDA_OK(cudaStreamSynchronize(g.comm));
        cudaStreamDestroy(g.comm);
        for (cudaEvent_t e : g.slabEv) cudaEventDestroy(e);
        for (cudaEvent_t e : g.xDone) cudaEventDestroy(e);
        cudaEventDestroy(g.commDone);
    }
    if (g.copy) {
        CUDA_OK(cudaStreamSynchronize(g.copy));
        cudaStreamDestroy(g.copy);
        for (cudaEvent_t e : g.upEv) cudaEventDestroy(e);
    }
    if (g.world > 1) comm::shutdown();
    io::release();
    cudaFree(g.dPartials);
    cudaFree(g.dTicket);
    cudaFree(g.dResult);
    cudaFreeHost(g.hResult);
    cudaFree(g.dWidePartials);
    cudaFree(g.dWide);
    cudaFreeHost(g.hWide);
    if (g.ownsStream) cudaStreamDestroy(g.stream);
    g = Engine();
}

void syncQuESTEnv(QuESTEnv env) {
    (void)env;
    requireEngine();
    flushAll();
    CUDA_OK(cudaStreamSynchronize(g.stream));
    if (g.world > 1) comm::hostBarrier();
}

int syncQuESTSuccess(int successCode) {
    if (g.world > 1) return comm::allAnd(successCode);
    return successCode;
}

void reportQuESTEnv(QuESTEnv env) {
    if (env.rank != 0) return;
    cudaDeviceProp prop;
    CUDA_OK(cudaGetDeviceProperties(&prop, g.device));
    std::printf("EXECUTION ENVIRONMENT:\n");
    std::printf("Running on %d GPU%s (quest-b200, one process per GPU), device %d: %s, %d SMs\n", env.numRanks,
                env.numRanks > 1 ? "s" : "", g.device, prop.name, prop.multiProcessorCount);
    std::printf("Number of ranks is %d\n", env.numRanks);
    std::printf("OpenMP disabled\n");
    std::printf("Precision: size of qreal is %ld bytes\n", sizeof(qreal));
}

void getEnvironmentString(QuESTEnv env, Qureg qureg, char str[200]) {
    std::snprintf(str, 200, "%dqubits_GPU_%dranks_b200", qureg.numQubitsInStateVec, env.numRanks);
}

// every rank must draw the same random numbers: rank 0's key is shared (QuEST_cpu_distributed.c:1294-1305)
void seedQuESTDefault(void) {
    unsigned long int key[2];
    getQuESTDefaultSeedKey(key);
    if (g.up && g.world > 1) comm::broadcastBytes(key, sizeof key, 0);
    init_by_array(key, 2);
}

// ---------------------------------------------------------------------------------------------------------
// registers                                                              (replaces QuEST_gpu.cu:156-232)
// ---------------------------------------------------------------------------------------------------------

void statevec_createQureg(Qureg* qureg, int numQubits, QuESTEnv env) {
    requireEngine();
    std::memset(qureg, 0, sizeof *qureg);
    const idx_t total = 1LL << numQubits;
    const idx_t per = total / env.numRanks;
    if (per < 1) {
        std::printf("!!!\nquest-b200: %d qubits cannot be sharded over %d GPUs\n!!!\nexiting..\n", numQubits, env.numRanks);
        std::exit(EXIT_FAILURE);
    }
    QuregRec* r = new QuregRec();
    r->n = per;
    r->localBits = log2i((u64)per);
    r->map.reset(numQubits, r->localBits);
    // no host mirror and no pair buffer: the shard (2 * per * sizeof(qreal) bytes) is the only allocation
    cudaError_t e1 = cudaMalloc(&r->re, per * sizeof(qreal));
    cudaError_t e2 = e1 == cudaSuccess ? cudaMalloc(&r->im, per * sizeof(qreal)) : e1;
    if (e1 != cudaSuccess || e2 != cudaSuccess) {
        std::printf("Could not allocate memory!");  // reference text, QuEST_cpu.c:1216
        std::printf(" (%lld amplitudes of %zu bytes on device %d: %s)\n", per, 2 * sizeof(qreal), g.device,
                    cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
        std::fflush(stdout);
        std::exit(EXIT_FAILURE);
    }
    qureg->numQubitsInStateVec = numQubits;
    qureg->numAmpsTotal = total;
    qureg->numAmpsPerChunk = per;
    qureg->chunkId = env.rank;
    qureg->numChunks = env.numRanks;
    qureg->isDensityMatrix = 0;
    qureg->deviceStateVec.real = r->re;
    qureg->deviceStateVec.imag = r->im;
    qureg->pairStateVec.real = reinterpret_cast<qreal*>(r);
    qureg->firstLevelReduction = reinterpret_cast<qreal*>(g.dPartials);
    liveRegisters.push_back(r);
    if (g.world > 1) comm::registerShard(*qureg, r->re, r->im, per);
}

void statevec_destroyQureg(Qureg qureg, QuESTEnv env) {
    (void)env;
    QuregRec* r = rec(qureg);
    if (!r) return;
    r->queue.clear();  // gates never observed need not run
    if (g.pendingReg == r) settle();
    settleUpload(r);
    liveRegisters.erase(std::remove(liveRegisters.begin(), liveRegisters.end(), r), liveRegisters.end());
    CUDA_OK(cudaStreamSynchronize(g.stream));
    if (g.world > 1) comm::unregisterShard(qureg);
    CUDA_OK(cudaFree(r->re));
    CUDA_OK(cudaFree(r->im));
    delete r;
}

// host -> device slice copy; complete on return, as in the reference (QuEST_gpu.cu:34-49)
void statevec_setAmps(Qureg qureg, idx_t startInd, qreal* reals, qreal* imags, idx_t numAmps) {
    QuregRec* r = (startInd == 0 && numAmps == qureg.numAmpsTotal) ? fresh(qureg) : canon(qureg);
    dirty(r);
    const idx_t off = (idx_t)qureg.chunkId * r->n;
    idx_t lo = std::max<idx_t>(startInd - off, 0), hi = std::min<idx_t>(startInd - off + numAmps, r->n);
    if (hi > lo) {
        const idx_t src = lo + off - startInd;
        if (uploadDeferred(r, lo, hi, reals + src, imags + src)) return;
        io::upload(r->re + lo, reals + src, (hi - lo) * sizeof(qreal), g.stream);
        io::upload(r->im + lo, imags + src, (hi - lo) * sizeof(qreal), g.stream);
    }
    CUDA_OK(cudaStreamSynchronize(g.stream));
}

// Pending gates are executed under the old setting first.
void qb200_setReordering(int enabled) {
    requireEngine();
    flushAll();
    g.reorder = enabled != 0;
}
int qb200_getReordering(void) { return g.reorder ? 1 : 0; }

void qb200_setDeferredUpload(int enabled) {
    requireEngine();
    flushAll();
    g.deferUpload = enabled != 0;
}
unsigned long long qb200_getDeferredUploadCount(void) { return g.deferredUploads; }
unsigned long long qb200_getPassesUnderUploadCount(void) { return g.passesUnderUpload; }

void qb200_copyAmpsToHost(Qureg qureg, idx_t start, idx_t n, qreal* reals, qreal* imags) {
    QuregRec* r = canon(qureg);
    io::download(reals, r->re + start, n * sizeof(qreal), g.stream);
    io::download(imags, r->im + start, n * sizeof(qreal), g.stream);
    CUDA_OK(cudaStreamSynchronize(g.stream));
}

// Pending gates are executed under the old setting first.
void qb200_setOverlap(int enabled) {
    requireEngine();
    flushAll();
    g.overlap = enabled != 0;
}
unsigned long long qb200_getOverlappedExchangeCount(void) { return g.overlapped; }

unsigned long long qb200_getPipelinedTransferCount(void) { return io::pipelinedTransfers(); }

void statevec_cloneQureg(Qureg dst, Qureg src) {
    QuregRec *d = fresh(dst), *s = live(src);
    d->map = s->map;  // same layout as the source
    CUDA_OK(cudaMemcpyAsync(d->re, s->re, d->n * sizeof(qreal), cudaMemcpyDeviceToDevice, g.stream));
    CUDA_OK(cudaMemcpyAsync(d->im, s->im, d->n * sizeof(qreal), cudaMemcpyDeviceToDevice, g.stream));
}


qreal statevec_getRealAmp(Qureg qureg, idx_t index) { return readOne(qureg, index, false); }
qreal statevec_getImagAmp(Qureg qureg, idx_t index) { return readOne(qureg, index, true); }

int statevec_compareStates(Qureg a, Qureg b, qreal precision) {
    QuregRec *ra = canon(a), *rb = canon(b);
    const int vec = pickVec(ra->n);
    const u64 nVec = (u64)ra->n / vec;
    dispatchV(vec, [&](auto V) {
        k_countDiffs<decltype(V)::value><<<reduceBlocks(nVec), kBlock, 0, g.stream>>>(
            ra->re, ra->im, rb->re, rb->im, nVec, precision, g.dPartials, g.dTicket, g.dResult);
    });
    launched();
    double cnt;
    fetchResult(1, &cnt);
    return cnt == 0.0;
}

// ---------------------------------------------------------------------------------------------------------
// initial states                                                          (replaces QuEST_gpu.cu:346-466)
// ---------------------------------------------------------------------------------------------------------

void statevec_initZeroState(Qureg q) {
    QuregRec* r = fresh(q);  // the whole state is overwritten: pending gates and the qubit layout are dropped
    CUDA_OK(cudaMemsetAsync(r->re, 0, r->n * sizeof(qreal), g.stream));
    CUDA_OK(cudaMemsetAsync(r->im, 0, r->n * sizeof(qreal), g.stream));
    setOneGlobal(q, 0, 1, 0);
}

void statevec_initPlusState(Qureg q) {
    fresh(q);
    const idx_t total = q.numAmpsPerChunk * q.numChunks;
    fillShard(q, 1.0 / sqrt((double)(qreal)total), 0);
}

void statevec_initClassicalState(Qureg q, idx_t stateInd) {
    QuregRec* r = fresh(q);
    CUDA_OK(cudaMemsetAsync(r->re, 0, r->n * sizeof(qreal), g.stream));
    CUDA_OK(cudaMemsetAsync(r->im, 0, r->n * sizeof(qreal), g.stream));
    setOneGlobal(q, stateInd, 1, 0);
}

void statevec_initStateDebug(Qureg q) {
    QuregRec* r = fresh(q);
    k_initDebug<<<blocksFor((u64)r->n), kBlock, 0, g.stream>>>(r->re, r->im, (u64)r->n, (u64)q.chunkId * r->n);
    launched();
}

void statevec_initStateOfSingleQubit(Qureg* q, int qubitId, int outcome) {
    QuregRec* r = fresh(*q);
    const idx_t total = q->numAmpsPerChunk * q->numChunks;
    const qreal norm = 1.0 / sqrt((qreal)total / 2.0);
    k_initSingleQubit<<<blocksFor((u64)r->n), kBlock, 0, g.stream>>>(r->re, r->im, (u64)r->n, (u64)q->chunkId * r->n,
                                                                    qubitId, outcome, norm);
    launched();
}

void densmatr_initPlusState(Qureg q) {
    fresh(q);
    const idx_t dim = 1LL << q.numQubitsRepresented;
    fillShard(q, 1.0 / ((qreal)dim), 0);
}

void densmatr_initClassicalState(Qureg q, idx_t stateInd) {
    QuregRec* r = fresh(q);
    const idx_t dim = 1LL << q.numQubitsRepresented;
    CUDA_OK(cudaMemsetAsync(r->re, 0, r->n * sizeof(qreal), g.stream));
    CUDA_OK(cudaMemsetAsync(r->im, 0, r->n * sizeof(qreal), g.stream));
    setOneGlobal(q, (dim + 1) * stateInd, 1, 0);
}

// A whole copy of a (much smaller) state-vector next to a density-matrix shard.  One GPU: the register itself.
// Sharded: an ncclAllGather of the shards into a temporary - the exchange the reference makes with MPI_Allgather
// into pairStateVec (copyVecIntoMatrixPairState, QuEST_cpu_distributed.c:373-403).
struct WholeState {
    const qreal *re, *im;
    qreal* owned;
};
static WholeState wholeStateOf(const Qureg& psi) {
    QuregRec* p = canon(psi);
    if (g.world == 1) return WholeState{p->re, p->im, nullptr};
    const size_t shardBytes = (size_t)p->n * sizeof(qreal), total = shardBytes * g.world;
    qreal* buf = nullptr;
    CUDA_OK(cudaMalloc(&buf, 2 * total));
    qreal* bufIm = buf + (size_t)p->n * g.world;
    comm::allGatherBytes(p->re, buf, shardBytes, g.stream);
    comm::allGatherBytes(p->im, bufIm, shardBytes, g.stream);
    return WholeState{buf, bufIm, buf};
}
static void release(WholeState& w) {
    if (!w.owned) return;
    CUDA_OK(cudaStreamSynchronize(g.stream));
    CUDA_OK(cudaFree(w.owned));
    w.owned = nullptr;
}

// this rank's shard holds whole columns [chunkId * n / dim, ...) of the density matrix (QuEST_cpu.c:1107-1150)
void densmatr_initPureState(Qureg rho, Qureg psi) {
    if (rho.numAmpsPerChunk < psi.numAmpsTotal) comm::unsupported("a density-matrix shard smaller than one column");
    WholeState w = wholeStateOf(psi);
    QuregRec* d = fresh(rho);
    k_outerProduct<<<blocksFor((u64)d->n), kBlock, 0, g.stream>>>(d->re, d->im, w.re, w.im, (u64)psi.numAmpsTotal,
                                                                 (u64)d->n, (u64)rho.chunkId * (u64)d->n);
    launched();
    release(w);
}

// ---------------------------------------------------------------------------------------------------------
// gates                                                                  (replaces QuEST_gpu.cu:531-1330)
// ---------------------------------------------------------------------------------------------------------

void statevec_compactUnitary(Qureg q, const int t, Complex alpha, Complex beta) { submit(q, compactOp(t, 0, alpha, beta)); }
void statevec_controlledCompactUnitary(Qureg q, const int c, const int t, Complex alpha, Complex beta) {
    submit(q, compactOp(t, 1ull << c, alpha, beta));
}
void statevec_unitary(Qureg q, const int t, ComplexMatrix2 u) { submit(q, matrixOp(t, 0, u)); }
void statevec_controlledUnitary(Qureg q, const int c, const int t, ComplexMatrix2 u) { submit(q, matrixOp(t, 1ull << c, u)); }
void statevec_multiControlledUnitary(Qureg q, int* ctrls, const int n, const int t, ComplexMatrix2 u) {
    submit(q, matrixOp(t, maskOf(ctrls, n), u));
}
void statevec_hadamard(Qureg q, const int t) {
    PendingOp op = denseOp(kFHadamard, t, 0);
    op.c[0] = (qreal)(1.0 / sqrt(2));
    submit(q, op);
}
void statevec_pauliX(Qureg q, const int t) { submit(q, denseOp(kFSwap, t, 0)); }
void statevec_controlledNot(Qureg q, const int c, const int t) { submit(q, denseOp(kFSwap, t, 1ull << c)); }
static void submitY(Qureg q, int t, u64 ctrl, qreal fac) {
    PendingOp op = denseOp(kFPauliY, t, ctrl);
    op.c[0] = fac;
    submit(q, op);
}
void statevec_pauliY(Qureg q, const int t) { submitY(q, t, 0, 1); }
void statevec_pauliYConj(Qureg q, const int t) { submitY(q, t, 0, -1); }
void statevec_controlledPauliY(Qureg q, const int c, const int t) { submitY(q, t, 1ull << c, 1); }
void statevec_controlledPauliYConj(Qureg q, const int c, const int t) { submitY(q, t, 1ull << c, -1); }

void statevec_phaseShiftByTerm(Qureg q, const int t, Complex term) { submit(q, phaseOp(1ull << t, term.real, term.imag)); }
void statevec_controlledPhaseShift(Qureg q, const int q1, const int q2, qreal angle) {
    submit(q, phaseOp((1ull << q1) | (1ull << q2), (qreal)cos((double)angle), (qreal)sin((double)angle)));
}
void statevec_multiControlledPhaseShift(Qureg q, int* ctrls, int n, qreal angle) {
    submit(q, phaseOp(maskOf(ctrls, n), (qreal)cos((double)angle), (qreal)sin((double)angle)));
}
void statevec_controlledPhaseFlip(Qureg q, const int q1, const int q2) { submit(q, denseOp(kFFlip, -1, (1ull << q1) | (1ull << q2))); }
void statevec_multiControlledPhaseFlip(Qureg q, int* ctrls, int n) { submit(q, denseOp(kFFlip, -1, maskOf(ctrls, n))); }

// ---------------------------------------------------------------------------------------------------------
// reductions                                                             (replaces QuEST_gpu.cu:1111-1164,1335-1840)
// ---------------------------------------------------------------------------------------------------------


// The reference sums on the host (serial Kahan, QuEST_cpu_local.c:123-151; its GPU build copies the whole state
// back, QuEST_gpu.cu:1134-1164).  Here it is one read-only pass in HBM with double accumulators.
qreal statevec_calcTotalProb(Qureg q) { return (qreal)sumSquares(q); }
qreal densmatr_calcPurity(Qureg q) { return (qreal)sumSquares(q); }


qreal statevec_calcProbOfOutcome(Qureg q, const int t, int outcome) {
    qreal p = (qreal)probOfZero(q, t);
    if (outcome == 1) p = 1.0 - p;  // defined this way also for unnormalised states (QuEST_cpu_local.c:271)
    return p;
}

Complex statevec_calcInnerProduct(Qureg bra, Qureg ket) {
    QuregRec *b = canon(bra), *k = canon(ket);
    const int vec = pickVec(b->n);
    const u64 nVec = (u64)b->n / vec;
    dispatchV(vec, [&](auto V) {
        k_innerProduct<decltype(V)::value><<<reduceBlocks(nVec), kBlock, 0, g.stream>>>(
            b->re, b->im, k->re, k->im, nVec, g.dPartials, g.dTicket, g.dResult);
    });
    launched();
    double s[2];
    fetchResult(2, s);
    Complex out;
    out.real = (qreal)s[0];
    out.imag = (qreal)s[1];
    return out;
}


qreal densmatr_calcTotalProb(Qureg q) { return (qreal)densDiagSum(q, -1); }

qreal densmatr_calcProbOfOutcome(Qureg q, const int t, int outcome) {
    qreal p = (qreal)densDiagSum(q, t);
    if (outcome == 1) p = 1.0 - p;
    return p;
}

// each rank sums over the columns it holds, the partial sums meet in fetchResult (QuEST_cpu_distributed.c:407-420)
qreal densmatr_calcFidelity(Qureg rho, Qureg psi) {
    if (rho.numAmpsPerChunk < psi.numAmpsTotal) comm::unsupported("a density-matrix shard smaller than one column");
    WholeState w = wholeStateOf(psi);
    QuregRec* d = canon(rho);
    const u64 dim = (u64)psi.numAmpsTotal, nCols = (u64)d->n / dim;
    k_densFidelity<<<reduceBlocks(dim), kBlock, 0, g.stream>>>(d->re, d->im, w.re, w.im, dim, (u64)rho.chunkId * nCols,
                                                              nCols, g.dPartials, g.dTicket, g.dResult);
    launched();
    double s;
    fetchResult(1, &s);
    release(w);
    return (qreal)s;
}

// ---------------------------------------------------------------------------------------------------------
// collapse                                                               (replaces QuEST_gpu.cu:1847-1966)
// ---------------------------------------------------------------------------------------------------------

void statevec_collapseToKnownProbOutcome(Qureg q, const int tLogical, int outcome, qreal prob) {
    QuregRec* r = live(q);
    dirty(r);
    const int t = r->map.phys[tLogical];
    const qreal renorm = 1 / sqrt((double)prob);
    const int vec = pickVec(r->n);
    if (t >= r->localBits) {
        // whole shard is either kept (renormalised) or cleared (QuEST_cpu_distributed.c:1274-1292)
        const int bit = (q.chunkId >> (t - r->localBits)) & 1;
        if (bit == outcome) {
            const u64 nVec = (u64)r->n / vec;
            dispatchV(vec, [&](auto V) {
                k_scaleAll<decltype(V)::value><<<blocksFor(nVec), kBlock, 0, g.stream>>>(r->re, r->im, nVec, renorm);
            });
            launched();
        } else {
            CUDA_OK(cudaMemsetAsync(r->re, 0, r->n * sizeof(qreal), g.stream));
            CUDA_OK(cudaMemsetAsync(r->im, 0, r->n * sizeof(qreal), g.stream));
        }
        return;
    }
    const Planned p = makePlan(r->n, vec, {t}, 0);
    const int tvb = t - log2i(vec);
    dispatchVT(vec, t, [&](auto V, auto T) {
        k_collapse<decltype(V)::value, decltype(T)::value>
            <<<blocksFor(p.nTasks), kBlock, 0, g.stream>>>(r->re, r->im, p.nTasks, p.plan, tvb, outcome, renorm);
    });
    launched();
}

// recorded like a noise channel: it acts on the (row bit, column bit) block structure of the measured qubit
void densmatr_collapseToKnownProbOutcome(Qureg q, const int t, int outcome, qreal prob) {
    PendingOp op = denseOp(kFDensCollapse, t, 0);
    op.target2 = t + q.numQubitsRepresented;
    op.c[0] = prob;
    op.c[1] = (qreal)outcome;
    submit(q, op);
}

// ---------------------------------------------------------------------------------------------------------
// density-matrix noise                                                   (replaces QuEST_gpu.cu:1997-2275)
// ---------------------------------------------------------------------------------------------------------


static void submitNoise(Qureg q, int kind, int t, qreal c0, qreal c1, qreal c2) {
    PendingOp op = denseOp(kind, t, 0);
    op.target2 = t + q.numQubitsRepresented;  // the qubit's column bit in the flattened density matrix
    op.c[0] = c0;
    op.c[1] = c1;
    op.c[2] = c2;
    submit(q, op);
}

// scalars formed exactly as the reference does on the host (QuEST_cpu.c:67-70,119,168-169)
void densmatr_oneQubitDephase(Qureg q, const int t, qreal dephase) {
    const qreal retain = 1 - dephase;
    submitNoise(q, kFDephase, t, retain, 0, 0);
}

void densmatr_oneQubitDepolarise(Qureg q, const int t, qreal level) {
    if (level == 0) return;
    const qreal retain = 1 - level;
    submitNoise(q, kFDepolarise, t, retain, level, retain);
}

void densmatr_oneQubitDamping(Qureg q, const int t, qreal damping) {
    if (damping == 0) return;
    const qreal retain = 1 - damping;
    const qreal dephase = sqrt((double)retain);
    submitNoise(q, kFDamping, t, dephase, damping, retain);
}

// Diagonal in the amplitude index, so it runs in whatever layout the register is in: the four bits are looked up
// in the qubit map and tested on the global index.
static void dephase2Now(const Qureg& q, QuregRec* r, int q1, int q2, qreal retain) {
    const int n = q.numQubitsRepresented;
    const QubitMap& m = r->map;
    const int vec = pickVec(r->n);
    const u64 nVec = (u64)r->n / vec;
    dispatchV(vec, [&](auto V) {
        k_dephase2<decltype(V)::value><<<blocksFor(nVec), kBlock, 0, g.stream>>>(
            r->re, r->im, nVec, (u64)q.chunkId * r->n, 1ull << m.phys[q1], 1ull << m.phys[q1 + n], 1ull << m.phys[q2],
            1ull << m.phys[q2 + n], retain);
    });
    launched();
}

void densmatr_twoQubitDephase(Qureg q, const int q1, const int q2, qreal dephase) {
    QuregRec* r = live(q);
    dirty(r);
    const qreal retain = 1 - dephase;
    dephase2Now(q, r, q1, q2, retain);
}

// Make every listed qubit of a sharded register local (half-shard slot exchanges), none evicting another.
static void bringLocal(const Qureg& q, QuregRec* r, std::initializer_list<int> qubits, const char* what) {
    if (g.world == 1) return;
    u64 keep = 0;
    for (int t : qubits) keep |= 1ull << t;
    for (int t : qubits) {
        if (r->map.phys[t] >= r->localBits) {
            const SlotSwap sw = g.remap ? planSwapIn(r->map, t, keep) : SlotSwap{-1, -1};
            if (sw.globalSlot < 0) comm::unsupported(what);
            exchangeSlots(q, sw);
        }
        r->map.touch(t);
    }
}

// The reference's distributed version exchanges whole shards up to three times (QuEST_cpu_distributed.c:724-806).
// Here the column bits of the two qubits are first made local (at most log2(G) half-shard exchanges, none when an
// earlier gate or channel already brought them in), after which the three sweeps are the single-GPU ones.
void densmatr_twoQubitDepolarise(Qureg q, int q1, int q2, qreal level) {
    if (level == 0) return;
    QuregRec* r = live(q);
    dirty(r);
    const int n = q.numQubitsRepresented;
    bringLocal(q, r, {q1, q2, q1 + n, q2 + n},
               "two-qubit depolarising across shards (with QUEST_B200_REMAP=0, or a shard too small to re-home its qubits)");
    // host scalars exactly as QuEST_cpu_local.c:43-48
    qreal eta = 2 / level;
    qreal delta = eta - 1 - sqrt((double)((eta - 1) * (eta - 1) - 1));
    qreal gamma = 1 + delta;
    gamma = 1 / (gamma * gamma * gamma);
    // all four bits are local now: ONE pass over the 16-element blocks they span (k_depol2)
    const QubitMap& m = r->map;
    const int bits[4] = {m.phys[q1], m.phys[q1 + n], m.phys[q2], m.phys[q2 + n]};
    Depol2 d;
    for (int e = 0; e < 16; ++e) {
        d.offs[e] = 0;
        for (int j = 0; j < 4; ++j)
            if ((e >> j) & 1) d.offs[e] |= 1ull << bits[j];
    }
    d.retain = 1 - level;
    d.delta = delta;
    d.gamma = gamma;
    const Planned p = makePlan(r->n, 1, {bits[0], bits[1], bits[2], bits[3]}, 0);
    k_depol2<<<blocksFor(p.nTasks), kBlock, 0, g.stream>>>(r->re, r->im, p.nTasks, p.plan, d);
    launched();
}

void densmatr_addDensityMatrix(Qureg a, qreal p, Qureg b) {
    QuregRec *ra = canon(a), *rb = canon(b);
    dirty(ra);
    const int vec = pickVec(ra->n);
    const u64 nVec = (u64)ra->n / vec;
    dispatchV(vec, [&](auto V) {
        k_mixDensity<decltype(V)::value><<<blocksFor(nVec), kBlock, 0, g.stream>>>(ra->re, ra->im, rb->re, rb->im, nVec, p);
    });
    launched();
}

// ---------------------------------------------------------------------------------------------------------
// quest_b200.h extension entry points
// ---------------------------------------------------------------------------------------------------------

void* qb200_getStream(void) { return (void*)g.stream; }

// Pending gates were recorded under the old setting: run them first.
void qb200_setArithmetic(int contracted) {
    requireEngine();
    flushAll();
    g.contract = contracted != 0;
}
int qb200_getArithmetic(void) { return g.contract ? 1 : 0; }

void qb200_setStream(void* stream) {
    requireEngine();
    flushAll();
    CUDA_OK(cudaStreamSynchronize(g.stream));
    if (g.ownsStream) cudaStreamDestroy(g.stream);
    g.stream = (cudaStream_t)stream;
    g.ownsStream = false;
}

unsigned long long qb200_getLaunchCount(void) { return g.launches; }

void qb200_setFusion(int enabled) {
    flushAll();
    g.fuse = enabled != 0;
}

int qb200_getFusion(void) { return g.fuse ? 1 : 0; }

// Sharded registers only.  Off: every dense gate on a shard-selecting qubit runs as a cross-shard pass
// (k_pairsPeer).  Switching drains the queues and restores every register's identity layout first.
void qb200_setRemap(int enabled) {
    for (QuregRec* r : liveRegisters) {
        Qureg q;
        std::memset(&q, 0, sizeof q);
        q.pairStateVec.real = reinterpret_cast<qreal*>(r);
        q.numAmpsPerChunk = r->n;
        q.chunkId = g.rank;
        q.numChunks = g.world;
        canon(q);
    }
    g.remap = enabled != 0;
}

unsigned long long qb200_getSlotExchangeCount(void) { return g.slotSwaps; }

void qb200_flush(void) { flushAll(); }

void qb200_setKernelTiming(int enabled) {
    g.timeFused = enabled != 0;
}

// total device time (ms) and count of the fused launches recorded since timing was enabled; clears the record
void qb200_getFusedKernelTime(double* totalMs, unsigned long long* launches) {
    CUDA_OK(cudaStreamSynchronize(g.stream));
    double sum = 0;
    const bool show = std::getenv("QUEST_B200_PRINT_PASS_TIMES") != nullptr;  // diagnostics: every pass's own device time
    unsigned long long passes = 0;
    for (auto& ev : g.fusedEvents) {
        float ms = 0;
        CUDA_OK(cudaEventElapsedTime(&ms, ev.e0, ev.e1));
        if (show) std::printf(ev.opensPass ? "[pass %.2f ms] " : "[+slab %.2f ms] ", ms);
        sum += ms;
        passes += ev.opensPass ? 1 : 0;
        cudaEventDestroy(ev.e0);
        cudaEventDestroy(ev.e1);
    }
    *totalMs = sum;
    *launches = passes;
    g.fusedEvents.clear();
}

void qb200_getFusionStats(unsigned long long* fusedLaunches, unsigned long long* fusedGates) {
    *fusedLaunches = g.fusedLaunches;
    *fusedGates = g.fusedGates;
}

int qb200_getDevice(void) { return g.device; }

void qb200_getDevicePointers(Qureg qureg, void** real, void** imag) {
    QuregRec* r = canon(qureg);
    dirty(r);
    r->exposed = true;  // the caller may write through these pointers at any time: no result is cached from now on
    *real = r->re;
    *imag = r->im;
}

// Host planning logic exposed for CPU-only tests (no CUDA call is made): what would rank `chunkId` of a register
// with 2^localBits amplitudes per shard do for a one-target gate?  out[0..7] = action, partner, amUpper, launches,
// nTasks, nIns, setMask, loMask; out[8..15] = insertion positions; out[16] = skipMask; out[17] = local ctrlMask.
void qb200_debugPlanGate(int localBits, int chunkId, int target, unsigned long long ctrlMask, int vec,
                         unsigned long long* out) {
    const ShardedGate sg = planShardedGate(localBits, chunkId, target, ctrlMask, vec);
    out[0] = sg.action; out[1] = sg.partner; out[2] = sg.amUpper; out[3] = sg.launches;
    out[4] = sg.planned.nTasks; out[5] = sg.planned.plan.nIns; out[6] = sg.planned.plan.setMask;
    out[7] = sg.planned.plan.loMask;
    for (int i = 0; i < kMaxInsert; ++i) out[8 + i] = sg.planned.plan.pos[i];
    out[16] = sg.planned.plan.skipMask;
    out[17] = sg.ctrlMask;
}

// Host qubit-layout logic, exposed for CPU-only tests (no CUDA call): replays the slot exchanges submit() would
// perform for a sequence of dense gates (logical target, logical control mask) on a register of nQubits sharded into
// 2^localBits-amplitude shards.  eventsOut: 3 ints per gate (globalSlot or -1, localSlot, physical target after the
// exchange); physOut[nQubits]: final logical->physical map; canonOut: pairs (globalSlot, localSlot) that restore the
// identity layout.  Returns the number of canonicalising exchanges.
int qb200_debugRemap(int nQubits, int localBits, int nOps, const int* targets, const unsigned long long* masks,
                     int* eventsOut, int* physOut, int* canonOut) {
    QubitMap m;
    m.reset(nQubits, localBits);
    for (int i = 0; i < nOps; ++i) {
        SlotSwap sw{-1, -1};
        if (localBits >= 3 && m.phys[targets[i]] >= localBits) {
            sw = planSwapIn(m, targets[i], masks[i]);
            if (sw.globalSlot >= 0) m.applySwap(sw.globalSlot, sw.localSlot);
        }
        m.touch(targets[i]);
        eventsOut[3 * i] = sw.globalSlot;
        eventsOut[3 * i + 1] = sw.localSlot;
        eventsOut[3 * i + 2] = m.phys[targets[i]];
    }
    for (int q = 0; q < nQubits; ++q) physOut[q] = m.phys[q];
    SlotSwap back[192];
    const int n = planCanonicalize(m, back);
    for (int i = 0; i < n; ++i) {
        canonOut[2 * i] = back[i].globalSlot;
        canonOut[2 * i + 1] = back[i].localSlot;
    }
    return n;
}

// Host walk of a recorded queue of logical gates (qb::planWalk), exposed for CPU-only tests (no CUDA call).  The layout
// and use history travel in and out through mapState (phys[64], lastUse[64], prevUse[64], tick = 193 words), so that a
// test can replay several flushes in a row.  stepsOut: 3 ints per step - exchange: (1, globalSlot, localSlot); segment:
// (0, first, count).  Returns the number of steps.
int qb200_debugWalk(int nQubits, int localBits, int nOps, const int* targets, const int* targets2,
                    const unsigned long long* masks, unsigned long long* mapState, int* stepsOut) {
    QubitMap m;
    m.reset(nQubits, localBits);
    if (mapState[192]) {
        for (int q = 0; q < 64; ++q) {
            m.phys[q] = (int)mapState[q];
            m.logicalAt[m.phys[q]] = q;
            m.lastUse[q] = mapState[64 + q];
            m.prevUse[q] = mapState[128 + q];
        }
        m.tick = mapState[192];
    }
    std::vector<WalkOp> ops(nOps);
    for (int i = 0; i < nOps; ++i) ops[i] = WalkOp{targets[i], targets2[i], masks[i]};
    const std::vector<WalkStep> steps = planWalk(m, ops);
    for (size_t k = 0; k < steps.size(); ++k) {
        stepsOut[3 * k] = steps[k].isExchange;
        stepsOut[3 * k + 1] = steps[k].isExchange ? steps[k].sw.globalSlot : steps[k].first;
        stepsOut[3 * k + 2] = steps[k].isExchange ? steps[k].sw.localSlot : steps[k].count;
    }
    for (int q = 0; q < 64; ++q) {
        mapState[q] = (unsigned long long)m.phys[q];
        mapState[64 + q] = m.lastUse[q];
        mapState[128 + q] = m.prevUse[q];
    }
    mapState[192] = m.tick;
    return (int)steps.size();
}

// The flush-time peephole (qb::fuseAdjacentPairs), exposed for CPU-only tests (no CUDA call): kindsOut / firstOut receive,
// per op of the fused list, its kind and the index of the (first) input op it came from; returns the fused length.
int qb200_debugPeephole(int nOps, const int* kinds, const int* targets, const unsigned long long* masks, int* kindsOut,
                        int* firstOut) {
    std::vector<PendingOp> ops(nOps);
    for (int i = 0; i < nOps; ++i) {
        std::memset(&ops[i], 0, sizeof(PendingOp));
        ops[i].kind = kinds[i];
        ops[i].target = targets[i];
        ops[i].target2 = -1;
        ops[i].mask = masks[i];
        ops[i].c[0] = (qreal)i;  // a tag: which input op the coefficients came from
    }
    const std::vector<PendingOp> fused = fuseAdjacentPairs(ops);
    for (size_t k = 0; k < fused.size(); ++k) {
        kindsOut[k] = fused[k].kind;
        firstOut[k] = fused[k].kind == kFHadMatrix ? (int)fused[k].c[8] : (int)fused[k].c[0];
    }
    return (int)fused.size();
}

// Host canonicalisation logic (qb::planCanonicalize) for ANY layout, exposed for CPU-only tests (no CUDA call):
// phys[nQubits] = logical -> bit position; out = pairs (globalSlot, localSlot); returns the number of exchanges.
int qb200_debugCanon(int nQubits, int localBits, const int* phys, int* out) {
    QubitMap m;
    m.reset(nQubits, localBits);
    for (int q = 0; q < nQubits; ++q) {
        m.phys[q] = phys[q];
        m.logicalAt[phys[q]] = q;
    }
    SlotSwap back[192];
    const int n = planCanonicalize(m, back);
    for (int i = 0; i < n; ++i) {
        out[2 * i] = back[i].globalSlot;
        out[2 * i + 1] = back[i].localSlot;
    }
    return n;
}

// Host slab logic (qb::planSlabs, qb::planSlotSwap), exposed for CPU-only tests (no CUDA call).  Returns the number of
// slab bits chosen (bitsOut[3]); swapOut, per slab and per rank of the pair (upper first): nTasks, nIns, setMask,
// pos[8] = 11 words each.
int qb200_debugSlabs(int localBits, unsigned long long busyMask, int want, int lowestAllowed, int vec, int localSlot,
                     int* bitsOut, unsigned long long* swapOut) {
    const SlabPlan sp = planSlabs(localBits, busyMask, want, lowestAllowed);
    for (int j = 0; j < sp.nBits; ++j) bitsOut[j] = sp.bits[j];
    unsigned long long* o = swapOut;
    for (int s = 0; s < (1 << sp.nBits); ++s)
        for (int upper = 1; upper >= 0; --upper) {
            const SwapLaunch sl = planSlotSwap(localBits, vec, localSlot, upper != 0, sp.nBits ? &sp : nullptr, s);
            o[0] = sl.active ? sl.planned.nTasks : 0;
            o[1] = sl.planned.plan.nIns;
            o[2] = sl.planned.plan.setMask;
            for (int i = 0; i < kMaxInsert; ++i) o[3 + i] = sl.planned.plan.pos[i];
            o += 11;
        }
    return sp.nBits;
}

// Host pass-planning logic of the fusion queue, exposed for CPU-only tests (no CUDA call): plans the given gate list
// exactly as flush() would for a shard of 2^localBits amplitudes.  Gates: kind (qb_fuse.h FusedKind), target, second target
// (-1 if none), mask.  passesOut, per pass: offset of its gates in orderOut, count, nTile, nSub, tileBits[kTileBits]
// (= 4 + 11 ints); orderOut: the gates of all passes in execution order (indices into the input list); subsOut, per
// sub-group in pass order: pass index, first (position in orderOut), count, rb[3], spectator (7 ints); descOut, per input
// gate: opcode, regMask, tileMask, outerMask (4 words).  Returns the number of passes; *nSubsOut receives the number of
// sub-groups.
int qb200_debugTilePlan(int localBits, int nOps, const int* kinds, const int* targets, const int* targets2,
                        const unsigned long long* masks, int reorder, int* passesOut, int* orderOut, int* subsOut,
                        unsigned long long* descOut, int* nSubsOut) {
    std::vector<PendingOp> ops(nOps);
    for (int i = 0; i < nOps; ++i) {
        std::memset(&ops[i], 0, sizeof(PendingOp));
        ops[i].kind = kinds[i];
        ops[i].target = targets[i];
        ops[i].target2 = targets2[i];
        ops[i].mask = masks[i];
    }
    const std::vector<TilePass> passes = planPasses(ops, localBits, 0, reorder != 0);
    int nSubs = 0, at = 0;
    for (size_t pi = 0; pi < passes.size(); ++pi) {
        const TilePass& p = passes[pi];
        int* o = passesOut + (4 + kTileBits) * pi;
        o[0] = at; o[1] = p.count; o[2] = p.nTile; o[3] = (int)p.subs.size();
        for (int j = 0; j < kTileBits; ++j) o[4 + j] = j < p.nTile ? p.tileBits[j] : -1;
        for (int j = 0; j < p.count; ++j) orderOut[at + j] = p.order[j];
        for (const TileSub& sg : p.subs) {
            int* so = subsOut + 7 * nSubs++;
            so[0] = (int)pi; so[1] = at + sg.first; so[2] = sg.count;
            for (int j = 0; j < kSubBits; ++j) so[3 + j] = sg.rb[j];
            so[6] = sg.spect;
            if (p.count == 1) continue;  // single-gate passes run through the single-gate kernels: no descriptor
            for (int j = sg.first; j < sg.first + sg.count; ++j) {
                const int k = p.order[j];
                const TileGate tg = describe(ops[k], p, sg);
                unsigned long long* d = descOut + 4 * k;
                d[0] = (unsigned long long)tg.opcode; d[1] = tg.regMask; d[2] = tg.tileMask; d[3] = tg.outerMask;
            }
        }
        at += p.count;
    }
    *nSubsOut = nSubs;
    return (int)passes.size();
}

}  // extern "C"

// shared with qb_comm.cu
namespace qb {
namespace hostapi {
cudaStream_t stream() { return g.stream; }
int numSMs() { return g.numSMs; }
int vecFor(long long n) { return pickVec(n); }
void countLaunch() { launched(); }
}  // namespace hostapi
}  // namespace qb
