// quest-b200 fused pass (sm_100a): ONE read and ONE write of the state for a whole run of recorded gates.
//
// Planning is done on the host (qb_fuse.h): a PASS is a run of gates whose dense targets, together with the lowest
// kTileLow index bits, fall on at most kTileBits index bits - the tile bits.  Here, every CTA
//   1. stages the 2^kTileBits amplitudes that differ only in the tile bits in shared memory.  They lie in HBM as
//      2^(kTileBits-kTileLow) rows of 2^kTileLow contiguous amplitudes per array, so each row is a whole 128-byte line
//      (more when padding bits extend the low run); the copies are asynchronous (LDGSTS), all in flight at once;
//   2. meanwhile decides, once per CTA, which gates of the pass its tile takes part in (controls / phase-mask bits
//      outside the tile are the same for all of its amplitudes);
//   3. runs the pass SUB-GROUP by sub-group: each thread pulls the 2^kSubBits amplitudes spanned by the sub-group's
//      register bits out of shared memory, applies the sub-group's gates in program order - the same per-pair
//      expressions as the single-gate kernels (qb_kernels.cuh), so results are bit-identical to gate-at-a-time
//      execution -, and puts them back; a barrier, and the next sub-group re-blocks the tile along other bits;
//   4. writes the tile back.
// Shared memory is addressed through qb::tilePlace, an XOR placement that keeps the re-blocking loads of any register
// bit choice (nearly) free of bank conflicts.  With 3 register bits the gate bodies are 4 pairs long and exist in only
// 3 positions each: the whole dispatch table is ~14 KB of code and stays in the 32 KB instruction cache, which the
// register-resident 4-qubit design it replaces (16 amplitudes per thread, ~55 KB of gate bodies) did not.
#pragma once
#include <cstddef>

#include "qb_fuse.h"
#include "qb_kernels.cuh"

namespace qb {

constexpr int kRegBits = kSubBits + kSpectBits;        // index bits a thread holds in registers
constexpr int kSubElems = 1 << kRegBits;                // 16 amplitudes per thread
constexpr int kTileThreads = 1 << (kTileBits - kRegBits);  // one register block per thread: 128 threads per CTA
constexpr int kTileCopies = (1 << kTileBits) / kTileThreads;  // amplitudes per thread when staging the tile

struct TileSubDesc {
    int first, count;                  // gates [first, first + count) of the launch (count <= kMaxSubGates)
    int ins[kRegBits];                 // register bits + spectator bit (tile-index space), ascending: held at 0 in a thread's base index
    unsigned int placeBit[kRegBits];   // BYTE offset in shared memory, as an XOR term, of bit j of the register element index
                                       // (bits 0..2 = the register bits in ascending order, bit 3 = the spectator bit):
                                       // sizeof(qreal) * tilePlace(1 << position); placement is linear, so element e sits at
                                       // the XOR of the terms of its set bits with the thread's base placement
};

struct TileGateCtl {                   // 4 bytes: one constant-bank load per gate
    unsigned short tileMask;           // selecting bits among the tile bits that are not register bits (tile-index space)
    unsigned char regMask;
    unsigned char opcode;              // top byte: one shift extracts it
};
static_assert(kTileBits <= 16, "tileMask is 16 bits wide");
static_assert(sizeof(TileGateCtl) == 4, "k_tile reads a control word as one unsigned int");

constexpr int kMaxRuns = kTileBits + kMaxSlabBits + 1;

struct TileLaunch {
    int nTile;                         // tile bits of this pass (<= kTileBits)
    int nLow;                          // min(kTileLow, nTile): tile-index bits [0, nLow) are amplitude-index bits [0, nLow)
    int nSub, nGates;
    // CTA index -> tile base: the index bits that are neither tile bits nor slab bits form nRuns contiguous runs; run i
    // takes runLen[i] bits of the CTA index (from bit runSrc[i]) to amplitude-index bit runDst[i]
    int nRuns;
    unsigned char runSrc[kMaxRuns], runDst[kMaxRuns], runLen[kMaxRuns];
    unsigned long long slabMask;       // value of the slab bits for this launch (0 for a whole-shard pass)
    unsigned long long rowOff[1 << (kTileBits - kTileLow)];  // amplitude offset of tile index (r << nLow) from the tile base
    unsigned long long stepOff[kTileCopies];                 // amplitude offset of tile index j * kTileThreads
    TileSubDesc sub[kMaxTileSubs];
    TileGateCtl ctl[kMaxTileGates];
    unsigned long long outerMask[kMaxTileGates];  // control/selecting bits outside the tile
    qreal c[kMaxTileGates][9];         // coefficients (meaning per kind: qb_fuse.h PendingOp::c)
};

// OpMatrix with the four outputs built up side by side: every product that reads an old value is issued before the
// last partial sum of that value's replacement, so the new values can be allocated to the registers of the old ones and
// the gate bodies need no register moves where control flow rejoins after the dispatch.  Each output is still
// ((a*x - b*y) + c*z) - d*w, evaluated left to right: the same IEEE bits as OpMatrix / the reference's expressions.
struct TileMatrix {
    qreal r00, i00, r01, i01, r10, i10, r11, i11;
    __device__ __forceinline__ void operator()(qreal& uR, qreal& uI, qreal& lR, qreal& lI) const {
        qreal a = r00 * uR, b = r00 * uI, c = r10 * uR, d = r10 * uI;
        a = a - i00 * uI; b = b + i00 * uR; c = c - i10 * uI; d = d + i10 * uR;
        a = a + r01 * lR; b = b + r01 * lI; c = c + r11 * lR; d = d + r11 * lI;
        const qreal t0 = i01 * lI, t1 = i01 * lR, t2 = i11 * lI, t3 = i11 * lR;
        uR = a - t0; uI = b + t1; lR = c - t2; lI = d + t3;
    }
};
struct TileMatrixFma {
    qreal r00, i00, r01, i01, r10, i10, r11, i11;
    __device__ __forceinline__ void operator()(qreal& uR, qreal& uI, qreal& lR, qreal& lI) const {
        qreal a = r00 * uR, b = r00 * uI, c = r10 * uR, d = r10 * uI;
        a = fma(-i00, uI, a); b = fma(i00, uR, b); c = fma(-i10, uI, c); d = fma(i10, uR, d);
        a = fma(r01, lR, a); b = fma(r01, lI, b); c = fma(r11, lR, c); d = fma(r11, lI, d);
        uR = fma(-i01, lI, a);
        uI = fma(i01, lR, b);
        const qreal nlR = fma(-i11, lI, c);
        lI = fma(i11, lR, d);
        lR = nlR;
    }
};

// Exchange of the two halves of a pair IN PLACE (three XORs per 64-bit word, as opaque instructions).  Written as
// `t = u; u = l; l = t` the compiler turns the exchange into a renaming of registers and then needs a move for every word
// where control flow rejoins after the dispatch - or, with a run-time control mask, a pair of selects for every word of
// every pair; when the bodies are specialised on their control mask, ptxas spills at those joins.  XORs leave every value
// in its own register, so the specialised bodies cost exactly their 6 instructions per complex amplitude.
struct OpSwapInPlace {
    __device__ __forceinline__ void operator()(qreal& uR, qreal& uI, qreal& lR, qreal& lI) const {
        if constexpr (sizeof(qreal) == 8) {
            long long a = __double_as_longlong((double)uR), b = __double_as_longlong((double)lR);
            long long c = __double_as_longlong((double)uI), d = __double_as_longlong((double)lI);
            asm volatile("xor.b64 %0, %0, %1;\n\txor.b64 %1, %1, %0;\n\txor.b64 %0, %0, %1;" : "+l"(a), "+l"(b));
            asm volatile("xor.b64 %0, %0, %1;\n\txor.b64 %1, %1, %0;\n\txor.b64 %0, %0, %1;" : "+l"(c), "+l"(d));
            uR = (qreal)__longlong_as_double(a); lR = (qreal)__longlong_as_double(b);
            uI = (qreal)__longlong_as_double(c); lI = (qreal)__longlong_as_double(d);
        } else {
            int a = __float_as_int((float)uR), b = __float_as_int((float)lR), c = __float_as_int((float)uI), d = __float_as_int((float)lI);
            asm volatile("xor.b32 %0, %0, %1;\n\txor.b32 %1, %1, %0;\n\txor.b32 %0, %0, %1;" : "+r"(a), "+r"(b));
            asm volatile("xor.b32 %0, %0, %1;\n\txor.b32 %1, %1, %0;\n\txor.b32 %0, %0, %1;" : "+r"(c), "+r"(d));
            uR = (qreal)__int_as_float(a); lR = (qreal)__int_as_float(b);
            uI = (qreal)__int_as_float(c); lI = (qreal)__int_as_float(d);
        }
    }
};

// no control among the register bits: straight-line arithmetic on all pairs
template <int P, int R, class PairFn>
__device__ __forceinline__ void forEveryPair(qreal (&xr)[1 << R], qreal (&xi)[1 << R], PairFn fn) {
#pragma unroll
    for (int e = 0; e < (1 << R); ++e) {
        if ((e >> P) & 1) continue;
        fn(xr[e], xi[e], xr[e | (1 << P)], xi[e | (1 << P)]);
    }
}

template <int P, int R, class PairFn>
__device__ __forceinline__ void forEachPair(qreal (&xr)[1 << R], qreal (&xi)[1 << R], unsigned int regMask, PairFn fn) {
#pragma unroll
    for (int e = 0; e < (1 << R); ++e) {
        if ((e >> P) & 1) continue;
        if ((e & regMask) == regMask) fn(xr[e], xi[e], xr[e | (1 << P)], xi[e | (1 << P)]);
    }
}

// Diagonal gates touch only the register elements whose register bits contain the mask: compile-time masks (the caller
// switches on the 8 possible values), so exactly the selected elements are computed.
template <int MASK, int R, class ElemFn>
__device__ __forceinline__ void forSelected(qreal (&xr)[1 << R], qreal (&xi)[1 << R], ElemFn fn) {
#pragma unroll
    for (int e = 0; e < (1 << R); ++e)
        if ((e & MASK) == MASK) fn(xr[e], xi[e]);
}
// Pair operations with the control mask as a compile-time constant (one opcode per (target, control) position, see
// kOpSwapCtl): exactly the selected pairs are touched.  MASK never contains bit P.
template <int P, int MASK, int R, class PairFn>
__device__ __forceinline__ void forMaskedPairs(qreal (&xr)[1 << R], qreal (&xi)[1 << R], PairFn fn) {
#pragma unroll
    for (int e = 0; e < (1 << R); ++e) {
        if ((e >> P) & 1) continue;
        if ((e & MASK) == MASK) fn(xr[e], xi[e], xr[e | (1 << P)], xi[e | (1 << P)]);
    }
}
template <int R, class ElemFn>
__device__ __forceinline__ void forSelectedByMask(qreal (&xr)[1 << R], qreal (&xi)[1 << R], unsigned int regMask, ElemFn fn) {
    static_assert(kSubBits == 3, "eight register masks: the spectator bit is never part of a mask");
    switch (regMask) {
        case 0: forSelected<0, R>(xr, xi, fn); break;
        case 1: forSelected<1, R>(xr, xi, fn); break;
        case 2: forSelected<2, R>(xr, xi, fn); break;
        case 3: forSelected<3, R>(xr, xi, fn); break;
        case 4: forSelected<4, R>(xr, xi, fn); break;
        case 5: forSelected<5, R>(xr, xi, fn); break;
        case 6: forSelected<6, R>(xr, xi, fn); break;
        default: forSelected<7, R>(xr, xi, fn); break;
    }
}

// One-qubit noise channel on the (row bit P, column bit Q) block structure, P < Q register bits.  Same arithmetic as
// k_noise1 (reference QuEST_cpu.c:36-65, 118-215): coherences (1,0),(0,1) scaled by c0; populations (0,0),(1,1)
// mixed (depolarising) or decayed (damping).  The channels are symmetric under exchanging the two bits.
template <int KIND, int P, int Q, int R>
__device__ __forceinline__ void noiseQuads(qreal (&xr)[1 << R], qreal (&xi)[1 << R], qreal c0, qreal c1, qreal c2) {
#pragma unroll
    for (int e = 0; e < (1 << R); ++e) {
        if (((e >> P) & 1) || ((e >> Q) & 1)) continue;
        const int e10 = e | (1 << P), e01 = e | (1 << Q), e11 = e10 | e01;
        if (KIND == kFDensCollapse) {  // reference QuEST_cpu.c:778-852: kept block divided by the outcome probability
            const bool one = c1 != 0;
            const qreal kR = (one ? xr[e11] : xr[e]) / c0, kI = (one ? xi[e11] : xi[e]) / c0;
            xr[e] = one ? (qreal)0 : kR;   xi[e] = one ? (qreal)0 : kI;
            xr[e11] = one ? kR : (qreal)0; xi[e11] = one ? kI : (qreal)0;
            xr[e10] = 0; xi[e10] = 0;
            xr[e01] = 0; xi[e01] = 0;
            continue;
        }
        xr[e10] = c0 * xr[e10]; xi[e10] = c0 * xi[e10];
        xr[e01] = c0 * xr[e01]; xi[e01] = c0 * xi[e01];
        if (KIND == kFDepolarise) {
            const qreal avR = (xr[e] + xr[e11]) / 2, avI = (xi[e] + xi[e11]) / 2;
            xr[e] = c2 * xr[e] + c1 * avR;     xi[e] = c2 * xi[e] + c1 * avI;
            xr[e11] = c2 * xr[e11] + c1 * avR; xi[e11] = c2 * xi[e11] + c1 * avI;
        } else if (KIND == kFDamping) {
            xr[e] = xr[e] + c1 * xr[e11];      xi[e] = xi[e] + c1 * xi[e11];
            xr[e11] = c2 * xr[e11];            xi[e11] = c2 * xi[e11];
        }
    }
}

enum TileTier : int { kTierCore = 0,    // matrix, hadamard, swap / controlledNot, phase (and the fused hadamard + matrix)
                      kTierGates = 1,   // every state-vector gate
                      kTierNoise = 2 }; // + density-matrix noise channels and collapse

#define QB_POS3(BASE, ...)                                             \
    case (BASE) + 0: { constexpr int P = 0; __VA_ARGS__; } break;      \
    case (BASE) + 1: { constexpr int P = 1; __VA_ARGS__; } break;      \
    case (BASE) + 2: { constexpr int P = 2; __VA_ARGS__; } break;
#define QB_NOISE3(KIND)                                                                                             \
    case kOpNoise + ((KIND) - kFDephase) * 3 + 0: noiseQuads<KIND, 0, 1, kRegBits>(xr, xi, C(0), C(1), C(2)); break; \
    case kOpNoise + ((KIND) - kFDephase) * 3 + 1: noiseQuads<KIND, 0, 2, kRegBits>(xr, xi, C(0), C(1), C(2)); break; \
    case kOpNoise + ((KIND) - kFDephase) * 3 + 2: noiseQuads<KIND, 1, 2, kRegBits>(xr, xi, C(0), C(1), C(2)); break;

// One recorded gate on the 8 amplitudes a thread holds in registers: a single jump on the host-made opcode into code
// for that (kind, register position).  Coefficients come straight from the constant bank (kernel parameters).
template <int TIER, bool FMA>
__device__ __forceinline__ void applyTileGate(const TileLaunch& tl, int g, int opcode, unsigned int regMask,
                                              qreal (&xr)[kSubElems], qreal (&xi)[kSubElems]) {
    using Matrix = typename std::conditional<FMA, TileMatrixFma, TileMatrix>::type;
    using Phase = typename std::conditional<FMA, DiagPhaseFma, DiagPhase>::type;
#define C(i) tl.c[g][i]
    switch (opcode) {
        QB_POS3(kOpMatrix, {
            const Matrix m{C(0), C(1), C(2), C(3), C(4), C(5), C(6), C(7)};
            if (regMask == 0) forEveryPair<P, kRegBits>(xr, xi, m);
            else forEachPair<P, kRegBits>(xr, xi, regMask, m);
        })
        QB_POS3(kOpHadamard, forEveryPair<P, kRegBits>(xr, xi, OpHadamard{C(0)}))  // the API has no controlled hadamard
        QB_POS3(kOpHadMatrix, {  // hadamard, then a matrix gate, on the same uncontrolled target (qb::fuseAdjacentPairs)
            const Matrix m{C(0), C(1), C(2), C(3), C(4), C(5), C(6), C(7)};
            const OpHadamard h{C(8)};
            forEveryPair<P, kRegBits>(xr, xi, [&](qreal& uR, qreal& uI, qreal& lR, qreal& lI) {
                h(uR, uI, lR, lI);
                m(uR, uI, lR, lI);
            });
        })
        QB_POS3(kOpSwap, {
            if (regMask == 0) forEveryPair<P, kRegBits>(xr, xi, OpSwapInPlace{});  // pauliX, or a control outside the registers
            else forEachPair<P, kRegBits>(xr, xi, regMask, OpSwap{});              // two controls among the register bits
        })
        // controlledNot whose control is a register bit: (target, control) = (0,1) (0,2) (1,0) (1,2) (2,0) (2,1)
        case kOpSwapCtl + 0: forMaskedPairs<0, 2, kRegBits>(xr, xi, OpSwapInPlace{}); break;
        case kOpSwapCtl + 1: forMaskedPairs<0, 4, kRegBits>(xr, xi, OpSwapInPlace{}); break;
        case kOpSwapCtl + 2: forMaskedPairs<1, 1, kRegBits>(xr, xi, OpSwapInPlace{}); break;
        case kOpSwapCtl + 3: forMaskedPairs<1, 4, kRegBits>(xr, xi, OpSwapInPlace{}); break;
        case kOpSwapCtl + 4: forMaskedPairs<2, 1, kRegBits>(xr, xi, OpSwapInPlace{}); break;
        case kOpSwapCtl + 5: forMaskedPairs<2, 2, kRegBits>(xr, xi, OpSwapInPlace{}); break;
        case kOpPhase + 0: forSelected<0, kRegBits>(xr, xi, Phase{C(0), C(1)}); break;
        case kOpPhase + 1: forSelected<1, kRegBits>(xr, xi, Phase{C(0), C(1)}); break;
        case kOpPhase + 2: forSelected<2, kRegBits>(xr, xi, Phase{C(0), C(1)}); break;
        case kOpPhase + 3: forSelected<3, kRegBits>(xr, xi, Phase{C(0), C(1)}); break;
        case kOpPhase + 4: forSelected<4, kRegBits>(xr, xi, Phase{C(0), C(1)}); break;
        case kOpPhase + 5: forSelected<5, kRegBits>(xr, xi, Phase{C(0), C(1)}); break;
        case kOpPhase + 6: forSelected<6, kRegBits>(xr, xi, Phase{C(0), C(1)}); break;
        case kOpPhase + 7: forSelected<7, kRegBits>(xr, xi, Phase{C(0), C(1)}); break;
        QB_POS3(kOpPauliY, if constexpr (TIER >= kTierGates) forEachPair<P, kRegBits>(xr, xi, regMask, OpPauliY{C(0)}); else __builtin_unreachable())
        case kOpFlip:
            if constexpr (TIER >= kTierGates) forSelectedByMask<kRegBits>(xr, xi, regMask, DiagFlip{});
            else __builtin_unreachable();
            break;
        default:
            if constexpr (TIER == kTierNoise) {
                switch (opcode) {
                    QB_NOISE3(kFDephase)
                    QB_NOISE3(kFDepolarise)
                    QB_NOISE3(kFDamping)
                    QB_NOISE3(kFDensCollapse)
                    default: __builtin_unreachable();
                }
            } else {
                __builtin_unreachable();  // the host never sends a noise opcode to this tier (qb::tierOf)
            }
            break;
    }
#undef C
}
#undef QB_POS3
#undef QB_NOISE3

// asynchronous global -> shared copy of one amplitude component (LDGSTS)
__device__ __forceinline__ void copyAsync(qreal* smemDst, const qreal* src) {
    const unsigned int d = (unsigned int)__cvta_generic_to_shared(smemDst);
    if constexpr (sizeof(qreal) == 8)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(src) : "memory");
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src) : "memory");
}

__host__ __device__ constexpr int lowestSetBit(int e) { return (e & 1) ? 0 : (e & 2) ? 1 : (e & 4) ? 2 : 3; }

__device__ __forceinline__ unsigned int insertZero(unsigned int k, int b) {
    const unsigned int low = k & ((1u << b) - 1u);
    return ((k >> b) << (b + 1)) | low;
}

// Four CTAs of 128 threads per SM.  The register budget is set a little under a quarter of the register file (120, not
// 128): on a sharded register the qubit-slot exchange (k_swapPeer, one small CTA per SM on the communication stream) has
// to fit NEXT to four resident pass CTAs - with 128 registers per thread the pass owns the whole file, an exchange CTA
// can only take the place of a pass CTA, and every pass that overlaps an exchange loses a quarter of its occupancy.
#ifndef QB_TILE_MAXREG
#define QB_TILE_MAXREG 120
#endif

template <int TIER, bool FMA>
__global__ void __maxnreg__(QB_TILE_MAXREG) k_tile(qreal* __restrict__ re, qreal* __restrict__ im,
                                                                   const __grid_constant__ TileLaunch tl) {
    constexpr unsigned int kTileElems = 1u << kTileBits;
    __shared__ qreal sAmp[2 * kTileElems];          // real parts, then imaginary parts: one address register serves both
    __shared__ unsigned int sSubAct[kMaxTileSubs];  // per sub-group: which of its gates this tile takes part in
    const unsigned int tid = threadIdx.x;

    // tile base: this CTA's amplitude index with the tile bits at 0 (and the slab bits at their fixed value)
    unsigned long long base = tl.slabMask;
    {
        const unsigned long long k = blockIdx.x;
        for (int i = 0; i < tl.nRuns; ++i)
            base |= ((k >> tl.runSrc[i]) & ((1ull << tl.runLen[i]) - 1ull)) << tl.runDst[i];
    }

    // 1. stage the tile: consecutive threads take consecutive amplitudes of a row.  Thread t copies tile indices
    // t, t + 128, t + 256, ...: both the amplitude offset and the shared-memory placement are linear in the index, so
    // index t + 128 j costs one add / one xor (with constants) on top of what index t needs.
    const unsigned int nElems = 1u << tl.nTile;
    const int nLow = tl.nLow;
    const unsigned int p0 = tilePlace(tid);
    qreal* const gRe = re + base + tl.rowOff[tid >> nLow] + (tid & ((1u << nLow) - 1u));
    qreal* const gIm = im + (gRe - re);
    if (tid < nElems) {
#pragma unroll
        for (unsigned int j = 0; j < kTileCopies; ++j) {
            if (j * kTileThreads < nElems) {
                const unsigned int p = p0 ^ tilePlace(j * kTileThreads);  // compile-time constant per j
                copyAsync(&sAmp[p], gRe + tl.stepOff[j]);
                copyAsync(&sAmp[kTileElems + p], gIm + tl.stepOff[j]);
            }
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    // 2. which gates does this tile take part in?  One 32-bit word per sub-group (bit j = its j-th gate): controls and
    // phase-mask bits outside the tile are the same for every amplitude of the tile.  Warp w serves sub-groups w, w+4, ...
    for (int s = (int)(tid >> 5); s < tl.nSub; s += kTileThreads / 32) {
        const int j = (int)(tid & 31u);
        const bool mine = j < tl.sub[s].count;
        const unsigned long long m = mine ? tl.outerMask[tl.sub[s].first + j] : ~0ull;
        const unsigned int word = __ballot_sync(0xffffffffu, mine && (base & m) == m);
        if (j == 0) sSubAct[s] = word;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // 3. the sub-groups
    const unsigned int nBlocks = nElems >> kRegBits;
    for (int s = 0; s < tl.nSub; ++s) {
        const TileSubDesc& sg = tl.sub[s];
        const int gFirst = sg.first;
        const unsigned int act = sSubAct[s];
        if (tid < nBlocks && act != 0u) {
            // this thread's block: the register and spectator bits of its base tile index are 0
            const unsigned int idx = insertZero(insertZero(insertZero(insertZero(tid, sg.ins[0]), sg.ins[1]), sg.ins[2]), sg.ins[3]);
            const unsigned int pb = tilePlace(idx) * (unsigned int)sizeof(qreal);
            char* const sBase = reinterpret_cast<char*>(sAmp);
            constexpr unsigned int kImag = kTileElems * sizeof(qreal);
            qreal xr[kSubElems], xi[kSubElems];
            {
                unsigned int q[kSubElems];
                q[0] = pb;
#pragma unroll
                for (int e = 1; e < kSubElems; ++e) q[e] = q[e & (e - 1)] ^ sg.placeBit[lowestSetBit(e)];
#pragma unroll
                for (int e = 0; e < kSubElems; ++e) {
                    xr[e] = *reinterpret_cast<const qreal*>(sBase + q[e]);
                    xi[e] = *reinterpret_cast<const qreal*>(sBase + q[e] + kImag);
                }
            }
            // gate loop: the control word of the next gate is fetched while the current gate computes
            unsigned int todo = act;
            int g = gFirst + __ffs((int)todo) - 1;
            // the control word is fetched as ONE 32-bit constant-bank load (tileMask | regMask << 16 | opcode << 24): read
            // through the struct's fields the compiler issues three narrow loads and a chain of byte permutes per gate.
            // After the last gate the look-ahead index is gFirst - 1: still a word of the kernel parameters (for the
            // first sub-group the end of tl.sub), fetched and never used - cheaper than selecting a valid index.
            const unsigned int* const ctlWords = reinterpret_cast<const unsigned int*>(&tl) + offsetof(TileLaunch, ctl) / sizeof(unsigned int);
            unsigned int ctl = ctlWords[g];
            while (true) {
                todo &= todo - 1u;
                const int gNext = gFirst + __ffs((int)todo) - 1;
                const unsigned int next = ctlWords[gNext];
                if ((ctl & 0xffffu & ~idx) == 0u) applyTileGate<TIER, FMA>(tl, g, (int)(ctl >> 24), (ctl >> 16) & 0xffu, xr, xi);
                if (!todo) break;
                g = gNext;
                ctl = next;
            }
            {  // (addresses recomputed: keeping them alive costs registers the gates need)
                unsigned int q[kSubElems];
                q[0] = pb;
#pragma unroll
                for (int e = 1; e < kSubElems; ++e) q[e] = q[e & (e - 1)] ^ sg.placeBit[lowestSetBit(e)];
#pragma unroll
                for (int e = 0; e < kSubElems; ++e) {
                    *reinterpret_cast<qreal*>(sBase + q[e]) = xr[e];
                    *reinterpret_cast<qreal*>(sBase + q[e] + kImag) = xi[e];
                }
            }
        }
        __syncthreads();
    }

    // 4. write the tile back
    if (tid < nElems) {
#pragma unroll
        for (unsigned int j = 0; j < kTileCopies; ++j) {
            if (j * kTileThreads < nElems) {
                const unsigned int p = p0 ^ tilePlace(j * kTileThreads);
                gRe[tl.stepOff[j]] = sAmp[p];
                gIm[tl.stepOff[j]] = sAmp[kTileElems + p];
            }
        }
    }
}

}  // namespace qb
