// quest-b200 kernels (sm_100a).  Every gate is ONE streaming pass over the amplitudes it changes:
//   - a dense task index is expanded by bit insertion (qb::expand) so that only selected vectors are visited
//     (a controlled gate touches half the state, a k-controlled phase 1/2^k of it);
//   - each thread moves whole 8/16/32-byte vectors of real[] and imag[] with one LDG/STG each, all loads issued
//     before the first dependent use, consecutive threads on consecutive vectors (fully coalesced whenever the
//     selecting bits are above the vector);
//   - target or control bits that fall INSIDE a vector are resolved in registers (template parameter TLO / the
//     lane predicate), so low targets cost the same DRAM traffic as high ones;
//   - arithmetic is written with the operand order of the reference CPU loops and compiled with -fmad=false, so
//     amplitudes match the reference build bit for bit.
// Reductions accumulate in double per thread, fold with warp shuffles, and finish in the last block to arrive.
#pragma once

#include <type_traits>

#include "qb_device.cuh"
#include "qb_fuse.h"

namespace qb {

constexpr int kBlock = 256;

// ======================================================================================================
// 2x2 operators on one amplitude pair (up = target bit 0, lo = target bit 1)
// ======================================================================================================

// general 2x2 (reference QuEST_cpu.c:1672-1681).  Also serves compactUnitary [[a, -conj(b)], [b, conj(a)]]
// (QuEST_cpu.c:1614-1623): with r01=-bR, i01=bI, r10=bR, i10=bI, r11=aR, i11=-aI every expression below evaluates to
// the same IEEE bits as the reference's compact form, because x + (-b)*y == x - b*y and x - (-a)*y == x + a*y exactly.
struct OpMatrix {
    qreal r00, i00, r01, i01, r10, i10, r11, i11;
    __device__ __forceinline__ void operator()(qreal& uR, qreal& uI, qreal& lR, qreal& lI) const {
        const qreal nuR = r00 * uR - i00 * uI + r01 * lR - i01 * lI;
        const qreal nuI = r00 * uI + i00 * uR + r01 * lI + i01 * lR;
        const qreal nlR = r10 * uR - i10 * uI + r11 * lR - i11 * lI;
        const qreal nlI = r10 * uI + i10 * uR + r11 * lI + i11 * lR;
        uR = nuR; uI = nuI; lR = nlR; lI = nlI;
    }
};

// Same operator with the products contracted into fused multiply-adds: 4 fp64 instructions per output instead of 7,
// one rounding per term instead of two.  NOT bit-identical to the reference's CPU arithmetic (differences of an
// ulp or two per gate); used only by the fused pass and only when the host opted in (qb200_setArithmetic).
struct OpMatrixFma {
    qreal r00, i00, r01, i01, r10, i10, r11, i11;
    __device__ __forceinline__ void operator()(qreal& uR, qreal& uI, qreal& lR, qreal& lI) const {
        const qreal nuR = fma(-i01, lI, fma(r01, lR, fma(-i00, uI, r00 * uR)));
        const qreal nuI = fma(i01, lR, fma(r01, lI, fma(i00, uR, r00 * uI)));
        const qreal nlR = fma(-i11, lI, fma(r11, lR, fma(-i10, uI, r10 * uR)));
        const qreal nlI = fma(i11, lR, fma(r11, lI, fma(i10, uR, r10 * uI)));
        uR = nuR; uI = nuI; lR = nlR; lI = nlI;
    }
};

// (up +- lo) * (1/sqrt 2)                                 (reference QuEST_cpu.c:2600-2604)
struct OpHadamard {
    qreal rec;
    __device__ __forceinline__ void operator()(qreal& uR, qreal& uI, qreal& lR, qreal& lI) const {
        const qreal nuR = rec * (uR + lR), nuI = rec * (uI + lI);
        const qreal nlR = rec * (uR - lR), nlI = rec * (uI - lI);
        uR = nuR; uI = nuI; lR = nlR; lI = nlI;
    }
};

// pauliX / controlledNot: exchange                        (reference QuEST_cpu.c:2219-2226)
struct OpSwap {
    __device__ __forceinline__ void operator()(qreal& uR, qreal& uI, qreal& lR, qreal& lI) const {
        const qreal tR = uR, tI = uI;
        uR = lR; uI = lI; lR = tR; lI = tI;
    }
};

// fac * [[0,-i],[i,0]], fac = +1 (Y) or -1 (conj Y)        (reference QuEST_cpu.c:2406-2409)
struct OpPauliY {
    qreal fac;
    __device__ __forceinline__ void operator()(qreal& uR, qreal& uI, qreal& lR, qreal& lI) const {
        const qreal nuR = fac * lI, nuI = fac * -lR;
        const qreal nlR = fac * -uI, nlI = fac * uR;
        uR = nuR; uI = nuI; lR = nlR; lI = nlI;
    }
};

// One thread = one task.  TLO < 0: target bit above the vector -> thread owns an "up" and a "lo" vector
// (tvb = target bit position in vector-index space).  TLO >= 0: target is lane bit TLO -> pairs live inside
// one vector.
template <class Op, int VEC, int TLO>
__global__ void __launch_bounds__(kBlock) k_pairs(qreal* __restrict__ re, qreal* __restrict__ im,
                                                 unsigned long long nTasks, BitPlan plan, int tvb, Op op) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nTasks) return;
    const unsigned long long vi = expand(k, plan);
    if ((vi & plan.skipMask) != plan.skipMask) return;
    if constexpr (TLO < 0) {
        const unsigned long long iu = vi * VEC, il = iu + ((unsigned long long)VEC << tvb);
        Vec<qreal, VEC> uR = load<VEC>(re + iu), uI = load<VEC>(im + iu);
        Vec<qreal, VEC> lR = load<VEC>(re + il), lI = load<VEC>(im + il);
#pragma unroll
        for (int e = 0; e < VEC; ++e)
            if ((e & plan.loMask) == plan.loMask) op(uR.v[e], uI.v[e], lR.v[e], lI.v[e]);
        store<VEC>(re + iu, uR);
        store<VEC>(im + iu, uI);
        store<VEC>(re + il, lR);
        store<VEC>(im + il, lI);
    } else {
        const unsigned long long i = vi * VEC;
        Vec<qreal, VEC> R = load<VEC>(re + i), I = load<VEC>(im + i);
#pragma unroll
        for (int e = 0; e < VEC; ++e)
            if (!((e >> TLO) & 1) && (e & plan.loMask) == plan.loMask)
                op(R.v[e], I.v[e], R.v[e | (1 << TLO)], I.v[e | (1 << TLO)]);
        store<VEC>(re + i, R);
        store<VEC>(im + i, I);
    }
}

// Cross-shard variant: the target qubit selects the GPU, so the "up" half of every pair lives in one shard and the
// "lo" half at the same local index in the partner shard.  Instead of exchanging shards into a pair buffer and then
// combining (reference QuEST_cpu_distributed.c:451-479 + QuEST_cpu.c:1699-1740), each GPU applies the full 2x2
// update to HALF of the pairs, reading the partner's amplitudes through its NVLink peer mapping and writing both
// results in place (its own locally, the partner's remotely).  Every pair is owned by exactly one thread on one
// GPU, so no staging memory and no second pass are needed; NVLink carries 16 bytes per pair each way, the same
// as the exchange.  `up*`/`lo*` are already ordered (upper shard first) by the host.
template <class Op, int VEC>
__global__ void __launch_bounds__(kBlock) k_pairsPeer(qreal* __restrict__ upRe, qreal* __restrict__ upIm,
                                                     qreal* __restrict__ loRe, qreal* __restrict__ loIm,
                                                     unsigned long long nTasks, BitPlan plan, Op op) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nTasks) return;
    const unsigned long long vi = expand(k, plan);
    if ((vi & plan.skipMask) != plan.skipMask) return;
    const unsigned long long i = vi * VEC;
    Vec<qreal, VEC> uR = load<VEC>(upRe + i), uI = load<VEC>(upIm + i);
    Vec<qreal, VEC> lR = load<VEC>(loRe + i), lI = load<VEC>(loIm + i);
#pragma unroll
    for (int e = 0; e < VEC; ++e)
        if ((e & plan.loMask) == plan.loMask) op(uR.v[e], uI.v[e], lR.v[e], lI.v[e]);
    store<VEC>(upRe + i, uR);
    store<VEC>(upIm + i, uI);
    store<VEC>(loRe + i, lR);
    store<VEC>(loIm + i, lI);
}

// Qubit-slot exchange between a shard-selecting bit and local bit l (qb_plan.h: QubitMap).  `a*` is the shard whose
// global bit is 0, `b*` the one whose bit is 1; the amplitudes with (global=0, l=1) and (global=1, l=0) trade
// places: a[j | 2^l] <-> b[j] for every j with bit l clear.  Half a shard crosses NVLink in each direction.  As in
// k_pairsPeer the two GPUs split the index range between them and each performs complete swaps in place.
// Grid-stride: the launch decides how much of the GPU the exchange occupies.  Alone it takes every SM; pipelined with
// a fused pass (backend_cuda.cu: flushThenExchange) it runs as one CTA per SM - enough loads in flight to keep NVLink
// busy - and leaves the other resident-CTA slots to the pass.
template <int VEC>
__global__ void __launch_bounds__(kBlock) k_swapPeer(qreal* __restrict__ aRe, qreal* __restrict__ aIm,
                                                    qreal* __restrict__ bRe, qreal* __restrict__ bIm,
                                                    unsigned long long nTasks, BitPlan plan, int lvb) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < nTasks; k += stride) {
        const unsigned long long vi = expand(k, plan);
        const unsigned long long ib = vi * VEC, ia = ib + ((unsigned long long)VEC << lvb);
        const Vec<qreal, VEC> ar = load<VEC>(aRe + ia), ai = load<VEC>(aIm + ia);
        const Vec<qreal, VEC> br = load<VEC>(bRe + ib), bi = load<VEC>(bIm + ib);
        store<VEC>(aRe + ia, br);
        store<VEC>(aIm + ia, bi);
        store<VEC>(bRe + ib, ar);
        store<VEC>(bIm + ib, ai);
    }
}

// The same exchange for running NEXT TO a fused pass (backend_cuda.cu: flushSharded): held to 32 registers per thread so
// that a 128-thread CTA fits into the 4096 registers four resident pass CTAs (120 registers x 128 threads each) leave
// free on an SM.  16-byte vectors at most: four of them in flight per thread are 16 of those registers.
template <int VEC>
__global__ void __maxnreg__(32) k_swapPeerSlim(qreal* __restrict__ aRe, qreal* __restrict__ aIm, qreal* __restrict__ bRe,
                                               qreal* __restrict__ bIm, unsigned long long nTasks, BitPlan plan, int lvb) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < nTasks; k += stride) {
        const unsigned long long vi = expand(k, plan);
        const unsigned long long ib = vi * VEC, ia = ib + ((unsigned long long)VEC << lvb);
        const Vec<qreal, VEC> ar = load<VEC>(aRe + ia), ai = load<VEC>(aIm + ia);
        const Vec<qreal, VEC> br = load<VEC>(bRe + ib), bi = load<VEC>(bIm + ib);
        store<VEC>(aRe + ia, br);
        store<VEC>(aIm + ia, bi);
        store<VEC>(bRe + ib, ar);
        store<VEC>(bIm + ib, ai);
    }
}

// ======================================================================================================
// diagonal operators on single amplitudes
// ======================================================================================================

struct DiagPhase {  // x *= (c + i s)                        (reference QuEST_cpu.c:2699-2700)
    qreal c, s;
    __device__ __forceinline__ void operator()(qreal& r, qreal& m) const {
        const qreal nr = c * r - s * m, nm = s * r + c * m;
        r = nr; m = nm;
    }
};
struct DiagPhaseFma {  // contracted form, see OpMatrixFma
    qreal c, s;
    __device__ __forceinline__ void operator()(qreal& r, qreal& m) const {
        const qreal nr = fma(-s, m, c * r), nm = fma(c, m, s * r);
        r = nr; m = nm;
    }
};
struct DiagFlip {  // x = -x                                 (reference QuEST_cpu.c:3014-3015)
    __device__ __forceinline__ void operator()(qreal& r, qreal& m) const { r = -r; m = -m; }
};
struct DiagScale {  // x = f * x
    qreal f;
    __device__ __forceinline__ void operator()(qreal& r, qreal& m) const { r = f * r; m = f * m; }
};

template <class Op, int VEC>
__global__ void __launch_bounds__(kBlock) k_diag(qreal* __restrict__ re, qreal* __restrict__ im,
                                                unsigned long long nTasks, BitPlan plan, Op op) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nTasks) return;
    const unsigned long long vi = expand(k, plan);
    if ((vi & plan.skipMask) != plan.skipMask) return;
    const unsigned long long i = vi * VEC;
    Vec<qreal, VEC> R = load<VEC>(re + i), I = load<VEC>(im + i);
#pragma unroll
    for (int e = 0; e < VEC; ++e)
        if ((e & plan.loMask) == plan.loMask) op(R.v[e], I.v[e]);
    store<VEC>(re + i, R);
    store<VEC>(im + i, I);
}

// ======================================================================================================
// reductions.  scratch layout: partials[j * gridDim.x + block] for j < NV, one ticket counter, result[NV].
// The last block to take a ticket adds the per-block partials in index order (deterministic for a fixed grid)
// and the counter wraps to 0 by itself, so the scratch needs no clearing between launches.
// ======================================================================================================

template <int NV>
__device__ __forceinline__ void finishReduce(double (&acc)[NV], double* partials, unsigned int* ticket, double* result) {
    __shared__ double smem[NV * 32];
    __shared__ bool amLast;
    blockSum<NV>(acc, smem);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int j = 0; j < NV; ++j) partials[(size_t)j * gridDim.x + blockIdx.x] = acc[j];
        __threadfence();
        const unsigned int t = atomicInc(ticket, gridDim.x - 1);
        amLast = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (amLast) {
        __threadfence();
        double s[NV];
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            s[j] = 0;
            for (unsigned int b = threadIdx.x; b < gridDim.x; b += blockDim.x)
                s[j] += __ldcg(partials + (size_t)j * gridDim.x + b);
        }
        __syncthreads();
        blockSum<NV>(s, smem);
        if (threadIdx.x == 0) {
#pragma unroll
            for (int j = 0; j < NV; ++j) result[j] = s[j];
        }
    }
}

// sum |amp|^2 over amplitudes whose target bit is 0          (reference QuEST_cpu.c:2880-2887)
template <int VEC, int TLO>
__global__ void __launch_bounds__(kBlock) k_probZero(const qreal* __restrict__ re, const qreal* __restrict__ im,
                                                    unsigned long long nTasks, BitPlan plan, double* partials,
                                                    unsigned int* ticket, double* result) {
    double acc[1] = {0.0};
    const unsigned long long stride = (unsigned long long)gridDim.x * kBlock;
#pragma unroll 4
    for (unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x; k < nTasks; k += stride) {
        const unsigned long long i = expand(k, plan) * VEC;
        const Vec<qreal, VEC> R = load<VEC>(re + i), I = load<VEC>(im + i);
#pragma unroll
        for (int e = 0; e < VEC; ++e)
            if (TLO < 0 || !((e >> (TLO < 0 ? 0 : TLO)) & 1))
                acc[0] += (double)(R.v[e] * R.v[e] + I.v[e] * I.v[e]);
    }
    finishReduce<1>(acc, partials, ticket, result);
}

// P(bit b = 0) for EVERY index bit b of the shard in one pass, plus the total: acc[b] for b < nBits, acc[kProbAll-1].
// Serves the second and later calcProbOfOutcome calls on an unchanged state (backend_cuda.cu: probOfZero), e.g. the
// per-qubit read-out after a QFT: one pass instead of one per qubit.
constexpr int kProbAll = 40;

template <int VEC>
__global__ void __launch_bounds__(kBlock) k_probAllZero(const qreal* __restrict__ re, const qreal* __restrict__ im,
                                                       unsigned long long nVec, int nBits, double* partials,
                                                       unsigned int* ticket, double* result) {
    constexpr int LOW = VEC == 4 ? 2 : (VEC == 2 ? 1 : 0);  // index bits that live inside one vector
    double acc[kProbAll];
#pragma unroll
    for (int b = 0; b < kProbAll; ++b) acc[b] = 0.0;
    const unsigned long long stride = (unsigned long long)gridDim.x * kBlock;
#pragma unroll 2
    for (unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x; k < nVec; k += stride) {
        const unsigned long long i = k * VEC;
        const Vec<qreal, VEC> R = load<VEC>(re + i), I = load<VEC>(im + i);
        double p[VEC], sum = 0.0;
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            p[e] = (double)(R.v[e] * R.v[e] + I.v[e] * I.v[e]);
            sum += p[e];
        }
#pragma unroll
        for (int b = 0; b < LOW; ++b)
#pragma unroll
            for (int e = 0; e < VEC; ++e)
                if (!((e >> b) & 1)) acc[b] += p[e];
#pragma unroll
        for (int b = LOW; b < kProbAll - 1; ++b)
            if (b < nBits) acc[b] += ((i >> b) & 1ull) ? 0.0 : sum;
        acc[kProbAll - 1] += sum;
    }
    finishReduce<kProbAll>(acc, partials, ticket, result);
}

// Bits that select the shard: a shard contributes everything or nothing (QuEST_cpu_distributed.c:1227-1247).
__global__ void k_probAllShardBits(double* result, int localBits, int nShardBits, int chunkId) {
    for (int gb = 0; gb < nShardBits && localBits + gb < kProbAll - 1; ++gb)
        result[localBits + gb] = ((chunkId >> gb) & 1) ? 0.0 : result[kProbAll - 1];
}

// sum |amp|^2 over every amplitude: calcTotalProb (state-vector) and calcPurity (density)
template <int VEC>
__global__ void __launch_bounds__(kBlock) k_sumSquares(const qreal* __restrict__ re, const qreal* __restrict__ im,
                                                      unsigned long long nVec, double* partials, unsigned int* ticket,
                                                      double* result) {
    double acc[1] = {0.0};
    const unsigned long long stride = (unsigned long long)gridDim.x * kBlock;
#pragma unroll 4
    for (unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x; k < nVec; k += stride) {
        const Vec<qreal, VEC> R = load<VEC>(re + k * VEC), I = load<VEC>(im + k * VEC);
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[0] += (double)(R.v[e] * R.v[e] + I.v[e] * I.v[e]);
    }
    finishReduce<1>(acc, partials, ticket, result);
}

// sum conj(bra_i) ket_i, real and imaginary part in one pass  (reference QuEST_cpu.c:1025-1026)
template <int VEC>
__global__ void __launch_bounds__(kBlock) k_innerProduct(const qreal* __restrict__ bR, const qreal* __restrict__ bI,
                                                        const qreal* __restrict__ kR, const qreal* __restrict__ kI,
                                                        unsigned long long nVec, double* partials,
                                                        unsigned int* ticket, double* result) {
    double acc[2] = {0.0, 0.0};
    const unsigned long long stride = (unsigned long long)gridDim.x * kBlock;
#pragma unroll 2
    for (unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x; k < nVec; k += stride) {
        const Vec<qreal, VEC> br = load<VEC>(bR + k * VEC), bi = load<VEC>(bI + k * VEC);
        const Vec<qreal, VEC> kr = load<VEC>(kR + k * VEC), ki = load<VEC>(kI + k * VEC);
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            acc[0] += (double)(br.v[e] * kr.v[e] + bi.v[e] * ki.v[e]);
            acc[1] += (double)(br.v[e] * ki.v[e] - bi.v[e] * kr.v[e]);
        }
    }
    finishReduce<2>(acc, partials, ticket, result);
}

// sum of Re rho_dd over diagonal entries d held by this shard; bit < 0 sums all, else only d with that bit 0
// (reference QuEST_cpu.c:2820-2826, QuEST_cpu_local.c:110-116)
__global__ void __launch_bounds__(kBlock) k_densDiagSum(const qreal* __restrict__ re, unsigned long long dim, int bit,
                                                       unsigned long long shardOffset, unsigned long long shardLen,
                                                       double* partials, unsigned int* ticket, double* result) {
    double acc[1] = {0.0};
    const unsigned long long stride = (unsigned long long)gridDim.x * kBlock;
    for (unsigned long long d = (unsigned long long)blockIdx.x * kBlock + threadIdx.x; d < dim; d += stride) {
        const unsigned long long g = d * (dim + 1);
        if (g < shardOffset || g - shardOffset >= shardLen) continue;
        if (bit >= 0 && ((d >> bit) & 1)) continue;
        acc[0] += (double)re[g - shardOffset];
    }
    finishReduce<1>(acc, partials, ticket, result);
}

// number of amplitude components that differ by more than `precision` (compareStates)
template <int VEC>
__global__ void __launch_bounds__(kBlock) k_countDiffs(const qreal* __restrict__ aR, const qreal* __restrict__ aI,
                                                      const qreal* __restrict__ bR, const qreal* __restrict__ bI,
                                                      unsigned long long nVec, qreal precision, double* partials,
                                                      unsigned int* ticket, double* result) {
    double acc[1] = {0.0};
    const unsigned long long stride = (unsigned long long)gridDim.x * kBlock;
    for (unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x; k < nVec; k += stride) {
        const Vec<qreal, VEC> ar = load<VEC>(aR + k * VEC), ai = load<VEC>(aI + k * VEC);
        const Vec<qreal, VEC> br = load<VEC>(bR + k * VEC), bi = load<VEC>(bI + k * VEC);
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            if (fabs(ar.v[e] - br.v[e]) > precision) acc[0] += 1.0;
            if (fabs(ai.v[e] - bi.v[e]) > precision) acc[0] += 1.0;
        }
    }
    finishReduce<1>(acc, partials, ticket, result);
}

// <psi| rho |psi>: thread per row r accumulates sum_c rho(r,c) psi_c over the columns [col0, col0+nCols) this
// shard holds, then multiplies conj(psi_r); psi is a whole state  (reference QuEST_cpu.c:916-988)
__global__ void __launch_bounds__(kBlock) k_densFidelity(const qreal* __restrict__ dR, const qreal* __restrict__ dI,
                                                        const qreal* __restrict__ pR, const qreal* __restrict__ pI,
                                                        unsigned long long dim, unsigned long long col0,
                                                        unsigned long long nCols, double* partials,
                                                        unsigned int* ticket, double* result) {
    double acc[1] = {0.0};
    const unsigned long long stride = (unsigned long long)gridDim.x * kBlock;
    for (unsigned long long r = (unsigned long long)blockIdx.x * kBlock + threadIdx.x; r < dim; r += stride) {
        double rowR = 0, rowI = 0;
        for (unsigned long long c = 0; c < nCols; ++c) {
            const double eR = dR[r + dim * c], eI = dI[r + dim * c];
            const double vR = pR[col0 + c], vI = pI[col0 + c];
            rowR += eR * vR - eI * vI;
            rowI += eR * vI + eI * vR;
        }
        acc[0] += rowR * (double)pR[r] - rowI * (-(double)pI[r]);
    }
    finishReduce<1>(acc, partials, ticket, result);
}

// ======================================================================================================
// collapse
// ======================================================================================================

// state-vector: kept half *= renorm, other half = 0          (reference QuEST_cpu.c:3067-3087)
template <int VEC, int TLO>
__global__ void __launch_bounds__(kBlock) k_collapse(qreal* __restrict__ re, qreal* __restrict__ im,
                                                    unsigned long long nTasks, BitPlan plan, int tvb, int outcome,
                                                    qreal renorm) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nTasks) return;
    const unsigned long long vi = expand(k, plan);
    if constexpr (TLO < 0) {
        const unsigned long long i0 = vi * VEC, i1 = i0 + ((unsigned long long)VEC << tvb);
        const unsigned long long keep = outcome ? i1 : i0, kill = outcome ? i0 : i1;
        Vec<qreal, VEC> R = load<VEC>(re + keep), I = load<VEC>(im + keep);
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            R.v[e] = R.v[e] * renorm;
            I.v[e] = I.v[e] * renorm;
        }
        store<VEC>(re + keep, R);
        store<VEC>(im + keep, I);
        store<VEC>(re + kill, zeros<VEC>());
        store<VEC>(im + kill, zeros<VEC>());
    } else {
        const unsigned long long i = vi * VEC;
        Vec<qreal, VEC> R = load<VEC>(re + i), I = load<VEC>(im + i);
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            const bool kept = ((e >> TLO) & 1) == outcome;
            R.v[e] = kept ? R.v[e] * renorm : (qreal)0;
            I.v[e] = kept ? I.v[e] * renorm : (qreal)0;
        }
        store<VEC>(re + i, R);
        store<VEC>(im + i, I);
    }
}

// whole-shard variants used when the measured qubit selects the shard (multi-GPU): scale or clear everything
template <int VEC>
__global__ void __launch_bounds__(kBlock) k_scaleAll(qreal* __restrict__ re, qreal* __restrict__ im,
                                                    unsigned long long nVec, qreal f) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nVec) return;
    Vec<qreal, VEC> R = load<VEC>(re + k * VEC), I = load<VEC>(im + k * VEC);
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
        R.v[e] = R.v[e] * f;
        I.v[e] = I.v[e] * f;
    }
    store<VEC>(re + k * VEC, R);
    store<VEC>(im + k * VEC, I);
}

// density matrix: block (outcome,outcome) of the (row bit t, column bit t+n) quadrants is divided by prob, the
// other three are cleared                                   (reference QuEST_cpu.c:778-852, :737-746)
// The column bit is always above the vector (tcb in vector-index space); rowBitsSet/colBitSet describe bits
// that are fixed by the shard offset when they are global (multi-GPU) - here both are local.
template <int VEC, int TLO>
__global__ void __launch_bounds__(kBlock) k_densCollapse(qreal* __restrict__ re, qreal* __restrict__ im,
                                                        unsigned long long nTasks, BitPlan plan, int trb, int tcb,
                                                        int outcome, qreal prob) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nTasks) return;
    const unsigned long long vi = expand(k, plan);
    if constexpr (TLO < 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int a = q & 1, b = q >> 1;
            const unsigned long long i = (vi | ((unsigned long long)a << trb) | ((unsigned long long)b << tcb)) * VEC;
            if (a == outcome && b == outcome) {
                Vec<qreal, VEC> R = load<VEC>(re + i), I = load<VEC>(im + i);
#pragma unroll
                for (int e = 0; e < VEC; ++e) {
                    R.v[e] /= prob;
                    I.v[e] /= prob;
                }
                store<VEC>(re + i, R);
                store<VEC>(im + i, I);
            } else {
                store<VEC>(re + i, zeros<VEC>());
                store<VEC>(im + i, zeros<VEC>());
            }
        }
    } else {
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const unsigned long long i = (vi | ((unsigned long long)b << tcb)) * VEC;
            if (b == outcome) {
                Vec<qreal, VEC> R = load<VEC>(re + i), I = load<VEC>(im + i);
#pragma unroll
                for (int e = 0; e < VEC; ++e) {
                    const bool kept = ((e >> TLO) & 1) == outcome;
                    R.v[e] = kept ? R.v[e] / prob : (qreal)0;
                    I.v[e] = kept ? I.v[e] / prob : (qreal)0;
                }
                store<VEC>(re + i, R);
                store<VEC>(im + i, I);
            } else {
                store<VEC>(re + i, zeros<VEC>());
                store<VEC>(im + i, zeros<VEC>());
            }
        }
    }
}

// ======================================================================================================
// one-qubit noise channels on a density matrix, ONE pass each (the reference GPU backend needs two).
// Quadrant (a,b): a = row bit t, b = column bit t+n.  (1,0) and (0,1) are the coherences; (0,0),(1,1) the
// populations, paired as (i, i | both).
// ======================================================================================================

enum NoiseKind { kDephase = 0, kDepolarise = 1, kDamping = 2 };

template <int KIND>
__device__ __forceinline__ void noiseOffDiag(qreal& x, qreal p0) { x = p0 * x; }

// p0 = factor on coherences (1-lambda, or sqrt(1-d)); p1 = lambda or d; p2 = 1-lambda or 1-d
template <int KIND>
__device__ __forceinline__ void noiseDiag(qreal& x00, qreal& x11, qreal p1, qreal p2) {
    if (KIND == kDepolarise) {  // reference QuEST_cpu.c:153-160
        const qreal av = (x00 + x11) / 2;
        x00 = p2 * x00 + p1 * av;
        x11 = p2 * x11 + p1 * av;
    } else if (KIND == kDamping) {  // reference QuEST_cpu.c:206-210
        x00 = x00 + p1 * x11;
        x11 = p2 * x11;
    }
}

template <int KIND, int VEC, int TLO>
__global__ void __launch_bounds__(kBlock) k_noise1(qreal* __restrict__ re, qreal* __restrict__ im,
                                                  unsigned long long nTasks, BitPlan plan, int trb, int tcb, qreal p0,
                                                  qreal p1, qreal p2) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nTasks) return;
    const unsigned long long vi = expand(k, plan);
    if constexpr (TLO < 0) {
        const unsigned long long i00 = vi * VEC;
        const unsigned long long i10 = i00 + ((unsigned long long)VEC << trb);  // row bit set
        const unsigned long long i01 = i00 + ((unsigned long long)VEC << tcb);  // column bit set
        const unsigned long long i11 = i10 + ((unsigned long long)VEC << tcb);
        Vec<qreal, VEC> R10 = load<VEC>(re + i10), I10 = load<VEC>(im + i10);
        Vec<qreal, VEC> R01 = load<VEC>(re + i01), I01 = load<VEC>(im + i01);
        if constexpr (KIND != kDephase) {
            Vec<qreal, VEC> R00 = load<VEC>(re + i00), I00 = load<VEC>(im + i00);
            Vec<qreal, VEC> R11 = load<VEC>(re + i11), I11 = load<VEC>(im + i11);
#pragma unroll
            for (int e = 0; e < VEC; ++e) {
                noiseDiag<KIND>(R00.v[e], R11.v[e], p1, p2);
                noiseDiag<KIND>(I00.v[e], I11.v[e], p1, p2);
            }
            store<VEC>(re + i00, R00);
            store<VEC>(im + i00, I00);
            store<VEC>(re + i11, R11);
            store<VEC>(im + i11, I11);
        }
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            noiseOffDiag<KIND>(R10.v[e], p0);
            noiseOffDiag<KIND>(I10.v[e], p0);
            noiseOffDiag<KIND>(R01.v[e], p0);
            noiseOffDiag<KIND>(I01.v[e], p0);
        }
        store<VEC>(re + i10, R10);
        store<VEC>(im + i10, I10);
        store<VEC>(re + i01, R01);
        store<VEC>(im + i01, I01);
    } else {
        // row bit is lane bit TLO: vector b=0 holds (0,0) and (1,0) lanes, vector b=1 holds (0,1) and (1,1)
        const unsigned long long i0 = vi * VEC, i1 = i0 + ((unsigned long long)VEC << tcb);
        Vec<qreal, VEC> R0 = load<VEC>(re + i0), I0 = load<VEC>(im + i0);
        Vec<qreal, VEC> R1 = load<VEC>(re + i1), I1 = load<VEC>(im + i1);
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            if ((e >> TLO) & 1) continue;
            const int f = e | (1 << TLO);
            noiseOffDiag<KIND>(R0.v[f], p0);  // (1,0)
            noiseOffDiag<KIND>(I0.v[f], p0);
            noiseOffDiag<KIND>(R1.v[e], p0);  // (0,1)
            noiseOffDiag<KIND>(I1.v[e], p0);
            noiseDiag<KIND>(R0.v[e], R1.v[f], p1, p2);  // (0,0) with (1,1)
            noiseDiag<KIND>(I0.v[e], I1.v[f], p1, p2);
        }
        store<VEC>(re + i0, R0);
        store<VEC>(im + i0, I0);
        store<VEC>(re + i1, R1);
        store<VEC>(im + i1, I1);
    }
}

// ---- second-tier density operations: simple one-amplitude-per-thread kernels --------------------------

// two-qubit dephase: scale every element where either qubit's row and column bits differ
// (reference QuEST_cpu.c:102-114).  One vector per thread; the four bits are tested on the GLOBAL index of every lane.
template <int VEC>
__global__ void __launch_bounds__(kBlock) k_dephase2(qreal* __restrict__ re, qreal* __restrict__ im,
                                                    unsigned long long nVec, unsigned long long off,
                                                    unsigned long long in1, unsigned long long out1,
                                                    unsigned long long in2, unsigned long long out2, qreal retain) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nVec) return;
    const unsigned long long i = k * VEC;
    Vec<qreal, VEC> R = load<VEC>(re + i), I = load<VEC>(im + i);
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
        const unsigned long long p1 = (i + e + off) & (in1 | out1), p2 = (i + e + off) & (in2 | out2);
        if (p1 == in1 || p1 == out1 || p2 == in2 || p2 == out2) {
            R.v[e] = retain * R.v[e];
            I.v[e] = retain * I.v[e];
        }
    }
    store<VEC>(re + i, R);
    store<VEC>(im + i, I);
}

// Two-qubit depolarising channel in ONE pass.  The reference makes four: a two-qubit dephase and three pair sweeps
// (QuEST_cpu_local.c:40-51, QuEST_cpu.c:77-116 and :379-477).  All four only ever combine the 16 elements that differ in
// the row and column bits of the two qubits, so a thread takes such a block into registers - element e: bit 0 = row bit
// of qubit 1, bit 1 = its column bit, bits 2, 3 = the same for qubit 2 - and applies the four stages one after the other
// with the reference's expressions: identical bits, a quarter of the HBM traffic.
struct Depol2 {
    unsigned long long offs[16];  // amplitude offset of element e from the thread's base index
    qreal retain, delta, gamma;
};

__device__ __forceinline__ void depol2Sweep(qreal (&xr)[16], qreal (&xi)[16], int sel, int other, int pair, qreal delta, qreal gamma) {
    // pairs (e, (e | sel) ^ pair) over the elements whose `sel` bits are 0 and whose `other` bits are equal (QuEST_cpu.c:406-477)
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        if ((e & sel) != 0 || !((e & other) == 0 || (e & other) == other)) continue;
        const int f = (e | sel) ^ pair;
        const qreal r0 = xr[e], i0 = xi[e], r1 = xr[f], i1 = xi[f];
        xr[e] = gamma * (r0 + delta * r1);
        xi[e] = gamma * (i0 + delta * i1);
        xr[f] = gamma * (r1 + delta * r0);
        xi[f] = gamma * (i1 + delta * i0);
    }
}

__global__ void __launch_bounds__(kBlock) k_depol2(qreal* __restrict__ re, qreal* __restrict__ im, unsigned long long nTasks,
                                                  BitPlan plan, const __grid_constant__ Depol2 d) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nTasks) return;
    const unsigned long long base = expand(k, plan);
    qreal xr[16], xi[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        xr[e] = re[base + d.offs[e]];
        xi[e] = im[base + d.offs[e]];
    }
    // stage 1: two-qubit dephase - elements where either qubit's row and column bits differ
#pragma unroll
    for (int e = 0; e < 16; ++e)
        if (((e ^ (e >> 1)) & 1) || (((e >> 2) ^ (e >> 3)) & 1)) {
            xr[e] = d.retain * xr[e];
            xi[e] = d.retain * xi[e];
        }
    // stages 2-4: the three sweeps, in the reference's order (tot1 = bits 0,1; tot2 = bits 2,3)
    depol2Sweep(xr, xi, 3, 12, 0, d.delta, (qreal)1);
    depol2Sweep(xr, xi, 12, 3, 0, d.delta, (qreal)1);
    depol2Sweep(xr, xi, 12, 3, 3, d.delta, d.gamma);
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        re[base + d.offs[e]] = xr[e];
        im[base + d.offs[e]] = xi[e];
    }
}

// rho = (1-p) rho + p sigma, in the reference's two-step form  (reference QuEST_cpu.c:905-911)
template <int VEC>
__global__ void __launch_bounds__(kBlock) k_mixDensity(qreal* __restrict__ aR, qreal* __restrict__ aI,
                                                      const qreal* __restrict__ bR, const qreal* __restrict__ bI,
                                                      unsigned long long nVec, qreal p) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nVec) return;
    Vec<qreal, VEC> ar = load<VEC>(aR + k * VEC), ai = load<VEC>(aI + k * VEC);
    const Vec<qreal, VEC> br = load<VEC>(bR + k * VEC), bi = load<VEC>(bI + k * VEC);
    const qreal keep = 1 - p;
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
        ar.v[e] = ar.v[e] * keep;
        ai.v[e] = ai.v[e] * keep;
        ar.v[e] = ar.v[e] + p * br.v[e];
        ai.v[e] = ai.v[e] + p * bi.v[e];
    }
    store<VEC>(aR + k * VEC, ar);
    store<VEC>(aI + k * VEC, ai);
}

// rho(r,c) = psi_r conj(psi_c), column-major                  (mathematically correct form)
__global__ void __launch_bounds__(kBlock) k_outerProduct(qreal* __restrict__ dR, qreal* __restrict__ dI,
                                                        const qreal* __restrict__ pR, const qreal* __restrict__ pI,
                                                        unsigned long long dim, unsigned long long n,
                                                        unsigned long long off) {
    const unsigned long long i = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (i >= n) return;
    const unsigned long long r = (i + off) % dim, c = (i + off) / dim;
    const qreal kr = pR[r], ki = pI[r], br = pR[c], bi = -pI[c];
    dR[i] = kr * br - ki * bi;
    dI[i] = kr * bi + ki * br;
}

// ======================================================================================================
// initial states
// ======================================================================================================

template <int VEC>
__global__ void __launch_bounds__(kBlock) k_fill(qreal* __restrict__ re, qreal* __restrict__ im, unsigned long long nVec,
                                                qreal vr, qreal vi) {
    const unsigned long long k = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= nVec) return;
    Vec<qreal, VEC> R, I;
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
        R.v[e] = vr;
        I.v[e] = vi;
    }
    store<VEC>(re + k * VEC, R);
    store<VEC>(im + k * VEC, I);
}

__global__ void k_setOne(qreal* re, qreal* im, unsigned long long i, qreal vr, qreal vi) {
    re[i] = vr;
    im[i] = vi;
}

// amp_k = (2k/10, (2k+1)/10) with k the global index          (reference QuEST_cpu.c:1500-1501)
__global__ void __launch_bounds__(kBlock) k_initDebug(qreal* __restrict__ re, qreal* __restrict__ im,
                                                     unsigned long long n, unsigned long long off) {
    const unsigned long long i = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (i >= n) return;
    re[i] = ((off + i) * 2.0) / 10.0;
    im[i] = ((off + i) * 2.0 + 1.0) / 10.0;
}

// one qubit fixed to `outcome`, the rest uniform                (reference QuEST_cpu.c:1454-1463)
__global__ void __launch_bounds__(kBlock) k_initSingleQubit(qreal* __restrict__ re, qreal* __restrict__ im,
                                                           unsigned long long n, unsigned long long off, int qubit,
                                                           int outcome, qreal norm) {
    const unsigned long long i = (unsigned long long)blockIdx.x * kBlock + threadIdx.x;
    if (i >= n) return;
    re[i] = ((int)(((off + i) >> qubit) & 1) == outcome) ? norm : (qreal)0;
    im[i] = 0;
}

}  // namespace qb
