/* quest-b200 ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C CPU restatement of the reference's backend for the gate-application path
 * (reference: QuEST/src/CPU/QuEST_cpu.c and QuEST_cpu_local.c).  It implements the backend seam declared in
 * quest_b200/csrc/quest_internal.h on host arrays, so the product's own host front-end can be linked on top of
 * it (oracle/Makefile -> oracle/_build/libQuEST_oracle.so) and driven through the same QuEST.h calls as the
 * CUDA library.
 *
 * Who may use it: tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, and only
 * as the checker or the timed CPU baseline.  The product library (quest_b200/lib/libQuEST.so) never links,
 * loads or calls anything in this directory and has no CPU fallback.
 *
 * Parity status: PINNED.  tests/test_oracle_cpu.py checks this file against
 *   (1) every golden vector the reference's suite holds for the path (tests/golden/ref_unit_tests.json,
 *       converted from /root/reference/tests/{unit,essential} by tests/golden/make_golden.py), and
 *   (2) the reference itself compiled here (oracle/_ref/libQuEST.so) on seeded random circuits, bit for bit.
 *
 * Arithmetic notes.  The reference is compiled -O2 -mavx (no FMA contraction, no reassociation), so each
 * amplitude update below is written with the reference's operand order and evaluates to the same IEEE bits.
 * Index arithmetic is restated with bit insertion instead of the reference's divide/modulo; the visited
 * (up, lo) pairs are the same.  Every function cites the reference lines it follows.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#include "quest_internal.h"

typedef long long int idx_t;

/* index of the k-th amplitude pair's |0> member for target bit t: insert a 0 at bit t
 * (== thisBlock*sizeBlock + thisTask%sizeHalfBlock, QuEST_cpu.c:1602-1603) */
static inline idx_t insert_zero(idx_t k, int t) {
    idx_t low = k & ((1LL << t) - 1);
    return ((k >> t) << (t + 1)) | low;
}

static inline idx_t global_offset(Qureg q) { return q.numAmpsPerChunk * (idx_t)q.chunkId; }

static idx_t mask_of(const int* qubits, int n) {
    idx_t m = 0;
    for (int i = 0; i < n; i++) m |= 1LL << qubits[i];
    return m;
}

/* ---------------------------------------------------------------------------------------------------
 * environment (QuEST_cpu_local.c:154-189, 288-296)
 * ------------------------------------------------------------------------------------------------- */

QuESTEnv createQuESTEnv(void) {
    QuESTEnv env;
    env.rank = 0;
    env.numRanks = 1;
    seedQuESTDefault();
    return env;
}

void syncQuESTEnv(QuESTEnv env) { (void)env; }
int syncQuESTSuccess(int successCode) { return successCode; }
void destroyQuESTEnv(QuESTEnv env) { (void)env; }

void reportQuESTEnv(QuESTEnv env) {
    printf("EXECUTION ENVIRONMENT:\n");
    printf("quest-b200 CPU oracle (test infrastructure), one process\n");
    printf("Number of ranks is %d\n", env.numRanks);
#ifdef _OPENMP
    printf("OpenMP enabled\n");
    printf("Number of threads available is %d\n", omp_get_max_threads());
#else
    printf("OpenMP disabled\n");
#endif
    printf("Precision: size of qreal is %ld bytes\n", sizeof(qreal));
}

void getEnvironmentString(QuESTEnv env, Qureg qureg, char str[200]) {
    int threads = 1;
#ifdef _OPENMP
    threads = omp_get_max_threads();
#endif
    sprintf(str, "%dqubits_CPU_%dranksx%dthreads", qureg.numQubitsInStateVec, env.numRanks, threads);
}

void seedQuESTDefault(void) {
    unsigned long int key[2];
    getQuESTDefaultSeedKey(key);
    init_by_array(key, 2);
}

/* ---------------------------------------------------------------------------------------------------
 * registers (QuEST_cpu.c:1202-1250, 1160-1200, 1388-1419, 1557-1568; QuEST_cpu_local.c:191-197)
 * ------------------------------------------------------------------------------------------------- */

void statevec_createQureg(Qureg* qureg, int numQubits, QuESTEnv env) {
    idx_t total = 1LL << numQubits;
    idx_t per = total / env.numRanks;
    memset(qureg, 0, sizeof *qureg);
    qureg->stateVec.real = malloc(per * sizeof(qreal));
    qureg->stateVec.imag = malloc(per * sizeof(qreal));
    if ((!qureg->stateVec.real || !qureg->stateVec.imag) && per) {
        printf("Could not allocate memory!");
        exit(EXIT_FAILURE);
    }
    qureg->numQubitsInStateVec = numQubits;
    qureg->numAmpsTotal = total;
    qureg->numAmpsPerChunk = per;
    qureg->chunkId = env.rank;
    qureg->numChunks = env.numRanks;
    qureg->isDensityMatrix = 0;
}

void statevec_destroyQureg(Qureg qureg, QuESTEnv env) {
    (void)env;
    free(qureg.stateVec.real);
    free(qureg.stateVec.imag);
}

void statevec_setAmps(Qureg q, idx_t startInd, qreal* reals, qreal* imags, idx_t numAmps) {
    idx_t off = global_offset(q);
    idx_t lo = startInd - off, hi = lo + numAmps;
    if (lo < 0) lo = 0;
    if (hi > q.numAmpsPerChunk) hi = q.numAmpsPerChunk;
    for (idx_t i = lo; i < hi; i++) {
        q.stateVec.real[i] = reals[i + off - startInd];
        q.stateVec.imag[i] = imags[i + off - startInd];
    }
}

void qb200_copyAmpsToHost(Qureg q, idx_t start, idx_t n, qreal* reals, qreal* imags) {
    memcpy(reals, q.stateVec.real + start, n * sizeof(qreal));
    memcpy(imags, q.stateVec.imag + start, n * sizeof(qreal));
}

void statevec_cloneQureg(Qureg dst, Qureg src) {
    memcpy(dst.stateVec.real, src.stateVec.real, dst.numAmpsPerChunk * sizeof(qreal));
    memcpy(dst.stateVec.imag, src.stateVec.imag, dst.numAmpsPerChunk * sizeof(qreal));
}

int statevec_compareStates(Qureg a, Qureg b, qreal precision) {
    for (idx_t i = 0; i < a.numAmpsPerChunk; i++) {
        if (absReal(a.stateVec.real[i] - b.stateVec.real[i]) > precision) return 0;
        if (absReal(a.stateVec.imag[i] - b.stateVec.imag[i]) > precision) return 0;
    }
    return 1;
}

qreal statevec_getRealAmp(Qureg q, idx_t index) { return q.stateVec.real[index - global_offset(q)]; }
qreal statevec_getImagAmp(Qureg q, idx_t index) { return q.stateVec.imag[index - global_offset(q)]; }

/* ---------------------------------------------------------------------------------------------------
 * initial states (QuEST_cpu.c:1284-1386, 1427-1504, 1038-1105, 1107-1158)
 * ------------------------------------------------------------------------------------------------- */

static void fill(Qureg q, qreal re, qreal im) {
    idx_t n = q.numAmpsPerChunk;
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        q.stateVec.real[i] = re;
        q.stateVec.imag[i] = im;
    }
}

static void set_one_global(Qureg q, idx_t globalInd) {
    if (q.chunkId == globalInd / q.numAmpsPerChunk) {
        q.stateVec.real[globalInd % q.numAmpsPerChunk] = 1.0;
        q.stateVec.imag[globalInd % q.numAmpsPerChunk] = 0.0;
    }
}

void statevec_initZeroState(Qureg q) {
    fill(q, 0.0, 0.0);
    set_one_global(q, 0);
}

void statevec_initPlusState(Qureg q) {
    idx_t total = q.numAmpsPerChunk * q.numChunks;
    fill(q, 1.0 / sqrt((qreal)total), 0.0);
}

void statevec_initClassicalState(Qureg q, idx_t stateInd) {
    fill(q, 0.0, 0.0);
    set_one_global(q, stateInd);
}

void statevec_initStateDebug(Qureg q) {
    idx_t n = q.numAmpsPerChunk, off = global_offset(q);
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        q.stateVec.real[i] = ((off + i) * 2.0) / 10.0;
        q.stateVec.imag[i] = ((off + i) * 2.0 + 1.0) / 10.0;
    }
}

void statevec_initStateOfSingleQubit(Qureg* q, int qubitId, int outcome) {
    idx_t n = q->numAmpsPerChunk, off = global_offset(*q);
    idx_t total = n * q->numChunks;
    qreal norm = 1.0 / sqrt((qreal)total / 2.0);
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        int bit = (int)(((off + i) >> qubitId) & 1);
        q->stateVec.real[i] = (bit == outcome) ? norm : 0.0;
        q->stateVec.imag[i] = 0.0;
    }
}

void densmatr_initPlusState(Qureg q) {
    idx_t dim = 1LL << q.numQubitsRepresented;
    fill(q, 1.0 / ((qreal)dim), 0.0);
}

void densmatr_initClassicalState(Qureg q, idx_t stateInd) {
    idx_t dim = 1LL << q.numQubitsRepresented;
    fill(q, 0.0, 0.0);
    set_one_global(q, (dim + 1) * stateInd);
}

/* rho = |psi><psi|: element (r,c) = psi_r * conj(psi_c).  The reference's CPU loop has a sign slip in exactly
 * this product (QuEST_cpu.c:1153-1154; its GPU kernel QuEST_gpu.cu:79-82 has the right signs), so the
 * mathematically correct form is used and this function is NOT parity-tested against the CPU reference. */
void densmatr_initPureState(Qureg rho, Qureg psi) {
    idx_t dim = psi.numAmpsTotal;
#pragma omp parallel for schedule(static)
    for (idx_t c = 0; c < dim; c++)
        for (idx_t r = 0; r < dim; r++) {
            qreal kr = psi.stateVec.real[r], ki = psi.stateVec.imag[r];
            qreal br = psi.stateVec.real[c], bi = -psi.stateVec.imag[c];
            rho.stateVec.real[r + c * dim] = kr * br - ki * bi;
            rho.stateVec.imag[r + c * dim] = kr * bi + ki * br;
        }
}

/* ---------------------------------------------------------------------------------------------------
 * dense one-target gates.  One loop shape for all of them: for each pair (up, lo = up + 2^t) whose global
 * index contains every bit of ctrlMask, replace (up, lo) by M (up, lo)^T.
 * ------------------------------------------------------------------------------------------------- */

#define FOR_EACH_PAIR(q, t, ctrlMask, ...)                                           \
    do {                                                                               \
        idx_t nPairs_ = (q).numAmpsPerChunk >> 1, off_ = global_offset(q);             \
        idx_t half_ = 1LL << (t), cm_ = (ctrlMask);                                    \
        qreal* re = (q).stateVec.real;                                                 \
        qreal* im = (q).stateVec.imag;                                                 \
        _Pragma("omp parallel for schedule(static)") for (idx_t k_ = 0; k_ < nPairs_; k_++) { \
            idx_t up = insert_zero(k_, (t)), lo = up + half_;                          \
            if (((up + off_) & cm_) != cm_) continue;                                  \
            __VA_ARGS__                                                                \
        }                                                                              \
    } while (0)

/* [[alpha, -conj(beta)], [beta, conj(alpha)]]; operand order of QuEST_cpu.c:1614-1623 */
static void apply_compact(Qureg q, int t, idx_t cm, Complex alpha, Complex beta) {
    qreal aR = alpha.real, aI = alpha.imag, bR = beta.real, bI = beta.imag;
    FOR_EACH_PAIR(q, t, cm, {
        qreal uR = re[up], uI = im[up], lR = re[lo], lI = im[lo];
        re[up] = aR * uR - aI * uI - bR * lR - bI * lI;
        im[up] = aR * uI + aI * uR - bR * lI + bI * lR;
        re[lo] = bR * uR - bI * uI + aR * lR + aI * lI;
        im[lo] = bR * uI + bI * uR + aR * lI - aI * lR;
    });
}

/* general 2x2; operand order of QuEST_cpu.c:1672-1681 */
static void apply_matrix(Qureg q, int t, idx_t cm, ComplexMatrix2 u) {
    FOR_EACH_PAIR(q, t, cm, {
        qreal uR = re[up], uI = im[up], lR = re[lo], lI = im[lo];
        re[up] = u.r0c0.real * uR - u.r0c0.imag * uI + u.r0c1.real * lR - u.r0c1.imag * lI;
        im[up] = u.r0c0.real * uI + u.r0c0.imag * uR + u.r0c1.real * lI + u.r0c1.imag * lR;
        re[lo] = u.r1c0.real * uR - u.r1c0.imag * uI + u.r1c1.real * lR - u.r1c1.imag * lI;
        im[lo] = u.r1c0.real * uI + u.r1c0.imag * uR + u.r1c1.real * lI + u.r1c1.imag * lR;
    });
}

void statevec_compactUnitary(Qureg q, const int t, Complex alpha, Complex beta) { apply_compact(q, t, 0, alpha, beta); }
void statevec_controlledCompactUnitary(Qureg q, const int c, const int t, Complex alpha, Complex beta) {
    apply_compact(q, t, 1LL << c, alpha, beta); /* QuEST_cpu.c:1800-1865 */
}
void statevec_unitary(Qureg q, const int t, ComplexMatrix2 u) { apply_matrix(q, t, 0, u); }
void statevec_controlledUnitary(Qureg q, const int c, const int t, ComplexMatrix2 u) {
    apply_matrix(q, t, 1LL << c, u); /* QuEST_cpu.c:1930-1995 */
}
void statevec_multiControlledUnitary(Qureg q, int* ctrls, const int n, const int t, ComplexMatrix2 u) {
    apply_matrix(q, t, mask_of(ctrls, n), u); /* QuEST_cpu_local.c:219-226, QuEST_cpu.c:1867-1928 */
}

/* (up +- lo)/sqrt2 with recRoot2 = 1.0/sqrt(2) as a multiplier; QuEST_cpu.c:2577,2600-2604 */
void statevec_hadamard(Qureg q, const int t) {
    qreal r = 1.0 / sqrt(2);
    FOR_EACH_PAIR(q, t, 0, {
        qreal uR = re[up], uI = im[up], lR = re[lo], lI = im[lo];
        re[up] = r * (uR + lR);
        im[up] = r * (uI + lI);
        re[lo] = r * (uR - lR);
        im[lo] = r * (uI - lI);
    });
}

static void apply_swap(Qureg q, int t, idx_t cm) { /* QuEST_cpu.c:2214-2227, 2306-2321 */
    FOR_EACH_PAIR(q, t, cm, {
        qreal uR = re[up], uI = im[up];
        re[up] = re[lo];
        im[up] = im[lo];
        re[lo] = uR;
        im[lo] = uI;
    });
}
void statevec_pauliX(Qureg q, const int t) { apply_swap(q, t, 0); }
void statevec_controlledNot(Qureg q, const int c, const int t) { apply_swap(q, t, 1LL << c); }

/* conjFac * [[0,-i],[i,0]]; QuEST_cpu.c:2406-2409, 2496-2506 */
static void apply_y(Qureg q, int t, idx_t cm, int conjFac) {
    FOR_EACH_PAIR(q, t, cm, {
        qreal uR = re[up], uI = im[up];
        re[up] = conjFac * im[lo];
        im[up] = conjFac * -re[lo];
        re[lo] = conjFac * -uI;
        im[lo] = conjFac * uR;
    });
}
void statevec_pauliY(Qureg q, const int t) { apply_y(q, t, 0, 1); }
void statevec_pauliYConj(Qureg q, const int t) { apply_y(q, t, 0, -1); }
void statevec_controlledPauliY(Qureg q, const int c, const int t) { apply_y(q, t, 1LL << c, 1); }
void statevec_controlledPauliYConj(Qureg q, const int c, const int t) { apply_y(q, t, 1LL << c, -1); }

/* ---------------------------------------------------------------------------------------------------
 * diagonal gates: multiply every amplitude whose global index contains `mask` by (c + i s)
 * (QuEST_cpu.c:2666-2704, 2706-2743, 2745-2786, 2941-3021)
 * ------------------------------------------------------------------------------------------------- */

static void apply_phase(Qureg q, idx_t mask, qreal c, qreal s) {
    idx_t n = q.numAmpsPerChunk, off = global_offset(q);
    qreal* re = q.stateVec.real;
    qreal* im = q.stateVec.imag;
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        if (((i + off) & mask) != mask) continue;
        qreal r = re[i], m = im[i];
        re[i] = c * r - s * m;
        im[i] = s * r + c * m;
    }
}

static void apply_flip(Qureg q, idx_t mask) {
    idx_t n = q.numAmpsPerChunk, off = global_offset(q);
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        if (((i + off) & mask) != mask) continue;
        q.stateVec.real[i] = -q.stateVec.real[i];
        q.stateVec.imag[i] = -q.stateVec.imag[i];
    }
}

void statevec_phaseShiftByTerm(Qureg q, const int t, Complex term) { apply_phase(q, 1LL << t, term.real, term.imag); }
void statevec_controlledPhaseShift(Qureg q, const int q1, const int q2, qreal angle) {
    apply_phase(q, (1LL << q1) | (1LL << q2), cos(angle), sin(angle));
}
void statevec_multiControlledPhaseShift(Qureg q, int* ctrls, int n, qreal angle) {
    apply_phase(q, mask_of(ctrls, n), cos(angle), sin(angle));
}
void statevec_controlledPhaseFlip(Qureg q, const int q1, const int q2) { apply_flip(q, (1LL << q1) | (1LL << q2)); }
void statevec_multiControlledPhaseFlip(Qureg q, int* ctrls, int n) { apply_flip(q, mask_of(ctrls, n)); }

/* ---------------------------------------------------------------------------------------------------
 * reductions (QuEST_cpu_local.c:100-151, 267-281; QuEST_cpu.c:854-881, 994-1034, 2789-2899)
 * ------------------------------------------------------------------------------------------------- */

/* serial Kahan sum over re^2 then im^2 of each amplitude, single thread, as the reference */
qreal statevec_calcTotalProb(Qureg q) {
    qreal total = 0, c = 0, y, t;
    for (idx_t i = 0; i < q.numAmpsPerChunk; i++) {
        y = q.stateVec.real[i] * q.stateVec.real[i] - c;
        t = total + y;
        c = (t - total) - y;
        total = t;
        y = q.stateVec.imag[i] * q.stateVec.imag[i] - c;
        t = total + y;
        c = (t - total) - y;
        total = t;
    }
    return total;
}

qreal densmatr_calcTotalProb(Qureg q) {
    qreal total = 0, c = 0, y, t;
    idx_t dim = 1LL << q.numQubitsRepresented;
    for (idx_t col = 0; col < dim; col++) {
        y = q.stateVec.real[col * (dim + 1)] - c;
        t = total + y;
        c = (t - total) - y;
        total = t;
    }
    return total;
}

static qreal prob_of_zero(Qureg q, int t) {
    idx_t nPairs = q.numAmpsPerChunk >> 1;
    qreal* re = q.stateVec.real;
    qreal* im = q.stateVec.imag;
    qreal sum = 0;
#pragma omp parallel for schedule(static) reduction(+ : sum)
    for (idx_t k = 0; k < nPairs; k++) {
        idx_t i = insert_zero(k, t);
        sum += re[i] * re[i] + im[i] * im[i];
    }
    return sum;
}

/* P(1) is DEFINED as 1 - P(0), also for unnormalised states (QuEST_cpu_local.c:271) */
qreal statevec_calcProbOfOutcome(Qureg q, const int t, int outcome) {
    qreal p = prob_of_zero(q, t);
    return outcome == 1 ? 1.0 - p : p;
}

qreal densmatr_calcProbOfOutcome(Qureg q, const int t, int outcome) {
    idx_t dim = 1LL << q.numQubitsRepresented;
    qreal sum = 0;
#pragma omp parallel for schedule(static) reduction(+ : sum)
    for (idx_t d = 0; d < dim; d++)
        if (((d >> t) & 1) == 0) sum += q.stateVec.real[d * (dim + 1)];
    return outcome == 1 ? 1.0 - sum : sum;
}

Complex statevec_calcInnerProduct(Qureg bra, Qureg ket) {
    qreal sr = 0, si = 0;
    idx_t n = bra.numAmpsPerChunk;
#pragma omp parallel for schedule(static) reduction(+ : sr, si)
    for (idx_t i = 0; i < n; i++) {
        qreal bR = bra.stateVec.real[i], bI = bra.stateVec.imag[i];
        qreal kR = ket.stateVec.real[i], kI = ket.stateVec.imag[i];
        sr += bR * kR + bI * kI;
        si += bR * kI - bI * kR;
    }
    Complex out;
    out.real = sr;
    out.imag = si;
    return out;
}

qreal densmatr_calcPurity(Qureg q) {
    qreal sum = 0;
    idx_t n = q.numAmpsPerChunk;
#pragma omp parallel for schedule(static) reduction(+ : sum)
    for (idx_t i = 0; i < n; i++) sum += q.stateVec.real[i] * q.stateVec.real[i] + q.stateVec.imag[i] * q.stateVec.imag[i];
    return sum;
}

/* <psi| rho |psi>, real part (QuEST_cpu.c:916-992) */
qreal densmatr_calcFidelity(Qureg rho, Qureg psi) {
    idx_t dim = psi.numAmpsTotal;
    qreal total = 0;
#pragma omp parallel for schedule(static) reduction(+ : total)
    for (idx_t r = 0; r < dim; r++) {
        qreal pr = psi.stateVec.real[r], pi = -psi.stateVec.imag[r];
        qreal rowR = 0, rowI = 0;
        for (idx_t c = 0; c < dim; c++) {
            qreal dR = rho.stateVec.real[r + dim * c], dI = rho.stateVec.imag[r + dim * c];
            qreal vR = psi.stateVec.real[c], vI = psi.stateVec.imag[c];
            rowR += dR * vR - dI * vI;
            rowI += dR * vI + dI * vR;
        }
        total += rowR * pr - rowI * pi;
    }
    return total;
}

/* ---------------------------------------------------------------------------------------------------
 * collapse (QuEST_cpu.c:3023-3090 state-vector: kept half *= 1/sqrt(p); :727-852 density: kept block /= p)
 * ------------------------------------------------------------------------------------------------- */

void statevec_collapseToKnownProbOutcome(Qureg q, const int t, int outcome, qreal prob) {
    qreal renorm = 1 / sqrt(prob);
    idx_t n = q.numAmpsPerChunk, off = global_offset(q);
    qreal* re = q.stateVec.real;
    qreal* im = q.stateVec.imag;
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        if ((((i + off) >> t) & 1) == outcome) {
            re[i] = re[i] * renorm;
            im[i] = im[i] * renorm;
        } else {
            re[i] = 0;
            im[i] = 0;
        }
    }
}

void densmatr_collapseToKnownProbOutcome(Qureg q, const int t, int outcome, qreal prob) {
    idx_t n = q.numAmpsPerChunk, off = global_offset(q);
    int tCol = t + q.numQubitsRepresented;
    qreal* re = q.stateVec.real;
    qreal* im = q.stateVec.imag;
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        idx_t g = i + off;
        if (((g >> t) & 1) == outcome && ((g >> tCol) & 1) == outcome) {
            re[i] /= prob;
            im[i] /= prob;
        } else {
            re[i] = 0;
            im[i] = 0;
        }
    }
}

/* ---------------------------------------------------------------------------------------------------
 * density-matrix noise (QuEST_cpu.c:36-215, 379-479; QuEST_cpu_local.c:26-51)
 * A row bit t and its column partner t+n select four quadrants of rho: 00, 01, 10, 11.
 * ------------------------------------------------------------------------------------------------- */

static void scale_offdiag(Qureg q, int t, qreal retain) { /* QuEST_cpu.c:36-65 */
    idx_t n = q.numAmpsPerChunk, off = global_offset(q);
    idx_t inner = 1LL << t, outer = 1LL << (t + q.numQubitsRepresented), both = inner | outer;
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        idx_t pat = (i + off) & both;
        if (pat == inner || pat == outer) {
            q.stateVec.real[i] = retain * q.stateVec.real[i];
            q.stateVec.imag[i] = retain * q.stateVec.imag[i];
        }
    }
}

void densmatr_oneQubitDephase(Qureg q, const int t, qreal dephase) { scale_offdiag(q, t, 1 - dephase); }

void densmatr_twoQubitDephase(Qureg q, const int q1, const int q2, qreal dephase) { /* QuEST_cpu.c:77-116 */
    qreal retain = 1 - dephase;
    idx_t n = q.numAmpsPerChunk, off = global_offset(q);
    int nq = q.numQubitsRepresented;
    idx_t in1 = 1LL << q1, out1 = 1LL << (q1 + nq), tot1 = in1 | out1;
    idx_t in2 = 1LL << q2, out2 = 1LL << (q2 + nq), tot2 = in2 | out2;
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        idx_t p1 = (i + off) & tot1, p2 = (i + off) & tot2;
        if (p1 == in1 || p1 == out1 || p2 == in2 || p2 == out2) {
            q.stateVec.real[i] = retain * q.stateVec.real[i];
            q.stateVec.imag[i] = retain * q.stateVec.imag[i];
        }
    }
}

void densmatr_oneQubitDepolarise(Qureg q, const int t, qreal level) { /* QuEST_cpu_local.c:26-31, QuEST_cpu.c:118-165 */
    if (level == 0) return;
    qreal retain = 1 - level;
    idx_t n = q.numAmpsPerChunk, off = global_offset(q);
    idx_t inner = 1LL << t, outer = 1LL << (t + q.numQubitsRepresented), both = inner | outer;
    qreal* re = q.stateVec.real;
    qreal* im = q.stateVec.imag;
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        idx_t pat = (i + off) & both;
        if (pat == inner || pat == outer) {
            re[i] = retain * re[i];
            im[i] = retain * im[i];
        } else if ((i & both) == 0) {
            idx_t j = i | both;
            qreal avR = (re[i] + re[j]) / 2, avI = (im[i] + im[j]) / 2;
            re[i] = retain * re[i] + level * avR;
            im[i] = retain * im[i] + level * avI;
            re[j] = retain * re[j] + level * avR;
            im[j] = retain * im[j] + level * avI;
        }
    }
}

void densmatr_oneQubitDamping(Qureg q, const int t, qreal damping) { /* QuEST_cpu_local.c:33-38, QuEST_cpu.c:167-215 */
    if (damping == 0) return;
    qreal retain = 1 - damping;
    qreal dephase = sqrt(retain);
    idx_t n = q.numAmpsPerChunk, off = global_offset(q);
    idx_t inner = 1LL << t, outer = 1LL << (t + q.numQubitsRepresented), both = inner | outer;
    qreal* re = q.stateVec.real;
    qreal* im = q.stateVec.imag;
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        idx_t pat = (i + off) & both;
        if (pat == inner || pat == outer) {
            re[i] = dephase * re[i];
            im[i] = dephase * im[i];
        } else if ((i & both) == 0) {
            idx_t j = i | both;
            re[i] = re[i] + damping * re[j];
            im[i] = im[i] + damping * im[j];
            re[j] = retain * re[j];
            im[j] = retain * im[j];
        }
    }
}

/* three sweeps over the 16 two-qubit quadrants (QuEST_cpu.c:379-479), after a two-qubit dephase */
static void depol2_sweep(Qureg q, idx_t totSel, idx_t totOther, idx_t pairMask, qreal delta, qreal gamma) {
    idx_t n = q.numAmpsPerChunk, off = global_offset(q);
    qreal* re = q.stateVec.real;
    qreal* im = q.stateVec.imag;
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        idx_t pSel = (i + off) & totSel, pOther = (i + off) & totOther;
        if (pSel == 0 && (pOther == 0 || pOther == totOther)) {
            idx_t j = (i | totSel) ^ pairMask;
            qreal r0 = re[i], i0 = im[i];
            re[i] = gamma * (re[i] + delta * re[j]);
            im[i] = gamma * (im[i] + delta * im[j]);
            re[j] = gamma * (re[j] + delta * r0);
            im[j] = gamma * (im[j] + delta * i0);
        }
    }
}

void densmatr_twoQubitDepolarise(Qureg q, int q1, int q2, qreal level) { /* QuEST_cpu_local.c:40-51 */
    if (level == 0) return;
    qreal eta = 2 / level;
    qreal delta = eta - 1 - sqrt((eta - 1) * (eta - 1) - 1);
    qreal gamma = 1 + delta;
    gamma = 1 / (gamma * gamma * gamma);
    densmatr_twoQubitDephase(q, q1, q2, level);
    int nq = q.numQubitsRepresented;
    idx_t tot1 = (1LL << q1) | (1LL << (q1 + nq)), tot2 = (1LL << q2) | (1LL << (q2 + nq));
    /* steps one and two apply no gamma; writing x + delta*y as 1*(x + delta*y) is exact */
    depol2_sweep(q, tot1, tot2, 0, delta, 1);
    depol2_sweep(q, tot2, tot1, 0, delta, 1);
    depol2_sweep(q, tot2, tot1, tot1, delta, gamma);
}

void densmatr_addDensityMatrix(Qureg a, qreal p, Qureg b) { /* QuEST_cpu.c:883-913 */
    idx_t n = a.numAmpsPerChunk;
#pragma omp parallel for schedule(static)
    for (idx_t i = 0; i < n; i++) {
        a.stateVec.real[i] *= 1 - p;
        a.stateVec.imag[i] *= 1 - p;
        a.stateVec.real[i] += p * b.stateVec.real[i];
        a.stateVec.imag[i] += p * b.stateVec.imag[i];
    }
}
