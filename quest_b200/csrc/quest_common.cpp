// quest-b200: hardware-agnostic host glue between the API layer and a backend.
//
// What lives here is the part of the gate path that the reference also keeps on the host
// (QuEST/src/QuEST_common.c): turning rotations into (alpha, beta) pairs, turning Z-family gates into one
// phase term, and the measurement decision.  All trigonometry is done here with the host libm in double,
// exactly where the reference does it (QuEST_common.c:62-69,195-235), so that the device only ever sees
// finished coefficients and CPU/GPU results can be compared bit for bit.  libm is always called on doubles (explicit
// casts): the reference is C99, where cos(float) means cos((double)x), and some C++ tool-chains add float overloads.
#include <sys/time.h>
#include <unistd.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "quest_internal.h"

extern "C" {

// ---- small helpers --------------------------------------------------------------------------------

void ensureIndsIncrease(int* ind1, int* ind2) {
    if (*ind1 > *ind2) {
        int t = *ind1;
        *ind1 = *ind2;
        *ind2 = t;
    }
}

qreal getVectorMagnitude(Vector v) { return sqrt((double)(v.x * v.x + v.y * v.y + v.z * v.z)); }

Complex getConjugateScalar(Complex s) {
    Complex c;
    c.real = s.real;
    c.imag = -s.imag;
    return c;
}

ComplexMatrix2 getConjugateMatrix(ComplexMatrix2 m) {
    ComplexMatrix2 c;
    c.r0c0 = getConjugateScalar(m.r0c0);
    c.r0c1 = getConjugateScalar(m.r0c1);
    c.r1c0 = getConjugateScalar(m.r1c0);
    c.r1c1 = getConjugateScalar(m.r1c1);
    return c;
}

void shiftIndices(int* indices, int numIndices, int shift) {
    for (int j = 0; j < numIndices; ++j) indices[j] += shift;
}

// djb2, as the reference (QuEST_common.c:123-131)
unsigned long int hashString(char* str) {
    unsigned long int h = 5381;
    for (int c; (c = *str++);) h = h * 33 + c;
    return h;
}

// R(angle, axis) = exp(-i angle/2 axis.sigma) = [[alpha, -conj(beta)], [beta, conj(alpha)]]
// Follows QuEST_common.c:62-69: the axis is normalised first, then each component is formed as
// sin(angle/2)*component (product evaluated in double, stored as qreal).
void getComplexPairFromRotation(qreal angle, Vector axis, Complex* alpha, Complex* beta) {
    qreal mag = getVectorMagnitude(axis);
    Vector unit;
    unit.x = axis.x / mag;
    unit.y = axis.y / mag;
    unit.z = axis.z / mag;
    alpha->real = cos(angle / 2.0);
    alpha->imag = -sin(angle / 2.0) * unit.z;
    beta->real = sin(angle / 2.0) * unit.y;
    beta->imag = -sin(angle / 2.0) * unit.x;
}

// U(alpha, beta) = Rz(rz2) Ry(ry) Rz(rz1); used only for the QASM log (QuEST_common.c:72-81)
void getZYZRotAnglesFromComplexPair(Complex alpha, Complex beta, qreal* rz2, qreal* ry, qreal* rz1) {
    qreal alphaMag = sqrt((double)(alpha.real * alpha.real + alpha.imag * alpha.imag));
    *ry = 2.0 * acos((double)alphaMag);
    qreal alphaPhase = atan2((double)alpha.imag, (double)alpha.real);
    qreal betaPhase = atan2((double)beta.imag, (double)beta.real);
    *rz2 = -alphaPhase + betaPhase;
    *rz1 = -alphaPhase - betaPhase;
}

// u = exp(i phase) U(alpha, beta); used only for the QASM log (QuEST_common.c:84-96)
void getComplexPairAndPhaseFromUnitary(ComplexMatrix2 u, Complex* alpha, Complex* beta, qreal* globalPhase) {
    qreal p00 = atan2((double)u.r0c0.imag, (double)u.r0c0.real);
    qreal p11 = atan2((double)u.r1c1.imag, (double)u.r1c1.real);
    *globalPhase = (p00 + p11) / 2.0;
    qreal c = cos((double)*globalPhase);
    qreal s = sin((double)*globalPhase);
    alpha->real = u.r0c0.real * c + u.r0c0.imag * s;
    alpha->imag = u.r0c0.imag * c - u.r0c0.real * s;
    beta->real = u.r1c0.real * c + u.r1c0.imag * s;
    beta->imag = u.r1c0.imag * c - u.r1c0.real * s;
}

// ---- seeding --------------------------------------------------------------------------------------

// key = {wall-clock milliseconds, pid}  (QuEST_common.c:133-148)
void getQuESTDefaultSeedKey(unsigned long int* key) {
    struct timeval tv;
    gettimeofday(&tv, NULL);
    double ms = (tv.tv_sec) * 1000 + (tv.tv_usec) / 1000;
    key[0] = (unsigned long int)ms;
    key[1] = (unsigned long int)getpid();
}

void seedQuEST(unsigned long int* seedArray, int numSeeds) { init_by_array(seedArray, numSeeds); }

// ---- measurement decision -------------------------------------------------------------------------

// Branch order is load-bearing: no random draw is consumed when the outcome is (numerically) certain,
// otherwise outcome = (u > P(0)) with u on [0,1]  (QuEST_common.c:103-121).
int generateMeasurementOutcome(qreal zeroProb, qreal* outcomeProb) {
    int outcome;
    if (zeroProb < REAL_EPS)
        outcome = 1;
    else if (1 - zeroProb < REAL_EPS)
        outcome = 0;
    else
        outcome = (genrand_real1() > zeroProb);
    *outcomeProb = outcome ? 1 - zeroProb : zeroProb;
    return outcome;
}

int statevec_measureWithStats(Qureg qureg, int measureQubit, qreal* outcomeProb) {
    qreal zeroProb = statevec_calcProbOfOutcome(qureg, measureQubit, 0);
    int outcome = generateMeasurementOutcome(zeroProb, outcomeProb);
    statevec_collapseToKnownProbOutcome(qureg, measureQubit, outcome, *outcomeProb);
    return outcome;
}

int densmatr_measureWithStats(Qureg qureg, int measureQubit, qreal* outcomeProb) {
    qreal zeroProb = densmatr_calcProbOfOutcome(qureg, measureQubit, 0);
    int outcome = generateMeasurementOutcome(zeroProb, outcomeProb);
    densmatr_collapseToKnownProbOutcome(qureg, measureQubit, outcome, *outcomeProb);
    return outcome;
}

// ---- scalar conveniences --------------------------------------------------------------------------

qreal statevec_getProbAmp(Qureg qureg, long long int index) {
    qreal re = statevec_getRealAmp(qureg, index);
    qreal im = statevec_getImagAmp(qureg, index);
    return re * re + im * im;
}

qreal statevec_calcFidelity(Qureg qureg, Qureg pureState) {
    Complex ip = statevec_calcInnerProduct(qureg, pureState);
    return ip.real * ip.real + ip.imag * ip.imag;
}

// ---- Z-family gates: one multiplicative term on the |1> half (QuEST_common.c:195-235) ------------

static void phaseByTerm(Qureg q, int target, qreal re, qreal im) {
    Complex term;
    term.real = re;
    term.imag = im;
    statevec_phaseShiftByTerm(q, target, term);
}

void statevec_phaseShift(Qureg q, const int t, qreal angle) { phaseByTerm(q, t, cos((double)angle), sin((double)angle)); }
void statevec_pauliZ(Qureg q, const int t) { phaseByTerm(q, t, -1, 0); }
void statevec_sGate(Qureg q, const int t) { phaseByTerm(q, t, 0, 1); }
void statevec_sGateConj(Qureg q, const int t) { phaseByTerm(q, t, 0, -1); }
void statevec_tGate(Qureg q, const int t) { phaseByTerm(q, t, 1 / sqrt(2), 1 / sqrt(2)); }
void statevec_tGateConj(Qureg q, const int t) { phaseByTerm(q, t, 1 / sqrt(2), -1 / sqrt(2)); }

// ---- rotations: all reduce to (controlled) compactUnitary (QuEST_common.c:237-303) ----------------

static const Vector kX = {1, 0, 0}, kY = {0, 1, 0}, kZ = {0, 0, 1};

static void rotationPair(qreal angle, Vector axis, bool conj, Complex* a, Complex* b) {
    getComplexPairFromRotation(angle, axis, a, b);
    if (conj) {
        a->imag *= -1;
        b->imag *= -1;
    }
}

void statevec_rotateAroundAxis(Qureg q, const int t, qreal angle, Vector axis) {
    Complex a, b;
    rotationPair(angle, axis, false, &a, &b);
    statevec_compactUnitary(q, t, a, b);
}

void statevec_rotateAroundAxisConj(Qureg q, const int t, qreal angle, Vector axis) {
    Complex a, b;
    rotationPair(angle, axis, true, &a, &b);
    statevec_compactUnitary(q, t, a, b);
}

void statevec_controlledRotateAroundAxis(Qureg q, const int c, const int t, qreal angle, Vector axis) {
    Complex a, b;
    rotationPair(angle, axis, false, &a, &b);
    statevec_controlledCompactUnitary(q, c, t, a, b);
}

void statevec_controlledRotateAroundAxisConj(Qureg q, const int c, const int t, qreal angle, Vector axis) {
    Complex a, b;
    rotationPair(angle, axis, true, &a, &b);
    statevec_controlledCompactUnitary(q, c, t, a, b);
}

void statevec_rotateX(Qureg q, const int t, qreal angle) { statevec_rotateAroundAxis(q, t, angle, kX); }
void statevec_rotateY(Qureg q, const int t, qreal angle) { statevec_rotateAroundAxis(q, t, angle, kY); }
void statevec_rotateZ(Qureg q, const int t, qreal angle) { statevec_rotateAroundAxis(q, t, angle, kZ); }
void statevec_controlledRotateX(Qureg q, const int c, const int t, qreal angle) {
    statevec_controlledRotateAroundAxis(q, c, t, angle, kX);
}
void statevec_controlledRotateY(Qureg q, const int c, const int t, qreal angle) {
    statevec_controlledRotateAroundAxis(q, c, t, angle, kY);
}
void statevec_controlledRotateZ(Qureg q, const int c, const int t, qreal angle) {
    statevec_controlledRotateAroundAxis(q, c, t, angle, kZ);
}

}  // extern "C"
