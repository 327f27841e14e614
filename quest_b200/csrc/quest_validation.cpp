// quest-b200: argument validation for the QuEST C API.
//
// Error convention kept from the reference (QuEST/src/QuEST_validation.c:82-92): an invalid argument prints
//   "!!!\nQuEST Error in function <api name>: <message>\n!!!\nexiting..\n"
// and calls exit(<code>), where <code> is the position of the error in the reference's ErrorCode enum
// (QuEST_validation.c:19-49) and <message> is the reference's text (:51-80).  No error returns, no errno.
// Numerical checks use REAL_EPS exactly as the reference does (:94-128).
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "quest_internal.h"

namespace {

struct Fault {
    int code;
    const char* text;
};

// code = exit status = index in the reference's ErrorCode enum
constexpr Fault kBadNumQubits{1, "Invalid number of qubits. Must create >0."};
constexpr Fault kBadTarget{2, "Invalid target qubit. Note qubits are zero indexed."};
constexpr Fault kBadControl{3, "Invalid control qubit. Note qubits are zero indexed."};
constexpr Fault kBadStateIndex{4, "Invalid state index. Must be >=0 and <2^numQubits."};
constexpr Fault kBadNumAmps{5, "Invalid number of amplitudes. Must be >=0 and <=2^numQubits."};
constexpr Fault kBadOffsetNumAmps{6, "More amplitudes given than exist in the statevector from the given starting index."};
constexpr Fault kTargetIsControl{7, "Control qubit cannot equal target qubit."};
constexpr Fault kTargetInControls{8, "Control qubits cannot include target qubit."};
constexpr Fault kTargetsNotUnique{9, "The two target qubits must be unique."};
constexpr Fault kBadNumControls{10, "Invalid number of control qubits. Must be >0 and <numQubits."};
constexpr Fault kNonUnitaryMatrix{11, "Matrix is not unitary."};
constexpr Fault kNonUnitaryPair{12, "Compact matrix formed by given complex numbers is not unitary."};
constexpr Fault kZeroVector{13, "Invalid axis vector. Must be non-zero."};
constexpr Fault kCollapseZeroProb{15, "Can't collapse to state with zero probability."};
constexpr Fault kBadOutcome{16, "Invalid measurement outcome -- must be either 0 or 1."};
constexpr Fault kCannotOpenFile{17, "Could not open file"};
constexpr Fault kSecondMustBeStatevec{18, "Second argument must be a state-vector."};
constexpr Fault kDimsMismatch{19, "Dimensions of the qubit registers don't match."};
constexpr Fault kTypesMismatch{20, "Registers must both be state-vectors or both be density matrices."};
constexpr Fault kOnlyStatevecs{21, "Operation valid only for state-vectors."};
constexpr Fault kOnlyDensmatrs{22, "Operation valid only for density matrices."};
constexpr Fault kBadProb{23, "Probabilities must be in [0, 1]."};
constexpr Fault kBadOneQubitDephase{25, "The probability of a single qubit dephase error cannot exceed 1/2, which maximally mixes."};
constexpr Fault kBadTwoQubitDephase{26, "The probability of a two-qubit qubit dephase error cannot exceed 3/4, which maximally mixes."};
constexpr Fault kBadOneQubitDepol{27, "The probability of a single qubit depolarising error cannot exceed 3/4, which maximally mixes."};
constexpr Fault kBadTwoQubitDepol{28, "The probability of a two-qubit depolarising error cannot exceed 15/16, which maximally mixes."};

void require(bool ok, const Fault& f, const char* caller) {
    if (ok) return;
    std::printf("!!!\n");
    std::printf("QuEST Error in function %s: %s\n", caller, f.text);
    std::printf("!!!\n");
    std::printf("exiting..\n");
    std::fflush(stdout);
    std::exit(f.code);
}

bool pairIsUnitary(Complex a, Complex b) {
    return absReal(-1 + a.real * a.real + a.imag * a.imag + b.real * b.real + b.imag * b.imag) < REAL_EPS;
}

// columns have unit norm and are orthogonal (real and imaginary part of the inner product)
bool matrixIsUnitary(ComplexMatrix2 u) {
    if (absReal(u.r0c0.real * u.r0c0.real + u.r0c0.imag * u.r0c0.imag + u.r1c0.real * u.r1c0.real +
                u.r1c0.imag * u.r1c0.imag - 1) > REAL_EPS)
        return false;
    if (absReal(u.r0c1.real * u.r0c1.real + u.r0c1.imag * u.r0c1.imag + u.r1c1.real * u.r1c1.real +
                u.r1c1.imag * u.r1c1.imag - 1) > REAL_EPS)
        return false;
    if (absReal(u.r0c0.real * u.r0c1.real + u.r0c0.imag * u.r0c1.imag + u.r1c0.real * u.r1c1.real +
                u.r1c0.imag * u.r1c1.imag) > REAL_EPS)
        return false;
    if (absReal(u.r0c1.real * u.r0c0.imag - u.r0c0.real * u.r0c1.imag + u.r1c1.real * u.r1c0.imag -
                u.r1c0.real * u.r1c1.imag) > REAL_EPS)
        return false;
    return true;
}

bool qubitInRange(Qureg q, int qubit) { return qubit >= 0 && qubit < q.numQubitsRepresented; }

}  // namespace

extern "C" {

void validateCreateNumQubits(int numQubits, const char* caller) { require(numQubits > 0, kBadNumQubits, caller); }

void validateStateIndex(Qureg qureg, long long int stateInd, const char* caller) {
    long long int limit = 1LL << qureg.numQubitsRepresented;
    require(stateInd >= 0 && stateInd < limit, kBadStateIndex, caller);
}

void validateNumAmps(Qureg qureg, long long int startInd, long long int numAmps, const char* caller) {
    validateStateIndex(qureg, startInd, caller);
    require(numAmps >= 0 && numAmps <= qureg.numAmpsTotal, kBadNumAmps, caller);
    require(numAmps + startInd <= qureg.numAmpsTotal, kBadOffsetNumAmps, caller);
}

void validateTarget(Qureg qureg, int targetQubit, const char* caller) {
    require(qubitInRange(qureg, targetQubit), kBadTarget, caller);
}

void validateControlTarget(Qureg qureg, int controlQubit, int targetQubit, const char* caller) {
    validateTarget(qureg, targetQubit, caller);
    require(qubitInRange(qureg, controlQubit), kBadControl, caller);
    require(controlQubit != targetQubit, kTargetIsControl, caller);
}

void validateUniqueTargets(Qureg qureg, int qubit1, int qubit2, const char* caller) {
    validateTarget(qureg, qubit1, caller);
    validateTarget(qureg, qubit2, caller);
    require(qubit1 != qubit2, kTargetsNotUnique, caller);
}

void validateMultiControls(Qureg qureg, int* controlQubits, const int numControlQubits, const char* caller) {
    require(numControlQubits > 0 && numControlQubits <= qureg.numQubitsRepresented, kBadNumControls, caller);
    for (int i = 0; i < numControlQubits; ++i) require(qubitInRange(qureg, controlQubits[i]), kBadControl, caller);
}

void validateMultiControlsTarget(Qureg qureg, int* controlQubits, const int numControlQubits, const int targetQubit,
                                 const char* caller) {
    validateTarget(qureg, targetQubit, caller);
    validateMultiControls(qureg, controlQubits, numControlQubits, caller);
    for (int i = 0; i < numControlQubits; ++i) require(controlQubits[i] != targetQubit, kTargetInControls, caller);
}

void validateUnitaryMatrix(ComplexMatrix2 u, const char* caller) { require(matrixIsUnitary(u), kNonUnitaryMatrix, caller); }

void validateUnitaryComplexPair(Complex alpha, Complex beta, const char* caller) {
    require(pairIsUnitary(alpha, beta), kNonUnitaryPair, caller);
}

void validateVector(Vector vec, const char* caller) { require(getVectorMagnitude(vec) > REAL_EPS, kZeroVector, caller); }

void validateStateVecQureg(Qureg qureg, const char* caller) { require(!qureg.isDensityMatrix, kOnlyStatevecs, caller); }

void validateDensityMatrQureg(Qureg qureg, const char* caller) { require(qureg.isDensityMatrix, kOnlyDensmatrs, caller); }

void validateOutcome(int outcome, const char* caller) { require(outcome == 0 || outcome == 1, kBadOutcome, caller); }

void validateMeasurementProb(qreal prob, const char* caller) { require(prob > REAL_EPS, kCollapseZeroProb, caller); }

void validateMatchingQuregDims(Qureg a, Qureg b, const char* caller) {
    require(a.numQubitsRepresented == b.numQubitsRepresented, kDimsMismatch, caller);
}

void validateMatchingQuregTypes(Qureg a, Qureg b, const char* caller) {
    require(a.isDensityMatrix == b.isDensityMatrix, kTypesMismatch, caller);
}

void validateSecondQuregStateVec(Qureg qureg2, const char* caller) {
    require(!qureg2.isDensityMatrix, kSecondMustBeStatevec, caller);
}

void validateFileOpened(int found, const char* caller) { require(found, kCannotOpenFile, caller); }

void validateProb(qreal prob, const char* caller) { require(prob >= 0 && prob <= 1, kBadProb, caller); }

void validateOneQubitDephaseProb(qreal prob, const char* caller) {
    validateProb(prob, caller);
    require(prob <= 1 / 2.0, kBadOneQubitDephase, caller);
}

void validateTwoQubitDephaseProb(qreal prob, const char* caller) {
    validateProb(prob, caller);
    require(prob <= 3 / 4.0, kBadTwoQubitDephase, caller);
}

void validateOneQubitDepolProb(qreal prob, const char* caller) {
    validateProb(prob, caller);
    require(prob <= 3 / 4.0, kBadOneQubitDepol, caller);
}

// the reference reuses the one-qubit depolarising message for damping (QuEST_validation.c:255-258)
void validateOneQubitDampingProb(qreal prob, const char* caller) {
    validateProb(prob, caller);
    require(prob <= 1.0, kBadOneQubitDepol, caller);
}

void validateTwoQubitDepolProb(qreal prob, const char* caller) {
    validateProb(prob, caller);
    require(prob <= 15 / 16.0, kBadTwoQubitDepol, caller);
}

}  // extern "C"
