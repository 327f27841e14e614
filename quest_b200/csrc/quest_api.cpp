// quest-b200: the QuEST.h entry points.
//
// Every call is: validate -> one or two backend passes -> QASM record, the same three steps the reference
// front-end performs (QuEST/src/QuEST.c).  A density matrix rho of n qubits is stored as a 2n-qubit state-vector
// and U rho U^dagger is applied as U on qubit t followed by conj(U) on qubit t+n (QuEST.c:8-10,167-481); which
// gates need an explicit conjugate (negated angle, conjugated coefficients, "Conj" kernel) is encoded per gate
// below exactly as the reference chooses it.  Nothing in this file touches amplitudes or the GPU.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "quest_internal.h"

namespace {

// Runs `pass(shift, conjugate)` once for a state-vector, and a second time on the column qubits for a density
// matrix.  `shift` is what must be added to every qubit index of the second pass.
template <class Pass>
inline void onBothSides(const Qureg& q, Pass&& pass) {
    pass(0, false);
    if (q.isDensityMatrix) pass(q.numQubitsRepresented, true);
}

// multi-control lists are caller-owned; the reference shifts them in place for the second pass and shifts
// them back (QuEST.c:279-281) - same here, so the caller's array is observably unchanged on return.
struct ShiftedControls {
    int* ctrls;
    int n, shift;
    ShiftedControls(int* c, int n_, int s) : ctrls(c), n(n_), shift(s) { shiftIndices(ctrls, n, shift); }
    ~ShiftedControls() { shiftIndices(ctrls, n, -shift); }
};

}  // namespace

extern "C" {

// ---- registers ------------------------------------------------------------------------------------ (QuEST.c:28-59)

Qureg createQureg(int numQubits, QuESTEnv env) {
    validateCreateNumQubits(numQubits, __func__);
    Qureg q;
    statevec_createQureg(&q, numQubits, env);
    q.isDensityMatrix = 0;
    q.numQubitsRepresented = numQubits;
    q.numQubitsInStateVec = numQubits;
    qasm_setup(&q);
    initZeroState(q);
    return q;
}

Qureg createDensityQureg(int numQubits, QuESTEnv env) {
    validateCreateNumQubits(numQubits, __func__);
    Qureg q;
    statevec_createQureg(&q, 2 * numQubits, env);
    q.isDensityMatrix = 1;
    q.numQubitsRepresented = numQubits;
    q.numQubitsInStateVec = 2 * numQubits;
    qasm_setup(&q);
    initZeroState(q);
    return q;
}

void destroyQureg(Qureg qureg, QuESTEnv env) {
    statevec_destroyQureg(qureg, env);
    qasm_free(qureg);
}

int getNumQubits(Qureg qureg) { return qureg.numQubitsRepresented; }

int getNumAmps(Qureg qureg) {
    validateStateVecQureg(qureg, __func__);
    return (int)qureg.numAmpsTotal;
}

// ---- QASM ----------------------------------------------------------------------------------------- (QuEST.c:66-85)

void startRecordingQASM(Qureg qureg) { qasm_startRecording(qureg); }
void stopRecordingQASM(Qureg qureg) { qasm_stopRecording(qureg); }
void clearRecordedQASM(Qureg qureg) { qasm_clearRecorded(qureg); }
void printRecordedQASM(Qureg qureg) { qasm_printRecorded(qureg); }
void writeRecordedQASMToFile(Qureg qureg, char* filename) {
    validateFileOpened(qasm_writeRecordedToFile(qureg, filename), __func__);
}

// ---- initialisation ------------------------------------------------------------------------------- (QuEST.c:92-160)

void initZeroState(Qureg qureg) {
    statevec_initZeroState(qureg);  // |0..0><0..0| flattened is again "amplitude 0 is one"
    qasm_recordInitZero(qureg);
}

void initPlusState(Qureg qureg) {
    if (qureg.isDensityMatrix)
        densmatr_initPlusState(qureg);
    else
        statevec_initPlusState(qureg);
    qasm_recordInitPlus(qureg);
}

void initClassicalState(Qureg qureg, long long int stateInd) {
    validateStateIndex(qureg, stateInd, __func__);
    if (qureg.isDensityMatrix)
        densmatr_initClassicalState(qureg, stateInd);
    else
        statevec_initClassicalState(qureg, stateInd);
    qasm_recordInitClassical(qureg, stateInd);
}

void initPureState(Qureg qureg, Qureg pure) {
    validateSecondQuregStateVec(pure, __func__);
    validateMatchingQuregDims(qureg, pure, __func__);
    if (qureg.isDensityMatrix)
        densmatr_initPureState(qureg, pure);
    else
        statevec_cloneQureg(qureg, pure);
    qasm_recordComment(qureg, "Here, the register was initialised to an undisclosed given pure state.");
}

void initStateFromAmps(Qureg qureg, qreal* reals, qreal* imags) {
    validateStateVecQureg(qureg, __func__);
    statevec_setAmps(qureg, 0, reals, imags, qureg.numAmpsTotal);
    qasm_recordComment(qureg, "Here, the register was initialised to an undisclosed given pure state.");
}

void setAmps(Qureg qureg, long long int startInd, qreal* reals, qreal* imags, long long int numAmps) {
    validateStateVecQureg(qureg, __func__);
    validateNumAmps(qureg, startInd, numAmps, __func__);
    statevec_setAmps(qureg, startInd, reals, imags, numAmps);
    qasm_recordComment(qureg, "Here, some amplitudes in the statevector were manually edited.");
}

void setDensityAmps(Qureg qureg, qreal* reals, qreal* imags) {
    statevec_setAmps(qureg, 0, reals, imags, qureg.numAmpsTotal);
    qasm_recordComment(qureg, "Here, some amplitudes in the density matrix were manually edited.");
}

void cloneQureg(Qureg targetQureg, Qureg copyQureg) {
    validateMatchingQuregTypes(targetQureg, copyQureg, __func__);
    validateMatchingQuregDims(targetQureg, copyQureg, __func__);
    statevec_cloneQureg(targetQureg, copyQureg);
}

// ---- one-target dense gates ----------------------------------------------------------------------- (QuEST.c:167-346)

void hadamard(Qureg qureg, const int targetQubit) {
    validateTarget(qureg, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool) { statevec_hadamard(qureg, targetQubit + s); });  // H is real
    qasm_recordGate(qureg, GATE_HADAMARD, targetQubit);
}

void pauliX(Qureg qureg, const int targetQubit) {
    validateTarget(qureg, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool) { statevec_pauliX(qureg, targetQubit + s); });
    qasm_recordGate(qureg, GATE_SIGMA_X, targetQubit);
}

void pauliY(Qureg qureg, const int targetQubit) {
    validateTarget(qureg, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        if (conj)
            statevec_pauliYConj(qureg, targetQubit + s);
        else
            statevec_pauliY(qureg, targetQubit + s);
    });
    qasm_recordGate(qureg, GATE_SIGMA_Y, targetQubit);
}

// conj(Rx(a)) = Rx(-a), conj(Ry(a)) = Ry(a), conj(Rz(a)) = Rz(-a)                        (QuEST.c:178-209)
void rotateX(Qureg qureg, const int targetQubit, qreal angle) {
    validateTarget(qureg, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool conj) { statevec_rotateX(qureg, targetQubit + s, conj ? -angle : angle); });
    qasm_recordParamGate(qureg, GATE_ROTATE_X, targetQubit, angle);
}

void rotateY(Qureg qureg, const int targetQubit, qreal angle) {
    validateTarget(qureg, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool) { statevec_rotateY(qureg, targetQubit + s, angle); });
    qasm_recordParamGate(qureg, GATE_ROTATE_Y, targetQubit, angle);
}

void rotateZ(Qureg qureg, const int targetQubit, qreal angle) {
    validateTarget(qureg, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool conj) { statevec_rotateZ(qureg, targetQubit + s, conj ? -angle : angle); });
    qasm_recordParamGate(qureg, GATE_ROTATE_Z, targetQubit, angle);
}

void rotateAroundAxis(Qureg qureg, const int rotQubit, qreal angle, Vector axis) {
    validateTarget(qureg, rotQubit, __func__);
    validateVector(axis, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        if (conj)
            statevec_rotateAroundAxisConj(qureg, rotQubit + s, angle, axis);
        else
            statevec_rotateAroundAxis(qureg, rotQubit + s, angle, axis);
    });
    qasm_recordAxisRotation(qureg, angle, axis, rotQubit);
}

void compactUnitary(Qureg qureg, const int targetQubit, Complex alpha, Complex beta) {
    validateTarget(qureg, targetQubit, __func__);
    validateUnitaryComplexPair(alpha, beta, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        statevec_compactUnitary(qureg, targetQubit + s, conj ? getConjugateScalar(alpha) : alpha,
                                conj ? getConjugateScalar(beta) : beta);
    });
    qasm_recordCompactUnitary(qureg, alpha, beta, targetQubit);
}

void unitary(Qureg qureg, const int targetQubit, ComplexMatrix2 u) {
    validateTarget(qureg, targetQubit, __func__);
    validateUnitaryMatrix(u, __func__);
    onBothSides(qureg, [&](int s, bool conj) { statevec_unitary(qureg, targetQubit + s, conj ? getConjugateMatrix(u) : u); });
    qasm_recordUnitary(qureg, u, targetQubit);
}

// ---- one-control dense gates ---------------------------------------------------------------------- (QuEST.c:211-313,407-429)

void controlledNot(Qureg qureg, const int controlQubit, const int targetQubit) {
    validateControlTarget(qureg, controlQubit, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool) { statevec_controlledNot(qureg, controlQubit + s, targetQubit + s); });
    qasm_recordControlledGate(qureg, GATE_SIGMA_X, controlQubit, targetQubit);
}

void controlledPauliY(Qureg qureg, const int controlQubit, const int targetQubit) {
    validateControlTarget(qureg, controlQubit, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        if (conj)
            statevec_controlledPauliYConj(qureg, controlQubit + s, targetQubit + s);
        else
            statevec_controlledPauliY(qureg, controlQubit + s, targetQubit + s);
    });
    qasm_recordControlledGate(qureg, GATE_SIGMA_Y, controlQubit, targetQubit);
}

void controlledRotateX(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle) {
    validateControlTarget(qureg, controlQubit, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        statevec_controlledRotateX(qureg, controlQubit + s, targetQubit + s, conj ? -angle : angle);
    });
    qasm_recordControlledParamGate(qureg, GATE_ROTATE_X, controlQubit, targetQubit, angle);
}

void controlledRotateY(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle) {
    validateControlTarget(qureg, controlQubit, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool) { statevec_controlledRotateY(qureg, controlQubit + s, targetQubit + s, angle); });
    qasm_recordControlledParamGate(qureg, GATE_ROTATE_Y, controlQubit, targetQubit, angle);
}

void controlledRotateZ(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle) {
    validateControlTarget(qureg, controlQubit, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        statevec_controlledRotateZ(qureg, controlQubit + s, targetQubit + s, conj ? -angle : angle);
    });
    qasm_recordControlledParamGate(qureg, GATE_ROTATE_Z, controlQubit, targetQubit, angle);
}

void controlledRotateAroundAxis(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle, Vector axis) {
    validateControlTarget(qureg, controlQubit, targetQubit, __func__);
    validateVector(axis, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        if (conj)
            statevec_controlledRotateAroundAxisConj(qureg, controlQubit + s, targetQubit + s, angle, axis);
        else
            statevec_controlledRotateAroundAxis(qureg, controlQubit + s, targetQubit + s, angle, axis);
    });
    qasm_recordControlledAxisRotation(qureg, angle, axis, controlQubit, targetQubit);
}

void controlledCompactUnitary(Qureg qureg, const int controlQubit, const int targetQubit, Complex alpha, Complex beta) {
    validateControlTarget(qureg, controlQubit, targetQubit, __func__);
    validateUnitaryComplexPair(alpha, beta, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        statevec_controlledCompactUnitary(qureg, controlQubit + s, targetQubit + s,
                                          conj ? getConjugateScalar(alpha) : alpha,
                                          conj ? getConjugateScalar(beta) : beta);
    });
    qasm_recordControlledCompactUnitary(qureg, alpha, beta, controlQubit, targetQubit);
}

void controlledUnitary(Qureg qureg, const int controlQubit, const int targetQubit, ComplexMatrix2 u) {
    validateControlTarget(qureg, controlQubit, targetQubit, __func__);
    validateUnitaryMatrix(u, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        statevec_controlledUnitary(qureg, controlQubit + s, targetQubit + s, conj ? getConjugateMatrix(u) : u);
    });
    qasm_recordControlledUnitary(qureg, u, controlQubit, targetQubit);
}

void multiControlledUnitary(Qureg qureg, int* controlQubits, const int numControlQubits, const int targetQubit,
                            ComplexMatrix2 u) {
    validateMultiControlsTarget(qureg, controlQubits, numControlQubits, targetQubit, __func__);
    validateUnitaryMatrix(u, __func__);
    statevec_multiControlledUnitary(qureg, controlQubits, numControlQubits, targetQubit, u);
    if (qureg.isDensityMatrix) {
        int shift = qureg.numQubitsRepresented;
        ShiftedControls guard(controlQubits, numControlQubits, shift);
        statevec_multiControlledUnitary(qureg, controlQubits, numControlQubits, targetQubit + shift, getConjugateMatrix(u));
    }
    qasm_recordMultiControlledUnitary(qureg, u, controlQubits, numControlQubits, targetQubit);
}

// ---- diagonal gates ------------------------------------------------------------------------------- (QuEST.c:337-455)

void pauliZ(Qureg qureg, const int targetQubit) {
    validateTarget(qureg, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool) { statevec_pauliZ(qureg, targetQubit + s); });
    qasm_recordGate(qureg, GATE_SIGMA_Z, targetQubit);
}

void sGate(Qureg qureg, const int targetQubit) {
    validateTarget(qureg, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        if (conj)
            statevec_sGateConj(qureg, targetQubit + s);
        else
            statevec_sGate(qureg, targetQubit + s);
    });
    qasm_recordGate(qureg, GATE_S, targetQubit);
}

void tGate(Qureg qureg, const int targetQubit) {
    validateTarget(qureg, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        if (conj)
            statevec_tGateConj(qureg, targetQubit + s);
        else
            statevec_tGate(qureg, targetQubit + s);
    });
    qasm_recordGate(qureg, GATE_T, targetQubit);
}

void phaseShift(Qureg qureg, const int targetQubit, qreal angle) {
    validateTarget(qureg, targetQubit, __func__);
    onBothSides(qureg, [&](int s, bool conj) { statevec_phaseShift(qureg, targetQubit + s, conj ? -angle : angle); });
    qasm_recordParamGate(qureg, GATE_PHASE_SHIFT, targetQubit, angle);
}

void controlledPhaseShift(Qureg qureg, const int idQubit1, const int idQubit2, qreal angle) {
    validateControlTarget(qureg, idQubit1, idQubit2, __func__);
    onBothSides(qureg, [&](int s, bool conj) {
        statevec_controlledPhaseShift(qureg, idQubit1 + s, idQubit2 + s, conj ? -angle : angle);
    });
    qasm_recordControlledParamGate(qureg, GATE_PHASE_SHIFT, idQubit1, idQubit2, angle);
}

void multiControlledPhaseShift(Qureg qureg, int* controlQubits, int numControlQubits, qreal angle) {
    validateMultiControls(qureg, controlQubits, numControlQubits, __func__);
    statevec_multiControlledPhaseShift(qureg, controlQubits, numControlQubits, angle);
    if (qureg.isDensityMatrix) {
        ShiftedControls guard(controlQubits, numControlQubits, qureg.numQubitsRepresented);
        statevec_multiControlledPhaseShift(qureg, controlQubits, numControlQubits, -angle);
    }
    qasm_recordMultiControlledParamGate(qureg, GATE_PHASE_SHIFT, controlQubits, numControlQubits - 1,
                                        controlQubits[numControlQubits - 1], angle);
}

void controlledPhaseFlip(Qureg qureg, const int idQubit1, const int idQubit2) {
    validateControlTarget(qureg, idQubit1, idQubit2, __func__);
    onBothSides(qureg, [&](int s, bool) { statevec_controlledPhaseFlip(qureg, idQubit1 + s, idQubit2 + s); });
    qasm_recordControlledGate(qureg, GATE_SIGMA_Z, idQubit1, idQubit2);
}

void multiControlledPhaseFlip(Qureg qureg, int* controlQubits, int numControlQubits) {
    validateMultiControls(qureg, controlQubits, numControlQubits, __func__);
    statevec_multiControlledPhaseFlip(qureg, controlQubits, numControlQubits);
    if (qureg.isDensityMatrix) {
        ShiftedControls guard(controlQubits, numControlQubits, qureg.numQubitsRepresented);
        statevec_multiControlledPhaseFlip(qureg, controlQubits, numControlQubits);
    }
    qasm_recordMultiControlledGate(qureg, GATE_SIGMA_Z, controlQubits, numControlQubits - 1,
                                   controlQubits[numControlQubits - 1]);
}

// ---- amplitude access ----------------------------------------------------------------------------- (QuEST.c:498-539)

qreal getRealAmp(Qureg qureg, long long int index) {
    validateStateVecQureg(qureg, __func__);
    validateStateIndex(qureg, index, __func__);
    return statevec_getRealAmp(qureg, index);
}

qreal getImagAmp(Qureg qureg, long long int index) {
    validateStateVecQureg(qureg, __func__);
    validateStateIndex(qureg, index, __func__);
    return statevec_getImagAmp(qureg, index);
}

qreal getProbAmp(Qureg qureg, long long int index) {
    validateStateVecQureg(qureg, __func__);
    validateStateIndex(qureg, index, __func__);
    return statevec_getProbAmp(qureg, index);
}

Complex getAmp(Qureg qureg, long long int index) {
    validateStateVecQureg(qureg, __func__);
    validateStateIndex(qureg, index, __func__);
    Complex amp;
    amp.real = statevec_getRealAmp(qureg, index);
    amp.imag = statevec_getImagAmp(qureg, index);
    return amp;
}

Complex getDensityAmp(Qureg qureg, long long int row, long long int col) {
    validateDensityMatrQureg(qureg, __func__);
    validateStateIndex(qureg, row, __func__);
    validateStateIndex(qureg, col, __func__);
    long long ind = row + col * (1LL << qureg.numQubitsRepresented);
    Complex amp;
    amp.real = statevec_getRealAmp(qureg, ind);
    amp.imag = statevec_getImagAmp(qureg, ind);
    return amp;
}

// ---- measurement ---------------------------------------------------------------------------------- (QuEST.c:546-599)

qreal collapseToOutcome(Qureg qureg, const int measureQubit, int outcome) {
    validateTarget(qureg, measureQubit, __func__);
    validateOutcome(outcome, __func__);
    qreal prob;
    if (qureg.isDensityMatrix) {
        prob = densmatr_calcProbOfOutcome(qureg, measureQubit, outcome);
        validateMeasurementProb(prob, __func__);
        densmatr_collapseToKnownProbOutcome(qureg, measureQubit, outcome, prob);
    } else {
        prob = statevec_calcProbOfOutcome(qureg, measureQubit, outcome);
        validateMeasurementProb(prob, __func__);
        statevec_collapseToKnownProbOutcome(qureg, measureQubit, outcome, prob);
    }
    qasm_recordMeasurement(qureg, measureQubit);
    return prob;
}

int measureWithStats(Qureg qureg, int measureQubit, qreal* outcomeProb) {
    validateTarget(qureg, measureQubit, __func__);
    int outcome = qureg.isDensityMatrix ? densmatr_measureWithStats(qureg, measureQubit, outcomeProb)
                                        : statevec_measureWithStats(qureg, measureQubit, outcomeProb);
    qasm_recordMeasurement(qureg, measureQubit);
    return outcome;
}

int measure(Qureg qureg, int measureQubit) {
    validateTarget(qureg, measureQubit, __func__);
    qreal unusedProb;
    int outcome = qureg.isDensityMatrix ? densmatr_measureWithStats(qureg, measureQubit, &unusedProb)
                                        : statevec_measureWithStats(qureg, measureQubit, &unusedProb);
    qasm_recordMeasurement(qureg, measureQubit);
    return outcome;
}

// ---- calculations --------------------------------------------------------------------------------- (QuEST.c:606-645)

qreal calcTotalProb(Qureg qureg) {
    return qureg.isDensityMatrix ? densmatr_calcTotalProb(qureg) : statevec_calcTotalProb(qureg);
}

Complex calcInnerProduct(Qureg bra, Qureg ket) {
    validateStateVecQureg(bra, __func__);
    validateStateVecQureg(ket, __func__);
    validateMatchingQuregDims(bra, ket, __func__);
    return statevec_calcInnerProduct(bra, ket);
}

qreal calcProbOfOutcome(Qureg qureg, const int measureQubit, int outcome) {
    validateTarget(qureg, measureQubit, __func__);
    validateOutcome(outcome, __func__);
    return qureg.isDensityMatrix ? densmatr_calcProbOfOutcome(qureg, measureQubit, outcome)
                                 : statevec_calcProbOfOutcome(qureg, measureQubit, outcome);
}

qreal calcPurity(Qureg qureg) {
    validateDensityMatrQureg(qureg, __func__);
    return densmatr_calcPurity(qureg);
}

qreal calcFidelity(Qureg qureg, Qureg pureState) {
    validateSecondQuregStateVec(pureState, __func__);
    validateMatchingQuregDims(qureg, pureState, __func__);
    return qureg.isDensityMatrix ? densmatr_calcFidelity(qureg, pureState) : statevec_calcFidelity(qureg, pureState);
}

// ---- decoherence: the API probability is rescaled to the channel parameter here -------------------- (QuEST.c:592-692)

void addDensityMatrix(Qureg combineQureg, qreal otherProb, Qureg otherQureg) {
    validateDensityMatrQureg(combineQureg, __func__);
    validateDensityMatrQureg(otherQureg, __func__);
    validateMatchingQuregDims(combineQureg, otherQureg, __func__);
    validateProb(otherProb, __func__);
    densmatr_addDensityMatrix(combineQureg, otherProb, otherQureg);
}

static void oneQubitDephase(Qureg qureg, int targetQubit, qreal prob, const char* name) {
    validateDensityMatrQureg(qureg, name);
    validateTarget(qureg, targetQubit, name);
    validateOneQubitDephaseProb(prob, name);
    densmatr_oneQubitDephase(qureg, targetQubit, 2 * prob);
}

static void oneQubitDepolarise(Qureg qureg, int targetQubit, qreal prob, const char* name) {
    validateDensityMatrQureg(qureg, name);
    validateTarget(qureg, targetQubit, name);
    validateOneQubitDepolProb(prob, name);
    densmatr_oneQubitDepolarise(qureg, targetQubit, (4 * prob) / 3.0);
}

static void oneQubitDamping(Qureg qureg, int targetQubit, qreal prob, const char* name) {
    validateDensityMatrQureg(qureg, name);
    validateTarget(qureg, targetQubit, name);
    validateOneQubitDampingProb(prob, name);
    densmatr_oneQubitDamping(qureg, targetQubit, prob);
}

void applyOneQubitDephaseError(Qureg qureg, const int targetQubit, qreal prob) { oneQubitDephase(qureg, targetQubit, prob, __func__); }
void applyOneQubitDepolariseError(Qureg qureg, const int targetQubit, qreal prob) { oneQubitDepolarise(qureg, targetQubit, prob, __func__); }
void applyOneQubitDampingError(Qureg qureg, const int targetQubit, qreal prob) { oneQubitDamping(qureg, targetQubit, prob, __func__); }
void mixDephasing(Qureg qureg, const int targetQubit, qreal prob) { oneQubitDephase(qureg, targetQubit, prob, __func__); }
void mixDepolarising(Qureg qureg, const int targetQubit, qreal prob) { oneQubitDepolarise(qureg, targetQubit, prob, __func__); }
void mixDamping(Qureg qureg, const int targetQubit, qreal prob) { oneQubitDamping(qureg, targetQubit, prob, __func__); }

void applyTwoQubitDephaseError(Qureg qureg, int qubit1, int qubit2, qreal prob) {
    validateDensityMatrQureg(qureg, __func__);
    validateUniqueTargets(qureg, qubit1, qubit2, __func__);
    validateTwoQubitDephaseProb(prob, __func__);
    ensureIndsIncrease(&qubit1, &qubit2);
    densmatr_twoQubitDephase(qureg, qubit1, qubit2, (4 * prob) / 3.0);
}

void applyTwoQubitDepolariseError(Qureg qureg, int qubit1, int qubit2, qreal prob) {
    validateDensityMatrQureg(qureg, __func__);
    validateUniqueTargets(qureg, qubit1, qubit2, __func__);
    validateTwoQubitDepolProb(prob, __func__);
    ensureIndsIncrease(&qubit1, &qubit2);
    densmatr_twoQubitDepolarise(qureg, qubit1, qubit2, (16 * prob) / 15.0);
}

// ---- reporting and debug -------------------------------------------------------------------------- (QuEST_common.c:166-193, QuEST.c:699-726)

// state_rank_<chunkId>.csv, "real, imag" header on rank 0, %.12f per component.  The reference reads its host
// arrays; here the shard is streamed out of HBM in bounded slabs.
void reportState(Qureg qureg) {
    char filename[100];
    std::snprintf(filename, sizeof filename, "state_rank_%d.csv", qureg.chunkId);
    FILE* f = std::fopen(filename, "w");
    if (!f) return;
    if (qureg.chunkId == 0) std::fprintf(f, "real, imag\n");
    const long long slab = 1LL << 20;
    std::vector<qreal> re(slab), im(slab);
    for (long long at = 0; at < qureg.numAmpsPerChunk; at += slab) {
        long long n = qureg.numAmpsPerChunk - at < slab ? qureg.numAmpsPerChunk - at : slab;
        qb200_copyAmpsToHost(qureg, at, n, re.data(), im.data());
        for (long long i = 0; i < n; ++i) std::fprintf(f, "%.12f, %.12f\n", (double)re[i], (double)im[i]);
    }
    std::fclose(f);
}

void reportStateToScreen(Qureg qureg, QuESTEnv env, int reportRank) {
    if (qureg.numQubitsInStateVec > 5) {
        std::printf("Error: reportStateToScreen will not print output for systems of more than 5 qubits.\n");
        return;
    }
    std::vector<qreal> re(qureg.numAmpsPerChunk), im(qureg.numAmpsPerChunk);
    // Every rank reads its shard back BEFORE the rank-by-rank printing: on a sharded register the copy may first have
    // to restore the identity qubit layout, which is a collective exchange with partner ranks - it must not sit inside
    // the one-rank-at-a-time loop below, whose host barrier the partner would already be waiting in.
    qb200_copyAmpsToHost(qureg, 0, qureg.numAmpsPerChunk, re.data(), im.data());
    for (int rank = 0; rank < qureg.numChunks; ++rank) {
        if (qureg.chunkId == rank) {
            if (reportRank) {
                std::printf("Reporting state from rank %d [\n", qureg.chunkId);
                std::printf("real, imag\n");
            } else if (rank == 0) {
                std::printf("Reporting state [\n");
                std::printf("real, imag\n");
            }
            for (long long i = 0; i < qureg.numAmpsPerChunk; ++i)
                std::printf(REAL_STRING_FORMAT ", " REAL_STRING_FORMAT "\n", re[i], im[i]);
            if (reportRank || rank == qureg.numChunks - 1) std::printf("]\n");
            std::fflush(stdout);
        }
        syncQuESTEnv(env);
    }
}

void reportQuregParams(Qureg qureg) {
    long long int numAmps = 1LL << qureg.numQubitsInStateVec;
    if (qureg.chunkId == 0) {
        std::printf("QUBITS:\n");
        std::printf("Number of qubits is %d.\n", qureg.numQubitsInStateVec);
        std::printf("Number of amps is %lld.\n", numAmps);
        std::printf("Number of amps per rank is %lld.\n", numAmps / qureg.numChunks);
    }
}

int compareStates(Qureg qureg1, Qureg qureg2, qreal precision) {
    validateMatchingQuregDims(qureg1, qureg2, __func__);
    return statevec_compareStates(qureg1, qureg2, precision);
}

void initStateDebug(Qureg qureg) { statevec_initStateDebug(qureg); }

// Every rank parses the whole file and keeps the lines that fall in its shard (QuEST_cpu.c:1507-1555); lines
// starting with '#' are skipped, the first line that does not scan as "re, im" leaves that amplitude untouched.
void initStateFromSingleFile(Qureg* qureg, char filename[200], QuESTEnv env) {
    (void)env;
    FILE* fp = std::fopen(filename, "r");
    validateFileOpened(fp != NULL, __func__);
    long long chunk = qureg->numAmpsPerChunk, total = chunk * qureg->numChunks;
    std::vector<qreal> re(chunk), im(chunk);
    qb200_copyAmpsToHost(*qureg, 0, chunk, re.data(), im.data());
    char line[200];
    long long global = 0, local = 0;
    while (std::fgets(line, sizeof line, fp) != NULL && global < total) {
        if (line[0] == '#') continue;
        if (global / chunk == qureg->chunkId) {
            double a = re[local], b = im[local];
            std::sscanf(line, "%lf, %lf", &a, &b);
            re[local] = (qreal)a;
            im[local] = (qreal)b;
            ++local;
        }
        ++global;
    }
    std::fclose(fp);
    statevec_setAmps(*qureg, (long long)qureg->chunkId * chunk, re.data(), im.data(), chunk);
}

void initStateOfSingleQubit(Qureg* qureg, int qubitId, int outcome) {
    validateStateVecQureg(*qureg, __func__);
    validateTarget(*qureg, qubitId, __func__);
    validateOutcome(outcome, __func__);
    statevec_initStateOfSingleQubit(qureg, qubitId, outcome);
}

int getQuEST_PREC(void) { return sizeof(qreal) / 4; }
int QuESTPrecision(void) { return QuEST_PREC; }

}  // extern "C"
