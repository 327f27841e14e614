/* quest-b200 internal seam between the host front-end (quest_api.cpp, quest_common.cpp, ...) and a
 * backend that owns the amplitudes.
 *
 * The names and argument meanings follow the reference's backend contract
 * (QuEST/src/QuEST_internal.h:41-173) so that the same front-end can sit on
 *   - the product backend  : quest_b200/csrc/backend_cuda.cu (sm_100a kernels, HBM-resident state), or
 *   - the test-only oracle : oracle/quest_oracle.c (CPU restatement of QuEST_cpu.c / QuEST_cpu_local.c).
 * A backend defines every function in the "backend" sections; the front-end defines the rest.
 */
#ifndef QUEST_B200_INTERNAL_H
#define QUEST_B200_INTERNAL_H

#include "QuEST.h"
#include "QuEST_debug.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- host helpers defined by the front-end (quest_common.cpp) ---------------------------------- */
unsigned long int hashString(char* str);
qreal getVectorMagnitude(Vector vec);
Complex getConjugateScalar(Complex scalar);
ComplexMatrix2 getConjugateMatrix(ComplexMatrix2 matr);
void ensureIndsIncrease(int* ind1, int* ind2);
void getComplexPairFromRotation(qreal angle, Vector axis, Complex* alpha, Complex* beta);
void getZYZRotAnglesFromComplexPair(Complex alpha, Complex beta, qreal* rz2, qreal* ry, qreal* rz1);
void getComplexPairAndPhaseFromUnitary(ComplexMatrix2 u, Complex* alpha, Complex* beta, qreal* globalPhase);
void shiftIndices(int* indices, int numIndices, int shift);
void getQuESTDefaultSeedKey(unsigned long int* key);
int generateMeasurementOutcome(qreal zeroProb, qreal* outcomeProb);

/* compositions over the backend (quest_common.cpp; reference QuEST_common.c:160-326) */
qreal statevec_getProbAmp(Qureg qureg, long long int index);
void statevec_phaseShift(Qureg qureg, const int targetQubit, qreal angle);
void statevec_pauliZ(Qureg qureg, const int targetQubit);
void statevec_sGate(Qureg qureg, const int targetQubit);
void statevec_tGate(Qureg qureg, const int targetQubit);
void statevec_sGateConj(Qureg qureg, const int targetQubit);
void statevec_tGateConj(Qureg qureg, const int targetQubit);
void statevec_rotateX(Qureg qureg, const int rotQubit, qreal angle);
void statevec_rotateY(Qureg qureg, const int rotQubit, qreal angle);
void statevec_rotateZ(Qureg qureg, const int rotQubit, qreal angle);
void statevec_rotateAroundAxis(Qureg qureg, const int rotQubit, qreal angle, Vector axis);
void statevec_rotateAroundAxisConj(Qureg qureg, const int rotQubit, qreal angle, Vector axis);
void statevec_controlledRotateX(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle);
void statevec_controlledRotateY(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle);
void statevec_controlledRotateZ(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle);
void statevec_controlledRotateAroundAxis(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle, Vector axis);
void statevec_controlledRotateAroundAxisConj(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle, Vector axis);
int statevec_measureWithStats(Qureg qureg, int measureQubit, qreal* outcomeProb);
int densmatr_measureWithStats(Qureg qureg, int measureQubit, qreal* outcomeProb);
qreal statevec_calcFidelity(Qureg qureg, Qureg pureState);

/* ---- backend: registers and amplitude transfer -------------------------------------------------- */
void statevec_createQureg(Qureg* qureg, int numQubits, QuESTEnv env);
void statevec_destroyQureg(Qureg qureg, QuESTEnv env);
void statevec_initZeroState(Qureg qureg);
void statevec_initPlusState(Qureg qureg);
void statevec_initStateDebug(Qureg qureg);
void statevec_initClassicalState(Qureg qureg, long long int stateInd);
void statevec_initStateOfSingleQubit(Qureg* qureg, int qubitId, int outcome);
void statevec_setAmps(Qureg qureg, long long int startInd, qreal* reals, qreal* imags, long long int numAmps);
void statevec_cloneQureg(Qureg targetQureg, Qureg copyQureg);
int statevec_compareStates(Qureg mq1, Qureg mq2, qreal precision);
qreal statevec_getRealAmp(Qureg qureg, long long int index);
qreal statevec_getImagAmp(Qureg qureg, long long int index);
/* quest-b200 addition: copy n amplitudes of THIS rank's shard, starting at local index `start`, to host
 * arrays.  Replaces the reference's direct reads of qureg.stateVec (QuEST_common.c:174-177,
 * QuEST_cpu.c:1266-1269), which has no host mirror here. */
void qb200_copyAmpsToHost(Qureg qureg, long long int start, long long int n, qreal* reals, qreal* imags);

void densmatr_initPlusState(Qureg targetQureg);
void densmatr_initClassicalState(Qureg qureg, long long int stateInd);
void densmatr_initPureState(Qureg targetQureg, Qureg copyQureg);

/* ---- backend: gates ----------------------------------------------------------------------------- */
void statevec_compactUnitary(Qureg qureg, const int targetQubit, Complex alpha, Complex beta);
void statevec_unitary(Qureg qureg, const int targetQubit, ComplexMatrix2 u);
void statevec_controlledCompactUnitary(Qureg qureg, const int controlQubit, const int targetQubit, Complex alpha, Complex beta);
void statevec_controlledUnitary(Qureg qureg, const int controlQubit, const int targetQubit, ComplexMatrix2 u);
void statevec_multiControlledUnitary(Qureg qureg, int* controlQubits, const int numControlQubits, const int targetQubit, ComplexMatrix2 u);
void statevec_hadamard(Qureg qureg, const int targetQubit);
void statevec_pauliX(Qureg qureg, const int targetQubit);
void statevec_pauliY(Qureg qureg, const int targetQubit);
void statevec_pauliYConj(Qureg qureg, const int targetQubit);
void statevec_controlledNot(Qureg qureg, const int controlQubit, const int targetQubit);
void statevec_controlledPauliY(Qureg qureg, const int controlQubit, const int targetQubit);
void statevec_controlledPauliYConj(Qureg qureg, const int controlQubit, const int targetQubit);
void statevec_phaseShiftByTerm(Qureg qureg, const int targetQubit, Complex term);
void statevec_controlledPhaseShift(Qureg qureg, const int idQubit1, const int idQubit2, qreal angle);
void statevec_multiControlledPhaseShift(Qureg qureg, int* controlQubits, int numControlQubits, qreal angle);
void statevec_controlledPhaseFlip(Qureg qureg, const int idQubit1, const int idQubit2);
void statevec_multiControlledPhaseFlip(Qureg qureg, int* controlQubits, int numControlQubits);

/* ---- backend: reductions, collapse -------------------------------------------------------------- */
qreal statevec_calcTotalProb(Qureg qureg);
qreal statevec_calcProbOfOutcome(Qureg qureg, const int measureQubit, int outcome);
Complex statevec_calcInnerProduct(Qureg bra, Qureg ket);
void statevec_collapseToKnownProbOutcome(Qureg qureg, const int measureQubit, int outcome, qreal outcomeProb);
qreal densmatr_calcTotalProb(Qureg qureg);
qreal densmatr_calcPurity(Qureg qureg);
qreal densmatr_calcFidelity(Qureg qureg, Qureg pureState);
qreal densmatr_calcProbOfOutcome(Qureg qureg, const int measureQubit, int outcome);
void densmatr_collapseToKnownProbOutcome(Qureg qureg, const int measureQubit, int outcome, qreal outcomeProb);

/* ---- backend: density-matrix noise -------------------------------------------------------------- */
void densmatr_oneQubitDephase(Qureg qureg, const int targetQubit, qreal dephase);
void densmatr_twoQubitDephase(Qureg qureg, const int qubit1, const int qubit2, qreal dephase);
void densmatr_oneQubitDepolarise(Qureg qureg, const int targetQubit, qreal depolLevel);
void densmatr_oneQubitDamping(Qureg qureg, const int targetQubit, qreal damping);
void densmatr_twoQubitDepolarise(Qureg qureg, int qubit1, int qubit2, qreal depolLevel);
void densmatr_addDensityMatrix(Qureg combineQureg, qreal otherProb, Qureg otherQureg);

/* ---- front-end: QASM log (quest_qasm.cpp) -------------------------------------------------------- */
typedef enum {
    GATE_SIGMA_X, GATE_SIGMA_Y, GATE_SIGMA_Z, GATE_T, GATE_S, GATE_HADAMARD,
    GATE_ROTATE_X, GATE_ROTATE_Y, GATE_ROTATE_Z, GATE_ROTATE_AROUND_AXIS, GATE_UNITARY, GATE_PHASE_SHIFT
} TargetGate;

void qasm_setup(Qureg* qureg);
void qasm_free(Qureg qureg);
void qasm_startRecording(Qureg qureg);
void qasm_stopRecording(Qureg qureg);
void qasm_clearRecorded(Qureg qureg);
void qasm_printRecorded(Qureg qureg);
int qasm_writeRecordedToFile(Qureg qureg, char* filename);
void qasm_recordComment(Qureg qureg, const char* comment);
void qasm_recordGate(Qureg qureg, TargetGate gate, int targetQubit);
void qasm_recordParamGate(Qureg qureg, TargetGate gate, int targetQubit, qreal param);
void qasm_recordCompactUnitary(Qureg qureg, Complex alpha, Complex beta, int targetQubit);
void qasm_recordUnitary(Qureg qureg, ComplexMatrix2 u, int targetQubit);
void qasm_recordAxisRotation(Qureg qureg, qreal angle, Vector axis, const int targetQubit);
void qasm_recordControlledGate(Qureg qureg, TargetGate gate, int controlQubit, int targetQubit);
void qasm_recordControlledParamGate(Qureg qureg, TargetGate gate, int controlQubit, int targetQubit, qreal param);
void qasm_recordControlledCompactUnitary(Qureg qureg, Complex alpha, Complex beta, int controlQubit, int targetQubit);
void qasm_recordControlledUnitary(Qureg qureg, ComplexMatrix2 u, int controlQubit, int targetQubit);
void qasm_recordControlledAxisRotation(Qureg qureg, qreal angle, Vector axis, int controlQubit, int targetQubit);
void qasm_recordMultiControlledGate(Qureg qureg, TargetGate gate, int* controlQubits, const int numControlQubits, const int targetQubit);
void qasm_recordMultiControlledParamGate(Qureg qureg, TargetGate gate, int* controlQubits, const int numControlQubits, const int targetQubit, qreal param);
void qasm_recordMultiControlledUnitary(Qureg qureg, ComplexMatrix2 u, int* controlQubits, const int numControlQubits, const int targetQubit);
void qasm_recordMeasurement(Qureg qureg, const int measureQubit);
void qasm_recordInitZero(Qureg qureg);
void qasm_recordInitPlus(Qureg qureg);
void qasm_recordInitClassical(Qureg qureg, long long int stateInd);

/* ---- front-end: validation (quest_validation.cpp; reference QuEST_validation.c) ------------------ */
void validateCreateNumQubits(int numQubits, const char* caller);
void validateStateIndex(Qureg qureg, long long int stateInd, const char* caller);
void validateNumAmps(Qureg qureg, long long int startInd, long long int numAmps, const char* caller);
void validateTarget(Qureg qureg, int targetQubit, const char* caller);
void validateControlTarget(Qureg qureg, int controlQubit, int targetQubit, const char* caller);
void validateUniqueTargets(Qureg qureg, int qubit1, int qubit2, const char* caller);
void validateMultiControls(Qureg qureg, int* controlQubits, const int numControlQubits, const char* caller);
void validateMultiControlsTarget(Qureg qureg, int* controlQubits, const int numControlQubits, const int targetQubit, const char* caller);
void validateUnitaryMatrix(ComplexMatrix2 u, const char* caller);
void validateUnitaryComplexPair(Complex alpha, Complex beta, const char* caller);
void validateVector(Vector vector, const char* caller);
void validateStateVecQureg(Qureg qureg, const char* caller);
void validateDensityMatrQureg(Qureg qureg, const char* caller);
void validateOutcome(int outcome, const char* caller);
void validateMeasurementProb(qreal prob, const char* caller);
void validateMatchingQuregDims(Qureg qureg1, Qureg qureg2, const char* caller);
void validateMatchingQuregTypes(Qureg qureg1, Qureg qureg2, const char* caller);
void validateSecondQuregStateVec(Qureg qureg2, const char* caller);
void validateFileOpened(int found, const char* caller);
void validateProb(qreal prob, const char* caller);
void validateOneQubitDephaseProb(qreal prob, const char* caller);
void validateTwoQubitDephaseProb(qreal prob, const char* caller);
void validateOneQubitDepolProb(qreal prob, const char* caller);
void validateTwoQubitDepolProb(qreal prob, const char* caller);
void validateOneQubitDampingProb(qreal prob, const char* caller);

/* ---- front-end: RNG (mt19937.cpp; same exported names as reference mt19937ar.h) ------------------ */
void init_genrand(unsigned long s);
void init_by_array(unsigned long init_key[], int key_length);
unsigned long genrand_int32(void);
long genrand_int31(void);
double genrand_real1(void);
double genrand_real2(void);
double genrand_real3(void);
double genrand_res53(void);

#ifdef __cplusplus
}
#endif

#endif /* QUEST_B200_INTERNAL_H */
