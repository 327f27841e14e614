/* quest-b200: the QuEST C API, served by a B200 (sm_100a) gate-application engine.
 *
 * This header declares exactly the types and entry points of the reference's public header
 * (QuEST/include/QuEST.h) so that existing C programs and the reference's ctypes harness relink
 * against quest-b200's libQuEST.so unchanged.  Struct layouts are ABI-identical (x86-64):
 * sizeof(Qureg)=112, QuESTEnv=8, Complex=2*sizeof(qreal), ComplexMatrix2=8*sizeof(qreal),
 * Vector=3*sizeof(qreal).  Each declaration cites the reference line it replaces as [ref:NNN].
 *
 * Semantics that differ from the reference's CPU build only in where the state lives:
 *   - amplitudes are held in B200 HBM as two SoA arrays (real[], imag[]) of qreal;
 *     Qureg.stateVec.{real,imag} are NULL (no host mirror is ever allocated);
 *   - gate calls return once the kernel is enqueued; every value-returning call and
 *     syncQuESTEnv() synchronise;
 *   - invalid input prints the reference's error text and exit()s with the reference's code.
 */
#ifndef QUEST_H
#define QUEST_H

#include "QuEST_precision.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- private-by-convention structures (field order is ABI) -------------------- [ref:18-45] */

enum phaseGateType { SIGMA_Z = 0, S_GATE = 1, T_GATE = 2 };

typedef struct {
    char* buffer;   /* NUL-terminated QASM text            */
    int bufferSize; /* capacity in chars                    */
    int bufferFill; /* chars used                           */
    int isLogging;  /* gates are appended only when nonzero */
} QASMLogger;

typedef struct ComplexArray {
    qreal* real;
    qreal* imag;
} ComplexArray;

/* ---- public structures ------------------------------------------------------- [ref:57-121] */

typedef struct Complex {
    qreal real;
    qreal imag;
} Complex;

typedef struct ComplexMatrix2 {
    Complex r0c0, r0c1;
    Complex r1c0, r1c1;
} ComplexMatrix2;

typedef struct Vector {
    qreal x, y, z;
} Vector;

/* A register of qubits: a state-vector of 2^n amplitudes, or a density matrix stored as a
 * 2n-qubit state-vector (column-major: element (r,c) at r + c*2^n).
 * quest-b200 use of the pointer fields:
 *   stateVec            NULL (never mirrored on the host)
 *   deviceStateVec      this rank's shard in HBM: real[], imag[] (numAmpsPerChunk each)
 *   pairStateVec.real   opaque per-register engine record (peer mappings, scratch)
 *   firstLevelReduction device scratch for block partial sums
 *   qasmLog             valid QASMLogger*, as in the reference                                   */
typedef struct Qureg {
    int isDensityMatrix;
    int numQubitsRepresented;
    int numQubitsInStateVec;
    long long int numAmpsPerChunk;
    long long int numAmpsTotal;
    int chunkId;
    int numChunks;
    ComplexArray stateVec;
    ComplexArray pairStateVec;
    ComplexArray deviceStateVec;
    qreal *firstLevelReduction, *secondLevelReduction;
    QASMLogger* qasmLog;
} Qureg;

/* Execution environment: rank = this process's shard index, numRanks = number of GPUs the
 * state is sharded over (one process per GPU; 1 unless launched with WORLD_SIZE>1). */
typedef struct QuESTEnv {
    int rank;
    int numRanks;
} QuESTEnv;

/* ---- registers --------------------------------------------------------------------------- */
Qureg createQureg(int numQubits, QuESTEnv env);                         /* [ref:139] */
Qureg createDensityQureg(int numQubits, QuESTEnv env);                  /* [ref:151] */
void destroyQureg(Qureg qureg, QuESTEnv env);                           /* [ref:160] */
void reportState(Qureg qureg);                                          /* [ref:183] */
void reportStateToScreen(Qureg qureg, QuESTEnv env, int reportRank);    /* [ref:189] */
void reportQuregParams(Qureg qureg);                                    /* [ref:195] */
int getNumQubits(Qureg qureg);                                          /* [ref:199] */
int getNumAmps(Qureg qureg);      /* int, as in the reference: overflows at >=31 qubits [ref:203] */

/* ---- initialisation ---------------------------------------------------------------------- */
void initZeroState(Qureg qureg);                                        /* [ref:211] */
void initPlusState(Qureg qureg);                                        /* [ref:222] */
void initClassicalState(Qureg qureg, long long int stateInd);           /* [ref:232] */
void initPureState(Qureg qureg, Qureg pure);                            /* [ref:241] */
void initStateFromAmps(Qureg qureg, qreal* reals, qreal* imags);        /* [ref:254] */
void setAmps(Qureg qureg, long long int startInd, qreal* reals, qreal* imags,
             long long int numAmps);                                    /* [ref:269] */
void cloneQureg(Qureg targetQureg, Qureg copyQureg);                    /* [ref:279] */

/* ---- diagonal (phase) gates: never communicate between shards ------------------------------ */
void phaseShift(Qureg qureg, const int targetQubit, qreal angle);                               /* [ref:311] */
void controlledPhaseShift(Qureg qureg, const int idQubit1, const int idQubit2, qreal angle);    /* [ref:352] */
void multiControlledPhaseShift(Qureg qureg, int* controlQubits, int numControlQubits,
                               qreal angle);                                                    /* [ref:392] */
void controlledPhaseFlip(Qureg qureg, const int idQubit1, const int idQubit2);                  /* [ref:428] */
void multiControlledPhaseFlip(Qureg qureg, int* controlQubits, int numControlQubits);           /* [ref:472] */
void sGate(Qureg qureg, const int targetQubit);                                                 /* [ref:501] */
void tGate(Qureg qureg, const int targetQubit);                                                 /* [ref:530] */

/* ---- environment -------------------------------------------------------------------------- */
QuESTEnv createQuESTEnv(void);                                          /* [ref:540] */
void destroyQuESTEnv(QuESTEnv env);                                     /* [ref:548] */
void syncQuESTEnv(QuESTEnv env);                                        /* [ref:554] */
int syncQuESTSuccess(int successCode);                                  /* [ref:562] */
void reportQuESTEnv(QuESTEnv env);                                      /* [ref:568] */
void getEnvironmentString(QuESTEnv env, Qureg qureg, char str[200]);    /* [ref:570] */

/* ---- amplitude access (each is a device->host scalar read; synchronises) ------------------- */
Complex getAmp(Qureg qureg, long long int index);                       /* [ref:580] */
qreal getRealAmp(Qureg qureg, long long int index);                     /* [ref:591] */
qreal getImagAmp(Qureg qureg, long long int index);                     /* [ref:602] */
qreal getProbAmp(Qureg qureg, long long int index);                     /* [ref:612] */
Complex getDensityAmp(Qureg qureg, long long int row, long long int col); /* [ref:623] */
qreal calcTotalProb(Qureg qureg);                                       /* [ref:634] */

/* ---- dense one-target gates ---------------------------------------------------------------- */
void compactUnitary(Qureg qureg, const int targetQubit, Complex alpha, Complex beta);           /* [ref:670] */
void unitary(Qureg qureg, const int targetQubit, ComplexMatrix2 u);                             /* [ref:696] */
void rotateX(Qureg qureg, const int rotQubit, qreal angle);                                     /* [ref:727] */
void rotateY(Qureg qureg, const int rotQubit, qreal angle);                                     /* [ref:758] */
void rotateZ(Qureg qureg, const int rotQubit, qreal angle);                                     /* [ref:789] */
void rotateAroundAxis(Qureg qureg, const int rotQubit, qreal angle, Vector axis);               /* [ref:805] */
void controlledRotateX(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle); /* [ref:837] */
void controlledRotateY(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle); /* [ref:868] */
void controlledRotateZ(Qureg qureg, const int controlQubit, const int targetQubit, qreal angle); /* [ref:899] */
void controlledRotateAroundAxis(Qureg qureg, const int controlQubit, const int targetQubit,
                                qreal angle, Vector axis);                                      /* [ref:935] */
void controlledCompactUnitary(Qureg qureg, const int controlQubit, const int targetQubit,
                              Complex alpha, Complex beta);                                     /* [ref:979] */
void controlledUnitary(Qureg qureg, const int controlQubit, const int targetQubit,
                       ComplexMatrix2 u);                                                       /* [ref:1021] */
void multiControlledUnitary(Qureg qureg, int* controlQubits, const int numControlQubits,
                            const int targetQubit, ComplexMatrix2 u);                           /* [ref:1077] */
void pauliX(Qureg qureg, const int targetQubit);                                                /* [ref:1106] */
void pauliY(Qureg qureg, const int targetQubit);                                                /* [ref:1136] */
void pauliZ(Qureg qureg, const int targetQubit);                                                /* [ref:1166] */
void hadamard(Qureg qureg, const int targetQubit);                                              /* [ref:1198] */
void controlledNot(Qureg qureg, const int controlQubit, const int targetQubit);                 /* [ref:1237] */
void controlledPauliY(Qureg qureg, const int controlQubit, const int targetQubit);              /* [ref:1278] */

/* ---- probabilities, measurement ------------------------------------------------------------ */
qreal calcProbOfOutcome(Qureg qureg, const int measureQubit, int outcome);  /* P(1) is 1-P(0) [ref:1291] */
qreal collapseToOutcome(Qureg qureg, const int measureQubit, int outcome);  /* [ref:1309] */
int measure(Qureg qureg, int measureQubit);                                 /* [ref:1321] */
int measureWithStats(Qureg qureg, int measureQubit, qreal* outcomeProb);    /* [ref:1335] */
Complex calcInnerProduct(Qureg bra, Qureg ket);                             /* [ref:1338] */

/* ---- RNG seeding (host-side mt19937ar, process-global state) ------------------------------- */
void seedQuESTDefault(void);                                                /* [ref:1349] */
void seedQuEST(unsigned long int* seedArray, int numSeeds);                 /* [ref:1364] */

/* ---- QASM log ------------------------------------------------------------------------------ */
void startRecordingQASM(Qureg qureg);                                       /* [ref:1369] */
void stopRecordingQASM(Qureg qureg);                                        /* [ref:1375] */
void clearRecordedQASM(Qureg qureg);                                        /* [ref:1379] */
void printRecordedQASM(Qureg qureg);                                        /* [ref:1382] */
void writeRecordedQASMToFile(Qureg qureg, char* filename);                  /* [ref:1385] */

/* ---- density-matrix noise and maths -------------------------------------------------------- */
void applyOneQubitDephaseError(Qureg qureg, const int targetQubit, qreal prob);     /* [ref:1405] */
void applyTwoQubitDephaseError(Qureg qureg, const int qubit1, const int qubit2, qreal prob); /* [ref:1431] */
void applyOneQubitDepolariseError(Qureg qureg, const int targetQubit, qreal prob);  /* [ref:1455] */
void applyOneQubitDampingError(Qureg qureg, const int targetQubit, qreal prob);     /* [ref:1477] */
void applyTwoQubitDepolariseError(Qureg qureg, const int qubit1, const int qubit2, qreal prob); /* [ref:1526] */
void addDensityMatrix(Qureg combineQureg, qreal prob, Qureg otherQureg);            /* [ref:1539] */
qreal calcPurity(Qureg qureg);                                                      /* [ref:1556] */
qreal calcFidelity(Qureg qureg, Qureg pureState);                                   /* [ref:1571] */

/* QuEST v3 spellings named by this project's target API list; each is the same operation as the
 * applyOneQubit*Error call above it (same validation, same arithmetic). */
void mixDephasing(Qureg qureg, const int targetQubit, qreal prob);
void mixDepolarising(Qureg qureg, const int targetQubit, qreal prob);
void mixDamping(Qureg qureg, const int targetQubit, qreal prob);

/* sizeof(qreal)/4: lets a binding discover the precision of a loaded library [ref: QuEST.c:724] */
int getQuEST_PREC(void);

#ifdef __cplusplus
}
#endif

#endif /* QUEST_H */
