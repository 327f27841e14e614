// quest-b200: OpenQASM 2.0 log attached to each register.
//
// Pure host string code, outside the accelerated path; kept because createQureg must hand back a valid
// Qureg.qasmLog (reference QuEST.c:37) and the public API exposes start/stop/clear/print/write.
// The text produced matches the reference logger line for line (QuEST/src/QuEST_qasm.c): header
// "OPENQASM 2.0;\nqreg q[n];\ncreg c[n];\n", gate mnemonics x y z t s h Rx Ry Rz U, one 'c' prefix per control,
// parameters printed with REAL_QASM_FORMAT, and the explanatory comments / Rz phase fixes the reference adds
// after controlled phase gates and controlled unitaries.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "quest_internal.h"

namespace {

constexpr int kInitialCapacity = 1000;
constexpr int kMaxLine = 200;

const char* mnemonic(TargetGate g) {
    switch (g) {
        case GATE_SIGMA_X: return "x";
        case GATE_SIGMA_Y: return "y";
        case GATE_SIGMA_Z: return "z";
        case GATE_T: return "t";
        case GATE_S: return "s";
        case GATE_HADAMARD: return "h";
        case GATE_ROTATE_X: return "Rx";
        case GATE_ROTATE_Y: return "Ry";
        case GATE_ROTATE_Z: return "Rz";
        case GATE_UNITARY: return "U";
        case GATE_PHASE_SHIFT: return "Rz";  // phase fixed by a following Rz when controlled
        default: return "?";
    }
}

[[noreturn]] void logFailure() {
    std::printf("!!!\nINTERNAL ERROR: QASM line buffer filled!\n!!!");
    std::exit(1);
}

bool logging(const Qureg& q) { return q.qasmLog && q.qasmLog->isLogging; }

void appendText(Qureg q, const char* text, int len) {
    QASMLogger* log = q.qasmLog;
    int need = log->bufferFill + len + 1;
    if (need > log->bufferSize) {
        int cap = log->bufferSize;
        while (cap < need) cap *= 2;
        char* grown = static_cast<char*>(std::realloc(log->buffer, cap));
        if (!grown) logFailure();
        log->buffer = grown;
        log->bufferSize = cap;
    }
    std::memcpy(log->buffer + log->bufferFill, text, len);
    log->bufferFill += len;
    log->buffer[log->bufferFill] = '\0';
}

void appendf(Qureg q, const char* fmt, ...) {
    char line[kMaxLine + 1];
    va_list ap;
    va_start(ap, fmt);
    int len = std::vsnprintf(line, sizeof line, fmt, ap);
    va_end(ap);
    if (len < 0 || len >= kMaxLine) logFailure();
    appendText(q, line, len);
}

// "<c..c><gate>[(p0,p1,..)] q[c0],..,q[t];\n"
void appendGate(Qureg q, TargetGate gate, const int* controls, int numControls, int target, const qreal* params,
                int numParams) {
    char line[kMaxLine + 1];
    int len = 0;
    auto put = [&](const char* fmt, auto... args) {
        if (len < kMaxLine) len += std::snprintf(line + len, kMaxLine - len, fmt, args...);
    };
    for (int i = 0; i < numControls; ++i) put("%s", "c");
    put("%s", mnemonic(gate));
    if (numParams > 0) {
        put("%s", "(");
        for (int i = 0; i < numParams; ++i) {
            put(REAL_QASM_FORMAT, params[i]);
            if (i != numParams - 1) put("%s", ",");
        }
        put("%s", ")");
    }
    put("%s", " ");
    for (int i = 0; i < numControls; ++i) put("q[%d],", controls[i]);
    put("q[%d];\n", target);
    if (len >= kMaxLine) logFailure();
    appendText(q, line, len);
}

void zyzParams(Complex alpha, Complex beta, qreal out[3]) {
    getZYZRotAnglesFromComplexPair(alpha, beta, &out[0], &out[1], &out[2]);
}

}  // namespace

extern "C" {

void qasm_setup(Qureg* qureg) {
    QASMLogger* log = static_cast<QASMLogger*>(std::malloc(sizeof(QASMLogger)));
    if (!log) logFailure();
    log->buffer = static_cast<char*>(std::malloc(kInitialCapacity));
    if (!log->buffer) logFailure();
    log->bufferSize = kInitialCapacity;
    log->isLogging = 0;
    log->bufferFill = std::snprintf(log->buffer, log->bufferSize, "OPENQASM 2.0;\nqreg q[%d];\ncreg c[%d];\n",
                                    qureg->numQubitsRepresented, qureg->numQubitsRepresented);
    qureg->qasmLog = log;
}

void qasm_free(Qureg qureg) {
    if (!qureg.qasmLog) return;
    std::free(qureg.qasmLog->buffer);
    std::free(qureg.qasmLog);
}

void qasm_startRecording(Qureg qureg) { qureg.qasmLog->isLogging = 1; }
void qasm_stopRecording(Qureg qureg) { qureg.qasmLog->isLogging = 0; }

void qasm_clearRecorded(Qureg qureg) {
    qureg.qasmLog->buffer[0] = '\0';
    qureg.qasmLog->bufferFill = 0;
}

void qasm_printRecorded(Qureg qureg) { std::printf("%s", qureg.qasmLog->buffer); }

int qasm_writeRecordedToFile(Qureg qureg, char* filename) {
    FILE* f = std::fopen(filename, "w");
    if (!f) return 0;
    std::fprintf(f, "%s", qureg.qasmLog->buffer);
    std::fclose(f);
    return 1;
}

void qasm_recordComment(Qureg qureg, const char* comment) {
    if (logging(qureg)) appendf(qureg, "// %s\n", comment);
}

void qasm_recordGate(Qureg qureg, TargetGate gate, int targetQubit) {
    if (logging(qureg)) appendGate(qureg, gate, nullptr, 0, targetQubit, nullptr, 0);
}

void qasm_recordParamGate(Qureg qureg, TargetGate gate, int targetQubit, qreal param) {
    if (logging(qureg)) appendGate(qureg, gate, nullptr, 0, targetQubit, &param, 1);
}

void qasm_recordCompactUnitary(Qureg qureg, Complex alpha, Complex beta, int targetQubit) {
    if (!logging(qureg)) return;
    qreal p[3];
    zyzParams(alpha, beta, p);
    appendGate(qureg, GATE_UNITARY, nullptr, 0, targetQubit, p, 3);
}

void qasm_recordUnitary(Qureg qureg, ComplexMatrix2 u, int targetQubit) {
    if (!logging(qureg)) return;
    Complex alpha, beta;
    qreal phase, p[3];
    getComplexPairAndPhaseFromUnitary(u, &alpha, &beta, &phase);
    zyzParams(alpha, beta, p);
    appendGate(qureg, GATE_UNITARY, nullptr, 0, targetQubit, p, 3);
}

void qasm_recordAxisRotation(Qureg qureg, qreal angle, Vector axis, const int targetQubit) {
    if (!logging(qureg)) return;
    Complex alpha, beta;
    qreal p[3];
    getComplexPairFromRotation(angle, axis, &alpha, &beta);
    zyzParams(alpha, beta, p);
    appendGate(qureg, GATE_UNITARY, nullptr, 0, targetQubit, p, 3);
}

void qasm_recordControlledGate(Qureg qureg, TargetGate gate, int controlQubit, int targetQubit) {
    if (logging(qureg)) appendGate(qureg, gate, &controlQubit, 1, targetQubit, nullptr, 0);
}

void qasm_recordControlledParamGate(Qureg qureg, TargetGate gate, int controlQubit, int targetQubit, qreal param) {
    if (!logging(qureg)) return;
    appendGate(qureg, gate, &controlQubit, 1, targetQubit, &param, 1);
    if (gate == GATE_PHASE_SHIFT) {
        qasm_recordComment(qureg, "Restoring the discarded global phase of the previous controlled phase gate");
        qreal fix = param / 2.0;
        appendGate(qureg, GATE_ROTATE_Z, nullptr, 0, targetQubit, &fix, 1);
    }
}

void qasm_recordControlledCompactUnitary(Qureg qureg, Complex alpha, Complex beta, int controlQubit, int targetQubit) {
    if (!logging(qureg)) return;
    qreal p[3];
    zyzParams(alpha, beta, p);
    appendGate(qureg, GATE_UNITARY, &controlQubit, 1, targetQubit, p, 3);
}

void qasm_recordControlledUnitary(Qureg qureg, ComplexMatrix2 u, int controlQubit, int targetQubit) {
    if (!logging(qureg)) return;
    Complex alpha, beta;
    qreal phase, p[3];
    getComplexPairAndPhaseFromUnitary(u, &alpha, &beta, &phase);
    zyzParams(alpha, beta, p);
    appendGate(qureg, GATE_UNITARY, &controlQubit, 1, targetQubit, p, 3);
    qasm_recordComment(qureg, "Restoring the discarded global phase of the previous controlled unitary");
    appendGate(qureg, GATE_ROTATE_Z, nullptr, 0, targetQubit, &phase, 1);
}

void qasm_recordControlledAxisRotation(Qureg qureg, qreal angle, Vector axis, int controlQubit, int targetQubit) {
    if (!logging(qureg)) return;
    Complex alpha, beta;
    qreal p[3];
    getComplexPairFromRotation(angle, axis, &alpha, &beta);
    zyzParams(alpha, beta, p);
    appendGate(qureg, GATE_UNITARY, &controlQubit, 1, targetQubit, p, 3);
}

void qasm_recordMultiControlledGate(Qureg qureg, TargetGate gate, int* controlQubits, const int numControlQubits,
                                    const int targetQubit) {
    if (logging(qureg)) appendGate(qureg, gate, controlQubits, numControlQubits, targetQubit, nullptr, 0);
}

void qasm_recordMultiControlledParamGate(Qureg qureg, TargetGate gate, int* controlQubits, const int numControlQubits,
                                         const int targetQubit, qreal param) {
    if (!logging(qureg)) return;
    appendGate(qureg, gate, controlQubits, numControlQubits, targetQubit, &param, 1);
    if (gate == GATE_PHASE_SHIFT) {
        qasm_recordComment(qureg, "Restoring the discarded global phase of the previous multicontrolled phase gate");
        qreal fix = param / 2.0;
        appendGate(qureg, GATE_ROTATE_Z, nullptr, 0, targetQubit, &fix, 1);
    }
}

void qasm_recordMultiControlledUnitary(Qureg qureg, ComplexMatrix2 u, int* controlQubits, const int numControlQubits,
                                       const int targetQubit) {
    if (!logging(qureg)) return;
    Complex alpha, beta;
    qreal phase, p[3];
    getComplexPairAndPhaseFromUnitary(u, &alpha, &beta, &phase);
    zyzParams(alpha, beta, p);
    appendGate(qureg, GATE_UNITARY, controlQubits, numControlQubits, targetQubit, p, 3);
    appendGate(qureg, GATE_ROTATE_Z, nullptr, 0, targetQubit, &phase, 1);
}

void qasm_recordMeasurement(Qureg qureg, const int measureQubit) {
    if (logging(qureg)) appendf(qureg, "measure q[%d] -> c[%d];\n", measureQubit, measureQubit);
}

void qasm_recordInitZero(Qureg qureg) {
    if (logging(qureg)) appendf(qureg, "reset q;\n");
}

void qasm_recordInitPlus(Qureg qureg) {
    if (!logging(qureg)) return;
    qasm_recordComment(qureg, "Initialising state |+>");
    qasm_recordInitZero(qureg);
    appendf(qureg, "h q;\n");
}

void qasm_recordInitClassical(Qureg qureg, long long int stateInd) {
    if (!logging(qureg)) return;
    char cmt[kMaxLine + 1];
    std::snprintf(cmt, sizeof cmt, "Initialising state |%lld>", stateInd);
    qasm_recordComment(qureg, cmt);
    qasm_recordInitZero(qureg);
    for (int q = 0; q < qureg.numQubitsRepresented; ++q)
        if ((stateInd >> q) & 1) qasm_recordGate(qureg, GATE_SIGMA_X, q);
}

}  // extern "C"
