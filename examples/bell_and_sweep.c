/* A plain C program written against QuEST.h only - the kind of program the reference builds as its `demo`
 * (reference CMakeLists.txt:11,44).  It links against quest-b200's libQuEST.so without any source change:
 *
 *   gcc -O2 -I include examples/bell_and_sweep.c -o demo -L quest_b200/lib -lQuEST -lm -Wl,-rpath,$PWD/quest_b200/lib
 *   ./demo [numQubits]
 *
 * Prints values a reader can check by hand (GHZ probabilities, a measurement under a fixed seed, purity after
 * noise) and the wall-clock time of a gate sweep, synchronising before reading the clock as any GPU backend needs.
 */
#include <stdio.h>
#include <stdlib.h>
#include <time.h>

#include "QuEST.h"

static double now(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

int main(int argc, char** argv) {
    int n = argc > 1 ? atoi(argv[1]) : 20;
    QuESTEnv env = createQuESTEnv();
    Qureg q = createQureg(n, env);
    if (env.rank == 0) reportQuregParams(q);

    /* GHZ state: H on qubit 0, then a chain of CNOTs */
    hadamard(q, 0);
    for (int t = 1; t < n; t++) controlledNot(q, t - 1, t);
    Complex first = getAmp(q, 0), last = getAmp(q, (1LL << n) - 1);
    if (env.rank == 0) {
        printf("GHZ: amp[0] = %.12f, amp[2^n-1] = %.12f (expect 0.707106781187 each)\n", (double)first.real, (double)last.real);
        printf("GHZ: P(qubit %d = 1) = %.12f, total probability = %.12f\n", n - 1, (double)calcProbOfOutcome(q, n - 1, 1),
               (double)calcTotalProb(q));
    }
    unsigned long int seed[1] = {1};
    seedQuEST(seed, 1);
    int m0 = measure(q, 0);
    int m1 = measure(q, n - 1);
    if (env.rank == 0) printf("GHZ: measured qubit 0 -> %d, qubit %d -> %d (always equal)\n", m0, n - 1, m1);

    /* a sweep of rotations and controlled phases over every qubit, timed */
    initPlusState(q);
    syncQuESTEnv(env);
    double t0 = now();
    int gates = 0;
    for (int layer = 0; layer < 4; layer++)
        for (int t = 0; t < n; t++) {
            rotateY(q, t, 0.1 * (t + 1) + 0.01 * layer);
            Complex alpha = {0.6, 0.0}, beta = {0.0, 0.8};
            compactUnitary(q, t, alpha, beta);
            controlledPhaseShift(q, t, (t + 1) % n, 0.3);
            gates += 3;
        }
    syncQuESTEnv(env);
    double dt = now() - t0;
    if (env.rank == 0)
        printf("sweep: %d gates on %d qubits in %.4f s (%.1f gates/s), total probability = %.12f\n", gates, n, dt, gates / dt,
               (double)calcTotalProb(q));
    destroyQureg(q, env);

    /* a small noisy density matrix */
    int nd = n / 2 < 12 ? n / 2 : 12;
    Qureg rho = createDensityQureg(nd, env);
    initPlusState(rho);
    for (int t = 0; t < nd; t++) {
        rotateX(rho, t, 0.2 * (t + 1));
        applyOneQubitDephaseError(rho, t, 0.05);
        applyOneQubitDepolariseError(rho, t, 0.02);
        applyOneQubitDampingError(rho, t, 0.1);
    }
    if (env.rank == 0)
        printf("density: %d qubits, purity = %.12f, trace = %.12f\n", nd, (double)calcPurity(rho), (double)calcTotalProb(rho));
    destroyQureg(rho, env);
    destroyQuESTEnv(env);
    return 0;
}
