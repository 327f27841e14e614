// Micro-benchmark: sustained DMUL/DADD (no FMA) and DFMA lane-operations per second of one B200, as a function of the
// operand DATA.  Every thread keeps 8 independent complex pairs in registers and applies a fixed 2x2 complex matrix to
// them over and over (the arithmetic of a dense one-qubit gate, 28 fp64 instructions per pair).  The operand values
// come from memory (pattern chosen on the command line) so that nothing folds at compile time.
//   build/fp64_peak <pattern> [iters]      pattern: 0 zeros, 1 real-only random, 2 complex random, 3 ones
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) k(const double* __restrict__ in, double* __restrict__ out, const double* __restrict__ m, int iters) {
    double ur[4], ui[4], lr[4], li[4];
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        ur[j] = in[(t * 16 + 4 * j + 0)]; ui[j] = in[(t * 16 + 4 * j + 1)];
        lr[j] = in[(t * 16 + 4 * j + 2)]; li[j] = in[(t * 16 + 4 * j + 3)];
    }
    const double r00 = m[0], i00 = m[1], r01 = m[2], i01 = m[3], r10 = m[4], i10 = m[5], r11 = m[6], i11 = m[7];
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const double nuR = r00 * ur[j] - i00 * ui[j] + r01 * lr[j] - i01 * li[j];
            const double nuI = r00 * ui[j] + i00 * ur[j] + r01 * li[j] + i01 * lr[j];
            const double nlR = r10 * ur[j] - i10 * ui[j] + r11 * lr[j] - i11 * li[j];
            const double nlI = r10 * ui[j] + i10 * ur[j] + r11 * li[j] + i11 * lr[j];
            ur[j] = nuR; ui[j] = nuI; lr[j] = nlR; li[j] = nlI;
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        out[t * 16 + 4 * j + 0] = ur[j]; out[t * 16 + 4 * j + 1] = ui[j];
        out[t * 16 + 4 * j + 2] = lr[j]; out[t * 16 + 4 * j + 3] = li[j];
    }
}

int main(int argc, char** argv) {
    const int pattern = argc > 1 ? atoi(argv[1]) : 2;
    const int iters = argc > 2 ? atoi(argv[2]) : 20000;
    const int blocks = 148 * 8, threads = 256;
    const size_t n = (size_t)blocks * threads * 16;
    double* h = (double*)malloc(n * sizeof(double));
    srand(1);
    for (size_t i = 0; i < n; ++i) {
        const double r = (rand() / (double)RAND_MAX - 0.5) * 1e-4;
        h[i] = pattern == 0 ? 0.0 : pattern == 3 ? 1.0 : (pattern == 1 && (i & 1)) ? 0.0 : r;
    }
    // a unitary (compactUnitary of the benchmark): keeps magnitudes constant over the iterations
    const double a = 1.232, p1 = 0.423, p2 = -0.6523;
    double hm[8] = {cos(a) * cos(p1), cos(a) * sin(p1), -sin(a) * cos(p2), sin(a) * sin(p2),
                    sin(a) * cos(p2), sin(a) * sin(p2), cos(a) * cos(p1), -cos(a) * sin(p1)};
    if (pattern == 1) { hm[1] = hm[3] = hm[5] = hm[7] = 0; hm[0] = hm[6] = cos(a); hm[2] = -sin(a); hm[4] = sin(a); }  // real rotation keeps imag = 0
    double *din, *dout, *dm;
    cudaMalloc(&din, n * 8); cudaMalloc(&dout, n * 8); cudaMalloc(&dm, 64);
    cudaMemcpy(din, h, n * 8, cudaMemcpyHostToDevice); cudaMemcpy(dm, hm, 64, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<blocks, threads>>>(din, dout, dm, 100);
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        k<<<blocks, threads>>>(din, dout, dm, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double ops = (double)blocks * threads * 4 * 28 * iters;
        printf("pattern %d: %.2f ms, %.2f T fp64 lane-ops/s = %.1f lanes/clk/SM at 1.965 GHz\n", pattern, ms, ops / ms / 1e9, ops / (ms * 1e-3) / 148 / 1.965e9);
    }
    return cudaGetLastError() != cudaSuccess;
}
