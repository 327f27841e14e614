// quest-b200: total probability of a state held in device memory, computed WITHOUT allocating anything.
//
// Who this is for.  calcTotalProb() and the other reductions of this library keep their per-block partial sums in
// scratch that createQuESTEnv() obtains with cudaMalloc.  Some hosts of a simulator cannot do that:
//   * a process in which another runtime (a tensor framework's caching allocator, a memory-partitioned MPS/MIG
//     client with a fixed quota) already owns all of HBM, so a further cudaMalloc fails or is forbidden by policy;
//   * a latency-critical service loop, where cudaMalloc/cudaFree are unwelcome because they synchronise the
//     whole device;
//   * code that only holds raw device pointers to real[]/imag[] (for instance from qb200_getDevicePointers, or
//     amplitudes produced by its own kernels) and never creates a QuESTEnv at all.
// For them this file offers one entry point, qb200_calcTotalProbNoAlloc(), whose only working memory is a
// fixed array that is part of the library image itself.
//
// Limits and rules a caller must know:
//   * at most QB200_NOALLOC_MAX_AMPS = 2^31 amplitudes per call (a 31-qubit state-vector, or a 15-qubit density
//     matrix when used for purity); larger sizes are refused with an error, nothing is launched;
//   * the scratch is 256 KiB + 256 B of device memory per GPU that loads the library, held until the process exits;
//   * every call on a GPU uses the same scratch and finds the previous call's contents in it.  The call clears the
//     one word that has to start at zero (the arrival counter) itself; the partial sums are fully overwritten
//     before they are read.  Two calls that overlap in time on the same GPU (different streams, or different host
//     threads) race on that memory and give wrong sums - serialise them;
//   * the first call on each GPU looks up the device address of the scratch (cudaGetSymbolAddress).  Make that
//     first call outside of any stream capture.
// The function calls no cudaMalloc, cudaFree, cudaHostAlloc or any library routine that allocates.
#include <cuda_runtime.h>

#include <atomic>
#include <cstdio>

#include "qb_device.cuh"
#include "quest_b200.h"

static constexpr int kThreads = 256;
static constexpr long long kSlabAmps = 1LL << 16;             // amplitudes reduced by one block (1 MiB of state)
static constexpr int kMaxBlocks = 1 << 15;                    // => 2^31 amplitudes at most
static constexpr long long kMaxAmps = kSlabAmps * kMaxBlocks; // QB200_NOALLOC_MAX_AMPS
static constexpr int kHeaderBytes = 256;                      // arrival counter lives here, partial sums follow
static constexpr int kMaxDevices = 64;

// The whole working memory of this entry point: 256 B header + 2^15 partial sums = 256 KiB + 256 B per GPU.
// In host code this name is only a handle for the runtime, not a device address: the launcher below obtains the
// real address once per device with cudaGetSymbolAddress and hands that to the kernel.
static __device__ __align__(256) unsigned char g_scratch[kHeaderBytes + kMaxBlocks * sizeof(double)];

namespace {

// One block sums |amp|^2 over its slab; the last block to finish adds the per-block sums in index order.
template <int VEC>
__global__ void __launch_bounds__(kThreads) k_totalProbSlabs(const qreal* __restrict__ re, const qreal* __restrict__ im,
                                                            long long numAmps, unsigned char* scratch, double* out) {
    unsigned int* arrivals = reinterpret_cast<unsigned int*>(scratch);
    double* partial = reinterpret_cast<double*>(scratch + kHeaderBytes);
    __shared__ double smem[32];
    __shared__ bool amLast;

    const long long begin = (long long)blockIdx.x * kSlabAmps;
    const long long end = begin + kSlabAmps < numAmps ? begin + kSlabAmps : numAmps;
    double acc[1] = {0.0};
#pragma unroll 4
    for (long long i = begin + (long long)threadIdx.x * VEC; i < end; i += (long long)kThreads * VEC) {
        const qb::Vec<qreal, VEC> r = qb::load<VEC>(re + i), m = qb::load<VEC>(im + i);
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[0] += (double)(r.v[e] * r.v[e] + m.v[e] * m.v[e]);
    }
    qb::blockSum<1>(acc, smem);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = acc[0];
        __threadfence();
        amLast = (atomicAdd(arrivals, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!amLast) return;
    __threadfence();
    double total[1] = {0.0};
    for (unsigned int b = threadIdx.x; b < gridDim.x; b += kThreads) total[0] += __ldcg(partial + b);
    __syncthreads();
    qb::blockSum<1>(total, smem);
    if (threadIdx.x == 0) *out = total[0];
}

// Per-device address of g_scratch, looked up on the first call on that device.  Atomic: two host threads may make
// their first call on one GPU at the same time (both then store the same value).
std::atomic<unsigned char*> g_scratchAddr[kMaxDevices];

int fail(const char* what, cudaError_t e = cudaSuccess) {
    if (e != cudaSuccess)
        std::fprintf(stderr, "qb200_calcTotalProbNoAlloc: %s: %s\n", what, cudaGetErrorString(e));
    else
        std::fprintf(stderr, "qb200_calcTotalProbNoAlloc: %s\n", what);
    return e != cudaSuccess ? (int)e : -1;
}

}  // namespace

extern "C" int qb200_calcTotalProbNoAlloc(const qreal* dReal, const qreal* dImag, long long int numAmps,
                                          double* dSumOut, void* cudaStream) {
    if (!dReal || !dImag || !dSumOut) return fail("null device pointer");
    if (numAmps < 1) return fail("numAmps must be at least 1");
    if (numAmps > kMaxAmps) {
        std::fprintf(stderr,
                     "qb200_calcTotalProbNoAlloc: %lld amplitudes exceed the fixed scratch of this entry point "
                     "(QB200_NOALLOC_MAX_AMPS = %lld); use calcTotalProb(), which allocates its scratch\n",
                     numAmps, kMaxAmps);
        return -2;
    }
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail("cudaGetDevice", e);
    if (dev < 0 || dev >= kMaxDevices) return fail("device ordinal out of range");
    // The scratch belongs to the CURRENT device, so the amplitudes must live there too: a pointer into another GPU's
    // memory would be paired with the wrong scratch (and a kernel launched on the wrong device).
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, dReal) == cudaSuccess && attr.type == cudaMemoryTypeDevice && attr.device != dev)
        return fail("dReal lives on another GPU than the current device; call cudaSetDevice for the GPU that holds the state");
    unsigned char* scratch = g_scratchAddr[dev].load(std::memory_order_acquire);
    if (!scratch) {  // set-up, once per device, with that device current
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_scratch);
        if (e != cudaSuccess || !p) return fail("cudaGetSymbolAddress(scratch)", e);
        scratch = static_cast<unsigned char*>(p);
        g_scratchAddr[dev].store(scratch, std::memory_order_release);
    }
    cudaStream_t stream = static_cast<cudaStream_t>(cudaStream);
    // the scratch still holds whatever the previous call left: only the arrival counter has to start at zero
    e = cudaMemsetAsync(scratch, 0, kHeaderBytes, stream);
    if (e != cudaSuccess) return fail("cudaMemsetAsync(arrival counter)", e);
    const unsigned int blocks = (unsigned int)((numAmps + kSlabAmps - 1) / kSlabAmps);
    const bool wide = (numAmps % 4 == 0) && ((reinterpret_cast<uintptr_t>(dReal) | reinterpret_cast<uintptr_t>(dImag)) % 32 == 0);
    if (wide)
        k_totalProbSlabs<4><<<blocks, kThreads, 0, stream>>>(dReal, dImag, numAmps, scratch, dSumOut);
    else
        k_totalProbSlabs<1><<<blocks, kThreads, 0, stream>>>(dReal, dImag, numAmps, scratch, dSumOut);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail("kernel launch", e);
    return 0;
}
