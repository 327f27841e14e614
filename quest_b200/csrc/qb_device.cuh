// quest-b200 device building blocks: wide vector loads/stores on the SoA amplitude arrays, and the bit-insertion
// plan that maps a dense task index onto the amplitudes a gate actually touches.
//
// Layout in HBM: one shard = two arrays real[n], imag[n] of qreal, n = 2^L, each 256-byte aligned.  A "vector" is
// VEC consecutive amplitudes of one array (VEC*sizeof(qreal) = 8, 16 or 32 bytes; 32 B = one DRAM sector, moved by
// a single LDG.E.ENL2.256 / STG.E.ENL2.256 on sm_100a).  Bits [0, log2 VEC) of an amplitude index select the lane
// inside a vector ("lo" bits); the remaining bits select the vector ("hi" bits).
#pragma once

#include <cstdint>

#include "QuEST_precision.h"
#include "qb_plan.h"

namespace qb {

__device__ __forceinline__ unsigned long long expand(unsigned long long k, const BitPlan& p) {
#pragma unroll
    for (int i = 0; i < kMaxInsert; ++i) {
        if (i < p.nIns) {
            const int b = p.pos[i];
            const unsigned long long low = k & ((1ull << b) - 1ull);
            k = ((k >> b) << (b + 1)) | low;
        }
    }
    return k | p.setMask;
}

template <typename T, int N>
struct Vec {
    T v[N];
};

// ---- double ------------------------------------------------------------------------------------------
__device__ __forceinline__ Vec<double, 1> ldv(const double* p, Vec<double, 1>*) {
    Vec<double, 1> r;
    asm volatile("ld.global.f64 %0, [%1];" : "=d"(r.v[0]) : "l"(p));
    return r;
}
__device__ __forceinline__ Vec<double, 2> ldv(const double* p, Vec<double, 2>*) {
    Vec<double, 2> r;
    asm volatile("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(r.v[0]), "=d"(r.v[1]) : "l"(p));
    return r;
}
__device__ __forceinline__ Vec<double, 4> ldv(const double* p, Vec<double, 4>*) {
    Vec<double, 4> r;
    asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(r.v[0]), "=d"(r.v[1]), "=d"(r.v[2]), "=d"(r.v[3])
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void stv(double* p, const Vec<double, 1>& r) {
    asm volatile("st.global.f64 [%0], %1;" ::"l"(p), "d"(r.v[0]) : "memory");
}
__device__ __forceinline__ void stv(double* p, const Vec<double, 2>& r) {
    asm volatile("st.global.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(r.v[0]), "d"(r.v[1]) : "memory");
}
__device__ __forceinline__ void stv(double* p, const Vec<double, 4>& r) {
    asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(r.v[0]), "d"(r.v[1]), "d"(r.v[2]), "d"(r.v[3])
                 : "memory");
}

// ---- float -------------------------------------------------------------------------------------------
__device__ __forceinline__ Vec<float, 1> ldv(const float* p, Vec<float, 1>*) {
    Vec<float, 1> r;
    asm volatile("ld.global.f32 %0, [%1];" : "=f"(r.v[0]) : "l"(p));
    return r;
}
__device__ __forceinline__ Vec<float, 2> ldv(const float* p, Vec<float, 2>*) {
    Vec<float, 2> r;
    asm volatile("ld.global.v2.f32 {%0,%1}, [%2];" : "=f"(r.v[0]), "=f"(r.v[1]) : "l"(p));
    return r;
}
__device__ __forceinline__ Vec<float, 4> ldv(const float* p, Vec<float, 4>*) {
    Vec<float, 4> r;
    asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3])
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void stv(float* p, const Vec<float, 1>& r) {
    asm volatile("st.global.f32 [%0], %1;" ::"l"(p), "f"(r.v[0]) : "memory");
}
__device__ __forceinline__ void stv(float* p, const Vec<float, 2>& r) {
    asm volatile("st.global.v2.f32 [%0], {%1,%2};" ::"l"(p), "f"(r.v[0]), "f"(r.v[1]) : "memory");
}
__device__ __forceinline__ void stv(float* p, const Vec<float, 4>& r) {
    asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3])
                 : "memory");
}

template <int N>
__device__ __forceinline__ Vec<qreal, N> load(const qreal* p) {
    return ldv(p, (Vec<qreal, N>*)nullptr);
}
template <int N>
__device__ __forceinline__ void store(qreal* p, const Vec<qreal, N>& r) {
    stv(p, r);
}
template <int N>
__device__ __forceinline__ Vec<qreal, N> zeros() {
    Vec<qreal, N> r;
#pragma unroll
    for (int e = 0; e < N; ++e) r.v[e] = 0;
    return r;
}

// ---- block-wide sum of up to NV doubles per thread: shuffle tree inside each warp, then across warps ----
template <int NV>
__device__ __forceinline__ void blockSum(double (&x)[NV], double* smem /* >= NV*32 doubles */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nWarps = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x[j] += __shfl_xor_sync(0xffffffffu, x[j], o);
        if (lane == 0) smem[j * 32 + warp] = x[j];
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            double y = lane < nWarps ? smem[j * 32 + lane] : 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) y += __shfl_xor_sync(0xffffffffu, y, o);
            x[j] = y;
        }
    }
}

}  // namespace qb
