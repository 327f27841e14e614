/* quest-b200 extension entry points: what a host needs beyond QuEST.h to integrate the engine with its own
 * CUDA streams, timers and device buffers.  None of these exist in the reference; each comment names the
 * reference mechanism it stands in for.  Plain C ABI: pointers and sizes only.
 */
#ifndef QUEST_B200_H
#define QUEST_B200_H

#include "QuEST.h"

#ifdef __cplusplus
extern "C" {
#endif

/* The CUDA stream (cudaStream_t) every kernel and copy of this process is enqueued on.  The reference GPU
 * backend uses the legacy default stream (QuEST_gpu.cu:591); a dedicated stream lets the caller bracket work
 * with its own CUDA events. */
void* qb200_getStream(void);

/* Make the engine enqueue on a caller-owned stream from now on (drains the current stream first). */
void qb200_setStream(void* cudaStream);

/* Number of kernels this process has launched since createQuESTEnv(). */
unsigned long long qb200_getLaunchCount(void);

/* CUDA device ordinal this process drives. */
int qb200_getDevice(void);

/* Device addresses of this rank's shard: real[numAmpsPerChunk], imag[numAmpsPerChunk] of qreal.
 * Stands in for the reference's Qureg.stateVec host arrays (QuEST.h:100), which this library never allocates. */
void qb200_getDevicePointers(Qureg qureg, void** real, void** imag);

/* Copy n amplitudes of this rank's shard, from local index start, to host arrays.  Synchronises.
 * Stands in for direct reads of Qureg.stateVec (QuEST_common.c:174-177). */
void qb200_copyAmpsToHost(Qureg qureg, long long int start, long long int n, qreal* reals, qreal* imags);

#ifdef __cplusplus
}
#endif

#endif /* QUEST_B200_H */
