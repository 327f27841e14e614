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

/* Gate fusion.  Gate calls may return before the GPU has run them (the reference GPU backend launches
 * asynchronously too, QuEST_gpu.cu:591); with fusion enabled (default; QUEST_B200_FUSE=0 disables) they are recorded
 * and executed at the next call that needs the amplitudes, consecutive gates whose targets fall on up to eleven index
 * bits (the four lowest included) sharing ONE pass over HBM through a shared-memory tile.  Results are bit-identical to
 * gate-at-a-time execution.
 * qb200_setFusion(0) drains every queue and returns to one kernel per gate call; qb200_flush() drains the queues
 * without waiting for the GPU; qb200_getFusionStats reports how many fused passes ran and how many gates they held. */
void qb200_setFusion(int enabled);
int qb200_getFusion(void);
void qb200_flush(void);
void qb200_getFusionStats(unsigned long long* fusedLaunches, unsigned long long* fusedGates);

/* Sharded registers: a slot exchange (see qb200_setRemap) and the fused passes on both sides of it are cut into slabs and
 * pipelined on two streams - the exchange of a slab runs over NVLink while the pass before it computes the next slab, and
 * the pass after it follows slab by slab (default; QUEST_B200_OVERLAP=0 / qb200_setOverlap(0) run them back to back).
 * Results are identical either way.  The counter reports how many exchanges were pipelined.  No reference counterpart:
 * the reference exchanges whole shards with a blocking MPI_Sendrecv (QuEST_cpu_distributed.c:451-479). */
void qb200_setOverlap(int enabled);
unsigned long long qb200_getOverlappedExchangeCount(void);

/* calcProbOfOutcome on state-vectors: the second query on an unchanged state computes P(qubit = 0) for EVERY qubit in
 * one pass over HBM; further queries are answered from that result until any call changes the state (e.g. the per-qubit
 * read-out after a QFT costs two passes instead of one per qubit).  QUEST_B200_PROB_CACHE=0 disables it.  Registers whose
 * device pointers were handed out (qb200_getDevicePointers) are never cached.  The reference recomputes every query
 * (QuEST_cpu_local.c:267, QuEST_cpu.c:2844).  Numerical note: the all-qubit pass adds the same terms in another order
 * than the one-qubit pass, so the first and a later query on an unchanged state may differ in the last bits (~1e-16
 * relative); measure() draws against whichever value it is given, so an outcome can differ from the uncached path only
 * when the random draw lands within that distance of P(0). */

/* Bulk state I/O.  setAmps / initStateFromAmps / setDensityAmps and qb200_copyAmpsToHost accept any host pointer.  From
 * pinned memory the copy runs at the PCIe rate as it is; from pageable memory (malloc, numpy) transfers of 8 MiB and more
 * are staged through pinned buffers by several worker threads (QUEST_B200_IO_THREADS, default min(8, cores/2)).  The
 * counter reports how many transfers took that path.  Reference counterpart: statevec_setAmps, QuEST_gpu.cu:34-44
 * (two plain cudaMemcpy). */
unsigned long long qb200_getPipelinedTransferCount(void);

/* Gate reordering (off by default; QUEST_B200_REORDER=1 / qb200_setReordering(1)).  When recorded gates are executed, a
 * gate may move forward past gates it commutes with (every index bit the two share is a control or phase-mask bit of
 * both) so that it can join an earlier pass: layered circuits - a gate on every qubit, then the next kind of gate on
 * every qubit - run qubit block by qubit block (BASELINE configs[0]: 12 -> 4 passes per layer, 1.4x faster).  The
 * mathematics is unchanged, the ORDER OF ROUNDINGS is not: amplitudes differ from the reference's in the last bit or
 * two (measured 4e-18 absolute at 24 qubits), well inside the 1e-12 bar, but no longer bit-identical - hence opt-in.
 * No reference counterpart: the reference executes every call at once. */
void qb200_setReordering(int enabled);
int qb200_getReordering(void);

/* Deferred uploads (off by default; single-GPU registers).  The reference's setAmps / initStateFromAmps are complete on
 * return (QuEST_gpu.cu:34-44: two blocking cudaMemcpy), which leaves the GPU idle for the whole PCIe transfer - 0.3 s for
 * a 30-qubit state.  With qb200_setDeferredUpload(1) a call that overwrites the WHOLE register from PINNED host memory
 * returns as soon as the copies are enqueued (eight slabs on a copy stream); gate calls made afterwards are recorded as
 * usual, and when they are executed the leading passes follow the upload slab by slab, so the transfer and the
 * arithmetic overlap.  The price: the caller must not modify or free the host arrays before the next synchronising
 * call (any value-returning call such as calcTotalProb / getAmp / measure, or syncQuESTEnv).  Transfers that do not
 * qualify (pageable memory, partial ranges, small or sharded registers) are copied the blocking way as before.
 * The counters report how many uploads were deferred and how many passes ran underneath one. */
void qb200_setDeferredUpload(int enabled);
unsigned long long qb200_getDeferredUploadCount(void);
unsigned long long qb200_getPassesUnderUploadCount(void);

/* Arithmetic of the fused pass.  0 (default): every product and sum is rounded separately, exactly as the reference's
 * CPU build does (gcc without FMA), so amplitudes are bit-identical to it.  1 (or QUEST_B200_FMA=1): products feeding
 * a sum may be contracted into fused multiply-adds - fewer fp64 instructions per gate, results within an ulp or two
 * per gate of the exact mode (inside the 1e-12 bar, not bit-identical).  Pending gates run under the old setting.
 * No reference counterpart: the reference's precision is a compile-time choice (QuEST_precision.h). */
void qb200_setArithmetic(int contracted);
int qb200_getArithmetic(void);

/* Sharded registers (WORLD_SIZE > 1).  With remapping enabled (default; QUEST_B200_REMAP=0 disables) a dense gate on a
 * qubit that selects the GPU is served by making that qubit local: it is exchanged with a local qubit - half a shard
 * over NVLink - and later gates on it cost no communication; the reference exchanges whole shards for every such gate
 * (QuEST_cpu_distributed.c:451-479).  In the default mode the exchanges are planned when the recorded gates are
 * executed, from the whole queue: the qubit that leaves is the one whose next use lies furthest away (known from the
 * queue, else predicted from its last two uses), which costs one exchange per shard-selecting qubit per sweep of all
 * qubits.  Every index-addressed call (getAmp, setAmps, ...) sees the logical layout.  qb200_setRemap(0) restores the
 * identity layout and runs such gates as cross-shard passes. */
void qb200_setRemap(int enabled);
unsigned long long qb200_getSlotExchangeCount(void);

/* Measurement aid: with timing enabled every fused pass is bracketed by CUDA events on the engine stream;
 * qb200_getFusedKernelTime synchronises, returns their summed device time and count, and clears the record. */
void qb200_setKernelTiming(int enabled);
void qb200_getFusedKernelTime(double* totalMs, unsigned long long* launches);

/* CUDA device ordinal this process drives. */
int qb200_getDevice(void);

/* Device addresses of this rank's shard: real[numAmpsPerChunk], imag[numAmpsPerChunk] of qreal.
 * Stands in for the reference's Qureg.stateVec host arrays (QuEST.h:100), which this library never allocates.
 * The call first executes every gate recorded for the register (and restores the identity qubit layout of a sharded
 * one).  From then on the register takes no part in gate fusion: each later gate call enqueues its own kernel on the
 * engine stream at once, so work the caller enqueues on that stream (qb200_getStream) afterwards sees the amplitudes
 * all earlier API calls produce, and nothing recorded can later overwrite what the caller wrote.  Work on OTHER
 * streams must be ordered against the engine stream by the caller (events, or syncQuESTEnv). */
void qb200_getDevicePointers(Qureg qureg, void** real, void** imag);

/* Copy n amplitudes of this rank's shard, from local index start, to host arrays.  Synchronises.
 * Stands in for direct reads of Qureg.stateVec (QuEST_common.c:174-177). */
void qb200_copyAmpsToHost(Qureg qureg, long long int start, long long int n, qreal* reals, qreal* imags);

/* Sum of |amplitude|^2 over numAmps amplitudes held in device arrays dReal/dImag, written as one double to device
 * memory at dSumOut, enqueued on cudaStream (0 = default stream).  Returns 0 on success, non-zero after printing the
 * reason to stderr.  This entry point has no counterpart in the reference: it exists for hosts that must not
 * allocate device memory (another runtime owns all of HBM, a fixed MPS/MIG quota, a service loop that cannot
 * afford cudaMalloc's device-wide synchronisation, or callers that hold only raw amplitude pointers and no QuESTEnv).
 * It calls no allocator; its working memory is a fixed 256 KiB + 256 B array inside the library image, one per GPU,
 * kept until the process exits.  Limits: numAmps <= QB200_NOALLOC_MAX_AMPS (2^31), larger sizes are refused.  All
 * calls on a GPU share that array (the call clears the counter it needs itself): calls that overlap in time on
 * one GPU race and must be serialised by the caller.  The first call on each GPU performs a one-time address
 * look-up and should be made outside of any stream capture.  The GPU that holds dReal/dImag must be the current
 * device (cudaSetDevice) and own cudaStream; a state on another GPU is refused. */
#define QB200_NOALLOC_MAX_AMPS (1LL << 31)
int qb200_calcTotalProbNoAlloc(const qreal* dReal, const qreal* dImag, long long int numAmps, double* dSumOut,
                               void* cudaStream);

#ifdef __cplusplus
}
#endif

#endif /* QUEST_B200_H */
