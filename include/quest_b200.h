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
 * and executed at the next call that needs the amplitudes, consecutive gates on up to four distinct target qubits
 * sharing ONE pass over HBM.  Results are bit-identical to gate-at-a-time execution.
 * qb200_setFusion(0) drains every queue and returns to one kernel per gate call; qb200_flush() drains the queues
 * without waiting for the GPU; qb200_getFusionStats reports how many fused passes ran and how many gates they held. */
void qb200_setFusion(int enabled);
int qb200_getFusion(void);
void qb200_flush(void);
void qb200_getFusionStats(unsigned long long* fusedLaunches, unsigned long long* fusedGates);

/* Sharded registers: when a slot exchange (see qb200_setRemap) follows recorded gates, the last fused pass and the
 * exchange are cut into slabs and pipelined on two streams - the exchange of a slab runs over NVLink while the pass
 * computes the next one (default; QUEST_B200_OVERLAP=0 / qb200_setOverlap(0) run them back to back).  Results are
 * identical either way.  The counter reports how many exchanges were pipelined.  No reference counterpart: the
 * reference exchanges whole shards with a blocking MPI_Sendrecv (QuEST_cpu_distributed.c:451-479). */
void qb200_setOverlap(int enabled);
unsigned long long qb200_getOverlappedExchangeCount(void);

/* Sharded registers, opt-in (QUEST_B200_DISTRIBUTED=1 / qb200_setDistributed(1); off by default: validated on 2 GPUs
 * only, and faster than remapping only when a shard-selecting qubit is not reused by later groups): a run of gates whose fused qubits include g shard-selecting qubits runs as ONE pass over the 2^g GPUs
 * that differ in those qubits, each GPU processing its slab of the pairs over the peer mappings, instead of one slot
 * exchange per such qubit.  The counter reports how many such passes ran.  No reference counterpart. */
void qb200_setDistributed(int mode);  /* 0 off, 1 always distributed, 2 make a qubit local first if the recorded queue uses it again */
unsigned long long qb200_getDistributedPassCount(void);

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

/* Arithmetic of the fused pass.  0 (default): every product and sum is rounded separately, exactly as the reference's
 * CPU build does (gcc without FMA), so amplitudes are bit-identical to it.  1 (or QUEST_B200_FMA=1): products feeding
 * a sum may be contracted into fused multiply-adds - fewer fp64 instructions per gate, results within an ulp or two
 * per gate of the exact mode (inside the 1e-12 bar, not bit-identical).  Pending gates run under the old setting.
 * No reference counterpart: the reference's precision is a compile-time choice (QuEST_precision.h). */
void qb200_setArithmetic(int contracted);
int qb200_getArithmetic(void);

/* Sharded registers (WORLD_SIZE > 1).  With remapping enabled (default; QUEST_B200_REMAP=0 disables) a dense gate on a
 * qubit that selects the GPU first makes that qubit local by exchanging it with the least recently used local qubit
 * - half a shard over NVLink - and later gates on it cost no communication; the reference exchanges whole shards for
 * every such gate (QuEST_cpu_distributed.c:451-479).  Every index-addressed call (getAmp, setAmps, ...) sees the
 * logical layout.  qb200_setRemap(0) restores the identity layout and runs such gates as cross-shard passes. */
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

/* Test hook (host logic only, makes no CUDA call): the plan rank `chunkId` of a register with 2^localBits amplitudes
 * per shard would execute for a one-target gate.  Stands in for the reference's per-gate local/distributed decision
 * (halfMatrixBlockFitsInChunk / chunkIsUpper / getChunkPairId, QuEST_cpu_distributed.c:231-365).  Layout of `out`
 * (18 words) is documented at the definition. */
void qb200_debugPlanGate(int localBits, int chunkId, int target, unsigned long long ctrlMask, int vec,
                         unsigned long long* out);

/* Test hook (host logic only, makes no CUDA call): the qubit-slot exchanges a sharded register would perform for a
 * sequence of dense gates, and the exchanges that restore the identity layout afterwards.  Layouts at the definition.
 * No counterpart in the reference, which exchanges whole shards for every gate on a shard-selecting qubit
 * (QuEST_cpu_distributed.c:451-479). */
int qb200_debugRemap(int nQubits, int localBits, int nOps, const int* targets, const unsigned long long* masks,
                     int* eventsOut, int* physOut, int* canonOut);

/* Test hook (host logic only, makes no CUDA call): how the distributed mode cuts a queue of logical gates into groups and
 * which slot exchanges it places before each.  Layout at the definition. */
int qb200_debugWalk(int nQubits, int localBits, int nOps, const int* kinds, const int* targets, const int* targets2,
                    const unsigned long long* masks, int lookahead, int* stepsOut);

/* Test hook (host logic only, makes no CUDA call): the plan of a distributed fused pass for rank chunkId.  Layout at the
 * definition. */
int qb200_debugDistributed(int localBits, int chunkId, const int* bits, int* out);

/* Test hook (host logic only, makes no CUDA call): the slab bits chosen for pipelining a fused pass with a slot
 * exchange, and every k_swapPeer launch plan of that exchange.  Layouts at the definition. */
int qb200_debugSlabs(int localBits, unsigned long long busyMask, int want, int lowestAllowed, int vec, int localSlot,
                     int* bitsOut, unsigned long long* swapOut);

/* Test hook (host logic only, makes no CUDA call): how the fusion queue would group a gate list for a shard of
 * 2^localBits amplitudes.  kinds: 0 matrix, 1 hadamard, 2 swap, 3 pauliY, 4 phase, 5 flip.  Layouts at the definition. */
int qb200_debugFuse(int localBits, int nOps, const int* kinds, const int* targets, const unsigned long long* masks,
                    int* groupsOut, unsigned long long* descOut);

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
