/* quest-b200 test hooks.  NOT part of the installed interface (include/): these symbols exist so that the host-side
 * planning logic - which rank does what for a sharded gate, where qubits live after slot exchanges, how a pipelined
 * exchange is cut into slabs, how the fusion queue is cut into passes and sub-groups - can be unit-tested on a machine
 * without a GPU (tests/test_shard_plan.py, tests/test_fusion_plan.py).  None of them makes a CUDA call.  Layouts of the
 * output arrays are documented at the definitions in backend_cuda.cu. */
#ifndef QB_DEBUG_H
#define QB_DEBUG_H

#ifdef __cplusplus
extern "C" {
#endif

/* The plan rank `chunkId` of a register with 2^localBits amplitudes per shard would execute for a one-target gate.
 * Stands in for the reference's per-gate local/distributed decision (halfMatrixBlockFitsInChunk / chunkIsUpper /
 * getChunkPairId, QuEST_cpu_distributed.c:231-365). */
void qb200_debugPlanGate(int localBits, int chunkId, int target, unsigned long long ctrlMask, int vec,
                         unsigned long long* out);

/* The qubit-slot exchanges a sharded register would perform for a sequence of dense gates executed one by one, and the
 * exchanges that restore the identity layout afterwards. */
int qb200_debugRemap(int nQubits, int localBits, int nOps, const int* targets, const unsigned long long* masks,
                     int* eventsOut, int* physOut, int* canonOut);

/* How a recorded queue of logical gates on a sharded register is cut into local segments and slot exchanges, with the
 * layout and use history carried from one call to the next. */
int qb200_debugWalk(int nQubits, int localBits, int nOps, const int* targets, const int* targets2,
                    const unsigned long long* masks, unsigned long long* mapState, int* stepsOut);

/* The flush-time peephole: which adjacent gates of a queue are given one dispatch. */
int qb200_debugPeephole(int nOps, const int* kinds, const int* targets, const unsigned long long* masks, int* kindsOut,
                        int* firstOut);

/* The exchanges that restore the identity layout from any layout. */
int qb200_debugCanon(int nQubits, int localBits, const int* phys, int* out);

/* The slab bits chosen for pipelining a fused pass with a slot exchange, and every k_swapPeer launch plan of that
 * exchange. */
int qb200_debugSlabs(int localBits, unsigned long long busyMask, int want, int lowestAllowed, int vec, int localSlot,
                     int* bitsOut, unsigned long long* swapOut);

/* How the fusion queue cuts a gate list into passes (tile bits) and sub-groups (register bits), and the descriptor the
 * kernel receives for every gate. */
int qb200_debugTilePlan(int localBits, int nOps, const int* kinds, const int* targets, const int* targets2,
                        const unsigned long long* masks, int reorder, int* passesOut, int* orderOut, int* subsOut,
                        unsigned long long* descOut, int* nSubsOut);

#ifdef __cplusplus
}
#endif

#endif /* QB_DEBUG_H */
