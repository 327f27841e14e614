// quest-b200 multi-GPU layer: one process per GPU, the state sharded by its top log2(G) qubits
// (reference model: QuEST/src/CPU/QuEST_cpu_distributed.c).  Declarations only; see qb_comm.cu.
#pragma once

#include <cuda_runtime.h>

#include "QuEST.h"

namespace qb {
namespace comm {

struct Launch {
    int rank = 0;
    int world = 1;
    int device = 0;
    char masterAddr[64] = "127.0.0.1";
    int port = 0;
};

Launch readLaunchEnv();
void init(const Launch& ln, int device, cudaStream_t stream);
void shutdown();
void hostBarrier();
int allAnd(int flag);
void broadcastBytes(void* buf, size_t n, int root);
void allReduceSum(double* devBuf, int n, cudaStream_t stream);
void allGatherBytes(const void* devIn, void* devOut, size_t bytesPerRank, cudaStream_t stream);
void registerShard(Qureg& q, void* re, void* im, long long n);
void unregisterShard(const Qureg& q);
[[noreturn]] void unsupported(const char* what);

// Peer-shard addresses of the partner that differs in global bit `gbit` of the rank index, plus the stream-ordered
// handshakes that bracket a kernel touching both shards.
struct Partner {
    qreal* re;
    qreal* im;
    int rank;
};
Partner partnerOf(const Qureg& q, int gbit);
Partner shardOf(const Qureg& q, int rank);  // any rank's shard as mapped into this process (own rank: own pointers)
void pairBarrier(int partnerRank, cudaStream_t stream, int channel = 0);  // channel: 0 engine stream, 1 communication stream

}  // namespace comm
}  // namespace qb
