// quest-b200 multi-GPU layer (placeholder until the NVLink path lands): single-process, single-GPU only.
#include <cstdio>
#include <cstdlib>

#include "qb_comm.h"

namespace qb {
namespace comm {

Launch readLaunchEnv() { return Launch(); }
void init(const Launch&, int, cudaStream_t) {}
void shutdown() {}
void hostBarrier() {}
int allAnd(int flag) { return flag; }
void broadcastBytes(void*, size_t, int) {}
void allReduceSum(double*, int, cudaStream_t) {}
void registerShard(Qureg&, void*, void*, long long) {}
void unregisterShard(const Qureg&) {}
void unsupported(const char* what) {
    std::printf("!!!\nquest-b200: %s is not implemented\n!!!\nexiting..\n", what);
    std::fflush(stdout);
    std::exit(EXIT_FAILURE);
}
Partner partnerOf(const Qureg&, int) { unsupported("cross-shard gate without a multi-GPU environment"); }
void pairBarrier(int, cudaStream_t) {}

}  // namespace comm
}  // namespace qb
