// quest-b200 multi-GPU layer: one process per GPU, state sharded by its top log2(G) qubits.
//
// Replaces the MPI plumbing of the reference's distributed CPU backend (QuEST/src/CPU/QuEST_cpu_distributed.c):
//   MPI_Init / rank / size          -> launcher environment (torchrun: RANK, WORLD_SIZE, LOCAL_RANK, MASTER_*)
//   MPI_Bcast / MPI_Barrier / LAND  -> a small TCP star rooted at rank 0 (host-side control traffic only)
//   MPI_Sendrecv of whole shards    -> nothing is sent: partner shards are mapped into this process with CUDA IPC
//                                      and read/written directly over NVLink by the gate kernel (k_pairsPeer)
//   MPI_Allreduce of one scalar     -> ncclAllReduce on the engine stream
//   (pair handshakes)               -> a sequence counter in the partner's memory + cuStreamWaitValue32, no kernel spins
// NCCL is loaded with dlopen (the copy already in the process if a framework brought one, else libnccl.so.2), so a
// single-GPU user never needs it.
#include <cuda.h>  // driver types only: entry points are fetched with cudaGetDriverEntryPoint, libcuda is not linked

#include <arpa/inet.h>
#include <dlfcn.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <thread>
#include <vector>

#include "qb_comm.h"

namespace qb {
namespace comm {

namespace {

[[noreturn]] void die(const char* what, const char* detail = "") {
    std::printf("!!!\nquest-b200 multi-GPU failure: %s %s\n!!!\nexiting..\n", what, detail);
    std::fflush(stdout);
    std::exit(EXIT_FAILURE);
}

#define COMM_CUDA_OK(call)                                   \
    do {                                                     \
        cudaError_t e_ = (call);                             \
        if (e_ != cudaSuccess) die(#call, cudaGetErrorString(e_)); \
    } while (0)

// ---- TCP star --------------------------------------------------------------------------------------------
struct Star {
    int rank = 0, world = 1;
    int up = -1;              // non-root: socket to rank 0
    std::vector<int> down;    // root: socket of each rank (index = rank; [0] unused)
} star;

void sendAll(int fd, const void* buf, size_t n) {
    const char* p = static_cast<const char*>(buf);
    while (n) {
        ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
        if (k <= 0) die("bootstrap send failed");
        p += k;
        n -= (size_t)k;
    }
}

void recvAll(int fd, void* buf, size_t n) {
    char* p = static_cast<char*>(buf);
    while (n) {
        ssize_t k = ::recv(fd, p, n, 0);
        if (k <= 0) die("bootstrap recv failed (a peer process exited?)");
        p += k;
        n -= (size_t)k;
    }
}

void starConnect(const Launch& ln) {
    star.rank = ln.rank;
    star.world = ln.world;
    int one = 1;
    if (ln.rank == 0) {
        int srv = ::socket(AF_INET, SOCK_STREAM, 0);
        ::setsockopt(srv, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
        sockaddr_in addr{};
        addr.sin_family = AF_INET;
        addr.sin_addr.s_addr = htonl(INADDR_ANY);
        addr.sin_port = htons((uint16_t)ln.port);
        if (::bind(srv, reinterpret_cast<sockaddr*>(&addr), sizeof addr) != 0) die("cannot bind bootstrap port (set QUEST_B200_PORT)");
        if (::listen(srv, ln.world) != 0) die("listen failed");
        star.down.assign(ln.world, -1);
        for (int i = 1; i < ln.world; ++i) {
            int fd = ::accept(srv, nullptr, nullptr);
            if (fd < 0) die("accept failed");
            ::setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
            int peerRank = -1;
            recvAll(fd, &peerRank, sizeof peerRank);
            if (peerRank < 1 || peerRank >= ln.world || star.down[peerRank] != -1) die("bad rank in bootstrap hello");
            star.down[peerRank] = fd;
        }
        ::close(srv);
    } else {
        addrinfo hints{}, *res = nullptr;
        hints.ai_family = AF_INET;
        hints.ai_socktype = SOCK_STREAM;
        char port[16];
        std::snprintf(port, sizeof port, "%d", ln.port);
        if (::getaddrinfo(ln.masterAddr, port, &hints, &res) != 0 || !res) die("cannot resolve MASTER_ADDR", ln.masterAddr);
        int fd = -1;
        for (int attempt = 0; attempt < 1200; ++attempt) {  // up to ~120 s for rank 0 to come up
            fd = ::socket(AF_INET, SOCK_STREAM, 0);
            if (::connect(fd, res->ai_addr, res->ai_addrlen) == 0) break;
            ::close(fd);
            fd = -1;
            std::this_thread::sleep_for(std::chrono::milliseconds(100));
        }
        ::freeaddrinfo(res);
        if (fd < 0) die("cannot reach rank 0's bootstrap port");
        ::setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
        sendAll(fd, &ln.rank, sizeof ln.rank);
        star.up = fd;
    }
}

// every rank contributes n bytes; every rank receives world*n bytes ordered by rank
void allGather(const void* mine, size_t n, void* all) {
    if (star.world == 1) {
        std::memcpy(all, mine, n);
        return;
    }
    char* out = static_cast<char*>(all);
    if (star.rank == 0) {
        std::memcpy(out, mine, n);
        for (int r = 1; r < star.world; ++r) recvAll(star.down[r], out + (size_t)r * n, n);
        for (int r = 1; r < star.world; ++r) sendAll(star.down[r], out, n * star.world);
    } else {
        sendAll(star.up, mine, n);
        recvAll(star.up, out, n * star.world);
    }
}

// ---- NCCL through dlopen -----------------------------------------------------------------------------------
struct NcclId {
    char internal[128];
};
typedef void* NcclComm;
struct Nccl {
    void* lib = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    NcclComm comm = nullptr;
} nccl;
constexpr int kNcclInt8 = 0, kNcclFloat64 = 8, kNcclSum = 0;

#define NCCL_OK(call)                                                            \
    do {                                                                         \
        int r_ = (call);                                                         \
        if (r_ != 0) die(#call, nccl.GetErrorString ? nccl.GetErrorString(r_) : ""); \
    } while (0)

void loadNccl() {
    const char* env = std::getenv("QUEST_B200_NCCL_LIB");
    if (env) nccl.lib = ::dlopen(env, RTLD_NOW | RTLD_LOCAL);
    if (!nccl.lib) nccl.lib = ::dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // already in the process?
    if (!nccl.lib) nccl.lib = ::dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!nccl.lib) die("cannot load libnccl.so.2 (set QUEST_B200_NCCL_LIB):", ::dlerror());
    auto sym = [&](const char* name) {
        void* p = ::dlsym(nccl.lib, name);
        if (!p) die("libnccl lacks symbol", name);
        return p;
    };
    nccl.GetUniqueId = reinterpret_cast<decltype(nccl.GetUniqueId)>(sym("ncclGetUniqueId"));
    nccl.CommInitRank = reinterpret_cast<decltype(nccl.CommInitRank)>(sym("ncclCommInitRank"));
    nccl.CommDestroy = reinterpret_cast<decltype(nccl.CommDestroy)>(sym("ncclCommDestroy"));
    nccl.AllReduce = reinterpret_cast<decltype(nccl.AllReduce)>(sym("ncclAllReduce"));
    nccl.AllGather = reinterpret_cast<decltype(nccl.AllGather)>(sym("ncclAllGather"));
    nccl.GetErrorString = reinterpret_cast<decltype(nccl.GetErrorString)>(sym("ncclGetErrorString"));
}

// ---- per-register peer mappings ----------------------------------------------------------------------------
struct PeerMap {
    std::vector<qreal*> re, im;  // [rank]; own entry = own pointers
};
std::map<const void*, PeerMap> peers;  // key: the register's opaque record pointer (Qureg.pairStateVec.real)

cudaStream_t engineStream = nullptr;

// ---- pair handshake ----------------------------------------------------------------------------------------
// Each rank owns an array of sequence counters, one 64-byte slot per (peer, channel), mapped into every peer with
// CUDA IPC.  signal = a one-thread kernel that stores the next sequence number into MY slot of the PARTNER's array
// over NVLink; wait = a stream memory operation (cuStreamWaitValue32, >=) on the partner's slot of MY array.  Nothing
// spins on an SM and nothing needs an SM to be free, so a handshake can run while a compute kernel fills the GPU.
// Channels keep handshakes issued on different streams from being mistaken for one another.
constexpr int kChannels = 2, kSlotWords = 16;
struct Flags {
    unsigned int* mine = nullptr;
    std::vector<unsigned int*> peer;   // [rank]
    std::vector<unsigned int> seq;     // [rank * kChannels + channel]: handshakes issued so far
    CUresult (*waitValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int) = nullptr;
    unsigned int waitFlags = 0;
} flags;

__global__ void k_signal(unsigned int* slot, unsigned int value) {
    __threadfence_system();  // everything this stream did before (stream order) is visible before the counter moves
    *reinterpret_cast<volatile unsigned int*>(slot) = value;
}

void setupFlags() {
    const size_t bytes = (size_t)star.world * kChannels * kSlotWords * sizeof(unsigned int);
    COMM_CUDA_OK(cudaMalloc(&flags.mine, bytes));
    COMM_CUDA_OK(cudaMemset(flags.mine, 0, bytes));
    COMM_CUDA_OK(cudaDeviceSynchronize());
    cudaIpcMemHandle_t h;
    COMM_CUDA_OK(cudaIpcGetMemHandle(&h, flags.mine));
    std::vector<cudaIpcMemHandle_t> all(star.world);
    allGather(&h, sizeof h, all.data());
    flags.peer.assign(star.world, nullptr);
    for (int r = 0; r < star.world; ++r) {
        if (r == star.rank) {
            flags.peer[r] = flags.mine;
            continue;
        }
        void* p = nullptr;
        COMM_CUDA_OK(cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess));
        flags.peer[r] = static_cast<unsigned int*>(p);
    }
    flags.seq.assign((size_t)star.world * kChannels, 0u);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult found;
    COMM_CUDA_OK(cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &found));
    if (!fn || found != cudaDriverEntryPointSuccess) die("the CUDA driver lacks cuStreamWaitValue32");
    flags.waitValue32 = reinterpret_cast<decltype(flags.waitValue32)>(fn);
    int dev = 0, canFlush = 0;
    COMM_CUDA_OK(cudaGetDevice(&dev));
    if (cudaDeviceGetAttribute(&canFlush, cudaDevAttrCanFlushRemoteWrites, dev) != cudaSuccess) {
        cudaGetLastError();
        canFlush = 0;
    }
    flags.waitFlags = CU_STREAM_WAIT_VALUE_GEQ | (canFlush ? CU_STREAM_WAIT_VALUE_FLUSH : 0u);
}

void teardownFlags() {
    for (int r = 0; r < (int)flags.peer.size(); ++r)
        if (r != star.rank && flags.peer[r]) cudaIpcCloseMemHandle(flags.peer[r]);
    hostBarrier();  // peers have dropped their mappings
    if (flags.mine) cudaFree(flags.mine);
    flags = Flags();
}

}  // namespace

// ---- launcher environment --------------------------------------------------------------------------------------

static int envInt(const char* a, const char* b, int dflt) {
    const char* s = std::getenv(a);
    if (!s && b) s = std::getenv(b);
    return s ? std::atoi(s) : dflt;
}

Launch readLaunchEnv() {
    Launch ln;
    ln.rank = envInt("QUEST_B200_RANK", "RANK", 0);
    ln.world = envInt("QUEST_B200_WORLD", "WORLD_SIZE", 1);
    ln.device = envInt("QUEST_B200_DEVICE", "LOCAL_RANK", 0);
    const char* addr = std::getenv("MASTER_ADDR");
    std::snprintf(ln.masterAddr, sizeof ln.masterAddr, "%s", addr ? addr : "127.0.0.1");
    ln.port = envInt("QUEST_B200_PORT", nullptr, envInt("MASTER_PORT", nullptr, 29500) + 17);
    if (ln.world < 1 || (ln.world & (ln.world - 1)) != 0) die("WORLD_SIZE must be a power of two");
    if (ln.rank < 0 || ln.rank >= ln.world) die("RANK out of range");
    return ln;
}

void init(const Launch& ln, int device, cudaStream_t stream) {
    (void)device;
    engineStream = stream;
    starConnect(ln);
    loadNccl();
    NcclId id;
    std::memset(&id, 0, sizeof id);
    if (ln.rank == 0) NCCL_OK(nccl.GetUniqueId(&id));
    broadcastBytes(&id, sizeof id, 0);
    NCCL_OK(nccl.CommInitRank(&nccl.comm, ln.world, id, ln.rank));
    setupFlags();
    hostBarrier();
}

void shutdown() {
    hostBarrier();
    if (nccl.comm) nccl.CommDestroy(nccl.comm);
    nccl.comm = nullptr;
    teardownFlags();
    if (star.up >= 0) ::close(star.up);
    for (int fd : star.down)
        if (fd >= 0) ::close(fd);
    star = Star();
}

void hostBarrier() {
    if (star.world == 1) return;
    char mine = 1;
    std::vector<char> all(star.world);
    allGather(&mine, 1, all.data());
}

int allAnd(int flag) {
    std::vector<int> all(star.world);
    allGather(&flag, sizeof flag, all.data());
    int out = 1;
    for (int v : all) out = out && v;
    return out;
}

void broadcastBytes(void* buf, size_t n, int root) {
    if (star.world == 1) return;
    std::vector<char> all(n * star.world);
    allGather(buf, n, all.data());
    std::memcpy(buf, all.data() + (size_t)root * n, n);
}

void allReduceSum(double* devBuf, int n, cudaStream_t stream) {
    NCCL_OK(nccl.AllReduce(devBuf, devBuf, (size_t)n, kNcclFloat64, kNcclSum, nccl.comm, stream));
}

// every rank's `bytesPerRank` bytes, concatenated in rank order, into devOut on every rank (stream-ordered)
void allGatherBytes(const void* devIn, void* devOut, size_t bytesPerRank, cudaStream_t stream) {
    NCCL_OK(nccl.AllGather(devIn, devOut, bytesPerRank, kNcclInt8, nccl.comm, stream));
}

// Exchange CUDA IPC handles of this register's shard and map every peer's shard into this process.
void registerShard(Qureg& q, void* re, void* im, long long n) {
    (void)n;
    struct Handles {
        cudaIpcMemHandle_t re, im;
    } mine;
    COMM_CUDA_OK(cudaIpcGetMemHandle(&mine.re, re));
    COMM_CUDA_OK(cudaIpcGetMemHandle(&mine.im, im));
    std::vector<Handles> all(star.world);
    allGather(&mine, sizeof mine, all.data());
    PeerMap pm;
    pm.re.assign(star.world, nullptr);
    pm.im.assign(star.world, nullptr);
    for (int r = 0; r < star.world; ++r) {
        if (r == star.rank) {
            pm.re[r] = static_cast<qreal*>(re);
            pm.im[r] = static_cast<qreal*>(im);
            continue;
        }
        void *pr = nullptr, *pi = nullptr;
        COMM_CUDA_OK(cudaIpcOpenMemHandle(&pr, all[r].re, cudaIpcMemLazyEnablePeerAccess));
        COMM_CUDA_OK(cudaIpcOpenMemHandle(&pi, all[r].im, cudaIpcMemLazyEnablePeerAccess));
        pm.re[r] = static_cast<qreal*>(pr);
        pm.im[r] = static_cast<qreal*>(pi);
    }
    peers[q.pairStateVec.real] = pm;
    hostBarrier();
}

// Peers must drop their mappings before the owner frees the allocation.
void unregisterShard(const Qureg& q) {
    hostBarrier();  // every rank has drained its stream (caller synchronised) and arrived here
    auto it = peers.find(q.pairStateVec.real);
    if (it != peers.end()) {
        for (int r = 0; r < star.world; ++r) {
            if (r == star.rank) continue;
            cudaIpcCloseMemHandle(it->second.re[r]);
            cudaIpcCloseMemHandle(it->second.im[r]);
        }
        peers.erase(it);
    }
    hostBarrier();
}

void unsupported(const char* what) {
    std::printf("!!!\nquest-b200: %s is not implemented\n!!!\nexiting..\n", what);
    std::fflush(stdout);
    std::exit(EXIT_FAILURE);
}

Partner partnerOf(const Qureg& q, int gbit) {
    auto it = peers.find(q.pairStateVec.real);
    if (it == peers.end()) die("cross-shard gate on a register without peer mappings");
    Partner p;
    p.rank = q.chunkId ^ (1 << gbit);
    p.re = it->second.re[p.rank];
    p.im = it->second.im[p.rank];
    return p;
}

Partner shardOf(const Qureg& q, int rank) {
    auto it = peers.find(q.pairStateVec.real);
    if (it == peers.end()) die("distributed pass on a register without peer mappings");
    Partner p;
    p.rank = rank;
    p.re = it->second.re[rank];
    p.im = it->second.im[rank];
    return p;
}

// Stream-ordered handshake with one partner: work enqueued on `stream` after this call starts only after the partner's
// stream has reached its matching call (same partner pair, same channel), i.e. after everything the partner enqueued
// before it has completed and is visible here.
void pairBarrier(int partnerRank, cudaStream_t stream, int channel) {
    const unsigned int seq = ++flags.seq[(size_t)partnerRank * kChannels + channel];
    unsigned int* theirs = flags.peer[partnerRank] + ((size_t)star.rank * kChannels + channel) * kSlotWords;
    unsigned int* ours = flags.mine + ((size_t)partnerRank * kChannels + channel) * kSlotWords;
    k_signal<<<1, 1, 0, stream>>>(theirs, seq);
    COMM_CUDA_OK(cudaGetLastError());
    const CUresult r = flags.waitValue32(reinterpret_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(ours), seq,
                                         flags.waitFlags);
    if (r != CUDA_SUCCESS) die("cuStreamWaitValue32 failed in the pair handshake");
}

}  // namespace comm
}  // namespace qb
