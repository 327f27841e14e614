// quest-b200 bulk state I/O: host <-> HBM copies of amplitude arrays.
//
// The QuEST API hands over plain host pointers (setAmps, initStateFromAmps, setDensityAmps; the read-back used by
// reportState and by bindings).  From pinned memory one cudaMemcpyAsync already runs at the PCIe rate.  From ordinary
// pageable memory - what a C program's malloc or a numpy array is - the driver stages through a small pinned
// buffer on one thread and reaches a fraction of that.  Here large pageable transfers are cut into slices, one per
// worker thread; each worker moves its slice through two pinned 4 MiB buffers on its own stream, so the host-side
// memcpy of chunk k+1 overlaps the DMA of chunk k and several cores feed the link at once.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "qb_io.h"

namespace qb {
namespace io {
namespace {

constexpr size_t kChunk = 4u << 20;      // staging buffer size
constexpr size_t kDirectBelow = 8u << 20;  // smaller transfers: one plain copy
constexpr int kMaxWorkers = 8;

#define IO_CUDA_OK(call)                                                                              \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) {                                                                      \
            std::printf("!!!\nquest-b200: CUDA error in bulk I/O: %s (%s)\n!!!\nexiting..\n", cudaGetErrorString(e_), #call); \
            std::fflush(stdout);                                                                      \
            std::exit(EXIT_FAILURE);                                                                  \
        }                                                                                             \
    } while (0)

struct Lane {
    cudaStream_t stream = nullptr;
    cudaEvent_t done[2] = {nullptr, nullptr};
    char* stage[2] = {nullptr, nullptr};
};

struct Pool {
    int device = -1;
    std::vector<Lane> lanes;
} pool;

unsigned long long pipelinedCalls = 0;

int workerCount() {
    if (const char* s = std::getenv("QUEST_B200_IO_THREADS")) return std::max(1, std::min(kMaxWorkers, std::atoi(s)));
    const unsigned hw = std::thread::hardware_concurrency();
    return (int)std::max(1u, std::min((unsigned)kMaxWorkers, hw / 2));
}

void ensurePool(int device, int workers) {
    if (pool.device == device && (int)pool.lanes.size() >= workers) return;
    release();
    pool.device = device;
    pool.lanes.resize(workers);
    for (Lane& l : pool.lanes) {
        IO_CUDA_OK(cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking));
        for (int b = 0; b < 2; ++b) {
            IO_CUDA_OK(cudaEventCreateWithFlags(&l.done[b], cudaEventDisableTiming));
            IO_CUDA_OK(cudaHostAlloc((void**)&l.stage[b], kChunk, cudaHostAllocDefault));
        }
    }
}

bool isPageable(const void* host) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, host) != cudaSuccess) {
        cudaGetLastError();  // older drivers report unregistered memory as an error
        return true;
    }
    return a.type == cudaMemoryTypeUnregistered;
}

// one worker: its slice [begin, end) of the transfer through its lane's two staging buffers
void moveSlice(int device, Lane* lane, char* dev, char* host, size_t begin, size_t end, bool toDevice) {
    IO_CUDA_OK(cudaSetDevice(device));
    int b = 0;
    size_t pendingOff[2] = {0, 0}, pendingLen[2] = {0, 0};
    for (size_t off = begin; off < end; off += kChunk, b ^= 1) {
        const size_t len = std::min(kChunk, end - off);
        // the buffer's previous DMA must have finished before it is reused
        IO_CUDA_OK(cudaEventSynchronize(lane->done[b]));
        if (toDevice) {
            std::memcpy(lane->stage[b], host + off, len);
            IO_CUDA_OK(cudaMemcpyAsync(dev + off, lane->stage[b], len, cudaMemcpyHostToDevice, lane->stream));
        } else {
            if (pendingLen[b]) std::memcpy(host + pendingOff[b], lane->stage[b], pendingLen[b]);
            IO_CUDA_OK(cudaMemcpyAsync(lane->stage[b], dev + off, len, cudaMemcpyDeviceToHost, lane->stream));
            pendingOff[b] = off;
            pendingLen[b] = len;
        }
        IO_CUDA_OK(cudaEventRecord(lane->done[b], lane->stream));
    }
    IO_CUDA_OK(cudaStreamSynchronize(lane->stream));
    if (!toDevice)
        for (int k = 0; k < 2; ++k, b ^= 1)  // drain in issue order
            if (pendingLen[b]) std::memcpy(host + pendingOff[b], lane->stage[b], pendingLen[b]);
}

void transfer(void* dev, void* host, size_t bytes, cudaStream_t engineStream, bool toDevice) {
    if (bytes == 0) return;
    if (bytes < kDirectBelow || !isPageable(host)) {
        if (toDevice)
            IO_CUDA_OK(cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, engineStream));
        else
            IO_CUDA_OK(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, engineStream));
        return;
    }
    int device = 0;
    IO_CUDA_OK(cudaGetDevice(&device));
    const int workers = workerCount();
    ensurePool(device, workers);
    // everything enqueued so far reads or writes the old contents: let it finish before the lanes touch the array
    IO_CUDA_OK(cudaStreamSynchronize(engineStream));
    const size_t nChunks = (bytes + kChunk - 1) / kChunk;
    const size_t perWorker = (nChunks + workers - 1) / workers * kChunk;
    std::vector<std::thread> threads;
    for (int w = 0; w < workers; ++w) {
        const size_t begin = (size_t)w * perWorker, end = std::min(bytes, begin + perWorker);
        if (begin >= end) break;
        threads.emplace_back(moveSlice, device, &pool.lanes[w], static_cast<char*>(dev), static_cast<char*>(host), begin,
                             end, toDevice);
    }
    for (std::thread& t : threads) t.join();  // every lane has synchronised its stream: the data is in place
    ++pipelinedCalls;
}

}  // namespace

void upload(void* dev, const void* host, size_t bytes, cudaStream_t engineStream) {
    transfer(dev, const_cast<void*>(host), bytes, engineStream, true);
}

void download(void* host, const void* dev, size_t bytes, cudaStream_t engineStream) {
    transfer(const_cast<void*>(dev), host, bytes, engineStream, false);
}

unsigned long long pipelinedTransfers() { return pipelinedCalls; }

void release() {
    for (Lane& l : pool.lanes) {
        if (l.stream) cudaStreamDestroy(l.stream);
        for (int b = 0; b < 2; ++b) {
            if (l.done[b]) cudaEventDestroy(l.done[b]);
            if (l.stage[b]) cudaFreeHost(l.stage[b]);
        }
    }
    pool.lanes.clear();
    pool.device = -1;
}

}  // namespace io
}  // namespace qb
