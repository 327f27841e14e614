// quest-b200 bulk state I/O: see qb_io.cu.
#pragma once

#include <cstddef>

#include <cuda_runtime.h>

namespace qb {
namespace io {

// Copy `bytes` between a host array and device memory.  Pinned or small: enqueued on `engineStream` (the caller
// synchronises).  Large and pageable: drains `engineStream`, then moves the data through pinned staging buffers on
// several worker threads and returns when it is in place.
void upload(void* dev, const void* host, size_t bytes, cudaStream_t engineStream);
void download(void* host, const void* dev, size_t bytes, cudaStream_t engineStream);

unsigned long long pipelinedTransfers();  // how many transfers took the staged multi-thread path (tests)
void release();                            // frees the staging pool (destroyQuESTEnv)

}  // namespace io
}  // namespace qb
