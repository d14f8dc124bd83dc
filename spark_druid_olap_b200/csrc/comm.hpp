// NCCL, reached through dlopen (the library has no link-time dependency on it): the few entry points the sharded
// execution path calls.  One communicator per engine = per GPU = per process.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <string>

namespace sdo {

constexpr int COMM_ID_BYTES = 128;  // sizeof(ncclUniqueId)

// Loads libnccl (an already loaded copy first -- PyTorch brings its own --, then `path`, then the default search path).
// Throws std::runtime_error when it cannot.
void nccl_load(const char* path_or_null);
void nccl_unique_id(void* id_out);
void* nccl_comm_init(const void* id, int rank, int world);  // the calling thread's current device
void nccl_comm_destroy(void* comm);
void nccl_all_gather_bytes(void* comm, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t stream);

}  // namespace sdo
