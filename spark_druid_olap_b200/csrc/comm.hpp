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
enum class NcclType : int { U8 = 1, I64 = 4, F64 = 8 };
enum class NcclOp : int { Sum = 0, Max = 2, Min = 3 };
void nccl_reduce(void* comm, const void* send, void* recv, size_t count, NcclType type, NcclOp op, int root, cudaStream_t stream);
// point-to-point pieces of the hash-partitioned exchange; bracket a round of them with nccl_group_start / nccl_group_end
void nccl_send_bytes(void* comm, const void* send, size_t bytes, int peer, cudaStream_t stream);
void nccl_recv_bytes(void* comm, void* recv, size_t bytes, int peer, cudaStream_t stream);
void nccl_group_start();
void nccl_group_end();

}  // namespace sdo
