#include "comm.hpp"

#include <dlfcn.h>

#include <cstring>
#include <mutex>
#include <stdexcept>

namespace sdo {
namespace {

struct UniqueId { char internal[COMM_ID_BYTES]; };  // ncclUniqueId: passed by value

struct Api {
  void* lib = nullptr;
  int (*GetUniqueId)(UniqueId*) = nullptr;
  int (*CommInitRank)(void**, int, UniqueId, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  int (*Reduce)(const void*, void*, size_t, int, int, int, void*, cudaStream_t) = nullptr;
  int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
Api g_api;
std::mutex g_mu;

Api& api() {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_api.lib) throw std::runtime_error("NCCL is not loaded: call sdo_comm_unique_id / sdo_engine_comm_init first");
  return g_api;
}

void check(int rc, const char* what) {
  if (rc != 0) {
    const char* msg = g_api.GetErrorString ? g_api.GetErrorString(rc) : "?";
    throw std::runtime_error(std::string("NCCL error in ") + what + ": " + msg);
  }
}

}  // namespace

void nccl_load(const char* path) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_api.lib) return;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // the copy the host process already uses, if any
  if (!h && path && *path) h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
  if (!h) throw std::runtime_error(std::string("cannot load libnccl.so.2: ") + dlerror());
  Api a;
  a.lib = h;
  *(void**)&a.GetUniqueId = dlsym(h, "ncclGetUniqueId");
  *(void**)&a.CommInitRank = dlsym(h, "ncclCommInitRank");
  *(void**)&a.CommDestroy = dlsym(h, "ncclCommDestroy");
  *(void**)&a.AllGather = dlsym(h, "ncclAllGather");
  *(void**)&a.GetErrorString = dlsym(h, "ncclGetErrorString");
  *(void**)&a.Reduce = dlsym(h, "ncclReduce");
  *(void**)&a.Send = dlsym(h, "ncclSend");
  *(void**)&a.Recv = dlsym(h, "ncclRecv");
  *(void**)&a.GroupStart = dlsym(h, "ncclGroupStart");
  *(void**)&a.GroupEnd = dlsym(h, "ncclGroupEnd");
  if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllGather || !a.Reduce || !a.Send || !a.Recv || !a.GroupStart || !a.GroupEnd) throw std::runtime_error("libnccl.so.2 lacks the expected entry points");
  g_api = a;
}

void nccl_unique_id(void* id_out) {
  UniqueId id;
  memset(&id, 0, sizeof id);
  check(api().GetUniqueId(&id), "ncclGetUniqueId");
  memcpy(id_out, &id, sizeof id);
}

void* nccl_comm_init(const void* id_bytes, int rank, int world) {
  UniqueId id;
  memcpy(&id, id_bytes, sizeof id);
  void* comm = nullptr;
  check(api().CommInitRank(&comm, world, id, rank), "ncclCommInitRank");
  return comm;
}

void nccl_comm_destroy(void* comm) {
  if (comm && g_api.CommDestroy) g_api.CommDestroy(comm);
}

void nccl_all_gather_bytes(void* comm, const void* send, void* recv, size_t bytes, cudaStream_t stream) {
  check(api().AllGather(send, recv, bytes, /*ncclInt8*/ 0, comm, stream), "ncclAllGather");
}


// ncclDataType_t / ncclRedOp_t values (nccl.h): uint8 = 1, int64 = 4, float64 = 8; sum = 0, max = 2, min = 3
void nccl_reduce(void* comm, const void* send, void* recv, size_t count, NcclType type, NcclOp op, int root, cudaStream_t stream) {
  check(api().Reduce(send, recv, count, (int)type, (int)op, root, comm, stream), "ncclReduce");
}
void nccl_send_bytes(void* comm, const void* send, size_t bytes, int peer, cudaStream_t stream) {
  check(api().Send(send, bytes, /*ncclInt8*/ 0, peer, comm, stream), "ncclSend");
}
void nccl_recv_bytes(void* comm, void* recv, size_t bytes, int peer, cudaStream_t stream) {
  check(api().Recv(recv, bytes, /*ncclInt8*/ 0, peer, comm, stream), "ncclRecv");
}
void nccl_group_start() { check(api().GroupStart(), "ncclGroupStart"); }
void nccl_group_end() { check(api().GroupEnd(), "ncclGroupEnd"); }

}  // namespace sdo
