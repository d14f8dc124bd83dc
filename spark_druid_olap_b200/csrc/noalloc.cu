// sdo_groupby_sum_noalloc: a dense group-by (count + double sum per key) that never allocates.
//
// Who this is for: code that is handed a GPU it may compute on but not manage memory on -- a
// plugin called inside another runtime's memory budget (a Spark executor whose device memory is
// fully pre-reserved by a pool the plugin cannot grow), a callback running where the CUDA context
// forbids allocation, or any host that must guarantee "no cudaMalloc after start-up".  Such callers
// cannot use sdo_query_execute, which takes per-query scratch from the stream-ordered pool.
//
// This is the lane-private scan of kernels.cu (scan_kernel<STRAT_LANE> + reduce_partials) reduced to
// one id column and one metric.  Its scratch -- the per-block partial table and the block ticket --
// lives in the fixed array below instead of pool memory.  Every GPU that loads this library carries
// the array (about 0.4 MB) until the process exits.
//
// Rules for callers (also in include/sdo_engine.h):
//  * at most SDO_NOALLOC_MAX_KEYS keys and SDO_NOALLOC_MAX_ROWS rows; more is refused with SDO_INVALID;
//  * the first call on each device looks up the array's device address; make that first call
//    before any stream capture begins;
//  * the call makes `device` current only for its own duration: the calling thread's current device is put back
//    before it returns, whatever the outcome (a plugin must not change its host's device behind its back);
//  * all calls on a device share the one array, which still holds the previous call's partials:
//    the kernel rewrites every word it later reads and the ticket is zeroed on entry, but two calls
//    that overlap in time on the same GPU (different streams or threads) race on it -- serialise them.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <mutex>
#include <string>

#include "../../include/sdo_engine.h"

namespace {

constexpr int NA_THREADS = 256;
constexpr int NA_ROWS_PER_THREAD = 4;
constexpr int NA_MAX_KEYS = SDO_NOALLOC_MAX_KEYS;  // 24 keys x 2 slots x 256 lanes x 8 B = 96 KB of shared memory
constexpr int NA_MAX_BLOCKS = 1024;
constexpr int NA_SLOTS = 2;  // count, sum
// words: [0] block ticket (low 32 bits used), [32 ...] partials[block][key][slot]
constexpr size_t NA_SCRATCH_WORDS = 32 + (size_t)NA_MAX_BLOCKS * NA_MAX_KEYS * NA_SLOTS;

static __device__ __align__(256) unsigned long long g_noalloc_scratch[NA_SCRATCH_WORDS];

// `scratch` is the device address of g_noalloc_scratch, resolved on the host with
// cudaGetSymbolAddress (the bare symbol name is only a host-side handle there).
__global__ void __launch_bounds__(NA_THREADS) groupby_sum_noalloc_kernel(const uint8_t* __restrict__ keys, const double* __restrict__ vals,
                                                                         long long n_rows, int n_keys, long long* __restrict__ out_counts,
                                                                         double* __restrict__ out_sums, unsigned long long* scratch) {
  extern __shared__ __align__(16) unsigned long long sm[];  // [key][slot][lane]
  __shared__ unsigned int s_last;
  const int tid = threadIdx.x;
  const int cells = n_keys * NA_SLOTS;
  for (int x = 0; x < cells; x++) sm[(size_t)x * NA_THREADS + tid] = 0ull;
  unsigned long long* mine = sm + tid;

  // rows in groups of 4 per thread; n_rows need not be a multiple of 4 (tail handled row by row)
  const long long n4 = n_rows / 4;
  for (long long g = (long long)blockIdx.x * NA_THREADS + tid; g < n4; g += (long long)gridDim.x * NA_THREADS) {
    const unsigned int k4 = __ldcs((const unsigned int*)keys + g);
    const double2 a = __ldcs((const double2*)vals + 2 * g), b = __ldcs((const double2*)vals + 2 * g + 1);
    const double v[4] = {a.x, a.y, b.x, b.y};
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const unsigned int k = (k4 >> (8 * i)) & 0xffu;
      if (k < (unsigned int)n_keys) {
        mine[(size_t)(k * NA_SLOTS) * NA_THREADS] += 1ull;
        *(double*)&mine[(size_t)(k * NA_SLOTS + 1) * NA_THREADS] += v[i];
      }
    }
  }
  if (blockIdx.x == 0 && tid < (int)(n_rows - n4 * 4)) {
    const long long r = n4 * 4 + tid;
    const unsigned int k = keys[r];
    if (k < (unsigned int)n_keys) {
      mine[(size_t)(k * NA_SLOTS) * NA_THREADS] += 1ull;
      *(double*)&mine[(size_t)(k * NA_SLOTS + 1) * NA_THREADS] += vals[r];
    }
  }
  __syncthreads();

  // block partial: fixed-order fold of the 256 lanes, one warp per cell
  unsigned long long* partials = scratch + 32;
  const int warp = tid >> 5, lane = tid & 31;
  for (int x = warp; x < cells; x += NA_THREADS / 32) {
    const unsigned long long* src = sm + (size_t)x * NA_THREADS;
    const bool is_sum = (x & 1) != 0;
    unsigned long long v = src[lane];
    for (int j = 1; j < NA_THREADS / 32; j++) {
      unsigned long long w = src[lane + 32 * j];
      v = is_sum ? (unsigned long long)__double_as_longlong(__longlong_as_double((long long)v) + __longlong_as_double((long long)w)) : v + w;
    }
    for (int off = 16; off > 0; off >>= 1) {
      unsigned long long w = __shfl_down_sync(0xffffffffu, v, off);
      v = is_sum ? (unsigned long long)__double_as_longlong(__longlong_as_double((long long)v) + __longlong_as_double((long long)w)) : v + w;
    }
    if (lane == 0) partials[(size_t)blockIdx.x * cells + x] = v;
  }
  // last block to finish folds the partials in block order (bit-reproducible) and re-arms the ticket
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd((unsigned int*)scratch, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int x = warp; x < cells; x += NA_THREADS / 32) {
    const bool is_sum = (x & 1) != 0;
    unsigned long long v = 0ull;
    for (unsigned int b = lane; b < gridDim.x; b += 32) {
      unsigned long long w = *(volatile unsigned long long*)&partials[(size_t)b * cells + x];
      v = is_sum ? (unsigned long long)__double_as_longlong(__longlong_as_double((long long)v) + __longlong_as_double((long long)w)) : v + w;
    }
    for (int off = 16; off > 0; off >>= 1) {
      unsigned long long w = __shfl_down_sync(0xffffffffu, v, off);
      v = is_sum ? (unsigned long long)__double_as_longlong(__longlong_as_double((long long)v) + __longlong_as_double((long long)w)) : v + w;
    }
    if (lane == 0) {
      if (is_sum) out_sums[x >> 1] = __longlong_as_double((long long)v);
      else out_counts[x >> 1] = (long long)v;
    }
  }
  if (tid == 0) *(unsigned int*)scratch = 0u;
}

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_mu;
unsigned long long* g_scratch_addr[NA_MAX_DEVICES] = {nullptr};
int g_blocks_per_sm[NA_MAX_DEVICES] = {0};
int g_sms[NA_MAX_DEVICES] = {0};
}  // namespace

namespace sdo {
void set_last_error(const std::string& m);  // capi.cpp: the message sdo_last_error() returns on this thread
}

extern "C" int sdo_groupby_sum_noalloc(int device, const uint8_t* keys_dev, const double* values_dev, int64_t n_rows, int n_keys,
                                       int64_t* counts_out_dev, double* sums_out_dev, void* cuda_stream) {
  auto fail = [](int code, const std::string& m) { sdo::set_last_error(m); return code; };
  sdo::set_last_error("");
  if (n_keys < 1 || n_keys > NA_MAX_KEYS)
    return fail(SDO_INVALID, "sdo_groupby_sum_noalloc: n_keys = " + std::to_string(n_keys) + " is outside 1.." + std::to_string(NA_MAX_KEYS) +
                                 " (the fixed scratch array holds no more)");
  if (n_rows < 0 || n_rows > (int64_t)SDO_NOALLOC_MAX_ROWS)
    return fail(SDO_INVALID, "sdo_groupby_sum_noalloc: n_rows = " + std::to_string((long long)n_rows) + " is outside 0.." + std::to_string((long long)SDO_NOALLOC_MAX_ROWS));
  if (device < 0 || device >= NA_MAX_DEVICES) return fail(SDO_INVALID, "sdo_groupby_sum_noalloc: bad device ordinal");
  if (!counts_out_dev || !sums_out_dev || (n_rows > 0 && (!keys_dev || !values_dev))) return fail(SDO_INVALID, "sdo_groupby_sum_noalloc: NULL pointer");
  if (((uintptr_t)keys_dev & 3) || ((uintptr_t)values_dev & 15)) return fail(SDO_INVALID, "sdo_groupby_sum_noalloc: keys must be 4-byte and values 16-byte aligned");
  // the caller's current device is restored on every way out of this function
  struct DeviceRestore {
    int prev = -1;
    ~DeviceRestore() { if (prev >= 0) cudaSetDevice(prev); }
  } restore;
  int prev = -1;
  cudaError_t e = cudaGetDevice(&prev);
  if (e != cudaSuccess) return fail(SDO_NO_DEVICE, std::string("sdo_groupby_sum_noalloc: cudaGetDevice: ") + cudaGetErrorString(e));
  if (prev != device) {
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(SDO_NO_DEVICE, std::string("sdo_groupby_sum_noalloc: cudaSetDevice: ") + cudaGetErrorString(e));
    restore.prev = prev;
  }
  const size_t smem = (size_t)n_keys * NA_SLOTS * NA_THREADS * 8;
  {
    // once per device, with that device current: resolve the array's device address
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_addr[device]) {
      unsigned long long* p = nullptr;
      e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
      if (e != cudaSuccess || !p) return fail(SDO_ERROR, std::string("sdo_groupby_sum_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
      e = cudaFuncSetAttribute(groupby_sum_noalloc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NA_MAX_KEYS * NA_SLOTS * NA_THREADS * 8);
      if (e != cudaSuccess) return fail(SDO_ERROR, std::string("sdo_groupby_sum_noalloc: cudaFuncSetAttribute: ") + cudaGetErrorString(e));
      int sms = 0;
      e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
      if (e != cudaSuccess) return fail(SDO_ERROR, std::string("sdo_groupby_sum_noalloc: ") + cudaGetErrorString(e));
      g_sms[device] = sms;
      g_scratch_addr[device] = p;
    }
  }
  int per_sm = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, groupby_sum_noalloc_kernel, NA_THREADS, smem);
  if (e != cudaSuccess) return fail(SDO_ERROR, std::string("sdo_groupby_sum_noalloc: occupancy query: ") + cudaGetErrorString(e));
  int grid = std::max(1, per_sm) * g_sms[device];
  const long long groups = n_rows / 4;
  const long long want = (groups + NA_THREADS - 1) / NA_THREADS;
  if (grid > want) grid = (int)std::max<long long>(1, want);
  if (grid > NA_MAX_BLOCKS) grid = NA_MAX_BLOCKS;
  cudaStream_t st = (cudaStream_t)cuda_stream;
  unsigned long long* scratch = g_scratch_addr[device];
  // the ticket must start at zero even if an earlier launch was aborted half-way
  e = cudaMemsetAsync(scratch, 0, 256, st);
  if (e != cudaSuccess) return fail(SDO_ERROR, std::string("sdo_groupby_sum_noalloc: cudaMemsetAsync: ") + cudaGetErrorString(e));
  groupby_sum_noalloc_kernel<<<grid, NA_THREADS, smem, st>>>(keys_dev, values_dev, (long long)n_rows, n_keys, (long long*)counts_out_dev, sums_out_dev, scratch);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(SDO_ERROR, std::string("sdo_groupby_sum_noalloc: launch: ") + cudaGetErrorString(e));
  return SDO_OK;
}
