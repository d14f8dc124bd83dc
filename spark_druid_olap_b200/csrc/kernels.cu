// CUDA kernels of the OLAP-index hot path (sm_100a).  The fused scan -> filter -> group-by aggregate
// kernel lives in scan_kernel.cuh (shared with the run-time specialiser, jit.cpp); this file instantiates
// its ahead-of-time (interpreting) variants and holds the small kernels around it: table initialisation,
// partial reduction, compaction of occupied groups, top-k, HLL gather, loader encode / re-encode,
// multi-GPU partition / merge.
#include <algorithm>
#include <cstdio>
#include <mutex>

#include "common.hpp"
#include "kernels.hpp"
#include "scan_kernel.cuh"
#include "xxh64.hpp"

namespace sdo {

__host__ __device__ __forceinline__ long long f64_ord(long long bits) { return bits ^ ((bits >> 63) & 0x7FFFFFFFFFFFFFFFll); }

// SM count of the calling thread's current device (grid sizes of the helper kernels are multiples of it)
static int sm_count() {
  static std::mutex mu;
  static int cached[64] = {};
  int dev = 0;
  SDO_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(mu);
  if (dev < 0 || dev >= 64) return 148;
  if (!cached[dev]) SDO_CUDA(cudaDeviceGetAttribute(&cached[dev], cudaDevAttrMultiProcessorCount, dev));
  return cached[dev];
}

template <int STRAT>
__global__ void __launch_bounds__(SCAN_THREADS, 4) scan_kernel(const __grid_constant__ ScanParams p, const uint32_t* __restrict__ coarse) {
  scan_body<STRAT>(p, coarse);
}

size_t scan_hll_smem_bytes(int hll_mode, bool wide, uint64_t n_keys, int n_hll) {
  const size_t per_set = wide ? (size_t)HLL_M * 4 : (size_t)HLL_M;
  if (hll_mode == HLL_IN_SMEM_ALL) return (size_t)n_keys * n_hll * per_set;
  if (hll_mode == HLL_IN_SMEM_SPAN) return (size_t)n_hll * per_set * (SCAN_THREADS / 32);  // one cached group per warp
  return 0;
}
size_t scan_smem_bytes(int strategy, uint64_t n_keys, int n_slots, uint32_t replicas, uint32_t rep_stride) {
  if (strategy == STRAT_LANE) return (size_t)n_keys * n_slots * SCAN_THREADS * 8;
  if (strategy == STRAT_SPAN) return (size_t)n_slots * SCAN_THREADS * 8;
  if (strategy == STRAT_BLOCK) return replicas > 1 ? (size_t)replicas * rep_stride * 8 : (size_t)n_keys * n_slots * 8;
  return 0;
}

template <int STRAT>
static void opt_in_shared_memory() {
  // Opt in to the largest dynamic shared memory once per device: the attribute belongs to the function, not to a
  // launch, so setting it per query would let one host thread shrink it under another thread's pending launch.
  static std::mutex mu;
  static bool done[64] = {};
  {
    int dev = 0;
    SDO_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    if (dev >= 0 && dev < 64 && !done[dev]) {
      SDO_CUDA(cudaFuncSetAttribute(scan_kernel<STRAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 64));
      done[dev] = true;
    }
  }
}
template <int STRAT>
static int occupancy(size_t smem) {
  opt_in_shared_memory<STRAT>();
  int nb = 0;
  SDO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, scan_kernel<STRAT>, SCAN_THREADS, smem));
  return nb;
}

int scan_grid_size(int strategy, size_t smem_bytes, int sm_count) {
  int nb = 0;
  switch (strategy) {
    case STRAT_LANE: nb = occupancy<STRAT_LANE>(smem_bytes); break;
    case STRAT_BLOCK: nb = occupancy<STRAT_BLOCK>(smem_bytes); break;
    case STRAT_GLOBAL: nb = occupancy<STRAT_GLOBAL>(smem_bytes); break;
    case STRAT_SPAN: nb = occupancy<STRAT_SPAN>(smem_bytes); break;
    case STRAT_EMIT: nb = occupancy<STRAT_EMIT>(smem_bytes); break;
    default: nb = occupancy<STRAT_HASH>(smem_bytes); break;
  }
  if (nb < 1) nb = 1;
  return nb * sm_count;
}

void launch_scan(int strategy, const ScanParams& p, int grid, size_t smem, cudaStream_t st) {
  // the coarse span index follows the n_spans + 1 span records in the same device blob
  const uint32_t* coarse = (const uint32_t*)(p.spans + p.n_spans + 1);
  switch (strategy) {
    case STRAT_LANE: opt_in_shared_memory<STRAT_LANE>(); scan_kernel<STRAT_LANE><<<grid, SCAN_THREADS, smem, st>>>(p, coarse); break;
    case STRAT_BLOCK: opt_in_shared_memory<STRAT_BLOCK>(); scan_kernel<STRAT_BLOCK><<<grid, SCAN_THREADS, smem, st>>>(p, coarse); break;
    case STRAT_GLOBAL: opt_in_shared_memory<STRAT_GLOBAL>(); scan_kernel<STRAT_GLOBAL><<<grid, SCAN_THREADS, smem, st>>>(p, coarse); break;
    case STRAT_SPAN: opt_in_shared_memory<STRAT_SPAN>(); scan_kernel<STRAT_SPAN><<<grid, SCAN_THREADS, smem, st>>>(p, coarse); break;
    case STRAT_EMIT: opt_in_shared_memory<STRAT_EMIT>(); scan_kernel<STRAT_EMIT><<<grid, SCAN_THREADS, smem, st>>>(p, coarse); break;
    default: opt_in_shared_memory<STRAT_HASH>(); scan_kernel<STRAT_HASH><<<grid, SCAN_THREADS, smem, st>>>(p, coarse); break;
  }
  SDO_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------
// table initialisation / partial reduction
// ---------------------------------------------------------------------------------------------
__global__ void fill_slots_kernel(unsigned long long* table, uint64_t n, int ns, const __grid_constant__ ScanParams p) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) table[i] = p.slot_init[i % ns];
}
void launch_fill_slots(unsigned long long* table, uint64_t n_rows, int n_slots, const uint64_t*, const ScanParams& p, cudaStream_t st) {
  uint64_t n = n_rows * (uint64_t)n_slots;
  if (!n) return;
  int grid = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count() * 16);
  fill_slots_kernel<<<grid, 256, 0, st>>>(table, n, n_slots, p);
  SDO_CUDA(cudaGetLastError());
}
// empty hash table: every bucket = [HASH_EMPTY | slot_init[0..ns) | 0 ...]
__global__ void fill_hash_kernel(unsigned long long* table, uint64_t cap, uint32_t stride, int ns, const __grid_constant__ ScanParams p) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x, n = cap * stride;
  for (; i < n; i += step) {
    const uint32_t w = (uint32_t)(i & (stride - 1));  // stride is a power of two
    table[i] = w == 0 ? HASH_EMPTY : (w <= (uint32_t)ns ? p.slot_init[w - 1] : 0ull);
  }
}
void launch_fill_hash(unsigned long long* table, uint64_t cap, uint32_t stride, int n_slots, const ScanParams& p, cudaStream_t st) {
  if (!cap) return;
  int grid = (int)std::min<uint64_t>((cap * stride + 255) / 256, (uint64_t)sm_count() * 16);
  fill_hash_kernel<<<grid, 256, 0, st>>>(table, cap, stride, n_slots, p);
  SDO_CUDA(cudaGetLastError());
}
__global__ void fill_u64_kernel(unsigned long long* dst, uint64_t n, unsigned long long v) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) dst[i] = v;
}
void launch_fill_u64(unsigned long long* dst, uint64_t n, unsigned long long v, cudaStream_t st) {
  if (!n) return;
  int grid = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count() * 16);
  fill_u64_kernel<<<grid, 256, 0, st>>>(dst, n, v);
  SDO_CUDA(cudaGetLastError());
}

// out[x] = fold over blocks (fixed order) of partials[b][x]; one warp per x
__global__ void reduce_partials_kernel(const unsigned long long* __restrict__ partials, int n_blocks, uint32_t n, int ns,
                                       const __grid_constant__ ScanParams p, unsigned long long* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if ((uint32_t)warp >= n) return;
  const uint32_t x = warp;
  const int op = p.slot_op[x % ns];
  unsigned long long v = p.slot_init[x % ns];
  for (int b = lane; b < n_blocks; b += 32) {
    unsigned long long w = partials[(size_t)b * n + x];
    if (op == SLOT_SUM_I64) v += w;
    else if (op == SLOT_SUM_F64) v = __double_as_longlong(__longlong_as_double(v) + __longlong_as_double(w));
    else if (op == SLOT_MIN_I64) v = (unsigned long long)min((long long)v, (long long)w);
    else v = (unsigned long long)max((long long)v, (long long)w);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    unsigned long long w = __shfl_down_sync(0xffffffffu, v, off);
    if (op == SLOT_SUM_I64) v += w;
    else if (op == SLOT_SUM_F64) v = __double_as_longlong(__longlong_as_double(v) + __longlong_as_double(w));
    else if (op == SLOT_MIN_I64) v = (unsigned long long)min((long long)v, (long long)w);
    else v = (unsigned long long)max((long long)v, (long long)w);
  }
  if (lane == 0) out[x] = v;
}
void launch_reduce_partials(const unsigned long long* partials, int n_blocks, uint32_t n, int n_slots, const ScanParams& p,
                            unsigned long long* out, cudaStream_t st) {
  if (!n) return;
  int threads = 128;
  int grid = (int)(((uint64_t)n * 32 + threads - 1) / threads);
  reduce_partials_kernel<<<grid, threads, 0, st>>>(partials, n_blocks, n, n_slots, p, out);
  SDO_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------
// compaction of occupied groups
// ---------------------------------------------------------------------------------------------
// dense: group i = table[i * ns ..], occupied when its row count (slot 0) is non-zero;
// hash: bucket i = table[i * stride ..] = [key | slots], occupied when the key is set
template <bool HASH>
__global__ void compact_kernel(const unsigned long long* __restrict__ table, uint64_t n, int ns, uint32_t stride,
                               unsigned long long* __restrict__ out_keys, unsigned long long* __restrict__ out_slots,
                               unsigned long long* __restrict__ out_src, unsigned long long* counter, unsigned long long key_off, unsigned long long key_mul) {
  __shared__ unsigned long long s_base;
  __shared__ unsigned int s_warp[8];
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool occ = false;
  if (i < n) occ = HASH ? (table[i * stride] != HASH_EMPTY) : (table[i * ns] != 0ull);
  const unsigned int bal = __ballot_sync(0xffffffffu, occ);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) s_warp[warp] = __popc(bal);
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int tot = 0;
    for (int w = 0; w < 8; w++) { unsigned int c = s_warp[w]; s_warp[w] = tot; tot += c; }
    s_base = tot ? atomicAdd(counter, (unsigned long long)tot) : 0ull;
  }
  __syncthreads();
  if (!occ) return;
  const unsigned long long o = s_base + s_warp[warp] + __popc(bal & ((1u << lane) - 1u));
  out_keys[o] = HASH ? table[i * stride] : i * key_mul + key_off;
  if (out_src) out_src[o] = i;
  const unsigned long long* row = HASH ? table + i * stride + 1 : table + i * ns;
  for (int s = 0; s < ns; s++) out_slots[o * ns + s] = row[s];
}
void launch_compact_dense(const unsigned long long* table, uint64_t n_keys, int n_slots, unsigned long long* out_keys,
                          unsigned long long* out_slots, unsigned long long* out_src, unsigned long long* counter, cudaStream_t st, unsigned long long key_off, unsigned long long key_mul) {
  if (!n_keys) return;
  compact_kernel<false><<<(unsigned)((n_keys + 255) / 256), 256, 0, st>>>(table, n_keys, n_slots, 0, out_keys, out_slots, out_src, counter, key_off, key_mul ? key_mul : 1);
  SDO_CUDA(cudaGetLastError());
}
void launch_compact_hash(const unsigned long long* table, uint64_t cap, int n_slots, uint32_t stride,
                         unsigned long long* out_keys, unsigned long long* out_slots, unsigned long long* out_src,
                         unsigned long long* counter, cudaStream_t st) {
  if (!cap) return;
  compact_kernel<true><<<(unsigned)((cap + 255) / 256), 256, 0, st>>>(table, cap, n_slots, stride, out_keys, out_slots, out_src, counter, 0ull, 1ull);
  SDO_CUDA(cudaGetLastError());
}
__global__ void gather_hll_kernel(const uint8_t* __restrict__ hll, const unsigned long long* __restrict__ src, uint64_t n, int n_hll,
                                  uint8_t* __restrict__ out) {
  // one block per output row; copies n_hll * 512 bytes as 32-bit words
  const uint64_t row = blockIdx.x;
  if (row >= n) return;
  const uint32_t words = (uint32_t)n_hll * (HLL_M / 4);
  const uint32_t* s = (const uint32_t*)(hll + src[row] * (uint64_t)n_hll * HLL_M);
  uint32_t* d = (uint32_t*)(out + row * (uint64_t)n_hll * HLL_M);
  for (uint32_t w = threadIdx.x; w < words; w += blockDim.x) d[w] = s[w];
}
// Per sketch: sum of 2^-register and the number of zero registers, the two inputs of the HLL++ estimator
// (Spark HyperLogLogPlusPlusHelper.query).  One warp per sketch, 16 registers per lane.  Every term is a power of two
// and the partial sums of a real sketch stay inside 53 bits, so the sum is exact in any order.
__global__ void hll_stats_kernel(const uint8_t* __restrict__ regs, uint64_t n_sets, double* __restrict__ out) {
  const uint64_t set = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (set >= n_sets) return;
  const int lane = threadIdx.x & 31;
  const uint4 v = __ldg((const uint4*)(regs + set * HLL_M) + lane);
  const unsigned w[4] = {v.x, v.y, v.z, v.w};
  double z = 0.0;
  int zeros = 0;
#pragma unroll
  for (int k = 0; k < 4; k++)
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const unsigned r = (w[k] >> (8 * b)) & 63u;
      z += __longlong_as_double((long long)(1023u - r) << 52);  // 2^-r
      zeros += r == 0u;
    }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    z += __shfl_xor_sync(0xffffffffu, z, off);
    zeros += __shfl_xor_sync(0xffffffffu, zeros, off);
  }
  if (lane == 0) {
    out[set * 2] = z;
    out[set * 2 + 1] = (double)zeros;
  }
}
void launch_hll_stats(const uint8_t* regs, uint64_t n_sets, double* out, cudaStream_t st) {
  if (!n_sets) return;
  hll_stats_kernel<<<(unsigned)((n_sets + 7) / 8), 256, 0, st>>>(regs, n_sets, out);
  SDO_CUDA(cudaGetLastError());
}
void launch_gather_hll(const uint8_t* hll, const unsigned long long* src, uint64_t n, int n_hll, uint8_t* out, cudaStream_t st) {
  if (!n) return;
  gather_hll_kernel<<<(unsigned)n, 128, 0, st>>>(hll, src, n, n_hll, out);
  SDO_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------
// top-k: rounds of block-local bitonic sorts over (ord, key) pairs; ties broken by key ascending
// ---------------------------------------------------------------------------------------------
constexpr int TOPK_TILE = 2048;
struct TopkItem { unsigned long long ord, key, idx; };
__device__ __forceinline__ bool topk_before(const TopkItem& a, const TopkItem& b) {
  // "a ranks before b": larger ord first, then smaller key
  return a.ord != b.ord ? a.ord > b.ord : a.key < b.key;
}
// round 0 reads the aggregate rows; later rounds read candidate indices from `in_idx`
__global__ void __launch_bounds__(1024) topk_round_kernel(const unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ slots,
                                                          int ns, int slot, int is_f64, int desc, const unsigned long long* __restrict__ in_idx,
                                                          uint64_t n, int k, unsigned long long* __restrict__ out_idx) {
  __shared__ TopkItem it[TOPK_TILE];
  const uint64_t base = (uint64_t)blockIdx.x * TOPK_TILE;
  for (int t = threadIdx.x; t < TOPK_TILE; t += blockDim.x) {
    uint64_t j = base + t;
    TopkItem x;
    unsigned long long idx = ~0ull;
    if (j < n) idx = in_idx ? in_idx[j] : j;
    if (idx != ~0ull) {
      unsigned long long raw = slots[idx * ns + slot];
      long long o = is_f64 ? f64_ord((long long)raw) : (long long)raw;
      unsigned long long u = (unsigned long long)o ^ 0x8000000000000000ull;  // signed -> unsigned order
      x.ord = desc ? u : ~u;
      x.key = keys[idx];
      x.idx = idx;
    } else {
      x.ord = 0; x.key = ~0ull; x.idx = ~0ull;  // padding: ranks last
    }
    it[t] = x;
  }
  __syncthreads();
  for (int size = 2; size <= TOPK_TILE; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < TOPK_TILE / 2; t += blockDim.x) {
        int lo = 2 * t - (t & (stride - 1));
        int hi = lo + stride;
        bool up = ((lo & size) == 0);
        TopkItem a = it[lo], b = it[hi];
        bool pad_a = a.idx == ~0ull, pad_b = b.idx == ~0ull;
        bool a_first = pad_b ? true : (pad_a ? false : topk_before(a, b));
        if (a_first != up) { it[lo] = b; it[hi] = a; }
      }
      __syncthreads();
    }
  }
  // padding sorts last, so the first k entries are the tile's best (or ~0 when it holds fewer)
  for (int t = threadIdx.x; t < k; t += blockDim.x) out_idx[(uint64_t)blockIdx.x * k + t] = it[t].idx;
}
size_t topk_scratch_elems(uint64_t n, int k) { return (size_t)((n + TOPK_TILE - 1) / TOPK_TILE) * (size_t)k + 16; }
void launch_topk(const unsigned long long* keys, const unsigned long long* slots, uint64_t n, int n_slots, int slot, int slot_is_f64,
                 int descending, int k, unsigned long long* scratch_a, unsigned long long* scratch_b, unsigned long long* out_idx,
                 cudaStream_t st) {
  // k <= TOPK_TILE / 2 so that every round shrinks the candidate set
  const unsigned long long* in = nullptr;
  uint64_t cur = n;
  unsigned long long* bufs[2] = {scratch_a, scratch_b};
  int which = 0;
  for (;;) {
    uint64_t blocks = (cur + TOPK_TILE - 1) / TOPK_TILE;
    unsigned long long* out = blocks == 1 ? out_idx : bufs[which];
    topk_round_kernel<<<(unsigned)blocks, 1024, 0, st>>>(keys, slots, n_slots, slot, slot_is_f64, descending, in, cur, k, out);
    SDO_CUDA(cudaGetLastError());
    if (blocks == 1) break;
    // padded entries (~0) from short tiles sort last and are ignored by `remain` only in the
    // last tile, so later rounds treat ~0 indices as padding
    in = out;
    cur = blocks * (uint64_t)k;
    which ^= 1;
  }
}

__global__ void gather_rows_kernel(const unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ slots,
                                   const unsigned long long* __restrict__ idx, uint64_t k, int ns, unsigned long long* __restrict__ out_keys,
                                   unsigned long long* __restrict__ out_slots) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= k) return;
  unsigned long long j = idx[i];
  out_keys[i] = j == ~0ull ? ~0ull : keys[j];
  for (int s = 0; s < ns; s++) out_slots[i * ns + s] = j == ~0ull ? 0ull : slots[j * ns + s];
}
void launch_gather_rows(const unsigned long long* keys, const unsigned long long* slots, const unsigned long long* idx, uint64_t k, int ns,
                        unsigned long long* out_keys, unsigned long long* out_slots, cudaStream_t st) {
  if (!k) return;
  gather_rows_kernel<<<(unsigned)((k + 127) / 128), 128, 0, st>>>(keys, slots, idx, k, ns, out_keys, out_slots);
  SDO_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------
// loader kernels
// ---------------------------------------------------------------------------------------------
__global__ void encode_dim_kernel(const void* __restrict__ src, bool src64, const int32_t* __restrict__ perm, int64_t r0, int64_t n,
                                  const uint32_t* __restrict__ lut, void* __restrict__ dst, int w) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) {
    int64_t s = perm ? (int64_t)perm[r0 + i] : r0 + i;
    uint32_t v = src64 ? (uint32_t)((const int64_t*)src)[s] : (uint32_t)((const int32_t*)src)[s];
    if (lut) v = lut[v];
    if (w == 1) ((uint8_t*)dst)[i] = (uint8_t)v;
    else if (w == 2) ((uint16_t*)dst)[i] = (uint16_t)v;
    else ((uint32_t*)dst)[i] = v;
  }
}
void launch_encode_dim(const void* src, bool src64, const int32_t* perm, int64_t r0, int64_t n, const uint32_t* lut, void* dst, int w,
                       cudaStream_t st) {
  if (!n) return;
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 16);
  encode_dim_kernel<<<grid, 256, 0, st>>>(src, src64, perm, r0, n, lut, dst, w);
  SDO_CUDA(cudaGetLastError());
}
__global__ void reencode_kernel(const void* __restrict__ src, int sw, void* __restrict__ dst, int dw, int64_t n, const uint32_t* __restrict__ lut) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) {
    uint32_t v = sw == 1 ? ((const uint8_t*)src)[i] : (sw == 2 ? ((const uint16_t*)src)[i] : ((const uint32_t*)src)[i]);
    if (lut) v = lut[v];
    if (dw == 1) ((uint8_t*)dst)[i] = (uint8_t)v;
    else if (dw == 2) ((uint16_t*)dst)[i] = (uint16_t)v;
    else ((uint32_t*)dst)[i] = v;
  }
}
void launch_reencode(const void* src, int sw, void* dst, int dw, int64_t n, const uint32_t* lut, cudaStream_t st) {
  if (!n) return;
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 16);
  reencode_kernel<<<grid, 256, 0, st>>>(src, sw, dst, dw, n, lut);
  SDO_CUDA(cudaGetLastError());
}
__global__ void gather64_kernel(const unsigned long long* __restrict__ src, const int32_t* __restrict__ perm, int64_t r0, int64_t n,
                                unsigned long long* __restrict__ dst) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) dst[i] = src[perm ? (int64_t)perm[r0 + i] : r0 + i];
}
void launch_gather64(const void* src, const int32_t* perm, int64_t r0, int64_t n, void* dst, cudaStream_t st) {
  if (!n) return;
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 16);
  gather64_kernel<<<grid, 256, 0, st>>>((const unsigned long long*)src, perm, r0, n, (unsigned long long*)dst);
  SDO_CUDA(cudaGetLastError());
}
__global__ void hll_lut_strings_kernel(const char* __restrict__ bytes, const int64_t* __restrict__ offsets, int64_t n, uint16_t* __restrict__ lut) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  lut[i] = hll_pack(xxh64_bytes(bytes + offsets[i], (size_t)(offsets[i + 1] - offsets[i]), HLL_SEED));
}
void launch_hll_lut_strings(const char* bytes, const int64_t* offsets, int64_t n, uint16_t* lut, cudaStream_t st) {
  if (!n) return;
  hll_lut_strings_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(bytes, offsets, n, lut);
  SDO_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------
// sparse multi-GPU merge: split rows by owner rank, merge received rows into the local hash table
// ---------------------------------------------------------------------------------------------
// Owner rank of a key.  Rows are counted / placed with warp-aggregated shared-memory counters: one global atomic per
// destination per block (count) or per 1024-row tile (scatter) -- per-row global atomics on n_ranks hot addresses
// serialised at the L2 (14 ms for 18.75 M rows; 0.3 ms now).
// key_range > 0: rank r owns keys [r * key_range, (r + 1) * key_range) -- its share of the groups is then a direct-indexed table
__device__ __forceinline__ unsigned owner_rank(unsigned long long key, int n_ranks, unsigned long long key_range) {
  if (key_range) return (unsigned)min((unsigned long long)(n_ranks - 1), key / key_range);
  return shard_owner(key, (unsigned)n_ranks, 0);
}
__global__ void __launch_bounds__(256) partition_count_kernel(const unsigned long long* __restrict__ keys, uint64_t n, int n_ranks, unsigned long long key_range, unsigned long long* counts) {
  __shared__ unsigned int s_hist[64];
  if (threadIdx.x < 64) s_hist[threadIdx.x] = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n_round = (n + step - 1) / step * step;  // whole warps take every iteration (match is a warp collective)
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += step) {
    const bool valid = i < n;
    const unsigned r = valid ? owner_rank(keys[i], n_ranks, key_range) : 0xFFFFFFFFu;
    const unsigned peers = __match_any_sync(0xffffffffu, r);
    if (valid && lane == (unsigned)__ffs(peers) - 1u) atomicAdd(&s_hist[r], (unsigned)__popc(peers));
  }
  __syncthreads();
  if (threadIdx.x < (unsigned)n_ranks && s_hist[threadIdx.x]) atomicAdd(counts + threadIdx.x, (unsigned long long)s_hist[threadIdx.x]);
}
constexpr int PART_ROWS = 4;  // rows per thread per tile
__global__ void __launch_bounds__(256) partition_scatter_kernel(const unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ slots, uint64_t n,
                                                                int ns, int n_ranks, unsigned long long key_range, unsigned long long* cursors,
                                                                unsigned long long* __restrict__ out_keys, unsigned long long* __restrict__ out_slots) {
  __shared__ unsigned int s_hist[64];
  __shared__ unsigned long long s_base[64];
  const unsigned lane = threadIdx.x & 31;
  const uint64_t tile = 256ull * PART_ROWS;
  for (uint64_t t0 = (uint64_t)blockIdx.x * tile; t0 < n; t0 += (uint64_t)gridDim.x * tile) {
    if (threadIdx.x < 64) s_hist[threadIdx.x] = 0;
    __syncthreads();
    unsigned dest[PART_ROWS], off[PART_ROWS];
#pragma unroll
    for (int k = 0; k < PART_ROWS; k++) {
      const uint64_t i = t0 + (uint64_t)k * 256 + threadIdx.x;
      const bool valid = i < n;
      const unsigned r = valid ? owner_rank(keys[i], n_ranks, key_range) : 0xFFFFFFFFu;
      const unsigned peers = __match_any_sync(0xffffffffu, r);
      const int leader = __ffs(peers) - 1;
      unsigned base = 0;
      if (valid && (int)lane == leader) base = atomicAdd(&s_hist[r], (unsigned)__popc(peers));  // the tile-local run of this warp's rows for r
      base = __shfl_sync(0xffffffffu, base, leader);
      dest[k] = r;
      off[k] = base + (unsigned)__popc(peers & ((1u << lane) - 1u));
    }
    __syncthreads();
    if (threadIdx.x < (unsigned)n_ranks) s_base[threadIdx.x] = s_hist[threadIdx.x] ? atomicAdd(cursors + threadIdx.x, (unsigned long long)s_hist[threadIdx.x]) : 0ull;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < PART_ROWS; k++) {
      const uint64_t i = t0 + (uint64_t)k * 256 + threadIdx.x;
      if (i < n) {
        const unsigned long long o = s_base[dest[k]] + off[k];
        out_keys[o] = keys[i];
        for (int s = 0; s < ns; s++) out_slots[o * ns + s] = slots[i * ns + s];
      }
    }
    __syncthreads();
  }
}
void launch_partition_rows(const unsigned long long* keys, const unsigned long long* slots, uint64_t n, int n_slots, int n_ranks,
                           unsigned long long* out_keys, unsigned long long* out_slots, unsigned long long* counts_dev,
                           unsigned long long* cursors_dev, cudaStream_t st, unsigned long long key_range) {
  // counts_dev must be zeroed by the caller; cursors_dev receives the exclusive prefix (host computes it between the two launches)
  if (!n) return;
  int grid = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count() * 8);
  if (!cursors_dev) {
    partition_count_kernel<<<grid, 256, 0, st>>>(keys, n, n_ranks, key_range, counts_dev);
  } else {
    grid = (int)std::min<uint64_t>((n + 256 * PART_ROWS - 1) / (256 * PART_ROWS), (uint64_t)sm_count() * 8);
    partition_scatter_kernel<<<grid, 256, 0, st>>>(keys, slots, n, n_slots, n_ranks, key_range, cursors_dev, out_keys, out_slots);
  }
  SDO_CUDA(cudaGetLastError());
}
__global__ void merge_rows_kernel(const unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ slots, uint64_t n,
                                  const __grid_constant__ ScanParams p, int implicit_count) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  const int ns = p.n_slots, nsr = implicit_count ? ns - 1 : ns;
  for (; i < n; i += step) {
    uint64_t h = hash_claim(p.hkeys, p.hmask, p.hstride, keys[i]);
    if (h == HASH_EMPTY) { p.counters[1] = 1; continue; }
    unsigned long long* row = p.table + h * p.hstride + 1;
    if (implicit_count) atomicAdd(row, 1ull);
    for (int s = implicit_count ? 1 : 0; s < ns; s++) {
      unsigned long long v = slots[i * nsr + (implicit_count ? s - 1 : s)];
      unsigned long long* q = row + s;
      const int op = p.slot_op[s];
      if (op == SLOT_SUM_I64) atomicAdd(q, v);
      else if (op == SLOT_SUM_F64) atomicAdd((double*)q, __longlong_as_double(v));
      else if (op == SLOT_MIN_I64) atomicMin((long long*)q, (long long)v);
      else atomicMax((long long*)q, (long long)v);
    }
  }
}
// rows merged into a direct-indexed table (no key storage, no probing): row of key k = (k - key_off) / key_mul
__global__ void merge_rows_dense_kernel(const unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ slots, uint64_t n,
                                        const __grid_constant__ ScanParams p, unsigned long long key_off, unsigned long long key_mul, int implicit_count) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  const int ns = p.n_slots, nsr = implicit_count ? ns - 1 : ns;
  for (; i < n; i += step) {
    const unsigned long long d = keys[i] - key_off, g = d / key_mul;
    if (keys[i] < key_off || g * key_mul != d || g >= p.n_keys) { p.counters[1] = 1; continue; }
    unsigned long long* row = p.table + g * ns;
    if (implicit_count) atomicAdd(row, 1ull);
    for (int s = implicit_count ? 1 : 0; s < ns; s++) {
      const unsigned long long v = slots[i * nsr + (implicit_count ? s - 1 : s)];
      unsigned long long* q = row + s;
      const int op = p.slot_op[s];
      if (op == SLOT_SUM_I64) atomicAdd(q, v);
      else if (op == SLOT_SUM_F64) atomicAdd((double*)q, __longlong_as_double(v));
      else if (op == SLOT_MIN_I64) atomicMin((long long*)q, (long long)v);
      else atomicMax((long long*)q, (long long)v);
    }
  }
}
void launch_merge_rows_dense(const unsigned long long* keys, const unsigned long long* slots, uint64_t n, const ScanParams& p, unsigned long long key_off,
                             unsigned long long key_mul, bool implicit_count, cudaStream_t st) {
  if (!n) return;
  int grid = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count() * 8);
  merge_rows_dense_kernel<<<grid, 256, 0, st>>>(keys, slots, n, p, key_off, key_mul ? key_mul : 1, implicit_count ? 1 : 0);
  SDO_CUDA(cudaGetLastError());
}
void launch_merge_rows(const unsigned long long* keys, const unsigned long long* slots, uint64_t n, const ScanParams& p, cudaStream_t st, bool implicit_count) {
  if (!n) return;
  int grid = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count() * 8);
  merge_rows_kernel<<<grid, 256, 0, st>>>(keys, slots, n, p, implicit_count ? 1 : 0);
  SDO_CUDA(cudaGetLastError());
}

}  // namespace sdo

// ---------------------------------------------------------------------------------------------
// large results: rows -> result columns on the device
// ---------------------------------------------------------------------------------------------
namespace sdo {
__global__ void materialize_kernel(const unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ slots,
                                   const __grid_constant__ MatParams mp) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  for (; r < mp.n; r += step) {
    unsigned long long k = keys[r];
    if (mp.has_tk) {
      const unsigned long long tk = k / mp.tk_stride;
      k -= tk * mp.tk_stride;
      if (mp.tk_out) mp.tk_out[r] = (uint32_t)tk;
      if (mp.ts_out) mp.ts_out[r] = mp.bucket_lut[tk];
    }
    for (int d = 0; d < mp.n_dims; d++) {
      const unsigned long long id = (k / mp.stride[d]) % mp.card[d];
      if (mp.dim_is64[d]) ((long long*)mp.dim_out[d])[r] = (long long)id;
      else ((uint32_t*)mp.dim_out[d])[r] = (uint32_t)id;
    }
    for (int a = 0; a < mp.n_aggs; a++) {
      const unsigned long long raw = slots[r * mp.ns + mp.agg_slot[a]];
      switch (mp.agg_mode[a]) {
        case MAT_I64: ((long long*)mp.agg_out[a])[r] = (long long)raw; break;
        case MAT_F64_BITS: ((unsigned long long*)mp.agg_out[a])[r] = raw; break;
        case MAT_I64_TIMES: ((long long*)mp.agg_out[a])[r] = (long long)raw * mp.agg_mul[a]; break;
        default: {  // order-preserving int64 back to a double; untouched min / max slots become +/- infinity
          const long long o = (long long)raw;
          double v;
          if (o == 0x7FFFFFFFFFFFFFFFll) v = __longlong_as_double(0x7FF0000000000000ll);
          else if (o == (long long)0x8000000000000000ull) v = __longlong_as_double((long long)0xFFF0000000000000ull);
          else v = __longlong_as_double(f64_ord(o));
          ((double*)mp.agg_out[a])[r] = v;
        }
      }
    }
  }
}
void launch_materialize(const unsigned long long* keys, const unsigned long long* slots, const MatParams& mp, cudaStream_t st) {
  if (!mp.n) return;
  int grid = (int)std::min<uint64_t>((mp.n + 255) / 256, (uint64_t)sm_count() * 16);
  materialize_kernel<<<grid, 256, 0, st>>>(keys, slots, mp);
  SDO_CUDA(cudaGetLastError());
}
}  // namespace sdo

// ---------------------------------------------------------------------------------------------
// dense multi-GPU merge of all-gathered partial blobs
// ---------------------------------------------------------------------------------------------
namespace sdo {
__global__ void merge_gathered_kernel(const unsigned long long* __restrict__ gathered, int n_ranks, uint64_t blob_words, uint64_t table_words,
                                      uint64_t cells, int ns, const __grid_constant__ ScanParams p, unsigned long long* __restrict__ table) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  for (; i < blob_words; i += step) {
    if (i < table_words) {
      if (i >= cells) continue;  // alignment padding
      const int op = p.slot_op[i % ns];
      unsigned long long v = gathered[i];
      for (int r = 1; r < n_ranks; r++) v = slot_merge(op, v, gathered[(uint64_t)r * blob_words + i]);  // rank order: reproducible
      table[i] = v;
    } else {
      // HLL registers: byte-wise max, 8 registers per word
      unsigned long long v = gathered[i];
      for (int r = 1; r < n_ranks; r++) {
        const unsigned long long w = gathered[(uint64_t)r * blob_words + i];
        v = ((unsigned long long)__vmaxu4((unsigned)(v >> 32), (unsigned)(w >> 32)) << 32) | __vmaxu4((unsigned)v, (unsigned)w);
      }
      table[i] = v;
    }
  }
}
// dense tables are [key][slot]; a collective reduction wants one operator over a contiguous run: [slot][key] and back
__global__ void slot_transpose_kernel(const unsigned long long* __restrict__ in, unsigned long long* __restrict__ out, uint64_t n_keys, int ns, int to_slot_major) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x, n = n_keys * (uint64_t)ns;
  for (; i < n; i += step) {
    // i walks the OUTPUT so that writes are coalesced
    if (to_slot_major) { const uint64_t s = i / n_keys, k = i - s * n_keys; out[i] = in[k * ns + s]; }
    else { const uint64_t k = i / ns, s = i - k * ns; out[i] = in[s * n_keys + k]; }
  }
}
void launch_slot_transpose(const unsigned long long* in, unsigned long long* out, uint64_t n_keys, int n_slots, bool to_slot_major, cudaStream_t st) {
  const uint64_t n = n_keys * (uint64_t)n_slots;
  if (!n) return;
  int grid = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count() * 16);
  slot_transpose_kernel<<<grid, 256, 0, st>>>(in, out, n_keys, n_slots, to_slot_major ? 1 : 0);
  SDO_CUDA(cudaGetLastError());
}
void launch_merge_gathered(const unsigned long long* gathered, int n_ranks, size_t blob_bytes, size_t table_bytes, uint64_t cells, int n_slots,
                           const ScanParams& p, unsigned long long* table, cudaStream_t st) {
  const uint64_t words = blob_bytes / 8;
  if (!words) return;
  int grid = (int)std::min<uint64_t>((words + 255) / 256, (uint64_t)sm_count() * 8);
  merge_gathered_kernel<<<grid, 256, 0, st>>>(gathered, n_ranks, words, table_bytes / 8, cells, n_slots, p, table);
  SDO_CUDA(cudaGetLastError());
}
}  // namespace sdo

// ---------------------------------------------------------------------------------------------
// HLL sketch column: one packed (register index, rho) per row, built once per (segment, dimension)
// ---------------------------------------------------------------------------------------------
namespace sdo {
__global__ void hll_column_kernel(const void* __restrict__ ids, int width, int mode, const uint16_t* __restrict__ lut, int64_t n,
                                  uint16_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) {
    const uint32_t id = width == 1 ? ((const uint8_t*)ids)[i] : (width == 2 ? ((const uint16_t*)ids)[i] : ((const uint32_t*)ids)[i]);
    out[i] = (uint16_t)hll_value(mode, lut, id);
  }
}
void launch_hll_column(const void* ids, int width, int mode, const uint16_t* lut, int64_t n, uint16_t* out, cudaStream_t st) {
  if (!n) return;
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)sm_count() * 16);
  hll_column_kernel<<<grid, 256, 0, st>>>(ids, width, mode, lut, n, out);
  SDO_CUDA(cudaGetLastError());
}
}  // namespace sdo

// ---------------------------------------------------------------------------------------------
// Select queries: which rows pass, in row order, and their column values
// ---------------------------------------------------------------------------------------------
namespace sdo {
namespace {

// the query's filter on one row (same postfix program as the scan's, one bit per stack entry)
__device__ __forceinline__ bool row_passes(const ScanParams& p, const void* const* cols, uint32_t r) {
  uint32_t stack = 0;  // bit 0 = top
  for (int k = 0; k < p.n_ops; k++) {
    const uint32_t op = p.op_code[k];
    if (op == FOP_LEAF) stack = (stack << 1) | leaf_test<1>(p, p.op_arg[k], cols, r);
    else if (op == FOP_TRUE) stack = (stack << 1) | 1u;
    else if (op == FOP_FALSE) stack <<= 1;
    else if (op == FOP_NOT) stack ^= 1u;
    else {
      const uint32_t a = stack & 1u;
      stack >>= 1;
      stack = op == FOP_AND ? ((stack & ~1u) | (stack & a)) : (stack | a);
    }
  }
  return p.n_ops == 0 || (stack & 1u);
}

// span of unit u (block-uniform): coarse window, then the threads probe it
__device__ __forceinline__ uint32_t span_of_unit(const ScanParams& p, const uint32_t* coarse, uint32_t u, uint32_t* s_span) {
  const uint32_t lo = coarse[u / 64], hi = coarse[u / 64 + 1];
  for (uint32_t s = lo + threadIdx.x; s <= hi && s < p.n_spans; s += blockDim.x)
    if (p.spans[s].unit_begin <= u && u < p.spans[s + 1].unit_begin) *s_span = s;
  __syncthreads();
  return *s_span;
}

}  // namespace

// counts[u] = rows of unit u that pass the filter
__global__ void __launch_bounds__(256) select_count_kernel(const __grid_constant__ ScanParams p, const uint32_t* __restrict__ coarse, uint32_t* __restrict__ counts) {
  __shared__ uint32_t s_span, s_count;
  for (uint32_t u = blockIdx.x; u < p.n_units; u += gridDim.x) {
    if (threadIdx.x == 0) s_count = 0;
    const DevSpan sp = p.spans[span_of_unit(p, coarse, u, &s_span)];
    const void* const* cols = p.segptr + (size_t)sp.seg * p.n_cols;
    const uint32_t first = (sp.row_begin & ~3u) + (u - sp.unit_begin) * UNIT_ROWS;
    uint32_t mine = 0;
    for (uint32_t r = first + threadIdx.x; r < first + UNIT_ROWS && r < sp.row_end; r += 256)
      if (r >= sp.row_begin && row_passes(p, cols, r)) mine++;
    mine = __reduce_add_sync(0xffffffffu, mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&s_count, mine);
    __syncthreads();
    if (threadIdx.x == 0) counts[u] = s_count;
    __syncthreads();
  }
}

// One block per selected unit: the passing rows with rank (in row order) in [skip, skip + take) go to out[out_base ...].
__global__ void __launch_bounds__(256) select_emit_kernel(const __grid_constant__ ScanParams p, const uint32_t* __restrict__ coarse, const SelectUnit* __restrict__ units,
                                                          uint32_t* __restrict__ out_seg, uint32_t* __restrict__ out_row) {
  __shared__ uint32_t s_span, s_warp[8];
  const SelectUnit su = units[blockIdx.x];
  const DevSpan sp = p.spans[span_of_unit(p, coarse, su.unit, &s_span)];
  const void* const* cols = p.segptr + (size_t)sp.seg * p.n_cols;
  const uint32_t first = (sp.row_begin & ~3u) + (su.unit - sp.unit_begin) * UNIT_ROWS;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t seen = 0;  // passing rows of the earlier rounds
  for (uint32_t r0 = first; r0 < first + UNIT_ROWS && r0 < sp.row_end; r0 += 256) {
    const uint32_t r = r0 + threadIdx.x;
    const bool ok = r >= sp.row_begin && r < sp.row_end && row_passes(p, cols, r);
    const unsigned b = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) s_warp[warp] = __popc(b);
    __syncthreads();
    uint32_t before = 0, total = 0;
    for (uint32_t w = 0; w < 8; w++) {
      if (w < warp) before += s_warp[w];
      total += s_warp[w];
    }
    const uint32_t rank = seen + before + __popc(b & ((1u << lane) - 1u));
    if (ok && rank >= su.skip && rank < su.skip + su.take) {
      out_seg[su.out_base + rank - su.skip] = sp.seg;
      out_row[su.out_base + rank - su.skip] = r;
    }
    seen += total;
    __syncthreads();
  }
}

// out[i] = column `col` of row (seg[i], row[i]), widened: ids -> uint32, metrics -> 8 bytes
__global__ void select_gather_kernel(const void* const* __restrict__ segptr, uint32_t n_cols, int col, int width, const uint32_t* __restrict__ seg,
                                     const uint32_t* __restrict__ row, uint32_t n, void* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const void* base = segptr[(size_t)seg[i] * n_cols + col];
  const uint32_t r = row[i];
  if (width == 8) ((unsigned long long*)out)[i] = ((const unsigned long long*)base)[r];
  else ((uint32_t*)out)[i] = width == 1 ? ((const uint8_t*)base)[r] : (width == 2 ? ((const uint16_t*)base)[r] : ((const uint32_t*)base)[r]);
}

void launch_select_count(const ScanParams& p, uint32_t* counts, cudaStream_t st) {
  if (!p.n_units) return;
  const uint32_t* coarse = (const uint32_t*)(p.spans + p.n_spans + 1);
  select_count_kernel<<<(unsigned)std::min<uint32_t>(p.n_units, (uint32_t)sm_count() * 8), 256, 0, st>>>(p, coarse, counts);
  SDO_CUDA(cudaGetLastError());
}
void launch_select_emit(const ScanParams& p, const SelectUnit* units, uint32_t n_units, uint32_t* out_seg, uint32_t* out_row, cudaStream_t st) {
  if (!n_units) return;
  const uint32_t* coarse = (const uint32_t*)(p.spans + p.n_spans + 1);
  select_emit_kernel<<<n_units, 256, 0, st>>>(p, coarse, units, out_seg, out_row);
  SDO_CUDA(cudaGetLastError());
}
void launch_select_gather(const ScanParams& p, int col, int width, const uint32_t* seg, const uint32_t* row, uint32_t n, void* out, cudaStream_t st) {
  if (!n) return;
  select_gather_kernel<<<(n + 255) / 256, 256, 0, st>>>(p.segptr, p.n_cols, col, width, seg, row, n, out);
  SDO_CUDA(cudaGetLastError());
}
}  // namespace sdo
