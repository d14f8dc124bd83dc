// CUDA kernels of the OLAP-index hot path (sm_100a): fused scan -> filter -> group-by aggregate
// over dictionary-encoded segment columns, plus the small kernels around it (loader encode,
// partial reduction, compaction, top-k, HLL lookup tables, multi-GPU partition/merge).
//
// What these replace: the per-segment work of a Druid historical behind
// DruidClient.performQuery (/root/reference/src/main/scala/org/sparklinedata/druid/client/
// DruidClient.scala:228-285): bitmap-index filter -> cursor -> per-row dictionary-id key ->
// BufferAggregator.aggregate -> merge ("hot loop on the Druid side").
//
// Shape of the scan kernel (all strategies):
//   * persistent grid, blocks pull chunks of units (1 unit = 1024 rows = 256 threads x 4 rows)
//     from an atomic ticket; a unit never crosses a span, spans never cross segments;
//   * a thread owns 4 consecutive rows, so it loads a 1-byte id column with one 32-bit load,
//     a 2-byte one with one 64-bit load, a 4-byte one with one 128-bit load and an 8-byte
//     metric with two 128-bit loads -- one 32-byte DRAM sector per metric per thread, skipped
//     entirely when none of its 4 rows passes the filter;
//   * the filter is a postfix program over leaf masks (4 bits per thread), evaluated on a
//     bit-packed stack held in one register;
//   * the key is span.key_base + sum(id_d * stride_d);
//   * aggregation: STRAT_LANE  lane-private accumulators in shared memory (no atomics, fixed
//                              reduction order => bit-reproducible sums),
//                  STRAT_BLOCK block-shared table in shared memory, shared atomics, flushed once,
//                  STRAT_GLOBAL direct-indexed table in HBM, global atomics,
//                  STRAT_HASH  open-addressing table in HBM, atomicCAS claim + global atomics.
#include <cstdio>

#include "common.hpp"
#include "kernels.hpp"
#include "xxh64.hpp"

namespace sdo {

// ---------------------------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void load_ids4(const void* base, int width, uint32_t r0, uint32_t (&out)[4]) {
  if (width == 1) {
    uint32_t w = __ldcs((const uint32_t*)((const uint8_t*)base + r0));
    out[0] = w & 0xffu; out[1] = (w >> 8) & 0xffu; out[2] = (w >> 16) & 0xffu; out[3] = w >> 24;
  } else if (width == 2) {
    uint2 w = __ldcs((const uint2*)((const uint16_t*)base + r0));
    out[0] = w.x & 0xffffu; out[1] = w.x >> 16; out[2] = w.y & 0xffffu; out[3] = w.y >> 16;
  } else {
    uint4 w = __ldcs((const uint4*)((const uint32_t*)base + r0));
    out[0] = w.x; out[1] = w.y; out[2] = w.z; out[3] = w.w;
  }
}
__device__ __forceinline__ void load_f64x4(const void* base, uint32_t r0, double (&v)[4]) {
  const double2* p = (const double2*)((const double*)base + r0);
  double2 a = __ldcs(p), b = __ldcs(p + 1);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void load_i64x4(const void* base, uint32_t r0, long long (&v)[4]) {
  const longlong2* p = (const longlong2*)((const long long*)base + r0);
  longlong2 a = __ldcs(p), b = __ldcs(p + 1);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
// order-preserving map double <-> int64 (its own inverse)
__host__ __device__ __forceinline__ long long f64_ord(long long bits) { return bits ^ ((bits >> 63) & 0x7FFFFFFFFFFFFFFFll); }
__device__ __forceinline__ long long f64_to_ord(double d) { return f64_ord(__double_as_longlong(d)); }

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}

__device__ __forceinline__ uint32_t leaf_mask(const DevLeaf& lf, const void* const* cols, uint32_t r0) {
  uint32_t ids[4];
  load_ids4(cols[lf.col], lf.width, r0, ids);
  uint32_t m = 0;
  if (lf.kind == LEAF_RANGE) {
    const uint32_t lo = lf.lo, span = lf.hi - lf.lo;
#pragma unroll
    for (int i = 0; i < 4; i++) m |= (uint32_t)((ids[i] - lo) < span) << i;
  } else {
    const uint32_t* bits = lf.bits;
#pragma unroll
    for (int i = 0; i < 4; i++) m |= ((__ldg(bits + (ids[i] >> 5)) >> (ids[i] & 31)) & 1u) << i;
  }
  return m;
}

__device__ __forceinline__ uint32_t eval_filter(const ScanParams& p, const void* const* cols, uint32_t r0) {
  uint64_t st = 0;  // stack of 4-bit masks, top in the low nibble
  const int n = p.n_ops;
  for (int k = 0; k < n; k++) {
    const uint32_t op = p.op_code[k];
    if (op == FOP_LEAF) {
      st = (st << 4) | leaf_mask(p.leaves[p.op_arg[k]], cols, r0);
    } else if (op == FOP_AND) {
      uint64_t a = st & 0xF;
      st >>= 4;
      st = (st & ~0xFull) | (st & a);
    } else if (op == FOP_OR) {
      uint64_t a = st & 0xF;
      st >>= 4;
      st |= a;
    } else if (op == FOP_NOT) {
      st ^= 0xF;
    } else if (op == FOP_TRUE) {
      st = (st << 4) | 0xF;
    } else {
      st = st << 4;
    }
  }
  return (uint32_t)(st & 0xF);
}

// HLL++ register update: byte registers packed 4 per word; plain read first (most updates are
// no-ops once a register has warmed up), compare-and-swap on the word only when it must grow.
__device__ __forceinline__ void hll_update(uint8_t* regs, uint32_t packed) {
  const uint32_t idx = packed >> 6, rho = packed & 63u;
  if (regs[idx] >= rho) return;
  unsigned int* word = (unsigned int*)(regs + (idx & ~3u));
  const int sh = (idx & 3u) * 8;
  unsigned int old = *(volatile unsigned int*)word;
  for (;;) {
    if (((old >> sh) & 0xffu) >= rho) return;
    unsigned int nw = (old & ~(0xffu << sh)) | (rho << sh);
    unsigned int prev = atomicCAS(word, old, nw);
    if (prev == old) return;
    old = prev;
  }
}
// same update on a shared-memory sketch
__device__ __forceinline__ void hll_update_sm(unsigned int* words, uint32_t packed) {
  const uint32_t idx = packed >> 6, rho = packed & 63u;
  unsigned int* word = words + (idx >> 2);
  const int sh = (idx & 3u) * 8;
  unsigned int old = *(volatile unsigned int*)word;
  for (;;) {
    if (((old >> sh) & 0xffu) >= rho) return;
    unsigned int nw = (old & ~(0xffu << sh)) | (rho << sh);
    unsigned int prev = atomicCAS(word, old, nw);
    if (prev == old) return;
    old = prev;
  }
}
// merge `n_words` shared-memory sketch words into the HBM registers (byte-wise max) and clear them
__device__ __forceinline__ void hll_flush_sm(unsigned int* sm_words, unsigned int* global_words, uint32_t n_words) {
  for (uint32_t w = threadIdx.x; w < n_words; w += SCAN_THREADS) {
    const unsigned int mine = sm_words[w];
    if (!mine) continue;
    sm_words[w] = 0u;
    unsigned int old = *(volatile unsigned int*)(global_words + w);
    for (;;) {
      const unsigned int nw = __vmaxu4(old, mine);
      if (nw == old) break;
      const unsigned int prev = atomicCAS(global_words + w, old, nw);
      if (prev == old) break;
      old = prev;
    }
  }
}
__device__ __forceinline__ uint32_t hll_value(const DevAgg& a, uint32_t id) {
  if (a.hll_mode == HLL_LUT) return __ldg(a.hll_lut + id);
  if (a.hll_mode == HLL_DIRECT_INT) return hll_pack(xxh64_int((int32_t)id, HLL_SEED));
  return hll_pack(xxh64_long((int64_t)id, HLL_SEED));
}

// ---------------------------------------------------------------------------------------------
// accumulation policies
// ---------------------------------------------------------------------------------------------
template <int STRAT>
struct Acc;

template <>
struct Acc<STRAT_LANE> {
  typedef uint32_t key_t;
  unsigned long long* base;  // smem + tid
  int ns;
  unsigned int* hll_sm;  // shared-memory HLL sketches (HLL_IN_SMEM_*), as words
  static constexpr int SLOT_STRIDE = SCAN_THREADS;  // elements between consecutive slots of one group
  __device__ __forceinline__ unsigned long long* at(key_t g, int slot) const { return base + (size_t)(g * ns + slot) * SCAN_THREADS; }
  // a thread's own count stays far below 2^32: bump the low word only
  __device__ __forceinline__ void count1(unsigned long long* q) const { *(unsigned int*)q += 1u; }
  __device__ __forceinline__ void countn(unsigned long long* q, int c) const { *(unsigned int*)q += (unsigned int)c; }
  __device__ __forceinline__ void add_i64(unsigned long long* q, long long v) const { *q += (unsigned long long)v; }
  __device__ __forceinline__ void add_f64(unsigned long long* q, double v) const { *(double*)q += v; }
  __device__ __forceinline__ void min_i64(unsigned long long* q, long long v) const { long long c = *(long long*)q; if (v < c) *(long long*)q = v; }
  __device__ __forceinline__ void max_i64(unsigned long long* q, long long v) const { long long c = *(long long*)q; if (v > c) *(long long*)q = v; }
};
// shared-memory or global atomics; the pointer's address space is visible after inlining
struct AtomicOps {
  __device__ __forceinline__ void add_i64(unsigned long long* q, long long v) const { atomicAdd(q, (unsigned long long)v); }
  __device__ __forceinline__ void add_f64(unsigned long long* q, double v) const { atomicAdd((double*)q, v); }
  __device__ __forceinline__ void min_i64(unsigned long long* q, long long v) const { atomicMin((long long*)q, v); }
  __device__ __forceinline__ void max_i64(unsigned long long* q, long long v) const { atomicMax((long long*)q, v); }
};
template <>
struct Acc<STRAT_BLOCK> : AtomicOps {
  typedef uint32_t key_t;
  unsigned long long* base;  // smem table
  int ns;
  unsigned int* hll_sm;  // shared-memory HLL sketches (HLL_IN_SMEM_*), as words
  static constexpr int SLOT_STRIDE = 1;
  __device__ __forceinline__ unsigned long long* at(key_t g, int slot) const { return base + (size_t)g * ns + slot; }
  // one launch scans fewer than 2^32 rows (checked on the host), so a block's count fits the low
  // word: a native 32-bit shared atomic instead of the 64-bit compare-and-swap loop
  __device__ __forceinline__ void count1(unsigned long long* q) const { atomicAdd((unsigned int*)q, 1u); }
  __device__ __forceinline__ void countn(unsigned long long* q, int c) const { atomicAdd((unsigned int*)q, (unsigned int)c); }
};
template <>
struct Acc<STRAT_GLOBAL> : AtomicOps {
  typedef uint64_t key_t;
  unsigned long long* base;
  int ns;
  unsigned int* hll_sm;  // shared-memory HLL sketches (HLL_IN_SMEM_*), as words
  static constexpr int SLOT_STRIDE = 1;
  __device__ __forceinline__ unsigned long long* at(key_t g, int slot) const { return base + g * (uint64_t)ns + slot; }
  __device__ __forceinline__ void count1(unsigned long long* q) const { atomicAdd(q, 1ull); }
  __device__ __forceinline__ void countn(unsigned long long* q, int c) const { atomicAdd(q, (unsigned long long)c); }
};
template <>
struct Acc<STRAT_HASH> : AtomicOps {
  typedef uint64_t key_t;
  unsigned long long* base;
  int ns;
  unsigned int* hll_sm;  // shared-memory HLL sketches (HLL_IN_SMEM_*), as words
  static constexpr int SLOT_STRIDE = 1;
  __device__ __forceinline__ unsigned long long* at(key_t g, int slot) const { return base + g * (uint64_t)ns + slot; }
  __device__ __forceinline__ void count1(unsigned long long* q) const { atomicAdd(q, 1ull); }
  __device__ __forceinline__ void countn(unsigned long long* q, int c) const { atomicAdd(q, (unsigned long long)c); }
};

// find-or-claim the slot of `key` (linear probing).  Returns HASH_EMPTY if the table is full.
__device__ __forceinline__ uint64_t hash_claim(unsigned long long* hkeys, uint64_t mask, uint64_t key) {
  uint64_t h = mix64(key) & mask;
  for (uint64_t probes = 0; probes <= mask; probes++) {
    unsigned long long cur = *(volatile unsigned long long*)(hkeys + h);
    if (cur == key) return h;
    if (cur == HASH_EMPTY) {
      unsigned long long prev = atomicCAS(hkeys + h, HASH_EMPTY, (unsigned long long)key);
      if (prev == HASH_EMPTY || prev == key) return h;
    }
    h = (h + 1) & mask;
  }
  return HASH_EMPTY;
}

// one unit of the scan (filter -> key -> aggregate over 4096 rows)
#include "scan_unit.cuh"

// coarse index over spans lives right after the sentinel span: see plan.cpp (layout of the
// per-query device blob).  COARSE_UNITS units per coarse entry.
constexpr uint32_t COARSE_UNITS = 64;

template <int STRAT>
__global__ void __launch_bounds__(SCAN_THREADS, 4) scan_kernel(const __grid_constant__ ScanParams p, const uint32_t* __restrict__ coarse) {
  extern __shared__ __align__(16) unsigned long long smem[];
  __shared__ uint32_t s_chunk, s_span;
  const int tid = threadIdx.x;
  const int ns = p.n_slots;
  Acc<STRAT> acc;
  acc.ns = ns;
  // shared-memory HLL sketches sit behind the aggregate table
  const uint32_t hll_sm_words = p.hll_mode == HLL_IN_SMEM_ALL ? (uint32_t)p.n_keys * p.n_hll * (HLL_M / 4)
                                : (p.hll_mode == HLL_IN_SMEM_SPAN ? (uint32_t)p.n_hll * (HLL_M / 4) : 0u);
  {
    size_t table_words = STRAT == STRAT_LANE ? (size_t)p.n_keys * ns * SCAN_THREADS : (STRAT == STRAT_BLOCK ? (size_t)p.n_keys * ns : 0);
    acc.hll_sm = (unsigned int*)(smem + table_words);
    for (uint32_t w = tid; w < hll_sm_words; w += SCAN_THREADS) acc.hll_sm[w] = 0u;
  }
  uint64_t hll_key = ~0ull;  // HLL_IN_SMEM_SPAN: group whose sketch is cached (block-uniform)
  if (STRAT == STRAT_LANE) {
    const uint32_t n = (uint32_t)p.n_keys * ns;
    for (uint32_t x = 0; x < n; x++) smem[(size_t)x * SCAN_THREADS + tid] = p.slot_init[x % ns];
    acc.base = smem + tid;
  } else if (STRAT == STRAT_BLOCK) {
    const uint32_t n = (uint32_t)p.n_keys * ns;
    for (uint32_t x = tid; x < n; x += SCAN_THREADS) smem[x] = p.slot_init[x % ns];
    acc.base = smem;
  } else {
    acc.base = p.table;
  }
  __syncthreads();

  for (;;) {
    if (tid == 0) s_chunk = atomicAdd(&p.counters[0], 1u);
    __syncthreads();
    const uint32_t u0 = s_chunk * p.units_per_chunk;
    if (u0 >= p.n_units) break;
    const uint32_t u1 = min(p.n_units, u0 + p.units_per_chunk);
    {
      const uint32_t lo = coarse[u0 / COARSE_UNITS], hi = coarse[u0 / COARSE_UNITS + 1];
      for (uint32_t s = lo + tid; s <= hi && s < p.n_spans; s += SCAN_THREADS)
        if (p.spans[s].unit_begin <= u0 && u0 < p.spans[s + 1].unit_begin) s_span = s;
    }
    __syncthreads();
    uint32_t s = s_span;
    DevSpan sp = p.spans[s];
    uint32_t next_begin = p.spans[s + 1].unit_begin;
    for (uint32_t u = u0; u < u1; u++) {
      while (u >= next_begin) {
        s++;
        sp = p.spans[s];
        next_begin = p.spans[s + 1].unit_begin;
      }
      if (p.hll_mode == HLL_IN_SMEM_SPAN && sp.key_base != hll_key) {
        // the cached sketch belongs to another group: merge it into HBM, start an empty one
        __syncthreads();
        if (hll_key != ~0ull) hll_flush_sm(acc.hll_sm, (unsigned int*)(p.hll + hll_key * (uint64_t)p.n_hll * HLL_M), hll_sm_words);
        hll_key = sp.key_base;
        __syncthreads();
      }
      process_unit<STRAT>(p, sp, u - sp.unit_begin, acc);
    }
  }
  if (hll_sm_words) {
    __syncthreads();
    if (p.hll_mode == HLL_IN_SMEM_ALL) hll_flush_sm(acc.hll_sm, (unsigned int*)p.hll, hll_sm_words);
    else if (hll_key != ~0ull) hll_flush_sm(acc.hll_sm, (unsigned int*)(p.hll + hll_key * (uint64_t)p.n_hll * HLL_M), hll_sm_words);
  }

  if (STRAT == STRAT_LANE) {
    // fixed-order reduction of the 256 lane-private copies: 8 strided values per lane, then a
    // shuffle tree; one warp per (key, slot)
    __syncthreads();
    const uint32_t n = (uint32_t)p.n_keys * ns;
    const int warp = tid >> 5, lane = tid & 31;
    unsigned long long* out = p.table + (size_t)blockIdx.x * n;
    for (uint32_t x = warp; x < n; x += SCAN_THREADS / 32) {
      const unsigned long long* src = smem + (size_t)x * SCAN_THREADS;
      const int op = p.slot_op[x % ns];
      unsigned long long v = src[lane];
#pragma unroll
      for (int j = 1; j < SCAN_THREADS / 32; j++) {
        unsigned long long w = src[lane + 32 * j];
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
  } else if (STRAT == STRAT_BLOCK) {
    __syncthreads();
    const uint32_t n = (uint32_t)p.n_keys * ns;
    for (uint32_t x = tid; x < n; x += SCAN_THREADS) {
      const unsigned long long v = smem[x];
      const int sl = x % ns;
      if (v == p.slot_init[sl]) continue;
      const int op = p.slot_op[sl];
      if (op == SLOT_SUM_I64) atomicAdd(p.table + x, v);
      else if (op == SLOT_SUM_F64) atomicAdd((double*)(p.table + x), __longlong_as_double(v));
      else if (op == SLOT_MIN_I64) atomicMin((long long*)(p.table + x), (long long)v);
      else atomicMax((long long*)(p.table + x), (long long)v);
    }
  }
}

size_t scan_hll_smem_bytes(int hll_mode, uint64_t n_keys, int n_hll) {
  if (hll_mode == HLL_IN_SMEM_ALL) return (size_t)n_keys * n_hll * HLL_M;
  if (hll_mode == HLL_IN_SMEM_SPAN) return (size_t)n_hll * HLL_M;
  return 0;
}
size_t scan_smem_bytes(int strategy, uint64_t n_keys, int n_slots) {
  if (strategy == STRAT_LANE) return (size_t)n_keys * n_slots * SCAN_THREADS * 8;
  if (strategy == STRAT_BLOCK) return (size_t)n_keys * n_slots * 8;
  return 0;
}

template <int STRAT>
static int occupancy(size_t smem) {
  // static shared memory counts against the 48 KB default too, so opt in well before the edge
  if (smem > 32 * 1024) SDO_CUDA(cudaFuncSetAttribute(scan_kernel<STRAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
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
    default: nb = occupancy<STRAT_HASH>(smem_bytes); break;
  }
  if (nb < 1) nb = 1;
  return nb * sm_count;
}

void launch_scan(int strategy, const ScanParams& p, int grid, size_t smem, cudaStream_t st) {
  // the coarse span index follows the n_spans + 1 span records in the same device blob
  const uint32_t* coarse = (const uint32_t*)(p.spans + p.n_spans + 1);
  switch (strategy) {
    case STRAT_LANE: scan_kernel<STRAT_LANE><<<grid, SCAN_THREADS, smem, st>>>(p, coarse); break;
    case STRAT_BLOCK: scan_kernel<STRAT_BLOCK><<<grid, SCAN_THREADS, smem, st>>>(p, coarse); break;
    case STRAT_GLOBAL: scan_kernel<STRAT_GLOBAL><<<grid, SCAN_THREADS, smem, st>>>(p, coarse); break;
    default: scan_kernel<STRAT_HASH><<<grid, SCAN_THREADS, smem, st>>>(p, coarse); break;
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
  int grid = (int)std::min<uint64_t>((n + 255) / 256, 148 * 16);
  fill_slots_kernel<<<grid, 256, 0, st>>>(table, n, n_slots, p);
  SDO_CUDA(cudaGetLastError());
}
__global__ void fill_u64_kernel(unsigned long long* dst, uint64_t n, unsigned long long v) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) dst[i] = v;
}
void launch_fill_u64(unsigned long long* dst, uint64_t n, unsigned long long v, cudaStream_t st) {
  if (!n) return;
  int grid = (int)std::min<uint64_t>((n + 255) / 256, 148 * 16);
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
template <bool HASH>
__global__ void compact_kernel(const unsigned long long* __restrict__ hkeys, const unsigned long long* __restrict__ table, uint64_t n,
                               int ns, unsigned long long* __restrict__ out_keys, unsigned long long* __restrict__ out_slots,
                               unsigned long long* __restrict__ out_src, unsigned long long* counter) {
  __shared__ unsigned long long s_base;
  __shared__ unsigned int s_warp[8];
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool occ = false;
  if (i < n) occ = HASH ? (hkeys[i] != HASH_EMPTY) : (table[i * ns] != 0ull);
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
  out_keys[o] = HASH ? hkeys[i] : i;
  if (out_src) out_src[o] = i;
  for (int s = 0; s < ns; s++) out_slots[o * ns + s] = table[i * ns + s];
}
void launch_compact_dense(const unsigned long long* table, uint64_t n_keys, int n_slots, unsigned long long* out_keys,
                          unsigned long long* out_slots, unsigned long long* out_src, unsigned long long* counter, cudaStream_t st) {
  if (!n_keys) return;
  compact_kernel<false><<<(unsigned)((n_keys + 255) / 256), 256, 0, st>>>(nullptr, table, n_keys, n_slots, out_keys, out_slots, out_src, counter);
  SDO_CUDA(cudaGetLastError());
}
void launch_compact_hash(const unsigned long long* hkeys, const unsigned long long* table, uint64_t cap, int n_slots,
                         unsigned long long* out_keys, unsigned long long* out_slots, unsigned long long* out_src,
                         unsigned long long* counter, cudaStream_t st) {
  if (!cap) return;
  compact_kernel<true><<<(unsigned)((cap + 255) / 256), 256, 0, st>>>(hkeys, table, cap, n_slots, out_keys, out_slots, out_src, counter);
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
  int grid = (int)std::min<int64_t>((n + 255) / 256, 148 * 16);
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
  int grid = (int)std::min<int64_t>((n + 255) / 256, 148 * 16);
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
  int grid = (int)std::min<int64_t>((n + 255) / 256, 148 * 16);
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
__global__ void partition_count_kernel(const unsigned long long* __restrict__ keys, uint64_t n, int n_ranks, unsigned long long* counts) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) atomicAdd(counts + (mix64(keys[i] * 0x9E3779B97F4A7C15ull) >> 32) % (unsigned)n_ranks, 1ull);
}
__global__ void partition_scatter_kernel(const unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ slots, uint64_t n,
                                         int ns, int n_ranks, unsigned long long* cursors, unsigned long long* __restrict__ out_keys,
                                         unsigned long long* __restrict__ out_slots) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) {
    unsigned r = (unsigned)((mix64(keys[i] * 0x9E3779B97F4A7C15ull) >> 32) % (unsigned)n_ranks);
    unsigned long long o = atomicAdd(cursors + r, 1ull);
    out_keys[o] = keys[i];
    for (int s = 0; s < ns; s++) out_slots[o * ns + s] = slots[i * ns + s];
  }
}
void launch_partition_rows(const unsigned long long* keys, const unsigned long long* slots, uint64_t n, int n_slots, int n_ranks,
                           unsigned long long* out_keys, unsigned long long* out_slots, unsigned long long* counts_dev,
                           unsigned long long* cursors_dev, cudaStream_t st) {
  // counts_dev must be zeroed by the caller; cursors_dev receives the exclusive prefix (host computes it between the two launches)
  if (!n) return;
  int grid = (int)std::min<uint64_t>((n + 255) / 256, 148 * 8);
  if (!cursors_dev) {
    partition_count_kernel<<<grid, 256, 0, st>>>(keys, n, n_ranks, counts_dev);
  } else {
    partition_scatter_kernel<<<grid, 256, 0, st>>>(keys, slots, n, n_slots, n_ranks, cursors_dev, out_keys, out_slots);
  }
  SDO_CUDA(cudaGetLastError());
}
__global__ void merge_rows_kernel(const unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ slots, uint64_t n,
                                  const __grid_constant__ ScanParams p) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  const int ns = p.n_slots;
  for (; i < n; i += step) {
    uint64_t h = hash_claim(p.hkeys, p.hmask, keys[i]);
    if (h == HASH_EMPTY) { p.counters[1] = 1; continue; }
    for (int s = 0; s < ns; s++) {
      unsigned long long v = slots[i * ns + s];
      unsigned long long* q = p.table + h * ns + s;
      const int op = p.slot_op[s];
      if (op == SLOT_SUM_I64) atomicAdd(q, v);
      else if (op == SLOT_SUM_F64) atomicAdd((double*)q, __longlong_as_double(v));
      else if (op == SLOT_MIN_I64) atomicMin((long long*)q, (long long)v);
      else atomicMax((long long*)q, (long long)v);
    }
  }
}
void launch_merge_rows(const unsigned long long* keys, const unsigned long long* slots, uint64_t n, const ScanParams& p, cudaStream_t st) {
  if (!n) return;
  int grid = (int)std::min<uint64_t>((n + 255) / 256, 148 * 8);
  merge_rows_kernel<<<grid, 256, 0, st>>>(keys, slots, n, p);
  SDO_CUDA(cudaGetLastError());
}

}  // namespace sdo
