// Device code of the fused scan -> filter -> group-by aggregate kernel (sm_100a).
//
// What it replaces: the per-segment work of a Druid historical behind DruidClient.performQuery
// (/root/reference/src/main/scala/org/sparklinedata/druid/client/DruidClient.scala:228-285):
// bitmap-index filter -> cursor -> per-row dictionary-id key -> BufferAggregator.aggregate -> merge
// ("hot loop on the Druid side").
//
// This header is compiled two ways:
//   * ahead of time by nvcc (kernels.cu): the query shape (filter program, column widths, aggregator
//     list) is read from ScanParams at run time -- an interpreter whose dispatch is warp-uniform;
//   * at run time by NVRTC (jit.cpp) with SDO_JIT defined and the shape spelled out as constexpr
//     tables in a generated prelude: every loop over leaves / dimensions / aggregators unrolls and
//     every dispatch folds away, leaving straight-line code for that one query shape.
// Only the SHAPE is baked in; values (id ranges, bitsets, strides, pointers, key-space size) always
// come from ScanParams, so one compiled shape serves every query of that shape.
//
// Layout of the work:
//   * persistent grid; every warp draws its own tickets (512 .. 4096 rows, 16 interleaved counters) and walks them in
//     passes of 512 rows = 32 lanes x 4 vectors x 4 consecutive rows, the vectors 128 rows apart so that every load
//     instruction of a warp is fully coalesced; no block-wide barrier in the main loop.  Rows are counted in units of
//     4096; a unit never crosses a span, a span never a segment;
//   * a thread's 4 consecutive rows make every column load a natural vector: 1-byte ids -> 32-bit
//     load, 2-byte -> 64-bit, 4-byte -> 128-bit, 8-byte metric -> two 128-bit loads = one 32-byte DRAM
//     sector, skipped entirely when none of the 4 rows survives the filter;
//   * filter = postfix program over per-leaf 16-bit masks on a register stack; each leaf issues its 4
//     loads back to back before any compare;
//   * the warp looks at its densest lane: few survivors -> every lane walks its own surviving rows with
//     scalar loads (the aggregate path runs max-popcount times per warp instead of 16 times with most
//     lanes idle); otherwise vector by vector, all 32 lanes together, and a warp whose surviving rows
//     share one key folds them with shuffles into one atomic per aggregate;
//   * aggregation: LANE lane-private shared-memory accumulators (no atomics, fixed reduction order), SPAN the same for
//     the warp's current time bucket only (queries without grouping dimension), BLOCK block-shared replicated
//     shared-memory table, GLOBAL direct-indexed HBM table, HASH open addressing (key and slots in one bucket).
#pragma once

#include "kernels.hpp"
#include "xxh64.hpp"

namespace sdo {

// ---- query-shape accessors: run-time fields (AOT) or constexpr tables (JIT) ------------------------
#ifdef SDO_JIT
#define Q_UNROLL _Pragma("unroll")
#define Q_N_OPS jit::n_ops
#define Q_OP_CODE(k) jit::op_code[k]
#define Q_OP_ARG(k) jit::op_arg[k]
#define Q_LEAF_KIND(l) jit::leaf_kind[l]
#define Q_LEAF_WIDTH(l) jit::leaf_width[l]
#define Q_LEAF_COL(l) jit::leaf_col[l]
#define Q_N_DIMS jit::n_dims
#define Q_DIM_WIDTH(d) jit::dim_width[d]
#define Q_DIM_COL(d) jit::dim_col[d]
#define Q_DIM_REMAP(d) jit::dim_remap[d]
#define Q_N_AGGS jit::n_aggs
#define Q_AGG_OP(a) jit::agg_op[a]
#define Q_AGG_COL(a) jit::agg_col[a]
#define Q_AGG_SLOT(a) jit::agg_slot[a]
#define Q_AGG_LEAF(a) jit::agg_leaf[a]
#define Q_AGG_WIDTH(a) jit::agg_width[a]
#define Q_AGG_HLL_INDEX(a) jit::agg_hll_index[a]
#define Q_AGG_HLL_MODE(a) jit::agg_hll_mode[a]
#define Q_N_SLOTS jit::n_slots
#define Q_N_HLL jit::n_hll
#define Q_HLL_PLACE jit::hll_place
#define Q_HLL_WIDE jit::hll_wide
#define Q_TICKET_SHIFT jit::ticket_shift
#define Q_IGNORE_KEY_BASE jit::ignore_key_base
#define Q_N_COLS jit::n_cols
#define Q_SLOT_OP(s) jit::slot_op[s]
#define Q_EXPR_N(e) jit::expr_n[e]
#define Q_EXPR_CODE(e, k) jit::expr_code[e][k]
#define Q_EXPR_ARG(e, k) jit::expr_arg[e][k]
#else
#define Q_UNROLL
#define Q_N_OPS p.n_ops
#define Q_OP_CODE(k) p.op_code[k]
#define Q_OP_ARG(k) p.op_arg[k]
#define Q_LEAF_KIND(l) p.leaves[l].kind
#define Q_LEAF_WIDTH(l) p.leaves[l].width
#define Q_LEAF_COL(l) p.leaves[l].col
#define Q_N_DIMS p.n_dims
#define Q_DIM_WIDTH(d) p.dims[d].width
#define Q_DIM_COL(d) p.dims[d].col
#define Q_DIM_REMAP(d) (p.dims[d].remap != nullptr)
#define Q_N_AGGS p.n_aggs
#define Q_AGG_OP(a) p.aggs[a].op
#define Q_AGG_COL(a) p.aggs[a].col
#define Q_AGG_SLOT(a) p.aggs[a].slot
#define Q_AGG_LEAF(a) p.aggs[a].leaf
#define Q_AGG_WIDTH(a) p.aggs[a].width
#define Q_AGG_HLL_INDEX(a) p.aggs[a].hll_index
#define Q_AGG_HLL_MODE(a) p.aggs[a].hll_mode
#define Q_N_SLOTS p.n_slots
#define Q_N_HLL p.n_hll
#define Q_HLL_PLACE p.hll_mode
#define Q_HLL_WIDE p.hll_wide
#define Q_TICKET_SHIFT p.ticket_shift
#define Q_IGNORE_KEY_BASE p.ignore_key_base
#define Q_N_COLS p.n_cols
#define Q_SLOT_OP(s) p.slot_op[s]
#define Q_EXPR_N(e) p.exprs[e].n_ops
#define Q_EXPR_CODE(e, k) p.exprs[e].code[k]
#define Q_EXPR_ARG(e, k) p.exprs[e].arg[k]
#endif

constexpr int VECS = UNIT_ROWS / (SCAN_THREADS * ROWS_PER_THREAD);  // vectors per thread per unit
// Work distribution.  Warp tickets (default): every warp draws its own tickets (512 .. 4096 rows, see ticket_shift) and
// walks them in passes of 512 rows -- 32 lanes x 4 vectors x 4 rows, the vectors 128 rows apart -- so the main loop has
// no block-wide barrier and warps drift apart freely.  Block tickets (SDO_WARP_TICKETS 0): the block draws chunks of two
// units and all 256 threads walk a unit together, vectors 1024 rows apart.
#ifndef SDO_WARP_TICKETS
#define SDO_WARP_TICKETS 1
#endif
constexpr bool WARP_TICKETS = SDO_WARP_TICKETS != 0;
constexpr uint32_t VEC_STRIDE = (WARP_TICKETS ? 32 : SCAN_THREADS) * ROWS_PER_THREAD;  // rows between a thread's vectors
constexpr uint32_t PASS_ROWS = VEC_STRIDE * VECS;            // rows one pass of a warp (or of the block) covers
// rows per ticket = 1 << ticket_shift (512 .. 4096), chosen per query shape by the host: long tickets where a warp pays
// per ticket (flushing its cached sketch), short ones where work per row varies (selective filters, hash inserts)
constexpr uint32_t UNIT_SHIFT = 12;
static_assert((1u << UNIT_SHIFT) == UNIT_ROWS, "unit size");
constexpr uint32_t TICKET_BATCH = 8;                         // consecutive tickets that belong to one region
#ifndef SDO_DENSE_NV
#define SDO_DENSE_NV 1
#endif
constexpr int DENSE_NV = SDO_DENSE_NV;  // vectors handed to one accumulate call on the dense path (1, 2 or 4)
constexpr int SPARSE_MAX = 5;           // densest lane has at most this many survivors -> row-by-row path
constexpr uint32_t COARSE_UNITS = 64;   // units per coarse span-index entry (plan.cpp builds it)

// ---- loads --------------------------------------------------------------------------------------------
__device__ __forceinline__ void unpack(uint32_t v, uint32_t (&id)[4]) { id[0] = v & 0xffu; id[1] = (v >> 8) & 0xffu; id[2] = (v >> 16) & 0xffu; id[3] = v >> 24; }
__device__ __forceinline__ void unpack(uint2 v, uint32_t (&id)[4]) { id[0] = v.x & 0xffffu; id[1] = v.x >> 16; id[2] = v.y & 0xffffu; id[3] = v.y >> 16; }
__device__ __forceinline__ void unpack(uint4 v, uint32_t (&id)[4]) { id[0] = v.x; id[1] = v.y; id[2] = v.z; id[3] = v.w; }

template <int N>
__device__ __forceinline__ void load_ids(const void* base, int width, uint32_t r0, uint32_t (&out)[N]);
template <>
__device__ __forceinline__ void load_ids<4>(const void* base, int width, uint32_t r0, uint32_t (&out)[4]) {
  if (width == 1) unpack(__ldcs((const uint32_t*)((const uint8_t*)base + r0)), out);
  else if (width == 2) unpack(__ldcs((const uint2*)((const uint16_t*)base + r0)), out);
  else unpack(__ldcs((const uint4*)((const uint32_t*)base + r0)), out);
}
template <>
__device__ __forceinline__ void load_ids<1>(const void* base, int width, uint32_t r0, uint32_t (&out)[1]) {
  if (width == 1) out[0] = __ldg((const uint8_t*)base + r0);
  else if (width == 2) out[0] = __ldg((const uint16_t*)base + r0);
  else out[0] = __ldg((const uint32_t*)base + r0);
}
template <int N>
__device__ __forceinline__ void load_f64(const void* base, uint32_t r0, double (&v)[N]);
template <>
__device__ __forceinline__ void load_f64<4>(const void* base, uint32_t r0, double (&v)[4]) {
  const double2* q = (const double2*)((const double*)base + r0);
  double2 a = __ldcs(q), b = __ldcs(q + 1);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
template <>
__device__ __forceinline__ void load_f64<1>(const void* base, uint32_t r0, double (&v)[1]) { v[0] = __ldg((const double*)base + r0); }
template <int N>
__device__ __forceinline__ void load_i64(const void* base, uint32_t r0, long long (&v)[N]);
template <>
__device__ __forceinline__ void load_i64<4>(const void* base, uint32_t r0, long long (&v)[4]) {
  const longlong2* q = (const longlong2*)((const long long*)base + r0);
  longlong2 a = __ldcs(q), b = __ldcs(q + 1);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
template <>
__device__ __forceinline__ void load_i64<1>(const void* base, uint32_t r0, long long (&v)[1]) { v[0] = __ldg((const long long*)base + r0); }

// order-preserving map double <-> int64 (its own inverse)
__device__ __forceinline__ long long f64_to_ord(double d) {
  long long b = __double_as_longlong(d);
  return b ^ ((b >> 63) & 0x7FFFFFFFFFFFFFFFll);
}
__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}

// ---- HLL++ registers -------------------------------------------------------------------------------------
// One byte per register, 4 per word: plain read first (most updates are no-ops once a register has
// warmed up), compare-and-swap on the word only when it must grow.
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
// wide layout: one 32-bit word per register, so an update is a single native shared-memory atomicMax (no read, no
// compare-and-swap loop); costs 4x the shared memory of the packed layout
#ifndef SDO_HLL_READ_FIRST
#define SDO_HLL_READ_FIRST 0  // measured on SF100 day / month timeseries: 1.36 ms with the read, 1.26 ms without
#endif
__device__ __forceinline__ void hll_update_sm_wide(unsigned int* regs, uint32_t packed) {
  // once a register has warmed up most updates change nothing: a plain read filters them out before the atomic
  // (registers only grow, so a stale read can only cause a redundant atomic, never a lost update)
  if (SDO_HLL_READ_FIRST && *(volatile unsigned int*)(regs + (packed >> 6)) >= (packed & 63u)) return;
  atomicMax(regs + (packed >> 6), packed & 63u);
}
// merge `n_words` packed words (4 registers each) of shared-memory sketches into the HBM registers (byte-wise max) and
// clear them; WIDE: the shared-memory copy holds one word per register
// (threads `first`, `first + step`, ... of the group that owns the sketches share the words)
template <bool WIDE>
__device__ __forceinline__ void hll_flush_sm(unsigned int* sm_words, unsigned int* global_words, uint32_t n_words, uint32_t first, uint32_t step) {
  // a warp flushing its own sketch makes 4 rounds per register set: unrolled so the 4 global reads overlap
#pragma unroll 4
  for (uint32_t w = first; w < n_words; w += step) {
    unsigned int mine;
    if (WIDE) {
      uint4* q = (uint4*)sm_words + w;
      const uint4 r = *q;
      mine = r.x | (r.y << 8) | (r.z << 16) | (r.w << 24);
      if (!mine) continue;
      *q = make_uint4(0u, 0u, 0u, 0u);
    } else {
      mine = sm_words[w];
      if (!mine) continue;
      sm_words[w] = 0u;
    }
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
__device__ __forceinline__ uint32_t hll_value(int mode, const uint16_t* lut, uint32_t id) {
  if (mode == HLL_PACKED) return id;  // the column already holds idx << 6 | rho
  if (mode == HLL_LUT) return __ldg(lut + id);
  if (mode == HLL_DIRECT_INT) return hll_pack(xxh64_int((int32_t)id, HLL_SEED));
  return hll_pack(xxh64_long((int64_t)id, HLL_SEED));
}

// ---- accumulation policies -----------------------------------------------------------------------------------
template <int STRAT>
struct Acc;

template <>
struct Acc<STRAT_LANE> {
  typedef uint32_t key_t;
  unsigned long long* base;  // smem + tid
  int ns;
  unsigned int* hll_sm;  // shared-memory HLL sketches (HLL_IN_SMEM_*), as words
  static constexpr int SLOT_STRIDE = SCAN_THREADS;  // elements between consecutive slots of one group
  __device__ __forceinline__ unsigned long long* at(key_t g) const { return base + (size_t)(g * ns) * SCAN_THREADS; }
  // a thread's own count stays far below 2^32: bump the low word only
  __device__ __forceinline__ void count1(unsigned long long* q) const { *(unsigned int*)q += 1u; }
  __device__ __forceinline__ void countn(unsigned long long* q, int c) const { *(unsigned int*)q += (unsigned int)c; }
  __device__ __forceinline__ void add_i64(unsigned long long* q, long long v) const { *q += (unsigned long long)v; }
  __device__ __forceinline__ void add_f64(unsigned long long* q, double v) const { *(double*)q += v; }
  __device__ __forceinline__ void min_i64(unsigned long long* q, long long v) const { long long c = *(long long*)q; if (v < c) *(long long*)q = v; }
  __device__ __forceinline__ void max_i64(unsigned long long* q, long long v) const { long long c = *(long long*)q; if (v > c) *(long long*)q = v; }
};
template <>
struct Acc<STRAT_SPAN> {
  typedef uint32_t key_t;
  unsigned long long* base;  // smem + tid
  int ns;
  unsigned int* hll_sm;  // shared-memory HLL sketches (HLL_IN_SMEM_*), as words
  static constexpr int SLOT_STRIDE = SCAN_THREADS;  // elements between consecutive slots of one group
  // one set of slots per thread, for the warp's current key (scan_body flushes it when the key changes)
  __device__ __forceinline__ unsigned long long* at(key_t) const { return base; }
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
  unsigned int* hll_sm;
  static constexpr int SLOT_STRIDE = 1;
  __device__ __forceinline__ unsigned long long* at(key_t g) const { return base + (size_t)g * ns; }
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
  unsigned int* hll_sm;
  static constexpr int SLOT_STRIDE = 1;
  __device__ __forceinline__ unsigned long long* at(key_t g) const { return base + g * (uint64_t)ns; }
  __device__ __forceinline__ void count1(unsigned long long* q) const { atomicAdd(q, 1ull); }
  __device__ __forceinline__ void countn(unsigned long long* q, int c) const { atomicAdd(q, (unsigned long long)c); }
};
template <>
struct Acc<STRAT_HASH> : AtomicOps {
  typedef uint64_t key_t;
  unsigned long long* base;
  int ns;
  unsigned int* hll_sm;
  static constexpr int SLOT_STRIDE = 1;
  __device__ __forceinline__ unsigned long long* at(key_t g) const { return base + g * (uint64_t)ns; }
  __device__ __forceinline__ void count1(unsigned long long* q) const { atomicAdd(q, 1ull); }
  __device__ __forceinline__ void countn(unsigned long long* q, int c) const { atomicAdd(q, (unsigned long long)c); }
};

template <>
struct Acc<STRAT_EMIT> : AtomicOps {  // nothing is accumulated: accumulate() hands the rows to emit_rows()
  typedef uint64_t key_t;
  unsigned long long* base;
  int ns;
  unsigned int* hll_sm;
  static constexpr int SLOT_STRIDE = 1;
  __device__ __forceinline__ unsigned long long* at(key_t) const { return base; }
  __device__ __forceinline__ void count1(unsigned long long*) const {}
  __device__ __forceinline__ void countn(unsigned long long*, int) const {}
};

// find-or-claim the slot of `key` (linear probing).  Returns HASH_EMPTY if the table is full.
__device__ __forceinline__ uint64_t hash_claim(unsigned long long* hkeys, uint64_t mask, uint32_t stride, uint64_t key) {
  uint64_t h = mix64(key) & mask;
  for (uint64_t probes = 0; probes <= mask; probes++) {
    unsigned long long* slot = hkeys + h * stride;
    unsigned long long cur = *(volatile unsigned long long*)slot;
    if (cur == key) return h;
    if (cur == HASH_EMPTY) {
      unsigned long long prev = atomicCAS(slot, HASH_EMPTY, (unsigned long long)key);
      if (prev == HASH_EMPTY || prev == key) return h;
    }
    h = (h + 1) & mask;
  }
  return HASH_EMPTY;
}

// ---- filter ----------------------------------------------------------------------------------------------------
// mask (N bits) of rows r0 .. r0+N-1 whose id satisfies leaf `l` (aggregator-level filters)
template <int N>
__device__ __forceinline__ uint32_t leaf_test(const ScanParams& p, int l, const void* const* cols, uint32_t r0) {
  uint32_t ids[N];
  load_ids<N>(cols[Q_LEAF_COL(l)], Q_LEAF_WIDTH(l), r0, ids);
  uint32_t m = 0;
  if (Q_LEAF_KIND(l) == LEAF_RANGE) {
    const uint32_t lo = p.leaves[l].lo, span = p.leaves[l].hi - p.leaves[l].lo;
#pragma unroll
    for (int i = 0; i < N; i++) m |= (uint32_t)((ids[i] - lo) < span) << i;
  } else {
    const uint32_t* bits = p.leaves[l].bits;
#pragma unroll
    for (int i = 0; i < N; i++) m |= ((__ldg(bits + (ids[i] >> 5)) >> (ids[i] & 31)) & 1u) << i;
  }
  return m;
}

template <int W> struct IdVec;
template <> struct IdVec<1> { typedef uint32_t type; };
template <> struct IdVec<2> { typedef uint2 type; };
template <> struct IdVec<4> { typedef uint4 type; };

// 16-row leaf evaluation, specialised by id width.  Vectors outside the span (vmask bit clear) re-read
// vector 0 instead of branching; their bits are cleared by the caller's validity mask.
template <int W>
__device__ __forceinline__ uint32_t leaf_mask16_w(const ScanParams& p, int l, const void* col, uint32_t base, uint32_t vmask) {
  typedef typename IdVec<W>::type V;
  const V* q = (const V*)col + (base >> 2);
  V v[VECS];
#pragma unroll
  for (int j = 0; j < VECS; j++) v[j] = __ldcs(q + (((vmask >> j) & 1u) ? j * (int)(VEC_STRIDE / 4) : 0));
  uint32_t m = 0;
  if (Q_LEAF_KIND(l) == LEAF_RANGE) {
    const uint32_t lo = p.leaves[l].lo, span = p.leaves[l].hi - p.leaves[l].lo;
#pragma unroll
    for (int j = 0; j < VECS; j++) {
      uint32_t id[4];
      unpack(v[j], id);
#pragma unroll
      for (int i = 0; i < 4; i++) m |= (uint32_t)((id[i] - lo) < span) << (4 * j + i);
    }
  } else {
    const uint32_t* bits = p.leaves[l].bits;
#pragma unroll
    for (int j = 0; j < VECS; j++) {
      uint32_t id[4];
      unpack(v[j], id);
#pragma unroll
      for (int i = 0; i < 4; i++) m |= ((__ldg(bits + (id[i] >> 5)) >> (id[i] & 31)) & 1u) << (4 * j + i);
    }
  }
  return m;
}
__device__ __forceinline__ uint32_t leaf_mask16(const ScanParams& p, int l, const void* const* cols, uint32_t base, uint32_t vmask) {
  const void* col = cols[Q_LEAF_COL(l)];
  if (Q_LEAF_WIDTH(l) == 1) return leaf_mask16_w<1>(p, l, col, base, vmask);
  if (Q_LEAF_WIDTH(l) == 2) return leaf_mask16_w<2>(p, l, col, base, vmask);
  return leaf_mask16_w<4>(p, l, col, base, vmask);
}

// postfix filter program over 16-bit masks; stack of 8 entries in two registers (top = low 16 bits of lo)
__device__ __forceinline__ uint32_t eval_filter16(const ScanParams& p, const void* const* cols, uint32_t base, uint32_t vmask) {
  uint64_t lo = 0, hi = 0;
  Q_UNROLL
  for (int k = 0; k < Q_N_OPS; k++) {
    const uint32_t op = Q_OP_CODE(k);
    if (op == FOP_LEAF || op == FOP_TRUE || op == FOP_FALSE) {
      uint64_t v = op == FOP_LEAF ? leaf_mask16(p, Q_OP_ARG(k), cols, base, vmask) : (op == FOP_TRUE ? 0xFFFFull : 0ull);
      hi = (hi << 16) | (lo >> 48);
      lo = (lo << 16) | v;
    } else if (op == FOP_NOT) {
      lo ^= 0xFFFFull;
    } else {
      const uint64_t a = lo & 0xFFFFull;
      lo = (lo >> 16) | (hi << 48);
      hi >>= 16;
      if (op == FOP_AND) lo = (lo & ~0xFFFFull) | (lo & a);
      else lo |= a;
    }
  }
  return (uint32_t)(lo & 0xFFFFull);
}

// ---- aggregation -------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_add(double v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}
__device__ __forceinline__ long long warp_add(long long v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}
__device__ __forceinline__ long long warp_min(long long v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}
__device__ __forceinline__ long long warp_max(long long v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}

// Argument expression `e` of an AGG_*_EXPR aggregation for N consecutive rows: a postfix program over metric columns and
// constants.  Round-to-nearest intrinsics keep the compiler from contracting a * b + c into one fused operation, so each
// row's value is bit-identical to the CPU's.  With the program spelled out as constants (JIT) the loop unrolls and the
// stack lives in registers; the interpreter (AOT) indexes it dynamically.
template <int N>
__device__ __forceinline__ void eval_expr(const ScanParams& p, int e, const void* const* cols, uint32_t r0, double (&out)[N]) {
  double st[EXPR_STACK][N];
  int sp = 0;
  Q_UNROLL
  for (int k = 0; k < Q_EXPR_N(e); k++) {
    const int op = Q_EXPR_CODE(e, k), arg = Q_EXPR_ARG(e, k);
    if (op == EX_COL_F64) {
      load_f64<N>(cols[arg], r0, st[sp]);
      sp++;
    } else if (op == EX_COL_I64) {
      long long v[N];
      load_i64<N>(cols[arg], r0, v);
#pragma unroll
      for (int i = 0; i < N; i++) st[sp][i] = (double)v[i];
      sp++;
    } else if (op == EX_CONST) {
      const double c = p.exprs[e].consts[arg];
#pragma unroll
      for (int i = 0; i < N; i++) st[sp][i] = c;
      sp++;
    } else if (op == EX_NEG || op == EX_ABS || op == EX_FLOOR || op == EX_CEIL || op == EX_SQRT) {
#pragma unroll
      for (int i = 0; i < N; i++) {
        const double x = st[sp - 1][i];
        st[sp - 1][i] = op == EX_NEG ? -x : op == EX_ABS ? fabs(x) : op == EX_FLOOR ? floor(x) : op == EX_CEIL ? ceil(x) : __dsqrt_rn(x);
      }
    } else {
      sp--;
#pragma unroll
      for (int i = 0; i < N; i++) {
        const double x = st[sp - 1][i], y = st[sp][i];
        st[sp - 1][i] = op == EX_ADD ? __dadd_rn(x, y) : op == EX_SUB ? __dsub_rn(x, y) : op == EX_MUL ? __dmul_rn(x, y)
                        : op == EX_DIV ? __ddiv_rn(x, y) : op == EX_MIN ? fmin(x, y) : op == EX_MAX ? fmax(x, y) : fmod(x, y);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < N; i++) out[i] = st[0][i];
}

// Row set of one accumulate call: NV vectors of N consecutive rows each; row (v, i) = r0[v] + i.  Handing several
// vectors to one call puts all their loads in flight together and folds a uniform-key warp once for all of them.
template <int NV>
struct RowSet { uint32_t r0[NV]; };
template <int N, int NV>
__device__ __forceinline__ void load_ids_m(const void* base, int width, const RowSet<NV>& rs, uint32_t (&out)[N * NV]) {
#pragma unroll
  for (int v = 0; v < NV; v++) {
    uint32_t t[N];
    load_ids<N>(base, width, rs.r0[v], t);
#pragma unroll
    for (int i = 0; i < N; i++) out[v * N + i] = t[i];
  }
}
template <int N, int NV>
__device__ __forceinline__ void load_f64_m(const void* base, const RowSet<NV>& rs, double (&out)[N * NV]) {
#pragma unroll
  for (int v = 0; v < NV; v++) {
    double t[N];
    load_f64<N>(base, rs.r0[v], t);
#pragma unroll
    for (int i = 0; i < N; i++) out[v * N + i] = t[i];
  }
}
template <int N, int NV>
__device__ __forceinline__ void load_i64_m(const void* base, const RowSet<NV>& rs, long long (&out)[N * NV]) {
#pragma unroll
  for (int v = 0; v < NV; v++) {
    long long t[N];
    load_i64<N>(base, rs.r0[v], t);
#pragma unroll
    for (int i = 0; i < N; i++) out[v * N + i] = t[i];
  }
}
template <int N, int NV>
__device__ __forceinline__ uint32_t leaf_test_m(const ScanParams& p, int l, const void* const* cols, const RowSet<NV>& rs) {
  uint32_t m = 0;
#pragma unroll
  for (int v = 0; v < NV; v++) m |= leaf_test<N>(p, l, cols, rs.r0[v]) << (v * N);
  return m;
}
template <int N, int NV>
__device__ __forceinline__ void eval_expr_m(const ScanParams& p, int e, const void* const* cols, const RowSet<NV>& rs, double (&out)[N * NV]) {
#pragma unroll
  for (int v = 0; v < NV; v++) {
    double t[N];
    eval_expr<N>(p, e, cols, rs.r0[v], t);
#pragma unroll
    for (int i = 0; i < N; i++) out[v * N + i] = t[i];
  }
}

// STRAT_EMIT: write the rows of `r0` with a bit in m out as groups of one row each: the key and, for every aggregator,
// the value its slot would have received (its identity where the aggregator's own filter fails).  Slot 0, the row count,
// is 1 by construction and is not stored: a row is n_slots - 1 words.  With emit_owners > 1 every row goes to the region
// of the rank that owns its key.  WARP = all 32 lanes are here: cursor bumps are aggregated over the warp (per owner);
// otherwise (row-by-row path under divergence) one bump per row.
template <int N, int NV, bool WARP>
__device__ __forceinline__ void emit_rows(const ScanParams& p, const void* const* cols, const uint64_t (&key)[N * NV], const RowSet<NV>& r0, uint32_t m) {
  constexpr int R = N * NV;
  const int nsr = Q_N_SLOTS - 1;
  unsigned long long pos[R];  // row index inside the output arrays (owner region included)
  if (p.emit_owners > 1) {
    const unsigned W = (unsigned)p.emit_owners;
    const unsigned lane = threadIdx.x & 31u;
    unsigned own[R];
#pragma unroll
    for (int i = 0; i < R; i++) own[i] = ((m >> i) & 1u) ? shard_owner(key[i], W, p.emit_interleave) : 0xFFFFFFFFu;
    if (WARP && W <= 32) {
      // One cursor bump per owner for all the rows of the call (16 per lane on the specialised kernel): lane o counts owner
      // o's rows with ballots while every row learns its rank among them; W bumps per 512 rows instead of W per 32.
      unsigned rank[R];
      unsigned mine = 0;  // lane o: rows of owner o in this call
      for (unsigned o = 0; o < W; o++) {
        unsigned run = 0;
#pragma unroll
        for (int i = 0; i < R; i++) {
          const unsigned b = __ballot_sync(0xffffffffu, own[i] == o);
          if (own[i] == o) rank[i] = run + (unsigned)__popc(b & ((1u << lane) - 1u));
          run += (unsigned)__popc(b);
        }
        if (lane == o) mine = run;
      }
      unsigned long long base = 0;
      if (lane < W && mine) base = atomicAdd(p.emit_cursors + lane, (unsigned long long)mine);
#pragma unroll
      for (int i = 0; i < R; i++) {
        const unsigned long long q = __shfl_sync(0xffffffffu, base, own[i] & 31u) + rank[i];
        if (own[i] != 0xFFFFFFFFu && q >= p.emit_cap) { p.counters[1] = 1; m &= ~(1u << i); }
        pos[i] = (unsigned long long)own[i] * p.emit_cap + q;
      }
    } else {
#pragma unroll
      for (int i = 0; i < R; i++) {
        const bool on = own[i] != 0xFFFFFFFFu;
        unsigned long long q = 0;
        if (WARP) {
          const unsigned peers = __match_any_sync(0xffffffffu, own[i]);
          const int leader = __ffs(peers) - 1;
          if (on && (int)lane == leader) q = atomicAdd(p.emit_cursors + own[i], (unsigned long long)__popc(peers));
          q = __shfl_sync(0xffffffffu, q, leader);
          q += (unsigned long long)__popc(peers & ((1u << lane) - 1u));
        } else if (on) {
          q = atomicAdd(p.emit_cursors + own[i], 1ull);
        }
        if (on && q >= p.emit_cap) { p.counters[1] = 1; m &= ~(1u << i); }
        pos[i] = (unsigned long long)own[i] * p.emit_cap + q;
      }
    }
    if (!m) return;
  } else {
    const int c = __popc(m);
    unsigned long long pos0;
    if (WARP) {
      int incl = c;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, off);
        if ((int)(threadIdx.x & 31u) >= off) incl += t;
      }
      const int total = __shfl_sync(0xffffffffu, incl, 31);
      unsigned long long base = 0;
      if ((threadIdx.x & 31u) == 0 && total) base = atomicAdd(p.emit_cursors, (unsigned long long)total);
      base = __shfl_sync(0xffffffffu, base, 0);
      pos0 = base + (unsigned long long)(incl - c);
    } else {
      if (!c) return;
      pos0 = atomicAdd(p.emit_cursors, (unsigned long long)c);
    }
    if (!c) return;
    if (pos0 + (unsigned long long)c > p.emit_cap) { p.counters[1] = 1; return; }
    unsigned long long q = pos0;
#pragma unroll
    for (int i = 0; i < R; i++) { pos[i] = q; q += (m >> i) & 1u; }
  }
#pragma unroll
  for (int i = 0; i < R; i++)
    if ((m >> i) & 1u) p.emit_keys[pos[i]] = key[i];
  Q_UNROLL
  for (int a = 0; a < Q_N_AGGS; a++) {
    uint32_t ma = m;
    if (Q_AGG_LEAF(a) >= 0) ma &= leaf_test_m<N, NV>(p, Q_AGG_LEAF(a), cols, r0);
    const int slot = Q_AGG_SLOT(a);
    const unsigned long long ident = p.slot_init[slot];
    const void* col = Q_AGG_OP(a) >= AGG_SUM_EXPR ? nullptr : cols[Q_AGG_COL(a)];
    unsigned long long out[R];
    const int op = Q_AGG_OP(a);
    const bool from_f64 = op == AGG_SUM_F64 || op == AGG_MIN_F64 || op == AGG_MAX_F64 || op == AGG_SUM_I64_F64 || op == AGG_MIN_I64_F64 || op == AGG_MAX_I64_F64;
    const bool from_i64 = op == AGG_SUM_I64 || op == AGG_MIN_I64 || op == AGG_MAX_I64 || op == AGG_SUM_F64_I64 || op == AGG_MIN_F64_I64 || op == AGG_MAX_F64_I64;
    double vf[R];
    long long vi[R];
#pragma unroll
    for (int i = 0; i < R; i++) { vf[i] = 0.0; vi[i] = 0; }
    if (from_f64) load_f64_m<N, NV>(col, r0, vf);
    else if (from_i64) load_i64_m<N, NV>(col, r0, vi);
    else if (op >= AGG_SUM_EXPR) eval_expr_m<N, NV>(p, Q_AGG_COL(a), cols, r0, vf);
#pragma unroll
    for (int i = 0; i < R; i++) {
      long long w;
      switch (op) {
        case AGG_SUM_F64: case AGG_SUM_EXPR: w = __double_as_longlong(vf[i]); break;            // the double itself
        case AGG_MIN_F64: case AGG_MAX_F64: case AGG_MIN_EXPR: case AGG_MAX_EXPR: w = f64_to_ord(vf[i]); break;  // order-preserving int64
        case AGG_SUM_I64: case AGG_MIN_I64: case AGG_MAX_I64: w = vi[i]; break;
        case AGG_SUM_F64_I64: w = __double_as_longlong((double)vi[i]); break;
        case AGG_MIN_F64_I64: case AGG_MAX_F64_I64: w = f64_to_ord((double)vi[i]); break;
        case AGG_SUM_I64_F64: case AGG_MIN_I64_F64: case AGG_MAX_I64_F64: w = (long long)vf[i]; break;
        case AGG_COUNT_IF: w = 1; break;
        default: w = (long long)ident; break;
      }
      out[i] = (unsigned long long)w;
    }
#pragma unroll
    for (int i = 0; i < R; i++)
      if ((m >> i) & 1u) p.emit_slots[pos[i] * nsr + (slot - 1)] = ((ma >> i) & 1u) ? out[i] : ident;
  }
}

// Aggregate the rows of `r0` (N = 4: vector loads; N = 1: scalar loads); bit v * N + i of m = row (v, i) takes part.
// WARP = called by all 32 lanes together (lanes without rows pass m = 0): lets a warp whose surviving
// rows all share one key (time buckets, sorted low-cardinality keys) fold them with shuffles and
// issue one atomic per aggregate instead of one per row.
template <int STRAT, int N, int NV, bool WARP>
__device__ __forceinline__ void accumulate(const ScanParams& p, const void* const* cols, const Acc<STRAT>& acc, uint64_t key_base,
                                           const RowSet<NV>& r0, uint32_t m) {
  constexpr int R = N * NV;  // rows handled by this call
  typedef typename Acc<STRAT>::key_t key_t;
  key_t key[R];
#pragma unroll
  for (int i = 0; i < R; i++) key[i] = (key_t)key_base;
  Q_UNROLL
  for (int d = 0; d < Q_N_DIMS; d++) {
    uint32_t ids[R];
    load_ids_m<N, NV>(cols[Q_DIM_COL(d)], Q_DIM_WIDTH(d), r0, ids);
    if (Q_DIM_REMAP(d)) {
      const uint32_t* lut = p.dims[d].remap;
#pragma unroll
      for (int i = 0; i < R; i++) ids[i] = __ldg(lut + ids[i]);
    }
    const key_t stride = (key_t)p.dims[d].stride;
#pragma unroll
    for (int i = 0; i < R; i++) key[i] += (key_t)ids[i] * stride;
  }
  if constexpr (STRAT == STRAT_EMIT) {
    emit_rows<N, NV, WARP>(p, cols, key, r0, m);
    return;
  }
  if (STRAT == STRAT_HASH) {
#pragma unroll
    for (int i = 0; i < R; i++) {
      if ((m >> i) & 1u) {
        uint64_t h = hash_claim(p.hkeys, p.hmask, p.hstride, (uint64_t)key[i]);
        if (h == HASH_EMPTY) { p.counters[1] = 1; m &= ~(1u << i); }
        key[i] = (key_t)h;
      }
    }
  }
  // row pointers: slot s of row i lives at row[i] + s * Acc::SLOT_STRIDE
  unsigned long long* row[R];
#pragma unroll
  for (int i = 0; i < R; i++) row[i] = acc.at(key[i]);
  bool uni = false, is_leader = false;
  unsigned long long* rowL = row[0];
  if (WARP && STRAT != STRAT_LANE && STRAT != STRAT_SPAN) {
    key_t lk = key[0];
#pragma unroll
    for (int i = R - 1; i >= 0; i--)
      if ((m >> i) & 1u) { lk = key[i]; rowL = row[i]; }
    bool same = true;
#pragma unroll
    for (int i = 0; i < R; i++)
      if (((m >> i) & 1u) && key[i] != lk) same = false;
    const unsigned have = __ballot_sync(0xffffffffu, m != 0u);
    const int leader = __ffs(have) - 1;
    const key_t kl = __shfl_sync(0xffffffffu, lk, leader);
    uni = __all_sync(0xffffffffu, m == 0u || (same && lk == kl));
    is_leader = (int)(threadIdx.x & 31u) == leader;
  }
  // STRAT_SPAN: every row of the call lands in the thread's one private cell, so the rows are folded in registers and the
  // shared-memory cell is touched once per call (a read-modify-write per row was a chain of dependent shared-memory round
  // trips: the short-scoreboard stall that held the timeseries shapes at 0.72 of the HBM roofline)
  constexpr bool ONE_CELL = STRAT == STRAT_SPAN;
  if (uni) {
    const int c = __reduce_add_sync(0xffffffffu, __popc(m));
    if (is_leader) acc.countn(rowL, c);
  } else if (ONE_CELL) {
    if (m) acc.countn(row[0], __popc(m));
  } else {
#pragma unroll
    for (int i = 0; i < R; i++)
      if ((m >> i) & 1u) acc.count1(row[i]);  // slot 0: row count of the group
  }

  Q_UNROLL
  for (int a = 0; a < Q_N_AGGS; a++) {
    uint32_t ma = m;
    if (Q_AGG_LEAF(a) >= 0) {
      ma &= leaf_test_m<N, NV>(p, Q_AGG_LEAF(a), cols, r0);
      if (!WARP && !ma) continue;
    }
    const int so = Q_AGG_SLOT(a) * Acc<STRAT>::SLOT_STRIDE;
    const void* col = Q_AGG_OP(a) >= AGG_SUM_EXPR ? nullptr : cols[Q_AGG_COL(a)];  // *_EXPR: col is an expression index
#define SDO_ROWS(EXPR)                                     \
  _Pragma("unroll") for (int i = 0; i < R; i++) {          \
    if ((ma >> i) & 1u) {                                  \
      unsigned long long* q = row[i] + so;                 \
      EXPR;                                                \
    }                                                      \
  }
// uniform-key warp: fold the rows with shuffles, one atomic by the leader lane
#define SDO_SUMF(VAL)                                                                   \
  if (uni) {                                                                            \
    double t = 0.0;                                                                     \
    _Pragma("unroll") for (int i = 0; i < R; i++) if ((ma >> i) & 1u) t += (VAL);       \
    t = warp_add(t);                                                                    \
    if (is_leader) acc.add_f64(rowL + so, t);                                           \
  } else if (ONE_CELL) {                                                                \
    double t = 0.0;                                                                     \
    _Pragma("unroll") for (int i = 0; i < R; i++) if ((ma >> i) & 1u) t += (VAL);       \
    if (ma) acc.add_f64(row[0] + so, t);                                                \
  } else { SDO_ROWS(acc.add_f64(q, (VAL))) }
#define SDO_SUMI(VAL)                                                                   \
  if (uni) {                                                                            \
    long long t = 0;                                                                    \
    _Pragma("unroll") for (int i = 0; i < R; i++) if ((ma >> i) & 1u) t += (VAL);       \
    t = warp_add(t);                                                                    \
    if (is_leader) acc.add_i64(rowL + so, t);                                           \
  } else if (ONE_CELL) {                                                                \
    long long t = 0;                                                                    \
    _Pragma("unroll") for (int i = 0; i < R; i++) if ((ma >> i) & 1u) t += (VAL);       \
    if (ma) acc.add_i64(row[0] + so, t);                                                \
  } else { SDO_ROWS(acc.add_i64(q, (VAL))) }
#define SDO_MINI(VAL)                                                                   \
  if (uni) {                                                                            \
    long long t = 0x7FFFFFFFFFFFFFFFll;                                                 \
    _Pragma("unroll") for (int i = 0; i < R; i++) if ((ma >> i) & 1u) t = min(t, (long long)(VAL)); \
    t = warp_min(t);                                                                    \
    if (is_leader) acc.min_i64(rowL + so, t);                                           \
  } else if (ONE_CELL) {                                                                \
    long long t = 0x7FFFFFFFFFFFFFFFll;                                                 \
    _Pragma("unroll") for (int i = 0; i < R; i++) if ((ma >> i) & 1u) t = min(t, (long long)(VAL)); \
    if (ma) acc.min_i64(row[0] + so, t);                                                \
  } else { SDO_ROWS(acc.min_i64(q, (VAL))) }
#define SDO_MAXI(VAL)                                                                   \
  if (uni) {                                                                            \
    long long t = (long long)0x8000000000000000ull;                                     \
    _Pragma("unroll") for (int i = 0; i < R; i++) if ((ma >> i) & 1u) t = max(t, (long long)(VAL)); \
    t = warp_max(t);                                                                    \
    if (is_leader) acc.max_i64(rowL + so, t);                                           \
  } else if (ONE_CELL) {                                                                \
    long long t = (long long)0x8000000000000000ull;                                     \
    _Pragma("unroll") for (int i = 0; i < R; i++) if ((ma >> i) & 1u) t = max(t, (long long)(VAL)); \
    if (ma) acc.max_i64(row[0] + so, t);                                                \
  } else { SDO_ROWS(acc.max_i64(q, (VAL))) }
    switch (Q_AGG_OP(a)) {
      case AGG_SUM_F64: { double v[R]; load_f64_m<N, NV>(col, r0, v); SDO_SUMF(v[i]) break; }
      case AGG_SUM_I64: { long long v[R]; load_i64_m<N, NV>(col, r0, v); SDO_SUMI(v[i]) break; }
      case AGG_COUNT_IF: { SDO_SUMI(1ll) break; }
      case AGG_MIN_F64: { double v[R]; load_f64_m<N, NV>(col, r0, v); SDO_MINI(f64_to_ord(v[i])) break; }
      case AGG_MAX_F64: { double v[R]; load_f64_m<N, NV>(col, r0, v); SDO_MAXI(f64_to_ord(v[i])) break; }
      case AGG_MIN_I64: { long long v[R]; load_i64_m<N, NV>(col, r0, v); SDO_MINI(v[i]) break; }
      case AGG_MAX_I64: { long long v[R]; load_i64_m<N, NV>(col, r0, v); SDO_MAXI(v[i]) break; }
      case AGG_SUM_F64_I64: { long long v[R]; load_i64_m<N, NV>(col, r0, v); SDO_SUMF((double)v[i]) break; }
      case AGG_MIN_F64_I64: { long long v[R]; load_i64_m<N, NV>(col, r0, v); SDO_MINI(f64_to_ord((double)v[i])) break; }
      case AGG_MAX_F64_I64: { long long v[R]; load_i64_m<N, NV>(col, r0, v); SDO_MAXI(f64_to_ord((double)v[i])) break; }
      case AGG_SUM_I64_F64: { double v[R]; load_f64_m<N, NV>(col, r0, v); SDO_SUMI((long long)v[i]) break; }
      case AGG_MIN_I64_F64: { double v[R]; load_f64_m<N, NV>(col, r0, v); SDO_MINI((long long)v[i]) break; }
      case AGG_MAX_I64_F64: { double v[R]; load_f64_m<N, NV>(col, r0, v); SDO_MAXI((long long)v[i]) break; }
      case AGG_SUM_EXPR: { double v[R]; eval_expr_m<N, NV>(p, Q_AGG_COL(a), cols, r0, v); SDO_SUMF(v[i]) break; }
      case AGG_MIN_EXPR: { double v[R]; eval_expr_m<N, NV>(p, Q_AGG_COL(a), cols, r0, v); SDO_MINI(f64_to_ord(v[i])) break; }
      case AGG_MAX_EXPR: { double v[R]; eval_expr_m<N, NV>(p, Q_AGG_COL(a), cols, r0, v); SDO_MAXI(f64_to_ord(v[i])) break; }
      case AGG_HLL: {
        uint32_t ids[R];
        load_ids_m<N, NV>(col, Q_AGG_WIDTH(a), r0, ids);
#pragma unroll
        for (int i = 0; i < R; i++)
          if ((ma >> i) & 1u) {
            const uint32_t packed = hll_value(Q_AGG_HLL_MODE(a), p.aggs[a].hll_lut, ids[i]);
            if (Q_HLL_PLACE == HLL_IN_GLOBAL) {
              uint8_t* regs = p.hll + (((uint64_t)key[i] * (uint64_t)Q_N_HLL + (uint64_t)Q_AGG_HLL_INDEX(a)) << HLL_P);
              hll_update(regs, packed);
            } else {
              const uint32_t set = (Q_HLL_PLACE == HLL_IN_SMEM_ALL ? (uint32_t)key[i] * (uint32_t)Q_N_HLL : 0u) + (uint32_t)Q_AGG_HLL_INDEX(a);
              if (Q_HLL_WIDE) hll_update_sm_wide(acc.hll_sm + set * HLL_M, packed);
              else hll_update_sm(acc.hll_sm + set * (HLL_M / 4), packed);
            }
          }
        break;
      }
      default: break;
    }
#undef SDO_ROWS
#undef SDO_SUMF
#undef SDO_SUMI
#undef SDO_MINI
#undef SDO_MAXI
  }
}

// One pass: 16 rows per thread starting at `base` (this thread's first row), vectors VEC_STRIDE rows apart.  Must be
// called by all 32 lanes of a warp together (uses full-warp collectives).
template <int STRAT>
__device__ __forceinline__ void process_pass(const ScanParams& p, const DevSpan& sp, uint32_t base, const Acc<STRAT>& acc) {
  uint32_t m = 0, vmask = 0;
  if (base < sp.row_end) {
    const uint32_t left = sp.row_end - base;  // rows from this thread's first row to the span end
#pragma unroll
    for (int j = 0; j < VECS; j++) {
      if (left > j * VEC_STRIDE) {
        const uint32_t n = min(4u, left - j * VEC_STRIDE);
        m |= ((1u << n) - 1u) << (4 * j);
        vmask |= 1u << j;
      }
    }
    if (base < sp.row_begin) m &= 0xFFF0u | (0xFu << (sp.row_begin - base));  // the span's first vector may start up to 3 rows early
  }
  const void* const* cols = p.segptr + (size_t)sp.seg * Q_N_COLS;
  if (Q_N_OPS && vmask) m &= eval_filter16(p, cols, base, vmask);
  const int densest = __reduce_max_sync(0xffffffffu, __popc(m));
  if (densest == 0) return;
  const uint64_t key_base = Q_IGNORE_KEY_BASE ? 0ull : sp.key_base;
  if (densest <= SPARSE_MAX) {
    for (int it = 0; it < densest; it++) {
      if (m) {
        const int b = __ffs(m) - 1;
        m &= m - 1;
        RowSet<1> one;
        one.r0[0] = base + (uint32_t)(b >> 2) * VEC_STRIDE + (uint32_t)(b & 3);
        accumulate<STRAT, 1, 1, false>(p, cols, acc, key_base, one, 1u);
      }
    }
  } else {
#pragma unroll 1
    for (int j = 0; j < VECS; j += DENSE_NV) {
      const uint32_t mj = (m >> (4 * j)) & ((1u << (4 * DENSE_NV)) - 1u);
      if (!__any_sync(0xffffffffu, mj != 0u)) continue;
      // lanes whose vector lies outside the span join with m = 0 and read a harmless in-span address
      RowSet<DENSE_NV> rs;
#pragma unroll
      for (int v = 0; v < DENSE_NV; v++) rs.r0[v] = ((vmask >> (j + v)) & 1u) ? base + (j + v) * VEC_STRIDE : (sp.row_begin & ~3u);
      accumulate<STRAT, 4, DENSE_NV, true>(p, cols, acc, key_base, rs, mj);
    }
  }
}

__device__ __forceinline__ unsigned long long slot_merge(int op, unsigned long long v, unsigned long long w) {
  if (op == SLOT_SUM_I64) return v + w;
  if (op == SLOT_SUM_F64) return (unsigned long long)__double_as_longlong(__longlong_as_double((long long)v) + __longlong_as_double((long long)w));
  if (op == SLOT_MIN_I64) return (unsigned long long)min((long long)v, (long long)w);
  return (unsigned long long)max((long long)v, (long long)w);
}

// STRAT_SPAN: fold the warp's 32 private copies of every slot, merge them into row `key` of the HBM table, start over
__device__ __forceinline__ void span_flush(const ScanParams& p, unsigned long long* smem, int tid, int ns, uint64_t key) {
  for (int s = 0; s < ns; s++) {
    unsigned long long* cell = smem + (size_t)s * SCAN_THREADS + tid;
    const unsigned long long init = p.slot_init[s];
    const int op = p.slot_op[s];
    unsigned long long v = *cell;
    *cell = init;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v = slot_merge(op, v, __shfl_xor_sync(0xffffffffu, v, off));
    if ((tid & 31) == 0 && v != init) {
      unsigned long long* q = p.table + key * (uint64_t)ns + s;
      if (op == SLOT_SUM_I64) atomicAdd(q, v);
      else if (op == SLOT_SUM_F64) atomicAdd((double*)q, __longlong_as_double((long long)v));
      else if (op == SLOT_MIN_I64) atomicMin((long long*)q, (long long)v);
      else atomicMax((long long*)q, (long long)v);
    }
  }
}

// ---- the kernel body ------------------------------------------------------------------------------------------------
// `coarse` = coarse span index, stored right after the n_spans + 1 span records in the per-query device blob.
template <int STRAT>
__device__ __forceinline__ void scan_body(const ScanParams& p, const uint32_t* __restrict__ coarse) {
  extern __shared__ __align__(16) unsigned long long smem[];
  __shared__ uint32_t s_chunk, s_span;
  const int tid = threadIdx.x;
  const uint32_t lane = tid & 31u, warp = tid >> 5;
  const int ns = Q_N_SLOTS;
  Acc<STRAT> acc;
  acc.ns = ns;
  // shared-memory HLL sketches sit behind the aggregate table
  // hll_words = packed words (4 registers each) the block's sketches merge into; the shared copy is 4x that when wide
  // HLL_IN_SMEM_SPAN caches the sketches of one group at a time: per warp with warp tickets, per block otherwise
  const uint32_t hll_words = Q_HLL_PLACE == HLL_IN_SMEM_ALL ? (uint32_t)p.n_keys * Q_N_HLL * (HLL_M / 4)
                             : (Q_HLL_PLACE == HLL_IN_SMEM_SPAN ? (uint32_t)Q_N_HLL * (HLL_M / 4) : 0u);
  const uint32_t hll_copies = Q_HLL_PLACE == HLL_IN_SMEM_SPAN && WARP_TICKETS ? SCAN_THREADS / 32 : 1;
  const uint32_t hll_copy_words = Q_HLL_WIDE ? hll_words * 4u : hll_words;
  const uint32_t hll_sm_words = hll_copy_words * hll_copies;
  {
    size_t table_words = STRAT == STRAT_SPAN ? (size_t)ns * SCAN_THREADS : STRAT == STRAT_LANE ? (size_t)p.n_keys * ns * SCAN_THREADS
                         : (STRAT == STRAT_BLOCK ? (p.rep_mask ? (size_t)(p.rep_mask + 1u) * p.rep_stride : (size_t)p.n_keys * ns) : 0);
    acc.hll_sm = (unsigned int*)(smem + ((table_words + 1) & ~(size_t)1));  // 16-byte aligned (the host pads likewise)
    for (uint32_t w = tid; w < hll_sm_words; w += SCAN_THREADS) acc.hll_sm[w] = 0u;
    if (hll_copies > 1) acc.hll_sm += warp * hll_copy_words;
  }
  const uint32_t flush_first = hll_copies > 1 ? lane : (uint32_t)tid, flush_step = hll_copies > 1 ? 32u : (uint32_t)SCAN_THREADS;
  uint64_t hll_key = ~0ull;  // HLL_IN_SMEM_SPAN: group whose sketch is cached (uniform over the warp / block)
  uint64_t slot_key = ~0ull; // STRAT_SPAN: group the private accumulators belong to
  if (STRAT == STRAT_SPAN) {
    for (int x = 0; x < ns; x++) smem[(size_t)x * SCAN_THREADS + tid] = p.slot_init[x];
    acc.base = smem + tid;
  } else if (STRAT == STRAT_LANE) {
    const uint32_t n = (uint32_t)p.n_keys * ns;
    for (uint32_t x = 0; x < n; x++) smem[(size_t)x * SCAN_THREADS + tid] = p.slot_init[x % ns];
    acc.base = smem + tid;
  } else if (STRAT == STRAT_BLOCK) {
    const uint32_t n = (uint32_t)p.n_keys * ns;
    if (p.rep_mask) {
      const uint32_t total = (p.rep_mask + 1u) * p.rep_stride;
      for (uint32_t x = tid; x < total; x += SCAN_THREADS) smem[x] = p.slot_init[(x % p.rep_stride) % ns];
      acc.base = smem + (size_t)(lane & p.rep_mask) * p.rep_stride;
    } else {
      for (uint32_t x = tid; x < n; x += SCAN_THREADS) smem[x] = p.slot_init[x % ns];
      acc.base = smem;
    }
  } else if (STRAT == STRAT_HASH) {
    acc.base = p.table + 1;  // word 0 of a bucket is its key
    acc.ns = (int)p.hstride;
  } else if (STRAT == STRAT_EMIT) {
    acc.base = nullptr;
  } else {
    acc.base = p.table;
  }
  __syncthreads();

  if (WARP_TICKETS) {
    // Ticket draws of ~6000 warps would serialise on one address, so there are TICKET_REGIONS counters.  Region g owns
    // the batches (runs of TICKET_BATCH consecutive tickets) whose index is g modulo TICKET_REGIONS: all regions advance
    // through the rows at the same pace, so the GPU as a whole still walks one moving window (L2 locality of the group
    // tables).  A warp starts at its block's region and moves on, never back, when a region runs dry.
    const uint32_t ticket_rows = 1u << Q_TICKET_SHIFT, unit_ticket_shift = UNIT_SHIFT - Q_TICKET_SHIFT;
    const uint32_t n_tickets = p.n_units << unit_ticket_shift;
    const uint32_t round = TICKET_BATCH * TICKET_REGIONS, full = n_tickets / round, rem = n_tickets % round;
    uint32_t region = blockIdx.x % TICKET_REGIONS, dry = 0;
    DevSpan sp = p.spans[0];
    uint32_t next_begin = 0;  // units [sp.unit_begin, next_begin) belong to the cached span
    for (;;) {
      uint32_t t = 0xFFFFFFFFu;
      if (lane == 0) {
        while (dry < TICKET_REGIONS) {
          unsigned int* c = p.counters + 32 * (region + 1);
          const uint32_t tail = rem > region * TICKET_BATCH ? min(TICKET_BATCH, rem - region * TICKET_BATCH) : 0u;
          const uint32_t lim = full * TICKET_BATCH + tail;  // tickets this region owns
          if (*(volatile unsigned int*)c < lim) {
            const uint32_t k = atomicAdd(c, 1u);
            if (k < lim) { t = ((k / TICKET_BATCH) * TICKET_REGIONS + region) * TICKET_BATCH + k % TICKET_BATCH; break; }
          }
          region = (region + 1) % TICKET_REGIONS;
          dry++;
        }
      }
      t = __shfl_sync(0xffffffffu, t, 0);
      if (t == 0xFFFFFFFFu) break;
      const uint32_t u = t >> unit_ticket_shift;
      if (u < sp.unit_begin || u >= next_begin) {
        // the span that holds unit u: the coarse index narrows it to a window, the lanes test 32 candidates at a time
        const uint32_t lo = coarse[u / COARSE_UNITS], hi = coarse[u / COARSE_UNITS + 1];
        uint32_t s = lo;
        for (uint32_t s0 = lo; s0 <= hi; s0 += 32) {
          const uint32_t c = s0 + lane;
          const bool hit = c <= hi && c < p.n_spans && p.spans[c].unit_begin <= u && u < p.spans[c + 1].unit_begin;
          const unsigned found = __ballot_sync(0xffffffffu, hit);
          if (found) { s = s0 + (uint32_t)__ffs(found) - 1u; break; }
        }
        sp = p.spans[s];
        next_begin = p.spans[s + 1].unit_begin;
      }
      const uint32_t row0 = (sp.row_begin & ~3u) + (u - sp.unit_begin) * UNIT_ROWS + (t & ((1u << unit_ticket_shift) - 1u)) * ticket_rows;
      if (row0 >= sp.row_end) continue;  // the span ends before this quarter of its last unit
      if (STRAT == STRAT_SPAN && sp.key_base != slot_key) {
        if (slot_key != ~0ull) span_flush(p, smem, tid, ns, slot_key);
        slot_key = sp.key_base;
      }
      if (Q_HLL_PLACE == HLL_IN_SMEM_SPAN && sp.key_base != hll_key) {
        // the warp's cached sketch belongs to another group: merge it into HBM, start an empty one
        __syncwarp();
        if (hll_key != ~0ull) {
          unsigned int* dst = (unsigned int*)(p.hll + hll_key * (uint64_t)Q_N_HLL * HLL_M);
          if (Q_HLL_WIDE) hll_flush_sm<true>(acc.hll_sm, dst, hll_words, flush_first, flush_step);
          else hll_flush_sm<false>(acc.hll_sm, dst, hll_words, flush_first, flush_step);
        }
        hll_key = sp.key_base;
        __syncwarp();
      }
#pragma unroll 1
      for (uint32_t r = row0; r < row0 + ticket_rows && r < sp.row_end; r += PASS_ROWS) process_pass<STRAT>(p, sp, r + lane * ROWS_PER_THREAD, acc);
    }
  } else
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
      if (STRAT == STRAT_SPAN && sp.key_base != slot_key) {
        if (slot_key != ~0ull) span_flush(p, smem, tid, ns, slot_key);
        slot_key = sp.key_base;
      }
      if (Q_HLL_PLACE == HLL_IN_SMEM_SPAN && sp.key_base != hll_key) {
        // the cached sketch belongs to another group: merge it into HBM, start an empty one
        __syncthreads();
        if (hll_key != ~0ull) {
          unsigned int* dst = (unsigned int*)(p.hll + hll_key * (uint64_t)Q_N_HLL * HLL_M);
          if (Q_HLL_WIDE) hll_flush_sm<true>(acc.hll_sm, dst, hll_words, flush_first, flush_step);
          else hll_flush_sm<false>(acc.hll_sm, dst, hll_words, flush_first, flush_step);
        }
        hll_key = sp.key_base;
        __syncthreads();
      }
      process_pass<STRAT>(p, sp, (sp.row_begin & ~3u) + (u - sp.unit_begin) * UNIT_ROWS + tid * ROWS_PER_THREAD, acc);
    }
  }
  if (hll_sm_words) {
    __syncthreads();
    unsigned int* dst = Q_HLL_PLACE == HLL_IN_SMEM_ALL ? (unsigned int*)p.hll : (unsigned int*)(p.hll + hll_key * (uint64_t)Q_N_HLL * HLL_M);
    if (Q_HLL_PLACE == HLL_IN_SMEM_ALL || hll_key != ~0ull) {
      if (Q_HLL_WIDE) hll_flush_sm<true>(acc.hll_sm, dst, hll_words, flush_first, flush_step);
      else hll_flush_sm<false>(acc.hll_sm, dst, hll_words, flush_first, flush_step);
    }
  }

  if (STRAT == STRAT_SPAN) {
    if (slot_key != ~0ull) span_flush(p, smem, tid, ns, slot_key);
  } else if (STRAT == STRAT_LANE) {
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
      for (int j = 1; j < SCAN_THREADS / 32; j++) v = slot_merge(op, v, src[lane + 32 * j]);
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) v = slot_merge(op, v, __shfl_down_sync(0xffffffffu, v, off));
      if (lane == 0) out[x] = v;
    }
  } else if (STRAT == STRAT_BLOCK) {
    __syncthreads();
    const uint32_t n = (uint32_t)p.n_keys * ns;
    for (uint32_t x = tid; x < n; x += SCAN_THREADS) {
      const int sl = x % ns;
      const int op = p.slot_op[sl];
      unsigned long long v = smem[x];
      for (uint32_t r = 1; r <= p.rep_mask; r++) v = slot_merge(op, v, smem[(size_t)r * p.rep_stride + x]);
      if (v == p.slot_init[sl]) continue;
      if (op == SLOT_SUM_I64) atomicAdd(p.table + x, v);
      else if (op == SLOT_SUM_F64) atomicAdd((double*)(p.table + x), __longlong_as_double((long long)v));
      else if (op == SLOT_MIN_I64) atomicMin((long long*)(p.table + x), (long long)v);
      else atomicMax((long long*)(p.table + x), (long long)v);
    }
  }
}

}  // namespace sdo
