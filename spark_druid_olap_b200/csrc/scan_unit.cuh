// One unit of the scan: 256 threads x 16 rows (4 vectors of 4 consecutive rows per thread, the
// vectors 1024 rows apart so every load instruction of a warp is fully coalesced).
// Included by kernels.cu after the accumulation policies.
//
// Shape:
//   1. validity mask (16 bits per thread) from the span bounds;
//   2. filter: the postfix program is decoded once per unit and applied to all 4 vectors of the
//      thread (16-bit masks on a register stack); each leaf issues its 4 loads back to back
//      (width-specialised) before any compare, so a warp keeps 4 x 128..512 B in flight per leaf;
//   3. the warp looks at its densest lane: if few rows survive (selective filters), every lane
//      walks its own surviving rows one by one with scalar loads -- the aggregate path then runs
//      max-popcount times per warp instead of 16 times with most lanes idle; otherwise each vector
//      with a surviving row is aggregated with vector loads (one 32-byte sector per metric).
//   The aggregator dispatch (switch on op) sits outside the per-row loops.
#pragma once

constexpr int VECS = UNIT_ROWS / (SCAN_THREADS * ROWS_PER_THREAD);  // vectors per thread per unit
constexpr uint32_t VEC_STRIDE = SCAN_THREADS * ROWS_PER_THREAD;     // rows between a thread's vectors
constexpr int SPARSE_MAX = 5;  // densest lane has at most this many survivors -> row-by-row path

template <int N>
__device__ __forceinline__ void load_ids(const void* base, int width, uint32_t r0, uint32_t (&out)[N]);
template <>
__device__ __forceinline__ void load_ids<4>(const void* base, int width, uint32_t r0, uint32_t (&out)[4]) {
  load_ids4(base, width, r0, out);
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
__device__ __forceinline__ void load_f64<4>(const void* base, uint32_t r0, double (&v)[4]) { load_f64x4(base, r0, v); }
template <>
__device__ __forceinline__ void load_f64<1>(const void* base, uint32_t r0, double (&v)[1]) { v[0] = __ldg((const double*)base + r0); }
template <int N>
__device__ __forceinline__ void load_i64(const void* base, uint32_t r0, long long (&v)[N]);
template <>
__device__ __forceinline__ void load_i64<4>(const void* base, uint32_t r0, long long (&v)[4]) { load_i64x4(base, r0, v); }
template <>
__device__ __forceinline__ void load_i64<1>(const void* base, uint32_t r0, long long (&v)[1]) { v[0] = __ldg((const long long*)base + r0); }

// mask (N bits) of rows r0 .. r0+N-1 whose id satisfies the leaf (aggregator-level filters)
template <int N>
__device__ __forceinline__ uint32_t leaf_test(const DevLeaf& lf, const void* const* cols, uint32_t r0) {
  uint32_t ids[N];
  load_ids<N>(cols[lf.col], lf.width, r0, ids);
  uint32_t m = 0;
  if (lf.kind == LEAF_RANGE) {
    const uint32_t lo = lf.lo, span = lf.hi - lf.lo;
#pragma unroll
    for (int i = 0; i < N; i++) m |= (uint32_t)((ids[i] - lo) < span) << i;
  } else {
    const uint32_t* bits = lf.bits;
#pragma unroll
    for (int i = 0; i < N; i++) m |= ((__ldg(bits + (ids[i] >> 5)) >> (ids[i] & 31)) & 1u) << i;
  }
  return m;
}

// ---- 16-row leaf evaluation, specialised by id width ------------------------------------------------
// Vectors outside the span (vmask bit clear) re-read vector 0 instead of branching; their bits are
// cleared by the caller's validity mask.
template <int W> struct IdVec;
template <> struct IdVec<1> { typedef uint32_t type; };
template <> struct IdVec<2> { typedef uint2 type; };
template <> struct IdVec<4> { typedef uint4 type; };

__device__ __forceinline__ void unpack(uint32_t v, uint32_t (&id)[4]) { id[0] = v & 0xffu; id[1] = (v >> 8) & 0xffu; id[2] = (v >> 16) & 0xffu; id[3] = v >> 24; }
__device__ __forceinline__ void unpack(uint2 v, uint32_t (&id)[4]) { id[0] = v.x & 0xffffu; id[1] = v.x >> 16; id[2] = v.y & 0xffffu; id[3] = v.y >> 16; }
__device__ __forceinline__ void unpack(uint4 v, uint32_t (&id)[4]) { id[0] = v.x; id[1] = v.y; id[2] = v.z; id[3] = v.w; }

template <int W>
__device__ __forceinline__ uint32_t leaf_mask16_w(const DevLeaf& lf, const void* col, uint32_t base, uint32_t vmask) {
  typedef typename IdVec<W>::type V;
  const V* p = (const V*)col + (base >> 2);
  V v[VECS];
#pragma unroll
  for (int j = 0; j < VECS; j++) v[j] = __ldcs(p + (((vmask >> j) & 1u) ? j * (int)(VEC_STRIDE / 4) : 0));
  uint32_t m = 0;
  if (lf.kind == LEAF_RANGE) {
    const uint32_t lo = lf.lo, span = lf.hi - lf.lo;
#pragma unroll
    for (int j = 0; j < VECS; j++) {
      uint32_t id[4];
      unpack(v[j], id);
#pragma unroll
      for (int i = 0; i < 4; i++) m |= (uint32_t)((id[i] - lo) < span) << (4 * j + i);
    }
  } else {
    const uint32_t* bits = lf.bits;
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

__device__ __forceinline__ uint32_t leaf_mask16(const DevLeaf& lf, const void* const* cols, uint32_t base, uint32_t vmask) {
  const void* col = cols[lf.col];
  if (lf.width == 1) return leaf_mask16_w<1>(lf, col, base, vmask);
  if (lf.width == 2) return leaf_mask16_w<2>(lf, col, base, vmask);
  return leaf_mask16_w<4>(lf, col, base, vmask);
}

// postfix filter program over 16-bit masks; stack of 8 entries in two registers (top = low 16 bits of lo)
__device__ __forceinline__ uint32_t eval_filter16(const ScanParams& p, const void* const* cols, uint32_t base, uint32_t vmask) {
  uint64_t lo = 0, hi = 0;
  const int n = p.n_ops;
  for (int k = 0; k < n; k++) {
    const uint32_t op = p.op_code[k];
    if (op == FOP_LEAF || op == FOP_TRUE || op == FOP_FALSE) {
      uint64_t v = op == FOP_LEAF ? leaf_mask16(p.leaves[p.op_arg[k]], cols, base, vmask) : (op == FOP_TRUE ? 0xFFFFull : 0ull);
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

// aggregate N consecutive rows starting at r0 (N = 4: vector loads; N = 1: scalar loads); m = which rows
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

// WARP = called by all 32 lanes together (lanes without rows pass m = 0): lets a warp whose surviving
// rows all share one key (time buckets, sorted low-cardinality keys) fold them with shuffles and
// issue one atomic per aggregate instead of one per row.
template <int STRAT, int N, bool WARP>
__device__ __forceinline__ void accumulate(const ScanParams& p, const void* const* cols, const Acc<STRAT>& acc, uint64_t key_base,
                                           uint32_t r0, uint32_t m) {
  typedef typename Acc<STRAT>::key_t key_t;
  key_t key[N];
#pragma unroll
  for (int i = 0; i < N; i++) key[i] = (key_t)key_base;
  for (int d = 0; d < p.n_dims; d++) {
    uint32_t ids[N];
    load_ids<N>(cols[p.dims[d].col], p.dims[d].width, r0, ids);
    const key_t stride = (key_t)p.dims[d].stride;
#pragma unroll
    for (int i = 0; i < N; i++) key[i] += (key_t)ids[i] * stride;
  }
  if (STRAT == STRAT_HASH) {
#pragma unroll
    for (int i = 0; i < N; i++) {
      if ((m >> i) & 1u) {
        uint64_t h = hash_claim(p.hkeys, p.hmask, (uint64_t)key[i]);
        if (h == HASH_EMPTY) { p.counters[1] = 1; m &= ~(1u << i); }
        key[i] = (key_t)h;
      }
    }
  }
  // row pointers: slot s of row i lives at row[i] + s * Acc::SLOT_STRIDE
  unsigned long long* row[N];
#pragma unroll
  for (int i = 0; i < N; i++) row[i] = acc.at(key[i], 0);
  bool uni = false, is_leader = false;
  unsigned long long* rowL = row[0];
  if (WARP && STRAT != STRAT_LANE) {
    key_t lk = key[0];
#pragma unroll
    for (int i = N - 1; i >= 0; i--)
      if ((m >> i) & 1u) { lk = key[i]; rowL = row[i]; }
    bool same = true;
#pragma unroll
    for (int i = 0; i < N; i++)
      if (((m >> i) & 1u) && key[i] != lk) same = false;
    const unsigned have = __ballot_sync(0xffffffffu, m != 0u);
    const int leader = __ffs(have) - 1;
    const key_t kl = __shfl_sync(0xffffffffu, lk, leader);
    uni = __all_sync(0xffffffffu, m == 0u || (same && lk == kl));
    is_leader = (int)(threadIdx.x & 31u) == leader;
  }
  if (uni) {
    const int c = __reduce_add_sync(0xffffffffu, __popc(m));
    if (is_leader) acc.countn(rowL, c);
  } else {
#pragma unroll
    for (int i = 0; i < N; i++)
      if ((m >> i) & 1u) acc.count1(row[i]);  // slot 0: row count of the group
  }

  const int n_aggs = p.n_aggs;
  for (int a = 0; a < n_aggs; a++) {
    const DevAgg& ag = p.aggs[a];
    uint32_t ma = m;
    if (ag.leaf >= 0) {
      ma &= leaf_test<N>(p.leaves[ag.leaf], cols, r0);
      if (!WARP && !ma) continue;
    }
    const int so = ag.slot * Acc<STRAT>::SLOT_STRIDE;
#define SDO_ROWS(EXPR)                                     \
  _Pragma("unroll") for (int i = 0; i < N; i++) {          \
    if ((ma >> i) & 1u) {                                  \
      unsigned long long* q = row[i] + so;                 \
      EXPR;                                                \
    }                                                      \
  }
// uniform-key warp: fold the rows with shuffles, one atomic by the leader lane
#define SDO_SUMF(VAL)                                                                   \
  if (uni) {                                                                            \
    double t = 0.0;                                                                     \
    _Pragma("unroll") for (int i = 0; i < N; i++) if ((ma >> i) & 1u) t += (VAL);       \
    t = warp_add(t);                                                                    \
    if (is_leader) acc.add_f64(rowL + so, t);                                           \
  } else { SDO_ROWS(acc.add_f64(q, (VAL))) }
#define SDO_SUMI(VAL)                                                                   \
  if (uni) {                                                                            \
    long long t = 0;                                                                    \
    _Pragma("unroll") for (int i = 0; i < N; i++) if ((ma >> i) & 1u) t += (VAL);       \
    t = warp_add(t);                                                                    \
    if (is_leader) acc.add_i64(rowL + so, t);                                           \
  } else { SDO_ROWS(acc.add_i64(q, (VAL))) }
#define SDO_MINI(VAL)                                                                   \
  if (uni) {                                                                            \
    long long t = 0x7FFFFFFFFFFFFFFFll;                                                 \
    _Pragma("unroll") for (int i = 0; i < N; i++) if ((ma >> i) & 1u) t = min(t, (long long)(VAL)); \
    t = warp_min(t);                                                                    \
    if (is_leader) acc.min_i64(rowL + so, t);                                           \
  } else { SDO_ROWS(acc.min_i64(q, (VAL))) }
#define SDO_MAXI(VAL)                                                                   \
  if (uni) {                                                                            \
    long long t = (long long)0x8000000000000000ull;                                     \
    _Pragma("unroll") for (int i = 0; i < N; i++) if ((ma >> i) & 1u) t = max(t, (long long)(VAL)); \
    t = warp_max(t);                                                                    \
    if (is_leader) acc.max_i64(rowL + so, t);                                           \
  } else { SDO_ROWS(acc.max_i64(q, (VAL))) }
    switch (ag.op) {
      case AGG_SUM_F64: { double v[N]; load_f64<N>(cols[ag.col], r0, v); SDO_SUMF(v[i]) break; }
      case AGG_SUM_I64: { long long v[N]; load_i64<N>(cols[ag.col], r0, v); SDO_SUMI(v[i]) break; }
      case AGG_COUNT_IF: { SDO_SUMI(1ll) break; }
      case AGG_MIN_F64: { double v[N]; load_f64<N>(cols[ag.col], r0, v); SDO_MINI(f64_to_ord(v[i])) break; }
      case AGG_MAX_F64: { double v[N]; load_f64<N>(cols[ag.col], r0, v); SDO_MAXI(f64_to_ord(v[i])) break; }
      case AGG_MIN_I64: { long long v[N]; load_i64<N>(cols[ag.col], r0, v); SDO_MINI(v[i]) break; }
      case AGG_MAX_I64: { long long v[N]; load_i64<N>(cols[ag.col], r0, v); SDO_MAXI(v[i]) break; }
      case AGG_SUM_F64_I64: { long long v[N]; load_i64<N>(cols[ag.col], r0, v); SDO_SUMF((double)v[i]) break; }
      case AGG_MIN_F64_I64: { long long v[N]; load_i64<N>(cols[ag.col], r0, v); SDO_MINI(f64_to_ord((double)v[i])) break; }
      case AGG_MAX_F64_I64: { long long v[N]; load_i64<N>(cols[ag.col], r0, v); SDO_MAXI(f64_to_ord((double)v[i])) break; }
      case AGG_SUM_I64_F64: { double v[N]; load_f64<N>(cols[ag.col], r0, v); SDO_SUMI((long long)v[i]) break; }
      case AGG_MIN_I64_F64: { double v[N]; load_f64<N>(cols[ag.col], r0, v); SDO_MINI((long long)v[i]) break; }
      case AGG_MAX_I64_F64: { double v[N]; load_f64<N>(cols[ag.col], r0, v); SDO_MAXI((long long)v[i]) break; }
      case AGG_HLL: {
        uint32_t ids[N];
        load_ids<N>(cols[ag.col], ag.width, r0, ids);
#pragma unroll
        for (int i = 0; i < N; i++)
          if ((ma >> i) & 1u) {
            const uint32_t packed = hll_value(ag, ids[i]);
            if (p.hll_mode == HLL_IN_GLOBAL) {
              uint8_t* regs = p.hll + (((uint64_t)key[i] * (uint64_t)p.n_hll + (uint64_t)ag.hll_index) << HLL_P);
              hll_update(regs, packed);
            } else {
              const uint32_t set = (p.hll_mode == HLL_IN_SMEM_ALL ? (uint32_t)key[i] * (uint32_t)p.n_hll : 0u) + (uint32_t)ag.hll_index;
              hll_update_sm(acc.hll_sm + set * (HLL_M / 4), packed);
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

// Must be called by all 256 threads of the block (uses full-warp collectives).
template <int STRAT>
__device__ __forceinline__ void process_unit(const ScanParams& p, const DevSpan& sp, uint32_t unit_in_span, const Acc<STRAT>& acc) {
  const uint32_t base = (sp.row_begin & ~3u) + unit_in_span * UNIT_ROWS + threadIdx.x * ROWS_PER_THREAD;
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
  const void* const* cols = p.segptr + (size_t)sp.seg * p.n_cols;
  if (p.n_ops && vmask) m &= eval_filter16(p, cols, base, vmask);
  const int densest = __reduce_max_sync(0xffffffffu, __popc(m));
  if (densest == 0) return;
  const uint64_t key_base = p.ignore_key_base ? 0ull : sp.key_base;
  if (densest <= SPARSE_MAX) {
    for (int it = 0; it < densest; it++) {
      if (m) {
        const int b = __ffs(m) - 1;
        m &= m - 1;
        accumulate<STRAT, 1, false>(p, cols, acc, key_base, base + (uint32_t)(b >> 2) * VEC_STRIDE + (uint32_t)(b & 3), 1u);
      }
    }
  } else {
#pragma unroll 1
    for (int j = 0; j < VECS; j++) {
      const uint32_t mj = (m >> (4 * j)) & 0xFu;
      if (!__any_sync(0xffffffffu, mj != 0u)) continue;
      // lanes whose vector j lies outside the span join with m = 0 and read a harmless in-span address
      const uint32_t r0 = ((vmask >> j) & 1u) ? base + j * VEC_STRIDE : (sp.row_begin & ~3u);
      accumulate<STRAT, 4, true>(p, cols, acc, key_base, r0, mj);
    }
  }
}
