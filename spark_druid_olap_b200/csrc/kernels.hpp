// Device-side contract of the scan / filter / aggregate kernels (kernels.cu) and their launchers.
#pragma once
#ifdef __CUDACC_RTC__
// NVRTC (run-time specialisation): no host headers; fixed-width types by hand
typedef unsigned char uint8_t;
typedef unsigned short uint16_t;
typedef unsigned int uint32_t;
typedef unsigned long long uint64_t;
typedef signed char int8_t;
typedef short int16_t;
typedef int int32_t;
typedef long long int64_t;
#else
#include <cuda_runtime.h>

#include <cstdint>
#endif

namespace sdo {

constexpr int MAX_LEAVES = 12;
constexpr int MAX_OPS = 32;
constexpr int FILTER_STACK = 8;     // masks the device evaluator's register stack holds (eval_filter16: 8 x 16 bits in two registers)
constexpr int MAX_DIMS = 8;
constexpr int MAX_AGGS = 16;
constexpr int MAX_SLOTS = MAX_AGGS + 1;
constexpr int MAX_EXPRS = 8;        // aggregations whose argument is an arithmetic expression over metrics
constexpr int MAX_EXPR_OPS = 24;
constexpr int MAX_EXPR_CONSTS = 8;
constexpr int EXPR_STACK = 6;
constexpr int SCAN_THREADS = 256;
constexpr int ROWS_PER_THREAD = 4;
constexpr int UNIT_ROWS = SCAN_THREADS * ROWS_PER_THREAD * 4;  // 4096 rows: one pass of a block (4 vectors per thread)
constexpr uint64_t HASH_EMPTY = 0xFFFFFFFFFFFFFFFFull;
// ScanParams::counters: word 0 = block-ticket counter, word 1 = hash overflow flag, words 4-5 = compaction cursor,
// word 32 * (g + 1) = ticket counter of region g (128 bytes apart: one L2 atomic unit is not a bottleneck for the others)
constexpr int TICKET_REGIONS = 16;
constexpr int COUNTER_BYTES = 128 * (TICKET_REGIONS + 1);

// A span is a run of rows of one segment that shares its time-derived key part; interval pruning
// and time bucketing are resolved to spans on the host, so kernels never read a time column.
struct DevSpan {
  uint32_t seg;         // index into segptr rows
  uint32_t row_begin;   // first row (inclusive), segment-relative
  uint32_t row_end;     // last row (exclusive)
  uint32_t unit_begin;  // units (of UNIT_ROWS rows, from row_begin & ~3) before this span
  uint64_t key_base;    // time-derived key component
};

enum LeafKind : int32_t { LEAF_RANGE = 0, LEAF_BITSET = 1 };
struct DevLeaf {
  int32_t col;    // column slot in segptr row
  int32_t width;  // id bytes: 1, 2, 4
  int32_t kind;
  uint32_t lo, hi;       // RANGE: lo <= id < hi
  const uint32_t* bits;  // BITSET over ids
};
enum FilterOp : uint8_t { FOP_LEAF = 0, FOP_AND = 1, FOP_OR = 2, FOP_NOT = 3, FOP_TRUE = 4, FOP_FALSE = 5 };

struct DevDim {
  int32_t col;
  int32_t width;
  uint64_t stride;
  const uint32_t* remap;  // extraction function on the dimension: id -> id of the extracted value (else null)
};

// how a row value reaches its 8-byte slot
enum AggOp : int32_t {
  AGG_SUM_I64 = 0,      // int64 metric
  AGG_SUM_F64 = 1,      // float64 metric
  AGG_MIN_I64 = 2,
  AGG_MAX_I64 = 3,
  AGG_MIN_F64 = 4,      // kept as order-preserving int64
  AGG_MAX_F64 = 5,
  AGG_SUM_F64_I64 = 6,  // doubleSum over an int64 metric
  AGG_SUM_I64_F64 = 7,  // longSum over a float64 metric (truncating cast per row, as Druid)
  AGG_COUNT_IF = 8,     // filtered count: adds 1 where the aggregator's own filter matches
  AGG_HLL = 9,          // cardinality: HLL++ registers, no slot
  AGG_MIN_F64_I64 = 10,
  AGG_MAX_F64_I64 = 11,
  AGG_MIN_I64_F64 = 12,
  AGG_MAX_I64_F64 = 13,
  AGG_SUM_EXPR = 14,    // doubleSum / doubleMin / doubleMax of an arithmetic expression over metric columns
  AGG_MIN_EXPR = 15,    // (col = index into ScanParams::exprs)
  AGG_MAX_EXPR = 16,
};
// postfix program over doubles: the argument of a JavascriptAggregationSpec's fnAggregate (JSAggGenerator.scala:36-62)
// or of an aggregator's `expression` attribute.  Evaluated per row with IEEE round-to-nearest operations and no
// multiply-add contraction, so a row's value is the one a CPU computes.
enum ExprOp : uint8_t {
  EX_COL_F64 = 0,  // arg = column slot in the segptr row
  EX_COL_I64 = 1,  // int64 metric, converted to double (JavaScript numbers are doubles)
  EX_CONST = 2,    // arg = index into consts
  EX_ADD = 3, EX_SUB = 4, EX_MUL = 5, EX_DIV = 6, EX_NEG = 7, EX_ABS = 8, EX_MIN = 9, EX_MAX = 10, EX_FLOOR = 11,
  EX_CEIL = 12, EX_SQRT = 13, EX_MOD = 14,
};
struct DevExpr {
  int32_t n_ops;
  int32_t pad;
  uint8_t code[MAX_EXPR_OPS];
  uint8_t arg[MAX_EXPR_OPS];
  double consts[MAX_EXPR_CONSTS];
};
// how two partial slots merge (PostAggregate.scala:39-60: count/sum -> sum, min -> min, max -> max)
enum SlotOp : int32_t { SLOT_SUM_I64 = 0, SLOT_SUM_F64 = 1, SLOT_MIN_I64 = 2, SLOT_MAX_I64 = 3 };

// how a row's (register index << 6 | rho) is obtained: hashed from an integer id, looked up per dictionary id, or
// read ready-made from the dimension's sketch column (what the scan kernel uses)
enum HllMode : int32_t { HLL_DIRECT_INT = 0, HLL_DIRECT_LONG = 1, HLL_LUT = 2, HLL_PACKED = 3 };
// where HLL registers are updated during the scan
enum HllPlace : int32_t {
  HLL_IN_GLOBAL = 0,     // straight into the HBM register table
  HLL_IN_SMEM_ALL = 1,   // every group's sketch lives in shared memory, merged into HBM at block end
  HLL_IN_SMEM_SPAN = 2,  // no grouping dimension: the key is constant per span, one cached sketch per block
};
struct DevAgg {
  int32_t op;
  int32_t col;        // metric column slot (or dimension slot for HLL)
  int32_t slot;       // 8-byte slot index within the group row (slot 0 = row count)
  int32_t leaf;       // filtered aggregation: leaf index, else -1
  int32_t width;      // HLL: id width
  int32_t hll_index;  // HLL: which register set of the group
  int32_t hll_mode;
  int32_t pad;
  const uint16_t* hll_lut;  // HLL_LUT: id -> (idx << 6 | rho)
};

enum Strategy : int32_t {
  STRAT_LANE = 0,   // lane-private shared-memory accumulators (tiny key spaces)
  STRAT_BLOCK = 1,  // block-shared shared-memory table with shared atomics
  STRAT_HASH = 2,   // open-addressing global hash table (atomicCAS)
  STRAT_GLOBAL = 3, // direct-indexed global table (dense key space too big for shared memory)
  STRAT_SPAN = 4,   // no grouping dimension (timeseries): the key is constant over a span, so a thread keeps private
                    // accumulators for its warp's current key only and the warp merges them into the direct-indexed
                    // HBM table when it moves to another key -- no atomics, votes or shuffles per row
  STRAT_EMIT = 5,   // no aggregation at all: every surviving row is written out as a group of its own, [key | 1 | slot inputs],
                    // for the cross-GPU exchange of high-cardinality group-bys (rows ~ groups: aggregating locally first
                    // would cost a second pass of atomics at the owner and shrink nothing)
};

struct ScanParams {
  const DevSpan* spans;  // n_spans + 1 entries (sentinel carries unit_begin = n_units)
  uint32_t n_spans;
  uint32_t n_units;
  uint32_t units_per_chunk;     // block-ticket variant only (SDO_WARP_TICKETS 0)
  uint32_t n_cols;
  const void* const* segptr;  // [segment][n_cols] column base pointers
  int32_t n_leaves, n_ops;
  DevLeaf leaves[MAX_LEAVES];
  uint8_t op_code[MAX_OPS];
  uint8_t op_arg[MAX_OPS];
  int32_t n_dims;
  DevDim dims[MAX_DIMS];
  int32_t n_aggs;
  DevAgg aggs[MAX_AGGS];
  int32_t n_slots;  // per group, slot 0 = row count
  int32_t n_hll;
  int32_t ignore_key_base;  // count-only pass: all rows fall into key 0
  int32_t hll_mode;         // HllPlace: where the registers live while scanning
  uint64_t n_keys;  // dense key-space size (LANE / BLOCK / GLOBAL)
  int32_t slot_op[MAX_SLOTS];
  uint64_t slot_init[MAX_SLOTS];
  unsigned long long* table;    // LANE: partials [grid][n_keys*n_slots]; BLOCK/GLOBAL/SPAN: [n_keys*n_slots]; HASH: [cap*hstride]
  uint8_t* hll;                 // [n_keys or cap][n_hll][512]
  // HASH: one bucket = hstride words: [key | slot 0 .. slot n_slots-1 | padding]; key and aggregates of a group share a
  // 32-byte sector (hstride is a power of two), so claiming the bucket and updating it touch one sector
  unsigned long long* hkeys;    // HASH: bucket array base (== table); bucket h starts at hkeys + h * hstride
  uint64_t hmask;               // HASH: cap - 1
  uint32_t hstride;             // HASH: words per bucket
  int32_t n_exprs;
  DevExpr exprs[MAX_EXPRS];
  int32_t hll_wide;  // shared-memory sketches keep one 32-bit word per register (native atomicMax) instead of one byte
  int32_t ticket_shift;  // log2 of the rows a warp draws per ticket (9 .. 12)
  // BLOCK: the shared-memory table is replicated; lane l works on replica l & rep_mask (rep_stride words apart), so lanes
  // of a warp whose rows share a key rarely retry each other's 64-bit compare-and-swap adds.  Folded at block end.
  uint32_t rep_mask;
  uint32_t rep_stride;
  unsigned int* counters;       // [0] chunk ticket, [1] overflow flag
  // EMIT: output rows (same layout as compacted groups: keys[n], slots[n][n_slots]); the cursor is the u64 at counters + 4
  // Rows carry n_slots - 1 words (slot 0, the row count, is 1 by construction and is not stored).  With emit_owners > 1 the
  // rows are split on the way out by the rank that owns their key (interleaved: key mod owners, or by hash): owner o's rows
  // go to keys[o * emit_cap ...] / slots[o * emit_cap * (n_slots - 1) ...] under cursor emit_cursors[o].
  unsigned long long* emit_keys;
  unsigned long long* emit_slots;
  unsigned long long* emit_cursors;  // [max(emit_owners, 1)] rows written per owner
  uint64_t emit_cap;                 // rows per owner region
  int32_t emit_owners;               // 0 / 1: one stream
  int32_t emit_interleave;           // 1: owner = key mod owners; 0: owner = mix(key) mod owners
};
// who owns a key in a sharded exchange (the same on device and host)
#ifdef __CUDACC__
__host__ __device__
#endif
inline unsigned shard_owner(unsigned long long key, unsigned owners, int interleave) {
  if (interleave) return (unsigned)(key % owners);
  unsigned long long x = key * 0x9E3779B97F4A7C15ull;
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
  return (unsigned)((x >> 32) % owners);
}

#ifndef __CUDACC_RTC__
// ---- launchers (host) ---------------------------------------------------------------------------
// loader helpers
void launch_encode_dim(const void* src, bool src64, const int32_t* perm, int64_t r0, int64_t n, const uint32_t* lut,
                       void* dst, int dst_width, cudaStream_t st);
void launch_reencode(const void* src, int src_width, void* dst, int dst_width, int64_t n, const uint32_t* lut, cudaStream_t st);
void launch_gather64(const void* src, const int32_t* perm, int64_t r0, int64_t n, void* dst, cudaStream_t st);
// HLL lookup table of a string dictionary: lut[i] = hll_pack(xxh64(bytes of entry i))
void launch_hll_lut_strings(const char* bytes, const int64_t* offsets, int64_t n, uint16_t* lut, cudaStream_t st);
// sketch column of a dimension: out[row] = hll_pack(XxHash64(value of row)), from the id column (mode HLL_DIRECT_* hashes
// the integer id, HLL_LUT looks the dictionary id up)
void launch_hll_column(const void* ids, int width, int mode, const uint16_t* lut, int64_t n, uint16_t* out, cudaStream_t st);

// scan
size_t scan_smem_bytes(int strategy, uint64_t n_keys, int n_slots, uint32_t replicas = 1, uint32_t rep_stride = 0);
size_t scan_hll_smem_bytes(int hll_mode, bool wide, uint64_t n_keys, int n_hll);
int scan_grid_size(int strategy, size_t smem_bytes, int sm_count);
void launch_scan(int strategy, const ScanParams& p, int grid, size_t smem_bytes, cudaStream_t st);
void launch_fill_slots(unsigned long long* table, uint64_t n_rows, int n_slots, const uint64_t* slot_init_dev_or_null,
                       const ScanParams& p, cudaStream_t st);
void launch_fill_u64(unsigned long long* dst, uint64_t n, unsigned long long v, cudaStream_t st);
void launch_fill_hash(unsigned long long* table, uint64_t cap, uint32_t stride, int n_slots, const ScanParams& p, cudaStream_t st);
void launch_reduce_partials(const unsigned long long* partials, int n_blocks, uint32_t n, int n_slots, const ScanParams& p,
                            unsigned long long* out, cudaStream_t st);
// compaction of non-empty groups: dense tables (keyed by slot 0 > 0) and hash tables
void launch_compact_dense(const unsigned long long* table, uint64_t n_keys, int n_slots, unsigned long long* out_keys,
                          unsigned long long* out_slots, unsigned long long* out_src, unsigned long long* counter, cudaStream_t st,
                          unsigned long long key_off = 0, unsigned long long key_mul = 1);  // out key = index * key_mul + key_off
void launch_compact_hash(const unsigned long long* table, uint64_t cap, int n_slots, uint32_t stride,
                         unsigned long long* out_keys, unsigned long long* out_slots, unsigned long long* out_src,
                         unsigned long long* counter, cudaStream_t st);
// out[2 * set] = sum over the sketch's registers of 2^-register, out[2 * set + 1] = count of zero registers
void launch_hll_stats(const uint8_t* regs, uint64_t n_sets, double* out, cudaStream_t st);
void launch_gather_hll(const uint8_t* hll, const unsigned long long* src, uint64_t n, int n_hll, uint8_t* out, cudaStream_t st);
// top-k of (metric, key) rows: ord = order-preserving u64 of the metric (descending = larger first)
void launch_topk(const unsigned long long* keys, const unsigned long long* slots, uint64_t n, int n_slots, int slot, int slot_is_f64,
                 int descending, int k, unsigned long long* scratch_a, unsigned long long* scratch_b, unsigned long long* out_idx,
                 cudaStream_t st);
size_t topk_scratch_elems(uint64_t n, int k);
// Radix select in front of the tournament (select_kernels.cu): finds the groups at or above the k-th order value of any
// group table IN PLACE -- compacted rows, a direct-indexed table (live = row count != 0, key = index) or a hash table
// (live = key set) -- and runs the tournament over those candidates only.
enum SelKind : int32_t { SEL_ROWS = 0, SEL_DENSE = 1, SEL_HASH = 2 };
struct SelView {
  const unsigned long long* base;  // ROWS: slots [n][stride]; DENSE: table [n][stride]; HASH: buckets [n][stride] = [key | slots ...]
  const unsigned long long* keys;  // ROWS only
  uint64_t n;
  uint32_t stride;     // words per row / bucket
  uint32_t slot_off;   // word of the ordering aggregate inside a row / bucket
  uint32_t slot0_off;  // word of slot 0 inside a row / bucket (ROWS / DENSE: 0, HASH: 1)
  unsigned long long key_off, key_mul;  // DENSE: key of row i = i * key_mul + key_off (a rank's interleaved share of a key space); key_mul 0 = 1
  int32_t kind, is_f64, desc;
};
struct SelState {  // device-resident search state, zeroed before the first pass
  unsigned long long prefix;   // digits fixed so far (most significant first)
  unsigned long long k_rem;    // rank still to find below the prefix
  unsigned long long live;     // live groups in the table
  unsigned long long n_cand;   // candidates appended
  unsigned long long above;    // live groups whose fixed digits rank strictly above the prefix
  int32_t shift;               // candidates are the live groups with (ord >> shift) >= (prefix >> shift)
  int32_t done;
  int32_t overflow;            // more candidates than SEL_CAP (ties): the caller falls back to the full tournament
  int32_t pad;
};
constexpr uint32_t SEL_CAP = 1u << 16;
size_t select_scratch_bytes(int k);
// Queues: select passes, candidate collection, tournament over the candidates, gather of the k winners into
// out_keys[k] / out_slots[k][n_slots] (padding: key ~0).  `scratch` holds select_scratch_bytes(k); its first bytes are the
// SelState (read `overflow` and `live` after the stream has drained).
void launch_select_topk(const SelView& v, int n_slots, int k, void* scratch, unsigned long long* out_keys, unsigned long long* out_slots, cudaStream_t st);
void launch_gather_rows(const unsigned long long* keys, const unsigned long long* slots, const unsigned long long* idx, uint64_t k, int ns,
                        unsigned long long* out_keys, unsigned long long* out_slots, cudaStream_t st);
// Large results: turn compacted (key, slots) rows into the result's columns on the device, so the host
// only receives finished column arrays (dimension ids / integer values, timestamps, aggregate columns).
enum MatMode : int32_t { MAT_I64 = 0, MAT_F64_BITS = 1, MAT_F64_ORD = 2, MAT_I64_TIMES = 3 };
struct MatParams {
  uint64_t n;
  int32_t ns, n_dims, n_aggs, has_tk;
  uint64_t tk_stride;
  uint64_t stride[MAX_DIMS], card[MAX_DIMS];
  void* dim_out[MAX_DIMS];      // uint32 ids or int64 values
  int32_t dim_is64[MAX_DIMS];
  uint32_t* tk_out;             // time key per row (ids of time-format dimensions), or NULL
  long long* ts_out;            // bucket start per row, or NULL
  const long long* bucket_lut;  // [n_tk]
  int32_t agg_slot[MAX_AGGS], agg_mode[MAX_AGGS];
  long long agg_mul[MAX_AGGS];
  void* agg_out[MAX_AGGS];      // int64 or double columns
};
void launch_materialize(const unsigned long long* keys, const unsigned long long* slots, const MatParams& mp, cudaStream_t st);

// sparse multi-GPU merge helpers
void launch_partition_rows(const unsigned long long* keys, const unsigned long long* slots, uint64_t n, int n_slots, int n_ranks,
                           unsigned long long* out_keys, unsigned long long* out_slots, unsigned long long* counts_dev,
                           unsigned long long* cursors_dev, cudaStream_t st, unsigned long long key_range = 0);  // key_range > 0: owner = key / key_range, else by hash
// dense multi-GPU merge: fold n_ranks gathered blobs (table + HLL registers each) into this rank's table
void launch_merge_gathered(const unsigned long long* gathered, int n_ranks, size_t blob_bytes, size_t table_bytes, uint64_t cells, int n_slots,
                           const ScanParams& p, unsigned long long* table, cudaStream_t st);
// [key][slot] <-> [slot][key] (the layout a per-slot ncclReduce needs)
void launch_slot_transpose(const unsigned long long* in, unsigned long long* out, uint64_t n_keys, int n_slots, bool to_slot_major, cudaStream_t st);
// implicit_count: the rows are raw rows (STRAT_EMIT): n_slots - 1 words each, slot 0 += 1 per row
void launch_merge_rows(const unsigned long long* keys, const unsigned long long* slots, uint64_t n, const ScanParams& p, cudaStream_t st, bool implicit_count = false);
// the same into a direct-indexed table of p.n_keys rows: row of key k = (k - key_off) / key_mul (keys that do not map raise the overflow flag)
void launch_merge_rows_dense(const unsigned long long* keys, const unsigned long long* slots, uint64_t n, const ScanParams& p, unsigned long long key_off,
                             unsigned long long key_mul, bool implicit_count, cudaStream_t st);

// select queries: per-unit counts of passing rows; then, for the units that hold the wanted page, the passing rows with
// rank in [skip, skip + take) in row order; then a gather of each requested column
struct SelectUnit { uint32_t unit, skip, take, out_base; };
void launch_select_count(const ScanParams& p, uint32_t* counts, cudaStream_t st);
void launch_select_emit(const ScanParams& p, const SelectUnit* units, uint32_t n_units, uint32_t* out_seg, uint32_t* out_row, cudaStream_t st);
void launch_select_gather(const ScanParams& p, int col, int width, const uint32_t* seg, const uint32_t* row, uint32_t n, void* out, cudaStream_t st);

#endif  // !__CUDACC_RTC__

}  // namespace sdo
