/*
 * sdo_engine.h -- C ABI of the B200 OLAP-index engine (libsdo_engine.so).
 *
 * This is the drop-in boundary for the one call the Sparkline accelerator makes into its
 * OLAP index: DruidRDD.compute -> mQry.executeQuery(client)
 *   (/root/reference/src/main/scala/org/sparklinedata/druid/DruidRDD.scala:180,
 *    QuerySpec.executeQuery, DruidQuerySpec.scala:598-599,
 *    DruidQueryServerClient.executeQueryAsStream, client/DruidClient.scala:448-451).
 * Everything behind that call in the reference (HTTP to a Druid broker/historical, the Druid
 * scan/filter/group-by engines, JSON result streaming) is replaced by these entry points.
 * A JNI shim binds them 1:1; tests and bench bind them with ctypes.
 *
 * Conventions: plain pointers and sizes only; no exceptions cross the ABI; every function
 * returns an sdo_status (0 = OK) unless stated; strings are UTF-8, NUL-terminated unless a
 * length is passed; all returned pointers stay owned by the engine/result they came from.
 * Threading: one engine handle may be used from many threads at once; sdo_query_execute is
 * re-entrant (each call runs on its own CUDA stream with its own scratch), as many Spark tasks
 * share one executor JVM (reference pool: 100 connections, DruidPlanner.scala:83-90).
 */
#ifndef SDO_ENGINE_H
#define SDO_ENGINE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sdo_engine sdo_engine;
typedef struct sdo_result sdo_result;
typedef struct sdo_partial sdo_partial;

/* Status codes.  Any non-zero status maps to DruidDataSourceException on the Scala side
 * (/root/reference/src/main/scala/org/sparklinedata/druid/DruidDataSourceException.scala),
 * except SDO_UNSUPPORTED, which tells the caller to leave the plan un-rewritten (spatial filters,
 * descending select, javascript outside the expression subset: SURVEY.md 8b "Errors"), and
 * SDO_CANCELLED, which the caller
 * swallows as DruidRDD does (DruidRDD.scala:181-186). */
typedef enum {
  SDO_OK = 0,
  SDO_ERROR = 1,
  SDO_UNSUPPORTED = 2,
  SDO_INVALID = 3,
  SDO_NOT_FOUND = 4,
  SDO_CANCELLED = 5,
  SDO_NO_DEVICE = 6,
  SDO_OUT_OF_MEMORY = 7
} sdo_status;

/* ---- engine lifetime ---------------------------------------------------------------------
 * Replaces the per-task DruidQueryServerClient construction (DruidRDD.scala:175-177) and the
 * connection pool set-up in DruidPlanner.apply (DruidPlanner.scala:54-58).
 * device >= 0: CUDA device ordinal; queries run there.  device < 0: planning-only engine (spec
 * parsing, segment metadata, sdo_query_explain); sdo_query_execute then fails with
 * SDO_NO_DEVICE -- there is no CPU execution path. */
int sdo_engine_create(int device, sdo_engine** out);
void sdo_engine_destroy(sdo_engine* e);
/* Last error message of the calling thread ("" if none).  Replaces the message carried by
 * DruidDataSourceException (client/DruidClient.scala:206-224,266-283). */
const char* sdo_last_error(void);
/* Engine build/version string; "0.9.1-b200" style so that DruidDataSource.supportsBoundFilter
 * (metadata/DruidDataSource.scala:94-97, needs >= 0.9.0) holds. */
const char* sdo_version(void);

/* ---- load path ----------------------------------------------------------------------------
 * New (the reference delegates ingestion to a Druid index task, client/DruidOverlordClient.scala:
 * 51-129; the segment schema follows src/test/resources/tpch_index_task.json.template).
 * sdo_datasource_create takes options as JSON:
 *   {"segmentGranularity":"MONTH"|"DAY"|"YEAR"|"ALL", "queryGranularity":"DAY"|"NONE"|"HOUR"|...}
 * (defaults MONTH / DAY, as tpch_index_task.json.template:175-176). */
int sdo_datasource_create(sdo_engine* e, const char* name, const char* options_json);
int sdo_datasource_drop(sdo_engine* e, const char* name);

typedef enum { SDO_COL_TIME = 0, SDO_COL_DIM = 1, SDO_COL_METRIC = 2 } sdo_col_kind;
typedef enum {
  SDO_T_INT64 = 0,   /* data: int64[n_rows]   (time in epoch ms; integer dimension; long metric) */
  SDO_T_FLOAT64 = 1, /* data: double[n_rows]  (double metric) */
  SDO_T_STRING = 2,  /* data: int64 offsets[n_rows+1] into bytes (string dimension, row-wise) */
  SDO_T_CODES = 3,   /* data: int32 codes[n_rows] into a caller dictionary (dict_offsets[n_dict+1], bytes) */
  SDO_T_INT32 = 4    /* data: int32[n_rows]   (integer dimension) */
} sdo_col_type;
/* Spark type of a dimension's source column; decides what approx_count_distinct hashes
 * (XxHash64 of int vs long vs UTF-8 bytes), SURVEY.md 7 "hard parts". */
typedef enum { SDO_SRC_STRING = 0, SDO_SRC_INT32 = 1, SDO_SRC_INT64 = 2 } sdo_src_type;

typedef struct {
  const char* name;
  int kind;                    /* sdo_col_kind */
  int type;                    /* sdo_col_type */
  const void* data;
  const char* bytes;           /* STRING: row bytes; CODES: dictionary bytes */
  const int64_t* dict_offsets; /* CODES only */
  int64_t n_dict;              /* CODES only */
  int src_type;                /* sdo_src_type (dimensions) */
} sdo_column;

/* Append n_rows rows (any time order, any time range) to a datasource: the engine truncates
 * time to the query granularity, sorts by time, cuts month segments, dictionary-encodes the
 * dimensions (ids 1/2/4 bytes wide by cardinality) and moves the columns into HBM.  Every call
 * creates one new partition ("shard") per touched segment interval; `partition_num` is recorded
 * in the segment metadata (DruidSegmentInfo.shardSpec.partitionNum,
 * metadata/DruidMetadataCache.scala:64-76).  Column buffers are borrowed for the call only. */
int sdo_segment_load(sdo_engine* e, const char* datasource, int64_t n_rows, int n_cols,
                     const sdo_column* cols, int partition_num);

/* ---- metadata -----------------------------------------------------------------------------
 * Replace DruidClient.metadata / timeBoundary / dataSourceInfo / serverStatus
 * (client/DruidClient.scala:337-364,402-415,495-499).  JSON shapes follow MetadataResponse and
 * DruidSegmentInfo (client/DruidMessages.scala:25-53, metadata/DruidMetadataCache.scala:64-76).
 * The returned string is owned by the calling thread's slot and valid until its next call. */
int sdo_datasource_metadata(sdo_engine* e, const char* name, const char** json_out);
int sdo_datasource_segments(sdo_engine* e, const char* name, const char** json_out);

/* ---- query --------------------------------------------------------------------------------
 * sdo_query_execute replaces QuerySpec.executeQuery (DruidQuerySpec.scala:598-599): the JSON is
 * exactly what DruidClient.executeQueryAsStream renders (client/DruidClient.scala:328-331) with
 * Utils.jsonFormat (Utils.scala:38-105).  queryType in {groupBy, timeseries, topN, search, select}.
 *   search (SearchQuerySpec, DruidQuerySpec.scala:870-927; what the planner emits for a group-by on one
 *     dimension without aggregations): result columns dimension, value, count.
 *   select (SelectSpec, :980-1098): one page of rows per call; result columns timestamp, segmentId,
 *     offset, then the requested dimensions and metrics; the pagingIdentifiers for the next page are in
 *     sdo_result_stats_json (pass them back in pagingSpec with fromNext = true).
 *   javascript filters / extraction functions / aggregations are accepted when the function is a single
 *     `return <expression>` (comparisons, string methods, arithmetic); nothing of it runs per row.
 *   context keys read: queryId, skipEmptyBuckets, gpuTimeBoundary, and the test knobs gpuStrategy,
 *     gpuJit (true: wait for the shape's specialised kernel instead of running ahead-of-time code while it compiles),
 *     gpuHllPlace, gpuTicketRows, gpuShardMerge; gpuStrategy emit = exchange raw rows in a sharded query; unknown keys are ignored.
 *     skipEmptyBuckets defaults to TRUE here (Druid: false): the parity target is Spark running the
 *     un-rewritten plan, which has no rows for empty buckets.  An explicit false is honoured for
 *     timeseries queries (one row per granularity bucket between the datasource's first and last
 *     timestamp inside the query intervals, empty ones with the aggregators' identities; at most
 *     65536 buckets); groupBy never zero-fills, as in Druid. */
int sdo_query_execute(sdo_engine* e, const char* query_json, size_t len, sdo_result** out);
/* Query shapes whose specialised scan kernel is still compiling on a background thread.  Queries of such a shape run on
 * the ahead-of-time kernel meanwhile (no query waits for the compiler unless its context says gpuJit: true); a benchmark
 * that wants steady-state numbers waits for 0.  No reference counterpart. */
int sdo_jit_pending(void);
/* Plan only (no device work): returns a JSON description of the spans, compiled filter,
 * key layout, strategy and algorithmic bytes.  No reference counterpart beyond
 * EXPLAIN DRUID REWRITE (sparklinedata/commands/DruidMetadataCommands.scala:49). */
int sdo_query_explain(sdo_engine* e, const char* query_json, size_t len, const char** json_out);
/* Cancel a running query by context.queryId from another thread.  Replaces
 * CancellableHolder / TaskCancelHandler (client/DruidClient.scala:82-103, DruidRDD.scala:428-491).
 * The query fails with SDO_CANCELLED at its next poll: before its scan is queued, or after the scan
 * has been queued and before results are read back (a launched scan kernel is not interrupted).
 * A cancel for an id that is not running is remembered for 5 seconds (the canceller may race ahead
 * of the execute call) and then forgotten, and a mark dies with the query it names, so a finished
 * query's id can be reused by a retry (DruidRDD.scala:80 reuses <queryId>-<partition>). */
int sdo_query_cancel(sdo_engine* e, const char* query_id);

/* ---- result -------------------------------------------------------------------------------
 * Replaces the ResultRow stream (client/DruidMessages.scala:67-153) and the iterators
 * (sources/druid/DruidQueryResultIterator.scala:58-90, DruidTopNResultIterator.scala:32-74).
 * Columnar, in pinned host memory; column order = [timestamp if granularity != all], dimensions
 * (outputName), aggregations (name), post-aggregations (name)  (DruidQuerySpec.scala:614-627). */
typedef enum {
  SDO_RES_INT64 = 0,   /* values: int64[n_rows] */
  SDO_RES_FLOAT64 = 1, /* values: double[n_rows] */
  SDO_RES_STRING = 2,  /* values: uint32 ids[n_rows] into (dict_offsets, dict_bytes) */
  SDO_RES_DIM_INT64 = 3, /* integer-valued dimension: values int64[n_rows]; its Druid value is the decimal string */
  SDO_RES_TIMESTAMP = 4  /* bucket start, epoch ms, int64[n_rows] */
} sdo_res_type;

typedef struct {
  const char* name;
  int type;                    /* sdo_res_type */
  const void* values;
  const int64_t* dict_offsets; /* STRING */
  const char* dict_bytes;      /* STRING */
  int64_t n_dict;              /* STRING */
} sdo_result_column;

int64_t sdo_result_num_rows(const sdo_result* r);
int sdo_result_num_columns(const sdo_result* r);
int sdo_result_column_get(const sdo_result* r, int i, sdo_result_column* out);
/* HyperLogLog++ registers (p = 9: 512 bytes per row) behind cardinality column `i`, so the
 * Scala side can merge partitions or run Spark's own estimator; NULL for other columns. */
const uint8_t* sdo_result_hll_registers(const sdo_result* r, int i);
/* Per-query execution facts as JSON: strategy, kernel/total times (CUDA events), rows scanned,
 * algorithmic bytes, launches.  Feeds DruidQueryExecutionView (DruidRDD.scala:194-216). */
const char* sdo_result_stats_json(const sdo_result* r);
/* Release a result.  Replaces closing the HTTP entity (client/DruidClient.scala:160-166). */
void sdo_result_free(sdo_result* r);

/* ---- sharded execution (one engine per GPU, partial aggregates merged by the host layer) ---
 * Replaces per-historical partial queries + Spark re-aggregation (DruidRDD.scala:244-269,
 * sources/druid/DruidStrategy.scala:353-359, sources/druid/PostAggregate.scala:39-60).
 * sdo_query_partial runs the scan/aggregate on this engine's segments and leaves the partial
 * aggregate state on the device; the host layer merges the exposed device buffers across ranks
 * (NCCL sum/min/max on dense tables; hash-partitioned all-to-all for sparse ones), then
 * sdo_partial_finish turns the merged state into a result. */
typedef struct {
  int dense;            /* 1: fixed-layout table indexed by packed key; 0: (key, value) rows */
  int64_t n_keys;       /* dense: key-space size; sparse: number of local groups */
  int n_slots;          /* 8-byte aggregate slots per key */
  void* keys;           /* sparse: device uint64[n_keys] packed keys; dense: NULL */
  void* slots;          /* device [n_keys * n_slots] 8-byte slots, slot-major per key */
  const int* slot_ops;  /* host int[n_slots]: 0 = int64 sum, 1 = float64 sum, 2 = int64 min, 3 = int64 max */
  void* hll;            /* device uint8[n_keys * n_hll * 512] registers or NULL */
  int n_hll;
  int64_t blob_bytes;   /* dense: table and registers are one contiguous block of this many bytes starting at `slots` */
} sdo_partial_view;

/* Shards of one datasource that live in different engines must share one id space per dimension before
 * their partials can be merged: a string dimension needs the same sorted dictionary everywhere, an integer
 * dimension the same cardinality.  (A Druid broker merges on string values, so the reference needs no such
 * step; here the merge is on packed ids.)  sdo_dimension_dictionary exposes a string dimension's dictionary
 * (n entries, offsets[n+1] into bytes; integer dimensions report n = 0; pointers stay valid until the
 * dimension is next extended or the datasource dropped); sdo_dimension_extend merges `n_values` strings
 * into it -- re-encoding resident segments on the device when ids move -- or raises an integer dimension's
 * cardinality to at least `min_cardinality`. */
int sdo_dimension_dictionary(sdo_engine* e, const char* datasource, const char* dimension, int64_t* n_out,
                             const int64_t** offsets_out, const char** bytes_out, int64_t* cardinality_out);
int sdo_dimension_extend(sdo_engine* e, const char* datasource, const char* dimension, int64_t n_values,
                         const int64_t* value_offsets, const char* value_bytes, int64_t min_cardinality);

int sdo_query_partial(sdo_engine* e, const char* query_json, size_t len, sdo_partial** out);
int sdo_partial_view_get(sdo_partial* p, sdo_partial_view* out);
/* Dense path, small tables: copy this rank's blob (blob_bytes) into a caller-owned device buffer (so it can be
 * all-gathered in one collective whatever mix of sum / min / max slots it holds), then fold the n_ranks gathered
 * blobs, in rank order, back into this rank's table.  Both only enqueue work on the partial's stream. */
int sdo_partial_export(sdo_partial* p, void* dst_dev, int64_t dst_bytes);
int sdo_partial_merge_gathered(sdo_partial* p, const void* gathered_dev, int n_ranks);
/* Sparse path: split local groups by owner = mix(key) % n_ranks into n_ranks contiguous runs
 * (counts_out[n_ranks], host), reordering keys/slots in place on the device. */
int sdo_partial_partition(sdo_partial* p, int n_ranks, int64_t* counts_out);
/* Sparse path: replace the local state by the merge of `n` received (key, slots) rows. */
int sdo_partial_merge_rows(sdo_partial* p, const void* keys_dev, const void* slots_dev, int64_t n);
/* The CUDA stream (cudaStream_t) the partial's device work was queued on; the host layer must
 * order its collectives after it. */
void* sdo_partial_stream(sdo_partial* p);
int sdo_partial_finish(sdo_partial* p, sdo_result** out);
void sdo_partial_free(sdo_partial* p);

/* ---- persistence ------------------------------------------------------------------------------------
 * Write a datasource as it sits in HBM -- dictionaries, encoded id / metric columns, time run index -- to one file, and
 * bring it back under `name` (which must not exist) without dictionary encoding or time sort.  Stands in for the index
 * the reference's Druid keeps on deep storage between restarts (written once by the indexing service,
 * SLD/client/DruidOverlordClient.scala:51-129).  The file is private to this library version (magic "SDOSEG01"). */
int sdo_datasource_save(sdo_engine* e, const char* name, const char* path);
int sdo_datasource_restore(sdo_engine* e, const char* name, const char* path);

/* ---- sharded execution with the merge inside the library ---------------------------------------
 * One engine per GPU, one process per engine.  Replaces the Spark-side re-aggregation of the per-historical partial
 * results (SQD/DruidStrategy.scala:353-359, SQD/PostAggregate.scala:39-103; partitions from SLD/DruidRDD.scala:244-269):
 * scan, the merge of the partial aggregates over NCCL and the finish are queued on one stream by one call, without a
 * round trip through the host language between them.  The merge is chosen from the query's key space (the same on every
 * rank once dictionaries are aligned), SURVEY 8e:
 *   gather : dense table + HLL registers up to 256 KB -- one ncclAllGather of the raw partial blobs, folded in rank
 *            order on `root` (bit-reproducible);
 *   reduce : larger dense tables (up to 64 MB) -- the table is transposed to [slot][key] and every slot reduced to `root`
 *            with its own operator (ncclReduce int64 / float64 sum, int64 min / max), HLL registers as uint8 max, one
 *            NCCL group;
 *   sparse : key spaces beyond that.  Such key spaces have about as many groups as rows on every rank, so nothing is
 *            aggregated locally: the scan writes every surviving row (key + the aggregators' inputs, count implicit)
 *            straight into the region of the rank that owns its key -- keys are dealt round-robin (key mod world) when a
 *            rank's share of the key space fits a direct-indexed table, by hash otherwise --, the regions are exchanged
 *            with grouped ncclSend / ncclRecv and every rank aggregates what it owns once.  A query that asks for the top
 *            K <= 1024 groups by one aggregate (TopN with a numeric metric, limitSpec over one aggregate column, no
 *            having) takes the top K of every share, all-gathers the K * world candidates and finishes on `root`.  Any
 *            other sparse query returns, on EVERY rank, that rank's share of the groups (*out non-NULL everywhere; the
 *            caller unions them, as Spark unions partitions); ordering / limit / having then apply per share.
 *            Cardinality aggregations have no sparse merge (SDO_UNSUPPORTED before anything is queued).
 *            Stats of a sparse merge carry shardPhasesMs (split / exchange / merge / topk device times) and exchangeBytes.
 *
 * sdo_comm_unique_id: on one rank; hand the 128 bytes to every rank (any channel).  `nccl_library` = path of
 *   libnccl.so.2 to load if the process has none loaded yet (NULL: default search path).
 * sdo_engine_comm_init: on every rank, collectively; the engine's device must be the calling thread's current one.
 * sdo_query_execute_sharded: on every rank, collectively, with the same query (ids and time keys must mean the same on
 *   every rank: same dictionaries, context.gpuTimeBoundary set -- what sharded.align_datasource arranges).  On `root`
 *   *out is the merged result; elsewhere *out is NULL (except the sparse shares described above).  The caller issues
 *   sharded queries of one engine one at a time and in the same order on every rank.
 * sdo_query_execute_sharded_at: the same for callers that run several sharded queries at once on one engine
 *   (DruidPlanner.scala:83-90 allows 100 connections): `turn` is the query's position in the communicator's collective
 *   order -- the same number on every rank, consecutive from sdo_engine_comm_next_turn (0 after comm_init).  Scans of
 *   different turns overlap; their collectives are issued in turn order, so two ranks can never pair mismatched
 *   collectives.  A turn whose query fails before its first collective is given up; a turn that never arrives makes the
 *   later ones fail after 120 s.
 * context.gpuShardMerge = gather | reduce | sparse forces the merge (tests). */
#define SDO_COMM_ID_BYTES 128
int sdo_comm_unique_id(void* id_out, const char* nccl_library);
int sdo_engine_comm_init(sdo_engine* e, const void* id, int rank, int world, const char* nccl_library);
int sdo_query_execute_sharded(sdo_engine* e, const char* query_json, size_t len, int root, sdo_result** out);
int sdo_query_execute_sharded_at(sdo_engine* e, const char* query_json, size_t len, int root, int64_t turn, sdo_result** out);
int sdo_engine_comm_next_turn(sdo_engine* e, int64_t* turn_out);

/* ---- allocation-free group-by ----------------------------------------------------------------
 * No counterpart in the reference (it has no native code and no notion of device memory).
 *
 * For callers that may launch work on a GPU but must not allocate on it: a plugin living inside
 * another runtime's fixed device-memory reservation, a host that promises "no cudaMalloc after
 * start-up", code running where allocation calls are not permitted.  sdo_query_execute does not
 * fit them, because every query takes scratch from the CUDA memory pool.
 *
 * Computes, for id column keys_dev[n_rows] (one byte per row, ids < n_keys; larger ids are ignored)
 * and metric values_dev[n_rows]:  counts_out_dev[k] = rows with id k,  sums_out_dev[k] = sum of their
 * values -- the same lane-private shared-memory aggregation a dense GroupBy uses, but with its
 * scratch (per-block partials and a block ticket) in a fixed array compiled into the library, about
 * 0.4 MB that every GPU loading the library holds until exit.  All five buffers are device memory
 * owned by the caller (keys 4-byte aligned, values 16-byte aligned); `cuda_stream` is a cudaStream_t
 * (NULL = default stream); the call only enqueues work and returns.  It performs no cudaMalloc /
 * cudaFree and calls no allocating library routine.
 *
 * Limits: 1 <= n_keys <= SDO_NOALLOC_MAX_KEYS and 0 <= n_rows <= SDO_NOALLOC_MAX_ROWS; anything
 * beyond is refused with SDO_INVALID and a message in sdo_last_error().
 * `device` is made current for the duration of the call only; the calling thread's current device
 * is the same after the call as before it.
 * First call on a device: resolves the scratch array's device address (once per device, under a
 * lock); make that first call before starting any stream capture.
 * Sharing: every call on a device uses the same array, which still holds the last call's partials.
 * The kernel overwrites all words it reads back and the ticket is reset on entry, so calls issued
 * one after another are independent -- but calls that overlap on the GPU (two streams, two
 * threads) race on the array and must be serialised by the caller. */
#define SDO_NOALLOC_MAX_KEYS 24
#define SDO_NOALLOC_MAX_ROWS 4294967296LL
int sdo_groupby_sum_noalloc(int device, const uint8_t* keys_dev, const double* values_dev, int64_t n_rows, int n_keys,
                            int64_t* counts_out_dev, double* sums_out_dev, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* SDO_ENGINE_H */
