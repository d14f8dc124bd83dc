/* CPU oracle, C restatement (TEST / BENCH INFRASTRUCTURE ONLY -- never linked into the product).
 *
 * What Spark's un-rewritten plan computes for one GroupBy/Timeseries/TopN query over the raw flattened
 * table: interval + predicate filter -> hash aggregate (HashAggregateExec) -> per-group count / sum / min /
 * max / HyperLogLog++.  Restated from the same sources as oracle/druid_oracle.py (see its header for the
 * reference file:line list); cross-checked against it in tests/test_c_oracle_cpu.py; used by bench.py as the
 * timed CPU arm (`kind: "port"`) because neither Spark nor a JVM exists on the box.
 *
 * Threads partition the rows; every thread aggregates into its own open-addressing table keyed by the packed
 * group key; tables are merged at the end.  Doubles are summed with Neumaier compensation (error far below
 * the 1e-9 parity tolerance), longs exactly.  HLL++: p = 9, XXH64 seed 42 of the typed value.
 *
 * PARITY: unpinned for aggregate values (no reference-run outputs exist offline); XXH64 pinned by vectors.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define HLL_M 512
#define P1 0x9E3779B185EBCA87ULL
#define P2 0xC2B2AE3D27D4EB4FULL
#define P3 0x165667B19E3779F9ULL
#define P4 0x85EBCA77C2B2AE63ULL
#define P5 0x27D4EB2F165667C5ULL

static inline uint64_t rotl(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
static inline uint64_t fmix(uint64_t h) { h ^= h >> 33; h *= P2; h ^= h >> 29; h *= P3; h ^= h >> 32; return h; }
/* Spark XXH64.hashInt / hashLong = XXH64 of the 4 / 8 little-endian bytes */
uint64_t oc_xxh64_int(int32_t v, uint64_t seed) {
  uint64_t h = seed + P5 + 4;
  h ^= (uint64_t)(uint32_t)v * P1;
  h = rotl(h, 23) * P2 + P3;
  return fmix(h);
}
uint64_t oc_xxh64_long(int64_t v, uint64_t seed) {
  uint64_t h = seed + P5 + 8;
  h ^= rotl((uint64_t)v * P2, 31) * P1;
  h = rotl(h, 27) * P1 + P4;
  return fmix(h);
}
static inline void hll_idx_rho(uint64_t h, int* idx, int* rho) {
  *idx = (int)(h >> 55);
  uint64_t w = (h << 9) | (1ULL << 8);
  *rho = __builtin_clzll(w) + 1;
}

/* ---- query description (filled from Python) ------------------------------------------------------- */
enum { LEAF_TABLE = 0, LEAF_RANGE = 1 };           /* table[code] != 0, or lo <= value <= hi */
enum { COL_I32 = 0, COL_I64 = 1 };
enum { OP_LEAF = 0, OP_AND = 1, OP_OR = 2, OP_NOT = 3 };
enum { AGG_COUNT = 0, AGG_LSUM = 1, AGG_DSUM = 2, AGG_LMIN = 3, AGG_LMAX = 4, AGG_DMIN = 5, AGG_DMAX = 6, AGG_HLL = 7 };

typedef struct {
  int kind, col_type;
  const void* col;
  const uint8_t* table;
  int64_t lo, hi;
} oc_leaf;
typedef struct {
  int col_type;           /* COL_I32 codes or COL_I64 values */
  const void* col;
  int64_t base;           /* value - base is the key component */
  int64_t card;           /* component range */
} oc_dim;
typedef struct {
  int kind;
  int col_is_f64;         /* metric column type */
  const void* col;        /* metric, or the dimension column for HLL */
  int hll_col_type;       /* COL_I32 (codes -> packed table) or COL_I64 values */
  int hll_src;            /* 1 = hash as int, 2 = hash as long (integer dimensions) */
  const uint16_t* hll_table; /* per-code (idx << 6 | rho) for string dimensions */
  int filter_leaf;        /* filtered aggregator: leaf index or -1 */
} oc_agg;
typedef struct {
  const int64_t* time;
  int64_t n_rows;
  int n_intervals;
  const int64_t* intervals;   /* pairs [start, end) */
  int n_leaves, n_ops;
  const oc_leaf* leaves;
  const uint8_t* op_code;
  const int32_t* op_arg;
  int n_dims;
  const oc_dim* dims;
  /* time bucket: bucket id = lut[(t - t0) / unit], or no bucket */
  int has_bucket;
  int64_t t0, unit, n_lut, n_buckets;
  const int32_t* lut;
  int n_aggs;
  const oc_agg* aggs;
  int n_threads;
} oc_query;

typedef struct {
  uint64_t key;
  int64_t count;
} oc_slot_hdr;

typedef struct {
  uint64_t cap, used;
  uint64_t* keys;       /* ~0 = empty */
  int64_t* counts;
  double* vals;         /* [cap][n_aggs][2]  (value, compensation) ; longs stored in the bits of value */
  uint8_t* hll;         /* [cap][n_hll][512] */
  int n_aggs, n_hll;
} oc_table;

static uint64_t mix(uint64_t x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33; return x; }

static void table_init(oc_table* t, uint64_t cap, int n_aggs, int n_hll) {
  t->cap = cap; t->used = 0; t->n_aggs = n_aggs; t->n_hll = n_hll;
  t->keys = (uint64_t*)malloc(cap * 8);
  memset(t->keys, 0xff, cap * 8);
  t->counts = (int64_t*)calloc(cap, 8);
  t->vals = (double*)calloc(cap * (size_t)(n_aggs ? n_aggs : 1) * 2, 8);
  t->hll = n_hll ? (uint8_t*)calloc(cap * (size_t)n_hll, HLL_M) : NULL;
}
static void table_free(oc_table* t) { free(t->keys); free(t->counts); free(t->vals); free(t->hll); }

static void slot_init(oc_table* t, uint64_t s, const oc_query* q) {
  for (int a = 0; a < q->n_aggs; a++) {
    double* v = t->vals + (s * q->n_aggs + a) * 2;
    int64_t iv = 0;
    switch (q->aggs[a].kind) {
      case AGG_LMIN: iv = INT64_MAX; memcpy(v, &iv, 8); break;
      case AGG_LMAX: iv = INT64_MIN; memcpy(v, &iv, 8); break;
      case AGG_DMIN: v[0] = INFINITY; break;
      case AGG_DMAX: v[0] = -INFINITY; break;
      default: v[0] = 0; break;
    }
    v[1] = 0;
  }
}
static uint64_t table_find(oc_table* t, uint64_t key, const oc_query* q);
static void table_grow(oc_table* t, const oc_query* q) {
  oc_table n;
  table_init(&n, t->cap * 2, t->n_aggs, t->n_hll);
  for (uint64_t s = 0; s < t->cap; s++) {
    if (t->keys[s] == ~0ULL) continue;
    uint64_t h = mix(t->keys[s]) & (n.cap - 1);
    while (n.keys[h] != ~0ULL) h = (h + 1) & (n.cap - 1);
    n.keys[h] = t->keys[s];
    n.counts[h] = t->counts[s];
    memcpy(n.vals + h * (size_t)t->n_aggs * 2, t->vals + s * (size_t)t->n_aggs * 2, (size_t)t->n_aggs * 16);
    if (t->n_hll) memcpy(n.hll + h * (size_t)t->n_hll * HLL_M, t->hll + s * (size_t)t->n_hll * HLL_M, (size_t)t->n_hll * HLL_M);
    n.used++;
  }
  table_free(t);
  *t = n;
  (void)q;
}
static uint64_t table_find(oc_table* t, uint64_t key, const oc_query* q) {
  if (t->used * 2 >= t->cap) table_grow(t, q);
  uint64_t h = mix(key) & (t->cap - 1);
  for (;;) {
    if (t->keys[h] == key) return h;
    if (t->keys[h] == ~0ULL) {
      t->keys[h] = key;
      t->used++;
      slot_init(t, h, q);
      return h;
    }
    h = (h + 1) & (t->cap - 1);
  }
}

static inline void neumaier(double* v, double x) {
  double s = v[0] + x;
  if (fabs(v[0]) >= fabs(x)) v[1] += (v[0] - s) + x;
  else v[1] += (x - s) + v[0];
  v[0] = s;
}

static inline int leaf_ok(const oc_leaf* l, int64_t r) {
  int64_t v = l->col_type == COL_I32 ? ((const int32_t*)l->col)[r] : ((const int64_t*)l->col)[r];
  if (l->kind == LEAF_TABLE) return l->table[v] != 0;
  return v >= l->lo && v <= l->hi;
}

typedef struct {
  const oc_query* q;
  int64_t r0, r1;
  oc_table t;
} oc_work;

static void* worker(void* arg) {
  oc_work* w = (oc_work*)arg;
  const oc_query* q = w->q;
  int n_hll = 0;
  for (int a = 0; a < q->n_aggs; a++) n_hll += q->aggs[a].kind == AGG_HLL;
  table_init(&w->t, 1024, q->n_aggs, n_hll);
  for (int64_t r = w->r0; r < w->r1; r++) {
    const int64_t t = q->time[r];
    int in = q->n_intervals == 0;
    for (int i = 0; i < q->n_intervals && !in; i++) in = t >= q->intervals[2 * i] && t < q->intervals[2 * i + 1];
    if (!in) continue;
    if (q->n_ops) {
      int stack[32], sp = 0;
      for (int k = 0; k < q->n_ops; k++) {
        switch (q->op_code[k]) {
          case OP_LEAF: stack[sp++] = leaf_ok(&q->leaves[q->op_arg[k]], r); break;
          case OP_AND: sp--; stack[sp - 1] = stack[sp - 1] && stack[sp]; break;
          case OP_OR: sp--; stack[sp - 1] = stack[sp - 1] || stack[sp]; break;
          default: stack[sp - 1] = !stack[sp - 1]; break;
        }
      }
      if (!stack[0]) continue;
    }
    uint64_t key = 0;
    if (q->has_bucket) key = (uint64_t)q->lut[(t - q->t0) / q->unit];
    for (int d = 0; d < q->n_dims; d++) {
      const oc_dim* dm = &q->dims[d];
      int64_t v = dm->col_type == COL_I32 ? ((const int32_t*)dm->col)[r] : ((const int64_t*)dm->col)[r];
      key = key * (uint64_t)dm->card + (uint64_t)(v - dm->base);
    }
    uint64_t s = table_find(&w->t, key, q);
    w->t.counts[s]++;
    int hi = 0;
    for (int a = 0; a < q->n_aggs; a++) {
      const oc_agg* ag = &q->aggs[a];
      int this_hll = ag->kind == AGG_HLL ? hi++ : -1;
      if (ag->filter_leaf >= 0 && !leaf_ok(&q->leaves[ag->filter_leaf], r)) continue;
      double* v = w->t.vals + (s * q->n_aggs + a) * 2;
      double dv = 0;
      int64_t lv = 0;
      if (ag->kind >= AGG_LSUM && ag->kind <= AGG_DMAX) {
        if (ag->col_is_f64) { dv = ((const double*)ag->col)[r]; lv = (int64_t)dv; }
        else { lv = ((const int64_t*)ag->col)[r]; dv = (double)lv; }
      }
      int64_t cur;
      switch (ag->kind) {
        case AGG_COUNT: memcpy(&cur, v, 8); cur++; memcpy(v, &cur, 8); break;
        case AGG_LSUM: memcpy(&cur, v, 8); cur += lv; memcpy(v, &cur, 8); break;
        case AGG_DSUM: neumaier(v, dv); break;
        case AGG_LMIN: memcpy(&cur, v, 8); if (lv < cur) memcpy(v, &lv, 8); break;
        case AGG_LMAX: memcpy(&cur, v, 8); if (lv > cur) memcpy(v, &lv, 8); break;
        case AGG_DMIN: if (dv < v[0]) v[0] = dv; break;
        case AGG_DMAX: if (dv > v[0]) v[0] = dv; break;
        case AGG_HLL: {
          int idx, rho;
          if (ag->hll_col_type == COL_I32 && ag->hll_table) {
            uint16_t pk = ag->hll_table[((const int32_t*)ag->col)[r]];
            idx = pk >> 6; rho = pk & 63;
          } else {
            int64_t x = ag->hll_col_type == COL_I32 ? ((const int32_t*)ag->col)[r] : ((const int64_t*)ag->col)[r];
            hll_idx_rho(ag->hll_src == 1 ? oc_xxh64_int((int32_t)x, 42) : oc_xxh64_long(x, 42), &idx, &rho);
          }
          uint8_t* reg = w->t.hll + (s * (size_t)w->t.n_hll + this_hll) * HLL_M + idx;
          if (*reg < rho) *reg = (uint8_t)rho;
          break;
        }
      }
    }
  }
  return NULL;
}

/* Result handle */
typedef struct {
  oc_table t;
  int n_aggs, n_hll;
  int kinds[64];
} oc_result;

static void merge_into(oc_table* dst, const oc_table* src, const oc_query* q) {
  for (uint64_t s = 0; s < src->cap; s++) {
    if (src->keys[s] == ~0ULL) continue;
    uint64_t d = table_find(dst, src->keys[s], q);
    dst->counts[d] += src->counts[s];
    int hi = 0;
    for (int a = 0; a < q->n_aggs; a++) {
      double* dv = dst->vals + (d * q->n_aggs + a) * 2;
      const double* sv = src->vals + (s * q->n_aggs + a) * 2;
      int64_t x, y;
      switch (q->aggs[a].kind) {
        case AGG_COUNT: case AGG_LSUM: memcpy(&x, dv, 8); memcpy(&y, sv, 8); x += y; memcpy(dv, &x, 8); break;
        case AGG_DSUM: neumaier(dv, sv[0]); dv[1] += sv[1]; break;
        case AGG_LMIN: memcpy(&x, dv, 8); memcpy(&y, sv, 8); if (y < x) memcpy(dv, &y, 8); break;
        case AGG_LMAX: memcpy(&x, dv, 8); memcpy(&y, sv, 8); if (y > x) memcpy(dv, &y, 8); break;
        case AGG_DMIN: if (sv[0] < dv[0]) dv[0] = sv[0]; break;
        case AGG_DMAX: if (sv[0] > dv[0]) dv[0] = sv[0]; break;
        case AGG_HLL: {
          uint8_t* dr = dst->hll + (d * (size_t)dst->n_hll + hi) * HLL_M;
          const uint8_t* sr = src->hll + (s * (size_t)src->n_hll + hi) * HLL_M;
          for (int j = 0; j < HLL_M; j++) if (sr[j] > dr[j]) dr[j] = sr[j];
          hi++;
          break;
        }
      }
    }
  }
}

oc_result* oc_execute(const oc_query* q) {
  int nt = q->n_threads < 1 ? 1 : (q->n_threads > 256 ? 256 : q->n_threads);
  if ((int64_t)nt > q->n_rows / 1024 + 1) nt = (int)(q->n_rows / 1024 + 1);
  oc_work* w = (oc_work*)calloc((size_t)nt, sizeof(oc_work));
  pthread_t* th = (pthread_t*)calloc((size_t)nt, sizeof(pthread_t));
  int64_t per = (q->n_rows + nt - 1) / nt;
  for (int i = 0; i < nt; i++) {
    w[i].q = q;
    w[i].r0 = per * i < q->n_rows ? per * i : q->n_rows;
    w[i].r1 = per * (i + 1) < q->n_rows ? per * (i + 1) : q->n_rows;
    pthread_create(&th[i], NULL, worker, &w[i]);
  }
  for (int i = 0; i < nt; i++) pthread_join(th[i], NULL);
  oc_result* res = (oc_result*)calloc(1, sizeof(oc_result));
  res->t = w[0].t;
  for (int i = 1; i < nt; i++) { merge_into(&res->t, &w[i].t, q); table_free(&w[i].t); }
  res->n_aggs = q->n_aggs;
  res->n_hll = res->t.n_hll;
  for (int a = 0; a < q->n_aggs && a < 64; a++) res->kinds[a] = q->aggs[a].kind;
  free(w);
  free(th);
  return res;
}
int64_t oc_result_groups(const oc_result* r) { return (int64_t)r->t.used; }
/* keys[n], counts[n], vals[n][n_aggs] (doubles; long aggregates as int64 bits), hll[n][n_hll][512] */
void oc_result_fetch(const oc_result* r, uint64_t* keys, int64_t* counts, double* vals, uint8_t* hll) {
  uint64_t o = 0;
  for (uint64_t s = 0; s < r->t.cap; s++) {
    if (r->t.keys[s] == ~0ULL) continue;
    keys[o] = r->t.keys[s];
    counts[o] = r->t.counts[s];
    for (int a = 0; a < r->n_aggs; a++) {
      const double* v = r->t.vals + (s * r->n_aggs + a) * 2;
      vals[o * r->n_aggs + a] = r->kinds[a] == AGG_DSUM ? v[0] + v[1] : v[0];
    }
    if (r->n_hll && hll) memcpy(hll + o * (size_t)r->n_hll * HLL_M, r->t.hll + s * (size_t)r->n_hll * HLL_M, (size_t)r->n_hll * HLL_M);
    o++;
  }
}
void oc_result_free(oc_result* r) { table_free(&r->t); free(r); }
