// Synthetic flattened TPC-H ("orderLineItemPartSupplier") generator.
//
// The reference ships no data (its fixture blob is stripped), so tests and the
// bench feed the engine and the oracle from this deterministic generator.  The
// schema is the 53-column flat table of
//   /root/reference/src/main/scala/org/sparklinedata/druid/tools/TpchBenchMark.scala:43-97
// minus the free-text columns (comments, addresses, phones, p_name) that no
// query on the hot path references.  Distributions follow SURVEY.md §8d.
//
// Output is columnar and already sorted by ship day (the index's time column,
// tpch_index_task.json.template: timestampSpec l_shipdate), stable in
// (order index, line number), so month segments are contiguous row ranges.
// Every value is a pure function of (seed, order index, line number): the
// result does not depend on the number of threads or on how orders are
// sharded (a shard is a contiguous order-index range, which mirrors the
// index's hashed shards inside a month segment).
//
// This is input production for tests/bench, not part of the query path.

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

typedef uint64_t u64;
typedef int64_t i64;

inline u64 mix64(u64 x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
inline u64 rnd(u64 seed, u64 stream, u64 idx) {
  return mix64(mix64(seed ^ (stream * 0xD6E8FEB86659FD93ull)) + idx);
}
// uniform integer in [lo, hi]
inline i64 uni(u64 r, i64 lo, i64 hi) {
  u64 range = (u64)(hi - lo + 1);
  return lo + (i64)(((__uint128_t)r * range) >> 64);
}

// days since 1970-01-01 for a civil date (proleptic Gregorian)
i64 days_from_civil(int y, int m, int d) {
  y -= m <= 2;
  const i64 era = (y >= 0 ? y : y - 399) / 400;
  const unsigned yoe = (unsigned)(y - era * 400);
  const unsigned doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
  const unsigned doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  return era * 146097 + (i64)doe - 719468;
}
void civil_from_days(i64 z, int* y, int* m, int* d) {
  z += 719468;
  const i64 era = (z >= 0 ? z : z - 146096) / 146097;
  const unsigned doe = (unsigned)(z - era * 146097);
  const unsigned yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  const i64 yy = (i64)yoe + era * 400;
  const unsigned doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  const unsigned mp = (5 * doy + 2) / 153;
  *d = (int)(doy - (153 * mp + 2) / 5 + 1);
  *m = (int)(mp < 10 ? mp + 3 : mp - 9);
  *y = (int)(yy + (*m <= 2));
}

const i64 DAY0 = 8035;            // 1992-01-01 as days since epoch
const int N_ORDER_DAYS = 2406;    // 1992-01-01 .. 1998-08-02
const int N_DAYS = N_ORDER_DAYS + 121 + 30;  // ship/receipt span
const i64 CUTOFF = 9298;          // 1995-06-17 as days since epoch

const char* NATIONS[25] = {"ALGERIA", "ARGENTINA", "BRAZIL", "CANADA", "EGYPT", "ETHIOPIA", "FRANCE",
                           "GERMANY", "INDIA", "INDONESIA", "IRAN", "IRAQ", "JAPAN", "JORDAN", "KENYA",
                           "MOROCCO", "MOZAMBIQUE", "PERU", "CHINA", "ROMANIA", "SAUDI ARABIA", "VIETNAM",
                           "RUSSIA", "UNITED KINGDOM", "UNITED STATES"};
const int NATION_REGION[25] = {0, 1, 1, 1, 4, 0, 3, 3, 2, 2, 4, 4, 2, 4, 0, 0, 0, 1, 2, 3, 4, 2, 3, 3, 1};
const char* REGIONS[5] = {"AFRICA", "AMERICA", "ASIA", "EUROPE", "MIDDLE EAST"};
const char* SEGMENTS[5] = {"AUTOMOBILE", "BUILDING", "FURNITURE", "MACHINERY", "HOUSEHOLD"};
const char* PRIORITIES[5] = {"1-URGENT", "2-HIGH", "3-MEDIUM", "4-NOT SPECIFIED", "5-LOW"};
const char* SHIPMODES[7] = {"REG AIR", "AIR", "RAIL", "SHIP", "TRUCK", "MAIL", "FOB"};
const char* INSTRUCTS[4] = {"DELIVER IN PERSON", "COLLECT COD", "NONE", "TAKE BACK RETURN"};
const char* TYPE_S1[6] = {"STANDARD", "SMALL", "MEDIUM", "LARGE", "ECONOMY", "PROMO"};
const char* TYPE_S2[5] = {"ANODIZED", "BURNISHED", "PLATED", "POLISHED", "BRUSHED"};
const char* TYPE_S3[5] = {"TIN", "NICKEL", "BRASS", "STEEL", "COPPER"};
const char* CONT_S1[5] = {"SM", "LG", "MED", "JUMBO", "WRAP"};
const char* CONT_S2[8] = {"CASE", "BOX", "BAG", "JAR", "PKG", "PACK", "CAN", "DRUM"};
const char* FLAGS[3] = {"R", "A", "N"};
const char* STATUSES[2] = {"F", "O"};
const char* OSTATUS[3] = {"F", "O", "P"};

enum ColId {
  C_L_SHIPDATE_MS = 0,  // time, int64 ms
  C_O_ORDERKEY, C_O_CUSTKEY, C_L_PARTKEY, C_L_SUPPKEY, C_L_LINENUMBER, C_P_SIZE, C_O_SHIPPRIORITY,
  C_PS_PARTKEY, C_PS_SUPPKEY,                                   // int dims
  C_L_RETURNFLAG, C_L_LINESTATUS, C_L_SHIPMODE, C_L_SHIPINSTRUCT, C_O_ORDERSTATUS, C_O_ORDERPRIORITY,
  C_O_ORDERDATE, C_L_COMMITDATE, C_L_RECEIPTDATE, C_L_SHIPDATE, C_ORDER_YEAR, C_C_MKTSEGMENT,
  C_C_NATION, C_C_REGION, C_S_NATION, C_S_REGION, C_P_TYPE, C_P_BRAND, C_P_CONTAINER, C_P_MFGR,
  C_O_CLERK, C_C_NAME, C_S_NAME,                                // string dims (codes)
  C_L_QUANTITY, C_L_EXTENDEDPRICE, C_L_DISCOUNT, C_L_TAX, C_O_TOTALPRICE, C_PS_SUPPLYCOST, C_S_ACCTBAL,
  C_C_ACCTBAL, C_P_RETAILPRICE,                                 // f64 metrics
  C_PS_AVAILQTY, C_COUNT,                                       // i64 metrics
  C_NCOLS
};

// kind: 0 = time (int64 ms), 1 = int dimension (int64 values), 2 = string dimension (int32 codes +
// dictionary), 3 = f64 metric, 4 = i64 metric
struct ColInfo { const char* name; int kind; };
const ColInfo COLS[C_NCOLS] = {
    {"l_shipdate_ms", 0},
    {"o_orderkey", 1}, {"o_custkey", 1}, {"l_partkey", 1}, {"l_suppkey", 1}, {"l_linenumber", 1},
    {"p_size", 1}, {"o_shippriority", 1}, {"ps_partkey", 1}, {"ps_suppkey", 1},
    {"l_returnflag", 2}, {"l_linestatus", 2}, {"l_shipmode", 2}, {"l_shipinstruct", 2},
    {"o_orderstatus", 2}, {"o_orderpriority", 2}, {"o_orderdate", 2}, {"l_commitdate", 2},
    {"l_receiptdate", 2}, {"l_shipdate", 2}, {"order_year", 2}, {"c_mktsegment", 2}, {"c_nation", 2},
    {"c_region", 2}, {"s_nation", 2}, {"s_region", 2}, {"p_type", 2}, {"p_brand", 2},
    {"p_container", 2}, {"p_mfgr", 2}, {"o_clerk", 2}, {"c_name", 2}, {"s_name", 2},
    {"l_quantity", 3}, {"l_extendedprice", 3}, {"l_discount", 3}, {"l_tax", 3}, {"o_totalprice", 3},
    {"ps_supplycost", 3}, {"s_acctbal", 3}, {"c_acctbal", 3}, {"p_retailprice", 3},
    {"ps_availqty", 4}, {"count", 4}};

struct Scale {
  i64 n_orders, n_cust, n_supp, n_part, n_clerk;
};
Scale scale_of(double sf) {
  Scale s;
  s.n_orders = std::max<i64>(8, (i64)llround(1500000.0 * sf));
  s.n_cust = std::max<i64>(30, (i64)llround(150000.0 * sf));
  s.n_supp = std::max<i64>(20, (i64)llround(10000.0 * sf));
  s.n_part = std::max<i64>(40, (i64)llround(200000.0 * sf));
  s.n_clerk = std::max<i64>(5, (i64)llround(1000.0 * sf));
  return s;
}

struct Line {
  i64 partkey, suppkey;
  int qty, disc, tax;         // disc/tax in cents of a unit (0..10, 0..8)
  i64 ext_cents;              // extendedprice in cents
  int ship, commit, receipt;  // days since epoch
  int flag, status, mode, instruct;
};
struct Order {
  i64 key, custkey;
  int odate, prio, clerk, nlines, ostatus;
  i64 total_cents_x10000;  // sum ext*(1+tax)*(1-disc) scaled
  Line l[7];
};

inline i64 retail_cents(i64 partkey) { return 90000 + ((partkey / 10) % 20001) + 100 * (partkey % 1000); }

void gen_order(u64 seed, const Scale& sc, i64 i, Order* o) {
  o->key = ((i >> 3) << 5) + (i & 7) + 1;
  // customers with custkey % 3 == 0 never order (TPC-H 4.2.3)
  i64 c = uni(rnd(seed, 1, (u64)i), 1, sc.n_cust - sc.n_cust / 3);
  o->custkey = c + (c - 1) / 2;  // skips multiples of 3
  o->odate = (int)(DAY0 + uni(rnd(seed, 2, (u64)i), 0, N_ORDER_DAYS - 1));
  o->prio = (int)uni(rnd(seed, 3, (u64)i), 0, 4);
  o->clerk = (int)uni(rnd(seed, 4, (u64)i), 1, sc.n_clerk);
  o->nlines = (int)uni(rnd(seed, 5, (u64)i), 1, 7);
  int nf = 0;
  i64 total = 0;
  for (int j = 0; j < o->nlines; j++) {
    Line& L = o->l[j];
    u64 li = (u64)i * 8 + (u64)j;
    L.partkey = uni(rnd(seed, 10, li), 1, sc.n_part);
    i64 k = uni(rnd(seed, 11, li), 0, 3);
    i64 S = sc.n_supp;
    L.suppkey = (L.partkey + k * (S / 4 + (L.partkey - 1) / S)) % S + 1;
    L.qty = (int)uni(rnd(seed, 12, li), 1, 50);
    L.disc = (int)uni(rnd(seed, 13, li), 0, 10);
    L.tax = (int)uni(rnd(seed, 14, li), 0, 8);
    L.ext_cents = (i64)L.qty * retail_cents(L.partkey);
    L.ship = o->odate + (int)uni(rnd(seed, 15, li), 1, 121);
    L.commit = o->odate + (int)uni(rnd(seed, 16, li), 30, 90);
    L.receipt = L.ship + (int)uni(rnd(seed, 17, li), 1, 30);
    L.flag = (L.receipt <= CUTOFF) ? (int)uni(rnd(seed, 18, li), 0, 1) : 2;
    L.status = (L.ship > CUTOFF) ? 1 : 0;
    L.mode = (int)uni(rnd(seed, 19, li), 0, 6);
    L.instruct = (int)uni(rnd(seed, 20, li), 0, 3);
    nf += (L.status == 0);
    total += L.ext_cents * (100 + L.tax) * (100 - L.disc);
  }
  o->ostatus = nf == o->nlines ? 0 : (nf == 0 ? 1 : 2);
  o->total_cents_x10000 = total;
}

inline int cust_nation(u64 seed, i64 custkey) { return (int)uni(rnd(seed, 30, (u64)custkey), 0, 24); }
inline int cust_segment(u64 seed, i64 custkey) { return (int)uni(rnd(seed, 31, (u64)custkey), 0, 4); }
inline i64 cust_acctbal_cents(u64 seed, i64 custkey) { return uni(rnd(seed, 32, (u64)custkey), -99999, 999999); }
inline int supp_nation(u64 seed, i64 suppkey) { return (int)uni(rnd(seed, 40, (u64)suppkey), 0, 24); }
inline i64 supp_acctbal_cents(u64 seed, i64 suppkey) { return uni(rnd(seed, 41, (u64)suppkey), -99999, 999999); }
inline int part_type(u64 seed, i64 partkey) { return (int)uni(rnd(seed, 50, (u64)partkey), 0, 149); }
inline int part_brand(u64 seed, i64 partkey) { return (int)uni(rnd(seed, 51, (u64)partkey), 0, 24); }
inline int part_container(u64 seed, i64 partkey) { return (int)uni(rnd(seed, 52, (u64)partkey), 0, 39); }
inline int part_size(u64 seed, i64 partkey) { return (int)uni(rnd(seed, 53, (u64)partkey), 1, 50); }
inline i64 ps_availqty(u64 seed, i64 partkey, i64 suppkey) {
  return uni(rnd(seed, 60, (u64)partkey * 1000003ull + (u64)suppkey), 1, 9999);
}
inline i64 ps_supplycost_cents(u64 seed, i64 partkey, i64 suppkey) {
  return uni(rnd(seed, 61, (u64)partkey * 1000003ull + (u64)suppkey), 100, 100000);
}

const i64 CHUNK = 16384;  // orders per work unit; fixed so output is thread-count independent

struct Job {
  double sf;
  u64 seed;
  i64 o_begin, o_end;
  Scale sc;
};

Job make_job(double sf, u64 seed, int shard, int n_shards) {
  Job j;
  j.sf = sf;
  j.seed = seed;
  j.sc = scale_of(sf);
  if (n_shards < 1) n_shards = 1;
  if (shard < 0) shard = 0;
  if (shard >= n_shards) shard = n_shards - 1;
  i64 per = (j.sc.n_orders + n_shards - 1) / n_shards;
  j.o_begin = std::min<i64>(j.sc.n_orders, per * shard);
  j.o_end = std::min<i64>(j.sc.n_orders, j.o_begin + per);
  return j;
}

int n_threads() {
  const char* e = getenv("SDO_GEN_THREADS");
  int n = e ? atoi(e) : (int)std::thread::hardware_concurrency();
  return std::max(1, std::min(n, 64));
}

template <class F>
void parallel_chunks(i64 n_chunks, F f) {
  int nt = (int)std::min<i64>(n_threads(), std::max<i64>(1, n_chunks));
  std::atomic<i64> next(0);
  std::vector<std::thread> th;
  for (int t = 0; t < nt; t++)
    th.emplace_back([&]() {
      for (;;) {
        i64 c = next.fetch_add(1);
        if (c >= n_chunks) break;
        f(c);
      }
    });
  for (auto& t : th) t.join();
}

// counts[chunk * N_DAYS + day]
void count_pass(const Job& job, std::vector<uint32_t>& counts, i64 n_chunks) {
  counts.assign((size_t)n_chunks * N_DAYS, 0);
  parallel_chunks(n_chunks, [&](i64 c) {
    uint32_t* cnt = &counts[(size_t)c * N_DAYS];
    i64 b = job.o_begin + c * CHUNK, e = std::min(job.o_end, b + CHUNK);
    Order o;
    for (i64 i = b; i < e; i++) {
      gen_order(job.seed, job.sc, i, &o);
      for (int j = 0; j < o.nlines; j++) cnt[o.l[j].ship - DAY0]++;
    }
  });
}

int find_col(const char* name) {
  for (int i = 0; i < C_NCOLS; i++)
    if (!strcmp(COLS[i].name, name)) return i;
  return -1;
}

std::string date_str(i64 days) {
  int y, m, d;
  civil_from_days(days, &y, &m, &d);
  char b[16];
  snprintf(b, sizeof b, "%04d-%02d-%02d", y, m, d);
  return b;
}

}  // namespace

extern "C" {

// Number of generator columns and their (name, kind) descriptors.
int sdo_tpch_num_columns() { return C_NCOLS; }
const char* sdo_tpch_column_name(int i) { return (i >= 0 && i < C_NCOLS) ? COLS[i].name : nullptr; }
int sdo_tpch_column_kind(const char* name) {
  int c = find_col(name);
  return c < 0 ? -1 : COLS[c].kind;
}

// Rows (line items) of one shard.
int64_t sdo_tpch_count_rows(double sf, uint64_t seed, int shard, int n_shards) {
  Job job = make_job(sf, seed, shard, n_shards);
  i64 n_chunks = (job.o_end - job.o_begin + CHUNK - 1) / CHUNK;
  std::vector<i64> per(std::max<i64>(1, n_chunks), 0);
  parallel_chunks(n_chunks, [&](i64 c) {
    i64 b = job.o_begin + c * CHUNK, e = std::min(job.o_end, b + CHUNK), n = 0;
    for (i64 i = b; i < e; i++) n += uni(rnd(job.seed, 5, (u64)i), 1, 7);
    per[c] = n;
  });
  i64 tot = 0;
  for (i64 v : per) tot += v;
  return tot;
}

// Fill caller-allocated column arrays (each n_rows long; element type by kind:
// time/int-dim/i64-metric -> int64, string-dim -> int32 codes, f64 metric -> double).
// Rows come out sorted by ship day, ties in (order index, line number) order.
// Returns rows written, or -1 on an unknown column.
int64_t sdo_tpch_fill(double sf, uint64_t seed, int shard, int n_shards, int n_cols,
                      const char* const* col_names, void* const* out) {
  Job job = make_job(sf, seed, shard, n_shards);
  std::vector<int> cols(n_cols);
  for (int k = 0; k < n_cols; k++) {
    cols[k] = find_col(col_names[k]);
    if (cols[k] < 0) return -1;
  }
  i64 n_chunks = (job.o_end - job.o_begin + CHUNK - 1) / CHUNK;
  std::vector<uint32_t> counts;
  count_pass(job, counts, n_chunks);
  // exclusive prefix in (day-major, chunk-minor) order -> start offset of each (chunk, day)
  std::vector<i64> start((size_t)std::max<i64>(1, n_chunks) * N_DAYS, 0);
  i64 run = 0;
  for (int d = 0; d < N_DAYS; d++)
    for (i64 c = 0; c < n_chunks; c++) {
      start[(size_t)c * N_DAYS + d] = run;
      run += counts[(size_t)c * N_DAYS + d];
    }
  const u64 seed_ = job.seed;
  parallel_chunks(n_chunks, [&](i64 c) {
    i64* pos = &start[(size_t)c * N_DAYS];
    i64 b = job.o_begin + c * CHUNK, e = std::min(job.o_end, b + CHUNK);
    Order o;
    for (i64 i = b; i < e; i++) {
      gen_order(seed_, job.sc, i, &o);
      int cn = cust_nation(seed_, o.custkey);
      for (int j = 0; j < o.nlines; j++) {
        const Line& L = o.l[j];
        i64 r = pos[L.ship - DAY0]++;
        int sn = supp_nation(seed_, L.suppkey);
        for (int k = 0; k < n_cols; k++) {
          void* p = out[k];
          switch (cols[k]) {
            case C_L_SHIPDATE_MS: ((i64*)p)[r] = (i64)L.ship * 86400000ll; break;
            case C_O_ORDERKEY: ((i64*)p)[r] = o.key; break;
            case C_O_CUSTKEY: ((i64*)p)[r] = o.custkey; break;
            case C_L_PARTKEY: case C_PS_PARTKEY: ((i64*)p)[r] = L.partkey; break;
            case C_L_SUPPKEY: case C_PS_SUPPKEY: ((i64*)p)[r] = L.suppkey; break;
            case C_L_LINENUMBER: ((i64*)p)[r] = j + 1; break;
            case C_P_SIZE: ((i64*)p)[r] = part_size(seed_, L.partkey); break;
            case C_O_SHIPPRIORITY: ((i64*)p)[r] = 0; break;
            case C_L_RETURNFLAG: ((int32_t*)p)[r] = L.flag; break;
            case C_L_LINESTATUS: ((int32_t*)p)[r] = L.status; break;
            case C_L_SHIPMODE: ((int32_t*)p)[r] = L.mode; break;
            case C_L_SHIPINSTRUCT: ((int32_t*)p)[r] = L.instruct; break;
            case C_O_ORDERSTATUS: ((int32_t*)p)[r] = o.ostatus; break;
            case C_O_ORDERPRIORITY: ((int32_t*)p)[r] = o.prio; break;
            case C_O_ORDERDATE: ((int32_t*)p)[r] = (int32_t)(o.odate - DAY0); break;
            case C_L_COMMITDATE: ((int32_t*)p)[r] = (int32_t)(L.commit - DAY0); break;
            case C_L_RECEIPTDATE: ((int32_t*)p)[r] = (int32_t)(L.receipt - DAY0); break;
            case C_L_SHIPDATE: ((int32_t*)p)[r] = (int32_t)(L.ship - DAY0); break;
            case C_ORDER_YEAR: {
              int y, m, d;
              civil_from_days(o.odate, &y, &m, &d);
              ((int32_t*)p)[r] = y - 1992;
              break;
            }
            case C_C_MKTSEGMENT: ((int32_t*)p)[r] = cust_segment(seed_, o.custkey); break;
            case C_C_NATION: ((int32_t*)p)[r] = cn; break;
            case C_C_REGION: ((int32_t*)p)[r] = NATION_REGION[cn]; break;
            case C_S_NATION: ((int32_t*)p)[r] = sn; break;
            case C_S_REGION: ((int32_t*)p)[r] = NATION_REGION[sn]; break;
            case C_P_TYPE: ((int32_t*)p)[r] = part_type(seed_, L.partkey); break;
            case C_P_BRAND: ((int32_t*)p)[r] = part_brand(seed_, L.partkey); break;
            case C_P_CONTAINER: ((int32_t*)p)[r] = part_container(seed_, L.partkey); break;
            case C_P_MFGR: ((int32_t*)p)[r] = part_brand(seed_, L.partkey) / 5; break;
            case C_O_CLERK: ((int32_t*)p)[r] = o.clerk - 1; break;
            case C_C_NAME: ((int32_t*)p)[r] = (int32_t)(o.custkey - 1); break;
            case C_S_NAME: ((int32_t*)p)[r] = (int32_t)(L.suppkey - 1); break;
            case C_L_QUANTITY: ((double*)p)[r] = (double)L.qty; break;
            case C_L_EXTENDEDPRICE: ((double*)p)[r] = (double)L.ext_cents / 100.0; break;
            case C_L_DISCOUNT: ((double*)p)[r] = (double)L.disc / 100.0; break;
            case C_L_TAX: ((double*)p)[r] = (double)L.tax / 100.0; break;
            case C_O_TOTALPRICE: ((double*)p)[r] = (double)((o.total_cents_x10000 + 5000) / 10000) / 100.0; break;
            case C_PS_SUPPLYCOST: ((double*)p)[r] = (double)ps_supplycost_cents(seed_, L.partkey, L.suppkey) / 100.0; break;
            case C_S_ACCTBAL: ((double*)p)[r] = (double)supp_acctbal_cents(seed_, L.suppkey) / 100.0; break;
            case C_C_ACCTBAL: ((double*)p)[r] = (double)cust_acctbal_cents(seed_, o.custkey) / 100.0; break;
            case C_P_RETAILPRICE: ((double*)p)[r] = (double)retail_cents(L.partkey) / 100.0; break;
            case C_PS_AVAILQTY: ((i64*)p)[r] = ps_availqty(seed_, L.partkey, L.suppkey); break;
            case C_COUNT: ((i64*)p)[r] = 1; break;
            default: break;
          }
        }
      }
    }
  });
  return run;
}

// Dictionary (code -> UTF-8 string) of a string dimension at scale `sf`, as a
// concatenated byte buffer plus n+1 offsets.  Call with bytes == NULL to size.
// Returns the number of entries, or -1 for a non-string column.
int64_t sdo_tpch_dictionary(const char* name, double sf, int64_t* offsets, char* bytes, int64_t* n_bytes) {
  int c = find_col(name);
  if (c < 0 || COLS[c].kind != 2) return -1;
  Scale sc = scale_of(sf);
  std::vector<std::string> v;
  char b[64];
  switch (c) {
    case C_L_RETURNFLAG: for (auto s : FLAGS) v.push_back(s); break;
    case C_L_LINESTATUS: for (auto s : STATUSES) v.push_back(s); break;
    case C_L_SHIPMODE: for (auto s : SHIPMODES) v.push_back(s); break;
    case C_L_SHIPINSTRUCT: for (auto s : INSTRUCTS) v.push_back(s); break;
    case C_O_ORDERSTATUS: for (auto s : OSTATUS) v.push_back(s); break;
    case C_O_ORDERPRIORITY: for (auto s : PRIORITIES) v.push_back(s); break;
    case C_O_ORDERDATE: case C_L_COMMITDATE: case C_L_RECEIPTDATE: case C_L_SHIPDATE:
      for (int d = 0; d < N_DAYS; d++) v.push_back(date_str(DAY0 + d));
      break;
    case C_ORDER_YEAR: for (int y = 1992; y <= 1998; y++) v.push_back(std::to_string(y)); break;
    case C_C_MKTSEGMENT: for (auto s : SEGMENTS) v.push_back(s); break;
    case C_C_NATION: case C_S_NATION: for (auto s : NATIONS) v.push_back(s); break;
    case C_C_REGION: case C_S_REGION: for (auto s : REGIONS) v.push_back(s); break;
    case C_P_TYPE:
      for (int t = 0; t < 150; t++)
        v.push_back(std::string(TYPE_S1[t / 25]) + " " + TYPE_S2[(t / 5) % 5] + " " + TYPE_S3[t % 5]);
      break;
    case C_P_BRAND:
      for (int t = 0; t < 25; t++) { snprintf(b, sizeof b, "Brand#%d%d", t / 5 + 1, t % 5 + 1); v.push_back(b); }
      break;
    case C_P_CONTAINER:
      for (int t = 0; t < 40; t++) v.push_back(std::string(CONT_S1[t / 8]) + " " + CONT_S2[t % 8]);
      break;
    case C_P_MFGR:
      for (int t = 0; t < 5; t++) { snprintf(b, sizeof b, "Manufacturer#%d", t + 1); v.push_back(b); }
      break;
    case C_O_CLERK:
      for (i64 t = 0; t < sc.n_clerk; t++) { snprintf(b, sizeof b, "Clerk#%09lld", (long long)t + 1); v.push_back(b); }
      break;
    case C_C_NAME:
      for (i64 t = 0; t < sc.n_cust; t++) { snprintf(b, sizeof b, "Customer#%09lld", (long long)t + 1); v.push_back(b); }
      break;
    case C_S_NAME:
      for (i64 t = 0; t < sc.n_supp; t++) { snprintf(b, sizeof b, "Supplier#%09lld", (long long)t + 1); v.push_back(b); }
      break;
    default: return -1;
  }
  i64 tot = 0;
  for (auto& s : v) tot += (i64)s.size();
  if (n_bytes) *n_bytes = tot;
  if (offsets && bytes) {
    i64 off = 0;
    for (size_t i = 0; i < v.size(); i++) {
      offsets[i] = off;
      memcpy(bytes + off, v[i].data(), v[i].size());
      off += (i64)v[i].size();
    }
    offsets[v.size()] = off;
  }
  return (i64)v.size();
}

}  // extern "C"
