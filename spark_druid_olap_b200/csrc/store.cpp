#include "store.hpp"

#include <algorithm>
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <memory>
#include <numeric>
#include <unordered_map>

#include "kernels.hpp"

namespace sdo {

int utf8_cmp(const char* a, size_t na, const char* b, size_t nb) {
  int c = memcmp(a, b, std::min(na, nb));  // memcmp compares as unsigned char
  if (c) return c;
  return na < nb ? -1 : (na > nb ? 1 : 0);
}
int64_t Dictionary::lower_bound(const std::string& s) const {
  int64_t lo = 0, hi = size();
  while (lo < hi) {
    int64_t mid = (lo + hi) / 2;
    if (utf8_cmp(data(mid), len(mid), s.data(), s.size()) < 0) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}
int64_t Dictionary::upper_bound(const std::string& s) const {
  int64_t lo = 0, hi = size();
  while (lo < hi) {
    int64_t mid = (lo + hi) / 2;
    if (utf8_cmp(data(mid), len(mid), s.data(), s.size()) <= 0) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}
int64_t Dictionary::find(const std::string& s) const {
  int64_t i = lower_bound(s);
  if (i < size() && len(i) == s.size() && !memcmp(data(i), s.data(), s.size())) return i;
  return -1;
}

int DataSource::dim_index(const std::string& n) const {
  for (size_t i = 0; i < dims.size(); i++)
    if (dims[i].name == n) return (int)i;
  return -1;
}
int DataSource::metric_index(const std::string& n) const {
  for (size_t i = 0; i < metrics.size(); i++)
    if (metrics[i].name == n) return (int)i;
  return -1;
}

static std::string upper(std::string s) {
  for (auto& c : s) c = (char)toupper((unsigned char)c);
  return s;
}

int64_t truncate_time(const std::string& gran, int64_t t) {
  if (gran == "NONE") return t;
  if (gran == "ALL") return t;
  Granularity g = simple_granularity(gran);
  return bucket_start(g, t);
}
int64_t segment_start(const std::string& gran, int64_t t) {
  if (gran == "ALL") return INT64_MIN / 2;
  Granularity g = simple_granularity(gran);
  return bucket_start(g, t);
}
int64_t segment_end(const std::string& gran, int64_t start) {
  if (gran == "ALL") return INT64_MAX / 2;
  Granularity g = simple_granularity(gran);
  if (g.months) {
    Civil c = civil_from_ms(start, 0);
    int64_t midx = c.year * 12 + (c.month - 1) + g.months;
    return days_from_civil(floordiv(midx, 12), (int)floormod(midx, 12) + 1, 1) * MS_DAY;
  }
  return start + g.period_ms;
}

Store::~Store() {
  for (auto& kv : sources_) {
    for (auto& s : kv.second->segments) free_segment(*s);
    for (auto& d : kv.second->dims)
      if (d.hll_lut) cudaFree(d.hll_lut);
  }
}

void Store::free_segment(Segment& s) {
  if (device_ < 0) return;
  for (void* p : s.dim_data)
    if (p) cudaFree(p);
  for (void* p : s.metric_data)
    if (p) cudaFree(p);
  for (void* p : s.dim_hll)
    if (p) cudaFree(p);
  s.dim_data.clear();
  s.metric_data.clear();
  s.dim_hll.clear();
}

void Store::create(const std::string& name, const std::string& options_json) {
  std::lock_guard<std::mutex> lk(mu_);
  if (sources_.count(name)) throw Invalid("datasource '" + name + "' already exists");
  auto ds = std::make_shared<DataSource>();
  ds->name = name;
  if (!options_json.empty()) {
    Json o;
    try {
      o = json_parse(options_json.data(), options_json.size());
    } catch (const JsonError& e) {
      throw Invalid(e.what());
    }
    ds->segment_granularity = upper(o.get_str("segmentGranularity", "MONTH"));
    ds->query_granularity = upper(o.get_str("queryGranularity", "DAY"));
  }
  // validate names early
  segment_start(ds->segment_granularity, 0);
  truncate_time(ds->query_granularity, 0);
  sources_[name] = ds;
}

void Store::drop(const std::string& name) {
  std::shared_ptr<DataSource> ds;
  {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = sources_.find(name);
    if (it == sources_.end()) throw NotFound("unknown datasource '" + name + "'");
    ds = it->second;
    sources_.erase(it);
  }
  std::unique_lock<std::shared_mutex> lk(ds->mu);  // waits for running queries
  if (device_ >= 0) SDO_CUDA(cudaSetDevice(device_));
  for (auto& s : ds->segments) free_segment(*s);
  ds->segments.clear();
  for (auto& d : ds->dims)
    if (d.hll_lut) { cudaFree(d.hll_lut); d.hll_lut = nullptr; }
}

std::shared_ptr<DataSource> Store::get(const std::string& name) const {
  std::lock_guard<std::mutex> lk(mu_);
  auto it = sources_.find(name);
  if (it == sources_.end()) throw NotFound("unknown datasource '" + name + "'");
  return it->second;
}

static int width_for(int64_t card) { return card <= 256 ? 1 : (card <= 65536 ? 2 : 4); }

namespace {
struct DevBuf {
  void* p = nullptr;
  explicit DevBuf(size_t n) { SDO_CUDA(cudaMalloc(&p, std::max<size_t>(n, 256))); }
  ~DevBuf() { if (p) cudaFree(p); }
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
};
// segment column arrays are padded so that 16-byte vector loads past n_rows stay in bounds
size_t padded_bytes(int64_t n_rows, int width) { return round_up((size_t)n_rows * width + 64, 256); }
}  // namespace


// Re-encode dimension `di` of every resident segment: ids through `lut` (nullptr = unchanged) into `new_width` bytes.
void Store::reencode_segments(DataSource& ds, int di, const std::vector<uint32_t>* lut, int new_width) {
  if (device_ < 0) {
    for (auto& s : ds.segments)
      if ((int)s->dim_width.size() > di) s->dim_width[di] = new_width;
    return;
  }
  std::unique_ptr<DevBuf> dl;
  if (lut) {
    dl.reset(new DevBuf(lut->size() * 4));
    SDO_CUDA(cudaMemcpy(dl->p, lut->data(), lut->size() * 4, cudaMemcpyHostToDevice));
  }
  for (auto& s : ds.segments) {
    if ((int)s->dim_data.size() <= di || !s->dim_data[di]) continue;
    int ow = s->dim_width[di];
    void* nb = nullptr;
    SDO_CUDA(cudaMalloc(&nb, padded_bytes(s->n_rows, new_width)));
    SDO_CUDA(cudaMemset(nb, 0, padded_bytes(s->n_rows, new_width)));
    launch_reencode(s->dim_data[di], ow, nb, new_width, s->n_rows, dl ? (const uint32_t*)dl->p : nullptr, 0);
    SDO_CUDA(cudaDeviceSynchronize());
    s->device_bytes += padded_bytes(s->n_rows, new_width);
    s->device_bytes -= padded_bytes(s->n_rows, ow);
    cudaFree(s->dim_data[di]);
    s->dim_data[di] = nb;
    s->dim_width[di] = new_width;
  }
}

// Merge the values of `bd` into the sorted, de-duplicated datasource dictionary of dimension `di`; resident
// segments are re-encoded when ids move or widen.  code_to_id (optional) maps bd's entries to the new ids.
void Store::merge_dictionary(DataSource& ds, int di, const Dictionary& bd, std::vector<uint32_t>* code_to_id) {
  DimMeta& dm = ds.dims[di];
  const Dictionary& od = *dm.dict;
  int64_t nb = bd.size(), no = od.size();
  std::vector<int64_t> border((size_t)nb);
  std::iota(border.begin(), border.end(), 0);
  std::sort(border.begin(), border.end(), [&](int64_t a, int64_t b) { return utf8_cmp(bd.data(a), bd.len(a), bd.data(b), bd.len(b)) < 0; });
  Dictionary merged;
  std::vector<uint32_t> old_to_new((size_t)no);
  std::vector<uint32_t> c2i((size_t)nb, 0);
  int64_t i = 0, j = 0;
  while (i < no || j < nb) {
    int cmp;
    if (i >= no) cmp = 1;
    else if (j >= nb) cmp = -1;
    else cmp = utf8_cmp(od.data(i), od.len(i), bd.data(border[j]), bd.len(border[j]));
    uint32_t id = (uint32_t)merged.size();
    if (cmp < 0) { merged.push(od.data(i), od.len(i)); old_to_new[i++] = id; }
    else if (cmp > 0) {
      const char* p = bd.data(border[j]);
      size_t n = bd.len(border[j]);
      merged.push(p, n);
      c2i[border[j++]] = id;
      // duplicates inside the batch dictionary map to the same id
      while (j < nb && utf8_cmp(bd.data(border[j]), bd.len(border[j]), p, n) == 0) c2i[border[j++]] = id;
    } else {
      merged.push(od.data(i), od.len(i));
      old_to_new[i++] = id;
      const char* p = merged.data(id);
      size_t n = merged.len(id);
      while (j < nb && utf8_cmp(bd.data(border[j]), bd.len(border[j]), p, n) == 0) c2i[border[j++]] = id;
    }
  }
  bool changed = merged.size() != no;
  int w = std::max(dm.width, width_for(merged.size()));
  if (changed && no > 0) {
    bool identity = true;
    for (int64_t k = 0; k < no; k++)
      if (old_to_new[k] != (uint32_t)k) { identity = false; break; }
    if (!identity || w != dm.width) reencode_segments(ds, di, identity ? nullptr : &old_to_new, w);
    if (dm.hll_lut) { cudaFree(dm.hll_lut); dm.hll_lut = nullptr; dm.hll_lut_card = 0; }
  }
  dm.dict = std::make_shared<Dictionary>(std::move(merged));
  dm.cardinality = dm.dict->size();
  dm.width = w;
  if (code_to_id) code_to_id->swap(c2i);
}

// Make a dimension's id space at least as large as asked: merge `values` into a string dictionary, or
// raise an integer dimension's cardinality.  Used to align shards that live in different engines.
void Store::extend_dimension(const std::string& name, const std::string& dim, const Dictionary* values, int64_t min_cardinality) {
  auto ds = get(name);
  std::unique_lock<std::shared_mutex> lk(ds->mu);
  ds->version++;
  if (device_ >= 0) SDO_CUDA(cudaSetDevice(device_));
  int di = ds->dim_index(dim);
  if (di < 0) throw NotFound("unknown dimension '" + dim + "' in datasource '" + name + "'");
  DimMeta& dm = ds->dims[di];
  if (dm.enc == DimEncoding::Dict) {
    if (values && values->size() > 0) merge_dictionary(*ds, di, *values, nullptr);
  } else {
    if (min_cardinality > 4294967295ll) throw Unsupported("integer dimension cardinality beyond 2^32 - 1");
    if (min_cardinality > dm.cardinality) {
      dm.cardinality = min_cardinality;
      int w = width_for(dm.cardinality);
      if (w > dm.width) { reencode_segments(*ds, di, nullptr, w); dm.width = w; }
    }
  }
}

std::shared_ptr<Dictionary> Store::dictionary(const std::string& name, const std::string& dim) const {
  auto ds = get(name);
  std::shared_lock<std::shared_mutex> lk(ds->mu);
  int di = ds->dim_index(dim);
  if (di < 0) throw NotFound("unknown dimension '" + dim + "' in datasource '" + name + "'");
  return ds->dims[di].dict;
}

void Store::load(const std::string& name, int64_t n_rows, const std::vector<LoadColumn>& cols, int partition) {
  auto ds = get(name);
  std::unique_lock<std::shared_mutex> lk(ds->mu);
  if (n_rows < 0 || n_rows > 2147483647ll) throw Invalid("sdo_segment_load: n_rows must be in [0, 2^31)");
  ds->version++;
  if (n_rows == 0) return;
  const bool dev = device_ >= 0;
  if (dev) SDO_CUDA(cudaSetDevice(device_));

  // ---- time: truncate, order, cut -----------------------------------------------------------
  const LoadColumn* tc = nullptr;
  for (auto& c : cols)
    if (c.kind == 0) {
      if (tc) throw Invalid("more than one time column");
      tc = &c;
    }
  if (!tc || tc->type != 0 || !tc->data) throw Invalid("sdo_segment_load needs one SDO_COL_TIME column of type SDO_T_INT64");
  const int64_t* traw = (const int64_t*)tc->data;
  std::vector<int64_t> tt((size_t)n_rows);
  bool sorted = true;
  {
    const std::string& qg = ds->query_granularity;
    if (qg == "NONE" || qg == "ALL") {
      memcpy(tt.data(), traw, (size_t)n_rows * 8);
    } else {
      Granularity g = simple_granularity(qg);
      // fast path for fixed periods aligned to the epoch (DAY, HOUR, ...)
      if (!g.months && g.period_ms > 0 && g.period_ms % (7 * MS_DAY) != 0) {
        const int64_t p = g.period_ms;
        for (int64_t i = 0; i < n_rows; i++) tt[i] = traw[i] - floormod(traw[i], p);
      } else {
        for (int64_t i = 0; i < n_rows; i++) tt[i] = bucket_start(g, traw[i]);
      }
    }
    for (int64_t i = 1; i < n_rows; i++)
      if (tt[i] < tt[i - 1]) { sorted = false; break; }
  }
  std::vector<int32_t> perm;
  if (!sorted) {
    perm.resize((size_t)n_rows);
    std::iota(perm.begin(), perm.end(), 0);
    std::stable_sort(perm.begin(), perm.end(), [&](int32_t a, int32_t b) { return tt[a] < tt[b]; });
    std::vector<int64_t> t2((size_t)n_rows);
    for (int64_t i = 0; i < n_rows; i++) t2[i] = tt[perm[i]];
    tt.swap(t2);
  }
  struct Cut { int64_t r0, r1, start_ms; };
  std::vector<Cut> cuts;
  {
    int64_t r0 = 0;
    int64_t cur = segment_start(ds->segment_granularity, tt[0]);
    int64_t cur_end = segment_end(ds->segment_granularity, cur);
    for (int64_t i = 1; i < n_rows; i++) {
      if (tt[i] >= cur_end) {
        cuts.push_back({r0, i, cur});
        r0 = i;
        cur = segment_start(ds->segment_granularity, tt[i]);
        cur_end = segment_end(ds->segment_granularity, cur);
      }
    }
    cuts.push_back({r0, n_rows, cur});
  }

  // ---- dimensions: dictionaries, widths -------------------------------------------------------
  struct DimLoad {
    int di;
    const LoadColumn* col;
    std::vector<int32_t> owned_codes;   // STRING input converted to codes
    std::vector<uint32_t> code_to_id;   // Dict
  };
  std::vector<DimLoad> dloads;
  for (auto& c : cols) {
    if (c.kind != 1) continue;
    if (!c.data) throw Invalid("dimension '" + c.name + "' has no data");
    int di = ds->dim_index(c.name);
    bool is_int = (c.type == 0 || c.type == 4);
    if (c.type != 0 && c.type != 2 && c.type != 3 && c.type != 4) throw Invalid("dimension '" + c.name + "': bad column type");
    if (ds->metric_index(c.name) >= 0) throw Invalid("'" + c.name + "' is already a metric");
    if (di < 0) {
      DimMeta dm;
      dm.name = c.name;
      dm.enc = is_int ? DimEncoding::Direct : DimEncoding::Dict;
      dm.src_type = c.src_type;
      if (is_int && dm.src_type == 0) dm.src_type = (c.type == 4) ? 1 : 2;
      ds->dims.push_back(std::move(dm));
      di = (int)ds->dims.size() - 1;
    }
    DimMeta& dm = ds->dims[di];
    if ((dm.enc == DimEncoding::Direct) != is_int) throw Invalid("dimension '" + c.name + "' changes between integer and string encoding");
    DimLoad dl;
    dl.di = di;
    dl.col = &c;
    if (is_int) {
      int64_t mn = INT64_MAX, mx = INT64_MIN;
      if (c.type == 0) {
        const int64_t* v = (const int64_t*)c.data;
        for (int64_t i = 0; i < n_rows; i++) { mn = std::min(mn, v[i]); mx = std::max(mx, v[i]); }
      } else {
        const int32_t* v = (const int32_t*)c.data;
        for (int64_t i = 0; i < n_rows; i++) { mn = std::min<int64_t>(mn, v[i]); mx = std::max<int64_t>(mx, v[i]); }
      }
      if (mn < 0 || mx >= 4294967295ll) throw Unsupported("integer dimension '" + c.name + "' has values outside [0, 2^32-2]");
      dm.cardinality = std::max(dm.cardinality, mx + 1);
      int w = width_for(dm.cardinality);
      if (w > dm.width) { reencode_segments(*ds, di, nullptr, w); dm.width = w; }
    } else {
      // batch dictionary
      Dictionary bd;
      if (c.type == 3) {
        if (!c.dict_offsets || (!c.bytes && c.n_dict > 0 && c.dict_offsets[c.n_dict] > 0)) throw Invalid("dimension '" + c.name + "': CODES column needs dict_offsets and bytes");
        for (int64_t k = 0; k < c.n_dict; k++) bd.push(c.bytes + c.dict_offsets[k], (size_t)(c.dict_offsets[k + 1] - c.dict_offsets[k]));
        const int32_t* codes = (const int32_t*)c.data;
        for (int64_t i = 0; i < n_rows; i++)
          if (codes[i] < 0 || codes[i] >= c.n_dict) throw Invalid("dimension '" + c.name + "': code out of dictionary range");
      } else {
        const int64_t* off = (const int64_t*)c.data;
        std::unordered_map<std::string, int32_t> seen;
        dl.owned_codes.resize((size_t)n_rows);
        for (int64_t i = 0; i < n_rows; i++) {
          std::string s(c.bytes + off[i], (size_t)(off[i + 1] - off[i]));
          auto it = seen.find(s);
          if (it == seen.end()) {
            it = seen.emplace(s, (int32_t)bd.size()).first;
            bd.push(s.data(), s.size());
          }
          dl.owned_codes[i] = it->second;
        }
      }
      merge_dictionary(*ds, di, bd, &dl.code_to_id);
    }
    dloads.push_back(std::move(dl));
  }

  // ---- metrics ----------------------------------------------------------------------------------
  struct MetLoad { int mi; const LoadColumn* col; };
  std::vector<MetLoad> mloads;
  for (auto& c : cols) {
    if (c.kind != 2) continue;
    if (!c.data) throw Invalid("metric '" + c.name + "' has no data");
    if (c.type != 0 && c.type != 1) throw Invalid("metric '" + c.name + "' must be SDO_T_INT64 or SDO_T_FLOAT64");
    if (ds->dim_index(c.name) >= 0) throw Invalid("'" + c.name + "' is already a dimension");
    int mi = ds->metric_index(c.name);
    if (mi < 0) {
      MetricMeta mm;
      mm.name = c.name;
      mm.is_float = c.type == 1;
      ds->metrics.push_back(mm);
      mi = (int)ds->metrics.size() - 1;
    }
    MetricMeta& mm = ds->metrics[mi];
    if (mm.is_float != (c.type == 1)) throw Invalid("metric '" + c.name + "' changes type between loads");
    if (!mm.is_float) {
      const int64_t* v = (const int64_t*)c.data;
      int64_t first = mm.seen ? mm.const_value : v[0];
      bool cst = mm.constant;
      for (int64_t i = 0; cst && i < n_rows; i++) cst = v[i] == first;
      mm.constant = cst;
      mm.const_value = first;
    } else {
      mm.constant = false;
    }
    mm.seen = true;
    mloads.push_back({mi, &c});
  }

  // ---- build segments ---------------------------------------------------------------------------
  std::vector<std::unique_ptr<Segment>> segs;
  for (auto& ct : cuts) {
    std::unique_ptr<Segment> s(new Segment());
    s->start_ms = ct.start_ms;
    s->end_ms = segment_end(ds->segment_granularity, ct.start_ms);
    s->partition = partition;
    s->n_rows = ct.r1 - ct.r0;
    s->dim_data.assign(ds->dims.size(), nullptr);
    s->dim_hll.assign(ds->dims.size(), nullptr);
    s->dim_width.assign(ds->dims.size(), 0);
    s->metric_data.assign(ds->metrics.size(), nullptr);
    for (int64_t r = ct.r0; r < ct.r1; r++) {
      if (r == ct.r0 || tt[r] != tt[r - 1]) {
        s->time_vals.push_back(tt[r]);
        s->time_rows.push_back(r - ct.r0);
      }
    }
    s->time_rows.push_back(s->n_rows);
    segs.push_back(std::move(s));
  }

  if (dev) {
    std::unique_ptr<DevBuf> dperm;
    if (!sorted) {
      dperm.reset(new DevBuf((size_t)n_rows * 4));
      SDO_CUDA(cudaMemcpy(dperm->p, perm.data(), (size_t)n_rows * 4, cudaMemcpyHostToDevice));
    }
    for (auto& dl : dloads) {
      DimMeta& dm = ds->dims[dl.di];
      bool src64 = dl.col->type == 0;
      const void* src = dl.owned_codes.empty() ? dl.col->data : (const void*)dl.owned_codes.data();
      size_t esz = src64 ? 8 : 4;
      DevBuf raw((size_t)n_rows * esz);
      SDO_CUDA(cudaMemcpy(raw.p, src, (size_t)n_rows * esz, cudaMemcpyHostToDevice));
      std::unique_ptr<DevBuf> dlut;
      if (dm.enc == DimEncoding::Dict) {
        dlut.reset(new DevBuf(dl.code_to_id.size() * 4));
        SDO_CUDA(cudaMemcpy(dlut->p, dl.code_to_id.data(), dl.code_to_id.size() * 4, cudaMemcpyHostToDevice));
      }
      for (size_t k = 0; k < cuts.size(); k++) {
        Segment& s = *segs[k];
        size_t nb = padded_bytes(s.n_rows, dm.width);
        SDO_CUDA(cudaMalloc(&s.dim_data[dl.di], nb));
        SDO_CUDA(cudaMemset(s.dim_data[dl.di], 0, nb));
        s.dim_width[dl.di] = dm.width;
        s.device_bytes += nb;
        launch_encode_dim(raw.p, src64, dperm ? (const int32_t*)dperm->p : nullptr, cuts[k].r0, s.n_rows,
                          dlut ? (const uint32_t*)dlut->p : nullptr, s.dim_data[dl.di], dm.width, 0);
      }
      SDO_CUDA(cudaDeviceSynchronize());
    }
    for (auto& ml : mloads) {
      DevBuf raw((size_t)n_rows * 8);
      SDO_CUDA(cudaMemcpy(raw.p, ml.col->data, (size_t)n_rows * 8, cudaMemcpyHostToDevice));
      for (size_t k = 0; k < cuts.size(); k++) {
        Segment& s = *segs[k];
        size_t nb = padded_bytes(s.n_rows, 8);
        SDO_CUDA(cudaMalloc(&s.metric_data[ml.mi], nb));
        SDO_CUDA(cudaMemset(s.metric_data[ml.mi], 0, nb));
        s.device_bytes += nb;
        launch_gather64(raw.p, dperm ? (const int32_t*)dperm->p : nullptr, cuts[k].r0, s.n_rows, s.metric_data[ml.mi], 0);
      }
      SDO_CUDA(cudaDeviceSynchronize());
    }
  } else {
    for (auto& dl : dloads)
      for (auto& s : segs) s->dim_width[dl.di] = ds->dims[dl.di].width;
  }

  for (auto& s : segs) {
    ds->n_rows += s->n_rows;
    auto pos = std::upper_bound(ds->segments.begin(), ds->segments.end(), s, [](const std::unique_ptr<Segment>& a, const std::unique_ptr<Segment>& b) {
      return a->start_ms != b->start_ms ? a->start_ms < b->start_ms : a->partition < b->partition;
    });
    ds->segments.insert(pos, std::move(s));
  }
  // older segments may lack columns that arrived later
  for (auto& s : ds->segments) {
    s->dim_data.resize(ds->dims.size(), nullptr);
    s->dim_hll.resize(ds->dims.size(), nullptr);
    s->dim_width.resize(ds->dims.size(), 0);
    s->metric_data.resize(ds->metrics.size(), nullptr);
  }
}

// ---- metadata -----------------------------------------------------------------------------------

std::string Store::metadata_json(const std::string& name) const {
  auto ds = get(name);
  std::shared_lock<std::shared_mutex> lk(ds->mu);
  // MetadataResponse shape: client/DruidMessages.scala:25-53 -> {id, intervals, columns{name->{type,size,cardinality}}, size, numRows}
  std::string o = "{";
  o += "\"id\":";
  json_escape(o, "merged");
  o += ",\"intervals\":[";
  int64_t lo = INT64_MAX, hi = INT64_MIN;
  size_t bytes = 0;
  for (auto& s : ds->segments) {
    if (s->time_vals.empty()) continue;
    lo = std::min(lo, s->time_vals.front());
    hi = std::max(hi, s->time_vals.back());
    bytes += s->device_bytes;
  }
  if (lo <= hi) {
    o += "\"" + format_iso_instant(lo) + "/" + format_iso_instant(hi + 1) + "\"";
  }
  o += "],\"columns\":{";
  o += "\"__time\":{\"type\":\"LONG\",\"size\":0,\"cardinality\":null}";
  for (auto& d : ds->dims) {
    o += ",";
    json_escape(o, d.name);
    o += ":{\"type\":\"STRING\",\"size\":" + std::to_string((long long)ds->n_rows * d.width) + ",\"cardinality\":" + std::to_string((long long)d.cardinality) + "}";
  }
  for (auto& m : ds->metrics) {
    o += ",";
    json_escape(o, m.name);
    o += std::string(":{\"type\":\"") + (m.is_float ? "FLOAT" : "LONG") + "\",\"size\":" + std::to_string((long long)ds->n_rows * 8) + ",\"cardinality\":null}";
  }
  o += "},\"size\":" + std::to_string((unsigned long long)bytes) + ",\"numRows\":" + std::to_string((long long)ds->n_rows);
  o += ",\"queryGranularity\":";
  json_escape(o, ds->query_granularity);
  o += ",\"segmentGranularity\":";
  json_escape(o, ds->segment_granularity);
  if (lo <= hi)
    o += ",\"timeBoundary\":{\"minTime\":\"" + format_iso_instant(lo) + "\",\"maxTime\":\"" + format_iso_instant(hi) + "\",\"minMs\":" + std::to_string((long long)lo) +
         ",\"maxMs\":" + std::to_string((long long)hi) + "}";
  o += "}";
  return o;
}

std::string segment_identifier(const DataSource& ds, const Segment& s) {
  return ds.name + "_" + format_iso_instant(s.start_ms) + "/" + format_iso_instant(s.end_ms) + "_v1_" + std::to_string(s.partition);
}

std::string Store::segments_json(const std::string& name) const {
  auto ds = get(name);
  std::shared_lock<std::shared_mutex> lk(ds->mu);
  // DruidSegmentInfo shape: metadata/DruidMetadataCache.scala:64-76 {identifier, interval, version, size, shardSpec}
  std::string o = "[";
  bool first = true;
  for (auto& s : ds->segments) {
    if (!first) o += ",";
    first = false;
    std::string itv = format_iso_instant(s->start_ms) + "/" + format_iso_instant(s->end_ms);
    o += "{\"identifier\":";
    json_escape(o, segment_identifier(*ds, *s));
    o += ",\"interval\":\"" + itv + "\",\"version\":\"v1\",\"size\":" + std::to_string((unsigned long long)s->device_bytes);
    o += ",\"numRows\":" + std::to_string((long long)s->n_rows);
    o += ",\"shardSpec\":{\"type\":\"linear\",\"partitionNum\":" + std::to_string(s->partition) + "}";
    o += ",\"server\":\"gpu:" + std::to_string(device_) + "\"}";
  }
  o += "]";
  return o;
}

}  // namespace sdo

// ---- persistence: encoded segments on disk --------------------------------------------------------------------------------
// The reference re-reads its OLAP index from Druid's deep storage after a restart (the indexing service wrote it once,
// SLD/client/DruidOverlordClient.scala:51-129); here the already encoded, time-sorted segments are written as they sit in
// HBM and read back without going through dictionary encoding or the time sort again.
namespace sdo {
namespace {

constexpr char SAVE_MAGIC[8] = {'S', 'D', 'O', 'S', 'E', 'G', '0', '1'};

struct Writer {
  FILE* f;
  const std::string& path;
  void raw(const void* p, size_t n) {
    if (n && fwrite(p, 1, n, f) != n) throw std::runtime_error("write to '" + path + "' failed: " + strerror(errno));
  }
  template <class T> void pod(T v) { raw(&v, sizeof v); }
  void str(const std::string& s) { pod<uint64_t>(s.size()); raw(s.data(), s.size()); }
  template <class T> void vec(const std::vector<T>& v) { pod<uint64_t>(v.size()); raw(v.data(), v.size() * sizeof(T)); }
};
struct Reader {
  FILE* f;
  const std::string& path;
  void raw(void* p, size_t n) {
    if (n && fread(p, 1, n, f) != n) throw Invalid("'" + path + "' is truncated or not a saved datasource");
  }
  template <class T> T pod() { T v; raw(&v, sizeof v); return v; }
  std::string str() {
    uint64_t n = pod<uint64_t>();
    if (n > (1ull << 32)) throw Invalid("'" + path + "' is corrupt (string length)");
    std::string s(n, '\0');
    raw(&s[0], n);
    return s;
  }
  template <class T> std::vector<T> vec() {
    uint64_t n = pod<uint64_t>();
    if (n > (1ull << 36) / sizeof(T)) throw Invalid("'" + path + "' is corrupt (array length)");
    std::vector<T> v(n);
    raw(v.data(), n * sizeof(T));
    return v;
  }
};
struct FileCloser { void operator()(FILE* f) const { if (f) fclose(f); } };

}  // namespace

void Store::save(const std::string& name, const std::string& path) const {
  auto ds = get(name);
  std::shared_lock<std::shared_mutex> lk(ds->mu);
  std::unique_ptr<FILE, FileCloser> fp(fopen(path.c_str(), "wb"));
  if (!fp) throw Invalid("cannot open '" + path + "' for writing: " + strerror(errno));
  Writer w{fp.get(), path};
  w.raw(SAVE_MAGIC, 8);
  w.str(ds->segment_granularity);
  w.str(ds->query_granularity);
  w.pod<uint64_t>(ds->dims.size());
  for (auto& d : ds->dims) {
    w.str(d.name);
    w.pod<int32_t>(d.enc == DimEncoding::Dict ? 0 : 1);
    w.pod<int32_t>(d.src_type);
    w.pod<int32_t>(d.width);
    w.pod<int64_t>(d.cardinality);
    w.vec(d.dict->offsets);
    w.str(d.dict->bytes);
  }
  w.pod<uint64_t>(ds->metrics.size());
  for (auto& m : ds->metrics) {
    w.str(m.name);
    w.pod<int32_t>((m.is_float ? 1 : 0) | (m.constant ? 2 : 0) | (m.seen ? 4 : 0));
    w.pod<int64_t>(m.const_value);
  }
  w.pod<uint64_t>(ds->segments.size());
  if (device_ >= 0) SDO_CUDA(cudaSetDevice(device_));
  std::vector<char> host;
  auto column = [&](const void* dev, size_t bytes) {
    w.pod<uint64_t>(dev ? bytes : 0);
    if (!dev || !bytes) return;
    host.resize(bytes);
    SDO_CUDA(cudaMemcpy(host.data(), dev, bytes, cudaMemcpyDeviceToHost));
    w.raw(host.data(), bytes);
  };
  for (auto& sp : ds->segments) {
    const Segment& s = *sp;
    w.pod<int64_t>(s.start_ms);
    w.pod<int64_t>(s.end_ms);
    w.pod<int32_t>(s.partition);
    w.pod<int64_t>(s.n_rows);
    w.vec(s.time_vals);
    w.vec(s.time_rows);
    for (size_t d = 0; d < ds->dims.size(); d++) {
      w.pod<int32_t>(d < s.dim_width.size() ? s.dim_width[d] : 0);
      column(d < s.dim_data.size() ? s.dim_data[d] : nullptr, (size_t)s.n_rows * (d < s.dim_width.size() ? s.dim_width[d] : 0));
    }
    for (size_t m = 0; m < ds->metrics.size(); m++) column(m < s.metric_data.size() ? s.metric_data[m] : nullptr, (size_t)s.n_rows * 8);
  }
  if (fflush(fp.get()) != 0) throw std::runtime_error("write to '" + path + "' failed: " + strerror(errno));
}

void Store::restore(const std::string& name, const std::string& path) {
  if (device_ < 0) throw NoDevice("restoring a datasource needs a CUDA device: segments live in HBM");
  std::unique_ptr<FILE, FileCloser> fp(fopen(path.c_str(), "rb"));
  if (!fp) throw NotFound("cannot open '" + path + "': " + strerror(errno));
  Reader r{fp.get(), path};
  char magic[8];
  r.raw(magic, 8);
  if (memcmp(magic, SAVE_MAGIC, 8) != 0) throw Invalid("'" + path + "' is not a saved datasource (bad magic)");
  auto ds = std::make_shared<DataSource>();
  ds->name = name;
  ds->segment_granularity = r.str();
  ds->query_granularity = r.str();
  const uint64_t n_dims = r.pod<uint64_t>();
  if (n_dims > 4096) throw Invalid("'" + path + "' is corrupt (dimension count)");
  for (uint64_t i = 0; i < n_dims; i++) {
    DimMeta d;
    d.name = r.str();
    d.enc = r.pod<int32_t>() == 0 ? DimEncoding::Dict : DimEncoding::Direct;
    d.src_type = r.pod<int32_t>();
    d.width = r.pod<int32_t>();
    d.cardinality = r.pod<int64_t>();
    auto dict = std::make_shared<Dictionary>();
    dict->offsets = r.vec<int64_t>();
    dict->bytes = r.str();
    if (dict->offsets.empty() || dict->offsets.front() != 0 || (uint64_t)dict->offsets.back() != dict->bytes.size() ||
        (d.width != 1 && d.width != 2 && d.width != 4))
      throw Invalid("'" + path + "' is corrupt (dictionary of '" + d.name + "')");
    d.dict = dict;
    ds->dims.push_back(std::move(d));
  }
  const uint64_t n_metrics = r.pod<uint64_t>();
  if (n_metrics > 4096) throw Invalid("'" + path + "' is corrupt (metric count)");
  for (uint64_t i = 0; i < n_metrics; i++) {
    MetricMeta m;
    m.name = r.str();
    const int32_t flags = r.pod<int32_t>();
    m.is_float = flags & 1;
    m.constant = flags & 2;
    m.seen = flags & 4;
    m.const_value = r.pod<int64_t>();
    ds->metrics.push_back(std::move(m));
  }
  const uint64_t n_segs = r.pod<uint64_t>();
  if (n_segs > (1u << 24)) throw Invalid("'" + path + "' is corrupt (segment count)");
  SDO_CUDA(cudaSetDevice(device_));
  std::vector<char> host;
  try {
    for (uint64_t si = 0; si < n_segs; si++) {
      std::unique_ptr<Segment> s(new Segment());
      s->start_ms = r.pod<int64_t>();
      s->end_ms = r.pod<int64_t>();
      s->partition = r.pod<int32_t>();
      s->n_rows = r.pod<int64_t>();
      if (s->n_rows < 0 || s->n_rows > 0x7fffffffll) throw Invalid("'" + path + "' is corrupt (segment rows)");
      s->time_vals = r.vec<int64_t>();
      s->time_rows = r.vec<int64_t>();
      if (s->time_rows.size() != s->time_vals.size() + 1 || (!s->time_rows.empty() && s->time_rows.back() != s->n_rows))
        throw Invalid("'" + path + "' is corrupt (time index)");
      s->dim_data.assign(n_dims, nullptr);
      s->dim_hll.assign(n_dims, nullptr);
      s->dim_width.assign(n_dims, 0);
      s->metric_data.assign(n_metrics, nullptr);
      ds->segments.push_back(std::move(s));
      Segment& sg = *ds->segments.back();
      auto column = [&](size_t expect) -> void* {
        const uint64_t bytes = r.pod<uint64_t>();
        if (!bytes) return nullptr;
        if (bytes != expect) throw Invalid("'" + path + "' is corrupt (column size)");
        host.resize(bytes);
        r.raw(host.data(), bytes);
        // same slack as a loaded column: vector loads may run past the last row
        const size_t padded = round_up((size_t)bytes + 64, 256);
        void* dev = nullptr;
        SDO_CUDA(cudaMalloc(&dev, padded));
        SDO_CUDA(cudaMemset((char*)dev + bytes, 0, padded - bytes));
        SDO_CUDA(cudaMemcpy(dev, host.data(), bytes, cudaMemcpyHostToDevice));
        sg.device_bytes += padded;
        return dev;
      };
      for (uint64_t d = 0; d < n_dims; d++) {
        sg.dim_width[d] = r.pod<int32_t>();
        sg.dim_data[d] = column((size_t)sg.n_rows * sg.dim_width[d]);
      }
      for (uint64_t m = 0; m < n_metrics; m++) sg.metric_data[m] = column((size_t)sg.n_rows * 8);
      ds->n_rows += sg.n_rows;
    }
  } catch (...) {
    for (auto& sp : ds->segments) free_segment(*sp);
    throw;
  }
  ds->version = 1;
  std::lock_guard<std::mutex> lk(mu_);
  if (sources_.count(name)) {
    for (auto& sp : ds->segments) free_segment(*sp);
    throw Invalid("datasource '" + name + "' already exists");
  }
  sources_[name] = ds;
}

}  // namespace sdo
