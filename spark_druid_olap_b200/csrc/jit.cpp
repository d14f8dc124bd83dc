// Run-time specialisation of the scan kernel: the same hand-written CUDA source (scan_kernel.cuh) is
// compiled by NVRTC for sm_100a with the query SHAPE (filter program, column widths, aggregator list,
// strategy) spelled out as constexpr tables, so the interpreter's loops unroll and its dispatch folds away.
// Values (id ranges, bitsets, strides, pointers) stay run-time parameters: one compiled shape serves every
// query of that shape.  Shapes are cached for the life of the process and their cubins on disk (SDO_JIT_CACHE_DIR, default
// ~/.cache/sdo_engine/jit, keyed by a hash of shape + kernel sources), and a new shape compiles on a background thread
// while the queries that need it run on the ahead-of-time kernel: no query waits for the compiler unless it asks to
// (context.gpuJit = true).
//
// NVRTC and the driver API are loaded with dlopen, so the library itself has no link-time dependency on
// them (it must load on a machine without a GPU for planning-only use).  If either is missing or a
// compile fails, the caller falls back to the ahead-of-time kernels -- still the GPU path.
#include "jit.hpp"

#include <cuda.h>
#include <dlfcn.h>
#include <nvrtc.h>

#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <thread>

#include "xxh64.hpp"

namespace sdo {
namespace {

struct Api {
  bool ok = false;
  std::string why;
  // nvrtc
  nvrtcResult (*CreateProgram)(nvrtcProgram*, const char*, const char*, int, const char* const*, const char* const*) = nullptr;
  nvrtcResult (*CompileProgram)(nvrtcProgram, int, const char* const*) = nullptr;
  nvrtcResult (*GetCUBINSize)(nvrtcProgram, size_t*) = nullptr;
  nvrtcResult (*GetCUBIN)(nvrtcProgram, char*) = nullptr;
  nvrtcResult (*GetProgramLogSize)(nvrtcProgram, size_t*) = nullptr;
  nvrtcResult (*GetProgramLog)(nvrtcProgram, char*) = nullptr;
  nvrtcResult (*DestroyProgram)(nvrtcProgram*) = nullptr;
  // driver
  CUresult (*ModuleLoadData)(CUmodule*, const void*) = nullptr;
  CUresult (*ModuleGetFunction)(CUfunction*, CUmodule, const char*) = nullptr;
  CUresult (*FuncSetAttribute)(CUfunction, CUfunction_attribute, int) = nullptr;
  CUresult (*Occupancy)(int*, CUfunction, int, size_t) = nullptr;
  CUresult (*LaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, CUstream, void**, void**) = nullptr;
  std::string src_kernel, src_params, src_xxh;
};

void* open_any(std::initializer_list<const char*> names) {
  for (const char* n : names)
    if (void* h = dlopen(n, RTLD_NOW | RTLD_LOCAL)) return h;
  return nullptr;
}
bool read_file(const std::string& path, std::string* out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::stringstream ss;
  ss << f.rdbuf();
  *out = ss.str();
  return true;
}

Api& api() {
  static Api& a = *new Api;  // never destroyed: a background compilation may still read it while the process exits
  static std::once_flag once;
  std::call_once(once, [] {
    void* rtc = open_any({"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so"});
    void* drv = open_any({"libcuda.so.1", "libcuda.so"});
    if (!rtc) { a.why = "libnvrtc not found"; return; }
    if (!drv) { a.why = "libcuda not found"; return; }
#define RTC(sym) *(void**)&a.sym = dlsym(rtc, "nvrtc" #sym); if (!a.sym) { a.why = "nvrtc" #sym " missing"; return; }
    RTC(CreateProgram) RTC(CompileProgram) RTC(GetCUBINSize) RTC(GetCUBIN) RTC(GetProgramLogSize) RTC(GetProgramLog) RTC(DestroyProgram)
#undef RTC
    *(void**)&a.ModuleLoadData = dlsym(drv, "cuModuleLoadData");
    *(void**)&a.ModuleGetFunction = dlsym(drv, "cuModuleGetFunction");
    *(void**)&a.FuncSetAttribute = dlsym(drv, "cuFuncSetAttribute");
    *(void**)&a.Occupancy = dlsym(drv, "cuOccupancyMaxActiveBlocksPerMultiprocessor");
    *(void**)&a.LaunchKernel = dlsym(drv, "cuLaunchKernel");
    if (!a.ModuleLoadData || !a.ModuleGetFunction || !a.FuncSetAttribute || !a.Occupancy || !a.LaunchKernel) { a.why = "driver API symbols missing"; return; }
    // kernel sources sit next to the library: <dir>/csrc/*.cuh|hpp
    Dl_info info;
    if (!dladdr((void*)&api, &info) || !info.dli_fname) { a.why = "cannot locate the library on disk"; return; }
    std::string dir = info.dli_fname;
    size_t k = dir.find_last_of('/');
    dir = k == std::string::npos ? "." : dir.substr(0, k);
    if (!read_file(dir + "/csrc/scan_kernel.cuh", &a.src_kernel) || !read_file(dir + "/csrc/kernels.hpp", &a.src_params) ||
        !read_file(dir + "/csrc/xxh64.hpp", &a.src_xxh)) { a.why = "kernel sources not found under " + dir + "/csrc"; return; }
    a.ok = true;
  });
  return a;
}

template <class T, class F>
std::string table(const char* name, int n, F f) {
  std::string s = std::string("constexpr int ") + name + "[" + std::to_string(n + 1) + "] = {";
  for (int i = 0; i < n; i++) s += std::to_string((long long)(T)f(i)) + ",";
  s += "0};\n";
  return s;
}

std::string prelude(int strategy, const ScanParams& p) {
  const char* nv_env = getenv("SDO_JIT_NV");  // experiment knob: vectors per dense-path accumulate call
  const char* wt_env = getenv("SDO_JIT_WARP_TICKETS");  // experiment knob: 0 = block tickets
  const char* rf_env = getenv("SDO_JIT_HLL_READ_FIRST");  // experiment knob: 0 = always issue the shared-memory atomicMax
  std::string s = std::string("#define SDO_JIT 1\n") + (rf_env && *rf_env == '0' ? "#define SDO_HLL_READ_FIRST 0\n" : "") + "#define SDO_WARP_TICKETS " + (wt_env && *wt_env == '0' ? "0" : "1") + "\n#define SDO_DENSE_NV " + std::string(nv_env && *nv_env ? nv_env : (strategy == STRAT_SPAN || strategy == STRAT_EMIT ? "4" : "1")) + "\nnamespace sdo { namespace jit {\n";
  s += "constexpr int n_ops = " + std::to_string(p.n_ops) + ";\n";
  s += table<int>("op_code", p.n_ops, [&](int i) { return p.op_code[i]; });
  s += table<int>("op_arg", p.n_ops, [&](int i) { return p.op_arg[i]; });
  s += table<int>("leaf_kind", p.n_leaves, [&](int i) { return p.leaves[i].kind; });
  s += table<int>("leaf_width", p.n_leaves, [&](int i) { return p.leaves[i].width; });
  s += table<int>("leaf_col", p.n_leaves, [&](int i) { return p.leaves[i].col; });
  s += "constexpr int n_dims = " + std::to_string(p.n_dims) + ";\n";
  s += table<int>("dim_width", p.n_dims, [&](int i) { return p.dims[i].width; });
  s += table<int>("dim_col", p.n_dims, [&](int i) { return p.dims[i].col; });
  s += table<int>("dim_remap", p.n_dims, [&](int i) { return p.dims[i].remap != nullptr; });
  s += "constexpr int n_aggs = " + std::to_string(p.n_aggs) + ";\n";
  s += table<int>("agg_op", p.n_aggs, [&](int i) { return p.aggs[i].op; });
  s += table<int>("agg_col", p.n_aggs, [&](int i) { return p.aggs[i].col; });
  s += table<int>("agg_slot", p.n_aggs, [&](int i) { return p.aggs[i].slot; });
  s += table<int>("agg_leaf", p.n_aggs, [&](int i) { return p.aggs[i].leaf; });
  s += table<int>("agg_width", p.n_aggs, [&](int i) { return p.aggs[i].width; });
  s += table<int>("agg_hll_index", p.n_aggs, [&](int i) { return p.aggs[i].hll_index; });
  s += table<int>("agg_hll_mode", p.n_aggs, [&](int i) { return p.aggs[i].hll_mode; });
  s += "constexpr int n_slots = " + std::to_string(p.n_slots) + ";\n";
  s += table<int>("slot_op", p.n_slots, [&](int i) { return p.slot_op[i]; });
  s += table<int>("expr_n", p.n_exprs, [&](int i) { return p.exprs[i].n_ops; });
  for (const char* which : {"expr_code", "expr_arg"}) {
    s += std::string("constexpr int ") + which + "[" + std::to_string(p.n_exprs + 1) + "][" + std::to_string(MAX_EXPR_OPS) + "] = {";
    for (int e = 0; e < p.n_exprs; e++) {
      s += "{";
      for (int k = 0; k < MAX_EXPR_OPS; k++) s += std::to_string(k < p.exprs[e].n_ops ? (which[5] == 'c' ? p.exprs[e].code[k] : p.exprs[e].arg[k]) : 0) + ",";
      s += "},";
    }
    s += "{0}};\n";
  }
  s += "constexpr int n_hll = " + std::to_string(p.n_hll) + ";\n";
  s += "constexpr int hll_place = " + std::to_string(p.hll_mode) + ";\n";
  s += "constexpr int hll_wide = " + std::to_string(p.hll_wide) + ";\n";
  s += "constexpr unsigned ticket_shift = " + std::to_string(p.ticket_shift) + ";\n";
  s += "constexpr int ignore_key_base = " + std::to_string(p.ignore_key_base) + ";\n";
  s += "constexpr int n_cols = " + std::to_string(p.n_cols) + ";\n";
  s += "}}\n#include \"scan_kernel.cuh\"\n";
  const char* mb_env = getenv("SDO_JIT_MINBLOCKS");  // experiment knob: resident blocks per SM the compiler aims for
  // Registers per thread follow from the blocks per SM the compiler is told to fit (measured on SF10): wide shapes
  // (seven aggregates, six of them expressions) are fastest at 3 blocks / 85 registers -- 0.71 ms vs 0.86 at 4 and 1.17
  // at 6, the rest being spills; narrow ones (one or two aggregates, no sketch) gain a few percent at 6 blocks / 40 registers.
  const int weight = p.n_aggs + 2 * p.n_exprs + p.n_dims;
  // The span strategy folds a call's rows in registers and touches its shared-memory cell once per call: all 16 rows of a
  // pass in one call at 6 blocks per SM (SF100 day timeseries + HLL: 1.26 ms before, 0.99 ms folded 4 rows at a time,
  // 0.90 ms = the HBM roofline with 16).  A narrow filtered shape (TopN over one sum) gains 4 % at 8 blocks.
  const std::string min_blocks = mb_env && *mb_env ? mb_env
                                 : strategy == STRAT_EMIT ? "3"  // 16 rows per call keep keys, owners, ranks and positions live: 300 M rows 4.7 ms at 6 blocks (spills), 2.3 ms at 3
                                 : strategy == STRAT_SPAN && weight < 8 ? "6"
                                 : (weight >= 8 ? "3" : (weight >= 3 || p.n_hll > 0) ? "4" : (p.n_ops > 0 ? "8" : "6"));
  s += "extern \"C\" __global__ void __launch_bounds__(256, " + min_blocks + ") sdo_scan_jit(const __grid_constant__ sdo::ScanParams p, const uint32_t* __restrict__ coarse) {\n";
  s += "  sdo::scan_body<" + std::to_string(strategy) + ">(p, coarse);\n}\n";
  return s;
}

// One cache entry per shape (keyed by the prelude text).  A shape is compiled once per process -- by the query that first
// needs it (wait = true) or by a background thread while queries of that shape run on the ahead-of-time kernel -- and its
// cubin is kept on disk, keyed by a hash of everything that went into it, so a later process loads it in about a millisecond.
struct Entry {
  enum State { Compiling, Compiled, Ready, Failed } state = Compiling;
  std::string cubin;   // Compiled: waiting for a query thread (which has the CUDA context) to load it
  JitKernel kernel;    // Ready
  double compile_ms = 0;
};
// (heap objects that are never destroyed: a detached compile thread may outlive static destruction at process exit)
std::mutex& g_mu = *new std::mutex;
std::condition_variable& g_cv = *new std::condition_variable;
std::map<std::string, std::shared_ptr<Entry>>& g_cache = *new std::map<std::string, std::shared_ptr<Entry>>;
bool g_warned = false;

void warn_once(const std::string& m) {
  if (!g_warned) {
    fprintf(stderr, "[sdo] run-time kernel specialisation unavailable (%s); using the ahead-of-time kernels\n", m.c_str());
    g_warned = true;
  }
}

std::string cache_dir() {
  const char* d = getenv("SDO_JIT_CACHE_DIR");
  if (d && *d) return d;
  const char* x = getenv("XDG_CACHE_HOME");
  if (x && *x) return std::string(x) + "/sdo_engine/jit";
  const char* h = getenv("HOME");
  return std::string(h && *h ? h : "/tmp") + "/.cache/sdo_engine/jit";
}
void make_dirs(const std::string& path) {
  for (size_t i = 1; i <= path.size(); i++)
    if (i == path.size() || path[i] == '/') mkdir(path.substr(0, i).c_str(), 0755);
}
std::string cache_file(const Api& a, const std::string& src) {
  // everything the cubin depends on: the shape, the kernel sources, the target
  uint64_t h = xxh64_bytes(src.data(), src.size(), 1);
  h = xxh64_bytes(a.src_kernel.data(), a.src_kernel.size(), h);
  h = xxh64_bytes(a.src_params.data(), a.src_params.size(), h);
  h = xxh64_bytes(a.src_xxh.data(), a.src_xxh.size(), h);
  char name[64];
  snprintf(name, sizeof name, "/sm_100a-%016llx.cubin", (unsigned long long)h);
  return cache_dir() + name;
}

// NVRTC only: needs no CUDA context, so it can run on any thread
bool compile_cubin(Api& a, const std::string& src, std::string* cubin, std::string* err) {
  nvrtcProgram prog = nullptr;
  const char* hdr_src[] = {a.src_kernel.c_str(), a.src_params.c_str(), a.src_xxh.c_str()};
  const char* hdr_name[] = {"scan_kernel.cuh", "kernels.hpp", "xxh64.hpp"};
  if (a.CreateProgram(&prog, src.c_str(), "sdo_scan_jit.cu", 3, hdr_src, hdr_name) != NVRTC_SUCCESS) { *err = "nvrtcCreateProgram"; return false; }
  const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo"};
  if (a.CompileProgram(prog, 3, opts) != NVRTC_SUCCESS) {
    size_t n = 0;
    a.GetProgramLogSize(prog, &n);
    std::string log(n, 0);
    if (n) a.GetProgramLog(prog, &log[0]);
    a.DestroyProgram(&prog);
    *err = "nvrtc compile error: " + log.substr(0, 2000);
    return false;
  }
  size_t n = 0;
  if (a.GetCUBINSize(prog, &n) != NVRTC_SUCCESS || !n) { a.DestroyProgram(&prog); *err = "nvrtcGetCUBINSize"; return false; }
  cubin->assign(n, 0);
  a.GetCUBIN(prog, &(*cubin)[0]);
  a.DestroyProgram(&prog);
  return true;
}

// cubin from the disk cache, else compiled (and written to the cache); fills the entry, takes g_mu only to publish
void produce(Api& a, const std::string& src, std::shared_ptr<Entry> e) {
  const auto t0 = std::chrono::steady_clock::now();
  static const bool no_disk = [] { const char* v = getenv("SDO_JIT_DISK_CACHE"); return v && v[0] == '0'; }();
  const std::string file = cache_file(a, src);
  std::string cubin, err;
  bool ok = !no_disk && read_file(file, &cubin) && cubin.size() > 64;
  if (!ok) {
    ok = compile_cubin(a, src, &cubin, &err);
    if (ok && !no_disk) {
      make_dirs(cache_dir());
      const std::string tmp = file + "." + std::to_string((long long)getpid()) + ".tmp";
      std::ofstream f(tmp, std::ios::binary);
      if (f && f.write(cubin.data(), (std::streamsize)cubin.size()) && (f.close(), true)) rename(tmp.c_str(), file.c_str());
    }
  }
  std::lock_guard<std::mutex> lk(g_mu);
  e->compile_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  if (ok) { e->cubin = std::move(cubin); e->state = Entry::Compiled; }
  else { e->state = Entry::Failed; warn_once(err); }
  g_cv.notify_all();
}

}  // namespace

bool jit_available(std::string* why) {
  Api& a = api();
  if (why) *why = a.why;
  return a.ok;
}

// wait = true: compile now if the shape is new (the caller pays, *compile_ms says how much).  wait = false: a new shape
// starts compiling on a background thread and nullptr comes back -- the caller runs the ahead-of-time kernel this time.
const JitKernel* jit_get(int strategy, const ScanParams& p, double* compile_ms, bool wait) {
  if (compile_ms) *compile_ms = 0;
  Api& a = api();
  if (!a.ok) return nullptr;
  const std::string src = prelude(strategy, p);
  std::unique_lock<std::mutex> lk(g_mu);
  std::shared_ptr<Entry> e;
  auto it = g_cache.find(src);
  if (it == g_cache.end()) {
    e = std::make_shared<Entry>();
    g_cache[src] = e;
    if (wait) {
      lk.unlock();
      produce(a, src, e);
      lk.lock();
      if (compile_ms) *compile_ms = e->compile_ms;
    } else {
      std::thread([&a, src, e] { produce(a, src, e); }).detach();
      return nullptr;
    }
  } else {
    e = it->second;
    if (e->state == Entry::Compiling) {
      if (!wait) return nullptr;
      g_cv.wait(lk, [&] { return e->state != Entry::Compiling; });
    }
  }
  if (e->state == Entry::Compiled) {
    // load the module on this thread: it has the device's context current
    CUmodule mod = nullptr;
    CUfunction fn = nullptr;
    if (a.ModuleLoadData(&mod, e->cubin.data()) != CUDA_SUCCESS || a.ModuleGetFunction(&fn, mod, "sdo_scan_jit") != CUDA_SUCCESS) {
      e->state = Entry::Failed;
      warn_once("cuModuleLoadData / cuModuleGetFunction");
    } else {
      a.FuncSetAttribute(fn, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, 227 * 1024 - 64);
      e->kernel.func = (void*)fn;
      e->kernel.compile_ms = e->compile_ms;
      e->state = Entry::Ready;
    }
    e->cubin.clear();
    e->cubin.shrink_to_fit();
  }
  return e->state == Entry::Ready ? &e->kernel : nullptr;
}

int jit_pending() {
  std::lock_guard<std::mutex> lk(g_mu);
  int n = 0;
  for (auto& kv : g_cache) n += kv.second->state == Entry::Compiling;
  return n;
}

int jit_blocks_per_sm(const JitKernel* k, size_t smem) {
  int n = 0;
  if (api().Occupancy(&n, (CUfunction)k->func, SCAN_THREADS, smem) != CUDA_SUCCESS) return 0;
  return n;
}

bool jit_launch(const JitKernel* k, const ScanParams& p, int grid, size_t smem, cudaStream_t st) {
  const uint32_t* coarse = (const uint32_t*)(p.spans + p.n_spans + 1);
  void* args[] = {(void*)&p, (void*)&coarse};
  return api().LaunchKernel((CUfunction)k->func, (unsigned)grid, 1, 1, SCAN_THREADS, 1, 1, (unsigned)smem, (CUstream)st, args, nullptr) == CUDA_SUCCESS;
}

}  // namespace sdo
