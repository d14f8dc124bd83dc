// Run-time specialisation of the scan kernel: the same hand-written CUDA source (scan_kernel.cuh) is
// compiled by NVRTC for sm_100a with the query SHAPE (filter program, column widths, aggregator list,
// strategy) spelled out as constexpr tables, so the interpreter's loops unroll and its dispatch folds away.
// Values (id ranges, bitsets, strides, pointers) stay run-time parameters: one compiled shape serves every
// query of that shape, and shapes are cached for the life of the process.
//
// NVRTC and the driver API are loaded with dlopen, so the library itself has no link-time dependency on
// them (it must load on a machine without a GPU for planning-only use).  If either is missing or a
// compile fails, the caller falls back to the ahead-of-time kernels -- still the GPU path.
#include "jit.hpp"

#include <cuda.h>
#include <dlfcn.h>
#include <nvrtc.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <map>
#include <mutex>
#include <sstream>

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
  static Api a;
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
  static const char* nv_env = getenv("SDO_JIT_NV");  // experiment knob: vectors per dense-path accumulate call
  static const char* wt_env = getenv("SDO_JIT_WARP_TICKETS");  // experiment knob: 0 = block tickets
  std::string s = std::string("#define SDO_JIT 1\n") + "#define SDO_WARP_TICKETS " + (wt_env && *wt_env == '0' ? "0" : "1") + "\n#define SDO_DENSE_NV " + std::string(nv_env && *nv_env ? nv_env : "1") + "\nnamespace sdo { namespace jit {\n";
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
  static const char* mb_env = getenv("SDO_JIT_MINBLOCKS");  // experiment knob: resident blocks per SM the compiler aims for
  // Registers per thread follow from the blocks per SM the compiler is told to fit (measured on SF10): wide shapes
  // (seven aggregates, six of them expressions) are fastest at 3 blocks / 85 registers -- 0.71 ms vs 0.86 at 4 and 1.17
  // at 6, the rest being spills; narrow ones (one or two aggregates, no sketch) gain a few percent at 6 blocks / 40 registers.
  const int weight = p.n_aggs + 2 * p.n_exprs + p.n_dims;
  const std::string min_blocks = mb_env && *mb_env ? mb_env : (weight >= 8 ? "3" : (weight >= 3 || p.n_hll > 0) ? "4" : "6");
  s += "extern \"C\" __global__ void __launch_bounds__(256, " + min_blocks + ") sdo_scan_jit(const __grid_constant__ sdo::ScanParams p, const uint32_t* __restrict__ coarse) {\n";
  s += "  sdo::scan_body<" + std::to_string(strategy) + ">(p, coarse);\n}\n";
  return s;
}

std::mutex g_mu;
std::map<std::string, JitKernel> g_cache;  // keyed by the prelude text (= the shape)
bool g_warned = false;

}  // namespace

bool jit_available(std::string* why) {
  Api& a = api();
  if (why) *why = a.why;
  return a.ok;
}

const JitKernel* jit_get(int strategy, const ScanParams& p, double* compile_ms) {
  if (compile_ms) *compile_ms = 0;
  Api& a = api();
  if (!a.ok) return nullptr;
  const std::string src = prelude(strategy, p);
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_cache.find(src);
  if (it != g_cache.end()) return it->second.func ? &it->second : nullptr;
  auto t0 = std::chrono::steady_clock::now();
  JitKernel k;
  auto fail = [&](const std::string& m) -> const JitKernel* {
    if (!g_warned) {
      fprintf(stderr, "[sdo] run-time kernel specialisation unavailable (%s); using the ahead-of-time kernels\n", m.c_str());
      g_warned = true;
    }
    g_cache[src] = JitKernel();  // remember the failure: do not retry this shape
    return nullptr;
  };
  nvrtcProgram prog = nullptr;
  const char* hdr_src[] = {a.src_kernel.c_str(), a.src_params.c_str(), a.src_xxh.c_str()};
  const char* hdr_name[] = {"scan_kernel.cuh", "kernels.hpp", "xxh64.hpp"};
  if (a.CreateProgram(&prog, src.c_str(), "sdo_scan_jit.cu", 3, hdr_src, hdr_name) != NVRTC_SUCCESS) return fail("nvrtcCreateProgram");
  const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo"};
  nvrtcResult rc = a.CompileProgram(prog, 3, opts);
  if (rc != NVRTC_SUCCESS) {
    size_t n = 0;
    a.GetProgramLogSize(prog, &n);
    std::string log(n, 0);
    if (n) a.GetProgramLog(prog, &log[0]);
    a.DestroyProgram(&prog);
    return fail("nvrtc compile error: " + log.substr(0, 2000));
  }
  size_t n = 0;
  if (a.GetCUBINSize(prog, &n) != NVRTC_SUCCESS || !n) { a.DestroyProgram(&prog); return fail("nvrtcGetCUBINSize"); }
  std::string cubin(n, 0);
  a.GetCUBIN(prog, &cubin[0]);
  a.DestroyProgram(&prog);
  CUmodule mod = nullptr;
  CUfunction fn = nullptr;
  if (a.ModuleLoadData(&mod, cubin.data()) != CUDA_SUCCESS) return fail("cuModuleLoadData");
  if (a.ModuleGetFunction(&fn, mod, "sdo_scan_jit") != CUDA_SUCCESS) return fail("cuModuleGetFunction");
  a.FuncSetAttribute(fn, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, 227 * 1024 - 64);
  k.func = (void*)fn;
  k.compile_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  if (compile_ms) *compile_ms = k.compile_ms;
  auto ins = g_cache.emplace(src, k);
  return &ins.first->second;
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
