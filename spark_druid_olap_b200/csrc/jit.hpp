// Run-time specialisation of the scan kernel by query shape (see jit.cpp).
#pragma once
#include <string>

#include "kernels.hpp"

namespace sdo {

struct JitKernel {
  void* func = nullptr;  // CUfunction
  double compile_ms = 0;
};

bool jit_available(std::string* why);
// Kernel for (strategy, shape of p); nullptr if specialisation is not possible -- or, with wait = false, not ready yet:
// a new shape then compiles on a background thread and the caller uses the ahead-of-time kernel meanwhile.
const JitKernel* jit_get(int strategy, const ScanParams& p, double* compile_ms, bool wait);
int jit_pending();  // shapes whose background compilation has not finished
int jit_blocks_per_sm(const JitKernel* k, size_t smem_bytes);
bool jit_launch(const JitKernel* k, const ScanParams& p, int grid, size_t smem_bytes, cudaStream_t st);

}  // namespace sdo
