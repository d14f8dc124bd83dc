// Shared small utilities: status codes (mirrors include/sdo_engine.h), error types, CUDA checks.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <stdexcept>
#include <string>

#include "timeutil.hpp"  // Unsupported / Invalid

namespace sdo {

enum Status {
  OK = 0,
  ERROR = 1,        // generic failure (maps to DruidDataSourceException)
  UNSUPPORTED = 2,  // spec uses something the engine rejects; caller falls back to plain Spark
  INVALID = 3,      // malformed JSON / spec / argument
  NOT_FOUND = 4,    // unknown datasource / column
  CANCELLED = 5,
  NO_DEVICE = 6,    // engine created without a device (planning only) or CUDA unavailable
  OUT_OF_MEMORY = 7,
};

struct NotFound : std::runtime_error {
  explicit NotFound(const std::string& m) : std::runtime_error(m) {}
};
struct NoDevice : std::runtime_error {
  explicit NoDevice(const std::string& m) : std::runtime_error(m) {}
};
struct Cancelled : std::runtime_error {
  explicit Cancelled(const std::string& m) : std::runtime_error(m) {}
};
struct DeviceError : std::runtime_error {
  cudaError_t code;
  DeviceError(cudaError_t c, const std::string& m) : std::runtime_error(m), code(c) {}
};

inline void cuda_check(cudaError_t e, const char* what, const char* file, int line) {
  if (e == cudaSuccess) return;
  std::string m = std::string("CUDA error ") + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ") at " + file + ":" + std::to_string(line) + " in " + what;
  throw DeviceError(e, m);
}
#define SDO_CUDA(expr) ::sdo::cuda_check((expr), #expr, __FILE__, __LINE__)

inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace sdo
