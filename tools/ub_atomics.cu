// Micro-benchmark behind the high-cardinality group-by design: throughput of one count + one double-sum
// update per row into a direct-indexed table whose key window is W bytes wide, keys uniformly random inside the window.
//   variants: red   = two global reductions per row (red.global.add.u64 + red.global.add.f64), what scan_kernel<GLOBAL> issues
//             smem  = block-private shared-memory table (keys pre-partitioned to the block's range): u32 ATOMS + f64 CAS loop
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/ub_atomics.bin tools/ub_atomics.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mix64(uint64_t x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33; return x; }

__global__ void gen_keys(uint32_t* keys, double* vals, uint64_t n, uint64_t window_keys, uint64_t n_windows) {
  // rows [w * n / n_windows, (w + 1) * n / n_windows) fall into key window w (what a partition pass would produce)
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t w = i / ((n + n_windows - 1) / n_windows);
    keys[i] = (uint32_t)(w * window_keys + mix64(i) % window_keys);
    vals[i] = 1.0 + (double)(i & 1023);
  }
}
__global__ void __launch_bounds__(256) red_kernel(const uint32_t* __restrict__ keys, const double* __restrict__ vals, uint64_t n, unsigned long long* table) {
  const uint64_t n4 = n / 4;
  for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < n4; g += (uint64_t)gridDim.x * blockDim.x) {
    const uint4 k = __ldcs((const uint4*)keys + g);
    const double2 a = __ldcs((const double2*)vals + 2 * g), b = __ldcs((const double2*)vals + 2 * g + 1);
    const uint32_t kk[4] = {k.x, k.y, k.z, k.w};
    const double vv[4] = {a.x, a.y, b.x, b.y};
#pragma unroll
    for (int i = 0; i < 4; i++) {
      unsigned long long* q = table + (uint64_t)kk[i] * 2;
      atomicAdd(q, 1ull);
      atomicAdd((double*)(q + 1), vv[i]);
    }
  }
}
// one 128-bit bucket per key: [count u64 | sum f64]; one row = one red.v2? (not available for 64-bit) -> count in u32 array + f64 array
__global__ void __launch_bounds__(256) red_split_kernel(const uint32_t* __restrict__ keys, const double* __restrict__ vals, uint64_t n, unsigned int* cnt, double* sum) {
  const uint64_t n4 = n / 4;
  for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < n4; g += (uint64_t)gridDim.x * blockDim.x) {
    const uint4 k = __ldcs((const uint4*)keys + g);
    const double2 a = __ldcs((const double2*)vals + 2 * g), b = __ldcs((const double2*)vals + 2 * g + 1);
    const uint32_t kk[4] = {k.x, k.y, k.z, k.w};
    const double vv[4] = {a.x, a.y, b.x, b.y};
#pragma unroll
    for (int i = 0; i < 4; i++) {
      atomicAdd(cnt + kk[i], 1u);
      atomicAdd(sum + kk[i], vv[i]);
    }
  }
}
// shared-memory variant: block b owns rows [b * rows_per_block, ...) whose keys all lie in a 8192-key range (ideal second-level partition)
__global__ void __launch_bounds__(256) smem_kernel(const uint32_t* __restrict__ keys, const double* __restrict__ vals, uint64_t n, uint32_t range, unsigned long long* out) {
  extern __shared__ unsigned long long sm[];  // [range] sums then [range] counts (u32)
  double* ssum = (double*)sm;
  unsigned int* scnt = (unsigned int*)(sm + range);
  const uint64_t n_tiles = (n + range - 1) / range;
  for (uint64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    for (uint32_t x = threadIdx.x; x < range; x += 256) { ssum[x] = 0.0; scnt[x] = 0; }
    __syncthreads();
    const uint64_t r0 = t * range, r1 = min(n, r0 + range);
    for (uint64_t i = r0 + threadIdx.x; i < r1; i += 256) {
      const uint32_t k = keys[i] % range;
      atomicAdd(scnt + k, 1u);
      atomicAdd(ssum + k, vals[i]);
    }
    __syncthreads();
    for (uint32_t x = threadIdx.x; x < range; x += 256) { out[(r0 + x) * 2] = scnt[x]; out[(r0 + x) * 2 + 1] = __double_as_longlong(ssum[x]); }
    __syncthreads();
  }
}

int main(int argc, char** argv) {
  const uint64_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 200000000ull;
  uint32_t* keys; double* vals; unsigned long long* table;
  const uint64_t table_keys = 600000000ull;
  CK(cudaMalloc(&keys, n * 4)); CK(cudaMalloc(&vals, n * 8)); CK(cudaMalloc(&table, table_keys * 16));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const uint64_t windows[] = {1ull << 20, 2ull << 20, 4ull << 20, 8ull << 20, 64ull << 20, table_keys};  // keys per window (x16 B)
  for (uint64_t wk : windows) {
    const uint64_t n_windows = table_keys / wk ? table_keys / wk : 1;
    gen_keys<<<sms * 8, 256>>>(keys, vals, n, wk, n_windows);
    CK(cudaMemset(table, 0, table_keys * 16));
    for (int split = 0; split < 2; split++) {
      float best = 1e9;
      for (int it = 0; it < 3; it++) {
        CK(cudaEventRecord(e0));
        if (split) red_split_kernel<<<sms * 8, 256>>>(keys, vals, n, (unsigned int*)table, (double*)(table + table_keys / 2));
        else red_kernel<<<sms * 8, 256>>>(keys, vals, n, table);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
      }
      printf("%s window %6.0f MB (%llu windows): %8.3f ms for %llu rows = %6.1f G rows/s, %6.1f G atomics/s\n", split ? "red_split" : "red      ",
             wk * (split ? 12.0 : 16.0) / 1e6, (unsigned long long)n_windows, best, (unsigned long long)n, n / best / 1e6, 2.0 * n / best / 1e6);
    }
  }
  for (uint32_t range : {4096u, 8192u, 12288u}) {
    CK(cudaFuncSetAttribute(smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    gen_keys<<<sms * 8, 256>>>(keys, vals, n, table_keys, 1);
    float best = 1e9;
    for (int it = 0; it < 3; it++) {
      CK(cudaEventRecord(e0));
      smem_kernel<<<sms * (range <= 4096 ? 3 : range <= 8192 ? 2 : 1), 256, range * 12>>>(keys, vals, n, range, table);
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
    }
    printf("smem range %5u keys: %8.3f ms for %llu rows = %6.1f G rows/s\n", range, best, (unsigned long long)n, n / best / 1e6);
  }
  return 0;
}
