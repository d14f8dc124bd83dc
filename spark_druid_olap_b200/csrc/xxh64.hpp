// XXH64 as Spark uses it for approx_count_distinct (HyperLogLogPlusPlus hashes every value with
// XxHash64, seed 42: hashInt / hashLong / hashUnsafeBytes), and the HLL++ register mapping for
// relativeSD = 0.05 => p = 9 (512 registers).  Spark 2.0.2 is a "provided" dependency of the
// reference (build.sbt:17,27-29), not vendored; this restates the published XXH64 algorithm, and
// tests pin it to the public XXH64 test vectors.  Call sites in the reference:
//   sources/druid/AggregateTransform.scala:454-479 (approx_count_distinct -> cardinality aggregator).
#pragma once
#ifndef __CUDACC_RTC__
#include <cstddef>
#include <cstdint>
#endif

#if defined(__CUDACC__)
#define SDO_HD __host__ __device__ __forceinline__
#else
#define SDO_HD inline
#endif

namespace sdo {

constexpr uint64_t XXP1 = 0x9E3779B185EBCA87ull;
constexpr uint64_t XXP2 = 0xC2B2AE3D27D4EB4Full;
constexpr uint64_t XXP3 = 0x165667B19E3779F9ull;
constexpr uint64_t XXP4 = 0x85EBCA77C2B2AE63ull;
constexpr uint64_t XXP5 = 0x27D4EB2F165667C5ull;
constexpr uint64_t HLL_SEED = 42;
constexpr int HLL_P = 9;
constexpr int HLL_M = 1 << HLL_P;  // 512 registers, one byte each in our layout

SDO_HD uint64_t xx_rotl(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
SDO_HD uint64_t xx_fmix(uint64_t h) {
  h ^= h >> 33;
  h *= XXP2;
  h ^= h >> 29;
  h *= XXP3;
  h ^= h >> 32;
  return h;
}
// XXH64 of the 4 little-endian bytes of an int (Spark XXH64.hashInt)
SDO_HD uint64_t xxh64_int(int32_t v, uint64_t seed) {
  uint64_t h = seed + XXP5 + 4ull;
  h ^= (uint64_t)(uint32_t)v * XXP1;
  h = xx_rotl(h, 23) * XXP2 + XXP3;
  return xx_fmix(h);
}
// XXH64 of the 8 little-endian bytes of a long (Spark XXH64.hashLong)
SDO_HD uint64_t xxh64_long(int64_t v, uint64_t seed) {
  uint64_t h = seed + XXP5 + 8ull;
  h ^= xx_rotl((uint64_t)v * XXP2, 31) * XXP1;
  h = xx_rotl(h, 27) * XXP1 + XXP4;
  return xx_fmix(h);
}
SDO_HD uint64_t xx_read64(const unsigned char* p) {
  uint64_t v = 0;
  for (int i = 7; i >= 0; i--) v = (v << 8) | p[i];
  return v;
}
SDO_HD uint32_t xx_read32(const unsigned char* p) {
  return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
SDO_HD uint64_t xx_round(uint64_t acc, uint64_t input) {
  acc += input * XXP2;
  acc = xx_rotl(acc, 31);
  return acc * XXP1;
}
SDO_HD uint64_t xx_merge(uint64_t h, uint64_t v) {
  v = xx_round(0, v);
  h ^= v;
  return h * XXP1 + XXP4;
}
// XXH64 of a byte string (Spark XXH64.hashUnsafeBytes on UTF8String bytes)
SDO_HD uint64_t xxh64_bytes(const void* data, size_t len, uint64_t seed) {
  const unsigned char* p = (const unsigned char*)data;
  const unsigned char* end = p + len;
  uint64_t h;
  if (len >= 32) {
    const unsigned char* limit = end - 32;
    uint64_t v1 = seed + XXP1 + XXP2, v2 = seed + XXP2, v3 = seed, v4 = seed - XXP1;
    do {
      v1 = xx_round(v1, xx_read64(p));
      v2 = xx_round(v2, xx_read64(p + 8));
      v3 = xx_round(v3, xx_read64(p + 16));
      v4 = xx_round(v4, xx_read64(p + 24));
      p += 32;
    } while (p <= limit);
    h = xx_rotl(v1, 1) + xx_rotl(v2, 7) + xx_rotl(v3, 12) + xx_rotl(v4, 18);
    h = xx_merge(h, v1);
    h = xx_merge(h, v2);
    h = xx_merge(h, v3);
    h = xx_merge(h, v4);
  } else {
    h = seed + XXP5;
  }
  h += (uint64_t)len;
  while (p + 8 <= end) {
    uint64_t k1 = xx_round(0, xx_read64(p));
    h ^= k1;
    h = xx_rotl(h, 27) * XXP1 + XXP4;
    p += 8;
  }
  if (p + 4 <= end) {
    h ^= (uint64_t)xx_read32(p) * XXP1;
    h = xx_rotl(h, 23) * XXP2 + XXP3;
    p += 4;
  }
  while (p < end) {
    h ^= (uint64_t)(*p) * XXP5;
    h = xx_rotl(h, 11) * XXP1;
    p++;
  }
  return xx_fmix(h);
}

SDO_HD int xx_clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return __clzll((long long)x);
#else
  return x ? __builtin_clzll(x) : 64;
#endif
}
// HLL++ register update of Spark's HyperLogLogPlusPlusHelper.update: idx = top p bits,
// rho = leading zeros of the remaining bits (padded) + 1.  Packed as idx << 6 | rho.
SDO_HD uint16_t hll_pack(uint64_t hash) {
  uint32_t idx = (uint32_t)(hash >> (64 - HLL_P));
  uint64_t w = (hash << HLL_P) | (1ull << (HLL_P - 1));
  uint32_t rho = (uint32_t)xx_clz64(w) + 1;
  return (uint16_t)((idx << 6) | rho);
}

}  // namespace sdo
