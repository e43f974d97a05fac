// Shared device/host helpers for the resolution-pde B200 spectral path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <limits.h>

namespace rpde {

// Two-level index -> element offset:  off(i) = (i / d) * sh + (i % d) * sl.
// A plain stride is {INT_MAX, 0, s}.  This mirrors what a TMA tensor map can address, so
// the tensor-core path can turn the same descriptors into CUtensorMaps.
struct Idx2 {
  int d;
  long long sh, sl;
};
__host__ __device__ __forceinline__ long long off2(const Idx2& a, int i) {
  if (a.d == INT_MAX) return (long long)i * a.sl;
  return (long long)(i / a.d) * a.sh + (long long)(i % a.d) * a.sl;
}
static inline Idx2 lin(long long s) { return Idx2{INT_MAX, 0, s}; }
static inline Idx2 two(int d, long long sh, long long sl) { return Idx2{d, sh, sl}; }

struct MatDesc {
  Idx2 b, r, c;  // batch, row, column index maps
};

enum Act { ACT_NONE = 0, ACT_GELU = 1, ACT_RELU = 2 };

__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float dgelu_f(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * expf(-0.5f * x * x);
  return cdf + x * pdf;
}
// h = gelu(x), g = gelu'(x) from one erf evaluation
__device__ __forceinline__ void gelu_both(float x, float& h, float& g) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * expf(-0.5f * x * x);
  h = x * cdf;
  g = cdf + x * pdf;
}
__device__ __forceinline__ float act_f(int act, float v) {
  if (act == ACT_GELU) return gelu_f(v);
  if (act == ACT_RELU) return v > 0.f ? v : 0.f;
  return v;
}
__device__ __forceinline__ float dact_f(int act, float z) {
  if (act == ACT_GELU) return dgelu_f(z);
  if (act == ACT_RELU) return z > 0.f ? 1.f : 0.f;
  return 1.f;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace rpde
