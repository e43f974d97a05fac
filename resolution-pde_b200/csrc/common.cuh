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

// GELU (exact erf form, custom_layer.py:60 nn.GELU / F.gelu) and its derivative, branch-free, from one ex2 and one rcp:
//   erfc(t) = s*P7(s) * exp(-t^2),  s = 1/(1 + t/2),  t = |x|/sqrt(2)      (minimax fit, 1.1e-8 absolute in exact
//   arithmetic, 2.7e-7 evaluated in fp32 -> 1.4e-7 absolute on Phi);  exp(-t^2) = exp(-x^2/2) is also the pdf factor.
// About 22 instructions per element instead of ~70 for erff + expf: the GEMM epilogues that apply it are issue-bound.
__device__ __forceinline__ void gelu_both(float x, float& h, float& g) {
  const float ax = fabsf(x);
  float s, e;  // MUFU.RCP / MUFU.EX2 (1-2 ulp): single instructions, no slow paths
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(fmaf(ax, 0.35355339059327376f, 1.0f)));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * x * -0.72134752044448170f));  // exp(-x^2/2)
  float q = 0.0588021083f;
  q = fmaf(q, s, -0.218240167f);
  q = fmaf(q, s, 0.12218285f);
  q = fmaf(q, s, 0.307675597f);
  q = fmaf(q, s, 0.130846771f);
  q = fmaf(q, s, 0.321697673f);
  q = fmaf(q, s, 0.277035179f);
  const float tail = 0.5f * q * s * e;               // Phi(-|x|)
  const float cdf = x > 0.f ? 1.0f - tail : tail;
  h = x * cdf;
  g = fmaf(x * 0.39894228040143267794f, e, cdf);
}
__device__ __forceinline__ float gelu_f(float x) {
  float h, g;
  gelu_both(x, h, g);
  return h;
}
__device__ __forceinline__ float dgelu_f(float x) {
  float h, g;
  gelu_both(x, h, g);
  return g;
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
