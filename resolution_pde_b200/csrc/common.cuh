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

// Two elements at a time on the packed fp32x2 pipe (FFMA2 / FMUL2 / FADD2, sm_100): the same fit, 11.5 instead of 19
// floating-point instructions per element.  The GEMM epilogues that apply GELU are instruction-issue bound.
typedef unsigned long long f32x2_t;
__device__ __forceinline__ f32x2_t f2_pack(float a, float b) {
  f32x2_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void f2_unpack(f32x2_t r, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(r)); }
__device__ __forceinline__ f32x2_t f2_fma(f32x2_t a, f32x2_t b, f32x2_t c) {
  f32x2_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f32x2_t f2_mul(f32x2_t a, f32x2_t b) {
  f32x2_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2_t f2_add(f32x2_t a, f32x2_t b) {
  f32x2_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2_t f2_c(float c) { return f2_pack(c, c); }
__device__ __forceinline__ void gelu_both2(float x0, float x1, float& h0, float& h1, float& g0, float& g1) {
  float s0, s1, e0, e1, a0, a1, d0, d1;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(s0) : "f"(fmaf(fabsf(x0), 0.35355339059327376f, 1.0f)));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(s1) : "f"(fmaf(fabsf(x1), 0.35355339059327376f, 1.0f)));
  const f32x2_t x = f2_pack(x0, x1);
  f2_unpack(f2_mul(f2_mul(x, x), f2_c(-0.72134752044448170f)), a0, a1);
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(a0));  // exp(-x^2/2)
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(a1));
  const f32x2_t s = f2_pack(s0, s1), e = f2_pack(e0, e1);
  f32x2_t q = f2_c(0.5f * 0.0588021083f);  // the 1/2 of Phi(-|x|) = q(s) s e / 2 rides in the coefficients (exact)
  q = f2_fma(q, s, f2_c(0.5f * -0.218240167f));
  q = f2_fma(q, s, f2_c(0.5f * 0.12218285f));
  q = f2_fma(q, s, f2_c(0.5f * 0.307675597f));
  q = f2_fma(q, s, f2_c(0.5f * 0.130846771f));
  q = f2_fma(q, s, f2_c(0.5f * 0.321697673f));
  q = f2_fma(q, s, f2_c(0.5f * 0.277035179f));
  const f32x2_t tail = f2_mul(f2_mul(q, s), e);                      // Phi(-|x|) <= 1/2
  f2_unpack(f2_fma(tail, f2_c(-1.0f), f2_c(0.5f)), d0, d1);          // 1/2 - Phi(-|x|) >= 0
  const f32x2_t cdf = f2_add(f2_pack(copysignf(d0, x0), copysignf(d1, x1)), f2_c(0.5f));
  f2_unpack(f2_mul(x, cdf), h0, h1);
  f2_unpack(f2_fma(f2_mul(x, e), f2_c(0.39894228040143267794f), cdf), g0, g1);
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
