// Fused F-FNO FeedForward layer for sm_100a, BF16 tensor cores (tcgen05.mma kind::f16, fp32 accumulate in TMEM):
//
//     y = res + LayerNorm( W3 gelu( W2 gelu( W1 x + b1 ) + b2 ) + b3 )            models/custom_layer.py:49-68 (FeedForward, 3 layers)
//                                                                                  models/spectral_convolution.py:248-254 (backcast_ff)
//                                                                                  models/ffno.py:228-235 (x = x + layer(x))
//
// for C = 64 channels and a hidden width of 4C = 256 (conf/model/ffno_2d/ffno_2d.yaml).  The 4C-wide hidden activations
// never leave the SM: they live in TMEM as the accumulator of one layer and, converted to BF16 in place, as the A
// operand of the next one (tcgen05.mma with A from TMEM).  HBM traffic of the forward pass is read x + read res + write y.
//
// One persistent CTA per SM; the three weight matrices (192 KB as BF16) are loaded once per CTA into shared memory in the
// K-major SWIZZLE_128B image the UMMA descriptors expect (built once per weight update by ff_pack_weights_kernel), and the
// same image serves the backward pass as the MN-major (transposed) operand.  Tile = 128 rows (TMEM lanes).
//
//   warp 0        UMMA issuer (one lane) + TMEM owner
//   warp 1        weight image: one cp.async.bulk sequence at kernel start
//   warps 4-7     x loader: fp32 rows -> BF16 -> swizzled shared-memory A tile (and, backward, the BF16 copy the
//                 weight-gradient kernel reads)
//   warps 8-23    epilogues, four warps per TMEM lane quadrant (16-column chunks c = 4j + sub): the epilogues are
//                 latency- and issue-bound CUDA-core work, four warps per scheduler hide each other's waits
//
// TMEM (512 columns):   [0,256) D1 = x W1^T  -> gelu -> H1 (BF16 pairs, [0,128))           [256,512) D2 = H1 W2^T -> H2 ([256,384))
//                       [128,192) D3 = H2 W3^T (read by the LayerNorm epilogue)
//   backward (same kernel, BWD = true; the forward part is recomputed, nothing but x is saved by the forward pass):
//                       [192,224) dZ3 (BF16)   [256,512) dH2 = dZ3 W3 -> * gelu'(Z2) -> dZ2 ([256,384))
//                       [0,256) dH1 = dZ2 W2 -> * gelu'(Z1) -> dZ1 ([0,128))              [256,320) dS = dZ1 W1
//   gelu'(Z1), gelu'(Z2) wait in a per-CTA scratch (L2-resident, written and re-read by the same thread); H1, H2, dZ1,
//   dZ2, dZ3 and x leave as BF16 for the weight-gradient kernel (ff_wgrad_kernel below: dW = dZ^T H with K = rows).
//
// In-place conversion: the result of 16-column chunk c (8 packed columns) lands in columns that belong to chunk c/2.
// The four warps of a quadrant own chunks 4j + sub and meet at a 128-thread named barrier after every chunk load, so
// chunk c/2 <= 2j + 1 has always been read by its owner before it is overwritten.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/rpde_b200.h"
#include "internal.h"
#include "common.cuh"
#include "tc_ptx.cuh"

namespace rpde {
namespace ff {
using namespace tc;

constexpr int C = 64, H = 256, TM = 128;
constexpr uint32_t W1_OFF = 0, W1_BYTES = H * 128;                                    // [256 rows][64 bf16]
constexpr uint32_t W2_OFF = W1_OFF + W1_BYTES, W2_PLANE = H * 128, W2_BYTES = 4 * W2_PLANE;  // 4 k-planes of [256 rows][64 bf16]
constexpr uint32_t W3_OFF = W2_OFF + W2_BYTES, W3_PLANE = C * 128, W3_BYTES = 4 * W3_PLANE;  // 4 k-planes of [64 rows][64 bf16]
constexpr uint32_t WBLOB = W3_OFF + W3_BYTES;                                          // 196608
constexpr uint32_t X_OFF = WBLOB, X_BYTES = TM * 128;                                  // A tile of layer 1: [128 rows][64 bf16]
constexpr int kEpiWarps = 16;                                                          // four per TMEM lane quadrant
constexpr uint32_t STG_OFF = X_OFF + X_BYTES;                                          // one 1 KB staging box per epilogue warp (TMA stores)
constexpr uint32_t STG_BYTES = kEpiWarps * 1024;
constexpr uint32_t PAR_OFF = STG_OFF + STG_BYTES;                                      // b1[256] b2[256] b3[64] gamma[64] beta[64]
constexpr uint32_t PAR_BYTES = (2 * H + 3 * C) * 4;
constexpr uint32_t BAR_OFF = PAR_OFF + PAR_BYTES;
constexpr uint32_t SMEM_BYTES = BAR_OFF + 256;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
constexpr int kThreads = (8 + kEpiWarps) * 32;   // warp 0 UMMA, 1 weights, 2-3 idle, 4-7 x loaders, 8-23 epilogues
// TMEM columns
constexpr uint32_t D1 = 0, H1 = 0, D3 = 128, DZ3 = 192, D2 = 256, H2 = 256, D4 = 256, DZ2 = 256, D5 = 0, DZ1 = 0, D6 = 256;
constexpr uint32_t ST_A = 224, ST_B = 232, ST_C = 240;   // row-statistic exchange between the four warps of a quadrant
// barriers
enum { B_WFULL = 0, B_XFULL, B_XEMPTY, B_D1, B_D2, B_D3, B_TILE, B_H1, B_H2 = B_H1 + 4, B_DZ3 = B_H2 + 4, B_D4, B_DZ2,
       B_D5 = B_DZ2 + 4, B_DZ1, B_D6 = B_DZ1 + 4, B_COUNT };
static_assert(B_COUNT * 8 + 8 <= 256, "barrier area");

struct Params {
  const float* x;       // [T][64] FeedForward input (spectral output)
  const float* res;     // [T][64] residual or NULL (forward)
  const uint8_t* blob;  // packed BF16 weight image (WBLOB bytes)
  const float *b1, *b2, *b3, *gamma, *beta;  // gamma == NULL: no LayerNorm
  long long T;
  float eps;
  const float* dy;      // backward: [T][64]
  uint16_t* xb;         // backward: BF16 copy of x for the weight-gradient kernel
  uint8_t* gscratch;    // backward: gridDim.x * 128 KB (8 KB per epilogue warp)
  float *dgamma, *dbeta;   // zero-initialised by the caller, accumulated atomically
};
// TMA maps of everything the epilogue warps store: boxes of 32 rows x 32 bytes, SWIZZLE_32B
struct Maps {
  CUtensorMap out;                     // forward: y; backward: ds   ([T][64] fp32, box 8 x 32)
  CUtensorMap h1, h2, dz1, dz2, dz3;   // backward: BF16 copies ([T][256] / [T][64], box 16 x 32)
};

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

__host__ __device__ constexpr uint32_t idesc_bf16(int M, int N, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_ts(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a_tmem), "l"(b), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
      "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st2(uint32_t taddr, float a, float b) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(taddr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void tmem_ld2(uint32_t taddr, float& a, float& b) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=f"(a), "=f"(b) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void pair_sync(int quad) { asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory"); }
__device__ __forceinline__ void stg128(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 ldg128(const void* p) {
  uint4 v;
  asm volatile("ld.global.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}

// GELU and its derivative for the BF16 variant, two elements at a time on the packed fp32x2 pipe, ONE special-function
// instruction per element:  Phi(x) = 1/2 + 1/2 tanh(x (a0 + a1 x^2 + a2 x^4)),  the odd polynomial fitted to
// atanh(erf(x / sqrt 2)) on [0, 6]: |gelu error| < 3e-5, |gelu' error| < 1.3e-4 in exact arithmetic, plus tanh.approx
// (2^-11 relative) -- an order of magnitude below the BF16 rounding (2^-9) applied to the result right after.
//   gelu(x) = x Phi(x),   gelu'(x) = Phi + x/2 (1 - t^2)(a0 + 3 a1 x^2 + 5 a2 x^4).
// The exact erf form (common.cuh: gelu_both2, two MUFU + 11.5 FP instructions per element) stays with the fp32-parity
// and TF32 engines; this kernel's epilogues are instruction-issue bound.
template <bool WITH_G>
__device__ __forceinline__ void gelu_tanh2(float x0, float x1, float& h0, float& h1, float& g0, float& g1) {
  constexpr float A0 = 7.97458471e-01f, A1 = 3.70503451e-02f, A2 = -3.58732361e-04f;
  const f32x2_t x = f2_pack(x0, x1);
  float u0, u1;
  f2_unpack(f2_mul(x, x), u0, u1);
  // beyond |x| = 8 the fitted polynomial would turn over: freeze its argument there (q(64) = 1.70 > 0, so x q stays
  // monotone and tanh saturates; gelu is x or 0 to 1e-15 out there)
  const f32x2_t u = f2_pack(fminf(u0, 64.f), fminf(u1, 64.f));
  f32x2_t q = f2_fma(u, f2_c(A2), f2_c(A1));
  q = f2_fma(q, u, f2_c(A0));
  float p0, p1, t0, t1;
  f2_unpack(f2_mul(q, x), p0, p1);
  asm("tanh.approx.f32 %0, %1;" : "=f"(t0) : "f"(p0));
  asm("tanh.approx.f32 %0, %1;" : "=f"(t1) : "f"(p1));
  const f32x2_t t = f2_pack(t0, t1), cdf = f2_fma(t, f2_c(0.5f), f2_c(0.5f));
  f2_unpack(f2_mul(x, cdf), h0, h1);
  if (WITH_G) {
    f32x2_t dq = f2_fma(u, f2_c(5.f * A2), f2_c(3.f * A1));
    dq = f2_fma(dq, u, f2_c(A0));
    const f32x2_t s = f2_fma(t, f2_mul(t, f2_c(-1.f)), f2_c(1.f));      // 1 - t^2
    f2_unpack(f2_fma(f2_mul(f2_mul(x, dq), s), f2_c(0.5f), cdf), g0, g1);
  }
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t* v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(v[0]), "r"(v[1]),
               "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}
// the four epilogue warps of a TMEM lane quadrant
__device__ __forceinline__ void quad_sync(int quad) { asm volatile("bar.sync %0, 128;" ::"r"(1 + quad) : "memory"); }

// The four warps of a quadrant share per-row partial sums through eight spare TMEM columns (every warp of the quadrant
// may access its lanes): write mine, barrier, read all four pairs and add them in a fixed order.
__device__ __forceinline__ void quad_exchange(uint32_t tlane, uint32_t col, int sub, int quad, float a, float b, float& sa, float& sb) {
  tmem_st2(tlane + col + 2 * sub, a, b);
  tmem_st_wait();
  tc_fence_before_sync();
  quad_sync(quad);
  tc_fence_after_sync();
  uint32_t r[8];
  tmem_ld8(tlane + col, r);
  tmem_ld_wait();
  sa = (__uint_as_float(r[0]) + __uint_as_float(r[2])) + (__uint_as_float(r[4]) + __uint_as_float(r[6]));
  sb = (__uint_as_float(r[1]) + __uint_as_float(r[3])) + (__uint_as_float(r[5]) + __uint_as_float(r[7]));
}

// per-warp staging box for bulk tensor stores: 32 rows (lanes) of 32 bytes, 16-byte chunk q of row r at q ^ ((r >> 2) & 1)
// (SWIZZLE_32B: the warp's two 128-bit stores are bank-conflict free)
struct Stage {
  uint32_t base;
  int lane;
  __device__ __forceinline__ void wait_free() const {   // the previous bulk store has finished reading the box
    if (lane == 0) bulk_wait_read0();
    __syncwarp();
  }
  __device__ __forceinline__ void write(const uint32_t* p) const {   // 8 words = this lane's 32-byte row
    const uint32_t row = base + (uint32_t)lane * 32u, sw = ((uint32_t)lane >> 2) & 1u;
#pragma unroll
    for (uint32_t q = 0; q < 2; ++q)
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((q ^ sw) << 4)), "r"(p[4 * q]), "r"(p[4 * q + 1]),
                   "r"(p[4 * q + 2]), "r"(p[4 * q + 3])
                   : "memory");
  }
  __device__ __forceinline__ void store(const CUtensorMap* m, int col, int row) const {
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      const int c[2] = {col, row};
      tma_store(base, m, 2, c, false);
      bulk_commit();
    }
  }
};

// lanes l and l + 16 receive the sum over the warp's 32 lanes of v[l & 15] (16 shuffles): the column sums of a 32 x 16
// block whose rows are the lanes.  v is destroyed.
__device__ __forceinline__ float colsum16(float (&v)[16], int lane) {
#pragma unroll
  for (int off = 8; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int k = 0; k < off; ++k) {
      const float send = up ? v[k] : v[k + off];
      const float keep = up ? v[k + off] : v[k];
      v[k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 16);
}

// One hidden-layer epilogue over this warp's four 16-column chunks (c = 4j + sub), converted in place.
//   MODE 0: h = gelu(acc + bias)                                         -> BF16 A operand                    (inference / forward)
//   MODE 1: the same, plus gelu'(acc + bias) -> scratch and h -> spill                                        (backward: recompute)
//   MODE 2: dz = acc * gelu' (from scratch) -> BF16 A operand, dz -> spill                                     (backward)
// scratch: this warp's private 4 KB for the layer, laid out [j][q][lane] in 16-byte units (coalesced, same thread reads back)
template <int MODE>
__device__ __forceinline__ void hidden_epi(uint32_t tlane, uint32_t src_col, uint32_t dst_col, int sub, int quad, int lane,
                                           uint32_t ready_bar0, uint32_t bias_smem, const Stage& stg, const CUtensorMap* spill_map,
                                           int row0, uint8_t* scratch) {
  uint4 g4[2];
  if (MODE == 2) {
#pragma unroll
    for (int q = 0; q < 2; ++q) g4[q] = ldg128(scratch + ((0 * 2 + q) * 32 + lane) * 16);
  }
#pragma unroll 1
  for (int j = 0; j < 4; ++j) {
    const int c = 4 * j + sub;
    uint32_t r[16];
    tmem_ld16(tlane + src_col + 16 * c, r);
    uint4 gn[2];
    if (MODE == 2 && j < 3) {   // the next chunk's gelu' is in flight while this one is processed
#pragma unroll
      for (int q = 0; q < 2; ++q) gn[q] = ldg128(scratch + (((j + 1) * 2 + q) * 32 + lane) * 16);
    }
    float v[16];
    if (MODE != 2) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 b4 = lds_f4(bias_smem + (16 * c + 4 * q) * 4);
        v[4 * q] = b4.x; v[4 * q + 1] = b4.y; v[4 * q + 2] = b4.z; v[4 * q + 3] = b4.w;
      }
    }
    tmem_ld_wait();
    tc_fence_before_sync();
    quad_sync(quad);   // the four warps of the quadrant have read chunks 4j .. 4j + 3: columns of chunks <= 2j + 1 may be overwritten
    tc_fence_after_sync();
    uint32_t p[8];
    if (MODE != 2) {
      uint32_t gp[8];
#pragma unroll
      for (int k = 0; k < 16; k += 2) {
        float h0, h1, g0, g1;
        float z0, z1;   // accumulator + bias on the packed pipe
        f2_unpack(f2_add(f2_pack(__uint_as_float(r[k]), __uint_as_float(r[k + 1])), f2_pack(v[k], v[k + 1])), z0, z1);
        gelu_tanh2<MODE == 1>(z0, z1, h0, h1, g0, g1);
        p[k >> 1] = pack_bf16(h0, h1);
        if (MODE == 1) gp[k >> 1] = pack_bf16(g0, g1);
      }
      tmem_st8(tlane + dst_col + 8 * c, p);
      if (MODE == 1) {
#pragma unroll
        for (int q = 0; q < 2; ++q) stg128(scratch + ((j * 2 + q) * 32 + lane) * 16, gp[4 * q], gp[4 * q + 1], gp[4 * q + 2], gp[4 * q + 3]);
      }
    } else {
      const uint32_t* gw = reinterpret_cast<const uint32_t*>(g4);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        p[k] = pack_bf16(__uint_as_float(r[2 * k]) * bf16_lo(gw[k]), __uint_as_float(r[2 * k + 1]) * bf16_hi(gw[k]));
      tmem_st8(tlane + dst_col + 8 * c, p);
    }
    if (MODE != 0) {   // BF16 copy for the weight-gradient kernel: staged, then one bulk tensor store per chunk
      stg.wait_free();
      stg.write(p);
      stg.store(spill_map, 16 * c, row0);
    }
    if (MODE == 2 && j < 3) {
      g4[0] = gn[0];
      g4[1] = gn[1];
    }
    tmem_st_wait();
    tc_fence_before_sync();
    __syncwarp();
    if (lane == 0) mbar_arrive(ready_bar0 + 8 * j);
  }
}

template <bool BWD>
__global__ void __launch_bounds__(kThreads, 1) ff3_kernel(const __grid_constant__ Params P, const __grid_constant__ Maps M) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t sbase = smem_u32(smem_raw);
  const uint32_t w1s = sbase + W1_OFF, w2s = sbase + W2_OFF, w3s = sbase + W3_OFF, xs = sbase + X_OFF;
  const uint32_t par = sbase + PAR_OFF;
  const uint32_t pb1 = par, pb2 = par + H * 4, pb3 = par + 2 * H * 4, pgam = pb3 + C * 4, pbet = pgam + C * 4;
  const uint32_t bars = sbase + BAR_OFF;
  auto bar = [&](int i) { return bars + 8u * (uint32_t)i; };
  const uint32_t tmem_slot = bars + 8u * B_COUNT;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long ntiles = (P.T + TM - 1) / TM;
  const bool ln = P.gamma != nullptr;

  {   // biases / LayerNorm parameters -> shared memory
    float* pf = reinterpret_cast<float*>(smem_raw + PAR_OFF);
    for (int i = threadIdx.x; i < 2 * H + 3 * C; i += kThreads) {
      float v;
      if (i < H) v = P.b1[i];
      else if (i < 2 * H) v = P.b2[i - H];
      else if (i < 2 * H + C) v = P.b3[i - 2 * H];
      else if (i < 2 * H + 2 * C) v = ln ? P.gamma[i - 2 * H - C] : 1.f;
      else v = (ln && P.beta) ? P.beta[i - 2 * H - 2 * C] : 0.f;
      pf[i] = v;
    }
  }
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&M.out);
    mbar_init(bar(B_WFULL), 1);
    mbar_init(bar(B_XFULL), 4);
    mbar_init(bar(B_XEMPTY), 1);
    mbar_init(bar(B_D1), 1);
    mbar_init(bar(B_D2), 1);
    mbar_init(bar(B_D3), 1);
    mbar_init(bar(B_TILE), kEpiWarps);
    for (int j = 0; j < 4; ++j) {
      mbar_init(bar(B_H1 + j), kEpiWarps);
      mbar_init(bar(B_H2 + j), kEpiWarps);
      mbar_init(bar(B_DZ2 + j), kEpiWarps);
      mbar_init(bar(B_DZ1 + j), kEpiWarps);
    }
    mbar_init(bar(B_DZ3), kEpiWarps);
    mbar_init(bar(B_D4), 1);
    mbar_init(bar(B_D5), 1);
    mbar_init(bar(B_D6), 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  uint32_t tmem;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(tmem_slot));

  if (warp == 0) {
    // ------------------------------------------------------------------------------------------------ UMMA issuer
    if (lane == 0) {
      constexpr uint32_t I256 = idesc_bf16(128, 256, 0), I64 = idesc_bf16(128, 64, 0);
      constexpr uint32_t I256T = idesc_bf16(128, 256, 1), I64T = idesc_bf16(128, 64, 1);
      const uint64_t dK = umma_desc(0, 16, 1024, UMMA_LAYOUT_SW128);                  // K-major, 8-row groups 1024 B apart
      const uint64_t dT3 = umma_desc(0, W3_PLANE, 1024, UMMA_LAYOUT_SW128);           // MN-major views of the same images:
      const uint64_t dT2 = umma_desc(0, W2_PLANE, 1024, UMMA_LAYOUT_SW128);           //   64-column groups one plane apart,
      const uint64_t dT1 = umma_desc(0, W1_BYTES, 1024, UMMA_LAYOUT_SW128);           //   8-k-row groups 1024 B apart
      mbar_wait(bar(B_WFULL), 0);
      uint32_t it = 0;
      for (long long t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
        const uint32_t ph = it & 1;
        mbar_wait(bar(B_XFULL), ph);
        if (it > 0) mbar_wait(bar(B_TILE), ph ^ 1);   // the last epilogue of the previous tile has left TMEM
        tc_fence_after_sync();
#pragma unroll
        for (int s = 0; s < 4; ++s)
          umma_bf16_ss(tmem + D1, dK + ((xs + 32u * s) >> 4), dK + ((w1s + 32u * s) >> 4), I256, s > 0);
        umma_commit(bar(B_XEMPTY));
        umma_commit(bar(B_D1));
        for (int j = 0; j < 4; ++j) {
          mbar_wait(bar(B_H1 + j), ph);
          tc_fence_after_sync();
#pragma unroll
          for (int s = 0; s < 4; ++s)
            umma_bf16_ts(tmem + D2, tmem + H1 + 32 * j + 8 * s, dK + ((w2s + j * W2_PLANE + 32u * s) >> 4), I256, (j | s) > 0);
        }
        umma_commit(bar(B_D2));
        for (int j = 0; j < 4; ++j) {
          mbar_wait(bar(B_H2 + j), ph);
          tc_fence_after_sync();
#pragma unroll
          for (int s = 0; s < 4; ++s)
            umma_bf16_ts(tmem + D3, tmem + H2 + 32 * j + 8 * s, dK + ((w3s + j * W3_PLANE + 32u * s) >> 4), I64, (j | s) > 0);
        }
        umma_commit(bar(B_D3));
        if (BWD) {
          mbar_wait(bar(B_DZ3), ph);
          tc_fence_after_sync();
#pragma unroll
          for (int s = 0; s < 4; ++s)   // dH2 = dZ3 W3: K = 64 output features of layer 3 (rows of the W3 image)
            umma_bf16_ts(tmem + D4, tmem + DZ3 + 8 * s, dT3 + ((w3s + 2048u * s) >> 4), I256T, s > 0);
          umma_commit(bar(B_D4));
          for (int j = 0; j < 4; ++j) {   // dH1 = dZ2 W2: K = 256 output features of layer 2
            mbar_wait(bar(B_DZ2 + j), ph);
            tc_fence_after_sync();
#pragma unroll
            for (int s = 0; s < 4; ++s)
              umma_bf16_ts(tmem + D5, tmem + DZ2 + 32 * j + 8 * s, dT2 + ((w2s + 8192u * j + 2048u * s) >> 4), I256T, (j | s) > 0);
          }
          umma_commit(bar(B_D5));
          for (int j = 0; j < 4; ++j) {   // dS = dZ1 W1
            mbar_wait(bar(B_DZ1 + j), ph);
            tc_fence_after_sync();
#pragma unroll
            for (int s = 0; s < 4; ++s)
              umma_bf16_ts(tmem + D6, tmem + DZ1 + 32 * j + 8 * s, dT1 + ((w1s + 8192u * j + 2048u * s) >> 4), I64T, (j | s) > 0);
          }
          umma_commit(bar(B_D6));
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------------------------------------ weight image
    if (lane == 0) {
      mbar_arrive_expect_tx(bar(B_WFULL), WBLOB);
      for (uint32_t o = 0; o < WBLOB; o += 32768) bulk_g2s(sbase + o, P.blob + o, 32768, bar(B_WFULL));
    }
  } else if (warp >= 4 && warp < 8) {
    // ------------------------------------------------------------------------------------------------ x loader
    const int lw = warp - 4, hl = lane >> 4, q = lane & 15;
    uint32_t it = 0;
    for (long long t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
#pragma unroll 1
      for (int pass = 0; pass < 2; ++pass) {
        float4 v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const long long row = t * TM + lw * 32 + 16 * pass + 2 * i + hl;
          v[i] = row < P.T ? __ldg(reinterpret_cast<const float4*>(P.x + row * C) + q) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (pass == 0) mbar_wait(bar(B_XEMPTY), (it & 1) ^ 1);   // layer 1 of the previous tile has consumed the buffer
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = lw * 32 + 16 * pass + 2 * i + hl;
          const uint32_t lo = pack_bf16(v[i].x, v[i].y), hi = pack_bf16(v[i].z, v[i].w);
          const uint32_t a = xs + r * 128 + ((((uint32_t)q >> 1) ^ ((uint32_t)r & 7u)) << 4) + ((uint32_t)q & 1u) * 8u;
          asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(a), "r"(lo), "r"(hi) : "memory");
          if (BWD) {
            const long long row = t * TM + r;
            if (row < P.T) asm volatile("st.global.v2.b32 [%0], {%1, %2};" ::"l"(P.xb + row * C + 4 * q), "r"(lo), "r"(hi) : "memory");
          }
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(B_XFULL));
    }
  } else if (warp >= 8) {
    // ------------------------------------------------------------------------------------------------ epilogues
    const int ew = warp - 8, quad = warp & 3, sub = ew >> 2;
    const uint32_t tlane = tmem + ((uint32_t)(quad * 32) << 16);
    const int rl = quad * 32 + lane;                       // row of the tile = TMEM lane
    const Stage stg{sbase + STG_OFF + (uint32_t)ew * 1024u, lane};
    float csg = 0.f, csb = 0.f;   // dgamma / dbeta column sums (the bias gradients come from the weight-gradient kernel)
    uint8_t* scr = BWD ? P.gscratch + ((size_t)blockIdx.x * kEpiWarps + ew) * 8192 : nullptr;   // [layer][4 KB] per warp
    uint32_t it = 0;
    for (long long t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
      const uint32_t ph = it & 1;
      const long long row = t * TM + rl;
      const int row0 = (int)(t * TM) + quad * 32;
      const bool valid = row < P.T;
      // ---- layer 1 and 2: bias + GELU, converted in place into the next layer's A operand
      mbar_wait(bar(B_D1), ph);
      tc_fence_after_sync();
      hidden_epi<BWD ? 1 : 0>(tlane, D1, H1, sub, quad, lane, bar(B_H1), pb1, stg, &M.h1, row0, scr);
      mbar_wait(bar(B_D2), ph);
      tc_fence_after_sync();
      hidden_epi<BWD ? 1 : 0>(tlane, D2, H2, sub, quad, lane, bar(B_H2), pb2, stg, &M.h2, row0, BWD ? scr + 4096 : nullptr);
      // ---- layer 3: bias, LayerNorm (row statistics shared by the four warps of the quadrant), 16 columns per warp
      float in[16];   // forward: residual; backward: dy
      {
        const float* src = BWD ? P.dy : P.res;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 a = (src && valid) ? __ldg(reinterpret_cast<const float4*>(src + row * C + 16 * sub) + q) : make_float4(0.f, 0.f, 0.f, 0.f);
          in[4 * q] = a.x; in[4 * q + 1] = a.y; in[4 * q + 2] = a.z; in[4 * q + 3] = a.w;
        }
      }
      mbar_wait(bar(B_D3), ph);
      tc_fence_after_sync();
      float v[16];
      {
        uint32_t r[16];
        tmem_ld16(tlane + D3 + 16 * sub, r);
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 b4 = lds_f4(pb3 + (16 * sub + 4 * q) * 4);
          v[4 * q] = __uint_as_float(r[4 * q]) + b4.x;
          v[4 * q + 1] = __uint_as_float(r[4 * q + 1]) + b4.y;
          v[4 * q + 2] = __uint_as_float(r[4 * q + 2]) + b4.z;
          v[4 * q + 3] = __uint_as_float(r[4 * q + 3]) + b4.w;
        }
      }
      float mean = 0.f, rstd = 1.f;
      if (ln) {
        float s = 0.f, tot, unused;
#pragma unroll
        for (int k = 0; k < 16; ++k) s += v[k];
        quad_exchange(tlane, ST_A, sub, quad, s, 0.f, tot, unused);
        mean = tot * (1.0f / C);
        float qv = 0.f;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const float d = v[k] - mean;
          qv = fmaf(d, d, qv);
        }
        quad_exchange(tlane, ST_B, sub, quad, qv, 0.f, tot, unused);
        rstd = 1.0f / sqrtf(tot * (1.0f / C) + P.eps);
      }
      if (!BWD) {
        // every TMEM access of this tile is done: the next tile may start
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(B_TILE));
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 g4 = lds_f4(pgam + (16 * sub + 4 * q) * 4), b4 = lds_f4(pbet + (16 * sub + 4 * q) * 4);
          if (ln) {
            v[4 * q] = (v[4 * q] - mean) * rstd * g4.x + b4.x;
            v[4 * q + 1] = (v[4 * q + 1] - mean) * rstd * g4.y + b4.y;
            v[4 * q + 2] = (v[4 * q + 2] - mean) * rstd * g4.z + b4.z;
            v[4 * q + 3] = (v[4 * q + 3] - mean) * rstd * g4.w + b4.w;
          }
          v[4 * q] += in[4 * q]; v[4 * q + 1] += in[4 * q + 1]; v[4 * q + 2] += in[4 * q + 2]; v[4 * q + 3] += in[4 * q + 3];
        }
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {   // two boxes of 8 fp32 columns
          stg.wait_free();
          stg.write(reinterpret_cast<const uint32_t*>(v) + 8 * hh);
          stg.store(&M.out, 16 * sub + 8 * hh, row0);
        }
      } else {
        // ---- LayerNorm backward: dz3 = rstd (g - mean(g) - xhat mean(g xhat)), g = dy gamma
        float dz[16];
        if (ln) {
          float s1 = 0.f, s2 = 0.f, t1, t2;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 g4 = lds_f4(pgam + (16 * sub + 4 * q) * 4);
            const float gg[4] = {g4.x, g4.y, g4.z, g4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int k = 4 * q + e;
              const float xh = (v[k] - mean) * rstd;
              const float g = in[k] * gg[e];
              s1 += g;
              s2 = fmaf(g, xh, s2);
              dz[k] = g;
              v[k] = xh;
            }
          }
          quad_exchange(tlane, ST_C, sub, quad, s1, s2, t1, t2);
          s1 = t1 * (1.0f / C);
          s2 = t2 * (1.0f / C);
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            dz[k] = rstd * (dz[k] - s1 - v[k] * s2);
            v[k] *= in[k];           // dy * xhat: dgamma
          }
          csg += colsum16(v, lane);
        } else {
#pragma unroll
          for (int k = 0; k < 16; ++k) dz[k] = in[k];
        }
        uint32_t p[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) p[k] = pack_bf16(dz[2 * k], dz[2 * k + 1]);
        tmem_st8(tlane + DZ3 + 8 * sub, p);
        stg.wait_free();
        stg.write(p);
        stg.store(&M.dz3, 16 * sub, row0);
        if (ln) csb += colsum16(in, lane);
        tmem_st_wait();
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(B_DZ3));
        // ---- dZ2 = dH2 * gelu'(Z2), dZ1 = dH1 * gelu'(Z1)
        mbar_wait(bar(B_D4), ph);
        tc_fence_after_sync();
        hidden_epi<2>(tlane, D4, DZ2, sub, quad, lane, bar(B_DZ2), 0, stg, &M.dz2, row0, scr + 4096);
        mbar_wait(bar(B_D5), ph);
        tc_fence_after_sync();
        hidden_epi<2>(tlane, D5, DZ1, sub, quad, lane, bar(B_DZ1), 0, stg, &M.dz1, row0, scr);
        // ---- dS
        mbar_wait(bar(B_D6), ph);
        tc_fence_after_sync();
        uint32_t r[16];
        tmem_ld16(tlane + D6 + 16 * sub, r);
        tmem_ld_wait();
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(B_TILE));
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          stg.wait_free();
          stg.write(r + 8 * hh);
          stg.store(&M.out, 16 * sub + 8 * hh, row0);
        }
      }
    }
    if (BWD && ln && lane < 16) {
      atomicAdd(P.dgamma + 16 * sub + lane, csg);
      atomicAdd(P.dbeta + 16 * sub + lane, csb);
    }
    if (lane == 0) bulk_wait0();   // the staging box must outlive the bulk stores that read it
    __syncwarp();
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after_sync();
    tmem_dealloc(tmem, 512);
  }
}

// fp32 weights (out, in) row-major -> the BF16 shared-memory image: K-major rows of 128 bytes (64 input features), the
// 16-byte chunk c of row r stored at chunk c ^ (r & 7) (SWIZZLE_128B), one 1024-byte-aligned plane per 64 input features
__global__ void ff_pack_weights_kernel(const float* __restrict__ w1, const float* __restrict__ w2, const float* __restrict__ w3,
                                       uint8_t* __restrict__ blob) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;   // one 16-byte chunk (8 elements) per thread
  if (i >= (int)(WBLOB / 16)) return;
  const uint32_t off = (uint32_t)i * 16u;
  const float* w;
  int rows, in, local;
  if (off < W2_OFF) { w = w1; rows = H; in = C; local = off; }
  else if (off < W3_OFF) { w = w2; rows = H; in = H; local = off - W2_OFF; }
  else { w = w3; rows = C; in = H; local = off - W3_OFF; }
  const int plane = local / (rows * 128), rem = local % (rows * 128);
  const int r = rem / 128, cpos = (rem % 128) / 16;
  const int c = cpos ^ (r & 7);
  const float* src = w + (size_t)r * in + plane * 64 + c * 8;
  uint4 o;
  o.x = pack_bf16(src[0], src[1]);
  o.y = pack_bf16(src[2], src[3]);
  o.z = pack_bf16(src[4], src[5]);
  o.w = pack_bf16(src[6], src[7]);
  *reinterpret_cast<uint4*>(blob + off) = o;
}

// ====================================================================================================================
// Weight gradients of the fused layer: dW2 = dZ2^T H1 (job 0) and dW1 = dZ1^T Xb, dW3^T = H2^T dZ3 (job 1), K = rows.
// Both operands are BF16 row-major [T][features] in global memory = MN-major operands: one TMA box of 64 rows x 64
// features lands as a SWIZZLE_128B atom stack ([k rows][128 B]).  Each CTA owns a contiguous slab of 64-row k-blocks,
// accumulates in TMEM (fp32) and adds its partial result to the zero-initialised gradient with red.global.add.
//   warp 0: TMA producer, warp 1: UMMA issuer + TMEM owner, warps 2-5: column sums of the dZ tiles while they sit in
//   shared memory (the bias gradients), then the final TMEM -> global reduction.
// ====================================================================================================================
constexpr int WG_THREADS = 6 * 32;
constexpr uint32_t WG_BOX = 64 * 128;   // one 64 x 64 BF16 box

template <int JOB>
struct WgCfg {
  // job 0: A = dZ2 (4 boxes), B = H1 (4 boxes)               -> D[2 x 128 lanes][256 cols] = 512 TMEM columns
  // job 1: A0 = dZ1 (4 boxes), B0 = Xb (1 box) -> cols [0,128);  A1 = H2 (4 boxes), B1 = dZ3 (1 box) -> cols [128,256)
  static constexpr int boxes = JOB == 0 ? 8 : 10;
  static constexpr uint32_t stage_bytes = boxes * WG_BOX;
  static constexpr int stages = JOB == 0 ? 3 : 2;
  static constexpr uint32_t smem_bytes = stages * stage_bytes + 1024 + 256;
  static constexpr int tmem_cols = JOB == 0 ? 512 : 256;
};

struct WgParams {
  long long T;
  float *dw1, *dw2, *dw3;   // (256,64), (256,256), (64,256) fp32, zero-initialised
  float *db1, *db2, *db3;   // bias gradients = column sums of dZ1 / dZ2 / dZ3, zero-initialised
};

template <int JOB>
__global__ void __launch_bounds__(WG_THREADS, 1)
ff_wgrad_kernel(const __grid_constant__ WgParams P, const __grid_constant__ CUtensorMap mA0, const __grid_constant__ CUtensorMap mB0,
                const __grid_constant__ CUtensorMap mA1, const __grid_constant__ CUtensorMap mB1) {
  using G = WgCfg<JOB>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = sbase + G::stages * G::stage_bytes;
  auto full = [&](int s) { return bars + 8u * s; };
  auto empty = [&](int s) { return bars + 8u * (G::stages + s); };
  const uint32_t done = bars + 8u * (2 * G::stages), tmem_slot = done + 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long kblocks = (P.T + 63) / 64;
  const long long per = (kblocks + gridDim.x - 1) / gridDim.x;
  const long long kb0 = (long long)blockIdx.x * per, kb1 = kb0 + per < kblocks ? kb0 + per : kblocks;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mA0);
    tma_prefetch_desc(&mB0);
    for (int s = 0; s < G::stages; ++s) {
      mbar_init(full(s), 1);
      mbar_init(empty(s), 5);   // the UMMA commit + the four column-sum warps
    }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, G::tmem_cols);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  uint32_t tmem;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(tmem_slot));

  if (warp == 0) {
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (long long kb = kb0; kb < kb1; ++kb) {
        mbar_wait(empty(s), ph ^ 1);
        mbar_arrive_expect_tx(full(s), G::stage_bytes);
        const uint32_t st = sbase + s * G::stage_bytes;
        const int row = (int)(kb * 64);
        if (JOB == 0) {
          for (int b = 0; b < 4; ++b) {
            const int c[2] = {b * 64, row};
            tma_load(st + b * WG_BOX, &mA0, full(s), 2, c);
            tma_load(st + (4 + b) * WG_BOX, &mB0, full(s), 2, c);
          }
        } else {
          for (int b = 0; b < 4; ++b) {
            const int c[2] = {b * 64, row};
            tma_load(st + b * WG_BOX, &mA0, full(s), 2, c);
            tma_load(st + (5 + b) * WG_BOX, &mA1, full(s), 2, c);
          }
          const int c0[2] = {0, row};
          tma_load(st + 4 * WG_BOX, &mB0, full(s), 2, c0);
          tma_load(st + 9 * WG_BOX, &mB1, full(s), 2, c0);
        }
        if (++s == G::stages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // both operands MN-major: 64-feature groups one box (8192 B) apart, 8-k-row groups 1024 B apart, a k-step of 16
      // rows advances the start address by 2048 B
      const uint64_t dMN = umma_desc(0, WG_BOX, 1024, UMMA_LAYOUT_SW128);
      constexpr uint32_t IA = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(128 >> 4) << 24);
      constexpr uint32_t I256 = IA | ((uint32_t)(256 >> 3) << 17), I64 = IA | ((uint32_t)(64 >> 3) << 17);
      int s = 0;
      uint32_t ph = 0;
      for (long long kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full(s), ph);
        tc_fence_after_sync();
        const uint32_t st = sbase + s * G::stage_bytes;
        const uint32_t acc0 = kb > kb0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint32_t acc = acc0 | (k > 0);
          if (JOB == 0) {
            const uint64_t b = dMN + ((st + 4 * WG_BOX + 2048u * k) >> 4);
            umma_bf16_ss(tmem, dMN + ((st + 2048u * k) >> 4), b, I256, acc);                       // output features 0..127
            umma_bf16_ss(tmem + 256, dMN + ((st + 2 * WG_BOX + 2048u * k) >> 4), b, I256, acc);    // 128..255
          } else {
            const uint64_t b0 = dMN + ((st + 4 * WG_BOX + 2048u * k) >> 4), b1 = dMN + ((st + 9 * WG_BOX + 2048u * k) >> 4);
            umma_bf16_ss(tmem, dMN + ((st + 2048u * k) >> 4), b0, I64, acc);
            umma_bf16_ss(tmem + 64, dMN + ((st + 2 * WG_BOX + 2048u * k) >> 4), b0, I64, acc);
            umma_bf16_ss(tmem + 128, dMN + ((st + 5 * WG_BOX + 2048u * k) >> 4), b1, I64, acc);
            umma_bf16_ss(tmem + 192, dMN + ((st + 7 * WG_BOX + 2048u * k) >> 4), b1, I64, acc);
          }
        }
        umma_commit(empty(s));
        if (++s == G::stages) { s = 0; ph ^= 1; }
      }
      umma_commit(done);
    }
  } else if (kb1 > kb0) {
    {
      // bias gradients: thread (box b = warp - 2, word w = lane) owns features 64 b + 2 w, + 1 of the 256-wide dZ operand
      // (boxes 0..3 of the stage); the 16-byte chunk c of k-row k sits at c ^ (k & 7) (SWIZZLE_128B)
      const int b = warp - 2;
      float a0 = 0.f, a1 = 0.f, c0 = 0.f, c1 = 0.f;
      int s = 0;
      uint32_t ph = 0;
      for (long long kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full(s), ph);
        const uint32_t box = sbase + s * G::stage_bytes + b * WG_BOX + (((uint32_t)lane & 3u) << 2);
        const uint32_t chunk = (uint32_t)lane >> 2;
#pragma unroll 8
        for (uint32_t k = 0; k < 64; ++k) {
          uint32_t w;
          asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(box + k * 128u + ((chunk ^ (k & 7u)) << 4)));
          a0 += bf16_lo(w);
          a1 += bf16_hi(w);
        }
        if (JOB == 1 && b == 0) {   // dZ3 (box 9, 64 features)
          const uint32_t box3 = sbase + s * G::stage_bytes + 9 * WG_BOX + (((uint32_t)lane & 3u) << 2);
#pragma unroll 8
          for (uint32_t k = 0; k < 64; ++k) {
            uint32_t w;
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(box3 + k * 128u + ((chunk ^ (k & 7u)) << 4)));
            c0 += bf16_lo(w);
            c1 += bf16_hi(w);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty(s));
        if (++s == G::stages) { s = 0; ph ^= 1; }
      }
      float* db = JOB == 0 ? P.db2 : P.db1;
      atomicAdd(db + 64 * b + 2 * lane, a0);
      atomicAdd(db + 64 * b + 2 * lane + 1, a1);
      if (JOB == 1 && b == 0) {
        atomicAdd(P.db3 + 2 * lane, c0);
        atomicAdd(P.db3 + 2 * lane + 1, c1);
      }
    }
    // TMEM lane quadrant = warp % 4: warps 2..5 cover quadrants 2, 3, 0, 1.  Lane = feature m within the 128-row block.
    const int quad = warp & 3;
    const uint32_t tlane = tmem + ((uint32_t)(quad * 32) << 16);
    const int m = quad * 32 + lane;
    mbar_wait(done, 0);
    tc_fence_after_sync();
    constexpr int NCH = G::tmem_cols / 32;
#pragma unroll 1
    for (int ch = 0; ch < NCH; ++ch) {
      uint32_t r[32];
      tmem_ld32(tlane + 32 * ch, r);
      tmem_ld_wait();
      float* dst;
      int stride = 1;
      if (JOB == 0) {
        // columns [0,256): dW2[m][n]; [256,512): dW2[128 + m][n]
        const int blk = ch >> 3, n0 = (ch & 7) * 32;
        dst = P.dw2 + (size_t)(blk * 128 + m) * H + n0;
      } else if (ch < 4) {
        // [0,64): dW1[m][i]; [64,128): dW1[128 + m][i]
        const int blk = ch >> 1, n0 = (ch & 1) * 32;
        dst = P.dw1 + (size_t)(blk * 128 + m) * C + n0;
      } else {
        // [128,192): dW3^T[m][o]; [192,256): dW3^T[128 + m][o]  ->  dW3[o][blk * 128 + m]
        const int blk = (ch - 4) >> 1, n0 = ((ch - 4) & 1) * 32;
        dst = P.dw3 + (size_t)n0 * H + blk * 128 + m;
        stride = H;
      }
#pragma unroll
      for (int k = 0; k < 32; ++k) atomicAdd(dst + (size_t)k * stride, __uint_as_float(r[k]));
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc(tmem, G::tmem_cols);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess) p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}
// [T][cols] BF16 row-major, boxes of 64 rows x 64 columns, 128B swizzle, rows beyond T read as zero
static bool bf16_map(CUtensorMap* m, const void* base, long long T, int cols) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return false;
  const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)T};
  const cuuint64_t gstr[1] = {(cuuint64_t)cols * 2};
  const cuuint32_t box[2] = {64, 64}, estr[2] = {1, 1};
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// [T][cols] row-major (elem_bytes per element), boxes of 32 rows x 32 bytes, 32B swizzle: the epilogue warps' staging boxes
static bool stage_map(CUtensorMap* m, const void* base, long long T, int cols, int elem_bytes) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return false;
  const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)T};
  const cuuint64_t gstr[1] = {(cuuint64_t)cols * elem_bytes};
  const cuuint32_t box[2] = {(cuuint32_t)(32 / elem_bytes), 32}, estr[2] = {1, 1};
  return enc(m, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), gdim,
             gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <typename K>
static cudaError_t set_smem(K kernel, int bytes) {
  // per-device attribute: set on every call (cheap) so that a process driving several GPUs configures each of them
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

}  // namespace ff
}  // namespace rpde

using namespace rpde;
using namespace rpde::ff;

static inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }
static int sm_count() {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms;
}
#define FF_CUDA(expr)                                                                  \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) return rpde_internal_fail(RPDE_E_CUDA, cudaGetErrorString(_e)); \
  } while (0)

extern "C" {

size_t rpde_ff3_weight_image_bytes(void) { return WBLOB; }

int rpde_ff3_supported(int c, int hidden) { return c == C && hidden == H; }

int rpde_ff3_pack_weights(const float* w1, const float* w2, const float* w3, void* image, rpde_stream_t stream) {
  if (!w1 || !w2 || !w3 || !image || (reinterpret_cast<uintptr_t>(image) & 15))
    return rpde_internal_fail(RPDE_E_INVALID, "rpde_ff3_pack_weights: bad arguments");
  ff_pack_weights_kernel<<<(WBLOB / 16 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w1, w2, w3, (uint8_t*)image);
  rpde_internal_count_launch(0);
  FF_CUDA(cudaGetLastError());
  return RPDE_OK;
}

int rpde_ff3_fwd(const float* x, const float* res, const void* image, const float* b1, const float* b2, const float* b3,
                 const float* gamma, const float* beta, float* y, long long T, float eps, rpde_stream_t stream) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!x || !image || !b1 || !b2 || !b3 || !y || T <= 0 || (gamma && !beta) || !al(x) || !al(res) || !al(image) || !al(b1) || !al(b2) ||
      !al(b3) || !al(gamma) || !al(beta) || !al(y))
    return rpde_internal_fail(RPDE_E_INVALID, "rpde_ff3_fwd: bad arguments (NULL or not 16-byte aligned)");
  if (T > 0x7fffffffLL) return rpde_internal_fail(RPDE_E_INVALID, "rpde_ff3_fwd: T too large");
  Params P{};
  P.x = x; P.res = res; P.blob = (const uint8_t*)image; P.b1 = b1; P.b2 = b2; P.b3 = b3; P.gamma = gamma; P.beta = beta;
  P.T = T; P.eps = eps;
  Maps M;
  memset(&M, 0, sizeof(M));
  if (!stage_map(&M.out, y, T, C, 4)) return rpde_internal_fail(RPDE_E_CUDA, "rpde_ff3_fwd: cuTensorMapEncodeTiled failed");
  FF_CUDA(set_smem(ff3_kernel<false>, SMEM_BYTES));
  const long long tiles = (T + TM - 1) / TM;
  const int sms = sm_count();
  const unsigned grid = (unsigned)(tiles < sms ? tiles : sms);
  ff3_kernel<false><<<grid, kThreads, SMEM_BYTES, (cudaStream_t)stream>>>(P, M);
  rpde_internal_count_launch(1);
  FF_CUDA(cudaGetLastError());
  return RPDE_OK;
}

size_t rpde_ff3_bwd_workspace_bytes(long long T) {
  const size_t t = (size_t)T;
  return up256(t * C * 2) * 2 + up256(t * H * 2) * 4 + up256((size_t)256 * 2 * TM * 512);
}

int rpde_ff3_bwd(const float* x, const float* dy, const void* image, const float* b1, const float* b2, const float* b3,
                 const float* gamma, float* ds, float* dw1, float* db1, float* dw2, float* db2, float* dw3, float* db3,
                 float* dgamma, float* dbeta, long long T, float eps, void* ws, size_t ws_bytes, rpde_stream_t stream) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!x || !dy || !image || !b1 || !b2 || !b3 || !ds || !dw1 || !db1 || !dw2 || !db2 || !dw3 || !db3 || T <= 0 || T > 0x7fffffffLL ||
      (gamma && (!dgamma || !dbeta)) || !al(x) || !al(dy) || !al(image) || !al(b1) || !al(b2) || !al(b3) || !al(gamma) || !al(ds) || !al(ws))
    return rpde_internal_fail(RPDE_E_INVALID, "rpde_ff3_bwd: bad arguments (NULL or not 16-byte aligned)");
  if (!ws || ws_bytes < rpde_ff3_bwd_workspace_bytes(T)) return rpde_internal_fail(RPDE_E_WORKSPACE, "rpde_ff3_bwd: workspace too small");
  cudaStream_t s = (cudaStream_t)stream;
  const int sms = sm_count();
  if (sms > 256) return rpde_internal_fail(RPDE_E_INVALID, "rpde_ff3_bwd: more than 256 SMs");
  const size_t t = (size_t)T;
  uint8_t* w = (uint8_t*)ws;
  Params P{};
  P.x = x; P.dy = dy; P.blob = (const uint8_t*)image; P.b1 = b1; P.b2 = b2; P.b3 = b3; P.gamma = gamma; P.T = T; P.eps = eps;
  uint16_t *dz3s, *h1s, *h2s, *dz1s, *dz2s;
  P.xb = (uint16_t*)w; w += up256(t * C * 2);
  dz3s = (uint16_t*)w; w += up256(t * C * 2);
  h1s = (uint16_t*)w;  w += up256(t * H * 2);
  h2s = (uint16_t*)w;  w += up256(t * H * 2);
  dz1s = (uint16_t*)w; w += up256(t * H * 2);
  dz2s = (uint16_t*)w; w += up256(t * H * 2);
  P.gscratch = w;
  Maps M;
  memset(&M, 0, sizeof(M));
  if (!stage_map(&M.out, ds, T, C, 4) || !stage_map(&M.h1, h1s, T, H, 2) || !stage_map(&M.h2, h2s, T, H, 2) ||
      !stage_map(&M.dz1, dz1s, T, H, 2) || !stage_map(&M.dz2, dz2s, T, H, 2) || !stage_map(&M.dz3, dz3s, T, C, 2))
    return rpde_internal_fail(RPDE_E_CUDA, "rpde_ff3_bwd: cuTensorMapEncodeTiled failed");
  P.dgamma = dgamma; P.dbeta = dbeta;
  FF_CUDA(cudaMemsetAsync(db1, 0, H * sizeof(float), s));
  FF_CUDA(cudaMemsetAsync(db2, 0, H * sizeof(float), s));
  FF_CUDA(cudaMemsetAsync(db3, 0, C * sizeof(float), s));
  if (gamma) {
    FF_CUDA(cudaMemsetAsync(dgamma, 0, C * sizeof(float), s));
    FF_CUDA(cudaMemsetAsync(dbeta, 0, C * sizeof(float), s));
  }
  FF_CUDA(cudaMemsetAsync(dw1, 0, (size_t)H * C * sizeof(float), s));
  FF_CUDA(cudaMemsetAsync(dw2, 0, (size_t)H * H * sizeof(float), s));
  FF_CUDA(cudaMemsetAsync(dw3, 0, (size_t)C * H * sizeof(float), s));
  FF_CUDA(set_smem(ff3_kernel<true>, SMEM_BYTES));
  const long long tiles = (T + TM - 1) / TM;
  const unsigned grid = (unsigned)(tiles < sms ? tiles : sms);
  ff3_kernel<true><<<grid, kThreads, SMEM_BYTES, s>>>(P, M);
  rpde_internal_count_launch(1);
  FF_CUDA(cudaGetLastError());
  // weight gradients from the BF16 copies
  CUtensorMap mdz2, mh1, mdz1, mxb, mh2, mdz3;
  if (!bf16_map(&mdz2, dz2s, T, H) || !bf16_map(&mh1, h1s, T, H) || !bf16_map(&mdz1, dz1s, T, H) || !bf16_map(&mxb, P.xb, T, C) ||
      !bf16_map(&mh2, h2s, T, H) || !bf16_map(&mdz3, dz3s, T, C))
    return rpde_internal_fail(RPDE_E_CUDA, "rpde_ff3_bwd: cuTensorMapEncodeTiled failed");
  WgParams W{T, dw1, dw2, dw3, db1, db2, db3};
  const long long kblocks = (T + 63) / 64;
  const unsigned wgrid = (unsigned)(kblocks < sms ? kblocks : sms);
  FF_CUDA(set_smem(ff_wgrad_kernel<0>, WgCfg<0>::smem_bytes));
  ff_wgrad_kernel<0><<<wgrid, WG_THREADS, WgCfg<0>::smem_bytes, s>>>(W, mdz2, mh1, mdz2, mh1);
  rpde_internal_count_launch(1);
  FF_CUDA(cudaGetLastError());
  FF_CUDA(set_smem(ff_wgrad_kernel<1>, WgCfg<1>::smem_bytes));
  ff_wgrad_kernel<1><<<wgrid, WG_THREADS, WgCfg<1>::smem_bytes, s>>>(W, mdz1, mxb, mh2, mdz3);
  rpde_internal_count_launch(1);
  FF_CUDA(cudaGetLastError());
  return RPDE_OK;
}

}  // extern "C"
