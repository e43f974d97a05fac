// Tensor-core contraction engine for sm_100a: TMA -> 128B-swizzled shared memory -> tcgen05.mma (kind::tf32,
// accumulators in TMEM) -> tcgen05.ld epilogue.  Same GemmP contract as gemm_simt.cuh.
//
// Precision variants (rpde_set_precision):
//   3xTF32 : every fp32 operand tile is split on chip into hi = tf32(x) and lo = x - hi (two-deep lo ring beside the raw
//            TMA ring); hi*hi accumulates in one TMEM column block, hi*lo + lo*hi in a second one, the epilogue adds
//            them -> fp32-level accuracy (the parity engine).
//   TF32   : one pass on the raw fp32 bits (~5e-4 relative), the fast variant with its own stated tolerance.
//
// One persistent CTA per SM, warp-specialised:
//   warp 0, 14  TMA producers (one lane each: A tiles / B tiles) filling a ring of raw fp32 stages, full[s] <- complete_tx
//   warp 1      UMMA issuer (one lane) + TMEM owner: 4 (x3) tcgen05.mma per stage, commit -> empty[s]; last -> acc_full
//   warps 2-5   splitter (3xTF32 only): lo = x - tf32(x) of the landed tiles into a two-deep lo ring, fence.proxy.async,
//               one arrival per warp -> ready[s]
//   warps 6-13  epilogue (two warps per TMEM lane quadrant, half of the columns each): tcgen05.ld 32x32b.x32, bias /
//               activation / act' in registers, rows staged in a 128B-swizzled 32x32 box per warp and written with bulk tensor
//               stores (Z and C), old C / aux fetched into the same box with bulk tensor loads, beta-only as a bulk
//               reduce-add; outputs TMA cannot address (split-K atomics, row-contiguous or unaligned C) take the
//               register/transposition path.  Two TMEM accumulators: the epilogue of tile i overlaps the main loop of i+1.
// Operands are addressed through TMA tensor maps derived from the GemmP index maps (K-major or MN-major, up to two
// levels per index), so the channels-last / channels-first / interleaved layouts of the spectral path need no copies.
#pragma once
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <mutex>
#include <type_traits>

#include "common.cuh"
#include "gemm_simt.cuh"
#include "tc_ptx.cuh"

namespace rpde {

struct TcOpnd {
  int major;    // 0: K-major (K contiguous), 1: MN-major (M or N contiguous)
  int ndim;     // tensor-map rank
  int boxes;    // TMA issues per stage (1 unless the MN-major slab fallback is in use)
  int box_mn;   // MN elements covered by one issue
  int box_bytes;
  int grouped;  // MN-major: the 32-element MN groups are a tensor dimension of their own (one box per stage)
  int mn_d, k_d, b_d;                            // level divisors (INT_MAX: single level)
  signed char dm0, dm1, dk0, dk1, db0, db1, dg;  // tensor-map dimension of each level, -1 if absent
};

struct TcParams {
  GemmP g;
  TcOpnd a, b, c;   // c: the output viewed as rows x contiguous columns (TMA epilogue), boxes of 32 x 32
  int epi_tma;      // 1: C / Z leave through bulk tensor stores from the staging boxes, aux / old C arrive the same way
  int epi_reduce;   // 1: beta is realised as a bulk tensor reduce-add (no activation, no Z)
  int n_fast;       // tile order: n runs fastest
  int l2_prefetch;  // bit 0 / 1: prefetch the next tile's A / B boxes into L2; bit 2: the epilogue's input boxes
  int tiles_m, tiles_n, kblocks, kchunk_blocks, total_tiles;
  int kc;       // k-blocks per TMEM accumulation chain (chunk); chunks are summed in registers with RN adds
  int c_vec;    // C / Z / aux / bias base pointers are 16-byte aligned
  int dbg;
  int epi_mode;  // EM_* bits when the epilogue matches a compiled mode, else -1
  int skip;  // bring-up only (RPDE_TC_SKIP bitmask): 1 epilogue emit, 2 operand prep, 4 UMMA issue, 8 TMA loads
};

namespace tc {

// -DRPDE_TC_PROFILE: cycle counters around the waits / issues of the two single-thread roles (bring-up only)
#ifdef RPDE_TC_PROFILE
#define TCP_DECL(...) long long __VA_ARGS__
#define TCP_T0(v) const long long v = clock64()
#define TCP_ADD(acc, v) acc += clock64() - v
#else
#define TCP_DECL(...)
#define TCP_T0(v)
#define TCP_ADD(acc, v)
#endif

constexpr int BM = 128, BK = 32;
// compile-time epilogue modes of the TMA epilogue (TcParams::epi_mode; -1 = every flag read at run time)
constexpr int EM_IN_MUL = 1, EM_BIAS = 2, EM_GELU_G = 4, EM_COLSUM = 8;
constexpr int A_TILE_BYTES = BM * BK * 4;
constexpr int kEpiWarps = 8;                         // two warps per TMEM lane quadrant, each owning half of the tile's columns
constexpr int kProducerBWarp = 6 + kEpiWarps;         // second TMA producer (B operand) behind the epilogue warps
constexpr int kThreads = 32 * (7 + kEpiWarps);
constexpr int kEpiBytes = kEpiWarps * 8192;          // two 128B-swizzled 32x32 fp32 staging boxes per epilogue warp
constexpr int kSmemBudget = 227 * 1024 - 2048 - kEpiBytes;

template <int BN, bool SPLIT>
struct Cfg {
  static constexpr int b_tile = BN * BK * 4;
  // The TMA ring holds raw fp32 tiles only.  3xTF32: the lo tiles live in a separate two-deep ring that is occupied just
  // from the split until the UMMAs that read it retire, so the bytes in flight behind the tensor core are not halved.
  // 3xTF32 feeds A from TMEM (hi and lo written there by the splitter warps), so only B_lo needs shared memory.
  static constexpr int stage_bytes = A_TILE_BYTES + b_tile;
  static constexpr int lo_bufs = SPLIT ? 2 : 0;
  static constexpr int lo_bytes = lo_bufs * b_tile;
  static constexpr int raw = (kSmemBudget - lo_bytes) / stage_bytes;
  static constexpr int stages = raw > 6 ? 6 : raw;
  static constexpr int smem_bytes = stages * stage_bytes + lo_bytes + kEpiBytes + 2048;  // + alignment slack + barriers
  // TMEM columns.  TF32: two accumulators [acc0 | acc1].  3xTF32: [big0 | big1 | small | A ring: 2 x (hi 32 | lo 32)]:
  // hi*hi accumulates in big (double-buffered across tiles), hi*lo + lo*hi in small (single: the epilogue folds it into
  // big first thing and hands it back), and the A operand of the next two k-blocks lives in the ring.
  static constexpr int acc_cols = BN;
  static constexpr int small_col = 2 * BN;
  static constexpr int aring_col = 3 * BN;
  static constexpr int need_cols = SPLIT ? 3 * BN + 128 : 2 * BN;
  static constexpr int tmem_cols = need_cols <= 32 ? 32 : need_cols <= 64 ? 64 : need_cols <= 128 ? 128 : need_cols <= 256 ? 256 : 512;
  static_assert(need_cols <= 512, "TMEM holds 512 columns");
};

// Coordinates live in registers: slots are written with compare-selects, never with dynamic indexing (which would
// push the array to local memory on the producer's critical path).
__device__ __forceinline__ void put(int (&c)[5], int slot, int val) {
#pragma unroll
  for (int i = 0; i < 5; ++i) c[i] = (slot == i) ? val : c[i];
}
// coordinates of the box at (mn, k = 0, bz)
__device__ __forceinline__ void opnd_coords(const TcOpnd& o, int mn, int bz, int (&c)[5]) {
#pragma unroll
  for (int i = 0; i < 5; ++i) c[i] = 0;
  if (o.grouped) {
    put(c, o.dg, ((o.mn_d == INT_MAX) ? mn : mn % o.mn_d) >> 5);
    if (o.dm1 >= 0) put(c, o.dm1, mn / o.mn_d);
  } else {
    if (o.dm0 >= 0) put(c, o.dm0, (o.mn_d == INT_MAX) ? mn : mn % o.mn_d);
    if (o.dm1 >= 0) put(c, o.dm1, mn / o.mn_d);
  }
  if (o.db0 >= 0) put(c, o.db0, (o.b_d == INT_MAX) ? bz : bz % o.b_d);
  if (o.db1 >= 0) put(c, o.db1, bz / o.b_d);
}
__device__ __forceinline__ void opnd_set_k(int k_d, int dk0, int dk1, int k, int (&c)[5]) {
  if (k_d == INT_MAX) {
    put(c, dk0, k);
  } else {
    put(c, dk0, k % k_d);
    if (dk1 >= 0) put(c, dk1, k / k_d);
  }
}
__device__ __forceinline__ void opnd_load(const TcOpnd& o, const CUtensorMap* map, uint32_t dst, uint32_t bar, int (&c)[5], int mn0) {
  if (o.boxes == 1) {
    tma_load(dst, map, bar, o.ndim, c);
    return;
  }
  for (int i = 0; i < o.boxes; ++i) {  // MN-major slab fallback (ragged / deeply nested MN index)
    const int mn = mn0 + i * o.box_mn;
    if (o.dm0 >= 0) put(c, o.dm0, (o.mn_d == INT_MAX) ? mn : mn % o.mn_d);
    if (o.dm1 >= 0) put(c, o.dm1, mn / o.mn_d);
    tma_load(dst + i * o.box_bytes, map, bar, o.ndim, c);
  }
}

// tile id -> (tm, tn, z): the dimension with fewer tiles runs fastest so that concurrently resident CTAs share the
// larger operand's tiles through L2
__device__ __forceinline__ void tile_coords(const TcParams& P, int t, int& tm, int& tn, int& z) {
  const int tiles_mn = P.tiles_m * P.tiles_n;
  z = t / tiles_mn;
  const int r = t - z * tiles_mn;
  if (P.n_fast) {
    tn = r % P.tiles_n;
    tm = r / P.tiles_n;
  } else {
    tm = r % P.tiles_m;
    tn = r / P.tiles_m;
  }
}

// warp-uniform activation over one 32-column slice (branches are uniform, the unrolled bodies give the ILP)
__device__ __forceinline__ void act32(int act, float (&v)[32]) {
  if (act == ACT_GELU) {
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
      float g0, g1;
      gelu_both2(v[j], v[j + 1], v[j], v[j + 1], g0, g1);
    }
  } else if (act == ACT_RELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = v[j] > 0.f ? v[j] : 0.f;
  }
}
__device__ __forceinline__ void dact32_mul(int act, float (&v)[32], const float (&z)[32]) {
  if (act == 3) {  // z already holds act'(pre-activation)
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= z[j];
  } else if (act == ACT_GELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= dgelu_f(z[j]);
  } else if (act == ACT_RELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = z[j] > 0.f ? v[j] : 0.f;
  }
}
__device__ __forceinline__ float4 act4(int act, float4 x) {
  x.x = act_f(act, x.x); x.y = act_f(act, x.y); x.z = act_f(act, x.z); x.w = act_f(act, x.w);
  return x;
}
__device__ __forceinline__ float4 dact4(int act, float4 z) {
  z.x = dact_f(act, z.x); z.y = dact_f(act, z.y); z.z = dact_f(act, z.z); z.w = dact_f(act, z.w);
  return z;
}

__device__ __forceinline__ float tf32_trunc(float v) { return __uint_as_float(__float_as_uint(v) & 0xffffe000u); }
__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t a, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// in-place preparation of one landed operand region: optional activation-on-load, optional hi/lo split.
// The tensor core truncates fp32 operands to TF32 itself, so the hi tile is the raw tile; lo = x - trunc(x) is exact.
// n16 (float4 count) is a multiple of 512: every thread handles whole batches of four, loads first, so the shared-memory
// latency is paid once per batch instead of once per element.
template <bool SPLIT>
__device__ __forceinline__ void prep_region(uint32_t hi0, uint32_t lo0, int n16, int act, int ct) {
  for (int i0 = ct; i0 < n16; i0 += 512) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = lds128(hi0 + 16 * (i0 + 128 * u));
    if (act) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        v[u] = act4(act, v[u]);
        sts128(hi0 + 16 * (i0 + 128 * u), v[u]);
      }
    }
    if (SPLIT) {
#pragma unroll
      for (int u = 0; u < 4; ++u)
        sts128(lo0 + 16 * (i0 + 128 * u), make_float4(v[u].x - tf32_trunc(v[u].x), v[u].y - tf32_trunc(v[u].y),
                                                      v[u].z - tf32_trunc(v[u].z), v[u].w - tf32_trunc(v[u].w)));
    }
  }
}

// 3xTF32, A operand: thread (quad, lane) owns row quad*32 + lane of the landed 128 x 32 tile.  It gathers the row from the
// swizzled stage (K-major: 8 x 16 B chunks, chunk c at c ^ (row & 7); MN-major: 32 words, one per k, the 32-byte chunk
// (lane / 8) ^ (k & 3) of row k inside group quad), applies the optional activation, and writes tf32(x) and x - tf32(x)
// into the TMEM ring slot (lane = row, 32 + 32 columns).  The UMMAs then fetch A from TMEM, B from shared memory.
__device__ __forceinline__ void prep_a_tmem(uint32_t a_smem, uint32_t taddr, int quad, int lane, int mn_major, int act) {
  float x[32];
  if (!mn_major) {
    const uint32_t row = a_smem + (uint32_t)(quad * 32 + lane) * 128u, sw = lane & 7;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const float4 v = lds128(row + ((c ^ sw) << 4));
      x[4 * c] = v.x; x[4 * c + 1] = v.y; x[4 * c + 2] = v.z; x[4 * c + 3] = v.w;
    }
  } else {
    const uint32_t base = a_smem + (uint32_t)quad * 4096u + (((uint32_t)lane & 7u) << 2), hi3 = (uint32_t)lane >> 3;
#pragma unroll
    for (int k = 0; k < 32; ++k)
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x[k]) : "r"(base + k * 128u + ((hi3 ^ (k & 3u)) << 5)));
  }
  if (act) {
#pragma unroll
    for (int j = 0; j < 32; ++j) x[j] = act_f(act, x[j]);
  }
  uint32_t hi[32], lo[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const float h = tf32_trunc(x[j]);
    hi[j] = __float_as_uint(h);
    lo[j] = __float_as_uint(x[j] - h);
  }
  tmem_st32(taddr, hi);
  tmem_st32(taddr + 32, lo);  // the caller overlaps the B split with these stores and then waits (tmem_st_wait)
}

// CH = true: a tile's K range is cut into chunks of P.kc k-blocks, each accumulated in TMEM from zero and then added
// (round-to-nearest) into fp32 registers of the epilogue warps.  The tensor core truncates its accumulator after
// every UMMA, so the chain length bounds the rounding bias (measured 2.2e-8 relative per UMMA).
template <int BN, bool SPLIT, bool CH>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ TcParams P, const __grid_constant__ CUtensorMap mapA,
               const __grid_constant__ CUtensorMap mapB, const __grid_constant__ CUtensorMap mapC,
               const __grid_constant__ CUtensorMap mapZ, const __grid_constant__ CUtensorMap mapI) {
  using C = Cfg<BN, SPLIT>;
  constexpr int S = C::stages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // stage s: [A][B] raw fp32 (the tensor core truncates to TF32 itself: raw = hi); lo buffer l: [A_lo][B_lo]
  auto a_hi = [&](int s) { return smem_base + s * C::stage_bytes; };
  auto b_hi = [&](int s) { return smem_base + s * C::stage_bytes + A_TILE_BYTES; };
  const uint32_t lo_base = smem_base + S * C::stage_bytes;
  auto b_lo = [&](int l) { return lo_base + l * C::b_tile; };
  const uint32_t epi_base = lo_base + C::lo_bytes;  // 1024-aligned: the staging boxes use the TMA 128B swizzle
  const uint32_t bar_base = epi_base + kEpiBytes;
  auto full = [&](int s) { return bar_base + 8 * s; };
  auto ready = [&](int s) { return bar_base + 8 * (S + s); };
  auto empty = [&](int s) { return bar_base + 8 * (2 * S + s); };
  auto acc_full = [&](int a) { return bar_base + 8 * (3 * S + a); };
  auto acc_empty = [&](int a) { return bar_base + 8 * (3 * S + 2 + a); };
  auto epi_in = [&](int w) { return bar_base + 8 * (3 * S + 4 + w); };
  auto lo_empty = [&](int l) { return bar_base + 8 * (3 * S + 4 + kEpiWarps + l); };
  const uint32_t small_empty = bar_base + 8 * (3 * S + 4 + kEpiWarps + 2);
  const uint32_t tmem_slot = bar_base + 8 * (3 * S + 4 + kEpiWarps + 3);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const GemmP& p = P.g;
  const bool use_prep = SPLIT || p.a_act || p.b_act;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapA);
    tma_prefetch_desc(&mapB);
    if (P.epi_tma) tma_prefetch_desc(&mapC);
    for (int s = 0; s < S; ++s) {
      mbar_init(full(s), 2);  // one arrive.expect_tx per producer
      mbar_init(ready(s), 4);
      mbar_init(empty(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(acc_full(a), 1);
      mbar_init(acc_empty(a), kEpiWarps);
    }
    for (int w = 0; w < kEpiWarps; ++w) mbar_init(epi_in(w), 1);
    for (int l = 0; l < 2; ++l) mbar_init(lo_empty(l), 1);
    mbar_init(small_empty, kEpiWarps);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, C::tmem_cols);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  // Programmatic dependent launch: everything above (barriers, TMEM, descriptor prefetch) touches no global data and may
  // run while the previous kernel of the stream drains; from here on its results are needed.  The next kernel may
  // begin its own prologue as soon as every CTA of this one has passed this point.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  if (warp == 0 || warp == kProducerBWarp) {
    // ------------------------------------------------------------------ TMA producers: warp 0 loads the A tiles, the last
    // warp the B tiles (issuing one bulk tensor load costs a thread ~200 cycles; two threads keep the ring ahead of the
    // tensor core).  The k loop is a handful of integer multiply-adds: coordinates advance as base + k0*e0 + k1*e1 with
    // 0/1 masks chosen once per kernel; ring slot and phase are carried.
    if (lane == 0) {
      const bool isA = warp == 0;
      const TcOpnd& o = isA ? P.a : P.b;
      const CUtensorMap* map = isA ? &mapA : &mapB;
      const int kd = o.k_d, nd = o.ndim;
      const int tile_mn = isA ? BM : BN;
      const uint32_t bytes = isA ? (uint32_t)A_TILE_BYTES : (uint32_t)C::b_tile;
      const uint32_t dst_off = isA ? 0u : (uint32_t)A_TILE_BYTES;
      int e0[5], e1[5];
#pragma unroll
      for (int j = 0; j < 5; ++j) {
        e0[j] = (j == o.dk0) ? 1 : 0;
        e1[j] = (kd != INT_MAX && j == o.dk1) ? 1 : 0;
      }
      const bool simple = o.boxes == 1;
      // the B producer also pulls the NEXT tile's A boxes into L2 (bit 0 of l2_prefetch), one box per k-block
      const bool pf = !isA && (P.l2_prefetch & 1) && P.a.boxes == 1;
      const int a_kd = P.a.k_d;
      int s = 0;
      uint32_t ph = 0;
      TCP_DECL(pw = 0, pi = 0, pall = clock64(), pn = 0);
      for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
        int tm, tn, z;
        tile_coords(P, t, tm, tn, z);
        const int sk = z % p.splitk, bz = z / p.splitk;
        const int kb0 = sk * P.kchunk_blocks, kb1 = min(P.kblocks, kb0 + P.kchunk_blocks);
        const int mn0 = (isA ? tm : tn) * tile_mn;
        int cbase[5];
        opnd_coords(o, mn0, bz, cbase);
        const int k_start = kb0 * BK;
        int k0 = (kd == INT_MAX) ? k_start : k_start % kd, k1 = (kd == INT_MAX) ? 0 : k_start / kd;
        // next tile (prefetch target)
        int pbase[5], pk0 = 0, pk1 = 0, pkn = 0;
        bool pf_on = false;
        if (pf && t + (int)gridDim.x < P.total_tiles) {
          int tm2, tn2, z2;
          tile_coords(P, t + gridDim.x, tm2, tn2, z2);
          const int sk2 = z2 % p.splitk, bz2 = z2 / p.splitk;
          const int q0 = sk2 * P.kchunk_blocks;
          pkn = min(P.kblocks, q0 + P.kchunk_blocks) - q0;
          pf_on = (tm2 != tm || bz2 != bz || sk2 != sk);
          opnd_coords(P.a, tm2 * BM, bz2, pbase);
          const int ks = q0 * BK;
          pk0 = (a_kd == INT_MAX) ? ks : ks % a_kd;
          pk1 = (a_kd == INT_MAX) ? 0 : ks / a_kd;
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          int x[5];
#pragma unroll
          for (int j = 0; j < 5; ++j) x[j] = cbase[j] + k0 * e0[j] + k1 * e1[j];
          const uint32_t fb = full(s), dst = a_hi(s) + dst_off;
          TCP_T0(q0);
          mbar_wait(empty(s), ph ^ 1);
          TCP_ADD(pw, q0);
          TCP_T0(q1);
          if (P.skip & 8) {
            mbar_arrive(fb);
          } else {
            mbar_arrive_expect_tx(fb, bytes);
            if (simple) tma_load(dst, map, fb, nd, x);
            else opnd_load(o, map, dst, fb, x, mn0);
          }
          if (pf_on && kb - kb0 < pkn) {
            int y[5];
#pragma unroll
            for (int j = 0; j < 5; ++j)
              y[j] = pbase[j] + ((j == P.a.dk0) ? pk0 : 0) + ((a_kd != INT_MAX && j == P.a.dk1) ? pk1 : 0);
            tma_prefetch_l2(&mapA, P.a.ndim, y);
            pk0 += BK;
            if (pk0 >= a_kd) { pk0 = 0; ++pk1; }
          }
          TCP_ADD(pi, q1);
#ifdef RPDE_TC_PROFILE
          ++pn;
#endif
          k0 += BK;
          if (k0 >= kd) { k0 = 0; ++k1; }
          if (++s == S) { s = 0; ph ^= 1; }
        }
      }
#ifdef RPDE_TC_PROFILE
      if (blockIdx.x == 0) printf("[tc producer %c] iters %lld total %lld wait_empty %lld issue %lld\n", isA ? 'A' : 'B', pn, clock64() - pall, pw, pi);
#endif
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ UMMA issuer
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_tf32(BM, BN, P.a.major, P.b.major);
      const uint32_t idesc_ts = umma_idesc_tf32(BM, BN, 0, P.b.major);  // A from TMEM is never transposed
      const uint32_t a_step = (P.a.major ? 1024u : 32u) >> 4, b_step = (P.b.major ? 1024u : 32u) >> 4;
      const uint32_t a_lbo = P.a.major ? BK * 128u : 16u, b_lbo = P.b.major ? BK * 128u : 16u;
      const uint32_t a_sbo = P.a.major ? 512u : 1024u, b_sbo = P.b.major ? 512u : 1024u;
      const uint32_t a_lay = P.a.major ? UMMA_LAYOUT_SW128_BASE32B : UMMA_LAYOUT_SW128;
      const uint32_t b_lay = P.b.major ? UMMA_LAYOUT_SW128_BASE32B : UMMA_LAYOUT_SW128;
      // descriptors with a zero start address: the 14-bit address field (smem offset >> 4 < 2^14) is added per use
      const uint64_t dA = umma_desc(0, a_lbo, a_sbo, a_lay), dB = umma_desc(0, b_lbo, b_sbo, b_lay);
      int s = 0, l = 0;
      uint32_t ph = 0, cit = 0;
      TCP_DECL(mw = 0, ma = 0, mi = 0, mc = 0, mall = clock64(), mn = 0);
      for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
        int tm, tn, z;
        tile_coords(P, t, tm, tn, z);
        const int sk = z % p.splitk;
        const int kb0 = sk * P.kchunk_blocks, kb1 = min(P.kblocks, kb0 + P.kchunk_blocks);
        for (int c0 = kb0; c0 < kb1; c0 += P.kc, ++cit) {
          const int c1 = min(kb1, c0 + P.kc);
          const int buf = cit & 1;
          TCP_T0(r0);
          mbar_wait(acc_empty(buf), ((cit >> 1) & 1) ^ 1);
          if (SPLIT) mbar_wait(small_empty, (cit & 1) ^ 1);  // the epilogue has folded the previous chain's small block
          TCP_ADD(ma, r0);
          tc_fence_after_sync();
          const uint32_t d = tmem_base + buf * C::acc_cols;
          const uint32_t dsm = tmem_base + C::small_col;
          for (int kb = c0; kb < c1; ++kb) {
            TCP_T0(r1);
            mbar_wait(use_prep ? ready(s) : full(s), ph);
            TCP_ADD(mw, r1);
            TCP_T0(r2);
            tc_fence_after_sync();
            const uint64_t ah0 = dA + (a_hi(s) >> 4), bh0 = dB + (b_hi(s) >> 4);
            const uint64_t bl0 = dB + (b_lo(l) >> 4);
            const uint32_t at = tmem_base + C::aring_col + l * 64;  // A of this k-block in TMEM: [hi 32 cols | lo 32 cols]
            if (!(P.skip & 4)) {
#pragma unroll
              for (int k4 = 0; k4 < BK / 8; ++k4) {
                const uint32_t acc = (kb > c0 || k4 > 0) ? 1u : 0u;
                const uint64_t ah = ah0 + k4 * a_step, bh = bh0 + k4 * b_step;
                if (SPLIT) {
                  // big (+)= A_hi B_hi ; small (+)= A_hi B_lo + A_lo B_hi.  The large products meet one accumulator
                  // truncation per k-step; the 2^-11 corrections live in their own TMEM columns.  A comes from TMEM.
                  umma_tf32_ts(d, at + k4 * 8, bh, idesc_ts, acc);
                  umma_tf32_ts(dsm, at + k4 * 8, bl0 + k4 * b_step, idesc_ts, acc);
                  umma_tf32_ts(dsm, at + 32 + k4 * 8, bh, idesc_ts, 1u);
                } else {
                  umma_tf32(d, ah, bh, idesc, acc);
                }
              }
            }
            TCP_ADD(mi, r2);
            TCP_T0(r3);
            umma_commit(empty(s));
            if (SPLIT) {
              umma_commit(lo_empty(l));
              l ^= 1;
            }
            TCP_ADD(mc, r3);
#ifdef RPDE_TC_PROFILE
            ++mn;
#endif
            if (++s == S) { s = 0; ph ^= 1; }
          }
          umma_commit(acc_full(buf));
        }
      }
#ifdef RPDE_TC_PROFILE
      if (blockIdx.x == 0)
        printf("[tc mma] iters %lld total %lld wait_ready %lld wait_acc_empty %lld issue %lld commit %lld\n", mn, clock64() - mall, mw, ma, mi, mc);
#endif
    }
  } else if (warp < 6) {
    // ------------------------------------------------------------------ operand preparation (activation-on-load, hi/lo split)
    if (use_prep) {
      const int ct = threadIdx.x - 64;
      int s = 0, l = 0;
      uint32_t ph = 0, lph = 0;
      TCP_DECL(qf = 0, ql = 0, qa = 0, qb = 0, qt = 0, qall = clock64(), qn = 0);
      for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
        int tm, tn, z;
        tile_coords(P, t, tm, tn, z);
        const int sk = z % p.splitk;
        const int kb0 = sk * P.kchunk_blocks, kb1 = min(P.kblocks, kb0 + P.kchunk_blocks);
        for (int kb = kb0; kb < kb1; ++kb) {
          TCP_T0(w0);
          mbar_wait(full(s), ph);
          TCP_ADD(qf, w0);
          TCP_T0(w1);
          if (SPLIT) mbar_wait(lo_empty(l), lph ^ 1);  // the UMMAs of two iterations ago have released lo buffer l
          TCP_ADD(ql, w1);
          TCP_T0(w2);
          if (!(P.skip & 2)) {
            if (SPLIT) {
              // A: this thread's row (TMEM lane = tile row) -> hi / lo -> TMEM ring slot l
              const uint32_t at = tmem_base + C::aring_col + l * 64 + ((uint32_t)((warp & 3) * 32) << 16);
              prep_a_tmem(a_hi(s), at, warp & 3, lane, P.a.major, p.a_act);
            } else if (p.a_act) {
              prep_region<false>(a_hi(s), 0, A_TILE_BYTES / 16, p.a_act, ct);
            }
            TCP_ADD(qa, w2);
            TCP_T0(w3);
            if (SPLIT || p.b_act) prep_region<SPLIT>(b_hi(s), b_lo(l), C::b_tile / 16, p.b_act, ct);
            TCP_ADD(qb, w3);
          }
          TCP_T0(w4);
          if (SPLIT && !(P.skip & 2)) tmem_st_wait();
          if (SPLIT) tc_fence_before_sync();
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(ready(s));  // one arrival per warp
          TCP_ADD(qt, w4);
#ifdef RPDE_TC_PROFILE
          ++qn;
#endif
          if (SPLIT) {
            if (l == 1) lph ^= 1;
            l ^= 1;
          }
          if (++s == S) { s = 0; ph ^= 1; }
        }
      }
#ifdef RPDE_TC_PROFILE
      if (blockIdx.x == 0 && threadIdx.x == 64)
        printf("[tc prep] iters %lld total %lld wait_full %lld wait_lo_empty %lld A %lld B %lld tail %lld\n", qn, clock64() - qall, qf, ql, qa, qb, qt);
#endif
    }
  } else if (warp < kProducerBWarp) {
    // ------------------------------------------------------------------ epilogue: warps 6..13.  A warp may only touch the
    // TMEM lane quadrant (warp % 4); the two warps of a quadrant split the tile's columns in halves.
    const int ew = warp - 6;
    const int quad = warp & 3, half = ew >> 2;
    constexpr int CPW = (BN / 32) / 2;  // 32-column slices per epilogue warp
    // Two staging boxes per warp.  Box 1 receives the bulk loads (saved act' / old C), prefetched one slice ahead, or
    // the second output stream (Z); box 0 stages C.  With a single output stream and no input the boxes alternate, so
    // a slice never waits for the bulk store of the one before it.
    const uint32_t stg = epi_base + ew * 8192;
    const uint32_t stg1 = stg + 4096;
    const uint32_t sw = lane & 7;
    const uint32_t inbar = epi_in(ew);
    uint32_t in_phase = 0;
    const bool col_contig = (p.c.c.sl == 1) && (p.c.c.d == INT_MAX || (p.c.c.d % 32) == 0) && P.dbg != 3;
    const bool tma_in = P.epi_tma && (p.dact || (p.beta && !P.epi_reduce));
    uint32_t cit = 0;
    // bulk-store bookkeeping (lane 0 owns the groups; the counters are warp-uniform): ncommit groups committed so far,
    // groups <= done have finished reading shared memory, last[b] = the last group that reads box b
    uint32_t ncommit = 0, done = 0, last0 = 0, last1 = 0, flip = 0;
    float cs_acc[CPW];  // running column sums (bias gradient) of the column slices this warp owns
    int cs_n0[CPW];
#pragma unroll
    for (int i = 0; i < CPW; ++i) { cs_acc[i] = 0.f; cs_n0[i] = -1; }
    TCP_DECL(ea = 0, ef = 0, ee = 0, ei = 0, eb = 0, eall = clock64(), en = 0, e1 = 0, e2 = 0, e3 = 0, g1 = 0, g2 = 0, g3 = 0, g4 = 0);
    bool pf_valid = false;  // a prefetched input box is in flight ...
    int pf_t = -1, pf_n0 = -1;  // ... for this (tile, first column)

    auto box_free = [&](uint32_t last) {
      if (last > done) {
        TCP_T0(b0);
        const bool one = (ncommit - last) >= 1;
        if (lane == 0) {
          if (one) bulk_wait_read1();
          else bulk_wait_read0();
        }
        __syncwarp();
        done = ncommit - (one ? 1u : 0u);
        TCP_ADD(eb, b0);
      }
    };
    auto stage_write = [&](uint32_t box, const float (&v)[32]) {
      const uint32_t row_addr = box + lane * 128;
#pragma unroll
      for (int q = 0; q < 8; ++q) sts128(row_addr + ((q ^ sw) << 4), make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]));
    };
    // bulk load of the input box of (tile tt, slice ci) into box 1; false when that slice does not exist
    auto in_issue = [&](int tt, int ci) -> bool {
      if (tt >= P.total_tiles) return false;
      int tm2, tn2, z2;
      tile_coords(P, tt, tm2, tn2, z2);
      const int mb = tm2 * BM + quad * 32, nn = tn2 * BN + (half * CPW + ci) * 32;
      if (mb >= p.M || nn >= p.N) return false;
      int c5[5];
      opnd_coords(P.c, mb, z2 / p.splitk, c5);
      opnd_set_k(P.c.k_d, P.c.dk0, P.c.dk1, nn, c5);
      if (lane == 0) {
        mbar_arrive_expect_tx(inbar, 4096);
        tma_load(stg1, &mapI, inbar, P.c.ndim, c5);
      }
      pf_valid = true;
      pf_t = tt;
      pf_n0 = nn;
      return true;
    };

    // The tile loop is instantiated once per common epilogue mode (MODE >= 0: the flag tests fold at compile time, which
    // matters because an epilogue warp is a serial instruction stream with little to hide behind) and once with every
    // flag read at run time (MODE < 0).
    auto run_tiles = [&](auto mode_tag) {
    constexpr int MODE = decltype(mode_tag)::value;
    constexpr bool RT = MODE < 0;
    const bool f_tma = RT ? (P.epi_tma != 0) : true;
    const bool f_in = RT ? tma_in : ((MODE & EM_IN_MUL) != 0);
    const int f_bias = RT ? p.bias_mode : ((MODE & EM_BIAS) ? 1 : 0);
    const bool f_Z = RT ? (p.Z != nullptr) : ((MODE & EM_GELU_G) != 0);
    const bool f_zgrad = RT ? (p.z_grad != 0) : true;
    const int f_act = RT ? p.act : ((MODE & EM_GELU_G) ? (int)ACT_GELU : 0);
    const int f_dact = RT ? p.dact : ((MODE & EM_IN_MUL) ? 3 : 0);
    const bool f_colsum = RT ? (p.colsum != nullptr) : ((MODE & EM_COLSUM) != 0);
    for (int t = blockIdx.x; t < P.total_tiles; t += gridDim.x) {
      int tm, tn, z;
      tile_coords(P, t, tm, tn, z);
      const int sk = z % p.splitk, bz = z / p.splitk;
      const int kb0 = sk * P.kchunk_blocks, kb1 = min(P.kblocks, kb0 + P.kchunk_blocks);
      const int m_base = tm * BM + quad * 32;
      const int m = m_base + lane;
      const bool mval = m < p.M;
      const long long ro = off2(p.c.b, bz) + off2(p.c.r, mval ? m : 0);
      const float bias_m = (p.bias_mode == 2 && mval) ? __ldg(p.bias + m) : 0.f;
      int cc5[5];
      if (f_tma) opnd_coords(P.c, m_base, bz, cc5);
      if (RT && tma_in && (P.l2_prefetch & 4)) {
        // pull the input boxes of this warp's NEXT tile into L2 now: the bulk load one slice ahead then has L2 latency
        const int tt = t + (int)gridDim.x;
        if (tt < P.total_tiles) {
          int tm2, tn2, z2;
          tile_coords(P, tt, tm2, tn2, z2);
          const int mb = tm2 * BM + quad * 32;
          if (mb < p.M) {
            int c5[5];
            opnd_coords(P.c, mb, z2 / p.splitk, c5);
#pragma unroll
            for (int ci = 0; ci < CPW; ++ci) {
              const int nn = tn2 * BN + (half * CPW + ci) * 32;
              if (nn < p.N) {
                opnd_set_k(P.c.k_d, P.c.dk0, P.c.dk1, nn, c5);
                if (lane == 0) tma_prefetch_l2(&mapI, P.c.ndim, c5);
              }
            }
          }
        }
      }

      // ---- one 32-column slice of this warp's 32 rows: v[j] = accumulator of (row m, column n0 + j)
      auto emit = [&](float (&v)[32], int n0, int ci) {
        if (RT && (P.dbg == 5 || (P.skip & 1))) return;
        if (f_tma) {
          // TMA epilogue: rows are staged in the swizzled box and leave (or arrive) as one bulk tensor copy
          if (m_base >= p.M) return;  // warp-uniform
          TCP_T0(h0);
          opnd_set_k(P.c.k_d, P.c.dk0, P.c.dk1, n0, cc5);
          TCP_ADD(g1, h0);
          TCP_T0(h1);
          if (f_in) {
            if (pf_valid && (pf_t != t || pf_n0 != n0)) {  // cannot happen with the fixed slice order; drain defensively
              mbar_wait(inbar, in_phase);
              in_phase ^= 1;
              pf_valid = false;
            }
            if (!pf_valid) {
              __syncwarp();
              if (lane == 0) {
                mbar_arrive_expect_tx(inbar, 4096);
                tma_load(stg1, &mapI, inbar, P.c.ndim, cc5);
              }
            }
            pf_valid = false;
          }
          if (RT) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] *= p.alpha;
          }
          if (f_bias == 1) {
            if (!RT || (n0 + 32 <= p.N && P.c_vec)) {
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + n0) + q);
                v[4 * q] += b4.x; v[4 * q + 1] += b4.y; v[4 * q + 2] += b4.z; v[4 * q + 3] += b4.w;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] += (n0 + j < p.N) ? __ldg(p.bias + n0 + j) : 0.f;
            }
          } else if (f_bias == 2) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] += bias_m;
          }
          TCP_ADD(g2, h1);
          if (f_in) {
            TCP_T0(i0);
            mbar_wait(inbar, in_phase);
            TCP_ADD(ei, i0);
            in_phase ^= 1;
            float u[32];
            const uint32_t row1 = stg1 + lane * 128;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 x = lds128(row1 + ((q ^ sw) << 4));
              u[4 * q] = x.x; u[4 * q + 1] = x.y; u[4 * q + 2] = x.z; u[4 * q + 3] = x.w;
            }
            __syncwarp();  // every lane has read its row: box 1 may be refilled
            // prefetch the input of the slice this warp emits next (same tile, or the first slice of its next tile)
            if (ci + 1 < CPW && n0 + 32 < p.N) {
              int c5[5];
#pragma unroll
              for (int i = 0; i < 5; ++i) c5[i] = cc5[i];
              opnd_set_k(P.c.k_d, P.c.dk0, P.c.dk1, n0 + 32, c5);
              if (lane == 0) {
                mbar_arrive_expect_tx(inbar, 4096);
                tma_load(stg1, &mapI, inbar, P.c.ndim, c5);
              }
              pf_valid = true;
              pf_t = t;
              pf_n0 = n0 + 32;
            } else {
              in_issue(t + (int)gridDim.x, 0);
            }
            if (f_dact) {
              dact32_mul(f_dact, v, u);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] += u[j];
            }
          }
          TCP_T0(h4);
          bool acted = false;
          // box of the C slice: box 0, or alternating when nothing else uses box 1
          const bool alt = !f_in && !f_Z && (flip & 1);
          flip ^= 1;
          const uint32_t boxc = alt ? stg1 : stg;
          if (f_Z) {
            const bool zb1 = !f_in;  // Z rides in box 1 unless the input loads own it
            const uint32_t boxz = zb1 ? stg1 : stg;
            box_free(zb1 ? last1 : last0);
            if (f_zgrad) {
              float w[32];
              if (RT && P.dbg == 2) {
#pragma unroll
                for (int j = 0; j < 32; ++j) w[j] = v[j];
              } else if (f_act == ACT_GELU) {
#pragma unroll
                for (int j = 0; j < 32; j += 2) gelu_both2(v[j], v[j + 1], v[j], v[j + 1], w[j], w[j + 1]);
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                  w[j] = dact_f(f_act, v[j]);
                  v[j] = act_f(f_act, v[j]);
                }
              }
              acted = true;
              stage_write(boxz, w);
            } else {
              stage_write(boxz, v);
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store(boxz, &mapZ, P.c.ndim, cc5, false);
              bulk_commit();
            }
            ++ncommit;
            if (zb1) last1 = ncommit;
            else last0 = ncommit;
          }
          TCP_ADD(g4, h4);
          TCP_T0(h3);
          if (!acted) act32(f_act, v);
          TCP_ADD(g3, h3);
          box_free(alt ? last1 : last0);
          TCP_T0(s2);
          stage_write(boxc, v);
          fence_proxy_async_smem();
          __syncwarp();
          TCP_ADD(e2, s2);
          TCP_T0(s3);
          if (lane == 0) {
            if (!RT || P.dbg != 4) tma_store(boxc, &mapC, P.c.ndim, cc5, P.epi_reduce != 0);
            bulk_commit();
          }
          TCP_ADD(e3, s3);
          ++ncommit;
          if (alt) last1 = ncommit;
          else last0 = ncommit;
          if (f_colsum) {
            // bias gradient of the producing layer: column sums of the rows just staged (valid rows only)
            const int rows = min(32, p.M - m_base);
            float cs = 0.f;
            const uint32_t cb = boxc + (((uint32_t)lane & 3u) << 2), chk = (uint32_t)lane >> 2;
            auto ldcol = [&](int r) {
              float x;
              asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(cb + r * 128 + ((chk ^ ((uint32_t)r & 7u)) << 4)));
              return x;
            };
            if (rows == 32) {
              float c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
#pragma unroll
              for (int r = 0; r < 32; r += 4) {
                c0 += ldcol(r);
                c1 += ldcol(r + 1);
                c2 += ldcol(r + 2);
                c3 += ldcol(r + 3);
              }
              cs = (c0 + c1) + (c2 + c3);
            } else {
              for (int r = 0; r < rows; ++r) cs += ldcol(r);
            }
            // a CTA usually keeps its column block from tile to tile: accumulate in a register, flush on change / at the end
#pragma unroll
            for (int i = 0; i < CPW; ++i) {
              if (i == ci) {
                if (cs_n0[i] != n0) {
                  if (cs_n0[i] >= 0 && cs_n0[i] + lane < p.N) atomicAdd(p.colsum + cs_n0[i] + lane, cs_acc[i]);
                  cs_acc[i] = 0.f;
                  cs_n0[i] = n0;
                }
                cs_acc[i] += cs;
              }
            }
          }
          return;
        }
        if (!RT) return;
        if (col_contig) {
          // transpose through the staging box so that global accesses run along the contiguous columns
          stage_write(stg, v);
          __syncwarp();
          const long long coff = off2(p.c.c, n0);
          const bool aligned = !mval || (((ro + coff) & 3) == 0);
          const bool vec = P.c_vec && (n0 + 32 <= p.N) && __all_sync(0xffffffffu, aligned);
          if (vec) {
            const int rs = lane >> 3, c4 = (lane & 7) * 4;
            float4 bn = make_float4(0.f, 0.f, 0.f, 0.f);
            if (p.bias_mode == 1) bn = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + c4));
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int r = i * 4 + rs;
              const long long ror = __shfl_sync(0xffffffffu, ro, r);
              const float bm = __shfl_sync(0xffffffffu, bias_m, r);
              const bool mv = (m_base + r) < p.M;
              float4 x = lds128(stg + r * 128 + ((((uint32_t)lane & 7u) ^ ((uint32_t)r & 7u)) << 4));
              if (!mv) continue;
              const long long o = ror + coff + c4;
              x.x *= p.alpha; x.y *= p.alpha; x.z *= p.alpha; x.w *= p.alpha;
              if (p.splitk > 1) {
                atomicAdd(p.C + o, x.x); atomicAdd(p.C + o + 1, x.y); atomicAdd(p.C + o + 2, x.z); atomicAdd(p.C + o + 3, x.w);
                continue;
              }
              if (p.bias_mode == 1) { x.x += bn.x; x.y += bn.y; x.z += bn.z; x.w += bn.w; }
              else if (p.bias_mode == 2) { x.x += bm; x.y += bm; x.z += bm; x.w += bm; }
              if (p.beta) {
                const float4 cv = *reinterpret_cast<const float4*>(p.C + o);
                x.x += cv.x; x.y += cv.y; x.z += cv.z; x.w += cv.w;
              }
              if (p.Z) *reinterpret_cast<float4*>(p.Z + o) = p.z_grad ? dact4(p.act, x) : x;
              if (p.act) x = act4(p.act, x);
              if (p.dact) {
                const float4 av = __ldg(reinterpret_cast<const float4*>(p.aux + o));
                const float4 dv = (p.dact == 3) ? av : dact4(p.dact, av);
                x.x *= dv.x; x.y *= dv.y; x.z *= dv.z; x.w *= dv.w;
              }
              if (P.dbg != 4) *reinterpret_cast<float4*>(p.C + o) = x;
              if (p.colsum) {
                atomicAdd(p.colsum + n0 + c4, x.x); atomicAdd(p.colsum + n0 + c4 + 1, x.y);
                atomicAdd(p.colsum + n0 + c4 + 2, x.z); atomicAdd(p.colsum + n0 + c4 + 3, x.w);
              }
            }
          } else {
            const int n = n0 + lane;
            const bool nval = n < p.N;
            const long long co = off2(p.c.c, nval ? n : 0);
            const float bn = (p.bias_mode == 1 && nval) ? __ldg(p.bias + n) : 0.f;
#pragma unroll 4
            for (int r = 0; r < 32; ++r) {
              const long long ror = __shfl_sync(0xffffffffu, ro, r);
              const float bm = __shfl_sync(0xffffffffu, bias_m, r);
              float x;
              asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(stg + r * 128 + (((((uint32_t)lane >> 2) ^ ((uint32_t)r & 7u)) << 4) | (((uint32_t)lane & 3u) << 2))));
              if (!nval || (m_base + r) >= p.M) continue;
              const long long o = ror + co;
              x *= p.alpha;
              if (p.splitk > 1) {
                atomicAdd(p.C + o, x);
                continue;
              }
              if (p.bias_mode == 1) x += bn;
              else if (p.bias_mode == 2) x += bm;
              if (p.beta) x += p.C[o];
              if (p.Z) p.Z[o] = p.z_grad ? dact_f(p.act, x) : x;
              if (p.act) x = act_f(p.act, x);
              if (p.dact) x *= (p.dact == 3) ? __ldg(p.aux + o) : dact_f(p.dact, __ldg(p.aux + o));
              p.C[o] = x;
              if (p.colsum) atomicAdd(p.colsum + n, x);
            }
          }
          __syncwarp();
        } else if (mval) {
          // rows are the contiguous (or an arbitrary) direction: one row per lane is already the coalesced order
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int n = n0 + j;
            if (n >= p.N) break;
            const long long o = ro + off2(p.c.c, n);
            float x = p.alpha * v[j];
            if (p.splitk > 1) {
              atomicAdd(p.C + o, x);
              continue;
            }
            if (p.bias_mode == 1) x += __ldg(p.bias + n);
            else if (p.bias_mode == 2) x += bias_m;
            if (p.beta) x += p.C[o];
            if (p.Z) p.Z[o] = p.z_grad ? dact_f(p.act, x) : x;
            if (p.act) x = act_f(p.act, x);
            if (p.dact) x *= (p.dact == 3) ? __ldg(p.aux + o) : dact_f(p.dact, __ldg(p.aux + o));
            p.C[o] = x;
            if (p.colsum) atomicAdd(p.colsum + n, x);
          }
        }
      };

      if (!CH) {
        const int buf = cit & 1;
        TCP_T0(a0);
        mbar_wait(acc_full(buf), (cit >> 1) & 1);
        TCP_ADD(ea, a0);
        tc_fence_after_sync();
        const uint32_t tq = tmem_base + ((uint32_t)(quad * 32) << 16);
        TCP_T0(f0);
        if (SPLIT) {
          // fold the small block into big (in TMEM) and hand it back, so the next tile's main loop starts at once
#pragma unroll 1
          for (int ci = 0; ci < CPW; ++ci) {
            const int cc = half * CPW + ci;
            uint32_t r[32], r2[32];
            tmem_ld32(tq + buf * C::acc_cols + cc * 32, r);
            tmem_ld32(tq + C::small_col + cc * 32, r2);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
            tmem_st32(tq + buf * C::acc_cols + cc * 32, r);
          }
          tmem_st_wait();
          tc_fence_before_sync();
          __syncwarp();
          if (lane == 0) mbar_arrive(small_empty);
        }
        TCP_ADD(ef, f0);
        TCP_T0(e0);
#pragma unroll 1
        for (int ci = 0; ci < CPW; ++ci) {
          const int cc = half * CPW + ci;
          const int n0 = tn * BN + cc * 32;
          if (n0 >= p.N) break;  // warp-uniform
          uint32_t r[32];
          float v[32];
          TCP_T0(s1);
          tmem_ld32(tq + buf * C::acc_cols + cc * 32, r);
          tmem_ld_wait();
          TCP_ADD(e1, s1);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          emit(v, n0, ci);
        }
        TCP_ADD(ee, e0);
#ifdef RPDE_TC_PROFILE
        ++en;
#endif
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty(buf));
        ++cit;
      } else {
        float acc[CPW * 32];
#pragma unroll
        for (int j = 0; j < CPW * 32; ++j) acc[j] = 0.f;
        for (int c0 = kb0; c0 < kb1; c0 += P.kc, ++cit) {
          const int buf = cit & 1;
          mbar_wait(acc_full(buf), (cit >> 1) & 1);
          tc_fence_after_sync();
#pragma unroll
          for (int ci = 0; ci < CPW; ++ci) {
            uint32_t r[32];
            const uint32_t tq = tmem_base + ((uint32_t)(quad * 32) << 16);
            const uint32_t ta = tq + buf * C::acc_cols + (half * CPW + ci) * 32;
            tmem_ld32(ta, r);
            if (SPLIT) {
              uint32_t r2[32];
              tmem_ld32(tq + C::small_col + (half * CPW + ci) * 32, r2);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) acc[ci * 32 + j] += __uint_as_float(r[j]) + __uint_as_float(r2[j]);
            } else {
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) acc[ci * 32 + j] += __uint_as_float(r[j]);
            }
          }
          tc_fence_before_sync();
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(acc_empty(buf));
            if (SPLIT) mbar_arrive(small_empty);
          }
        }
#pragma unroll
        for (int ci = 0; ci < CPW; ++ci) {
          const int n0 = tn * BN + (half * CPW + ci) * 32;
          if (n0 < p.N) {
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = acc[ci * 32 + j];
            emit(v, n0, ci);
          }
        }
      }
    }
    };  // run_tiles
    if constexpr (CH) {  // chunked chains emit once per many accumulator hand-overs: one instantiation is enough
      run_tiles(std::integral_constant<int, -1>{});
    } else
    switch (P.epi_mode) {
      case 0: run_tiles(std::integral_constant<int, 0>{}); break;
      case EM_BIAS: run_tiles(std::integral_constant<int, EM_BIAS>{}); break;
      case EM_GELU_G: run_tiles(std::integral_constant<int, EM_GELU_G>{}); break;
      case EM_BIAS | EM_GELU_G: run_tiles(std::integral_constant<int, EM_BIAS | EM_GELU_G>{}); break;
      case EM_IN_MUL: run_tiles(std::integral_constant<int, EM_IN_MUL>{}); break;
      case EM_IN_MUL | EM_COLSUM: run_tiles(std::integral_constant<int, EM_IN_MUL | EM_COLSUM>{}); break;
      default: run_tiles(std::integral_constant<int, -1>{}); break;
    }
    if (pf_valid) mbar_wait(inbar, in_phase);  // never leave a bulk load in flight
#ifdef RPDE_TC_PROFILE
    if (blockIdx.x == 0 && lane == 0 && (ew == 0 || ew == 7))
      printf("[tc epi %d] tiles %lld total %lld wait_acc_full %lld fold %lld emit %lld (wait_in %lld box_free %lld tmem_ld %lld stage %lld store %lld setk %lld alpha_bias %lld act %lld zblock %lld)\n", ew, en, clock64() - eall, ea, ef, ee, ei, eb, e1, e2, e3, g1, g2, g3, g4);
#endif
    if (p.colsum) {
#pragma unroll
      for (int i = 0; i < CPW; ++i)
        if (cs_n0[i] >= 0 && cs_n0[i] + lane < p.N) atomicAdd(p.colsum + cs_n0[i] + lane, cs_acc[i]);
    }
    if (lane == 0) bulk_wait0();  // the staging boxes must outlive the bulk stores that read them
    __syncwarp();
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc(tmem_base, C::tmem_cols);
  }
}

}  // namespace tc

// ================================================================================================
// Host side: GemmP index maps -> TMA tensor maps, tile/split-K choice, launch.
// ================================================================================================
namespace tc {

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

struct Lvl {
  long long size, stride;  // elements
};

// index map + extent -> one or two (size, stride) levels; *d = divisor of the two-level form
static int split_levels(const Idx2& m, long long extent, Lvl out[2], int* d) {
  if (m.d == INT_MAX || m.d >= extent) {
    out[0] = Lvl{extent, m.sl};
    *d = INT_MAX;
    return 1;
  }
  if (m.d == 1) {  // degenerate split: i / 1 * sh
    out[0] = Lvl{extent, m.sh};
    *d = INT_MAX;
    return 1;
  }
  if (extent % m.d) return -1;
  out[0] = Lvl{m.d, m.sl};
  out[1] = Lvl{extent / m.d, m.sh};
  *d = m.d;
  return 2;
}

// Builds the tensor map + device recipe of one operand.  Returns false when TMA cannot address it (-> CUDA-core engine).
static bool build_operand(const float* base, const Idx2& mn_map, long long MN, const Idx2& k_map, long long K, const Idx2& b_map,
                          long long BT, int tile_mn, CUtensorMap* map, TcOpnd* o) {
  EncodeTiledFn enc = encode_fn();
  if (!enc || (reinterpret_cast<uintptr_t>(base) & 15)) return false;
  Lvl lm[2], lk[2], lb[2];
  const int nm = split_levels(mn_map, MN, lm, &o->mn_d);
  const int nk = split_levels(k_map, K, lk, &o->k_d);
  const int nb = split_levels(b_map, BT, lb, &o->b_d);
  if (nm < 0 || nk < 0 || nb < 0) return false;
  if (lk[0].stride == 1 && (nk == 1 || lk[0].size % BK == 0)) o->major = 0;
  else if (lm[0].stride == 1 && (nm == 1 || lm[0].size % 32 == 0)) o->major = 1;
  else return false;

  struct Dim {
    long long size, stride;
    int box;
    signed char* slot;
  };
  Dim dims[8];
  int nd = 0;
  o->dm0 = o->dm1 = o->dk0 = o->dk1 = o->db0 = o->db1 = o->dg = -1;
  o->grouped = 0;
  if (o->major == 0) {
    int bm0 = tile_mn, bm1 = 1;
    if (nm == 2) {
      const long long d = lm[0].size;
      if (d % tile_mn == 0) { bm0 = tile_mn; bm1 = 1; }
      else if (tile_mn % d == 0) { bm0 = (int)d; bm1 = (int)(tile_mn / d); }
      else return false;
    }
    dims[nd++] = Dim{lk[0].size, lk[0].stride, BK, &o->dk0};
    if (nk == 2) dims[nd++] = Dim{lk[1].size, lk[1].stride, 1, &o->dk1};
    dims[nd++] = Dim{lm[0].size, lm[0].stride, bm0, &o->dm0};
    if (nm == 2) dims[nd++] = Dim{lm[1].size, lm[1].stride, bm1, &o->dm1};
    o->boxes = 1;
    o->box_mn = tile_mn;
    o->box_bytes = tile_mn * BK * 4;
  } else {
    if (nk == 2 && lk[0].size % BK) return false;
    // preferred: the 32-element MN groups become a tensor dimension, so one box brings the whole tile
    const long long lo_ext = lm[0].size;
    const bool can_group = (lo_ext % 32 == 0) && (nm == 1 || lo_ext % tile_mn == 0 || tile_mn % lo_ext == 0);
    const int nb_live = (lb[0].size > 1 && lb[0].stride != 0 ? 1 : 0) + (nb == 2 && lb[1].size > 1 && lb[1].stride != 0 ? 1 : 0);
    const int rank_grouped = 1 + nk + 1 + (nm == 2 ? 1 : 0) + nb_live;
    if (can_group && rank_grouped <= 5) {
      int bg = tile_mn / 32, bm1 = 1;
      if (nm == 2 && lo_ext < tile_mn) { bg = (int)(lo_ext / 32); bm1 = (int)(tile_mn / lo_ext); }
      static signed char dummy;
      dims[nd++] = Dim{32, 1, 32, &dummy};
      dims[nd++] = Dim{lk[0].size, lk[0].stride, BK, &o->dk0};
      if (nk == 2) dims[nd++] = Dim{lk[1].size, lk[1].stride, 1, &o->dk1};
      dims[nd++] = Dim{lo_ext / 32, 32, bg, &o->dg};
      if (nm == 2) dims[nd++] = Dim{lm[1].size, lm[1].stride, bm1, &o->dm1};
      o->grouped = 1;
      o->boxes = 1;
      o->box_mn = tile_mn;
      o->box_bytes = tile_mn * BK * 4;
    } else {
      dims[nd++] = Dim{lm[0].size, lm[0].stride, 32, &o->dm0};
      dims[nd++] = Dim{lk[0].size, lk[0].stride, BK, &o->dk0};
      if (nk == 2) dims[nd++] = Dim{lk[1].size, lk[1].stride, 1, &o->dk1};
      if (nm == 2) dims[nd++] = Dim{lm[1].size, lm[1].stride, 1, &o->dm1};
      o->boxes = tile_mn / 32;
      o->box_mn = 32;
      o->box_bytes = 32 * BK * 4;
    }
  }
  dims[nd++] = Dim{lb[0].size, lb[0].stride, 1, &o->db0};
  if (nb == 2) dims[nd++] = Dim{lb[1].size, lb[1].stride, 1, &o->db1};

  cuuint64_t gdim[5], gstr[5];
  cuuint32_t box[5], estr[5] = {1, 1, 1, 1, 1};
  int r = 0;
  for (int i = 0; i < nd; ++i) {
    Dim& dm = dims[i];
    if (i > 0 && dm.box == 1 && (dm.size == 1 || dm.stride == 0)) continue;  // coordinate contributes nothing
    if (i > 0) {
      if (dm.size == 1 && (dm.stride <= 0 || (dm.stride & 3))) dm.stride = 1LL << 20;  // never dereferenced beyond index 0
      if (dm.stride <= 0 || (dm.stride & 3)) return false;                              // TMA: 16-byte multiples
    }
    if (r == 5 || dm.size <= 0 || dm.size > 0xffffffffLL) return false;
    gdim[r] = (cuuint64_t)dm.size;
    gstr[r] = (cuuint64_t)dm.stride * 4;
    box[r] = (cuuint32_t)dm.box;
    *dm.slot = (signed char)r;
    ++r;
  }
  o->ndim = r;
  if (r < 2) return false;
  const CUresult rc = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)r, const_cast<float*>(base), gdim, gstr + 1, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, o->major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return rc == CUDA_SUCCESS;
}

struct EpiMaps {
  CUtensorMap c, z, i;
};

template <int BN, bool SPLIT, bool CH>
static cudaError_t launch_cfg(const TcParams& P, const CUtensorMap& ma, const CUtensorMap& mb, const EpiMaps& em, int sms,
                              cudaStream_t s) {
  using C = Cfg<BN, SPLIT>;
  // cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: one flag per device, so a single process that
  // drives several GPUs (nn.DataParallel: one thread per device) configures the kernel on each of them
  static std::atomic<unsigned char> attr_done[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !attr_done[dev].load(std::memory_order_acquire)) {
    const cudaError_t attr_rc =
        cudaFuncSetAttribute(gemm_tc_kernel<BN, SPLIT, CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::smem_bytes);
    if (attr_rc != cudaSuccess) return attr_rc;
    if (dev >= 0 && dev < 64) attr_done[dev].store(1, std::memory_order_release);
  }
  const int grid = P.total_tiles < sms ? P.total_tiles : sms;
  static const bool pdl = !(getenv("RPDE_TC_PDL") && atoi(getenv("RPDE_TC_PDL")) == 0);
  if (pdl) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = C::smem_bytes;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, gemm_tc_kernel<BN, SPLIT, CH>, P, ma, mb, em.c, em.z, em.i);
  }
  gemm_tc_kernel<BN, SPLIT, CH><<<grid, kThreads, C::smem_bytes, s>>>(P, ma, mb, em.c, em.z, em.i);
  return cudaGetLastError();
}

}  // namespace tc

// Tries the tensor-core engine; *launched = false means the shape is not TMA-addressable (caller falls back).
static cudaError_t gemm_tc_try_launch(const GemmP& g, int prec, cudaStream_t s, bool* launched) {
  using namespace tc;
  *launched = false;
  if (g.M <= 0 || g.N <= 0 || g.K <= 0 || g.batch <= 0) return cudaSuccess;
  // launch-bound sizes: the CUDA-core kernel starts faster than a 231 KB-smem persistent CTA with a TMEM allocation
  static const long long min_macs = getenv("RPDE_TC_MIN_MACS") ? atoll(getenv("RPDE_TC_MIN_MACS")) : (1LL << 18);
  if ((long long)g.M * g.N * g.K * g.batch < min_macs) return cudaSuccess;
  TcParams P;
  P.g = g;
  {
    const char* e = getenv("RPDE_TC_DBG");
    P.dbg = e ? atoi(e) : 0;
    const char* sk = getenv("RPDE_TC_SKIP");
    P.skip = sk ? atoi(sk) : 0;
  }
  CUtensorMap ma, mb;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  static const int kc_env = getenv("RPDE_TC_KC") ? atoi(getenv("RPDE_TC_KC")) : 0;
  // k-blocks (x32) per TMEM accumulation chain: 8 = 256 k = 32 UMMAs into the hi*hi accumulator (~7e-7 relative
  // truncation); longer reductions are chunked and summed in registers.
  const int kc = kc_env > 0 ? kc_env : 8;
  const int kblocks = (g.K + BK - 1) / BK;
  const int tiles_m = (g.M + BM - 1) / BM;
  // split-K (only when the caller allows atomically accumulated output)
  auto choose_split = [&](int BN) {
    int splitk = g.splitk;
    if (splitk == 0) {
      const long long ctas = (long long)tiles_m * ((g.N + BN - 1) / BN) * g.batch;
      long long want = (sms + ctas - 1) / ctas;
      const long long maxs = (kblocks + 1) / 2;  // at least 64 k per split
      if (want > maxs) want = maxs;
      if (want < 1) want = 1;
      splitk = (int)want;
    }
    return splitk;
  };
  const bool split3 = (prec == 1);
  // 3xTF32 keeps two accumulator column blocks per tile (and two tiles in flight): 4 * BN <= 512 TMEM columns
  int BN = (g.N > 128 && !split3) ? 256 : (g.N > 64 ? 128 : 64);
  // long reductions are chunked and summed in registers: that variant holds BN accumulators per thread
  {
    const int sk = choose_split(BN);
    const int per_tile = (kblocks + sk - 1) / sk;
    if (per_tile > kc && BN > 128) BN = 128;
  }
  bool ok = false;
  for (; BN >= 64; BN >>= 1) {
    ok = build_operand(g.A, g.a.r, g.M, g.a.c, g.K, g.a.b, g.batch, BM, &ma, &P.a) &&
         build_operand(g.B, g.b.c, g.N, g.b.r, g.K, g.b.b, g.batch, BN, &mb, &P.b);
    if (ok) break;
  }
  if (!ok) return cudaSuccess;
  P.tiles_m = tiles_m;
  P.tiles_n = (g.N + BN - 1) / BN;
  P.kblocks = kblocks;
  int splitk = choose_split(BN);
  P.kchunk_blocks = (P.kblocks + splitk - 1) / splitk;
  splitk = (P.kblocks + P.kchunk_blocks - 1) / P.kchunk_blocks;
  P.g.splitk = splitk;
  const bool chunked = P.kchunk_blocks > kc;
  P.kc = chunked ? kc : P.kchunk_blocks;
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  P.c_vec = al16(g.C) && al16(g.Z) && al16(g.aux) && al16(g.bias);
  const long long total = (long long)P.tiles_m * P.tiles_n * g.batch * splitk;
  if (total > 0x7fffffffLL) return cudaSuccess;
  P.total_tiles = (int)total;
  P.n_fast = P.tiles_n <= P.tiles_m;
  {
    static const int pf_env = getenv("RPDE_TC_L2PF") ? atoi(getenv("RPDE_TC_L2PF")) : 4;  // measured: the extra TMA requests cost more than the latency they hide
    P.l2_prefetch = pf_env;
  }
  // TMA epilogue: the output (and Z / aux / the old C) addressed as 32 x 32 boxes of rows x contiguous columns
  EpiMaps em;
  memset(&em, 0, sizeof(em));
  P.epi_tma = P.epi_reduce = 0;
  memset(&P.c, 0, sizeof(P.c));
  static const bool no_epi_tma = getenv("RPDE_TC_NO_EPI_TMA") != nullptr;
  if (splitk == 1 && !no_epi_tma && P.dbg != 3) {
    TcOpnd co, tmp;
    bool okc = build_operand(g.C, g.c.r, g.M, g.c.c, g.N, g.c.b, g.batch, 32, &em.c, &co) && co.major == 0 && co.boxes == 1;
    if (okc && g.Z) okc = build_operand(g.Z, g.c.r, g.M, g.c.c, g.N, g.c.b, g.batch, 32, &em.z, &tmp) && tmp.major == 0;
    if (okc && g.dact) {
      okc = build_operand(g.aux, g.c.r, g.M, g.c.c, g.N, g.c.b, g.batch, 32, &em.i, &tmp) && tmp.major == 0;
    } else if (okc && g.beta) {
      if (g.act || g.Z) em.i = em.c;
      else P.epi_reduce = 1;
    }
    if (okc) {
      P.epi_tma = 1;
      P.c = co;
    } else {
      P.epi_reduce = 0;
    }
  }
  // compiled epilogue modes (see EM_*): everything else reads its flags at run time
  P.epi_mode = -1;
  static const bool no_modes = getenv("RPDE_TC_NO_EPI_MODES") != nullptr;
  if (P.epi_tma && !no_modes && g.alpha == 1.f && P.dbg == 0 && P.skip == 0 && (g.N % 32) == 0 &&
      (g.bias_mode == 0 || (g.bias_mode == 1 && P.c_vec))) {
    const int bbit = g.bias_mode == 1 ? EM_BIAS : 0;
    if (!g.dact && !g.act && !g.Z && !g.colsum && (!g.beta || P.epi_reduce)) P.epi_mode = bbit;
    else if (g.act == ACT_GELU && g.Z && g.z_grad && !g.dact && !g.beta && !g.colsum) P.epi_mode = EM_GELU_G | bbit;
    else if (g.dact == 3 && !g.act && !g.Z && !g.beta && !bbit) P.epi_mode = EM_IN_MUL | (g.colsum ? EM_COLSUM : 0);
  }
  cudaError_t rc;
  if (chunked) {
    if (BN == 128) rc = split3 ? launch_cfg<128, true, true>(P, ma, mb, em, sms, s) : launch_cfg<128, false, true>(P, ma, mb, em, sms, s);
    else rc = split3 ? launch_cfg<64, true, true>(P, ma, mb, em, sms, s) : launch_cfg<64, false, true>(P, ma, mb, em, sms, s);
  } else if (BN == 256) {
    rc = launch_cfg<256, false, false>(P, ma, mb, em, sms, s);
  } else if (BN == 128) {
    rc = split3 ? launch_cfg<128, true, false>(P, ma, mb, em, sms, s) : launch_cfg<128, false, false>(P, ma, mb, em, sms, s);
  } else {
    rc = split3 ? launch_cfg<64, true, false>(P, ma, mb, em, sms, s) : launch_cfg<64, false, false>(P, ma, mb, em, sms, s);
  }
  *launched = (rc == cudaSuccess);
  if (P.dbg == 9)
    fprintf(stderr, "[rpde tc] M=%d N=%d K=%d batch=%d BN=%d splitk=%d majorA=%d majorB=%d tiles=%d beta=%d act=%d dact=%d Z=%d epi_tma=%d red=%d ch=%d\n",
            g.M, g.N, g.K, g.batch, BN, splitk, P.a.major, P.b.major, P.total_tiles, g.beta, g.act, g.dact, g.Z != nullptr, P.epi_tma,
            P.epi_reduce, (int)chunked);
  return rc;
}

}  // namespace rpde
