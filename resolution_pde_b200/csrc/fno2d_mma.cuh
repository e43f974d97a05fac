// FNOBlock2d W-axis kernels on the warp-level tensor path (mma.sync m16n8k8 TF32 with the 3xTF32 split: fp32 parity).
//
// These passes are HBM-bound (reduction lengths 24..64 per output), but as fp32-FMA register tiles they were instruction-
// bound (fno1d_synth_bypass_kernel<128, 32>: 266 us against a 62 us HBM floor at 16 x 32 x 256^2).  One m16n8k8 MMA replaces
// 32 FMAs per lane; with three MMAs per product the issue slots drop ~5x and the pass becomes a stream over x / y / z.
//
// fno2d_synth_mma_kernel (spectral_convolution.py:97 irfft2's last axis + fno_blocks.py:69-71):
//   y[b,o,h,:] = act( sum_r Z[r][(b,h)][o] * tab[r][:]  +  sum_i wb[o,i] * x[b,i,h,:]  +  bb[o] ),   z = the pre-activation
// as D[o][n] = A[o][k] * B[k][n] per image row (b, h): k runs over the 2*m2 spectral rows, then over the input channels.
//   * persistent CTAs of eight warps; a warp owns 32 consecutive grid points of one image row (256 / W rows per iteration)
//   * resident in shared memory, already split into TF32 hi / lo planes: the synthesis table (2 m2 x W) and the bypass weight
//   * per image row: Z (2 m2 x O) staged hi / lo through a two-deep ring; x goes global -> registers as float4 (each element
//     is used by exactly one warp), split in registers, prefetched one iteration ahead
//   * the MMA column <-> grid point permutation (column c of n-tile j <-> point 4*pi(c) + j, pi(c) = (c >> 1) + 4 * (c & 1))
//     makes every x load, table load and y / z store a float4 on 64..128 contiguous bytes per channel row
// adjoint = 1 (backward): dx = analysis adjoint + wb^T dpre through the same code (table rows as stored, wb read transposed).
#pragma once
#include "common.cuh"

namespace rpde {

constexpr int F2M_LDA = 40;   // A-operand planes [k][o]: stride = 8 mod 32 -> the (t, g) fragment loads hit 32 distinct banks
constexpr int F2M_MAXC = 32;  // channels (both sides) covered by the register tiles
constexpr int F2M_MAXR = 32;  // spectral rows 2 * m2, rounded up to a multiple of 8

__device__ __forceinline__ uint32_t tf32_hi(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return r;
}
// resident operands (split once): round to nearest
__device__ __forceinline__ void tf32_split_rn(float v, uint32_t& hi, uint32_t& lo) {
  hi = tf32_hi(v);
  lo = __float_as_uint(v - __uint_as_float(hi));
}
// streamed operands: hi = v rounded to 10 mantissa bits, half away from zero, by an integer add and a mask (cvt.rna.tf32
// costs four instructions on sm_100: compare against inf, add, mask, select), lo = v - hi exactly.  Rounding, not truncating:
// a truncated hi leaves lo with the sign of v, and the tensor core's own truncation of lo then biases every product the same
// way -- measured 2.4e-5 on a cancelling 262144-term weight-gradient sum; with a rounded hi the residual errors have random sign.
__device__ __forceinline__ void tf32_split(float v, uint32_t& hi, uint32_t& lo) {
  hi = (__float_as_uint(v) + 0x1000u) & 0xffffe000u;
  lo = __float_as_uint(v - __uint_as_float(hi));
}
// the same, pinned where it is written: inside a loop over output chunks the compiler would otherwise hoist the split of
// loop-invariant operands and keep hi and lo planes live (2x the registers of the raw operand)
__device__ __forceinline__ void tf32_split_here(float v, uint32_t& hi, uint32_t& lo) {
  asm volatile("and.b32 %0, %1, 0xffffe000;" : "=r"(hi) : "r"(__float_as_uint(v) + 0x1000u));
  lo = __float_as_uint(v - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// d += a * b to fp32 accuracy: the two cross terms first, the leading term last
__device__ __forceinline__ void mma_3xtf32(float (&d)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4], uint32_t bh0,
                                           uint32_t bh1, uint32_t bl0, uint32_t bl1) {
  mma_tf32(d, al, bh0, bh1);
  mma_tf32(d, ah, bl0, bl1);
  mma_tf32(d, ah, bh0, bh1);
}

__host__ __device__ constexpr size_t f2m_synth_smem(int W, int RP) {
  return ((size_t)2 * (RP / 2) * (2 * W + 4)            // table hi / lo, rows t and t + 4 of a k-step interleaved per point
          + (size_t)2 * F2M_MAXC * F2M_LDA              // bypass weight hi / lo
          + (size_t)2 * 2 * (W >= 256 ? 1 : 256 / W) * RP * F2M_LDA   // Z ring: 2 stages x hi / lo x rows per iteration
          + F2M_MAXC) * sizeof(float);                  // bias
}

// Table layout in shared memory.  An HMMA takes its B operand from a register PAIR (k = t, k = t + 4) and a lane needs that
// pair for four points, so rows t and t + 4 of each k-step are interleaved per point: one float4 = the pairs of two points,
// loaded straight into the registers the MMA reads (assembling the pairs from two row-major float4 cost two MOVs per HMMA).
// Row (ks, t) holds, per 32-point warp segment, eight 8-float groups (four points x two rows); group pg sits at slot
// sigma(pg) so that the two point groups read in one shared-memory phase (lanes g = 2 q, 2 q + 1 -> groups q, 4 + q) are 16
// floats apart, and with a row stride of 4 mod 32 the eight lanes of a phase cover all 32 banks.
__device__ __forceinline__ int f2m_slot(int pg) { return pg < 4 ? (pg & 1) + 4 * (pg >> 1) : ((pg - 4) & 1) + 4 * ((pg - 4) >> 1) + 2; }

template <int W>
__global__ void __launch_bounds__(256, W <= 256 ? 2 : 1) fno2d_synth_mma_kernel(const float* __restrict__ x, const float* __restrict__ Z,
                                                                 const float* __restrict__ table, const float* __restrict__ wb,
                                                                 const float* __restrict__ bb, float* __restrict__ y,
                                                                 float* __restrict__ z, int Bn, int I, int O, int K, int act,
                                                                 int adjoint, int Hd) {
  static_assert(W == 64 || W == 128 || W == 256 || W == 512 || W == 1024, "one warp per 32 grid points, eight warps");
  constexpr int TW = W > 256 ? 256 : W;   // points per iteration of one image row (longer rows: W / 256 chunks, whole table resident)
  constexpr int CH = W / TW;
  constexpr int WPR = TW / 32;       // warps per image row (chunk)
  constexpr int RPI = 8 / WPR;       // image rows per iteration
  constexpr int LDP = 2 * W + 4;     // paired table rows: stride = 4 mod 32
  extern __shared__ __align__(16) float smem[];
  const int R = 2 * K, RP = (R + 7) & ~7, KS_T = RP / 8, KS_X = wb ? (I + 7) / 8 : 0;
  float* tabH = smem;                                   // [ks * 4 + t][LDP]
  float* tabL = tabH + (RP / 2) * LDP;
  float* wbH = tabL + (RP / 2) * LDP;
  float* wbL = wbH + F2M_MAXC * F2M_LDA;
  float* zring = wbL + F2M_MAXC * F2M_LDA;            // [stage][hi|lo][row][r][LDA]
  float* biasS = zring + 2 * 2 * RPI * RP * F2M_LDA;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int pg = (g >> 1) + 4 * (g & 1);               // pi(g): this lane's float4 group inside the warp's 32 points
  const int wrow = warp / WPR, nw = (warp % WPR) * 32;   // + 256 * chunk = the warp's first point

  // ---- resident operands -------------------------------------------------------------------------------------------
  for (int e = tid; e < RP * W; e += 256) {
    const int r = e / W, n = e - r * W;
    float v = 0.f;
    if (r < R) v = __ldg(table + (adjoint ? (long long)r : (long long)(r & 1) * K + (r >> 1)) * W + n);
    uint32_t hi, lo;
    tf32_split_rn(v, hi, lo);
    const int ks = r >> 3, tt = r & 3, half = (r >> 2) & 1;
    const int off = (ks * 4 + tt) * LDP + (n >> 5) * 64 + f2m_slot((n >> 2) & 7) * 8 + (n & 3) * 2 + half;
    tabH[off] = __uint_as_float(hi);
    tabL[off] = __uint_as_float(tf32_hi(__uint_as_float(lo)));
  }
  for (int e = tid; e < F2M_MAXC * F2M_MAXC; e += 256) {
    const int i = e / F2M_MAXC, o = e - i * F2M_MAXC;  // A[o][k = i]
    float v = 0.f;
    if (wb && i < I && o < O) v = __ldg(adjoint ? wb + (long long)i * O + o : wb + (long long)o * I + i);
    uint32_t hi, lo;
    tf32_split_rn(v, hi, lo);
    wbH[i * F2M_LDA + o] = __uint_as_float(hi);
    wbL[i * F2M_LDA + o] = __uint_as_float(tf32_hi(__uint_as_float(lo)));
  }
  if (tid < F2M_MAXC) biasS[tid] = (bb && tid < O) ? __ldg(bb + tid) : 0.f;

  const int ntiles = ((Bn + RPI - 1) / RPI) * CH;         // tile = (row group, chunk), chunk fastest
  constexpr int ZPT = (RPI * F2M_MAXR * F2M_MAXC + 255) / 256;   // Z elements per thread and iteration (upper bound)
  const int zcount = RPI * RP * F2M_MAXC;
  const long long chan = (long long)Hd * W;                      // elements between two channels of one image
  // per-lane table fragment base (row t of k-step 0, this lane's point group) and A fragment bases
  const int tab_off0 = t * LDP + (nw >> 5) * 64 + f2m_slot(pg) * 8;
  const int a_off = t * F2M_LDA + g;

  // this warp's x rows (two float4 per k-step): the address of channel t, point group pg of image row `row`
  auto x_ptr = [&](int row, int n0) {
    const int bh = row / Hd, hh = row - bh * Hd;
    return x + ((long long)bh * I * Hd + hh) * W + n0 + 4 * pg + (long long)t * chan;
  };
  auto load_x = [&](int row, int n0, float4 (&xb)[8]) {
    const float* p = x_ptr(row, n0);
    const bool ok = row < Bn;
#pragma unroll
    for (int q = 0; q < 8; ++q) {    // channel t + 4 q: k-step q >> 1, fragment row t (+ 4)
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ok && (q >> 1) < KS_X && t + 4 * q < I) v = __ldg(reinterpret_cast<const float4*>(p + 4 * q * chan));
      xb[q] = v;
    }
  };
  // the next iteration's x rows are only pulled into L2 (no registers held across the MMA loop; the eight lanes g = 0..7
  // of equal t share a 128-byte line, lane g = 0 asks for it)
  auto prefetch_x = [&](int row, int n0) {
    if (row >= Bn || g != 0) return;
    const float* p = x_ptr(row, n0);
#pragma unroll
    for (int q = 0; q < 8; ++q)
      if ((q >> 1) < KS_X && t + 4 * q < I) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 4 * q * chan));
  };
  auto load_z = [&](int tile, float (&zb)[ZPT]) {
#pragma unroll
    for (int q = 0; q < ZPT; ++q) {
      const int e = tid + 256 * q;
      const int o = e % F2M_MAXC, rr = e / F2M_MAXC;       // rr = rowslot * RP + r
      const int slot = RPI == 1 ? 0 : rr / RP, r = rr - slot * RP;
      const int row = (tile / CH) * RPI + slot;
      float v = 0.f;
      if (e < zcount && row < Bn && r < R && o < O) v = __ldg(Z + ((long long)r * Bn + row) * O + o);
      zb[q] = v;
    }
  };
  auto store_z = [&](int stage, const float (&zb)[ZPT]) {
    float* zh = zring + (size_t)stage * 2 * RPI * RP * F2M_LDA;
    float* zl = zh + RPI * RP * F2M_LDA;
#pragma unroll
    for (int q = 0; q < ZPT; ++q) {
      const int e = tid + 256 * q;
      if (e >= zcount) continue;
      const int o = e % F2M_MAXC, rr = e / F2M_MAXC;
      uint32_t hi, lo;
      tf32_split(zb[q], hi, lo);
      zh[rr * F2M_LDA + o] = __uint_as_float(hi);
      zl[rr * F2M_LDA + o] = __uint_as_float(lo);
    }
  };
  float zb[ZPT];
  int tile = blockIdx.x;
  if (tile < ntiles) {
    load_z(tile, zb);
    store_z(0, zb);
  }
  __syncthreads();

  int stage = 0;
  for (; tile < ntiles; tile += gridDim.x) {
    const int next = tile + gridDim.x;
    const int row = (tile / CH) * RPI + wrow, n0 = (tile % CH) * TW + nw;
    const int tab_off = tab_off0 + (tile % CH) * (TW / 32) * 64;
    float4 xb[8];
    load_x(row, n0, xb);
    if (next < ntiles) {
      prefetch_x((next / CH) * RPI + wrow, (next % CH) * TW + nw);
      load_z(next, zb);
    }
    if (row < Bn) {
      float acc[2][4][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[m][j][c] = 0.f;
      const float* zh = zring + (size_t)stage * 2 * RPI * RP * F2M_LDA + (size_t)wrow * RP * F2M_LDA + a_off;
      const float* zl = zh + RPI * RP * F2M_LDA;
      // spectral rows
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        if (ks < KS_T) {
          const float4 h0 = *reinterpret_cast<const float4*>(tabH + tab_off + ks * 4 * LDP);        // points j = 0, 1
          const float4 h1 = *reinterpret_cast<const float4*>(tabH + tab_off + ks * 4 * LDP + 4);    // points j = 2, 3
          const float4 l0 = *reinterpret_cast<const float4*>(tabL + tab_off + ks * 4 * LDP);
          const float4 l1 = *reinterpret_cast<const float4*>(tabL + tab_off + ks * 4 * LDP + 4);
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            uint32_t ah[4], al[4];
            const int o = ks * 8 * F2M_LDA + 16 * m;
            ah[0] = __float_as_uint(zh[o]);
            ah[1] = __float_as_uint(zh[o + 8]);
            ah[2] = __float_as_uint(zh[o + 4 * F2M_LDA]);
            ah[3] = __float_as_uint(zh[o + 4 * F2M_LDA + 8]);
            al[0] = __float_as_uint(zl[o]);
            al[1] = __float_as_uint(zl[o + 8]);
            al[2] = __float_as_uint(zl[o + 4 * F2M_LDA]);
            al[3] = __float_as_uint(zl[o + 4 * F2M_LDA + 8]);
            mma_3xtf32(acc[m][0], ah, al, __float_as_uint(h0.x), __float_as_uint(h0.y), __float_as_uint(l0.x), __float_as_uint(l0.y));
            mma_3xtf32(acc[m][1], ah, al, __float_as_uint(h0.z), __float_as_uint(h0.w), __float_as_uint(l0.z), __float_as_uint(l0.w));
            mma_3xtf32(acc[m][2], ah, al, __float_as_uint(h1.x), __float_as_uint(h1.y), __float_as_uint(l1.x), __float_as_uint(l1.y));
            mma_3xtf32(acc[m][3], ah, al, __float_as_uint(h1.z), __float_as_uint(h1.w), __float_as_uint(l1.z), __float_as_uint(l1.w));
          }
        }
      }
      // bypass conv: input channels
      const float* wh = wbH + a_off;
      const float* wl = wbL + a_off;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        if (ks < KS_X) {
          const float v0[4] = {xb[2 * ks].x, xb[2 * ks].y, xb[2 * ks].z, xb[2 * ks].w};
          const float v1[4] = {xb[2 * ks + 1].x, xb[2 * ks + 1].y, xb[2 * ks + 1].z, xb[2 * ks + 1].w};
          uint32_t bh0[4], bh1[4], bl0[4], bl1[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            tf32_split(v0[j], bh0[j], bl0[j]);
            tf32_split(v1[j], bh1[j], bl1[j]);
          }
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            uint32_t ah[4], al[4];
            const int o = ks * 8 * F2M_LDA + 16 * m;
            ah[0] = __float_as_uint(wh[o]);
            ah[1] = __float_as_uint(wh[o + 8]);
            ah[2] = __float_as_uint(wh[o + 4 * F2M_LDA]);
            ah[3] = __float_as_uint(wh[o + 4 * F2M_LDA + 8]);
            al[0] = __float_as_uint(wl[o]);
            al[1] = __float_as_uint(wl[o + 8]);
            al[2] = __float_as_uint(wl[o + 4 * F2M_LDA]);
            al[3] = __float_as_uint(wl[o + 4 * F2M_LDA + 8]);
#pragma unroll
            for (int j = 0; j < 4; ++j) mma_3xtf32(acc[m][j], ah, al, bh0[j], bh1[j], bl0[j], bl1[j]);
          }
        }
      }
      // epilogue: lane holds, for channels o = 16 m + g (+ 8), the points n0 + 4 t + {0..3} (accumulator columns 2 t of the
      // four n-tiles) and n0 + 16 + 4 t + {0..3} (columns 2 t + 1); the channel pointer advances by eight channels per step
      const int bh = row / Hd, hh = row - bh * Hd;
      float* yp = y + ((long long)bh * O * Hd + hh) * W + n0 + 4 * t + (long long)g * chan;
      const long long zd = z ? z - y : 0;
#pragma unroll
      for (int m = 0; m < 2; ++m) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int o = 16 * m + g + 8 * half;
          if (o < O) {
            const float bias = biasS[o];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              float pre[4], out[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) pre[j] = acc[m][j][2 * half + c] + bias;
              if (act == ACT_GELU) {   // two elements per instruction on the packed fp32x2 pipe
                float d0, d1;
                gelu_both2(pre[0], pre[1], out[0], out[1], d0, d1);
                gelu_both2(pre[2], pre[3], out[2], out[3], d0, d1);
              } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) out[j] = act_f(act, pre[j]);
              }
              *reinterpret_cast<float4*>(yp + 16 * c) = make_float4(out[0], out[1], out[2], out[3]);
              if (z) *reinterpret_cast<float4*>(yp + zd + 16 * c) = make_float4(pre[0], pre[1], pre[2], pre[3]);
            }
          }
          yp += 8 * chan;
        }
      }
    }
    if (next < ntiles) store_z(stage ^ 1, zb);
    __syncthreads();
    stage ^= 1;
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// fno1d_synth_mma_kernel: FNOBlock1d synthesis + 1x1 bypass conv + bias + activation (fno_blocks.py:25-33 around
// spectral_convolution.py:54), and with adjoint = 1 the complete input gradient (analysis adjoint + wb^T dpre), for up to 64
// channels per side:   y[b,o,:] = act( sum_r Z[r][b][o] * tab[r][:] + sum_i wb[o,i] * x[b,i,:] + bb[o] ).
// CTA = (b, 128 points): four point segments x two halves of the output channels over its eight warps.  Nothing is resident
// across tiles (at batch 16 there are as many tiles as SMs): the CTA stages the table slice of its 128 points (paired layout
// of fno2d_synth_mma_kernel), the row's Z and the bypass weight as TF32 hi / lo planes, while its x fragments are already in
// flight.  N % 128 == 0, I, O <= 64, 2 K <= 32.
constexpr int F1M_LDA = 72;    // A-operand planes [k][o], o < 64: stride = 8 mod 32
__host__ __device__ constexpr size_t f1m_synth_smem(int RP) {
  return ((size_t)2 * (RP / 2) * (2 * 128 + 4) + (size_t)2 * 64 * F1M_LDA + (size_t)2 * RP * F1M_LDA + 64) * sizeof(float);
}

__global__ void __launch_bounds__(256, 1) fno1d_synth_mma_kernel(const float* __restrict__ x, const float* __restrict__ Z,
                                                                 const float* __restrict__ table, const float* __restrict__ wb,
                                                                 const float* __restrict__ bb, float* __restrict__ y,
                                                                 float* __restrict__ z, int Bn, int I, int O, int N, int K, int act,
                                                                 int adjoint) {
  constexpr int LDP = 2 * 128 + 4;
  extern __shared__ __align__(16) float smem[];
  const int R = 2 * K, RP = (R + 7) & ~7, KS_T = RP / 8, KS_X = wb ? (I + 7) / 8 : 0;
  float* tabH = smem;
  float* tabL = tabH + (RP / 2) * LDP;
  float* wbH = tabL + (RP / 2) * LDP;
  float* wbL = wbH + 64 * F1M_LDA;
  float* zH = wbL + 64 * F1M_LDA;
  float* zL = zH + RP * F1M_LDA;
  float* biasS = zL + RP * F1M_LDA;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int pg = (g >> 1) + 4 * (g & 1);
  const int nb = blockIdx.x * 128;
  const int seg = warp & 3, oh = warp >> 2, n0 = nb + 32 * seg;
  // staging in batches of eight independent loads per thread (one L2 round trip per batch instead of one per element)
  for (int e0 = tid; e0 < RP * 128; e0 += 8 * 256) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int e = e0 + 256 * u, r = e >> 7, n = e & 127;
      v[u] = (e < RP * 128 && r < R) ? __ldg(table + (adjoint ? (long long)r : (long long)(r & 1) * K + (r >> 1)) * N + nb + n) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int e = e0 + 256 * u, r = e >> 7, n = e & 127;
      if (e >= RP * 128) continue;
      uint32_t hi, lo;
      tf32_split_rn(v[u], hi, lo);
      const int ks = r >> 3, tt = r & 3, half = (r >> 2) & 1;
      const int off = (ks * 4 + tt) * LDP + (n >> 5) * 64 + f2m_slot((n >> 2) & 7) * 8 + (n & 3) * 2 + half;
      tabH[off] = __uint_as_float(hi);
      tabL[off] = __uint_as_float(tf32_hi(__uint_as_float(lo)));
    }
  }
  for (int e0 = tid; e0 < 64 * 64; e0 += 8 * 256) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int e = e0 + 256 * u, i = e >> 6, o = e & 63;   // A[o][k = i]
      v[u] = (wb && i < I && o < O) ? __ldg(adjoint ? wb + (long long)i * O + o : wb + (long long)o * I + i) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int e = e0 + 256 * u, i = e >> 6, o = e & 63;
      uint32_t hi, lo;
      tf32_split_rn(v[u], hi, lo);
      wbH[i * F1M_LDA + o] = __uint_as_float(hi);
      wbL[i * F1M_LDA + o] = __uint_as_float(tf32_hi(__uint_as_float(lo)));
    }
  }
  if (tid < 64) biasS[tid] = (bb && tid < O) ? __ldg(bb + tid) : 0.f;
  // The table slice of this CTA's 128 points, the bypass weight and the bias stay for all of its rows: blockIdx.y strides over b
  // (one row per CTA at batch 16, ~14 at batch 256); per row only Z is re-staged, between two barriers.
  for (int b = blockIdx.y; b < Bn; b += gridDim.y) {
  // x fragments first: their latency hides behind the staging below
  float4 xb[16];
  {
    const float* p = x + ((long long)b * I + t) * N + n0 + 4 * pg;
#pragma unroll
    for (int q = 0; q < 16; ++q) {      // channel t + 4 q: k-step q >> 1, fragment row t (+ 4)
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if ((q >> 1) < KS_X && t + 4 * q < I) v = __ldg(reinterpret_cast<const float4*>(p + (long long)4 * q * N));
      xb[q] = v;
    }
  }
  __syncthreads();      // the previous row's MMAs are done with zH / zL (and the resident operands are staged)
  {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int e = tid + 256 * u, r = e >> 6, o = e & 63;
      v[u] = (e < RP * 64 && r < R && o < O) ? __ldg(Z + ((long long)r * Bn + b) * O + o) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int e = tid + 256 * u, r = e >> 6, o = e & 63;
      if (e >= RP * 64) continue;
      uint32_t hi, lo;
      tf32_split_rn(v[u], hi, lo);
      zH[r * F1M_LDA + o] = __uint_as_float(hi);
      zL[r * F1M_LDA + o] = __uint_as_float(tf32_hi(__uint_as_float(lo)));
    }
  }
  __syncthreads();

  float acc[2][4][4];
#pragma unroll
  for (int m = 0; m < 2; ++m)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[m][j][c] = 0.f;
  const int tab_off = t * LDP + seg * 64 + f2m_slot(pg) * 8;
  const int a_off = t * F1M_LDA + 32 * oh + g;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    if (ks < KS_T) {
      const float4 h0 = *reinterpret_cast<const float4*>(tabH + tab_off + ks * 4 * LDP);
      const float4 h1 = *reinterpret_cast<const float4*>(tabH + tab_off + ks * 4 * LDP + 4);
      const float4 l0 = *reinterpret_cast<const float4*>(tabL + tab_off + ks * 4 * LDP);
      const float4 l1 = *reinterpret_cast<const float4*>(tabL + tab_off + ks * 4 * LDP + 4);
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        uint32_t ah[4], al[4];
        const int o = a_off + ks * 8 * F1M_LDA + 16 * m;
        ah[0] = __float_as_uint(zH[o]);
        ah[1] = __float_as_uint(zH[o + 8]);
        ah[2] = __float_as_uint(zH[o + 4 * F1M_LDA]);
        ah[3] = __float_as_uint(zH[o + 4 * F1M_LDA + 8]);
        al[0] = __float_as_uint(zL[o]);
        al[1] = __float_as_uint(zL[o + 8]);
        al[2] = __float_as_uint(zL[o + 4 * F1M_LDA]);
        al[3] = __float_as_uint(zL[o + 4 * F1M_LDA + 8]);
        mma_3xtf32(acc[m][0], ah, al, __float_as_uint(h0.x), __float_as_uint(h0.y), __float_as_uint(l0.x), __float_as_uint(l0.y));
        mma_3xtf32(acc[m][1], ah, al, __float_as_uint(h0.z), __float_as_uint(h0.w), __float_as_uint(l0.z), __float_as_uint(l0.w));
        mma_3xtf32(acc[m][2], ah, al, __float_as_uint(h1.x), __float_as_uint(h1.y), __float_as_uint(l1.x), __float_as_uint(l1.y));
        mma_3xtf32(acc[m][3], ah, al, __float_as_uint(h1.z), __float_as_uint(h1.w), __float_as_uint(l1.z), __float_as_uint(l1.w));
      }
    }
  }
#pragma unroll
  for (int ks = 0; ks < 8; ++ks) {
    if (ks < KS_X) {
      const float v0[4] = {xb[2 * ks].x, xb[2 * ks].y, xb[2 * ks].z, xb[2 * ks].w};
      const float v1[4] = {xb[2 * ks + 1].x, xb[2 * ks + 1].y, xb[2 * ks + 1].z, xb[2 * ks + 1].w};
      uint32_t bh0[4], bh1[4], bl0[4], bl1[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        tf32_split(v0[j], bh0[j], bl0[j]);
        tf32_split(v1[j], bh1[j], bl1[j]);
      }
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        uint32_t ah[4], al[4];
        const int o = a_off + ks * 8 * F1M_LDA + 16 * m;
        ah[0] = __float_as_uint(wbH[o]);
        ah[1] = __float_as_uint(wbH[o + 8]);
        ah[2] = __float_as_uint(wbH[o + 4 * F1M_LDA]);
        ah[3] = __float_as_uint(wbH[o + 4 * F1M_LDA + 8]);
        al[0] = __float_as_uint(wbL[o]);
        al[1] = __float_as_uint(wbL[o + 8]);
        al[2] = __float_as_uint(wbL[o + 4 * F1M_LDA]);
        al[3] = __float_as_uint(wbL[o + 4 * F1M_LDA + 8]);
#pragma unroll
        for (int j = 0; j < 4; ++j) mma_3xtf32(acc[m][j], ah, al, bh0[j], bh1[j], bl0[j], bl1[j]);
      }
    }
  }
  float* yp = y + ((long long)b * O + 32 * oh + g) * N + n0 + 4 * t;
  const long long zd = z ? z - y : 0;
#pragma unroll
  for (int m = 0; m < 2; ++m) {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int o = 32 * oh + 16 * m + g + 8 * half;
      if (o < O) {
        const float bias = biasS[o];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          float pre[4], out[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) pre[j] = acc[m][j][2 * half + c] + bias;
          if (act == ACT_GELU) {
            float d0, d1;
            gelu_both2(pre[0], pre[1], out[0], out[1], d0, d1);
            gelu_both2(pre[2], pre[3], out[2], out[3], d0, d1);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) out[j] = act_f(act, pre[j]);
          }
          *reinterpret_cast<float4*>(yp + 16 * c) = make_float4(out[0], out[1], out[2], out[3]);
          if (z) *reinterpret_cast<float4*>(yp + zd + 16 * c) = make_float4(pre[0], pre[1], pre[2], pre[3]);
        }
      }
      yp += (long long)8 * N;
    }
  }
  }   // rows
}

// ---------------------------------------------------------------------------------------------------------------------
// conv1x1_wgrad_mma_kernel: parameter gradients of a channels-first 1x1 convolution (fno_blocks.py:29,67 bypass conv and
// the lifting / projection layers),  dw[o][i] = sum_(b,s) dy[b,o,s] x[b,i,s],  db[o] = sum_(b,s) dy[b,o,s],
// as D[o][i] = A[o][k] B[k][i] with k = grid point.  It replaces a split-K contraction with a 10^6-long reduction
// (140 us for 268 MB at 32 channels) plus a separate channel reduction for the bias.
//   * a warp owns 32 consecutive points of one batch element per step and a 32 x 32 block of (o, i); with more than 32
//     channels on a side the blocks are spread over the warps of the CTA (NB = blocks, 8 / NB point streams per CTA)
//   * both operands come straight from global memory as float4 (k-step j of the MMA <-> points 4 t' + j, so a lane's float4
//     at point offset 4 t (+ 16) feeds the four k-steps), split to TF32 hi / lo in registers
//   * accumulators stay in registers over the whole persistent loop; one shared-memory reduction over the CTA's streams
//     and one atomicAdd per element and CTA at the end (dw / db are zero-filled by the caller)
// S % 32 == 0 is required (float4 alignment and a warp's 32 points inside one row).
__global__ void __launch_bounds__(256, 2) conv1x1_wgrad_mma_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                                   float* __restrict__ dw, float* __restrict__ db, int Bn, int I,
                                                                   int O, int S) {
  // per-warp sums [32 o][40]: columns 0..31 the weight block, 32 the bias; stride 40 keeps the float2 updates conflict-free
  __shared__ __align__(16) float red[8][32 * 40];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int nbi = (I + 31) / 32, nbo = (O + 31) / 32, NB = nbi * nbo;      // 1, 2 or 4 blocks
  const int blk = warp % NB, stream = warp / NB, nstreams = 8 / NB;
  const int o0 = (blk / nbi) * 32, i0 = (blk % nbi) * 32;
  const int segs = S / 32;
  const long long ntiles = (long long)Bn * segs;
  float bsum[2][2];
#pragma unroll
  for (int m = 0; m < 2; ++m) bsum[m][0] = bsum[m][1] = 0.f;
  float* mine = red[warp];
  for (int e = lane; e < 32 * 40; e += 32) mine[e] = 0.f;
  __syncwarp();
  for (long long tile = (long long)blockIdx.x * nstreams + stream; tile < ntiles; tile += (long long)gridDim.x * nstreams) {
    const int b = (int)(tile / segs), n0 = (int)(tile - (long long)b * segs) * 32;
    float4 dv[2][2][2], xv[4][2];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int o = o0 + 16 * m + g + 8 * hh;
#pragma unroll
        for (int c = 0; c < 2; ++c)
          dv[m][hh][c] = o < O ? __ldg(reinterpret_cast<const float4*>(dy + ((long long)b * O + o) * S + n0 + 16 * c + 4 * t))
                               : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int i = i0 + 8 * q + g;
#pragma unroll
      for (int c = 0; c < 2; ++c)
        xv[q][c] = i < I ? __ldg(reinterpret_cast<const float4*>(x + ((long long)b * I + i) * S + n0 + 16 * c + 4 * t))
                         : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
      for (int hh = 0; hh < 2; ++hh)
#pragma unroll
        for (int c = 0; c < 2; ++c) bsum[m][hh] += (dv[m][hh][c].x + dv[m][hh][c].y) + (dv[m][hh][c].z + dv[m][hh][c].w);
    // the tile's MMA chain starts from zero and is added to the warp's sums with round-to-nearest adds: the tensor core
    // truncates when it adds into its accumulator, and a chain over all tiles would bias cancelling sums by ~1e-5
    float acc[2][4][4];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[m][q][c] = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      auto comp = [&](const float4& v) { return j == 0 ? v.x : j == 1 ? v.y : j == 2 ? v.z : v.w; };
      uint32_t bh[4][2], bl[4][2];
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int c = 0; c < 2; ++c) tf32_split(comp(xv[q][c]), bh[q][c], bl[q][c]);
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        uint32_t ah[4], al[4];
        tf32_split(comp(dv[m][0][0]), ah[0], al[0]);
        tf32_split(comp(dv[m][1][0]), ah[1], al[1]);
        tf32_split(comp(dv[m][0][1]), ah[2], al[2]);
        tf32_split(comp(dv[m][1][1]), ah[3], al[3]);
#pragma unroll
        for (int q = 0; q < 4; ++q) mma_3xtf32(acc[m][q], ah, al, bh[q][0], bh[q][1], bl[q][0], bl[q][1]);
      }
    }
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float2* p0 = reinterpret_cast<float2*>(mine + (16 * m + g) * 40 + 8 * q + 2 * t);
        float2* p1 = reinterpret_cast<float2*>(mine + (16 * m + g + 8) * 40 + 8 * q + 2 * t);
        float2 a = *p0, b = *p1;
        a.x += acc[m][q][0]; a.y += acc[m][q][1];
        b.x += acc[m][q][2]; b.y += acc[m][q][3];
        *p0 = a;
        *p1 = b;
      }
  }
  // CTA reduction: the streams of one block add up in shared memory, then one atomic per element
  // bias partial sums: the four lanes of a group hold the same rows; column 32 of the staging tile carries them
#pragma unroll
  for (int m = 0; m < 2; ++m)
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      float v = bsum[m][hh];
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      if (t == 0) mine[(16 * m + g + 8 * hh) * 40 + 32] = v;
    }
  __syncthreads();
  for (int e = tid; e < NB * 32 * 33; e += 256) {
    const int bk = e / (32 * 33), r = e - bk * 32 * 33;
    const int row = r / 33, col = r - row * 33;
    float v = 0.f;
    for (int st = 0; st < nstreams; ++st) v += red[st * NB + bk][row * 40 + col];
    const int o = (bk / nbi) * 32 + row;
    if (o >= O) continue;
    if (col == 32) {
      if (db && bk % nbi == 0) atomicAdd(db + o, v);
    } else {
      const int i = (bk % nbi) * 32 + col;
      if (i < I) atomicAdd(dw + (long long)o * I + i, v);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Projection MLP of FNO1d / FNO2d (fno_blocks.py:35-45, 73-83; fno.py:20,45): y = W2 gelu(W1 x + b1) + b2, channels first,
// x (B, I, S) -> hidden (Hd = 4 * width) -> y (B, O, S) with O = out_channels (1 for every configured dataset).
// Composed, the hidden tensor and gelu'(z) are 2 x 537 MB at C4-256 (0.63 ms forward, 1.05 ms backward).  Fused, the hidden
// activations live in MMA accumulators: forward reads x and writes y; backward recomputes them from x.
//
// mlp_fwd_mma_kernel: a warp owns 32 consecutive points; per 32-row chunk of the hidden layer it forms z = W1 x (3xTF32,
// D[h][n] = A[h][i] B[i][n]; W1 resident in shared memory as interleaved (hi, lo) pairs, x in registers as float4 B
// fragments with the same column <-> point permutation as fno2d_synth_mma_kernel), applies bias + GELU and folds the rows
// into y[o][n] += w2[o][h] gelu(z) on the CUDA cores; one 8-lane shuffle reduction per tile, float4 stores.
constexpr int MLP_MAXO = 4;
__host__ __device__ constexpr int mlp_ip(int I) { return I <= 32 ? 32 : 64; }          // input channels as the register tiles see them
__host__ __device__ constexpr int mlp_ldw(int I) { return mlp_ip(I) + 4; }             // W1 plane row stride, = 4 mod 32
__host__ __device__ constexpr size_t mlp_fwd_smem(int I, int Hd) {
  return ((size_t)2 * ((Hd + 31) & ~31) * mlp_ldw(I) + (size_t)(MLP_MAXO + 1) * ((Hd + 31) & ~31)) * sizeof(float);
}
// W1 lives in shared memory as two planes (TF32 hi, lo) [h][LDW], read with scalar loads at compile-time offsets from one
// per-lane base (no address arithmetic and no register shuffling in front of the MMA).  LDW = 4 mod 32: the fragments of
// z = W1 x (A[m = h][k = i]: lanes (g, t) -> row h0 + g, column i0 + t) hit 32 distinct banks, those of dx = W1^T dz
// (A[m = i][k = h]: lanes (g, t) -> row h0 + t, column i0 + g) take a 2-way conflict (cheaper than a second copy).
__device__ __forceinline__ void mlp_stage_w1(const float* __restrict__ w1, float* w1h, float* w1l, int I, int Hd, int HP, int IP) {
  const int LDW = IP + 4;
  // eight independent loads per thread and round trip (a plain element loop serialises on L2 latency: 64 round trips for the
  // 256 x 64 weight of the FNO1d projection, most of the kernel's time on a small batch)
  for (int e0 = threadIdx.x; e0 < HP * IP; e0 += 8 * blockDim.x) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int e = e0 + u * blockDim.x, h = e / IP, i = e - h * IP;
      v[u] = (e < HP * IP && h < Hd && i < I) ? __ldg(w1 + (long long)h * I + i) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int e = e0 + u * blockDim.x, h = e / IP, i = e - h * IP;
      if (e >= HP * IP) continue;
      uint32_t hi, lo;
      tf32_split_rn(v[u], hi, lo);
      w1h[h * LDW + i] = __uint_as_float(hi);
      w1l[h * LDW + i] = __uint_as_float(tf32_hi(__uint_as_float(lo)));
    }
  }
}
// A fragment of the z orientation at (h0, i0): bz = plane + g * LDW + t
template <int LDW>
__device__ __forceinline__ void mlp_afrag_z(const float* bzh, const float* bzl, int h0, int i0, uint32_t (&ah)[4], uint32_t (&al)[4]) {
  const int o = h0 * LDW + i0;
  ah[0] = __float_as_uint(bzh[o]); ah[1] = __float_as_uint(bzh[o + 8 * LDW]);
  ah[2] = __float_as_uint(bzh[o + 4]); ah[3] = __float_as_uint(bzh[o + 8 * LDW + 4]);
  al[0] = __float_as_uint(bzl[o]); al[1] = __float_as_uint(bzl[o + 8 * LDW]);
  al[2] = __float_as_uint(bzl[o + 4]); al[3] = __float_as_uint(bzl[o + 8 * LDW + 4]);
}
// A fragment of the transposed orientation at (h0, i0): rows (m) are input channels, columns (k) hidden rows; bx = plane + t * LDW + g
template <int LDW>
__device__ __forceinline__ void mlp_afrag_x(const float* bxh, const float* bxl, int h0, int i0, uint32_t (&ah)[4], uint32_t (&al)[4]) {
  const int o = h0 * LDW + i0;
  ah[0] = __float_as_uint(bxh[o]); ah[1] = __float_as_uint(bxh[o + 8]);
  ah[2] = __float_as_uint(bxh[o + 4 * LDW]); ah[3] = __float_as_uint(bxh[o + 4 * LDW + 8]);
  al[0] = __float_as_uint(bxl[o]); al[1] = __float_as_uint(bxl[o + 8]);
  al[2] = __float_as_uint(bxl[o + 4 * LDW]); al[3] = __float_as_uint(bxl[o + 4 * LDW + 8]);
}

template <int KS, int NO>   // k-steps of eight input channels: 4 (I <= 32) or 8 (I <= 64); NO: outputs held (1 or MLP_MAXO)
__global__ void __launch_bounds__(256, KS == 4 ? 2 : 1) mlp_fwd_mma_kernel(const float* __restrict__ x, const float* __restrict__ w1,
                                                                          const float* __restrict__ b1, const float* __restrict__ w2,
                                                                          const float* __restrict__ b2, float* __restrict__ y, int Bn,
                                                                          int I, int Hd, int O, int S) {
  extern __shared__ __align__(16) float smem[];
  constexpr int IP = 8 * KS, LDW = IP + 4;
  const int HP = (Hd + 31) & ~31;
  float* w1h = smem;                  // [HP][LDW]
  float* w1l = w1h + HP * LDW;
  float* b1s = w1l + HP * LDW;        // [HP]
  float* w2s = b1s + HP;              // [MLP_MAXO][HP]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int pg = (g >> 1) + 4 * (g & 1);
  mlp_stage_w1(w1, w1h, w1l, I, Hd, HP, IP);
  for (int e = tid; e < HP; e += 256) {
    b1s[e] = (b1 && e < Hd) ? __ldg(b1 + e) : 0.f;
#pragma unroll
    for (int o = 0; o < MLP_MAXO; ++o) w2s[o * HP + e] = (o < O && e < Hd) ? __ldg(w2 + (long long)o * Hd + e) : 0.f;
  }
  __syncthreads();
  const int segs = S / 32;
  const long long ntiles = (long long)Bn * segs;
  for (long long tile = (long long)blockIdx.x * 8 + warp; tile < ntiles; tile += (long long)gridDim.x * 8) {
    const int b = (int)(tile / segs), n0 = (int)(tile - (long long)b * segs) * 32;
    float4 xb[2 * KS];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int i = 8 * ks + t + 4 * h;
        xb[2 * ks + h] = i < I ? __ldg(reinterpret_cast<const float4*>(x + ((long long)b * I + i) * S + n0 + 4 * pg))
                               : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    float ysum[NO][4][2];
#pragma unroll
    for (int o = 0; o < NO; ++o)
#pragma unroll
      for (int j = 0; j < 4; ++j) ysum[o][j][0] = ysum[o][j][1] = 0.f;
    for (int hc = 0; hc < HP; hc += 32) {
      const float* bzh = w1h + (hc + g) * LDW + t;
      const float* bzl = w1l + (hc + g) * LDW + t;
      float acc[2][4][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[m][j][c] = 0.f;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        const float v0[4] = {xb[2 * ks].x, xb[2 * ks].y, xb[2 * ks].z, xb[2 * ks].w};
        const float v1[4] = {xb[2 * ks + 1].x, xb[2 * ks + 1].y, xb[2 * ks + 1].z, xb[2 * ks + 1].w};
        uint32_t bh0[4], bh1[4], bl0[4], bl1[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          tf32_split_here(v0[j], bh0[j], bl0[j]);
          tf32_split_here(v1[j], bh1[j], bl1[j]);
        }
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          uint32_t ah[4], al[4];
          mlp_afrag_z<LDW>(bzh, bzl, 16 * m, 8 * ks, ah, al);
#pragma unroll
          for (int j = 0; j < 4; ++j) mma_3xtf32(acc[m][j], ah, al, bh0[j], bh1[j], bl0[j], bl1[j]);
        }
      }
      // bias + GELU, rows folded into the outputs
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const int h = hc + 16 * m + g + 8 * hh;
          const float bias = b1s[h];
          float hv[4][2];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float d0, d1;
            gelu_both2(acc[m][j][2 * hh] + bias, acc[m][j][2 * hh + 1] + bias, hv[j][0], hv[j][1], d0, d1);
          }
#pragma unroll
          for (int o = 0; o < NO; ++o) {
            if (o < O) {
              const float w = w2s[o * HP + h];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                ysum[o][j][0] = fmaf(w, hv[j][0], ysum[o][j][0]);
                ysum[o][j][1] = fmaf(w, hv[j][1], ysum[o][j][1]);
              }
            }
          }
        }
    }
    // the eight lanes of equal t hold the partial sums of the same 8 points: butterfly over g, lane g == o stores output o
#pragma unroll
    for (int o = 0; o < NO; ++o) {
      if (o < O) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            float v = ysum[o][j][c];
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            ysum[o][j][c] = v;
          }
        if (g == o) {
          const float bo = b2 ? __ldg(b2 + o) : 0.f;
          float* yp = y + ((long long)b * O + o) * S + n0 + 4 * t;
          *reinterpret_cast<float4*>(yp) = make_float4(ysum[o][0][0] + bo, ysum[o][1][0] + bo, ysum[o][2][0] + bo, ysum[o][3][0] + bo);
          *reinterpret_cast<float4*>(yp + 16) = make_float4(ysum[o][0][1] + bo, ysum[o][1][1] + bo, ysum[o][2][1] + bo, ysum[o][3][1] + bo);
        }
      }
    }
  }
}

// mlp_bwd_mma_kernel: every gradient of the projection MLP in one pass over x and dy.  A warp owns (32 points, one 32-row
// chunk of the hidden layer); the NCH = Hd / 32 chunks of a point tile sit in NCH warps of the CTA (8 / NCH point tiles per
// iteration).  Per iteration and warp:
//   1. z = W1[chunk] x  (3xTF32; the x tile arrives through a cp.async double buffer shared by the chunk warps)
//   2. h = gelu(z + b1), dz = (sum_o w2[o] dy[o]) * gelu'(z + b1) in the accumulator registers; dw2 / db1 partial sums
//   3. dw1[chunk] += dz x^T: dz is the A operand straight from the accumulator layout (k <-> the accumulator's columns)
//   4. dx partial = W1[chunk]^T dz: dz goes through the warp's shared-memory tile to reach the B-operand layout
//   5. the chunk warps add their dx partials through shared memory, 16 input channels at a time; float4 stores
// dw1 / db1 / dw2 / db2 leave the CTA as atomics at the end of the persistent loop (zero-filled by the caller).
constexpr int MLP_LDX = 36;   // x / dz tiles [row][32 points]: stride = 4 mod 32
__host__ __device__ constexpr size_t mlp_bwd_smem(int I, int Hd) {
  const int HP = (Hd + 31) & ~31, IP = mlp_ip(I);
  return ((size_t)2 * HP * (IP + 4) + (size_t)(MLP_MAXO + 1) * HP + (size_t)2 * (256 / HP) * IP * MLP_LDX + (size_t)8 * 32 * MLP_LDX) *
         sizeof(float);
}

template <int KS, int NO>
__global__ void __launch_bounds__(256, KS == 4 ? 2 : 1) mlp_bwd_mma_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                             const float* __restrict__ w1, const float* __restrict__ b1,
                                                             const float* __restrict__ w2, float* __restrict__ dx,
                                                             float* __restrict__ dw1, float* __restrict__ db1,
                                                             float* __restrict__ dw2, float* __restrict__ db2, int Bn, int I, int Hd,
                                                             int O, int S) {
  constexpr int IP = 8 * KS, IT = KS / 2, LDW = IP + 4;
  extern __shared__ __align__(16) float smem[];
  const int HP = (Hd + 31) & ~31, NCH = HP / 32, PT = 8 / NCH;
  float* w1h = smem;
  float* w1l = w1h + HP * LDW;
  float* b1s = w1l + HP * LDW;
  float* w2s = b1s + HP;
  float* xs = w2s + MLP_MAXO * HP;                     // [2][PT][IP][MLP_LDX]
  float* bufs = xs + 2 * PT * IP * MLP_LDX;            // [8 warps][32][MLP_LDX]: dz tile, then dx partials
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int pg = (g >> 1) + 4 * (g & 1);
  const int ptid = warp / NCH, hc = (warp % NCH) * 32;
  float* mybuf = bufs + warp * 32 * MLP_LDX;
  const float* bzh = w1h + (hc + g) * LDW + t;       // fragment bases of this warp's chunk, both orientations
  const float* bzl = w1l + (hc + g) * LDW + t;
  const float* bxh = w1h + (hc + t) * LDW + g;
  const float* bxl = w1l + (hc + t) * LDW + g;
  mlp_stage_w1(w1, w1h, w1l, I, Hd, HP, IP);
  for (int e = tid; e < HP; e += 256) {
    b1s[e] = (b1 && e < Hd) ? __ldg(b1 + e) : 0.f;
#pragma unroll
    for (int o = 0; o < MLP_MAXO; ++o) w2s[o * HP + e] = (o < O && e < Hd) ? __ldg(w2 + (long long)o * Hd + e) : 0.f;
  }
  for (int e = tid; e < 2 * PT * IP * MLP_LDX; e += 256) xs[e] = 0.f;   // rows i >= I stay zero
  __syncthreads();

  const int segs = S / 32;
  const long long ntiles = (long long)Bn * segs, ngroups = (ntiles + PT - 1) / PT;
  auto stage_x = [&](long long group, int bufi) {
    // PT tiles x I rows x eight 16-byte pieces
    for (int e = tid; e < PT * I * 8; e += 256) {
      const int piece = e & 7, r = e >> 3, p = r / I, i = r - p * I;
      const long long tile = group * PT + p;
      if (tile < ntiles) {
        const int b = (int)(tile / segs), n0 = (int)(tile - (long long)b * segs) * 32;
        const float* src = x + ((long long)b * I + i) * S + n0 + 4 * piece;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(xs + ((bufi * PT + p) * IP + i) * MLP_LDX + 4 * piece);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src));
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  float dw1acc[2][KS][4];
  float db1p[2][2], dw2p[NO][2][2], db2p[NO];
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    db1p[m][0] = db1p[m][1] = 0.f;
#pragma unroll
    for (int o = 0; o < NO; ++o) dw2p[o][m][0] = dw2p[o][m][1] = 0.f;
#pragma unroll
    for (int q = 0; q < KS; ++q)
#pragma unroll
      for (int c = 0; c < 4; ++c) dw1acc[m][q][c] = 0.f;
  }
#pragma unroll
  for (int o = 0; o < NO; ++o) db2p[o] = 0.f;

  long long group = blockIdx.x;
  int bufi = 0;
  if (group < ngroups) stage_x(group, 0);
  for (; group < ngroups; group += gridDim.x) {
    const long long nextg = group + gridDim.x;
    if (nextg < ngroups) {
      stage_x(nextg, bufi ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const long long tile = group * PT + ptid;
    const bool live = tile < ntiles;
    const int b = live ? (int)(tile / segs) : 0, n0 = live ? (int)(tile - (long long)b * segs) * 32 : 0;
    const float* xt = xs + (bufi * PT + ptid) * IP * MLP_LDX;
    float dxacc[IT][4][4];
#pragma unroll
    for (int mt = 0; mt < IT; ++mt)
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int c = 0; c < 4; ++c) dxacc[mt][j][c] = 0.f;
    if (live) {
      // ---- 1. z chunk ------------------------------------------------------------------------------------------
      float acc[2][4][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[m][j][c] = 0.f;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        const float4 x0 = *reinterpret_cast<const float4*>(xt + (8 * ks + t) * MLP_LDX + 4 * pg);
        const float4 x1 = *reinterpret_cast<const float4*>(xt + (8 * ks + t + 4) * MLP_LDX + 4 * pg);
        const float v0[4] = {x0.x, x0.y, x0.z, x0.w}, v1[4] = {x1.x, x1.y, x1.z, x1.w};
        uint32_t bh0[4], bh1[4], bl0[4], bl1[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          tf32_split(v0[j], bh0[j], bl0[j]);
          tf32_split(v1[j], bh1[j], bl1[j]);
        }
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          uint32_t ah[4], al[4];
          mlp_afrag_z<LDW>(bzh, bzl, 16 * m, 8 * ks, ah, al);
#pragma unroll
          for (int j = 0; j < 4; ++j) mma_3xtf32(acc[m][j], ah, al, bh0[j], bh1[j], bl0[j], bl1[j]);
        }
      }
      // ---- 2. dz in place; column (j, c) of the accumulators <-> point n0 + 16 c + 4 t + j ----------------------------
      float dyv[NO][4][2];
#pragma unroll
      for (int o = 0; o < NO; ++o) {
        if (o < O) {
          const float* dp = dy + ((long long)b * O + o) * S + n0 + 4 * t;
          const float4 d0 = __ldg(reinterpret_cast<const float4*>(dp)), d1 = __ldg(reinterpret_cast<const float4*>(dp + 16));
          dyv[o][0][0] = d0.x; dyv[o][1][0] = d0.y; dyv[o][2][0] = d0.z; dyv[o][3][0] = d0.w;
          dyv[o][0][1] = d1.x; dyv[o][1][1] = d1.y; dyv[o][2][1] = d1.z; dyv[o][3][1] = d1.w;
          if (hc == 0 && g == 0) db2p[o] += ((d0.x + d0.y) + (d0.z + d0.w)) + ((d1.x + d1.y) + (d1.z + d1.w));
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) dyv[o][j][0] = dyv[o][j][1] = 0.f;
        }
      }
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const int hl = 16 * m + g + 8 * hh, h = hc + hl;
          const float bias = b1s[h];
          float w2v[NO];
#pragma unroll
          for (int o = 0; o < NO; ++o) w2v[o] = w2s[o * HP + h];
          float dzr[4][2];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float h0, h1, g0, g1;
            gelu_both2(acc[m][j][2 * hh] + bias, acc[m][j][2 * hh + 1] + bias, h0, h1, g0, g1);
            float dh0 = 0.f, dh1 = 0.f;
#pragma unroll
            for (int o = 0; o < NO; ++o) {
              dh0 = fmaf(w2v[o], dyv[o][j][0], dh0);
              dh1 = fmaf(w2v[o], dyv[o][j][1], dh1);
              dw2p[o][m][hh] = fmaf(dyv[o][j][0], h0, fmaf(dyv[o][j][1], h1, dw2p[o][m][hh]));
            }
            dzr[j][0] = dh0 * g0;
            dzr[j][1] = dh1 * g1;
            db1p[m][hh] += dzr[j][0] + dzr[j][1];
            acc[m][j][2 * hh] = dzr[j][0];
            acc[m][j][2 * hh + 1] = dzr[j][1];
          }
          *reinterpret_cast<float4*>(mybuf + hl * MLP_LDX + 4 * t) = make_float4(dzr[0][0], dzr[1][0], dzr[2][0], dzr[3][0]);
          *reinterpret_cast<float4*>(mybuf + hl * MLP_LDX + 16 + 4 * t) = make_float4(dzr[0][1], dzr[1][1], dzr[2][1], dzr[3][1]);
        }
      __syncwarp();
      // ---- 3. dw1[chunk] += dz x^T: k-step j, k = t <-> point 4 t + j, k = t + 4 <-> point 16 + 4 t + j -----------------
      // The tensor core adds into its accumulator with truncation: a chain over all of a warp's tiles (~170 MMAs) leaves
      // a systematic 1e-5 error on cancelling sums.  Each tile's 12-MMA chain therefore starts from zero and is added to the
      // persistent sums on the CUDA cores (round to nearest).  Half of the input channels at a time (register budget); the
      // x values are re-read per (j, q) as scalars (4-way bank conflicts on 32 loads cost less than holding 2 KS float4).
#pragma unroll
      for (int qh = 0; qh < 2; ++qh) {
        float tmp[2][KS / 2][4];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int q = 0; q < KS / 2; ++q)
#pragma unroll
            for (int c = 0; c < 4; ++c) tmp[m][q][c] = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t ah[2][4], al[2][4];
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            tf32_split_here(acc[m][j][0], ah[m][0], al[m][0]);   // pinned: not merged across the two halves
            tf32_split_here(acc[m][j][2], ah[m][1], al[m][1]);
            tf32_split_here(acc[m][j][1], ah[m][2], al[m][2]);
            tf32_split_here(acc[m][j][3], ah[m][3], al[m][3]);
          }
#pragma unroll
          for (int q = 0; q < KS / 2; ++q) {
            const int qq = qh * (KS / 2) + q;
            const float va = xt[(8 * qq + g) * MLP_LDX + 4 * t + j];
            const float vc = xt[(8 * qq + g) * MLP_LDX + 16 + 4 * t + j];
            uint32_t bh0, bl0, bh1, bl1;
            tf32_split(va, bh0, bl0);
            tf32_split(vc, bh1, bl1);
            mma_3xtf32(tmp[0][q], ah[0], al[0], bh0, bh1, bl0, bl1);
            mma_3xtf32(tmp[1][q], ah[1], al[1], bh0, bh1, bl0, bl1);
          }
        }
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int q = 0; q < KS / 2; ++q)
#pragma unroll
            for (int c = 0; c < 4; ++c) dw1acc[m][qh * (KS / 2) + q][c] += tmp[m][q][c];
      }
      // ---- 4. dx partial = W1[chunk]^T dz --------------------------------------------------------------------------
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const float4 z0 = *reinterpret_cast<const float4*>(mybuf + (8 * ks + t) * MLP_LDX + 4 * pg);
        const float4 z1 = *reinterpret_cast<const float4*>(mybuf + (8 * ks + t + 4) * MLP_LDX + 4 * pg);
        const float v0[4] = {z0.x, z0.y, z0.z, z0.w}, v1[4] = {z1.x, z1.y, z1.z, z1.w};
        uint32_t bh0[4], bh1[4], bl0[4], bl1[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          tf32_split(v0[j], bh0[j], bl0[j]);
          tf32_split(v1[j], bh1[j], bl1[j]);
        }
#pragma unroll
        for (int mt = 0; mt < IT; ++mt) {
          uint32_t ah[4], al[4];
          mlp_afrag_x<LDW>(bxh, bxl, 8 * ks, 16 * mt, ah, al);
#pragma unroll
          for (int j = 0; j < 4; ++j) mma_3xtf32(dxacc[mt][j], ah, al, bh0[j], bh1[j], bl0[j], bl1[j]);
        }
      }
      __syncwarp();
    }
    // ---- 5. dx = sum over the chunk warps, 16 input channels per round --------------------------------------------
#pragma unroll
    for (int mt = 0; mt < IT; ++mt) {
      if (live) {
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          float* row = mybuf + (g + 8 * hh) * MLP_LDX;
          *reinterpret_cast<float4*>(row + 4 * t) =
              make_float4(dxacc[mt][0][2 * hh], dxacc[mt][1][2 * hh], dxacc[mt][2][2 * hh], dxacc[mt][3][2 * hh]);
          *reinterpret_cast<float4*>(row + 16 + 4 * t) =
              make_float4(dxacc[mt][0][2 * hh + 1], dxacc[mt][1][2 * hh + 1], dxacc[mt][2][2 * hh + 1], dxacc[mt][3][2 * hh + 1]);
        }
      }
      __syncthreads();
      if (dx) {
        // the point tile's NCH warps share its 16 rows x 8 float4
        for (int e = tid - ptid * NCH * 32; live && e < 128; e += NCH * 32) {
          const int r = e >> 3, p4 = e & 7, i = 16 * mt + r;
          if (i < I) {
            float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int w = 0; w < NCH; ++w) {
              const float4 v = *reinterpret_cast<const float4*>(bufs + ((ptid * NCH + w) * 32 + r) * MLP_LDX + 4 * p4);
              sum.x += v.x; sum.y += v.y; sum.z += v.z; sum.w += v.w;
            }
            *reinterpret_cast<float4*>(dx + ((long long)b * I + i) * S + n0 + 4 * p4) = sum;
          }
        }
      }
      __syncthreads();
    }
    bufi ^= 1;
  }
  // ---- parameter gradients ---------------------------------------------------------------------------------------
#pragma unroll
  for (int m = 0; m < 2; ++m) {
#pragma unroll
    for (int q = 0; q < KS; ++q)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int h = hc + 16 * m + g + 8 * (c >> 1), i = 8 * q + 2 * t + (c & 1);
        if (h < Hd && i < I) atomicAdd(dw1 + (long long)h * I + i, dw1acc[m][q][c]);
      }
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int h = hc + 16 * m + g + 8 * hh;
      float v = db1p[m][hh];
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      if (t == 0 && h < Hd && db1) atomicAdd(db1 + h, v);
#pragma unroll
      for (int o = 0; o < NO; ++o) {
        float u = dw2p[o][m][hh];
        u += __shfl_xor_sync(0xffffffffu, u, 1);
        u += __shfl_xor_sync(0xffffffffu, u, 2);
        if (t == 0 && h < Hd && o < O) atomicAdd(dw2 + (long long)o * Hd + h, u);
      }
    }
  }
#pragma unroll
  for (int o = 0; o < NO; ++o) {
    float v = db2p[o];   // non-zero only in lanes g == 0 of the chunk-0 warps: the four t lanes hold disjoint points
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    if (lane == 0 && hc == 0 && o < O && db2) atomicAdd(db2 + o, v);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// fno2d_waxis_mma_kernel: the W-axis truncated DFT of the image rows (spectral_convolution.py:82 rfft2's last axis) and, in
// the backward pass, the adjoint of the W-axis synthesis with act'(z) applied on load and dpre written once:
//   T[row][r] = sum_w x[row][w] * tab[r][w],   row = (b, c, h),  r < R = 2 m2 <= 32
// as D[row][r] = A[row][k = w] B[k = w][r]: a warp owns 16 image rows, walks the row in groups of 32 points (four k-steps;
// k-step j <-> points 4 t' + j, so the lane's two float4 per row feed all four), the table is resident in shared memory as
// TF32 hi / lo planes.  Each group's 12-MMA chain starts from zero and is added to the row sums on the CUDA cores (the tensor
// core's truncating accumulation would bias the cancelling spectral sums).  The next group's loads are issued before the
// current group's MMAs.  Same arguments as fno2d_waxis_kernel; W % 32 == 0.
// PRESPLIT = false (long rows: FNO1d at N = 1024 would need 266 KB for two planes): the table stays in shared memory as one
// fp32 plane and its fragments are split when they are read.  The output element (row, r) goes to T[row * s_row + r * s_r]:
// (R, 1) for the 2-D half transform, (1, rows) for the FNO1d spectrum layout [mode][re|im][b][i] (fno_fused.cuh).
__host__ __device__ constexpr size_t f2m_waxis_smem(int W, int RP, bool presplit) {
  return (size_t)(presplit ? 2 : 1) * RP * (W + 4) * sizeof(float);
}

template <bool PRESPLIT>
__global__ void __launch_bounds__(256, PRESPLIT ? 2 : 1) fno2d_waxis_mma_kernel(const float* __restrict__ x, const float* __restrict__ tab,
                                                                 float* __restrict__ T, long long rows, int W, int R,
                                                                 const float* __restrict__ z, int act, float* __restrict__ xout,
                                                                 int kplanes, long long s_row, long long s_r) {
  extern __shared__ __align__(16) float smem[];
  // Table rows keep each 32-point group as [t][j][k = t | k = t + 4]: the lane's eight floats at 8 t are the B operand pairs
  // of the four k-steps (no register shuffling in front of the MMA); row stride = 4 mod 32: conflict-free float4 loads
  // gridDim.y > 1 (few rows: FNO1d at batch 16 has 64 row tiles): the CTAs of one blockIdx.y take W / gridDim.y points of
  // every row, stage only that slice of the table, and add their partial sums into T with atomics (T zero-filled by the host)
  const int Wc = W / gridDim.y, w0 = blockIdx.y * Wc;
  const int RP = (R + 7) & ~7, NT = RP / 8, LDT = Wc + 4;
  float* tabH = smem;
  float* tabL = tabH + (PRESPLIT ? RP * LDT : 0);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  for (int e = tid; e < RP * Wc; e += 256) {
    const int r = e / Wc, n = e - r * Wc;
    const int tr = kplanes ? (r & 1) * kplanes + (r >> 1) : r;
    const float v = r < R ? __ldg(tab + (long long)tr * W + w0 + n) : 0.f;
    // point n = 32 gi + 16 half + 4 tt + j  ->  32 gi + 8 tt + 2 j + half
    const int off = r * LDT + (n & ~31) + 8 * ((n >> 2) & 3) + 2 * (n & 3) + ((n >> 4) & 1);
    if (PRESPLIT) {
      uint32_t hi, lo;
      tf32_split_rn(v, hi, lo);
      tabH[off] = __uint_as_float(hi);
      tabL[off] = __uint_as_float(tf32_hi(__uint_as_float(lo)));
    } else {
      tabH[off] = v;
    }
  }
  __syncthreads();
  const int groups = Wc / 32;
  const long long mtiles = (rows + 15) / 16;
  for (long long mt = (long long)blockIdx.x * 8 + warp; mt < mtiles; mt += (long long)gridDim.x * 8) {
    const long long row0 = mt * 16 + g, row1 = row0 + 8;
    const bool ok0 = row0 < rows, ok1 = row1 < rows;
    const float* x0 = x + row0 * W + w0 + 4 * t;
    const float* x1 = x + row1 * W + w0 + 4 * t;
    float sum[4][4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int c = 0; c < 4; ++c) sum[nt][c] = 0.f;
    // raw loads of one group: v[0..1] row g at points 4 t and 16 + 4 t, v[2..3] row g + 8 (zz: the pre-activations).  They are
    // issued one group ahead; act'(z), the dpre store and the TF32 split happen when the group is consumed.
    auto load = [&](int gi, float4 (&v)[4], float4 (&zz)[4]) {
      const int p = 32 * gi;
      const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
      v[0] = ok0 ? __ldg(reinterpret_cast<const float4*>(x0 + p)) : zero;
      v[1] = ok0 ? __ldg(reinterpret_cast<const float4*>(x0 + p + 16)) : zero;
      v[2] = ok1 ? __ldg(reinterpret_cast<const float4*>(x1 + p)) : zero;
      v[3] = ok1 ? __ldg(reinterpret_cast<const float4*>(x1 + p + 16)) : zero;
      if (z) {
        const float* z0 = z + row0 * W + w0 + 4 * t + p;
        const float* z1 = z + row1 * W + w0 + 4 * t + p;
        zz[0] = ok0 ? __ldg(reinterpret_cast<const float4*>(z0)) : zero;
        zz[1] = ok0 ? __ldg(reinterpret_cast<const float4*>(z0 + 16)) : zero;
        zz[2] = ok1 ? __ldg(reinterpret_cast<const float4*>(z1)) : zero;
        zz[3] = ok1 ? __ldg(reinterpret_cast<const float4*>(z1 + 16)) : zero;
      }
    };
    float4 nv[4], nz[4];
    load(0, nv, nz);
    for (int gi = 0; gi < groups; ++gi) {
      float4 cur[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) cur[q] = nv[q];
      if (z) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float d0, d1, d2, d3;
          if (act == ACT_GELU) {   // two derivatives per instruction on the packed fp32x2 pipe
            float h0, h1;
            gelu_both2(nz[q].x, nz[q].y, h0, h1, d0, d1);
            gelu_both2(nz[q].z, nz[q].w, h0, h1, d2, d3);
          } else {
            d0 = dact_f(act, nz[q].x); d1 = dact_f(act, nz[q].y); d2 = dact_f(act, nz[q].z); d3 = dact_f(act, nz[q].w);
          }
          cur[q].x *= d0; cur[q].y *= d1; cur[q].z *= d2; cur[q].w *= d3;
        }
        if (xout) {
          float* o0 = xout + row0 * W + w0 + 4 * t + 32 * gi;
          float* o1 = xout + row1 * W + w0 + 4 * t + 32 * gi;
          if (ok0) {
            *reinterpret_cast<float4*>(o0) = cur[0];
            *reinterpret_cast<float4*>(o0 + 16) = cur[1];
          }
          if (ok1) {
            *reinterpret_cast<float4*>(o1) = cur[2];
            *reinterpret_cast<float4*>(o1 + 16) = cur[3];
          }
        }
      }
      if (gi + 1 < groups) load(gi + 1, nv, nz);
      // A fragments of the four k-steps: a0 = (row g, k = t), a1 = (row g + 8, t), a2 = (row g, t + 4), a3 = (row g + 8, t + 4)
      uint32_t ah[4][4], al[4][4];
      {
        const float a0[4] = {cur[0].x, cur[0].y, cur[0].z, cur[0].w}, a2[4] = {cur[1].x, cur[1].y, cur[1].z, cur[1].w};
        const float a1[4] = {cur[2].x, cur[2].y, cur[2].z, cur[2].w}, a3[4] = {cur[3].x, cur[3].y, cur[3].z, cur[3].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          tf32_split(a0[j], ah[j][0], al[j][0]);
          tf32_split(a1[j], ah[j][1], al[j][1]);
          tf32_split(a2[j], ah[j][2], al[j][2]);
          tf32_split(a3[j], ah[j][3], al[j][3]);
        }
      }
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        if (nt < NT) {
          const float* th = tabH + (8 * nt + g) * LDT + 32 * gi + 8 * t;
          const float* tl = tabL + (8 * nt + g) * LDT + 32 * gi + 8 * t;
          const float4 h0 = *reinterpret_cast<const float4*>(th), h1 = *reinterpret_cast<const float4*>(th + 4);
          const float bv0[4] = {h0.x, h0.z, h1.x, h1.z}, bv1[4] = {h0.y, h0.w, h1.y, h1.w};   // k = t | k = t + 4 of k-steps j = 0..3
          uint32_t bh0[4], bh1[4], bl0[4], bl1[4];
          if (PRESPLIT) {
            const float4 l0 = *reinterpret_cast<const float4*>(tl), l1 = *reinterpret_cast<const float4*>(tl + 4);
            const float lv0[4] = {l0.x, l0.z, l1.x, l1.z}, lv1[4] = {l0.y, l0.w, l1.y, l1.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              bh0[j] = __float_as_uint(bv0[j]); bh1[j] = __float_as_uint(bv1[j]);
              bl0[j] = __float_as_uint(lv0[j]); bl1[j] = __float_as_uint(lv1[j]);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              tf32_split(bv0[j], bh0[j], bl0[j]);
              tf32_split(bv1[j], bh1[j], bl1[j]);
            }
          }
          float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < 4; ++j) mma_3xtf32(acc, ah[j], al[j], bh0[j], bh1[j], bl0[j], bl1[j]);
#pragma unroll
          for (int c = 0; c < 4; ++c) sum[nt][c] += acc[c];
        }
      }
    }
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      const int r = 8 * nt + 2 * t;
      if (nt < NT) {
        if (gridDim.y > 1) {
          if (ok0) {
            if (r < R) atomicAdd(T + row0 * s_row + r * s_r, sum[nt][0]);
            if (r + 1 < R) atomicAdd(T + row0 * s_row + (r + 1) * s_r, sum[nt][1]);
          }
          if (ok1) {
            if (r < R) atomicAdd(T + row1 * s_row + r * s_r, sum[nt][2]);
            if (r + 1 < R) atomicAdd(T + row1 * s_row + (r + 1) * s_r, sum[nt][3]);
          }
          continue;
        }
        if (ok0) {
          if (r < R) T[row0 * s_row + r * s_r] = sum[nt][0];
          if (r + 1 < R) T[row0 * s_row + (r + 1) * s_r] = sum[nt][1];
        }
        if (ok1) {
          if (r < R) T[row1 * s_row + r * s_r] = sum[nt][2];
          if (r + 1 < R) T[row1 * s_row + (r + 1) * s_r] = sum[nt][3];
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// H-axis kernels of FNOBlock2d on the warp-level tensor path.  Both read their table as [(h, r)][a], a < A = 4 m1 (the C2C
// analysis / synthesis of the kept rows [0, m1) U [H - m1, H), spectral_convolution.py:82,97); the fp32-FMA versions
// (fno2d_rows_kernel / fno2d_expand_kernel) were latency-bound at 31 / 42 us for 150 MFLOP and 12.6 MB.
//
// fno2d_expand_mma_kernel:  Z[(2 ky + r)][b * H + h][c] = sum_a tab[(h * 2 + r)][a] * in[(b, c)][a][ky]
// as D[(h, r)][c] = A[(h, r)][a] B[a][c] per (b, ky).  A warp owns 16 table rows (8 image rows x {re, im}) with its A
// fragments resident in registers as TF32 hi / lo; it loops over channel tiles and ky quads, reading B straight from global
// memory: for a fixed (c, a) the ky values are contiguous, so three float4 per lane carry all twelve ky of a fragment element.
// CTA = four warps = 32 image rows of one batch element.  KS = k-steps of eight a (A <= 8 KS).  Ky % 4 == 0, C % 8 == 0.
template <int KS>
__global__ void __launch_bounds__(128) fno2d_expand_mma_kernel(const float* __restrict__ in, const float* __restrict__ tab,
                                                               float* __restrict__ Z, int Bn, int C, int H, int Ky, int A) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int b = blockIdx.y;
  const int row_base = (blockIdx.x * 4 + warp) * 16;          // first table row (h * 2 + r) of this warp
  if (row_base >= 2 * H) return;
  const long long BH = (long long)Bn * H;
  uint32_t ah[KS][4], al[KS][4];
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int row = row_base + g + 8 * (q & 1), a = 8 * ks + t + 4 * (q >> 1);
      const float v = (row < 2 * H && a < A) ? __ldg(tab + (long long)row * A + a) : 0.f;
      tf32_split_rn(v, ah[ks][q], al[ks][q]);
    }
  }
  // rows of the accumulator fragment: g -> table row row_base + g, g + 8 -> row_base + g + 8
  const int tr0 = row_base + g, tr1 = tr0 + 8;
  const int h0 = tr0 >> 1, r0 = tr0 & 1, h1 = tr1 >> 1, r1 = tr1 & 1;
  // gridDim.z splits the channel tiles (small images have few table rows: more CTAs instead of longer loops)
  for (int c0 = 8 * blockIdx.z; c0 < C; c0 += 8 * gridDim.z) {
    const int c = c0 + g;                                      // B fragment column (n = g)
    const float* src = in + ((long long)b * C + c) * A * Ky;
    for (int kq = 0; kq < Ky; kq += 4) {
      float acc[4][4];
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[k][q] = 0.f;
      float4 b0[KS], b1[KS];
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        const int a0 = 8 * ks + t, a1 = a0 + 4;
        b0[ks] = a0 < A ? __ldg(reinterpret_cast<const float4*>(src + (long long)a0 * Ky + kq)) : make_float4(0.f, 0.f, 0.f, 0.f);
        b1[ks] = a1 < A ? __ldg(reinterpret_cast<const float4*>(src + (long long)a1 * Ky + kq)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        const float v0[4] = {b0[ks].x, b0[ks].y, b0[ks].z, b0[ks].w}, v1[4] = {b1[ks].x, b1[ks].y, b1[ks].z, b1[ks].w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          uint32_t bh0, bl0, bh1, bl1;
          tf32_split(v0[k], bh0, bl0);
          tf32_split(v1[k], bh1, bl1);
          mma_3xtf32(acc[k], ah[ks], al[ks], bh0, bh1, bl0, bl1);
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int ky = kq + k;
        if (ky < Ky) {
          if (h0 < H)
            *reinterpret_cast<float2*>(Z + ((long long)(2 * ky + r0) * BH + (long long)b * H + h0) * C + c0 + 2 * t) =
                make_float2(acc[k][0], acc[k][1]);
          if (h1 < H)
            *reinterpret_cast<float2*>(Z + ((long long)(2 * ky + r1) * BH + (long long)b * H + h1) * C + c0 + 2 * t) =
                make_float2(acc[k][2], acc[k][3]);
        }
      }
    }
  }
}

// fno2d_rows_mma_kernel:  out[p][a][ky] = sum_(h, r) tab[(h * 2 + r)][a] * T[(p * H + h)][2 ky + r]     p = (b, c) plane
// as D[a][(p, ky)] = A[a][k] B[k][(p, ky)] with k = (h, r).  CTA = two planes; its eight warps split the 2 H reduction
// (every warp: all three 16-row tiles of a x both planes x its k range), add their partial tiles in shared memory with
// round-to-nearest adds and write the two planes' (a, ky) tiles as contiguous rows.  A <= 48 + ... : MT = ceil(A / 16) <= 4
// row tiles; Ky <= 12 (n-tile pairs per plane: 16 columns, ky >= Ky padded with zero).
template <int MT>
__global__ void __launch_bounds__(256) fno2d_rows_mma_kernel(const float* __restrict__ T, const float* __restrict__ tab,
                                                             float* __restrict__ out, long long planes, int H, int Ky, int A) {
  extern __shared__ float red_raw[];                    // per warp: [a][(plane, ky16)], row stride 33
  float (*red)[MT * 16][33] = reinterpret_cast<float (*)[MT * 16][33]>(red_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const long long p0 = (long long)blockIdx.x * 2;
  const int R = 2 * Ky, K = 2 * H, ksteps = (K + 7) / 8;
  const int per = (ksteps + 7) / 8;                      // k-steps per warp
  float acc[MT][4][4];
#pragma unroll
  for (int m = 0; m < MT; ++m)
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[m][n][q] = 0.f;
  // raw operands of one k-step: every global load is issued before the first value is used, and the next k-step's loads go
  // out before the current k-step's MMAs (the loop is a chain of L2 round trips otherwise)
  const int ks_end = min(ksteps, (warp + 1) * per);
  // per-lane operand bases, so that a load is one multiply-add away (the address arithmetic of 20 predicated scalar loads per
  // k-step was most of this kernel's instructions):
  //   B: k = 8 ks + t + 4 q -> row h = 4 ks + 2 q + (t >> 1), r = t & 1; column n = g of tile nt -> plane nt >> 1, ky = 8 (nt & 1) + g
  //   A: a = 16 m + g + 8 (q & 1), k = 8 ks + t + 4 (q >> 1)
  const float* Tb[4];
  bool okb[4];
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) {
    const long long p = p0 + (nt >> 1);
    const int ky = 8 * (nt & 1) + g;
    okb[nt] = p < planes && ky < Ky;
    Tb[nt] = T + (okb[nt] ? (p * H + (t >> 1)) * R + 2 * ky + (t & 1) : 0);
  }
  int ao[MT][4];
  bool oka[MT][2];
#pragma unroll
  for (int m = 0; m < MT; ++m) {
#pragma unroll
    for (int q = 0; q < 4; ++q) ao[m][q] = (t + 4 * (q >> 1)) * A + 16 * m + g + 8 * (q & 1);
    oka[m][0] = 16 * m + g < A;
    oka[m][1] = 16 * m + g + 8 < A;
  }
  auto load = [&](int ks, float (&braw)[4][2], float (&araw)[MT][4]) {
    const bool k0 = 8 * ks + t < K, k1 = 8 * ks + t + 4 < K;
    const long long rb = (long long)(4 * ks) * R;
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      braw[nt][0] = (okb[nt] && k0) ? __ldg(Tb[nt] + rb) : 0.f;
      braw[nt][1] = (okb[nt] && k1) ? __ldg(Tb[nt] + rb + 2 * R) : 0.f;
    }
    const float* ta = tab + (long long)(8 * ks) * A;
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
      for (int q = 0; q < 4; ++q) araw[m][q] = (oka[m][q & 1] && ((q >> 1) ? k1 : k0)) ? __ldg(ta + ao[m][q]) : 0.f;
  };
  float bnext[4][2], anext[MT][4];
  if (warp * per < ks_end) load(warp * per, bnext, anext);
  for (int ks = warp * per; ks < ks_end; ++ks) {
    float braw[4][2], araw[MT][4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      braw[nt][0] = bnext[nt][0];
      braw[nt][1] = bnext[nt][1];
    }
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
      for (int q = 0; q < 4; ++q) araw[m][q] = anext[m][q];
    if (ks + 1 < ks_end) load(ks + 1, bnext, anext);
    uint32_t bh[4][2], bl[4][2];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      tf32_split(braw[nt][0], bh[nt][0], bl[nt][0]);
      tf32_split(braw[nt][1], bh[nt][1], bl[nt][1]);
    }
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      uint32_t ah[4], al[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) tf32_split(araw[m][q], ah[q], al[q]);
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) mma_3xtf32(acc[m][nt], ah, al, bh[nt][0], bh[nt][1], bl[nt][0], bl[nt][1]);
    }
  }
#pragma unroll
  for (int m = 0; m < MT; ++m)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      red[warp][16 * m + g][8 * nt + 2 * t] = acc[m][nt][0];
      red[warp][16 * m + g][8 * nt + 2 * t + 1] = acc[m][nt][1];
      red[warp][16 * m + g + 8][8 * nt + 2 * t] = acc[m][nt][2];
      red[warp][16 * m + g + 8][8 * nt + 2 * t + 1] = acc[m][nt][3];
    }
  __syncthreads();
  for (int e = threadIdx.x; e < 2 * A * Ky; e += 256) {
    const int pl = e / (A * Ky), rem = e - pl * (A * Ky);
    const int a = rem / Ky, ky = rem - a * Ky;
    if (p0 + pl >= planes) continue;
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) v += red[w][a][16 * pl + ky];
    out[(p0 + pl) * A * Ky + rem] = v;
  }
}

}  // namespace rpde
