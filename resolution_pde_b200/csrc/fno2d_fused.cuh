// Fused FNOBlock2d (fno_blocks.py:63-71 around spectral_convolution.py:79-98): the separable truncated transform
//   W axis: R2C keeping m2 modes           H axis: C2C keeping the rows [0, m1) U [H - m1, H)   (J = 2 m1 rows)
// as five fp32-FMA launches forward instead of ~19 tensor-engine launches of 8..130-row tiles (each a latency-bound
// 60-190 us contraction at the reference's sizes).  The K-space tile of one (b, c) plane is J x m2 complex numbers (2.3 KB at
// 24 x 12), so nothing but the W-axis half transform (2 m2 floats per image row, 9 % of the activation) and that tile ever
// leaves the SM between the read of x and the write of y:
//
//   1. fno2d_waxis_kernel over the B*C*H image rows (persistent, table in smem): T[(b,c,h)][(ky,ri)] = sum_w x[b,c,h,w] fa[(ky,ri)][w]
//   2. fno2d_rows_kernel, CTA = one (b, c) plane:                            spec[(b,c)][(j,ro)][ky] = sum_(h,ri) g2[(j,ro)][(ri,h)] T[(b,c,h)][(ky,ri)]
//   3. fno2d_mix_kernel, thread = (b, o, j, ky):                             om[(b,o)][(j,ro)][ky] = sum_i spec[(b,i)] (*) w[i,o,j,ky]   (weights1 | weights2)
//   4. fno2d_expand_kernel, thread = (b, h, o):                              Z[(ky,ro)][(b,h)][o] = sum_(j,ri) g3[(h,ro)][(j,ri)] om[(b,o)][(j,ri)][ky]
//   5. fno1d_synth_bypass_kernel with one "batch" per image row (b, h):      y[b,o,h,w] = act( sum_(ky,ro) Z fs[(ky,ro)][w] + sum_i wb[o,i] x[b,i,h,w] + bb[o] )
//
// The backward pass is the same five kernels with the transposed tables (the adjoint of 5 is 1 applied to dy * act'(z),
// the adjoint of 1 is 5 with the bypass weight transposed), plus two mode-space kernels for dspec and the weight gradient,
// which is written straight into the complex (I, O, m1, m2) parameter layout: no packed-weight workspace in either direction.
#pragma once
#include "common.cuh"

namespace rpde {

// lanes l and l + 16 receive the sum over the warp's 32 lanes of v[l & 15] (16 shuffles instead of 16 x 5); v is destroyed
__device__ __forceinline__ float f2_colsum16(float (&v)[16], int lane) {
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

// W-axis pass over image rows:  T[row][r] = sum_w x[row][w] * tab[r][w]   for r < R = 2 Ky <= 32.
// Persistent CTAs: the whole table sits in shared memory (loaded once per CTA), every warp takes two image rows at a time
// (each table value read from shared memory feeds two FMAs) and walks the table rows in groups of eight: 2 x 8 partial
// sums per lane, reduced by one 16-shuffle transpose-reduction per group.
// Backward use (adjoint of the W-axis synthesis): z / act = pre-activation and activation of the block, the row is multiplied by
// act'(z) on load and written once to xout; kplanes > 0: the table is the plane-major synthesis table [ro][k][w] (kind 4) and
// table row ro * K + k feeds output r = 2 k + ro.
template <int NPL>
__global__ void __launch_bounds__(256, (NPL <= 8 ? 4 : (NPL <= 16 ? 2 : 1))) fno2d_waxis_kernel(const float* __restrict__ x, const float* __restrict__ tab, float* __restrict__ T,
                                                             long long rows, int N, int R, const float* __restrict__ z, int act,
                                                             float* __restrict__ xout, int kplanes) {
  extern __shared__ float st[];                         // [ceil(R / 8) * 8][N], rows beyond R zero
  const int R8 = (R + 7) & ~7;
  for (int e = threadIdx.x; e < R8 * N; e += 256) {
    const int r = e / N, n = e - r * N;
    const int tr = kplanes ? (r & 1) * kplanes + (r >> 1) : r;
    st[e] = r < R ? __ldg(tab + (long long)tr * N + n) : 0.f;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (long long r0 = ((long long)blockIdx.x * 8 + warp) * 2; r0 < rows; r0 += (long long)gridDim.x * 16) {
    float v[2][NPL];
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      const long long row = r0 + s;
      const bool ok = row < rows;
#pragma unroll
      for (int j = 0; j < NPL; ++j) {
        const int n = lane + 32 * j;
        v[s][j] = (ok && n < N) ? __ldg(x + row * N + n) : 0.f;
      }
      if (z) {
        float zz[NPL];
#pragma unroll
        for (int j = 0; j < NPL; ++j) {
          const int n = lane + 32 * j;
          zz[j] = (ok && n < N) ? __ldg(z + row * N + n) : 0.f;
        }
#pragma unroll
        for (int j = 0; j < NPL; ++j) {
          const int n = lane + 32 * j;
          v[s][j] *= dact_f(act, zz[j]);
          if (xout && ok && n < N) xout[row * N + n] = v[s][j];
        }
      }
    }
    for (int g = 0; g < R8; g += 8) {
      float p[16];
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float* t = st + (g + r) * N + lane;
        float a0 = 0.f, a1 = 0.f;
#pragma unroll
        for (int j = 0; j < NPL; ++j) {
          const float tv = (lane + 32 * j < N) ? t[32 * j] : 0.f;
          a0 = fmaf(v[0][j], tv, a0);
          a1 = fmaf(v[1][j], tv, a1);
        }
        p[r] = a0;
        p[8 + r] = a1;
      }
      const float sum = f2_colsum16(p, lane);          // lanes 0-7: row r0, table rows g..g+7; lanes 8-15: row r0 + 1
      const int l = lane & 15, s = l >> 3, r = g + (l & 7);
      if (lane < 16 && r < R && r0 + s < rows) T[(r0 + s) * R + r] = sum;
    }
  }
}

// H-axis contraction per plane:  out[p][a][ky] = sum_{h, r} tab[(h * 2 + r)][a] * T[(p * H + h)][2 ky + r]      a < A (= 2J)
// CTA = one plane p.  Both operands are staged through shared memory in chunks of F2_RC image rows (the plane's half
// transform: F2_RC x 2 Ky contiguous floats; the table slice: 2 F2_RC x A contiguous floats), so the inner loop has no global
// load: thread = (a, group of KG ky values) with a fastest reads its table value conflict-free and the staged values as broadcasts.
constexpr int F2_RC = 64;
template <int KG>
__global__ void __launch_bounds__(256) fno2d_rows_kernel(const float* __restrict__ T, const float* __restrict__ tab,
                                                         float* __restrict__ out, int H, int Ky, int A) {
  extern __shared__ float sm[];                        // tile [F2_RC][2 Ky] | table slice [2 F2_RC][A]
  const long long p = blockIdx.x;
  const int R = 2 * Ky;
  float* tile = sm;
  float* tb = sm + F2_RC * R;
  const int ngroups = (Ky + KG - 1) / KG;
  const int a = threadIdx.x % A, g = threadIdx.x / A;   // A <= 256
  const bool active = g < ngroups;                      // requires A * ngroups <= 256 (checked on the host)
  const int k0 = g * KG;
  float acc[KG];
#pragma unroll
  for (int k = 0; k < KG; ++k) acc[k] = 0.f;
  for (int h0 = 0; h0 < H; h0 += F2_RC) {
    const int hc = min(F2_RC, H - h0);
    __syncthreads();
    const float* src = T + (p * H + h0) * R;
    for (int e = threadIdx.x; e < hc * R; e += 256) tile[e] = __ldg(src + e);
    const float* ts = tab + (long long)(h0 * 2) * A;
    for (int e = threadIdx.x; e < 2 * hc * A; e += 256) tb[e] = __ldg(ts + e);
    __syncthreads();
    if (active) {
#pragma unroll 4
      for (int h = 0; h < hc; ++h) {
        const float u0 = tb[(2 * h) * A + a], u1 = tb[(2 * h + 1) * A + a];
        const float* s = tile + h * R + 2 * k0;
#pragma unroll
        for (int k = 0; k < KG; ++k) {
          if (k0 + k < Ky) acc[k] = fmaf(u0, s[2 * k], fmaf(u1, s[2 * k + 1], acc[k]));
        }
      }
    }
  }
  if (active) {
#pragma unroll
    for (int k = 0; k < KG; ++k)
      if (k0 + k < Ky) out[(p * A + a) * Ky + k0 + k] = acc[k];
  }
}

// Mode-space products over the channel index.  Spectra are laid out [plane (b, c)][(j, re|im)][ky]; the weights keep the
// parameter layout (I, O, m1, m2) complex, weights1 for rows j < m1 and weights2 for j >= m1.
//   mode 0 (forward):   out[(b,o)] = sum_i in[(b,i)] * w[i,o]                 threads over (b, o, j, ky), loop i
//   mode 1 (adjoint):   out[(b,i)] = sum_o in[(b,o)] * conj(w[i,o])           threads over (b, i, j, ky), loop o
__global__ void __launch_bounds__(256) fno2d_mix_kernel(const float* __restrict__ in, const float* __restrict__ w1,
                                                        const float* __restrict__ w2, float* __restrict__ out, int Bn, int I, int O,
                                                        int m1, int Ky, int mode) {
  const int J = 2 * m1;
  const int Co = mode ? I : O, Cl = mode ? O : I;        // output channels, looped channels
  const long long total = (long long)Bn * Co * J * Ky;
  for (long long t = (long long)blockIdx.x * 256 + threadIdx.x; t < total; t += (long long)gridDim.x * 256) {
    const int ky = (int)(t % Ky);
    long long r = t / Ky;
    const int j = (int)(r % J);
    r /= J;
    const int co = (int)(r % Co);
    const int b = (int)(r / Co);
    const float* w = j < m1 ? w1 : w2;
    const int jj = j < m1 ? j : j - m1;
    float re = 0.f, im = 0.f;
#pragma unroll 8   // eight channels' loads in flight per round trip (the loop is a pure L2-latency chain otherwise)
    for (int cl = 0; cl < Cl; ++cl) {
      const int i = mode ? co : cl, o = mode ? cl : co;
      const float* sp = in + (((long long)b * Cl + cl) * (2 * J) + 2 * j) * Ky + ky;
      const float sr = __ldg(sp), si = __ldg(sp + Ky);
      const float2 wv = __ldg(reinterpret_cast<const float2*>(w + ((((long long)i * O + o) * m1 + jj) * Ky + ky) * 2));
      if (!mode) {
        re = fmaf(sr, wv.x, re); re = fmaf(-si, wv.y, re);
        im = fmaf(sr, wv.y, im); im = fmaf(si, wv.x, im);
      } else {
        re = fmaf(sr, wv.x, re); re = fmaf(si, wv.y, re);
        im = fmaf(si, wv.x, im); im = fmaf(-sr, wv.y, im);
      }
    }
    float* op = out + (((long long)b * Co + co) * (2 * J) + 2 * j) * Ky + ky;
    op[0] = re;
    op[Ky] = im;
  }
}

// dw[i,o,j,ky] = sum_b conj(spec[(b,i)][j][ky]) * dom[(b,o)][j][ky], written into the complex parameter layouts of weights1 / weights2
__global__ void __launch_bounds__(256) fno2d_wgrad_kernel(const float* __restrict__ spec, const float* __restrict__ dom,
                                                          float* __restrict__ dw1, float* __restrict__ dw2, int Bn, int I, int O,
                                                          int m1, int Ky) {
  const int J = 2 * m1;
  const long long total = (long long)I * O * J * Ky;
  for (long long t = (long long)blockIdx.x * 256 + threadIdx.x; t < total; t += (long long)gridDim.x * 256) {
    const int ky = (int)(t % Ky);
    long long r = t / Ky;
    const int j = (int)(r % J);
    r /= J;
    const int o = (int)(r % O);
    const int i = (int)(r / O);
    float re = 0.f, im = 0.f;
#pragma unroll 8
    for (int b = 0; b < Bn; ++b) {
      const float* sp = spec + (((long long)b * I + i) * (2 * J) + 2 * j) * Ky + ky;
      const float* dp = dom + (((long long)b * O + o) * (2 * J) + 2 * j) * Ky + ky;
      const float sr = __ldg(sp), si = __ldg(sp + Ky), dr = __ldg(dp), di = __ldg(dp + Ky);
      re = fmaf(sr, dr, re); re = fmaf(si, di, re);
      im = fmaf(sr, di, im); im = fmaf(-si, dr, im);
    }
    float* dw = j < m1 ? dw1 : dw2;
    const int jj = j < m1 ? j : j - m1;
    float* op = dw + ((((long long)i * O + o) * m1 + jj) * Ky + ky) * 2;
    op[0] = re;
    op[1] = im;
  }
}

// ---- the same two mode-space products, tiled through shared memory ---------------------------------------------------------
// The thread-per-output kernels above re-read every spectrum value once per output channel and every weight once per batch
// element through L1: 2.85 M sectors (91 MB) for 3.6 MB of data, 17-23 us whatever the image size -- a quarter of the
// mixed-resolution pass (BASELINE config 4).  Here a CTA owns one kept row j, all ky, a chunk of 8 output channels (and 16
// batch elements); operands are staged once per CTA with 48-96-byte runs, the inner loop reads only shared memory.
constexpr int F2X_KY = 16;   // ky slots per row (Ky <= 16)
constexpr int F2X_OC = 8;    // output channels per CTA
constexpr int F2X_BC = 16;   // batch elements per CTA (mix) / per staging round (weight gradient)
constexpr int F2X_CL = 32;   // looped channels per staging round

// mode 0: out[(b,o)] = sum_i in[(b,i)] * w[i,o];  mode 1: out[(b,i)] = sum_o in[(b,o)] * conj(w[i,o])
// grid (J, ceil(Co / 8), ceil(B / 8)); thread = (ky, b of 8, half of the 8 output channels), four outputs per thread;
// smem: S[cl][ri][b][16] | Wt[cl][c][ky][re|im].  VEC: Ky % 4 == 0 and 16-byte aligned operands -> float4 staging
// (a spectrum row is Ky floats, a weight row Ky complex numbers: 3 and 6 float4 at Ky = 12).
constexpr int F2X_MB = 8;    // batch elements per CTA of the mixing kernel
constexpr int F2X_MCL = 32;  // looped channels per staging round of the mixing kernel (64 KB of dynamic shared memory: one round at width 32)
template <bool VEC>
__global__ void __launch_bounds__(256) fno2d_mix_tiled_kernel(const float* __restrict__ in, const float* __restrict__ w1,
                                                              const float* __restrict__ w2, float* __restrict__ out, int Bn, int I,
                                                              int O, int m1, int Ky, int mode) {
  extern __shared__ __align__(16) float mixsm[];
  float* S = mixsm;                                         // F2X_MCL * 2 * F2X_MB * F2X_KY   (32 KB)
  float* Wt = mixsm + F2X_MCL * 2 * F2X_MB * F2X_KY;        // F2X_MCL * F2X_OC * F2X_KY * 2   (32 KB)
  const int J = 2 * m1;
  const int Co = mode ? I : O, Cl = mode ? O : I;
  const int j = blockIdx.x, co0 = blockIdx.y * F2X_OC, b0 = blockIdx.z * F2X_MB;
  const float* w = j < m1 ? w1 : w2;
  const int jj = j < m1 ? j : j - m1;
  const int ky = threadIdx.x & 15, bl = (threadIdx.x >> 4) & 7, ch = threadIdx.x >> 7;
  float re[4], im[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) re[c] = im[c] = 0.f;
  for (int cl0 = 0; cl0 < Cl; cl0 += F2X_MCL) {
    const int ncl = min(F2X_MCL, Cl - cl0);
    __syncthreads();
    if (VEC) {
      const int q4 = Ky / 4;                                   // float4 per spectrum row; 2 q4 per weight row
      for (int e = threadIdx.x; e < ncl * 2 * F2X_MB * q4; e += 256) {
        const int q = e % q4, row = e / q4, b = row & 7, ri = (row >> 3) & 1, cl = row >> 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (b0 + b < Bn)
          v = __ldg(reinterpret_cast<const float4*>(in + ((((long long)(b0 + b) * Cl + cl0 + cl) * J + j) * 2 + ri) * Ky) + q);
        *reinterpret_cast<float4*>(S + row * F2X_KY + 4 * q) = v;
      }
      for (int e = threadIdx.x; e < ncl * F2X_OC * 2 * q4; e += 256) {
        const int q = e % (2 * q4), row = e / (2 * q4), c = row & 7, cl = row >> 3;
        const int i = mode ? co0 + c : cl0 + cl, o = mode ? cl0 + cl : co0 + c;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (co0 + c < Co) v = __ldg(reinterpret_cast<const float4*>(w + (((long long)i * O + o) * m1 + jj) * Ky * 2) + q);
        *reinterpret_cast<float4*>(Wt + row * F2X_KY * 2 + 4 * q) = v;
      }
    } else {
      for (int e = threadIdx.x; e < ncl * 2 * F2X_MB * F2X_KY; e += 256) {
        const int k = e & 15, row = e >> 4, b = row & 7, ri = (row >> 3) & 1, cl = row >> 4;
        float v = 0.f;
        if (k < Ky && b0 + b < Bn) v = __ldg(in + ((((long long)(b0 + b) * Cl + cl0 + cl) * J + j) * 2 + ri) * Ky + k);
        S[e] = v;
      }
      for (int e = threadIdx.x; e < ncl * F2X_OC * F2X_KY * 2; e += 256) {
        const int k2 = e & 31, row = e >> 5, c = row & 7, cl = row >> 3;
        const int i = mode ? co0 + c : cl0 + cl, o = mode ? cl0 + cl : co0 + c;
        float v = 0.f;
        if (k2 < 2 * Ky && co0 + c < Co) v = __ldg(w + (((long long)i * O + o) * m1 + jj) * Ky * 2 + k2);
        Wt[e] = v;
      }
    }
    __syncthreads();
#pragma unroll 4
    for (int cl = 0; cl < ncl; ++cl) {
      const float sr = S[((cl * 2 + 0) * F2X_MB + bl) * F2X_KY + ky], si = S[((cl * 2 + 1) * F2X_MB + bl) * F2X_KY + ky];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float2 wv = *reinterpret_cast<const float2*>(Wt + ((cl * F2X_OC + 4 * ch + c) * F2X_KY + ky) * 2);
        if (!mode) {
          re[c] = fmaf(sr, wv.x, re[c]); re[c] = fmaf(-si, wv.y, re[c]);
          im[c] = fmaf(sr, wv.y, im[c]); im[c] = fmaf(si, wv.x, im[c]);
        } else {
          re[c] = fmaf(sr, wv.x, re[c]); re[c] = fmaf(si, wv.y, re[c]);
          im[c] = fmaf(si, wv.x, im[c]); im[c] = fmaf(-sr, wv.y, im[c]);
        }
      }
    }
  }
  if (ky < Ky && b0 + bl < Bn) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int co = co0 + 4 * ch + c;
      if (co < Co) {
        float* op = out + ((((long long)(b0 + bl) * Co + co) * J + j) * 2) * Ky + ky;
        op[0] = re[c];
        op[Ky] = im[c];
      }
    }
  }
}

// dw[i,o,j,ky] = sum_b conj(spec[(b,i)][j][ky]) * dom[(b,o)][j][ky]
// grid (J, ceil(I / 8), ceil(O / 8)); thread = (ky, i of 8, half of the 8 o); smem: S[b][i][ri][16] | D[b][o][ro][16]
__global__ void __launch_bounds__(256) fno2d_wgrad_tiled_kernel(const float* __restrict__ spec, const float* __restrict__ dom,
                                                                float* __restrict__ dw1, float* __restrict__ dw2, int Bn, int I,
                                                                int O, int m1, int Ky) {
  __shared__ float S[F2X_BC * F2X_OC * 2 * F2X_KY];
  __shared__ float D[F2X_BC * F2X_OC * 2 * F2X_KY];
  const int J = 2 * m1;
  const int j = blockIdx.x, i0 = blockIdx.y * F2X_OC, o0 = blockIdx.z * F2X_OC;
  const int ky = threadIdx.x & 15, il = (threadIdx.x >> 4) & 7, oh = threadIdx.x >> 7;
  float re[4], im[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) re[c] = im[c] = 0.f;
  for (int b0 = 0; b0 < Bn; b0 += F2X_BC) {
    const int nb = min(F2X_BC, Bn - b0);
    __syncthreads();
    for (int e = threadIdx.x; e < nb * F2X_OC * 2 * F2X_KY; e += 256) {
      const int k = e & 15, ri = (e >> 4) & 1, c = (e >> 5) & 7, b = e >> 8;
      float vs = 0.f, vd = 0.f;
      if (k < Ky && i0 + c < I) vs = __ldg(spec + ((((long long)(b0 + b) * I + i0 + c) * J + j) * 2 + ri) * Ky + k);
      if (k < Ky && o0 + c < O) vd = __ldg(dom + ((((long long)(b0 + b) * O + o0 + c) * J + j) * 2 + ri) * Ky + k);
      S[e] = vs;
      D[e] = vd;
    }
    __syncthreads();
#pragma unroll 4
    for (int b = 0; b < nb; ++b) {
      const float sr = S[((b * F2X_OC + il) * 2 + 0) * F2X_KY + ky], si = S[((b * F2X_OC + il) * 2 + 1) * F2X_KY + ky];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float dr = D[((b * F2X_OC + 4 * oh + c) * 2 + 0) * F2X_KY + ky], di = D[((b * F2X_OC + 4 * oh + c) * 2 + 1) * F2X_KY + ky];
        re[c] = fmaf(sr, dr, re[c]); re[c] = fmaf(si, di, re[c]);
        im[c] = fmaf(sr, di, im[c]); im[c] = fmaf(-si, dr, im[c]);
      }
    }
  }
  const int i = i0 + il;
  if (ky < Ky && i < I) {
    float* dw = j < m1 ? dw1 : dw2;
    const int jj = j < m1 ? j : j - m1;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int o = o0 + 4 * oh + c;
      if (o < O) *reinterpret_cast<float2*>(dw + ((((long long)i * O + o) * m1 + jj) * Ky + ky) * 2) = make_float2(re[c], im[c]);
    }
  }
}

// Z[(2 ky + r)][b * H + h][c] = sum_{a} tab[(h * 2 + r)][a] * in[(b, c)][a][ky]         a < A (= 2J), r in {0, 1}
// grid (ceil(H / 16), B), 256 threads = 8 image rows x 32 channels, two row groups per CTA.  The CTA's 32 table rows sit in
// shared memory (broadcast reads in the inner loop, no global load there); the plane tiles in[(b, c)] are staged transposed
// to [a][ky][c] (row stride 33: the transposing stores are conflict-free) in chunks of F2_AC values of a, so the inner loop
// reads conflict-free shared memory (channel = lane) and the results leave as coalesced 128-byte stores.
constexpr int F2_AC = 16;
constexpr int F2_EH = 16;   // image rows per CTA
template <int KYT>
__global__ void __launch_bounds__(256) fno2d_expand_kernel(const float* __restrict__ in, const float* __restrict__ tab,
                                                           float* __restrict__ Z, int Bn, int C, int H, int Ky, int A) {
  extern __shared__ float sm[];                        // [F2_AC * KYT][33] | table rows [2 F2_EH][A]
  float* ts = sm + F2_AC * KYT * 33;
  const int b = blockIdx.y, h0 = blockIdx.x * F2_EH;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long BH = (long long)Bn * H;
  for (int e = threadIdx.x; e < 2 * F2_EH * A; e += 256) {
    const int r = e / A;
    ts[e] = (h0 + r / 2 < H) ? __ldg(tab + (long long)h0 * 2 * A + e) : 0.f;
  }
  for (int c0 = 0; c0 < C; c0 += 32) {
    const int c = c0 + lane;
    for (int k0 = 0; k0 < Ky; k0 += KYT) {
      float a0[2][KYT], a1[2][KYT];
#pragma unroll
      for (int q = 0; q < 2; ++q)
#pragma unroll
        for (int k = 0; k < KYT; ++k) a0[q][k] = a1[q][k] = 0.f;
      for (int ac = 0; ac < A; ac += F2_AC) {
        const int na = min(F2_AC, A - ac);
        __syncthreads();
        // in[(b, c)][a][ky]: for a fixed channel the (a, ky) block is contiguous; transpose to [a][ky][c]
        for (int e = threadIdx.x; e < na * KYT * 32; e += 256) {
          const int cc = e / (na * KYT), rem = e - cc * (na * KYT);
          const int aa = rem / KYT, k = rem - aa * KYT;
          float v = 0.f;
          if (c0 + cc < C && k0 + k < Ky) v = __ldg(in + (((long long)b * C + c0 + cc) * A + ac + aa) * Ky + k0 + k);
          sm[(aa * KYT + k) * 33 + cc] = v;
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const float* t0 = ts + (2 * (q * 8 + warp)) * A + ac;
          const float* t1 = t0 + A;
#pragma unroll 4
          for (int aa = 0; aa < na; ++aa) {
            const float u0 = t0[aa], u1 = t1[aa];
            const float* s = sm + aa * KYT * 33 + lane;
#pragma unroll
            for (int k = 0; k < KYT; ++k) {
              const float v = s[k * 33];
              a0[q][k] = fmaf(u0, v, a0[q][k]);
              a1[q][k] = fmaf(u1, v, a1[q][k]);
            }
          }
        }
      }
      if (c < C) {
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const int h = h0 + q * 8 + warp;
          if (h >= H) continue;
#pragma unroll
          for (int k = 0; k < KYT; ++k) {
            if (k0 + k >= Ky) break;
            Z[((long long)(2 * (k0 + k)) * BH + (long long)b * H + h) * C + c] = a0[q][k];
            Z[((long long)(2 * (k0 + k) + 1) * BH + (long long)b * H + h) * C + c] = a1[q][k];
          }
        }
      }
    }
  }
}

}  // namespace rpde
