// Fused FNOBlock2d (fno_blocks.py:63-71 around spectral_convolution.py:79-98): the separable truncated transform
//   W axis: R2C keeping m2 modes           H axis: C2C keeping the rows [0, m1) U [H - m1, H)   (J = 2 m1 rows)
// as five fp32-FMA launches forward instead of ~19 tensor-engine launches of 8..130-row tiles (each a latency-bound
// 60-190 us contraction at the reference's sizes).  The K-space tile of one (b, c) plane is J x m2 complex numbers (2.3 KB at
// 24 x 12), so nothing but the W-axis half transform (2 m2 floats per image row, 9 % of the activation) and that tile ever
// leaves the SM between the read of x and the write of y:
//
//   1. fno1d_analysis_kernel (fno_fused.cuh) over the B*C*H image rows:      T[(ky,ri)][(b,c,h)] = sum_w x[b,c,h,w] fa[(ky,ri)][w]
//   2. fno2d_rows_kernel, CTA = one (b, c) plane:                            spec[(b,c)][(j,ro)][ky] = sum_(ri,h) g2[(j,ro)][(ri,h)] T[(ky,ri)][(b,c,h)]
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

constexpr int F2_HC = 256;   // image rows per shared-memory chunk of the H-axis kernels

// out[p][a][ky] = sum_{c = (r, h)} tab[a][r * H + h] * in[(2 ky + r)][p * H + h]      a < A (= 2J), r in {0, 1}
// CTA = one plane p; the plane's half transform (2 Ky rows of H values) is staged through shared memory in chunks of F2_HC rows.
__global__ void __launch_bounds__(256) fno2d_rows_kernel(const float* __restrict__ in, const float* __restrict__ tab,
                                                         float* __restrict__ out, long long planes, int H, int Ky, int A) {
  extern __shared__ float tile[];                      // [2 Ky][F2_HC + 1]
  const long long p = blockIdx.x;
  const int ld = F2_HC + 1;
  const int items = A * Ky;
  const long long rows = planes * H;
  // every thread owns up to 4 (a, ky) outputs (A * Ky <= 1024)
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int h0 = 0; h0 < H; h0 += F2_HC) {
    const int hc = min(F2_HC, H - h0);
    __syncthreads();
    for (int e = threadIdx.x; e < 2 * Ky * hc; e += 256) {
      const int r = e / hc, h = e - r * hc;
      tile[r * ld + h] = __ldg(in + (long long)r * rows + p * H + h0 + h);
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int item = threadIdx.x + 256 * q;
      if (item >= items) break;
      const int a = item / Ky, ky = item - a * Ky;
      const float* t0 = tab + (long long)a * 2 * H + h0;      // r = 0 half of the table row
      const float* t1 = t0 + H;                                // r = 1
      const float* s0 = tile + (2 * ky) * ld;
      const float* s1 = s0 + ld;
      float x0 = 0.f, x1 = 0.f;
#pragma unroll 4
      for (int h = 0; h < hc; ++h) {
        x0 = fmaf(__ldg(t0 + h), s0[h], x0);
        x1 = fmaf(__ldg(t1 + h), s1[h], x1);
      }
      acc[q] += x0 + x1;
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int item = threadIdx.x + 256 * q;
    if (item < items) out[p * items + item] = acc[q];
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

// Z[(2 ky + r)][b * H + h][c] = sum_{a} tab[(h * 2 + r)][a] * in[(b, c)][a][ky]         a < A (= 2J), r in {0, 1}
// grid (ceil(H / 8), B), 256 threads = 8 image rows x 32 channels (channels beyond 32 are looped); KYT = ky values per pass
template <int KYT>
__global__ void __launch_bounds__(256) fno2d_expand_kernel(const float* __restrict__ in, const float* __restrict__ tab,
                                                           float* __restrict__ Z, int Bn, int C, int H, int Ky, int A) {
  const int b = blockIdx.y;
  const int h = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (h >= H) return;
  const long long BH = (long long)Bn * H;
  const float* t0 = tab + (long long)(h * 2) * A;
  const float* t1 = t0 + A;
  for (int c = threadIdx.x & 31; c < C; c += 32) {
    const float* src = in + ((long long)b * C + c) * A * Ky;
    for (int k0 = 0; k0 < Ky; k0 += KYT) {
      float a0[KYT], a1[KYT];
#pragma unroll
      for (int k = 0; k < KYT; ++k) a0[k] = a1[k] = 0.f;
      for (int a = 0; a < A; ++a) {
        const float u0 = __ldg(t0 + a), u1 = __ldg(t1 + a);
        const float* s = src + (long long)a * Ky + k0;
#pragma unroll
        for (int k = 0; k < KYT; ++k) {
          const float v = (k0 + k < Ky) ? __ldg(s + k) : 0.f;
          a0[k] = fmaf(u0, v, a0[k]);
          a1[k] = fmaf(u1, v, a1[k]);
        }
      }
#pragma unroll
      for (int k = 0; k < KYT; ++k) {
        if (k0 + k >= Ky) break;
        Z[((long long)(2 * (k0 + k)) * BH + (long long)b * H + h) * C + c] = a0[k];
        Z[((long long)(2 * (k0 + k) + 1) * BH + (long long)b * H + h) * C + c] = a1[k];
      }
    }
  }
}

}  // namespace rpde
