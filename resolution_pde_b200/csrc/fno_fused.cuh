// Fused FNOBlock1d forward (fno_blocks.py:25-33 around spectral_convolution.py:38-55): three launches instead of
// memset + weight pack + three contractions + the bypass conv.  At the reference's sizes (C2: B=16, width 64, N=1024,
// 16 modes) every one of those contractions is a latency-bound 8..128-tile problem (~10 us each); here the retained
// spectrum never takes a detour through a packed-weight workspace and x is read from L2 by both kernels.
//
//   kernel 1 (analysis), warp = one (b, i) row held in registers:
//       spec[k][ri][b][i] = sum_n x[b,i,n] fa[(k,ri)][n]                 (kept for the backward pass)
//   kernel 2 (mode mixing), CTA = (8 output channels, b):
//       om[k][ro][b][o]   = sum_i spec[k][.][b][i] (*) w[i,o,k]          (complex product, w in the parameter layout)
//   kernel 3 (synthesis + bypass 1x1 conv + bias + activation), CTA = (tile of TN grid points, b):
//       y[b,o,n] = act( sum_(k,ro) om[k][ro][b][o] fs[n][(k,ro)] + sum_i wb[o,i] x[b,i,n] + bb[o] )
//
// fp32 FMA throughout (no tensor cores: the K dimensions are 2*modes = 32 and width = 64, and the problem is bound by
// launch latency and L2, not arithmetic), so the result is engine-independent and within 1e-6 of the composed path.
#pragma once
#include "common.cuh"

namespace rpde {

constexpr int FNO_OB = 64;     // output channels per register-tile pass in the synthesis kernel
constexpr int FNO_MO = 8;      // output channels per CTA in the mixing kernel

// Analysis: CTA = (8 rows (b, i), chunk of FNO_RC table rows).  Each warp keeps its x row in registers (NPL values per lane,
// n = lane + 32 j); the table chunk is staged in shared memory by the whole CTA, so every global load of the kernel is
// issued up front (one L2 round trip), then FNO_RC dot products per warp run out of registers and shared memory.
constexpr int FNO_RC = 8;
template <int NPL>
// Backward use (adjoint of the synthesis): x = dy, z / act = the block's pre-activation and activation (the row is multiplied
// by act'(z) on load and written once to xout = dL/d(pre-activation) by the CTAs of the first table chunk), fa = the plane-major
// synthesis table [ro][k][n] (rowmap = 1: table row t = ro*K + k lands in output row 2k + ro).
__global__ void __launch_bounds__(256) fno1d_analysis_kernel(const float* __restrict__ x, const float* __restrict__ fa,
                                                             float* __restrict__ spec, int Bn, int I, int N, int K,
                                                             const float* __restrict__ z, int act, float* __restrict__ xout,
                                                             int rowmap) {
  extern __shared__ float tab[];                                              // [FNO_RC][N]
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);      // b * I + i
  const bool valid = row < (long long)Bn * I;
  const int r0 = blockIdx.y * FNO_RC, rc = min(FNO_RC, 2 * K - r0);
  // the table chunk goes to shared memory with cp.async (no register dependency), so it is in flight together with the row loads
  const float* src = fa + (long long)r0 * N;
  const int total = rc * N;
  const bool vec = (N & 3) == 0;
  if (vec) {
    for (int e = threadIdx.x * 4; e < total; e += 1024) {
      const unsigned dst = (unsigned)__cvta_generic_to_shared(tab + e);
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src + e) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  float v[NPL];
  if (valid) {
    const float* xr = x + row * N;
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
      const int n = lane + 32 * j;
      v[j] = n < N ? __ldg(xr + n) : 0.f;
    }
    if (z) {
      const float* zr = z + row * N;
      float zz[NPL];
#pragma unroll
      for (int j = 0; j < NPL; ++j) {
        const int n = lane + 32 * j;
        zz[j] = n < N ? __ldg(zr + n) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < NPL; ++j) v[j] *= dact_f(act, zz[j]);
    }
    if (xout && blockIdx.y == 0) {
      float* o = xout + row * N;
#pragma unroll
      for (int j = 0; j < NPL; ++j) {
        const int n = lane + 32 * j;
        if (n < N) o[n] = v[j];
      }
    }
  }
  if (vec) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  } else {
    for (int e = threadIdx.x; e < total; e += 256) tab[e] = __ldg(src + e);
  }
  __syncthreads();
  if (!valid) return;
  const int b = (int)(row / I), i = (int)(row - (long long)b * I);
  for (int rr = 0; rr < rc; ++rr) {
    const float* t = tab + rr * N;
    float a0 = 0.f, a1 = 0.f;
#pragma unroll
    for (int j = 0; j < NPL; j += 2) {
      const int n = lane + 32 * j;
      a0 = fmaf(v[j], n < N ? t[n] : 0.f, a0);
      a1 = fmaf(v[j + 1], n + 32 < N ? t[n + 32] : 0.f, a1);
    }
    const float sum = warp_sum(a0 + a1);
    const int r = r0 + rr;
    const int orow = rowmap ? 2 * (r % K) + r / K : r;
    if (lane == 0) spec[((long long)orow * Bn + b) * I + i] = sum;
  }
}

// Mixing: CTA = (block of FNO_MO output channels, b); thread pairs (oo, k).  For a fixed i the weights of the CTA's
// (o, k) pairs are one contiguous run of the parameter tensor (I, O, K, 2), so the loads coalesce without a packing pass.
// smem: sp[2K][I + 1]
__global__ void __launch_bounds__(256) fno1d_mix_kernel(const float* __restrict__ spec, const float* __restrict__ w,
                                                        float* __restrict__ om, int Bn, int I, int O, int K) {
  extern __shared__ float sp[];
  const int b = blockIdx.y, o0 = blockIdx.x * FNO_MO;
  const int ob = min(FNO_MO, O - o0);
  const int ld = I + 1;
  for (int e = threadIdx.x; e < 2 * K * I; e += 256) {
    const int r = e / I, i = e - r * I;
    sp[r * ld + i] = spec[((long long)r * Bn + b) * I + i];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < ob * K; e += 256) {
    const int oo = e / K, k = e - oo * K;
    const float* sr = sp + (2 * k) * ld;
    const float* si = sr + ld;
    const float* wp = w + ((long long)(o0 + oo) * K + k) * 2;
    float re = 0.f, im = 0.f;
#pragma unroll 16
    for (int i = 0; i < I; ++i) {
      const float2 wv = __ldg(reinterpret_cast<const float2*>(wp + (long long)i * O * K * 2));
      re = fmaf(sr[i], wv.x, re);
      re = fmaf(-si[i], wv.y, re);
      im = fmaf(sr[i], wv.y, im);
      im = fmaf(si[i], wv.x, im);
    }
    om[((long long)(2 * k) * Bn + b) * O + o0 + oo] = re;
    om[((long long)(2 * k + 1) * Bn + b) * O + o0 + oo] = im;
  }
}

// smem: A[KT][FNO_OB] (om slice then wbT slice), Bm[KT][TN] (fs^T slice then x slice); KT = chunk of the reduction
constexpr int FNO_KT = 32;

// TN = grid points per CTA (128: 4x8 outputs per thread; 64: 4x4 per thread, twice the CTAs -- chosen by the host so that the
// grid fills the SMs at least twice and neighbouring CTAs overlap their load and FMA phases)
// OB = output channels per register-tile pass (64: 4 x NT outputs per thread over 16 x 16 threads; 32: 8 x 32 threads, for
// layers with at most 32 channels, where a 64-wide tile would spend half of its FMAs on padding)
template <int TN, int OB = FNO_OB>
__global__ void __launch_bounds__(256) fno1d_synth_bypass_kernel(const float* __restrict__ x, const float* __restrict__ om,
                                                                 const float* __restrict__ fs4, const float* __restrict__ wb,
                                                                 const float* __restrict__ bb, float* __restrict__ y,
                                                                 float* __restrict__ z, int Bn, int I, int O, int N, int K,
                                                                 int act, int adjoint, int Hd) {
  // Hd > 1 (FNOBlock2d, fno2d_fused.cuh): the "batch" index is an image row (b, h) of a (B, C, Hd, N) tensor, Bn = B * Hd:
  // spectral rows are indexed by (b, h), the rows of x / y / z by ((b * C + c) * Hd + h)
  // adjoint = 1 (backward, dx = analysis adjoint of dsp + bypass dgrad): the table is the analysis table fa[(k,ri)][n] (row r
  // as is) and the bypass weight is read transposed, wb[(reduction channel)][output channel]
  constexpr int OG = OB / 4, NG = 256 / OG;
  constexpr int NT = TN / NG;            // consecutive grid points per thread (8 or 4)
  static_assert(NT % 4 == 0, "float4 stores");
  __shared__ __align__(16) float As[FNO_KT][OB + 4];
  __shared__ __align__(16) float Bs[FNO_KT][TN + 4];
  const int b = blockIdx.y, n0 = blockIdx.x * TN;
  const int bh = b / Hd, hh = b - bh * Hd;
  const int tid = threadIdx.x;
  const int tn = (tid % NG) * NT;
  const int to = (tid / NG) * 4;        // 4 consecutive output channels
  const int R = 2 * K;
  for (int o0 = 0; o0 < O; o0 += OB) {
    float acc[4][NT];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < NT; ++c) acc[a][c] = 0.f;
    // reduction index r runs over the 2K spectral rows, then (when there is a bypass conv) over the I input channels
    const int RT = R + (wb ? I : 0);
    for (int r0 = 0; r0 < RT; r0 += FNO_KT) {
      __syncthreads();
      // every global load of the chunk is issued before the first shared-memory store (one L2 round trip per chunk)
      float ra[FNO_KT * OB / 256], rb[FNO_KT * TN / 256];
#pragma unroll
      for (int j = 0; j < FNO_KT * OB / 256; ++j) {
        const int e = tid + 256 * j;
        const int rr = e / OB, oo = e - rr * OB;
        const int r = r0 + rr, o = o0 + oo;
        const bool ok = r < RT && o < O;
        const float* p = r < R ? om + ((long long)r * Bn + b) * O + o
                               : (adjoint ? wb + (long long)(r - R) * O + o : wb + (long long)o * I + (r - R));
        ra[j] = ok ? __ldg(p) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < FNO_KT * TN / 256; ++j) {
        const int e = tid + 256 * j;
        const int rr = e / TN, nn = e - rr * TN;
        const int r = r0 + rr, n = n0 + nn;
        const bool ok = r < RT && n < N;      // spectral rows r = (k, ro) come from the plane-major synthesis table fs4[ro][k][n]
        const float* p = r < R ? fs4 + (adjoint ? (long long)r : (long long)(r & 1) * K + (r >> 1)) * N + n
                               : x + (((long long)bh * I + (r - R)) * Hd + hh) * N + n;
        rb[j] = ok ? __ldg(p) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < FNO_KT * OB / 256; ++j) {
        const int e = tid + 256 * j;
        As[e / OB][e % OB] = ra[j];
      }
#pragma unroll
      for (int j = 0; j < FNO_KT * TN / 256; ++j) {
        const int e = tid + 256 * j;
        Bs[e / TN][e % TN] = rb[j];
      }
      __syncthreads();
#pragma unroll 8
      for (int rr = 0; rr < FNO_KT; ++rr) {
        const float4 a = *reinterpret_cast<const float4*>(&As[rr][to]);
        const float av[4] = {a.x, a.y, a.z, a.w};
        float bv[NT];
#pragma unroll
        for (int q = 0; q < NT / 4; ++q) {
          const float4 p = *reinterpret_cast<const float4*>(&Bs[rr][tn + 4 * q]);
          bv[4 * q] = p.x; bv[4 * q + 1] = p.y; bv[4 * q + 2] = p.z; bv[4 * q + 3] = p.w;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < NT; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int o = o0 + to + i;
      if (o >= O) continue;
      const float bias = bb ? bb[o] : 0.f;
      const long long row = (((long long)bh * O + o) * Hd + hh) * N;
      if ((N & 3) == 0 && n0 + tn + NT <= N) {
        float pre[NT], out[NT];
#pragma unroll
        for (int j = 0; j < NT; ++j) {
          pre[j] = acc[i][j] + bias;
          out[j] = act_f(act, pre[j]);
        }
        float4* yp = reinterpret_cast<float4*>(y + row + n0 + tn);
        float4* zp = reinterpret_cast<float4*>(z + row + n0 + tn);
#pragma unroll
        for (int q = 0; q < NT / 4; ++q) {
          yp[q] = make_float4(out[4 * q], out[4 * q + 1], out[4 * q + 2], out[4 * q + 3]);
          if (z) zp[q] = make_float4(pre[4 * q], pre[4 * q + 1], pre[4 * q + 2], pre[4 * q + 3]);
        }
        continue;
      }
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        const int n = n0 + tn + j;
        if (n >= N) continue;
        const float pre = acc[i][j] + bias;
        if (z) z[row + n] = pre;
        y[row + n] = act_f(act, pre);
      }
    }
  }
}

// ---- backward mode-space kernels of SpectralConv1d (spectral_convolution.py:34-36 compl_mul1d, its two adjoints) ------------
// Spectra are mode-major, [k][re|im][b][c]; the weights keep the parameter layout (I, O, K) complex.  They replace
// pack_weights + two batched contractions + memset + unpack (five launches, 26 us per block at batch 16): at these sizes
// (K x B x C = 16 x 16 x 64) every launch is latency, so fewer and simpler launches win.
constexpr int M1_IC = 8;     // input channels per CTA
constexpr int M1_BC = 16;    // batch elements per staging round

// adjoint = 1:  out[k][.][b][i] = sum_o in[k][.][b][o] * conj(w[i,o,k])     (mixing adjoint, Co = I, Cl = O)
// adjoint = 0:  out[k][.][b][o] = sum_i in[k][.][b][i] * w[i,o,k]           (forward mixing,  Co = O, Cl = I)
// grid (K, ceil(Co / 8), ceil(B / 16)), 128 threads = (b, output channel of 8); smem: D[2][M1_BC][Cl + 1] | W[8][2 Cl + 2]
__global__ void __launch_bounds__(128) fno1d_mix_modes_kernel(const float* __restrict__ in, const float* __restrict__ w,
                                                              float* __restrict__ out, int Bn, int I, int O, int K, int adjoint) {
  extern __shared__ float sm1[];                      // padded rows: no bank conflicts over b / over the output channel
  const int Co = adjoint ? I : O, Cl = adjoint ? O : I;
  float* D = sm1;
  float* Wc = sm1 + 2 * M1_BC * (Cl + 1);
  const int k = blockIdx.x, c0 = blockIdx.y * M1_IC, b0 = blockIdx.z * M1_BC;
  const int nb = min(M1_BC, Bn - b0);
  for (int e = threadIdx.x; e < 2 * nb * Cl; e += 128) {
    const int cl = e % Cl, rb = e / Cl, b = rb % nb, ro = rb / nb;
    D[(ro * M1_BC + b) * (Cl + 1) + cl] = __ldg(in + ((long long)(k * 2 + ro) * Bn + b0 + b) * Cl + cl);
  }
  for (int e = threadIdx.x; e < M1_IC * Cl; e += 128) {
    const int cl = e % Cl, c = e / Cl;
    const int i = adjoint ? c0 + c : cl, o = adjoint ? cl : c0 + c;
    float2 v = make_float2(0.f, 0.f);
    if (c0 + c < Co) v = __ldg(reinterpret_cast<const float2*>(w + (((long long)i * O + o) * K + k) * 2));
    Wc[c * (2 * Cl + 2) + 2 * cl] = v.x;
    Wc[c * (2 * Cl + 2) + 2 * cl + 1] = v.y;
  }
  __syncthreads();
  const int c = threadIdx.x & 7, b = threadIdx.x >> 3;
  if (b >= nb || c0 + c >= Co) return;
  const float* dr = D + b * (Cl + 1);
  const float* di = D + (M1_BC + b) * (Cl + 1);
  const float* wp = Wc + c * (2 * Cl + 2);
  float re = 0.f, im = 0.f;
  if (adjoint) {
#pragma unroll 4
    for (int cl = 0; cl < Cl; ++cl) {
      const float wr = wp[2 * cl], wi = wp[2 * cl + 1], a = dr[cl], q = di[cl];
      re = fmaf(a, wr, re); re = fmaf(q, wi, re);
      im = fmaf(q, wr, im); im = fmaf(-a, wi, im);
    }
  } else {
#pragma unroll 4
    for (int cl = 0; cl < Cl; ++cl) {
      const float wr = wp[2 * cl], wi = wp[2 * cl + 1], a = dr[cl], q = di[cl];
      re = fmaf(a, wr, re); re = fmaf(-q, wi, re);
      im = fmaf(a, wi, im); im = fmaf(q, wr, im);
    }
  }
  out[((long long)(k * 2 + 0) * Bn + b0 + b) * Co + c0 + c] = re;
  out[((long long)(k * 2 + 1) * Bn + b0 + b) * Co + c0 + c] = im;
}

// dw[i,o,k] = sum_b conj(spec[k][.][b][i]) * dom[k][.][b][o]       grid (K, ceil(I / 8), ceil(O / 32)), 256 threads = (o, i)
__global__ void __launch_bounds__(256) fno1d_wgrad_kernel(const float* __restrict__ spec, const float* __restrict__ dom,
                                                          float* __restrict__ dw, int Bn, int I, int O, int K) {
  __shared__ float S[2][M1_BC][M1_IC];
  __shared__ float D[2][M1_BC][32];
  const int k = blockIdx.x, i0 = blockIdx.y * M1_IC, o0 = blockIdx.z * 32;
  const int ol = threadIdx.x & 31, il = threadIdx.x >> 5;
  float re = 0.f, im = 0.f;
  for (int b0 = 0; b0 < Bn; b0 += M1_BC) {
    const int nb = min(M1_BC, Bn - b0);
    __syncthreads();
    for (int e = threadIdx.x; e < 2 * nb * M1_IC; e += 256) {
      const int i = e % M1_IC, rb = e / M1_IC, b = rb % nb, ri = rb / nb;
      S[ri][b][i] = i0 + i < I ? __ldg(spec + ((long long)(k * 2 + ri) * Bn + b0 + b) * I + i0 + i) : 0.f;
    }
    for (int e = threadIdx.x; e < 2 * nb * 32; e += 256) {
      const int o = e & 31, rb = e >> 5, b = rb % nb, ro = rb / nb;
      D[ro][b][o] = o0 + o < O ? __ldg(dom + ((long long)(k * 2 + ro) * Bn + b0 + b) * O + o0 + o) : 0.f;
    }
    __syncthreads();
#pragma unroll 4
    for (int b = 0; b < nb; ++b) {
      const float sr = S[0][b][il], si = S[1][b][il], dr = D[0][b][ol], di = D[1][b][ol];
      re = fmaf(sr, dr, re); re = fmaf(si, di, re);
      im = fmaf(sr, di, im); im = fmaf(-si, dr, im);
    }
  }
  if (i0 + il < I && o0 + ol < O)
    *reinterpret_cast<float2*>(dw + (((long long)(i0 + il) * O + o0 + ol) * K + k) * 2) = make_float2(re, im);
}

}  // namespace rpde
