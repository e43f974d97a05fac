// Descriptor-driven batched fp32 contraction on the CUDA cores (exact fp32 FMA path).
//
// Every stage of the spectral path (truncated DFT analysis / synthesis, per-mode complex
// channel mixing written as a real 2Cx2C block product, the pointwise linears and their
// adjoints) is a dense contraction  C[b] = epi(alpha * A[b] * B[b]).  The operands are
// addressed through two-level index maps (common.cuh: Idx2), so channels-first / channels-last
// layouts, interleaved (re,im) planes and transposes need no copy passes through HBM.
//
// This kernel is the any-shape / any-alignment path and the numerical yardstick of the
// tensor-core path (gemm_tc.cuh), which takes over for TMA-aligned shapes.
#pragma once
#include "common.cuh"

namespace rpde {

struct GemmP {
  const float* A;
  const float* B;
  float* C;
  MatDesc a, b, c;  // A: rows=M cols=K ; B: rows=K cols=N ; C: rows=M cols=N
  int M, N, K, batch;
  int splitk, kchunk;  // splitk>1: partial products are atomically added into C (epilogue = alpha only)
  float alpha;
  int beta;           // 1: add the existing C before the activation
  const float* bias;  // bias_mode 1: per column n, 2: per row m
  int bias_mode;
  int act;            // activation applied after bias/beta
  float* Z;           // optional pre-activation output (same layout as C)
  const float* aux;   // with dact: result *= act'(aux) (same layout as C); dact == 3: aux already holds act'
  int dact;
  int z_grad;         // 1: Z receives act'(pre-activation) instead of the pre-activation
  float* colsum;      // optional: += column sums of the final result (bias gradient of the producing layer), N floats
  int a_act, b_act;   // activation applied to the elements of A / B as they are loaded (operand = act(stored tensor))
  // filled by gemm_launch
  int a_kfast, b_kfast, a_kuni, b_kuni, tiles_m, tiles_n;
};

template <int BM, int BN>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const GemmP p) {
  constexpr int BK = 16;
  constexpr int TM = BM / 16, TN = BN / 16;
  constexpr int LA = BM * BK / 256, LB = BN * BK / 256;
  __shared__ __align__(16) float As[2][BK][BM + 4];
  __shared__ __align__(16) float Bs[2][BK][BN + 4];

  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  long long bid = blockIdx.x;
  const int tn = (int)(bid % p.tiles_n);
  bid /= p.tiles_n;
  const int tm = (int)(bid % p.tiles_m);
  bid /= p.tiles_m;
  const int z = (int)bid;
  const int bz = z / p.splitk, sk = z - bz * p.splitk;
  const int kbeg = sk * p.kchunk;
  const int kend = min(p.K, kbeg + p.kchunk);
  const int m0 = tm * BM, n0 = tn * BN;
  const float* __restrict__ A = p.A + off2(p.a.b, bz);
  const float* __restrict__ B = p.B + off2(p.b.b, bz);

  // ---- loader maps: kfast = 16 consecutive k per half-warp, else consecutive rows/cols per warp
  long long arow[LA], bcol[LB];
  bool aval[LA], bval[LB];
  if (p.a_kfast) {
#pragma unroll
    for (int j = 0; j < LA; ++j) {
      const int m = m0 + (t >> 4) + 16 * j;
      aval[j] = m < p.M;
      arow[j] = off2(p.a.r, aval[j] ? m : 0);
    }
  } else {
    const int m = m0 + (t % BM);
    aval[0] = m < p.M;
    arow[0] = off2(p.a.r, aval[0] ? m : 0);
  }
  if (p.b_kfast) {
#pragma unroll
    for (int j = 0; j < LB; ++j) {
      const int n = n0 + (t >> 4) + 16 * j;
      bval[j] = n < p.N;
      bcol[j] = off2(p.b.c, bval[j] ? n : 0);
    }
  } else {
    const int n = n0 + (t % BN);
    bval[0] = n < p.N;
    bcol[0] = off2(p.b.c, bval[0] ? n : 0);
  }

  float ra[LA], rb[LB];
  auto load_tile = [&](int k0) {
    if (p.a_kfast) {
      const int kk = t & 15, k = k0 + kk;
      const long long ko = p.a_kuni ? off2(p.a.c, k0) + (long long)kk * p.a.c.sl : off2(p.a.c, k);
      const bool kin = k < kend;
#pragma unroll
      for (int j = 0; j < LA; ++j) ra[j] = (kin && aval[j]) ? act_f(p.a_act, __ldg(A + arow[j] + ko)) : 0.f;
    } else {
      const long long kb = p.a_kuni ? off2(p.a.c, k0) : 0;
#pragma unroll
      for (int j = 0; j < LA; ++j) {
        const int kk = t / BM + (256 / BM) * j, k = k0 + kk;
        const long long ko = p.a_kuni ? kb + (long long)kk * p.a.c.sl : off2(p.a.c, k);
        ra[j] = (k < kend && aval[0]) ? act_f(p.a_act, __ldg(A + arow[0] + ko)) : 0.f;
      }
    }
    if (p.b_kfast) {
      const int kk = t & 15, k = k0 + kk;
      const long long ko = p.b_kuni ? off2(p.b.r, k0) + (long long)kk * p.b.r.sl : off2(p.b.r, k);
      const bool kin = k < kend;
#pragma unroll
      for (int j = 0; j < LB; ++j) rb[j] = (kin && bval[j]) ? act_f(p.b_act, __ldg(B + bcol[j] + ko)) : 0.f;
    } else {
      const long long kb = p.b_kuni ? off2(p.b.r, k0) : 0;
#pragma unroll
      for (int j = 0; j < LB; ++j) {
        const int kk = t / BN + (256 / BN) * j, k = k0 + kk;
        const long long ko = p.b_kuni ? kb + (long long)kk * p.b.r.sl : off2(p.b.r, k);
        rb[j] = (k < kend && bval[0]) ? act_f(p.b_act, __ldg(B + bcol[0] + ko)) : 0.f;
      }
    }
  };
  auto store_tile = [&](int buf) {
    if (p.a_kfast) {
#pragma unroll
      for (int j = 0; j < LA; ++j) As[buf][t & 15][(t >> 4) + 16 * j] = ra[j];
    } else {
#pragma unroll
      for (int j = 0; j < LA; ++j) As[buf][t / BM + (256 / BM) * j][t % BM] = ra[j];
    }
    if (p.b_kfast) {
#pragma unroll
      for (int j = 0; j < LB; ++j) Bs[buf][t & 15][(t >> 4) + 16 * j] = rb[j];
    } else {
#pragma unroll
      for (int j = 0; j < LB; ++j) Bs[buf][t / BN + (256 / BN) * j][t % BN] = rb[j];
    }
  };

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  const int nk = kend > kbeg ? (kend - kbeg + BK - 1) / BK : 0;
  if (nk > 0) {
    load_tile(kbeg);
    store_tile(0);
  }
  __syncthreads();
  for (int it = 0; it < nk; ++it) {
    const int buf = it & 1;
    if (it + 1 < nk) load_tile(kbeg + (it + 1) * BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
      {
        const float4 v = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
        a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = v.w;
        if (TM == 8) {
          const float4 w = *reinterpret_cast<const float4*>(&As[buf][kk][BM / 2 + ty * 4]);
          a[TM - 4] = w.x; a[TM - 3] = w.y; a[TM - 2] = w.z; a[TM - 1] = w.w;
        }
      }
      {
        const float4 v = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
        b[0] = v.x; b[1] = v.y; b[2] = v.z; b[3] = v.w;
        if (TN == 8) {
          const float4 w = *reinterpret_cast<const float4*>(&Bs[buf][kk][BN / 2 + tx * 4]);
          b[TN - 4] = w.x; b[TN - 3] = w.y; b[TN - 2] = w.z; b[TN - 1] = w.w;
        }
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (it + 1 < nk) store_tile(buf ^ 1);
    __syncthreads();
  }

  // ---- epilogue
  const long long cb = off2(p.c.b, bz);
  long long ccol[TN];
  int ncol[TN];
#pragma unroll
  for (int j = 0; j < TN; ++j) {
    ncol[j] = n0 + ((TN == 8 && j >= 4) ? BN / 2 + tx * 4 + (j - 4) : tx * 4 + j);
    ccol[j] = off2(p.c.c, ncol[j] < p.N ? ncol[j] : 0);
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + ((TM == 8 && i >= 4) ? BM / 2 + ty * 4 + (i - 4) : ty * 4 + i);
    if (m >= p.M) continue;
    const long long ro = cb + off2(p.c.r, m);
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      if (ncol[j] >= p.N) continue;
      const long long o = ro + ccol[j];
      float v = p.alpha * acc[i][j];
      if (p.splitk > 1) {
        atomicAdd(p.C + o, v);
        continue;
      }
      if (p.bias_mode == 1) v += __ldg(p.bias + ncol[j]);
      else if (p.bias_mode == 2) v += __ldg(p.bias + m);
      if (p.beta) v += p.C[o];
      if (p.Z) p.Z[o] = p.z_grad ? dact_f(p.act, v) : v;
      v = act_f(p.act, v);
      if (p.dact) v *= (p.dact == 3) ? __ldg(p.aux + o) : dact_f(p.dact, __ldg(p.aux + o));
      p.C[o] = v;
      if (p.colsum) atomicAdd(p.colsum + ncol[j], v);
    }
  }
}

static inline GemmP gemm_default() {
  GemmP p;
  p.A = nullptr; p.B = nullptr; p.C = nullptr;
  p.a = MatDesc{lin(0), lin(0), lin(0)};
  p.b = p.a; p.c = p.a;
  p.M = p.N = p.K = 0; p.batch = 1;
  p.splitk = 1; p.kchunk = 0;
  p.alpha = 1.f; p.beta = 0;
  p.bias = nullptr; p.bias_mode = 0;
  p.act = ACT_NONE; p.Z = nullptr; p.aux = nullptr; p.dact = 0; p.a_act = p.b_act = 0; p.z_grad = 0; p.colsum = nullptr;
  p.a_kfast = p.b_kfast = p.a_kuni = p.b_kuni = 0; p.tiles_m = p.tiles_n = 0;
  return p;
}

// splitk: 1 = no split; 0 = choose automatically (caller guarantees C is zero-filled or is meant to be
// accumulated into, and uses no bias/beta/act epilogue).
static cudaError_t gemm_simt_launch(GemmP p, cudaStream_t s) {
  if (p.M <= 0 || p.N <= 0 || p.batch <= 0) return cudaSuccess;
  constexpr int BK = 16;
  const int BM = p.M > 64 ? 128 : 64, BN = p.N > 64 ? 128 : 64;
  p.tiles_m = (p.M + BM - 1) / BM;
  p.tiles_n = (p.N + BN - 1) / BN;
  if (p.splitk == 0) {
    const long long ctas = (long long)p.tiles_m * p.tiles_n * p.batch;
    int want = (int)((2 * 148 + ctas - 1) / ctas);            // fill ~2 waves
    const int by_acc = (p.K + 4095) / 4096;                     // keep fp32 chains short
    if (want < by_acc) want = by_acc;
    const int maxs = (p.K + 255) / 256;                         // at least 256 k per split
    if (want > maxs) want = maxs;
    if (want < 1) want = 1;
    p.splitk = want;
  }
  p.kchunk = ((p.K + p.splitk - 1) / p.splitk + BK - 1) / BK * BK;
  if (p.kchunk <= 0) p.kchunk = BK;
  p.splitk = p.K > 0 ? (p.K + p.kchunk - 1) / p.kchunk : 1;
  p.a_kfast = (p.a.c.sl == 1);
  p.b_kfast = (p.b.r.sl == 1) && !(p.b.c.sl == 1);
  p.a_kuni = (p.a.c.d == INT_MAX) || (p.a.c.d % BK == 0);
  p.b_kuni = (p.b.r.d == INT_MAX) || (p.b.r.d % BK == 0);
  const long long grid = (long long)p.tiles_m * p.tiles_n * p.batch * p.splitk;
  if (grid > 0x7fffffffLL) return cudaErrorInvalidValue;
  if (BM == 128 && BN == 128) gemm_simt_kernel<128, 128><<<(unsigned)grid, 256, 0, s>>>(p);
  else if (BM == 128) gemm_simt_kernel<128, 64><<<(unsigned)grid, 256, 0, s>>>(p);
  else if (BN == 128) gemm_simt_kernel<64, 128><<<(unsigned)grid, 256, 0, s>>>(p);
  else gemm_simt_kernel<64, 64><<<(unsigned)grid, 256, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace rpde
