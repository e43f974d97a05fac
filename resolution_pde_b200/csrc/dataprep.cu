// Data-preparation kernels of SURVEY section 8f rank 3 that are not contractions:
//   * global mean / standard deviation of a training set (the reference flattens every sample into a Python list of
//     floats and calls torch.tensor(...).mean() / .std(): dataloaders/ks_naive_markov.py:377-390,
//     dataloaders/ns_naive_true_multires.py:513-525),
//   * naive stride down-sampling (dataloaders/ns_naive_true_multires.py:363-366: data[:, :, ::f, ::f, :]),
//   * binned sums of a power spectrum (utils/frequency_error.py: the radial-band totals of the Parseval route).
// HBM-bound streaming kernels; sums are carried in fp64 so the statistics do not depend on the sample order.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rpde_b200.h"
#include "internal.h"

namespace {

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// acc[0] += n, acc[1] += sum x, acc[2] += sum x^2  (fp64)
__global__ void moments_kernel(const float* __restrict__ x, long long n, double* __restrict__ acc) {
  double s = 0.0, q = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool vec = (reinterpret_cast<uintptr_t>(x) & 15) == 0;
  if (vec) {
    const long long n4 = n >> 2;
    for (long long j = i; j < n4; j += stride) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(x) + j);
      float ps = (v.x + v.y) + (v.z + v.w);                    // four-element partials in fp32, the long sums in fp64
      float pq = fmaf(v.x, v.x, v.y * v.y) + fmaf(v.z, v.z, v.w * v.w);
      s += ps;
      q += pq;
    }
    for (long long j = (n4 << 2) + i; j < n; j += stride) {
      const float v = x[j];
      s += v;
      q += (double)v * v;
    }
  } else {
    for (long long j = i; j < n; j += stride) {
      const float v = x[j];
      s += v;
      q += (double)v * v;
    }
  }
  s = warp_sum_d(s);
  q = warp_sum_d(q);
  __shared__ double rs[8], rq[8];
  if ((threadIdx.x & 31) == 0) {
    rs[threadIdx.x >> 5] = s;
    rq[threadIdx.x >> 5] = q;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) {
      a += rs[k];
      b += rq[k];
    }
    atomicAdd(acc + 1, a);
    atomicAdd(acc + 2, b);
    if (blockIdx.x == 0) atomicAdd(acc, (double)n);
  }
}

// out = {mean, unbiased std} as torch.Tensor.mean() / .std() define them
__global__ void moments_finalize_kernel(const double* __restrict__ acc, float* __restrict__ out) {
  const double n = acc[0], mean = acc[1] / n;
  double var = (acc[2] - n * mean * mean) / (n - 1.0);
  if (var < 0.0) var = 0.0;
  out[0] = (float)mean;
  out[1] = (float)sqrt(var);
}

// y[o][i][j][c] = x[o][i * f][j * f][c]   (rows H -> ceil(H/f), columns W -> ceil(W/f), C trailing channels; 1-D: H = 1)
__global__ void stride_downsample_kernel(const float* __restrict__ x, float* __restrict__ y, long long outer, int H, int W, int C,
                                         int fh, int fw, int Ho, int Wo) {
  const long long total = outer * Ho * Wo * C;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(t % C);
    long long r = t / C;
    const int j = (int)(r % Wo);
    r /= Wo;
    const int i = (int)(r % Ho);
    const long long o = r / Ho;
    y[t] = __ldg(x + ((o * H + (long long)i * fh) * W + (long long)j * fw) * C + c);
  }
}

// out[r][b] = sum_j p[r][j] * (bin[j] == b)   (fp64; one block per (bin, row))
__global__ void binned_sums_kernel(const double* __restrict__ p, const int* __restrict__ bin, double* __restrict__ out, int nbins,
                                   long long n) {
  const int b = blockIdx.x, r = blockIdx.y;
  double s = 0.0;
  for (long long j = threadIdx.x; j < n; j += blockDim.x)
    if (bin[j] == b) s += p[(long long)r * n + j];
  s = warp_sum_d(s);
  __shared__ double red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) a += red[k];
    out[(long long)r * nbins + b] = a;
  }
}

}  // namespace

#define DP_CHECK(what)                                                                         \
  do {                                                                                         \
    rpde_internal_count_launch(0);                                                             \
    cudaError_t _e = cudaGetLastError();                                                       \
    if (_e != cudaSuccess) return rpde_internal_fail(RPDE_E_CUDA, cudaGetErrorString(_e));     \
  } while (0)

extern "C" {

int rpde_moments_accumulate(const float* x, long long n, double* acc3, rpde_stream_t stream) {
  if (!x || !acc3 || n <= 0) return rpde_internal_fail(RPDE_E_INVALID, "rpde_moments_accumulate: bad arguments");
  long long blocks = (n / 4 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  moments_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, n, acc3);
  DP_CHECK();
  return RPDE_OK;
}

int rpde_moments_finalize(const double* acc3, float* mean_std, rpde_stream_t stream) {
  if (!acc3 || !mean_std) return rpde_internal_fail(RPDE_E_INVALID, "rpde_moments_finalize: bad arguments");
  moments_finalize_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(acc3, mean_std);
  DP_CHECK();
  return RPDE_OK;
}

int rpde_stride_downsample(const float* x, float* y, long long outer, int H, int W, int C, int fh, int fw, rpde_stream_t stream) {
  if (!x || !y || outer <= 0 || H <= 0 || W <= 0 || C <= 0 || fh <= 0 || fw <= 0)
    return rpde_internal_fail(RPDE_E_INVALID, "rpde_stride_downsample: bad arguments");
  const int Ho = (H + fh - 1) / fh, Wo = (W + fw - 1) / fw;
  const long long total = outer * Ho * Wo * C;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  stride_downsample_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, y, outer, H, W, C, fh, fw, Ho, Wo);
  DP_CHECK();
  return RPDE_OK;
}

int rpde_binned_sums_f64(const double* p, const int* bin, double* out, int rows, int nbins, long long n, rpde_stream_t stream) {
  if (!p || !bin || !out || rows <= 0 || nbins <= 0 || n <= 0 || rows > 65535)
    return rpde_internal_fail(RPDE_E_INVALID, "rpde_binned_sums_f64: bad arguments");
  binned_sums_kernel<<<dim3((unsigned)nbins, (unsigned)rows), 256, 0, (cudaStream_t)stream>>>(p, bin, out, nbins, n);
  DP_CHECK();
  return RPDE_OK;
}

}  // extern "C"
