// C ABI of the B200-native spectral path (see include/rpde_b200.h for the contract).
// Host-side operator orchestration: every operator is a short sequence of descriptor-driven
// contractions (gemm_*.cuh) and pointwise kernels (pointwise.cuh) enqueued on the caller's stream.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <atomic>
#include <mutex>
#include <vector>

#include "../../include/rpde_b200.h"
#include "internal.h"
#include "common.cuh"
#include "gemm_simt.cuh"
#include "pointwise.cuh"
#include "fno_fused.cuh"
#include "fno2d_fused.cuh"
#include "fno2d_mma.cuh"
#ifdef RPDE_WITH_TC
#include "gemm_tc.cuh"
#endif

using namespace rpde;

static thread_local char g_err[512] = "";
static thread_local int g_prec = RPDE_PREC_FP32;

static int fail(int code, const char* fmt, const char* a = "") {
  snprintf(g_err, sizeof(g_err), fmt, a);
  return code;
}
#define RPDE_CUDA(expr)                                                        \
  do {                                                                         \
    cudaError_t _e = (expr);                                                   \
    if (_e != cudaSuccess) {                                                   \
      snprintf(g_err, sizeof(g_err), "CUDA error: %s (rpde.cu:%d)", cudaGetErrorString(_e), __LINE__); \
      return RPDE_E_CUDA;                                                      \
    }                                                                          \
  } while (0)
// launch statistics only (monotonic counter of kernels this library enqueued; read by bench.py)
static std::atomic<unsigned long long> g_launches{0};
static std::atomic<unsigned long long> g_tc_launches{0};  // contractions that ran on the tcgen05 engine
static std::atomic<unsigned long long> g_simt_launches{0};  // contractions that ran on the CUDA-core FMA engine
#define RPDE_LAUNCH_CHECK()        \
  do {                             \
    g_launches.fetch_add(1);       \
    RPDE_CUDA(cudaGetLastError()); \
  } while (0)

// Optional launch trace (rpde_trace_*): shape and engine of every contraction in launch order, so that a profiler's kernel
// list (names only) can be joined with the shapes.  Off by default; a mutex-protected host vector, touched only when enabled.
struct TraceRec {
  long long v[8];   // M, N, K, batch, split-K, engine (0 CUDA-core, 1 tcgen05), epilogue flags, 0
};
static std::mutex g_trace_mu;
static std::vector<TraceRec> g_trace;
static std::atomic<int> g_trace_on{0};
static void trace_gemm(const GemmP& p, int engine) {
  if (!g_trace_on.load(std::memory_order_relaxed)) return;
  const long long flags = (p.beta ? 1 : 0) | (p.act ? 2 : 0) | (p.dact ? 4 : 0) | (p.Z ? 8 : 0) | (p.colsum ? 16 : 0) | (p.bias_mode ? 32 : 0);
  std::lock_guard<std::mutex> lk(g_trace_mu);
  g_trace.push_back(TraceRec{{p.M, p.N, p.K, p.batch, p.splitk, engine, flags, 0}});
}

static cudaError_t gemm(const GemmP& p, cudaStream_t s) {
  if (p.M > 0 && p.N > 0 && p.batch > 0) g_launches.fetch_add(1);
#ifdef RPDE_WITH_TC
  if (g_prec != RPDE_PREC_FP32) {
    bool launched = false;
    const cudaError_t rc = gemm_tc_try_launch(p, g_prec, s, &launched);
    if (rc != cudaSuccess) return rc;
    if (launched) {
      g_tc_launches.fetch_add(1);
      trace_gemm(p, 1);
      return cudaSuccess;
    }
  }
#endif
  {
    static const bool log = getenv("RPDE_TC_DBG") && atoi(getenv("RPDE_TC_DBG")) == 9;
    if (log) fprintf(stderr, "[rpde simt] M=%d N=%d K=%d batch=%d splitk=%d\n", p.M, p.N, p.K, p.batch, p.splitk);
  }
  if (p.M > 0 && p.N > 0 && p.batch > 0) {
    g_simt_launches.fetch_add(1);
    trace_gemm(p, 0);
  }
  return gemm_simt_launch(p, s);
}

int rpde_internal_fail(int code, const char* msg) { return fail(code, "%s", msg); }
void rpde_internal_count_launch(int tensor_core) {
  g_launches.fetch_add(1);
  if (tensor_core) g_tc_launches.fetch_add(1);
}
int rpde_internal_precision(void) { return g_prec; }

static inline size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }
struct Carver {
  char* base;
  size_t off, cap;
  float* take(size_t floats) {
    float* p = reinterpret_cast<float*>(base + off);
    off += align_up(floats * sizeof(float));
    return p;
  }
};

extern "C" {

int rpde_version(void) { return 100; }
const char* rpde_last_error(void) { return g_err; }
int rpde_set_precision(int prec) {
  if (prec < 0 || prec > 3) return fail(RPDE_E_INVALID, "rpde_set_precision: unknown precision%s");
#ifndef RPDE_WITH_TC
  if (prec != RPDE_PREC_FP32) return fail(RPDE_E_INVALID, "rpde_set_precision: library built without the tensor-core engine%s");
#endif
  g_prec = prec;
  return RPDE_OK;
}
int rpde_get_precision(void) { return g_prec; }
unsigned long long rpde_launch_count(void) { return g_launches.load(); }
unsigned long long rpde_tc_launch_count(void) { return g_tc_launches.load(); }
unsigned long long rpde_simt_launch_count(void) { return g_simt_launches.load(); }
int rpde_trace_enable(int on) {
  std::lock_guard<std::mutex> lk(g_trace_mu);
  g_trace.clear();
  g_trace_on.store(on ? 1 : 0);
  return RPDE_OK;
}
int rpde_trace_count(void) {
  std::lock_guard<std::mutex> lk(g_trace_mu);
  return (int)g_trace.size();
}
int rpde_trace_get(int i, long long* out8) {
  std::lock_guard<std::mutex> lk(g_trace_mu);
  if (!out8 || i < 0 || i >= (int)g_trace.size()) return fail(RPDE_E_INVALID, "rpde_trace_get: bad index%s");
  for (int k = 0; k < 8; ++k) out8[k] = g_trace[i].v[k];
  return RPDE_OK;
}

// ------------------------------------------------------------------------------------------------
int rpde_dft_table(float* out, int kind, int n, int k, int m1, double scale, rpde_stream_t stream) {
  if (!out || n <= 0 || k <= 0 || kind < 0 || kind > 4) return fail(RPDE_E_INVALID, "rpde_dft_table: bad arguments%s");
  cudaStream_t s = (cudaStream_t)stream;
  const long long total = ((kind <= 1 || kind == 4) ? 2LL : 4LL) * k * n;
  dft_table_kernel<<<grid_for(total), 256, 0, s>>>(out, kind, n, k, m1, scale);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

static int pack_w(const float* w, float* wp, int I, int O, int modes, int keep, int sub, int sub_keep, cudaStream_t s) {
  const long long total = (long long)keep * sub_keep * 4 * I * O;
  if (sub == 1 && sub_keep == 1 && keep <= 128 && I <= 65535 && (O + 31) / 32 <= 65535) {
    dim3 g((unsigned)I, (unsigned)((O + 31) / 32));
    pack_weights_tiled_kernel<<<g, 256, 0, s>>>(w, wp, I, O, modes, keep);
  } else {
    pack_weights_kernel<<<grid_for(total), 256, 0, s>>>(w, wp, I, O, modes, keep, sub, sub_keep);
  }
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}
static int unpack_g(const float* P, float* dw, int I, int O, int modes, int keep, int sub, int sub_keep, int acc,
                    cudaStream_t s) {
  const long long total = (long long)I * O * modes;
  if (sub == 1 && sub_keep == 1 && keep <= 128 && I <= 65535 && (O + 31) / 32 <= 65535) {
    dim3 g((unsigned)I, (unsigned)((O + 31) / 32));
    unpack_wgrad_tiled_kernel<<<g, 256, 0, s>>>(P, dw, I, O, modes, keep, acc);
  } else {
    unpack_wgrad_kernel<<<grid_for(total), 256, 0, s>>>(P, dw, I, O, modes, keep, sub, sub_keep, acc);
  }
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

// ------------------------------------------------------------------------------------------------
// FNO SpectralConv1d
// ------------------------------------------------------------------------------------------------
size_t rpde_spectral1d_workspace_bytes(int B, int I, int O, int N, int modes) {
  (void)N;
  const size_t K = modes;
  size_t f = 0;
  f += align_up(K * 4 * I * O * 4);       // packed weights
  f += align_up(K * B * 2 * O * 4);       // mixed spectrum / its gradient
  f += align_up(K * B * 2 * I * 4);       // gradient of the input spectrum
  f += align_up(K * 4 * I * O * 4);       // packed weight gradient
  return f;
}

int rpde_spectral1d_fwd(const float* x, const float* w, const float* fa, const float* fs, float* spec, float* y,
                        float* z, int B, int I, int O, int N, int modes, int beta, int act, void* ws,
                        size_t ws_bytes, rpde_stream_t stream) {
  if (!x || !w || !fa || !fs || !spec || !y || B <= 0 || I <= 0 || O <= 0 || N <= 0 || modes <= 0)
    return fail(RPDE_E_INVALID, "rpde_spectral1d_fwd: bad arguments%s");
  if (modes > N / 2 + 1) return fail(RPDE_E_MODES, "SpectralConv1d: modes1 exceeds N//2+1%s");
  if (!ws || ws_bytes < rpde_spectral1d_workspace_bytes(B, I, O, N, modes))
    return fail(RPDE_E_WORKSPACE, "rpde_spectral1d_fwd: workspace too small%s");
  cudaStream_t s = (cudaStream_t)stream;
  const int K = modes;
  Carver cv{(char*)ws, 0, ws_bytes};
  float* wp = cv.take((size_t)K * 4 * I * O);
  float* om = cv.take((size_t)K * B * 2 * O);
  int rc = pack_w(w, wp, I, O, modes, K, 1, 1, s);
  if (rc) return rc;
  // Layouts: spec[k][ri][b][i], om[k][ro][b][o] -- the (mode, re/im) pair is the slow index, so the synthesis reads its
  // A operand with a single-level K stride (B*O) and unit-stride rows: both are TMA-addressable.
  const long long BI = (long long)B * I, BO = (long long)B * O;
  // 1. analysis (split-K: only B*I/128 row tiles exist): spec[(k,ri)][(b,i)] = sum_n x[(b,i)][n] fa[(k,ri)][n]
  RPDE_CUDA(cudaMemsetAsync(spec, 0, (size_t)K * 2 * BI * sizeof(float), s));
  GemmP g = gemm_default();
  g.A = x; g.B = fa; g.C = spec;
  g.M = B * I; g.N = 2 * K; g.K = N;
  g.a = MatDesc{lin(0), lin(N), lin(1)};
  g.b = MatDesc{lin(0), lin(1), lin(N)};
  g.c = MatDesc{lin(0), lin(1), lin(BI)};
  g.splitk = 0;
  RPDE_CUDA(gemm(g, s));
  // 2. mixing per mode: om[k][ro][b][o] = sum_(ri,i) spec[k][ri][b][i] wp[k][(ri,i)][(ro,o)]
  g = gemm_default();
  g.A = spec; g.B = wp; g.C = om;
  g.batch = K; g.M = B; g.N = 2 * O; g.K = 2 * I;
  g.a = MatDesc{lin(2 * BI), lin(I), two(I, BI, 1)};
  g.b = MatDesc{lin(4LL * I * O), lin(2 * O), lin(1)};
  g.c = MatDesc{lin(2 * BO), lin(O), two(O, BO, 1)};
  RPDE_CUDA(gemm(g, s));
  // 3. synthesis: y[(b,o)][n] = sum_(k,ro) om[(k,ro)][(b,o)] fs[n][(k,ro)]
  g = gemm_default();
  g.A = om; g.B = fs; g.C = y;
  g.M = B * O; g.N = N; g.K = 2 * K;
  g.a = MatDesc{lin(0), lin(1), lin(BO)};
  g.b = MatDesc{lin(0), lin(1), lin(2 * K)};
  g.c = MatDesc{lin(0), lin(N), lin(1)};
  g.beta = beta; g.act = act; g.Z = z;
  RPDE_CUDA(gemm(g, s));
  return RPDE_OK;
}

// synthesis + bypass launch shared by the fused forward and (adjoint = 1) the fused input gradient
static int launch_fno1d_synth(const float* x, const float* om, const float* table, const float* wb, const float* bb, float* y,
                              float* z, int B, int I, int O, int N, int modes, int act, int adjoint, cudaStream_t s, int Hd = 1) {
  static const int tn_sel = getenv("RPDE_FNO_TN") ? atoi(getenv("RPDE_FNO_TN")) : 0;   // bring-up switch: 64 or 128
  static const bool use_mma = !(getenv("RPDE_FNO2D_MMA") && atoi(getenv("RPDE_FNO2D_MMA")) == 0);
  // FNOBlock2d image rows of 64 / 128 / 256 points, at most 32 channels: the tensor-path kernel (fno2d_mma.cuh)
  const int RPm = (2 * modes + 7) & ~7;
  if (use_mma && Hd > 1 && (N == 64 || N == 128 || N == 256 || N == 512 || N == 1024) && I <= F2M_MAXC && O <= F2M_MAXC &&
      2 * modes <= F2M_MAXR && f2m_synth_smem(N, RPm) <= 227 * 1024) {
    const int RP = RPm;
    const size_t smem = f2m_synth_smem(N, RP);
    const int rpi = N >= 256 ? 1 : 256 / N, ch = N > 256 ? N / 256 : 1;
    long long tiles = (((long long)B + rpi - 1) / rpi) * ch;
    const long long resident = N <= 256 ? 2 * 148 : 148;
    const unsigned grid = (unsigned)(tiles < resident ? tiles : resident);
#define RPDE_SYNTH_MMA(WW)                                                                                                   \
  do {                                                                                                                       \
    RPDE_CUDA(cudaFuncSetAttribute(fno2d_synth_mma_kernel<WW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));     \
    fno2d_synth_mma_kernel<WW><<<grid, 256, smem, s>>>(x, om, table, wb, bb, y, z, B, I, O, modes, act, adjoint, Hd);        \
  } while (0)
    if (N == 64) RPDE_SYNTH_MMA(64);
    else if (N == 128) RPDE_SYNTH_MMA(128);
    else if (N == 256) RPDE_SYNTH_MMA(256);
    else if (N == 512) RPDE_SYNTH_MMA(512);
    else RPDE_SYNTH_MMA(1024);
#undef RPDE_SYNTH_MMA
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  if (use_mma && Hd == 1 && N % 128 == 0 && I <= 64 && O <= 64 && 2 * modes <= F2M_MAXR && !tn_sel) {
    // FNOBlock1d: one CTA per (b, 128 points) on the warp-level tensor path (fno2d_mma.cuh)
    const size_t smem = f1m_synth_smem(RPm);
    RPDE_CUDA(cudaFuncSetAttribute(fno1d_synth_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int chunks = N / 128;
    int by = 148 / chunks > 0 ? 148 / chunks : 1;  // one resident CTA per SM (146 registers); more rows are walked by the CTA
    if (by > B) by = B;
    dim3 g1((unsigned)chunks, (unsigned)by);
    fno1d_synth_mma_kernel<<<g1, 256, smem, s>>>(x, om, table, wb, bb, y, z, B, I, O, N, modes, act, adjoint);
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  const bool wide = tn_sel ? tn_sel == 128 : (long long)((N + 127) / 128) * B >= 2 * 148;   // enough 128-point tiles to fill the SMs twice
  if (O <= 32 && (wide || Hd > 1)) {     // narrow layers (FNO2d width 32): a 32-channel register tile, no padded FMAs
    dim3 g2((unsigned)((N + 127) / 128), (unsigned)B);
    fno1d_synth_bypass_kernel<128, 32><<<g2, 256, 0, s>>>(x, om, table, wb, bb, y, z, B, I, O, N, modes, act, adjoint, Hd);
  } else if (wide) {
    dim3 g2((unsigned)((N + 127) / 128), (unsigned)B);
    fno1d_synth_bypass_kernel<128><<<g2, 256, 0, s>>>(x, om, table, wb, bb, y, z, B, I, O, N, modes, act, adjoint, Hd);
  } else {
    dim3 g2((unsigned)((N + 63) / 64), (unsigned)B);
    fno1d_synth_bypass_kernel<64><<<g2, 256, 0, s>>>(x, om, table, wb, bb, y, z, B, I, O, N, modes, act, adjoint, Hd);
  }
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

static int launch_waxis_mma(const float* x, const float* tab, float* T, long long rows, int N, int R, const float* z, int act,
                            float* xout, int kplanes, long long s_row, long long s_r, cudaStream_t s);

// analysis launch shared by the fused forward (x -> spec) and the fused backward prologue (dy, z -> dpre, dom)
static int launch_fno1d_analysis(const float* x, const float* table, float* out, int B, int rows_per_b, int N, int modes,
                                 const float* z, int act, float* xout, int rowmap, cudaStream_t s) {
  {
    // the tensor-path kernel (fno2d_mma.cuh); with few rows it splits the points of a row over several CTAs
    static const bool use_mma = !(getenv("RPDE_FNO2D_MMA") && atoi(getenv("RPDE_FNO2D_MMA")) == 0);
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    const long long rows = (long long)B * rows_per_b;
    const int R = 2 * modes;
    if (use_mma && rows >= 128 && N % 32 == 0 && R <= F2M_MAXR && f2m_waxis_smem(N, (R + 7) & ~7, false) <= 200 * 1024 &&
        al16(x) && al16(z) && al16(xout))
      return launch_waxis_mma(x, table, out, rows, N, R, z, act, xout, rowmap ? modes : 0, 1, rows, s);
  }
  dim3 g1((unsigned)(((long long)B * rows_per_b + 7) / 8), (unsigned)((2 * modes + FNO_RC - 1) / FNO_RC));
  const size_t smem1 = (size_t)FNO_RC * N * sizeof(float);
  if (smem1 > 48 * 1024)      // N in (1536, 2048]: 64 KB table chunk (the attribute is per device: set it on every call)
    RPDE_CUDA(cudaFuncSetAttribute(fno1d_analysis_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  if (N <= 256)
    fno1d_analysis_kernel<8><<<g1, 256, smem1, s>>>(x, table, out, B, rows_per_b, N, modes, z, act, xout, rowmap);
  else if (N <= 1024)
    fno1d_analysis_kernel<32><<<g1, 256, smem1, s>>>(x, table, out, B, rows_per_b, N, modes, z, act, xout, rowmap);
  else
    fno1d_analysis_kernel<64><<<g1, 256, smem1, s>>>(x, table, out, B, rows_per_b, N, modes, z, act, xout, rowmap);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_fno_block1d_fwd(const float* x, const float* w, const float* fa, const float* fs4, const float* wb,
                         const float* bb, float* spec, float* y, float* z, int B, int I, int O, int N, int modes, int act,
                         void* ws, size_t ws_bytes, rpde_stream_t stream) {
  if (!x || !w || !fa || !fs4 || !y || B <= 0 || I <= 0 || O <= 0 || N <= 0 || modes <= 0 || B > 65535 || (bb && !wb))
    return fail(RPDE_E_INVALID, "rpde_fno_block1d_fwd: bad arguments%s");
  if (modes > N / 2 + 1) return fail(RPDE_E_MODES, "SpectralConv1d: modes1 exceeds N//2+1%s");
  if (N > 2048) return fail(RPDE_E_INVALID, "rpde_fno_block1d_fwd: grid too long for the fused kernels (use rpde_spectral1d_fwd)%s");
  const size_t smem2 = (size_t)2 * modes * (I + 1) * sizeof(float);
  if (smem2 > 200 * 1024) return fail(RPDE_E_INVALID, "rpde_fno_block1d_fwd: modes x channels too large for the fused kernels%s");
  const size_t need = align_up((size_t)modes * B * 2 * O * 4) + (spec ? 0 : align_up((size_t)modes * B * 2 * I * 4));
  if (!ws || ws_bytes < need) return fail(RPDE_E_WORKSPACE, "rpde_fno_block1d_fwd: workspace too small%s");
  cudaStream_t s = (cudaStream_t)stream;
  float* om = (float*)ws;
  if (!spec) spec = (float*)((char*)ws + align_up((size_t)modes * B * 2 * O * 4));   // inference: spectrum stays in the workspace
  int rc1 = launch_fno1d_analysis(x, fa, spec, B, I, N, modes, nullptr, 0, nullptr, 0, s);
  if (rc1) return rc1;
  if (I <= 128 && B <= 64) {     // small batch: one CTA per (mode, 8 output channels), operands staged once (fno_fused.cuh)
    dim3 gk((unsigned)modes, (unsigned)((O + M1_IC - 1) / M1_IC), (unsigned)((B + M1_BC - 1) / M1_BC));
    const size_t smk = (size_t)(2 * M1_BC * (I + 1) + M1_IC * (2 * I + 2)) * sizeof(float);
    fno1d_mix_modes_kernel<<<gk, 128, smk, s>>>(spec, w, om, B, I, O, modes, 0);
  } else {
    if (smem2 > 48 * 1024)
      RPDE_CUDA(cudaFuncSetAttribute(fno1d_mix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    dim3 gm((unsigned)((O + FNO_MO - 1) / FNO_MO), (unsigned)B);
    fno1d_mix_kernel<<<gm, 256, smem2, s>>>(spec, w, om, B, I, O, modes);
  }
  RPDE_LAUNCH_CHECK();
  return launch_fno1d_synth(x, om, fs4, wb, bb, y, z, B, I, O, N, modes, act, 0, s);
}

// z / act / fs4 / dpre non-NULL select the fused prologue (rpde_fno_block1d_bwd): dpre = dy * act'(z) and its projection on the
// synthesis table in one launch instead of rpde_act_bwd + memset + split-K contraction
static int spectral1d_bwd_impl(const float* dy, const float* z, int act, const float* fs4, float* dpre, const float* wb,
                               int fuse_dx, const float* w,
                               const float* fa, const float* fs, const float* spec, float* dx, int dx_accumulate, float* dw,
                               int B, int I, int O, int N, int modes, void* ws, size_t ws_bytes, rpde_stream_t stream) {
  if (!dy || !w || !fa || !(fs || fs4) || !spec || B <= 0 || I <= 0 || O <= 0 || N <= 0 || modes <= 0)
    return fail(RPDE_E_INVALID, "rpde_spectral1d_bwd: bad arguments%s");
  if (modes > N / 2 + 1) return fail(RPDE_E_MODES, "SpectralConv1d: modes1 exceeds N//2+1%s");
  if (!ws || ws_bytes < rpde_spectral1d_workspace_bytes(B, I, O, N, modes))
    return fail(RPDE_E_WORKSPACE, "rpde_spectral1d_bwd: workspace too small%s");
  cudaStream_t s = (cudaStream_t)stream;
  const int K = modes;
  Carver cv{(char*)ws, 0, ws_bytes};
  float* wp = cv.take((size_t)K * 4 * I * O);
  float* dom = cv.take((size_t)K * B * 2 * O);
  float* dsp = cv.take((size_t)K * B * 2 * I);
  float* P = cv.take((size_t)K * 4 * I * O);
  // mode-space FMA kernels (fno_fused.cuh) for the widths the FNO configurations use: two launches instead of
  // pack + two batched contractions + memset + unpack
  static const bool use_modespace = !(getenv("RPDE_FNO1D_MODESPACE") && atoi(getenv("RPDE_FNO1D_MODESPACE")) == 0);
  // (small batches only: from B ~ 100 on the batched contractions of the tensor engine are faster -- measured 38 + 20 us against
  // 16 + 10 us at B = 256)
  const bool tiled = use_modespace && I <= 128 && O <= 128 && B <= 64;
  int rc = RPDE_OK;
  if (!tiled) {
    rc = pack_w(w, wp, I, O, modes, K, 1, 1, s);
    if (rc) return rc;
  }
  const long long BI = (long long)B * I, BO = (long long)B * O;
  // 3'. dom[(k,ro)][(b,o)] = sum_n dy[(b,o)][n] fs[n][(k,ro)]
  GemmP g = gemm_default();
  if (fs4) {
    rc = launch_fno1d_analysis(dy, fs4, dom, B, O, N, K, z, act, dpre, 1, s);
    if (rc) return rc;
    if (dpre) dy = dpre;
  } else {
    RPDE_CUDA(cudaMemsetAsync(dom, 0, (size_t)K * 2 * BO * sizeof(float), s));
    g.A = dy; g.B = fs; g.C = dom;
    g.M = B * O; g.N = 2 * K; g.K = N;
    g.a = MatDesc{lin(0), lin(N), lin(1)};
    g.b = MatDesc{lin(0), lin(2 * K), lin(1)};
    g.c = MatDesc{lin(0), lin(1), lin(BO)};
    g.splitk = 0;
    RPDE_CUDA(gemm(g, s));
  }
  if (dw && tiled) {
    dim3 gw((unsigned)K, (unsigned)((I + M1_IC - 1) / M1_IC), (unsigned)((O + 31) / 32));
    fno1d_wgrad_kernel<<<gw, 256, 0, s>>>(spec, dom, dw, B, I, O, K);
    RPDE_LAUNCH_CHECK();
  } else if (dw) {
    // dWp[k] = spec[k]^T dom[k]
    RPDE_CUDA(cudaMemsetAsync(P, 0, (size_t)K * 4 * I * O * sizeof(float), s));
    g = gemm_default();
    g.A = spec; g.B = dom; g.C = P;
    g.batch = K; g.M = 2 * I; g.N = 2 * O; g.K = B;
    g.a = MatDesc{lin(2 * BI), two(I, BI, 1), lin(I)};
    g.b = MatDesc{lin(2 * BO), lin(O), two(O, BO, 1)};
    g.c = MatDesc{lin(4LL * I * O), lin(2 * O), lin(1)};
    g.splitk = 0;
    RPDE_CUDA(gemm(g, s));
    rc = unpack_g(P, dw, I, O, modes, K, 1, 1, 0, s);
    if (rc) return rc;
  }
  if (dx) {
    if (tiled) {
      dim3 gm((unsigned)K, (unsigned)((I + M1_IC - 1) / M1_IC), (unsigned)((B + M1_BC - 1) / M1_BC));
      const size_t smem = (size_t)(2 * M1_BC * (O + 1) + M1_IC * (2 * O + 2)) * sizeof(float);
      fno1d_mix_modes_kernel<<<gm, 128, smem, s>>>(dom, w, dsp, B, I, O, K, 1);
      RPDE_LAUNCH_CHECK();
    } else {
      // 2'. dsp[k] = dom[k] wp[k]^T
      g = gemm_default();
      g.A = dom; g.B = wp; g.C = dsp;
      g.batch = K; g.M = B; g.N = 2 * I; g.K = 2 * O;
      g.a = MatDesc{lin(2 * BO), lin(O), two(O, BO, 1)};
      g.b = MatDesc{lin(4LL * I * O), lin(1), lin(2 * O)};
      g.c = MatDesc{lin(2 * BI), lin(I), two(I, BI, 1)};
      RPDE_CUDA(gemm(g, s));
    }
    // 1'. dx[(b,i)][n] (+)= sum_(k,ri) dsp[(k,ri)][(b,i)] fa[(k,ri)][n]
    if (fuse_dx && !dx_accumulate)   // (+ the bypass conv's input gradient wb^T dpre in the same kernel; dy == dpre here)
      return launch_fno1d_synth(dy, dsp, fa, wb, nullptr, dx, nullptr, B, /*reduction channels*/ O, /*outputs*/ I, N, K, ACT_NONE, 1, s);
    g = gemm_default();
    g.A = dsp; g.B = fa; g.C = dx;
    g.M = B * I; g.N = N; g.K = 2 * K;
    g.a = MatDesc{lin(0), lin(1), lin(BI)};
    g.b = MatDesc{lin(0), lin(N), lin(1)};
    g.c = MatDesc{lin(0), lin(N), lin(1)};
    g.beta = dx_accumulate;
    RPDE_CUDA(gemm(g, s));
  }
  return RPDE_OK;
}

int rpde_spectral1d_bwd(const float* dy, const float* w, const float* fa, const float* fs, const float* spec,
                        float* dx, int dx_accumulate, float* dw, int B, int I, int O, int N, int modes, void* ws,
                        size_t ws_bytes, rpde_stream_t stream) {
  if (!fs) return fail(RPDE_E_INVALID, "rpde_spectral1d_bwd: bad arguments%s");
  return spectral1d_bwd_impl(dy, nullptr, 0, nullptr, nullptr, nullptr, 0, w, fa, fs, spec, dx, dx_accumulate, dw, B, I, O, N,
                             modes, ws, ws_bytes, stream);
}

// 1: rpde_fno_block1d_bwd produces the complete input gradient (spectral + bypass conv) when it is given wb.
// The choice is the caller's and is made per call by passing wb or NULL; no environment switch is read here.
int rpde_fno_block1d_bwd_fuses_dx(void) { return 1; }

int rpde_fno_block1d_bwd(const float* dy, const float* z, int act, const float* w, const float* fa, const float* fs4,
                         const float* spec, const float* wb, float* dpre, float* dx, int dx_accumulate, float* dw, int B, int I,
                         int O, int N, int modes, void* ws, size_t ws_bytes, rpde_stream_t stream) {
  if (!fs4 || N > 2048 || (z && !dpre) || (!z && dpre))
    return fail(RPDE_E_INVALID, "rpde_fno_block1d_bwd: bad arguments (N <= 2048; z and dpre come together)%s");
  if (wb && (dx_accumulate || !dx))
    return fail(RPDE_E_INVALID, "rpde_fno_block1d_bwd: wb (fused bypass input gradient) needs dx and dx_accumulate == 0%s");
  if (wb && B > 65535) return fail(RPDE_E_INVALID, "rpde_fno_block1d_bwd: fused input gradient is limited to B <= 65535%s");
  return spectral1d_bwd_impl(dy, z, act, fs4, dpre, wb, wb ? 1 : 0, w, fa, nullptr, spec, dx, dx_accumulate, dw, B, I, O, N,
                             modes, ws, ws_bytes, stream);
}

// ------------------------------------------------------------------------------------------------
// Fused FNOBlock2d (fno2d_fused.cuh)
// ------------------------------------------------------------------------------------------------
static bool fno2d_fused_ok(int B, int I, int O, int H, int W, int m1, int m2) {
  // image rows as the synthesis kernel's grid.y, W-axis rows in registers (W <= 1024), 2 m2 <= 32 table rows, and the H-axis
  // kernel's thread = (a, ky group) mapping: 4 m1 <= 256 values of a, (4 m1) * ceil(m2 / 4) <= 256 threads
  return (long long)B * H <= 65535 && W <= 1024 && 2 * m2 <= 32 && 4 * m1 <= 256 && 4LL * m1 * ((m2 + 3) / 4) <= 256 &&
         (long long)B * (I > O ? I : O) * H <= 0x7fffffffLL / 4;
}
int rpde_fno_block2d_supported(int B, int I, int O, int H, int W, int m1, int m2) { return fno2d_fused_ok(B, I, O, H, W, m1, m2) ? 1 : 0; }

size_t rpde_fno_block2d_workspace_bytes(int B, int I, int O, int H, int W, int m1, int m2) {
  (void)W;
  const size_t C = (size_t)(I > O ? I : O), J = 2 * (size_t)m1, Ky = m2;
  size_t f = 0;
  f += align_up(2 * Ky * (size_t)B * C * H * 4);   // W-axis half transform T / its adjoint
  f += align_up((size_t)B * C * 2 * J * Ky * 4);   // mixed spectrum om / dom
  f += align_up((size_t)B * C * 2 * J * Ky * 4);   // dspec, or the input spectrum when the caller keeps none
  f += align_up(2 * Ky * (size_t)B * H * C * 4);   // Z: the H-axis synthesis per image row
  return f;
}

// W-axis / 1-D analysis on the warp-level tensor path (fno2d_mma.cuh): 16 rows per warp, the table resident in shared memory
// as TF32 hi / lo planes when both fit in 100 KB (two CTAs per SM), as one fp32 plane split on the fly otherwise
static int launch_waxis_mma(const float* x, const float* tab, float* T, long long rows, int N, int R, const float* z, int act,
                            float* xout, int kplanes, long long s_row, long long s_r, cudaStream_t s) {
  const int RP = (R + 7) & ~7;
  long long ctas = ((rows + 15) / 16 + 7) / 8;
  // few row tiles: split the points of a row over gridDim.y CTAs (partial sums meet in T through atomics) until ~128 CTAs exist
  int split = 1;
  // ... or until both TF32 planes of the table slice fit (two CTAs per SM, no splitting of table fragments in the loop)
  while ((ctas * split < 128 || f2m_waxis_smem(N / split, RP, true) > 100 * 1024) && N % (64 * split) == 0 && N / (2 * split) >= 64)
    split *= 2;
  const int Wc = N / split;
  const bool presplit = f2m_waxis_smem(Wc, RP, true) <= 100 * 1024;
  const size_t sm = f2m_waxis_smem(Wc, RP, presplit);
  const long long resident = presplit ? 2 * 148 : 148;
  if (ctas > resident) ctas = resident;
  if (split > 1) RPDE_CUDA(cudaMemsetAsync(T, 0, (size_t)rows * R * sizeof(float), s));   // T is dense: rows x R either way
  dim3 grid((unsigned)ctas, (unsigned)split);
  if (presplit) {
    RPDE_CUDA(cudaFuncSetAttribute(fno2d_waxis_mma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    fno2d_waxis_mma_kernel<true><<<grid, 256, sm, s>>>(x, tab, T, rows, N, R, z, act, xout, kplanes, s_row, s_r);
  } else {
    RPDE_CUDA(cudaFuncSetAttribute(fno2d_waxis_mma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    fno2d_waxis_mma_kernel<false><<<grid, 256, sm, s>>>(x, tab, T, rows, N, R, z, act, xout, kplanes, s_row, s_r);
  }
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

static int launch_fno2d_waxis(const float* x, const float* tab, float* T, long long rows, int N, int R, const float* z, int act,
                              float* xout, int kplanes, cudaStream_t s) {
  static const bool use_mma = !(getenv("RPDE_FNO2D_MMA") && atoi(getenv("RPDE_FNO2D_MMA")) == 0);
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if (use_mma && N % 32 == 0 && R <= F2M_MAXR && f2m_waxis_smem(N, (R + 7) & ~7, false) <= 200 * 1024 && al16(x) && al16(z) && al16(xout))
    return launch_waxis_mma(x, tab, T, rows, N, R, z, act, xout, kplanes, R, 1, s);
  const size_t smem = (size_t)((R + 7) & ~7) * N * sizeof(float);
  long long blocks = (rows + 15) / 16;
  if (blocks > 148 * 4) blocks = 148 * 4;
  const unsigned g = (unsigned)blocks;
#define RPDE_WAXIS(NPL)                                                                                                      \
  do {                                                                                                                       \
    if (smem > 48 * 1024)                                                                                                    \
      RPDE_CUDA(cudaFuncSetAttribute(fno2d_waxis_kernel<NPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));      \
    fno2d_waxis_kernel<NPL><<<g, 256, smem, s>>>(x, tab, T, rows, N, R, z, act, xout, kplanes);                            \
  } while (0)
  if (N <= 64) RPDE_WAXIS(2);
  else if (N <= 128) RPDE_WAXIS(4);
  else if (N <= 256) RPDE_WAXIS(8);
  else if (N <= 512) RPDE_WAXIS(16);
  else RPDE_WAXIS(32);
#undef RPDE_WAXIS
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}
static int launch_fno2d_rows(const float* T, const float* tab, float* out, long long planes, int H, int Ky, int A, cudaStream_t s) {
  static const bool use_mma = !(getenv("RPDE_FNO2D_MMA") && atoi(getenv("RPDE_FNO2D_MMA")) == 0);
  if (use_mma && A <= 48 && Ky <= 12) {   // two planes per CTA on the warp-level tensor path (fno2d_mma.cuh)
    const unsigned grid = (unsigned)((planes + 1) / 2);
#define RPDE_ROWS_MMA(MT)                                                                                                   \
  do {                                                                                                                       \
    const size_t sm = (size_t)8 * MT * 16 * 33 * sizeof(float);                                                             \
    if (sm > 48 * 1024)                                                                                                      \
      RPDE_CUDA(cudaFuncSetAttribute(fno2d_rows_mma_kernel<MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));      \
    fno2d_rows_mma_kernel<MT><<<grid, 256, sm, s>>>(T, tab, out, planes, H, Ky, A);                                         \
  } while (0)
    if (A <= 16) RPDE_ROWS_MMA(1);
    else if (A <= 32) RPDE_ROWS_MMA(2);
    else RPDE_ROWS_MMA(3);
#undef RPDE_ROWS_MMA
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  const size_t smem = (size_t)F2_RC * (2 * Ky + 2 * A) * sizeof(float);    // <= 64 * (32 + 512) * 4 = 136 KB, 30 KB at 12 x 12 modes
  if (smem > 48 * 1024) RPDE_CUDA(cudaFuncSetAttribute(fno2d_rows_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  fno2d_rows_kernel<4><<<(unsigned)planes, 256, smem, s>>>(T, tab, out, H, Ky, A);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}
static int launch_fno2d_expand(const float* in, const float* tab, float* Z, int B, int C, int H, int Ky, int A, cudaStream_t s) {
  static const bool use_mma = !(getenv("RPDE_FNO2D_MMA") && atoi(getenv("RPDE_FNO2D_MMA")) == 0);
  if (use_mma && A <= 64 && Ky % 4 == 0 && C % 8 == 0 && (reinterpret_cast<uintptr_t>(in) & 15) == 0 &&
      (reinterpret_cast<uintptr_t>(Z) & 7) == 0) {
    // 32 image rows per CTA (four warps x 16 table rows) on the warp-level tensor path (fno2d_mma.cuh)
    const unsigned mchunks = (unsigned)((2 * H + 63) / 64);
    unsigned zsplit = 1;                                  // enough CTAs for two waves of four-warp CTAs where the shape allows
    while (zsplit < (unsigned)(C / 8) && mchunks * (unsigned)B * zsplit < 4 * 148) zsplit *= 2;
    dim3 gm(mchunks, (unsigned)B, zsplit);
    if (A <= 32) fno2d_expand_mma_kernel<4><<<gm, 128, 0, s>>>(in, tab, Z, B, C, H, Ky, A);
    else if (A <= 48) fno2d_expand_mma_kernel<6><<<gm, 128, 0, s>>>(in, tab, Z, B, C, H, Ky, A);
    else fno2d_expand_mma_kernel<8><<<gm, 128, 0, s>>>(in, tab, Z, B, C, H, Ky, A);
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  dim3 g((unsigned)((H + F2_EH - 1) / F2_EH), (unsigned)B);
  // A <= 256 (fno2d_fused_ok): at most 16 * 12 * 33 + 32 * 256 floats = 57 KB
#define RPDE_EXPAND(KYT)                                                                                                     \
  do {                                                                                                                       \
    const size_t smem = ((size_t)F2_AC * KYT * 33 + (size_t)2 * F2_EH * A) * sizeof(float);                                  \
    if (smem > 48 * 1024)                                                                                                    \
      RPDE_CUDA(cudaFuncSetAttribute(fno2d_expand_kernel<KYT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));     \
    fno2d_expand_kernel<KYT><<<g, 256, smem, s>>>(in, tab, Z, B, C, H, Ky, A);                                             \
  } while (0)
  if (Ky <= 4) RPDE_EXPAND(4);
  else if (Ky <= 8) RPDE_EXPAND(8);
  else RPDE_EXPAND(12);
#undef RPDE_EXPAND
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

// per-mode channel mixing (mode 0) and its adjoint (mode 1): the shared-memory-tiled kernel when the ky row fits its 16 slots
static int launch_fno2d_mix(const float* in, const float* w1, const float* w2, float* out, int B, int I, int O, int m1, int Ky,
                            int mode, cudaStream_t s) {
  const int J = 2 * m1, Co = mode ? I : O;
  if (Ky <= F2X_KY && (long long)B <= 65535LL * F2X_MB) {
    dim3 g((unsigned)J, (unsigned)((Co + F2X_OC - 1) / F2X_OC), (unsigned)((B + F2X_MB - 1) / F2X_MB));
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    const size_t smem = (size_t)(F2X_MCL * 2 * F2X_MB * F2X_KY + F2X_MCL * F2X_OC * F2X_KY * 2) * sizeof(float);   // 64 KB
    if (Ky % 4 == 0 && al16(in) && al16(w1) && al16(w2)) {
      RPDE_CUDA(cudaFuncSetAttribute(fno2d_mix_tiled_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      fno2d_mix_tiled_kernel<true><<<g, 256, smem, s>>>(in, w1, w2, out, B, I, O, m1, Ky, mode);
    } else {
      RPDE_CUDA(cudaFuncSetAttribute(fno2d_mix_tiled_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      fno2d_mix_tiled_kernel<false><<<g, 256, smem, s>>>(in, w1, w2, out, B, I, O, m1, Ky, mode);
    }
  } else {
    const long long total = (long long)B * Co * J * Ky;
    fno2d_mix_kernel<<<grid_for(total), 256, 0, s>>>(in, w1, w2, out, B, I, O, m1, Ky, mode);
  }
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_fno_block2d_fwd(const float* x, const float* w1, const float* w2, const float* fa_y, const float* fs4_y, const float* g2t,
                         const float* g3, const float* wb, const float* bb, float* spec, float* y, float* z, int B, int I, int O,
                         int H, int W, int m1, int m2, int act, void* ws, size_t ws_bytes, rpde_stream_t stream) {
  if (!x || !w1 || !w2 || !fa_y || !fs4_y || !g2t || !g3 || !y || B <= 0 || I <= 0 || O <= 0 || H <= 0 || W <= 0 || m1 <= 0 ||
      m2 <= 0 || (bb && !wb))
    return fail(RPDE_E_INVALID, "rpde_fno_block2d_fwd: bad arguments%s");
  if (m2 > W / 2 + 1 || m1 > H) return fail(RPDE_E_MODES, "SpectralConv2d: modes exceed the available spectrum%s");
  if (!fno2d_fused_ok(B, I, O, H, W, m1, m2))
    return fail(RPDE_E_INVALID, "rpde_fno_block2d_fwd: shape outside the fused kernels (use rpde_conv1x1_fwd + rpde_spectral2d_fwd)%s");
  if (!ws || ws_bytes < rpde_fno_block2d_workspace_bytes(B, I, O, H, W, m1, m2))
    return fail(RPDE_E_WORKSPACE, "rpde_fno_block2d_fwd: workspace too small%s");
  cudaStream_t s = (cudaStream_t)stream;
  const int J = 2 * m1, Ky = m2;
  const size_t C = (size_t)(I > O ? I : O);
  Carver cv{(char*)ws, 0, ws_bytes};
  float* T = cv.take(2 * (size_t)Ky * B * C * H);
  float* om = cv.take((size_t)B * C * 2 * J * Ky);
  float* sp_ws = cv.take((size_t)B * C * 2 * J * Ky);
  float* Z = cv.take(2 * (size_t)Ky * B * H * C);
  if (!spec) spec = sp_ws;   // inference: the spectrum stays in the workspace
  int rc = launch_fno2d_waxis(x, fa_y, T, (long long)B * I * H, W, 2 * Ky, nullptr, 0, nullptr, 0, s);
  if (rc) return rc;
  rc = launch_fno2d_rows(T, g2t, spec, (long long)B * I, H, Ky, 2 * J, s);
  if (rc) return rc;
  rc = launch_fno2d_mix(spec, w1, w2, om, B, I, O, m1, Ky, 0, s);
  if (rc) return rc;
  rc = launch_fno2d_expand(om, g3, Z, B, O, H, Ky, 2 * J, s);
  if (rc) return rc;
  return launch_fno1d_synth(x, Z, fs4_y, wb, bb, y, z, B * H, I, O, W, Ky, act, 0, s, H);
}

// Both H-axis kernels read their table as [(h, r)][a]: forward 2 uses g2t (the kind-2 table transposed to [(h,ri)][(j,ro)]) and
// 4 uses g3 (kind 3 as generated); the adjoints swap them.  dpre (= dy * act'(z), written when z != NULL) is what the bypass conv's parameter gradient consumes; dx
// receives the complete input gradient (spectral adjoint + wb^T dpre when wb is given).
int rpde_fno_block2d_bwd(const float* dy, const float* z, int act, const float* w1, const float* w2, const float* fa_y,
                         const float* fs4_y, const float* g3, const float* g2t, const float* spec, const float* wb, float* dpre,
                         float* dx, float* dw1, float* dw2, int B, int I, int O, int H, int W, int m1, int m2, void* ws,
                         size_t ws_bytes, rpde_stream_t stream) {
  if (!dy || !w1 || !w2 || !fa_y || !fs4_y || !g3 || !g2t || !spec || B <= 0 || I <= 0 || O <= 0 || H <= 0 || W <= 0 || m1 <= 0 ||
      m2 <= 0 || (z && !dpre) || (!z && dpre))
    return fail(RPDE_E_INVALID, "rpde_fno_block2d_bwd: bad arguments (z and dpre come together)%s");
  if (!fno2d_fused_ok(B, I, O, H, W, m1, m2)) return fail(RPDE_E_INVALID, "rpde_fno_block2d_bwd: shape outside the fused kernels%s");
  if (!ws || ws_bytes < rpde_fno_block2d_workspace_bytes(B, I, O, H, W, m1, m2))
    return fail(RPDE_E_WORKSPACE, "rpde_fno_block2d_bwd: workspace too small%s");
  cudaStream_t s = (cudaStream_t)stream;
  const int J = 2 * m1, Ky = m2;
  const size_t C = (size_t)(I > O ? I : O);
  Carver cv{(char*)ws, 0, ws_bytes};
  float* T = cv.take(2 * (size_t)Ky * B * C * H);
  float* dom = cv.take((size_t)B * C * 2 * J * Ky);
  float* dsp = cv.take((size_t)B * C * 2 * J * Ky);
  float* U = cv.take(2 * (size_t)Ky * B * H * C);
  // 5'. dpre = dy * act'(z); T[(ky,ro)][(b,o,h)] = sum_w dpre fs[(ky,ro)][w]
  int rc = launch_fno2d_waxis(dy, fs4_y, T, (long long)B * O * H, W, 2 * Ky, z, act, dpre, Ky, s);
  if (rc) return rc;
  const float* g = dpre ? dpre : dy;
  // 4'. dom[(b,o)][(j,ri)][ky] = sum_(h,ro) g3[(h,ro)][(j,ri)] T[(b,o,h)][(ky,ro)]
  rc = launch_fno2d_rows(T, g3, dom, (long long)B * O, H, Ky, 2 * J, s);
  if (rc) return rc;
  if (dw1 && dw2) {
    if (Ky <= F2X_KY) {
      dim3 gw((unsigned)J, (unsigned)((I + F2X_OC - 1) / F2X_OC), (unsigned)((O + F2X_OC - 1) / F2X_OC));
      fno2d_wgrad_tiled_kernel<<<gw, 256, 0, s>>>(spec, dom, dw1, dw2, B, I, O, m1, Ky);
    } else {
      const long long total = (long long)I * O * J * Ky;
      fno2d_wgrad_kernel<<<grid_for(total), 256, 0, s>>>(spec, dom, dw1, dw2, B, I, O, m1, Ky);
    }
    RPDE_LAUNCH_CHECK();
  }
  if (dx) {
    rc = launch_fno2d_mix(dom, w1, w2, dsp, B, I, O, m1, Ky, 1, s);
    if (rc) return rc;
    // 2'. U[(ky,ri)][(b,h)][i] = sum_(j,ro) g2[(j,ro)][(ri,h)] dspec[(b,i)][(j,ro)][ky]
    rc = launch_fno2d_expand(dsp, g2t, U, B, I, H, Ky, 2 * J, s);
    if (rc) return rc;
    // 1'. dx[b,i,h,w] = sum_(ky,ri) U fa[(ky,ri)][w] + sum_o wb[o,i] dpre[b,o,h,w]
    return launch_fno1d_synth(g, U, fa_y, wb, nullptr, dx, nullptr, B * H, /*reduction channels*/ O, /*outputs*/ I, W, Ky, ACT_NONE, 1, s, H);
  }
  return RPDE_OK;
}

// ------------------------------------------------------------------------------------------------
// FNO SpectralConv2d: separable truncated transform  (W: R2C keep m2) x (H: C2C keep rows [0,m1) U [H-m1,H))
// ------------------------------------------------------------------------------------------------
size_t rpde_spectral2d_workspace_bytes(int B, int I, int O, int H, int W, int m1, int m2) {
  (void)W;
  const size_t J = 2 * (size_t)m1, Ky = m2, C = (size_t)(I > O ? I : O);
  size_t f = 0;
  f += align_up(J * Ky * 4 * I * O * 4);        // packed weights
  f += align_up((size_t)B * C * 2 * H * Ky * 4); // half-transformed planes S1 / Z
  f += align_up(J * Ky * B * 2 * C * 4);        // mixed spectrum or its gradient
  f += align_up(J * Ky * B * 2 * C * 4);        // gradient of the input spectrum
  f += align_up(J * Ky * 4 * I * O * 4);        // packed weight gradient
  return f;
}

int rpde_spectral2d_fwd(const float* x, const float* w1, const float* w2, const float* fa_y, const float* fs_y,
                        const float* gf, const float* gi, float* spec, float* y, float* z, int B, int I, int O,
                        int H, int W, int m1, int m2, int beta, int act, void* ws, size_t ws_bytes,
                        rpde_stream_t stream) {
  if (!x || !w1 || !w2 || !fa_y || !fs_y || !gf || !gi || !spec || !y || B <= 0 || I <= 0 || O <= 0 || H <= 0 ||
      W <= 0 || m1 <= 0 || m2 <= 0)
    return fail(RPDE_E_INVALID, "rpde_spectral2d_fwd: bad arguments%s");
  if (m2 > W / 2 + 1 || m1 > H) return fail(RPDE_E_MODES, "SpectralConv2d: modes exceed the available spectrum%s");
  if (!ws || ws_bytes < rpde_spectral2d_workspace_bytes(B, I, O, H, W, m1, m2))
    return fail(RPDE_E_WORKSPACE, "rpde_spectral2d_fwd: workspace too small%s");
  cudaStream_t s = (cudaStream_t)stream;
  const int J = 2 * m1, Ky = m2;
  const size_t Cm = (size_t)(I > O ? I : O);
  Carver cv{(char*)ws, 0, ws_bytes};
  float* wp = cv.take((size_t)J * Ky * 4 * I * O);
  float* pl = cv.take((size_t)B * Cm * 2 * H * Ky);
  float* om = cv.take((size_t)J * Ky * B * 2 * Cm);
  int rc = pack_w(w1, wp, I, O, m1 * m2, m1, m2, m2, s);
  if (rc) return rc;
  rc = pack_w(w2, wp + (size_t)m1 * Ky * 4 * I * O, I, O, m1 * m2, m1, m2, m2, s);
  if (rc) return rc;
  const long long HK = (long long)H * Ky;
  // 1. y-analysis: pl[(b,i)][ri][h][ky] = sum_w x[(b,i,h)][w] fa_y[(ky,ri)][w]
  GemmP g = gemm_default();
  g.A = x; g.B = fa_y; g.C = pl;
  g.M = B * I * H; g.N = 2 * Ky; g.K = W;
  g.a = MatDesc{lin(0), lin(W), lin(1)};
  g.b = MatDesc{lin(0), lin(1), lin(W)};
  g.c = MatDesc{lin(0), two(H, 2 * HK, Ky), two(2, 1, HK)};
  RPDE_CUDA(gemm(g, s));
  // 2. x-analysis: spec[j][ky][b][ro][i] = sum_(ri,h) gf[(j,ro)][(ri,h)] pl[(b,i)][ri][h][ky]
  g = gemm_default();
  g.A = gf; g.B = pl; g.C = spec;
  g.batch = B * I; g.M = 2 * J; g.N = Ky; g.K = 2 * H;
  g.a = MatDesc{lin(0), lin(2 * H), lin(1)};
  g.b = MatDesc{lin(2 * HK), lin(Ky), lin(1)};
  g.c = MatDesc{two(I, 2LL * I, 1), two(2, (long long)Ky * B * 2 * I, I), lin((long long)B * 2 * I)};
  RPDE_CUDA(gemm(g, s));
  // 3. mixing per retained mode, written as om[b][o][j][ri][ky]: the x-synthesis then reads a [(j,ri)][ky] matrix per
  //    (b,o) with single-level strides (TMA-addressable); the scattered store here touches only J*Ky*B*2O floats
  const long long JK2 = 2LL * J * Ky;
  g = gemm_default();
  g.A = spec; g.B = wp; g.C = om;
  g.batch = J * Ky; g.M = B; g.N = 2 * O; g.K = 2 * I;
  g.a = MatDesc{lin((long long)B * 2 * I), lin(2 * I), lin(1)};
  g.b = MatDesc{lin(4LL * I * O), lin(2 * O), lin(1)};
  g.c = MatDesc{two(Ky, 2LL * Ky, 1), lin((long long)O * JK2), two(O, Ky, JK2)};
  RPDE_CUDA(gemm(g, s));
  // 4. x-synthesis: pl[(b,o)][ro][h][ky] = sum_(j,ri) gi[(h,ro)][(j,ri)] om[(b,o)][(j,ri)][ky]
  g = gemm_default();
  g.A = gi; g.B = om; g.C = pl;
  g.batch = B * O; g.M = 2 * H; g.N = Ky; g.K = 2 * J;
  g.a = MatDesc{lin(0), lin(2 * J), lin(1)};
  g.b = MatDesc{lin(JK2), lin(Ky), lin(1)};
  g.c = MatDesc{lin(2 * HK), two(2, Ky, HK), lin(1)};
  RPDE_CUDA(gemm(g, s));
  // 5. y-synthesis: y[(b,o,h)][w] = sum_(ri,ky) pl[(b,o)][ri][h][ky] fs_y[ri][ky][w], one contraction per plane ri so
  //    that both operands keep a single-level, unit-stride reduction index (TMA-addressable): the re plane writes
  //    (on top of the bypass result when beta), the im plane accumulates and applies the activation.
  for (int ri = 0; ri < 2; ++ri) {
    g = gemm_default();
    g.A = pl + (size_t)ri * HK; g.B = fs_y + (size_t)ri * Ky * W; g.C = y;
    g.M = B * O * H; g.N = W; g.K = Ky;
    g.a = MatDesc{lin(0), two(H, 2 * HK, Ky), lin(1)};
    g.b = MatDesc{lin(0), lin(W), lin(1)};
    g.c = MatDesc{lin(0), lin(W), lin(1)};
    if (ri == 0) {
      g.beta = beta;
    } else {
      g.beta = 1; g.act = act; g.Z = z;
    }
    RPDE_CUDA(gemm(g, s));
  }
  return RPDE_OK;
}

int rpde_spectral2d_bwd(const float* dy, const float* w1, const float* w2, const float* fa_y, const float* fs_y,
                        const float* gf, const float* gi, const float* spec, float* dx, int dx_accumulate,
                        float* dw1, float* dw2, int B, int I, int O, int H, int W, int m1, int m2, void* ws,
                        size_t ws_bytes, rpde_stream_t stream) {
  if (!dy || !w1 || !w2 || !fa_y || !fs_y || !gf || !gi || !spec || B <= 0 || I <= 0 || O <= 0 || H <= 0 ||
      W <= 0 || m1 <= 0 || m2 <= 0)
    return fail(RPDE_E_INVALID, "rpde_spectral2d_bwd: bad arguments%s");
  if (m2 > W / 2 + 1 || m1 > H) return fail(RPDE_E_MODES, "SpectralConv2d: modes exceed the available spectrum%s");
  if (!ws || ws_bytes < rpde_spectral2d_workspace_bytes(B, I, O, H, W, m1, m2))
    return fail(RPDE_E_WORKSPACE, "rpde_spectral2d_bwd: workspace too small%s");
  cudaStream_t s = (cudaStream_t)stream;
  const int J = 2 * m1, Ky = m2;
  const size_t Cm = (size_t)(I > O ? I : O);
  Carver cv{(char*)ws, 0, ws_bytes};
  float* wp = cv.take((size_t)J * Ky * 4 * I * O);
  float* pl = cv.take((size_t)B * Cm * 2 * H * Ky);
  float* dom = cv.take((size_t)J * Ky * B * 2 * Cm);
  float* dsp = cv.take((size_t)J * Ky * B * 2 * Cm);
  float* P = cv.take((size_t)J * Ky * 4 * I * O);
  int rc = pack_w(w1, wp, I, O, m1 * m2, m1, m2, m2, s);
  if (rc) return rc;
  rc = pack_w(w2, wp + (size_t)m1 * Ky * 4 * I * O, I, O, m1 * m2, m1, m2, m2, s);
  if (rc) return rc;
  const long long HK = (long long)H * Ky;
  // 5'. pl[(b,o)][ri][h][ky] = sum_w dy[(b,o,h)][w] fs_y[ri][ky][w]      (columns n = ri*Ky + ky)
  GemmP g = gemm_default();
  g.A = dy; g.B = fs_y; g.C = pl;
  g.M = B * O * H; g.N = 2 * Ky; g.K = W;
  g.a = MatDesc{lin(0), lin(W), lin(1)};
  g.b = MatDesc{lin(0), lin(1), lin(W)};
  g.c = MatDesc{lin(0), two(H, 2 * HK, Ky), two(Ky, HK, 1)};
  RPDE_CUDA(gemm(g, s));
  // 4'. dom[j][ky][b][ri][o] = sum_(ro,h) gi[(h,ro)][(j,ri)] pl[(b,o)][ro][h][ky]
  g = gemm_default();
  g.A = gi; g.B = pl; g.C = dom;
  g.batch = B * O; g.M = 2 * J; g.N = Ky; g.K = 2 * H;
  g.a = MatDesc{lin(0), lin(1), two(H, 2LL * J, 4LL * J)};
  g.b = MatDesc{lin(2 * HK), lin(Ky), lin(1)};
  g.c = MatDesc{two(O, 2LL * O, 1), two(2, (long long)Ky * B * 2 * O, O), lin((long long)B * 2 * O)};
  RPDE_CUDA(gemm(g, s));
  if (dw1 || dw2) {
    RPDE_CUDA(cudaMemsetAsync(P, 0, (size_t)J * Ky * 4 * I * O * sizeof(float), s));
    g = gemm_default();
    g.A = spec; g.B = dom; g.C = P;
    g.batch = J * Ky; g.M = 2 * I; g.N = 2 * O; g.K = B;
    g.a = MatDesc{lin((long long)B * 2 * I), lin(1), lin(2 * I)};
    g.b = MatDesc{lin((long long)B * 2 * O), lin(2 * O), lin(1)};
    g.c = MatDesc{lin(4LL * I * O), lin(2 * O), lin(1)};
    g.splitk = 0;
    RPDE_CUDA(gemm(g, s));
    if (dw1) { rc = unpack_g(P, dw1, I, O, m1 * m2, m1, m2, m2, 0, s); if (rc) return rc; }
    if (dw2) { rc = unpack_g(P + (size_t)m1 * Ky * 4 * I * O, dw2, I, O, m1 * m2, m1, m2, m2, 0, s); if (rc) return rc; }
  }
  if (dx) {
    // 3'. dsp[mode] = dom[mode] wp[mode]^T, written as dsp[b][i][j][ro][ky] (see step 3 of the forward pass)
    const long long JK2 = 2LL * J * Ky;
    g = gemm_default();
    g.A = dom; g.B = wp; g.C = dsp;
    g.batch = J * Ky; g.M = B; g.N = 2 * I; g.K = 2 * O;
    g.a = MatDesc{lin((long long)B * 2 * O), lin(2 * O), lin(1)};
    g.b = MatDesc{lin(4LL * I * O), lin(1), lin(2 * O)};
    g.c = MatDesc{two(Ky, 2LL * Ky, 1), lin((long long)I * JK2), two(I, Ky, JK2)};
    RPDE_CUDA(gemm(g, s));
    // 2'. pl[(b,i)][ri][h][ky] = sum_(j,ro) gf[(j,ro)][(ri,h)] dsp[(b,i)][(j,ro)][ky]
    g = gemm_default();
    g.A = gf; g.B = dsp; g.C = pl;
    g.batch = B * I; g.M = 2 * H; g.N = Ky; g.K = 2 * J;
    g.a = MatDesc{lin(0), lin(1), lin(2 * H)};
    g.b = MatDesc{lin(JK2), lin(Ky), lin(1)};
    g.c = MatDesc{lin(2 * HK), lin(Ky), lin(1)};
    RPDE_CUDA(gemm(g, s));
    // 1'. dx[(b,i,h)][w] (+)= sum_(ri,ky) pl[(b,i)][ri][h][ky] fa_y[(ky,ri)][w], one contraction per plane ri
    for (int ri = 0; ri < 2; ++ri) {
      g = gemm_default();
      g.A = pl + (size_t)ri * HK; g.B = fa_y + (size_t)ri * W; g.C = dx;
      g.M = B * I * H; g.N = W; g.K = Ky;
      g.a = MatDesc{lin(0), two(H, 2 * HK, Ky), lin(1)};
      g.b = MatDesc{lin(0), lin(2LL * W), lin(1)};
      g.c = MatDesc{lin(0), lin(W), lin(1)};
      g.beta = ri == 0 ? dx_accumulate : 1;
      RPDE_CUDA(gemm(g, s));
    }
  }
  return RPDE_OK;
}

// ------------------------------------------------------------------------------------------------
// F-FNO factorised pass along one axis of a channels-last tensor  [outer][n][rest][C]
// ------------------------------------------------------------------------------------------------
size_t rpde_ffno_axis_workspace_bytes(long long outer, int n, long long rest, int C, int k) {
  (void)n;
  const size_t sp = (size_t)outer * 2 * k * rest * C;
  size_t f = 0;
  f += align_up((size_t)k * 4 * C * C * 4);  // packed weights
  f += align_up(sp * 4);                     // mixed spectrum / its gradient
  f += align_up(sp * 4);                     // gradient of the input spectrum
  f += align_up((size_t)k * 4 * C * C * 4);  // packed weight gradient
  return f;
}

int rpde_ffno_axis_fwd(const float* x, const float* w, const float* fa, const float* fs, float* spec, float* y,
                       int accumulate, long long outer, int n, long long rest, int C, int k, int w_modes, int mix,
                       void* ws, size_t ws_bytes, rpde_stream_t stream) {
  if (!x || !fa || !fs || !spec || !y || outer <= 0 || n <= 0 || rest <= 0 || C <= 0 || k <= 0 || k > n / 2 + 1 ||
      (mix == RPDE_MIX_FULL && (!w || k > w_modes)) || outer > INT_MAX || outer * rest > INT_MAX ||
      rest * C > INT_MAX)
    return fail(RPDE_E_INVALID, "rpde_ffno_axis_fwd: bad arguments%s");
  if (!ws || ws_bytes < rpde_ffno_axis_workspace_bytes(outer, n, rest, C, k))
    return fail(RPDE_E_WORKSPACE, "rpde_ffno_axis_fwd: workspace too small%s");
  cudaStream_t s = (cudaStream_t)stream;
  const long long inner = rest * C;
  const size_t sp = (size_t)outer * 2 * k * inner;
  Carver cv{(char*)ws, 0, ws_bytes};
  float* wp = cv.take((size_t)k * 4 * C * C);
  float* om = cv.take(sp);
  // 1. analysis: spec[o][(k,ri)][inner] = fa[(k,ri)][n] x[o][n][inner]
  GemmP g = gemm_default();
  g.A = fa; g.B = x; g.C = spec;
  g.batch = (int)outer; g.M = 2 * k; g.N = (int)inner; g.K = n;
  g.a = MatDesc{lin(0), lin(n), lin(1)};
  g.b = MatDesc{lin((long long)n * inner), lin(inner), lin(1)};
  g.c = MatDesc{lin(2LL * k * inner), lin(inner), lin(1)};
  RPDE_CUDA(gemm(g, s));
  const float* mixed = spec;
  if (mix == RPDE_MIX_FULL) {
    int rc = pack_w(w, wp, C, C, w_modes, k, 1, 1, s);
    if (rc) return rc;
    // 2. per mode: om[(o,r)][(ro,c')] = spec[(o,r)][(ri,c)] wp[k]
    g = gemm_default();
    g.A = spec; g.B = wp; g.C = om;
    g.batch = k; g.M = (int)(outer * rest); g.N = 2 * C; g.K = 2 * C;
    g.a = MatDesc{lin(2 * inner), two((int)rest, 2LL * k * inner, C), two(C, inner, 1)};
    g.b = MatDesc{lin(4LL * C * C), lin(2 * C), lin(1)};
    g.c = g.a;
    RPDE_CUDA(gemm(g, s));
    mixed = om;
  }
  // 3. synthesis: y[o][n][inner] (+)= fs[n][(k,ri)] mixed[o][(k,ri)][inner]
  g = gemm_default();
  g.A = fs; g.B = mixed; g.C = y;
  g.batch = (int)outer; g.M = n; g.N = (int)inner; g.K = 2 * k;
  g.a = MatDesc{lin(0), lin(2 * k), lin(1)};
  g.b = MatDesc{lin(2LL * k * inner), lin(inner), lin(1)};
  g.c = MatDesc{lin((long long)n * inner), lin(inner), lin(1)};
  g.beta = accumulate;
  RPDE_CUDA(gemm(g, s));
  return RPDE_OK;
}

int rpde_ffno_axis_bwd(const float* dy, const float* w, const float* fa, const float* fs, const float* spec,
                       float* dx, int dx_accumulate, float* dw, int dw_accumulate, long long outer, int n,
                       long long rest, int C, int k, int w_modes, int mix, void* ws, size_t ws_bytes,
                       rpde_stream_t stream) {
  if (!dy || !fa || !fs || outer <= 0 || n <= 0 || rest <= 0 || C <= 0 || k <= 0 || k > n / 2 + 1 ||
      (mix == RPDE_MIX_FULL && (!w || (dw && !spec) || k > w_modes)) || outer > INT_MAX || outer * rest > INT_MAX ||
      rest * C > INT_MAX)
    return fail(RPDE_E_INVALID, "rpde_ffno_axis_bwd: bad arguments%s");
  if (!ws || ws_bytes < rpde_ffno_axis_workspace_bytes(outer, n, rest, C, k))
    return fail(RPDE_E_WORKSPACE, "rpde_ffno_axis_bwd: workspace too small%s");
  cudaStream_t s = (cudaStream_t)stream;
  const long long inner = rest * C;
  const size_t sp = (size_t)outer * 2 * k * inner;
  Carver cv{(char*)ws, 0, ws_bytes};
  float* wp = cv.take((size_t)k * 4 * C * C);
  float* dom = cv.take(sp);
  float* dsp = cv.take(sp);
  float* P = cv.take((size_t)k * 4 * C * C);
  // 3'. dom[o] = fs^T dy[o]
  GemmP g = gemm_default();
  g.A = fs; g.B = dy; g.C = dom;
  g.batch = (int)outer; g.M = 2 * k; g.N = (int)inner; g.K = n;
  g.a = MatDesc{lin(0), lin(1), lin(2 * k)};
  g.b = MatDesc{lin((long long)n * inner), lin(inner), lin(1)};
  g.c = MatDesc{lin(2LL * k * inner), lin(inner), lin(1)};
  RPDE_CUDA(gemm(g, s));
  const float* dspec = dom;
  if (mix == RPDE_MIX_FULL) {
    int rc;
    if (dw) {
      RPDE_CUDA(cudaMemsetAsync(P, 0, (size_t)k * 4 * C * C * sizeof(float), s));
      // dWp[k] = spec_k^T dom_k  (contraction over (o,r))
      g = gemm_default();
      g.A = spec; g.B = dom; g.C = P;
      g.batch = k; g.M = 2 * C; g.N = 2 * C; g.K = (int)(outer * rest);
      g.a = MatDesc{lin(2 * inner), two(C, inner, 1), two((int)rest, 2LL * k * inner, C)};
      g.b = MatDesc{lin(2 * inner), two((int)rest, 2LL * k * inner, C), two(C, inner, 1)};
      g.c = MatDesc{lin(4LL * C * C), lin(2 * C), lin(1)};
      g.splitk = 0;
      RPDE_CUDA(gemm(g, s));
      rc = unpack_g(P, dw, C, C, w_modes, k, 1, 1, dw_accumulate, s);
      if (rc) return rc;
    }
    if (dx) {
      rc = pack_w(w, wp, C, C, w_modes, k, 1, 1, s);
      if (rc) return rc;
      // 2'. dsp_k = dom_k wp_k^T
      g = gemm_default();
      g.A = dom; g.B = wp; g.C = dsp;
      g.batch = k; g.M = (int)(outer * rest); g.N = 2 * C; g.K = 2 * C;
      g.a = MatDesc{lin(2 * inner), two((int)rest, 2LL * k * inner, C), two(C, inner, 1)};
      g.b = MatDesc{lin(4LL * C * C), lin(1), lin(2 * C)};
      g.c = g.a;
      RPDE_CUDA(gemm(g, s));
      dspec = dsp;
    }
  }
  if (dx) {
    // 1'. dx[o] (+)= fa^T dspec[o]
    g = gemm_default();
    g.A = fa; g.B = dspec; g.C = dx;
    g.batch = (int)outer; g.M = n; g.N = (int)inner; g.K = 2 * k;
    g.a = MatDesc{lin(0), lin(1), lin(n)};
    g.b = MatDesc{lin(2LL * k * inner), lin(inner), lin(1)};
    g.c = MatDesc{lin((long long)n * inner), lin(inner), lin(1)};
    g.beta = dx_accumulate;
    RPDE_CUDA(gemm(g, s));
  }
  return RPDE_OK;
}

// ------------------------------------------------------------------------------------------------
// channels-last Linear
// ------------------------------------------------------------------------------------------------
static MatDesc rows_cf(long long T, int ch, long long S) {
  // element (t, c) of a channels-first tensor (T/S, ch, S): t = b*S + s -> b*ch*S + c*S + s
  (void)T;
  return MatDesc{lin(0), two((int)S, (long long)ch * S, 1), lin(S)};
}

int rpde_linear_fwd(const float* x, const float* w, const float* bias, float* y, float* z, long long T, int in,
                    int out, int act, long long y_cf_spatial, rpde_stream_t stream) {
  if (!x || !w || !y || T <= 0 || in <= 0 || out <= 0 || T > INT_MAX || y_cf_spatial > INT_MAX ||
      (y_cf_spatial > 0 && T % y_cf_spatial != 0))
    return fail(RPDE_E_INVALID, "rpde_linear_fwd: bad arguments%s");
  const int save_grad = (act & RPDE_ACT_SAVE_GRAD) ? 1 : 0;
  act &= 0xff;
  if (in <= 4 && (out & 3) == 0 && out <= 1024 && act == RPDE_ACT_NONE && !z && y_cf_spatial == 0 &&
      (reinterpret_cast<uintptr_t>(y) & 15) == 0) {
    // a handful of input channels (in_proj): streaming kernel, one pass over the output
    cudaStream_t s = (cudaStream_t)stream;
    const size_t smem = (size_t)(out * in + out) * sizeof(float);
    const unsigned grid = grid_for(T * (out / 4));
    switch (in) {
      case 1: linear_smallk_fwd_kernel<1><<<grid, 256, smem, s>>>(x, w, bias, y, T, out); break;
      case 2: linear_smallk_fwd_kernel<2><<<grid, 256, smem, s>>>(x, w, bias, y, T, out); break;
      case 3: linear_smallk_fwd_kernel<3><<<grid, 256, smem, s>>>(x, w, bias, y, T, out); break;
      default: linear_smallk_fwd_kernel<4><<<grid, 256, smem, s>>>(x, w, bias, y, T, out); break;
    }
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  GemmP g = gemm_default();
  g.A = x; g.B = w; g.C = y;
  g.M = (int)T; g.N = out; g.K = in;
  g.a = MatDesc{lin(0), lin(in), lin(1)};
  g.b = MatDesc{lin(0), lin(1), lin(in)};
  g.c = y_cf_spatial > 0 ? rows_cf(T, out, y_cf_spatial) : MatDesc{lin(0), lin(out), lin(1)};
  g.bias = bias; g.bias_mode = bias ? 1 : 0;
  g.act = act; g.Z = z; g.z_grad = save_grad;
  RPDE_CUDA(gemm(g, (cudaStream_t)stream));
  return RPDE_OK;
}

static int reduce_channels(const float* X, float* out, long long outer, int C, long long inner, cudaStream_t s);

int rpde_linear_bwd_input(const float* dy, const float* w, const float* zprev, int dact, float* dx, int accumulate,
                          float* dbias_prev, long long T, int in, int out, long long dy_cf_spatial,
                          rpde_stream_t stream) {
  if (!dy || !w || !dx || T <= 0 || in <= 0 || out <= 0 || T > INT_MAX || (dact && !zprev) || dact < 0 || dact > 3 ||
      (dy_cf_spatial > 0 && T % dy_cf_spatial != 0) || (dbias_prev && accumulate))
    return fail(RPDE_E_INVALID, "rpde_linear_bwd_input: bad arguments%s");
  // The CUDA-core engine's fused column sums are one atomic per element: T fp32 additions per column in arrival order
  // (2.9e-5 relative at T = 262144, tests/test_gpu_baseline_configs.py).  The fp32 engine is the numerical yardstick, so
  // there the bias gradient comes from the hierarchical channel reduction below; the tensor-core engine sums per warp
  // and tile in registers first and keeps the fused form.
  const bool fused_colsum = dbias_prev && (g_prec != RPDE_PREC_FP32 || T <= 16384);
  if (fused_colsum) RPDE_CUDA(cudaMemsetAsync(dbias_prev, 0, (size_t)in * sizeof(float), (cudaStream_t)stream));
  GemmP g = gemm_default();
  g.A = dy; g.B = w; g.C = dx;
  g.M = (int)T; g.N = in; g.K = out;
  g.a = dy_cf_spatial > 0 ? rows_cf(T, out, dy_cf_spatial) : MatDesc{lin(0), lin(out), lin(1)};
  g.b = MatDesc{lin(0), lin(in), lin(1)};
  g.c = MatDesc{lin(0), lin(in), lin(1)};
  g.beta = accumulate;
  g.aux = zprev; g.dact = dact; g.colsum = fused_colsum ? dbias_prev : nullptr;
  if (accumulate && dact) return fail(RPDE_E_INVALID, "rpde_linear_bwd_input: accumulate with dact is ambiguous%s");
  RPDE_CUDA(gemm(g, (cudaStream_t)stream));
  if (dbias_prev && !fused_colsum) return reduce_channels(dx, dbias_prev, T, in, 1, (cudaStream_t)stream);
  return RPDE_OK;
}

static int reduce_channels(const float* X, float* out, long long outer, int C, long long inner, cudaStream_t s) {
  if (inner != 1 && outer * inner <= 65536) {
    // small problem (the 1-D configurations): one block per channel writes its sum directly, no memset, no atomics
    reduce_channels_kernel<<<dim3(1, (unsigned)C), 256, 0, s>>>(X, out, outer, C, inner, 8192);
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  RPDE_CUDA(cudaMemsetAsync(out, 0, (size_t)C * sizeof(float), s));
  if (inner == 1) {
    long long chunk = (outer + 148 * 4 - 1) / (148 * 4);
    if (chunk < 64) chunk = 64;
    dim3 grid((unsigned)((outer + chunk - 1) / chunk), (unsigned)((C + 31) / 32));
    reduce_channels_kernel<<<grid, 256, 0, s>>>(X, out, outer, C, 1, chunk);
  } else {
    long long chunk = 8192;  // floats of one contiguous inner run per block pass
    const long long cpi = (inner + chunk - 1) / chunk, pairs = outer * cpi;
    long long gx = (148LL * 16 + C - 1) / C;  // about 16 blocks per SM over all channels
    if (gx > pairs) gx = pairs;
    if (gx < 1) gx = 1;
    dim3 grid((unsigned)gx, (unsigned)C);
    reduce_channels_kernel<<<grid, 256, 0, s>>>(X, out, outer, C, inner, chunk);
  }
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_linear_bwd_params(const float* dy, const float* x, float* dw, float* db, long long T, int in, int out,
                           long long dy_cf_spatial, rpde_stream_t stream) {
  if (!dy || !x || !dw || T <= 0 || in <= 0 || out <= 0 || T > INT_MAX ||
      (dy_cf_spatial > 0 && T % dy_cf_spatial != 0))
    return fail(RPDE_E_INVALID, "rpde_linear_bwd_params: bad arguments%s");
  cudaStream_t s = (cudaStream_t)stream;
  RPDE_CUDA(cudaMemsetAsync(dw, 0, (size_t)out * in * sizeof(float), s));
  if (in <= 4 && dy_cf_spatial == 0 && out <= 256 && 256 % out == 0) {
    // a handful of input channels: one streaming pass over dy gives dw and db together
    if (db) RPDE_CUDA(cudaMemsetAsync(db, 0, (size_t)out * sizeof(float), s));
    long long rpb = (T + 148 * 4 - 1) / (148 * 4);
    if (rpb < 256) rpb = 256;
    const unsigned grid = (unsigned)((T + rpb - 1) / rpb);
    const size_t smem = (size_t)(in + 1) * 256 * sizeof(float);
    switch (in) {
      case 1: linear_smallk_wgrad_kernel<1><<<grid, 256, smem, s>>>(dy, x, dw, db, T, out, rpb); break;
      case 2: linear_smallk_wgrad_kernel<2><<<grid, 256, smem, s>>>(dy, x, dw, db, T, out, rpb); break;
      case 3: linear_smallk_wgrad_kernel<3><<<grid, 256, smem, s>>>(dy, x, dw, db, T, out, rpb); break;
      default: linear_smallk_wgrad_kernel<4><<<grid, 256, smem, s>>>(dy, x, dw, db, T, out, rpb); break;
    }
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  GemmP g = gemm_default();
  g.A = dy; g.B = x; g.C = dw;
  g.M = out; g.N = in; g.K = (int)T;
  if (dy_cf_spatial > 0) {
    const long long S = dy_cf_spatial;
    g.a = MatDesc{lin(0), lin(S), two((int)S, (long long)out * S, 1)};
  } else {
    g.a = MatDesc{lin(0), lin(1), lin(out)};
  }
  g.b = MatDesc{lin(0), lin(in), lin(1)};
  g.c = MatDesc{lin(0), lin(in), lin(1)};
  g.splitk = 0;
  RPDE_CUDA(gemm(g, s));
  if (db) {
    if (dy_cf_spatial > 0) return reduce_channels(dy, db, T / dy_cf_spatial, out, dy_cf_spatial, s);
    return reduce_channels(dy, db, T, out, 1, s);
  }
  return RPDE_OK;
}

// ------------------------------------------------------------------------------------------------
// channels-first 1x1 convolution
// ------------------------------------------------------------------------------------------------
int rpde_conv1x1_fwd(const float* x, const float* w, const float* bias, float* y, float* z, int B, int I, int O,
                     long long S, int act, int beta, rpde_stream_t stream) {
  if (!x || !w || !y || B <= 0 || I <= 0 || O <= 0 || S <= 0 || S > INT_MAX)
    return fail(RPDE_E_INVALID, "rpde_conv1x1_fwd: bad arguments%s");
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if (I <= 4 && !act && !z && !beta && (S % 4) == 0 && O <= 1024 && al16(x) && al16(y)) {
    // few input channels (lifting layer): one streaming kernel instead of a K = I contraction
    const long long total = (long long)B * (S / 4);
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    const size_t sm = (size_t)(O * I + O) * sizeof(float);
    cudaStream_t s = (cudaStream_t)stream;
    switch (I) {
      case 1: conv1x1_smallk_fwd_kernel<1><<<(unsigned)blocks, 256, sm, s>>>(x, w, bias, y, B, O, S); break;
      case 2: conv1x1_smallk_fwd_kernel<2><<<(unsigned)blocks, 256, sm, s>>>(x, w, bias, y, B, O, S); break;
      case 3: conv1x1_smallk_fwd_kernel<3><<<(unsigned)blocks, 256, sm, s>>>(x, w, bias, y, B, O, S); break;
      default: conv1x1_smallk_fwd_kernel<4><<<(unsigned)blocks, 256, sm, s>>>(x, w, bias, y, B, O, S); break;
    }
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  GemmP g = gemm_default();
  g.A = w; g.B = x; g.C = y;
  g.batch = B; g.M = O; g.N = (int)S; g.K = I;
  g.a = MatDesc{lin(0), lin(I), lin(1)};
  g.b = MatDesc{lin((long long)I * S), lin(S), lin(1)};
  g.c = MatDesc{lin((long long)O * S), lin(S), lin(1)};
  g.bias = bias; g.bias_mode = bias ? 2 : 0;
  g.act = act; g.Z = z; g.beta = beta;
  RPDE_CUDA(gemm(g, (cudaStream_t)stream));
  return RPDE_OK;
}

int rpde_conv1x1_bwd_input(const float* dy, const float* w, const float* zprev, int dact, float* dx, int accumulate,
                           int B, int I, int O, long long S, rpde_stream_t stream) {
  if (!dy || !w || !dx || B <= 0 || I <= 0 || O <= 0 || S <= 0 || S > INT_MAX || (dact && !zprev) ||
      (accumulate && dact))
    return fail(RPDE_E_INVALID, "rpde_conv1x1_bwd_input: bad arguments%s");
  GemmP g = gemm_default();
  g.A = w; g.B = dy; g.C = dx;
  g.batch = B; g.M = I; g.N = (int)S; g.K = O;
  g.a = MatDesc{lin(0), lin(1), lin(I)};
  g.b = MatDesc{lin((long long)O * S), lin(S), lin(1)};
  g.c = MatDesc{lin((long long)I * S), lin(S), lin(1)};
  g.beta = accumulate; g.aux = zprev; g.dact = dact;
  RPDE_CUDA(gemm(g, (cudaStream_t)stream));
  return RPDE_OK;
}

int rpde_conv1x1_bwd_params(const float* dy, const float* x, float* dw, float* db, int B, int I, int O,
                            long long S, rpde_stream_t stream) {
  if (!dy || !x || !dw || B <= 0 || I <= 0 || O <= 0 || S <= 0 || S > INT_MAX || (long long)B * S > INT_MAX)
    return fail(RPDE_E_INVALID, "rpde_conv1x1_bwd_params: bad arguments%s");
  cudaStream_t s = (cudaStream_t)stream;
  static const bool use_mma = !(getenv("RPDE_FNO2D_MMA") && atoi(getenv("RPDE_FNO2D_MMA")) == 0);
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  const bool mma_path = use_mma && S % 32 == 0 && I <= 64 && O <= 64 && al16(dy) && al16(x);
  if (mma_path && db == dw + (size_t)O * I) {      // dw | db back to back (the Python wrapper): one memset for both
    RPDE_CUDA(cudaMemsetAsync(dw, 0, ((size_t)O * I + O) * sizeof(float), s));
  } else {
    RPDE_CUDA(cudaMemsetAsync(dw, 0, (size_t)O * I * sizeof(float), s));
    if (mma_path && db) RPDE_CUDA(cudaMemsetAsync(db, 0, (size_t)O * sizeof(float), s));
  }
  if (mma_path) {
    // weight and bias gradient in one streaming pass on the warp-level tensor path (fno2d_mma.cuh)
    const int nb = ((I + 31) / 32) * ((O + 31) / 32), nstreams = 8 / nb;
    const long long tiles = (long long)B * (S / 32);
    // at least four tiles per warp: every CTA ends with one atomic per (o, i) element, so a small problem spread over
    // hundreds of CTAs is bound by a million contended atomics (19 us for 8 MB of operands at FNO1d batch 16)
    long long ctas = (tiles + 4 * nstreams - 1) / (4 * nstreams);
    if (ctas > 2 * 148) ctas = 2 * 148;
    conv1x1_wgrad_mma_kernel<<<(unsigned)ctas, 256, 0, s>>>(dy, x, dw, db, B, I, O, (int)S);
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  GemmP g = gemm_default();
  g.A = dy; g.B = x; g.C = dw;
  g.M = O; g.N = I; g.K = (int)(B * S);
  g.a = MatDesc{lin(0), lin(S), two((int)S, (long long)O * S, 1)};
  g.b = MatDesc{lin(0), two((int)S, (long long)I * S, 1), lin(S)};
  g.c = MatDesc{lin(0), lin(I), lin(1)};
  g.splitk = 0;
  RPDE_CUDA(gemm(g, s));
  if (db) return reduce_channels(dy, db, B, O, S, s);
  return RPDE_OK;
}

// ------------------------------------------------------------------------------------------------
// Fused projection MLP (fno2d_mma.cuh): the hidden layer never reaches HBM
// ------------------------------------------------------------------------------------------------
int rpde_mlp_supported(int I, int Hd, int O, long long S) {
  static const bool use_mma = !(getenv("RPDE_FNO2D_MMA") && atoi(getenv("RPDE_FNO2D_MMA")) == 0);
  const int HP = (Hd + 31) & ~31;
  return use_mma && I >= 1 && I <= 64 && Hd >= 1 && (HP == 32 || HP == 64 || HP == 128 || HP == 256) && O >= 1 && O <= MLP_MAXO &&
                 S > 0 && S % 32 == 0 && S <= INT_MAX
             ? 1 : 0;
}

int rpde_mlp_fwd(const float* x, const float* w1, const float* b1, const float* w2, const float* b2, float* y, int B, int I, int Hd,
                 int O, long long S, rpde_stream_t stream) {
  if (!x || !w1 || !w2 || !y || B <= 0) return fail(RPDE_E_INVALID, "rpde_mlp_fwd: bad arguments%s");
  if (!rpde_mlp_supported(I, Hd, O, S))
    return fail(RPDE_E_INVALID, "rpde_mlp_fwd: shape outside the fused kernel (I <= 64, hidden 32/64/128/256, O <= 4, S % 32 == 0)%s");
  cudaStream_t s = (cudaStream_t)stream;
  const size_t smem = mlp_fwd_smem(I, Hd);
  const long long tiles = (long long)B * (S / 32);
  const int per_sm = I <= 32 ? 2 : 1;
  long long ctas = (tiles + 7) / 8;
  if (ctas > 148LL * per_sm) ctas = 148LL * per_sm;
#define RPDE_MLP_FWD(KS, NO)                                                                                                    \
  do {                                                                                                                          \
    RPDE_CUDA(cudaFuncSetAttribute(mlp_fwd_mma_kernel<KS, NO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));        \
    mlp_fwd_mma_kernel<KS, NO><<<(unsigned)ctas, 256, smem, s>>>(x, w1, b1, w2, b2, y, B, I, Hd, O, (int)S);                    \
  } while (0)
  if (I <= 32) { if (O == 1) RPDE_MLP_FWD(4, 1); else RPDE_MLP_FWD(4, MLP_MAXO); }
  else { if (O == 1) RPDE_MLP_FWD(8, 1); else RPDE_MLP_FWD(8, MLP_MAXO); }
#undef RPDE_MLP_FWD
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_mlp_bwd(const float* x, const float* dy, const float* w1, const float* b1, const float* w2, float* dx, float* dw1,
                 float* db1, float* dw2, float* db2, int B, int I, int Hd, int O, long long S, rpde_stream_t stream) {
  if (!x || !dy || !w1 || !w2 || !dw1 || !dw2 || B <= 0) return fail(RPDE_E_INVALID, "rpde_mlp_bwd: bad arguments%s");
  if (!rpde_mlp_supported(I, Hd, O, S)) return fail(RPDE_E_INVALID, "rpde_mlp_bwd: shape outside the fused kernel%s");
  cudaStream_t s = (cudaStream_t)stream;
  // the four gradients are accumulated with atomics: zero-fill them, in one memset when the caller laid them out back to back
  // (dw1 | db1 | dw2 | db2, as the Python wrapper does: every launch counts on the latency-bound FNO1d step)
  if (db1 == dw1 + (size_t)Hd * I && dw2 == db1 + Hd && db2 == dw2 + (size_t)O * Hd) {
    RPDE_CUDA(cudaMemsetAsync(dw1, 0, ((size_t)Hd * I + Hd + (size_t)O * Hd + O) * sizeof(float), s));
  } else {
    RPDE_CUDA(cudaMemsetAsync(dw1, 0, (size_t)Hd * I * sizeof(float), s));
    RPDE_CUDA(cudaMemsetAsync(dw2, 0, (size_t)O * Hd * sizeof(float), s));
    if (db1) RPDE_CUDA(cudaMemsetAsync(db1, 0, (size_t)Hd * sizeof(float), s));
    if (db2) RPDE_CUDA(cudaMemsetAsync(db2, 0, (size_t)O * sizeof(float), s));
  }
  const size_t smem = mlp_bwd_smem(I, Hd);
  const int HP = (Hd + 31) & ~31, PT = 256 / HP;
  const long long tiles = (long long)B * (S / 32);
  long long ctas = (tiles + PT - 1) / PT;
  const long long resident = 148LL * (I <= 32 ? 2 : 1);
  if (ctas > resident) ctas = resident;
#define RPDE_MLP_BWD(KS, NO)                                                                                                    \
  do {                                                                                                                          \
    RPDE_CUDA(cudaFuncSetAttribute(mlp_bwd_mma_kernel<KS, NO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));        \
    mlp_bwd_mma_kernel<KS, NO><<<(unsigned)ctas, 256, smem, s>>>(x, dy, w1, b1, w2, dx, dw1, db1, dw2, db2, B, I, Hd, O, (int)S); \
  } while (0)
  if (I <= 32) { if (O == 1) RPDE_MLP_BWD(4, 1); else RPDE_MLP_BWD(4, MLP_MAXO); }
  else { if (O == 1) RPDE_MLP_BWD(8, 1); else RPDE_MLP_BWD(8, MLP_MAXO); }
#undef RPDE_MLP_BWD
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_act_bwd(const float* dy, const float* z, float* out, long long n, int act, rpde_stream_t stream) {
  if (!dy || !z || !out || n <= 0) return fail(RPDE_E_INVALID, "rpde_act_bwd: bad arguments%s");
  dact_mul_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(dy, z, out, n, act);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}
int rpde_act_fwd(const float* x, const float* res, float* out, long long n, int act, rpde_stream_t stream) {
  if (!x || !out || n <= 0) return fail(RPDE_E_INVALID, "rpde_act_fwd: bad arguments%s");
  act_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(x, res, out, n, act);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

// ------------------------------------------------------------------------------------------------
int rpde_layernorm_fwd(const float* x, const float* res, const float* gamma, const float* beta, float* y,
                       float* mean, float* rstd, long long T, int C, float eps, rpde_stream_t stream) {
  if (!x || !gamma || !beta || !y || T <= 0 || C <= 0 || (mean && !rstd))
    return fail(RPDE_E_INVALID, "rpde_layernorm_fwd: bad arguments%s");
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if ((C == 16 || C == 32 || C == 64 || C == 128) && al16(x) && al16(res) && al16(gamma) && al16(beta) && al16(y)) {
    // one float4 per lane, C / 4 lanes per row, four row groups in flight: streams at HBM rate
    const long long groups = (T * (C / 4) + 4 * 256 - 1) / (4 * 256);
    const unsigned grid = (unsigned)(groups < 148 * 8 ? (groups < 1 ? 1 : groups) : 148 * 8);
    cudaStream_t s = (cudaStream_t)stream;
    switch (C) {
      case 16: layernorm_fwd_vec_kernel<4><<<grid, 256, 0, s>>>(x, res, gamma, beta, y, mean, rstd, T, eps); break;
      case 32: layernorm_fwd_vec_kernel<8><<<grid, 256, 0, s>>>(x, res, gamma, beta, y, mean, rstd, T, eps); break;
      case 64: layernorm_fwd_vec_kernel<16><<<grid, 256, 0, s>>>(x, res, gamma, beta, y, mean, rstd, T, eps); break;
      default: layernorm_fwd_vec_kernel<32><<<grid, 256, 0, s>>>(x, res, gamma, beta, y, mean, rstd, T, eps); break;
    }
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  const long long threads = T * 32;
  layernorm_fwd_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, res, gamma, beta, y,
                                                                                           mean, rstd, T, C, eps);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}
int rpde_layernorm_bwd(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                       float* dx, float* dgamma, float* dbeta, float* dxsum, long long T, int C, rpde_stream_t stream) {
  if (!dy || !x || !gamma || !mean || !rstd || !dx || !dgamma || !dbeta || T <= 0 || C <= 0 || C > 4096)
    return fail(RPDE_E_INVALID, "rpde_layernorm_bwd: bad arguments%s");
  cudaStream_t s = (cudaStream_t)stream;
  RPDE_CUDA(cudaMemsetAsync(dgamma, 0, (size_t)C * sizeof(float), s));
  RPDE_CUDA(cudaMemsetAsync(dbeta, 0, (size_t)C * sizeof(float), s));
  if (dxsum) RPDE_CUDA(cudaMemsetAsync(dxsum, 0, (size_t)C * sizeof(float), s));
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if ((C == 16 || C == 32 || C == 64 || C == 128) && al16(dy) && al16(x) && al16(gamma) && al16(dx)) {
    const long long groups = (T * (C / 4) + 4 * 256 - 1) / (4 * 256);
    const unsigned grid = (unsigned)(groups < 148 * 4 ? (groups < 1 ? 1 : groups) : 148 * 4);
    switch (C) {
      case 16: layernorm_bwd_vec_kernel<4><<<grid, 256, 0, s>>>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, dxsum, T); break;
      case 32: layernorm_bwd_vec_kernel<8><<<grid, 256, 0, s>>>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, dxsum, T); break;
      case 64: layernorm_bwd_vec_kernel<16><<<grid, 256, 0, s>>>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, dxsum, T); break;
      default: layernorm_bwd_vec_kernel<32><<<grid, 256, 0, s>>>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, dxsum, T); break;
    }
    RPDE_LAUNCH_CHECK();
    return RPDE_OK;
  }
  long long rpb = (T + 148 * 8 - 1) / (148 * 8);
  if (rpb < 8) rpb = 8;
  const unsigned grid = (unsigned)((T + rpb - 1) / rpb);
  layernorm_bwd_kernel<<<grid, 256, 3 * C * sizeof(float), s>>>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, dxsum, T,
                                                               C, (int)rpb);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_weight_norm_fwd(const float* v, const float* g, float* w, float* norms, int out, int in,
                         rpde_stream_t stream) {
  if (!v || !g || !w || !norms || out <= 0 || in <= 0) return fail(RPDE_E_INVALID, "rpde_weight_norm_fwd: bad arguments%s");
  weight_norm_fwd_kernel<<<(out * 32 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(v, g, w, norms, out, in);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}
int rpde_weight_norm_bwd(const float* dw, const float* v, const float* g, const float* norms, float* dv, float* dg,
                         int out, int in, rpde_stream_t stream) {
  if (!dw || !v || !g || !norms || !dv || !dg || out <= 0 || in <= 0)
    return fail(RPDE_E_INVALID, "rpde_weight_norm_bwd: bad arguments%s");
  weight_norm_bwd_kernel<<<(out * 32 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(dw, v, g, norms, dv, dg, out, in);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_concat_grid_fwd(const float* x, const float* g0, const float* g1, float* out, int B, int Cin, int S0,
                         int S1, int channels_last, rpde_stream_t stream) {
  if (!x || !out || B <= 0 || Cin <= 0 || S0 <= 0 || S1 <= 0 || (g1 && !g0))
    return fail(RPDE_E_INVALID, "rpde_concat_grid_fwd: bad arguments%s");
  const int G = (g0 ? 1 : 0) + (g1 ? 1 : 0);
  const long long total = (long long)B * (Cin + G) * S0 * S1;
  concat_grid_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(x, g0, g1, out, B, Cin, S0, S1, G,
                                                                        channels_last);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}
int rpde_concat_grid_bwd(const float* dout, float* dx, int B, int Cin, int S0, int S1, int G, int channels_last,
                         rpde_stream_t stream) {
  if (!dout || !dx || B <= 0 || Cin <= 0 || S0 <= 0 || S1 <= 0 || G < 0)
    return fail(RPDE_E_INVALID, "rpde_concat_grid_bwd: bad arguments%s");
  const long long total = (long long)B * Cin * S0 * S1;
  concat_grid_bwd_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(dout, dx, B, Cin, S0, S1, G,
                                                                            channels_last);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_rel_l2_fwd(const float* x, const float* y, float* acc, float* per_sample, float* loss, int B, long long S,
                    int reduce_mode, rpde_stream_t stream) {
  if (!x || !y || !acc || !per_sample || B <= 0 || S <= 0 || (reduce_mode && !loss))
    return fail(RPDE_E_INVALID, "rpde_rel_l2_fwd: bad arguments%s");
  cudaStream_t s = (cudaStream_t)stream;
  RPDE_CUDA(cudaMemsetAsync(acc, 0, (size_t)2 * B * sizeof(float), s));
  long long per = (148LL * 8 + B - 1) / B;
  long long chunk = (S + per - 1) / per;
  if (chunk < 4096) chunk = 4096;
  dim3 grid((unsigned)((S + chunk - 1) / chunk), (unsigned)B);
  rel_l2_partial_kernel<<<grid, 256, 0, s>>>(x, y, acc, S, chunk);
  RPDE_LAUNCH_CHECK();
  rel_l2_finalize_kernel<<<1, 256, 0, s>>>(acc, per_sample, loss, B, reduce_mode);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}
int rpde_rel_l2_bwd(const float* x, const float* y, const float* acc, const float* gout, float* dx, int B,
                    long long S, int reduce_mode, rpde_stream_t stream) {
  if (!x || !y || !acc || !gout || !dx || B <= 0 || S <= 0) return fail(RPDE_E_INVALID, "rpde_rel_l2_bwd: bad arguments%s");
  long long gx = (S + 255) / 256;
  const long long cap = (148LL * 16 + B - 1) / B;
  if (gx > cap) gx = cap;
  dim3 grid((unsigned)gx, (unsigned)B);
  rel_l2_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, y, acc, gout, dx, S, B, reduce_mode);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

// ---- rpde_rel_l2_fwd_static: the same two kernels as rpde_rel_l2_fwd with the 2*B-float accumulator in a
// library-owned array instead of a caller buffer, for callers that may not allocate (CUDA graph capture, serving
// loops).  Every device that loads the library holds the 32 KB until exit.  All calls on a device share the array:
// it is cleared at the start of every call (it keeps the previous call's sums), and calls that overlap on the GPU race.
static __device__ __align__(256) float g_rel_l2_static_acc[2 * RPDE_REL_L2_STATIC_MAX_B];

// In host code g_rel_l2_static_acc is only the host-side stand-in of the symbol; the device address is looked up once
// per device (with that device current) and cached.
static int rel_l2_static_acc_ptr(float** out) {
  constexpr int kMaxDev = 64;
  static float* ptrs[kMaxDev] = {};
  static std::mutex mu;
  int dev = -1;
  RPDE_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDev) return fail(RPDE_E_INVALID, "rpde_rel_l2_fwd_static: device index out of range%s");
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    float* p = nullptr;
    RPDE_CUDA(cudaGetSymbolAddress((void**)&p, g_rel_l2_static_acc));
    if (!p) return fail(RPDE_E_CUDA, "rpde_rel_l2_fwd_static: cudaGetSymbolAddress returned a null address%s");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return RPDE_OK;
}

int rpde_rel_l2_fwd_static(const float* x, const float* y, float* per_sample, float* loss, int B, long long S,
                           int reduce_mode, rpde_stream_t stream) {
  if (!x || !y || !per_sample || B <= 0 || S <= 0 || (reduce_mode && !loss))
    return fail(RPDE_E_INVALID, "rpde_rel_l2_fwd_static: bad arguments%s");
  if (B > RPDE_REL_L2_STATIC_MAX_B) {
    snprintf(g_err, sizeof(g_err), "rpde_rel_l2_fwd_static: B = %d exceeds RPDE_REL_L2_STATIC_MAX_B = %d "
             "(use rpde_rel_l2_fwd with a caller-owned accumulator)", B, RPDE_REL_L2_STATIC_MAX_B);
    return RPDE_E_WORKSPACE;
  }
  float* acc = nullptr;
  const int rc = rel_l2_static_acc_ptr(&acc);
  if (rc != RPDE_OK) return rc;
  return rpde_rel_l2_fwd(x, y, acc, per_sample, loss, B, S, reduce_mode, stream);
}

int rpde_rollout_renorm(const float* pred, float* preds_out, float* next_state, int B, long long S,
                        long long pred_batch_stride, long long preds_batch_stride, float y_scale, float y_mean,
                        float x_mean, float x_scale, int renorm, rpde_stream_t stream) {
  if (!pred || !next_state || B <= 0 || S <= 0 || pred_batch_stride < S || (preds_out && preds_batch_stride < S) ||
      (renorm && x_scale == 0.f))
    return fail(RPDE_E_INVALID, "rpde_rollout_renorm: bad arguments%s");
  const bool vec = S % 4 == 0 && pred_batch_stride % 4 == 0 && (!preds_out || preds_batch_stride % 4 == 0) &&
                   (((uintptr_t)pred | (uintptr_t)preds_out | (uintptr_t)next_state) & 15) == 0;
  const long long n = vec ? S / 4 : S;
  long long gx = (n + 255) / 256;
  const long long cap = (148LL * 8 + B - 1) / B;
  if (gx > cap) gx = cap;
  dim3 grid((unsigned)gx, (unsigned)B);
  if (vec)
    rollout_renorm_kernel<4><<<grid, 256, 0, (cudaStream_t)stream>>>(pred, preds_out, next_state, S, pred_batch_stride,
                                                                     preds_batch_stride, y_scale, y_mean, x_mean,
                                                                     x_scale, renorm);
  else
    rollout_renorm_kernel<1><<<grid, 256, 0, (cudaStream_t)stream>>>(pred, preds_out, next_state, S, pred_batch_stride,
                                                                     preds_batch_stride, y_scale, y_mean, x_mean,
                                                                     x_scale, renorm);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_rollout_loss(const float* preds, const float* truth, float* acc, float* step_losses, float* batch_loss, int B,
                      int T, long long S, long long preds_batch_stride, long long preds_step_stride,
                      long long truth_batch_stride, long long truth_step_stride, float y_scale, float y_mean,
                      int decode, rpde_stream_t stream) {
  if (!preds || !truth || !acc || !step_losses || !batch_loss || B <= 0 || T <= 0 || S <= 0 ||
      (long long)B * T > 0x7fffffffLL)
    return fail(RPDE_E_INVALID, "rpde_rollout_loss: bad arguments%s");
  cudaStream_t s = (cudaStream_t)stream;
  const long long BT = (long long)B * T;
  RPDE_CUDA(cudaMemsetAsync(acc, 0, (size_t)2 * BT * sizeof(float), s));
  long long per = (148LL * 8 + BT - 1) / BT;
  long long chunk = (S + per - 1) / per;
  if (chunk < 4096) chunk = 4096;
  if ((S + chunk - 1) / chunk > 65535) chunk = (S + 65534) / 65535;
  dim3 grid((unsigned)BT, (unsigned)((S + chunk - 1) / chunk));
  rollout_loss_partial_kernel<<<grid, 256, 0, s>>>(preds, truth, acc, B, S, chunk, preds_batch_stride, preds_step_stride,
                                                   truth_batch_stride, truth_step_stride, y_scale, y_mean, decode);
  RPDE_LAUNCH_CHECK();
  rollout_loss_finalize_kernel<<<1, 256, 0, s>>>(acc, step_losses, batch_loss, B, T);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_adamw_step(const void* table, int ntensors, long long nchunks, float* hyper, double beta1, double beta2,
                    float eps, float weight_decay, rpde_stream_t stream) {
  if (!table || !hyper || ntensors <= 0 || nchunks <= 0 || !(beta1 >= 0.f && beta1 < 1.f) || !(beta2 >= 0.f && beta2 < 1.f))
    return fail(RPDE_E_INVALID, "rpde_adamw_step: bad arguments%s");
  static_assert(sizeof(AdamWTensor) == 48, "the table layout is part of the ABI (6 x 8 bytes per tensor)");
  cudaStream_t s = (cudaStream_t)stream;
  adamw_tick_kernel<<<1, 1, 0, s>>>(hyper, beta1, beta2);
  RPDE_LAUNCH_CHECK();
  const long long blocks = nchunks < 148LL * 16 ? nchunks : 148LL * 16;
  adamw_multi_kernel<<<(unsigned)blocks, 256, 0, s>>>(reinterpret_cast<const AdamWTensor*>(table), ntensors, hyper, (float)beta1,
                                                     (float)beta2, (float)(1.0 - beta1), (float)(1.0 - beta2), eps,
                                                     weight_decay, nchunks);
  RPDE_LAUNCH_CHECK();
  return RPDE_OK;
}

int rpde_matmul(const float* a, const float* b, float* c, int M, int N, int K, int trans_a, int trans_b,
                rpde_stream_t stream) {
  if (!a || !b || !c || M <= 0 || N <= 0 || K <= 0) return fail(RPDE_E_INVALID, "rpde_matmul: bad arguments%s");
  GemmP g = gemm_default();
  g.A = a; g.B = b; g.C = c;
  g.M = M; g.N = N; g.K = K;
  g.a = trans_a ? MatDesc{lin(0), lin(1), lin(M)} : MatDesc{lin(0), lin(K), lin(1)};
  g.b = trans_b ? MatDesc{lin(0), lin(1), lin(K)} : MatDesc{lin(0), lin(N), lin(1)};
  g.c = MatDesc{lin(0), lin(N), lin(1)};
  RPDE_CUDA(gemm(g, (cudaStream_t)stream));
  return RPDE_OK;
}

}  // extern "C"
