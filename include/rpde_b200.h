/* rpde_b200.h -- C ABI of the B200-native spectral-convolution path of resolution-pde.
 *
 * One shared library (librpde_b200.so), plain pointers and sizes, no torch types.  Every entry point
 * enqueues work on the given CUDA stream and returns immediately: 0 on success, a negative RPDE_E_* code
 * otherwise (text via rpde_last_error(), thread-local).  The library never allocates, frees or retains
 * device memory: the caller owns inputs, outputs and workspaces (sizes from the *_workspace_bytes
 * queries) and keeps them alive until the stream work has completed.  All tensors are contiguous fp32
 * in the documented layout; complex tensors are interleaved (re, im) pairs.  The calls are re-entrant,
 * hold no global mutable state, never synchronise and are CUDA-graph capturable.
 *
 * Each entry point replaces a piece of the reference's PyTorch operator path (file:line cited per
 * function; paths are relative to the reference repository root).
 */
#ifndef RPDE_B200_H
#define RPDE_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* rpde_stream_t; /* a cudaStream_t */

enum {
  RPDE_OK = 0,
  RPDE_E_INVALID = -1,     /* bad argument / unsupported shape */
  RPDE_E_WORKSPACE = -2,   /* workspace too small */
  RPDE_E_CUDA = -3,        /* CUDA launch error (text in rpde_last_error) */
  RPDE_E_MODES = -4        /* more modes requested than the spectrum holds (reference raises RuntimeError) */
};
enum { RPDE_ACT_NONE = 0, RPDE_ACT_GELU = 1, RPDE_ACT_RELU = 2 };
/* rpde_linear_fwd: or-ed into `act`: z receives act'(pre-activation) instead of the pre-activation, so the backward
 * pass multiplies by a stored factor (dact = RPDE_DACT_SAVED) and evaluates no transcendental. */
enum { RPDE_ACT_SAVE_GRAD = 0x100 };
enum { RPDE_DACT_SAVED = 3 };
enum { RPDE_MIX_FULL = 0, RPDE_MIX_LOWPASS = 1 };
/* DFT table kinds (rpde_dft_table) */
enum { RPDE_DFT_R2C_ANALYSIS = 0, RPDE_DFT_C2R_SYNTHESIS = 1, RPDE_DFT_C2C_ANALYSIS_ROWS = 2, RPDE_DFT_C2C_SYNTHESIS_ROWS = 3,
       RPDE_DFT_C2R_SYNTHESIS_PLANES = 4 };
/* precision of the contraction engine */
enum { RPDE_PREC_FP32 = 0, RPDE_PREC_3XTF32 = 1, RPDE_PREC_TF32 = 2,
       RPDE_PREC_BF16 = 3 /* fused BF16 FeedForward (rpde_ff3_*) where it applies, the TF32 engine elsewhere */ };

int rpde_version(void);
const char* rpde_last_error(void);
/* Selects the contraction engine for aligned shapes (per thread). FP32 = CUDA-core FMA everywhere. */
int rpde_set_precision(int prec);
int rpde_get_precision(void);
/* Number of kernels this library has enqueued since it was loaded (statistics for the benchmark). */
unsigned long long rpde_launch_count(void);
/* ... of which contractions executed by the tcgen05/TMA tensor-core engine (0 under RPDE_PREC_FP32). */
unsigned long long rpde_tc_launch_count(void);
/* ... and contractions that ran on the CUDA-core FMA engine (every one under RPDE_PREC_FP32; under the tensor-core
 * precisions only shapes TMA cannot address or launch-bound ones).  Tests assert on it so that a layout regression
 * cannot hide as a slowdown. */
unsigned long long rpde_simt_launch_count(void);

/* Launch trace (bench.py's in-situ kernel table): while enabled, every contraction records {M, N, K, batch, split-K,
 * engine (0 CUDA-core FMA, 1 tcgen05), epilogue flags (1 beta, 2 act, 4 dact, 8 Z, 16 colsum, 32 bias), 0} in launch order.
 * rpde_trace_enable(on) clears the list.  Host-side bookkeeping only; nothing is recorded while it is off. */
int rpde_trace_enable(int on);
int rpde_trace_count(void);
int rpde_trace_get(int i, long long* out8);

/* Twiddle / truncated-DFT tables, generated on the device in fp64 and rounded once to fp32.
 * Replaces the cuFFT/pocketfft plans behind torch.fft.rfft/irfft/rfft2/irfft2
 * (models/spectral_convolution.py:41,54,82,97,165,198,265,284,289,308).
 *   kind 0: out[k][2][n]   analysis, scale s          kind 1: out[n][k][2]   synthesis (c_k weights), scale s
 *   kind 2: out[j][2][2][n] C2C analysis of rows [0,m1) U [n-m1,n)   kind 3: out[n][2][j][2] its inverse
 *   kind 4: out[2][k][n]   the kind-1 coefficients as two planes (re, im) with unit stride along n  */
int rpde_dft_table(float* out, int kind, int n, int k, int m1, double scale, rpde_stream_t stream);

/* ---- FNO SpectralConv1d (models/spectral_convolution.py:24-55), channels-first (B,I,N) -> (B,O,N) --------
 * w: complex64 (I,O,modes) interleaved.  fa: kind-0 table [modes][2][N] scale 1, fs: kind-1 table scale 1/N.
 * y = act(spectral(x) + (beta ? y : 0)); z (optional) receives the pre-activation.  spec (B*2I*modes floats,
 * laid out [mode][re|im][b][i]) receives the retained spectrum for the backward pass.  modes > N/2+1 -> RPDE_E_MODES. */
size_t rpde_spectral1d_workspace_bytes(int B, int I, int O, int N, int modes);
int rpde_spectral1d_fwd(const float* x, const float* w, const float* fa, const float* fs, float* spec, float* y,
                        float* z, int B, int I, int O, int N, int modes, int beta, int act, void* ws,
                        size_t ws_bytes, rpde_stream_t stream);
/* Fused FNOBlock1d forward (models/fno_blocks.py:25-33: activation(spectral_conv(x) + bypass_conv(x)), around
 * spectral_convolution.py:38-55) in three launches: analysis (one register-resident row per warp), per-mode mixing straight
 * from the parameter layout, then synthesis + 1x1 bypass conv + bias + activation.  fp32 FMA (engine-independent).  fs4: kind-4 table [2][modes][N]
 * scale 1/N.  wb (O,I) / bb (O) may be NULL (plain SpectralConv1d).  spec ([mode][re|im][b][i]) and z (pre-activation)
 * may be NULL when no backward pass follows.  ws: modes*B*2*O floats (+ modes*B*2*I when spec is NULL), each rounded up
 * to 256 bytes.  Returns RPDE_E_INVALID when N > 2048 or modes*I exceeds the mixing kernel's shared memory (callers then use rpde_conv1x1_fwd + rpde_spectral1d_fwd). */
int rpde_fno_block1d_fwd(const float* x, const float* w, const float* fa, const float* fs4, const float* wb,
                         const float* bb, float* spec, float* y, float* z, int B, int I, int O, int N, int modes, int act,
                         void* ws, size_t ws_bytes, rpde_stream_t stream);
/* dx (+)= adjoint wrt x ; dw (complex64 (I,O,modes)) = gradient wrt w.  dx may be NULL. */
int rpde_spectral1d_bwd(const float* dy, const float* w, const float* fa, const float* fs, const float* spec,
                        float* dx, int dx_accumulate, float* dw, int B, int I, int O, int N, int modes, void* ws,
                        size_t ws_bytes, rpde_stream_t stream);

/* Backward of the fused block: rpde_spectral1d_bwd with a fused prologue.  dpre = dy * act'(z) (written when z != NULL; this
 * is the gradient the bypass conv's backward consumes) and its projection on the synthesis table come from one launch
 * instead of rpde_act_bwd + memset + split-K contraction; the rest (weight gradient, mixing adjoint, analysis adjoint) is
 * rpde_spectral1d_bwd's.  wb (O,I) is the explicit switch for the fused input gradient: when it is given (needs dx,
 * dx_accumulate = 0, B <= 65535, else RPDE_E_INVALID) dx = analysis adjoint + wb^T dpre (the bypass conv's input gradient)
 * comes from one kernel, so the caller must NOT run rpde_conv1x1_bwd_input as well; with wb = NULL dx is the spectral part
 * only.  No environment variable is read.  rpde_fno_block1d_bwd_fuses_dx() returns 1 when the library supports the wb
 * form.  fs4: kind-4 table scale 1/N.  z = dpre = NULL for act == none.  N <= 2048.  Workspace as
 * rpde_spectral1d_workspace_bytes. */
int rpde_fno_block1d_bwd_fuses_dx(void);
int rpde_fno_block1d_bwd(const float* dy, const float* z, int act, const float* w, const float* fa, const float* fs4,
                         const float* spec, const float* wb, float* dpre, float* dx, int dx_accumulate, float* dw, int B, int I,
                         int O, int N, int modes, void* ws, size_t ws_bytes, rpde_stream_t stream);

/* ---- FNO SpectralConv2d (models/spectral_convolution.py:61-98), (B,I,H,W) -> (B,O,H,W) ---------------------
 * w1, w2: complex64 (I,O,m1,m2).  fa_y/fs_y: kind 0/4 tables for axis W with k = m2 (scales 1, 1/W);
 * gf/gi: kind 2/3 tables for axis H with j = 2*m1 (scales 1, 1/H).  spec: (2*m1*m2*B*2I floats). */
size_t rpde_spectral2d_workspace_bytes(int B, int I, int O, int H, int W, int m1, int m2);
int rpde_spectral2d_fwd(const float* x, const float* w1, const float* w2, const float* fa_y, const float* fs_y,
                        const float* gf, const float* gi, float* spec, float* y, float* z, int B, int I, int O,
                        int H, int W, int m1, int m2, int beta, int act, void* ws, size_t ws_bytes,
                        rpde_stream_t stream);
int rpde_spectral2d_bwd(const float* dy, const float* w1, const float* w2, const float* fa_y, const float* fs_y,
                        const float* gf, const float* gi, const float* spec, float* dx, int dx_accumulate,
                        float* dw1, float* dw2, int B, int I, int O, int H, int W, int m1, int m2, void* ws,
                        size_t ws_bytes, rpde_stream_t stream);

/* Fused FNOBlock2d (models/fno_blocks.py:63-71 around spectral_convolution.py:79-98): y = act(SpectralConv2d(x) + bypass 1x1
 * conv(x) + bias) as five fp32-FMA launches (W-axis analysis over the image rows, H-axis analysis per plane, per-mode mixing
 * straight from the complex (I,O,m1,m2) parameters, H-axis synthesis, W-axis synthesis + bypass + bias + activation); only the
 * W-axis half transform (2*m2 floats per image row) and the J x m2 K-space tiles pass through the workspace.
 * fa_y: kind-0 table of W with m2 modes (scale 1); fs4_y: kind-4 table (scale 1/W); g2t: the kind-2 table of H with 2*m1 rows
 * (scale 1) transposed to [(h,ri)][(j,ro)]; g3: the kind-3 table (scale 1/H, [(h,ro)][(j,ri)] as generated).  spec (B*I*4*m1*m2 floats, [(b,i)][(j,re|im)][ky]) and z (pre-activation) may be NULL when no
 * backward pass follows; wb (O,I) / bb (O) may be NULL.  rpde_fno_block2d_supported() tells whether a shape is covered
 * (B*H <= 65535, W <= 1024, m2 <= 16, m1 <= 64 with 4*m1*ceil(m2/4) <= 256); otherwise callers compose rpde_conv1x1_fwd +
 * rpde_spectral2d_fwd.  Backward: the same two H-axis tables, used by the adjoint kernels the other way round.
 * dpre = dy * act'(z) (z = dpre = NULL for act none) is what the bypass conv's parameter gradient consumes; dx = the COMPLETE
 * input gradient (spectral adjoint + wb^T dpre when wb is given: do not run rpde_conv1x1_bwd_input as well); dw1 / dw2 are
 * written in the complex parameter layout.  Workspace: rpde_fno_block2d_workspace_bytes for both directions. */
int rpde_fno_block2d_supported(int B, int I, int O, int H, int W, int m1, int m2);
size_t rpde_fno_block2d_workspace_bytes(int B, int I, int O, int H, int W, int m1, int m2);
int rpde_fno_block2d_fwd(const float* x, const float* w1, const float* w2, const float* fa_y, const float* fs4_y, const float* g2t,
                         const float* g3, const float* wb, const float* bb, float* spec, float* y, float* z, int B, int I, int O,
                         int H, int W, int m1, int m2, int act, void* ws, size_t ws_bytes, rpde_stream_t stream);
int rpde_fno_block2d_bwd(const float* dy, const float* z, int act, const float* w1, const float* w2, const float* fa_y,
                         const float* fs4_y, const float* g3, const float* g2t, const float* spec, const float* wb, float* dpre,
                         float* dx, float* dw1, float* dw2, int B, int I, int O, int H, int W, int m1, int m2, void* ws,
                         size_t ws_bytes, rpde_stream_t stream);

/* ---- F-FNO factorised spectral pass along one axis (models/spectral_convolution.py:159-204, 256-318) -------
 * x is channels-last, viewed as [outer][n][rest][C]; the transform runs along n, the complex channel mixing
 * over C with w = fourier_weight (C,C,w_modes,2) restricted to its first k = min(n_modes, n/2+1) modes.
 * y (+)= irfft(mix(rfft(x)[:k])) ; spec ([outer][k][2][rest][C]) is kept for the backward pass.
 * 1-D layer: outer=B, rest=1.  2-D layer: y-pass outer=B*M, n=N, rest=1; x-pass outer=B, n=M, rest=N with
 * accumulate=1 (the reference's xx + xy, spectral_convolution.py:313). */
size_t rpde_ffno_axis_workspace_bytes(long long outer, int n, long long rest, int C, int k);
int rpde_ffno_axis_fwd(const float* x, const float* w, const float* fa, const float* fs, float* spec, float* y,
                       int accumulate, long long outer, int n, long long rest, int C, int k, int w_modes, int mix,
                       void* ws, size_t ws_bytes, rpde_stream_t stream);
/* dx (+)= ... ; dw (C,C,w_modes,2) (+)= ... (modes >= k receive exactly zero).  dw may be NULL (low-pass). */
int rpde_ffno_axis_bwd(const float* dy, const float* w, const float* fa, const float* fs, const float* spec,
                       float* dx, int dx_accumulate, float* dw, int dw_accumulate, long long outer, int n,
                       long long rest, int C, int k, int w_modes, int mix, void* ws, size_t ws_bytes,
                       rpde_stream_t stream);

/* ---- channels-last Linear with fused bias / activation (models/custom_layer.py:49-68 FeedForward layers,
 * ffno.py:55,75,160,179 in_proj/out_proj).  y[T,out] = act(x[T,in] W^T + b); z (optional) = pre-activation.
 * y_cf_spatial > 0: y is written channels-first (T/S, out, S) (ffno.py:123, 235 final permute). */
int rpde_linear_fwd(const float* x, const float* w, const float* bias, float* y, float* z, long long T, int in,
                    int out, int act, long long y_cf_spatial, rpde_stream_t stream);
/* dx (+)= (dy W) * act'(zprev)   (zprev/dact describe the activation that produced this layer's input;
 * dact = RPDE_DACT_SAVED: zprev already holds act').  dbias_prev (optional, `in` floats, overwritten) receives the
 * column sums of dx = the bias gradient of the layer that produced this layer's input. */
int rpde_linear_bwd_input(const float* dy, const float* w, const float* zprev, int dact, float* dx, int accumulate,
                          float* dbias_prev, long long T, int in, int out, long long dy_cf_spatial,
                          rpde_stream_t stream);
/* dw[out,in] = dy^T x ; db[out] = column sums of dy (db may be NULL) */
int rpde_linear_bwd_params(const float* dy, const float* x, float* dw, float* db, long long T, int in, int out,
                           long long dy_cf_spatial, rpde_stream_t stream);

/* ---- channels-first 1x1 convolution (models/fno.py:30,93 lifting; fno_blocks.py:29,38-39,67,76-77) ----------
 * y[b,o,s] = act(sum_i w[o,i] x[b,i,s] + bias[o] + (beta ? y : 0)) */
int rpde_conv1x1_fwd(const float* x, const float* w, const float* bias, float* y, float* z, int B, int I, int O,
                     long long S, int act, int beta, rpde_stream_t stream);
int rpde_conv1x1_bwd_input(const float* dy, const float* w, const float* zprev, int dact, float* dx, int accumulate,
                           int B, int I, int O, long long S, rpde_stream_t stream);
int rpde_conv1x1_bwd_params(const float* dy, const float* x, float* dw, float* db, int B, int I, int O,
                            long long S, rpde_stream_t stream);

/* ---- projection MLP of FNO1d / FNO2d as one kernel per direction (models/fno_blocks.py:35-45 MLP1d, :73-83 MLP2d;
 * fno.py:20,45: mid_channels = 4 * width): y = W2 gelu(W1 x + b1) + b2 on channels-first x (B, I, S), w1 (Hd, I), w2 (O, Hd).
 * The hidden activations stay in registers (forward) and are recomputed from x (backward): nothing but x, dy, y, dx and the
 * parameter gradients crosses HBM.  rpde_mlp_supported: I <= 64, Hd rounded up to 32 in {32, 64, 128, 256}, O <= 4,
 * S % 32 == 0; other shapes use rpde_conv1x1_fwd / _bwd_input / _bwd_params.  dx may be NULL; db1 / db2 may be NULL.
 * The parameter gradients are overwritten (zero-filled inside, accumulated with atomics). */
int rpde_mlp_supported(int I, int Hd, int O, long long S);
int rpde_mlp_fwd(const float* x, const float* w1, const float* b1, const float* w2, const float* b2, float* y, int B, int I,
                 int Hd, int O, long long S, rpde_stream_t stream);
int rpde_mlp_bwd(const float* x, const float* dy, const float* w1, const float* b1, const float* w2, float* dx, float* dw1,
                 float* db1, float* dw2, float* db2, int B, int I, int Hd, int O, long long S, rpde_stream_t stream);

/* out = dy * act'(z)   (backward of the FNO block activation, fno_blocks.py:33,71) */
int rpde_act_bwd(const float* dy, const float* z, float* out, long long n, int act, rpde_stream_t stream);
/* out = (res ? res : 0) + act(x)   (FSpectralConv1d activation + FFNO1D residual, spectral_convolution.py:155; ffno.py:118) */
int rpde_act_fwd(const float* x, const float* res, float* out, long long n, int act, rpde_stream_t stream);

/* ---- LayerNorm over the last dim with optional residual (custom_layer.py:61-62; ffno.py:118,230) ----------
 * y = (res ? res : 0) + LN(x) * gamma + beta ; mean/rstd (T floats each, optional) kept for backward. */
int rpde_layernorm_fwd(const float* x, const float* res, const float* gamma, const float* beta, float* y,
                       float* mean, float* rstd, long long T, int C, float eps, rpde_stream_t stream);
/* dxsum (optional, C floats): column sums of dx = bias gradient of the Linear that feeds the LayerNorm */
int rpde_layernorm_bwd(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                       float* dx, float* dgamma, float* dbeta, float* dxsum, long long T, int C, rpde_stream_t stream);

/* ---- weight-normed Linear weights (custom_layer.py:70-108, legacy weight_norm dim=0) ------------------- */
int rpde_weight_norm_fwd(const float* v, const float* g, float* w, float* norms, int out, int in,
                         rpde_stream_t stream);
int rpde_weight_norm_bwd(const float* dw, const float* v, const float* g, const float* norms, float* dv, float* dg,
                         int out, int in, rpde_stream_t stream);

/* ---- coordinate-grid concatenation (fno.py:56-68,131-141; ffno.py:98-109,210-222) ------------------------
 * x (B,Cin,S0,S1) channels-first -> out with G = (g0?1:0)+(g1?1:0) extra channels, channels-first or -last. */
int rpde_concat_grid_fwd(const float* x, const float* g0, const float* g1, float* out, int B, int Cin, int S0,
                         int S1, int channels_last, rpde_stream_t stream);
int rpde_concat_grid_bwd(const float* dout, float* dx, int B, int Cin, int S0, int S1, int G, int channels_last,
                         rpde_stream_t stream);

/* ---- RelativeL2Loss (utils/loss.py:31-59).  acc: 2*B floats scratch kept for backward; per_sample: B floats;
 * loss: 1 float (reduce_mode 1 = mean, 2 = sum, 0 = none -> only per_sample is written). */
int rpde_rel_l2_fwd(const float* x, const float* y, float* acc, float* per_sample, float* loss, int B, long long S,
                    int reduce_mode, rpde_stream_t stream);
int rpde_rel_l2_bwd(const float* x, const float* y, const float* acc, const float* gout, float* dx, int B,
                    long long S, int reduce_mode, rpde_stream_t stream);
/* rpde_rel_l2_fwd_static: no counterpart in the reference; added for callers
 * that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  Same kernels and results as rpde_rel_l2_fwd, but the 2*B-float accumulator lives in a file-scope
 * __device__ array of the library (2 * RPDE_REL_L2_STATIC_MAX_B floats = 32 KB, held by every GPU that loads the
 * library until exit) instead of a caller buffer; the call allocates nothing (no cudaMalloc / cudaFree / torch).
 * B > RPDE_REL_L2_STATIC_MAX_B returns RPDE_E_WORKSPACE with a message in rpde_last_error().  The device that owns
 * x / y / stream must be the current device.  All calls on one device share the array: it is cleared at the start of
 * each call, and two calls that overlap on the GPU (different streams, no dependency) race -- serialise them.  The
 * accumulator is not handed out, so there is no matching backward: use rpde_rel_l2_fwd when a gradient follows. */
#define RPDE_REL_L2_STATIC_MAX_B 4096
int rpde_rel_l2_fwd_static(const float* x, const float* y, float* per_sample, float* loss, int B, long long S,
                           int reduce_mode, rpde_stream_t stream);

/* ---- autoregressive rollout on the device (utils/autoregressive_step.py:284-309 perform_rollout_1d and the
 * per-step loss loop :170-196).  The state never leaves HBM between steps.
 * rpde_rollout_renorm: one pass over the model output of step t, pred[b] = S contiguous floats at
 * pred + b*pred_batch_stride (S = all non-batch elements: N in 1-D, H*W in 2-D; the stride lets the initial
 * condition trajectories[:, 0] be encoded with y_scale = 1, y_mean = 0 and no gather copy, :155-162).
 * preds_out + b*preds_batch_stride receives pred[b] unchanged (the reference's cat of normalised predictions;
 * the caller passes the pointer of step t; NULL = no such store).  next_state (B, S) receives the next
 * model input: renorm = 1 -> x_normalizer.encode(y_normalizer.decode(pred)) = ((pred*y_scale + y_mean) - x_mean)
 * / x_scale with y_scale = std_y + eps and x_scale = std_x + eps (SimpleNormalizer, ks_naive_markov.py:396-406;
 * four separately rounded fp32 operations like the reference's tensor-scalar ops); renorm = 0 -> pred (the
 * reference's branch when a normaliser is None, :305-306). */
int rpde_rollout_renorm(const float* pred, float* preds_out, float* next_state, int B, long long S,
                        long long pred_batch_stride, long long preds_batch_stride, float y_scale, float y_mean,
                        float x_mean, float x_scale, int renorm, rpde_stream_t stream);
/* rpde_rollout_loss: every step's RelativeL2Loss(size_average=True) of the decoded rollout against the ground
 * truth in one launch pair, instead of one loss call and one .item() per step (:192-196).  preds[b,t,:] at
 * preds + b*preds_batch_stride + t*preds_step_stride, truth likewise (strides in floats: truth may be the view
 * trajectories[:, 1:T+1] of a (B, T_all, S) tensor).  decode = 1 applies pred*y_scale + y_mean first (:176-181).
 * acc: 2*B*T floats scratch; step_losses: T floats, step_losses[t] = mean_b ||p-y|| / (||y|| + 1e-8);
 * batch_loss: 1 float = sum_t step_losses[t] / T.  B*T < 2^31 (no grid-dimension limit). */
int rpde_rollout_loss(const float* preds, const float* truth, float* acc, float* step_losses, float* batch_loss, int B,
                      int T, long long S, long long preds_batch_stride, long long preds_step_stride,
                      long long truth_batch_stride, long long truth_step_stride, float y_scale, float y_mean,
                      int decode, rpde_stream_t stream);

/* ---- data preparation that is not a contraction -------------------------------------------------
 * rpde_moments_accumulate: acc3 = {count, sum, sum of squares} (fp64, zero-initialised by the caller) += over n floats;
 * rpde_moments_finalize: mean_std = {mean, unbiased std} as torch's .mean() / .std() -- the reference's global normaliser
 * statistics, which it computes from a Python list of every scalar (dataloaders/ks_naive_markov.py:377-390,
 * dataloaders/ns_naive_true_multires.py:513-525).
 * rpde_stride_downsample: y[o][i][j][c] = x[o][i*fh][j*fw][c] (x: outer x H x W x C, y: outer x ceil(H/fh) x ceil(W/fw) x C),
 * the naive down-sampling data[:, :, ::f, ::f, :] of dataloaders/ns_naive_true_multires.py:363-366 (1-D: H = fh = 1).
 * rpde_binned_sums_f64: out[r][b] = sum_j p[r][j] [bin[j] == b] in fp64 (radial-band totals of utils/frequency_error.py:115-150;
 * bin = -1 selects nothing). */
int rpde_moments_accumulate(const float* x, long long n, double* acc3, rpde_stream_t stream);
int rpde_moments_finalize(const double* acc3, float* mean_std, rpde_stream_t stream);
int rpde_stride_downsample(const float* x, float* y, long long outer, int H, int W, int C, int fh, int fw, rpde_stream_t stream);
int rpde_binned_sums_f64(const double* p, const int* bin, double* out, int rows, int nbins, long long n, rpde_stream_t stream);

/* ---- fused multi-tensor AdamW (torch.optim.AdamW as the reference builds it: main_1d.py:144, main_2d.py:173;
 * SURVEY 8f rank 1).  table: device array of ntensors records {float* p; const float* g; float* m; float* v;
 * int64 n; int64 chunk0} (48 bytes each; chunk0 = index of the tensor's first 1024-element chunk, ascending);
 * nchunks = total chunks; hyper: device float[4] = {lr, step count, 1 - beta1^t, 1 - beta2^t} (the last two are written
 * by the call; betas are doubles because torch forms 1 - beta on the host in double).  The call advances the step count on the
 * device, then applies   p *= 1 - lr*wd;  m += (g - m)(1 - beta1);  v = beta2 v + (1 - beta2) g*g;
 * p -= lr / (1 - beta1^t) * m / (sqrt(v) / sqrt(1 - beta2^t) + eps)   to every element of every tensor in one
 * launch.  Complex parameters ride as their (re, im) fp32 view, as in torch.  Graph-capturable. */
int rpde_adamw_step(const void* table, int ntensors, long long nchunks, float* hyper, double beta1, double beta2,
                    float eps, float weight_decay, rpde_stream_t stream);

/* ---- fused F-FNO FeedForward layer, BF16 tensor cores (the fast variant; stated tolerance in DESIGN.md) --------------
 * y = res + LayerNorm(W3 gelu(W2 gelu(W1 x + b1) + b2) + b3) * gamma + beta for C = 64 channels and hidden width 256:
 * FeedForward with n_layers = 3, factor 4 (models/custom_layer.py:49-68) as FSpectralConv1d/2d call it
 * (models/spectral_convolution.py:248-254) plus the layer's residual (models/ffno.py:116-118, 228-235), in ONE kernel.
 * The hidden activations stay in TMEM (accumulator of one layer = BF16 A operand of the next); HBM traffic is
 * read x (+ res) and write y.  rpde_ff3_supported(C, hidden) tells whether a shape is covered (64, 256).
 * image: rpde_ff3_weight_image_bytes() bytes, the three weights (nn.Linear layout (out, in), fp32) as the BF16
 * shared-memory image the kernels load; rebuild it with rpde_ff3_pack_weights whenever a weight changes.
 * res, gamma/beta may be NULL (no residual / no LayerNorm).  All pointers 16-byte aligned, x/res/y are [T][64]. */
size_t rpde_ff3_weight_image_bytes(void);
int rpde_ff3_supported(int channels, int hidden);
int rpde_ff3_pack_weights(const float* w1, const float* w2, const float* w3, void* image, rpde_stream_t stream);
int rpde_ff3_fwd(const float* x, const float* res, const void* image, const float* b1, const float* b2, const float* b3,
                 const float* gamma, const float* beta, float* y, long long T, float eps, rpde_stream_t stream);
/* Backward of the same layer from x and dy alone: the forward pass is recomputed on chip (nothing but x is saved),
 * ds = gradient w.r.t. x; dw*, db*, dgamma, dbeta are overwritten.  The weight gradients come from a second kernel
 * pair over BF16 copies of the per-layer operands held in the workspace (rpde_ff3_bwd_workspace_bytes(T)).  The
 * gradient w.r.t. res is dy itself and is left to the caller. */
size_t rpde_ff3_bwd_workspace_bytes(long long T);
int rpde_ff3_bwd(const float* x, const float* dy, const void* image, const float* b1, const float* b2, const float* b3,
                 const float* gamma, float* ds, float* dw1, float* db1, float* dw2, float* db2, float* dw3, float* db3,
                 float* dgamma, float* dbeta, long long T, float eps, void* ws, size_t ws_bytes, rpde_stream_t stream);

/* ---- raw contraction (test hook): C[M,N] = A[M,K] B[K,N], row-major, through the selected engine -------- */
int rpde_matmul(const float* a, const float* b, float* c, int M, int N, int K, int trans_a, int trans_b,
                rpde_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* RPDE_B200_H */
