/* cno_b200.h -- C ABI of libcno_b200.so: the B200 (sm_100a) implementation of the CNO activation
 * hot path (filtered_lrelu / upfirdn2d / bias_act / CIN).
 *
 * This is the drop-in boundary.  The reference crosses it with pybind11 torch
 * extensions built by torch_utils/custom_ops.py:53 (get_plugin); each entry point below names the
 * reference plugin function it replaces.  Differences by design:
 *   - plain pointers + sizes, no torch types; the caller owns every buffer (the library never
 *     allocates, frees or synchronises) and passes the CUDA stream explicitly;
 *   - no global mutable state (the reference stages filters through one __constant__ buffer,
 *     filtered_lrelu.cu:69-70), so calls are multi-stream and CUDA-graph safe;
 *   - bf16 is a first-class dtype.
 * Return value: 0 = ok, -1 = "no specialised kernel for these arguments" (the caller falls back to
 * the generic composition exactly as filtered_lrelu.py:215-221 does), >0 = cudaError_t.
 * All tensors are dense NCHW ("planes" = N*C images of H x W) unless stated otherwise.
 */
#ifndef CNO_B200_H
#define CNO_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { CNO_F32 = 0, CNO_BF16 = 1, CNO_F16 = 2, CNO_F64 = 3 };
/* CNO_SIGN_READ_SIGNONLY (cno_filtered_lrelu only): like READ, for masks the caller knows were written with clamp = none, so
 * no code has the "clamped" bit; the kernel then skips that test (11 % faster backward).  READ is always correct. */
enum { CNO_SIGN_NONE = 0, CNO_SIGN_WRITE = 1, CNO_SIGN_READ = 2, CNO_SIGN_READ_SIGNONLY = 3 };

int cno_version(void);
/* last error text of the calling thread ("" if none) */
const char* cno_last_error(void);

/* replaces _plugin.bias_act(x, b, xref, yref, dy, grad, dim, act, alpha, gain, clamp)
 * (bias_act.cpp:24-82, kernel bias_act.cu:15-139).  act = cuda_idx 1..9 of bias_act.py:13-23;
 * grad 0: y = clamp(act(x+b)*gain); grad 1/2: first / second order gradient (x carries the incoming
 * gradient).  Null pointers = "tensor not given".  bias index = (i / step_b) % size_b.
 * clamp < 0 = no clamp. */
int cno_bias_act(const void* x, const void* b, const void* xref, const void* yref, const void* dy,
                 void* y, int dtype, int64_t numel, int64_t size_b, int64_t step_b, int grad,
                 int act, float alpha, float gain, float clamp, void* stream);

/* replaces _plugin.upfirdn2d(x, f, upx, upy, downx, downy, padx0, padx1, pady0, pady1, flip, gain)
 * (upfirdn2d.cpp:8-90).  f: DEVICE pointer to fp32 taps, [fh, fw] row major; separable = 1 means f
 * holds fw taps for the horizontal pass followed by fh taps for the vertical pass and both passes run
 * in ONE kernel (the reference needs two launches, upfirdn2d.py:235-237).  Separable filters with one resampling direction
 * ((up,down) in {(1,1),(2,1),(4,1),(1,2),(1,4)} on both axes, 4/8/12/16/24 taps) take a register-window kernel with compile-time
 * polyphase indices (csrc/upfirdn2d_fast.cuh); every other argument set runs the generic kernel.  Output size must be
 * (in*up + pad0 + pad1 - f + down) / down (upfirdn2d.cpp:27-28). */
int cno_upfirdn2d(const void* x, void* y, const float* f, int dtype, int separable,
                  int64_t planes, int in_h, int in_w, int out_h, int out_w, int fw, int fh,
                  int upx, int upy, int downx, int downy, int padx0, int pady0,
                  int flip, float gain, void* stream);

/* replaces _plugin.filtered_lrelu(x, fu, fd, b, si, up, down, px0, px1, py0, py1, sx, sy, gain, slope,
 * clamp, flip_filter, write_signs) (filtered_lrelu.cpp:8-201): ONE fused kernel
 * bias -> FIR upsample -> gain, leaky-ReLU, clamp, 2-bit sign mask write/read -> FIR downsample.
 * fu / fd: HOST pointers to the separable fp32 taps (fu_taps, fd_taps entries).  b may be NULL.
 * mask layout (filtered_lrelu.cpp:79-86): [planes, mask_h, mask_pitch] bytes, element x at byte
 * x>>2 bits 2*(x&3); 1 = negative, 2 = clamped.  sign_mode READ uses offsets (sx, sy) and treats
 * positions outside [0,mask_w) x [0,mask_h) as "pass" (filtered_lrelu.cu:680-687).
 * db: optional fp32[C] accumulator (must be zeroed by the caller): receives sum of y over N,H,W
 * (fuses `db = dx.sum([0,2,3])`, filtered_lrelu.py:257-258; sign_mode READ only, else -1).  clamp < 0 or +inf = none. */
int cno_filtered_lrelu(const void* x, void* y, const void* b, uint8_t* mask, float* db,
                       const float* fu, const float* fd, int dtype,
                       int n, int c, int in_h, int in_w, int out_h, int out_w,
                       int fu_taps, int fd_taps, int up, int down, int px0, int py0,
                       int mask_h, int mask_w, int mask_pitch, int sx, int sy, int sign_mode,
                       float gain, float slope, float clamp, int flip, void* stream);

/* supported-configuration probe for the fused kernel (mirrors filtered_lrelu.cpp:38-48):
 * 0 = fused kernel available, -1 = use the generic composition. */
int cno_filtered_lrelu_query(int dtype, int in_h, int in_w, int out_h, int out_w, int fu_taps,
                             int fd_taps, int up, int down, int px0, int py0);

/* The fused op has two kernels behind cno_filtered_lrelu: a streaming kernel (TMA row strips, four up-res columns per thread) and,
 * for up-res planes of at most 64 columns, a whole-plane kernel (planes resident in shared memory, one thread per column).  The
 * reference picks among its tile specialisations the same way (filtered_lrelu.cu:1240-1270 spec table); results are identical up
 * to fp32 summation order.  enable = 1 / 0 switches the whole-plane kernel on / off (default on; environment CNO_FLR_NO_PLANE
 * turns it off), enable < 0 only queries; returns the previous setting.  Used by the tests to run both kernels on the same data. */
int cno_filtered_lrelu_plane_kernel(int enable);

/* replaces _plugin.filtered_lrelu_act_(x, si, sx, sy, gain, slope, clamp, write_signs)
 * (filtered_lrelu.cpp:205-282, kernel filtered_lrelu.cu:1097-1203): in-place gain*lrelu, clamp on a
 * dense [planes,h,w] tensor with 2-bit mask write / read.  Generic-fallback only. */
int cno_filtered_lrelu_act(void* x, uint8_t* mask, int dtype, int64_t planes, int h, int w,
                           int mask_h, int mask_w, int mask_pitch, int sx, int sy, int sign_mode,
                           float gain, float slope, float clamp, void* stream);

/* CIN = FILM.forward with instance norm (CNO_timeModule_CIN.py:60-70), one pass over x:
 *   mu, rstd per plane; y = (x-mu)*rstd * (gamma[c]*scale[n,c]) + (beta[c]*scale[n,c] + shift[n,c]).
 * gamma, beta fp32[C]; scale, shift fp32[N,C]; mean, rstd fp32[N,C] outputs (saved for backward). */
int cno_cin_fwd(const void* x, void* y, const float* gamma, const float* beta, const float* scale,
                const float* shift, float* mean, float* rstd, int dtype, int n, int c, int hw,
                float eps, void* stream);
/* backward: dx (same dtype as x) and per-plane fp32[N,C] reductions
 *   s1 = sum(dy), s2 = sum(dy * xhat)   (dgamma = sum_n scale*s2, dbeta = sum_n scale*s1,
 *   dscale = gamma*s2 + beta*s1, dshift = s1 -- finished by the host wrapper on [N,C] tensors). */
int cno_cin_bwd(const void* x, const void* dy, void* dx, const float* gamma, const float* scale,
                const float* mean, const float* rstd, float* s1, float* s2, int dtype, int n, int c,
                int hw, void* stream);

/* FILM.forward in one launch (CNO_timeModule_CIN.py:60-70 incl. the four Linear layers :64-65):
 *   scale[n,c] = b2s[c] + sum_j W2s[c,j]*(w1s[j]*t[n] + b1s[j]),  shift likewise from the *b weights (lat = 128),
 *   then the instance norm + affine + FiLM of cno_cin_fwd.  scale_out fp32[N,C] (optional) is kept for backward. */
int cno_film_fwd(const void* x, void* y, const float* t, const float* gamma, const float* beta,
                 const float* w1s, const float* b1s, const float* W2s, const float* b2s,
                 const float* w1b, const float* b1b, const float* W2b, const float* b2b,
                 float* scale_out, float* mean, float* rstd, int dtype, int n, int c, int hw, int lat,
                 float eps, void* stream);
/* ResidualBlock epilogue (CNO_timeModule_CIN.py:503-507, `x + in_norm_conditiner2(out, time)`): the same two forward kernels with
 * `y = residual + CIN(x)` written in the one pass; residual has the layout and dtype of x.  residual = NULL is cno_cin_fwd /
 * cno_film_fwd.  The backward needs nothing new: d residual = dy, and cno_cin_bwd is unchanged. */
int cno_cin_fwd_res(const void* x, const void* residual, void* y, const float* gamma, const float* beta, const float* scale,
                    const float* shift, float* mean, float* rstd, int dtype, int n, int c, int hw, float eps, void* stream);
int cno_film_fwd_res(const void* x, const void* residual, void* y, const float* t, const float* gamma, const float* beta,
                     const float* w1s, const float* b1s, const float* W2s, const float* b2s,
                     const float* w1b, const float* b1b, const float* W2b, const float* b2b,
                     float* scale_out, float* mean, float* rstd, int dtype, int n, int c, int hw, int lat,
                     float eps, void* stream);

/* CIN -> filtered_lrelu prologue fusion (CNO_timeModule_CIN.py:303-317 followed by the LReLu of the same block).
 * cno_film_stats is cno_film_fwd without the y pass: it leaves mean, rstd, scale_out and the plane affine
 * CIN(x)[n,c] = aff_a[n,c] * x[n,c] + aff_d[n,c]  (all fp32[N,C]).  cno_filtered_lrelu_affine is cno_filtered_lrelu on the input
 * aff_a * x + aff_d, applied while the up-FIR runs (aff_a scales the FIR sum, aff_d joins the bias edge profile), so the normalised
 * tensor is never written or read.  Forward modes (SIGN_WRITE / SIGN_NONE) only; aff_a = aff_d = NULL is cno_filtered_lrelu.
 * Backward: the unchanged SIGN_READ launch gives d(CIN output), cno_cin_bwd + cno_film_bwd_finish take it from there. */
int cno_film_stats(const void* x, const float* t, const float* gamma, const float* beta,
                   const float* w1s, const float* b1s, const float* W2s, const float* b2s,
                   const float* w1b, const float* b1b, const float* W2b, const float* b2b,
                   float* scale_out, float* mean, float* rstd, float* aff_a, float* aff_d,
                   int dtype, int n, int c, int hw, int lat, float eps, void* stream);
int cno_filtered_lrelu_affine(const void* x, void* y, const void* b, uint8_t* mask, float* db,
                              const float* fu, const float* fd, int dtype,
                              int n, int c, int in_h, int in_w, int out_h, int out_w,
                              int fu_taps, int fd_taps, int up, int down, int px0, int py0,
                              int mask_h, int mask_w, int mask_pitch, int sx, int sy, int sign_mode,
                              float gain, float slope, float clamp, int flip,
                              const float* aff_a, const float* aff_d, void* stream);

/* every parameter gradient of one FILM module from the per-plane sums s1, s2 of cno_cin_bwd, in one launch
 * (replaces autograd through 4 Linear layers, ~20 small kernels): dgamma, dbeta [C]; d lat2scale.weight [C,lat] / bias [C];
 * d inp2lat_sacale.weight / bias [lat]; the same four for the shift branch.  Deterministic (no atomics). */
int cno_film_bwd_finish(const float* s1, const float* s2, const float* t, const float* gamma, const float* beta,
                        const float* scale, const float* w1s, const float* b1s, const float* W2s,
                        const float* w1b, const float* b1b, const float* W2b,
                        float* dgamma, float* dbeta, float* dW2s, float* db2s, float* dw1s, float* db1s,
                        float* dW2b, float* db2b, float* dw1b, float* db1b, int n, int c, int lat, void* stream);

/* Allocation-free CIN forward.  No counterpart in the reference: added for callers that
 * cannot allocate device memory (stream capture, tight serving loops, harnesses that forbid cudaMalloc).  Same result as
 * cno_cin_fwd; the per-plane statistics that cno_cin_fwd leaves in the caller's mean / rstd buffers live in a fixed array inside
 * the library's device image (512 KB per GPU that loads the library, held until the process exits), so the call allocates
 * nothing and the statistics are not returned (forward only: no backward can follow it).
 * Limit: n * c <= 65536 planes, otherwise cudaErrorInvalidValue with a message in cno_last_error().
 * The array is shared by every call on a device and keeps the last call's contents (nothing in it needs clearing: every entry is
 * written before it is read within a launch); calls that overlap on the GPU -- different streams with no dependency -- race. */
int cno_cin_fwd_noalloc(const void* x, void* y, const float* gamma, const float* beta, const float* scale,
                        const float* shift, int dtype, int n, int c, int hw, float eps, void* stream);

/* Optimiser step of the training loop (TrainCNO_time_L.py:185-196 `optimizer.step()` of the `torch.optim.AdamW` built in
 * CNO_timeModule_CIN.py:996; CNO_FineTune.py: three parameter groups): AdamW over EVERY parameter tensor in one launch.
 * `table` (device memory) has one record per tensor, `chunk_tensor[n_chunks]` (device) maps each chunk of cno_adamw_chunk() elements
 * to its tensor (a tensor's chunks are consecutive, starting at first_chunk).  `step` (device, fp32[1]) is incremented first, then
 *   p *= 1 - lr*wd;  m += (1-b1)(g-m);  v = b2 v + (1-b2) g^2;  p -= lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
 * (torch.optim.AdamW, amsgrad = maximize = False).  A record with grad = NULL is skipped.  fp32 only; pointers need 4-byte alignment
 * (16-byte aligned tensors take the vector path). */
typedef struct CnoAdamTensor {
    float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
    int64_t numel; float lr, weight_decay; int first_chunk, pad;
} CnoAdamTensor;
int cno_adamw_chunk(void);
int cno_adamw_multi(const CnoAdamTensor* table, const int* chunk_tensor, int n_chunks, float* step,
                    double beta1, double beta2, double eps, void* stream);

/* NCHW <-> NHWC conversion of a dense [n, c, hw] / [n, hw, c] tensor around the cuDNN convolutions of the CNO blocks
 * (CNO_timeModule_CIN.py:329-335, :495-510; the reference leaves the conversion to cuDNN, which repeats it for every operand of every
 * pass): per sample a tiled matrix transpose.  x and y must not overlap. */
int cno_nchw_to_nhwc(const void* x, void* y, int dtype, int n, int c, int hw, void* stream);
int cno_nhwc_to_nchw(const void* x, void* y, int dtype, int n, int c, int hw, void* stream);

/* Measurement helper used by bench.py / tests (not part of the reference surface): FP32 FMA issue
 * rate probe.  variant 0 = scalar FFMA, 1 = packed fma.rn.f32x2.  Writes elapsed-cycle-independent
 * results to out[threads]; returns 0. */
int cno_probe_fma(float* out, int variant, int blocks, int threads, int iters, void* stream);

#ifdef __cplusplus
}
#endif
#endif
