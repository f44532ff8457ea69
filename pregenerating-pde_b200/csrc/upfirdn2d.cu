// upfirdn2d: zero-insert upsample -> pad/crop -> FIR -> decimate for a batch of planes.
// Replaces the reference plugin upfirdn2d.cu:21-367 / upfirdn2d.cpp:8-90 (new kernel).  One CTA
// computes an output tile of one plane: the input footprint is staged in shared memory once, a
// separable filter runs horizontal-then-vertical inside the same CTA (one HBM read + one write; the
// reference launches two kernels with a round trip through HBM, upfirdn2d.py:235-237), a 2-D filter
// runs as a direct polyphase gather.  Only taps that meet a real sample are visited.
#include <stdlib.h>
#include "common.cuh"
#include "upfirdn2d_fast.cuh"

namespace {

struct UpParams {
    const void* x; void* y; const float* f;
    int64_t planes; int in_h, in_w, out_h, out_w, fw, fh, upx, upy, downx, downy, padx0, pady0, flip, separable;
    float gain;
    int tow, toh;        // output tile
    int itw, ith;        // input footprint (max) of a tile
    int tiles_x, tiles_y;
};

__device__ __forceinline__ int floor_div(int a, int b) { int q = a / b; return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q; }
__device__ __forceinline__ int pos_mod(int a, int b) { int r = a % b; return r < 0 ? r + b : r; }
__device__ __forceinline__ int ceil_div_s(int a, int b) { return -floor_div(-a, b); }

template <class T>
__global__ void __launch_bounds__(256) upfirdn2d_kernel(UpParams p) {
    extern __shared__ float smem[];
    const int ntaps = p.separable ? p.fw + p.fh : p.fw * p.fh;
    float* s_f = smem;                          // taps, already in correlation order
    float* s_in = s_f + ((ntaps + 3) & ~3);     // [ith][itw]
    float* s_h = s_in + p.ith * p.itw;          // separable: [ith][tow]
    const int tid = threadIdx.x;

    int64_t b = blockIdx.x;
    const int tx = (int)(b % p.tiles_x); b /= p.tiles_x;
    const int ty = (int)(b % p.tiles_y); const int64_t plane = b / p.tiles_y;
    const int ox0 = tx * p.tow, oy0 = ty * p.toh;
    const int tw = min(p.tow, p.out_w - ox0), th = min(p.toh, p.out_h - oy0);

    // taps in correlation order: g[t] = flip ? f[t] : f[T-1-t]  (upfirdn2d.py:190-191)
    if (p.separable) {
        for (int i = tid; i < p.fw; i += blockDim.x) s_f[i] = p.f[p.flip ? i : p.fw - 1 - i];
        for (int i = tid; i < p.fh; i += blockDim.x) s_f[p.fw + i] = p.f[p.fw + (p.flip ? i : p.fh - 1 - i)];
    } else {
        for (int i = tid; i < ntaps; i += blockDim.x) {
            const int r = i / p.fw, c = i % p.fw;
            s_f[i] = p.flip ? p.f[i] : p.f[(p.fh - 1 - r) * p.fw + (p.fw - 1 - c)];
        }
    }
    // input footprint of this tile: sample i sits at position i*up + pad0 of the padded signal
    const int ix0 = ceil_div_s(ox0 * p.downx - p.padx0, p.upx);
    const int iy0 = ceil_div_s(oy0 * p.downy - p.pady0, p.upy);
    const int ix1 = floor_div((ox0 + tw - 1) * p.downx + p.fw - 1 - p.padx0, p.upx);
    const int iy1 = floor_div((oy0 + th - 1) * p.downy + p.fh - 1 - p.pady0, p.upy);
    const int iw = ix1 - ix0 + 1, ih = iy1 - iy0 + 1;       // <= itw, ith by construction
    const T* xp = (const T*)p.x + plane * (int64_t)p.in_h * p.in_w;
    for (int i = tid; i < ih * iw; i += blockDim.x) {
        const int r = i / iw, c = i % iw;
        const int sy = iy0 + r, sx = ix0 + c;
        float v = 0.f;
        if (sy >= 0 && sy < p.in_h && sx >= 0 && sx < p.in_w) v = to_f32(xp[(int64_t)sy * p.in_w + sx]);
        s_in[r * p.itw + c] = v;
    }
    __syncthreads();
    T* yp = (T*)p.y + plane * (int64_t)p.out_h * p.out_w;
    if (p.separable) {
        // horizontal pass on every footprint row
        for (int i = tid; i < ih * tw; i += blockDim.x) {
            const int r = i / tw, ox = i % tw;
            const int px = (ox0 + ox) * p.downx;
            const int t0 = pos_mod(p.padx0 - px, p.upx);
            int src = (px + t0 - p.padx0) / p.upx - ix0;
            float acc = 0.f;
            for (int t = t0; t < p.fw; t += p.upx, src++) acc = fmaf(s_in[r * p.itw + src], s_f[t], acc);
            s_h[r * p.tow + ox] = acc;
        }
        __syncthreads();
        for (int i = tid; i < th * tw; i += blockDim.x) {
            const int oy = i / tw, ox = i % tw;
            const int py = (oy0 + oy) * p.downy;
            const int t0 = pos_mod(p.pady0 - py, p.upy);
            int src = (py + t0 - p.pady0) / p.upy - iy0;
            float acc = 0.f;
            for (int t = t0; t < p.fh; t += p.upy, src++) acc = fmaf(s_h[src * p.tow + ox], s_f[p.fw + t], acc);
            yp[(int64_t)(oy0 + oy) * p.out_w + ox0 + ox] = from_f32<T>(acc * p.gain);
        }
    } else {
        for (int i = tid; i < th * tw; i += blockDim.x) {
            const int oy = i / tw, ox = i % tw;
            const int px = (ox0 + ox) * p.downx, py = (oy0 + oy) * p.downy;
            const int tx0 = pos_mod(p.padx0 - px, p.upx), ty0 = pos_mod(p.pady0 - py, p.upy);
            const int sx0 = (px + tx0 - p.padx0) / p.upx - ix0;
            int sy = (py + ty0 - p.pady0) / p.upy - iy0;
            float acc = 0.f;
            for (int t_y = ty0; t_y < p.fh; t_y += p.upy, sy++) {
                int sx = sx0;
                for (int t_x = tx0; t_x < p.fw; t_x += p.upx, sx++) acc = fmaf(s_in[sy * p.itw + sx], s_f[t_y * p.fw + t_x], acc);
            }
            yp[(int64_t)(oy0 + oy) * p.out_w + ox0 + ox] = from_f32<T>(acc * p.gain);
        }
    }
}

}  // namespace

extern "C" int cno_upfirdn2d(const void* x, void* y, const float* f, int dtype, int separable,
                             int64_t planes, int in_h, int in_w, int out_h, int out_w, int fw, int fh,
                             int upx, int upy, int downx, int downy, int padx0, int pady0,
                             int flip, float gain, void* stream) {
    CNO_CHECK_ARG(x && y && f, "upfirdn2d: null pointer");
    CNO_CHECK_ARG(upx >= 1 && upy >= 1 && downx >= 1 && downy >= 1, "upfirdn2d: up/down factors must be >= 1");
    CNO_CHECK_ARG(fw >= 1 && fh >= 1 && fw <= 64 && fh <= 64, "upfirdn2d: filter size %dx%d not in 1..64", fw, fh);
    CNO_CHECK_ARG(in_h >= 1 && in_w >= 1, "upfirdn2d: empty input plane");
    if (planes <= 0 || out_h <= 0 || out_w <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    static EnvFlag env_generic;
    if (separable && upx == upy && downx == downy && fw == fh && !env_generic.on("CNO_UPFIRDN_GENERIC")) {
        // fast path: compile-time polyphase arithmetic, register windows (upfirdn2d_fast.cuh)
        upfast::Params q{};
        q.x = x; q.y = y; q.f = f; q.in_h = in_h; q.in_w = in_w; q.out_h = out_h; q.out_w = out_w;
        q.padx0 = padx0; q.pady0 = pady0; q.flip = flip; q.gain = gain;
        int rc = -1;
        switch (dtype) {
            case CNO_F32: rc = upfast::launch_cfg<float>(q, planes, upx, downx, fw, st); break;
            case CNO_BF16: rc = upfast::launch_cfg<__nv_bfloat16>(q, planes, upx, downx, fw, st); break;
            case CNO_F16: rc = upfast::launch_cfg<__half>(q, planes, upx, downx, fw, st); break;
            default: cno_set_error("upfirdn2d: unsupported dtype %d", dtype); return (int)cudaErrorInvalidValue;
        }
        if (rc >= 0) return rc;
    }
    UpParams p{};
    p.x = x; p.y = y; p.f = f; p.planes = planes; p.in_h = in_h; p.in_w = in_w; p.out_h = out_h; p.out_w = out_w;
    p.fw = fw; p.fh = fh; p.upx = upx; p.upy = upy; p.downx = downx; p.downy = downy; p.padx0 = padx0; p.pady0 = pady0;
    p.flip = flip; p.separable = separable; p.gain = gain;
    // tile: as wide as the plane up to 64 outputs, as tall as shared memory allows up to 32 rows
    p.tow = out_w < 64 ? out_w : 64;
    p.toh = out_h < 32 ? out_h : 32;
    size_t smem = 0;
    for (;;) {
        p.itw = ((p.tow - 1) * downx + fw - 1) / upx + 2;
        p.ith = ((p.toh - 1) * downy + fh - 1) / upy + 2;
        const int ntaps = separable ? fw + fh : fw * fh;
        smem = sizeof(float) * (size_t)(((ntaps + 3) & ~3) + p.ith * p.itw + (separable ? p.ith * p.tow : 0));
        if (smem <= 96 * 1024 || (p.toh == 1 && p.tow <= 8)) break;
        if (p.toh > 1) p.toh = (p.toh + 1) / 2; else p.tow = (p.tow + 1) / 2;
    }
    CNO_CHECK_ARG(smem <= 200 * 1024, "upfirdn2d: tile does not fit in shared memory");
    p.tiles_x = (out_w + p.tow - 1) / p.tow;
    p.tiles_y = (out_h + p.toh - 1) / p.toh;
    const int64_t blocks = planes * p.tiles_x * p.tiles_y;
    CNO_CHECK_ARG(blocks < (1ll << 31), "upfirdn2d: grid too large");
    cudaError_t e = cudaSuccess;
#define UP_LAUNCH(T) do { \
        if (smem > 48 * 1024) e = cudaFuncSetAttribute(upfirdn2d_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e == cudaSuccess) upfirdn2d_kernel<T><<<(unsigned)blocks, 256, smem, st>>>(p); } while (0)
    switch (dtype) {
        case CNO_F32: UP_LAUNCH(float); break;
        case CNO_BF16: UP_LAUNCH(__nv_bfloat16); break;
        case CNO_F16: UP_LAUNCH(__half); break;
        default: cno_set_error("upfirdn2d: unsupported dtype %d", dtype); return (int)cudaErrorInvalidValue;
    }
#undef UP_LAUNCH
    if (e != cudaSuccess) { cno_set_error("upfirdn2d: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e)); return (int)e; }
    CNO_LAUNCH_CHECK();
    return 0;
}
