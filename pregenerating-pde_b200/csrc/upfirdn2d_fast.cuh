// Fast path of upfirdn2d for separable filters with one resampling direction (up OR down, or neither):
//   (up, down) in {(1,1), (2,1), (4,1), (1,2), (1,4)}, the same on both axes, fw == fh == NT taps.
// These are the shapes upsample2d / downsample2d / filter2d and the generic filtered_lrelu composition
// produce (reference upfirdn2d.py:269-379, filtered_lrelu.py:215-221); everything else runs the generic
// kernel in upfirdn2d.cu.
//
// One CTA = one output tile of one plane, two passes through shared memory (one HBM read, one write):
//   stage : input footprint -> s_in (fp32, zero filled outside the plane = the padding)
//   h-pass: a thread owns ONE footprint row and a strip of R consecutive outputs; the WIN input samples of
//           the strip sit in registers, all polyphase index arithmetic is resolved at compile time because
//           strips start on a whole input sample ((o*DOWN - pad0) % UP == 0); taps live in registers
//   v-pass: a thread owns ONE column and a strip of R output rows (register window down the column), and
//           stores straight to y: a warp writes 32 consecutive elements of a row
// Row pitches are odd, so "consecutive threads = consecutive rows" (h-pass) and "consecutive threads =
// consecutive columns" (v-pass) are both bank-conflict free.
#pragma once
#include "common.cuh"

namespace upfast {

struct Params {
    const void* x; void* y; const float* f;      // f: NT horizontal taps then NT vertical taps (device memory)
    int in_h, in_w, out_h, out_w, padx0, pady0, flip;
    float gain;
    int tow, toh, tiles_x, tiles_y;
    int pin, ph;                                 // odd row pitches of s_in / s_h
    int ith;                                     // rows of the (maximal) input footprint
};

__host__ __device__ constexpr int strip_len(int up) { return up == 1 ? 4 : (up == 2 ? 8 : 16); }

template <int UP, int DOWN, int NT>
struct Fir {
    static constexpr int R = strip_len(UP);                               // outputs per strip
    static constexpr int WIN = ((R - 1) * DOWN + NT - 1) / UP + 1;        // input samples a strip touches
    static constexpr int STEP = R * DOWN / UP;                            // input samples between strips
    static_assert((R * DOWN) % UP == 0, "strips must start on a whole input sample");
    static_assert(NT % UP == 0, "every phase must have the same tap count");
    // out[j] = sum_k g[t0 + k*UP] * w[s0 + k],  t0 = first tap that meets a real sample
    static __device__ __forceinline__ void run(const float (&w)[WIN], const float (&g)[NT], float (&out)[R]) {
#pragma unroll
        for (int j = 0; j < R; j++) {
            const int t0 = (UP - (j * DOWN) % UP) % UP;
            const int s0 = (j * DOWN + t0) / UP;
            float a = 0.f;
#pragma unroll
            for (int k = 0; k < NT / UP; k++) a = fmaf(w[s0 + k], g[t0 + k * UP], a);
            out[j] = a;
        }
    }
};

__device__ __forceinline__ int floor_div(int a, int b) { int q = a / b; return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q; }
__device__ __forceinline__ int pos_mod(int a, int b) { int r = a % b; return r < 0 ? r + b : r; }

template <class T, int UP, int DOWN, int NT>
__global__ void __launch_bounds__(256) upfirdn2d_fast_kernel(const Params p) {
    using F = Fir<UP, DOWN, NT>;
    constexpr int R = F::R, WIN = F::WIN, STEP = F::STEP;
    extern __shared__ float smem[];
    __shared__ float s_f[2 * NT];
    float* s_in = smem;
    float* s_h = smem + p.ith * p.pin;
    const int tid = threadIdx.x, nthr = blockDim.x;

    int b = blockIdx.x;
    const int tx = b % p.tiles_x; b /= p.tiles_x;
    const int ty = b % p.tiles_y; const int plane = b / p.tiles_y;
    const int ox0 = tx * p.tow, oy0 = ty * p.toh;
    const int tw = min(p.tow, p.out_w - ox0), th = min(p.toh, p.out_h - oy0);
    // first strip starts ex (ey) outputs before the tile so that it starts on a whole input sample
    const int ex = (UP > 1) ? pos_mod(ox0 * DOWN - p.padx0, UP) : 0;
    const int ey = (UP > 1) ? pos_mod(oy0 * DOWN - p.pady0, UP) : 0;
    const int obx = ox0 - ex, oby = oy0 - ey;
    const int ibx = floor_div(obx * DOWN - p.padx0, UP), iby = floor_div(oby * DOWN - p.pady0, UP);
    const int nsx = (tw + ex + R - 1) / R, nsy = (th + ey + R - 1) / R;
    const int iw = (nsx - 1) * STEP + WIN, ih = (nsy - 1) * STEP + WIN;

    // taps in correlation order: g[t] = flip ? f[t] : f[NT-1-t]  (upfirdn2d.py:190-191)
    if (tid < 2 * NT) { const int a = tid < NT ? 0 : NT, t = tid - a; s_f[tid] = p.f[a + (p.flip ? t : NT - 1 - t)]; }

    const T* xp = (const T*)p.x + (size_t)plane * p.in_h * p.in_w;
    for (int r = tid >> 5; r < ih; r += nthr >> 5) {
        const int sy = iby + r;
        const bool rok = sy >= 0 && sy < p.in_h;
        const T* xr = xp + (size_t)(rok ? sy : 0) * p.in_w;
        for (int c = tid & 31; c < iw; c += 32) {
            const int sx = ibx + c;
            s_in[r * p.pin + c] = (rok && sx >= 0 && sx < p.in_w) ? to_f32(xr[sx]) : 0.f;
        }
    }
    __syncthreads();

    float g[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) g[t] = s_f[t];
    for (int task = tid; task < ih * nsx; task += nthr) {
        const int strip = task / ih, row = task - strip * ih;
        const float* src = s_in + row * p.pin + strip * STEP;
        float w[WIN], out[R];
#pragma unroll
        for (int i = 0; i < WIN; i++) w[i] = src[i];
        F::run(w, g, out);
        float* dst = s_h + row * p.ph + strip * R;
#pragma unroll
        for (int j = 0; j < R; j++) dst[j] = out[j];
    }
    __syncthreads();

#pragma unroll
    for (int t = 0; t < NT; t++) g[t] = s_f[NT + t] * p.gain;
    T* yp = (T*)p.y + (size_t)plane * p.out_h * p.out_w;
    const int ncols = nsx * R;
    for (int task = tid; task < nsy * ncols; task += nthr) {
        const int strip = task / ncols, col = task - strip * ncols;
        const float* src = s_h + strip * STEP * p.ph + col;
        float w[WIN], out[R];
#pragma unroll
        for (int i = 0; i < WIN; i++) w[i] = src[i * p.ph];
        F::run(w, g, out);
        const int ox = obx + col;
        if (ox < ox0 || ox >= ox0 + tw) continue;
        const int oyb = oby + strip * R;
#pragma unroll
        for (int j = 0; j < R; j++) {
            const int oy = oyb + j;
            if (oy >= oy0 && oy < oy0 + th) yp[(size_t)oy * p.out_w + ox] = from_f32<T>(out[j]);
        }
    }
}

// host: tile choice + launch.  Returns -1 when (up, down, taps) has no instantiation.
template <class T, int UP, int DOWN, int NT>
int launch(Params p, int64_t planes, cudaStream_t st) {
    using F = Fir<UP, DOWN, NT>;
    constexpr int R = F::R;
    auto roundup = [](int a, int b) { return (a + b - 1) / b * b; };
    p.tow = p.out_w < 128 ? roundup(p.out_w, R) : 128;
    p.toh = p.out_h < 64 ? roundup(p.out_h, R) : 64;
    if (DOWN > 1) { p.tow = p.tow > 64 ? 64 : p.tow; p.toh = p.toh > 32 ? 32 : p.toh; }
    size_t smem = 0; int tasks = 0;
    for (;;) {
        const int nsx = (p.tow + (UP - 1) + R - 1) / R, nsy = (p.toh + (UP - 1) + R - 1) / R;
        const int itw = (nsx - 1) * F::STEP + F::WIN;
        p.ith = (nsy - 1) * F::STEP + F::WIN;
        p.pin = itw | 1; p.ph = (nsx * R) | 1;
        smem = sizeof(float) * (size_t)p.ith * (p.pin + p.ph);
        tasks = nsy * nsx * R;
        if (smem <= 72 * 1024 || p.toh <= R) break;
        p.toh = roundup(p.toh / 2, R);
    }
    if (smem > 200 * 1024) return -1;
    p.tiles_x = (p.out_w + p.tow - 1) / p.tow;
    p.tiles_y = (p.out_h + p.toh - 1) / p.toh;
    const int64_t blocks = planes * p.tiles_x * p.tiles_y;
    if (blocks >= (1ll << 31)) { cno_set_error("upfirdn2d: grid too large"); return (int)cudaErrorInvalidValue; }
    int threads = (tasks + 31) / 32 * 32;
    threads = threads < 64 ? 64 : threads > 256 ? 256 : threads;
    auto kern = upfirdn2d_fast_kernel<T, UP, DOWN, NT>;
    static size_t configured = 48 * 1024;
    if (smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { cno_set_error("upfirdn2d: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e)); return (int)e; }
        configured = smem;
    }
    kern<<<(unsigned)blocks, threads, smem, st>>>(p);
    CNO_LAUNCH_CHECK();
    return 0;
}

template <class T, int UP, int DOWN>
int launch_taps(const Params& p, int64_t planes, int nt, cudaStream_t st) {
    switch (nt) {
        case 4: return launch<T, UP, DOWN, 4>(p, planes, st);
        case 8: return launch<T, UP, DOWN, 8>(p, planes, st);
        case 12: return launch<T, UP, DOWN, 12>(p, planes, st);
        case 16: return launch<T, UP, DOWN, 16>(p, planes, st);
        case 24: return launch<T, UP, DOWN, 24>(p, planes, st);
        default: return -1;
    }
}

template <class T>
int launch_cfg(const Params& p, int64_t planes, int up, int down, int nt, cudaStream_t st) {
    if (up == 1 && down == 1) return launch_taps<T, 1, 1>(p, planes, nt, st);
    if (up == 2 && down == 1) return launch_taps<T, 2, 1>(p, planes, nt, st);
    if (up == 4 && down == 1) return launch_taps<T, 4, 1>(p, planes, nt, st);
    if (up == 1 && down == 2) return launch_taps<T, 1, 2>(p, planes, nt, st);
    if (up == 1 && down == 4) return launch_taps<T, 1, 4>(p, planes, nt, st);
    return -1;
}

}  // namespace upfast
