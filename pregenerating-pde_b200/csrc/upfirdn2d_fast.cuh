// Fast path of upfirdn2d for separable filters with one resampling direction (up OR down, or neither):
//   (up, down) in {(1,1), (2,1), (4,1), (1,2), (1,4)}, the same on both axes, fw == fh == NT taps.
// These are the shapes upsample2d / downsample2d / filter2d and the generic filtered_lrelu composition
// produce (reference upfirdn2d.py:269-379, filtered_lrelu.py:215-221); everything else runs the generic
// kernel in upfirdn2d.cu.
//
// One CTA = one output tile of one plane, two passes through shared memory (one HBM read, one write).
// The whole tile geometry is compile time (tile size, footprint, row pitches), so the passes carry
// no index arithmetic beyond one constant division per task:
//   stage : input footprint -> s_in (fp32, zero filled outside the plane = the padding); a warp reads 32
//           consecutive columns of a row, four rows (4*K loads) in flight per thread
//   h-pass: a thread owns ONE footprint row and a strip of R consecutive outputs; the WIN input samples of
//           the strip are read with 16-byte LDS into registers, polyphase indices are compile time because
//           the tile grid is shifted so that every strip starts on a whole input sample
//           ((o*DOWN - pad0) % UP == 0); taps live in registers
//   v-pass: a thread owns ONE column and a strip of R output rows (register window down the column), and
//           stores straight to y: a warp writes 32 consecutive elements of a row
// Row pitches are 4 * odd: consecutive threads = consecutive rows makes the 16-byte accesses of the h-pass
// conflict free (8 threads of a quarter warp hit 8 different 16-byte bank groups), consecutive threads =
// consecutive columns makes the scalar accesses of the v-pass conflict free.
#pragma once
#include "common.cuh"

namespace upfast {

struct Params {
    const void* x; void* y; const float* f;      // f: NT horizontal taps then NT vertical taps (device memory)
    int in_h, in_w, out_h, out_w, padx0, pady0, flip;
    float gain;
    int tiles_x, tiles_y;
    int offx, offy;                              // tile grid shift: tile (tx, ty) starts at output (tx*TOW - offx, ty*TOH - offy)
};

__host__ __device__ constexpr int strip_len(int up) { return up == 1 ? 4 : (up == 2 ? 8 : 16); }
__host__ __device__ constexpr int pitch_for(int n) { int p = (n + 3) / 4; return 4 * (p | 1); }   // >= n, multiple of 4, p/4 odd

template <int UP, int DOWN, int NT>
struct Fir {
    static constexpr int R = strip_len(UP);                               // outputs per strip
    static constexpr int WIN = ((R - 1) * DOWN + NT - 1) / UP + 1;        // input samples a strip touches
    static constexpr int WIN4 = (WIN + 3) / 4;                            // ... in 16-byte loads
    static constexpr int STEP = R * DOWN / UP;                            // input samples between strips
    static_assert((R * DOWN) % UP == 0 && STEP % 4 == 0, "strips must start on a whole, 16-byte aligned input sample");
    static_assert(NT % UP == 0, "every phase must have the same tap count");
    // out[j] = sum_k g[t0 + k*UP] * w[s0 + k],  t0 = first tap that meets a real sample
    static __device__ __forceinline__ void run(const float (&w)[4 * WIN4], const float (&g)[NT], float (&out)[R]) {
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

// compile-time tile geometry.  SMALL = planes of at most ~64 outputs across
template <int UP, int DOWN, int NT, bool SMALL>
struct Geo {
    using F = Fir<UP, DOWN, NT>;
    static constexpr int TOW = SMALL ? (DOWN == 1 ? 32 : 16) : (DOWN == 1 ? 128 : DOWN == 2 ? 64 : 32);
    static constexpr int TOH = SMALL ? (DOWN == 1 ? 32 : 16) : (DOWN == 1 ? 64 : DOWN == 2 ? 32 : 16);
    static constexpr int NSX = TOW / F::R, NSY = TOH / F::R;
    static constexpr int ITW = (NSX - 1) * F::STEP + F::WIN, ITH = (NSY - 1) * F::STEP + F::WIN;
    static constexpr int PIN = pitch_for((NSX - 1) * F::STEP + 4 * F::WIN4);
    static constexpr int PH = pitch_for(TOW);
    static constexpr int THREADS = (TOW * NSY >= 256 && ITH * NSX >= 256) ? 256 : 128;
    static constexpr int SPAR = THREADS / TOW;                            // v-pass strips processed in parallel
    static constexpr int K = (ITW + 31) / 32;                             // 32-column chunks of a footprint row
    static constexpr size_t SMEM = sizeof(float) * (size_t)ITH * (PIN + PH);
    static_assert(TOW % F::R == 0 && TOH % F::R == 0 && THREADS % TOW == 0, "tile / strip / thread shapes");
};

// two packed 16-bit elements -> fp32 (low half = lower address)
template <class T> __device__ __forceinline__ float2 unpack2(uint32_t w);
template <> __device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t w) { return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u)); }
template <> __device__ __forceinline__ float2 unpack2<__half>(uint32_t w) { return __half22float2(*reinterpret_cast<const __half2*>(&w)); }
template <> __device__ __forceinline__ float2 unpack2<float>(uint32_t) { return make_float2(0.f, 0.f); }   // never used

template <class T, int UP, int DOWN, int NT, bool SMALL>
__global__ void __launch_bounds__(Geo<UP, DOWN, NT, SMALL>::THREADS)
upfirdn2d_fast_kernel(const Params p) {
    using F = Fir<UP, DOWN, NT>;
    using G = Geo<UP, DOWN, NT, SMALL>;
    constexpr int R = F::R, WIN4 = F::WIN4, STEP = F::STEP;
    constexpr int PIN = G::PIN, PH = G::PH, ITH = G::ITH, NW = G::THREADS / 32;
    extern __shared__ __align__(16) float smem[];
    __shared__ float s_f[2 * NT];
    float* s_in = smem;                 // [ITH][PIN]
    float* s_h = smem + ITH * PIN;      // [ITH][PH]
    const int tid = threadIdx.x;

    unsigned b = blockIdx.x;
    const int tx = b % (unsigned)p.tiles_x; b /= (unsigned)p.tiles_x;
    const int ty = b % (unsigned)p.tiles_y; const int plane = b / (unsigned)p.tiles_y;
    const int ox0 = tx * G::TOW - p.offx, oy0 = ty * G::TOH - p.offy;
    const int ibx = (ox0 * DOWN - p.padx0) / UP, iby = (oy0 * DOWN - p.pady0) / UP;      // exact by construction of offx / offy

    // taps in correlation order: g[t] = flip ? f[t] : f[NT-1-t]  (upfirdn2d.py:190-191)
    if (tid < 2 * NT) { const int a = tid < NT ? 0 : NT, t = tid - a; s_f[tid] = p.f[a + (p.flip ? t : NT - 1 - t)]; }

    // ---- stage: footprint -> s_in ------------------------------------------------------------------
    const T* xp = (const T*)p.x + (size_t)plane * p.in_h * p.in_w;
    const int warp = tid >> 5, lane = tid & 31;
    if (sizeof(T) == 2 && (p.in_w & 1) == 0 && ((uintptr_t)p.x & 3u) == 0) {
        // 16-bit elements: 4-byte loads of element pairs (even source column first), a warp covers 64 columns per load;
        // an even plane width keeps every pair inside one row and either both or none of its elements inside the plane
        constexpr int K2 = (G::ITW + 1 + 63) / 64, RB = 4;
        const int e = ibx & 1;                                   // the pair of lane 0 starts one column left of the footprint when ibx is odd
        unsigned cmask = 0;
#pragma unroll
        for (int k = 0; k < K2; k++) { const int sx = ibx - e + 2 * lane + 64 * k; if (sx >= 0 && sx < p.in_w) cmask |= 1u << k; }
        for (int r = warp; r < ITH; r += RB * NW) {
            uint32_t v[RB][K2];
#pragma unroll
            for (int u = 0; u < RB; u++) {
                const int sy = iby + r + u * NW;
                const bool rok = sy >= 0 && sy < p.in_h && r + u * NW < ITH;
                const T* xr = xp + (size_t)(rok ? sy : 0) * p.in_w + (ibx - e + 2 * lane);
#pragma unroll
                for (int k = 0; k < K2; k++)
                    v[u][k] = (rok && ((cmask >> k) & 1u)) ? *reinterpret_cast<const uint32_t*>(xr + 64 * k) : 0u;
            }
#pragma unroll
            for (int u = 0; u < RB; u++) {
                if (r + u * NW < ITH) {
#pragma unroll
                    for (int k = 0; k < K2; k++) {
                        const int c = 2 * lane + 64 * k - e;
                        const float2 f = unpack2<T>(v[u][k]);
                        float* dst = s_in + (r + u * NW) * PIN + c;
                        if (c >= 0 && c < PIN) dst[0] = f.x;
                        if (c + 1 < PIN) dst[1] = f.y;
                    }
                }
            }
        }
    } else {
        unsigned cmask = 0;              // bit k: column lane + 32k lies inside the plane
#pragma unroll
        for (int k = 0; k < G::K; k++) { const int sx = ibx + lane + 32 * k; if (sx >= 0 && sx < p.in_w) cmask |= 1u << k; }
        // footprint rows in flight per thread (RB * K independent loads before the first store); measured: 4 rows help fp32
        // (down2 65 -> 58 us) and hurt the 2-byte loads of bf16 / fp16 (128 -> 149 us)
        constexpr int RB = sizeof(T) == 2 ? 2 : 4;
        for (int r = warp; r < ITH; r += RB * NW) {
            float v[RB][G::K];
#pragma unroll
            for (int u = 0; u < RB; u++) {
                const int sy = iby + r + u * NW;
                const bool rok = sy >= 0 && sy < p.in_h && r + u * NW < ITH;
                const T* xr = xp + (size_t)(rok ? sy : 0) * p.in_w + (ibx + lane);
#pragma unroll
                for (int k = 0; k < G::K; k++) v[u][k] = (rok && ((cmask >> k) & 1u)) ? to_f32(xr[32 * k]) : 0.f;
            }
#pragma unroll
            for (int u = 0; u < RB; u++) {
                if (r + u * NW < ITH) {
#pragma unroll
                    for (int k = 0; k < G::K; k++)
                        if (lane + 32 * k < PIN) s_in[(r + u * NW) * PIN + lane + 32 * k] = v[u][k];
                }
            }
        }
    }
    __syncthreads();

    // ---- h-pass: s_in -> s_h ----------------------------------------------------------------------
    float g[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) g[t] = s_f[t];
    for (int task = tid; task < ITH * G::NSX; task += G::THREADS) {
        const int strip = task / ITH, row = task - strip * ITH;
        const float4* src = reinterpret_cast<const float4*>(s_in + row * PIN + strip * STEP);
        float w[4 * WIN4], out[R];
#pragma unroll
        for (int i = 0; i < WIN4; i++) { const float4 q = src[i]; w[4 * i] = q.x; w[4 * i + 1] = q.y; w[4 * i + 2] = q.z; w[4 * i + 3] = q.w; }
        F::run(w, g, out);
        float4* dst = reinterpret_cast<float4*>(s_h + row * PH + strip * R);
#pragma unroll
        for (int j = 0; j < R / 4; j++) dst[j] = make_float4(out[4 * j], out[4 * j + 1], out[4 * j + 2], out[4 * j + 3]);
    }
    __syncthreads();

    // ---- v-pass: s_h -> y -------------------------------------------------------------------------
#pragma unroll
    for (int t = 0; t < NT; t++) g[t] = s_f[NT + t] * p.gain;
    const int col = tid % G::TOW, ox = ox0 + col;
    if (ox < 0 || ox >= p.out_w) return;
    T* yp = (T*)p.y + (size_t)plane * p.out_h * p.out_w + ox;
    for (int strip = tid / G::TOW; strip < G::NSY; strip += G::SPAR) {
        const float* src = s_h + strip * STEP * PH + col;
        float w[4 * WIN4], out[R];
#pragma unroll
        for (int i = 0; i < F::WIN; i++) w[i] = src[i * PH];
        F::run(w, g, out);
        const int oyb = oy0 + strip * R;
        T* dst = yp + (ptrdiff_t)oyb * p.out_w;
        if (oyb >= 0 && oyb + R <= p.out_h) {
#pragma unroll
            for (int j = 0; j < R; j++) { *dst = from_f32<T>(out[j]); dst += p.out_w; }
        } else {
#pragma unroll
            for (int j = 0; j < R; j++) { if ((unsigned)(oyb + j) < (unsigned)p.out_h) *dst = from_f32<T>(out[j]); dst += p.out_w; }
        }
    }
}

// host: launch one instantiation
template <class T, int UP, int DOWN, int NT, bool SMALL>
int launch(Params p, int64_t planes, cudaStream_t st) {
    using G = Geo<UP, DOWN, NT, SMALL>;
    auto pos_mod = [](int a, int b) { int r = a % b; return r < 0 ? r + b : r; };
    p.offx = (UP > 1) ? pos_mod(-p.padx0, UP) : 0;
    p.offy = (UP > 1) ? pos_mod(-p.pady0, UP) : 0;
    p.tiles_x = (p.out_w + p.offx + G::TOW - 1) / G::TOW;
    p.tiles_y = (p.out_h + p.offy + G::TOH - 1) / G::TOH;
    const int64_t blocks = planes * p.tiles_x * p.tiles_y;
    if (blocks >= (1ll << 31)) { cno_set_error("upfirdn2d: grid too large"); return (int)cudaErrorInvalidValue; }
    auto kern = upfirdn2d_fast_kernel<T, UP, DOWN, NT, SMALL>;
    static DevSmemOptIn optin;                 // per instantiation, per device
    if (G::SMEM > 48 * 1024) {
        cudaError_t e = optin.ensure(kern, G::SMEM);
        if (e != cudaSuccess) { cno_set_error("upfirdn2d: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e)); return (int)e; }
    }
    kern<<<(unsigned)blocks, G::THREADS, G::SMEM, st>>>(p);
    CNO_LAUNCH_CHECK();
    return 0;
}

template <class T, int UP, int DOWN, int NT>
int launch_size(const Params& p, int64_t planes, cudaStream_t st) {
    // small tiles when a large tile would be mostly empty
    const bool small = p.out_w <= Geo<UP, DOWN, NT, false>::TOW / 2 || p.out_h <= Geo<UP, DOWN, NT, false>::TOH / 2;
    return small ? launch<T, UP, DOWN, NT, true>(p, planes, st) : launch<T, UP, DOWN, NT, false>(p, planes, st);
}

// Returns -1 when (up, down, taps) has no instantiation.
template <class T, int UP, int DOWN>
int launch_taps(const Params& p, int64_t planes, int nt, cudaStream_t st) {
    switch (nt) {
        case 4: return launch_size<T, UP, DOWN, 4>(p, planes, st);
        case 8: return launch_size<T, UP, DOWN, 8>(p, planes, st);
        case 12: return launch_size<T, UP, DOWN, 12>(p, planes, st);
        case 16: return launch_size<T, UP, DOWN, 16>(p, planes, st);
        case 24: return launch_size<T, UP, DOWN, 24>(p, planes, st);
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
