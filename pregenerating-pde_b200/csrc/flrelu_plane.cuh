// Fused filtered_lrelu for SMALL planes (the deep U-Net levels: 8^2 ... 32^2 inputs, up-res width <= 128): whole planes live in
// shared memory, one thread per up-sampled COLUMN.  Same C entry point and arguments as the streaming kernel
// (flrelu_fused_impl.cuh, which replaces filtered_lrelu.cu:131-1091 for large planes); the host picks this one when a plane
// fits (flr_plane::make_geom).
//
// Why a second kernel: the streaming kernel gives a thread four columns and keeps 15 warps per SM busy with long planes.  An 8x8
// plane has 26 up-res columns = 8 such threads, each paying a 12-row prologue and three 12-row bodies for 26 useful rows, in
// 2 waves of 1 CTA per SM: 22 us for 5.9 MB (0.04 of the HBM roofline, round-1 sweep).  Here a CTA holds P whole planes:
//   phase 0  x (+ bias, + CIN affine) -> X[P][H][XW] (fp32, zero padding = the FIR halo)          all threads, 16-byte loads
//   phase 1  one thread per column streams down its column in registers: horizontal polyphase up-FIR of each new input row
//            straight from X (the column's 6-7 taps in registers), vertical up-FIR (6 FFMA, taps in uniform registers), leaky
//            ReLU, sign byte -> S (WRITE) or code <- staged mask bytes (READ), vertical down-FIR scattered into 6 rotating
//            accumulators, finished rows -> D[P][OH][DP]
//   phase 2  horizontal down-FIR D -> y (16-byte stores), db partial sums (READ)
//   phase 3  WRITE: sign bytes -> 2-bit codes, 32-bit stores into the mask
// No warm-up rows, no column padding beyond the mask's own 16-element row padding, 4x the threads of the streaming mapping, and
// 18-23 KB of shared memory per 32^2 plane (2.6 KB per 8^2 plane), i.e. 20-64 resident warps per SM.  (A first version kept
// the h-up result A[P][H+12][UWp] in shared memory: 37 KB per 32^2 plane = 10 warps per SM, 37.6 us; latency-bound.)
#pragma once
#include "common.cuh"
#include "flrelu_host.h"

namespace flr_plane {
using namespace flr;

// division by a small runtime constant without the ~20-instruction integer divide: q = umulhi(x, ceil(2^32 / d)), exact for the
// index ranges of this kernel (x * d < 2^32); the magic numbers are computed on the host
struct FastDiv {
    uint32_t m; int d;
#if defined(__CUDACC__)
    __device__ __forceinline__ int div(int x) const { return d == 1 ? x : (int)__umulhi((uint32_t)x, m); }
#endif
};
static inline FastDiv make_fastdiv(int d) { FastDiv f; f.d = d < 1 ? 1 : d; f.m = f.d == 1 ? 0u : (uint32_t)((0x100000000ull + (uint64_t)f.d - 1) / (uint64_t)f.d); return f; }

struct Geom {
    FastDiv dUWp, dwv, dncq, dmw, dMPw;   // divisors: column threads per plane, 16-byte vectors per input row, output column quads,
                                          // mask words per row (G / 4), staged mask words per row (MP / 4)
    int P, NT;               // planes per CTA, threads per CTA
    int XW;                  // X row pitch (floats, multiple of 4); X column i <-> input column cminX + i
    int UWp;                 // column threads per plane = S row pitch = 4 * G
    int SR, MP, mb0;         // rows of S / staged mask rows per plane (= nbodies * PB), READ: row pitch (bytes, multiple of 4), first staged byte
    int DP;                  // D row pitch (floats, multiple of 4)
    int nbodies;
    int offS, offD;          // byte offsets inside dynamic shared memory (X at 0)
    size_t smem;
};

// returns 0 and fills g when the plane kernel covers the plan (make_plan already accepted the arguments)
static inline int make_geom(Geom& g, const Plan& plan, int rs_unused = 0) {
    (void)rs_unused;
    const Params& p = plan.p;
    const int up = plan.up, down = plan.down;
    const int wlx = 6 + (plan.rx < up - 1 ? 1 : 0), wly = 6 + (plan.ry < up - 1 ? 1 : 0);
    const int pb = 6 * down, pq = pb / up, ko = 6;
    g.UWp = 4 * p.G;
    // up-res planes of at most 64 columns (8^2 / 16^2 inputs of the I configuration, 16 -> 8 of D, 8 -> 16 of U): measured 1.3-1.9x over
    // the streaming kernel; at 32^2 (80 columns, 22 KB of shared memory per plane = 10 planes per SM) the two are equal
    if (g.UWp > 96 || p.H > 64 || p.W > 64) return -1;          // 65..96 columns: only when all planes are resident at once, see below
    g.nbodies = (p.OH + 5 + ko - 1) / ko;
    g.SR = g.nbodies * pb;
    (void)pq; (void)wly;
    const int nx = (g.UWp + up - 1) / up + wlx;                 // input columns cminX .. cminX + nx - 1 are read
    g.XW = (nx + 3) / 4 * 4;
    g.MP = (p.G + 4 + 3) / 4 * 4;
    {   // D row: float4 loads of the h-down reach float down * 4 * (ncq - 1) + 4 * NF4 - 1
        const int nf4 = (down * 3 + 6 * down + 3) / 4, need = down * 4 * (p.ncq - 1) + 4 * nf4;
        g.DP = (need > g.UWp ? need : g.UWp);
        g.DP = (g.DP + 3) / 4 * 4;
    }
    const size_t perX = (size_t)p.H * g.XW * 4, perD = (size_t)p.OH * g.DP * 4;
    size_t perS = (size_t)g.SR * g.UWp > (size_t)g.SR * g.MP ? (size_t)g.SR * g.UWp : (size_t)g.SR * g.MP;   // WRITE sign bytes / READ staged mask
    perS = (perS + 15) / 16 * 16;
    const size_t per = perX + perS + perD;
    if (per > 56 * 1024) return -1;
    // planes per SM by shared memory (1 KB per CTA reserved by the driver) and by threads; CTAs of 128-256 threads so that several
    // are resident and their phases overlap
    int per_sm = (int)((220 * 1024) / per);
    if (per_sm > 2048 / g.UWp) per_sm = 2048 / g.UWp;
    if (per_sm < 1) return -1;
    // 80-96 columns (32^2 inputs): 22 KB per plane = 10 planes per SM.  When the tensor fits the machine in one wave the kernel is one
    // plane latency (12-14 us, against a 20 us floor of the streaming kernel's 12-row bodies on short planes, which is slower than
    // the reference's kernel on [32,16,32,32]); with more planes than that the second wave ties it with the streaming kernel
    if (g.UWp > 64 && (long long)p.NC > (long long)per_sm * 148) return -1;
    int P = (96 + g.UWp - 1) / g.UWp;                           // >= 96 threads: small CTAs, many per SM, short tail
    if (P > per_sm) P = per_sm;
    if (P > p.NC) P = p.NC;
    if (P < 1) P = 1;
    g.P = P;
    g.NT = (P * g.UWp + 31) / 32 * 32;
    if (g.NT < 64) g.NT = 64;
    if (g.NT > 512) return -1;
    g.dUWp = make_fastdiv(g.UWp); g.dwv = make_fastdiv(p.W * plan.elem_size / 16); g.dncq = make_fastdiv(p.ncq);
    g.dmw = make_fastdiv(p.G / 4); g.dMPw = make_fastdiv(g.MP / 4);
    g.offS = (int)(P * perX);
    g.offD = g.offS + (int)((size_t)P * perS);
    g.smem = (size_t)g.offD + P * perD + 16;
    return 0;
}

template <class T> __device__ __forceinline__ void store4(T* dst, const float* o, bool vec, int nvalid);
template <> __device__ __forceinline__ void store4<float>(float* dst, const float* o, bool vec, int nvalid) {
    if (vec) *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
    else { for (int i = 0; i < 4; i++) if (i < nvalid) dst[i] = o[i]; }
}
template <> __device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* dst, const float* o, bool vec, int nvalid) {
    if (vec) {
        __nv_bfloat162 a = __floats2bfloat162_rn(o[0], o[1]), b = __floats2bfloat162_rn(o[2], o[3]);
        uint2 u; u.x = *reinterpret_cast<uint32_t*>(&a); u.y = *reinterpret_cast<uint32_t*>(&b);
        *reinterpret_cast<uint2*>(dst) = u;
    } else { for (int i = 0; i < 4; i++) if (i < nvalid) dst[i] = __float2bfloat16_rn(o[i]); }
}
template <> __device__ __forceinline__ void store4<__half>(__half* dst, const float* o, bool vec, int nvalid) {
    if (vec) {
        __half2 a = __floats2half2_rn(o[0], o[1]), b = __floats2half2_rn(o[2], o[3]);
        uint2 u; u.x = *reinterpret_cast<uint32_t*>(&a); u.y = *reinterpret_cast<uint32_t*>(&b);
        *reinterpret_cast<uint2*>(dst) = u;
    } else { for (int i = 0; i < 4; i++) if (i < nvalid) dst[i] = __float2half_rn(o[i]); }
}

template <class T, class C, int MODE, bool AFF>
__global__ void __launch_bounds__(512, 2)
flr_plane_kernel(const __grid_constant__ Params p, const __grid_constant__ Geom q, const T* __restrict__ x) {
    extern __shared__ __align__(16) uint8_t smem[];
    float* const X = reinterpret_cast<float*>(smem);
    uint8_t* const S = smem + q.offS;
    float* const D = reinterpret_cast<float*>(smem + q.offD);
    __shared__ float s_db[64];
    const int tid = threadIdx.x;
    const int plane0 = blockIdx.x * q.P;
    const int np = (p.NC - plane0) < q.P ? (p.NC - plane0) : q.P;
    const bool want_db = FLR_IS_READ(MODE) && p.db != nullptr;
    constexpr int EPV = 16 / (int)sizeof(T);

    // every phase is organised per plane slot: thread (u, c) = (tid / UWp, tid % UWp) works on plane u with index c, c + UWp, ...
    const int u = q.dUWp.div(tid), c = tid - u * q.UWp;
    const bool act = u < np;
    const int plane = plane0 + u;

    // ---- phase 0: x (+ bias, CIN affine) into X, zero halo around it; READ: stage the mask rows -------------------------------
    if (act) {                                                      // zero X (the FIR halo), 16 bytes at a time
        float4* X4 = reinterpret_cast<float4*>(X + (size_t)u * p.H * q.XW);
        for (int i = c; i < p.H * (q.XW >> 2); i += q.UWp) X4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (want_db && tid < 64) s_db[tid] = 0.f;
    __syncthreads();
    if (act) {
        float* Xu = X + (size_t)u * p.H * q.XW;
        const int wv = q.dwv.d, xl = -p.cminX;                       // input column col sits at X column col + xl
        float add = p.b ? to_f32<T>(reinterpret_cast<const T*>(p.b)[plane % p.C]) : 0.f, mul = 1.f;
        if (AFF) { mul = p.pa[plane]; add += p.pd[plane]; }
        const T* xb = x + (size_t)plane * p.H * p.W;
        for (int i = c; i < p.H * wv; i += q.UWp) {
            const int r = q.dwv.div(i), v4 = i - r * wv;
            const Vec16<T> v = ld16_stream(xb + (size_t)i * EPV);
            float* dst = Xu + r * q.XW + xl + v4 * EPV;
#pragma unroll
            for (int e = 0; e < EPV; e++) {
                const int xc = xl + v4 * EPV + e;
                if (xc >= 0 && xc < q.XW) dst[e] = fmaf(to_f32<T>(v.v[e]), mul, add);
            }
        }
        if (FLR_IS_READ(MODE)) {
            // staged mask: row jr <-> mask row jr + sy, byte k <-> mask byte mb0 + k; outside the mask = 0 ("pass")
            const int mw = q.dMPw.d;
            uint32_t* Su = reinterpret_cast<uint32_t*>(S + (size_t)u * q.SR * q.MP);
            const uint8_t* mb = p.mask + (int64_t)plane * p.mask_h * p.mask_pitch;
            for (int i = c; i < q.SR * mw; i += q.UWp) {
                const int jr = q.dMPw.div(i), k = i - jr * mw, my = jr + p.sy, byte = q.mb0 + 4 * k;
                uint32_t v = 0;
                if (my >= 0 && my < p.mask_h && byte >= 0 && byte < p.mask_pitch)
                    v = *reinterpret_cast<const uint32_t*>(mb + (int64_t)my * p.mask_pitch + byte);
                Su[i] = v;
            }
        }
    }
    __syncthreads();

    // ---- phase 1: one thread per up-res column --------------------------------------------------------------------------------
    {
        if (act) {
            // horizontal up-FIR of this column: phase psi = c mod UP, first input sample qx = c / UP (X column qx + k <-> input
            // column cminX + qx + k), taps of the phase in registers
            const int psi = c % C::UP, qx = c / C::UP;
            float hxr[C::WLX];
#pragma unroll
            for (int k = 0; k < C::WLX; k++) hxr[k] = p.hx[psi >> 1][k][psi & 1];
            const float* Xp = X + (size_t)u * p.H * q.XW + qx;
            auto hup = [&](int r) {                                               // r: input row (uniform across the CTA)
                float a = 0.f;
                if (r >= 0 && r < p.H) {
                    const float* xr = Xp + r * q.XW;
#pragma unroll
                    for (int k = 0; k < C::WLX; k++) a = fmaf(xr[k], hxr[k], a);
                }
                return a;
            };
            float* Dp = D + (size_t)u * p.OH * q.DP + c;
            uint8_t* Sp = S + (size_t)u * q.SR * q.UWp + c;                       // WRITE: sign byte of (row, column)
            // READ: code bits of (row j, column c) = bits [sh, sh + 2) of staged byte mbi of row j
            const int e = c + p.sx, eb = e >> 2;                                  // arithmetic shift = floor
            const int mbi = eb - q.mb0, sh = 2 * (e & 3);
            const bool ecol = e >= 0 && e < p.mask_w;
            const uint8_t* Mp = S + (size_t)u * q.SR * q.MP + mbi;
            // Window of the vertical up-FIR: the six h-up rows the NEXT phase needs, as a circular register buffer.  Phase phi of
            // input position qy reads rows qy + cminY + (phi > RY) + k, so one row is pushed per qy, right after phase RY; an
            // unrolled block holds a multiple of six pushes, hence the buffer index is a compile-time constant everywhere and no
            // register is ever moved.
            constexpr int NB2 = (C::PQ % 6 == 0) ? 1 : 2;                         // bodies per unrolled block
            float win[6], acc[C::ND];
#pragma unroll
            for (int r = 0; r < C::ND; r++) acc[r] = 0.f;
#pragma unroll
            for (int w = 0; w < 6; w++) win[w] = hup(p.cminY + w);                // input rows cminY .. cminY + 5
            int rin = p.cminY + 6;                                                // next input row
            const float slope = p.slope;
            for (int b = 0; b < q.nbodies; b += NB2) {
#pragma unroll
                for (int jj = 0; jj < NB2 * C::PB; jj++) {
                    const int jl = jj % C::PB, phi = jj % C::UP;
                    const int wb = (jj / C::UP + (phi > C::RY ? 1 : 0)) % 6;     // pushes done so far in this block
                    if (jj % C::DOWN == 0 && b * C::PB + jj >= p.UH) goto column_done;   // uniform: the rows below feed no output
                    float v = 0.f;
#pragma unroll
                    for (int k = 0; k < C::NTU; k++) v = fmaf(win[(wb + k) % 6], p.vy[phi][k], v);
                    if (phi == C::RY) { win[(jj / C::UP) % 6] = hup(rin); rin++; }    // replaces the oldest row: not needed by later phases
                    if (FLR_IS_READ(MODE)) {
                        uint32_t code = ecol ? ((uint32_t)Mp[0] >> sh) : 0u; Mp += q.MP;
                        float m = (code & 1u) ? slope : 1.f;
                        if (MODE == SIGN_READ && (code & 2u)) m = 0.f;
                        v *= m;
                    } else {
                        if (MODE == SIGN_WRITE) { *Sp = (uint8_t)(__float_as_uint(v) >> 31); Sp += q.UWp; }
                        v = fmaxf(v, v * slope);
                    }
#pragma unroll
                    for (int r = 0; r < C::ND; r++) {
                        const int t = (jl <= C::DOWN * r + C::DOWN - 1) ? jl + C::DOWN * (C::ND - 1 - r) : jl - C::DOWN * (r + 1);
                        acc[r] = fmaf(v, p.fd[t], acc[r]);
                    }
                    if (jl % C::DOWN == C::DOWN - 1) {
                        const int r = jl / C::DOWN, o = (b + jj / C::PB) * C::KO - (C::ND - 1) + r;
                        if (o >= 0 && o < p.OH) Dp[o * q.DP] = acc[r];
                        acc[r] = 0.f;
                    }
                }
            }
            column_done:;
        }
    }
    __syncthreads();

    // ---- phase 2: horizontal down-FIR, four outputs per task; db partial sums --------------------------------------------------
    if (act) {
        T* const yb = reinterpret_cast<T*>(p.y) + (size_t)plane * p.OH * p.OW;
        const float* Du = D + (size_t)u * p.OH * q.DP;
        const bool vec_ok = (p.OW & 3) == 0;
        float dbsum = 0.f;
        for (int i = c; i < p.OH * p.ncq; i += q.UWp) {
            const int o = q.dncq.div(i), cq = i - o * p.ncq;
            const float* dr = Du + o * q.DP + C::DOWN * 4 * cq;
            f4 d[C::NF4];
#pragma unroll
            for (int k = 0; k < C::NF4; k++) d[k] = *reinterpret_cast<const f4*>(dr + 4 * k);
            float out[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                f2 a = mk2(0.f, 0.f);
#pragma unroll
                for (int m = 0; m < C::TD / 2; m++) {
                    const int e = C::DOWN * k + 2 * m;
                    a = ffma2(mk2(f4c(d[e >> 2], e & 3), f4c(d[e >> 2], (e & 3) + 1)), mk2(p.fd[2 * m], p.fd[2 * m + 1]), a);
                }
                out[k] = a.x + a.y;
            }
            const int nvalid = p.OW - 4 * cq;
            store4<T>(yb + o * p.OW + 4 * cq, out, vec_ok, nvalid);
            if (want_db) {
#pragma unroll
                for (int k = 0; k < 4; k++) if (k < nvalid) dbsum += out[k];
            }
        }
        if (want_db) atomicAdd(&s_db[u], dbsum);
    }

    // ---- phase 3 (WRITE): sign bytes -> 2-bit codes; a task = 16 columns of one row = one 32-bit word of the mask ---------------
    if (MODE == SIGN_WRITE && act) {
        const int mw = q.dmw.d;                                      // G / 4 (mask rows are padded to 16 elements)
        const uint8_t* Su = S + (size_t)u * q.SR * q.UWp;
        uint8_t* mb = p.mask + (int64_t)plane * p.mask_h * p.mask_pitch;
        for (int i = c; i < p.UH * mw; i += q.UWp) {
            const int j = q.dmw.div(i), k = i - j * mw;
            const uint4 sb = *reinterpret_cast<const uint4*>(Su + j * q.UWp + 16 * k);
            // four sign bytes b0..b3 (0 / 1) of a word -> b0 | b1 << 2 | b2 << 4 | b3 << 6 in bits 24..31 of word * M (no carries)
            const uint32_t M = 0x01041040u;
            uint32_t code = ((sb.x * M) >> 24) | (((sb.y * M) >> 24) << 8) | (((sb.z * M) >> 24) << 16) | ((sb.w * M) & 0xff000000u);
            const int ncol = p.UW - 16 * k;                          // columns >= UW are padding: codes 0
            if (ncol < 16) code &= ncol <= 0 ? 0u : (0xffffffffu >> (32 - 2 * ncol));
            *reinterpret_cast<uint32_t*>(mb + (uint32_t)(j * p.mask_pitch) + 4 * k) = code;
        }
    }
    if (want_db) {
        __syncthreads();
        if (tid < np) atomicAdd(p.db + (plane0 + tid) % p.C, s_db[tid]);
    }
}

template <class T, class C, int MODE, bool AFF>
int launch_variant(const Plan& plan, const Geom& g, const void* x, cudaStream_t st) {
    auto kern = flr_plane_kernel<T, C, MODE, AFF>;
    static DevSmemOptIn optin;
    cudaError_t e = optin.ensure(kern, g.smem);
    if (e != cudaSuccess) { cno_set_error("filtered_lrelu (plane kernel): cannot reserve %zu bytes of shared memory: %s", g.smem, cudaGetErrorString(e)); return (int)e; }
    const int blocks = (plan.p.NC + g.P - 1) / g.P;
    kern<<<blocks, g.NT, g.smem, st>>>(plan.p, g, reinterpret_cast<const T*>(x));
    CNO_LAUNCH_CHECK();
    return 0;
}

template <class T, class C>
int launch_mode(const Plan& plan, const Geom& g, const void* x, int mode, cudaStream_t st) {
    if (plan.p.pa) {
        if (mode == CNO_SIGN_WRITE) return launch_variant<T, C, SIGN_WRITE, true>(plan, g, x, st);
        if (mode == CNO_SIGN_NONE) return launch_variant<T, C, SIGN_NONE, true>(plan, g, x, st);
        return -1;
    }
    if (mode == CNO_SIGN_WRITE) return launch_variant<T, C, SIGN_WRITE, false>(plan, g, x, st);
    if (mode == CNO_SIGN_READ) return launch_variant<T, C, SIGN_READ, false>(plan, g, x, st);
    if (mode == CNO_SIGN_READ_SIGNONLY) return launch_variant<T, C, SIGN_READ_SO, false>(plan, g, x, st);
    return launch_variant<T, C, SIGN_NONE, false>(plan, g, x, st);
}

template <class T>
int launch_cfg(const Plan& plan, const Geom& g, const void* x, int mode, cudaStream_t st) {
#define V(U, D, R) if (plan.up == U && plan.down == D && plan.rx == R && plan.ry == R) return launch_mode<T, Cfg<U, D, R, R>>(plan, g, x, mode, st);
#ifdef FLR_QUICK
    V(2, 2, 1)
#else
    V(2, 2, 0) V(2, 2, 1) V(2, 4, 0) V(2, 4, 1) V(4, 2, 0) V(4, 2, 1) V(4, 2, 2) V(4, 2, 3)
#endif
#undef V
    return -1;
}

}  // namespace flr_plane
