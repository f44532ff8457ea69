// CIN: time-conditioned instance norm (FILM.forward, CNO_timeModule_CIN.py:60-70) as ONE pass over x.
//   fwd:  mu, r = 1/sqrt(var+eps) per (n,c) plane;  y = (x-mu)*r*(gamma_c*s_nc) + (beta_c*s_nc + q_nc)
//   bwd:  S1 = sum dy, S2 = sum dy*xhat;  dx = gamma_c*s_nc*r*(dy - S1/HW - xhat*S2/HW)
// A plane is held in registers by GS threads (GS chosen on the host so that each thread keeps at most
// VPT 16-byte vectors); statistics use warp shuffles + one shared-memory hop per group, then the affine
// time conditioning is applied from registers, so HBM sees x exactly once (and dy once in backward).
// Planes that do not fit (hw not a multiple of the vector width, or > GS*VPT vectors) take the
// streaming path (three sweeps, second and third from L2).
#include <stdlib.h>
#include "common.cuh"
#include <cooperative_groups.h>
namespace cgrp = cooperative_groups;

namespace {

constexpr int CTA = 256;

// 16-byte vector held as four 32-bit registers; elements are unpacked with bit operations so that the per-thread copy of the
// plane stays in registers (an array of __nv_bfloat16 elements ends up in local memory: ptxas reported a 440-byte stack frame).
template <class T> struct Pack;
template <> struct Pack<float> {
    static constexpr int V = 4;
    static __device__ __forceinline__ void get(const uint4& u, float (&f)[4]) {
        f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
    }
    static __device__ __forceinline__ uint4 put(const float (&f)[4]) {
        return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
    }
};
template <> struct Pack<__nv_bfloat16> {
    static constexpr int V = 8;
    static __device__ __forceinline__ void get(const uint4& u, float (&f)[8]) {
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int i = 0; i < 4; i++) { f[2 * i] = __uint_as_float(w[i] << 16); f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
    }
    static __device__ __forceinline__ uint4 put(const float (&f)[8]) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; i++) { __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]); w[i] = *reinterpret_cast<uint32_t*>(&h); }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
};
template <> struct Pack<__half> {
    static constexpr int V = 8;
    static __device__ __forceinline__ void get(const uint4& u, float (&f)[8]) {
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int i = 0; i < 4; i++) { const float2 v = __half22float2(*reinterpret_cast<const __half2*>(&w[i])); f[2 * i] = v.x; f[2 * i + 1] = v.y; }
    }
    static __device__ __forceinline__ uint4 put(const float (&f)[8]) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; i++) { __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]); w[i] = *reinterpret_cast<uint32_t*>(&h); }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
};
__device__ __forceinline__ uint4 ldg16_stream(const void* p) {
    uint4 u;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(p));
    return u;
}

// sums of (a, b) over the `gs` threads that share a plane.  gs is a power of two, 8 <= gs <= 1024.
// red: shared scratch, 2 * blockDim/32 floats per slot; every call site uses its own slot, so one barrier per call suffices.
__device__ __forceinline__ void group_sum2(float& a, float& b, int gs, float* red, int slot) {
    const int lanes = gs < 32 ? gs : 32;
    for (int o = lanes >> 1; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
    if (gs <= 32) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nw_total = blockDim.x >> 5;
    const int wpg = gs >> 5;                 // warps per group
    float2* buf = reinterpret_cast<float2*>(red) + slot * nw_total;
    if (lane == 0) buf[warp] = make_float2(a, b);
    __syncthreads();
    const int w0 = (warp / wpg) * wpg;
    float sa = 0.f, sb = 0.f;
    for (int i = 0; i < wpg; i++) { const float2 v = buf[w0 + i]; sa += v.x; sb += v.y; }
    a = sa; b = sb;
}
__device__ __forceinline__ float group_sum(float v, int gs, float* red, int slot) {
    float z = 0.f;
    group_sum2(v, z, gs, red, slot);
    return v;
}

// FiLM conditioning computed inside the forward kernel (FILM.forward, CNO_timeModule_CIN.py:62-65):
//   scale[n,c] = lat2scale(inp2lat_sacale(t_n))_c = b2s[c] + sum_j W2s[c,j] * (w1s[j] * t_n + b1s[j]),  same for shift
struct FilmW {
    const float *t;                          // [N] lead times
    const float *w1s, *b1s, *W2s, *b2s;      // scale branch: Linear(1->lat) weight[lat] / bias[lat], Linear(lat->C) weight[C,lat] / bias[C]
    const float *w1b, *b1b, *W2b, *b2b;      // shift branch
    float* scale_out;                        // [N,C] scale saved for backward (may be null)
    int lat;
};

struct CinParams {
    const void* x; const void* dy; void* y;
    const void* res;                         // forward only, optional: y = res + CIN(x)  (ResidualBlock epilogue, CNO_timeModule_CIN.py:503-507)
    const float *gamma, *beta, *scale, *shift;
    float *mean, *rstd, *s1, *s2;
    float *aff_a, *aff_d;                    // forward only, optional: the plane's affine y = aff_a * x + aff_d; with y == nullptr the kernel
                                             // stops there (the consumer applies it: filtered_lrelu prologue)
    int64_t planes; int c, hw, gs; float eps;
    FilmW film;                              // film.t != nullptr: scale / shift come from the FiLM weights
};

// scale and shift of one plane, cooperatively by the `gs` threads of its group (all threads of the block must call)
__device__ __forceinline__ void film_affine(const CinParams& p, bool live, int64_t plane, int t, int gs, float* red, float& s, float& q);

template <class T, int VPT>
__global__ void __launch_bounds__(1024) cin_fwd_reg_kernel(CinParams p) {
    constexpr int V = Pack<T>::V;
    __shared__ float red[8 * 32];
    const int gs = p.gs, ppc = blockDim.x / gs;
    const int64_t plane = (int64_t)blockIdx.x * ppc + threadIdx.x / gs;
    const int t = threadIdx.x % gs;
    const bool live = plane < p.planes;
    const int nvec = p.hw / V;
    const T* xp = (const T*)p.x + (live ? plane : 0) * (int64_t)p.hw;
    // the whole plane goes into registers; all loads are issued before anything waits on them
    uint4 v[VPT];
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        v[k] = (live && i < nvec) ? ldg16_stream(xp + (int64_t)i * V) : make_uint4(0u, 0u, 0u, 0u);
    }
    // conditioning (independent of x): its weight loads and reduction overlap the x loads in flight
    float s = 0.f, q = 0.f;
    if (p.film.t) film_affine(p, live, plane, t, gs, red, s, q);
    else if (live) { s = p.scale[plane]; q = p.shift[plane]; }
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; k++) {        // padding vectors are zero: they add nothing
        float f[V]; Pack<T>::get(v[k], f);
#pragma unroll
        for (int e = 0; e < V; e++) sum += f[e];
    }
    const float inv_hw = 1.f / (float)p.hw;
    const float mu = group_sum(sum, gs, red, 0) * inv_hw;
    float sq = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        if (t + k * gs < nvec) {
            float f[V]; Pack<T>::get(v[k], f);
#pragma unroll
            for (int e = 0; e < V; e++) { const float d = f[e] - mu; sq = fmaf(d, d, sq); }
        }
    }
    const float var = group_sum(sq, gs, red, 1) * inv_hw;
    const float r = rsqrtf(var + p.eps);
    if (!live) return;
    const int ch = (int)(plane % p.c);
    const float a = p.gamma[ch] * s * r;
    const float d0 = p.beta[ch] * s + q - mu * a;
    if (t == 0) { p.mean[plane] = mu; p.rstd[plane] = r; if (p.aff_a) { p.aff_a[plane] = a; p.aff_d[plane] = d0; } }
    T* yp = (T*)p.y + plane * (int64_t)p.hw;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        if (p.y && i < nvec) {
            float f[V]; Pack<T>::get(v[k], f);
#pragma unroll
            for (int e = 0; e < V; e++) f[e] = fmaf(f[e], a, d0);
            if (p.res) {
                float rr[V]; Pack<T>::get(ldg16_stream((const T*)p.res + plane * (int64_t)p.hw + (int64_t)i * V), rr);
#pragma unroll
                for (int e = 0; e < V; e++) f[e] += rr[e];
            }
            *reinterpret_cast<uint4*>(yp + (int64_t)i * V) = Pack<T>::put(f);
        }
    }
}

template <class T, int VPT>
__global__ void __launch_bounds__(1024) cin_bwd_reg_kernel(CinParams p) {
    constexpr int V = Pack<T>::V;
    __shared__ float red[8 * 32];
    const int gs = p.gs, ppc = blockDim.x / gs;
    const int64_t plane = (int64_t)blockIdx.x * ppc + threadIdx.x / gs;
    const int t = threadIdx.x % gs;
    const bool live = plane < p.planes;
    const int nvec = p.hw / V;
    const int64_t base = (live ? plane : 0) * (int64_t)p.hw;
    const T* xp = (const T*)p.x + base; const T* gp = (const T*)p.dy + base;
    uint4 vx[VPT], vg[VPT];
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        const bool ok = live && i < nvec;
        vx[k] = ok ? ldg16_stream(xp + (int64_t)i * V) : make_uint4(0u, 0u, 0u, 0u);
        vg[k] = ok ? ldg16_stream(gp + (int64_t)i * V) : make_uint4(0u, 0u, 0u, 0u);
    }
    const float mu = live ? p.mean[plane] : 0.f, r = live ? p.rstd[plane] : 0.f;
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; k++) {        // padding vectors have dy = 0: they add nothing
        float fx[V], fg[V]; Pack<T>::get(vx[k], fx); Pack<T>::get(vg[k], fg);
#pragma unroll
        for (int e = 0; e < V; e++) { s1 += fg[e]; s2 = fmaf(fg[e], (fx[e] - mu) * r, s2); }
    }
    group_sum2(s1, s2, gs, red, 0);
    if (!live) return;
    const int ch = (int)(plane % p.c);
    if (t == 0) { p.s1[plane] = s1; p.s2[plane] = s2; }
    const float inv_hw = 1.f / (float)p.hw;
    const float a = p.gamma[ch] * p.scale[plane] * r;
    const float m1 = s1 * inv_hw, m2 = s2 * inv_hw;
    T* dp = (T*)p.y + base;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        if (i < nvec) {
            float fx[V], fg[V]; Pack<T>::get(vx[k], fx); Pack<T>::get(vg[k], fg);
#pragma unroll
            for (int e = 0; e < V; e++) fg[e] = a * (fg[e] - m1 - (fx[e] - mu) * r * m2);
            *reinterpret_cast<uint4*>(dp + (int64_t)i * V) = Pack<T>::put(fg);
        }
    }
}

// ---- planes larger than one CTA's registers (256^2: the wide fine-tune configuration): a thread-block CLUSTER holds the plane --
// CS CTAs of 512 threads each keep 1/CS of the plane in registers; the statistics are reduced inside each CTA (shuffles + one
// shared-memory hop) and the CS partial results are exchanged through distributed shared memory, so x (and dy) still cross HBM
// exactly once.  One exchange per kernel; the closing cluster barrier (arrive right after the remote reads, wait before exit)
// keeps a CTA's shared memory alive until its peers have read it and overlaps with the store phase.
constexpr int CLUSTER_CTA = 512;

__device__ __forceinline__ void cluster_sum2(cgrp::cluster_group& cluster, float& a, float& b, float2* part, float* red, int round) {
    group_sum2(a, b, (int)blockDim.x, red, round);             // CTA totals in every thread
    if (threadIdx.x == 0) part[round] = make_float2(a, b);
    cluster.sync();
    float sa = 0.f, sb = 0.f;
    const unsigned cs = cluster.num_blocks();
    for (unsigned r = 0; r < cs; r++) {                         // same order in every CTA: identical results
        const float2 v = *cluster.map_shared_rank(&part[round], r);
        sa += v.x; sb += v.y;
    }
    a = sa; b = sb;
    cluster.barrier_arrive();                                   // done reading the peers: the matching barrier_wait() sits at the kernel's end
}

template <class T, int VPT>
__global__ void __launch_bounds__(CLUSTER_CTA, 2) cin_fwd_cluster_kernel(CinParams p) {
    constexpr int V = Pack<T>::V;
    __shared__ float red[8 * 32];
    __shared__ float2 part[2];
    cgrp::cluster_group cluster = cgrp::this_cluster();
    const int cs = (int)cluster.num_blocks(), rank = (int)cluster.block_rank();
    const int64_t plane = blockIdx.x / cs;
    const int gs = cs * CLUSTER_CTA, t = rank * CLUSTER_CTA + threadIdx.x;
    const int nvec = p.hw / V;
    const T* xp = (const T*)p.x + plane * (int64_t)p.hw;
    uint4 v[VPT];
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        v[k] = i < nvec ? ldg16_stream(xp + (int64_t)i * V) : make_uint4(0u, 0u, 0u, 0u);
    }
    float s = 0.f, q = 0.f;                                     // conditioning: every CTA of the cluster computes it (tiny)
    if (p.film.t) film_affine(p, true, plane, threadIdx.x, CLUSTER_CTA, red, s, q);
    else { s = p.scale[plane]; q = p.shift[plane]; }
    // statistics with ONE cluster exchange: every CTA reduces its own part to (sum_r, M2_r about its own mean) with two cheap
    // CTA-level reductions, the cluster exchanges the pairs, and the parts are combined with the pairwise-variance identity
    //   M2 = sum_r [ M2_r + n_r * (mean_r - mu)^2 ]      (exact; no E[x^2] - mu^2 cancellation)
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        float f[V]; Pack<T>::get(v[k], f);
#pragma unroll
        for (int e = 0; e < V; e++) sum += f[e];
    }
    sum = group_sum(sum, CLUSTER_CTA, red, 0);
    auto count_of = [&](int rk) {                              // elements held by CTA rk
        int n = 0;
        for (int k = 0; k < VPT; k++) { const int m = nvec - k * gs - rk * CLUSTER_CTA; n += m < 0 ? 0 : (m > CLUSTER_CTA ? CLUSTER_CTA : m); }
        return n * V;
    };
    const int n_mine = count_of(rank);
    const float mean_mine = n_mine > 0 ? sum / (float)n_mine : 0.f;
    float sq = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        if (t + k * gs < nvec) {
            float f[V]; Pack<T>::get(v[k], f);
#pragma unroll
            for (int e = 0; e < V; e++) { const float d = f[e] - mean_mine; sq = fmaf(d, d, sq); }
        }
    }
    sq = group_sum(sq, CLUSTER_CTA, red, 1);
    if (threadIdx.x == 0) part[0] = make_float2(sum, sq);
    cluster.sync();
    float tot = 0.f, m2 = 0.f;
    float2 pr[8];
#pragma unroll
    for (int rk = 0; rk < 8; rk++) { pr[rk] = rk < cs ? *cluster.map_shared_rank(&part[0], rk) : make_float2(0.f, 0.f); tot += pr[rk].x; }
    cluster.barrier_arrive();                                   // done reading the peers; waited for before exit (split phase)
    const float inv_hw = 1.f / (float)p.hw;
    const float mu = tot * inv_hw;
#pragma unroll
    for (int rk = 0; rk < 8; rk++) {
        if (rk < cs) {
            const int n = count_of(rk);
            const float dm = (n > 0 ? pr[rk].x / (float)n : mu) - mu;
            m2 += pr[rk].y + (float)n * dm * dm;
        }
    }
    const float r = rsqrtf(m2 * inv_hw + p.eps);
    const int ch = (int)(plane % p.c);
    const float a = p.gamma[ch] * s * r;
    const float d0 = p.beta[ch] * s + q - mu * a;
    if (t == 0) { p.mean[plane] = mu; p.rstd[plane] = r; if (p.aff_a) { p.aff_a[plane] = a; p.aff_d[plane] = d0; } }
    T* yp = (T*)p.y + plane * (int64_t)p.hw;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        if (p.y && i < nvec) {
            float f[V]; Pack<T>::get(v[k], f);
#pragma unroll
            for (int e = 0; e < V; e++) f[e] = fmaf(f[e], a, d0);
            if (p.res) {
                float rr[V]; Pack<T>::get(ldg16_stream((const T*)p.res + plane * (int64_t)p.hw + (int64_t)i * V), rr);
#pragma unroll
                for (int e = 0; e < V; e++) f[e] += rr[e];
            }
            *reinterpret_cast<uint4*>(yp + (int64_t)i * V) = Pack<T>::put(f);
        }
    }
    cluster.barrier_wait();                                     // a CTA's shared memory must outlive its peers' reads
}

template <class T, int VPT>
__global__ void __launch_bounds__(CLUSTER_CTA, 2) cin_bwd_cluster_kernel(CinParams p) {
    constexpr int V = Pack<T>::V;
    __shared__ float red[8 * 32];
    __shared__ float2 part[2];
    cgrp::cluster_group cluster = cgrp::this_cluster();
    const int cs = (int)cluster.num_blocks(), rank = (int)cluster.block_rank();
    const int64_t plane = blockIdx.x / cs;
    const int gs = cs * CLUSTER_CTA, t = rank * CLUSTER_CTA + threadIdx.x;
    const int nvec = p.hw / V;
    const int64_t base = plane * (int64_t)p.hw;
    const T* xp = (const T*)p.x + base; const T* gp = (const T*)p.dy + base;
    uint4 vx[VPT], vg[VPT];
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        vx[k] = i < nvec ? ldg16_stream(xp + (int64_t)i * V) : make_uint4(0u, 0u, 0u, 0u);
        vg[k] = i < nvec ? ldg16_stream(gp + (int64_t)i * V) : make_uint4(0u, 0u, 0u, 0u);
    }
    const float mu = p.mean[plane], r = p.rstd[plane];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        float fx[V], fg[V]; Pack<T>::get(vx[k], fx); Pack<T>::get(vg[k], fg);
#pragma unroll
        for (int e = 0; e < V; e++) { s1 += fg[e]; s2 = fmaf(fg[e], (fx[e] - mu) * r, s2); }
    }
    cluster_sum2(cluster, s1, s2, part, red, 0);
    const int ch = (int)(plane % p.c);
    if (t == 0) { p.s1[plane] = s1; p.s2[plane] = s2; }
    const float inv_hw = 1.f / (float)p.hw;
    const float a = p.gamma[ch] * p.scale[plane] * r;
    const float m1 = s1 * inv_hw, m2 = s2 * inv_hw;
    T* dp = (T*)p.y + base;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        if (i < nvec) {
            float fx[V], fg[V]; Pack<T>::get(vx[k], fx); Pack<T>::get(vg[k], fg);
#pragma unroll
            for (int e = 0; e < V; e++) fg[e] = a * (fg[e] - m1 - (fx[e] - mu) * r * m2);
            *reinterpret_cast<uint4*>(dp + (int64_t)i * V) = Pack<T>::put(fg);
        }
    }
    cluster.barrier_wait();                                     // a CTA's shared memory must outlive its peers' reads
}

// streaming path: one CTA per plane, any hw
template <class T, bool BWD>
__global__ void __launch_bounds__(CTA) cin_stream_kernel(CinParams p) {
    __shared__ float red[8 * 32];
    const int64_t plane = blockIdx.x;
    const int64_t base = plane * (int64_t)p.hw;
    const T* xp = (const T*)p.x + base;
    const int ch = (int)(plane % p.c);
    const float inv_hw = 1.f / (float)p.hw;
    if (!BWD) {
        float sum = 0.f;
        for (int i = threadIdx.x; i < p.hw; i += CTA) sum += to_f32(xp[i]);
        const float mu = group_sum(sum, CTA, red, 0) * inv_hw;
        float sq = 0.f;
        for (int i = threadIdx.x; i < p.hw; i += CTA) { float d = to_f32(xp[i]) - mu; sq = fmaf(d, d, sq); }
        const float r = rsqrtf(group_sum(sq, CTA, red, 1) * inv_hw + p.eps);
        float s, q;
        if (p.film.t) film_affine(p, true, plane, threadIdx.x, CTA, red, s, q);
        else { s = p.scale[plane]; q = p.shift[plane]; }
        const float a = p.gamma[ch] * s * r, d0 = p.beta[ch] * s + q - mu * a;
        if (threadIdx.x == 0) { p.mean[plane] = mu; p.rstd[plane] = r; if (p.aff_a) { p.aff_a[plane] = a; p.aff_d[plane] = d0; } }
        T* yp = (T*)p.y + base;
        const T* rp = p.res ? (const T*)p.res + base : nullptr;
        if (p.y) for (int i = threadIdx.x; i < p.hw; i += CTA) yp[i] = from_f32<T>(fmaf(to_f32(xp[i]), a, d0) + (rp ? to_f32(rp[i]) : 0.f));
    } else {
        const T* gp = (const T*)p.dy + base;
        const float mu = p.mean[plane], r = p.rstd[plane];
        float s1 = 0.f, s2 = 0.f;
        for (int i = threadIdx.x; i < p.hw; i += CTA) { float g = to_f32(gp[i]); s1 += g; s2 = fmaf(g, (to_f32(xp[i]) - mu) * r, s2); }
        group_sum2(s1, s2, CTA, red, 0);
        if (threadIdx.x == 0) { p.s1[plane] = s1; p.s2[plane] = s2; }
        const float a = p.gamma[ch] * p.scale[plane] * r, m1 = s1 * inv_hw, m2 = s2 * inv_hw;
        T* dp = (T*)p.y + base;
        for (int i = threadIdx.x; i < p.hw; i += CTA) {
            const float xh = (to_f32(xp[i]) - mu) * r;
            dp[i] = from_f32<T>(a * (to_f32(gp[i]) - m1 - xh * m2));
        }
    }
}

__device__ __forceinline__ void film_affine(const CinParams& p, bool live, int64_t plane, int t, int gs, float* red, float& s, float& q) {
    const FilmW& f = p.film;
    float as = 0.f, aq = 0.f;
    if (live) {
        const int n = (int)(plane / p.c), ch = (int)(plane % p.c);
        const float tn = f.t[n];
        const float* rs = f.W2s + (int64_t)ch * f.lat; const float* rq = f.W2b + (int64_t)ch * f.lat;
#pragma unroll 8
        for (int j = t; j < f.lat; j += gs) {      // unrolled: small groups (gs = 8) run 16 iterations, their loads must overlap
            as = fmaf(rs[j], fmaf(f.w1s[j], tn, f.b1s[j]), as);
            aq = fmaf(rq[j], fmaf(f.w1b[j], tn, f.b1b[j]), aq);
        }
    }
    group_sum2(as, aq, gs, red, 2);
    s = 0.f; q = 0.f;
    if (live) {
        const int ch = (int)(plane % p.c);
        s = as + f.b2s[ch]; q = aq + f.b2b[ch];
        if (t == 0 && f.scale_out) f.scale_out[plane] = s;
    }
}

// ---- backward of the conditioning: every small gradient of one FILM module in ONE launch ----------------------
//   ds = gamma*S2 + beta*S1 (d scale), dq = S1 (d shift);  per channel  P = sum_n ds*t, Q = sum_n ds, P' = sum_n dq*t, Q' = sum_n dq
//   dgamma = sum_n scale*S2, dbeta = sum_n scale*S1, dW2s[c,j] = w1s[j]*P + b1s[j]*Q, db2s = Q, dw1s[j] = sum_c W2s[c,j]*P, db1s[j] = sum_c W2s[c,j]*Q
// Blocks 0..nb-1 own 8 channels each; ceil(lat/32) tail blocks own the [lat] vectors (deterministic: no atomics).
struct FilmBwd {
    const float *s1, *s2, *t, *gamma, *beta, *scale, *w1s, *b1s, *W2s, *w1b, *b1b, *W2b;
    float *dgamma, *dbeta, *dW2s, *db2s, *dw1s, *db1s, *dW2b, *db2b, *dw1b, *db1b;
    int n, c, lat;
};

__device__ __forceinline__ void film_channel_sums(const FilmBwd& p, int ch, int lane, float out[6]) {
    float dg = 0.f, db = 0.f, P = 0.f, Q = 0.f, Pq = 0.f, Qq = 0.f;
    const float g = p.gamma[ch], b = p.beta[ch];
    for (int n = lane; n < p.n; n += 32) {
        const int64_t i = (int64_t)n * p.c + ch;
        const float a1 = p.s1[i], a2 = p.s2[i], sc = p.scale[i], tn = p.t[n];
        const float ds = fmaf(g, a2, b * a1);
        dg = fmaf(sc, a2, dg); db = fmaf(sc, a1, db);
        P = fmaf(ds, tn, P); Q += ds; Pq = fmaf(a1, tn, Pq); Qq += a1;
    }
    out[0] = warp_sum(dg); out[1] = warp_sum(db); out[2] = warp_sum(P); out[3] = warp_sum(Q); out[4] = warp_sum(Pq); out[5] = warp_sum(Qq);
}

__global__ void __launch_bounds__(256) film_bwd_finish_kernel(FilmBwd p) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nb = (p.c + 7) / 8;
    if ((int)blockIdx.x < nb) {
        const int ch = blockIdx.x * 8 + warp;            // one warp per channel
        if (ch >= p.c) return;
        float r[6];
        film_channel_sums(p, ch, lane, r);
        if (lane == 0) { p.dgamma[ch] = r[0]; p.dbeta[ch] = r[1]; p.db2s[ch] = r[3]; p.db2b[ch] = r[5]; }
        for (int j = lane; j < p.lat; j += 32) {
            p.dW2s[(int64_t)ch * p.lat + j] = fmaf(p.w1s[j], r[2], p.b1s[j] * r[3]);
            p.dW2b[(int64_t)ch * p.lat + j] = fmaf(p.w1b[j], r[4], p.b1b[j] * r[5]);
        }
        return;
    }
    // tail blocks: dw1 / db1 of both branches = W2^T applied to the per-channel sums.  Tail block k owns 32 of the `lat`
    // outputs; its 256 threads are 32 columns x 8 channel slices, combined through shared memory (fixed order: deterministic).
    extern __shared__ float pq[];                          // [4][C] sums, then [8][32][4] partials
    float* part = pq + 4 * p.c;
    for (int ch = threadIdx.x; ch < p.c; ch += blockDim.x) {   // thread per channel, coalesced over channels
        const float g = p.gamma[ch], b = p.beta[ch];
        float P = 0.f, Q = 0.f, Pq = 0.f, Qq = 0.f;
#pragma unroll 16
        for (int n = 0; n < p.n; n++) {          // unrolled: 32 independent loads in flight instead of one L2 round trip per sample
            const float a1 = p.s1[(int64_t)n * p.c + ch], a2 = p.s2[(int64_t)n * p.c + ch], tn = p.t[n];
            const float ds = fmaf(g, a2, b * a1);
            P = fmaf(ds, tn, P); Q += ds; Pq = fmaf(a1, tn, Pq); Qq += a1;
        }
        pq[ch] = P; pq[p.c + ch] = Q; pq[2 * p.c + ch] = Pq; pq[3 * p.c + ch] = Qq;
    }
    __syncthreads();
    const int jl = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const int j = ((int)blockIdx.x - nb) * 32 + jl;
    float a = 0.f, b = 0.f, cc = 0.f, d = 0.f;
    if (j < p.lat) {
        const int per = (p.c + 7) / 8, c0 = sl * per, c1 = min(p.c, c0 + per);
#pragma unroll 8
        for (int ch = c0; ch < c1; ch++) {
            const float ws = p.W2s[(int64_t)ch * p.lat + j], wb = p.W2b[(int64_t)ch * p.lat + j];
            a = fmaf(ws, pq[ch], a); b = fmaf(ws, pq[p.c + ch], b);
            cc = fmaf(wb, pq[2 * p.c + ch], cc); d = fmaf(wb, pq[3 * p.c + ch], d);
        }
    }
    float* mine = part + (sl * 32 + jl) * 4;
    mine[0] = a; mine[1] = b; mine[2] = cc; mine[3] = d;
    __syncthreads();
    if (sl == 0 && j < p.lat) {
        float r[4] = {0.f, 0.f, 0.f, 0.f};
        for (int k = 0; k < 8; k++) for (int i = 0; i < 4; i++) r[i] += part[(k * 32 + jl) * 4 + i];
        p.dw1s[j] = r[0]; p.db1s[j] = r[1]; p.dw1b[j] = r[2]; p.db1b[j] = r[3];
    }
}

// ---- statistics-only forward of large planes (cno_film_stats) ------------------------------------------------------------------
// Nothing of the plane is kept, so it is streamed once instead of being parked in registers (the register kernels run 2 CTAs
// per SM and idle the memory system while they reduce: 39 us for fp32 [32,64,128,128]).  A thread folds chunks of 16 values,
// centred about the chunk's own mean, into a running (count, mean, M2) with the pairwise update
//   M2 = M2_a + M2_b + (mean_b - mean_a)^2 * n_a * n_b / (n_a + n_b)
// (no E[x^2] - mean^2 cancellation); lanes, then warps, merge the same way.  ~32 registers: 8 CTAs of 256 threads per SM.
struct Moments { float n, mean, m2; };
__device__ __forceinline__ void moments_merge(Moments& a, float bn, float bmean, float bm2) {
    const float n = a.n + bn;
    const float w = n > 0.f ? __fdividef(bn, n) : 0.f;
    const float d = bmean - a.mean;
    a.m2 = a.m2 + bm2 + d * d * a.n * w;
    a.mean = fmaf(d, w, a.mean);
    a.n = n;
}

constexpr int STATS_MIN_VEC = 1024;          // smaller planes: the register kernels with the store loop off (latency-bound anyway)

template <class T>
__global__ void __launch_bounds__(CTA) cin_stats_kernel(CinParams p) {
    constexpr int V = Pack<T>::V, U = 16 / V;           // 16 fp32 values per chunk
    __shared__ float red[8 * 32];
    __shared__ float wpart[3][CTA / 32];
    const int64_t plane = blockIdx.x;
    const int t = threadIdx.x;
    const int nvec = p.hw / V;
    const T* xp = (const T*)p.x + plane * (int64_t)p.hw;
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; u++) v[u] = (t + u * CTA < nvec) ? ldg16_stream(xp + (int64_t)(t + u * CTA) * V) : make_uint4(0u, 0u, 0u, 0u);
    float s = 0.f, q = 0.f;                  // conditioning under the first loads
    if (p.film.t) film_affine(p, true, plane, t, CTA, red, s, q);
    else { s = p.scale[plane]; q = p.shift[plane]; }
    Moments acc{0.f, 0.f, 0.f};
    for (int i0 = t; i0 < nvec; i0 += U * CTA) {
        float f[U][V];
        float sum = 0.f; int cnt = 0;
#pragma unroll
        for (int u = 0; u < U; u++) {        // padding vectors are zero: they add nothing to the sum
            Pack<T>::get(v[u], f[u]);
            if (i0 + u * CTA < nvec) cnt += V;
#pragma unroll
            for (int e = 0; e < V; e++) sum += f[u][e];
        }
        const int nx = i0 + U * CTA;         // next chunk's loads go out before this chunk's arithmetic retires
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = (nx + u * CTA < nvec) ? ldg16_stream(xp + (int64_t)(nx + u * CTA) * V) : make_uint4(0u, 0u, 0u, 0u);
        const float m = __fdividef(sum, (float)cnt);
        float m2 = 0.f;
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (i0 + u * CTA < nvec) {
#pragma unroll
                for (int e = 0; e < V; e++) { const float d = f[u][e] - m; m2 = fmaf(d, d, m2); }
            }
        }
        moments_merge(acc, (float)cnt, m, m2);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float bn = __shfl_xor_sync(0xffffffffu, acc.n, o), bm = __shfl_xor_sync(0xffffffffu, acc.mean, o),
                    b2 = __shfl_xor_sync(0xffffffffu, acc.m2, o);
        moments_merge(acc, bn, bm, b2);
    }
    if ((t & 31) == 0) { wpart[0][t >> 5] = acc.n; wpart[1][t >> 5] = acc.mean; wpart[2][t >> 5] = acc.m2; }
    __syncthreads();
    if (t == 0) {
        Moments tot{0.f, 0.f, 0.f};
#pragma unroll
        for (int w = 0; w < CTA / 32; w++) moments_merge(tot, wpart[0][w], wpart[1][w], wpart[2][w]);
        const float mu = tot.mean, r = rsqrtf(tot.m2 / (float)p.hw + p.eps);
        const int ch = (int)(plane % p.c);
        const float a = p.gamma[ch] * s * r;
        p.mean[plane] = mu; p.rstd[plane] = r;
        if (p.aff_a) { p.aff_a[plane] = a; p.aff_d[plane] = p.beta[ch] * s + q - mu * a; }
    }
}

template <class T, bool BWD>
int cin_launch(CinParams p, cudaStream_t st) {
    constexpr int V = 16 / sizeof(T);
    // 16-byte vectors per thread: bounded by the 64 registers a 1024-thread CTA leaves each thread (16-bit types unpack to twice
    // as many fp32 values per vector, so they hold half as many)
    static EnvFlag env_no_stats, env_no_cluster;
    constexpr int VPT = (BWD ? 4 : 8) / (sizeof(T) == 2 ? 2 : 1);
    const bool vec_ok = (p.hw % V == 0) && !(((uintptr_t)p.x | (uintptr_t)p.y | (uintptr_t)p.dy | (uintptr_t)p.res) & 15u);
    const int nvec = p.hw / V;
    int gs = 8;
    while (gs < 1024 && (int64_t)gs * VPT < nvec) gs <<= 1;
    if constexpr (!BWD) {
        if (!p.y && vec_ok && nvec >= STATS_MIN_VEC && !env_no_stats.on("CNO_CIN_NO_STATS_KERNEL")) {
            cin_stats_kernel<T><<<(unsigned)p.planes, CTA, 0, st>>>(p);
            CNO_LAUNCH_CHECK();
            return 0;
        }
    }
    if (vec_ok && (int64_t)gs * VPT >= nvec) {
        p.gs = gs;
        const int threads = gs > CTA ? gs : CTA;
        const int ppc = threads / gs;
        const int64_t blocks = (p.planes + ppc - 1) / ppc;
        if constexpr (BWD) cin_bwd_reg_kernel<T, VPT><<<(unsigned)blocks, threads, 0, st>>>(p);
        else cin_fwd_reg_kernel<T, VPT><<<(unsigned)blocks, threads, 0, st>>>(p);
    } else if (vec_ok && (int64_t)8 * CLUSTER_CTA * VPT >= nvec && p.planes * 8 < (1ll << 31) && !env_no_cluster.on("CNO_CIN_NO_CLUSTER")) {
        // thread-block cluster of 2 / 4 / 8 CTAs per plane (portable cluster size limit: 8)
        int cs = 2;
        while ((int64_t)cs * CLUSTER_CTA * VPT < nvec) cs <<= 1;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(p.planes * cs)); cfg.blockDim = dim3(CLUSTER_CTA); cfg.dynamicSmemBytes = 0; cfg.stream = st;
        cudaLaunchAttribute attr;
        attr.id = cudaLaunchAttributeClusterDimension;
        attr.val.clusterDim.x = (unsigned)cs; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
        cfg.attrs = &attr; cfg.numAttrs = 1;
        cudaError_t e;
        if constexpr (BWD) e = cudaLaunchKernelEx(&cfg, cin_bwd_cluster_kernel<T, VPT>, p);
        else e = cudaLaunchKernelEx(&cfg, cin_fwd_cluster_kernel<T, VPT>, p);
        if (e != cudaSuccess) { cno_set_error("cin: cluster launch failed: %s", cudaGetErrorString(e)); return (int)e; }
    } else {
        cin_stream_kernel<T, BWD><<<(unsigned)p.planes, CTA, 0, st>>>(p);
    }
    CNO_LAUNCH_CHECK();
    return 0;
}

template <bool BWD>
int cin_dispatch(const CinParams& p, int dtype, cudaStream_t st) {
    switch (dtype) {
        case CNO_F32: return cin_launch<float, BWD>(p, st);
        case CNO_BF16: return cin_launch<__nv_bfloat16, BWD>(p, st);
        case CNO_F16: return cin_launch<__half, BWD>(p, st);
    }
    cno_set_error("cin: unsupported dtype %d", dtype);
    return (int)cudaErrorInvalidValue;
}

}  // namespace

extern "C" int cno_cin_fwd_res(const void* x, const void* residual, void* y, const float* gamma, const float* beta, const float* scale,
                               const float* shift, float* mean, float* rstd, int dtype, int n, int c, int hw,
                               float eps, void* stream);
extern "C" int cno_cin_fwd(const void* x, void* y, const float* gamma, const float* beta, const float* scale,
                           const float* shift, float* mean, float* rstd, int dtype, int n, int c, int hw,
                           float eps, void* stream) {
    return cno_cin_fwd_res(x, nullptr, y, gamma, beta, scale, shift, mean, rstd, dtype, n, c, hw, eps, stream);
}

extern "C" int cno_cin_fwd_res(const void* x, const void* residual, void* y, const float* gamma, const float* beta, const float* scale,
                               const float* shift, float* mean, float* rstd, int dtype, int n, int c, int hw,
                               float eps, void* stream) {
    CNO_CHECK_ARG(x && y && gamma && beta && scale && shift && mean && rstd, "cin_fwd: null pointer");
    CNO_CHECK_ARG(n >= 0 && c > 0 && hw > 0, "cin_fwd: bad sizes n=%d c=%d hw=%d", n, c, hw);
    if (n == 0) return 0;
    CinParams p{}; p.x = x; p.y = y; p.res = residual; p.gamma = gamma; p.beta = beta; p.scale = scale; p.shift = shift;
    p.mean = mean; p.rstd = rstd; p.planes = (int64_t)n * c; p.c = c; p.hw = hw; p.eps = eps;
    return cin_dispatch<false>(p, dtype, (cudaStream_t)stream);
}

// ---- allocation-free CIN forward -----------------------------------------------------------------------------------------------
// cno_cin_fwd leaves the per-plane statistics (mean, rstd: 2 x N*C floats) in caller memory: the streaming kernel and the
// statistics kernel write them and read them back, and the backward wants them.  A forward-only caller (inference, a serving
// loop, code under stream capture) has no use for them and may not be able to allocate them.  This variant keeps them in a
// fixed array that lives in the library's device image: every GPU that loads the library holds CIN_NOALLOC_BYTES of it until the
// process exits.  Nothing in it has to start at zero (each plane's two entries are written before they are read in the same
// launch), so it is never cleared and always holds the previous call's statistics.  All calls on one device share it: two calls
// that overlap on the GPU (different streams, no dependency between them) race on it.
constexpr int CIN_NOALLOC_MAX_PLANES = 65536;                                   // N * C of one call
constexpr size_t CIN_NOALLOC_BYTES = 2 * sizeof(float) * (size_t)CIN_NOALLOC_MAX_PLANES;   // 512 KB
static __device__ __align__(256) float g_cin_noalloc_stats[2 * CIN_NOALLOC_MAX_PLANES];

// device address of the array on the CURRENT device (host code cannot use the symbol itself as a pointer), looked up once per device
static int cin_noalloc_scratch(float** out) {
    static std::atomic<float*> addr[CNO_MAX_DEVICES];
    int d = -1;
    cudaError_t e = cudaGetDevice(&d);
    if (e != cudaSuccess) { cno_set_error("cin_fwd_noalloc: cudaGetDevice failed: %s", cudaGetErrorString(e)); return (int)e; }
    const bool cached = d >= 0 && d < CNO_MAX_DEVICES;
    float* ptr = cached ? addr[d].load(std::memory_order_acquire) : nullptr;
    if (!ptr) {
        e = cudaGetSymbolAddress((void**)&ptr, g_cin_noalloc_stats);
        if (e != cudaSuccess || !ptr) {
            cno_set_error("cin_fwd_noalloc: cudaGetSymbolAddress(statistics scratch) failed on device %d: %s", d, cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : (int)cudaErrorInvalidSymbol;
        }
        if (cached) addr[d].store(ptr, std::memory_order_release);
    }
    *out = ptr;
    return 0;
}

extern "C" int cno_cin_fwd_noalloc(const void* x, void* y, const float* gamma, const float* beta, const float* scale,
                                   const float* shift, int dtype, int n, int c, int hw, float eps, void* stream) {
    CNO_CHECK_ARG(x && y && gamma && beta && scale && shift, "cin_fwd_noalloc: null pointer");
    CNO_CHECK_ARG(n >= 0 && c > 0 && hw > 0, "cin_fwd_noalloc: bad sizes n=%d c=%d hw=%d", n, c, hw);
    CNO_CHECK_ARG((int64_t)n * c <= CIN_NOALLOC_MAX_PLANES,
                  "cin_fwd_noalloc: n*c = %lld planes exceed the built-in statistics scratch of %d planes (use cno_cin_fwd with caller buffers)",
                  (long long)n * c, CIN_NOALLOC_MAX_PLANES);
    if (n == 0) return 0;
    float* stats = nullptr;
    if (int rc = cin_noalloc_scratch(&stats)) return rc;
    CinParams p{}; p.x = x; p.y = y; p.gamma = gamma; p.beta = beta; p.scale = scale; p.shift = shift;
    p.mean = stats; p.rstd = stats + CIN_NOALLOC_MAX_PLANES; p.planes = (int64_t)n * c; p.c = c; p.hw = hw; p.eps = eps;
    return cin_dispatch<false>(p, dtype, (cudaStream_t)stream);
}

extern "C" int cno_cin_bwd(const void* x, const void* dy, void* dx, const float* gamma, const float* scale,
                           const float* mean, const float* rstd, float* s1, float* s2, int dtype, int n, int c,
                           int hw, void* stream) {
    CNO_CHECK_ARG(x && dy && dx && gamma && scale && mean && rstd && s1 && s2, "cin_bwd: null pointer");
    CNO_CHECK_ARG(n >= 0 && c > 0 && hw > 0, "cin_bwd: bad sizes n=%d c=%d hw=%d", n, c, hw);
    if (n == 0) return 0;
    CinParams p{}; p.x = x; p.dy = dy; p.y = dx; p.gamma = gamma; p.scale = scale;
    p.mean = const_cast<float*>(mean); p.rstd = const_cast<float*>(rstd); p.s1 = s1; p.s2 = s2;
    p.planes = (int64_t)n * c; p.c = c; p.hw = hw;
    return cin_dispatch<true>(p, dtype, (cudaStream_t)stream);
}

extern "C" int cno_film_fwd_res(const void* x, const void* residual, void* y, const float* t, const float* gamma, const float* beta,
                                const float* w1s, const float* b1s, const float* W2s, const float* b2s,
                                const float* w1b, const float* b1b, const float* W2b, const float* b2b,
                                float* scale_out, float* mean, float* rstd, int dtype, int n, int c, int hw, int lat,
                                float eps, void* stream);
extern "C" int cno_film_fwd(const void* x, void* y, const float* t, const float* gamma, const float* beta,
                            const float* w1s, const float* b1s, const float* W2s, const float* b2s,
                            const float* w1b, const float* b1b, const float* W2b, const float* b2b,
                            float* scale_out, float* mean, float* rstd, int dtype, int n, int c, int hw, int lat,
                            float eps, void* stream) {
    return cno_film_fwd_res(x, nullptr, y, t, gamma, beta, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b, scale_out, mean, rstd, dtype, n, c, hw,
                            lat, eps, stream);
}

extern "C" int cno_film_fwd_res(const void* x, const void* residual, void* y, const float* t, const float* gamma, const float* beta,
                                const float* w1s, const float* b1s, const float* W2s, const float* b2s,
                                const float* w1b, const float* b1b, const float* W2b, const float* b2b,
                                float* scale_out, float* mean, float* rstd, int dtype, int n, int c, int hw, int lat,
                                float eps, void* stream) {
    CNO_CHECK_ARG(x && y && t && gamma && beta && w1s && b1s && W2s && b2s && w1b && b1b && W2b && b2b && mean && rstd, "film_fwd: null pointer");
    CNO_CHECK_ARG(n >= 0 && c > 0 && hw > 0 && lat > 0, "film_fwd: bad sizes n=%d c=%d hw=%d lat=%d", n, c, hw, lat);
    if (n == 0) return 0;
    CinParams p{}; p.x = x; p.y = y; p.res = residual; p.gamma = gamma; p.beta = beta; p.mean = mean; p.rstd = rstd;
    p.planes = (int64_t)n * c; p.c = c; p.hw = hw; p.eps = eps;
    p.film = FilmW{t, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b, scale_out, lat};
    return cin_dispatch<false>(p, dtype, (cudaStream_t)stream);
}

// Statistics-only FiLM forward: per-plane mean / rstd / scale and the folded affine (aff_a, aff_d) with CIN(x) = aff_a * x + aff_d,
// no y written.  cno_filtered_lrelu_affine consumes the pair, so the normalised tensor never exists in HBM.
extern "C" int cno_film_stats(const void* x, const float* t, const float* gamma, const float* beta,
                              const float* w1s, const float* b1s, const float* W2s, const float* b2s,
                              const float* w1b, const float* b1b, const float* W2b, const float* b2b,
                              float* scale_out, float* mean, float* rstd, float* aff_a, float* aff_d,
                              int dtype, int n, int c, int hw, int lat, float eps, void* stream) {
    CNO_CHECK_ARG(x && t && gamma && beta && w1s && b1s && W2s && b2s && w1b && b1b && W2b && b2b && mean && rstd && aff_a && aff_d,
                  "film_stats: null pointer");
    CNO_CHECK_ARG(n >= 0 && c > 0 && hw > 0 && lat > 0, "film_stats: bad sizes n=%d c=%d hw=%d lat=%d", n, c, hw, lat);
    if (n == 0) return 0;
    CinParams p{}; p.x = x; p.gamma = gamma; p.beta = beta; p.mean = mean; p.rstd = rstd; p.aff_a = aff_a; p.aff_d = aff_d;
    p.planes = (int64_t)n * c; p.c = c; p.hw = hw; p.eps = eps;
    p.film = FilmW{t, w1s, b1s, W2s, b2s, w1b, b1b, W2b, b2b, scale_out, lat};
    return cin_dispatch<false>(p, dtype, (cudaStream_t)stream);
}

extern "C" int cno_film_bwd_finish(const float* s1, const float* s2, const float* t, const float* gamma, const float* beta,
                                   const float* scale, const float* w1s, const float* b1s, const float* W2s,
                                   const float* w1b, const float* b1b, const float* W2b,
                                   float* dgamma, float* dbeta, float* dW2s, float* db2s, float* dw1s, float* db1s,
                                   float* dW2b, float* db2b, float* dw1b, float* db1b, int n, int c, int lat, void* stream) {
    CNO_CHECK_ARG(s1 && s2 && t && gamma && beta && scale && w1s && b1s && W2s && w1b && b1b && W2b, "film_bwd_finish: null input");
    CNO_CHECK_ARG(dgamma && dbeta && dW2s && db2s && dw1s && db1s && dW2b && db2b && dw1b && db1b, "film_bwd_finish: null output");
    CNO_CHECK_ARG(n > 0 && c > 0 && lat > 0 && c <= 8192, "film_bwd_finish: bad sizes n=%d c=%d lat=%d", n, c, lat);
    FilmBwd p{s1, s2, t, gamma, beta, scale, w1s, b1s, W2s, w1b, b1b, W2b, dgamma, dbeta, dW2s, db2s, dw1s, db1s, dW2b, db2b, dw1b, db1b, n, c, lat};
    const int nb = (c + 7) / 8;
    film_bwd_finish_kernel<<<nb + (lat + 31) / 32, 256, (size_t)(4 * c + 8 * 32 * 4) * sizeof(float), (cudaStream_t)stream>>>(p);
    CNO_LAUNCH_CHECK();
    return 0;
}
