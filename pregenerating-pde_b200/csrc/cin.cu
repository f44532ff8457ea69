// CIN: time-conditioned instance norm (FILM.forward, CNO_timeModule_CIN.py:60-70) as ONE pass over x.
//   fwd:  mu, r = 1/sqrt(var+eps) per (n,c) plane;  y = (x-mu)*r*(gamma_c*s_nc) + (beta_c*s_nc + q_nc)
//   bwd:  S1 = sum dy, S2 = sum dy*xhat;  dx = gamma_c*s_nc*r*(dy - S1/HW - xhat*S2/HW)
// A plane is held in registers by GS threads (GS chosen on the host so that each thread keeps at most
// VPT 16-byte vectors); statistics use warp shuffles + one shared-memory hop per group, then the affine
// time conditioning is applied from registers, so HBM sees x exactly once (and dy once in backward).
// Planes that do not fit (hw not a multiple of the vector width, or > GS*VPT vectors) take the
// streaming path (three sweeps, second and third from L2).
#include "common.cuh"

namespace {

constexpr int CTA = 256;

// sum over the `gs` threads that share a plane.  gs is a power of two, 8 <= gs <= 1024.
// red: shared scratch of blockDim/32 floats per slot.
__device__ __forceinline__ float group_sum(float v, int gs, float* red, int slot) {
    const int lanes = gs < 32 ? gs : 32;
    for (int o = lanes >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (gs <= 32) return v;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nw_total = blockDim.x >> 5;
    const int wpg = gs >> 5;                 // warps per group
    float* buf = red + slot * nw_total;
    __syncthreads();                         // previous use of this slot is finished
    if (lane == 0) buf[warp] = v;
    __syncthreads();
    const int w0 = (warp / wpg) * wpg;
    float s = 0.f;
    for (int i = 0; i < wpg; i++) s += buf[w0 + i];
    return s;
}

struct CinParams {
    const void* x; const void* dy; void* y;
    const float *gamma, *beta, *scale, *shift;
    float *mean, *rstd, *s1, *s2;
    int64_t planes; int c, hw, gs; float eps;
};

template <class T, int VPT>
__global__ void __launch_bounds__(1024) cin_fwd_reg_kernel(CinParams p) {
    constexpr int V = 16 / sizeof(T);
    __shared__ float red[2 * 32];
    const int gs = p.gs, ppc = blockDim.x / gs;
    const int64_t plane = (int64_t)blockIdx.x * ppc + threadIdx.x / gs;
    const int t = threadIdx.x % gs;
    const bool live = plane < p.planes;
    const int nvec = p.hw / V;
    const T* xp = (const T*)p.x + (live ? plane : 0) * (int64_t)p.hw;
    Vec16<T> v[VPT];
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        if (live && i < nvec) {
            v[k] = ld16_stream(xp + (int64_t)i * V);
#pragma unroll
            for (int e = 0; e < V; e++) sum += to_f32(v[k].v[e]);
        }
    }
    const float inv_hw = 1.f / (float)p.hw;
    const float mu = group_sum(sum, gs, red, 0) * inv_hw;
    float sq = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        if (live && i < nvec) {
#pragma unroll
            for (int e = 0; e < V; e++) { float d = to_f32(v[k].v[e]) - mu; sq = fmaf(d, d, sq); }
        }
    }
    const float var = group_sum(sq, gs, red, 1) * inv_hw;
    const float r = rsqrtf(var + p.eps);
    if (!live) return;
    const int ch = (int)(plane % p.c);
    const float s = p.scale[plane], q = p.shift[plane];
    const float a = p.gamma[ch] * s * r;
    const float d0 = p.beta[ch] * s + q - mu * a;
    if (t == 0) { p.mean[plane] = mu; p.rstd[plane] = r; }
    T* yp = (T*)p.y + plane * (int64_t)p.hw;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        if (i < nvec) {
            Vec16<T> o;
#pragma unroll
            for (int e = 0; e < V; e++) o.v[e] = from_f32<T>(fmaf(to_f32(v[k].v[e]), a, d0));
            st16(yp + (int64_t)i * V, o);
        }
    }
}

template <class T, int VPT>
__global__ void __launch_bounds__(1024) cin_bwd_reg_kernel(CinParams p) {
    constexpr int V = 16 / sizeof(T);
    __shared__ float red[2 * 32];
    const int gs = p.gs, ppc = blockDim.x / gs;
    const int64_t plane = (int64_t)blockIdx.x * ppc + threadIdx.x / gs;
    const int t = threadIdx.x % gs;
    const bool live = plane < p.planes;
    const int nvec = p.hw / V;
    const int64_t base = (live ? plane : 0) * (int64_t)p.hw;
    const T* xp = (const T*)p.x + base; const T* gp = (const T*)p.dy + base;
    const float mu = live ? p.mean[plane] : 0.f, r = live ? p.rstd[plane] : 0.f;
    Vec16<T> vx[VPT], vg[VPT];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; k++) {
        const int i = t + k * gs;
        if (live && i < nvec) {
            vx[k] = ld16_stream(xp + (int64_t)i * V);
            vg[k] = ld16_stream(gp + (int64_t)i * V);
#pragma unroll
            for (int e = 0; e < V; e++) {
                const float g = to_f32(vg[k].v[e]);
                s1 += g; s2 = fmaf(g, (to_f32(vx[k].v[e]) - mu) * r, s2);
            }
        }
    }
    s1 = group_sum(s1, gs, red, 0);
    s2 = group_sum(s2, gs, red, 1);
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
            Vec16<T> o;
#pragma unroll
            for (int e = 0; e < V; e++) {
                const float xh = (to_f32(vx[k].v[e]) - mu) * r;
                o.v[e] = from_f32<T>(a * (to_f32(vg[k].v[e]) - m1 - xh * m2));
            }
            st16(dp + (int64_t)i * V, o);
        }
    }
}

// streaming path: one CTA per plane, any hw
template <class T, bool BWD>
__global__ void __launch_bounds__(CTA) cin_stream_kernel(CinParams p) {
    __shared__ float red[2 * 32];
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
        const float s = p.scale[plane];
        const float a = p.gamma[ch] * s * r, d0 = p.beta[ch] * s + p.shift[plane] - mu * a;
        if (threadIdx.x == 0) { p.mean[plane] = mu; p.rstd[plane] = r; }
        T* yp = (T*)p.y + base;
        for (int i = threadIdx.x; i < p.hw; i += CTA) yp[i] = from_f32<T>(fmaf(to_f32(xp[i]), a, d0));
    } else {
        const T* gp = (const T*)p.dy + base;
        const float mu = p.mean[plane], r = p.rstd[plane];
        float s1 = 0.f, s2 = 0.f;
        for (int i = threadIdx.x; i < p.hw; i += CTA) { float g = to_f32(gp[i]); s1 += g; s2 = fmaf(g, (to_f32(xp[i]) - mu) * r, s2); }
        s1 = group_sum(s1, CTA, red, 0); s2 = group_sum(s2, CTA, red, 1);
        if (threadIdx.x == 0) { p.s1[plane] = s1; p.s2[plane] = s2; }
        const float a = p.gamma[ch] * p.scale[plane] * r, m1 = s1 * inv_hw, m2 = s2 * inv_hw;
        T* dp = (T*)p.y + base;
        for (int i = threadIdx.x; i < p.hw; i += CTA) {
            const float xh = (to_f32(xp[i]) - mu) * r;
            dp[i] = from_f32<T>(a * (to_f32(gp[i]) - m1 - xh * m2));
        }
    }
}

template <class T, bool BWD>
int cin_launch(CinParams p, cudaStream_t st) {
    constexpr int V = 16 / sizeof(T);
    constexpr int VPT = BWD ? 4 : 8;
    const bool vec_ok = (p.hw % V == 0) && !(((uintptr_t)p.x | (uintptr_t)p.y | (uintptr_t)p.dy) & 15u);
    const int nvec = p.hw / V;
    int gs = 8;
    while (gs < 1024 && (int64_t)gs * VPT < nvec) gs <<= 1;
    if (vec_ok && (int64_t)gs * VPT >= nvec) {
        p.gs = gs;
        const int threads = gs > CTA ? gs : CTA;
        const int ppc = threads / gs;
        const int64_t blocks = (p.planes + ppc - 1) / ppc;
        if constexpr (BWD) cin_bwd_reg_kernel<T, VPT><<<(unsigned)blocks, threads, 0, st>>>(p);
        else cin_fwd_reg_kernel<T, VPT><<<(unsigned)blocks, threads, 0, st>>>(p);
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

extern "C" int cno_cin_fwd(const void* x, void* y, const float* gamma, const float* beta, const float* scale,
                           const float* shift, float* mean, float* rstd, int dtype, int n, int c, int hw,
                           float eps, void* stream) {
    CNO_CHECK_ARG(x && y && gamma && beta && scale && shift && mean && rstd, "cin_fwd: null pointer");
    CNO_CHECK_ARG(n >= 0 && c > 0 && hw > 0, "cin_fwd: bad sizes n=%d c=%d hw=%d", n, c, hw);
    if (n == 0) return 0;
    CinParams p{}; p.x = x; p.y = y; p.gamma = gamma; p.beta = beta; p.scale = scale; p.shift = shift;
    p.mean = mean; p.rstd = rstd; p.planes = (int64_t)n * c; p.c = c; p.hw = hw; p.eps = eps;
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
