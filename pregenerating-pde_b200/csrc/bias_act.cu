// bias_act: y = clamp(act(x + b) * gain) and its first / second order gradients.
// Replaces the reference plugin bias_act.cu:15-139 / bias_act.cpp:24-82 (semantics restated, new
// kernel): 16-byte vectorised, one launch, fp32 / bf16 / fp16 / fp64 storage.
#include "common.cuh"
#include <math.h>

namespace {

template <class S> struct ActMath;   // S = compute type (float for 16/32-bit storage, double for fp64)

// f(u), f'(.) and f''(.) expressed in the quantity the reference keeps for backward
// (bias_act.py:13-23 `ref`): y/gain for everything except swish, which keeps x.
template <int ACT, int GRAD, class S>
__device__ __forceinline__ S act_eval(S g, S u /*x+b or xref+b*/, S yy /*yref/gain*/, S alpha) {
    const S one = (S)1, two = (S)2;
    const S selu_s = (S)1.0507009873554804934193349852946, selu_a = (S)1.6732632423543772848170429916717;
    if (ACT == 1) { return (GRAD == 2) ? (S)0 : g; }
    if (ACT == 2) { if (GRAD == 0) return g > 0 ? g : (S)0; if (GRAD == 1) return yy > 0 ? g : (S)0; return (S)0; }
    if (ACT == 3) { if (GRAD == 0) return g > 0 ? g : g * alpha; if (GRAD == 1) return yy > 0 ? g : g * alpha; return (S)0; }
    if (ACT == 4) { if (GRAD == 0) return tanh(g); S d1 = one - yy * yy; if (GRAD == 1) return g * d1; return g * d1 * (-two * yy); }
    if (ACT == 5) { if (GRAD == 0) return one / (one + exp(-g)); S d1 = yy * (one - yy); if (GRAD == 1) return g * d1; return g * d1 * (one - two * yy); }
    if (ACT == 6) { if (GRAD == 0) return g >= 0 ? g : expm1(g); if (GRAD == 1) return yy >= 0 ? g : g * (yy + one); return yy >= 0 ? (S)0 : g * (yy + one); }
    if (ACT == 7) { const S sa = selu_s * selu_a;
                    if (GRAD == 0) return g >= 0 ? selu_s * g : sa * expm1(g);
                    if (GRAD == 1) return yy >= 0 ? g * selu_s : g * (yy + sa); return yy >= 0 ? (S)0 : g * (yy + sa); }
    if (ACT == 8) { if (GRAD == 0) return g > (S)20 ? g : log1p(exp(g)); S e = exp(-yy); if (GRAD == 1) return g * (one - e); return g * e * (one - e); }
    if (ACT == 9) { if (GRAD == 0) return g / (one + exp(-g));
                    S sg = one / (one + exp(-u));               // sigmoid(x)
                    S d1 = sg * (one + u * (one - sg));          // swish'
                    if (GRAD == 1) return g * d1;
                    S d2 = sg * (one - sg) * (two + u * (one - two * sg));
                    return g * d2; }
    return (S)0;
}

template <class T> struct Compute { typedef float type; };
template <> struct Compute<double> { typedef double type; };
template <class T> __device__ __forceinline__ typename Compute<T>::type ldc(T v) { return to_f32(v); }
template <> __device__ __forceinline__ double ldc<double>(double v) { return v; }
template <class T, class S> __device__ __forceinline__ T stc(S v) { return from_f32<T>((float)v); }
template <> __device__ __forceinline__ double stc<double, double>(double v) { return v; }

struct BAParams {
    const void *x, *b, *xref, *yref, *dy; void* y;
    int64_t numel, size_b, step_b; float alpha, gain, clamp;
};

template <class T, int ACT, int GRAD>
__global__ void __launch_bounds__(256) bias_act_kernel(BAParams p) {
    typedef typename Compute<T>::type S;
    constexpr int V = 16 / sizeof(T);
    const T* x = (const T*)p.x; const T* b = (const T*)p.b; const T* xr = (const T*)p.xref;
    const T* yr = (const T*)p.yref; const T* dyp = (const T*)p.dy; T* y = (T*)p.y;
    const S alpha = (S)p.alpha, gain = (S)p.gain, clamp = (S)p.clamp;
    const S inv_gain = gain != (S)0 ? (S)1 / gain : (S)0;
    const int64_t nvec = p.numel / V;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    auto one = [&](S g, S bb, S xrv, S yrv, S dyv) -> S {
        S u = (GRAD == 0 ? g : xrv) + bb;
        S gin = GRAD == 0 ? u : g;
        S r = act_eval<ACT, GRAD, S>(gin, u, yrv * inv_gain, alpha) * gain * dyv;
        if (clamp >= (S)0) {
            if (GRAD == 0) r = r > clamp ? clamp : (r < -clamp ? -clamp : r);
            else {
                S yref_v = yrv;
                if (ACT == 9) yref_v = (u / ((S)1 + exp(-u))) * gain;   // swish keeps x, rebuild y
                r = (yref_v > -clamp && yref_v < clamp) ? r : (S)0;
            }
        }
        return r;
    };
    for (int64_t vi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; vi < nvec; vi += stride) {
        const int64_t i0 = vi * V;
        Vec16<T> vx = ld16_stream(x + i0), vxr, vyr, vdy, vo;
        if (xr) vxr = ld16_stream(xr + i0);
        if (yr) vyr = ld16_stream(yr + i0);
        if (dyp) vdy = ld16_stream(dyp + i0);
        // bias index: one 32-bit division per 16-byte vector when the whole vector lies in one bias run (step_b % V == 0,
        // the NCHW case); per element otherwise (channels_last: step_b == 1)
        const bool need_b = b != nullptr && (GRAD == 0 || ACT == 9);      // only swish reads x + b again in its gradients
        const uint32_t sb = (uint32_t)p.step_b, nb = (uint32_t)p.size_b;
        const bool uni = (sb % V) == 0;
        S bvec = (S)0;
        if (need_b && uni) bvec = ldc(b[((uint32_t)i0 / sb) % nb]);
#pragma unroll
        for (int k = 0; k < V; k++) {
            S bb = need_b ? (uni ? bvec : ldc(b[(((uint32_t)i0 + k) / sb) % nb])) : (S)0;
            vo.v[k] = stc<T, S>(one(ldc(vx.v[k]), bb, xr ? ldc(vxr.v[k]) : (S)0, yr ? ldc(vyr.v[k]) : (S)0, dyp ? ldc(vdy.v[k]) : (S)1));
        }
        st16(y + i0, vo);
    }
    // tail (numel not a multiple of the vector width)
    for (int64_t i = nvec * V + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.numel; i += stride) {
        S bb = b ? ldc(b[(i / p.step_b) % p.size_b]) : (S)0;
        y[i] = stc<T, S>(one(ldc(x[i]), bb, xr ? ldc(xr[i]) : (S)0, yr ? ldc(yr[i]) : (S)0, dyp ? ldc(dyp[i]) : (S)1));
    }
}

template <class T, int ACT>
int launch_grad(const BAParams& p, int grad, int blocks, cudaStream_t st) {
    switch (grad) {
        case 0: bias_act_kernel<T, ACT, 0><<<blocks, 256, 0, st>>>(p); break;
        case 1: bias_act_kernel<T, ACT, 1><<<blocks, 256, 0, st>>>(p); break;
        default: bias_act_kernel<T, ACT, 2><<<blocks, 256, 0, st>>>(p); break;
    }
    CNO_LAUNCH_CHECK();
    return 0;
}

template <class T>
int launch_act(const BAParams& p, int act, int grad, int blocks, cudaStream_t st) {
    switch (act) {
        case 1: return launch_grad<T, 1>(p, grad, blocks, st);
        case 2: return launch_grad<T, 2>(p, grad, blocks, st);
        case 3: return launch_grad<T, 3>(p, grad, blocks, st);
        case 4: return launch_grad<T, 4>(p, grad, blocks, st);
        case 5: return launch_grad<T, 5>(p, grad, blocks, st);
        case 6: return launch_grad<T, 6>(p, grad, blocks, st);
        case 7: return launch_grad<T, 7>(p, grad, blocks, st);
        case 8: return launch_grad<T, 8>(p, grad, blocks, st);
        default: return launch_grad<T, 9>(p, grad, blocks, st);
    }
}

}  // namespace

extern "C" int cno_bias_act(const void* x, const void* b, const void* xref, const void* yref, const void* dy,
                            void* y, int dtype, int64_t numel, int64_t size_b, int64_t step_b, int grad,
                            int act, float alpha, float gain, float clamp, void* stream) {
    CNO_CHECK_ARG(x && y, "bias_act: x and y must be given");
    CNO_CHECK_ARG(act >= 1 && act <= 9, "bias_act: act index %d out of range 1..9", act);
    CNO_CHECK_ARG(grad >= 0 && grad <= 2, "bias_act: grad must be 0, 1 or 2");
    CNO_CHECK_ARG(dtype >= CNO_F32 && dtype <= CNO_F64, "bias_act: bad dtype %d", dtype);
    CNO_CHECK_ARG(!b || (size_b > 0 && step_b > 0), "bias_act: bias needs size_b, step_b > 0");
    if (numel <= 0) return 0;
    CNO_CHECK_ARG(numel <= 0x7fffffffll, "bias_act: x is too large");      // 32-bit index maths (bias_act.cpp:37)
    const int es = cno_dtype_size(dtype);
    auto mis = [&](const void* q) { return q && ((uintptr_t)q & 15u); };
    CNO_CHECK_ARG(!(mis(x) || mis(xref) || mis(yref) || mis(dy) || mis(y)),
                  "bias_act: tensor base pointers must be 16-byte aligned");
    BAParams p{x, b, xref, yref, dy, y, numel, b ? size_b : 1, b ? step_b : 1, alpha, gain, clamp};
    // grid: enough 256-thread CTAs for one 16-byte vector per thread, capped at 16 resident waves
    // of the 148 SMs (grid-stride loop beyond that)
    int64_t want = (numel / (16 / es) + 255) / 256;
    int blocks = (int)(want < 1 ? 1 : (want > 148 * 16 ? 148 * 16 : want));
    cudaStream_t st = (cudaStream_t)stream;
    switch (dtype) {
        case CNO_F32: return launch_act<float>(p, act, grad, blocks, st);
        case CNO_BF16: return launch_act<__nv_bfloat16>(p, act, grad, blocks, st);
        case CNO_F16: return launch_act<__half>(p, act, grad, blocks, st);
        default: return launch_act<double>(p, act, grad, blocks, st);
    }
}
