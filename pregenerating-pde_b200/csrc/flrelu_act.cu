// filtered_lrelu_act: in-place gain * leaky-ReLU (+ clamp) on a dense [planes, h, w] tensor with the
// 2-bit sign mask written or read.  Only reached from the generic composition (the reference's
// filtered_lrelu.py:215-221 path); semantics follow filtered_lrelu.cu:1097-1203, kernel is new:
// each thread owns the four elements of one mask byte.
#include "common.cuh"

namespace {

struct ActParams {
    void* x; uint8_t* mask; int64_t planes; int h, w, mask_h, mask_w, mask_pitch, sx, sy;
    float gain, slope, clamp;
};

template <class T, int MODE>
__global__ void __launch_bounds__(256) flrelu_act_kernel(ActParams p) {
    const int quads = MODE == CNO_SIGN_WRITE ? p.mask_pitch : (p.w + 3) >> 2;   // WRITE also zeroes the pad bytes
    const int64_t total = p.planes * p.h * (int64_t)quads;
    const bool has_clamp = p.clamp >= 0.f && p.clamp < INFINITY;
    for (int64_t id = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; id < total; id += (int64_t)gridDim.x * blockDim.x) {
        const int qx = (int)(id % quads);
        const int64_t row = id / quads;             // plane * h + y
        const int yy = (int)(row % p.h);
        const int64_t plane = row / p.h;
        T* xp = (T*)p.x + row * (int64_t)p.w + qx * 4;
        const int nvalid = max(0, min(4, p.w - qx * 4));
        unsigned bits = 0;
        if (MODE == CNO_SIGN_READ) {
            const int my = yy + p.sy;
            if (my >= 0 && my < p.mask_h) {
                const uint8_t* mrow = p.mask + (plane * p.mask_h + my) * (int64_t)p.mask_pitch;
                const int e0 = qx * 4 + p.sx;          // first mask element (may be negative)
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int e = e0 + k;
                    if (e >= 0 && e < p.mask_w) bits |= ((mrow[e >> 2] >> ((e & 3) * 2)) & 3u) << (2 * k);
                }
            }
        }
        unsigned out_bits = 0;
        for (int k = 0; k < nvalid; k++) {
            float v = to_f32(xp[k]) * p.gain;
            if (MODE == CNO_SIGN_READ) {
                const unsigned s = (bits >> (2 * k)) & 3u;
                if (s & 1u) v *= p.slope;
                if (s & 2u) v = 0.f;
            } else {
                unsigned s = __float_as_uint(v) >> 31;
                if (s) v *= p.slope;
                if (has_clamp && fabsf(v) > p.clamp) { s = 2u; v = copysignf(p.clamp, v); }
                out_bits |= s << (2 * k);
            }
            xp[k] = from_f32<T>(v);
        }
        if (MODE == CNO_SIGN_WRITE) p.mask[(plane * p.mask_h + yy) * (int64_t)p.mask_pitch + qx] = (uint8_t)out_bits;
    }
}

template <class T>
int launch(const ActParams& p, int mode, cudaStream_t st) {
    const int64_t total = p.planes * p.h * (int64_t)(mode == CNO_SIGN_WRITE ? p.mask_pitch : (p.w + 3) >> 2);
    int64_t want = (total + 255) / 256;
    const int blocks = (int)(want < 1 ? 1 : (want > 148 * 32 ? 148 * 32 : want));
    if (mode == CNO_SIGN_WRITE) flrelu_act_kernel<T, CNO_SIGN_WRITE><<<blocks, 256, 0, st>>>(p);
    else if (mode == CNO_SIGN_READ) flrelu_act_kernel<T, CNO_SIGN_READ><<<blocks, 256, 0, st>>>(p);
    else flrelu_act_kernel<T, CNO_SIGN_NONE><<<blocks, 256, 0, st>>>(p);
    CNO_LAUNCH_CHECK();
    return 0;
}

}  // namespace

extern "C" int cno_filtered_lrelu_act(void* x, uint8_t* mask, int dtype, int64_t planes, int h, int w,
                                      int mask_h, int mask_w, int mask_pitch, int sx, int sy, int sign_mode,
                                      float gain, float slope, float clamp, void* stream) {
    CNO_CHECK_ARG(x, "filtered_lrelu_act: x is null");
    CNO_CHECK_ARG(sign_mode >= 0 && sign_mode <= 2, "filtered_lrelu_act: bad sign_mode %d", sign_mode);
    CNO_CHECK_ARG(sign_mode == CNO_SIGN_NONE || mask, "filtered_lrelu_act: mask required for sign_mode %d", sign_mode);
    if (sign_mode == CNO_SIGN_WRITE)
        CNO_CHECK_ARG(mask_h >= h && mask_pitch * 4 >= w, "filtered_lrelu_act: mask [%d x %d bytes] too small for [%d x %d]", mask_h, mask_pitch, h, w);
    if (planes <= 0 || h <= 0 || w <= 0) return 0;
    ActParams p{x, mask, planes, h, w, mask_h, mask_w, mask_pitch, sx, sy, gain, slope, clamp};
    switch (dtype) {
        case CNO_F32: return launch<float>(p, sign_mode, (cudaStream_t)stream);
        case CNO_BF16: return launch<__nv_bfloat16>(p, sign_mode, (cudaStream_t)stream);
        case CNO_F16: return launch<__half>(p, sign_mode, (cudaStream_t)stream);
    }
    cno_set_error("filtered_lrelu_act: unsupported dtype %d", dtype);
    return (int)cudaErrorInvalidValue;
}
