// error channel, version, and the FMA probe
#include "common.cuh"
#include <stdarg.h>

static thread_local char g_err[512] = "";

void cno_set_error(const char* fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}
extern "C" const char* cno_last_error(void) { return g_err; }
extern "C" int cno_version(void) { return 100; }

// ---- FP32 issue-rate probe: how fast does this B200 retire FFMA vs packed FFMA2? ---------------
template <int VARIANT>
__global__ void __launch_bounds__(256) probe_fma_kernel(float* out, int iters, float a, float b) {
    float2 acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (VARIANT == 0) { acc[i].x = fmaf(acc[i].x, a, b); acc[i].y = fmaf(acc[i].y, a, b); }
                else {
                    unsigned long long d, s = *reinterpret_cast<unsigned long long*>(&acc[i]);
                    float2 aa = make_float2(a, a), bb = make_float2(b, b);
                    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(s), "l"(*reinterpret_cast<unsigned long long*>(&aa)), "l"(*reinterpret_cast<unsigned long long*>(&bb)));
                    *reinterpret_cast<unsigned long long*>(&acc[i]) = d;
                }
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

extern "C" int cno_probe_fma(float* out, int variant, int blocks, int threads, int iters, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (variant == 0) probe_fma_kernel<0><<<blocks, threads, 0, st>>>(out, iters, 0.999f, 1e-3f);
    else probe_fma_kernel<1><<<blocks, threads, 0, st>>>(out, iters, 0.999f, 1e-3f);
    CNO_LAUNCH_CHECK();
    return 0;
}
