// NCHW <-> NHWC transposes around the cuDNN convolutions of the CNO blocks (CNO_timeModule_CIN.py:329-335, :495-510: conv -> FILM ->
// LReLu -> conv).  cuDNN's tensor-op convolution kernels are NHWC inside; fed NCHW tensors they convert every operand of every pass
// themselves (six activation-sized transposes per convolution and training step).  The host side (layout.py) makes the conversion
// explicit - x once (kept for the weight gradient), dy once, y and dx back - and these kernels do it: per sample a [R][S] -> [S][R]
// matrix transpose through a 32 x 33 shared-memory tile, 128-byte coalesced on both sides.  HBM-bound: 4 B read + 4 B written per element.
#include "common.cuh"

namespace {

// [R][S] -> [S][R] per sample; a CTA moves a TR x TS tile.  The wide tiles (128 along one dimension) keep 16 independent 4-byte loads
// per thread in flight before the barrier: 4.3 -> TB/s-class on the 128^2 activations (tools/time_layout.py).
template <class T, int TS, int TR>
__global__ void __launch_bounds__(256) transpose_kernel(const T* __restrict__ in, T* __restrict__ out, int R, int S) {
    __shared__ T tile[TR][TS + 1];
    const size_t plane = (size_t)R * S;
    in += blockIdx.z * plane; out += blockIdx.z * plane;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;                 // 32 x 8 threads
    const int s0 = blockIdx.x * TS, r0 = blockIdx.y * TR;
#pragma unroll
    for (int j = ty; j < TR; j += 8)
#pragma unroll
        for (int i = tx; i < TS; i += 32)
            if (s0 + i < S && r0 + j < R) tile[j][i] = in[(size_t)(r0 + j) * S + s0 + i];
    __syncthreads();
#pragma unroll
    for (int j = ty; j < TS; j += 8)
#pragma unroll
        for (int i = tx; i < TR; i += 32)
            if (r0 + i < R && s0 + j < S) out[(size_t)(s0 + j) * R + r0 + i] = tile[i][j];
}

template <class T, int TS, int TR> int launch_tile(const T* in, T* out, int n, int R, int S, cudaStream_t st) {
    dim3 grid((unsigned)((S + TS - 1) / TS), (unsigned)((R + TR - 1) / TR), (unsigned)n);
    transpose_kernel<T, TS, TR><<<grid, 256, 0, st>>>(in, out, R, S);
    CNO_LAUNCH_CHECK();
    return 0;
}

template <class T> int launch(const void* in, void* out, int n, int R, int S, cudaStream_t st) {
    const T* i = static_cast<const T*>(in); T* o = static_cast<T*>(out);
    if (S >= 128) return launch_tile<T, 128, 32>(i, o, n, R, S, st);
    if (R >= 128) return launch_tile<T, 32, 128>(i, o, n, R, S, st);
    return launch_tile<T, 32, 32>(i, o, n, R, S, st);
}

int dispatch(const void* in, void* out, int dtype, int n, int R, int S, void* stream) {
    CNO_CHECK_ARG(in && out, "layout: null pointer");
    CNO_CHECK_ARG(n >= 0 && R > 0 && S > 0 && n <= 65535 && (R + 31) / 32 <= 65535, "layout: bad sizes n=%d %d x %d", n, R, S);
    if (n == 0) return 0;
    switch (dtype) {
        case CNO_F32: return launch<float>(in, out, n, R, S, (cudaStream_t)stream);
        case CNO_BF16: case CNO_F16: return launch<uint16_t>(in, out, n, R, S, (cudaStream_t)stream);
    }
    cno_set_error("layout: unsupported dtype %d", dtype);
    return (int)cudaErrorInvalidValue;
}

}  // namespace

extern "C" int cno_nchw_to_nhwc(const void* x, void* y, int dtype, int n, int c, int hw, void* stream) { return dispatch(x, y, dtype, n, c, hw, stream); }
extern "C" int cno_nhwc_to_nchw(const void* x, void* y, int dtype, int n, int c, int hw, void* stream) { return dispatch(x, y, dtype, n, hw, c, stream); }
