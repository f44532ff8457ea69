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

// 16-bit elements: a thread moves 2 x 2 blocks - two 32-bit loads (two consecutive source rows, two consecutive columns each), two
// byte permutes, two 32-bit stores - so both sides move 128 bytes per warp request like the fp32 kernel (the element-wise version
// moves 64).  Tile: 64 x 64 elements; needs even R and S (rows stay 4-byte aligned).
__global__ void __launch_bounds__(256) transpose16_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int R, int S) {
    __shared__ uint32_t even[32][33], odd[32][33];        // [source column pair][source row pair]: columns 2j / 2j + 1 of the tile
    const size_t plane = (size_t)R * S / 2;                // in 32-bit words
    in += blockIdx.z * plane; out += blockIdx.z * plane;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int s0 = blockIdx.x * 64, r0 = blockIdx.y * 64, Sw = S >> 1, Rw = R >> 1;
    uint32_t a[4], b[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int rp = ty + 8 * k, r = r0 + 2 * rp, sw = (s0 >> 1) + tx;
        const bool ok = r < R && sw < Sw;                  // R even: r < R implies r + 1 < R
        a[k] = ok ? in[(size_t)r * Sw + sw] : 0u;
        b[k] = ok ? in[(size_t)(r + 1) * Sw + sw] : 0u;
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int rp = ty + 8 * k;
        even[tx][rp] = __byte_perm(a[k], b[k], 0x5410);   // column 2 tx:     rows 2 rp, 2 rp + 1
        odd[tx][rp] = __byte_perm(a[k], b[k], 0x7632);    // column 2 tx + 1
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const int sl = ty + 8 * k;                         // output row (source column) within the tile, 0 .. 63
        const int sg = s0 + sl, rw = (r0 >> 1) + tx;
        if (sg < S && rw < Rw) out[(size_t)sg * Rw + rw] = (sl & 1) ? odd[sl >> 1][tx] : even[sl >> 1][tx];
    }
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
        case CNO_BF16: case CNO_F16:
            if (!(R & 1) && !(S & 1) && !(((uintptr_t)in | (uintptr_t)out) & 3u) && (R + 63) / 64 <= 65535) {
                dim3 grid((unsigned)((S + 63) / 64), (unsigned)((R + 63) / 64), (unsigned)n);
                transpose16_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(static_cast<const uint32_t*>(in), static_cast<uint32_t*>(out), R, S);
                CNO_LAUNCH_CHECK();
                return 0;
            }
            return launch<uint16_t>(in, out, n, R, S, (cudaStream_t)stream);
    }
    cno_set_error("layout: unsupported dtype %d", dtype);
    return (int)cudaErrorInvalidValue;
}

}  // namespace

extern "C" int cno_nchw_to_nhwc(const void* x, void* y, int dtype, int n, int c, int hw, void* stream) { return dispatch(x, y, dtype, n, c, hw, stream); }
extern "C" int cno_nhwc_to_nchw(const void* x, void* y, int dtype, int n, int c, int hw, void* stream) { return dispatch(x, y, dtype, n, hw, c, stream); }
