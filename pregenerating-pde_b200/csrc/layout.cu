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

// 2 x 2 blocks per thread: two loads of two consecutive columns from two consecutive source rows, a re-pairing, two stores of two
// consecutive destination columns - 32-bit accesses for 16-bit elements (128 bytes per warp request instead of 64), 64-bit accesses for
// 32-bit elements.  Tile: 64 x 64 elements; needs even R and S and word-aligned pointers.
template <class W> struct Pair;
template <> struct Pair<uint32_t> {                        // two 16-bit elements per word
    static __device__ __forceinline__ uint32_t lo(uint32_t a, uint32_t b) { return __byte_perm(a, b, 0x5410); }
    static __device__ __forceinline__ uint32_t hi(uint32_t a, uint32_t b) { return __byte_perm(a, b, 0x7632); }
};
template <> struct Pair<uint2> {                           // two 32-bit elements per word
    static __device__ __forceinline__ uint2 lo(uint2 a, uint2 b) { return make_uint2(a.x, b.x); }
    static __device__ __forceinline__ uint2 hi(uint2 a, uint2 b) { return make_uint2(a.y, b.y); }
};
template <class W>
__global__ void __launch_bounds__(256) transpose2x2_kernel(const W* __restrict__ in, W* __restrict__ out, int R, int S) {
    __shared__ W even[32][33], odd[32][33];                // [source column pair][source row pair]: columns 2j / 2j + 1 of the tile
    const size_t plane = (size_t)R * S / 2;                // in words of two elements
    in += blockIdx.z * plane; out += blockIdx.z * plane;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int s0 = blockIdx.x * 64, r0 = blockIdx.y * 64, Sw = S >> 1, Rw = R >> 1;
    W a[4], b[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int rp = ty + 8 * k, r = r0 + 2 * rp, sw = (s0 >> 1) + tx;
        const bool ok = r < R && sw < Sw;                  // R even: r < R implies r + 1 < R
        a[k] = ok ? in[(size_t)r * Sw + sw] : W{};
        b[k] = ok ? in[(size_t)(r + 1) * Sw + sw] : W{};
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int rp = ty + 8 * k;
        even[tx][rp] = Pair<W>::lo(a[k], b[k]);            // column 2 tx:     rows 2 rp, 2 rp + 1
        odd[tx][rp] = Pair<W>::hi(a[k], b[k]);             // column 2 tx + 1
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const int sl = ty + 8 * k;                         // output row (source column) within the tile, 0 .. 63
        const int sg = s0 + sl, rw = (r0 >> 1) + tx;
        if (sg < S && rw < Rw) out[(size_t)sg * Rw + rw] = (sl & 1) ? odd[sl >> 1][tx] : even[sl >> 1][tx];
    }
}

template <class W> int launch2x2(const void* in, void* out, int n, int R, int S, cudaStream_t st) {
    dim3 grid((unsigned)((S + 63) / 64), (unsigned)((R + 63) / 64), (unsigned)n);
    transpose2x2_kernel<W><<<grid, 256, 0, st>>>(static_cast<const W*>(in), static_cast<W*>(out), R, S);
    CNO_LAUNCH_CHECK();
    return 0;
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
    static EnvFlag env_f32_2x2;
    const bool even = !(R & 1) && !(S & 1) && (R + 63) / 64 <= 65535;
    const bool use2x2_f32 = env_f32_2x2.on("CNO_LAYOUT_F32_2X2");
    switch (dtype) {
        case CNO_F32:
            if (even && !(((uintptr_t)in | (uintptr_t)out) & 7u) && use2x2_f32) return launch2x2<uint2>(in, out, n, R, S, (cudaStream_t)stream);
            return launch<float>(in, out, n, R, S, (cudaStream_t)stream);
        case CNO_BF16: case CNO_F16:
            if (even && !(((uintptr_t)in | (uintptr_t)out) & 3u)) return launch2x2<uint32_t>(in, out, n, R, S, (cudaStream_t)stream);
            return launch<uint16_t>(in, out, n, R, S, (cudaStream_t)stream);
    }
    cno_set_error("layout: unsupported dtype %d", dtype);
    return (int)cudaErrorInvalidValue;
}

}  // namespace

extern "C" int cno_nchw_to_nhwc(const void* x, void* y, int dtype, int n, int c, int hw, void* stream) { return dispatch(x, y, dtype, n, c, hw, stream); }
extern "C" int cno_nhwc_to_nchw(const void* x, void* y, int dtype, int n, int c, int hw, void* stream) { return dispatch(x, y, dtype, n, hw, c, stream); }
