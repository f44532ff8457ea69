// Bring-up probe: 3-D TMA tile load with the same PTX wrappers / encode call as flrelu_fused_impl.cuh.
// usage: tma_probe W H NC bw bh bp c0 c1 c2 [l2promo]
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void k(const __grid_constant__ CUtensorMap tmap, float* out, int c0, int c1, int c2, int n, uint32_t tx) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + ((n * 4 + 127) / 128) * 128);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(tx) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
            ::"r"(smem_u32(smem)), "l"(&tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
    }
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(0) : "memory");
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = reinterpret_cast<float*>(smem)[i];
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv) {
    if (argc < 10) return 2;
    int W = atoi(argv[1]), H = atoi(argv[2]), NC = atoi(argv[3]), bw = atoi(argv[4]), bh = atoi(argv[5]), bp = atoi(argv[6]);
    int c0 = atoi(argv[7]), c1 = atoi(argv[8]), c2 = atoi(argv[9]);
    int promo = argc > 10 ? atoi(argv[10]) : 1;
    void* ptr = nullptr; cudaDriverEntryPointQueryResult q;
    cudaFree(0);
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) { printf("no entry point\n"); return 3; }
    EncodeTiledFn enc = (EncodeTiledFn)ptr;
    size_t ne = (size_t)W * H * NC; float* hx = (float*)malloc(ne * 4);
    for (size_t i = 0; i < ne; i++) hx[i] = (float)(i % 1000) + 1.f;
    float *x, *out; cudaMalloc(&x, ne * 4); cudaMemcpy(x, hx, ne * 4, cudaMemcpyHostToDevice);
    int n = bw * bh * bp; cudaMalloc(&out, n * 4); cudaMemset(out, 0xff, n * 4);
    CUtensorMap map;
    cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)NC};
    cuuint64_t strides[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
    cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, (cuuint32_t)bp}, es[3] = {1, 1, 1};
    CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, x, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 4; }
    size_t smem = ((n * 4 + 127) / 128) * 128 + 64;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<<<1, 128, smem>>>(map, out, c0, c1, c2, n, (uint32_t)(n * 4));
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("FAULT: %s\n", cudaGetErrorString(e)); return 1; }
    float* ho = (float*)malloc(n * 4); cudaMemcpy(ho, out, n * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int p = 0; p < bp; p++) for (int y = 0; y < bh; y++) for (int xx = 0; xx < bw; xx++) {
        int gx = c0 + xx, gy = c1 + y, gp = c2 + p; float want = 0.f;
        if (gx >= 0 && gx < W && gy >= 0 && gy < H && gp >= 0 && gp < NC) want = hx[((size_t)gp * H + gy) * W + gx];
        if (ho[(p * bh + y) * bw + xx] != want) bad++;
    }
    printf("ok, mismatches=%d of %d\n", bad, n);
    return bad ? 5 : 0;
}
