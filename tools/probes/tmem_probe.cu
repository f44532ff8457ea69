// TMEM read-bandwidth probe (tensor-core question for the 16-bit path): a banded-Toeplitz tcgen05 version of
// the two vertical FIR passes would have to pull every up-sampled pixel (fp32 accumulator) out of TMEM with tcgen05.ld, apply the
// leaky ReLU / sign bits in registers and write bf16 back to shared memory for the second MMA.  How fast is that path?
// Measures tcgen05.ld (32x32b.x32: a warp reads its 32 lanes x 32 columns = 4 KB per instruction) with 4 warps per CTA (the
// four TMEM lane quarters), alone, with the epilogue arithmetic (max + sign pack + bf16 pack), and with the st.shared of the result.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_probe tmem_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

#define LD32(r, addr) asm volatile( \
    "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];" \
    : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),"=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15]), \
      "=r"(r[16]),"=r"(r[17]),"=r"(r[18]),"=r"(r[19]),"=r"(r[20]),"=r"(r[21]),"=r"(r[22]),"=r"(r[23]),"=r"(r[24]),"=r"(r[25]),"=r"(r[26]),"=r"(r[27]),"=r"(r[28]),"=r"(r[29]),"=r"(r[30]),"=r"(r[31]) \
    : "r"(addr))

// MODE 0: ld only | 1: ld + lrelu + sign bits + bf16 pack (registers) | 2: mode 1 + st.shared of the packed row segment
template <int MODE>
__global__ void __launch_bounds__(128, 1) probe(float* out, long long* cyc, int iters) {
    __shared__ uint32_t tmem_base;
    __shared__ __align__(16) uint32_t stage[128 * 20];         // 128 rows x 32 bf16 (+ pad): one column block of the B operand of the second MMA
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" :: "r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t taddr = tmem_base + ((uint32_t)(32 * warp) << 16);      // this warp's lane quarter
    float acc = 0.f; uint32_t bits = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll 1
        for (int c = 0; c < 256; c += 32) {                                // one 128 x 256 fp32 accumulator tile per iteration
            uint32_t r[32];
            LD32(r, taddr + c);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (MODE == 0) {
#pragma unroll
                for (int i = 0; i < 32; i++) bits ^= r[i];
            } else {
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    float a = __uint_as_float(r[i]), b = __uint_as_float(r[i + 1]);
                    bits = (bits << 2) | ((r[i] >> 31) << 1) | (r[i + 1] >> 31);
                    a = fmaxf(a, 0.2f * a); b = fmaxf(b, 0.2f * b);
                    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
                    pk[i >> 1] = *reinterpret_cast<uint32_t*>(&h);
                }
                if (MODE == 2) {
                    uint4* dst = reinterpret_cast<uint4*>(stage + (32 * warp + lane) * 20);
#pragma unroll
                    for (int i = 0; i < 4; i++) dst[i] = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
                } else {
#pragma unroll
                    for (int i = 0; i < 16; i++) bits ^= pk[i];
                }
            }
        }
    }
    const long long t1 = clock64();
    acc += __uint_as_float(bits & 0x3fffffffu) + (MODE == 2 ? __uint_as_float(stage[threadIdx.x] & 0x3fffffffu) : 0.f);
    out[blockIdx.x * 128 + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" :: "r"(tmem_base));
}

template <int MODE> void run(const char* name, float* out, long long* cyc) {
    const int iters = 2000, blocks = 148;
    probe<MODE><<<blocks, 128>>>(out, cyc, 10);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<MODE><<<blocks, 128>>>(out, cyc, iters);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < blocks; i++) avg += (double)h[i]; avg /= blocks;
    const double bytes = 128.0 * 256 * 4;                                   // per iteration and SM
    printf("%-46s %8.1f cycles per 128x256 fp32 tile = %6.1f B/clk/SM; whole launch %.3f ms -> %.0f GB/s per SM, %.1f TB/s chip  (%s)\n",
           name, avg / iters, bytes / (avg / iters), ms, bytes * iters / (ms * 1e-3) / 1e9, 148 * bytes * iters / (ms * 1e-3) / 1e12, cudaGetErrorString(e));
}

int main() {
    float* out; long long* cyc; cudaMalloc(&out, 148 * 128 * 4); cudaMalloc(&cyc, 148 * 8);
    run<0>("tcgen05.ld 32x32b.x32 only", out, cyc);
    run<1>("ld + lrelu + sign bits + bf16 pack", out, cyc);
    run<2>("ld + lrelu + sign + pack + st.shared (bf16)", out, cyc);
    printf("reference: 148 M up-res fp32 accumulators of [32,64,128,128] = 593 MB through this path\n");
    return 0;
}
