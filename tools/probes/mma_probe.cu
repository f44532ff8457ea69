// tcgen05.mma throughput probe for the banded-Toeplitz form of the vertical FIR passes: skinny-K bf16 MMAs,
// M = 128, N = 256, K = 80 (v-up: 64 + 6 input rows per 128 up-res rows, padded to 16) and K = 272 (v-down), operands in shared
// memory (no-swizzle K-major core-matrix layout, contents irrelevant: throughput only), accumulators in TMEM, one issuing thread.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_probe mma_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// shared-memory matrix descriptor, SWIZZLE_NONE: core matrices of 8 rows x 16 bytes; lbo = byte distance between core matrices
// adjacent in K, sbo = between core matrices adjacent in M / N
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;                       // descriptor version of sm_100
    return d;
}

__device__ __forceinline__ void mma_f16(uint32_t taddr, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        :: "r"(taddr), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

template <int KSTEPS>
__global__ void __launch_bounds__(128, 1) probe(long long* cyc, int* err, int tiles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tmem_base;
    __shared__ __align__(8) uint64_t bar;
    constexpr int M = 128, N = 256;
    uint8_t* A = smem;                                   // [KSTEPS][2 k-cores][M/8][128 B]
    uint8_t* B = smem + KSTEPS * M * 32;                 // [KSTEPS][2 k-cores][N/8][128 B]
    for (int i = threadIdx.x; i < KSTEPS * (M + N) * 32 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;   // finite bf16 pairs
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy writes of A / B visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    // instruction descriptor: D = F32 (bits 4-5 = 1), A = B = BF16 (bits 7-9, 10-12 = 1), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    long long t0 = 0, t1 = 0;
    if (threadIdx.x == 0) {
        const uint32_t a0 = smem_u32(A), b0 = smem_u32(B);
        t0 = clock64();
        for (int t = 0; t < tiles; t++) {
            const uint32_t taddr = tmem_base + (uint32_t)((t & 1) * 256);          // two accumulator stages
#pragma unroll
            for (int k = 0; k < KSTEPS; k++)
                mma_f16(taddr, make_desc(a0 + k * M * 32, M * 16, 128), make_desc(b0 + k * N * 32, N * 16, 128), idesc, k > 0);
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&bar)) : "memory");
        // bounded wait: a malformed descriptor must not hang the GPU box
        uint32_t done = 0;
        for (long long spin = 0; spin < 200000000ll && !done; spin++)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
        t1 = clock64();
        if (!done) *err = 1;
        cyc[blockIdx.x] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem_base));
}

template <int KSTEPS> void run(const char* name, long long* cyc, int* err) {
    const int tiles = 4000, blocks = 148;
    const size_t smem = (size_t)KSTEPS * (128 + 256) * 32;
    cudaFuncSetAttribute(probe<KSTEPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaMemset(err, 0, 4);
    probe<KSTEPS><<<blocks, 128, smem>>>(cyc, err, 16);
    cudaError_t e = cudaDeviceSynchronize();
    int herr = 0; cudaMemcpy(&herr, err, 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess || herr) { printf("%-40s warm-up failed: %s, timeout flag %d\n", name, cudaGetErrorString(e), herr); return; }
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<KSTEPS><<<blocks, 128, smem>>>(cyc, err, tiles);
    cudaEventRecord(e1);
    e = cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    cudaMemcpy(&herr, err, 4, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < blocks; i++) avg += (double)h[i]; avg /= blocks;
    const double mac = 128.0 * 256 * 16 * KSTEPS;
    printf("%-40s %8.1f cycles per tile (M128 N256 K%d) = %6.0f MAC/clk/SM; launch %.3f ms -> %.0f TFLOP/s chip  (%s, timeout flag %d)\n",
           name, avg / tiles, 16 * KSTEPS, mac / (avg / tiles), ms, 2.0 * mac * tiles * blocks / (ms * 1e-3) / 1e12, cudaGetErrorString(e), herr);
}


// ---- second probe (vbody_probe.cu follow-up): what does one M = 128 x N x K = 16 bf16 MMA cost as a function of N, of the B layout
// (K-major or MN-major no-swizzle core matrices) and of where A lives (shared memory descriptor or tensor memory)?
__device__ __forceinline__ void mma_f16_ta(uint32_t taddr, uint32_t a_taddr, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        :: "r"(taddr), "r"(a_taddr), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

template <int N, bool B_MN, bool A_TMEM>
__global__ void __launch_bounds__(128, 1) probe2(long long* cyc, int* err, int tiles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tmem_base;
    __shared__ __align__(8) uint64_t bar;
    constexpr int M = 128, KSTEPS = 4;
    uint8_t* A = smem;                                   // K-major: [KSTEPS][2 k-cores][M/8][128 B]
    uint8_t* B = smem + KSTEPS * M * 32;                 // K-major: [KSTEPS][2][N/8][128 B]; MN-major: [N/8][16 * KSTEPS rows][16 B]
    for (int i = threadIdx.x; i < KSTEPS * (M + N) * 32 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (B_MN ? (1u << 16) : 0u) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    if (threadIdx.x == 0) {
        const uint32_t a0 = smem_u32(A), b0 = smem_u32(B);
        const long long t0 = clock64();
        for (int t = 0; t < tiles; t++) {
            const uint32_t taddr = tmem_base + (uint32_t)((t & 1) * N);                 // accumulators at columns 0 .. 2N; A (tensor memory) at 480 ..
#pragma unroll
            for (int k = 0; k < KSTEPS; k++) {
                const uint64_t bdesc = B_MN ? make_desc(b0 + k * 256, 128, 16 * KSTEPS * 16) : make_desc(b0 + k * N * 32, N * 16, 128);
                if (A_TMEM) mma_f16_ta(taddr, tmem_base + 480 + 8 * k, bdesc, idesc, k > 0);
                else mma_f16(taddr, make_desc(a0 + k * M * 32, M * 16, 128), bdesc, idesc, k > 0);
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&bar)) : "memory");
        uint32_t done = 0;
        for (long long spin = 0; spin < 200000000ll && !done; spin++)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
        const long long t1 = clock64();
        if (!done) *err = 1;
        cyc[blockIdx.x] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem_base));
}

template <int N, bool B_MN, bool A_TMEM> void run2(long long* cyc, int* err) {
    const int tiles = 4000, blocks = 148;
    const size_t smem = (size_t)4 * (128 + N) * 32;
    cudaFuncSetAttribute(probe2<N, B_MN, A_TMEM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaMemset(err, 0, 4);
    probe2<N, B_MN, A_TMEM><<<blocks, 128, smem>>>(cyc, err, 16);
    cudaError_t e = cudaDeviceSynchronize();
    int herr = 0; cudaMemcpy(&herr, err, 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess || herr) { printf("N %3d B %s A %s: warm-up failed: %s, timeout flag %d\n", N, B_MN ? "MN" : "K ", A_TMEM ? "tmem" : "smem", cudaGetErrorString(e), herr); return; }
    probe2<N, B_MN, A_TMEM><<<blocks, 128, smem>>>(cyc, err, tiles);
    e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < blocks; i++) avg += (double)h[i]; avg /= blocks;
    const double per = avg / tiles / 4;
    printf("M 128 N %3d K 16, B %s-major, A in %s: %6.1f cycles per MMA = %5.0f MAC/clk/SM, operand bytes from shared memory %4.0f B/clk  (%s)\n",
           N, B_MN ? "MN" : "K ", A_TMEM ? "tensor memory" : "shared memory", per, 128.0 * N * 16 / per, ((A_TMEM ? 0 : 4096) + N * 32) / per, cudaGetErrorString(e));
}

int main() {
    long long* cyc; int* err; cudaMalloc(&cyc, 148 * 8); cudaMalloc(&err, 4);
    run<5>("v-up tile: K = 80", cyc, err);
    run<17>("v-down tile: K = 272", cyc, err);
    run<4>("K = 64", cyc, err);
    run2<256, false, false>(cyc, err); run2<256, true, false>(cyc, err); run2<256, false, true>(cyc, err); run2<256, true, true>(cyc, err);
    run2<144, false, false>(cyc, err); run2<144, true, false>(cyc, err); run2<144, false, true>(cyc, err); run2<144, true, true>(cyc, err);
    run2<64, false, false>(cyc, err); run2<64, true, false>(cyc, err); run2<64, true, true>(cyc, err);
    run2<32, true, false>(cyc, err); run2<32, true, true>(cyc, err); run2<80, true, false>(cyc, err); run2<80, true, true>(cyc, err);
    return 0;
}
