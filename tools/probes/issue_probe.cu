// Issue-slot probe for the fused FIR kernel's instruction diet: how many cycles does a scheduler spend on
// packed fma.rn.f32x2 (SASS FFMA2) alone and mixed with ALU-pipe instructions (SHF / LOP3 / FMNMX) or scalar FFMA?
// Decides whether the non-FMA instructions of the streaming body cost issue slots of their own (sum model) or hide in the
// second pipe cycle of an FFMA2 (max model).   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o issue_probe issue_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

struct f2 { float x, y; };
__device__ __forceinline__ f2 ffma2(f2 a, f2 b, f2 c) {
    f2 d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(*reinterpret_cast<unsigned long long*>(&d))
        : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c)));
    return d;
}
__device__ __forceinline__ float ffma(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ uint32_t shf(uint32_t a, uint32_t b) { uint32_t d; asm volatile("shf.l.wrap.b32 %0, %1, %2, 2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ float fmx(float a, float b) { float d; asm volatile("max.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }

struct Taps { float t[16]; };

// V: 0 FFMA2 reg-pair tap | 1 FFMA2 uniform tap | 2 = 1 + SHF 1:1 | 3 = 1 + SHF 1:2 | 4 = 1 + FMNMX 1:1 | 5 scalar FFMA uniform tap (2 per FFMA2 slot)
//    12 scalar FFMA acc = xr * uniform + acc | 13 = 12 + SHF per 2 FFMA | 14 scalar FFMA 3-reg | 15 = 12 + STS.U8 per 4 FFMA | 16 = 14 + SHF per 2
//    6 = 5 + SHF per 2 FFMA | 7 = 0 + SHF 1:1 | 8 = 1 + IMAD 1:1 | 9 = SHF only (2 per slot) | 10 = 1 + SHF 2:1 | 11 = 1 + LDS 1:4
template <int V>
__global__ void __launch_bounds__(512, 1) probe(const __grid_constant__ Taps tp, float* out, long long* cyc, int iters) {
    __shared__ float sm[1024];
    sm[threadIdx.x] = threadIdx.x; sm[threadIdx.x + 512] = 1.f;
    __syncthreads();
    f2 acc[8]; uint32_t r[8]; float s[8];
    for (int i = 0; i < 8; i++) { acc[i].x = threadIdx.x * 0.001f + i; acc[i].y = i * 0.5f; r[i] = threadIdx.x * 2654435761u + i; s[i] = i + threadIdx.x; }
    f2 x; x.x = 1.0001f; x.y = 0.9999f;
    f2 treg[8];
    for (int i = 0; i < 8; i++) { treg[i].x = tp.t[i] + threadIdx.x * 1e-9f; treg[i].y = tp.t[i + 8] + (V >= 17 ? threadIdx.x * 1e-9f : 0.f); }
    long long t0 = clock64();
    #pragma unroll 1
    for (int it = 0; it < iters; it++) {
        #pragma unroll
        for (int k = 0; k < 8; k++) {
            #pragma unroll
            for (int i = 0; i < 8; i++) {
                if (V == 0 || V == 7) acc[i] = ffma2(acc[i], treg[k], x);
                else if (V == 5 || V == 6) { acc[i].x = ffma(acc[i].x, tp.t[k], x.x); acc[i].y = ffma(acc[i].y, tp.t[k], x.y); }
                else if (V != 9 && V < 12) { f2 t; t.x = tp.t[k]; t.y = tp.t[k]; acc[i] = ffma2(acc[i], t, x); }
                if (V == 17) { f2 t; t.x = tp.t[k]; t.y = tp.t[k]; acc[i] = ffma2(treg[(i + k) & 7], t, acc[i]); }
                if (V == 18) { f2 t; t.x = tp.t[k]; t.y = tp.t[k + 8]; f2 xb; xb.x = treg[(i + k) & 7].x; xb.y = xb.x; acc[i] = ffma2(xb, t, acc[i]); }
                if (V == 19) { f2 t; t.x = tp.t[k]; t.y = tp.t[k]; acc[i] = ffma2(treg[k], t, acc[i]); }
                if (V == 12 || V == 13 || V == 15) { acc[i].x = ffma(treg[i].x, tp.t[k], acc[i].x); acc[i].y = ffma(treg[i].y, tp.t[k + 8], acc[i].y); }
                if (V == 14 || V == 16) { acc[i].x = ffma(treg[i].x, treg[k].y, acc[i].x); acc[i].y = ffma(treg[i].y, treg[k].x, acc[i].y); }
                if (V == 13 || V == 16) r[i] = shf(r[i], r[(i + 1) & 7]);
                if (V == 15 && (i & 1)) reinterpret_cast<volatile unsigned char*>(sm)[threadIdx.x + 512 * (i >> 1)] = (unsigned char)r[i];
                if (V == 2 || V == 7) r[i] = shf(r[i], r[(i + 1) & 7]);
                if (V == 3 && (i & 1)) r[i] = shf(r[i], r[(i + 1) & 7]);
                if (V == 4) s[i] = fmx(s[i], acc[(i + 3) & 7].x);
                if (V == 6) r[i] = shf(r[i], r[(i + 1) & 7]);
                if (V == 8) r[i] = imad(r[i], r[(i + 1) & 7], r[(i + 2) & 7]);
                if (V == 9) { r[i] = shf(r[i], r[(i + 1) & 7]); r[(i + 4) & 7] = shf(r[(i + 4) & 7], r[(i + 5) & 7]); }
                if (V == 10) { r[i] = shf(r[i], r[(i + 1) & 7]); r[(i + 4) & 7] = shf(r[(i + 4) & 7], r[(i + 5) & 7]); }
                if (V == 11 && (i & 3) == 0) s[i] += sm[(r[i] + k) & 1023];
            }
        }
    }
    long long t1 = clock64();
    float o = 0.f;
    for (int i = 0; i < 8; i++) o += acc[i].x + acc[i].y + __uint_as_float(r[i] & 0x3fffffffu) + s[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = o;
    if (threadIdx.x % 32 == 0) { cyc[2 * (blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32)] = t0; cyc[2 * (blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32) + 1] = t1; }
}

template <int V> void run(const char* name, int threads, float* out, long long* cyc, long long* h) {
    Taps tp; for (int i = 0; i < 16; i++) tp.t[i] = 0.999f + 1e-4f * i;
    const int iters = 2000, blocks = 148, warps = threads / 32;
    probe<V><<<blocks, threads>>>(tp, out, cyc, 10);
    cudaDeviceSynchronize();
    probe<V><<<blocks, threads>>>(tp, out, cyc, iters);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(h, cyc, sizeof(long long) * 2 * blocks * warps, cudaMemcpyDeviceToHost);
    double avg = 0;                     // per block (= per SM): last warp's end - first warp's start (schedulers may favour some warps)
    for (int b = 0; b < blocks; b++) {
        long long lo = h[2 * b * warps], hi = h[2 * b * warps + 1];
        for (int w = 0; w < warps; w++) { if (h[2 * (b * warps + w)] < lo) lo = h[2 * (b * warps + w)]; if (h[2 * (b * warps + w) + 1] > hi) hi = h[2 * (b * warps + w) + 1]; }
        avg += (double)(hi - lo);
    }
    avg /= blocks;
    // one iteration = 64 "slots" (a slot = one FFMA2 or its scalar pair) per warp; warps / 4 warps per scheduler
    const double per_slot = avg / iters / 64.0 / (warps / 4.0);
    printf("%-44s threads %3d  cycles/iter/warp %8.1f  -> scheduler cycles per slot %.3f   (%s)\n", name, threads, avg / iters, per_slot, cudaGetErrorString(e));
}

int main() {
    float* out; long long* cyc; cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 148 * 16 * 8 * 2);
    long long* h = new long long[148 * 16 * 2];
    for (int threads : {128, 256, 512}) {
        run<0>("FFMA2 reg tap", threads, out, cyc, h);
        run<1>("FFMA2 uniform tap", threads, out, cyc, h);
        run<2>("FFMA2 uniform + SHF 1:1", threads, out, cyc, h);
        run<3>("FFMA2 uniform + SHF 1 per 2", threads, out, cyc, h);
        run<10>("FFMA2 uniform + SHF 2 per 1", threads, out, cyc, h);
        run<4>("FFMA2 uniform + FMNMX 1:1", threads, out, cyc, h);
        run<8>("FFMA2 uniform + IMAD 1:1", threads, out, cyc, h);
        run<11>("FFMA2 uniform + LDS 1 per 4", threads, out, cyc, h);
        run<7>("FFMA2 reg + SHF 1:1", threads, out, cyc, h);
        run<5>("FFMA x2 uniform tap", threads, out, cyc, h);
        run<6>("FFMA x2 uniform + SHF 1 per 2 FFMA", threads, out, cyc, h);
        run<9>("SHF x2 only", threads, out, cyc, h);
        run<17>("FFMA2 acc[i] += x[(i+k)%8] * uniform", threads, out, cyc, h);
        run<18>("FFMA2 acc[i] += x.F32 bcast * uniform pair", threads, out, cyc, h);
        run<19>("FFMA2 acc[i] += x[k] * uniform (shared x)", threads, out, cyc, h);
        run<12>("FFMA x2 (R*R+R)", threads, out, cyc, h);
        run<13>("FFMA x2 (R*R+R) + SHF per 2 FFMA", threads, out, cyc, h);
        run<14>("FFMA x2 (R*UR+R)", threads, out, cyc, h);
        run<16>("FFMA x2 (R*UR+R) + SHF per 2 FFMA", threads, out, cyc, h);
        run<15>("FFMA x2 (R*R+R) + STS.U8 per 4 FFMA", threads, out, cyc, h);
    }
    return 0;
}
