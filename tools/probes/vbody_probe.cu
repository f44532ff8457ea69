// Tensor-core body probe for the 16-bit path of the fused filtered_lrelu (round-1 verdict item 6):
//   vertical up-FIR -> leaky ReLU + sign mask -> vertical down-FIR (+ the horizontal down-FIR straight out of TMEM)
// of the benchmark configuration [32,64,128,128] bf16 (up = down = 2, 12-tap filters, 266 x 266 up-sampled plane), with the two
// vertical passes as banded-Toeplitz tcgen05.mma tiles.  The horizontal up-FIR is NOT part of the probe: its result (bf16,
// 139 input rows x 266 up-res columns per plane) is the probe's input, read from a small L2-resident set of planes.
//
// Per plane, two overlapping column halves of 144 up-res columns ([0,144) and [128,272): each yields 64 output columns), and per half
// three steps t = 0, 1, 2 over up-res rows [128 t, 128 t + 128):
//   MMA1  U[128 x 144] (TMEM, fp32) = A1[128 x K] (constant band matrix: 6 taps per row, K = 80; K = 16 for the 10 rows of t = 2)
//                                     * AH[K x 144] (h-upsampled input rows 64 t .., bf16, MN-major core-matrix layout)
//   EPI1  8 warps: tcgen05.ld -> max(v, 0.2 v), 2-bit sign codes -> global mask, bf16 -> V[128 x 144] in shared memory (MN-major)
//   MMA2  D[128 x 144] (TMEM) += A2[128 x 64] (constant band matrix of the 12-tap down filter, shifted by 32 rows per 64-row chunk)
//                                * V chunk (two chunks per step)
//   EPI2  once per half: tcgen05.ld of D, 12-tap horizontal down-FIR in registers (a thread owns one output row), bf16 -> y
// Roles: warps 0-7 epilogue (TMEM lane quarter = warp & 3, column group = warp >> 2), warp 8 lane 0 issues MMA1, warp 10 lane 0 MMA2, warp 9 lane 0
// loads AH with cp.async.bulk.  U and V are double-buffered, so MMA1 of step g + 1 and MMA2 of step g run under EPI1.
// Every mbarrier wait is bounded (a wrong descriptor must not hang the box): on a timeout all roles leave and err is set.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o vbody_probe vbody_probe.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

constexpr int PLANES = 2048, NDIST = 64;           // planes of [32,64,...]; distinct input planes (5.3 MB: L2-resident)
constexpr int UH = 266, OH = 128, OW = 128;        // up-res rows, output size
constexpr int NH = 144, NG = NH / 8;               // up-res columns per half, 8-column groups
constexpr int KH = 144;                            // input rows held per half (139 used)
constexpr int AH_BYTES = NG * KH * 16;             // 41,472
constexpr int V_BYTES = NG * 128 * 16;             // 36,864
constexpr int A1_BYTES = 10 * 128 * 16, A2_BYTES = 8 * 256 * 16;
constexpr int MASK_PITCH = 68;                     // bytes per mask row (272 columns / 4)
constexpr int OFF_A1 = 0, OFF_A2 = OFF_A1 + A1_BYTES, OFF_AH = OFF_A2 + A2_BYTES, OFF_V = OFF_AH + 2 * AH_BYTES, OFF_END = OFF_V + 2 * V_BYTES;
constexpr int NTHREADS = 608;                     // 2 epilogue groups x 8 warps + MMA1 issuer + loader + MMA2 issuer
constexpr float SLOPE = 0.2f;

__constant__ float c_wup[2][6];                    // vertical up-FIR taps per phase (gain folded in)
__constant__ float c_fd[12];                       // down-FIR taps

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// SWIZZLE_NONE shared-memory matrix descriptor (core matrices of 8 x 16 bytes); K-major: lbo = distance between the two core matrices
// of one K = 16 step, sbo = between 8-row groups; MN-major: lbo = between 8-row groups of K, sbo = between 8-column groups
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((addr >> 4) & 0x3fff) | (uint64_t)((lbo >> 4) & 0x3fff) << 16 | (uint64_t)((sbo >> 4) & 0x3fff) << 32 | (uint64_t)1 << 46;
}
__device__ __forceinline__ void mma_bf16(uint32_t taddr, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(taddr), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_bf16_ta(uint32_t taddr, uint32_t a_taddr, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 :: "r"(taddr), "r"(a_taddr), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bar_arrive(uint64_t* bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory"); }
// bounded wait; false (and *abort = 1) on a timeout or when another role has given up
__device__ __forceinline__ bool bar_wait(uint64_t* bar, uint32_t parity, volatile int* abort) {
    const uint32_t a = smem_u32(bar);
    for (uint32_t spin = 0;; spin++) {
        uint32_t done;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(a), "r"(parity) : "memory");
        if (done) return true;
        __nanosleep(40);                                // back off: a spinning lane takes issue slots from the epilogue warps of its scheduler
        if ((spin & 255) == 255 && (*abort || spin > (1u << 20))) { *abort = 1; return false; }
    }
}
__device__ __forceinline__ bool warp_wait(uint64_t* bar, uint32_t parity, volatile int* abort, int lane) {
    (void)lane;                                     // measured: lane 0 polling + shuffle is SLOWER (167 vs 139 us), every lane polls
    const bool ok = bar_wait(bar, parity, abort);
    return __all_sync(0xffffffffu, ok);
}
#define ST8(addr, r) asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" \
    :: "r"(addr), "r"(r[0]),"r"(r[1]),"r"(r[2]),"r"(r[3]),"r"(r[4]),"r"(r[5]),"r"(r[6]),"r"(r[7]) : "memory")
#define LD16(r, addr) asm volatile( \
    "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];" \
    : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),"=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15]) \
    : "r"(addr))
#define TMEM_WAIT_LD() asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory")
#define FENCE_BEFORE() asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory")
#define FENCE_AFTER() asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory")

struct Bars { uint64_t ah_full[2], ah_empty[2], u_full[2], u_empty[2], v_full[2], v_empty[2], d_full, d_empty; };

__global__ void __launch_bounds__(NTHREADS, 1)
vbody(const uint8_t* __restrict__ in, __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ mask, int* err, long long* cyc, int swapB, int mode, long long* tl) {
    // tl: optional timeline of CTA 0, [step][8] clock64 stamps: 0 mma1 issue, 1 mma2 wait done, 2 mma2 issued, 3 epi u_full passed, 4 epi v_empty passed, 5 epi1 done, 6 epi2 done
    // mode bit 3: input loaded once per CTA (every plane computes on the first one's rows), bit 4: A1 from tensor memory instead of shared memory
    // mode bit 0: issue no MMA (commits only), bit 1: no EPI1 arithmetic / stores, bit 2: no EPI2 arithmetic / stores (timing ablations)
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ Bars bars;
    __shared__ uint32_t tmem_base;
    __shared__ int abort_flag;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long t_start = clock64();
    // ---- constant band matrices, K-major core-matrix layout [k / 8][row][k % 8]
    {
        __nv_bfloat16* A1 = reinterpret_cast<__nv_bfloat16*>(smem + OFF_A1);
        for (int i = tid; i < 10 * 128 * 8; i += NTHREADS) {
            const int kc = i / (128 * 8), m = (i >> 3) & 127, kk = 8 * kc + (i & 7), d = kk - (m >> 1);       // row m reads input rows (m >> 1) + 0..5
            A1[i] = __float2bfloat16_rn(d >= 0 && d < 6 ? c_wup[m & 1][d] : 0.f);
        }
        __nv_bfloat16* A2 = reinterpret_cast<__nv_bfloat16*>(smem + OFF_A2);
        for (int i = tid; i < 8 * 256 * 8; i += NTHREADS) {
            const int kc = i / (256 * 8), mx = (i >> 3) & 255, kk = 8 * kc + (i & 7), tt = kk - 2 * (mx - 128);   // extended row mx <-> output row (mx - 128) relative to the chunk
            A2[i] = __float2bfloat16_rn(tt >= 0 && tt < 12 ? c_fd[tt] : 0.f);
        }
        for (int i = tid; i < 2 * V_BYTES / 16; i += NTHREADS) reinterpret_cast<uint4*>(smem + OFF_V)[i] = make_uint4(0, 0, 0, 0);
    }
    if (tid == 0) {
        auto init = [](uint64_t* b, int n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(n)); };
        for (int i = 0; i < 2; i++) { init(&bars.ah_full[i], 1); init(&bars.ah_empty[i], 1); init(&bars.u_full[i], 1); init(&bars.u_empty[i], 8);
                                      init(&bars.v_full[i], 8); init(&bars.v_empty[i], 1); }
        init(&bars.d_full, 1); init(&bars.d_empty, 8);
        abort_flag = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    FENCE_BEFORE();
    __syncthreads();
    FENCE_AFTER();
    volatile int* abort = &abort_flag;
    constexpr uint32_t TA1 = 3 * NH;                                        // A1 in tensor memory: 128 lanes (rows) x 40 columns (two bf16 of K per column)
    if (warp < 4) {
        const int m = tid;                                                  // row of A1 = TMEM lane
#pragma unroll
        for (int j = 0; j < 5; j++) {
            uint32_t w[8];
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int k0 = 16 * j + 2 * i, d0 = k0 - (m >> 1), d1 = d0 + 1;
                __nv_bfloat162 hh = __floats2bfloat162_rn(d0 >= 0 && d0 < 6 ? c_wup[m & 1][d0] : 0.f, d1 >= 0 && d1 < 6 ? c_wup[m & 1][d1] : 0.f);
                w[i] = *reinterpret_cast<uint32_t*>(&hh);
            }
            ST8(tmem_base + ((uint32_t)(32 * warp) << 16) + TA1 + 8 * j, w);
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    FENCE_BEFORE();
    __syncthreads();
    FENCE_AFTER();
    const int nplanes = (PLANES - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int G = nplanes * 6;                                              // steps: (plane, half, row tile)
    const uint32_t tU = tmem_base, tD = tmem_base + 2 * NH;                 // U0 | U1 | D, 144 columns each
    const uint32_t sA1 = smem_u32(smem + OFF_A1), sA2 = smem_u32(smem + OFF_A2), sAH = smem_u32(smem + OFF_AH), sV = smem_u32(smem + OFF_V);

    if (warp == 16 || warp == 18) {
        if (lane == 0) {
            // D = F32, A = B = BF16, A K-major, B MN-major (bit 16), N = 144, M = 128
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((uint32_t)(NH >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            auto descB = [&](uint32_t addr, uint32_t group_stride) { return swapB ? make_desc(addr, group_stride, 128) : make_desc(addr, 128, group_stride); };
            auto mma1 = [&](int g) -> bool {
                const int t = g % 3, hc = g / 3, h = hc & 1, pl = hc >> 1, b = g & 1;
                if (t == 0 && (!(mode & 8) || pl == 0) && !bar_wait(&bars.ah_full[h], pl & 1, abort)) return false;
                if (!bar_wait(&bars.u_empty[b], ((g >> 1) & 1) ^ 1, abort)) return false;
                FENCE_AFTER();
                if (tl && blockIdx.x == 0 && g < 64) tl[g * 8 + 0] = clock64() - t_start;
                const int ks = t < 2 ? 5 : 1;
                if (!(mode & 1)) for (int s = 0; s < ks; s++)
                    if (mode & 16) mma_bf16_ta(tU + b * NH, tmem_base + TA1 + 8 * s, descB(sAH + h * AH_BYTES + (64 * t + 16 * s) * 16, KH * 16), idesc, s > 0);
                    else mma_bf16(tU + b * NH, make_desc(sA1 + s * 4096, 2048, 128), descB(sAH + h * AH_BYTES + (64 * t + 16 * s) * 16, KH * 16), idesc, s > 0);
                mma_commit(&bars.u_full[b]);
                if (t == 2) mma_commit(&bars.ah_empty[h]);
                return true;
            };
            auto mma2 = [&](int g) -> bool {
                const int t = g % 3, hc = g / 3, b = g & 1;
                if (!bar_wait(&bars.v_full[b], (g >> 1) & 1, abort)) return false;
                if (t == 0 && !bar_wait(&bars.d_empty, (hc & 1) ^ 1, abort)) return false;
                FENCE_AFTER();
                if (tl && blockIdx.x == 0 && g < 64) tl[g * 8 + 1] = clock64() - t_start;
                const int c0 = 2 * t, c1 = t < 2 ? 2 * t + 1 : 2 * t;
                if (!(mode & 1)) for (int c = c0; c <= c1; c++) {
                    const int ks = c == 4 ? 1 : 4;
                    for (int s = 0; s < ks; s++)
                        mma_bf16(tD, make_desc(sA2 + (128 - 32 * c) * 16 + s * 8192, 4096, 128), descB(sV + b * V_BYTES + (64 * (c & 1) + 16 * s) * 16, 128 * 16),
                                 idesc, !(c == 0 && s == 0));
                }
                mma_commit(&bars.v_empty[b]);
                if (tl && blockIdx.x == 0 && g < 64) tl[g * 8 + 2] = clock64() - t_start;
                if (t == 2) mma_commit(&bars.d_full);
                return true;
            };
            // two issuing threads (measured: commit + wait + fence cost one thread ~ 450 cycles per group, ~ 1,000 per step when one thread
            // issues both groups): warp 8 issues MMA1 (as far ahead as U allows: U[g & 1] is free once EPI1(g - 2) has LOADED it),
            // warp 10 issues MMA2(g) when EPI1(g) has stored V
            if (warp == 16) { bool ok = true; for (int g = 0; g < G && ok; g++) ok = mma1(g); }
            else { bool ok = true; for (int g = 0; g < G && ok; g++) ok = mma2(g); }
        }
    } else if (warp == 17) {
        if (lane == 0) {
            for (int i = 0; i < ((mode & 8) ? 2 : nplanes * 2); i++) {
                const int pl = i >> 1, h = i & 1, plane = (int)blockIdx.x + pl * (int)gridDim.x;
                if (!bar_wait(&bars.ah_empty[h], (pl & 1) ^ 1, abort)) break;
                const uint32_t bar = smem_u32(&bars.ah_full[h]);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(AH_BYTES) : "memory");
                const uint8_t* src = in + ((size_t)(plane % NDIST) * 2 + h) * AH_BYTES;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"(sAH + h * AH_BYTES), "l"(src), "r"(AH_BYTES), "r"(bar) : "memory");
            }
        }
    } else {
        // two epilogue groups of 8 warps: group 0 owns the even steps (U[0], V[0]), group 1 the odd ones - while one group sits in the
        // ~1,000-cycle barrier chain of its next step the other one computes
        const int grp = warp >> 3, q = warp & 3, cg = (warp >> 2) & 1;
        const uint32_t lane_off = (uint32_t)(32 * q) << 16;
        float fd[12];
#pragma unroll
        for (int i = 0; i < 12; i++) fd[i] = c_fd[i];
        // EPI2: D of half hc -> horizontal down-FIR -> y
        auto depi = [&](int hc) -> bool {
            const int h = hc & 1, pl = hc >> 1, plane = (int)blockIdx.x + pl * (int)gridDim.x, o = 32 * q + lane;
            if (!warp_wait(&bars.d_full, hc & 1, abort, lane)) return false;
            FENCE_AFTER();
            const int oc0 = 4 * cg;
            if (!(mode & 4)) {
                uint32_t cur[16], nxt[16];                             // sliding window over the column chunks oc0 .. oc0 + 4
                LD16(cur, tD + lane_off + 16 * oc0);
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    LD16(nxt, tD + lane_off + 16 * (oc0 + k + 1));
                    TMEM_WAIT_LD();
                    if (k == 3) {                                      // D is in registers: the next half's MMA2 may start under the last FIR block
                        FENCE_BEFORE();
                        __syncwarp();
                        if (lane == 0) bar_arrive(&bars.d_empty);
                    }
                    uint32_t pk[4];
#pragma unroll
                    for (int i = 0; i < 8; i += 2) {
                        float a0 = 0.f, a1 = 0.f;
#pragma unroll
                        for (int tt = 0; tt < 12; tt++) {
                            const int j0 = 2 * i + tt, j1 = 2 * i + 2 + tt;
                            a0 = fmaf(fd[tt], __uint_as_float(j0 < 16 ? cur[j0] : nxt[j0 - 16]), a0);
                            a1 = fmaf(fd[tt], __uint_as_float(j1 < 16 ? cur[j1] : nxt[j1 - 16]), a1);
                        }
                        __nv_bfloat162 hh = __floats2bfloat162_rn(a0, a1);
                        pk[i >> 1] = *reinterpret_cast<uint32_t*>(&hh);
                    }
                    if (!(mode & 64) || pk[0] == 0x12345678u) *reinterpret_cast<uint4*>(y + ((size_t)plane * OH + o) * OW + 64 * h + 8 * (oc0 + k)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
#pragma unroll
                    for (int i = 0; i < 16; i++) cur[i] = nxt[i];
                }
                return true;
            }
            FENCE_BEFORE();
            __syncwarp();
            if (lane == 0) bar_arrive(&bars.d_empty);
            return true;
        };
        bool ok = true;
        for (int g = grp; g < G && ok; g += 2) {
            const int t = g % 3, hc = g / 3, h = hc & 1, pl = hc >> 1, b = g & 1, plane = (int)blockIdx.x + pl * (int)gridDim.x;
            if (!warp_wait(&bars.u_full[b], (g >> 1) & 1, abort, lane)) { ok = false; break; }
            if (tl && blockIdx.x == 0 && g < 64 && tid == 0) tl[g * 8 + 3] = clock64() - t_start;
            FENCE_AFTER();
            const bool active = (t < 2 || q == 0) && !(mode & 2);
            const int ch0 = cg ? 5 : 0, nch = cg ? 4 : 5;
            // TMEM loads in batches of 2 + 2 + 1 chunks of 16 columns (608 threads leave 104 registers each), each batch processed under the next
            // one's loads; U is released after the last wait
            uint32_t ra[2][16], rb[2][16];
            const int krow = 32 * q + lane, r = 128 * t + krow;
            uint8_t* vdst = smem + OFF_V + b * V_BYTES + krow * 16;
            uint8_t* mrow = mask + ((size_t)plane * UH + (r < UH ? r : 0)) * MASK_PITCH + 32 * h;
            auto chunk = [&](const uint32_t* rr, int ch) {
                uint32_t pk[8], bq[4] = {0, 0, 0, 0};                                          // four independent chains, one per mask byte
#pragma unroll
                for (int i = 3; i >= 0; i--)
#pragma unroll
                    for (int j = 0; j < 4; j++) bq[j] = __funnelshift_l(rr[4 * j + i], bq[j], 2);       // (b << 2) | top two bits of the value
                const uint32_t bits = __byte_perm(__byte_perm(bq[0], bq[1], 0x0040), __byte_perm(bq[2], bq[3], 0x0040), 0x5410);
#pragma unroll
                for (int i = 0; i < 16; i += 2) {
                    float a = __uint_as_float(rr[i]), c = __uint_as_float(rr[i + 1]);
                    a = fmaxf(a, SLOPE * a); c = fmaxf(c, SLOPE * c);
                    __nv_bfloat162 hh = __floats2bfloat162_rn(a, c);
                    pk[i >> 1] = *reinterpret_cast<uint32_t*>(&hh);
                }
                *reinterpret_cast<uint4*>(vdst + (2 * ch) * 2048) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                *reinterpret_cast<uint4*>(vdst + (2 * ch + 1) * 2048) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                if (r < UH && !(h == 1 && ch == 0) && !(mode & 32)) *reinterpret_cast<uint32_t*>(mrow + 4 * ch) = (bits >> 1) & 0x55555555u;   // code 1 = negative
            };
            if (active) {
                LD16(ra[0], tU + b * NH + lane_off + 16 * ch0); LD16(ra[1], tU + b * NH + lane_off + 16 * (ch0 + 1));
                TMEM_WAIT_LD();
                LD16(rb[0], tU + b * NH + lane_off + 16 * (ch0 + 2)); LD16(rb[1], tU + b * NH + lane_off + 16 * (ch0 + 3));
            }
            if (!warp_wait(&bars.v_empty[b], ((g >> 1) & 1) ^ 1, abort, lane)) { ok = false; break; }
            if (tl && blockIdx.x == 0 && g < 64 && tid == 0) tl[g * 8 + 4] = clock64() - t_start;
            if (active) {
                chunk(ra[0], ch0); chunk(ra[1], ch0 + 1);                                  // under the second batch of loads
                TMEM_WAIT_LD();
                if (nch == 5) LD16(ra[0], tU + b * NH + lane_off + 16 * (ch0 + 4));
                chunk(rb[0], ch0 + 2); chunk(rb[1], ch0 + 3);
                TMEM_WAIT_LD();
            }
            FENCE_BEFORE();
            __syncwarp();
            if (lane == 0) bar_arrive(&bars.u_empty[b]);              // U[b] is in registers: MMA1 of step g + 2 may overwrite it
            if (active && nch == 5) chunk(ra[0], ch0 + 4);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) bar_arrive(&bars.v_full[b]);
            if (tl && blockIdx.x == 0 && g < 64 && tid == 0) tl[g * 8 + 5] = clock64() - t_start;
            if (t == 0 && g > 0) { ok = depi(hc - 1); if (tl && blockIdx.x == 0 && g < 64 && tid == 0) tl[g * 8 + 6] = clock64() - t_start; }
        }
        if (ok && grp == 1) depi(G / 3 - 1);            // G is even: the last step belongs to group 1
    }
    FENCE_BEFORE();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem_base));
    if (tid == 0) { if (abort_flag) atomicExch(err, 1); cyc[blockIdx.x] = clock64() - t_start; }
}

static float bf(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

int main(int argc, char** argv) {
    const int reps = argc > 1 ? atoi(argv[1]) : 10, tlmode = argc > 2 ? atoi(argv[2]) : 0;
    // 12-tap windowed-sinc low-pass (cutoff 0.25 of the up-res rate), the shape of the reference's design_lowpass_filter output
    float f[12], sum = 0.f;
    for (int i = 0; i < 12; i++) { const float x = (i - 5.5f) * 0.5f, w = 0.5f - 0.5f * cosf(2.f * 3.14159265f * (i + 0.5f) / 12.f); f[i] = (fabsf(x) < 1e-6f ? 1.f : sinf(3.14159265f * x) / (3.14159265f * x)) * w; sum += f[i]; }
    float wup[2][6], fd[12];
    for (int i = 0; i < 12; i++) { f[i] /= sum; fd[i] = bf(f[i]); }
    for (int ph = 0; ph < 2; ph++) for (int k = 0; k < 6; k++) wup[ph][k] = bf(f[2 * k + ph] * 2.f * sqrtf(2.f));       // up = 2, gain sqrt(2)
    cudaMemcpyToSymbol(c_wup, wup, sizeof(wup)); cudaMemcpyToSymbol(c_fd, fd, sizeof(fd));
    // input: NDIST planes of h-upsampled rows, [plane][row 0..143][col 0..271] bf16, packed per half as [col / 8][row][col % 8]
    std::vector<float> ah((size_t)NDIST * KH * 272);
    uint32_t seed = 12345u;
    for (auto& v : ah) { seed = seed * 1664525u + 1013904223u; v = bf(((seed >> 8) & 0xffff) / 32768.f - 1.f); }
    std::vector<__nv_bfloat16> packed((size_t)NDIST * 2 * AH_BYTES / 2);
    for (int p = 0; p < NDIST; p++) for (int h = 0; h < 2; h++) for (int g = 0; g < NG; g++) for (int k = 0; k < KH; k++) for (int e = 0; e < 8; e++)
        packed[((((size_t)p * 2 + h) * NG + g) * KH + k) * 8 + e] = __float2bfloat16_rn(ah[((size_t)p * KH + k) * 272 + 128 * h + 8 * g + e]);
    uint8_t *d_in, *d_mask; __nv_bfloat16* d_y; int* d_err; long long* d_cyc;
    const size_t ybytes = (size_t)PLANES * OH * OW * 2, mbytes = (size_t)PLANES * UH * MASK_PITCH;
    cudaMalloc(&d_in, packed.size() * 2); cudaMalloc(&d_y, ybytes); cudaMalloc(&d_mask, mbytes); cudaMalloc(&d_err, 4); cudaMalloc(&d_cyc, 148 * 8);
    long long* d_tl; cudaMalloc(&d_tl, 64 * 8 * 8); cudaMemset(d_tl, 0, 64 * 8 * 8);
    cudaMemcpy(d_in, packed.data(), packed.size() * 2, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(vbody, cudaFuncAttributeMaxDynamicSharedMemorySize, OFF_END);
    printf("vbody probe: %d planes, shared memory %d bytes per CTA, 148 CTAs x %d threads\n", PLANES, OFF_END, NTHREADS);

    // CPU restatement for a few planes (bf16 operands, fp32 accumulation, bf16 intermediate, as the tensor-core path computes)
    const int check[4] = {0, 1, 147, 2047};
    std::vector<float> yref(4 * OH * OW); std::vector<uint8_t> sref((size_t)4 * UH * 266); std::vector<float> uabs((size_t)4 * UH * 266);
    for (int ci = 0; ci < 4; ci++) {
        const float* a = &ah[(size_t)(check[ci] % NDIST) * KH * 272];
        std::vector<float> V((size_t)272 * 272, 0.f), D((size_t)OH * 272, 0.f);
        for (int r = 0; r < UH; r++) for (int c = 0; c < 266; c++) {
            float u = 0.f;
            for (int k = 0; k < 6; k++) u += wup[r & 1][k] * a[(size_t)((r >> 1) + k) * 272 + c];
            sref[((size_t)ci * UH + r) * 266 + c] = u < 0.f; uabs[((size_t)ci * UH + r) * 266 + c] = fabsf(u);
            V[(size_t)r * 272 + c] = bf(fmaxf(u, SLOPE * u));
        }
        for (int o = 0; o < OH; o++) for (int c = 0; c < 266; c++) { float d = 0.f; for (int t = 0; t < 12; t++) d += fd[t] * V[(size_t)(2 * o + t) * 272 + c]; D[(size_t)o * 272 + c] = d; }
        for (int o = 0; o < OH; o++) for (int x = 0; x < OW; x++) { float s = 0.f; for (int t = 0; t < 12; t++) s += fd[t] * D[(size_t)o * 272 + 2 * x + t]; yref[((size_t)ci * OH + o) * OW + x] = s; }
    }
    std::vector<__nv_bfloat16> hy(OH * OW); std::vector<uint8_t> hm((size_t)UH * MASK_PITCH);
    for (int swapB = 0; swapB < 2; swapB++) {
        cudaMemset(d_err, 0, 4); cudaMemset(d_y, 0xff, ybytes); cudaMemset(d_mask, 0xff, mbytes);
        vbody<<<148, NTHREADS, OFF_END>>>(d_in, d_y, d_mask, d_err, d_cyc, swapB, 0, nullptr);
        cudaError_t e = cudaDeviceSynchronize();
        int herr = 0; cudaMemcpy(&herr, d_err, 4, cudaMemcpyDeviceToHost);
        printf("B descriptor %s: launch %s, timeout flag %d\n", swapB ? "lbo = column-group stride, sbo = 128" : "lbo = 128 (K groups), sbo = column-group stride", cudaGetErrorString(e), herr);
        if (e != cudaSuccess) return 1;
        if (herr) continue;
        double maxerr = 0, maxref = 0; long long sign_bad = 0, sign_far = 0;
        for (int ci = 0; ci < 4; ci++) {
            cudaMemcpy(hy.data(), d_y + (size_t)check[ci] * OH * OW, OH * OW * 2, cudaMemcpyDeviceToHost);
            cudaMemcpy(hm.data(), d_mask + (size_t)check[ci] * UH * MASK_PITCH, hm.size(), cudaMemcpyDeviceToHost);
            for (int i = 0; i < OH * OW; i++) { const double r = yref[(size_t)ci * OH * OW + i], d = fabs((double)__bfloat162float(hy[i]) - r); if (!(d <= maxerr)) maxerr = d; if (fabs(r) > maxref) maxref = fabs(r); }
            for (int r = 0; r < UH; r++) for (int c = 0; c < 266; c++) {
                const int code = (hm[(size_t)r * MASK_PITCH + (c >> 2)] >> (2 * (c & 3))) & 3;
                if (code != sref[((size_t)ci * UH + r) * 266 + c]) { sign_bad++; if (uabs[((size_t)ci * UH + r) * 266 + c] > 1e-4f) sign_far++; }
            }
        }
        printf("  4 planes against the CPU restatement: max |y - ref| = %.3g (max |ref| %.3g, relative %.3g); sign codes differing: %lld (%lld at |u| > 1e-4)\n",
               maxerr, maxref, maxerr / maxref, sign_bad, sign_far);
        const bool good = maxerr / maxref < 2e-2 && sign_far == 0;
        printf("  => %s\n", good ? "PARITY OK (2e-2 of BASELINE.json for 16-bit I/O)" : "MISMATCH");
        if (!good) continue;
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        auto timed = [&](int mode, float& best, float& avg) {
            best = 1e9f; float tot = 0.f;
            for (int i = 0; i < reps; i++) {
                cudaEventRecord(e0);
                vbody<<<148, NTHREADS, OFF_END>>>(d_in, d_y, d_mask, d_err, d_cyc, swapB, mode, nullptr);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms = 0; cudaEventElapsedTime(&ms, e0, e1); tot += ms; if (ms < best) best = ms;
            }
            avg = tot / reps;
        };
        float best, avg;
        timed(0, best, avg);
        const float tot = avg * reps;
        long long hc[148]; cudaMemcpy(hc, d_cyc, sizeof(hc), cudaMemcpyDeviceToHost);
        long long cmax = 0; for (int i = 0; i < 148; i++) if (hc[i] > cmax) cmax = hc[i];
        printf("  %d launches: average %.1f us, best %.1f us per launch of %d planes; slowest CTA %lld cycles = %.0f cycles per plane (14 planes)\n",
               reps, tot / reps * 1e3, best * 1e3, PLANES, cmax, cmax / 14.0);
        printf("  written per launch: y %.1f MB + mask %.1f MB; read (L2-resident set) %.1f MB\n", ybytes / 1e6, mbytes / 1e6, (double)PLANES * 2 * AH_BYTES / 1e6);
        struct { int mode; const char* name; } abl[] = {{16, "A1 from tensor memory (instead of shared memory)"}, {8, "input resident (no reloads from L2)"}, {8 + 1, "resident, no MMA issued"},
            {8 + 2, "resident, no EPI1"}, {8 + 4, "resident, no EPI2"}, {8 + 6, "resident, MMA only"}, {8 + 7, "resident, barriers only"}, {8 + 32, "resident, no mask stores"}, {8 + 64, "resident, no y stores"}, {8 + 96, "resident, no global stores at all"}, {1, "no MMA issued"}, {6, "MMA only"}, {7, "barriers and loads only"}};
        for (auto& a : abl) { float b2, a2; timed(a.mode, b2, a2); printf("  ablation %2d  %-50s average %.1f us, best %.1f us\n", a.mode, a.name, a2 * 1e3, b2 * 1e3); }
        vbody<<<148, NTHREADS, OFF_END>>>(d_in, d_y, d_mask, d_err, d_cyc, swapB, tlmode, d_tl);
        cudaDeviceSynchronize();
        long long htl[64 * 8]; cudaMemcpy(htl, d_tl, sizeof(htl), cudaMemcpyDeviceToHost);
        printf("  timeline (mode %d) of CTA 0 (cycles since kernel start): step | mma1 issued | mma2: v_full seen, issued | epi: u_full seen, v_empty seen, epi1 done, epi2 done\n", tlmode);
        for (int g = 0; g < 30; g++) printf("   %2d (t=%d) | %7lld | %7lld %7lld | %7lld %7lld %7lld %7lld\n", g, g % 3, htl[g * 8], htl[g * 8 + 1], htl[g * 8 + 2], htl[g * 8 + 3], htl[g * 8 + 4], htl[g * 8 + 5], htl[g * 8 + 6]);
        int herr2 = 0; cudaMemcpy(&herr2, d_err, 4, cudaMemcpyDeviceToHost);
        printf("  timeout flag after the ablations: %d\n", herr2);
        break;
    }
    return 0;
}
