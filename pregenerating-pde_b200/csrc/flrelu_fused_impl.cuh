// Fused filtered_lrelu for sm_100a: bias -> FIR upsample -> gain/leaky-ReLU/clamp (+2-bit sign mask
// write or read) -> FIR downsample in ONE kernel.  Replaces the reference's filtered_lrelu_kernel
// (filtered_lrelu.cu:131-1091), setup_filters_kernel (:79-100) and the constant-memory copy
// (:103-109); the algorithm is described in flrelu_core.h.  This file adds the Blackwell plumbing:
// a 3-D TMA tensor map over x[planes, H, W] whose out-of-bounds zero fill supplies the padding halo,
// a 3-stage mbarrier ring of row strips, and the launch logic.
#pragma once
#include "common.cuh"
#include "flrelu_host.h"
#include <cuda.h>

namespace flr {
template <> struct Cvt<__nv_bfloat16> { static __device__ __forceinline__ float get(__nv_bfloat16 v) { return __bfloat162float(v); } };
template <> struct Cvt<__half> { static __device__ __forceinline__ float get(__half v) { return __half2float(v); } };
}

namespace flr_impl {
using namespace flr;

#ifdef FLR_TIMING     // diagnostic builds (tools/variants.sh): per-warp cycle accounting of the streaming loop, read by cno_flr_prof_read()
static __device__ long long g_flr_prof[148 * 16 * 8];
#define FLR_T(i) { const long long t_ = clock64(); tacc[i] += t_ - tlast; tlast = t_; }
#else
#define FLR_T(i)
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    // (a suspend-time hint on try_wait was measured twice, rounds 1 and 2: the poll rate and the launch time do not change)
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

template <class T> __device__ __forceinline__ void store_quad(T* dst, const float* o);
template <> __device__ __forceinline__ void store_quad<float>(float* dst, const float* o) {
    *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
}
template <> __device__ __forceinline__ void store_quad<__nv_bfloat16>(__nv_bfloat16* dst, const float* o) {
    __nv_bfloat162 a = __floats2bfloat162_rn(o[0], o[1]), b = __floats2bfloat162_rn(o[2], o[3]);
    uint2 u; u.x = *reinterpret_cast<uint32_t*>(&a); u.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(dst) = u;
}
template <> __device__ __forceinline__ void store_quad<__half>(__half* dst, const float* o) {
    __half2 a = __floats2half2_rn(o[0], o[1]), b = __floats2half2_rn(o[2], o[3]);
    uint2 u; u.x = *reinterpret_cast<uint32_t*>(&a); u.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(dst) = u;
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Synchronisation inside a CTA (no __syncthreads in the steady state):
//   x stages   : NSTAGE-slot TMA ring, one mbarrier per slot (transaction bytes)
//   D buffers  : NDBUF = 3 buffers of finished v-down rows.  Body k writes D[k % 3] and every thread arrives on full[k % 3];
//                the h-down of body k-1 runs AFTER body k, so its wait on full[k-1] has had a whole body to complete, and the
//                wait in front of body k+1 (guard for reusing D[(k+1) % 3] and the x slot / mask buffer of body k) comes after
//                that h-down work.  Warps therefore drift by up to one h-down phase without stalling each other.
template <class T, class C, int MODE, bool AFF>
__global__ void __launch_bounds__(FLR_MAXT, FLR_MINB)
flr_fused_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ Params p) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* xst = smem;
    float* D = reinterpret_cast<float*>(smem + (size_t)NSTAGE * p.stage_bytes);
    const size_t dbuf = (size_t)p.P * C::KO * p.dp;                      // floats per D buffer
    uint64_t* bars = reinterpret_cast<uint64_t*>(D + NDBUF * dbuf);
    uint64_t* full = bars + NSTAGE;
    uint8_t* M = reinterpret_cast<uint8_t*>(bars) + 64;                 // READ: 2 x [P][PB][mstride] staged mask rows;
    const size_t mbuf = (size_t)p.P * C::PB * p.mstride;                // WRITE: NDBUF x [PB][MSP] sign bytes on their way to HBM
    const size_t wbuf = (size_t)C::PB * stage_pitch<C>(p.msp);
    auto copy4 = [](uint8_t* dst, const uint8_t* src, bool ok) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(ok ? 4 : 0) : "memory");
    };
    const int tid = threadIdx.x;
    const uint32_t box_bytes = (uint32_t)(p.P * C::RS * p.bwx * (int)sizeof(T));
    const uint32_t stage_tx = box_bytes * (uint32_t)p.nbox;

    if (tid == 0) {
        for (int i = 0; i < NSTAGE; i++) mbar_init(&bars[i], 1);
        for (int i = 0; i < NDBUF; i++) mbar_init(&full[i], blockDim.x >> 5);     // one arrival per warp
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    const bool vec_ok = (p.OW & 3) == 0;
    T* const yb = reinterpret_cast<T*>(p.y);
    // fused `db = dx.sum([0,2,3])` (filtered_lrelu.py:257-258): the streaming threads sum their finished v-down rows against
    // per-column tap weights (Stream::dbw); partial sums meet in s_db[plane slot] at the end of a segment and leave with one
    // global atomic per plane and segment
    __shared__ float s_db[FLR_MAXT];
    const bool want_db = FLR_IS_READ(MODE) && p.db != nullptr;
    if (want_db) { for (int i = tid; i < FLR_MAXT; i += blockDim.x) s_db[i] = 0.f; }
    T* yplane = yb;                               // output plane of this thread (set per segment): h-down rows are written by the plane's own threads
    auto store4 = [&](int /*plane*/, int o, int ox, const float* out) {
        T* dst = yplane + (uint32_t)(o * p.OW + ox);
        if (vec_ok) store_quad<T>(dst, out);
        else {
#pragma unroll
            for (int i = 0; i < 4; i++) if (ox + i < p.OW) dst[i] = from_f32<T>(out[i]);
        }
    };

    // this CTA's range of the global row space, processed segment by segment
#ifdef FLR_TIMING
    long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0}; long long tlast = clock64(); const long long tstart = tlast;
#endif
    const int r0 = cta_cut(p, blockIdx.x, C::ALIGN, C::SNAP), r1 = cta_cut(p, blockIdx.x + 1, C::ALIGN, C::SNAP);
    int gs0 = 0;                                  // x stages issued by earlier segments (slot / parity bookkeeping)
    int db = 0; uint32_t dpar = 0;                // D buffer / full-barrier slot and phase parity of the NEXT body
    for (int gr = r0; gr < r1;) {
        const CtaGeom cg = seg_geom<C>(p, gr, r1);
        __syncthreads();                          // barrier init visible; the previous segment is completely finished
        int next_load = 0;
        auto issue = [&](int s) {
            const int slot = (gs0 + s) % NSTAGE;
            mbar_expect_tx(&bars[slot], stage_tx);
            for (int bx = 0; bx < p.nbox; bx++)
                tma_load_3d(xst + (size_t)slot * p.stage_bytes + (size_t)bx * box_bytes, &tmap, &bars[slot], p.tmaX + bx * p.bstep, cg.row0 + s * C::RS, cg.plane0);
        };
        if (tid == 0) while (next_load < cg.nstages_total && next_load < NSTAGE) issue(next_load++);

        Stream<T, C, MODE, AFF> st;
        st.init(p, cg, tid, [&](int ch) { return Cvt<T>::get(reinterpret_cast<const T*>(p.b)[ch]); });
        yplane = yb + (size_t)(cg.plane0 + st.unit) * p.OH * p.OW;
        // per-thread bases of the segment: a body only adds its slot / buffer offsets
        const T* const xbase = reinterpret_cast<const T*>(xst) + (size_t)st.unit * C::RS * p.bwx;
        float* const Dbase = D + (size_t)st.unit * C::KO * p.dp;
        const uint8_t* const Mrbase = M + (size_t)st.unit * C::PB * p.mstride;
        const uint32_t stage_el = (uint32_t)p.stage_bytes / (uint32_t)sizeof(T), pq_el = (uint32_t)(C::PQ * p.bwx);

        if (FLR_IS_READ(MODE) && st.active) stage_mask<C>(p, st.sm_lane, st.sm_dst, st.sm_src, cg.j_b0, M, copy4);
        mbar_wait(&bars[gs0 % NSTAGE], (uint32_t)((gs0 / NSTAGE) & 1));
        if (st.active) st.prologue(p, reinterpret_cast<const T*>(xst + (size_t)(gs0 % NSTAGE) * p.stage_bytes) + (size_t)st.unit * C::RS * p.bwx, cg.row0);
        if (FLR_IS_READ(MODE)) { asm volatile("cp.async.wait_all;" ::: "memory"); __syncthreads(); }

        FLR_T(0)                                  // 0: segment set-up (TMA issue, init, first stage wait, prologue)
        auto epilogue = [&](int kb, int buf) {     // h-down + sign-mask copy-out of finished body kb (D / staging buffer `buf`)
#ifndef FLR_X_NOHDOWN
            if (st.active) hdown_phase<T, C>(p, cg, cg.plane0 + st.unit, st.unit, st.g, kb, Dbase + buf * (uint32_t)dbuf, store4);
#endif
#ifndef FLR_X_NOMASKOUT
            if (MODE == SIGN_WRITE && st.active) mask_out<C>(p, st.mo_src, st.mo_dst, cg.j_b0 + kb * C::PB, M + buf * (uint32_t)wbuf);
#endif
        };
        int pb = 0; uint32_t ppar = 0;            // slot / parity of the previous body
        for (int k = 0; k < cg.nbodies; k++) {
            const int s = 1 + k / C::M, slot = (gs0 + s) % NSTAGE, sub = k % C::M;
            if (k > 0) {
#ifndef FLR_X_NOWAIT
                mbar_wait(&full[pb], ppar);       // every thread has finished body k-1 (and the h-down of body k-3)
#endif
                FLR_T(1)                          // 1: wait for the other warps (full barrier)
                if (tid == 0) {
                    const int consumed = 1 + k / C::M;
                    while (next_load < cg.nstages_total && next_load < consumed + NSTAGE) issue(next_load++);
                }
            }
            if (FLR_IS_READ(MODE) && k + 1 < cg.nbodies && st.active)    // next body's mask rows, in flight during this body
                stage_mask<C>(p, st.sm_lane, st.sm_dst, st.sm_src, cg.j_b0 + (k + 1) * C::PB, M + ((k + 1) & 1) * (uint32_t)mbuf, copy4);
            if (sub == 0) mbar_wait(&bars[slot], (uint32_t)(((gs0 + s) / NSTAGE) & 1));
            FLR_T(2)                              // 2: TMA issue, mask staging, x stage wait
            if (want_db) {                        // rows of this body that are real output rows (cf. hdown_phase: o = o0 - (ND-1) + k*KO + r)
                const int lo = (C::ND - 1) - k * C::KO, hi = (cg.o1 - cg.o0) + (C::ND - 1) - k * C::KO;
                st.db_lo = lo > 0 ? lo : 0; st.db_hi = hi < C::KO ? hi : C::KO;
            }
#ifndef FLR_X_NOBODY
            if (st.active) {
                const T* xt = xbase + slot * stage_el + sub * pq_el;
                const int grow0 = cg.row0 + s * C::RS + sub * C::PQ;
                float* Dk = Dbase + db * (uint32_t)dbuf;
                const uint8_t* Mk = Mrbase + (k & 1) * (uint32_t)mbuf;
                uint8_t* Mw = M + db * (uint32_t)wbuf + tid;
                // bodies whose input rows all lie inside the plane (all but the first / last of a plane) skip the row-validity selects
                const int jb = cg.j_b0 + k * C::PB;
                if (!(grow0 >= 0 && grow0 + C::PQ <= p.H)) st.template body<true, -1>(p, xt, grow0, jb, Dk, Mk, Mw);
                else if (C::UP == 2 && sizeof(T) == 4 && (p.xshift & 1)) st.template body<false, 1>(p, xt, grow0, jb, Dk, Mk, Mw);
                else st.template body<false, 0>(p, xt, grow0, jb, Dk, Mk, Mw);
            }
#endif
            if (FLR_IS_READ(MODE)) asm volatile("cp.async.wait_all;" ::: "memory");
            __syncwarp();
            FLR_T(3)                              // 3: body
            if ((tid & 31) == 0) mbar_arrive(&full[db]);
            if (k > 0) epilogue(k - 1, pb);
            FLR_T(4)                              // 4: h-down + mask copy-out
            pb = db; ppar = dpar;
            if (++db == NDBUF) { db = 0; dpar ^= 1u; }
        }
        mbar_wait(&full[pb], ppar);               // drain: h-down of the segment's last body
        FLR_T(1)
        epilogue(cg.nbodies - 1, pb);
        if (want_db) {
            // segmented warp reduction: lanes of one plane slot are contiguous
            const int lane = tid & 31;
            const int u = st.active ? st.unit : -1;
            float v = (st.dbacc[0].x + st.dbacc[0].y) + (st.dbacc[1].x + st.dbacc[1].y);
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const float ov = __shfl_down_sync(0xffffffffu, v, d);
                const int ou = __shfl_down_sync(0xffffffffu, u, d);
                if (lane + d < 32 && ou == u) v += ov;
            }
            const int pu = __shfl_up_sync(0xffffffffu, u, 1);
            if (u >= 0 && (lane == 0 || pu != u)) atomicAdd(&s_db[u], v);
            __syncthreads();
            if (tid < p.P) {
                if (cg.plane0 + tid < p.NC) atomicAdd(p.db + (cg.plane0 + tid) % p.C, s_db[tid]);
                s_db[tid] = 0.f;
            }
        }
        FLR_T(5)                                  // 5: drain h-down, db reduction
        gs0 += cg.nstages_total;
        gr += cg.o1 - cg.o0;
    }
#ifdef FLR_TIMING
    if ((tid & 31) == 0) {
        long long* o = g_flr_prof + ((size_t)blockIdx.x * 16 + (tid >> 5)) * 8;
        for (int i = 0; i < 6; i++) o[i] = tacc[i];
        o[6] = clock64() - tstart; o[7] = tstart;
    }
#endif
}

// ---- driver entry point for tensor-map encoding (no -lcuda link dependency) ----------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* ptr = nullptr; cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)ptr;
    }
    return fn;
}

inline int env_ctas() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("CNO_FLR_CTAS"); v = e ? atoi(e) : 0; if (v < 0) v = 0; }
    return v;
}

template <class T, class C, int MODE, bool AFF = false>
int launch_variant(const Plan& plan, const CUtensorMap& map, cudaStream_t st) {
    auto kern = flr_fused_kernel<T, C, MODE, AFF>;
    static DevSmemOptIn optin;         // per instantiation, per device
    {
        cudaError_t e = optin.ensure(kern, plan.smem_bytes);
        if (e != cudaSuccess) { cno_set_error("filtered_lrelu: cannot reserve %zu bytes of shared memory: %s", plan.smem_bytes, cudaGetErrorString(e)); return (int)e; }
    }
    // persistent grid: one CTA per resident slot (occupancy x SM count), cached per device for the last (threads, smem) asked
    struct Occ { std::atomic<uint64_t> key{0}; std::atomic<int> ctas{0}; };
    static Occ occ[CNO_MAX_DEVICES];
    const int dev = cno_current_device();
    const uint64_t key = ((uint64_t)plan.threads << 32) | (uint64_t)plan.smem_bytes;
    int occ_ctas = occ[dev].ctas.load(std::memory_order_acquire);
    if (occ[dev].key.load(std::memory_order_acquire) != key || occ_ctas == 0) {
        int per_sm = 0, sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, plan.threads, plan.smem_bytes) != cudaSuccess || per_sm < 1) per_sm = 1;
        occ_ctas = per_sm * sms;
        occ[dev].key.store(0, std::memory_order_release);      // invalidate, publish the value, then the key
        occ[dev].ctas.store(occ_ctas, std::memory_order_release);
        occ[dev].key.store(key, std::memory_order_release);
    }
    Plan pl = plan;
    if (env_ctas() == 0) set_partition(pl, occ_ctas);
    kern<<<pl.blocks, pl.threads, pl.smem_bytes, st>>>(map, pl.p);
    CNO_LAUNCH_CHECK();
    return 0;
}

template <class T, class C>
int launch_mode(const Plan& plan, const CUtensorMap& map, int mode, cudaStream_t st) {
    if (plan.p.pa) {                      // CIN affine folded into the x load: forward variants only
        if (mode == CNO_SIGN_WRITE) return launch_variant<T, C, SIGN_WRITE, true>(plan, map, st);
        if (mode == CNO_SIGN_NONE) return launch_variant<T, C, SIGN_NONE, true>(plan, map, st);
        return -1;
    }
    if (mode == CNO_SIGN_WRITE) return launch_variant<T, C, SIGN_WRITE>(plan, map, st);
    if (mode == CNO_SIGN_READ) return launch_variant<T, C, SIGN_READ>(plan, map, st);
    if (mode == CNO_SIGN_READ_SIGNONLY) return launch_variant<T, C, SIGN_READ_SO>(plan, map, st);
    return launch_variant<T, C, SIGN_NONE>(plan, map, st);
}

// The kernels of one storage type are spread over five translation units (flrelu_fused_parts.cuh: FLR_PART 0..4) so that the
// library builds in a few minutes on 8 cores; the unit that holds launch_cfg<T> only sees declarations.
#ifndef FLR_QUICK
#define FLR_EXTERN_MODE(T, U, D, R) extern template int launch_mode<T, Cfg<U, D, R, R>>(const Plan&, const CUtensorMap&, int, cudaStream_t);
#define FLR_EXTERN_ALL(T) FLR_EXTERN_MODE(T, 2, 2, 0) FLR_EXTERN_MODE(T, 2, 2, 1) FLR_EXTERN_MODE(T, 2, 4, 0) FLR_EXTERN_MODE(T, 2, 4, 1) \
    FLR_EXTERN_MODE(T, 4, 2, 0) FLR_EXTERN_MODE(T, 4, 2, 1) FLR_EXTERN_MODE(T, 4, 2, 2) FLR_EXTERN_MODE(T, 4, 2, 3)
#ifndef FLR_PART
FLR_EXTERN_ALL(float) FLR_EXTERN_ALL(__nv_bfloat16) FLR_EXTERN_ALL(__half)
#endif
#endif

template <class T>
int launch_cfg(const Plan& plan, const CUtensorMap& map, int mode, cudaStream_t st) {
#define V(U, D, R) if (plan.up == U && plan.down == D && plan.rx == R && plan.ry == R) return launch_mode<T, Cfg<U, D, R, R>>(plan, map, mode, st);
#ifdef FLR_QUICK      // development builds: only the CNO "I" variant
    V(2, 2, 1)
#else
    V(2, 2, 0) V(2, 2, 1) V(2, 4, 0) V(2, 4, 1) V(4, 2, 0) V(4, 2, 1) V(4, 2, 2) V(4, 2, 3)
#endif
#undef V
    return -1;
}

}  // namespace flr_impl
