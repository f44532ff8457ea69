// Core of the fused filtered_lrelu kernel, written so that the SAME source compiles
//   - under nvcc as the device code of flrelu_fused.cu (registers + shared memory + FFMA2), and
//   - under g++ as a sequential emulation used by tests/test_fused_emulation.py to check the index
//     mathematics on the CPU (test infrastructure only; the product path is the CUDA kernel).
//
// Algorithm (new design, not the reference's tile scheme of filtered_lrelu.cu:131-1091):
//   a thread owns FOUR adjacent columns of the up-sampled image (= one byte of the 2-bit sign mask)
//   and streams down the rows of its plane:
//     h-up   : polyphase horizontal up-FIR of the new input row from the TMA-staged x tile   (registers)
//     v-up   : polyphase vertical up-FIR from a register window of the last 6-7 h-up rows     (registers)
//     act    : gain (folded into the v-up taps), leaky ReLU, clamp, sign bits written / read  (registers)
//     v-down : each activated row is scattered into 6 rotating output-row accumulators        (registers)
//   finished rows go to shared memory once (D buffer), and a second phase applies the horizontal
//   down-FIR and writes y with 16-byte stores.  Zero halo recompute horizontally, the 266x266
//   intermediate of a 128x128 plane never leaves registers, all FIR arithmetic is packed
//   fma.rn.f32x2 with taps coming from kernel-parameter (uniform) registers.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define FLR_DEV __device__ __forceinline__
#define FLR_UNROLL _Pragma("unroll")
#define FLR_NOUNROLL _Pragma("unroll 1")
#else
#define FLR_DEV inline
#define FLR_UNROLL
#define FLR_NOUNROLL
#include <math.h>
#include <string.h>
#endif

namespace flr {

struct f2 { float x, y; };
struct alignas(16) f4 { float x, y, z, w; };

FLR_DEV f2 mk2(float a, float b) { f2 r; r.x = a; r.y = b; return r; }
FLR_DEV float f4c(const f4& v, int c) { return c == 0 ? v.x : c == 1 ? v.y : c == 2 ? v.z : v.w; }

// d = a * b + c on both lanes.  Device: one FFMA2 (ptxas folds broadcast operands to .F32 form).
FLR_DEV f2 ffma2(f2 a, f2 b, f2 c) {
#if defined(__CUDA_ARCH__) && !defined(FLR_NO_FFMA2)
    f2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;"
        : "=l"(*reinterpret_cast<unsigned long long*>(&d))
        : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)),
          "l"(*reinterpret_cast<unsigned long long*>(&c)));
    return d;
#else
    return mk2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y));
#endif
}

// (hi:lo) << s, upper word  (SHF.L.W on the device)
FLR_DEV uint32_t funnel_l(uint32_t lo, uint32_t hi, int s) {
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, hi, s);
#else
    return (hi << s) | (lo >> (32 - s));
#endif
}

// (hi:lo) >> s, lower word, 0 <= s < 32  (SHF.R.W on the device)
FLR_DEV uint32_t funnel_r(uint32_t lo, uint32_t hi, int s) {
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, s);
#else
    return s ? (lo >> s) | (hi << (32 - s)) : lo;
#endif
}

FLR_DEV uint32_t f_bits(float v) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(v);
#else
    uint32_t u; memcpy(&u, &v, 4); return u;
#endif
}

// SIGN_READ_SO: the caller guarantees the mask holds sign codes only (it was written with clamp = none): the "clamped" bit is not tested
enum { SIGN_NONE = 0, SIGN_WRITE = 1, SIGN_READ = 2, SIGN_READ_SO = 3 };
#define FLR_IS_READ(MODE) (((MODE) & 2) != 0)

// element -> fp32 (specialised for bf16 / fp16 by flrelu_fused.cu)
template <class T> struct Cvt;
template <> struct Cvt<float> { static FLR_DEV float get(float v) { return v; } };


// build-time tunables (defaults = the shipped configuration; tools/variants.sh builds others for A/B timing)
#ifndef FLR_NSTAGE
#define FLR_NSTAGE 2
#endif
#ifndef FLR_MAXT
#define FLR_MAXT 512            // threads per CTA (launch bound)
#endif
#ifndef FLR_MINB
#define FLR_MINB 1              // resident CTAs per SM the register budget is sized for
#endif
#ifndef FLR_SMEM_KB
#define FLR_SMEM_KB 222         // dynamic shared memory budget per CTA (227 KB per block on sm_100 minus the 2 KB static s_db and slack)
#endif
constexpr int NSTAGE = FLR_NSTAGE;
constexpr int NDBUF = 3;           // D buffers (finished v-down rows) in flight, see flr_fused_kernel
constexpr int MSP = FLR_MAXT;      // WRITE (DOWN = 2): row pitch of the staged sign bytes, layout [up-res row of the body][thread]: a thread's
                                   // byte of row jl sits at the compile-time offset jl * MSP from its own slot (one STS.U8, no
                                   // address arithmetic); mask_out() copies finished bodies to HBM with 32-bit stores.  DOWN = 4 bodies
                                   // have 24 rows: three such buffers would cost 36 KB and halve the planes per CTA of the wide D
                                   // configuration, so they use the runtime pitch Params::msp (one pointer step per row instead)
template <class C> FLR_DEV int stage_pitch(int msp) { return C::DOWN == 4 ? msp : MSP; }

// Everything the kernel needs besides the tensor map.  Filled by the host (flrelu_fused.cu).
struct Params {
    // tap tables first, 16-byte aligned, so that (lane0, lane1) pairs are aligned 64-bit operands
    alignas(16) float hx[2][8][2];   // h-up taps  [phase pair][k][lane], zero padded, scaled by UP
    alignas(16) float vy[4][8];      // v-up taps  [phase][k], scaled by UP * gain
    alignas(16) float fd[32];        // down taps in correlation order
    void* y; const void* b; uint8_t* mask; float* db;
    const float* pa; const float* pd;   // AFF variants: per-plane affine applied to x before the bias: input = pa[plane] * x + pd[plane]
                                        // (the CIN output); nullptr otherwise
    int NC, C, H, W, UH, UW, OH, OW;
    int G, P;                  // column groups per plane (ceil(UW/4)); planes per CTA
    int ncta, rows_total;      // persistent CTAs; rows_total = plane groups * OH ("global row space", see cta_rows)
    int cminX, cminY;          // first input column / row offset of the polyphase windows
    int tmaX, xshift;          // TMA start column = cminX rounded down to a 16-byte boundary (the hardware faults on an
                               // unaligned innermost coordinate; measured, tools/probes/tma_probe.cu); xshift = cminX - tmaX
    int bwx;                   // x tile row pitch (elements) = TMA box width
    int nbox, bstep;           // planes wider than one TMA box (256 elements) are staged as nbox boxes of bwx columns whose start
                               // columns are bstep = bwx - 16 apart (16 columns of overlap, so every thread finds its whole
                               // window inside ONE box); smem stage layout [nbox][P][RS][bwx]
    int dp, dq;                // D row pitch in floats = 4 * DOWN * dq.  A finished row is stored as DOWN interleaved classes:
                               // float4 f sits at float4 position (f / DOWN) + (f % DOWN) * dq, dq = 8 / DOWN (mod 8), so that the
                               // writers (consecutive f) and the h-down readers (f = DOWN*cq + i) are both bank-conflict free and
                               // a reader's NF4 loads are constant offsets from one address
    int ncq, ncq_shift;        // output column quads per row = ceil(OW/4); log2 when a power of two, else -1
    int msp;                   // WRITE, DOWN = 4: row pitch of the staged sign bytes = threads rounded up to 16 (stage_pitch)
    int hstep;                 // h-down: row lanes per plane slot = min(G / ncq, KO) (hdown_phase)
    int stage_bytes;           // bytes of one x stage (multiple of 128)
    int mask_h, mask_w, mask_pitch, sx, sy;
    int mlanes;                // READ: row lanes of stage_mask per plane slot = G / mnw (>= 1)
    int mws, mnw, mstride;     // READ: staged mask window per row: first byte (multiple of 4, may be < 0), 32-bit words, smem row stride
    float slope;
};

template <int UP_, int DOWN_, int RX_, int RY_>
struct Cfg {
    static constexpr int UP = UP_, DOWN = DOWN_, RX = RX_, RY = RY_;
    static constexpr int NTU = 6, ND = 6;                    // taps per up phase, rows in flight per down FIR
    static constexpr int TD = ND * DOWN;
    static constexpr int WLX = NTU + (RX < UP - 1 ? 1 : 0);
    static constexpr int WLY = NTU + (RY < UP - 1 ? 1 : 0);
    static constexpr int NXS = WLX + (UP == 2 ? 1 : 0);      // x samples a thread reads per input row
    static constexpr int PB = ND * DOWN;                     // up-res rows per body
    static constexpr int PQ = PB / UP;                       // input rows per body
    static constexpr int KO = ND;                            // output rows per body
    static constexpr int M = (WLY - 1 + PQ - 1) / PQ;        // bodies per x stage
    static constexpr int RS = M * PQ;                        // input rows per x stage (>= WLY-1)
    static constexpr int NF4 = (DOWN * 3 + TD + 3) / 4;      // float4 loads per h-down task
    static constexpr int ALIGN = (DOWN % UP == 0) ? 1 : UP / DOWN;   // a segment must start on a whole input row
    static constexpr int SNAP = 2 * KO;
    static_assert(PB % UP == 0, "body must hold whole input rows");
};

// Work partition.  The output rows of all plane groups form one global row space
//   gr = plane_group * OH + o,   0 <= gr < rows_total,
// cut into `ncta` nearly equal contiguous ranges, one per persistent CTA (the grid is sized to the
// resident CTA slots of the 148 SMs, so there is no wave tail).  A range is processed as a sequence of
// segments (one plane group, output rows [o0, o1)); a segment that starts inside a plane pays one
// warm-up body.  Cuts closer than SNAP rows to a plane boundary snap onto it.
FLR_DEV int cta_cut(const Params& p, int c, int align, int snap) {
    if (c >= p.ncta) return p.rows_total;
    const long long t = (long long)c * p.rows_total / p.ncta;
    int pg = (int)(t / p.OH), o = (int)(t - (long long)pg * p.OH);
    if (o < snap) o = 0;
    else if (p.OH - o < snap) { o = 0; pg++; }
    o -= o % align;
    return pg * p.OH + o;
}

// geometry of one segment (uniform values)
struct CtaGeom {
    int plane0, o0, o1, j_b0, q_b0, row0 /* global input row of stage 0, row 0 */, nbodies, nstages_total;
};

template <class C>
FLR_DEV CtaGeom seg_geom(const Params& p, int gr, int r1) {
    CtaGeom g;
    const int pg = gr / p.OH;
    g.plane0 = pg * p.P;
    g.o0 = gr - pg * p.OH;
    g.o1 = g.o0 + (r1 - gr) < p.OH ? g.o0 + (r1 - gr) : p.OH;
    g.j_b0 = C::DOWN * g.o0;
    g.q_b0 = g.j_b0 / C::UP;
    g.row0 = g.q_b0 + p.cminY + (C::WLY - 1) - C::RS;
    g.nbodies = (g.o1 - g.o0 + C::ND - 1 + C::KO - 1) / C::KO;
    g.nstages_total = 1 + (g.nbodies + C::M - 1) / C::M;
    return g;
}

template <class C> FLR_DEV int dpos(const Params& p, int f4idx) { return f4idx / C::DOWN + (f4idx % C::DOWN) * p.dq; }

// ------------------------------------------------------------------------------------------------
// per-thread state of the streaming phase (lives in registers on the device)
// ------------------------------------------------------------------------------------------------
template <class T, class C, int MODE, bool AFF = false>
struct Stream {
    f2 win[C::WLY][2];      // h-up rows: window of the vertical up-FIR, two column pairs
    f2 acc[C::ND][2];       // rotating output-row accumulators of the vertical down-FIR
    f2 bE[2];               // bias times the horizontal edge profile of the up-FIR
    f2 aff2;                // AFF: (a, a) of this thread's plane
    int unit, g;            // plane slot in the CTA, column group
    int xofs;               // element offset of this thread's first x sample inside its unit's tile row (includes the box offset)
    int dofs;               // float offset of this thread's float4 inside a D row
    bool active;
    uint32_t colmask;       // WRITE: valid-column bits; READ: valid mask-element bits
    int rd_word, rd_bits;   // READ: 32-bit word of the staged mask row that holds this thread's first code bit, and its bit position
    int sm_dst, sm_lane;    // READ: stage_mask role of this thread inside its plane slot: byte offset of its word column in the staged rows
                            // of the unit, first row it copies (rows sm_lane, sm_lane + mlanes, ...; >= PB: no role)
    const uint8_t* sm_src;  // READ: that word column in row 0 of the plane's mask (nullptr: outside the mask, zero-fill)
    int mo_src;             // WRITE: byte offset of this thread's first staged word inside a staging buffer (mask_out): jl0 * MSP + unit * G + 4 * wc
    uint8_t* mo_dst;        // WRITE: that word's place in the mask of the thread's plane, row jl0 of up-res row 0
    // fused db = y.sum([0,2,3]): sum over the outputs of a row = dot(finished v-down row, column weights), the weight of
    // column c being the sum of the down taps through which c reaches a valid output column
    f2 dbw[2], dbacc[2];
    int db_lo, db_hi;       // rows r of the current body with db_lo <= r < db_hi are real output rows of the segment (set by the caller)

    template <class LD>
    FLR_DEV void init(const Params& p, const CtaGeom& cg, int tid, LD ld_bias) {
        unit = tid / p.G; g = tid - unit * p.G;
        dofs = 4 * dpos<C>(p, g);
        {
            const int xoff = ((C::UP == 2) ? 2 * g : g) + p.xshift;
            int bi = xoff / p.bstep; if (bi > p.nbox - 1) bi = p.nbox - 1;
            xofs = bi * (p.P * C::RS * p.bwx) + (xoff - bi * p.bstep);
        }
        const int plane = cg.plane0 + unit;
        active = unit < p.P && plane < p.NC;
        FLR_UNROLL for (int r = 0; r < C::ND; r++) { acc[r][0] = mk2(0.f, 0.f); acc[r][1] = mk2(0.f, 0.f); }
        FLR_UNROLL for (int w = 0; w < C::WLY; w++) { win[w][0] = mk2(0.f, 0.f); win[w][1] = mk2(0.f, 0.f); }
        bE[0] = bE[1] = mk2(0.f, 0.f);
        colmask = 0; rd_word = 0; rd_bits = 0; mo_src = 0; mo_dst = nullptr; sm_dst = 0; sm_lane = C::PB; sm_src = nullptr;
        dbw[0] = dbw[1] = dbacc[0] = dbacc[1] = mk2(0.f, 0.f); db_lo = db_hi = 0; aff2 = mk2(1.f, 1.f);
        if (!active) return;
        if (FLR_IS_READ(MODE) && p.db) {       // db is a backward quantity: only the mask-reading variant carries the code
            float w[4];
            FLR_UNROLL for (int m = 0; m < 4; m++) {
                const int c = 4 * g + m;
                w[m] = 0.f;
                for (int t = c % C::DOWN; t < C::TD && t <= c; t += C::DOWN)
                    if ((c - t) / C::DOWN < p.OW) w[m] += p.fd[t];
            }
            dbw[0] = mk2(w[0], w[1]); dbw[1] = mk2(w[2], w[3]);
        }
        // bias edge profile: A = hup(x) + b * sum_k [column valid] * tap
        // a constant c added before the zero padding contributes c * (sum of in-bounds taps); the affine offset d joins the bias,
        // the affine scale a multiplies the FIR sum (hup)
        const float bias = (p.b ? ld_bias(plane % p.C) : 0.f) + (AFF ? p.pd[plane] : 0.f);
        if (AFF) { const float a_pl = p.pa[plane]; aff2 = mk2(a_pl, a_pl); }
        FLR_UNROLL for (int pr = 0; pr < 2; pr++) {
            const int qx = (C::UP == 2) ? 2 * g + pr : g;
            const int hp = (C::UP == 2) ? 0 : pr;
            float e0 = 0.f, e1 = 0.f;
            FLR_UNROLL for (int k = 0; k < C::WLX; k++) {
                const int col = qx + p.cminX + k;
                if (col >= 0 && col < p.W) { e0 += p.hx[hp][k][0]; e1 += p.hx[hp][k][1]; }
            }
            bE[pr] = mk2(e0 * bias, e1 * bias);
        }
        if (MODE == SIGN_WRITE) {
            FLR_UNROLL for (int m = 0; m < 4; m++) if (4 * g + m < p.UW) colmask |= 1u << (2 * m);   // bit 0 of each 2-bit code
            {   // mask_out role: G/4 word columns x 4 row lanes per plane slot
                const int nwr = p.G >> 2, jl0 = g / nwr, wc = g - jl0 * nwr;
                mo_src = jl0 * stage_pitch<C>(p.msp) + unit * p.G + 4 * wc;
                mo_dst = p.mask + ((int64_t)plane * p.mask_h + jl0) * p.mask_pitch + 4 * wc;
            }
        } else if (FLR_IS_READ(MODE)) {
            const int e0 = 4 * g + p.sx;
            const int fl = (e0 >= 0) ? e0 / 4 : -((-e0 + 3) / 4);
            rd_word = (fl - p.mws) >> 2; rd_bits = 8 * ((fl - p.mws) & 3) + 2 * (e0 - 4 * fl);     // fl >= mws by construction (set_mask_window)
            FLR_UNROLL for (int m = 0; m < 4; m++) { const int e = e0 + m; if (e >= 0 && e < p.mask_w) colmask |= 3u << (2 * m); }
            {   // stage_mask role: mnw word columns x mlanes row lanes per plane slot
                const int lane = g / p.mnw, wc = g - lane * p.mnw, byte = p.mws + 4 * wc;
                sm_lane = lane < p.mlanes ? lane : C::PB;
                sm_dst = unit * C::PB * p.mstride + 4 * wc;
                sm_src = (byte >= 0 && byte < p.mask_pitch) ? p.mask + (int64_t)plane * p.mask_h * p.mask_pitch + byte : nullptr;
            }
        }
    }

    // horizontal polyphase up-FIR of one staged input row -> two column pairs
    //   XPAR : parity of xshift when the caller knows it at compile time (0 / 1; the streaming body is instantiated for both so
    //          that its rows form one basic block), -1 = test p.xshift (prologue)
    template <int XPAR = -1>
    FLR_DEV void hup(const Params& p, const T* xrow /* tile row of this unit */, bool rowvalid, f2 out[2]) {
        float xs[8];
        const int xoff = xofs;
#if defined(__CUDA_ARCH__)
        if (C::UP == 2 && sizeof(T) == 4) {          // LDS.64 from the even column at or below xoff
            if (XPAR < 0 ? (p.xshift & 1) : XPAR) {  // uniform (the parity of xofs is that of xshift: 2g and bstep are even): LDS.32, 3 x LDS.64, LDS.32
                xs[0] = xrow[xoff];
                FLR_UNROLL for (int w = 0; w < 3; w++) { const float2 v = *reinterpret_cast<const float2*>(xrow + xoff + 1 + 2 * w); xs[1 + 2 * w] = v.x; xs[2 + 2 * w] = v.y; }
                xs[7] = xrow[xoff + 7];
            } else {
                FLR_UNROLL for (int w = 0; w < 4; w++) { const float2 v = *reinterpret_cast<const float2*>(xrow + xoff + 2 * w); xs[2 * w] = v.x; xs[2 * w + 1] = v.y; }
            }
        } else
#endif
        { FLR_UNROLL for (int w = 0; w < C::NXS; w++) xs[w] = Cvt<T>::get(xrow[xoff + w]); }
        FLR_UNROLL for (int pr = 0; pr < 2; pr++) {
            const f2 edge = rowvalid ? bE[pr] : mk2(0.f, 0.f);
            f2 a = AFF ? mk2(0.f, 0.f) : edge;
            const int qo = (C::UP == 2) ? pr : 0;
            const int hp = (C::UP == 2) ? 0 : pr;
            FLR_UNROLL for (int k = 0; k < C::WLX; k++)
                a = ffma2(mk2(xs[qo + k], xs[qo + k]), mk2(p.hx[hp][k][0], p.hx[hp][k][1]), a);
            out[pr] = AFF ? ffma2(a, aff2, edge) : a;
        }
    }

    FLR_DEV void push(const f2 a[2]) {
        FLR_UNROLL for (int w = 0; w + 1 < C::WLY; w++) { win[w][0] = win[w + 1][0]; win[w][1] = win[w + 1][1]; }
        win[C::WLY - 1][0] = a[0]; win[C::WLY - 1][1] = a[1];
    }

    // leaky ReLU with the sign bits of four columns written to / read from `code`.
    // (clamp is not handled by the fused kernel: the host plan refuses it and the caller composes
    //  the generic path, exactly like an unsupported filter shape.)
    FLR_DEV void act(const Params& p, f2 v[2], uint32_t& code) {
        if (FLR_IS_READ(MODE)) {
            float m[4];
            FLR_UNROLL for (int i = 0; i < 4; i++) m[i] = (code & (1u << (2 * i))) ? p.slope : 1.f;
            // "clamped" codes exist only in masks written by the generic forward path with clamp != none; the four extra
            // test / select pairs cost 11 % of the backward launch, so callers that know the mask's origin use SIGN_READ_SO
            if (MODE == SIGN_READ && (code & 0xaau)) {
                FLR_UNROLL for (int i = 0; i < 4; i++) if (code & (2u << (2 * i))) m[i] = 0.f;
            }
            const f2 z = mk2(0.f, 0.f);
            v[0] = ffma2(v[0], mk2(m[0], m[1]), z); v[1] = ffma2(v[1], mk2(m[2], m[3]), z);
        } else {
            if (MODE == SIGN_WRITE) {
                // sign bits of four floats -> bits 0, 2, 4, 6: one shift + three funnel shifts + shift + and
#ifndef FLR_X_NOPACK
                uint32_t r = f_bits(v[1].y) >> 30;
                r = funnel_l(f_bits(v[1].x), r, 2); r = funnel_l(f_bits(v[0].y), r, 2); r = funnel_l(f_bits(v[0].x), r, 2);
                code = (r >> 1) & colmask;
#else
                code = f_bits(v[0].x) >> 31;
#endif
            }
            // lrelu(v) = max(v, slope*v); the host plan only admits 0 <= slope <= 1
            const f2 sl = mk2(p.slope, p.slope), z = mk2(0.f, 0.f);
            const f2 s0 = ffma2(v[0], sl, z), s1 = ffma2(v[1], sl, z);
#ifndef FLR_X_NOMAX
            v[0] = mk2(fmaxf(v[0].x, s0.x), fmaxf(v[0].y, s0.y)); v[1] = mk2(fmaxf(v[1].x, s1.x), fmaxf(v[1].y, s1.y));
#else
            v[0] = s0; v[1] = s1;
#endif
        }
    }

    // READ: sign codes of one up-res row from the staged mask rows (stage_mask); bytes outside the mask are zero there
    FLR_DEV uint32_t read_code(const Params& p, const uint8_t* mrow) const {
        // the eight code bits may straddle two words: one funnel shift over the aligned pair (the staged buffer has slack for w[1])
        const uint32_t* w = reinterpret_cast<const uint32_t*>(mrow) + rd_word;
        return funnel_r(w[0], w[1], rd_bits) & colmask;
    }

    // One body: PQ input rows -> PB up-res rows -> KO finished rows of the vertical down-FIR.
    //   xtile : this unit's rows of the current x stage, first row of the body
    //   grow0 : global input row index of that first row
    //   jb    : global up-res row index of the body's first row
    //   Dbuf  : this unit's KO x dp floats of the D buffer being filled
    //   Mrows : READ: this unit's PB staged mask rows of the body
    //   Mst   : WRITE: this thread's byte of row 0 in the staging buffer of the body (rows MSP apart)
    //   EDGE  : false = the caller guarantees that every input row of the body lies inside the plane (no row-validity selects)
    template <bool EDGE, int XPAR>
    FLR_DEV void body(const Params& p, const T* xtile, int grow0, int jb, float* Dbuf, const uint8_t* Mrows, uint8_t* Mst) {
        FLR_UNROLL for (int jl = 0; jl < C::PB; jl++) {
            const int phi = jl % C::UP, qq = jl / C::UP;
            if (phi == 0) {
                const int grow = grow0 + qq;
                f2 a[2];
                this->template hup<XPAR>(p, xtile + qq * p.bwx, EDGE ? (grow >= 0 && grow < p.H) : true, a);
                push(a);
            }
            const int j = jb + jl;
            uint32_t code = 0;
            if (FLR_IS_READ(MODE)) code = read_code(p, Mrows + jl * p.mstride);
            // vertical polyphase up-FIR (taps carry UP * gain)
            f2 v[2] = {mk2(0.f, 0.f), mk2(0.f, 0.f)};
            FLR_UNROLL for (int k = 0; k < C::NTU; k++) {
                const f2 t = mk2(p.vy[phi][k], p.vy[phi][k]);
                v[0] = ffma2(win[(phi > C::RY ? 1 : 0) + k][0], t, v[0]);
                v[1] = ffma2(win[(phi > C::RY ? 1 : 0) + k][1], t, v[1]);
            }
            act(p, v, code);
            if (MODE == SIGN_WRITE) Mst[jl * stage_pitch<C>(p.msp)] = (uint8_t)code;      // staged; rows beyond UH are dropped by mask_out()
            (void)j;
            // vertical down-FIR, scatter form: row r of this body completes at jl = DOWN*r + DOWN-1
            FLR_UNROLL for (int r = 0; r < C::ND; r++) {
                const int t = (jl <= C::DOWN * r + C::DOWN - 1) ? jl + C::DOWN * (C::ND - 1 - r) : jl - C::DOWN * (r + 1);
                const f2 tt = mk2(p.fd[t], p.fd[t]);
                acc[r][0] = ffma2(v[0], tt, acc[r][0]);
                acc[r][1] = ffma2(v[1], tt, acc[r][1]);
            }
            if (jl % C::DOWN == C::DOWN - 1) {
                const int r = jl / C::DOWN;
                f4 o; o.x = acc[r][0].x; o.y = acc[r][0].y; o.z = acc[r][1].x; o.w = acc[r][1].y;
                *reinterpret_cast<f4*>(Dbuf + r * p.dp + dofs) = o;
                if (FLR_IS_READ(MODE) && r >= db_lo && r < db_hi) { dbacc[0] = ffma2(acc[r][0], dbw[0], dbacc[0]); dbacc[1] = ffma2(acc[r][1], dbw[1], dbacc[1]); }
                acc[r][0] = mk2(0.f, 0.f); acc[r][1] = mk2(0.f, 0.f);
            }
        }
    }

    // warm-up: the WLY-1 rows that precede the first body (last rows of stage 0)
    FLR_DEV void prologue(const Params& p, const T* xstage0_unit, int row0_global) {
        FLR_UNROLL for (int w = 0; w < C::WLY - 1; w++) {
            const int sr = C::RS - (C::WLY - 1) + w;
            const int grow = row0_global + sr;
            f2 a[2];
            hup(p, xstage0_unit + sr * p.bwx, grow >= 0 && grow < p.H, a);
            push(a);
        }
    }
};

// ------------------------------------------------------------------------------------------------
// READ mode: copy the mask rows of one body (up-res rows jb .. jb+PB-1 of every plane of the group, window
// [mws, mws + 4*mnw) bytes) into shared memory; positions outside the mask become zero (= "pass").
// copy4(dst, src, ok) moves one aligned 32-bit word (cp.async on the device) or zero-fills when !ok.
// ------------------------------------------------------------------------------------------------
template <class C, class CP>
FLR_DEV void stage_mask(const Params& p, int sm_lane, int sm_dst, const uint8_t* sm_src, int jb, uint8_t* Mdst, CP copy4) {
    // the G threads of a plane slot stage their own plane: a thread owns one 32-bit word column and every mlanes-th row of the body
    const int my0 = jb + p.sy;
    for (int jl = sm_lane; jl < C::PB; jl += p.mlanes) {
        const int row = my0 + jl;
        const bool ok = sm_src != nullptr && row >= 0 && row < p.mask_h;
        copy4(Mdst + sm_dst + jl * p.mstride, ok ? sm_src + (int64_t)row * p.mask_pitch : p.mask, ok);
    }
}

// ------------------------------------------------------------------------------------------------
// WRITE mode: copy the staged sign bytes of one finished body (layout [PB rows][MSP], thread t's byte at column t, i.e. plane
// slot u occupies bytes [u*G, u*G + G) of every row) to the mask in HBM with 32-bit words.  The G threads of a plane slot
// copy their own plane: G/4 word columns x 4 row lanes, a thread takes rows jl0, jl0 + 4, ... (immediate offsets in shared
// memory, one pointer step per row in HBM).  G is a multiple of 4 and the host checks mask_pitch % 4 == 0: all words aligned.
// ------------------------------------------------------------------------------------------------
template <class C>
FLR_DEV void mask_out(const Params& p, int mo_src, uint8_t* mo_dst, int jb, const uint8_t* Mbuf) {
    const int sp = stage_pitch<C>(p.msp);
    const int jl0 = mo_src / sp;                            // 0..3 (MSP is a power of two; DOWN = 4 divides once per body)
    int rows = p.UH - jb; if (rows > C::PB) rows = C::PB;
    const uint8_t* src = Mbuf + mo_src;
    uint32_t w[C::PB / 4];
    FLR_UNROLL for (int it = 0; it < C::PB / 4; it++) w[it] = *reinterpret_cast<const uint32_t*>(src + 4 * it * sp);
    uint8_t* dst = mo_dst + (uint32_t)(jb * p.mask_pitch);
    FLR_UNROLL for (int it = 0; it < C::PB / 4; it++) {
        if (jl0 + 4 * it < rows) *reinterpret_cast<uint32_t*>(dst) = w[it];
        dst += 4 * p.mask_pitch;
    }
}

// ------------------------------------------------------------------------------------------------
// horizontal down-FIR of the finished rows of one body; writes y.  The G threads of a plane slot process their own plane:
// thread g takes output column quad g % ncq and the rows g / ncq, g / ncq + hstep, ... of the KO finished rows
// (hstep = min(G / ncq, KO) row lanes, set by the host; G >= DOWN * ncq, so there are at least DOWN lanes).  No task index has to
// be decomposed per output and only D rows written by the plane's own threads are read.
// ------------------------------------------------------------------------------------------------
template <class T, class C, class ST>
FLR_DEV void hdown_phase(const Params& p, const CtaGeom& cg, int plane, int unit, int g, int body_idx,
                         const float* Dunit /* this plane slot's [KO][dp] */, ST store4) {
    int r, cq;
    if (p.ncq_shift >= 0) { cq = g & (p.ncq - 1); r = g >> p.ncq_shift; }
    else { r = g / p.ncq; cq = g - r * p.ncq; }
    if (r >= p.hstep) return;
    const int o_base = cg.o0 - (C::ND - 1) + body_idx * C::KO;
    const unsigned nrows_ok = (unsigned)(cg.o1 - cg.o0);
    for (; r < C::KO; r += p.hstep) {
        const int o = o_base + r;
        if ((unsigned)(o - cg.o0) >= nrows_ok) continue;
        const f4* drow = reinterpret_cast<const f4*>(Dunit + r * p.dp) + cq;
        f4 d[C::NF4];
        FLR_UNROLL for (int i = 0; i < C::NF4; i++) d[i] = drow[(i / C::DOWN) + (i % C::DOWN) * p.dq];
        float out[4];
        FLR_UNROLL for (int i = 0; i < 4; i++) {
            f2 a = mk2(0.f, 0.f);
            FLR_UNROLL for (int m = 0; m < C::TD / 2; m++) {
                const int e = C::DOWN * i + 2 * m;     // even: both lanes sit in the same float4
                a = ffma2(mk2(f4c(d[e >> 2], e & 3), f4c(d[e >> 2], (e & 3) + 1)), mk2(p.fd[2 * m], p.fd[2 * m + 1]), a);
            }
            out[i] = a.x + a.y;
        }
        store4(plane, o, 4 * cq, out);
    }
}

}  // namespace flr
