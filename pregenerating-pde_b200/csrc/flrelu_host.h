// Host-side planning for the fused filtered_lrelu kernel: polyphase tap tables, tile geometry and
// launch shape.  Pure C++ (shared by flrelu_fused.cu and the CPU emulation test).
#pragma once
#include "flrelu_core.h"
#include <stdlib.h>
#include <string.h>

namespace flr {

struct Plan {
    Params p;
    int threads;
    size_t smem_bytes;
    int blocks;
    int up, down, rx, ry;     // kernel variant
    int box_rows;             // TMA box: [P][box_rows][bwx]
    int elem_size;
};

static inline int pos_mod_i(int a, int b) { int r = a % b; return r < 0 ? r + b : r; }
static inline int floor_div_i(int a, int b) { return (a - pos_mod_i(a, b)) / b; }

// rows per x stage for a variant (must match Cfg<>::RS)
static inline int variant_rs(int up, int down, int ry) {
    const int wly = 6 + (ry < up - 1 ? 1 : 0);
    const int pq = 6 * down / up;
    const int m = (wly - 1 + pq - 1) / pq;
    return m * pq;
}

// which (up, down, rx, ry) variants are compiled (flrelu_fused.cu instantiates exactly these)
static inline bool variant_compiled(int up, int down, int rx, int ry) {
    if (up == 2 && down == 2) return rx == ry && (rx == 0 || rx == 1);
    if (up == 2 && down == 4) return rx == ry && (rx == 0 || rx == 1);
    if (up == 4 && down == 2) return rx == ry && rx >= 0 && rx <= 3;
    return false;
}

// number of persistent CTAs: every resident slot, but at least two bodies of output rows per CTA
static inline void set_partition(Plan& plan, int max_ctas, bool forced = false) {
    Params& p = plan.p;
    const int min_rows = p.OH < 12 ? p.OH : 12;            // at least two bodies of output rows (or one short plane) per CTA
    int n = forced ? max_ctas : p.rows_total / min_rows;
    if (n > max_ctas) n = max_ctas;
    if (n > p.rows_total) n = p.rows_total;
    if (n < 1) n = 1;
    p.ncta = n; plan.blocks = n;
}

// READ mode: window of mask bytes a row of threads touches, for the sign offset sx (call after make_plan)
static inline void set_mask_window(Plan& plan, int sx) {
    const int wb0 = floor_div_i(sx, 4);                // byte of column group 0
    plan.p.mws = floor_div_i(wb0, 4) * 4;
}

// returns 0 and fills `plan`, or -1 when the fused kernel does not cover the arguments
static inline int make_plan(Plan& plan, int elem_size, int n, int c, int in_h, int in_w, int out_h, int out_w,
                            const float* fu, int fu_taps, const float* fd, int fd_taps, int up, int down,
                            int px0, int py0, float gain, float slope, float clamp, int flip,
                            int forced_ctas /* 0 = heuristic */) {
    memset(&plan, 0, sizeof(plan));
    if (!((up == 2 || up == 4) && (down == 2 || down == 4))) return -1;
    if (fu_taps != 6 * up || fd_taps != 6 * down) return -1;
    const int rx = pos_mod_i(px0, up), ry = pos_mod_i(py0, up);
    if (!variant_compiled(up, down, rx, ry)) return -1;
    if (n <= 0 || c <= 0 || in_h <= 0 || in_w <= 0 || out_h <= 0 || out_w <= 0) return -1;
    if ((in_w * elem_size) % 16 != 0) return -1;                      // TMA global row stride
    if ((int64_t)n * c > 0x7fffffff / 4) return -1;
    Params& p = plan.p;
    p.NC = n * c; p.C = c; p.H = in_h; p.W = in_w; p.OH = out_h; p.OW = out_w;
    // up-res size implied by the output size: U = (out-1)*down + fd_taps  ..  up to +down-1
    // (the caller guarantees out = (U - (Td-1) + down-1)/down); we compute the up-res extent the
    // mask covers: sh = out*down - (down-1) + (Td-1)  (filtered_lrelu.cpp:81-85)
    p.UH = out_h * down - (down - 1) + (fd_taps - 1);
    p.UW = out_w * down - (down - 1) + (fd_taps - 1);
    p.G = (p.UW + 15) / 16 * 4;          // one thread per mask byte, mask rows padded to 16 elements
                                         // (filtered_lrelu.cpp:83-85): the pad bytes are written as zeros
    if (p.G > FLR_MAXT) return -1;
    p.cminX = -floor_div_i(px0, up);
    p.cminY = -floor_div_i(py0, up);
    const int epv = 16 / elem_size;                                    // elements per 16 bytes
    p.tmaX = floor_div_i(p.cminX, epv) * epv;
    p.xshift = p.cminX - p.tmaX;
    int bwx = ((up == 2) ? 2 * p.G + 6 : p.G + 6) + p.xshift + 1;     // +1: the odd-xshift path reads one float2 further
    bwx = (bwx + epv - 1) / epv * epv;
    p.nbox = 1; p.bstep = 1 << 30;
    if (bwx > 256) {                                                   // TMA box limit: several overlapping boxes per row strip
        p.bstep = 256 - 16;
        p.nbox = (bwx - 16 + p.bstep - 1) / p.bstep;
        bwx = 256;
    }
    p.bwx = bwx;
    p.ncq = (out_w + 3) / 4;
    {   // D row layout (flrelu_core.h Params::dq): DOWN classes of dq float4 each
        const int nf4 = (down * 3 + 6 * down + 3) / 4;                 // float4 loads of one h-down task (Cfg::NF4)
        int q = (p.G + down - 1) / down;
        const int need = p.ncq + (nf4 - 1) / down + 1;
        if (q < need) q = need;
        while (q % 8 != 8 / down) q++;
        p.dq = q; p.dp = 4 * down * q;
    }
    p.ncq_shift = -1;
    for (int s = 0; s < 16; s++) if ((1 << s) == p.ncq) p.ncq_shift = s;
    p.hstep = p.G / p.ncq; if (p.hstep > 6) p.hstep = 6; if (p.hstep < 1) return -1;
    const int rs = variant_rs(up, down, ry);
    const int mstride = ((p.G + 4 + 3) / 4) * 4;                      // staged mask row (READ mode): G + 1 bytes at any 4-byte phase
    p.mstride = mstride; p.mnw = mstride / 4; p.mws = 0;
    p.mlanes = p.G / p.mnw; if (p.mlanes < 1) p.mlanes = 1;
    // planes per CTA: fill up to FLR_MAXT threads (one big CTA per SM balances the four schedulers best), stay inside FLR_SMEM_KB
    int P = FLR_MAXT / p.G; if (P < 1) P = 1; if (P > p.NC) P = p.NC;
    {   // widest CTA (best scheduler balance, fewest wasted lanes) that still cuts the global row space into at least two
        // CTA-sized ranges per SM; short planes end up with fewer planes per CTA, i.e. more, smaller, concurrent CTAs
        static int forced = -1;
        if (forced < 0) { const char* e = getenv("CNO_FLR_P"); forced = e ? atoi(e) : 0; }
        const int p_min = (128 + p.G - 1) / p.G;
        const int min_rows = out_h < 12 ? out_h : 12;
        while (P > p_min && (int64_t)((p.NC + P - 1) / P) * out_h / min_rows < 2 * 148) P--;
        if (forced > 0 && forced < P) P = forced;
    }
    for (;; P--) {
        const size_t stage = ((size_t)p.nbox * P * rs * bwx * elem_size + 127) / 128 * 128;
        const size_t mread = (size_t)2 * P * (6 * down) * mstride + 16;      // READ staging, +16: read_code loads one word past a row
        const int msp = down == 4 ? (P * p.G + 15) / 16 * 16 : MSP;            // WRITE staging pitch (flrelu_core.h stage_pitch)
        const size_t mwrite = (size_t)NDBUF * (6 * down) * msp;
        const size_t total = NSTAGE * stage + (size_t)NDBUF * P * 6 * p.dp * 4 + 64 + (mread > mwrite ? mread : mwrite);
        if (total <= FLR_SMEM_KB * 1024 || P == 1) { p.stage_bytes = (int)stage; p.msp = msp; plan.smem_bytes = total; break; }
    }
    if (plan.smem_bytes > 227 * 1024) return -1;
    p.P = P;
    plan.threads = (P * p.G + 31) / 32 * 32;
    if (plan.threads > FLR_MAXT) return -1;
    // persistent CTAs over the global row space (flrelu_core.h cta_cut); finalised by set_partition()
    p.rows_total = ((p.NC + P - 1) / P) * out_h;
    set_partition(plan, forced_ctas > 0 ? forced_ctas : 2 * 148, forced_ctas > 0);
    if (clamp >= 0.f && clamp < 3.0e38f) return -1;                   // clamp: generic composition
    if (!(slope >= 0.f && slope <= 1.f)) return -1;                   // lrelu as max(v, slope*v)
    p.slope = slope;
    // ---- tap tables (correlation order; reference flips unless flip_filter, upfirdn2d.py:190-191)
    float gu[32], gd[32];
    for (int t = 0; t < fu_taps; t++) gu[t] = flip ? fu[t] : fu[fu_taps - 1 - t];
    for (int t = 0; t < fd_taps; t++) gd[t] = flip ? fd[t] : fd[fd_taps - 1 - t];
    for (int t = 0; t < 32; t++) p.fd[t] = t < fd_taps ? gd[t] : 0.f;
    // horizontal: phase psi = column mod up; first tap t0 = (px0 - psi) mod up; window offset off = psi > rx
    for (int hp = 0; hp < up / 2; hp++)
        for (int lane = 0; lane < 2; lane++) {
            const int psi = 2 * hp + lane;
            const int t0 = pos_mod_i(px0 - psi, up), off = psi > rx ? 1 : 0;
            for (int k = 0; k < 8; k++) {
                const int kk = k - off, t = t0 + up * kk;
                p.hx[hp][k][lane] = (kk >= 0 && kk < 6 && t < fu_taps) ? gu[t] * (float)up : 0.f;
            }
        }
    for (int phi = 0; phi < up; phi++) {
        const int t0 = pos_mod_i(py0 - phi, up);
        for (int k = 0; k < 8; k++) {
            const int t = t0 + up * k;
            p.vy[phi][k] = (k < 6 && t < fu_taps) ? gu[t] * (float)up * gain : 0.f;
        }
    }
    plan.up = up; plan.down = down; plan.rx = rx; plan.ry = ry; plan.box_rows = rs; plan.elem_size = elem_size;
    return 0;
}

}  // namespace flr
