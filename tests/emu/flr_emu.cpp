// Sequential CPU emulation of the fused filtered_lrelu kernel: compiles the SAME core header the
// CUDA kernel uses (pregenerating-pde_b200/csrc/flrelu_core.h) with g++, replaces TMA by a
// zero-filling copy and the thread grid by loops.  Test infrastructure: validates the polyphase /
// band / mask index mathematics without a GPU.  fp32 only.
#include "../../pregenerating-pde_b200/csrc/flrelu_host.h"
#include <vector>
#include <stdio.h>

using namespace flr;

template <class C, int MODE, bool AFF = false>
static void run(const Plan& plan, const float* x, float* y, const float* bias) {
    const Params& p = plan.p;
    const int nthreads = plan.threads;
    const size_t stage_el = (size_t)p.nbox * p.P * C::RS * p.bwx;
    std::vector<float> xst((size_t)NSTAGE * stage_el);
    std::vector<float> D((size_t)2 * p.P * C::KO * p.dp, NAN);   // NaN: catches reads of unwritten cells
    auto tma = [&](int slot, int row, int plane0) {
        for (int bx = 0; bx < p.nbox; bx++) {
            float* dst = xst.data() + (size_t)slot * stage_el + (size_t)bx * p.P * C::RS * p.bwx;
            for (int u = 0; u < p.P; u++) for (int r = 0; r < C::RS; r++) for (int cidx = 0; cidx < p.bwx; cidx++) {
                const int pl = plane0 + u, gy = row + r, gx = p.tmaX + bx * p.bstep + cidx;
                float v = 0.f;
                if (pl < p.NC && gy >= 0 && gy < p.H && gx >= 0 && gx < p.W) v = x[((size_t)pl * p.H + gy) * p.W + gx];
                dst[((size_t)u * C::RS + r) * p.bwx + cidx] = v;
            }
        }
    };
    const size_t mbuf = (size_t)p.P * C::PB * p.mstride;
    std::vector<uint8_t> M(2 * mbuf + 16, 0xAA);                     // 0xAA = "clamped" everywhere: catches reads of unstaged bytes
    std::vector<uint8_t> W((size_t)C::PB * stage_pitch<C>(p.msp), 0xFF);                // WRITE: staged sign bytes of one body
    auto copy4 = [](uint8_t* dst, const uint8_t* src, bool ok) { if (ok) memcpy(dst, src, 4); else memset(dst, 0, 4); };
    auto store4 = [&](int plane, int o, int ox, const float* out) {
        for (int i = 0; i < 4; i++) if (ox + i < p.OW) y[((size_t)plane * p.OH + o) * p.OW + ox + i] = out[i];
    };
    for (int blk = 0; blk < plan.blocks; blk++) {
        const int r0 = cta_cut(p, blk, C::ALIGN, C::SNAP), r1 = cta_cut(p, blk + 1, C::ALIGN, C::SNAP);
        int gs0 = 0;
        for (int gr = r0; gr < r1;) {
            const CtaGeom cg = seg_geom<C>(p, gr, r1);
            std::vector<Stream<float, C, MODE, AFF>> th(nthreads);
            int next_load = 0;
            auto issue = [&](int s) { tma((gs0 + s) % NSTAGE, cg.row0 + s * C::RS, cg.plane0); };
            while (next_load < cg.nstages_total && next_load < NSTAGE) issue(next_load++);
            for (int t = 0; t < nthreads; t++) {
                th[t].init(p, cg, t, [&](int ch) { return bias[ch]; });
                if (FLR_IS_READ(MODE) && th[t].active) stage_mask<C>(p, th[t].sm_lane, th[t].sm_dst, th[t].sm_src, cg.j_b0, M.data(), copy4);
                if (th[t].active) th[t].prologue(p, xst.data() + (size_t)(gs0 % NSTAGE) * stage_el + (size_t)th[t].unit * C::RS * p.bwx, cg.row0);
            }
            for (int k = 0; k < cg.nbodies; k++) {
                const int s = 1 + k / C::M, slot = (gs0 + s) % NSTAGE;
                if (FLR_IS_READ(MODE) && k + 1 < cg.nbodies) for (int t = 0; t < nthreads; t++) if (th[t].active)
                    stage_mask<C>(p, th[t].sm_lane, th[t].sm_dst, th[t].sm_src, cg.j_b0 + (k + 1) * C::PB, M.data() + (size_t)((k + 1) & 1) * mbuf, copy4);
                const int grow0 = cg.row0 + s * C::RS + (k % C::M) * C::PQ;
                const bool interior = grow0 >= 0 && grow0 + C::PQ <= p.H;
                for (int t = 0; t < nthreads; t++) if (th[t].active) {
                    const float* xt = xst.data() + (size_t)slot * stage_el + ((size_t)th[t].unit * C::RS + (k % C::M) * C::PQ) * p.bwx;
                    float* Dk = D.data() + ((size_t)(k & 1) * p.P + th[t].unit) * C::KO * p.dp;
                    const uint8_t* Mk = M.data() + (size_t)(k & 1) * mbuf + (size_t)th[t].unit * C::PB * p.mstride;
                    if (interior) th[t].template body<false, 0>(p, xt, grow0, cg.j_b0 + k * C::PB, Dk, Mk, W.data() + t);
                    else th[t].template body<true, -1>(p, xt, grow0, cg.j_b0 + k * C::PB, Dk, Mk, W.data() + t);
                }
                if (MODE == SIGN_WRITE) for (int t = 0; t < nthreads; t++) if (th[t].active) mask_out<C>(p, th[t].mo_src, th[t].mo_dst, cg.j_b0 + k * C::PB, W.data());
                const int consumed = 1 + (k + 1) / C::M;
                while (next_load < cg.nstages_total && next_load < consumed + NSTAGE) issue(next_load++);
                for (int t = 0; t < nthreads; t++) if (th[t].active)
                    hdown_phase<float, C>(p, cg, cg.plane0 + th[t].unit, th[t].unit, th[t].g, k,
                                          D.data() + ((size_t)(k & 1) * p.P + th[t].unit) * C::KO * p.dp, store4);
            }
            gs0 += cg.nstages_total;
            gr += cg.o1 - cg.o0;
        }
    }
}

template <class C>
static void run_mode(const Plan& plan, int mode, const float* x, float* y, const float* b) {
    if (plan.p.pa && mode == SIGN_WRITE) run<C, SIGN_WRITE, true>(plan, x, y, b);
    else if (plan.p.pa) run<C, SIGN_NONE, true>(plan, x, y, b);
    else if (mode == SIGN_WRITE) run<C, SIGN_WRITE>(plan, x, y, b);
    else if (mode == SIGN_READ) run<C, SIGN_READ>(plan, x, y, b);
    else if (mode == SIGN_READ_SO) run<C, SIGN_READ_SO>(plan, x, y, b);
    else run<C, SIGN_NONE>(plan, x, y, b);
}

extern "C" int flr_emulate_affine(const float* x, float* y, const float* b, uint8_t* mask, int n, int c, int in_h, int in_w,
                                  int out_h, int out_w, const float* fu, int fu_taps, const float* fd, int fd_taps, int up,
                                  int down, int px0, int py0, int mask_h, int mask_w, int mask_pitch, int sx, int sy,
                                  int mode, float gain, float slope, float clamp, int flip, int ctas, int* info,
                                  const float* pa, const float* pd);

extern "C" int flr_emulate(const float* x, float* y, const float* b, uint8_t* mask, int n, int c, int in_h, int in_w,
                           int out_h, int out_w, const float* fu, int fu_taps, const float* fd, int fd_taps, int up,
                           int down, int px0, int py0, int mask_h, int mask_w, int mask_pitch, int sx, int sy,
                           int mode, float gain, float slope, float clamp, int flip, int ctas, int* info) {
    return flr_emulate_affine(x, y, b, mask, n, c, in_h, in_w, out_h, out_w, fu, fu_taps, fd, fd_taps, up, down, px0, py0, mask_h, mask_w,
                              mask_pitch, sx, sy, mode, gain, slope, clamp, flip, ctas, info, nullptr, nullptr);
}

// the AFF variants (CIN affine folded into the x load): input = pa[plane] * x + pd[plane]
extern "C" int flr_emulate_affine(const float* x, float* y, const float* b, uint8_t* mask, int n, int c, int in_h, int in_w,
                                  int out_h, int out_w, const float* fu, int fu_taps, const float* fd, int fd_taps, int up,
                                  int down, int px0, int py0, int mask_h, int mask_w, int mask_pitch, int sx, int sy,
                                  int mode, float gain, float slope, float clamp, int flip, int ctas, int* info,
                                  const float* pa, const float* pd) {
    Plan plan;
    if (make_plan(plan, 4, n, c, in_h, in_w, out_h, out_w, fu, fu_taps, fd, fd_taps, up, down, px0, py0, gain, slope,
                  clamp, flip, ctas) != 0) return -1;
    plan.p.b = b; plan.p.mask = mask; plan.p.mask_h = mask_h; plan.p.mask_w = mask_w; plan.p.mask_pitch = mask_pitch;
    plan.p.sx = sx; plan.p.sy = sy; plan.p.pa = pa; plan.p.pd = pd;
    if (mode == SIGN_READ || mode == SIGN_READ_SO) set_mask_window(plan, sx);
    if (info) { info[0] = plan.p.P; info[1] = plan.threads; info[2] = plan.blocks; info[3] = plan.p.ncta; info[4] = (int)plan.smem_bytes; }
#define V(U, D, R) if (plan.up == U && plan.down == D && plan.rx == R && plan.ry == R) { run_mode<Cfg<U, D, R, R>>(plan, mode, x, y, b); return 0; }
    V(2, 2, 0) V(2, 2, 1) V(2, 4, 0) V(2, 4, 1) V(4, 2, 0) V(4, 2, 1) V(4, 2, 2) V(4, 2, 3)
#undef V
    return -1;
}
