// C entry points of the fused filtered_lrelu kernel (kernels are instantiated per dtype in
// flrelu_fused_{f32,bf16,f16}.cu so the translation units build in parallel).
#include "flrelu_fused_impl.cuh"
#include "flrelu_plane.cuh"

namespace flr_impl {
extern template int launch_cfg<float>(const Plan&, const CUtensorMap&, int, cudaStream_t);
extern template int launch_cfg<__nv_bfloat16>(const Plan&, const CUtensorMap&, int, cudaStream_t);
extern template int launch_cfg<__half>(const Plan&, const CUtensorMap&, int, cudaStream_t);
}
namespace flr_plane {
extern template int launch_cfg<float>(const Plan&, const Geom&, const void*, int, cudaStream_t);
extern template int launch_cfg<__nv_bfloat16>(const Plan&, const Geom&, const void*, int, cudaStream_t);
extern template int launch_cfg<__half>(const Plan&, const Geom&, const void*, int, cudaStream_t);
}
using namespace flr;
using namespace flr_impl;

// small planes take the whole-plane kernel (flrelu_plane.cuh) unless it is switched off: environment CNO_FLR_NO_PLANE at first use,
// or cno_filtered_lrelu_plane_kernel(0) (tests run both kernels on the same inputs)
static EnvFlag g_no_plane;
static std::atomic<int> g_plane_switch{-1};      // -1: follow the environment
extern "C" int cno_filtered_lrelu_plane_kernel(int enable) {
    const int prev = g_plane_switch.load() < 0 ? (g_no_plane.on("CNO_FLR_NO_PLANE") ? 0 : 1) : g_plane_switch.load();
    if (enable >= 0) g_plane_switch.store(enable ? 1 : 0);
    return prev;
}


extern "C" int cno_filtered_lrelu_query(int dtype, int in_h, int in_w, int out_h, int out_w, int fu_taps,
                                        int fd_taps, int up, int down, int px0, int py0) {
    if (dtype != CNO_F32 && dtype != CNO_BF16 && dtype != CNO_F16) return -1;
    float zeros[32] = {0};
    Plan plan;
    if (fu_taps > 32 || fd_taps > 32 || fu_taps < 1 || fd_taps < 1) return -1;
    return make_plan(plan, cno_dtype_size(dtype), 1, 1, in_h, in_w, out_h, out_w, zeros, fu_taps, zeros, fd_taps, up, down,
                     px0, py0, 1.f, 0.f, -1.f, 0, 0);
}

extern "C" int cno_filtered_lrelu_affine(const void* x, void* y, const void* b, uint8_t* mask, float* db,
                                         const float* fu, const float* fd, int dtype,
                                         int n, int c, int in_h, int in_w, int out_h, int out_w,
                                         int fu_taps, int fd_taps, int up, int down, int px0, int py0,
                                         int mask_h, int mask_w, int mask_pitch, int sx, int sy, int sign_mode,
                                         float gain, float slope, float clamp, int flip,
                                         const float* aff_a, const float* aff_d, void* stream);

extern "C" int cno_filtered_lrelu(const void* x, void* y, const void* b, uint8_t* mask, float* db,
                                  const float* fu, const float* fd, int dtype,
                                  int n, int c, int in_h, int in_w, int out_h, int out_w,
                                  int fu_taps, int fd_taps, int up, int down, int px0, int py0,
                                  int mask_h, int mask_w, int mask_pitch, int sx, int sy, int sign_mode,
                                  float gain, float slope, float clamp, int flip, void* stream) {
    return cno_filtered_lrelu_affine(x, y, b, mask, db, fu, fd, dtype, n, c, in_h, in_w, out_h, out_w, fu_taps, fd_taps, up, down,
                                     px0, py0, mask_h, mask_w, mask_pitch, sx, sy, sign_mode, gain, slope, clamp, flip,
                                     nullptr, nullptr, stream);
}

extern "C" int cno_filtered_lrelu_affine(const void* x, void* y, const void* b, uint8_t* mask, float* db,
                                         const float* fu, const float* fd, int dtype,
                                         int n, int c, int in_h, int in_w, int out_h, int out_w,
                                         int fu_taps, int fd_taps, int up, int down, int px0, int py0,
                                         int mask_h, int mask_w, int mask_pitch, int sx, int sy, int sign_mode,
                                         float gain, float slope, float clamp, int flip,
                                         const float* aff_a, const float* aff_d, void* stream) {
    CNO_CHECK_ARG((aff_a == nullptr) == (aff_d == nullptr), "filtered_lrelu: aff_a and aff_d go together");
    CNO_CHECK_ARG(x && y && fu && fd, "filtered_lrelu: null pointer");
    CNO_CHECK_ARG(sign_mode >= 0 && sign_mode <= 3, "filtered_lrelu: bad sign_mode %d", sign_mode);
    const bool reading = sign_mode == CNO_SIGN_READ || sign_mode == CNO_SIGN_READ_SIGNONLY;
    CNO_CHECK_ARG(sign_mode == CNO_SIGN_NONE || mask, "filtered_lrelu: sign_mode %d needs a mask", sign_mode);
    CNO_CHECK_ARG(gain > 0.f && slope >= 0.f, "filtered_lrelu: gain must be > 0 and slope >= 0");
    if (dtype != CNO_F32 && dtype != CNO_BF16 && dtype != CNO_F16) return -1;
    if (fu_taps > 32 || fd_taps > 32 || fu_taps < 1 || fd_taps < 1) return -1;
    if (db && !reading) return -1;     // db is accumulated by the mask-reading (backward) variant only
    if (((uintptr_t)x & 15u) || ((uintptr_t)y & 15u)) return -1;
    Plan plan;
    if (make_plan(plan, cno_dtype_size(dtype), n, c, in_h, in_w, out_h, out_w, fu, fu_taps, fd, fd_taps, up, down,
                  px0, py0, gain, slope, clamp, flip, env_ctas()) != 0) return -1;
    Params& p = plan.p;
    if (sign_mode == CNO_SIGN_WRITE)
        CNO_CHECK_ARG(mask_h >= p.UH && mask_pitch >= p.G, "filtered_lrelu: mask [%d rows x %d bytes] smaller than [%d x %d]", mask_h, mask_pitch, p.UH, p.G);
    p.y = y; p.b = b; p.mask = mask; p.db = db; p.pa = aff_a; p.pd = aff_d;
    if (aff_a && (sign_mode == CNO_SIGN_READ || sign_mode == CNO_SIGN_READ_SIGNONLY)) return -1;   // the affine belongs to the forward
    if (reading || sign_mode == CNO_SIGN_WRITE) {
        if (((uintptr_t)mask & 3u) || (mask_pitch & 3)) return -1;     // staged through shared memory with aligned 32-bit copies
        if (reading) set_mask_window(plan, sx);
    }
    p.mask_h = mask_h; p.mask_w = mask_w; p.mask_pitch = mask_pitch; p.sx = sx; p.sy = sy;

    cudaStream_t st = (cudaStream_t)stream;
    {   // small planes: whole planes in shared memory, one thread per up-res column (flrelu_plane.cuh)
        flr_plane::Geom g;
        if (cno_filtered_lrelu_plane_kernel(-1) && flr_plane::make_geom(g, plan) == 0) {
            g.mb0 = p.mws;
            switch (dtype) {
                case CNO_F32: return flr_plane::launch_cfg<float>(plan, g, x, sign_mode, st);
#ifndef FLR_QUICK
                case CNO_BF16: return flr_plane::launch_cfg<__nv_bfloat16>(plan, g, x, sign_mode, st);
                case CNO_F16: return flr_plane::launch_cfg<__half>(plan, g, x, sign_mode, st);
#endif
            }
            return -1;
        }
    }
    EncodeTiledFn enc = get_encode();
    if (!enc) { cno_set_error("filtered_lrelu: cuTensorMapEncodeTiled not available from the driver"); return (int)cudaErrorNotSupported; }
    CUtensorMap map;
    const cuuint64_t dims[3] = {(cuuint64_t)in_w, (cuuint64_t)in_h, (cuuint64_t)p.NC};
    const cuuint64_t strides[2] = {(cuuint64_t)in_w * plan.elem_size, (cuuint64_t)in_w * in_h * plan.elem_size};
    const cuuint32_t box[3] = {(cuuint32_t)p.bwx, (cuuint32_t)plan.box_rows, (cuuint32_t)p.P};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUtensorMapDataType dt = dtype == CNO_F32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                 : dtype == CNO_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    CUresult r = enc(&map, dt, 3, const_cast<void*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { cno_set_error("filtered_lrelu: cuTensorMapEncodeTiled failed (%d)", (int)r); return (int)cudaErrorInvalidValue; }
    switch (dtype) {
        case CNO_F32: return launch_cfg<float>(plan, map, sign_mode, st);
#ifndef FLR_QUICK
        case CNO_BF16: return launch_cfg<__nv_bfloat16>(plan, map, sign_mode, st);
        case CNO_F16: return launch_cfg<__half>(plan, map, sign_mode, st);
#endif
    }
    return -1;
}
