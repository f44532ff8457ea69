// Multi-tensor AdamW in one launch (the optimiser step of the training loop, TrainCNO_time_L.py:185-196 /
// CNO_timeModule_CIN.py:996 `torch.optim.AdamW`).  The CNO has ~500 small parameter tensors; torch's fused implementation packs tensor
// pointers into kernel arguments (4 KB) and needs ~32 launches = 0.8 ms per step for 73 MB of parameters.  Here the tensor table lives
// in device memory, so every tensor of every parameter group is updated by ONE grid: a CTA owns one 4,096-element chunk of one
// tensor (chunk -> tensor map precomputed by the host).  HBM-bound: 4 reads + 3 writes of 4 bytes per parameter element.
//
// Arithmetic = torch.optim.AdamW (amsgrad = False, maximize = False), all fp32:
//   p *= 1 - lr * wd;  m += (1 - b1) (g - m);  v = b2 v + (1 - b2) g^2;  p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// The step count t lives on the device (CUDA-graph replays advance it): cno_adamw_multi first increments it with a one-thread launch.
#include "common.cuh"

namespace {

constexpr int ADAM_CHUNK = 4096, ADAM_THREADS = 256;

struct AdamTick { float* step; };
__global__ void adamw_tick_kernel(AdamTick a) { a.step[0] += 1.f; }

__global__ void __launch_bounds__(ADAM_THREADS) adamw_multi_kernel(const CnoAdamTensor* __restrict__ table, const int* __restrict__ chunk_tensor,
                                                                    const float* __restrict__ step, double beta1d, double beta2d, float w1, float w2, float eps) {
    const int ti = chunk_tensor[blockIdx.x];
    const CnoAdamTensor e = table[ti];
    if (e.grad == nullptr) return;                                   // no gradient this step: torch skips the parameter as well
    const int64_t base = (int64_t)(blockIdx.x - e.first_chunk) * ADAM_CHUNK;
    const int64_t left = e.numel - base;
    const int n = left < ADAM_CHUNK ? (int)left : ADAM_CHUNK;
    // bias corrections in double, like torch's fused kernel (two pow per CTA); w1 = 1 - beta1, w2 = 1 - beta2 were rounded from double
    // by the host (1.f - 0.999f differs from (float)(1 - 0.999) by 5e-5, which is visible in exp_avg_sq)
    const double t = (double)step[0];
    const double bc1 = 1.0 - pow(beta1d, t), bc2 = 1.0 - pow(beta2d, t);
    const float beta2 = (float)beta2d;
    const float step_size = (float)((double)e.lr / bc1), inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2)), decay = (float)(1.0 - (double)e.lr * (double)e.weight_decay);
    float* p = e.param + base; const float* g = e.grad + base; float* m = e.exp_avg + base; float* v = e.exp_avg_sq + base;
    auto update = [&](float& pp, float gg, float& mm, float& vv) {
        pp *= decay;
        mm = fmaf(w1, gg - mm, mm);
        vv = fmaf(beta2, vv, w2 * gg * gg);
        const float denom = sqrtf(vv) * inv_sqrt_bc2 + eps;
        pp -= step_size * (mm / denom);
    };
    const bool vec = !(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15u);
    if (vec) {
        for (int i = 4 * threadIdx.x; i + 3 < n; i += 4 * ADAM_THREADS) {
            float4 P = *reinterpret_cast<float4*>(p + i), M = *reinterpret_cast<float4*>(m + i), V = *reinterpret_cast<float4*>(v + i);
            const float4 G = *reinterpret_cast<const float4*>(g + i);
            update(P.x, G.x, M.x, V.x); update(P.y, G.y, M.y, V.y); update(P.z, G.z, M.z, V.z); update(P.w, G.w, M.w, V.w);
            *reinterpret_cast<float4*>(p + i) = P; *reinterpret_cast<float4*>(m + i) = M; *reinterpret_cast<float4*>(v + i) = V;
        }
        for (int i = (n & ~3) + threadIdx.x; i < n; i += ADAM_THREADS) { float P = p[i], M = m[i], V = v[i]; update(P, g[i], M, V); p[i] = P; m[i] = M; v[i] = V; }
    } else {
        for (int i = threadIdx.x; i < n; i += ADAM_THREADS) { float P = p[i], M = m[i], V = v[i]; update(P, g[i], M, V); p[i] = P; m[i] = M; v[i] = V; }
    }
}

}  // namespace

extern "C" int cno_adamw_chunk(void) { return ADAM_CHUNK; }

extern "C" int cno_adamw_multi(const CnoAdamTensor* table, const int* chunk_tensor, int n_chunks, float* step,
                               double beta1, double beta2, double eps, void* stream) {
    CNO_CHECK_ARG(table && chunk_tensor && step, "adamw_multi: null pointer");
    CNO_CHECK_ARG(n_chunks >= 0, "adamw_multi: bad chunk count %d", n_chunks);
    CNO_CHECK_ARG(beta1 >= 0.0 && beta1 < 1.0 && beta2 >= 0.0 && beta2 < 1.0 && eps >= 0.0, "adamw_multi: bad hyper-parameters");
    adamw_tick_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(AdamTick{step});
    CNO_LAUNCH_CHECK();
    if (n_chunks == 0) return 0;
    adamw_multi_kernel<<<(unsigned)n_chunks, ADAM_THREADS, 0, (cudaStream_t)stream>>>(table, chunk_tensor, step, beta1, beta2, (float)(1.0 - beta1), (float)(1.0 - beta2), (float)eps);
    CNO_LAUNCH_CHECK();
    return 0;
}
