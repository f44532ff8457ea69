// fused filtered_lrelu kernels, float storage (see flrelu_fused_impl.cuh)
#include "flrelu_fused_impl.cuh"
namespace flr_impl { template int launch_cfg<float>(const Plan&, const CUtensorMap&, int, cudaStream_t); }
#ifdef FLR_TIMING
extern "C" int cno_flr_prof_read(long long* host, int n) {
    return (int)cudaMemcpyFromSymbol(host, flr_impl::g_flr_prof, sizeof(long long) * (size_t)n);
}
#endif
