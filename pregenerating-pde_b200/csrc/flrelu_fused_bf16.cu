// fused filtered_lrelu kernels, __nv_bfloat16 storage (see flrelu_fused_impl.cuh)
#include "flrelu_fused_impl.cuh"
namespace flr_impl { template int launch_cfg<__nv_bfloat16>(const Plan&, const CUtensorMap&, int, cudaStream_t); }
