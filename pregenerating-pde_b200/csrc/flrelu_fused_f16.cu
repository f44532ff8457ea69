// fused filtered_lrelu kernels, __half storage (see flrelu_fused_impl.cuh)
#include "flrelu_fused_impl.cuh"
namespace flr_impl { template int launch_cfg<__half>(const Plan&, const CUtensorMap&, int, cudaStream_t); }
