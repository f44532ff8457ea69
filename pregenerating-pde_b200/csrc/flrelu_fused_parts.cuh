// One fifth of the fused filtered_lrelu kernels of storage type FLR_T: compiled with -DFLR_PART=0..4 -DFLR_TYPE=... by the
// flrelu_fused_<type>_p<part>.cu units (parallel compilation; see launch_cfg in flrelu_fused_impl.cuh)
#include "flrelu_fused_impl.cuh"
namespace flr_impl {
#define FLR_INST(U, D, R) template int launch_mode<FLR_TYPE, Cfg<U, D, R, R>>(const Plan&, const CUtensorMap&, int, cudaStream_t);
#if FLR_PART == 0
FLR_INST(2, 2, 0)
#elif FLR_PART == 1
FLR_INST(2, 2, 1)
#elif FLR_PART == 2
FLR_INST(2, 4, 0) FLR_INST(2, 4, 1)
#elif FLR_PART == 3
FLR_INST(4, 2, 0) FLR_INST(4, 2, 1)
#else
FLR_INST(4, 2, 2) FLR_INST(4, 2, 3)
#endif
}
