// fused filtered_lrelu kernels, __nv_bfloat16 storage, part 0 of 5 (flrelu_fused_parts.cuh)
#define FLR_PART 0
#define FLR_TYPE __nv_bfloat16
#include "flrelu_fused_parts.cuh"
