// fused filtered_lrelu kernels, __nv_bfloat16 storage, part 1 of 5 (flrelu_fused_parts.cuh)
#define FLR_PART 1
#define FLR_TYPE __nv_bfloat16
#include "flrelu_fused_parts.cuh"
