// fused filtered_lrelu kernels, float storage, part 3 of 5 (flrelu_fused_parts.cuh)
#define FLR_PART 3
#define FLR_TYPE float
#include "flrelu_fused_parts.cuh"
