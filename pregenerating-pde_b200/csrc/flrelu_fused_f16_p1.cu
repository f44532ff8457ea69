// fused filtered_lrelu kernels, __half storage, part 1 of 5 (flrelu_fused_parts.cuh)
#define FLR_PART 1
#define FLR_TYPE __half
#include "flrelu_fused_parts.cuh"
