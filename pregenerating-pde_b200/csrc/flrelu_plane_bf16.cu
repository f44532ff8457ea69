// whole-plane filtered_lrelu kernels for small planes, __nv_bfloat16 storage (see flrelu_plane.cuh)
#include "flrelu_plane.cuh"
namespace flr_plane { template int launch_cfg<__nv_bfloat16>(const Plan&, const Geom&, const void*, int, cudaStream_t); }
