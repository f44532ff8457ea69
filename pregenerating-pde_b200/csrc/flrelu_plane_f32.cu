// whole-plane filtered_lrelu kernels for small planes, float storage (see flrelu_plane.cuh)
#include "flrelu_plane.cuh"
namespace flr_plane { template int launch_cfg<float>(const Plan&, const Geom&, const void*, int, cudaStream_t); }
