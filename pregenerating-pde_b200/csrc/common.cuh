// Shared device/host helpers for libcno_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include "../../include/cno_b200.h"

#ifndef __CUDA_ARCH__
#define CNO_HOST_ONLY 1
#endif

void cno_set_error(const char* fmt, ...);

#define CNO_CHECK_ARG(cond, ...) do { if (!(cond)) { cno_set_error(__VA_ARGS__); return (int)cudaErrorInvalidValue; } } while (0)
#define CNO_LAUNCH_CHECK() do { cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) { \
    cno_set_error("%s:%d CUDA launch failed: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); return (int)e_; } } while (0)

// ---- per-device launch caches ---------------------------------------------------------------------
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) and occupancy answers belong to ONE device: a process that drives several
// GPUs (torch.cuda.device(i) around the calls) needs them per ordinal, and the forward thread and the autograd thread may
// race on them, hence atomics.  Each kernel instantiation owns one DevSmemOptIn (function-local static).
#include <atomic>
constexpr int CNO_MAX_DEVICES = 64;
static inline int cno_current_device() { int d = 0; if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= CNO_MAX_DEVICES) d = 0; return d; }
struct DevSmemOptIn {
    std::atomic<size_t> bytes[CNO_MAX_DEVICES];
    // raises the kernel's dynamic shared memory limit on the current device when `need` exceeds what was configured there
    template <class K> cudaError_t ensure(K kern, size_t need) {
        const int d = cno_current_device();
        if (need <= bytes[d].load(std::memory_order_acquire)) return cudaSuccess;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need);
        if (e == cudaSuccess) { size_t cur = bytes[d].load(); while (cur < need && !bytes[d].compare_exchange_weak(cur, need)) {} }
        return e;
    }
};
// one cached environment switch (read once per process instead of a getenv() per launch)
struct EnvFlag {
    std::atomic<int> v{-1};
    bool on(const char* name) { int x = v.load(std::memory_order_relaxed); if (x < 0) { x = getenv(name) ? 1 : 0; v.store(x); } return x != 0; }
};

static inline int cno_dtype_size(int dt) { return dt == CNO_F32 ? 4 : dt == CNO_F64 ? 8 : 2; }

// element <-> fp32 conversion
template <class T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <class T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// 16-byte vector of elements
template <class T> struct Vec16 { static constexpr int N = 16 / sizeof(T); T v[N]; };

template <class T> __device__ __forceinline__ Vec16<T> ld16(const T* p) {
    Vec16<T> r; *reinterpret_cast<uint4*>(r.v) = *reinterpret_cast<const uint4*>(p); return r;
}
// streaming (read-once) 16-byte load: bypass L1 allocation
template <class T> __device__ __forceinline__ Vec16<T> ld16_stream(const T* p) {
    Vec16<T> r; uint4 u;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(p));
    *reinterpret_cast<uint4*>(r.v) = u; return r;
}
template <class T> __device__ __forceinline__ void st16(T* p, const Vec16<T>& r) {
    *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(r.v);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
