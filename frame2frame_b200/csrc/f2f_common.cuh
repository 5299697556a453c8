// f2f_common.cuh -- shared helpers of libf2f_b200 (error plumbing, reductions).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/f2f_b200.h"

namespace f2f {

void set_error(const char *fmt, ...);

#define F2F_CUDA_CHECK(expr)                                                        \
    do {                                                                            \
        cudaError_t _e = (expr);                                                    \
        if (_e != cudaSuccess) {                                                    \
            f2f::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,            \
                           cudaGetErrorString(_e));                                 \
            return F2F_ECUDA;                                                       \
        }                                                                           \
    } while (0)

#define F2F_LAUNCH_CHECK() F2F_CUDA_CHECK(cudaGetLastError())

// bench.py's measurement hook for the DnCNN kernels (dncnn.cu): while enabled, launches of class
// 0 (64->64 conv / dgrad) and 1 (64->64 wgrad) are bracketed with CUDA events on their stream.
void prof_begin(int cls, cudaStream_t st);
void prof_end(int cls, cudaStream_t st);

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

#ifdef __CUDACC__
__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// order-preserving float <-> uint mapping, for atomicMin/atomicMax on floats
__device__ __forceinline__ unsigned f2ord(float f)
{
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u)
{
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}
#endif

}  // namespace f2f
