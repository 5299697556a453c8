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

// Programmatic dependent launch (PDL).  Kernels of the per-layer DnCNN chain (64->64 convolution, input gradient, weight
// gradient and the small BatchNorm / reduction kernels between them) are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization: the successor's launch is processed while the predecessor still
// runs, its CTAs are scheduled as SMs free up, and they block in pdl_wait() until the predecessor grid has completed
// and flushed.  Every kernel launched this way MUST execute pdl_wait() before touching global memory, which makes the
// ordering transitive along the chain.  g_pdl is a bit mask (f2f_set_pdl / F2F_PDL in the environment): bit 0 the
// attribute on the big kernels, bit 1 on the small ones, bit 2 an early pdl_launch_dependents() in the big kernels,
// bit 3 in the small ones.  Measured at 1080p on one box (train iteration, profiles/r01_pdl_ab.log): 0 (fully
// serialised) 9.03 ms, 3 (default: attributes only) 8.69 ms, 11 8.69 ms, 15 (all early triggers) 9.28 ms -- releasing
// the dependents early from a persistent kernel that owns every SM only parks their CTAs, so it is off.
extern int g_conv_pair;   // 1: the 64->64 convolutions run as CTA pairs (cta_group::2), F2F_CONV_PAIR / f2f_set_conv_pair
extern int g_pdl;   // bit 0: attribute on the big kernels, 1: on the small ones, 2: early trigger in the big, 3: in the small
// Tile order of the NEXT 64->64 launch (conv, input gradient, weight gradient): 0 = top-down, 1 = bottom-up.  The
// layer-chain drivers alternate it so that every launch begins on the rows its predecessor touched last, which are
// still in L2 (f2f_set_serpentine / F2F_SERPENTINE: 0 leaves every launch top-down).  Results do not depend on it
// beyond FP32 summation order of the per-CTA weight-gradient partials and of the BatchNorm sums.
extern int g_serpentine, g_tile_rev;
inline void tile_dir(int rev) { g_tile_rev = g_serpentine ? rev : 0; }
template <class... KArgs, class... Args>
static inline cudaError_t launch_pdl(int big, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (g_pdl >> (big ? 0 : 1)) & 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// The same launch as a grid of CTA pairs (clusters of 2 along x: the two SMs of a TPC), for cta_group::2 kernels.
template <class... KArgs, class... Args>
static inline cudaError_t launch_pdl_pairs(int big, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                           Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (g_pdl >> (big ? 0 : 1)) & 1;
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = 2; attr[1].val.clusterDim.y = 1; attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 2;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// Multiprocessor count of the CURRENT device (cached per device ordinal; tvl1.cu).  Grids of the persistent kernels and
// every buffer sized per CTA derive from it -- nothing assumes 148.
int sm_count();

// One-time opt-in to > 48 KB of dynamic shared memory for `Kernel` on the current device.  The attribute is per device
// in CUDA, so the "done" state is a bit per device ordinal, not a process-wide flag (a second GPU in the same process
// -- `--gpu 1`, BatchF2F tests across devices -- would otherwise launch unconfigured).
template <auto Kernel>
static inline cudaError_t ensure_dyn_smem(int bytes)
{
    static unsigned long long done = 0;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 64 && ((done >> dev) & 1ull)) return cudaSuccess;
    e = cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess && dev < 64) done |= 1ull << dev;
    return e;
}

int conv64_f32(const void *d_in, const void *d_wprep, float *d_out32, float scale, int accumulate, int H, int W, cudaStream_t st);   // dncnn.cu

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
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
