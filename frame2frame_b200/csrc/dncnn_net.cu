// dncnn_net.cu -- the whole DnCNN-17 (lib/frame2frame/dncnn/net.py:6-39) as the reference
// trains it in blind_denoising.py:211-230: forward in train mode (BatchNorm batch statistics,
// running-stat update) or eval mode, backward to a flat FP32 gradient buffer, and Adam
// (blind_denoising.py:163-165).  The 64->64 convolutions and their input gradients are the
// tcgen05 kernels of dncnn.cu; this file adds
//   * the weight gradient of the 64->64 layers as a tcgen05 GEMM with the pixels as the
//     contraction (k_wgrad64), accumulating all nine taps of a layer in TMEM across the tiles
//     a CTA owns, with BatchNorm-backward applied while the operand is staged;
//   * the 1->64 first layer and 64->1 last layer (HBM-bound, SIMT), forward and backward;
//   * the per-channel BatchNorm bookkeeping between layers and the multi-tensor Adam step.
//
// Flat parameter vector (the order of DnCNN.state_dict(), 556 032 floats):
//   conv0.weight[64][1][3][3] | for k = 1..15: conv_k.weight[64][64][3][3], bn_k.weight[64],
//   bn_k.bias[64] | conv16.weight[1][64][3][3]
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdlib.h>
#include "f2f_common.cuh"
#include "tc05.cuh"

namespace f2f {
extern long g_dncnn_launches;
namespace dn { extern unsigned long long *g_conv_trace; }
namespace dnt {
int make_tensor_map(CUtensorMap *m, const void *base, int H, int W, int boxw, int boxh);   // dncnn.cu
}
namespace net {

using namespace tc05;
typedef tc05::h16 h16;

constexpr int NBN = 15;
constexpr int WSZ = 64 * 64 * 9;                   // 36864
__host__ __device__ constexpr size_t off_conv(int k) { return 576 + (size_t)(k - 1) * (WSZ + 128); }   // k = 1..15
__host__ __device__ constexpr size_t off_gamma(int k) { return off_conv(k) + WSZ; }
__host__ __device__ constexpr size_t off_beta(int k) { return off_conv(k) + WSZ + 64; }
constexpr size_t OFF_W16 = 576 + (size_t)NBN * (WSZ + 128);
constexpr size_t NPARAMS = OFF_W16 + 576;          // 556032
constexpr float BN_EPS = 1e-5f, BN_MOMENTUM = 0.1f;   // torch defaults, net.py:17

// per-layer coefficient block in the workspace (floats)
enum { C_SCALE = 0, C_SHIFT = 64, C_MEAN = 128, C_INVSTD = 192, C_A = 256, C_B = 320, C_C = 384, C_STRIDE = 448 };

// ----------------------------------------------------------------- small kernels

// all 30 tensor-core operand images (15 forward, 15 transposed/flipped) in one launch; same
// element mapping as k_prep_w64 in dncnn.cu
__global__ void k_prep_all(const float *__restrict__ params, h16 *__restrict__ wprep)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 30 * WSZ) return;
    const int which = idx / WSZ, e = idx % WSZ;
    const int transpose = which >= NBN, k = (which % NBN) + 1;
    const float *w = params + off_conv(k);
    const int kk = e & 7, n = (e >> 3) & 63, kc = (e >> 9) & 7, tap = e >> 12;
    const int kch = kc * 8 + kk;
    const float v = !transpose ? w[((size_t)n * 64 + kch) * 9 + tap] : w[((size_t)kch * 64 + n) * 9 + (8 - tap)];
    wprep[idx] = tc05::float_to_h16(v);
}

// Train-mode BatchNorm2d bookkeeping of one layer: batch mean / biased variance from
// the conv epilogue's sums, scale/shift for the next layer's operand staging, running-stat EMA
// with the unbiased variance, num_batches_tracked.
__global__ void k_bn_finalize(const double *__restrict__ stats, const float *__restrict__ gamma,
                              const float *__restrict__ beta, float *__restrict__ running, long long *nbt,
                              float *__restrict__ coef, double n, int trig)
{
    if (trig) pdl_launch_dependents();     // the next convolution sets itself up while this finishes (f2f_common.cuh)
    pdl_wait();
    const int c = threadIdx.x;
    if (c >= 64) return;
    const double mean = stats[c] / n;
    double var = stats[64 + c] / n - mean * mean;
    if (var < 0.0) var = 0.0;
    const float invstd = (float)(1.0 / sqrt(var + (double)BN_EPS));
    const float sc = gamma[c] * invstd;
    coef[C_SCALE + c] = sc;
    coef[C_SHIFT + c] = beta[c] - (float)mean * sc;
    coef[C_MEAN + c] = (float)mean;
    coef[C_INVSTD + c] = invstd;
    if (running) {
        const double unbiased = n > 1.0 ? var * n / (n - 1.0) : var;
        running[c] = (1.0f - BN_MOMENTUM) * running[c] + BN_MOMENTUM * (float)mean;
        running[64 + c] = (1.0f - BN_MOMENTUM) * running[64 + c] + BN_MOMENTUM * (float)unbiased;
    }
    if (nbt && c == 0) *nbt += 1;
}

// Eval mode: y = gamma*(z - running_mean)/sqrt(running_var + eps) + beta for all 15 layers.
__global__ void k_bn_eval_coeffs(const float *__restrict__ params, const float *__restrict__ running,
                                 float *__restrict__ coef)
{
    const int k = blockIdx.x + 1, c = threadIdx.x;
    if (c >= 64) return;
    const float *r = running + (size_t)(k - 1) * 128;
    const float invstd = 1.0f / sqrtf(r[64 + c] + BN_EPS);
    const float sc = params[off_gamma(k) + c] * invstd;
    float *cf = coef + (size_t)(k - 1) * C_STRIDE;
    cf[C_SCALE + c] = sc;
    cf[C_SHIFT + c] = params[off_beta(k) + c] - r[c] * sc;
    cf[C_MEAN + c] = r[c];
    cf[C_INVSTD + c] = invstd;
}

// Gradient scaling.  The backward pass is linear in the incoming gradient dy, and its 16-bit intermediates (masked
// gradients g_k, dz_k) have FP16's range when the library is built with FP16 storage: a mean-reduced loss at 1080p
// hands in |dy| ~ 5e-7, below FP16's normal range.  So dy is multiplied by the power of two s that brings max|dy| into
// [1, 2) on entry (k_last_bwd) and every parameter gradient is multiplied by 1/s on exit (k_wgrad_reduce,
// k_bn_bwd_coeffs, k_small_grads).  Powers of two are exact: for the reference's sum-reduced L1 loss (|dy| = 1,
// s = 1) and for BF16 storage nothing changes bit-wise.  absmax holds the bits of max|dy| (non-negative floats
// order like unsigned integers); 0, subnormal or non-finite -> s = 1.
__device__ __forceinline__ float grad_scale_from_bits(unsigned b, bool inverse)
{
    const unsigned E = b >> 23;
    if (E == 0 || E >= 250) return 1.0f;
    return __uint_as_float((inverse ? E : 254u - E) << 23);      // 2^(E-127) or 2^(127-E)
}
__device__ __forceinline__ float grad_scale(const unsigned *absmax) { return absmax ? grad_scale_from_bits(*absmax, false) : 1.0f; }
__device__ __forceinline__ float grad_unscale(const unsigned *absmax) { return absmax ? grad_scale_from_bits(*absmax, true) : 1.0f; }

__global__ void k_absmax(const float *__restrict__ v, size_t n, unsigned *__restrict__ absmax)
{
    unsigned m = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const unsigned b = __float_as_uint(fabsf(v[i]));
        m = b > m ? b : m;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned t = __shfl_xor_sync(0xffffffffu, m, o);
        m = t > m ? t : m;
    }
    if ((threadIdx.x & 31) == 0 && m) atomicMax(absmax, m);
}

// BatchNorm backward of one layer from the sums the dgrad epilogue produced (sum g, sum g*xhat):
// dgamma, dbeta into the gradient vector and dz = A*g + B*z + C for the operand staging of the
// layer's weight gradient.
__global__ void k_bn_bwd_coeffs(const double *__restrict__ stats, const float *__restrict__ gamma,
                                float *__restrict__ coef, float *__restrict__ dgamma, float *__restrict__ dbeta,
                                double n, int accumulate, int trig, const unsigned *__restrict__ absmax)
{
    if (trig) pdl_launch_dependents();
    pdl_wait();
    const int c = threadIdx.x;
    if (c >= 64) return;
    const double db = stats[c], dg = stats[64 + c];
    const float uns = grad_unscale(absmax);
    const double invstd = coef[C_INVSTD + c], mean = coef[C_MEAN + c], ga = gamma[c];
    coef[C_A + c] = (float)(ga * invstd);
    coef[C_B + c] = (float)(-ga * invstd * invstd * dg / n);
    coef[C_C + c] = (float)(ga * invstd * (-db / n + mean * invstd * dg / n));
    dgamma[c] = (accumulate ? dgamma[c] : 0.0f) + (float)dg * uns;
    dbeta[c] = (accumulate ? dbeta[c] : 0.0f) + (float)db * uns;
}

// first / last layer gradients (n = 576 * C values each) from their FP64 accumulators into the gradient vector
__global__ void k_small_grads(const double *__restrict__ dw0, const double *__restrict__ dw16, float *__restrict__ g_w0,
                              float *__restrict__ g_w16, int n, int accumulate, const unsigned *__restrict__ absmax)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * n) return;
    float *dst = i < n ? g_w0 + i : g_w16 + (i - n);
    const double v = i < n ? dw0[i] : dw16[i - n];
    *dst = (accumulate ? *dst : 0.0f) + (float)v * grad_unscale(absmax);
}

// torch.optim.Adam (amsgrad=False, L2 weight decay folded into the gradient), single flat tensor;
// operation order of torch/optim/adam.py::_single_tensor_adam (blind_denoising.py:163-165).
__global__ void k_adam(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v,
                       size_t n, float beta1, float beta2, float eps, float wd, float step_size, float bc2_sqrt)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float pi = p[i];
        const float gi = fmaf(wd, pi, g[i]);                 // grad.add(param, alpha=weight_decay)
        const float mi = fmaf(1.0f - beta1, gi - m[i], m[i]);   // exp_avg.lerp_(grad, 1 - beta1)
        const float vi = fmaf((1.0f - beta2) * gi, gi, v[i] * beta2);   // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        m[i] = mi;
        v[i] = vi;
        p[i] = pi - step_size * (mi / denom);                // param.addcdiv_(exp_avg, denom, value=-step_size)
    }
}

// ------------------------------------------------ first / last layer, row-streaming formulation
// The 1<->64-channel layers (net.py:13-14,19 and their backward) are HBM-bound by design: 128 B per pixel read or
// written.  A first formulation with 8 lanes per pixel and 8 channels per lane needed 72-88 accumulators per thread
// (8 warps per SM) and exposed one memory latency per pixel: 250 / 270 / 595 / 205 us at 1080p.
// Here a LANE OWNS TWO CHANNELS: a warp reads or writes one pixel's 128 contiguous bytes per step, the per-tap
// accumulators are 18 registers, and the 64-channel image streams through shared memory in 8-row x 64-pixel tiles
// fetched by the bulk-copy engine one tile ahead (64 KB in flight per SM without holding a register).
constexpr int LT = 16, LH = LT + 2, LHP = LH * LH;   // output tile / haloed tile of the 64->1 layer

namespace rs {
#ifndef F2F_LAST_SPLIT_A
#define F2F_LAST_SPLIT_A 0   // 1: the last layer's activation operand as a 16-bit hi + lo pair as well (see k_last_fwd)
#endif
constexpr int TW = 64, TH = 8, NT = 512;            // tile, threads per block (16 warps: row = warp / 2, half = warp & 1)
constexpr int ZROW = TW * 128, ZSTAGE = TH * ZROW;  // one staged row / tile of the NHWC BF16 image
constexpr int XW = TW + 2, XH = TH + 2;             // haloed FP32 tile of the 1-channel image
constexpr int OFF_X = 2 * ZSTAGE, OFF_BAR = OFF_X + ((XH * XW * 4 + 15) & ~15);
constexpr int SMEM = OFF_BAR + 16;

// thread 0: rows of tile (ty0, tx0) of `img` -> stage at `dst`, completion on `bar`
__device__ __forceinline__ void stage_tile(uint32_t dst, uint32_t bar, const h16 *img, int ty0, int tx0, int H, int W)
{
    const int rows = min(TH, H - ty0), nx = min(TW, W - tx0);
    fence_proxy_async();                                  // the stage was last read through the generic proxy
    mbar_arrive_expect_tx(bar, (uint32_t)(rows * nx * 128));
    for (int r = 0; r < rows; r++)
        bulk_g2s(dst + r * ZROW, img + ((size_t)(ty0 + r) * W + tx0) * 64, (uint32_t)(nx * 128), bar);
}

// haloed FP32 tile, zero outside the image, scaled by `mul`
__device__ __forceinline__ void load_halo(float *xS, const float *__restrict__ img, float mul, int ty0, int tx0, int H, int W)
{
    for (int i = threadIdx.x; i < XH * XW; i += NT) {
        const int yy = ty0 - 1 + i / XW, xx = tx0 - 1 + i % XW;
        xS[i] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? mul * __ldg(img + (size_t)yy * W + xx) : 0.0f;
    }
}

// sums NV per-lane values over the 16 warps (scratch: NT * NV floats) and adds channel pair sums to FP64 accumulators
template <int NV, class F>
__device__ __forceinline__ void block_pair_reduce(const float (&v)[NV], float *scratch, F dst)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; i++) scratch[(warp * NV + i) * 32 + lane] = v[i];
    __syncthreads();
    for (int e = threadIdx.x; e < 32 * NV; e += NT) {
        const int i = e >> 5, l = e & 31;
        double t = 0.0;
        for (int w = 0; w < NT / 32; w++) t += (double)scratch[(w * NV + i) * 32 + l];
        atomicAdd(dst(l, i), t);
    }
}

// Backward of the last layer: input gradient (1->64 transposed conv) masked by layer 15's ReLU, its
// BatchNorm-backward sums, and the last layer's weight gradient -- one pass over z15.
//   g15[q][ic]  = relu'(.) * sum_k dn[q - k + 1] * w16[ic][k]        dW16[ic][k] = sum_q a15[q][ic] * dn[q - k + 1]
__global__ void __launch_bounds__(NT, 1) k_last_bwd(const float *__restrict__ dy, float dy_sign, const h16 *__restrict__ z,
                                                    const float *__restrict__ coef, const float *__restrict__ w16,
                                                    h16 *__restrict__ g, double *__restrict__ stats,
                                                    double *__restrict__ dw16, int H, int W, int tiles_x, int ntiles,
                                                    const unsigned *__restrict__ absmax)
{
    extern __shared__ __align__(128) uint8_t smem[];
    dy_sign *= grad_scale(absmax);
    float *dnS = reinterpret_cast<float *>(smem + OFF_X);
    const uint32_t bar0 = smem_u32(smem + OFF_BAR);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, r = warp >> 1, xh = (warp & 1) * 32, c0 = 2 * lane;
    if (threadIdx.x == 0) { mbar_init(bar0, 1); mbar_init(bar0 + 8, 1); mbar_fence_init(); }
    float w[9][2], sc[2], sh[2], is[2], ms[2];
#pragma unroll
    for (int j = 0; j < 2; j++) {
#pragma unroll
        for (int t = 0; t < 9; t++) w[t][j] = w16[(c0 + j) * 9 + t];
        sc[j] = coef[C_SCALE + c0 + j]; sh[j] = coef[C_SHIFT + c0 + j];
        is[j] = coef[C_INVSTD + c0 + j]; ms[j] = -coef[C_MEAN + c0 + j] * is[j];     // xhat = z * is + ms
    }
    float acc[22];   // [0,18): dW16[tap][j]; [18,20): sum g; [20,22): sum g*xhat
#pragma unroll
    for (int i = 0; i < 22; i++) acc[i] = 0.0f;
    float2 w2[9], acc2[9];                                   // both channels of the lane per register pair
#pragma unroll
    for (int t = 0; t < 9; t++) { w2[t] = make_float2(w[t][0], w[t][1]); acc2[t] = make_float2(0.0f, 0.0f); }
    const float2 sc2 = make_float2(sc[0], sc[1]), sh2 = make_float2(sh[0], sh[1]);
    const float2 is2 = make_float2(is[0], is[1]), ms2 = make_float2(ms[0], ms[1]);
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x < ntiles)
        stage_tile(smem_u32(smem), bar0, z, (blockIdx.x / tiles_x) * TH, (blockIdx.x % tiles_x) * TW, H, W);
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, it++) {
        const int s = it & 1, ty0 = (tile / tiles_x) * TH, tx0 = (tile % tiles_x) * TW;
        const int next = tile + gridDim.x;
        if (threadIdx.x == 0 && next < ntiles)
            stage_tile(smem_u32(smem + (s ^ 1) * ZSTAGE), bar0 + 8 * (s ^ 1), z, (next / tiles_x) * TH, (next % tiles_x) * TW, H, W);
        load_halo(dnS, dy, dy_sign, ty0, tx0, H, W);
        __syncthreads();
        mbar_wait(bar0 + 8 * s, (it >> 1) & 1);
        const int y = ty0 + r, nx = min(TW, W - tx0);
        if (y < H) {
            const uint8_t *zrow = smem + s * ZSTAGE + r * ZROW + lane * 4;
            h16 *grow = g + ((size_t)y * W + tx0) * 64 + c0;
            const float *d2 = dnS + (r + 2) * XW + 2;       // dn[ky][kx] at d2[-ky * XW - kx + x]
            // four pixels per step: their 3 x 6 window of dn is read once (4.5 instead of 9 shared-memory loads per
            // pixel) and every multiply-add handles the lane's two channels at once (packed FP32 FMA; same
            // per-component operations in the same order as the scalar version)
            const int xend = min(xh + 32, nx);
            for (int x = xh; x < xend; x += 4) {
                float2 dv[3][6];                             // dv[ky][i] = dn at row -ky, column x - 2 + i (both halves)
#pragma unroll
                for (int ky = 0; ky < 3; ky++)
#pragma unroll
                    for (int i = 0; i < 6; i++) {
                        const float v = d2[x - 2 + i - ky * XW];
                        dv[ky][i] = make_float2(v, v);
                    }
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    if (x + i >= xend) break;
                    const float2 zz = unpack_h16(*reinterpret_cast<const uint32_t *>(zrow + (x + i) * 128));
                    const float2 yv = ffma2(zz, sc2, sh2);
                    const float2 av = make_float2(fmaxf(yv.x, 0.0f), fmaxf(yv.y, 0.0f));
                    float2 sv = make_float2(0.0f, 0.0f);
#pragma unroll
                    for (int t = 0; t < 9; t++) {
                        const float2 dn = dv[t / 3][i + 2 - (t % 3)];
                        sv = ffma2(dn, w2[t], sv);
                        acc2[t] = ffma2(av, dn, acc2[t]);
                    }
                    const uint32_t o = pack_h16(yv.x > 0.0f ? sv.x : 0.0f, yv.y > 0.0f ? sv.y : 0.0f);
                    *reinterpret_cast<uint32_t *>(grow + (size_t)(x + i) * 64) = o;
                    const float2 gr = unpack_h16(o);             // sums of the ROUNDED gradient, like the tcgen05 epilogue
                    acc[18] += gr.x; acc[19] += gr.y;
                    const float2 xh2 = ffma2(zz, is2, ms2);
                    acc[20] = fmaf(gr.x, xh2.x, acc[20]);
                    acc[21] = fmaf(gr.y, xh2.y, acc[21]);
                }
            }
        }
        __syncthreads();                                     // stage s and dnS are free again
    }
#pragma unroll
    for (int t = 0; t < 9; t++) { acc[2 * t] = acc2[t].x; acc[2 * t + 1] = acc2[t].y; }
    block_pair_reduce<22>(acc, reinterpret_cast<float *>(smem), [&](int l, int i) -> double * {
        if (i < 18) return dw16 + (2 * l + (i & 1)) * 9 + (i >> 1);
        if (i < 20) return stats + 2 * l + (i - 18);
        return stats + 64 + 2 * l + (i - 20);
    });
}

// Weight gradient of the first layer: dW0[oc][k] = sum_p g0[p][oc] * x[p + k - 1]
__global__ void __launch_bounds__(NT, 1) k_first_wgrad(const h16 *__restrict__ g0, const float *__restrict__ x,
                                                       double *__restrict__ dw0, int H, int W, int tiles_x, int ntiles)
{
    extern __shared__ __align__(128) uint8_t smem[];
    float *xS = reinterpret_cast<float *>(smem + OFF_X);
    const uint32_t bar0 = smem_u32(smem + OFF_BAR);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, r = warp >> 1, xh = (warp & 1) * 32;
    if (threadIdx.x == 0) { mbar_init(bar0, 1); mbar_init(bar0 + 8, 1); mbar_fence_init(); }
    float acc[18];
    float2 acc2[9];
#pragma unroll
    for (int t = 0; t < 9; t++) acc2[t] = make_float2(0.0f, 0.0f);
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x < ntiles)
        stage_tile(smem_u32(smem), bar0, g0, (blockIdx.x / tiles_x) * TH, (blockIdx.x % tiles_x) * TW, H, W);
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, it++) {
        const int s = it & 1, ty0 = (tile / tiles_x) * TH, tx0 = (tile % tiles_x) * TW;
        const int next = tile + gridDim.x;
        if (threadIdx.x == 0 && next < ntiles)
            stage_tile(smem_u32(smem + (s ^ 1) * ZSTAGE), bar0 + 8 * (s ^ 1), g0, (next / tiles_x) * TH, (next % tiles_x) * TW, H, W);
        load_halo(xS, x, 1.0f, ty0, tx0, H, W);
        __syncthreads();
        mbar_wait(bar0 + 8 * s, (it >> 1) & 1);
        const int y = ty0 + r, nx = min(TW, W - tx0);
        if (y < H) {
            const uint8_t *grow = smem + s * ZSTAGE + r * ZROW + lane * 4;
            const float *x0 = xS + r * XW;                   // x[p + (ky-1, kx-1)] at x0[ky * XW + kx + x]
            const int xend = min(xh + 32, nx);
            for (int x_ = xh; x_ < xend; x_ += 4) {          // four pixels per step, as in k_last_bwd
                float2 xv[3][6];
#pragma unroll
                for (int ky = 0; ky < 3; ky++)
#pragma unroll
                    for (int i = 0; i < 6; i++) {
                        const float v = x0[x_ + i + ky * XW];
                        xv[ky][i] = make_float2(v, v);
                    }
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    if (x_ + i >= xend) break;
                    const float2 gg = unpack_h16(*reinterpret_cast<const uint32_t *>(grow + (x_ + i) * 128));
#pragma unroll
                    for (int t = 0; t < 9; t++) acc2[t] = ffma2(gg, xv[t / 3][i + t % 3], acc2[t]);
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int t = 0; t < 9; t++) { acc[2 * t] = acc2[t].x; acc[2 * t + 1] = acc2[t].y; }
    block_pair_reduce<18>(acc, reinterpret_cast<float *>(smem),
                          [&](int l, int i) -> double * { return dw0 + (2 * l + (i & 1)) * 9 + (i >> 1); });
}

// Conv(1->64, 3x3, pad 1) + ReLU (net.py:13-14): FP32 image -> NHWC BF16 activation; a warp writes one pixel's 128 bytes per step.
constexpr int FT = 256;                              // 8 warps: warp = tile row
__global__ void __launch_bounds__(FT) k_first_fwd(const float *__restrict__ x, const float *__restrict__ w0, h16 *__restrict__ a0,
                                                  int H, int W, int tiles_x, int ntiles)
{
    __shared__ float xS[XH * XW];
    const int lane = threadIdx.x & 31, r = threadIdx.x >> 5, c0 = 2 * lane;
    float w[9][2];
#pragma unroll
    for (int j = 0; j < 2; j++)
#pragma unroll
        for (int t = 0; t < 9; t++) w[t][j] = w0[(c0 + j) * 9 + t];
    float2 w2[9];
#pragma unroll
    for (int t = 0; t < 9; t++) w2[t] = make_float2(w[t][0], w[t][1]);
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int ty0 = (tile / tiles_x) * TH, tx0 = (tile % tiles_x) * TW;
        __syncthreads();
        for (int i = threadIdx.x; i < XH * XW; i += FT) {
            const int yy = ty0 - 1 + i / XW, xx = tx0 - 1 + i % XW;
            xS[i] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(x + (size_t)yy * W + xx) : 0.0f;
        }
        __syncthreads();
        const int y = ty0 + r, nx = min(TW, W - tx0);
        if (y >= H) continue;
        const float *x0 = xS + r * XW;
        h16 *orow = a0 + ((size_t)y * W + tx0) * 64 + c0;
        for (int x_ = 0; x_ < nx; x_ += 4) {                 // four pixels per step: 3 x 6 window, packed FP32 FMAs
            float2 xv[3][6];
#pragma unroll
            for (int ky = 0; ky < 3; ky++)
#pragma unroll
                for (int i = 0; i < 6; i++) {
                    const float v = x0[x_ + i + ky * XW];
                    xv[ky][i] = make_float2(v, v);
                }
#pragma unroll
            for (int i = 0; i < 4; i++) {
                if (x_ + i >= nx) break;
                float2 sv = make_float2(0.0f, 0.0f);
#pragma unroll
                for (int t = 0; t < 9; t++) sv = ffma2(xv[t / 3][i + t % 3], w2[t], sv);
                *reinterpret_cast<uint32_t *>(orow + (size_t)(x_ + i) * 64) = pack_h16_relu(sv.x, sv.y);
            }
        }
    }
}

// Conv(64->1, 3x3, pad 1) (net.py:19) with layer 15's BatchNorm+ReLU on load; out = n (the noise estimate,
// SaoYan's forward) or x - n (net.py:31) when `residual`.  ONE LANE PER INPUT PIXEL of the
// haloed 18x18 tile: the lane reads its pixel's 128 bytes (eight 16-byte loads in flight), forms the nine per-tap
// dot products over all 64 channels with no cross-lane reduction (weights and coefficients are warp-uniform
// shared-memory reads), and the output pixel sums nine shifted partials from shared memory.
// Optional fusion of the frame-to-frame loss into the last layer (f2f_step_train): with `warped` set the
// kernel also forms d = m*out - m*warped, the loss sum / mean, the loss gradient wrt out (what k_masked_loss computes,
// kinds as f2f_masked_loss) and the bits of max|gradient| for the backward pass's power-of-two scaling (k_absmax) --
// three launches and one pass over the output fewer per inner iteration.
struct LossFuse {
    const float *warped, *mask;
    float *dout;
    double *loss;
    unsigned *absmax;
    int kind;
    float inv_n;
};

// MMA = true (the default, F2F_LAST_MMA=0 restores the CUDA-core version): the per-pixel dot products run on the tensor
// cores as warp-level mma.sync m16n8k16 -- P[16 pixels][taps] += A[16 pixels][16 channels] * Wt[16 channels][8 taps],
// FP32 accumulate.  The layer is HBM-bound (128 bytes in, 4 out per pixel; N = 9 is far too narrow for tcgen05), the
// point is to take the 640 FMAs and 150 warp-uniform shared-memory loads per pixel and lane off the CUDA cores, which
// held the scalar version at 2.5x its HBM time.  A lane (g = lane / 4, t = lane % 4) loads channels [16t, 16t+16)
// of pixels g and g+8 of a 16-pixel group (two 32-byte loads, every byte of the two 128-byte pixel rows used once),
// applies layer 15's BatchNorm+ReLU in FP32 and rounds it to the 16-bit storage type -- the operand rounding every
// 64->64 layer applies to its input -- while the WEIGHTS enter as a 16-bit value plus the 16-bit remainder (two
// products per k-step): a rounded weight is the same error at every pixel, and with single 16-bit weights the config0
// trajectory of the online loop moved by 0.06 dB on the second frame, past the 0.05 dB bar the default mode is held to
// there.  -DF2F_LAST_SPLIT_A=1 splits the activation as well (three products, FP32-class like the CUDA-core version).
// The words feed the A fragment as they lie: the contraction does not care which channel sits in which k slot as long
// as the B fragment uses the same assignment.  The BatchNorm coefficients of the lane's 16 channels stay in registers:
// read from shared memory (16 x 16 bytes per group) they made the L1/shared pipe the bound of the kernel (ncu: 83 %,
// 26 M shared-memory wavefronts per launch; 143 -> 105 us with them in registers).
__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1)
{
#ifndef F2F_STORE_BF16
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
#else
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
#endif
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <bool MMA>
__global__ void __launch_bounds__(256, MMA ? 2 : 1) k_last_fwd(const h16 *__restrict__ z, const float *__restrict__ coef,
                                                  const float *__restrict__ w16, const float *__restrict__ x,
                                                  float *__restrict__ out, int H, int W, int tiles_x, int ntiles, int residual,
                                                  LossFuse lf)
{
    double lacc = 0.0;
    unsigned amax = 0;
    constexpr int TS = MMA ? 10 : 9;                    // floats per pixel of T (10: 8-byte aligned pair stores)
    __shared__ __align__(16) float T[LHP * TS + 4];
    __shared__ __align__(16) float wS[MMA ? 4 : 8 * 72];          // [chunk][tap][8 channels]
    __shared__ __align__(16) float cS[MMA ? 4 : 128];             // scale[64], shift[64]
    // MMA: this lane's 16 channels are c(h, j) = 16t + 8h + j; k-step (h, e) puts channels 4e + {0,1} in k slots
    // 2t + {0,1} and channels 4e + {2,3} in k slots 2t + 8 + {0,1}
    __shared__ __align__(16) float w8T[MMA ? 4 * 16 : 4];        // MMA: [t][tap-8 weight of the lane's 16 channels]
    [[maybe_unused]] float sc[16], sh[16];                       // MMA: BatchNorm scale / shift of the lane's 16 channels
    // B fragments [k-step][register] of taps 0-7: w16 as a 16-bit hi part and the 16-bit remainder; tap 8 does not fit the
    // N = 8 tile and a second tile for one column doubles the mma.sync count, which bounds this kernel: it is an FP32
    // dot product over the lane's 16 channels and a 4-lane shuffle sum
    [[maybe_unused]] uint32_t bw[4][2], bl[4][2];
    if constexpr (MMA) {
        const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
#pragma unroll
        for (int i = 0; i < 16; i++) { sc[i] = coef[C_SCALE + 16 * t + i]; sh[i] = coef[C_SHIFT + 16 * t + i]; }
        // tap 8's weights come from shared memory (4 x 16 bytes per group): the registers go to the coefficients, whose
        // shared-memory reads (16 x 16 bytes per group when they lived there) had the kernel bound by the L1/shared pipe
        if (threadIdx.x < 64) w8T[threadIdx.x] = w16[threadIdx.x * 9 + 8];
#pragma unroll
        for (int ks = 0; ks < 4; ks++) {
            const int tap = g, c = 16 * t + 8 * (ks >> 1) + 4 * (ks & 1);
#pragma unroll
            for (int i = 0; i < 2; i++) {
                const float w0 = w16[(c + 2 * i) * 9 + tap], w1 = w16[(c + 2 * i + 1) * 9 + tap];
                const uint32_t hi = pack_h16(w0, w1);
                const float2 hf = unpack_h16(hi);
                bw[ks][i] = hi;
                bl[ks][i] = pack_h16(w0 - hf.x, w1 - hf.y);
            }
        }

    } else {
        for (int e = threadIdx.x; e < 8 * 72; e += 256) {
            const int c = e / 72, t = (e % 72) >> 3, j = e & 7;
            wS[e] = w16[(c * 8 + j) * 9 + t];
        }
        if (threadIdx.x < 64) { cS[threadIdx.x] = coef[C_SCALE + threadIdx.x]; cS[64 + threadIdx.x] = coef[C_SHIFT + threadIdx.x]; }
    }
    __syncthreads();
    const float4 *wv = reinterpret_cast<const float4 *>(wS), *cv = reinterpret_cast<const float4 *>(cS);
    // MMA: operand registers of two 16-pixel groups [slot][pixel g | g+8][channel half].  (Issuing the next tile's first
    // loads before this tile's output stage was tried and was slower, 167 against 143 us: the kernel is bound by the
    // mma.sync rate, not by load latency.)
    [[maybe_unused]] uint4 v[2][2][2];
    [[maybe_unused]] auto load_group = [&](int tl, int q, uint4 (&dst)[2][2]) {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
        const int by = (tl / tiles_x) * LT - 1, bx = (tl % tiles_x) * LT - 1;
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const int pix = (warp + 8 * q) * 16 + g + 8 * r;
            const int gy = by + pix / LH, gx = bx + pix % LH;
            dst[r][0] = dst[r][1] = make_uint4(0, 0, 0, 0);
            if (pix < LHP && gy >= 0 && gy < H && gx >= 0 && gx < W) {
                // one 32-byte load: the four lanes of a pixel cover its 128-byte line once (two 16-byte loads at a
                // 64-byte distance had every warp-wide load touch eight half-used lines: ncu showed L1 at 83 %)
                ldg256(z + ((size_t)gy * W + gx) * 64 + 16 * t, dst[r][0], dst[r][1]);
            }
        }
    };
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int ty0 = (tile / tiles_x) * LT, tx0 = (tile % tiles_x) * LT;
        // this thread's output pixel: its x / mask / warped values are fetched now, so that their latency passes under
        // the partial products instead of after the barrier
        const int oy = ty0 + (threadIdx.x >> 4), ox = tx0 + (threadIdx.x & 15);
        const bool oin = oy < H && ox < W;
        const size_t oi = (size_t)oy * W + ox;
        float xin = 0.0f, min_ = 0.0f, win = 0.0f;
        if (oin) {
            if (residual) xin = __ldg(x + oi);
            if (lf.warped) { min_ = __ldg(lf.mask + oi); win = __ldg(lf.warped + oi); }
        }
        if constexpr (MMA) {
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
            constexpr int NG = (LHP + 15) / 16;          // 16-pixel groups of the haloed tile; warp w takes w, w+8, w+16
            load_group(tile, 0, v[0]);
            load_group(tile, 1, v[1]);                   // two groups (8 x 16 bytes per lane) in flight
#pragma unroll
            for (int q = 0; q < 3; q++) {
                if (warp + 8 * q >= NG) break;           // warp-uniform (warps 5-7 have no third group)
                uint32_t aw[2][2][4];                    // [pixel][half][word]: BatchNorm + ReLU in FP32, rounded to 16 bits
                [[maybe_unused]] uint32_t al[2][2][4];   // F2F_LAST_SPLIT_A: the 16-bit remainder
                float p8[2] = {0.0f, 0.0f};              // tap 8 of the two pixels, this lane's 16 channels (CUDA cores)
                float w8v[16];
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const float4 w = reinterpret_cast<const float4 *>(w8T + 16 * t)[i];
                    w8v[4 * i] = w.x; w8v[4 * i + 1] = w.y; w8v[4 * i + 2] = w.z; w8v[4 * i + 3] = w.w;
                }
#pragma unroll
                for (int r = 0; r < 2; r++) {
                    const int pix = (warp + 8 * q) * 16 + g + 8 * r;
                    const int gy = ty0 - 1 + pix / LH, gx = tx0 - 1 + pix % LH;
                    const bool in = pix < LHP && gy >= 0 && gy < H && gx >= 0 && gx < W;   // zero padding is of a15, not of z15
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const uint32_t w4[4] = {v[q & 1][r][h].x, v[q & 1][r][h].y, v[q & 1][r][h].z, v[q & 1][r][h].w};
                        const float *scv = sc + h * 8, *shv = sh + h * 8;
                        const float *w8 = w8v + h * 8;
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const float2 f = unpack_h16(w4[j]);
                            const float a0 = fmaxf(fmaf(f.x, scv[2 * j], shv[2 * j]), 0.0f);
                            const float a1 = fmaxf(fmaf(f.y, scv[2 * j + 1], shv[2 * j + 1]), 0.0f);
                            const uint32_t pk = pack_h16(a0, a1);
                            aw[r][h][j] = in ? pk : 0u;
#if F2F_LAST_SPLIT_A
                            const float2 hf = unpack_h16(pk);
                            al[r][h][j] = in ? pack_h16(a0 - hf.x, a1 - hf.y) : 0u;
#endif
                            if (in) p8[r] = fmaf(a0, w8[2 * j], fmaf(a1, w8[2 * j + 1], p8[r]));
                        }
                    }
                }
                if (q == 0) load_group(tile, 2, v[0]);   // group 0's registers are free again
                float d0[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int ks = 0; ks < 4; ks++) {
                    const int h = ks >> 1, e = ks & 1;
                    // a * (wh + wl), the small product first (F2F_LAST_SPLIT_A: (ah + al) * (wh + wl) without al * wl)
#if F2F_LAST_SPLIT_A
                    mma16816(d0, al[0][h][2 * e], al[1][h][2 * e], al[0][h][2 * e + 1], al[1][h][2 * e + 1], bw[ks][0], bw[ks][1]);
#endif
                    mma16816(d0, aw[0][h][2 * e], aw[1][h][2 * e], aw[0][h][2 * e + 1], aw[1][h][2 * e + 1], bl[ks][0], bl[ks][1]);
                    mma16816(d0, aw[0][h][2 * e], aw[1][h][2 * e], aw[0][h][2 * e + 1], aw[1][h][2 * e + 1], bw[ks][0], bw[ks][1]);
                }
                // tap 8: the four lanes of a pixel row hold disjoint channel sets
#pragma unroll
                for (int r = 0; r < 2; r++) {
                    p8[r] += __shfl_xor_sync(0xffffffffu, p8[r], 1);
                    p8[r] += __shfl_xor_sync(0xffffffffu, p8[r], 2);
                }
                // accumulator (row g: d[0], d[1]; row g+8: d[2], d[3]) holds taps 2t, 2t+1 of its n tile
#pragma unroll
                for (int r = 0; r < 2; r++) {
                    const int pix = (warp + 8 * q) * 16 + g + 8 * r;
                    if (pix < LHP) {
                        *reinterpret_cast<float2 *>(&T[pix * TS + 2 * t]) = make_float2(d0[2 * r], d0[2 * r + 1]);
                        if (t == 0) T[pix * TS + 8] = p8[r];
                    }
                }
            }
        }
#pragma unroll 1
        for (int pix = threadIdx.x; !MMA && pix < LHP; pix += 256) {
            const int gy = ty0 - 1 + pix / LH, gx = tx0 - 1 + pix % LH;
            float part[9];
#pragma unroll
            for (int t = 0; t < 9; t++) part[t] = 0.0f;
            if (gy >= 0 && gy < H && gx >= 0 && gx < W) {
                const uint4 *src = reinterpret_cast<const uint4 *>(z + ((size_t)gy * W + gx) * 64);
                uint4 v[8];
#pragma unroll
                for (int c = 0; c < 8; c++) v[c] = __ldg(src + c);
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    const uint32_t vw[4] = {v[c].x, v[c].y, v[c].z, v[c].w};
                    const float4 s0 = cv[2 * c], s1 = cv[2 * c + 1], h0 = cv[16 + 2 * c], h1 = cv[16 + 2 * c + 1];
                    const float scv[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
                    const float shv[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
                    float a[8];
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const float2 f = unpack_h16(vw[q]);
                        a[2 * q] = fmaxf(fmaf(f.x, scv[2 * q], shv[2 * q]), 0.0f);
                        a[2 * q + 1] = fmaxf(fmaf(f.y, scv[2 * q + 1], shv[2 * q + 1]), 0.0f);
                    }
#pragma unroll
                    for (int t = 0; t < 9; t++) {
                        const float4 w0 = wv[c * 18 + 2 * t], w1 = wv[c * 18 + 2 * t + 1];
                        part[t] = fmaf(a[0], w0.x, fmaf(a[1], w0.y, fmaf(a[2], w0.z, fmaf(a[3], w0.w, part[t]))));
                        part[t] = fmaf(a[4], w1.x, fmaf(a[5], w1.y, fmaf(a[6], w1.z, fmaf(a[7], w1.w, part[t]))));
                    }
                }
            }
#pragma unroll
            for (int t = 0; t < 9; t++) T[pix * TS + t] = part[t];
        }
        __syncthreads();
        {
            const int y = threadIdx.x >> 4, xx = threadIdx.x & 15;
            float n = 0.0f;
#pragma unroll
            for (int t = 0; t < 9; t++) n += T[((y + t / 3) * LH + xx + t % 3) * TS + t];
            if (oin) {
                const size_t i = oi;
                const float o = residual ? xin - n : n;
                out[i] = o;
                if (lf.warped) {                       // blind_denoising.py:121 / warped_loss.py:227,239-248, as k_masked_loss
                    const float m = min_;
                    const float d = __fsub_rn(__fmul_rn(m, o), __fmul_rn(m, win));
                    float g;
                    if (lf.kind == 0) {
                        lacc += (double)fabsf(d);
                        g = m * ((d > 0.0f) ? 1.0f : ((d < 0.0f) ? -1.0f : 0.0f));
                    } else if (lf.kind == 1) {
                        lacc += (double)__fmul_rn(d, d);
                        g = __fmul_rn(m, __fmul_rn(__fmul_rn(2.0f, d), lf.inv_n));
                    } else {
                        const float r = sqrtf(__fadd_rn(__fmul_rn(d, d), 1e-6f));
                        lacc += (double)r;
                        g = __fmul_rn(m, __fmul_rn(__fdiv_rn(d, r), lf.inv_n));
                    }
                    lf.dout[i] = g;
                    const unsigned b = __float_as_uint(fabsf(g));
                    amax = b > amax ? b : amax;
                }
            }
        }
        __syncthreads();
    }
    if (lf.warped) {
        __shared__ double sl[8];
        __shared__ unsigned sa[8];
        lacc = warp_sum(lacc);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned v = __shfl_xor_sync(0xffffffffu, amax, o);
            amax = v > amax ? v : amax;
        }
        if ((threadIdx.x & 31) == 0) { sl[threadIdx.x >> 5] = lacc; sa[threadIdx.x >> 5] = amax; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double tsum = 0.0;
            unsigned m = 0;
            for (int k = 0; k < 8; k++) { tsum += sl[k]; m = sa[k] > m ? sa[k] : m; }
            atomicAdd(lf.loss, lf.kind == 0 ? tsum : tsum * (double)lf.inv_n);
            if (m) atomicMax(lf.absmax, m);
        }
    }
}
}  // namespace rs


// ------------------------------------------------ first / last layer, any channel count and storage type
// C image channels <-> 64 features (net.py:13-14,19 with `channels` = C; the reference's loader defaults to 3,
// lib/frame2frame/dncnn/io.py:64-67).  A warp per pixel, a lane owns two feature channels; the image is planar
// [C][H][W] FP32 like the reference's NCHW tensors, the feature maps NHWC in T = the 16-bit storage type (default mode)
// or float (precise mode).  The gray-scale default mode keeps its tuned kernels (namespace rs); these serve C != 1 and
// the precise mode.
namespace gen {
template <class T> struct St;
template <> struct St<h16> {
    static __device__ __forceinline__ float2 ld(const h16 *p) { return unpack_h16(__ldg(reinterpret_cast<const uint32_t *>(p))); }
    static __device__ __forceinline__ float2 st(h16 *p, float a, float b)        // -> the values as stored (rounded)
    {
        const uint32_t v = pack_h16(a, b);
        *reinterpret_cast<uint32_t *>(p) = v;
        return unpack_h16(v);
    }
};
template <> struct St<float> {
    static __device__ __forceinline__ float2 ld(const float *p) { return __ldg(reinterpret_cast<const float2 *>(p)); }
    static __device__ __forceinline__ float2 st(float *p, float a, float b)
    {
        *reinterpret_cast<float2 *>(p) = make_float2(a, b);
        return make_float2(a, b);
    }
};
constexpr int NT = 128;                                      // 4 warps per block

template <int NV, class F>
__device__ __forceinline__ void block_lane_reduce(const float (&v)[NV], float *scratch, F dst)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; i++) scratch[(warp * NV + i) * 32 + lane] = v[i];
    __syncthreads();
    for (int e = threadIdx.x; e < 32 * NV; e += NT) {
        const int i = e >> 5, l = e & 31;
        double tsum = 0.0;
        for (int w = 0; w < NT / 32; w++) tsum += (double)scratch[(w * NV + i) * 32 + l];
        atomicAdd(dst(l, i), tsum);
    }
}

template <int C, class T>
__global__ void __launch_bounds__(NT) k_first_fwd(const float *__restrict__ x, const float *__restrict__ w0, T *__restrict__ a0,
                                                  int H, int W)
{
    const int lane = threadIdx.x & 31, c0 = 2 * lane;
    float w[C * 9][2];
#pragma unroll
    for (int q = 0; q < C * 9; q++) { w[q][0] = w0[(size_t)c0 * C * 9 + q]; w[q][1] = w0[(size_t)(c0 + 1) * C * 9 + q]; }
    const size_t P = (size_t)H * W, nw = (size_t)gridDim.x * (NT / 32);
    for (size_t p = (size_t)blockIdx.x * (NT / 32) + (threadIdx.x >> 5); p < P; p += nw) {
        const int y = (int)(p / W), xx = (int)(p % W);
        float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
        for (int tp = 0; tp < 9; tp++) {
            const int yy = y + tp / 3 - 1, xq = xx + tp % 3 - 1;
            const bool in = yy >= 0 && yy < H && xq >= 0 && xq < W;
#pragma unroll
            for (int ci = 0; ci < C; ci++) {
                const float v = in ? __ldg(x + ci * P + (size_t)yy * W + xq) : 0.0f;
                s0 = fmaf(v, w[ci * 9 + tp][0], s0); s1 = fmaf(v, w[ci * 9 + tp][1], s1);
            }
        }
        St<T>::st(a0 + p * 64 + c0, fmaxf(s0, 0.0f), fmaxf(s1, 0.0f));
    }
}

template <int C, class T>
__global__ void __launch_bounds__(NT) k_last_fwd(const T *__restrict__ z, const float *__restrict__ coef, const float *__restrict__ w16,
                                                 const float *__restrict__ x, float *__restrict__ out, int H, int W, int residual)
{
    const int lane = threadIdx.x & 31, c0 = 2 * lane;
    float w[C][9][2];
#pragma unroll
    for (int co = 0; co < C; co++)
#pragma unroll
        for (int tp = 0; tp < 9; tp++) { w[co][tp][0] = w16[((size_t)co * 64 + c0) * 9 + tp]; w[co][tp][1] = w16[((size_t)co * 64 + c0 + 1) * 9 + tp]; }
    const float sc0 = coef[C_SCALE + c0], sc1 = coef[C_SCALE + c0 + 1], sh0 = coef[C_SHIFT + c0], sh1 = coef[C_SHIFT + c0 + 1];
    const size_t P = (size_t)H * W, nw = (size_t)gridDim.x * (NT / 32);
    for (size_t p = (size_t)blockIdx.x * (NT / 32) + (threadIdx.x >> 5); p < P; p += nw) {
        const int y = (int)(p / W), xx = (int)(p % W);
        float s[C];
#pragma unroll
        for (int co = 0; co < C; co++) s[co] = 0.0f;
#pragma unroll
        for (int tp = 0; tp < 9; tp++) {
            const int yy = y + tp / 3 - 1, xq = xx + tp % 3 - 1;
            if (yy >= 0 && yy < H && xq >= 0 && xq < W) {
                const float2 zz = St<T>::ld(z + ((size_t)yy * W + xq) * 64 + c0);
                const float a0v = fmaxf(fmaf(zz.x, sc0, sh0), 0.0f), a1v = fmaxf(fmaf(zz.y, sc1, sh1), 0.0f);
#pragma unroll
                for (int co = 0; co < C; co++) s[co] = fmaf(a1v, w[co][tp][1], fmaf(a0v, w[co][tp][0], s[co]));
            }
        }
#pragma unroll
        for (int co = 0; co < C; co++) {
            const float n = warp_sum(s[co]);
            if (lane == 0) out[co * P + p] = residual ? x[co * P + p] - n : n;
        }
    }
}

// g15 = relu'(.) * conv^T(dn, w16) (dn = +-scale * dy), its BatchNorm-backward sums, and dW16 -- one pass over z15
template <int C, class T>
__global__ void __launch_bounds__(NT) k_last_bwd(const float *__restrict__ dy, float dy_sign, const T *__restrict__ z,
                                                 const float *__restrict__ coef, const float *__restrict__ w16, T *__restrict__ g,
                                                 double *__restrict__ stats, double *__restrict__ dw16, int H, int W,
                                                 const unsigned *__restrict__ absmax)
{
    constexpr int NV = C * 18 + 4;
    __shared__ float scratch[(NT / 32) * NV * 32];
    dy_sign *= grad_scale(absmax);
    const int lane = threadIdx.x & 31, c0 = 2 * lane;
    float w[C][9][2], sc[2], sh[2], is[2], ms[2];
#pragma unroll
    for (int j = 0; j < 2; j++) {
#pragma unroll
        for (int co = 0; co < C; co++)
#pragma unroll
            for (int tp = 0; tp < 9; tp++) w[co][tp][j] = w16[((size_t)co * 64 + c0 + j) * 9 + tp];
        sc[j] = coef[C_SCALE + c0 + j]; sh[j] = coef[C_SHIFT + c0 + j];
        is[j] = coef[C_INVSTD + c0 + j]; ms[j] = -coef[C_MEAN + c0 + j] * is[j];
    }
    float acc[NV];
#pragma unroll
    for (int i = 0; i < NV; i++) acc[i] = 0.0f;
    const size_t P = (size_t)H * W, nw = (size_t)gridDim.x * (NT / 32);
    for (size_t p = (size_t)blockIdx.x * (NT / 32) + (threadIdx.x >> 5); p < P; p += nw) {
        const int y = (int)(p / W), xx = (int)(p % W);
        const float2 zz = St<T>::ld(z + p * 64 + c0);
        const float y0 = fmaf(zz.x, sc[0], sh[0]), y1 = fmaf(zz.y, sc[1], sh[1]);
        const float a0v = fmaxf(y0, 0.0f), a1v = fmaxf(y1, 0.0f);
        float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
        for (int tp = 0; tp < 9; tp++) {
            // the OUTPUT pixel whose tap tp reads this input pixel: q = p - (tap offset)
            const int yy = y - (tp / 3 - 1), xq = xx - (tp % 3 - 1);
            const bool in = yy >= 0 && yy < H && xq >= 0 && xq < W;
#pragma unroll
            for (int co = 0; co < C; co++) {
                const float dn = in ? dy_sign * __ldg(dy + co * P + (size_t)yy * W + xq) : 0.0f;
                s0 = fmaf(dn, w[co][tp][0], s0); s1 = fmaf(dn, w[co][tp][1], s1);
                acc[(co * 9 + tp) * 2] = fmaf(a0v, dn, acc[(co * 9 + tp) * 2]);
                acc[(co * 9 + tp) * 2 + 1] = fmaf(a1v, dn, acc[(co * 9 + tp) * 2 + 1]);
            }
        }
        const float2 gr = St<T>::st(g + p * 64 + c0, y0 > 0.0f ? s0 : 0.0f, y1 > 0.0f ? s1 : 0.0f);   // sums of the STORED gradient
        acc[C * 18] += gr.x; acc[C * 18 + 1] += gr.y;
        acc[C * 18 + 2] = fmaf(gr.x, fmaf(zz.x, is[0], ms[0]), acc[C * 18 + 2]);
        acc[C * 18 + 3] = fmaf(gr.y, fmaf(zz.y, is[1], ms[1]), acc[C * 18 + 3]);
    }
    block_lane_reduce<NV>(acc, scratch, [&](int l, int i) -> double * {
        if (i < C * 18) return dw16 + ((size_t)(i / 18) * 64 + 2 * l + (i & 1)) * 9 + (i % 18) / 2;
        if (i < C * 18 + 2) return stats + 2 * l + (i - C * 18);
        return stats + 64 + 2 * l + (i - C * 18 - 2);
    });
}

// dW0[oc][ci][tap] = sum_p g0[p][oc] * x[ci][p + tap - 1]
template <int C, class T>
__global__ void __launch_bounds__(NT) k_first_wgrad(const T *__restrict__ g0, const float *__restrict__ x, double *__restrict__ dw0,
                                                    int H, int W)
{
    constexpr int NV = C * 18;
    __shared__ float scratch[(NT / 32) * NV * 32];
    const int lane = threadIdx.x & 31, c0 = 2 * lane;
    float acc[NV];
#pragma unroll
    for (int i = 0; i < NV; i++) acc[i] = 0.0f;
    const size_t P = (size_t)H * W, nw = (size_t)gridDim.x * (NT / 32);
    for (size_t p = (size_t)blockIdx.x * (NT / 32) + (threadIdx.x >> 5); p < P; p += nw) {
        const int y = (int)(p / W), xx = (int)(p % W);
        const float2 gg = St<T>::ld(g0 + p * 64 + c0);
#pragma unroll
        for (int tp = 0; tp < 9; tp++) {
            const int yy = y + tp / 3 - 1, xq = xx + tp % 3 - 1;
            const bool in = yy >= 0 && yy < H && xq >= 0 && xq < W;
#pragma unroll
            for (int ci = 0; ci < C; ci++) {
                const float xv = in ? __ldg(x + ci * P + (size_t)yy * W + xq) : 0.0f;
                acc[(ci * 9 + tp) * 2] = fmaf(gg.x, xv, acc[(ci * 9 + tp) * 2]);
                acc[(ci * 9 + tp) * 2 + 1] = fmaf(gg.y, xv, acc[(ci * 9 + tp) * 2 + 1]);
            }
        }
    }
    block_lane_reduce<NV>(acc, scratch, [&](int l, int i) -> double * {
        return dw0 + ((size_t)(2 * l + (i & 1)) * C + i / 18) * 9 + (i % 18) / 2;
    });
}

static int bad_channels(int C)
{
    set_error("DnCNN channels = %d: the library is built for 1 (gray) and 3 (RGB) image channels", C);
    return F2F_EINVAL;
}
static int gen_grid() { return sm_count() * 16; }
template <class T>
static int first_fwd(int C, const float *x, const float *w0, T *a0, int H, int W, cudaStream_t st)
{
    if (C == 1) k_first_fwd<1, T><<<gen_grid(), NT, 0, st>>>(x, w0, a0, H, W);
    else if (C == 3) k_first_fwd<3, T><<<gen_grid(), NT, 0, st>>>(x, w0, a0, H, W);
    else return bad_channels(C);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}
template <class T>
static int last_fwd(int C, const T *z, const float *coef, const float *w16, const float *x, float *out, int H, int W,
                    int residual, cudaStream_t st)
{
    if (C == 1) k_last_fwd<1, T><<<gen_grid(), NT, 0, st>>>(z, coef, w16, x, out, H, W, residual);
    else if (C == 3) k_last_fwd<3, T><<<gen_grid(), NT, 0, st>>>(z, coef, w16, x, out, H, W, residual);
    else return bad_channels(C);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}
template <class T>
static int last_bwd(int C, const float *dy, float dy_sign, const T *z, const float *coef, const float *w16, T *g, double *stats,
                    double *dw16, int H, int W, const unsigned *absmax, cudaStream_t st)
{
    if (C == 1) k_last_bwd<1, T><<<gen_grid(), NT, 0, st>>>(dy, dy_sign, z, coef, w16, g, stats, dw16, H, W, absmax);
    else if (C == 3) k_last_bwd<3, T><<<gen_grid(), NT, 0, st>>>(dy, dy_sign, z, coef, w16, g, stats, dw16, H, W, absmax);
    else return bad_channels(C);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}
template <class T>
static int first_wgrad(int C, const T *g0, const float *x, double *dw0, int H, int W, cudaStream_t st)
{
    if (C == 1) k_first_wgrad<1, T><<<gen_grid(), NT, 0, st>>>(g0, x, dw0, H, W);
    else if (C == 3) k_first_wgrad<3, T><<<gen_grid(), NT, 0, st>>>(g0, x, dw0, H, W);
    else return bad_channels(C);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}
}  // namespace gen

// ------------------------------------------------------- tcgen05 weight gradient
// dW[oc][ic][ky][kx] = sum_p dz[p][oc] * a[p + (ky-1, kx-1)][ic]   (a = the layer's input activation)
//
// GEMM view per filter tap: D[ic][oc] += A[ic][p'] * B[oc][p] with the pixels as K -- both
// operands are "MN-major" (the contraction index is the strided one).  Shared-memory layout of a
// haloed activation tile: [row][8-channel chunk][pixel] with 16-byte elements, so that
//   * 8 consecutive pixels of one chunk are 8 consecutive 16-byte rows = one UMMA core matrix
//     (K = 8 pixels) and the next 8 pixels follow (LBO = 128 B): one MMA contracts 16 pixels of a row;
//   * the 8 chunks of a row are A_S apart (SBO) and the next halo row starts exactly 8 chunk
//     strides further, so a 16-group (M = 128) descriptor spans TWO halo rows: rows (y, y+1)
//     against output row r give the taps dy = y - r and dy + 1 in one instruction;
//   * a tap's dx is a 16-byte shift of the start address.
// Per output row r: 3 MMAs with halo rows (r, r+1) [taps (0,dx),(1,dx)] and 3 with rows
// (r+2, r+3) [taps (2,dx); upper half unused], each M=128 x N=64 x K=16.  The six 128x64 FP32
// accumulators (384 TMEM columns) stay resident over all tiles of the persistent CTA and are
// written once, as a per-CTA partial, at the end; k_wgrad_reduce sums the partials in a fixed
// order (deterministic) into the PyTorch weight-gradient layout.
//
// The dz operand is produced while staging: dz = A*g + B*z + C per channel (BatchNorm backward,
// k_bn_bwd_coeffs) from the ReLU-masked gradient g and the pre-BN output z of the layer, and is
// also written back over g so that the input-gradient convolution that follows reads dz directly.
namespace wg {
#ifndef F2F_WG_DZ_ROWS
#define F2F_WG_DZ_ROWS 4
#endif
// Tile = TR output rows x 16 pixels.  The producers are latency-bound (one DRAM round trip per tile and warp:
// ncu shows 47 % of the samples on the first use of the loaded registers), so the pipeline is made deep rather than
// wide: small tiles, FOUR stages, and TWO producer groups that fill alternate tiles -- every group has two tiles of
// slack before the MMA warp catches up with it.
constexpr int TR = 8;                   // output rows per tile
constexpr int HR = TR + 2;              // halo rows
constexpr int A_S = 19 * 16;            // 304 B between chunk lines (18 pixels + 1 pad: conflict-free 16-byte stores)
constexpr int A_R = 8 * A_S;            // 2432 B between halo rows
constexpr int A_ROWS = HR + 1;          // halo rows + the row the (r+2, r+3) pair touches beyond them
constexpr int A_BYTES = A_ROWS * A_R;
constexpr int D_S = 17 * 16;            // 272
constexpr int D_R = 8 * D_S;            // 2176
constexpr int D_BYTES = TR * D_R;
constexpr int STAGE = A_BYTES + D_BYTES;     // 44 160 B
constexpr int STAGES = 4;
constexpr int NGROUPS = 2;              // producer groups (tile it is staged by group it % NGROUPS into stage it % STAGES)
constexpr int OFF_BAR = STAGES * STAGE;
constexpr int SMEM_BYTES = OFF_BAR + 128;
constexpr int MMA_WARP = 8;
constexpr int PF_WARP = 9 + 8 * (NGROUPS - 1);           // L2 prefetch warp
constexpr int NTHREADS = (PF_WARP + 1) * 32;             // group 0: warps 0-7, MMA: warp 8, group 1: warps 9-16, prefetch: warp 17
constexpr int TMEM_COLS = 512;
constexpr int PARTIAL = 6 * 128 * 64;   // floats per CTA
constexpr int HPIX = HR * 18, DPIX = TR * 16;

struct Args {
    const h16 *act;       // layer input: z_{k-1} (BN+ReLU applied on load) or a_0 (raw)
    const float *coef_in;  // scale/shift of layer k-1, or NULL
    h16 *g;               // in: masked gradient g_k; out: dz_k
    const h16 *z;         // pre-BN output z_k
    const float *coef;     // A, B, C of layer k
    float *partial;        // [gridDim.x][PARTIAL]
    int H, W, tiles_x, tiles_y;
    int pdl_trigger;
    int m64;               // taps ky = 2 as M = 64 instructions (F2F_WGRAD_M64)
    int rev;               // tile order: 1 = bottom-up (f2f_common.cuh, g_tile_rev)
    int pf;                // L2 prefetch distance, in tiles of a producer group (0: off; F2F_WGRAD_PF)
    unsigned long long *trace;   // -DF2F_WG_DEBUG builds only: role trace, see WG_T below
    int dbg;               // -DF2F_WG_DEBUG builds only (tools/perf_wgrad.py): bit 0 = the MMA warp issues no MMAs
};
#ifdef F2F_WG_DEBUG
#define WG_DBG(bit) ((a.dbg >> (bit)) & 1)
// role trace of the debug build (f2f_dncnn_set_trace, tools/perf_wgrad.py): cycles summed over the CTAs, lane 0 of the first
// activation / dz producer warp of group 0 and the MMA issuer.  [0..4] activation producer: issuing loads, waiting
// for a free stage, loads landing (bit 6 set: all loaded registers are touched first), transform + stores, fence + arrive;
// [5..9] the same for the dz producer; [10] MMA warp waiting for a full stage, [11] issuing; [12..14] tiles of the three roles
#define WG_T(var) const long long var = clock64()
#define WG_ACC(slot, val) do { if (tr_on) tr_acc[slot] += (val); } while (0)
#else
#define WG_DBG(bit) 0
#define WG_T(var)
#define WG_ACC(slot, val)
#endif

// The producers below are register-staged (load -> transform -> one shared-memory store) and were bound by the DRAM
// latency of their loads.  The three tensor maps are used for NOTHING but cp.async.bulk.prefetch.tensor: before a
// producer group loads tile t it has the TMA unit pull the boxes of tile t + pf (activation halo box, g box, z box)
// into L2, clipped at the image edges by the map, so that the ld.global of that tile find their lines in L2.
__global__ void __launch_bounds__(NTHREADS, 1) k_wgrad64(Args a, const __grid_constant__ CUtensorMap tm_act,
                                                         const __grid_constant__ CUtensorMap tm_g,
                                                         const __grid_constant__ CUtensorMap tm_z)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform
    const int ntiles = a.tiles_x * a.tiles_y;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + OFF_BAR);
    const uint32_t bar0 = smem_u32(bars);
    auto BAR = [&](int i) { return bar0 + 8u * i; };   // full[STAGES] @0, empty[STAGES] @STAGES, done @2*STAGES; tmem ptr after
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(bars + 2 * STAGES + 2);
    volatile int *progress = reinterpret_cast<volatile int *>(tmem_ptr + 2);   // [group]: the tile (of the CTA) the group is staging

    if (threadIdx.x == 0) {
        // full: one arrival per producer thread (a per-warp arrival behind __syncwarp measured 20 % slower)
        for (int s = 0; s < STAGES; s++) { mbar_init(BAR(s), 256); mbar_init(BAR(STAGES + s), 1); }
        mbar_init(BAR(2 * STAGES), 1);
        progress[0] = progress[1] = 0;
        mbar_fence_init();
    }
    if (warp == MMA_WARP) tmem_alloc(smem_u32(tmem_ptr), TMEM_COLS);
    // the extra halo row of each stage is only ever multiplied into unused accumulator rows; keep it finite
    for (int i = threadIdx.x; i < STAGES * (A_R / 16); i += NTHREADS) {
        const int s = i / (A_R / 16), o = i % (A_R / 16);
        *reinterpret_cast<uint4 *>(smem + s * STAGE + HR * A_R + o * 16) = make_uint4(0, 0, 0, 0);
    }
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;
    pdl_wait();                  // set-up above overlapped the predecessor's tail (f2f_common.cuh)
    if (a.pdl_trigger) pdl_launch_dependents();

    // role of this warp: producer warp index pw in 0..7 of group grp, the MMA warp, or the prefetch warp
    const int grp = (warp > MMA_WARP && warp != PF_WARP) ? 1 : 0;
    const int pw = warp == PF_WARP ? 8 : warp > MMA_WARP ? warp - (MMA_WARP + 1) : warp;

    if (warp != MMA_WARP && pw < 4) {
        // ===== producer of the activation tile (HR x 18 halo) =====
        // A thread owns one 8-channel chunk of one of the 16 regular columns and walks that column's HR halo rows: a
        // row's address is the previous one plus an image row, the shared-memory destination a compile-time offset
        // from a per-thread base; tiles that lie inside the image (all but the frame's border) take a path without
        // any bounds predicate.  The 2 x HR x 8 elements of halo columns 16 and 17 are one item for every thread plus
        // one more for the first 32.  (The first version spent 65 instructions per 16-byte element, 20 of them
        // arithmetic on the data, and the role trace -- tools/perf_wgrad.py, -DF2F_WG_DEBUG -- showed the kernel bound
        // by its producers' instruction stream: knocking the MMAs out did not change its time.)
        const int ptid = pw * 32 + lane, chunk = ptid & 7, c = ptid >> 3;
        float2 sc[4], sh[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            sc[j] = a.coef_in ? make_float2(a.coef_in[C_SCALE + chunk * 8 + 2 * j], a.coef_in[C_SCALE + chunk * 8 + 2 * j + 1]) : make_float2(1.f, 1.f);
            sh[j] = a.coef_in ? make_float2(a.coef_in[C_SHIFT + chunk * 8 + 2 * j], a.coef_in[C_SHIFT + chunk * 8 + 2 * j + 1]) : make_float2(0.f, 0.f);
        }
        const bool bn = a.coef_in != nullptr;
        const ptrdiff_t rowpitch = (ptrdiff_t)a.W * 64;                 // elements between image rows
        const int xrow = ptid >> 4, xcol = 16 + ((ptid >> 3) & 1);      // extra items: rows xrow and (ptid < 32) 8 + xrow
        const bool x1 = ptid < (HR - 8) * 16;
        const uint32_t soff = chunk * A_S + c * 16, xoff = xrow * A_R + chunk * A_S + xcol * 16;
        // BatchNorm + ReLU of layer k-1 on the 8 channels of one element (two channels per packed FP32 FMA)
        auto xf = [&](const uint4 &v) {
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float2 r = ffma2(unpack_h16(w[j]), sc[j], sh[j]);
                o[j] = pack_h16_relu(r.x, r.y);
            }
            return make_uint4(o[0], o[1], o[2], o[3]);
        };
#ifdef F2F_WG_DEBUG
        const bool tr_on = a.trace && grp == 0 && pw == 0 && lane == 0;
        long long tr_acc[5] = {0, 0, 0, 0, 0};
        unsigned tr_tiles = 0;
#endif
        int it = grp;
        for (int tile = blockIdx.x + grp * gridDim.x; tile < ntiles; tile += NGROUPS * gridDim.x, it += NGROUPS) {
            const int s = it % STAGES;
            WG_T(t_a);
            const int tpos = a.rev ? ntiles - 1 - tile : tile;
            const int ty0 = (tpos / a.tiles_x) * TR - 1, tx0 = (tpos % a.tiles_x) * 16 - 1;
            const h16 *base = a.act + ((ptrdiff_t)ty0 * a.W + tx0) * 64 + chunk * 8;
            uint8_t *dstA = smem + s * STAGE;
            uint4 v[HR + 2];
            v[HR + 1] = make_uint4(0, 0, 0, 0);
            if (ty0 >= 0 && tx0 >= 0 && ty0 + HR <= a.H && tx0 + 18 <= a.W) {      // warp-uniform: the tile and its halo lie inside
#pragma unroll
                for (int u = 0; u < HR; u++) v[u] = __ldg(reinterpret_cast<const uint4 *>(base + u * rowpitch + c * 64));
                v[HR] = __ldg(reinterpret_cast<const uint4 *>(base + xrow * rowpitch + xcol * 64));
                if (x1) v[HR + 1] = __ldg(reinterpret_cast<const uint4 *>(base + (8 + xrow) * rowpitch + xcol * 64));
                WG_T(t_b);
                mbar_wait(BAR(STAGES + s), ((it / STAGES) & 1) ^ 1);      // the loads above are already in flight
                WG_T(t_c);
                if (bn) {
#pragma unroll
                    for (int u = 0; u < HR + 2; u++) v[u] = xf(v[u]);
                }
                WG_ACC(0, t_b - t_a); WG_ACC(1, t_c - t_b);
            } else {
                const bool colok = (unsigned)(tx0 + c) < (unsigned)a.W, xcolok = (unsigned)(tx0 + xcol) < (unsigned)a.W;
                uint32_t inb = 0;
#pragma unroll
                for (int u = 0; u < HR; u++) {
                    const bool ok = colok && (unsigned)(ty0 + u) < (unsigned)a.H;
                    v[u] = make_uint4(0, 0, 0, 0);
                    if (ok) v[u] = __ldg(reinterpret_cast<const uint4 *>(base + u * rowpitch + c * 64));
                    inb |= (uint32_t)ok << u;
                }
                const bool ok0 = xcolok && (unsigned)(ty0 + xrow) < (unsigned)a.H;
                const bool ok1 = x1 && xcolok && (unsigned)(ty0 + 8 + xrow) < (unsigned)a.H;
                v[HR] = make_uint4(0, 0, 0, 0);
                if (ok0) v[HR] = __ldg(reinterpret_cast<const uint4 *>(base + xrow * rowpitch + xcol * 64));
                if (ok1) v[HR + 1] = __ldg(reinterpret_cast<const uint4 *>(base + (8 + xrow) * rowpitch + xcol * 64));
                inb |= ((uint32_t)ok0 << HR) | ((uint32_t)ok1 << (HR + 1));
                mbar_wait(BAR(STAGES + s), ((it / STAGES) & 1) ^ 1);
                if (bn) {                                                  // positions outside the image stay zero
#pragma unroll
                    for (int u = 0; u < HR + 2; u++) v[u] = ((inb >> u) & 1u) ? xf(v[u]) : make_uint4(0, 0, 0, 0);
                }
            }
            WG_T(t_d);
#pragma unroll
            for (int u = 0; u < HR; u++) *reinterpret_cast<uint4 *>(dstA + soff + u * A_R) = v[u];
            *reinterpret_cast<uint4 *>(dstA + xoff) = v[HR];
            if (x1) *reinterpret_cast<uint4 *>(dstA + xoff + 8 * A_R) = v[HR + 1];
            WG_T(t_e);
            fence_proxy_async();
            mbar_arrive(BAR(s));
#ifdef F2F_WG_DEBUG
            { WG_T(t_f); WG_ACC(3, t_e - t_d); WG_ACC(4, t_f - t_e); tr_tiles++; }
#endif
        }
#ifdef F2F_WG_DEBUG
        if (tr_on) {
            for (int q = 0; q < 5; q++) atomicAdd(a.trace + q, (unsigned long long)tr_acc[q]);
            atomicAdd(a.trace + 12, (unsigned long long)tr_tiles);
        }
#endif
        if (grp == 0) {
            // ===== after the last tile: accumulators -> this CTA's partial (warps 0-3 = the four TMEM lane quadrants) =====
            mbar_wait(BAR(2 * STAGES), 0);
            fence_after_sync();
            float *dst = a.partial + (size_t)blockIdx.x * PARTIAL;
            const int row = warp * 32 + lane;
#pragma unroll 1
            for (int j = 0; j < 6; j++) {
#pragma unroll 1
                for (int part = 0; part < 2; part++) {
                    uint32_t r[32];
                    tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + j * 64 + part * 32, r);
                    tmem_ld_wait();
                    // M = 64 accumulators (a.m64, taps ky = 2): row i of D sits in TMEM lane 32 * (i / 16) + i % 16
                    const bool m64 = a.m64 && j >= 3;
                    const int orow = m64 ? warp * 16 + lane : row;
                    if (m64 ? lane < 16 : (j < 3 || row < 64)) {
                        float4 *o = reinterpret_cast<float4 *>(dst + ((size_t)j * 128 + orow) * 64 + part * 32);
#pragma unroll
                        for (int q = 0; q < 8; q++)
                            o[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                               __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
                    }
                }
            }
        }
    } else if (warp == PF_WARP) {
        // ===== L2 prefetch of the tiles ahead (one lane): the TMA unit pulls the activation halo box and the g / z boxes
        // of CTA tile q into L2 once the producer group that will stage it has started the tile pf group-tiles earlier.
        // A thread of its own, because a cp.async.bulk.prefetch.tensor can hold its issuing thread for ~1000 cycles
        // (measured with the role trace while producer lanes issued them).
        if (lane == 0 && a.pf > 0) {
            tma_prefetch_desc(&tm_act);
            tma_prefetch_desc(&tm_g);
            tma_prefetch_desc(&tm_z);
            for (int q = NGROUPS;; q++) {
                const int tf = blockIdx.x + q * gridDim.x;
                if (tf >= ntiles) break;
                while (*(progress + (q % NGROUPS)) < q - NGROUPS * a.pf) __nanosleep(64);
                const int tp = a.rev ? ntiles - 1 - tf : tf;
                const int px = (tp % a.tiles_x) * 16, py = (tp / a.tiles_x) * TR;
                tma_prefetch_l2_3d(&tm_g, 0, px, py);
                tma_prefetch_l2_3d(&tm_z, 0, px, py);
                tma_prefetch_l2_3d(&tm_act, 0, px - 1, py - 1);
            }
        }
    } else if (warp != MMA_WARP) {
        // ===== producer of the dz tile (TR x 16, no halo); writes dz back over g =====
        // same thread -> (column, chunk) assignment as the activation producer; rows in passes of F2F_WG_DZ_ROWS
        const int ptid = (pw - 4) * 32 + lane, chunk = ptid & 7, c = ptid >> 3;
        float2 cA[4], cB[4], cC[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            cA[j] = make_float2(a.coef[C_A + chunk * 8 + 2 * j], a.coef[C_A + chunk * 8 + 2 * j + 1]);
            cB[j] = make_float2(a.coef[C_B + chunk * 8 + 2 * j], a.coef[C_B + chunk * 8 + 2 * j + 1]);
            cC[j] = make_float2(a.coef[C_C + chunk * 8 + 2 * j], a.coef[C_C + chunk * 8 + 2 * j + 1]);
        }
        // dz = A * g + (B * z + C) on the 8 channels of one element: the fmaf(A, g, fmaf(B, z, C)) of the first version,
        // two channels per packed FP32 FMA (same roundings)
        auto xf = [&](const uint4 &g, const uint4 &z) {
            const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, zw[4] = {z.x, z.y, z.z, z.w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float2 r = ffma2(cA[j], unpack_h16(gw[j]), ffma2(cB[j], unpack_h16(zw[j]), cC[j]));
                o[j] = pack_h16(r.x, r.y);
            }
            return make_uint4(o[0], o[1], o[2], o[3]);
        };
        const ptrdiff_t rowpitch = (ptrdiff_t)a.W * 64;
        const uint32_t soff = chunk * D_S + c * 16;
#ifdef F2F_WG_DEBUG
        const bool tr_on = a.trace && grp == 0 && pw == 4 && lane == 0;
        long long tr_acc[5] = {0, 0, 0, 0, 0};
        unsigned tr_tiles = 0;
#endif
        int it = grp;
        for (int tile = blockIdx.x + grp * gridDim.x; tile < ntiles; tile += NGROUPS * gridDim.x, it += NGROUPS) {
            const int s = it % STAGES;
            WG_T(t_a);
            if (pw == 4 && lane == 0) *(progress + grp) = it;         // paces the prefetch warp
            const int tpos = a.rev ? ntiles - 1 - tile : tile;
            const int ty0 = (tpos / a.tiles_x) * TR, tx0 = (tpos % a.tiles_x) * 16;
            const ptrdiff_t o0 = ((ptrdiff_t)ty0 * a.W + tx0 + c) * 64 + chunk * 8;
            uint8_t *dstD = smem + s * STAGE + A_BYTES + soff;
            constexpr int U = F2F_WG_DZ_ROWS;     // rows in flight per thread (register budget: 96 registers per thread)
            if (ty0 + TR <= a.H && tx0 + 16 <= a.W) {      // warp-uniform: the tile lies inside the image
                // rolling window: U rows of g and z in flight; as soon as a row has been transformed its registers take
                // the loads of row + U, so one memory latency is exposed per tile instead of one per pass of U rows
                uint4 vg[U], vz[U];
#pragma unroll
                for (int u = 0; u < U; u++) {
                    vg[u] = *reinterpret_cast<const uint4 *>(a.g + o0 + u * rowpitch);
                    vz[u] = __ldg(reinterpret_cast<const uint4 *>(a.z + o0 + u * rowpitch));
                }
                WG_T(t_b);
                mbar_wait(BAR(STAGES + s), ((it / STAGES) & 1) ^ 1);      // the loads above are already in flight
                WG_T(t_c);
#pragma unroll
                for (int u = 0; u < TR; u++) {
                    const uint4 o = xf(vg[u % U], vz[u % U]);
                    if (u + U < TR) {
                        vg[u % U] = *reinterpret_cast<const uint4 *>(a.g + o0 + (u + U) * rowpitch);
                        vz[u % U] = __ldg(reinterpret_cast<const uint4 *>(a.z + o0 + (u + U) * rowpitch));
                    }
                    *reinterpret_cast<uint4 *>(a.g + o0 + u * rowpitch) = o;
                    *reinterpret_cast<uint4 *>(dstD + u * D_R) = o;
                }
#ifdef F2F_WG_DEBUG
                { WG_T(t_e); WG_ACC(0, t_b - t_a); WG_ACC(1, t_c - t_b); WG_ACC(3, t_e - t_c); }
#endif
            } else {
                const bool colok = tx0 + c < a.W;
#pragma unroll 1
                for (int r0 = 0; r0 < TR; r0 += U) {
                    uint4 vg[U], vz[U];
                    uint32_t inb = 0;
#pragma unroll
                    for (int u = 0; u < U; u++) {
                        const bool ok = colok && ty0 + r0 + u < a.H;
                        vg[u] = vz[u] = make_uint4(0, 0, 0, 0);
                        if (ok) {
                            vg[u] = *reinterpret_cast<const uint4 *>(a.g + o0 + (r0 + u) * rowpitch);
                            vz[u] = __ldg(reinterpret_cast<const uint4 *>(a.z + o0 + (r0 + u) * rowpitch));
                        }
                        inb |= (uint32_t)ok << u;
                    }
                    if (r0 == 0) mbar_wait(BAR(STAGES + s), ((it / STAGES) & 1) ^ 1);
#pragma unroll
                    for (int u = 0; u < U; u++) {
                        const bool ok = (inb >> u) & 1u;
                        const uint4 o = ok ? xf(vg[u], vz[u]) : make_uint4(0, 0, 0, 0);
                        if (ok) *reinterpret_cast<uint4 *>(a.g + o0 + (r0 + u) * rowpitch) = o;
                        *reinterpret_cast<uint4 *>(dstD + (r0 + u) * D_R) = o;
                    }
                }
            }
            WG_T(t_f);
            fence_proxy_async();
            mbar_arrive(BAR(s));
#ifdef F2F_WG_DEBUG
            { WG_T(t_g); WG_ACC(4, t_g - t_f); tr_tiles++; }
#endif
        }
#ifdef F2F_WG_DEBUG
        if (tr_on) {
            for (int q = 0; q < 5; q++) atomicAdd(a.trace + 5 + q, (unsigned long long)tr_acc[q]);
            atomicAdd(a.trace + 13, (unsigned long long)tr_tiles);
        }
#endif
    } else {
        // ===== MMA issuer: converged warp, the elected lane issues (see dncnn.cu) =====
        const uint32_t el = elect_one();
        constexpr uint32_t IDESC = idesc_h16(128, 64, 1, 1);
        // the third vertical tap has no partner row: as an M = 64 instruction it reads 2 KB of A instead of 4 KB
        constexpr uint32_t IDESC64 = idesc_h16(64, 64, 1, 1);
        // descriptors differ only in the start-address field (low word, 16-byte units)
        const uint64_t ad0 = smem_desc(0, 128, A_S), bd0 = smem_desc(0, 128, D_S);
        const uint32_t a_hi = (uint32_t)(ad0 >> 32), b_hi = (uint32_t)(bd0 >> 32);
#ifdef F2F_WG_DEBUG
        const bool tr_on = a.trace && el;
        long long tr_acc[2] = {0, 0};
        unsigned tr_tiles = 0;
#endif
        int it = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, it++) {
            const int s = it % STAGES;
            WG_T(t_a);
            mbar_wait(BAR(s), (it / STAGES) & 1);
            WG_T(t_b);
            fence_after_sync();
            const uint32_t a_base = smem_u32(smem + s * STAGE);
            const uint32_t a_lo = (uint32_t)ad0 | ((a_base & 0x3FFFFu) >> 4);
            const uint32_t b_lo = (uint32_t)bd0 | (((a_base + A_BYTES) & 0x3FFFFu) >> 4);
#pragma unroll
            for (int r = 0; r < TR; r++) {
                if (WG_DBG(0)) break;
                const uint64_t bd = ((uint64_t)b_hi << 32) | (b_lo + r * (D_R / 16));
#pragma unroll
                for (int j = 0; j < 6; j++) {
                    const int dy = j < 3 ? 0 : 2, dx = j % 3;
                    const uint64_t ad = ((uint64_t)a_hi << 32) | (a_lo + (r + dy) * (A_R / 16) + dx);
                    mma_ss_el(tmem_base + j * 64, ad, bd, (j >= 3 && a.m64) ? IDESC64 : IDESC, (it | r) != 0, el);
                }
            }
            if (el) mma_commit(BAR(STAGES + s));
            __syncwarp();
#ifdef F2F_WG_DEBUG
            { WG_T(t_c); WG_ACC(0, t_b - t_a); WG_ACC(1, t_c - t_b); tr_tiles++; }
#endif
        }
#ifdef F2F_WG_DEBUG
        if (tr_on) {
            atomicAdd(a.trace + 10, (unsigned long long)tr_acc[0]);
            atomicAdd(a.trace + 11, (unsigned long long)tr_acc[1]);
            atomicAdd(a.trace + 14, (unsigned long long)tr_tiles);
        }
#endif
        if (el) mma_commit(BAR(2 * STAGES));
        __syncwarp();
    }
    fence_before_sync();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, TMEM_COLS);
}

// dW[oc][ic][ky][kx] (+)= sum over CTA partials.  tap (ky,kx): ky < 2 -> accumulator kx, rows
// ky*64 + ic; ky == 2 -> accumulator 3 + kx, rows ic.
__global__ void k_wgrad_reduce(const float *__restrict__ partial, int nparts, float *__restrict__ dw, int accumulate, int trig,
                               const unsigned *__restrict__ absmax, float extra)
{
    if (trig) pdl_launch_dependents();
    pdl_wait();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 9 * 4096) return;
    const int oc = idx & 63, ic = (idx >> 6) & 63, tap = idx >> 12;
    const int ky = tap / 3, kx = tap % 3;
    const int j = ky < 2 ? kx : 3 + kx, row = ky < 2 ? ky * 64 + ic : ic;
    const float *src = partial + ((size_t)j * 128 + row) * 64 + oc;
    float t = 0.0f;
    for (int c = 0; c < nparts; c++) t += src[(size_t)c * PARTIAL];
    float *d = dw + ((size_t)oc * 64 + ic) * 9 + tap;
    *d = (accumulate ? *d : 0.0f) + t * (grad_unscale(absmax) * extra);
}
}  // namespace wg

// --------------------------------------------------------------- workspace plan
struct Plan {
    size_t act_bytes;       // one NHWC BF16 activation of the whole batch
    size_t o_x, o_a0, o_z, o_g0, o_g1, o_wprep, o_stats_f, o_stats_b, o_dw_small, o_absmax, o_coef, o_partial, total;
    int nz;
};

static int num_sms() { return sm_count(); }   // of the current device (f2f_common.cuh)

static Plan make_plan(int B, int H, int W, int train, int C = 1)
{
    Plan p;
    const size_t P = (size_t)H * W;
    p.act_bytes = align_up((size_t)B * P * 128, 1024);
    size_t o = 0;
    auto take = [&](size_t bytes) { const size_t r = o; o += align_up(bytes, 1024); return r; };
    p.o_x = take((size_t)B * P * 4 * C);
    p.o_a0 = take(p.act_bytes);
    p.nz = train ? NBN : 2;                       // eval: ping-pong
    p.o_z = take(p.act_bytes * p.nz);
    p.o_g0 = train ? take(p.act_bytes) : 0;
    p.o_g1 = train ? take(p.act_bytes) : 0;
    p.o_wprep = take((size_t)30 * WSZ * 2);
    p.o_stats_f = take((size_t)NBN * 128 * 8);
    p.o_stats_b = take((size_t)NBN * 128 * 8);     // contiguous with o_dw_small: one memset
    p.o_dw_small = take((size_t)1152 * C * 8);
    p.o_absmax = take(16);                         // bits of max|dy| (gradient scaling); zeroed by the same memset
    p.o_coef = take((size_t)NBN * C_STRIDE * 4);
    p.o_partial = train ? take((size_t)num_sms() * wg::PARTIAL * 4) : 0;   // one partial per CTA of the persistent grid
    p.total = o;
    return p;
}

static int simt_blocks() { return num_sms() * 4; }

static void *g_debug_dump = nullptr;   // see f2f_dncnn_set_debug_dump

// 0: the default mode (16-bit tensor-core operands); 1: FP32 on the CUDA cores (dncnn_precise.cuh); 2: FP32-class
// split operands on the tensor cores (dncnn_precise2.cuh) -- f2f_dncnn_set_precise
static int g_precise = getenv("F2F_PRECISE") ? atoi(getenv("F2F_PRECISE")) : 0;

// weight gradient of one 64->64 layer (f2f_dncnn_wgrad64) with the gradient-scale and an extra factor applied on reduction
static int wgrad64_scaled(const void *d_act, const float *d_coef_in, void *d_g, const void *d_z, const float *d_coef,
                          float *d_partial, float *d_dw, int H, int W, int accumulate, const unsigned *d_absmax, void *stream,
                          float extra_scale = 1.0f)
{
    if (!d_act || !d_g || !d_z || !d_coef || !d_partial || !d_dw || H < 1 || W < 1) {
        set_error("f2f_dncnn_wgrad64: bad arguments");
        return F2F_EINVAL;
    }
    F2F_CUDA_CHECK((ensure_dyn_smem<wg::k_wgrad64>(wg::SMEM_BYTES)));
    cudaStream_t st = (cudaStream_t)stream;
    wg::Args a;
    a.act = (const h16 *)d_act; a.coef_in = d_coef_in; a.g = (h16 *)d_g; a.z = (const h16 *)d_z; a.coef = d_coef;
    a.partial = d_partial; a.H = H; a.W = W; a.tiles_x = cdiv(W, 16); a.tiles_y = cdiv(H, wg::TR);
    const int ntiles = a.tiles_x * a.tiles_y;
    const int grid = ntiles < num_sms() ? ntiles : num_sms();
    prof_begin(1, st);
    a.pdl_trigger = (g_pdl >> 2) & 1;
    static const int m64 = getenv("F2F_WGRAD_M64") ? atoi(getenv("F2F_WGRAD_M64")) : 1;   // 0: M = 128 throughout (A/B, profiles/r01_m64_ab.log)
    a.m64 = m64;
    a.rev = g_tile_rev;
    // L2 prefetch distance of the producers (tiles of a producer group); 0 turns the prefetches off (A/B)
    static const int pf = getenv("F2F_WGRAD_PF") ? atoi(getenv("F2F_WGRAD_PF")) : 1;
    a.pf = pf < 0 ? 0 : pf > 8 ? 8 : pf;
    a.dbg = getenv("F2F_WGRAD_DBG") ? atoi(getenv("F2F_WGRAD_DBG")) : 0;
    a.trace = dn::g_conv_trace;
    CUtensorMap tm_act, tm_g, tm_z;
    int rc = dnt::make_tensor_map(&tm_act, d_act, H, W, 18, wg::HR);
    if (!rc) rc = dnt::make_tensor_map(&tm_g, d_g, H, W, 16, wg::TR);
    if (!rc) rc = dnt::make_tensor_map(&tm_z, d_z, H, W, 16, wg::TR);
    if (rc) return rc;
    F2F_CUDA_CHECK(launch_pdl(1, wg::k_wgrad64, dim3(grid), dim3(wg::NTHREADS), wg::SMEM_BYTES, st, a, tm_act, tm_g, tm_z));
    prof_end(1, st);
    F2F_CUDA_CHECK(launch_pdl(0, wg::k_wgrad_reduce, dim3(cdiv(9 * 4096, 256)), dim3(256), 0, st, (const float *)d_partial, grid,
                              d_dw, accumulate, (g_pdl >> 3) & 1, d_absmax, extra_scale));
    g_dncnn_launches += 2;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}


#include "dncnn_precise.cuh"
#include "dncnn_precise2.cuh"

}  // namespace net
}  // namespace f2f

using namespace f2f;
using namespace f2f::net;

extern "C" {

size_t f2f_dncnn_param_count(void) { return NPARAMS; }
size_t f2f_dncnn_param_count_c(int channels) { return channels >= 1 ? NPARAMS + (size_t)2 * 576 * (channels - 1) : 0; }

void f2f_dncnn_set_precise(int mode)
{
    g_precise = mode < 0 || mode > 2 ? 1 : mode;
    if (g_precise == 2 && tc05::H16_FORMAT) g_precise = 1;     // the hi / lo split needs FP16's 11 bits per half (BF16 build: mode 1)
}
int f2f_dncnn_get_precise(void) { return g_precise; }

size_t f2f_dncnn_workspace_bytes_c(int B, int H, int W, int train, int channels)
{
    if (B < 1 || H < 1 || W < 1 || channels < 1) return 0;
    if (g_precise == 2) return p2::make_plan2(B, H, W, train, channels).total;
    if (g_precise) return pr::make_plan_f(B, H, W, train, channels).total;
    return make_plan(B, H, W, train, channels).total;
}
size_t f2f_dncnn_workspace_bytes(int B, int H, int W, int train) { return f2f_dncnn_workspace_bytes_c(B, H, W, train, 1); }

void f2f_dncnn_set_debug_dump(void *d_buf) { g_debug_dump = d_buf; }

void f2f_dncnn_workspace_layout(int B, int H, int W, int train, size_t out[8])
{
    if (g_precise == 2) {
        const p2::Plan2 p = p2::make_plan2(B, H, W, train);
        out[0] = p.o_x; out[1] = p.o_a0; out[2] = p.o_z; out[3] = p.act32; out[4] = p.o_g0; out[5] = p.o_g1;
        out[6] = p.o_coef; out[7] = p.o_stats_f;
        return;
    }
    if (g_precise) {
        const pr::PlanF p = pr::make_plan_f(B, H, W, train);
        out[0] = p.o_x; out[1] = p.o_a0; out[2] = p.o_z; out[3] = p.act_bytes; out[4] = p.o_g0; out[5] = p.o_g1;
        out[6] = p.o_coef; out[7] = p.o_stats_f;
        return;
    }
    const Plan p = make_plan(B, H, W, train);
    out[0] = p.o_x; out[1] = p.o_a0; out[2] = p.o_z; out[3] = p.act_bytes; out[4] = p.o_g0; out[5] = p.o_g1;
    out[6] = p.o_coef; out[7] = p.o_stats_f;
}

static int forward_impl(const float *d_params, float *d_bn_running, long long *d_nbt, const float *d_x, float *d_y,
                        void *d_ws, size_t ws_bytes, int B, int H, int W, int train, int residual, void *stream,
                        const rs::LossFuse *fuse, int C = 1);

int f2f_dncnn_forward(const float *d_params, float *d_bn_running, long long *d_nbt, const float *d_x, float *d_y,
                      void *d_ws, size_t ws_bytes, int B, int H, int W, int train, int residual, void *stream)
{
    return forward_impl(d_params, d_bn_running, d_nbt, d_x, d_y, d_ws, ws_bytes, B, H, W, train, residual, stream, nullptr);
}

int f2f_dncnn_forward_c(const float *d_params, float *d_bn_running, long long *d_nbt, const float *d_x, float *d_y,
                        void *d_ws, size_t ws_bytes, int B, int H, int W, int train, int residual, int channels, void *stream)
{
    if (channels != 1 && channels != 3) return gen::bad_channels(channels);
    return forward_impl(d_params, d_bn_running, d_nbt, d_x, d_y, d_ws, ws_bytes, B, H, W, train, residual, stream, nullptr, channels);
}

// fuse != NULL (B = 1, default mode only): the loss of f2f_step_train is formed in the last layer's kernel; the caller has
// zeroed *fuse->loss, and fuse->absmax (filled in here: the workspace's own word) is zeroed below
static int forward_impl(const float *d_params, float *d_bn_running, long long *d_nbt, const float *d_x, float *d_y,
                        void *d_ws, size_t ws_bytes, int B, int H, int W, int train, int residual, void *stream,
                        const rs::LossFuse *fuse, int C)
{
    if (!d_params || !d_bn_running || !d_x || !d_y || !d_ws || B < 1 || H < 1 || W < 1) {
        set_error("f2f_dncnn_forward: bad arguments");
        return F2F_EINVAL;
    }
    if (g_precise == 2)
        return p2::forward(d_params, d_bn_running, d_nbt, d_x, d_y, d_ws, ws_bytes, B, H, W, train, residual, (cudaStream_t)stream, C);
    if (g_precise)
        return pr::forward(d_params, d_bn_running, d_nbt, d_x, d_y, d_ws, ws_bytes, B, H, W, train, residual, (cudaStream_t)stream, C);
    const Plan pl = make_plan(B, H, W, train, C);
    const float *d_w0 = d_params;
    d_params += (size_t)576 * (C - 1);              // off_conv / off_gamma / off_beta / OFF_W16 are written for C = 1
    if (ws_bytes < pl.total) {
        set_error("f2f_dncnn_forward: workspace too small (%zu bytes, need %zu)", ws_bytes, pl.total);
        return F2F_ENOMEM;
    }
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t *ws = (uint8_t *)d_ws;
    const size_t P = (size_t)H * W;
    h16 *wprep = (h16 *)(ws + pl.o_wprep);
    float *coef = (float *)(ws + pl.o_coef);
    double *stats_f = (double *)(ws + pl.o_stats_f);
    h16 *a0 = (h16 *)(ws + pl.o_a0);
    auto zbuf = [&](int k) { return (h16 *)(ws + pl.o_z + pl.act_bytes * (size_t)(train ? k - 1 : (k & 1))); };

    k_prep_all<<<cdiv(30 * WSZ, 256), 256, 0, st>>>(d_params, wprep);
    g_dncnn_launches++;
    if (train) {
        F2F_CUDA_CHECK(cudaMemsetAsync(stats_f, 0, (size_t)NBN * 128 * 8, st));
        F2F_CUDA_CHECK(cudaMemcpyAsync(ws + pl.o_x, d_x, (size_t)B * P * 4 * C, cudaMemcpyDeviceToDevice, st));
    } else {
        k_bn_eval_coeffs<<<NBN, 64, 0, st>>>(d_params, d_bn_running, coef);
        g_dncnn_launches++;
    }
    for (int b = 0; b < B; b++) {
        if (C != 1) {
            const int rc = gen::first_fwd<h16>(C, d_x + b * P * C, d_w0, a0 + b * P * 64, H, W, st);
            if (rc) return rc;
            continue;
        }
        const int rtx = cdiv(W, rs::TW), rnt = rtx * cdiv(H, rs::TH);
        rs::k_first_fwd<<<rnt < num_sms() * 6 ? rnt : num_sms() * 6, rs::FT, 0, st>>>(d_x + b * P, d_w0,
                                                                                    a0 + b * P * 64, H, W, rtx, rnt);
        g_dncnn_launches++;
    }
    F2F_LAUNCH_CHECK();
    for (int k = 1; k <= NBN; k++) {
        const h16 *in = k == 1 ? a0 : zbuf(k - 1);
        const float *cf = k == 1 ? nullptr : coef + (size_t)(k - 2) * C_STRIDE;
        for (int b = 0; b < B; b++) {
            tile_dir(k & 1);      // a0 / z_{k-1} was written in the opposite direction: start where it ended (L2)
            const int rc = f2f_dncnn_conv64(in + b * P * 64, zbuf(k) + b * P * 64, wprep + (size_t)(k - 1) * WSZ,
                                            cf ? cf + C_SCALE : nullptr, cf ? cf + C_SHIFT : nullptr,
                                            train ? stats_f + (size_t)(k - 1) * 128 : nullptr, H, W, stream);
            tile_dir(0);
            if (rc) return rc;
        }
        if (train) {
            F2F_CUDA_CHECK(launch_pdl(0, k_bn_finalize, dim3(1), dim3(64), 0, st, stats_f + (size_t)(k - 1) * 128,
                                      d_params + off_gamma(k), d_params + off_beta(k), d_bn_running + (size_t)(k - 1) * 128,
                                      d_nbt ? d_nbt + (k - 1) : nullptr, coef + (size_t)(k - 1) * C_STRIDE,
                                      (double)B * (double)P, (g_pdl >> 3) & 1));
            g_dncnn_launches++;
        }
    }
    const int tiles_x = cdiv(W, LT), ntiles = tiles_x * cdiv(H, LT);
    rs::LossFuse lf = {};
    if (fuse && train && B == 1) {
        lf = *fuse;
        lf.absmax = (unsigned *)(ws + pl.o_absmax);
        F2F_CUDA_CHECK(cudaMemsetAsync(lf.absmax, 0, 16, st));
    }
    for (int b = 0; b < B; b++) {
        if (C != 1) {
            const int rc = gen::last_fwd<h16>(C, zbuf(NBN) + b * P * 64, coef + (size_t)(NBN - 1) * C_STRIDE, d_params + OFF_W16,
                                              d_x + b * P * C, d_y + b * P * C, H, W, residual, st);
            if (rc) return rc;
            continue;
        }
        static const int last_mma = getenv("F2F_LAST_MMA") ? atoi(getenv("F2F_LAST_MMA")) : 1;   // 0: CUDA-core version (A/B)
        (last_mma ? rs::k_last_fwd<true> : rs::k_last_fwd<false>)<<<ntiles < simt_blocks() ? ntiles : simt_blocks(), 256, 0, st>>>(
            zbuf(NBN) + b * P * 64, coef + (size_t)(NBN - 1) * C_STRIDE, d_params + OFF_W16, d_x + b * P, d_y + b * P, H, W,
            tiles_x, ntiles, residual, lf);
        g_dncnn_launches++;
    }
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

int f2f_dncnn_wgrad64(const void *d_act, const float *d_coef_in, void *d_g, const void *d_z, const float *d_coef,
                      float *d_partial, float *d_dw, int H, int W, int accumulate, void *stream)
{
    return wgrad64_scaled(d_act, d_coef_in, d_g, d_z, d_coef, d_partial, d_dw, H, W, accumulate, nullptr, stream);
}

static int backward_impl(const float *d_params, float *d_grads, const float *d_dy, void *d_ws, size_t ws_bytes, int B, int H,
                         int W, int residual, int accumulate, void *stream, int absmax_ready, int C = 1);

int f2f_dncnn_backward(const float *d_params, float *d_grads, const float *d_dy, void *d_ws, size_t ws_bytes, int B, int H,
                       int W, int residual, int accumulate, void *stream)
{
    return backward_impl(d_params, d_grads, d_dy, d_ws, ws_bytes, B, H, W, residual, accumulate, stream, 0);
}

int f2f_dncnn_backward_c(const float *d_params, float *d_grads, const float *d_dy, void *d_ws, size_t ws_bytes, int B, int H,
                         int W, int residual, int accumulate, int channels, void *stream)
{
    if (channels != 1 && channels != 3) return gen::bad_channels(channels);
    return backward_impl(d_params, d_grads, d_dy, d_ws, ws_bytes, B, H, W, residual, accumulate, stream, 0, channels);
}

// absmax_ready: max|dy| already sits in the workspace (written by the fused last-layer kernel of forward_impl)
static int backward_impl(const float *d_params, float *d_grads, const float *d_dy, void *d_ws, size_t ws_bytes, int B, int H,
                         int W, int residual, int accumulate, void *stream, int absmax_ready, int C)
{
    if (!d_params || !d_grads || !d_dy || !d_ws || B < 1 || H < 1 || W < 1) {
        set_error("f2f_dncnn_backward: bad arguments");
        return F2F_EINVAL;
    }
    if (g_precise == 2)
        return p2::backward(d_params, d_grads, d_dy, d_ws, ws_bytes, B, H, W, residual, accumulate, (cudaStream_t)stream, C);
    if (g_precise)
        return pr::backward(d_params, d_grads, d_dy, d_ws, ws_bytes, B, H, W, residual, accumulate, (cudaStream_t)stream, C);
    const Plan pl = make_plan(B, H, W, 1, C);
    float *d_gw0 = d_grads;
    d_params += (size_t)576 * (C - 1);
    d_grads += (size_t)576 * (C - 1);
    if (ws_bytes < pl.total) {
        set_error("f2f_dncnn_backward: workspace too small (%zu bytes, need %zu)", ws_bytes, pl.total);
        return F2F_ENOMEM;
    }
    F2F_CUDA_CHECK((ensure_dyn_smem<rs::k_last_bwd>(rs::SMEM)));
    F2F_CUDA_CHECK((ensure_dyn_smem<rs::k_first_wgrad>(rs::SMEM)));
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t *ws = (uint8_t *)d_ws;
    const size_t P = (size_t)H * W;
    const double n = (double)B * (double)P;
    const h16 *wprep = (const h16 *)(ws + pl.o_wprep);
    float *coef = (float *)(ws + pl.o_coef);
    double *stats_b = (double *)(ws + pl.o_stats_b);
    double *dw0 = (double *)(ws + pl.o_dw_small), *dw16 = dw0 + 576 * C;
    const float *x = (const float *)(ws + pl.o_x);
    h16 *a0 = (h16 *)(ws + pl.o_a0);
    float *partial = (float *)(ws + pl.o_partial);
    auto zbuf = [&](int k) { return (h16 *)(ws + pl.o_z + pl.act_bytes * (size_t)(k - 1)); };
    h16 *gbuf[2] = {(h16 *)(ws + pl.o_g0), (h16 *)(ws + pl.o_g1)};

    // stats_b and the small double accumulators are adjacent (make_plan)
    unsigned *absmax = (unsigned *)(ws + pl.o_absmax);
    F2F_CUDA_CHECK(cudaMemsetAsync(stats_b, 0, (pl.o_absmax - pl.o_stats_b) + (absmax_ready ? 0 : 16), st));
    if (!absmax_ready) {
        k_absmax<<<num_sms(), 256, 0, st>>>(d_dy, (size_t)B * P * C, absmax);      // gradient scaling, see grad_scale()
        g_dncnn_launches++;
    }
    int cur = 0;   // gbuf[cur] holds g_k
    const int rtx = cdiv(W, rs::TW), rnt = rtx * cdiv(H, rs::TH);
    for (int b = 0; b < B; b++) {
        if (C != 1) {
            const int rc = gen::last_bwd<h16>(C, d_dy + b * P * C, residual ? -1.0f : 1.0f, zbuf(NBN) + b * P * 64,
                                              coef + (size_t)(NBN - 1) * C_STRIDE, d_params + OFF_W16, gbuf[cur] + b * P * 64,
                                              stats_b + (size_t)(NBN - 1) * 128, dw16, H, W, absmax, st);
            if (rc) return rc;
            continue;
        }
        rs::k_last_bwd<<<rnt < num_sms() ? rnt : num_sms(), rs::NT, rs::SMEM, st>>>(
            d_dy + b * P, residual ? -1.0f : 1.0f, zbuf(NBN) + b * P * 64, coef + (size_t)(NBN - 1) * C_STRIDE,
            d_params + OFF_W16, gbuf[cur] + b * P * 64, stats_b + (size_t)(NBN - 1) * 128, dw16, H, W, rtx, rnt, absmax);
        g_dncnn_launches++;
    }
    F2F_LAUNCH_CHECK();
    auto dump = [&](int slot, const void *src) {
        if (g_debug_dump)
            cudaMemcpyAsync((uint8_t *)g_debug_dump + (size_t)slot * pl.act_bytes, src, pl.act_bytes, cudaMemcpyDeviceToDevice, st);
    };
    for (int k = NBN; k >= 1; k--) {
        dump(2 * (k - 1), gbuf[cur]);
        float *cf = coef + (size_t)(k - 1) * C_STRIDE;
        F2F_CUDA_CHECK(launch_pdl(0, k_bn_bwd_coeffs, dim3(1), dim3(64), 0, st, (const double *)(stats_b + (size_t)(k - 1) * 128),
                                  d_params + off_gamma(k), cf, d_grads + off_gamma(k), d_grads + off_beta(k), n, accumulate, (g_pdl >> 3) & 1, (const unsigned *)absmax));
        g_dncnn_launches++;
        const h16 *act = k == 1 ? a0 : zbuf(k - 1);
        const float *cf_in = k == 1 ? nullptr : coef + (size_t)(k - 2) * C_STRIDE;
        for (int b = 0; b < B; b++) {
            tile_dir(1);          // g_k was written top-down (k_last_bwd / the input gradient below): start at its end (L2)
            int rc = wgrad64_scaled(act + b * P * 64, cf_in, gbuf[cur] + b * P * 64, zbuf(k) + b * P * 64, cf, partial,
                                    d_grads + off_conv(k), H, W, accumulate || b > 0, absmax, stream);
            if (rc) return rc;
            tile_dir(0);          // ... and the input gradient reads dz_k from where the weight gradient finished it
            if (b == B - 1) dump(2 * (k - 1) + 1, gbuf[cur]);
            // input gradient: g_{k-1} = relu'/BN-sums( conv(dz_k, W_k^T) )
            if (k >= 2)
                rc = f2f_dncnn_dgrad64(gbuf[cur] + b * P * 64, gbuf[cur ^ 1] + b * P * 64, wprep + (size_t)(NBN + k - 1) * WSZ,
                                       zbuf(k - 1) + b * P * 64, cf_in + C_SCALE, cf_in + C_SHIFT, cf_in + C_MEAN,
                                       cf_in + C_INVSTD, stats_b + (size_t)(k - 2) * 128, H, W, stream);
            else
                rc = f2f_dncnn_dgrad64(gbuf[cur] + b * P * 64, gbuf[cur ^ 1] + b * P * 64, wprep + (size_t)(NBN + k - 1) * WSZ,
                                       a0 + b * P * 64, nullptr, nullptr, nullptr, nullptr, nullptr, H, W, stream);
            if (rc) return rc;
        }
        cur ^= 1;
    }
    dump(30, gbuf[cur]);
    for (int b = 0; b < B; b++) {
        if (C != 1) {
            const int rc = gen::first_wgrad<h16>(C, gbuf[cur] + b * P * 64, x + b * P * C, dw0, H, W, st);
            if (rc) return rc;
            continue;
        }
        rs::k_first_wgrad<<<rnt < num_sms() ? rnt : num_sms(), rs::NT, rs::SMEM, st>>>(gbuf[cur] + b * P * 64, x + b * P, dw0, H,
                                                                                       W, rtx, rnt);
        g_dncnn_launches++;
    }
    k_small_grads<<<cdiv(1152 * C, 256), 256, 0, st>>>(dw0, dw16, d_gw0, d_grads + OFF_W16, 576 * C, accumulate, absmax);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

int f2f_adam_step(float *d_p, const float *d_g, float *d_m, float *d_v, size_t n, long long step, float lr, float beta1,
                  float beta2, float eps, float weight_decay, void *stream)
{
    if (!d_p || !d_g || !d_m || !d_v || step < 1) {
        set_error("f2f_adam_step: bad arguments");
        return F2F_EINVAL;
    }
    if (n == 0) return F2F_OK;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    const float step_size = (float)((double)lr / bc1);
    const float bc2_sqrt = (float)sqrt(bc2);
    const int blocks = (int)((n + 255) / 256 < (size_t)num_sms() * 8 ? (n + 255) / 256 : (size_t)num_sms() * 8);
    k_adam<<<blocks, 256, 0, (cudaStream_t)stream>>>(d_p, d_g, d_m, d_v, n, beta1, beta2, eps, weight_decay, step_size,
                                                     bc2_sqrt);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

// One inner iteration of blind_denoising.py:216-222 as a single call: train-mode forward
// (out = curr - net(curr)), masked loss against the pre-warped previous frame, backward, Adam.
int f2f_step_train(float *d_params, float *d_grads, float *d_m, float *d_v, float *d_bn_running, long long *d_nbt,
                   const float *d_curr, const float *d_warped, const float *d_mask, float *d_out, float *d_dout,
                   double *d_loss, void *d_ws, size_t ws_bytes, int H, int W, int loss_kind, long long step, float lr,
                   float beta1, float beta2, float eps, float weight_decay, void *stream)
{
    if (!d_out || !d_dout || !d_loss || !d_warped || !d_mask || loss_kind < 0 || loss_kind > 2) {
        set_error("f2f_step_train: bad arguments");
        return F2F_EINVAL;
    }
    int rc;
    static const int fuse_loss = getenv("F2F_FUSE_LOSS") ? atoi(getenv("F2F_FUSE_LOSS")) : 1;
    if (!g_precise && fuse_loss) {
        // masked loss, its gradient and max|gradient| in the last layer's kernel (rs::LossFuse)
        rs::LossFuse lf = {d_warped, d_mask, d_dout, d_loss, nullptr, loss_kind, 1.0f / (float)((size_t)H * W)};
        F2F_CUDA_CHECK(cudaMemsetAsync(d_loss, 0, sizeof(double), (cudaStream_t)stream));
        rc = forward_impl(d_params, d_bn_running, d_nbt, d_curr, d_out, d_ws, ws_bytes, 1, H, W, 1, 1, stream, &lf);
        if (rc) return rc;
        rc = backward_impl(d_params, d_grads, d_dout, d_ws, ws_bytes, 1, H, W, 1, 0, stream, 1);
        if (rc) return rc;
        return f2f_adam_step(d_params, d_grads, d_m, d_v, NPARAMS, step, lr, beta1, beta2, eps, weight_decay, stream);
    }
    rc = f2f_dncnn_forward(d_params, d_bn_running, d_nbt, d_curr, d_out, d_ws, ws_bytes, 1, H, W, 1, 1, stream);
    if (rc) return rc;
    rc = f2f_masked_loss(d_out, d_warped, d_mask, d_dout, d_loss, (size_t)H * W, loss_kind, stream);
    if (rc) return rc;
    rc = f2f_dncnn_backward(d_params, d_grads, d_dout, d_ws, ws_bytes, 1, H, W, 1, 0, stream);
    if (rc) return rc;
    return f2f_adam_step(d_params, d_grads, d_m, d_v, NPARAMS, step, lr, beta1, beta2, eps, weight_decay, stream);
}

}  // extern "C"
