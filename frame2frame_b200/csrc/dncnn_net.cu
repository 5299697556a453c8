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
#include <cuda_bf16.h>
#include "f2f_common.cuh"
#include "tc05.cuh"

namespace f2f {
extern long g_dncnn_launches;
namespace net {

using namespace tc05;
typedef __nv_bfloat16 bf16;

constexpr int NBN = 15;
constexpr int WSZ = 64 * 64 * 9;                   // 36864
constexpr size_t OFF_W0 = 0;
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
__global__ void k_prep_all(const float *__restrict__ params, bf16 *__restrict__ wprep)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 30 * WSZ) return;
    const int which = idx / WSZ, e = idx % WSZ;
    const int transpose = which >= NBN, k = (which % NBN) + 1;
    const float *w = params + off_conv(k);
    const int kk = e & 7, n = (e >> 3) & 63, kc = (e >> 9) & 7, tap = e >> 12;
    const int kch = kc * 8 + kk;
    const float v = !transpose ? w[((size_t)n * 64 + kch) * 9 + tap] : w[((size_t)kch * 64 + n) * 9 + (8 - tap)];
    wprep[idx] = __float2bfloat16(v);
}

// Conv(1->64, 3x3, pad 1) + ReLU (net.py:13-14): FP32 image -> NHWC BF16 activation.
// 8 lanes per pixel, each lane 8 output channels (one 16-byte store; a pixel's 128 B are contiguous).
__global__ void __launch_bounds__(256) k_conv_first(const float *__restrict__ x, const float *__restrict__ w0,
                                                    bf16 *__restrict__ a0, int H, int W)
{
    const int chunk = threadIdx.x & 7;
    float w[9][8];
#pragma unroll
    for (int t = 0; t < 9; t++)
#pragma unroll
        for (int j = 0; j < 8; j++) w[t][j] = w0[(chunk * 8 + j) * 9 + t];
    const size_t P = (size_t)H * W;
    constexpr int NP = 2;       // two pixels per thread and iteration, their 18 loads in flight together
    for (size_t p0 = (size_t)blockIdx.x * (32 * NP) + (threadIdx.x >> 3); p0 < P; p0 += (size_t)gridDim.x * (32 * NP)) {
        float xv[NP][9];
#pragma unroll
        for (int u = 0; u < NP; u++) {
            const size_t p = p0 + 32 * u;
            const int y = (int)(p / W), xx = (int)(p % W);
#pragma unroll
            for (int t = 0; t < 9; t++) {
                const int yy = y + t / 3 - 1, xq = xx + t % 3 - 1;
                xv[u][t] = (p < P && yy >= 0 && yy < H && xq >= 0 && xq < W) ? __ldg(x + (size_t)yy * W + xq) : 0.0f;
            }
        }
#pragma unroll
        for (int u = 0; u < NP; u++) {
            const size_t p = p0 + 32 * u;
            if (p >= P) continue;
            float acc[8];
#pragma unroll
            for (int j = 0; j < 8; j++) acc[j] = 0.0f;
#pragma unroll
            for (int t = 0; t < 9; t++)
#pragma unroll
                for (int j = 0; j < 8; j++) acc[j] = fmaf(xv[u][t], w[t][j], acc[j]);
            uint4 o;
            o.x = pack_bf16(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f));
            o.y = pack_bf16(fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
            o.z = pack_bf16(fmaxf(acc[4], 0.f), fmaxf(acc[5], 0.f));
            o.w = pack_bf16(fmaxf(acc[6], 0.f), fmaxf(acc[7], 0.f));
            *reinterpret_cast<uint4 *>(a0 + p * 64 + chunk * 8) = o;
        }
    }
}

// Train-mode BatchNorm2d bookkeeping of one layer: batch mean / biased variance from
// the conv epilogue's sums, scale/shift for the next layer's operand staging, running-stat EMA
// with the unbiased variance, num_batches_tracked.
__global__ void k_bn_finalize(const double *__restrict__ stats, const float *__restrict__ gamma,
                              const float *__restrict__ beta, float *__restrict__ running, long long *nbt,
                              float *__restrict__ coef, double n)
{
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

// BatchNorm backward of one layer from the sums the dgrad epilogue produced (sum g, sum g*xhat):
// dgamma, dbeta into the gradient vector and dz = A*g + B*z + C for the operand staging of the
// layer's weight gradient.
__global__ void k_bn_bwd_coeffs(const double *__restrict__ stats, const float *__restrict__ gamma,
                                float *__restrict__ coef, float *__restrict__ dgamma, float *__restrict__ dbeta,
                                double n, int accumulate)
{
    const int c = threadIdx.x;
    if (c >= 64) return;
    const double db = stats[c], dg = stats[64 + c];
    const double invstd = coef[C_INVSTD + c], mean = coef[C_MEAN + c], ga = gamma[c];
    coef[C_A + c] = (float)(ga * invstd);
    coef[C_B + c] = (float)(-ga * invstd * invstd * dg / n);
    coef[C_C + c] = (float)(ga * invstd * (-db / n + mean * invstd * dg / n));
    dgamma[c] = (accumulate ? dgamma[c] : 0.0f) + (float)dg;
    dbeta[c] = (accumulate ? dbeta[c] : 0.0f) + (float)db;
}

// Sums NV per-thread FP32 accumulators over the threads of the block that own the same
// 8-channel chunk (lane % 8) and adds the result to double accumulators in global memory.
// dst(i) maps value i of chunk `chunk` to its address.
template <int NV, class F>
__device__ __forceinline__ void block_chunk_reduce(float (&v)[NV], float *red /*[8 warps][8 chunks][NV]*/, F dst)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, chunk = threadIdx.x & 7;
#pragma unroll
    for (int i = 0; i < NV; i++) {
        v[i] += __shfl_xor_sync(0xffffffffu, v[i], 8);
        v[i] += __shfl_xor_sync(0xffffffffu, v[i], 16);
    }
    if (lane < 8) {
#pragma unroll
        for (int i = 0; i < NV; i++) red[(warp * 8 + chunk) * NV + i] = v[i];
    }
    __syncthreads();
    for (int e = threadIdx.x; e < 8 * NV; e += blockDim.x) {
        const int ch = e / NV, i = e % NV;
        double t = 0.0;
        for (int w = 0; w < 8; w++) t += (double)red[(w * 8 + ch) * NV + i];
        atomicAdd(dst(ch, i), t);
    }
}

// Conv(64->1, 3x3, pad 1) (net.py:19) with the BatchNorm+ReLU of layer 15 applied on load.
// Every input pixel's 64 channels are read once: 8 lanes form the nine per-tap partial dot
// products of that pixel, the output pixel then sums nine shifted partials from shared memory.
// out = n (the noise estimate, SaoYan's forward) or x - n (net.py:31) when `residual`.
constexpr int LT = 16, LH = LT + 2, LHP = LH * LH;
constexpr int LBW_STRIDE = 76;     // floats between the per-chunk weight tables in shared memory (conflict-free 16-byte reads)
__global__ void __launch_bounds__(256) k_conv_last(const bf16 *__restrict__ z, const float *__restrict__ coef,
                                                   const float *__restrict__ w16, const float *__restrict__ x,
                                                   float *__restrict__ out, int H, int W, int tiles_x, int ntiles,
                                                   int residual)
{
    __shared__ float T[LHP * 9 + 4];
    __shared__ __align__(16) float wS[8 * LBW_STRIDE];   // [chunk][tap][8 channels]: registers go to the loads in flight
    const int chunk = threadIdx.x & 7;
    for (int e = threadIdx.x; e < 8 * 72; e += 256) {
        const int c = e / 72, t = (e % 72) >> 3, j = e & 7;
        wS[c * LBW_STRIDE + t * 8 + j] = w16[(c * 8 + j) * 9 + t];
    }
    float sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { sc[j] = coef[C_SCALE + chunk * 8 + j]; sh[j] = coef[C_SHIFT + chunk * 8 + j]; }
    __syncthreads();
    const float4 *wv = reinterpret_cast<const float4 *>(wS + chunk * LBW_STRIDE);
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int ty0 = (tile / tiles_x) * LT, tx0 = (tile % tiles_x) * LT;
        // the whole haloed tile's loads are issued before any arithmetic (one exposed memory latency per tile, not eleven)
        constexpr int ROUNDS = (LHP * 8 + 255) / 256;
        uint4 v[ROUNDS];
        uint32_t okm = 0;
#pragma unroll
        for (int r = 0; r < ROUNDS; r++) {
            const int pix = (threadIdx.x + 256 * r) >> 3;
            const int gy = ty0 - 1 + pix / LH, gx = tx0 - 1 + pix % LH;
            v[r] = make_uint4(0, 0, 0, 0);
            if (pix < LHP && gy >= 0 && gy < H && gx >= 0 && gx < W) {
                okm |= 1u << r;
                v[r] = __ldg(reinterpret_cast<const uint4 *>(z + ((size_t)gy * W + gx) * 64 + chunk * 8));
            }
        }
#pragma unroll
        for (int r = 0; r < ROUNDS; r++) {
            const int pix = (threadIdx.x + 256 * r) >> 3;
            float part[9];
#pragma unroll
            for (int t = 0; t < 9; t++) part[t] = 0.0f;
            if ((okm >> r) & 1u) {
                const uint32_t vw[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
                float a[8];
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const float2 f = unpack_bf16(vw[q]);
                    a[2 * q] = fmaxf(fmaf(f.x, sc[2 * q], sh[2 * q]), 0.0f);
                    a[2 * q + 1] = fmaxf(fmaf(f.y, sc[2 * q + 1], sh[2 * q + 1]), 0.0f);
                }
#pragma unroll
                for (int t = 0; t < 9; t++) {
                    const float4 w0 = wv[2 * t], w1 = wv[2 * t + 1];
                    part[t] = fmaf(a[0], w0.x, fmaf(a[1], w0.y, fmaf(a[2], w0.z, fmaf(a[3], w0.w, 0.0f)))) +
                              fmaf(a[4], w1.x, fmaf(a[5], w1.y, fmaf(a[6], w1.z, fmaf(a[7], w1.w, 0.0f))));
                }
            }
            // every lane takes part in the shuffles (groups of 8 lanes share a pixel); only the store is guarded
#pragma unroll
            for (int t = 0; t < 9; t++) {
                part[t] += __shfl_xor_sync(0xffffffffu, part[t], 1);
                part[t] += __shfl_xor_sync(0xffffffffu, part[t], 2);
                part[t] += __shfl_xor_sync(0xffffffffu, part[t], 4);
            }
            // lane c of the group stores tap c (lane 0 also tap 8)
            float mine = part[0];
#pragma unroll
            for (int t = 1; t < 8; t++) mine = (chunk == t) ? part[t] : mine;
            if (pix < LHP) {
                T[pix * 9 + chunk] = mine;
                if (chunk == 0) T[pix * 9 + 8] = part[8];
            }
        }
        __syncthreads();
        {
            const int y = threadIdx.x >> 4, xx = threadIdx.x & 15;
            const int gy = ty0 + y, gx = tx0 + xx;
            float n = 0.0f;
#pragma unroll
            for (int t = 0; t < 9; t++) n += T[((y + t / 3) * LH + xx + t % 3) * 9 + t];
            if (gy < H && gx < W) {
                const size_t i = (size_t)gy * W + gx;
                out[i] = residual ? x[i] - n : n;
            }
        }
        __syncthreads();
    }
}

// Backward of the last layer: input gradient (1->64 transposed conv) masked by layer 15's ReLU,
// its BatchNorm-backward sums, and the last layer's weight gradient -- one pass over z15.
//   g15[q][ic]      = relu'(.) * sum_k dn[q - k + 1] * w16[ic][k]
//   dW16[ic][k]     = sum_q a15[q][ic] * dn[q - k + 1]
constexpr int LBW_SMEM = 64 * 88 * 4 + 8 * LBW_STRIDE * 4;
__global__ void __launch_bounds__(256, 1) k_conv_last_bwd(const float *__restrict__ dy, float dy_sign,
                                                          const bf16 *__restrict__ z, const float *__restrict__ coef,
                                                          const float *__restrict__ w16, bf16 *__restrict__ g,
                                                          double *__restrict__ stats, double *__restrict__ dw16, int H,
                                                          int W)
{
    extern __shared__ float red[];
    float *wS = red + 64 * 88;                   // [chunk][tap][8 channels]
    const int chunk = threadIdx.x & 7;
    for (int e = threadIdx.x; e < 8 * 72; e += 256) {
        const int c = e / 72, t = (e % 72) >> 3, j = e & 7;
        wS[c * LBW_STRIDE + t * 8 + j] = w16[(c * 8 + j) * 9 + t];
    }
    float sc[8], sh[8], mu[8], is[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        sc[j] = coef[C_SCALE + chunk * 8 + j]; sh[j] = coef[C_SHIFT + chunk * 8 + j];
        mu[j] = coef[C_MEAN + chunk * 8 + j]; is[j] = coef[C_INVSTD + chunk * 8 + j];
    }
    __syncthreads();
    const float4 *wv = reinterpret_cast<const float4 *>(wS + chunk * LBW_STRIDE);
    float acc[88];   // [0,72): dW16[tap][j]; [72,80): sum g; [80,88): sum g*xhat
#pragma unroll
    for (int i = 0; i < 88; i++) acc[i] = 0.0f;
    const size_t P = (size_t)H * W;
    // Two pixels per thread and iteration, all their loads issued before any arithmetic: with 8 warps per SM
    // (the 88 accumulators bound the occupancy) the loop is otherwise one exposed DRAM latency per pixel.
    constexpr int NP = 2;
    for (size_t p0 = (size_t)blockIdx.x * (32 * NP) + (threadIdx.x >> 3); p0 < P; p0 += (size_t)gridDim.x * (32 * NP)) {
        uint4 v[NP];
        float dn[NP][9];
        bool ok[NP];
#pragma unroll
        for (int u = 0; u < NP; u++) {
            const size_t p = p0 + 32 * u;
            ok[u] = p < P;
            v[u] = make_uint4(0, 0, 0, 0);
            if (ok[u]) v[u] = __ldg(reinterpret_cast<const uint4 *>(z + p * 64 + chunk * 8));
            const int y = (int)(p / W), xx = (int)(p % W);
#pragma unroll
            for (int t = 0; t < 9; t++) {
                const int yy = y - (t / 3) + 1, xq = xx - (t % 3) + 1;
                dn[u][t] = (ok[u] && yy >= 0 && yy < H && xq >= 0 && xq < W) ? dy_sign * __ldg(dy + (size_t)yy * W + xq) : 0.0f;
            }
        }
#pragma unroll
        for (int u = 0; u < NP; u++) {
            if (!ok[u]) continue;
            const size_t p = p0 + 32 * u;
            const uint32_t vw[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
            float zz[8], a[8], s[8];
#pragma unroll
            for (int q = 0; q < 4; q++) { const float2 f = unpack_bf16(vw[q]); zz[2 * q] = f.x; zz[2 * q + 1] = f.y; }
#pragma unroll
            for (int j = 0; j < 8; j++) { a[j] = fmaf(zz[j], sc[j], sh[j]); s[j] = 0.0f; }   // a = pre-ReLU BN output for now
#pragma unroll
            for (int t = 0; t < 9; t++) {
                const float4 w0 = wv[2 * t], w1 = wv[2 * t + 1];
                const float wt[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    s[j] = fmaf(dn[u][t], wt[j], s[j]);
                    acc[t * 8 + j] = fmaf(fmaxf(a[j], 0.0f), dn[u][t], acc[t * 8 + j]);
                }
            }
            float gg[8];
#pragma unroll
            for (int j = 0; j < 8; j++) gg[j] = a[j] > 0.0f ? s[j] : 0.0f;
            uint4 o;
            o.x = pack_bf16(gg[0], gg[1]); o.y = pack_bf16(gg[2], gg[3]);
            o.z = pack_bf16(gg[4], gg[5]); o.w = pack_bf16(gg[6], gg[7]);
            *reinterpret_cast<uint4 *>(g + p * 64 + chunk * 8) = o;
            const uint32_t ow[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
            for (int q = 0; q < 4; q++) {                  // sums of the ROUNDED gradient, like the tcgen05 epilogue
                const float2 f = unpack_bf16(ow[q]);
                acc[72 + 2 * q] += f.x; acc[72 + 2 * q + 1] += f.y;
                acc[80 + 2 * q] = fmaf(f.x, (zz[2 * q] - mu[2 * q]) * is[2 * q], acc[80 + 2 * q]);
                acc[80 + 2 * q + 1] = fmaf(f.y, (zz[2 * q + 1] - mu[2 * q + 1]) * is[2 * q + 1], acc[80 + 2 * q + 1]);
            }
        }
    }
    __syncthreads();
    block_chunk_reduce<88>(acc, red, [&](int ch, int i) -> double * {
        if (i < 72) return dw16 + (ch * 8 + (i & 7)) * 9 + (i >> 3);
        if (i < 80) return stats + ch * 8 + (i - 72);
        return stats + 64 + ch * 8 + (i - 80);
    });
}

// Weight gradient of the first layer: dW0[oc][k] = sum_p g0[p][oc] * x[p + k - 1]  (g0 already ReLU-masked)
__global__ void __launch_bounds__(256, 1) k_wgrad_first(const bf16 *__restrict__ g0, const float *__restrict__ x,
                                                        double *__restrict__ dw0, int H, int W)
{
    extern __shared__ float red[];
    const int chunk = threadIdx.x & 7;
    float acc[72];
#pragma unroll
    for (int i = 0; i < 72; i++) acc[i] = 0.0f;
    const size_t P = (size_t)H * W;
    constexpr int NP = 2;       // two pixels per thread and iteration, all their loads in flight together
    for (size_t p0 = (size_t)blockIdx.x * (32 * NP) + (threadIdx.x >> 3); p0 < P; p0 += (size_t)gridDim.x * (32 * NP)) {
        float xv[NP][9];
        uint4 v[NP];
#pragma unroll
        for (int u = 0; u < NP; u++) {
            const size_t p = p0 + 32 * u;
            const int y = (int)(p / W), xx = (int)(p % W);
            v[u] = make_uint4(0, 0, 0, 0);          // a zero gradient adds nothing: out-of-range pixels need no branch below
            if (p < P) v[u] = __ldg(reinterpret_cast<const uint4 *>(g0 + p * 64 + chunk * 8));
#pragma unroll
            for (int t = 0; t < 9; t++) {
                const int yy = y + t / 3 - 1, xq = xx + t % 3 - 1;
                xv[u][t] = (p < P && yy >= 0 && yy < H && xq >= 0 && xq < W) ? __ldg(x + (size_t)yy * W + xq) : 0.0f;
            }
        }
#pragma unroll
        for (int u = 0; u < NP; u++) {
            const uint32_t vw[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float2 f = unpack_bf16(vw[q]);
#pragma unroll
                for (int t = 0; t < 9; t++) {
                    acc[t * 8 + 2 * q] = fmaf(f.x, xv[u][t], acc[t * 8 + 2 * q]);
                    acc[t * 8 + 2 * q + 1] = fmaf(f.y, xv[u][t], acc[t * 8 + 2 * q + 1]);
                }
            }
        }
    }
    block_chunk_reduce<72>(acc, red, [&](int ch, int i) -> double * { return dw0 + (ch * 8 + (i & 7)) * 9 + (i >> 3); });
}

__global__ void k_small_grads(const double *__restrict__ dw0, const double *__restrict__ dw16, float *__restrict__ grads,
                              int accumulate)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 1152) return;
    float *dst = i < 576 ? grads + OFF_W0 + i : grads + OFF_W16 + (i - 576);
    const double v = i < 576 ? dw0[i] : dw16[i - 576];
    *dst = (accumulate ? *dst : 0.0f) + (float)v;
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
constexpr int NTHREADS = (9 + 8 * (NGROUPS - 1)) * 32;   // group 0: warps 0-7, MMA: warp 8, group 1: warps 9-16
constexpr int TMEM_COLS = 512;
constexpr int PARTIAL = 6 * 128 * 64;   // floats per CTA
constexpr int HPIX = HR * 18, DPIX = TR * 16;

struct Args {
    const bf16 *act;       // layer input: z_{k-1} (BN+ReLU applied on load) or a_0 (raw)
    const float *coef_in;  // scale/shift of layer k-1, or NULL
    bf16 *g;               // in: masked gradient g_k; out: dz_k
    const bf16 *z;         // pre-BN output z_k
    const float *coef;     // A, B, C of layer k
    float *partial;        // [gridDim.x][PARTIAL]
    int H, W, tiles_x, tiles_y;
};

__global__ void __launch_bounds__(NTHREADS, 1) k_wgrad64(Args a)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform
    const int ntiles = a.tiles_x * a.tiles_y;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + OFF_BAR);
    const uint32_t bar0 = smem_u32(bars);
    auto BAR = [&](int i) { return bar0 + 8u * i; };   // full[STAGES] @0, empty[STAGES] @STAGES, done @2*STAGES; tmem ptr after
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(bars + 2 * STAGES + 2);

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(BAR(s), 256); mbar_init(BAR(STAGES + s), 1); }
        mbar_init(BAR(2 * STAGES), 1);
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

    // role of this warp: producer warp index pw in 0..7 of group grp, or the MMA warp
    const int grp = warp > MMA_WARP ? 1 : 0;
    const int pw = warp > MMA_WARP ? warp - (MMA_WARP + 1) : warp;

    if (warp != MMA_WARP && pw < 4) {
        // ===== producer of the activation tile (HR x 18 halo) =====
        const int ptid = pw * 32 + lane, chunk = ptid & 7;
        float sc[8], sh[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            sc[j] = a.coef_in ? a.coef_in[C_SCALE + chunk * 8 + j] : 1.0f;
            sh[j] = a.coef_in ? a.coef_in[C_SHIFT + chunk * 8 + j] : 0.0f;
        }
        const bool bn = a.coef_in != nullptr;
        int it = grp;
        for (int tile = blockIdx.x + grp * gridDim.x; tile < ntiles; tile += NGROUPS * gridDim.x, it += NGROUPS) {
            const int s = it % STAGES;
            mbar_wait(BAR(STAGES + s), ((it / STAGES) & 1) ^ 1);
            uint8_t *dstA = smem + s * STAGE;
            const int ty0 = (tile / a.tiles_x) * TR - 1, tx0 = (tile % a.tiles_x) * 16 - 1;
            constexpr int U = (HPIX + 15) / 16;   // the whole tile's loads in flight at once
            uint4 v[U];
            uint32_t inb = 0;
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int pix = (ptid >> 3) + 16 * u;
                const int gy = ty0 + pix / 18, gx = tx0 + pix % 18;
                v[u] = make_uint4(0, 0, 0, 0);
                if (pix < HPIX && gy >= 0 && gy < a.H && gx >= 0 && gx < a.W) {
                    inb |= 1u << u;
                    v[u] = __ldg(reinterpret_cast<const uint4 *>(a.act + ((size_t)gy * a.W + gx) * 64 + chunk * 8));
                }
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int pix = (ptid >> 3) + 16 * u;
                if (pix < HPIX) {
                    uint4 o = v[u];
                    if (bn && ((inb >> u) & 1u)) {
                        uint32_t w[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const float2 f = unpack_bf16(w[j]);
                            w[j] = pack_bf16(fmaxf(fmaf(f.x, sc[2 * j], sh[2 * j]), 0.0f),
                                             fmaxf(fmaf(f.y, sc[2 * j + 1], sh[2 * j + 1]), 0.0f));
                        }
                        o = make_uint4(w[0], w[1], w[2], w[3]);
                    }
                    *reinterpret_cast<uint4 *>(dstA + (pix / 18) * A_R + chunk * A_S + (pix % 18) * 16) = o;
                }
            }
            fence_proxy_async();
            mbar_arrive(BAR(s));
        }
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
                    if (j < 3 || row < 64) {
                        float4 *o = reinterpret_cast<float4 *>(dst + ((size_t)j * 128 + row) * 64 + part * 32);
#pragma unroll
                        for (int q = 0; q < 8; q++)
                            o[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                               __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
                    }
                }
            }
        }
    } else if (warp != MMA_WARP) {
        // ===== producer of the dz tile (TR x 16, no halo); writes dz back over g =====
        const int ptid = (pw - 4) * 32 + lane, chunk = ptid & 7;
        float cA[8], cB[8], cC[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            cA[j] = a.coef[C_A + chunk * 8 + j]; cB[j] = a.coef[C_B + chunk * 8 + j]; cC[j] = a.coef[C_C + chunk * 8 + j];
        }
        int it = grp;
        for (int tile = blockIdx.x + grp * gridDim.x; tile < ntiles; tile += NGROUPS * gridDim.x, it += NGROUPS) {
            const int s = it % STAGES;
            mbar_wait(BAR(STAGES + s), ((it / STAGES) & 1) ^ 1);
            uint8_t *dstD = smem + s * STAGE + A_BYTES;
            const int ty0 = (tile / a.tiles_x) * TR, tx0 = (tile % a.tiles_x) * 16;
            constexpr int U = DPIX / 32;          // two passes of 4 rounds x 16 pixels (register budget: 17 warps -> 96 registers)
#pragma unroll 1
            for (int r0 = 0; r0 < DPIX / 16; r0 += U) {
            uint4 vg[U], vz[U];
            uint32_t inb = 0;
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int pix = (ptid >> 3) + 16 * (r0 + u);
                const int gy = ty0 + (pix >> 4), gx = tx0 + (pix & 15);
                vg[u] = vz[u] = make_uint4(0, 0, 0, 0);
                if (gy < a.H && gx < a.W) {
                    inb |= 1u << u;
                    const size_t o = ((size_t)gy * a.W + gx) * 64 + chunk * 8;
                    vg[u] = *reinterpret_cast<const uint4 *>(a.g + o);
                    vz[u] = __ldg(reinterpret_cast<const uint4 *>(a.z + o));
                }
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int pix = (ptid >> 3) + 16 * (r0 + u);
                uint4 o = make_uint4(0, 0, 0, 0);
                if ((inb >> u) & 1u) {
                    const uint32_t gw[4] = {vg[u].x, vg[u].y, vg[u].z, vg[u].w};
                    const uint32_t zw[4] = {vz[u].x, vz[u].y, vz[u].z, vz[u].w};
                    uint32_t w[4];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const float2 gg = unpack_bf16(gw[j]), zz = unpack_bf16(zw[j]);
                        w[j] = pack_bf16(fmaf(cA[2 * j], gg.x, fmaf(cB[2 * j], zz.x, cC[2 * j])),
                                         fmaf(cA[2 * j + 1], gg.y, fmaf(cB[2 * j + 1], zz.y, cC[2 * j + 1])));
                    }
                    o = make_uint4(w[0], w[1], w[2], w[3]);
                    const int gy = ty0 + (pix >> 4), gx = tx0 + (pix & 15);
                    *reinterpret_cast<uint4 *>(a.g + ((size_t)gy * a.W + gx) * 64 + chunk * 8) = o;
                }
                *reinterpret_cast<uint4 *>(dstD + (pix >> 4) * D_R + chunk * D_S + (pix & 15) * 16) = o;
            }
            }
            fence_proxy_async();
            mbar_arrive(BAR(s));
        }
    } else {
        // ===== MMA issuer: converged warp, the elected lane issues (see dncnn.cu) =====
        const uint32_t el = elect_one();
        constexpr uint32_t IDESC = idesc_bf16(128, 64, 1, 1);
        // descriptors differ only in the start-address field (low word, 16-byte units)
        const uint64_t ad0 = smem_desc(0, 128, A_S), bd0 = smem_desc(0, 128, D_S);
        const uint32_t a_hi = (uint32_t)(ad0 >> 32), b_hi = (uint32_t)(bd0 >> 32);
        int it = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, it++) {
            const int s = it % STAGES;
            mbar_wait(BAR(s), (it / STAGES) & 1);
            fence_after_sync();
            const uint32_t a_base = smem_u32(smem + s * STAGE);
            const uint32_t a_lo = (uint32_t)ad0 | ((a_base & 0x3FFFFu) >> 4);
            const uint32_t b_lo = (uint32_t)bd0 | (((a_base + A_BYTES) & 0x3FFFFu) >> 4);
#pragma unroll
            for (int r = 0; r < TR; r++) {
                const uint64_t bd = ((uint64_t)b_hi << 32) | (b_lo + r * (D_R / 16));
#pragma unroll
                for (int j = 0; j < 6; j++) {
                    const int dy = j < 3 ? 0 : 2, dx = j % 3;
                    const uint64_t ad = ((uint64_t)a_hi << 32) | (a_lo + (r + dy) * (A_R / 16) + dx);
                    mma_ss_el(tmem_base + j * 64, ad, bd, IDESC, (it | r) != 0, el);
                }
            }
            if (el) mma_commit(BAR(STAGES + s));
            __syncwarp();
        }
        if (el) mma_commit(BAR(2 * STAGES));
        __syncwarp();
    }
    fence_before_sync();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, TMEM_COLS);
}

// dW[oc][ic][ky][kx] (+)= sum over CTA partials.  tap (ky,kx): ky < 2 -> accumulator kx, rows
// ky*64 + ic; ky == 2 -> accumulator 3 + kx, rows ic.
__global__ void k_wgrad_reduce(const float *__restrict__ partial, int nparts, float *__restrict__ dw, int accumulate)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 9 * 4096) return;
    const int oc = idx & 63, ic = (idx >> 6) & 63, tap = idx >> 12;
    const int ky = tap / 3, kx = tap % 3;
    const int j = ky < 2 ? kx : 3 + kx, row = ky < 2 ? ky * 64 + ic : ic;
    const float *src = partial + ((size_t)j * 128 + row) * 64 + oc;
    float t = 0.0f;
    for (int c = 0; c < nparts; c++) t += src[(size_t)c * PARTIAL];
    float *d = dw + ((size_t)oc * 64 + ic) * 9 + tap;
    *d = (accumulate ? *d : 0.0f) + t;
}
}  // namespace wg

// --------------------------------------------------------------- workspace plan
struct Plan {
    size_t act_bytes;       // one NHWC BF16 activation of the whole batch
    size_t o_x, o_a0, o_z, o_g0, o_g1, o_wprep, o_stats_f, o_stats_b, o_dw_small, o_coef, o_partial, total;
    int nz;
};

static int num_sms()
{
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}

static Plan make_plan(int B, int H, int W, int train)
{
    Plan p;
    const size_t P = (size_t)H * W;
    p.act_bytes = align_up((size_t)B * P * 128, 1024);
    size_t o = 0;
    auto take = [&](size_t bytes) { const size_t r = o; o += align_up(bytes, 1024); return r; };
    p.o_x = take((size_t)B * P * 4);
    p.o_a0 = take(p.act_bytes);
    p.nz = train ? NBN : 2;                       // eval: ping-pong
    p.o_z = take(p.act_bytes * p.nz);
    p.o_g0 = train ? take(p.act_bytes) : 0;
    p.o_g1 = train ? take(p.act_bytes) : 0;
    p.o_wprep = take((size_t)30 * WSZ * 2);
    p.o_stats_f = take((size_t)NBN * 128 * 8);
    p.o_stats_b = take((size_t)NBN * 128 * 8);     // contiguous with o_dw_small: one memset
    p.o_dw_small = take((size_t)1152 * 8);
    p.o_coef = take((size_t)NBN * C_STRIDE * 4);
    p.o_partial = train ? take((size_t)148 * 2 * wg::PARTIAL * 4) : 0;
    p.total = o;
    return p;
}

static int simt_blocks() { return num_sms() * 4; }

static void *g_debug_dump = nullptr;   // see f2f_dncnn_set_debug_dump

}  // namespace net
}  // namespace f2f

using namespace f2f;
using namespace f2f::net;

extern "C" {

size_t f2f_dncnn_param_count(void) { return NPARAMS; }

size_t f2f_dncnn_workspace_bytes(int B, int H, int W, int train)
{
    if (B < 1 || H < 1 || W < 1) return 0;
    return make_plan(B, H, W, train).total;
}

void f2f_dncnn_set_debug_dump(void *d_buf) { g_debug_dump = d_buf; }

void f2f_dncnn_workspace_layout(int B, int H, int W, int train, size_t out[8])
{
    const Plan p = make_plan(B, H, W, train);
    out[0] = p.o_x; out[1] = p.o_a0; out[2] = p.o_z; out[3] = p.act_bytes; out[4] = p.o_g0; out[5] = p.o_g1;
    out[6] = p.o_coef; out[7] = p.o_stats_f;
}

int f2f_dncnn_forward(const float *d_params, float *d_bn_running, long long *d_nbt, const float *d_x, float *d_y,
                      void *d_ws, size_t ws_bytes, int B, int H, int W, int train, int residual, void *stream)
{
    if (!d_params || !d_bn_running || !d_x || !d_y || !d_ws || B < 1 || H < 1 || W < 1) {
        set_error("f2f_dncnn_forward: bad arguments");
        return F2F_EINVAL;
    }
    const Plan pl = make_plan(B, H, W, train);
    if (ws_bytes < pl.total) {
        set_error("f2f_dncnn_forward: workspace too small (%zu bytes, need %zu)", ws_bytes, pl.total);
        return F2F_ENOMEM;
    }
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t *ws = (uint8_t *)d_ws;
    const size_t P = (size_t)H * W;
    bf16 *wprep = (bf16 *)(ws + pl.o_wprep);
    float *coef = (float *)(ws + pl.o_coef);
    double *stats_f = (double *)(ws + pl.o_stats_f);
    bf16 *a0 = (bf16 *)(ws + pl.o_a0);
    auto zbuf = [&](int k) { return (bf16 *)(ws + pl.o_z + pl.act_bytes * (size_t)(train ? k - 1 : (k & 1))); };

    k_prep_all<<<cdiv(30 * WSZ, 256), 256, 0, st>>>(d_params, wprep);
    g_dncnn_launches++;
    if (train) {
        F2F_CUDA_CHECK(cudaMemsetAsync(stats_f, 0, (size_t)NBN * 128 * 8, st));
        F2F_CUDA_CHECK(cudaMemcpyAsync(ws + pl.o_x, d_x, (size_t)B * P * 4, cudaMemcpyDeviceToDevice, st));
    } else {
        k_bn_eval_coeffs<<<NBN, 64, 0, st>>>(d_params, d_bn_running, coef);
        g_dncnn_launches++;
    }
    for (int b = 0; b < B; b++) {
        k_conv_first<<<simt_blocks(), 256, 0, st>>>(d_x + b * P, d_params + OFF_W0, a0 + b * P * 64, H, W);
        g_dncnn_launches++;
    }
    F2F_LAUNCH_CHECK();
    for (int k = 1; k <= NBN; k++) {
        const bf16 *in = k == 1 ? a0 : zbuf(k - 1);
        const float *cf = k == 1 ? nullptr : coef + (size_t)(k - 2) * C_STRIDE;
        for (int b = 0; b < B; b++) {
            const int rc = f2f_dncnn_conv64(in + b * P * 64, zbuf(k) + b * P * 64, wprep + (size_t)(k - 1) * WSZ,
                                            cf ? cf + C_SCALE : nullptr, cf ? cf + C_SHIFT : nullptr,
                                            train ? stats_f + (size_t)(k - 1) * 128 : nullptr, H, W, stream);
            if (rc) return rc;
        }
        if (train) {
            k_bn_finalize<<<1, 64, 0, st>>>(stats_f + (size_t)(k - 1) * 128, d_params + off_gamma(k), d_params + off_beta(k),
                                            d_bn_running + (size_t)(k - 1) * 128, d_nbt ? d_nbt + (k - 1) : nullptr,
                                            coef + (size_t)(k - 1) * C_STRIDE, (double)B * (double)P);
            g_dncnn_launches++;
        }
    }
    const int tiles_x = cdiv(W, LT), ntiles = tiles_x * cdiv(H, LT);
    for (int b = 0; b < B; b++) {
        k_conv_last<<<ntiles < simt_blocks() ? ntiles : simt_blocks(), 256, 0, st>>>(
            zbuf(NBN) + b * P * 64, coef + (size_t)(NBN - 1) * C_STRIDE, d_params + OFF_W16, d_x + b * P, d_y + b * P, H, W,
            tiles_x, ntiles, residual);
        g_dncnn_launches++;
    }
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

int f2f_dncnn_wgrad64(const void *d_act, const float *d_coef_in, void *d_g, const void *d_z, const float *d_coef,
                      float *d_partial, float *d_dw, int H, int W, int accumulate, void *stream)
{
    if (!d_act || !d_g || !d_z || !d_coef || !d_partial || !d_dw || H < 1 || W < 1) {
        set_error("f2f_dncnn_wgrad64: bad arguments");
        return F2F_EINVAL;
    }
    static bool configured = false;
    if (!configured) {
        F2F_CUDA_CHECK(cudaFuncSetAttribute(wg::k_wgrad64, cudaFuncAttributeMaxDynamicSharedMemorySize, wg::SMEM_BYTES));
        configured = true;
    }
    cudaStream_t st = (cudaStream_t)stream;
    wg::Args a;
    a.act = (const bf16 *)d_act; a.coef_in = d_coef_in; a.g = (bf16 *)d_g; a.z = (const bf16 *)d_z; a.coef = d_coef;
    a.partial = d_partial; a.H = H; a.W = W; a.tiles_x = cdiv(W, 16); a.tiles_y = cdiv(H, wg::TR);
    const int ntiles = a.tiles_x * a.tiles_y;
    const int grid = ntiles < num_sms() ? ntiles : num_sms();
    prof_begin(1, st);
    wg::k_wgrad64<<<grid, wg::NTHREADS, wg::SMEM_BYTES, st>>>(a);
    prof_end(1, st);
    wg::k_wgrad_reduce<<<cdiv(9 * 4096, 256), 256, 0, st>>>(d_partial, grid, d_dw, accumulate);
    g_dncnn_launches += 2;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

int f2f_dncnn_backward(const float *d_params, float *d_grads, const float *d_dy, void *d_ws, size_t ws_bytes, int B, int H,
                       int W, int residual, int accumulate, void *stream)
{
    if (!d_params || !d_grads || !d_dy || !d_ws || B < 1 || H < 1 || W < 1) {
        set_error("f2f_dncnn_backward: bad arguments");
        return F2F_EINVAL;
    }
    const Plan pl = make_plan(B, H, W, 1);
    if (ws_bytes < pl.total) {
        set_error("f2f_dncnn_backward: workspace too small (%zu bytes, need %zu)", ws_bytes, pl.total);
        return F2F_ENOMEM;
    }
    static bool configured = false;
    if (!configured) {
        F2F_CUDA_CHECK(cudaFuncSetAttribute(k_conv_last_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, LBW_SMEM));
        F2F_CUDA_CHECK(cudaFuncSetAttribute(k_wgrad_first, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 72 * 4));
        configured = true;
    }
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t *ws = (uint8_t *)d_ws;
    const size_t P = (size_t)H * W;
    const double n = (double)B * (double)P;
    const bf16 *wprep = (const bf16 *)(ws + pl.o_wprep);
    float *coef = (float *)(ws + pl.o_coef);
    double *stats_b = (double *)(ws + pl.o_stats_b);
    double *dw0 = (double *)(ws + pl.o_dw_small), *dw16 = dw0 + 576;
    const float *x = (const float *)(ws + pl.o_x);
    bf16 *a0 = (bf16 *)(ws + pl.o_a0);
    float *partial = (float *)(ws + pl.o_partial);
    auto zbuf = [&](int k) { return (bf16 *)(ws + pl.o_z + pl.act_bytes * (size_t)(k - 1)); };
    bf16 *gbuf[2] = {(bf16 *)(ws + pl.o_g0), (bf16 *)(ws + pl.o_g1)};

    // stats_b and the small double accumulators are adjacent (make_plan)
    F2F_CUDA_CHECK(cudaMemsetAsync(stats_b, 0, (pl.o_dw_small - pl.o_stats_b) + 1152 * 8, st));
    int cur = 0;   // gbuf[cur] holds g_k
    for (int b = 0; b < B; b++) {
        k_conv_last_bwd<<<num_sms(), 256, LBW_SMEM, st>>>(d_dy + b * P, residual ? -1.0f : 1.0f,
                                                                 zbuf(NBN) + b * P * 64, coef + (size_t)(NBN - 1) * C_STRIDE,
                                                                 d_params + OFF_W16, gbuf[cur] + b * P * 64,
                                                                 stats_b + (size_t)(NBN - 1) * 128, dw16, H, W);
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
        k_bn_bwd_coeffs<<<1, 64, 0, st>>>(stats_b + (size_t)(k - 1) * 128, d_params + off_gamma(k), cf,
                                          d_grads + off_gamma(k), d_grads + off_beta(k), n, accumulate);
        g_dncnn_launches++;
        const bf16 *act = k == 1 ? a0 : zbuf(k - 1);
        const float *cf_in = k == 1 ? nullptr : coef + (size_t)(k - 2) * C_STRIDE;
        for (int b = 0; b < B; b++) {
            int rc = f2f_dncnn_wgrad64(act + b * P * 64, cf_in, gbuf[cur] + b * P * 64, zbuf(k) + b * P * 64, cf, partial,
                                       d_grads + off_conv(k), H, W, accumulate || b > 0, stream);
            if (rc) return rc;
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
        k_wgrad_first<<<simt_blocks() / 2, 256, 64 * 72 * 4, st>>>(gbuf[cur] + b * P * 64, x + b * P, dw0, H, W);
        g_dncnn_launches++;
    }
    k_small_grads<<<cdiv(1152, 256), 256, 0, st>>>(dw0, dw16, d_grads, accumulate);
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
    if (!d_out || !d_dout || !d_loss) { set_error("f2f_step_train: bad arguments"); return F2F_EINVAL; }
    int rc = f2f_dncnn_forward(d_params, d_bn_running, d_nbt, d_curr, d_out, d_ws, ws_bytes, 1, H, W, 1, 1, stream);
    if (rc) return rc;
    rc = f2f_masked_loss(d_out, d_warped, d_mask, d_dout, d_loss, (size_t)H * W, loss_kind, stream);
    if (rc) return rc;
    rc = f2f_dncnn_backward(d_params, d_grads, d_dout, d_ws, ws_bytes, 1, H, W, 1, 0, stream);
    if (rc) return rc;
    return f2f_adam_step(d_params, d_grads, d_m, d_v, NPARAMS, step, lr, beta1, beta2, eps, weight_decay, stream);
}

}  // extern "C"
