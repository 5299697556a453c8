// dncnn_precise2.cuh -- the SPLIT-OPERAND precise mode of the DnCNN path (f2f_dncnn_set_precise(2)): FP32-class
// arithmetic ON THE TENSOR CORES.  Every FP32 operand v of a 64->64 convolution, input gradient or weight gradient is
// written as v = hi + lo / 2048 with hi = fp16(v), lo = fp16((v - hi) * 2048) -- 22 significant bits -- and the product
// a * b is taken as  a_hi b_hi + (a_lo b_hi + a_hi b_lo) / 2048  (the dropped a_lo b_lo term is 2^-22 relative): three
// launches of the SAME tcgen05 kernels as the default mode (dnt::k_conv64t with an FP32-output epilogue,
// wg::k_wgrad64), FP16 x FP16 products exact in the FP32 accumulators.  Activations and gradients live in FP32 between
// layers (BatchNorm, ReLU masks and their sums are elementwise FP32 kernels here); only the tensor-core operands are
// 16-bit pairs.  About 5x the default mode's time instead of the ~40x of the CUDA-core mode (dncnn_precise.cuh), with
// the same parity against the FP32 oracle.
//
// Included by dncnn_net.cu (namespace f2f::net) after dncnn_precise.cuh.
#pragma once

namespace p2 {

constexpr float LO = 2048.0f, ILO = 1.0f / 2048.0f;

// two FP32 values -> packed 16-bit hi pair and lo pair
__device__ __forceinline__ void split2(float v0, float v1, uint32_t &hi, uint32_t &lo)
{
    hi = pack_h16(v0, v1);
    const float2 b = unpack_h16(hi);
    lo = pack_h16((v0 - b.x) * LO, (v1 - b.y) * LO);
}
__device__ __forceinline__ void split8(const float (&v)[8], uint4 &hi, uint4 &lo)
{
    split2(v[0], v[1], hi.x, lo.x); split2(v[2], v[3], hi.y, lo.y);
    split2(v[4], v[5], hi.z, lo.z); split2(v[6], v[7], hi.w, lo.w);
}
__device__ __forceinline__ void ld8(const float *p, float (&v)[8])
{
    const float4 a = __ldg(reinterpret_cast<const float4 *>(p)), b = __ldg(reinterpret_cast<const float4 *>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void st8(float *p, const float (&v)[8])
{
    reinterpret_cast<float4 *>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4 *>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}

// All elementwise kernels: a thread owns the 8-channel chunk (global thread id & 7) of pixels id >> 3, id >> 3 + stride, ...
// (the stride is a multiple of 8 threads, so the chunk -- and the per-channel coefficients in registers -- never changes).

// operand images of a layer input: a = in (raw) or relu(in * scale + shift) (BatchNorm + ReLU of the layer below)
__global__ void __launch_bounds__(256) k_split(const float *__restrict__ in, const float *__restrict__ coef, h16 *__restrict__ hi,
                                               h16 *__restrict__ lo, size_t npx)
{
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    const int c8 = (int)(tid & 7) * 8;
    float sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { sc[j] = coef ? coef[C_SCALE + c8 + j] : 1.0f; sh[j] = coef ? coef[C_SHIFT + c8 + j] : 0.0f; }
    for (size_t p = tid >> 3; p < npx; p += nthr >> 3) {
        float v[8];
        ld8(in + p * 64 + c8, v);
        if (coef) {
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = fmaxf(fmaf(v[j], sc[j], sh[j]), 0.0f);
        }
        uint4 h, l;
        split8(v, h, l);
        *reinterpret_cast<uint4 *>(hi + p * 64 + c8) = h;
        *reinterpret_cast<uint4 *>(lo + p * 64 + c8) = l;
    }
}

// per-thread channel sums -> one FP64 atomic per value and block (threads t, t + 8, ... of a block share their chunk)
__device__ __forceinline__ void block_chunk_reduce(const float (&s0)[8], const float (&s1)[8], double *stats)
{
    __shared__ float red[2][256][8];
    const int t = threadIdx.x;
#pragma unroll
    for (int j = 0; j < 8; j++) { red[0][t][j] = s0[j]; red[1][t][j] = s1[j]; }
    __syncthreads();
    if (t < 128) {
        const int which = t >> 6, c = t & 63, chunk = c >> 3, j = c & 7;
        double tot = 0.0;
        for (int k = chunk; k < 256; k += 8) tot += (double)red[which][k][j];
        atomicAdd(stats + which * 64 + c, tot);
    }
}

// train-mode BatchNorm batch sums of a layer's FP32 output: stats[0..63] += sum z, stats[64..127] += sum z^2
__global__ void __launch_bounds__(256) k_stats32(const float *__restrict__ z, double *__restrict__ stats, size_t npx)
{
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    const int c8 = (int)(tid & 7) * 8;
    float s0[8], s1[8];
#pragma unroll
    for (int j = 0; j < 8; j++) s0[j] = s1[j] = 0.0f;
    for (size_t p = tid >> 3; p < npx; p += nthr >> 3) {
        float v[8];
        ld8(z + p * 64 + c8, v);
#pragma unroll
        for (int j = 0; j < 8; j++) { s0[j] += v[j]; s1[j] = fmaf(v[j], v[j], s1[j]); }
    }
    block_chunk_reduce(s0, s1, stats);
}

// BatchNorm backward of a layer: dz = A * g + B * z + C (k_bn_bwd_coeffs) -> operand images of dz
__global__ void __launch_bounds__(256) k_dz_split(const float *__restrict__ g, const float *__restrict__ z, const float *__restrict__ coef,
                                                  h16 *__restrict__ hi, h16 *__restrict__ lo, size_t npx)
{
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    const int c8 = (int)(tid & 7) * 8;
    float cA[8], cB[8], cC[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { cA[j] = coef[C_A + c8 + j]; cB[j] = coef[C_B + c8 + j]; cC[j] = coef[C_C + c8 + j]; }
    for (size_t p = tid >> 3; p < npx; p += nthr >> 3) {
        float gv[8], zv[8], v[8];
        ld8(g + p * 64 + c8, gv);
        ld8(z + p * 64 + c8, zv);
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = fmaf(cA[j], gv[j], fmaf(cB[j], zv[j], cC[j]));
        uint4 h, l;
        split8(v, h, l);
        *reinterpret_cast<uint4 *>(hi + p * 64 + c8) = h;
        *reinterpret_cast<uint4 *>(lo + p * 64 + c8) = l;
    }
}

// ReLU mask of the layer below on an input gradient, and that layer's BatchNorm-backward sums (sum g, sum g * xhat).
// coef_p != NULL: zprev is the layer's pre-BN output, mask = z * scale + shift > 0; NULL: zprev is a post-ReLU activation.
__global__ void __launch_bounds__(256) k_mask_sums(const float *__restrict__ gin, const float *__restrict__ zprev,
                                                   const float *__restrict__ coef_p, float *__restrict__ gout,
                                                   double *__restrict__ stats, size_t npx)
{
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    const int c8 = (int)(tid & 7) * 8;
    float sc[8], sh[8], mu[8], is[8], s0[8], s1[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        sc[j] = coef_p ? coef_p[C_SCALE + c8 + j] : 1.0f; sh[j] = coef_p ? coef_p[C_SHIFT + c8 + j] : 0.0f;
        mu[j] = coef_p ? coef_p[C_MEAN + c8 + j] : 0.0f; is[j] = coef_p ? coef_p[C_INVSTD + c8 + j] : 1.0f;
        s0[j] = s1[j] = 0.0f;
    }
    for (size_t p = tid >> 3; p < npx; p += nthr >> 3) {
        float gv[8], zv[8];
        ld8(gin + p * 64 + c8, gv);
        ld8(zprev + p * 64 + c8, zv);
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (!(fmaf(zv[j], sc[j], sh[j]) > 0.0f)) gv[j] = 0.0f;
            s0[j] += gv[j];
            s1[j] = fmaf(gv[j], (zv[j] - mu[j]) * is[j], s1[j]);
        }
        st8(gout + p * 64 + c8, gv);
    }
    if (stats) block_chunk_reduce(s0, s1, stats);
}

// hi / lo operand images of the 30 weight tensors in the tensor-core layout of k_prep_all
__global__ void k_prep_split(const float *__restrict__ params, h16 *__restrict__ whi, h16 *__restrict__ wlo)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 30 * WSZ) return;
    const int which = idx / WSZ, e = idx % WSZ;
    const int transpose = which >= NBN, k = (which % NBN) + 1;
    const float *w = params + off_conv(k);
    const int kk = e & 7, n = (e >> 3) & 63, kc = (e >> 9) & 7, tap = e >> 12;
    const int kch = kc * 8 + kk;
    const float v = !transpose ? w[((size_t)n * 64 + kch) * 9 + tap] : w[((size_t)kch * 64 + n) * 9 + (8 - tap)];
    const h16 h = tc05::float_to_h16(v);
    whi[idx] = h;
    wlo[idx] = tc05::float_to_h16((v - unpack_h16((uint32_t)(*reinterpret_cast<const unsigned short *>(&h))).x) * LO);
}

__global__ void k_identity_coef(float *coef)      // A = 1, B = C = 0: the weight-gradient kernel then takes dz as it is
{
    const int c = threadIdx.x;
    if (c < 64) { coef[C_A + c] = 1.0f; coef[C_B + c] = 0.0f; coef[C_C + c] = 0.0f; }
}

// ---------------------------------------------------------------- workspace plan and drivers
struct Plan2 {
    size_t act32, act16;         // bytes of one FP32 / one 16-bit 64-channel image of the batch
    size_t o_x, o_a0, o_z, o_ahi, o_alo, o_g0, o_g1, o_dzhi, o_dzlo, o_whi, o_wlo, o_stats_f, o_stats_b, o_dw_small, o_absmax,
        o_coef, o_partial, total;
    int nz, na;
};

static Plan2 make_plan2(int B, int H, int W, int train, int C = 1)
{
    Plan2 p;
    const size_t P = (size_t)H * W;
    p.act32 = align_up((size_t)B * P * 256, 1024);
    p.act16 = align_up((size_t)B * P * 128, 1024);
    size_t o = 0;
    auto take = [&](size_t bytes) { const size_t r = o; o += align_up(bytes, 1024); return r; };
    p.o_x = take((size_t)B * P * 4 * C);
    p.o_a0 = take(p.act32);
    p.nz = train ? NBN : 2;
    p.o_z = take(p.act32 * p.nz);
    p.na = train ? NBN : 1;                        // operand images of a_0 .. a_14 (kept for the weight gradients)
    p.o_ahi = take(p.act16 * p.na);
    p.o_alo = take(p.act16 * p.na);
    p.o_g0 = train ? take(p.act32) : 0;
    p.o_g1 = train ? take(p.act32) : 0;
    p.o_dzhi = train ? take(p.act16) : 0;
    p.o_dzlo = train ? take(p.act16) : 0;
    p.o_whi = take((size_t)30 * WSZ * 2);
    p.o_wlo = take((size_t)30 * WSZ * 2);
    p.o_stats_f = take((size_t)NBN * 128 * 8);
    p.o_stats_b = take((size_t)NBN * 128 * 8);     // contiguous with o_dw_small and o_absmax: one memset
    p.o_dw_small = take((size_t)1152 * C * 8);
    p.o_absmax = take(16);
    p.o_coef = take((size_t)(NBN + 1) * C_STRIDE * 4);      // + one identity block for the weight-gradient kernel
    p.o_partial = train ? take((size_t)num_sms() * wg::PARTIAL * 4) : 0;
    p.total = o;
    return p;
}

static int ew_grid() { return num_sms() * 8; }

// out32 = conv3x3(a, W) from the operand images: a_hi W_hi + (a_lo W_hi + a_hi W_lo) / 2048
static int conv_split(const h16 *ahi, const h16 *alo, const h16 *whi, const h16 *wlo, float *out32, int H, int W, cudaStream_t st)
{
    int rc = conv64_f32(ahi, whi, out32, 1.0f, 0, H, W, st);
    if (!rc) rc = conv64_f32(alo, whi, out32, ILO, 1, H, W, st);
    if (!rc) rc = conv64_f32(ahi, wlo, out32, ILO, 1, H, W, st);
    return rc;
}

static int forward(const float *d_params, float *d_bn_running, long long *d_nbt, const float *d_x, float *d_y, void *d_ws,
                   size_t ws_bytes, int B, int H, int W, int train, int residual, cudaStream_t st, int C)
{
    const Plan2 pl = make_plan2(B, H, W, train, C);
    if (ws_bytes < pl.total) {
        set_error("f2f_dncnn_forward (precise 2): workspace too small (%zu bytes, need %zu)", ws_bytes, pl.total);
        return F2F_ENOMEM;
    }
    const float *d_w0 = d_params;
    d_params += (size_t)576 * (C - 1);
    uint8_t *ws = (uint8_t *)d_ws;
    const size_t P = (size_t)H * W;
    h16 *whi = (h16 *)(ws + pl.o_whi), *wlo = (h16 *)(ws + pl.o_wlo);
    float *coef = (float *)(ws + pl.o_coef), *a0 = (float *)(ws + pl.o_a0);
    double *stats_f = (double *)(ws + pl.o_stats_f);
    auto zbuf = [&](int k) { return (float *)(ws + pl.o_z + pl.act32 * (size_t)(train ? k - 1 : (k & 1))); };
    auto ahi = [&](int k) { return (h16 *)(ws + pl.o_ahi + pl.act16 * (size_t)(train ? k : 0)); };     // images of a_k
    auto alo = [&](int k) { return (h16 *)(ws + pl.o_alo + pl.act16 * (size_t)(train ? k : 0)); };
    k_prep_split<<<cdiv(30 * WSZ, 256), 256, 0, st>>>(d_params, whi, wlo);
    g_dncnn_launches++;
    if (train) {
        F2F_CUDA_CHECK(cudaMemsetAsync(stats_f, 0, (size_t)NBN * 128 * 8, st));
        F2F_CUDA_CHECK(cudaMemcpyAsync(ws + pl.o_x, d_x, (size_t)B * P * 4 * C, cudaMemcpyDeviceToDevice, st));
    } else {
        k_bn_eval_coeffs<<<NBN, 64, 0, st>>>(d_params, d_bn_running, coef);
        g_dncnn_launches++;
    }
    for (int b = 0; b < B; b++) {
        const int rc = gen::first_fwd<float>(C, d_x + b * P * C, d_w0, a0 + b * P * 64, H, W, st);
        if (rc) return rc;
    }
    for (int k = 1; k <= NBN; k++) {
        // operand images of this layer's input a_{k-1}: BatchNorm + ReLU of z_{k-1} (a_0 comes post-ReLU from the first layer)
        k_split<<<ew_grid(), 256, 0, st>>>(k == 1 ? a0 : zbuf(k - 1), k == 1 ? nullptr : coef + (size_t)(k - 2) * C_STRIDE,
                                           ahi(k - 1), alo(k - 1), (size_t)B * P);
        g_dncnn_launches++;
        for (int b = 0; b < B; b++) {
            const int rc = conv_split(ahi(k - 1) + b * P * 64, alo(k - 1) + b * P * 64, whi + (size_t)(k - 1) * WSZ,
                                      wlo + (size_t)(k - 1) * WSZ, zbuf(k) + b * P * 64, H, W, st);
            if (rc) return rc;
        }
        if (train) {
            k_stats32<<<ew_grid(), 256, 0, st>>>(zbuf(k), stats_f + (size_t)(k - 1) * 128, (size_t)B * P);
            k_bn_finalize<<<1, 64, 0, st>>>(stats_f + (size_t)(k - 1) * 128, d_params + off_gamma(k), d_params + off_beta(k),
                                            d_bn_running + (size_t)(k - 1) * 128, d_nbt ? d_nbt + (k - 1) : nullptr,
                                            coef + (size_t)(k - 1) * C_STRIDE, (double)B * (double)P, 0);
            g_dncnn_launches += 2;
        }
    }
    F2F_LAUNCH_CHECK();
    for (int b = 0; b < B; b++) {
        const int rc = gen::last_fwd<float>(C, zbuf(NBN) + b * P * 64, coef + (size_t)(NBN - 1) * C_STRIDE, d_params + OFF_W16,
                                            d_x + b * P * C, d_y + b * P * C, H, W, residual, st);
        if (rc) return rc;
    }
    return F2F_OK;
}

static int backward(const float *d_params, float *d_grads, const float *d_dy, void *d_ws, size_t ws_bytes, int B, int H, int W,
                    int residual, int accumulate, cudaStream_t st, int C)
{
    const Plan2 pl = make_plan2(B, H, W, 1, C);
    if (ws_bytes < pl.total) {
        set_error("f2f_dncnn_backward (precise 2): workspace too small (%zu bytes, need %zu)", ws_bytes, pl.total);
        return F2F_ENOMEM;
    }
    float *d_gw0 = d_grads;
    d_params += (size_t)576 * (C - 1);
    d_grads += (size_t)576 * (C - 1);
    uint8_t *ws = (uint8_t *)d_ws;
    const size_t P = (size_t)H * W;
    const double n = (double)B * (double)P;
    const h16 *whi = (const h16 *)(ws + pl.o_whi), *wlo = (const h16 *)(ws + pl.o_wlo);
    float *coef = (float *)(ws + pl.o_coef), *ident = coef + (size_t)NBN * C_STRIDE;
    double *stats_b = (double *)(ws + pl.o_stats_b);
    double *dw0 = (double *)(ws + pl.o_dw_small), *dw16 = dw0 + 576 * C;
    unsigned *absmax = (unsigned *)(ws + pl.o_absmax);
    const float *x = (const float *)(ws + pl.o_x);
    float *a0 = (float *)(ws + pl.o_a0), *partial = (float *)(ws + pl.o_partial);
    h16 *dzhi = (h16 *)(ws + pl.o_dzhi), *dzlo = (h16 *)(ws + pl.o_dzlo);
    auto zbuf = [&](int k) { return (float *)(ws + pl.o_z + pl.act32 * (size_t)(k - 1)); };
    auto ahi = [&](int k) { return (h16 *)(ws + pl.o_ahi + pl.act16 * (size_t)k); };
    auto alo = [&](int k) { return (h16 *)(ws + pl.o_alo + pl.act16 * (size_t)k); };
    float *gbuf[2] = {(float *)(ws + pl.o_g0), (float *)(ws + pl.o_g1)};
    F2F_CUDA_CHECK(cudaMemsetAsync(stats_b, 0, (pl.o_absmax - pl.o_stats_b) + 16, st));
    // the incoming gradient is brought into [1, 2) by a power of two (grad_scale) so that dz's images sit in FP16's range
    k_absmax<<<num_sms(), 256, 0, st>>>(d_dy, (size_t)B * P * C, absmax);
    k_identity_coef<<<1, 64, 0, st>>>(ident);
    g_dncnn_launches += 2;
    int cur = 0;
    for (int b = 0; b < B; b++) {
        const int rc = gen::last_bwd<float>(C, d_dy + b * P * C, residual ? -1.0f : 1.0f, zbuf(NBN) + b * P * 64,
                                            coef + (size_t)(NBN - 1) * C_STRIDE, d_params + OFF_W16, gbuf[cur] + b * P * 64,
                                            stats_b + (size_t)(NBN - 1) * 128, dw16, H, W, absmax, st);
        if (rc) return rc;
    }
    for (int k = NBN; k >= 1; k--) {
        float *cf = coef + (size_t)(k - 1) * C_STRIDE;
        k_bn_bwd_coeffs<<<1, 64, 0, st>>>((const double *)(stats_b + (size_t)(k - 1) * 128), d_params + off_gamma(k), cf,
                                          d_grads + off_gamma(k), d_grads + off_beta(k), n, accumulate, 0, (const unsigned *)absmax);
        k_dz_split<<<ew_grid(), 256, 0, st>>>(gbuf[cur], zbuf(k), cf, dzhi, dzlo, (size_t)B * P);
        g_dncnn_launches += 2;
        const float *cf_in = k == 1 ? nullptr : coef + (size_t)(k - 2) * C_STRIDE;
        for (int b = 0; b < B; b++) {
            const size_t o = b * P * 64;
            // dW_k = sum_p dz (x) a_{k-1}: three split terms through the tcgen05 weight-gradient kernel (dz taken as it is)
            int rc = wgrad64_scaled(ahi(k - 1) + o, nullptr, dzhi + o, dzhi + o, ident, partial, d_grads + off_conv(k), H, W,
                                    accumulate || b > 0, absmax, st, 1.0f);
            if (!rc) rc = wgrad64_scaled(alo(k - 1) + o, nullptr, dzhi + o, dzhi + o, ident, partial, d_grads + off_conv(k), H, W, 1,
                                         absmax, st, ILO);
            if (!rc) rc = wgrad64_scaled(ahi(k - 1) + o, nullptr, dzlo + o, dzlo + o, ident, partial, d_grads + off_conv(k), H, W, 1,
                                         absmax, st, ILO);
            // input gradient: conv3x3(dz, W^T flipped) -> FP32
            if (!rc) rc = conv_split(dzhi + o, dzlo + o, whi + (size_t)(NBN + k - 1) * WSZ, wlo + (size_t)(NBN + k - 1) * WSZ,
                                     gbuf[cur ^ 1] + o, H, W, st);
            if (rc) return rc;
        }
        // ReLU mask of the layer below and its BatchNorm-backward sums, in place
        k_mask_sums<<<ew_grid(), 256, 0, st>>>(gbuf[cur ^ 1], k >= 2 ? zbuf(k - 1) : a0, cf_in, gbuf[cur ^ 1],
                                               k >= 2 ? stats_b + (size_t)(k - 2) * 128 : nullptr, (size_t)B * P);
        g_dncnn_launches++;
        cur ^= 1;
    }
    F2F_LAUNCH_CHECK();
    for (int b = 0; b < B; b++) {
        const int rc = gen::first_wgrad<float>(C, gbuf[cur] + b * P * 64, x + b * P * C, dw0, H, W, st);
        if (rc) return rc;
    }
    k_small_grads<<<cdiv(1152 * C, 256), 256, 0, st>>>(dw0, dw16, d_gw0, d_grads + OFF_W16, 576 * C, accumulate, (const unsigned *)absmax);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

}  // namespace p2
