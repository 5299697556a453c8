// dncnn_precise.cuh -- the FP32 ("precise") mode of the DnCNN path: the same network, data flow and fusions as the
// tensor-core path of dncnn.cu / dncnn_net.cu (lib/frame2frame/dncnn/net.py:6-39 as trained by blind_denoising.py:
// 211-230), with every activation, gradient and operand kept in FP32 and every product accumulated in FP32 on the
// CUDA cores.  It exists for parity, not for speed: north_star asks for the reference's PSNR to 0.05 dB, the
// reference computes in FP32, and the 16-bit operands of the default mode are the one arithmetic difference between
// the two loops.  Running the loop in this mode separates "rounding of 16-bit operands" from "anything else": what is
// left against the FP32 CPU oracle is the loop's own sensitivity to summation order.  About 25x
// slower than the default mode (still ~30x the 16-core CPU reference); selected with f2f_dncnn_set_precise(1),
// F2F_PRECISE=1, `blind_denoising --precise`.
//
// Included by dncnn_net.cu after the small kernels it shares (BatchNorm bookkeeping, Adam, small-gradient scatter).
#pragma once

namespace pr {

constexpr int TR = 8, TC = 16;                       // output tile (rows x pixels) of the 64->64 kernels
constexpr int HR = TR + 2, HC = TC + 2;              // with the halo of a 3x3 filter
constexpr int PXS = 65;                              // floats between pixels of a staged tile (odd: conflict-free)
constexpr int CONV_SMEM = (9 * 64 * 64 + HR * HC * PXS) * 4;     // weights [tap][ic][oc] + haloed input tile
constexpr int CONV_THREADS = 128;                    // thread = 8 consecutive pixels of a row x 8 output channels

enum { PRO_RAW = 0, PRO_BN_RELU = 1 };
enum { EPI_STORE = 0, EPI_STATS = 1, EPI_DGRAD_BN = 2, EPI_DGRAD_RELU = 3 };

struct ConvF {
    const float *in;       // [H][W][64]
    float *out;            // [H][W][64]
    const float *w;        // [9][64 ic][64 oc]  (k_prep_f)
    const float *coef_in;  // PRO_BN_RELU: coefficient block of the layer below (scale, shift)
    double *stats;         // [2][64]
    const float *zprev;    // EPI_DGRAD_*: pre-activation (BN) / activation (RELU) of the layer below
    const float *coef_p;   // EPI_DGRAD_BN: coefficient block of the layer below
    int H, W, tiles_x, ntiles;
};

// FP32 operand images of the 15 forward and 15 transposed/flipped weight tensors: [which][tap][k][n]
//   forward  : n = oc, k = ic, w[oc][ic][tap]        input gradient: n = ic, k = oc, w[oc][ic][8 - tap]
__global__ void k_prep_f(const float *__restrict__ params, float *__restrict__ wprep)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 30 * WSZ) return;
    const int which = idx / WSZ, e = idx % WSZ;
    const int transpose = which >= NBN, k = (which % NBN) + 1;
    const float *w = params + off_conv(k);
    const int n = e & 63, kk = (e >> 6) & 63, tap = e >> 12;
    wprep[idx] = !transpose ? w[((size_t)n * 64 + kk) * 9 + tap] : w[((size_t)kk * 64 + n) * 9 + (8 - tap)];
}

// 3x3 64->64 convolution on FP32 NHWC, persistent over 8x16-pixel tiles; the whole weight tensor stays in shared memory.
template <int PRO, int EPI>
__global__ void __launch_bounds__(CONV_THREADS, 1) k_conv64f(ConvF a)
{
    extern __shared__ __align__(16) float smf[];
    float *sW = smf, *sIn = smf + 9 * 64 * 64;
    const int t = threadIdx.x, ocg = t & 7, pg = t >> 3;          // 8 channel groups x 16 pixel groups
    const int prow = pg >> 1, pcol = (pg & 1) * 8;
    for (int i = t; i < 9 * 64 * 64 / 4; i += CONV_THREADS)
        reinterpret_cast<float4 *>(sW)[i] = __ldg(reinterpret_cast<const float4 *>(a.w) + i);
    float s0[8], s1[8];                                            // per-channel sums over everything this thread stores
#pragma unroll
    for (int j = 0; j < 8; j++) s0[j] = s1[j] = 0.0f;
    float psc[8], psh[8], pmu[8], pis[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int c = ocg * 8 + j;
        psc[j] = EPI == EPI_DGRAD_BN ? a.coef_p[C_SCALE + c] : 1.0f;
        psh[j] = EPI == EPI_DGRAD_BN ? a.coef_p[C_SHIFT + c] : 0.0f;
        pmu[j] = EPI == EPI_DGRAD_BN ? a.coef_p[C_MEAN + c] : 0.0f;
        pis[j] = EPI == EPI_DGRAD_BN ? a.coef_p[C_INVSTD + c] : 1.0f;
    }
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
        const int ty0 = (tile / a.tiles_x) * TR, tx0 = (tile % a.tiles_x) * TC;
        __syncthreads();                                           // previous tile's readers are done (and sW is loaded)
        // haloed tile, BN + ReLU of the layer below applied on the way (zero padding applies to the ACTIVATION)
        for (int i = t; i < HR * HC * 16; i += CONV_THREADS) {
            const int q = i & 15, pix = i >> 4;
            const int gy = ty0 - 1 + pix / HC, gx = tx0 - 1 + pix % HC;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gy >= 0 && gy < a.H && gx >= 0 && gx < a.W) {
                v = __ldg(reinterpret_cast<const float4 *>(a.in + ((size_t)gy * a.W + gx) * 64) + q);
                if (PRO == PRO_BN_RELU) {
                    const float4 sc = *reinterpret_cast<const float4 *>(a.coef_in + C_SCALE + q * 4);
                    const float4 sh = *reinterpret_cast<const float4 *>(a.coef_in + C_SHIFT + q * 4);
                    v.x = fmaxf(fmaf(v.x, sc.x, sh.x), 0.f); v.y = fmaxf(fmaf(v.y, sc.y, sh.y), 0.f);
                    v.z = fmaxf(fmaf(v.z, sc.z, sh.z), 0.f); v.w = fmaxf(fmaf(v.w, sc.w, sh.w), 0.f);
                }
            }
            float *d = sIn + pix * PXS + q * 4;
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        }
        __syncthreads();
        float acc[8][8];
#pragma unroll
        for (int p = 0; p < 8; p++)
#pragma unroll
            for (int j = 0; j < 8; j++) acc[p][j] = 0.0f;
#pragma unroll 1
        for (int tap = 0; tap < 9; tap++) {
            const float *ip = sIn + ((prow + tap / 3) * HC + pcol + tap % 3) * PXS;
            const float *wp = sW + tap * 4096 + ocg * 8;
#pragma unroll 4
            for (int ic = 0; ic < 64; ic++) {
                const float4 w0 = *reinterpret_cast<const float4 *>(wp + ic * 64);
                const float4 w1 = *reinterpret_cast<const float4 *>(wp + ic * 64 + 4);
                const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const float av = ip[p * PXS + ic];
#pragma unroll
                    for (int j = 0; j < 8; j++) acc[p][j] = fmaf(av, wv[j], acc[p][j]);
                }
            }
        }
        const int gy = ty0 + prow;
        if (gy < a.H) {
#pragma unroll
            for (int p = 0; p < 8; p++) {
                const int gx = tx0 + pcol + p;
                if (gx >= a.W) continue;
                const size_t o = ((size_t)gy * a.W + gx) * 64 + ocg * 8;
                float v[8];
#pragma unroll
                for (int j = 0; j < 8; j++) v[j] = acc[p][j];
                if (EPI == EPI_DGRAD_BN || EPI == EPI_DGRAD_RELU) {
                    const float4 z0 = __ldg(reinterpret_cast<const float4 *>(a.zprev + o));
                    const float4 z1 = __ldg(reinterpret_cast<const float4 *>(a.zprev + o + 4));
                    const float z[8] = {z0.x, z0.y, z0.z, z0.w, z1.x, z1.y, z1.z, z1.w};
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        if (EPI == EPI_DGRAD_BN) {                // ReLU mask of the layer below + its BatchNorm-backward sums
                            if (!(fmaf(z[j], psc[j], psh[j]) > 0.0f)) v[j] = 0.0f;
                            s0[j] += v[j];
                            s1[j] = fmaf(v[j], (z[j] - pmu[j]) * pis[j], s1[j]);
                        } else if (!(z[j] > 0.0f)) {
                            v[j] = 0.0f;
                        }
                    }
                }
                if (EPI == EPI_STATS) {
#pragma unroll
                    for (int j = 0; j < 8; j++) { s0[j] += v[j]; s1[j] = fmaf(v[j], v[j], s1[j]); }
                }
                *reinterpret_cast<float4 *>(a.out + o) = make_float4(v[0], v[1], v[2], v[3]);
                *reinterpret_cast<float4 *>(a.out + o + 4) = make_float4(v[4], v[5], v[6], v[7]);
            }
        }
    }
    if ((EPI == EPI_STATS || EPI == EPI_DGRAD_BN) && a.stats) {
        // threads t, t+8, t+16, ... hold the same 8 channels: combine in shared memory, one FP64 atomic per value and CTA
        __syncthreads();
        float *red = sIn;                                          // 16 x 128 floats
#pragma unroll
        for (int j = 0; j < 8; j++) { red[pg * 128 + ocg * 8 + j] = s0[j]; red[pg * 128 + 64 + ocg * 8 + j] = s1[j]; }
        __syncthreads();
        double tot = 0.0;
        for (int g = 0; g < 16; g++) tot += (double)red[g * 128 + t];
        atomicAdd(a.stats + t, tot);
    }
}

// ---------------------------------------------------------------- weight gradient
// dW[oc][ic][ky][kx] = sum_p dz[p][oc] * act[p + (ky-1, kx-1)][ic].  blockIdx.y = ky; a thread keeps 3 (kx) x 4 (ic) x 4 (oc)
// sums over all tiles of its persistent CTA; dz = A*g + B*z + C (BatchNorm backward) is formed while the tile is staged
// and, by the ky = 0 CTAs, written to `dz_out` for the input-gradient convolution that follows.
constexpr int WG_THREADS = 256;
constexpr int WG_ACT = TR * HC * 68, WG_DZ = TR * TC * 64;        // floats: activation rows (pixel stride 68), dz tile
constexpr int WG_SMEM = (WG_ACT + WG_DZ) * 4;
constexpr int WG_PARTIAL = 3 * 64 * 64;                            // floats per CTA

struct WgF {
    const float *act, *coef_in;   // layer input: z_{k-1} (+ BN/ReLU coefficients) or a_0 (coef_in = NULL)
    const float *g, *z, *coef;    // masked gradient g_k, pre-BN output z_k, A/B/C of layer k
    float *dz_out;                // dz_k
    float *partial;               // [gridDim.y][gridDim.x][WG_PARTIAL]
    int H, W, tiles_x, ntiles;
};

__global__ void __launch_bounds__(WG_THREADS) k_wgrad64f(WgF a)
{
    extern __shared__ __align__(16) float smf[];
    float *sA = smf, *sD = smf + WG_ACT;
    const int t = threadIdx.x, ky = blockIdx.y;
    const int ic4 = (t & 15) * 4, oc4 = (t >> 4) * 4;
    float acc[3][4][4];
#pragma unroll
    for (int x = 0; x < 3; x++)
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int o = 0; o < 4; o++) acc[x][i][o] = 0.0f;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
        const int ty0 = (tile / a.tiles_x) * TR, tx0 = (tile % a.tiles_x) * TC;
        __syncthreads();
        // activation rows ty0 + ky - 1 .. + TR - 1, columns tx0 - 1 .. tx0 + TC
        for (int i = t; i < TR * HC * 16; i += WG_THREADS) {
            const int q = i & 15, pix = i >> 4;
            const int gy = ty0 + ky - 1 + pix / HC, gx = tx0 - 1 + pix % HC;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gy >= 0 && gy < a.H && gx >= 0 && gx < a.W) {
                v = __ldg(reinterpret_cast<const float4 *>(a.act + ((size_t)gy * a.W + gx) * 64) + q);
                if (a.coef_in) {
                    const float4 sc = *reinterpret_cast<const float4 *>(a.coef_in + C_SCALE + q * 4);
                    const float4 sh = *reinterpret_cast<const float4 *>(a.coef_in + C_SHIFT + q * 4);
                    v.x = fmaxf(fmaf(v.x, sc.x, sh.x), 0.f); v.y = fmaxf(fmaf(v.y, sc.y, sh.y), 0.f);
                    v.z = fmaxf(fmaf(v.z, sc.z, sh.z), 0.f); v.w = fmaxf(fmaf(v.w, sc.w, sh.w), 0.f);
                }
            }
            *reinterpret_cast<float4 *>(sA + pix * 68 + q * 4) = v;
        }
        for (int i = t; i < TR * TC * 16; i += WG_THREADS) {
            const int q = i & 15, pix = i >> 4;
            const int gy = ty0 + pix / TC, gx = tx0 + pix % TC;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gy < a.H && gx < a.W) {
                const size_t o = ((size_t)gy * a.W + gx) * 64;
                const float4 g = __ldg(reinterpret_cast<const float4 *>(a.g + o) + q);
                const float4 z = __ldg(reinterpret_cast<const float4 *>(a.z + o) + q);
                const float4 A = *reinterpret_cast<const float4 *>(a.coef + C_A + q * 4);
                const float4 B = *reinterpret_cast<const float4 *>(a.coef + C_B + q * 4);
                const float4 C = *reinterpret_cast<const float4 *>(a.coef + C_C + q * 4);
                v.x = fmaf(A.x, g.x, fmaf(B.x, z.x, C.x)); v.y = fmaf(A.y, g.y, fmaf(B.y, z.y, C.y));
                v.z = fmaf(A.z, g.z, fmaf(B.z, z.z, C.z)); v.w = fmaf(A.w, g.w, fmaf(B.w, z.w, C.w));
                if (ky == 0) reinterpret_cast<float4 *>(a.dz_out + o)[q] = v;
            }
            *reinterpret_cast<float4 *>(sD + pix * 64 + q * 4) = v;
        }
        __syncthreads();
#pragma unroll 1
        for (int r = 0; r < TR; r++) {
#pragma unroll 2
            for (int c = 0; c < TC; c++) {
                const float4 d = *reinterpret_cast<const float4 *>(sD + (r * TC + c) * 64 + oc4);
                const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                for (int x = 0; x < 3; x++) {
                    const float4 av = *reinterpret_cast<const float4 *>(sA + (r * HC + c + x) * 68 + ic4);
                    const float aa[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
                    for (int i = 0; i < 4; i++)
#pragma unroll
                        for (int o = 0; o < 4; o++) acc[x][i][o] = fmaf(aa[i], dv[o], acc[x][i][o]);
                }
            }
        }
    }
    float *dst = a.partial + ((size_t)ky * gridDim.x + blockIdx.x) * WG_PARTIAL;
#pragma unroll
    for (int x = 0; x < 3; x++)
#pragma unroll
        for (int i = 0; i < 4; i++)
            *reinterpret_cast<float4 *>(dst + (x * 64 + ic4 + i) * 64 + oc4) =
                make_float4(acc[x][i][0], acc[x][i][1], acc[x][i][2], acc[x][i][3]);
}

// partial[ky][cta][kx][ic][oc] summed over CTAs in a fixed order (deterministic) -> dW[oc][ic][ky][kx]
__global__ void k_wgrad_reduce_f(const float *__restrict__ partial, int nparts, float *__restrict__ dw, int accumulate)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 9 * 4096) return;
    const int oc = idx & 63, ic = (idx >> 6) & 63, tap = idx >> 12, ky = tap / 3, kx = tap % 3;
    const float *src = partial + (size_t)ky * nparts * WG_PARTIAL + (kx * 64 + ic) * 64 + oc;
    float s = 0.0f;
    for (int c = 0; c < nparts; c++) s += src[(size_t)c * WG_PARTIAL];
    float *d = dw + ((size_t)oc * 64 + ic) * 9 + tap;
    *d = (accumulate ? *d : 0.0f) + s;
}

// ---------------------------------------------------------------- workspace plan and drivers
struct PlanF {
    size_t act_bytes;
    size_t o_x, o_a0, o_z, o_g0, o_g1, o_dz, o_wprep, o_stats_f, o_stats_b, o_dw_small, o_coef, o_partial, total;
    int nz;
};

static PlanF make_plan_f(int B, int H, int W, int train, int C = 1)
{
    PlanF p;
    const size_t P = (size_t)H * W;
    p.act_bytes = align_up((size_t)B * P * 256, 1024);
    size_t o = 0;
    auto take = [&](size_t bytes) { const size_t r = o; o += align_up(bytes, 1024); return r; };
    p.o_x = take((size_t)B * P * 4 * C);
    p.o_a0 = take(p.act_bytes);
    p.nz = train ? NBN : 2;
    p.o_z = take(p.act_bytes * p.nz);
    p.o_g0 = train ? take(p.act_bytes) : 0;
    p.o_g1 = train ? take(p.act_bytes) : 0;
    p.o_dz = train ? take(p.act_bytes) : 0;
    p.o_wprep = take((size_t)30 * WSZ * 4);
    p.o_stats_f = take((size_t)NBN * 128 * 8);
    p.o_stats_b = take((size_t)NBN * 128 * 8);     // contiguous with o_dw_small: one memset
    p.o_dw_small = take((size_t)1152 * C * 8);
    p.o_coef = take((size_t)NBN * C_STRIDE * 4);
    p.o_partial = train ? take((size_t)3 * num_sms() * WG_PARTIAL * 4) : 0;
    p.total = o;
    return p;
}

template <int PRO, int EPI>
static int launch_conv(ConvF a, cudaStream_t st)
{
    F2F_CUDA_CHECK((ensure_dyn_smem<k_conv64f<PRO, EPI>>(CONV_SMEM)));
    a.tiles_x = cdiv(a.W, TC);
    a.ntiles = a.tiles_x * cdiv(a.H, TR);
    const int grid = a.ntiles < num_sms() ? a.ntiles : num_sms();
    k_conv64f<PRO, EPI><<<grid, CONV_THREADS, CONV_SMEM, st>>>(a);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

static int forward(const float *d_params, float *d_bn_running, long long *d_nbt, const float *d_x, float *d_y, void *d_ws,
                   size_t ws_bytes, int B, int H, int W, int train, int residual, cudaStream_t st, int C)
{
    const PlanF pl = make_plan_f(B, H, W, train, C);
    const float *d_w0 = d_params;
    d_params += (size_t)576 * (C - 1);              // off_conv / off_gamma / off_beta / OFF_W16 are written for C = 1
    if (ws_bytes < pl.total) {
        set_error("f2f_dncnn_forward (precise): workspace too small (%zu bytes, need %zu)", ws_bytes, pl.total);
        return F2F_ENOMEM;
    }
    uint8_t *ws = (uint8_t *)d_ws;
    const size_t P = (size_t)H * W;
    float *wprep = (float *)(ws + pl.o_wprep), *coef = (float *)(ws + pl.o_coef), *a0 = (float *)(ws + pl.o_a0);
    double *stats_f = (double *)(ws + pl.o_stats_f);
    auto zbuf = [&](int k) { return (float *)(ws + pl.o_z + pl.act_bytes * (size_t)(train ? k - 1 : (k & 1))); };
    k_prep_f<<<cdiv(30 * WSZ, 256), 256, 0, st>>>(d_params, wprep);
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
    F2F_LAUNCH_CHECK();
    for (int k = 1; k <= NBN; k++) {
        for (int b = 0; b < B; b++) {
            ConvF a = {};
            a.in = (k == 1 ? a0 : zbuf(k - 1)) + b * P * 64;
            a.out = zbuf(k) + b * P * 64;
            a.w = wprep + (size_t)(k - 1) * WSZ;
            a.coef_in = k == 1 ? nullptr : coef + (size_t)(k - 2) * C_STRIDE;
            a.stats = train ? stats_f + (size_t)(k - 1) * 128 : nullptr;
            a.H = H; a.W = W;
            int rc;
            if (k == 1) rc = train ? launch_conv<PRO_RAW, EPI_STATS>(a, st) : launch_conv<PRO_RAW, EPI_STORE>(a, st);
            else rc = train ? launch_conv<PRO_BN_RELU, EPI_STATS>(a, st) : launch_conv<PRO_BN_RELU, EPI_STORE>(a, st);
            if (rc) return rc;
        }
        if (train) {
            k_bn_finalize<<<1, 64, 0, st>>>(stats_f + (size_t)(k - 1) * 128, d_params + off_gamma(k), d_params + off_beta(k),
                                            d_bn_running + (size_t)(k - 1) * 128, d_nbt ? d_nbt + (k - 1) : nullptr,
                                            coef + (size_t)(k - 1) * C_STRIDE, (double)B * (double)P, 0);
            g_dncnn_launches++;
        }
    }
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
    const PlanF pl = make_plan_f(B, H, W, 1, C);
    float *d_gw0 = d_grads;
    d_params += (size_t)576 * (C - 1);
    d_grads += (size_t)576 * (C - 1);
    if (ws_bytes < pl.total) {
        set_error("f2f_dncnn_backward (precise): workspace too small (%zu bytes, need %zu)", ws_bytes, pl.total);
        return F2F_ENOMEM;
    }
    F2F_CUDA_CHECK((ensure_dyn_smem<k_wgrad64f>(WG_SMEM)));
    uint8_t *ws = (uint8_t *)d_ws;
    const size_t P = (size_t)H * W;
    const double n = (double)B * (double)P;
    const float *wprep = (const float *)(ws + pl.o_wprep);
    float *coef = (float *)(ws + pl.o_coef);
    double *stats_b = (double *)(ws + pl.o_stats_b);
    double *dw0 = (double *)(ws + pl.o_dw_small), *dw16 = dw0 + 576 * C;
    const float *x = (const float *)(ws + pl.o_x);
    float *a0 = (float *)(ws + pl.o_a0), *partial = (float *)(ws + pl.o_partial), *dz = (float *)(ws + pl.o_dz);
    auto zbuf = [&](int k) { return (float *)(ws + pl.o_z + pl.act_bytes * (size_t)(k - 1)); };
    float *gbuf[2] = {(float *)(ws + pl.o_g0), (float *)(ws + pl.o_g1)};
    F2F_CUDA_CHECK(cudaMemsetAsync(stats_b, 0, (pl.o_dw_small - pl.o_stats_b) + (size_t)1152 * C * 8, st));
    int cur = 0;
    for (int b = 0; b < B; b++) {
        const int rc = gen::last_bwd<float>(C, d_dy + b * P * C, residual ? -1.0f : 1.0f, zbuf(NBN) + b * P * 64,
                                            coef + (size_t)(NBN - 1) * C_STRIDE, d_params + OFF_W16, gbuf[cur] + b * P * 64,
                                            stats_b + (size_t)(NBN - 1) * 128, dw16, H, W, nullptr, st);
        if (rc) return rc;
    }
    const int tiles_x = cdiv(W, TC), ntiles = tiles_x * cdiv(H, TR);
    const int wg_grid = ntiles < num_sms() ? ntiles : num_sms();
    for (int k = NBN; k >= 1; k--) {
        float *cf = coef + (size_t)(k - 1) * C_STRIDE;
        k_bn_bwd_coeffs<<<1, 64, 0, st>>>((const double *)(stats_b + (size_t)(k - 1) * 128), d_params + off_gamma(k), cf,
                                          d_grads + off_gamma(k), d_grads + off_beta(k), n, accumulate, 0, (const unsigned *)nullptr);
        g_dncnn_launches++;
        const float *cf_in = k == 1 ? nullptr : coef + (size_t)(k - 2) * C_STRIDE;
        for (int b = 0; b < B; b++) {
            WgF wa;
            wa.act = (k == 1 ? a0 : zbuf(k - 1)) + b * P * 64; wa.coef_in = cf_in;
            wa.g = gbuf[cur] + b * P * 64; wa.z = zbuf(k) + b * P * 64; wa.coef = cf; wa.dz_out = dz + b * P * 64;
            wa.partial = partial; wa.H = H; wa.W = W; wa.tiles_x = tiles_x; wa.ntiles = ntiles;
            k_wgrad64f<<<dim3(wg_grid, 3), WG_THREADS, WG_SMEM, st>>>(wa);
            k_wgrad_reduce_f<<<cdiv(9 * 4096, 256), 256, 0, st>>>(partial, wg_grid, d_grads + off_conv(k), accumulate || b > 0);
            g_dncnn_launches += 2;
            F2F_LAUNCH_CHECK();
            ConvF a = {};
            a.in = dz + b * P * 64; a.out = gbuf[cur ^ 1] + b * P * 64; a.w = wprep + (size_t)(NBN + k - 1) * WSZ;
            a.zprev = (k >= 2 ? zbuf(k - 1) : a0) + b * P * 64; a.coef_p = cf_in;
            a.stats = k >= 2 ? stats_b + (size_t)(k - 2) * 128 : nullptr;
            a.H = H; a.W = W;
            const int rc = k >= 2 ? launch_conv<PRO_RAW, EPI_DGRAD_BN>(a, st) : launch_conv<PRO_RAW, EPI_DGRAD_RELU>(a, st);
            if (rc) return rc;
        }
        cur ^= 1;
    }
    for (int b = 0; b < B; b++) {
        const int rc = gen::first_wgrad<float>(C, gbuf[cur] + b * P * 64, x + b * P * C, dw0, H, W, st);
        if (rc) return rc;
    }
    k_small_grads<<<cdiv(1152 * C, 256), 256, 0, st>>>(dw0, dw16, d_gw0, d_grads + OFF_W16, 576 * C, accumulate, (const unsigned *)nullptr);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

}  // namespace pr
