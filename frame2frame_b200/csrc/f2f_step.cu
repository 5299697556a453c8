// f2f_step.cu -- the per-frame-pair frame-to-frame step for sm_100a:
// bilinear flow warp of the previous frame + validity mask + flow-consistency
// occlusion mask (one fused kernel, once per frame), and the masked loss with
// its gradient (one pass).
//
// Replaces WarpedLoss.warp / occlusion_mask / forward of the reference
// (blind_denoising.py:44-122; lib/frame2frame/warped_loss.py:145-248), which
// rebuilds a meshgrid on the host, runs grid_sample twice, and computes the
// occlusion map with numpy + scipy on the host every inner iteration.
//
// Arithmetic follows what the reference executes under torch 2.11 on CPU
// (probed, see oracle/f2f_oracle.py): grid = 2*(x+u)/(W-1)-1 with a true
// division; ATen's align_corners=False unnormalisation is ONE fused
// multiply-add ix = fma(g+1, W/2, -0.5); weights e = 1-w, s = 1-n; the blend is
// a chain of fused multiply-adds.  This file is compiled with -fmad=false and
// every fma below is explicit, so the roundings are exactly those.
#include "f2f_common.cuh"

namespace f2f {

// |du/dy + dv/dx| > 0.75 at (y,x) of the batch-0 flow (blind_denoising.py:92-94);
// F64: sum and compare in double like the numpy code, else float like
// warped_loss.py:188-190.
template <bool F64>
__device__ __forceinline__ bool occ_raw(const float *__restrict__ u, const float *__restrict__ v, int y, int x, int H,
                                        int W)
{
    if (y < 0 || y >= H || x < 0 || x >= W) return false;   // binary_dilation border_value = 0
    const size_t i = (size_t)y * W + x;
    const float a = (y < H - 1) ? u[i + W] - u[i] : 0.0f;
    const float b = (x < W - 1) ? v[i + 1] - v[i] : 0.0f;
    if (F64) return fabs((double)a + (double)b) > 0.75;
    return fabsf(a + b) > 0.75f;
}

// One thread = WM_ROWS vertically adjacent pixels, processed in phases over all of them: flow loads, then coordinates
// and the gathered taps, then the blend and the stores -- so that WM_ROWS independent chains of the two dependent DRAM
// round trips (flow, then taps) are in flight per thread.  (One pixel per thread ran seven waves of blocks, each exposing
// both latencies: 49 us at 1080p, 13 % of the HBM peak for 20 B/px.)  Per-pixel arithmetic is unchanged.
constexpr int WM_ROWS = 4;
template <bool F64>
__global__ void __launch_bounds__(256, 4) k_warp_mask(const float *__restrict__ xin, const float *__restrict__ flow,
                                                   const float *__restrict__ in_mask, float *__restrict__ warped,
                                                   float *__restrict__ valid, float *__restrict__ mask, int B, int C,
                                                   int H, int W, int align_corners)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, yb = blockIdx.y * WM_ROWS, b = blockIdx.z;
    if (x >= W) return;
    const size_t P = (size_t)H * W;
    const float *fu = flow + (size_t)b * 2 * P, *fv = fu + P;
    float du[WM_ROWS], dv[WM_ROWS];
#pragma unroll
    for (int k = 0; k < WM_ROWS; k++) {
        const int y = min(yb + k, H - 1);                  // rows past the image repeat the last one (not stored)
        const size_t i = (size_t)y * W + x;
        du[k] = fu[i];
        dv[k] = fv[i];
    }
    float nw[WM_ROWS], ne[WM_ROWS], sw[WM_ROWS], se[WM_ROWS], okv[WM_ROWS];
    size_t o00[WM_ROWS];
    unsigned tapm[WM_ROWS];                                // bits 0..3: taps 00, 01, 10, 11 inside the image
#pragma unroll
    for (int k = 0; k < WM_ROWS; k++) {
        const int y = min(yb + k, H - 1);
        // blind_denoising.py:62-69
        const float X = (float)x + du[k];
        const float Y = (float)y + dv[k];
        const float gx = (2.0f * X) / (float)max(W - 1, 1) - 1.0f;
        const float gy = (2.0f * Y) / (float)max(H - 1, 1) - 1.0f;
        float ix, iy;
        if (align_corners) {
            ix = (gx + 1.0f) * ((float)(W - 1) / 2);
            iy = (gy + 1.0f) * ((float)(H - 1) / 2);
        } else {   // as run by the reference: grid_sample default since torch 1.3
            ix = __fmaf_rn(gx + 1.0f, (float)W / 2, -0.5f);
            iy = __fmaf_rn(gy + 1.0f, (float)H / 2, -0.5f);
        }
        const float x0f = floorf(ix), y0f = floorf(iy);
        const float w = ix - x0f, e = 1.0f - w, n = iy - y0f, s_ = 1.0f - n;
        nw[k] = s_ * e; ne[k] = s_ * w; sw[k] = n * e; se[k] = n * w;
        // floorf of a float far outside int range: clamp before converting (everything out of range is padding)
        const int x0 = (int)fminf(fmaxf(x0f, -2.0f), (float)W + 1.0f), y0 = (int)fminf(fmaxf(y0f, -2.0f), (float)H + 1.0f);
        const bool okx0 = x0 >= 0 && x0 < W, okx1 = x0 + 1 >= 0 && x0 + 1 < W;
        const bool oky0 = y0 >= 0 && y0 < H, oky1 = y0 + 1 >= 0 && y0 + 1 < H;
        const bool nan_in = !(ix == ix) || !(iy == iy);
        const bool m00 = okx0 && oky0 && !nan_in, m01 = okx1 && oky0 && !nan_in, m10 = okx0 && oky1 && !nan_in,
                   m11 = okx1 && oky1 && !nan_in;
        tapm[k] = (m00 ? 1u : 0u) | (m01 ? 2u : 0u) | (m10 ? 4u : 0u) | (m11 ? 8u : 0u);
        o00[k] = (size_t)y0 * W + x0;
        // grid_sample of ones (blind_denoising.py:73-76)
        float wsum = m00 ? nw[k] : 0.0f;
        wsum = (m01 ? ne[k] : 0.0f) + wsum;
        wsum = (m10 ? sw[k] : 0.0f) + wsum;
        wsum = (m11 ? se[k] : 0.0f) + wsum;
        okv[k] = (wsum >= 0.9999f) ? 1.0f : 0.0f;
    }
    // occlusion map from the flow of batch item 0 (blind_denoising.py:92-93 index of[0, ...]): the raw test on rows
    // yb-1 .. yb+WM_ROWS at column x and on this thread's rows at columns x-1, x+1
    const float *u0 = flow, *v0 = flow + P;
    bool rawc[WM_ROWS + 2], dil[WM_ROWS];
#pragma unroll
    for (int k = 0; k < WM_ROWS + 2; k++) rawc[k] = occ_raw<F64>(u0, v0, yb + k - 1, x, H, W);
#pragma unroll
    for (int k = 0; k < WM_ROWS; k++)
        dil[k] = rawc[k + 1] || rawc[k] || rawc[k + 2] || occ_raw<F64>(u0, v0, yb + k, x - 1, H, W) ||
                 occ_raw<F64>(u0, v0, yb + k, x + 1, H, W);
    for (int c = 0; c < C; c++) {
        const size_t plane = ((size_t)b * C + c) * P;
        const float *img = xin + plane;
        float t00[WM_ROWS], t01[WM_ROWS], t10[WM_ROWS], t11[WM_ROWS];
#pragma unroll
        for (int k = 0; k < WM_ROWS; k++) {
            t00[k] = (tapm[k] & 1u) ? img[o00[k]] : 0.0f;
            t01[k] = (tapm[k] & 2u) ? img[o00[k] + 1] : 0.0f;
            t10[k] = (tapm[k] & 4u) ? img[o00[k] + W] : 0.0f;
            t11[k] = (tapm[k] & 8u) ? img[o00[k] + W + 1] : 0.0f;
        }
#pragma unroll
        for (int k = 0; k < WM_ROWS; k++) {
            const int y = yb + k;
            if (y >= H) break;
            const size_t i = (size_t)y * W + x;
            warped[plane + i] = __fmaf_rn(t11[k], se[k], __fmaf_rn(t10[k], sw[k], __fmaf_rn(t01[k], ne[k], t00[k] * nw[k])));
            if (valid) valid[plane + i] = okv[k];
            if (mask) {
                // only mask[0,0] is dilated by the reference (blind_denoising.py:103)
                const bool frame = (y == 0) || (y == H - 1) || (x == 0) || (x == W - 1);   // :106-109
                const bool occ = ((b == 0 && c == 0) ? dil[k] : rawc[k + 1]) || frame;
                float m = okv[k] * (occ ? 0.0f : 1.0f);                 // :112
                if (in_mask) m = in_mask[plane + i] * m;                // warped_loss.py:222-223
                mask[plane + i] = m;
            }
        }
    }
}

// occlusion_mask() alone, for the reference-shaped API (warped only gives the size)
template <bool F64>
__global__ void __launch_bounds__(256) k_occlusion(const float *__restrict__ flow, const float *__restrict__ old_mask,
                                                   float *__restrict__ mask, int B, int C, int H, int W)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const size_t P = (size_t)H * W, i = (size_t)y * W + x;
    const float *u0 = flow, *v0 = flow + P;
    const bool frame = (y == 0) || (y == H - 1) || (x == 0) || (x == W - 1);
    const bool raw = occ_raw<F64>(u0, v0, y, x, H, W);
    const bool dil = raw || occ_raw<F64>(u0, v0, y - 1, x, H, W) || occ_raw<F64>(u0, v0, y + 1, x, H, W) ||
                     occ_raw<F64>(u0, v0, y, x - 1, H, W) || occ_raw<F64>(u0, v0, y, x + 1, H, W);
    for (int c = 0; c < C; c++) {
        const size_t plane = ((size_t)b * C + c) * P;
        const bool occ = ((b == 0 && c == 0) ? dil : raw) || frame;
        mask[plane + i] = old_mask[plane + i] * (occ ? 0.0f : 1.0f);
    }
}

// Masked loss and its gradient wrt `in`, one pass.
//  kind 0: sum |m*in - m*w|                (blind_denoising.py:121, L1Loss(size_average=False))
//  kind 1: mean (m*in - m*w)^2             (warped_loss.py:227,244-246)
//  kind 2: mean sqrt((m*in - m*w)^2 + 1e-6) (warped_loss.py:240-243)
template <int KIND>
__global__ void __launch_bounds__(256) k_masked_loss(const float *__restrict__ in, const float *__restrict__ warped,
                                                     const float *__restrict__ mask, float *__restrict__ grad,
                                                     double *loss, size_t n, float inv_n)
{
    double acc = 0.0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float m = mask[i];
        const float d = m * in[i] - m * warped[i];
        float g;
        if (KIND == 0) {
            acc += (double)fabsf(d);
            g = m * ((d > 0.0f) ? 1.0f : ((d < 0.0f) ? -1.0f : 0.0f));
        } else if (KIND == 1) {
            acc += (double)(d * d);
            g = m * ((2.0f * d) * inv_n);
        } else {
            const float r = sqrtf(d * d + 1e-6f);
            acc += (double)r;
            g = m * ((d / r) * inv_n);
        }
        if (grad) grad[i] = g;
    }
    acc = warp_sum(acc);
    __shared__ double sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < (int)(blockDim.x >> 5); k++) t += sm[k];
        atomicAdd(loss, KIND == 0 ? t : t * (double)inv_n);
    }
}

__global__ void k_zero_double(double *p) { *p = 0.0; }

extern long g_f2f_step_launches;
long g_f2f_step_launches = 0;

}  // namespace f2f

using namespace f2f;

extern "C" {

int f2f_warp_mask(const float *d_x, const float *d_flow, const float *d_in_mask, float *d_warped, float *d_valid,
                  float *d_mask, int B, int C, int H, int W, int align_corners, int occ_f64, void *stream)
{
    if (B == 0) return F2F_OK;
    if (!d_x || !d_flow || !d_warped || B < 0 || C < 1 || H < 1 || W < 1) {
        set_error("f2f_warp_mask: bad arguments");
        return F2F_EINVAL;
    }
    dim3 grid(cdiv(W, 256), cdiv(H, WM_ROWS), B);
    if (occ_f64)
        k_warp_mask<true><<<grid, 256, 0, (cudaStream_t)stream>>>(d_x, d_flow, d_in_mask, d_warped, d_valid, d_mask, B, C,
                                                                  H, W, align_corners);
    else
        k_warp_mask<false><<<grid, 256, 0, (cudaStream_t)stream>>>(d_x, d_flow, d_in_mask, d_warped, d_valid, d_mask, B,
                                                                   C, H, W, align_corners);
    g_f2f_step_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

int f2f_occlusion_mask(const float *d_flow, const float *d_old_mask, float *d_mask, int B, int C, int H, int W,
                       int occ_f64, void *stream)
{
    if (B == 0) return F2F_OK;
    if (!d_flow || !d_old_mask || !d_mask || B < 0 || C < 1 || H < 1 || W < 1) {
        set_error("f2f_occlusion_mask: bad arguments");
        return F2F_EINVAL;
    }
    dim3 grid(cdiv(W, 256), H, B);
    if (occ_f64)
        k_occlusion<true><<<grid, 256, 0, (cudaStream_t)stream>>>(d_flow, d_old_mask, d_mask, B, C, H, W);
    else
        k_occlusion<false><<<grid, 256, 0, (cudaStream_t)stream>>>(d_flow, d_old_mask, d_mask, B, C, H, W);
    g_f2f_step_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

int f2f_masked_loss(const float *d_in, const float *d_warped, const float *d_mask, float *d_grad, double *d_loss,
                    size_t n, int kind, void *stream)
{
    if (!d_in || !d_warped || !d_mask || !d_loss || kind < 0 || kind > 2) {
        set_error("f2f_masked_loss: bad arguments");
        return F2F_EINVAL;
    }
    cudaStream_t st = (cudaStream_t)stream;
    k_zero_double<<<1, 1, 0, st>>>(d_loss);
    if (n) {
        const size_t cap = (size_t)sm_count() * 8;
        const int blocks = (int)((n + 255) / 256 < cap ? (n + 255) / 256 : cap);
        const float inv_n = 1.0f / (float)n;
        if (kind == 0) k_masked_loss<0><<<blocks, 256, 0, st>>>(d_in, d_warped, d_mask, d_grad, d_loss, n, inv_n);
        if (kind == 1) k_masked_loss<1><<<blocks, 256, 0, st>>>(d_in, d_warped, d_mask, d_grad, d_loss, n, inv_n);
        if (kind == 2) k_masked_loss<2><<<blocks, 256, 0, st>>>(d_in, d_warped, d_mask, d_grad, d_loss, n, inv_n);
    }
    g_f2f_step_launches += 2;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

}  // extern "C"
