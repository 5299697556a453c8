/*
 * oracle/tvl1_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the reference's TV-L1 optical flow (tvl1flow/*.c of
 * gauenk/frame2frame) used only as the parity checker for the CUDA path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * call it.  It is organised the way the CUDA kernels are organised (one fused
 * pass per primal-dual iteration, one fused pass per warp stage) instead of
 * the reference's one-loop-per-array structure, but every floating-point
 * operation is performed in the same type and order as the reference so that
 * the result is bit-identical to it (validated against oracle/_ref, the
 * reference's own sources compiled in place; see tests/test_oracle_tvl1.py).
 *
 * Parity pin: the reference ships no tests or golden vectors for this path
 * ("parity unpinned" by reference tests); this restatement is pinned by
 * direct comparison with the compiled reference (oracle/_ref/libtvl1_ref.so)
 * and by the fixtures under tests/golden/ generated from it.
 *
 * Build: gcc -std=c99 -O2 -ffp-contract=off -fopenmp -fPIC -shared
 * (-ffp-contract=off: the reference is built -std=c99 => no FMA contraction).
 *
 * Each function names the reference lines it follows (paths relative to the
 * reference root).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define TVO_MAX_ITER 300            /* tvl1flow_lib.c:23 */
#define TVO_PRESMOOTH_SIGMA 0.8     /* tvl1flow_lib.c:24 */
#define TVO_GRAD_IS_ZERO 1E-10      /* tvl1flow_lib.c:25 */
#define TVO_ZOOM_SIGMA_ZERO 0.6     /* zoom.c:17 */
#define TVO_MAX_SCALES 32

/* ---------------------------------------------------------------- sizes */

/* zoom.c:24-36 : (int)((float)n*factor + 0.5), the 0.5 is a double. */
void tvo_zoom_size(int nx, int ny, int *nxx, int *nyy, float factor)
{
    *nxx = (int)((float)nx * factor + 0.5);
    *nyy = (int)((float)ny * factor + 0.5);
}

/* main.c:159-163 : clamp nscales so the coarsest level stays >= ~16 px. */
void tvo_clamp_scales(int nx, int ny, float zfactor, int *nscales, int *fscale)
{
    const float N = 1 + log(hypot(nx, ny) / 16.0) / log(1 / zfactor);
    if (N < *nscales) *nscales = N;
    if (*nscales < *fscale) *fscale = *nscales;
}

/* ------------------------------------------------------------- gaussian */

/* mask.c:243-255 : taps B[0..size-1], size = (int)(5*sigma)+1. */
int tvo_gaussian_taps(double sigma, double *B, int maxtaps)
{
    const double den = 2 * sigma * sigma;
    const int size = (int)(5 * sigma) + 1;
    if (size > maxtaps) return -1;
    for (int i = 0; i < size; i++)
        B[i] = 1 / (sigma * sqrt(2.0 * 3.1415926)) * exp(-i * i / den);
    double norm = 0;
    for (int i = 0; i < size; i++) norm += B[i];
    norm *= 2;
    norm -= B[0];
    for (int i = 0; i < size; i++) B[i] /= norm;
    return size;
}

/* Index of the sample that the reference's padded line holds at logical
 * position c in [-size, n+size): mask.c:273-278 / 313-318.  Left pad mirrors
 * without repeating the edge, right pad mirrors repeating it. */
static inline int tvo_reflect(int c, int n)
{
    if (c < 0) return -c;
    if (c >= n) return 2 * n - 1 - c;
    return c;
}

/* mask.c:223-339 : separable smoothing, double accumulation, each pass
 * rounded to float.  Returns -1 where the reference would abort (:238-241). */
int tvo_gaussian(float *I, int xdim, int ydim, double sigma)
{
    double B[64];
    const int size = tvo_gaussian_taps(sigma, B, 64);
    if (size < 0 || size > xdim) return -1;
    float *tmp = malloc((size_t)xdim * ydim * sizeof *tmp);
    /* rows: I -> tmp */
#pragma omp parallel for
    for (int y = 0; y < ydim; y++) {
        const float *row = I + (size_t)y * xdim;
        for (int x = 0; x < xdim; x++) {
            double sum = B[0] * (double)row[x];
            for (int j = 1; j < size; j++)
                sum += B[j] * ((double)row[tvo_reflect(x - j, xdim)] +
                               (double)row[tvo_reflect(x + j, xdim)]);
            tmp[(size_t)y * xdim + x] = sum;
        }
    }
    /* columns: tmp -> I */
#pragma omp parallel for
    for (int y = 0; y < ydim; y++) {
        for (int x = 0; x < xdim; x++) {
            double sum = B[0] * (double)tmp[(size_t)y * xdim + x];
            for (int j = 1; j < size; j++)
                sum += B[j] *
                       ((double)tmp[(size_t)tvo_reflect(y - j, ydim) * xdim + x] +
                        (double)tmp[(size_t)tvo_reflect(y + j, ydim) * xdim + x]);
            I[(size_t)y * xdim + x] = sum;
        }
    }
    free(tmp);
    return 0;
}

/* -------------------------------------------------------------- bicubic */

/* bicubic_interpolation.c:102-110 : Keys a=-1/2 cubic, all in double. */
static inline double tvo_keys(double v0, double v1, double v2, double v3,
                              double t)
{
    return v1 + 0.5 * t * (v2 - v0 +
           t * (2.0 * v0 - 5.0 * v1 + 4.0 * v2 - v3 +
           t * (3.0 * (v1 - v2) + v3 - v0)));
}

static inline int tvo_clampi(int v, int n, int *out)
{
    if (v < 0) { *out = 1; return 0; }
    if (v >= n) { *out = 1; return n - 1; }
    return v;
}

/* bicubic_interpolation.c:138-234 (Neumann boundary compiled in, :15).
 * Truncating (int) casts; the row "minus" tap uses sx (sic, :159);
 * border_out returns 0 when any tap index was clamped (:197-198). */
float tvo_bicubic_at(const float *in, float uu, float vv, int nx, int ny,
                     int border_out)
{
    const int sx = (uu < 0) ? -1 : 1;
    const int sy = (vv < 0) ? -1 : 1;
    int out = 0;
    const int ix = (int)uu, iy = (int)vv;
    int cx[4], cy[4];
    cx[1] = tvo_clampi(ix, nx, &out);
    cy[1] = tvo_clampi(iy, ny, &out);
    cx[0] = tvo_clampi(ix - sx, nx, &out);
    cy[0] = tvo_clampi(iy - sx, ny, &out);
    cx[2] = tvo_clampi(ix + sx, nx, &out);
    cy[2] = tvo_clampi(iy + sy, ny, &out);
    cx[3] = tvo_clampi(ix + 2 * sx, nx, &out);
    cy[3] = tvo_clampi(iy + 2 * sy, ny, &out);
    if (out && border_out) return 0.0f;

    /* fractional parts are float subtractions of a float and an int (:232) */
    const double tx = uu - cx[1];
    const double ty = vv - cy[1];
    double col[4];
    for (int c = 0; c < 4; c++) /* along y inside each column first (:224-229) */
        col[c] = tvo_keys(in[cx[c] + (size_t)nx * cy[0]],
                          in[cx[c] + (size_t)nx * cy[1]],
                          in[cx[c] + (size_t)nx * cy[2]],
                          in[cx[c] + (size_t)nx * cy[3]], ty);
    return tvo_keys(col[0], col[1], col[2], col[3], tx);
}

/* zoom.c:43-81 */
int tvo_zoom_out(const float *I, float *Iout, int nx, int ny, float factor)
{
    float *Is = malloc((size_t)nx * ny * sizeof *Is);
    memcpy(Is, I, (size_t)nx * ny * sizeof *Is);
    int nxx, nyy;
    tvo_zoom_size(nx, ny, &nxx, &nyy, factor);
    const float sigma = TVO_ZOOM_SIGMA_ZERO * sqrt(1.0 / (factor * factor) - 1.0);
    if (tvo_gaussian(Is, nx, ny, sigma)) { free(Is); return -1; }
#pragma omp parallel for
    for (int i1 = 0; i1 < nyy; i1++)
        for (int j1 = 0; j1 < nxx; j1++) {
            const float i2 = (float)i1 / factor;
            const float j2 = (float)j1 / factor;
            Iout[(size_t)i1 * nxx + j1] = tvo_bicubic_at(Is, j2, i2, nx, ny, 0);
        }
    free(Is);
    return 0;
}

/* zoom.c:89-115 followed by the rescale of tvl1flow_lib.c:442-446, fused:
 * out = bicubic(I at (j/fx, i/fy)) * (float)(1.0/zfactor). */
void tvo_zoom_in_scaled(const float *I, float *Iout, int nx, int ny, int nxx,
                        int nyy, float zfactor)
{
    const float factorx = ((float)nxx / nx);
    const float factory = ((float)nyy / ny);
    const float mul = (float)1.0 / zfactor;
#pragma omp parallel for
    for (int i1 = 0; i1 < nyy; i1++)
        for (int j1 = 0; j1 < nxx; j1++) {
            const float i2 = (float)i1 / factory;
            const float j2 = (float)j1 / factorx;
            float g = tvo_bicubic_at(I, j2, i2, nx, ny, 0);
            g *= mul;
            Iout[(size_t)i1 * nxx + j1] = g;
        }
}

/* ------------------------------------------------------------- stencils */

/* mask.c:156-215 : one formula covers all the reference's edge cases:
 * dx = 0.5*(I[min(x+1,nx-1)] - I[max(x-1,0)]), same in y; the subtraction is
 * float, the product with 0.5 double, the store float. */
void tvo_centered_gradient(const float *I, float *dx, float *dy, int nx, int ny)
{
#pragma omp parallel for
    for (int y = 0; y < ny; y++)
        for (int x = 0; x < nx; x++) {
            const int xp = x + 1 < nx ? x + 1 : nx - 1, xm = x > 0 ? x - 1 : 0;
            const int yp = y + 1 < ny ? y + 1 : ny - 1, ym = y > 0 ? y - 1 : 0;
            const size_t k = (size_t)y * nx + x;
            dx[k] = 0.5 * (I[(size_t)y * nx + xp] - I[(size_t)y * nx + xm]);
            dy[k] = 0.5 * (I[(size_t)yp * nx + x] - I[(size_t)ym * nx + x]);
        }
}

/* mask.c:43-94 : backward-difference divergence with Chambolle's borders.
 * Written per pixel; the association of the float adds follows each of the
 * reference's nine cases. */
static inline float tvo_div_at(const float *v1, const float *v2, int x, int y,
                               int nx, int ny)
{
    const size_t p = (size_t)y * nx + x;
    const int xin = (x > 0 && x < nx - 1), yin = (y > 0 && y < ny - 1);
    if (xin && yin) {                       /* :55-68 */
        const float v1x = v1[p] - v1[p - 1];
        const float v2y = v2[p] - v2[p - nx];
        return v1x + v2y;
    }
    if (xin && y == 0) return v1[p] - v1[p - 1] + v2[p];          /* :75 */
    if (xin) return v1[p] - v1[p - 1] - v2[p - nx];               /* :76 */
    if (yin && x == 0) return v1[p] + v2[p] - v2[p - nx];         /* :85 */
    if (yin) return -v1[p - 1] + v2[p] - v2[p - nx];              /* :86 */
    if (x == 0 && y == 0) return v1[p] + v2[p];                   /* :90 */
    if (y == 0) return -v1[p - 1] + v2[p];                        /* :91 */
    if (x == 0) return v1[p] - v2[p - nx];                        /* :92 */
    return -v1[p - 1] - v2[p - nx];                               /* :93 */
}

/* ------------------------------------------------------- one-scale solve */

/* tvl1flow_lib.c:148-166 : three bicubic warps of (I1, I1x, I1y) by (u1,u2)
 * plus |grad|^2 and the constant part of rho, one pass. */
static void tvo_warp_stage(const float *I0, const float *I1, const float *I1x,
                           const float *I1y, const float *u1, const float *u2,
                           float *I1wx, float *I1wy, float *grad, float *rho_c,
                           int nx, int ny)
{
#pragma omp parallel for
    for (int y = 0; y < ny; y++)
        for (int x = 0; x < nx; x++) {
            const size_t i = (size_t)y * nx + x;
            const float uu = (float)(x + u1[i]);   /* bicubic_interpolation.c:259 */
            const float vv = (float)(y + u2[i]);
            const float w = tvo_bicubic_at(I1, uu, vv, nx, ny, 1);
            const float wx = tvo_bicubic_at(I1x, uu, vv, nx, ny, 1);
            const float wy = tvo_bicubic_at(I1y, uu, vv, nx, ny, 1);
            const float Ix2 = wx * wx;
            const float Iy2 = wy * wy;
            I1wx[i] = wx;
            I1wy[i] = wy;
            grad[i] = (Ix2 + Iy2);
            rho_c[i] = (w - wx * u1[i] - wy * u2[i] - I0[i]);
        }
}

/* tvl1flow_lib.c:178-234 for one pixel: thresholding step, div p, u update.
 * Returns the new (u1,u2) through pointers. */
static inline void tvo_primal_at(const float *rho_c, const float *I1wx,
                                 const float *I1wy, const float *grad,
                                 const float *u1, const float *u2,
                                 const float *p11, const float *p12,
                                 const float *p21, const float *p22, int x,
                                 int y, int nx, int ny, float l_t, float theta,
                                 float *n1, float *n2)
{
    const size_t i = (size_t)y * nx + x;
    const float rho = rho_c[i] + (I1wx[i] * u1[i] + I1wy[i] * u2[i]);
    float d1, d2;
    if (rho < -l_t * grad[i]) {
        d1 = l_t * I1wx[i];
        d2 = l_t * I1wy[i];
    } else if (rho > l_t * grad[i]) {
        d1 = -l_t * I1wx[i];
        d2 = -l_t * I1wy[i];
    } else if (grad[i] < TVO_GRAD_IS_ZERO) {
        d1 = d2 = 0;
    } else {
        const float fi = -rho / grad[i];
        d1 = fi * I1wx[i];
        d2 = fi * I1wy[i];
    }
    const float v1 = u1[i] + d1;
    const float v2 = u2[i] + d2;
    const float dp1 = tvo_div_at(p11, p12, x, y, nx, ny);
    const float dp2 = tvo_div_at(p21, p22, x, y, nx, ny);
    *n1 = v1 + theta * dp1;
    *n2 = v2 + theta * dp2;
}

/* tvl1flow_lib.c:96-286.  The reference's seven loops per iteration become
 * two passes over ping-pong buffers: (1) u_new and the squared change,
 * (2) forward gradient of u_new (mask.c:103-148) and the dual update
 * (:244-256: hypot in double rounded to float, 1.0+taut*g in double rounded
 * to float, float division).  The squared flow change is accumulated in
 * double here (the reference accumulates per-thread floats, which makes its
 * value -- not its flow -- depend on the thread count).
 * iters[w], errs[w] receive the count and last error of each warp. */
void tvo_solve_scale(const float *I0, const float *I1, float *u1, float *u2,
                     int nx, int ny, float tau, float lambda, float theta,
                     int warps, float epsilon, int *iters, float *errs)
{
    const size_t size = (size_t)nx * ny;
    const float l_t = lambda * theta;
    const size_t sf = sizeof(float);
    float *I1x = malloc(size * sf), *I1y = malloc(size * sf);
    float *I1wx = malloc(size * sf), *I1wy = malloc(size * sf);
    float *grad = malloc(size * sf), *rho_c = malloc(size * sf);
    float *n1 = malloc(size * sf), *n2 = malloc(size * sf);
    float *p11 = calloc(size, sf), *p12 = calloc(size, sf);
    float *p21 = calloc(size, sf), *p22 = calloc(size, sf);

    tvo_centered_gradient(I1, I1x, I1y, nx, ny);

    for (int w = 0; w < warps; w++) {
        tvo_warp_stage(I0, I1, I1x, I1y, u1, u2, I1wx, I1wy, grad, rho_c, nx, ny);
        int n = 0;
        float error = INFINITY;
        while (error > epsilon * epsilon && n < TVO_MAX_ITER) {
            n++;
            double acc = 0.0;
#pragma omp parallel for reduction(+ : acc)
            for (int y = 0; y < ny; y++)
                for (int x = 0; x < nx; x++) {
                    const size_t i = (size_t)y * nx + x;
                    float a, b;
                    tvo_primal_at(rho_c, I1wx, I1wy, grad, u1, u2, p11, p12,
                                  p21, p22, x, y, nx, ny, l_t, theta, &a, &b);
                    n1[i] = a;
                    n2[i] = b;
                    const float e = (a - u1[i]) * (a - u1[i]) +
                                    (b - u2[i]) * (b - u2[i]);
                    acc += e;
                }
            error = (float)(acc / (double)size);
            /* dual update reads only n1,n2 (neighbours) and its own p */
#pragma omp parallel for
            for (int y = 0; y < ny; y++)
                for (int x = 0; x < nx; x++) {
                    const size_t i = (size_t)y * nx + x;
                    const float u1x = x < nx - 1 ? n1[i + 1] - n1[i] : 0.0f;
                    const float u1y = y < ny - 1 ? n1[i + nx] - n1[i] : 0.0f;
                    const float u2x = x < nx - 1 ? n2[i + 1] - n2[i] : 0.0f;
                    const float u2y = y < ny - 1 ? n2[i + nx] - n2[i] : 0.0f;
                    const float taut = tau / theta;
                    const float g1 = hypot(u1x, u1y);
                    const float g2 = hypot(u2x, u2y);
                    const float ng1 = 1.0 + taut * g1;
                    const float ng2 = 1.0 + taut * g2;
                    p11[i] = (p11[i] + taut * u1x) / ng1;
                    p12[i] = (p12[i] + taut * u1y) / ng1;
                    p21[i] = (p21[i] + taut * u2x) / ng2;
                    p22[i] = (p22[i] + taut * u2y) / ng2;
                }
            memcpy(u1, n1, size * sf);
            memcpy(u2, n2, size * sf);
        }
        if (iters) iters[w] = n;
        if (errs) errs[w] = error;
    }
    free(I1x); free(I1y); free(I1wx); free(I1wy); free(grad); free(rho_c);
    free(n1); free(n2); free(p11); free(p12); free(p21); free(p22);
}

/* --------------------------------------------------------- normalisation */

/* tvl1flow_lib.c:293-348 : joint min/max; 255.0*(I-min)/den in double. */
void tvo_normalize(const float *I0, const float *I1, float *I0n, float *I1n,
                   int size)
{
    float mn = I0[0], mx = I0[0];
    for (int i = 0; i < size; i++) {
        if (I0[i] < mn) mn = I0[i];
        if (I0[i] > mx) mx = I0[i];
        if (I1[i] < mn) mn = I1[i];
        if (I1[i] > mx) mx = I1[i];
    }
    const float den = mx - mn;
    if (den > 0) {
        for (int i = 0; i < size; i++) {
            I0n[i] = 255.0 * (I0[i] - mn) / den;
            I1n[i] = 255.0 * (I1[i] - mn) / den;
        }
    } else {
        memcpy(I0n, I0, (size_t)size * sizeof(float));
        memcpy(I1n, I1, (size_t)size * sizeof(float));
    }
}

/* ------------------------------------------------------------ multiscale */

/* tvl1flow_lib.c:356-485.  iters/errs are [nscales][warps] (row s = scale s;
 * untouched rows for scales below fscale).  Returns 0, or -1 where the
 * reference would abort in gaussian(). */
int tvo_multiscale(const float *I0, const float *I1, float *u1, float *u2,
                   int nxx, int nyy, float tau, float lambda, float theta,
                   int nscales, int fscale, float zfactor, int warps,
                   float epsilon, int *iters, float *errs)
{
    if (nscales > TVO_MAX_SCALES || nscales < 1) return -2;
    const size_t size = (size_t)nxx * nyy;
    float *I0s[TVO_MAX_SCALES], *I1s[TVO_MAX_SCALES];
    float *u1s[TVO_MAX_SCALES], *u2s[TVO_MAX_SCALES];
    int nx[TVO_MAX_SCALES], ny[TVO_MAX_SCALES];
    int rc = 0;

    I0s[0] = malloc(size * sizeof(float));
    I1s[0] = malloc(size * sizeof(float));
    u1s[0] = u1; u2s[0] = u2; nx[0] = nxx; ny[0] = nyy;
    tvo_normalize(I0, I1, I0s[0], I1s[0], (int)size);
    if (tvo_gaussian(I0s[0], nxx, nyy, TVO_PRESMOOTH_SIGMA)) rc = -1;
    if (tvo_gaussian(I1s[0], nxx, nyy, TVO_PRESMOOTH_SIGMA)) rc = -1;

    int built = 1;
    for (int s = 1; s < nscales && !rc; s++, built++) {
        tvo_zoom_size(nx[s - 1], ny[s - 1], &nx[s], &ny[s], zfactor);
        const size_t sz = (size_t)nx[s] * ny[s];
        I0s[s] = malloc(sz * sizeof(float));
        I1s[s] = malloc(sz * sizeof(float));
        u1s[s] = malloc(sz * sizeof(float));
        u2s[s] = malloc(sz * sizeof(float));
        if (tvo_zoom_out(I0s[s - 1], I0s[s], nx[s - 1], ny[s - 1], zfactor)) rc = -1;
        if (tvo_zoom_out(I1s[s - 1], I1s[s], nx[s - 1], ny[s - 1], zfactor)) rc = -1;
    }
    if (!rc) {
        const size_t szc = (size_t)nx[nscales - 1] * ny[nscales - 1];
        for (size_t i = 0; i < szc; i++)
            u1s[nscales - 1][i] = u2s[nscales - 1][i] = 0.0f;
        for (int s = nscales - 1; s >= 0; s--) {
            if (s >= fscale)
                tvo_solve_scale(I0s[s], I1s[s], u1s[s], u2s[s], nx[s], ny[s],
                                tau, lambda, theta, warps, epsilon,
                                iters ? iters + s * warps : 0,
                                errs ? errs + s * warps : 0);
            if (!s) break;
            tvo_zoom_in_scaled(u1s[s], u1s[s - 1], nx[s], ny[s], nx[s - 1],
                               ny[s - 1], zfactor);
            tvo_zoom_in_scaled(u2s[s], u2s[s - 1], nx[s], ny[s], nx[s - 1],
                               ny[s - 1], zfactor);
        }
    }
    for (int s = 1; s < built; s++) {
        free(I0s[s]); free(I1s[s]); free(u1s[s]); free(u2s[s]);
    }
    free(I0s[0]); free(I1s[0]);
    return rc;
}
