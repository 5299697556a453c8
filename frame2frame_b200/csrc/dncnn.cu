// dncnn.cu -- DnCNN-17 (3x3 / 64-channel conv + BatchNorm + ReLU stack) forward
// and backward for sm_100a.  Replaces the cuDNN/ATen calls behind
// lib/frame2frame/dncnn/net.py:6-39 as driven by blind_denoising.py:211-230.
//
// The 64->64 convolutions are implicit GEMMs on the 5th-generation tensor cores:
// tcgen05.mma (kind::f16, BF16 x BF16 -> FP32) issued by one thread per CTA,
// accumulators in TMEM, weights staged once per CTA by the TMA engine
// (cp.async.bulk), activations in NHWC BF16 (one pixel = 64 channels = 128 B).
//
// Shared-memory operand layout ("chunk-planar", no swizzle): a haloed 18x18
// pixel tile is stored as 8 planes (one per 8-channel / 16-byte chunk) of
// consecutive pixels.  In that layout a UMMA core matrix (8 rows x 16 B) is 8
// horizontally adjacent pixels, so the A operand of filter tap (dy,dx) is the
// SAME tile read through a descriptor whose start address is advanced by
// ((dy+1)*18 + (dx+1)) * 16 bytes: nine taps, one tile, no im2col copies and no
// swizzle-phase constraints.  The same planes read as an MN-major operand give
// the pixel-contraction GEMM of the weight gradient.
//
// Train-mode BatchNorm needs the per-channel batch statistics of a layer's
// whole output before it can be normalised, so conv(k) stores the
// raw z_k (BF16) and reduces sum / sum-of-squares in its epilogue, and the
// normalise + ReLU is applied by conv(k+1)'s operand producer while it stages
// the tile -- no separate pass over the activations.
//
// CTA = 13 warps, persistent over 16x16-pixel tiles:
//   warps 0-3   epilogue  (TMEM -> registers -> smem transpose -> coalesced stores + statistics)
//   warps 4-11  producers (global -> BN/ReLU -> chunk-planar smem, 3-stage ring): two groups of four
//               warps take alternate tiles, each with the whole haloed tile's loads in flight at
//               once (the layer is within ~20% of being HBM-bound, so bytes in flight matter)
//   warp  12    MMA issuer (72 tcgen05.mma per tile into a double-buffered accumulator) + TMEM owner
#include <vector>
#include "f2f_common.cuh"
#include "tc05.cuh"

namespace f2f {

long g_dncnn_launches = 0;

// ---- per-launch CUDA-event timing of the tensor-core kernels (f2f_profile_dncnn) ----
namespace {
struct ProfClass {
    std::vector<cudaEvent_t> ev;   // pairs: begin, end
    size_t used = 0;
};
ProfClass g_prof_cls[2];
bool g_prof_on = false;
}  // namespace

void prof_begin(int cls, cudaStream_t st)
{
    if (!g_prof_on) return;
    ProfClass &p = g_prof_cls[cls];
    if (p.used + 2 > p.ev.size()) {
        for (int i = 0; i < 2; i++) {
            cudaEvent_t e;
            cudaEventCreate(&e);
            p.ev.push_back(e);
        }
    }
    cudaEventRecord(p.ev[p.used], st);
}
void prof_end(int cls, cudaStream_t st)
{
    if (!g_prof_on) return;
    ProfClass &p = g_prof_cls[cls];
    cudaEventRecord(p.ev[p.used + 1], st);
    p.used += 2;
}

namespace dn {

using namespace tc05;
typedef __nv_bfloat16 bf16;

constexpr int TW = 16, TH = 16;              // output tile
constexpr int HW = TW + 2, HH = TH + 2;      // with the 1-px halo of a 3x3 filter
constexpr int HPIX = HW * HH;                // 324
constexpr int CH_STRIDE = (HPIX + 1) * 16;   // bytes between 8-channel planes; odd multiple of 16 B: conflict-free staging
constexpr int A_STAGE = 8 * CH_STRIDE;       // 41600
constexpr int STAGES = 3;
constexpr int B_TAP = 64 * 64 * 2;           // one tap's 64x64 BF16 weights
constexpr int B_BYTES = 9 * B_TAP;           // 73728
constexpr int EPI_ROW = 144;                 // 128 B of channels + 16 B pad (bank spread)
constexpr int EPI_WARP = 32 * EPI_ROW;
constexpr int EPI_BYTES = 4 * EPI_WARP;
constexpr int OFF_B = 0;
constexpr int OFF_A = OFF_B + B_BYTES;
constexpr int OFF_EPI = OFF_A + STAGES * A_STAGE;
constexpr int OFF_BAR = OFF_EPI + EPI_BYTES;
constexpr int SMEM_BYTES = OFF_BAR + 256;
constexpr int NTHREADS = 416;
constexpr int MMA_WARP = 12;
constexpr int TMEM_COLS = 256;               // 2 tile buffers x 2 halves x 64 fp32 columns

enum Prologue { PRO_RAW = 0, PRO_BN_RELU = 1 };
enum Epilogue { EPI_STORE = 0, EPI_STATS = 1, EPI_DGRAD_BN = 2, EPI_DGRAD_RELU = 3 };

struct ConvArgs {
    const bf16 *in;        // [H][W][64]
    bf16 *out;             // [H][W][64]
    const bf16 *wprep;     // [9 taps][8 k-chunks][64 n][8 k] BF16 = the smem image of B
    const float *scale;    // prologue BN: a = relu(in*scale + shift), per input channel
    const float *shift;
    double *stats;         // [2][64] accumulators (EPI_STATS: sum z, sum z^2; EPI_DGRAD_BN: sum g, sum g*xhat)
    const bf16 *zprev;     // EPI_DGRAD_*: pre-activation (BN) or activation (RELU) of the layer below
    const float *p_scale;  // EPI_DGRAD_BN: BN of the layer below: y = z*p_scale + p_shift, xhat = (z - p_mean)*p_invstd
    const float *p_shift;
    const float *p_mean;
    const float *p_invstd;
    int H, W, tiles_x, tiles_y;
};

// ------------------------------------------------------------------ producer
// Stages one haloed tile: 324 pixels x 8 chunks of 16 B.  Thread t of the 128-thread group owns
// chunk t%8 of pixels t/8, t/8+16, ...: 8 consecutive lanes read one pixel's 128 B, a warp reads
// 4 adjacent pixels.  All 21 loads of a thread are issued before the first is consumed.
template <int PRO>
__device__ __forceinline__ void produce_tile(const ConvArgs &a, int tile, uint8_t *stage, int ptid, const float (&sc)[8],
                                             const float (&sh)[8])
{
    const int ty0 = (tile / a.tiles_x) * TH - 1, tx0 = (tile % a.tiles_x) * TW - 1;
    const int chunk = ptid & 7;
    uint8_t *dst_plane = stage + chunk * CH_STRIDE;
    constexpr int ROUNDS = (HPIX + 15) / 16;   // 21 rounds of 16 pixels (128 threads / 8 chunks)
    const int pix0 = ptid >> 3;
    uint4 v[ROUNDS];
    uint32_t inb = 0;
    {
        int hy = pix0 / HW, hx = pix0 % HW;
#pragma unroll
        for (int u = 0; u < ROUNDS; u++) {
            const int gy = ty0 + hy, gx = tx0 + hx;
            const bool ok = pix0 + 16 * u < HPIX && gy >= 0 && gy < a.H && gx >= 0 && gx < a.W;
            v[u] = make_uint4(0, 0, 0, 0);
            if (ok) {
                v[u] = __ldg(reinterpret_cast<const uint4 *>(a.in + ((size_t)gy * a.W + gx) * 64 + chunk * 8));
                inb |= 1u << u;
            }
            hx += 16;
            if (hx >= HW) { hx -= HW; hy += 1; }
        }
    }
#pragma unroll
    for (int u = 0; u < ROUNDS; u++) {
        const int pix = pix0 + 16 * u;
        if (pix < HPIX) {
            uint4 o = v[u];
            if (PRO == PRO_BN_RELU && ((inb >> u) & 1u)) {   // zero padding applies to the ACTIVATION: out-of-image stays 0
                uint32_t w[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const float2 f = unpack_bf16(w[j]);
                    const float y0 = fmaxf(fmaf(f.x, sc[2 * j], sh[2 * j]), 0.0f);
                    const float y1 = fmaxf(fmaf(f.y, sc[2 * j + 1], sh[2 * j + 1]), 0.0f);
                    w[j] = pack_bf16(y0, y1);
                }
                o = make_uint4(w[0], w[1], w[2], w[3]);
            }
            *reinterpret_cast<uint4 *>(dst_plane + pix * 16) = o;
        }
    }
}

// ------------------------------------------------------------------ epilogue
// One warp, one 32-pixel quarter of a tile half: TMEM row -> BF16 -> smem transpose -> coalesced
// 512-B stores.  In the transposed phase a lane always holds the same 8 channels (chunk lane%8),
// so the per-channel sums (batch statistics / BatchNorm-backward sums) are plain per-thread
// accumulators over everything the thread stores, reduced once at the end of the kernel.
struct EpiCoef { float sc[8], sh[8], mu[8], is[8]; };

template <int EPI>
__device__ __forceinline__ void epilogue_half(const ConvArgs &a, uint32_t tmem_addr, uint8_t *stg, int lane, int warp,
                                              int ty_base, int tx_base, float (&acc0)[8], float (&acc1)[8],
                                              const EpiCoef &cf)
{
    const int chunk = lane & 7;
    // the layer below's pre-activation for this lane's 8 pixels: issued first, consumed last
    uint4 zz[8];
    uint32_t inbm = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int q = (lane >> 3) + 4 * i;
        const int gy = ty_base + warp * 4 + (q >> 3), gx = tx_base + (q & 7);
        const bool ok = gy < a.H && gx < a.W;
        inbm |= (ok ? 1u : 0u) << i;
        if (EPI == EPI_DGRAD_BN || EPI == EPI_DGRAD_RELU) {
            zz[i] = make_uint4(0, 0, 0, 0);
            if (ok) zz[i] = __ldg(reinterpret_cast<const uint4 *>(a.zprev + ((size_t)gy * a.W + gx) * 64 + chunk * 8));
        }
    }
    // this thread's accumulator row m = 32*warp + lane  <->  pixel (tile row m/8, column m%8 of the half)
    uint8_t *myrow = stg + lane * EPI_ROW;
#pragma unroll
    for (int part = 0; part < 2; part++) {
        uint32_t r[32];
        tmem_ld32(tmem_addr + part * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; j++) {
            uint4 o;
            o.x = pack_bf16(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1]));
            o.y = pack_bf16(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
            o.z = pack_bf16(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
            o.w = pack_bf16(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
            *reinterpret_cast<uint4 *>(myrow + part * 64 + j * 16) = o;
        }
    }
    __syncwarp();
    // transposed phase: lane handles chunk (lane % 8) of pixels q = lane/8 + 4*i  (4 pixels = 512 contiguous bytes)
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int q = (lane >> 3) + 4 * i;           // pixel index within the warp's 32
        const int gy = ty_base + warp * 4 + (q >> 3), gx = tx_base + (q & 7);
        uint4 o = *reinterpret_cast<const uint4 *>(stg + q * EPI_ROW + chunk * 16);
        const bool inb = (inbm >> i) & 1u;
        uint32_t w[4] = {o.x, o.y, o.z, o.w};
        if (EPI == EPI_DGRAD_BN || EPI == EPI_DGRAD_RELU) {
            // ReLU mask of the layer below (and its BN-backward sums), net.py:16-18 backward
            const uint32_t zw[4] = {zz[i].x, zz[i].y, zz[i].z, zz[i].w};
#pragma unroll
            for (int j = 0; j < 4; j++) {
                float2 g = unpack_bf16(w[j]);
                const float2 z = unpack_bf16(zw[j]);
                if (EPI == EPI_DGRAD_BN) {
                    if (!(fmaf(z.x, cf.sc[2 * j], cf.sh[2 * j]) > 0.0f)) g.x = 0.0f;
                    if (!(fmaf(z.y, cf.sc[2 * j + 1], cf.sh[2 * j + 1]) > 0.0f)) g.y = 0.0f;
                    if (inb) {
                        acc0[2 * j] += g.x; acc0[2 * j + 1] += g.y;
                        acc1[2 * j] = fmaf(g.x, (z.x - cf.mu[2 * j]) * cf.is[2 * j], acc1[2 * j]);
                        acc1[2 * j + 1] = fmaf(g.y, (z.y - cf.mu[2 * j + 1]) * cf.is[2 * j + 1], acc1[2 * j + 1]);
                    }
                } else {
                    if (!(z.x > 0.0f)) g.x = 0.0f;
                    if (!(z.y > 0.0f)) g.y = 0.0f;
                }
                w[j] = pack_bf16(g.x, g.y);
            }
            o = make_uint4(w[0], w[1], w[2], w[3]);
        }
        if (EPI == EPI_STATS && inb) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float2 v = unpack_bf16(w[j]);
                acc0[2 * j] += v.x; acc0[2 * j + 1] += v.y;
                acc1[2 * j] = fmaf(v.x, v.x, acc1[2 * j]); acc1[2 * j + 1] = fmaf(v.y, v.y, acc1[2 * j + 1]);
            }
        }
        if (inb) *reinterpret_cast<uint4 *>(a.out + ((size_t)gy * a.W + gx) * 64 + chunk * 8) = o;
    }
    __syncwarp();
}

// ------------------------------------------------------------------- kernel
template <int PRO, int EPI>
__global__ void __launch_bounds__(NTHREADS, 1) k_conv64(ConvArgs a)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntiles = a.tiles_x * a.tiles_y;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + OFF_BAR);
    // barrier slots: full[3] @0, empty[3] @3, tfull[2] @6, tempty[2] @8, wbar @10; tmem base pointer @ +12
    const uint32_t bar0 = smem_u32(bars);
    auto BAR = [&](int i) { return bar0 + 8u * i; };
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(bars + 12);

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(BAR(s), 128); mbar_init(BAR(3 + s), 1); }
        for (int b = 0; b < 2; b++) { mbar_init(BAR(6 + b), 1); mbar_init(BAR(8 + b), 128); }
        mbar_init(BAR(10), 1);
        mbar_fence_init();
    }
    if (warp == MMA_WARP) tmem_alloc(smem_u32(tmem_ptr), TMEM_COLS);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp >= 4 && warp < MMA_WARP) {
        // ===== producers: group 0 (warps 4-7) stages tiles it = 0, 2, 4, ...; group 1 (warps 8-11) the odd ones =====
        const int group = (warp - 4) >> 2;
        const int ptid = (threadIdx.x - 128) & 127;
        float sc[8], sh[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            sc[j] = PRO == PRO_BN_RELU ? a.scale[(ptid & 7) * 8 + j] : 1.0f;
            sh[j] = PRO == PRO_BN_RELU ? a.shift[(ptid & 7) * 8 + j] : 0.0f;
        }
        int it = group;
        for (int tile = blockIdx.x + group * gridDim.x; tile < ntiles; tile += 2 * gridDim.x, it += 2) {
            const int s = it % STAGES;
            const uint32_t ph = (it / STAGES) & 1;
            mbar_wait(BAR(3 + s), ph ^ 1);                 // slot free (first pass returns immediately)
            produce_tile<PRO>(a, tile, smem + OFF_A + s * A_STAGE, ptid, sc, sh);
            fence_proxy_async();                           // st.shared -> visible to the tensor core's async proxy
            mbar_arrive(BAR(s));
        }
    } else if (warp == MMA_WARP) {
        // ===== MMA issuer =====
        if (lane == 0) {
            // weights: one bulk copy by the TMA engine, resident for the whole kernel
            mbar_arrive_expect_tx(BAR(10), B_BYTES);
            for (int t = 0; t < 9; t++) bulk_g2s(smem_u32(smem + OFF_B + t * B_TAP), a.wprep + (size_t)t * 4096, B_TAP, BAR(10));
            mbar_wait(BAR(10), 0);
            constexpr uint32_t IDESC = idesc_bf16(128, 64, 0, 0);
            int it = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, it++) {
                const int s = it % STAGES, b = it & 1;
                mbar_wait(BAR(8 + b), ((it >> 1) & 1) ^ 1);   // accumulator buffer drained by the epilogue
                mbar_wait(BAR(s), (it / STAGES) & 1);          // operand tile staged
                fence_after_sync();
                const uint32_t a_base = smem_u32(smem + OFF_A + s * A_STAGE);
                const uint32_t b_base = smem_u32(smem + OFF_B);
#pragma unroll 1
                for (int tap = 0; tap < 9; tap++) {
                    const int dy = tap / 3, dx = tap % 3;      // halo coordinates of the tap: (dy, dx) in 0..2
                    const uint32_t a_tap = a_base + (uint32_t)(dy * HW + dx) * 16u;
#pragma unroll
                    for (int j = 0; j < 4; j++) {              // K = 64 input channels = 4 x 16
                        const uint64_t bd = smem_desc(b_base + tap * B_TAP + j * 2 * 1024, 1024, 128);
#pragma unroll
                        for (int h = 0; h < 2; h++) {          // left / right 8 columns of the 16-wide tile
                            const uint64_t ad = smem_desc(a_tap + h * 8 * 16 + j * 2 * CH_STRIDE, CH_STRIDE, HW * 16);
                            mma_ss(tmem_base + b * 128 + h * 64, ad, bd, IDESC, (tap | j) != 0);
                        }
                    }
                }
                mma_commit(BAR(3 + s));    // operand slot reusable once these MMAs have read it
                mma_commit(BAR(6 + b));    // accumulators complete -> epilogue
            }
        }
        __syncwarp();
    } else {
        // ===== epilogue =====
        uint8_t *stg = smem + OFF_EPI + warp * EPI_WARP;
        float acc0[8], acc1[8];
        EpiCoef cf;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            acc0[j] = acc1[j] = 0.0f;
            const int c = (lane & 7) * 8 + j;
            cf.sc[j] = EPI == EPI_DGRAD_BN ? a.p_scale[c] : 1.0f;
            cf.sh[j] = EPI == EPI_DGRAD_BN ? a.p_shift[c] : 0.0f;
            cf.mu[j] = EPI == EPI_DGRAD_BN ? a.p_mean[c] : 0.0f;
            cf.is[j] = EPI == EPI_DGRAD_BN ? a.p_invstd[c] : 1.0f;
        }
        int it = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, it++) {
            const int b = it & 1;
            const int ty = (tile / a.tiles_x) * TH, tx = (tile % a.tiles_x) * TW;
            mbar_wait(BAR(6 + b), (it >> 1) & 1);
            fence_after_sync();
#pragma unroll 1
            for (int h = 0; h < 2; h++)
                epilogue_half<EPI>(a, tmem_base + ((uint32_t)(warp * 32) << 16) + b * 128 + h * 64, stg, lane, warp, ty,
                                   tx + h * 8, acc0, acc1, cf);
            fence_before_sync();
            mbar_arrive(BAR(8 + b));
        }
        if ((EPI == EPI_STATS || EPI == EPI_DGRAD_BN) && a.stats) {
            // lanes l, l+8, l+16, l+24 hold the same 8 channels
#pragma unroll
            for (int j = 0; j < 8; j++) {
                acc0[j] += __shfl_xor_sync(0xffffffffu, acc0[j], 8);
                acc0[j] += __shfl_xor_sync(0xffffffffu, acc0[j], 16);
                acc1[j] += __shfl_xor_sync(0xffffffffu, acc1[j], 8);
                acc1[j] += __shfl_xor_sync(0xffffffffu, acc1[j], 16);
            }
            if (lane < 8) {
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    atomicAdd(a.stats + lane * 8 + j, (double)acc0[j]);
                    atomicAdd(a.stats + 64 + lane * 8 + j, (double)acc1[j]);
                }
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, TMEM_COLS);
}

// Weight layout transform: PyTorch conv weight [oc][ic][3][3] fp32 -> the smem image of B for
//   forward (transpose = 0): B_tap[n = oc][k = ic] = w[oc][ic][tap]
//   dgrad   (transpose = 1): B_tap[n = ic][k = oc] = w[oc][ic][8 - tap]   (flipped filter, transposed channels)
// image order: [tap][k chunk of 8][n][8 k] bf16.
__global__ void k_prep_w64(const float *__restrict__ w, bf16 *__restrict__ out, int transpose)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 9 * 64 * 64) return;
    const int kk = idx & 7, n = (idx >> 3) & 63, kc = (idx >> 9) & 7, tap = idx >> 12;
    const int k = kc * 8 + kk;
    float v;
    if (!transpose) v = w[((size_t)n * 64 + k) * 9 + tap];
    else v = w[((size_t)k * 64 + n) * 9 + (8 - tap)];
    out[idx] = __float2bfloat16(v);
}

template <int PRO, int EPI>
static int launch_conv64(const ConvArgs &a, cudaStream_t st)
{
    static bool configured = false;
    if (!configured) {
        F2F_CUDA_CHECK(cudaFuncSetAttribute(k_conv64<PRO, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        configured = true;
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int ntiles = a.tiles_x * a.tiles_y;
    const int grid = ntiles < sms ? ntiles : sms;
    prof_begin(0, st);
    k_conv64<PRO, EPI><<<grid, NTHREADS, SMEM_BYTES, st>>>(a);
    prof_end(0, st);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

}  // namespace dn
}  // namespace f2f

using namespace f2f;
using namespace f2f::dn;

extern "C" {

void f2f_profile_dncnn(int on)
{
    for (auto &p : g_prof_cls) p.used = 0;
    g_prof_on = on != 0;
}

void f2f_profile_dncnn_read(double out[4])
{
    for (int c = 0; c < 2; c++) {
        ProfClass &p = g_prof_cls[c];
        double ms = 0.0;
        for (size_t i = 0; i + 1 < p.used; i += 2) {
            cudaEventSynchronize(p.ev[i + 1]);
            float t = 0.f;
            if (cudaEventElapsedTime(&t, p.ev[i], p.ev[i + 1]) == cudaSuccess) ms += t;
        }
        out[2 * c] = ms;
        out[2 * c + 1] = (double)(p.used / 2);
    }
}

int f2f_dncnn_prep_w64(const float *d_w, void *d_wprep, int transpose, void *stream)
{
    if (!d_w || !d_wprep) { set_error("f2f_dncnn_prep_w64: bad arguments"); return F2F_EINVAL; }
    k_prep_w64<<<cdiv(9 * 64 * 64, 256), 256, 0, (cudaStream_t)stream>>>(d_w, (bf16 *)d_wprep, transpose);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

int f2f_dncnn_conv64(const void *d_in, void *d_out, const void *d_wprep, const float *d_scale, const float *d_shift,
                     double *d_stats, int H, int W, void *stream)
{
    if (!d_in || !d_out || !d_wprep || H < 1 || W < 1) { set_error("f2f_dncnn_conv64: bad arguments"); return F2F_EINVAL; }
    ConvArgs a = {};
    a.in = (const bf16 *)d_in; a.out = (bf16 *)d_out; a.wprep = (const bf16 *)d_wprep;
    a.scale = d_scale; a.shift = d_shift; a.stats = d_stats;
    a.H = H; a.W = W; a.tiles_x = cdiv(W, TW); a.tiles_y = cdiv(H, TH);
    cudaStream_t st = (cudaStream_t)stream;
    const bool bn = d_scale != nullptr;
    if (bn && d_stats) return launch_conv64<PRO_BN_RELU, EPI_STATS>(a, st);
    if (bn) return launch_conv64<PRO_BN_RELU, EPI_STORE>(a, st);
    if (d_stats) return launch_conv64<PRO_RAW, EPI_STATS>(a, st);
    return launch_conv64<PRO_RAW, EPI_STORE>(a, st);
}

int f2f_dncnn_dgrad64(const void *d_dz, void *d_gout, const void *d_wprep_t, const void *d_zprev, const float *d_pscale,
                      const float *d_pshift, const float *d_pmean, const float *d_pinvstd, double *d_stats, int H, int W,
                      void *stream)
{
    if (!d_dz || !d_gout || !d_wprep_t || !d_zprev || H < 1 || W < 1) { set_error("f2f_dncnn_dgrad64: bad arguments"); return F2F_EINVAL; }
    ConvArgs a = {};
    a.in = (const bf16 *)d_dz; a.out = (bf16 *)d_gout; a.wprep = (const bf16 *)d_wprep_t;
    a.zprev = (const bf16 *)d_zprev; a.p_scale = d_pscale; a.p_shift = d_pshift; a.p_mean = d_pmean; a.p_invstd = d_pinvstd;
    a.stats = d_stats;
    a.H = H; a.W = W; a.tiles_x = cdiv(W, TW); a.tiles_y = cdiv(H, TH);
    cudaStream_t st = (cudaStream_t)stream;
    if (d_pscale) return launch_conv64<PRO_RAW, EPI_DGRAD_BN>(a, st);
    return launch_conv64<PRO_RAW, EPI_DGRAD_RELU>(a, st);
}

}  // extern "C"
