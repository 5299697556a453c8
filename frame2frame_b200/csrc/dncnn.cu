// dncnn.cu -- DnCNN-17 (3x3 / 64-channel conv + BatchNorm + ReLU stack) forward
// and backward for sm_100a.  Replaces the cuDNN/ATen calls behind
// lib/frame2frame/dncnn/net.py:6-39 as driven by blind_denoising.py:211-230.
//
// The 64->64 convolutions are implicit GEMMs on the 5th-generation tensor cores:
// tcgen05.mma (kind::f16, BF16 x BF16 -> FP32) issued by one thread per CTA,
// accumulators in TMEM, weights staged once per CTA by the TMA engine
// (cp.async.bulk), activations in NHWC BF16 (one pixel = 64 channels = 128 B).
//
// Train-mode BatchNorm needs the per-channel batch statistics of a layer's
// whole output before it can be normalised, so conv(k) stores the
// raw z_k (16-bit) and reduces sum / sum-of-squares in its epilogue, and the
// normalise + ReLU is applied by conv(k+1) on its operand tile in shared memory
// -- no separate pass over the activations.  The kernel is dnt::k_conv64t below.
#include <cuda.h>
#include <stdlib.h>
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

int g_conv_pair = getenv("F2F_CONV_PAIR") ? atoi(getenv("F2F_CONV_PAIR")) : 0;   // CTA-pair (cta_group::2) convolutions, experimental
int g_pdl = getenv("F2F_PDL") ? atoi(getenv("F2F_PDL")) : 3;   // programmatic dependent launch, see f2f_common.cuh
int g_serpentine = getenv("F2F_SERPENTINE") ? atoi(getenv("F2F_SERPENTINE")) : 1, g_tile_rev = 0;   // f2f_common.cuh

namespace dn {

using namespace tc05;
typedef tc05::h16 h16;

enum Prologue { PRO_RAW = 0, PRO_BN_RELU = 1 };
enum Epilogue { EPI_STORE = 0, EPI_STATS = 1, EPI_DGRAD_BN = 2, EPI_DGRAD_RELU = 3, EPI_F32 = 4 };

struct ConvArgs {
    const h16 *in;        // [H][W][64]
    h16 *out;             // [H][W][64]
    const h16 *wprep;     // [9 taps][8 k-chunks][64 n][8 k] BF16 = the smem image of B
    const float *scale;    // prologue BN: a = relu(in*scale + shift), per input channel
    const float *shift;
    double *stats;         // [2][64] accumulators (EPI_STATS: sum z, sum z^2; EPI_DGRAD_BN: sum g, sum g*xhat)
    const h16 *zprev;     // EPI_DGRAD_*: pre-activation (BN) or activation (RELU) of the layer below
    const float *p_scale;  // EPI_DGRAD_BN: BN of the layer below: y = z*p_scale + p_shift, xhat = (z - p_mean)*p_invstd
    const float *p_shift;
    const float *p_mean;
    const float *p_invstd;
    int H, W, tiles_x, tiles_y;
    float *out32;          // EPI_F32: FP32 accumulators go to [H][W][64] floats: out32 = (oacc ? out32 : 0) + oscale * acc
    float oscale;
    int oacc;
    unsigned long long *trace;   // f2f_dncnn_set_trace: per-role cycle counters (development aid) or NULL
    int pdl_trigger;             // release the dependent launch early (f2f_common.cuh)
    int rev;                     // tile order: 0 = top-down, 1 = bottom-up (L2 reuse between consecutive launches)
};

unsigned long long *g_conv_trace = nullptr;

// per-channel coefficients an epilogue thread keeps for its 8 channels
struct EpiCoef { float sc[8], sh[8], mu[8], is[8]; };


}  // namespace dn

// =====================================================================================
// dnt -- the TMA-fed variant of the 64->64 convolution, used whenever the operand needs no
// transform on load (every input-gradient convolution, and the first 64->64 layer).
//
// What the per-role cycle trace (f2f_dncnn_set_trace) and tools/ubench_mma.cu showed about the
// kernel above on B200: (1) an M128 N64 K16 MMA is bound by the shared-memory read of its
// operands (6 KB per instruction at 128 B/clk = 48 cycles, floor 32), and the SAME array serves
// every st.shared / ld.shared / global-load fill of the CTA, so bytes through shared memory per
// pixel, not tensor throughput, set the time; (2) the epilogue, not the MMA, was the critical
// path of the input-gradient launches: four warps waiting on the global loads of the layer
// below's pre-activation after every accumulator hand-over.
//
// Hence: tile = 16 rows x 8 pixels = ONE M128 block (accumulator = 64 TMEM columns,
// double-buffered); the haloed 18x10 operand tile is written by one TMA tensor load in the
// 128-byte-swizzled pixel-major layout (no register round trip, no st.shared, zero fill at the
// image border for free), taps are shifted views of it; the pre-activation tile the epilogue
// needs arrives the same way, two tiles ahead; eight epilogue warps drain two tiles at once.
//
//   warps 0-3 / 4-7  epilogue of even / odd tiles (TMEM lane quarter = warp % 4)
//   warp 8           MMA issuer (36 tcgen05.mma per tile) + TMEM owner
//   warp 9 / 10      TMA producer of the operand tiles / of the pre-activation tiles
namespace dnt {

using namespace tc05;
using dn::ConvArgs;
using dn::EpiCoef;
using dn::EPI_STORE;
using dn::EPI_STATS;
using dn::EPI_DGRAD_BN;
using dn::EPI_DGRAD_RELU;
using dn::EPI_F32;
typedef tc05::h16 h16;

constexpr int TW = 8, TH = 16;
constexpr int HW = TW + 2, HH = TH + 2, HPIX = HW * HH;   // 10 x 18 = 180 pixels
constexpr int A_TX = HPIX * 128;                           // bytes one operand load delivers
constexpr int A_STAGE = 23 * 1024;                         // 23040 rounded up to the swizzle period
constexpr int NA_MAX = 5;   // operand ring: 3 stages next to the pre-activation ring (input gradient), else 5
constexpr int Z_STAGE = TW * TH * 128;                     // 16384
constexpr int NZ = 3;
constexpr int B_TAP = 64 * 64 * 2, B_BYTES = 9 * B_TAP;
constexpr int EPI_ROW = 144, EPI_WARP = 32 * EPI_ROW, NEPI = 8;
constexpr int OFF_B = 0;
constexpr int OFF_A = OFF_B + B_BYTES;
constexpr int OFF_Z = OFF_A + 3 * A_STAGE;
constexpr int OFF_EPI = OFF_Z + NZ * Z_STAGE;
constexpr int OFF_BAR = OFF_EPI + NEPI * EPI_WARP;
constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;           // + slack to align the base to 1024 B
constexpr int NTHREADS_RAW = 11 * 32, NTHREADS_BN = 15 * 32;   // + four transform warps when the operand needs BN+ReLU
constexpr int MMA_WARP = 8, TMA_A_WARP = 9, TMA_Z_WARP = 10, XF_WARP0 = 11;
constexpr int TMEM_COLS = 128;
// barrier slots
constexpr int B_FULL_A = 0, B_EMPTY_A = 5, B_RAW_A = 10, B_FULL_Z = 15, B_EMPTY_Z = 18, B_TFULL = 21, B_TEMPTY = 23, B_W = 25, B_WPEER = 26;
static_assert(NA_MAX * A_STAGE <= 3 * A_STAGE + NZ * Z_STAGE, "the 5-stage operand ring reuses the pre-activation ring's space");
static_assert(SMEM_BYTES <= 232448, "shared memory budget");
static_assert(OFF_A % 1024 == 0 && OFF_Z % 1024 == 0 && A_STAGE % 1024 == 0, "swizzle alignment");

// One warp, its 32 pixels (4 rows x 8) of a tile: TMEM row -> BF16 -> smem transpose -> in the
// transposed phase a lane owns chunk lane%8 (8 channels) of pixels q = lane/8 + 4i, so per-channel
// sums are per-thread accumulators, and reads the layer below's pre-activation from the TMA-fed
// tile `zst` (128-byte rows, chunk c of pixel p at chunk c ^ (p & 7)).
template <int EPI>
__device__ __forceinline__ void epilogue_tile(const ConvArgs &a, uint32_t tmem_addr, uint8_t *stg, const uint8_t *zst,
                                              uint32_t zbar, uint32_t zphase, int lane, int wq, int ty, int tx,
                                              float (&acc0)[8], float (&acc1)[8], const EpiCoef &cf)
{
    const int chunk = lane & 7;
    if (EPI == EPI_F32) {
        // split-operand ("precise 2") passes: the FP32 accumulators themselves are the result.  Accumulator row
        // m = 32 wq + lane is pixel (row m / 8, column m % 8) of the tile; a lane owns its pixel's 64 floats.
        const int m = wq * 32 + lane;
        const int gy = ty + (m >> 3), gx = tx + (m & 7);
        const bool inb = gy < a.H && gx < a.W;
        float4 *dst = reinterpret_cast<float4 *>(a.out32 + ((size_t)gy * a.W + gx) * 64);
#pragma unroll
        for (int part = 0; part < 2; part++) {
            uint32_t r[32];
            tmem_ld32(tmem_addr + part * 32, r);
            tmem_ld_wait();
            if (inb) {
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    float4 v = make_float4(a.oscale * __uint_as_float(r[4 * q]), a.oscale * __uint_as_float(r[4 * q + 1]),
                                           a.oscale * __uint_as_float(r[4 * q + 2]), a.oscale * __uint_as_float(r[4 * q + 3]));
                    if (a.oacc) {
                        const float4 o = dst[part * 8 + q];
                        v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
                    }
                    dst[part * 8 + q] = v;
                }
            }
        }
        return;
    }
    uint8_t *myrow = stg + lane * EPI_ROW;
#pragma unroll
    for (int part = 0; part < 2; part++) {
        uint32_t r[32];
        tmem_ld32(tmem_addr + part * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; j++) {
            uint4 o;
            o.x = pack_h16(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1]));
            o.y = pack_h16(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
            o.z = pack_h16(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
            o.w = pack_h16(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
            *reinterpret_cast<uint4 *>(myrow + part * 64 + j * 16) = o;
        }
    }
    __syncwarp();
    if (EPI == EPI_DGRAD_BN || EPI == EPI_DGRAD_RELU) mbar_wait(zbar, zphase);
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int q = (lane >> 3) + 4 * i;           // pixel within the warp's 32
        const int gy = ty + wq * 4 + (q >> 3), gx = tx + (q & 7);
        const bool inb = gy < a.H && gx < a.W;
        uint4 o = *reinterpret_cast<const uint4 *>(stg + q * EPI_ROW + chunk * 16);
        uint32_t w[4] = {o.x, o.y, o.z, o.w};
        if (EPI == EPI_DGRAD_BN || EPI == EPI_DGRAD_RELU) {
            // ReLU mask of the layer below (and its BatchNorm-backward sums), net.py:16-18 backward
            const int zp = wq * 32 + q;              // pixel within the tile
            const uint4 zv = *reinterpret_cast<const uint4 *>(zst + zp * 128 + ((chunk ^ (zp & 7)) << 4));
            const uint32_t zw[4] = {zv.x, zv.y, zv.z, zv.w};
#pragma unroll
            for (int j = 0; j < 4; j++) {
                float2 g = unpack_h16(w[j]);
                const float2 z = unpack_h16(zw[j]);
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
                w[j] = pack_h16(g.x, g.y);
            }
            o = make_uint4(w[0], w[1], w[2], w[3]);
        }
        if (EPI == EPI_STATS && inb) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float2 v = unpack_h16(w[j]);
                acc0[2 * j] += v.x; acc0[2 * j + 1] += v.y;
                acc1[2 * j] = fmaf(v.x, v.x, acc1[2 * j]); acc1[2 * j + 1] = fmaf(v.y, v.y, acc1[2 * j + 1]);
            }
        }
        if (inb) *reinterpret_cast<uint4 *>(a.out + ((size_t)gy * a.W + gx) * 64 + chunk * 8) = o;
    }
    __syncwarp();
}

template <int PRO, int EPI, bool PAIR>
__global__ void __launch_bounds__(PRO == dn::PRO_BN_RELU ? NTHREADS_BN : NTHREADS_RAW, 1)
k_conv64t(ConvArgs a, const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ CUtensorMap tm_z)
{
    constexpr bool XF = PRO == dn::PRO_BN_RELU;
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte aligned base: the swizzle phase of a TMA-written row is then (row & 7)
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int ntiles = a.tiles_x * a.tiles_y;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + OFF_BAR);
    const uint32_t bar0 = smem_u32(bars);
    auto BAR = [&](int i) { return bar0 + 8u * i; };
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(bars + 27);
    volatile long long *tr_issue_t0 = reinterpret_cast<volatile long long *>(bars + 28);   // trace: issue start of the even tiles
    constexpr bool DGRAD = EPI == EPI_DGRAD_BN || EPI == EPI_DGRAD_RELU;
    constexpr int NA = DGRAD ? 3 : NA_MAX;
    // PAIR: this CTA and its neighbour (cluster of 2, one TPC) run their tiles in lock-step as ONE M = 256
    // cta_group::2 instruction stream issued by the leader (rank 0); each supplies its own operand tile and HALF of
    // the weights (32 of the 64 output channels), so an instruction reads 4 + 1 KB of shared memory per SM instead of
    // 4 + 2 KB.  Both run the leader's number of iterations; a tile index past the end is an all-zero tile (TMA zero
    // fill) whose outputs are masked like any pixel outside the frame.
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    const bool leader = rank == 0;
    const int first_tile = PAIR ? (int)(blockIdx.x & ~1u) : (int)blockIdx.x;
    const int nit = first_tile < ntiles ? (ntiles - first_tile + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    // a.rev: walk the image bottom-up.  Consecutive launches of the layer chain alternate the direction, so that a
    // launch starts on the part of its input the previous launch touched LAST -- still in the 126 MB L2 -- instead of
    // the part evicted first (a 1080p activation is 265 MB).
    auto tile_of = [&](int t) { return (!PAIR && a.rev) ? ntiles - 1 - t : t; };

    if (threadIdx.x == 0) {
        for (int s = 0; s < NA; s++) {
            mbar_init(BAR(B_FULL_A + s), XF ? (PAIR ? 8 : 128) : 1); mbar_init(BAR(B_EMPTY_A + s), 1); mbar_init(BAR(B_RAW_A + s), 1);
        }
        for (int s = 0; s < NZ; s++) { mbar_init(BAR(B_FULL_Z + s), 1); mbar_init(BAR(B_EMPTY_Z + s), 4); }
        for (int b = 0; b < 2; b++) { mbar_init(BAR(B_TFULL + b), 1); mbar_init(BAR(B_TEMPTY + b), PAIR ? 8 : 128); }
        mbar_init(BAR(B_W), 1);
        mbar_init(BAR(B_WPEER), 1);
        mbar_fence_init();
    }
    if (warp == MMA_WARP) {
        if (PAIR) tmem_alloc_pair(smem_u32(tmem_ptr), TMEM_COLS);
        else tmem_alloc(smem_u32(tmem_ptr), TMEM_COLS);
    }
    fence_before_sync();
    __syncthreads();
    if (PAIR) cluster_sync_all();          // the peer's barriers are initialised before anything arrives on them
    fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;
    pdl_wait();                  // everything above ran under the predecessor's tail (f2f_common.cuh)
    if (a.pdl_trigger) pdl_launch_dependents();

    if (warp == TMA_A_WARP) {
        // ===== operand tiles: one tensor load per tile (18 rows x 10 pixels x 128 B, zero fill outside) =====
        if (elect_one()) {
            tma_prefetch_desc(&tm_in);
            const bool tr = a.trace != nullptr;
            long long tr_wait = 0;
            for (int it = 0; it < nit; it++) {
                const int tile = tile_of(blockIdx.x + it * gridDim.x);
                const int s = it % NA;
                const int ty0 = (tile / a.tiles_x) * TH - 1, tx0 = (tile % a.tiles_x) * TW - 1;
                const long long t0 = tr ? clock64() : 0;
                mbar_wait(BAR(B_EMPTY_A + s), ((it / NA) & 1) ^ 1);
                if (tr) tr_wait += clock64() - t0;
                const uint32_t landed = BAR((XF ? B_RAW_A : B_FULL_A) + s);   // with a transform, the MMA waits for IT
                if (PAIR && !XF) {
                    // both CTAs' bytes are counted on the LEADER's barrier, which its MMA warp waits on
                    if (leader) mbar_arrive_expect_tx(landed, 2 * A_TX);
                    tma_load_3d_pair(smem_u32(smem + OFF_A + s * A_STAGE), &tm_in, mapa_u32(landed, 0), 0, tx0, ty0);
                } else {
                    mbar_arrive_expect_tx(landed, A_TX);
                    tma_load_3d(smem_u32(smem + OFF_A + s * A_STAGE), &tm_in, landed, 0, tx0, ty0);
                }
            }
            if (tr) atomicAdd(a.trace + 0, (unsigned long long)tr_wait);
        }
        __syncwarp();
    } else if (warp == TMA_Z_WARP) {
        // ===== the layer below's pre-activation for the epilogue (16 rows x 8 pixels x 128 B) =====
        if (DGRAD && elect_one()) {
            tma_prefetch_desc(&tm_z);
            for (int it = 0; it < nit; it++) {
                const int tile = tile_of(blockIdx.x + it * gridDim.x);
                const int s = it % NZ;
                const int ty0 = (tile / a.tiles_x) * TH, tx0 = (tile % a.tiles_x) * TW;
                mbar_wait(BAR(B_EMPTY_Z + s), ((it / NZ) & 1) ^ 1);
                mbar_arrive_expect_tx(BAR(B_FULL_Z + s), Z_STAGE);
                tma_load_3d(smem_u32(smem + OFF_Z + s * Z_STAGE), &tm_z, BAR(B_FULL_Z + s), 0, tx0, ty0);
            }
        }
        __syncwarp();
    } else if (XF && warp >= XF_WARP0) {
        // ===== operand transform, in place: a = relu(z * scale + shift), the BatchNorm + ReLU of the layer below
        // (net.py:17-18).  Pixels outside the image stay 0: zero padding applies to the ACTIVATION. =====
        const int xt = threadIdx.x - XF_WARP0 * 32, chunk = xt & 7;       // 128 threads; a thread keeps one channel chunk
        float sc[8], sh[8];
#pragma unroll
        for (int j = 0; j < 8; j++) { sc[j] = a.scale[chunk * 8 + j]; sh[j] = a.shift[chunk * 8 + j]; }
        for (int it = 0; it < nit; it++) {
            const int tile = tile_of(blockIdx.x + it * gridDim.x);
            const int s = it % NA;
            const int ty0 = (tile / a.tiles_x) * TH - 1, tx0 = (tile % a.tiles_x) * TW - 1;
            uint8_t *st = smem + OFF_A + s * A_STAGE;
            mbar_wait(BAR(B_RAW_A + s), (it / NA) & 1);
            constexpr int ROUNDS = (HPIX + 15) / 16;                      // 16 pixels per round of 128 threads
            if (ty0 >= 0 && tx0 >= 0 && ty0 + HH <= a.H && tx0 + HW <= a.W) {
                // the haloed tile lies inside the image (all tiles but the frame's border; warp-uniform): no bounds
                // arithmetic, and since 16 pixels further down the tile the swizzle phase (pix & 7) is the same, every
                // element of the thread sits at a compile-time offset from one per-thread address
                uint8_t *mine = st + (xt >> 3) * 128 + ((chunk ^ ((xt >> 3) & 7)) << 4);
                constexpr int FULL = HPIX / 16;                           // rounds every thread takes part in
#pragma unroll
                for (int u = 0; u < ROUNDS; u++) {
                    if (u < FULL || (xt >> 3) + 16 * u < HPIX) {
                        const uint4 q = *reinterpret_cast<const uint4 *>(mine + u * 2048);
                        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
                        uint32_t o[4];
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const float2 r = ffma2(unpack_h16(w[j]), make_float2(sc[2 * j], sc[2 * j + 1]), make_float2(sh[2 * j], sh[2 * j + 1]));
                            o[j] = pack_h16_relu(r.x, r.y);
                        }
                        *reinterpret_cast<uint4 *>(mine + u * 2048) = make_uint4(o[0], o[1], o[2], o[3]);
                    }
                }
                fence_proxy_async();
                if (PAIR) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(mapa_u32(BAR(B_FULL_A + s), 0));
                } else {
                    mbar_arrive(BAR(B_FULL_A + s));
                }
                continue;
            }
            uint4 v[ROUNDS];
            uint32_t ok = 0;
#pragma unroll
            for (int u = 0; u < ROUNDS; u++) {
                const int pix = (xt >> 3) + 16 * u;
                const int gy = ty0 + pix / HW, gx = tx0 + pix % HW;
                if (pix < HPIX && gy >= 0 && gy < a.H && gx >= 0 && gx < a.W) {
                    ok |= 1u << u;
                    v[u] = *reinterpret_cast<const uint4 *>(st + pix * 128 + ((chunk ^ (pix & 7)) << 4));
                }
            }
#pragma unroll
            for (int u = 0; u < ROUNDS; u++) {
                if ((ok >> u) & 1u) {
                    const int pix = (xt >> 3) + 16 * u;
                    uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const float2 f = unpack_h16(w[j]);
                        w[j] = pack_h16_relu(fmaf(f.x, sc[2 * j], sh[2 * j]), fmaf(f.y, sc[2 * j + 1], sh[2 * j + 1]));
                    }
                    *reinterpret_cast<uint4 *>(st + pix * 128 + ((chunk ^ (pix & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
                }
            }
            fence_proxy_async();                                          // st.shared -> visible to the tensor core's reads
            if (PAIR) {
                // one arrival per warp: 128 remote arrivals per tile serialise on the inter-SM path (measured)
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(mapa_u32(BAR(B_FULL_A + s), 0));
            } else {
                mbar_arrive(BAR(B_FULL_A + s));
            }
        }
    } else if (warp == MMA_WARP) {
        // ===== MMA issuer: the warp stays converged, the elected lane issues =====
        const uint32_t el = elect_one();
        if (el) {
            if (PAIR) {
                // this CTA's half of the weights: output channels 32*rank .. 32*rank+31 of every 8-channel K chunk
                // ([tap][ic/8][oc][8] in wprep), placed at rows 0..31 of the same no-swizzle layout
                mbar_arrive_expect_tx(BAR(B_W), B_BYTES / 2);
                for (int c = 0; c < 72; c++)
                    bulk_g2s(smem_u32(smem + OFF_B + c * 1024), a.wprep + (size_t)c * 512 + rank * 256, 512, BAR(B_W));
            } else {
                mbar_arrive_expect_tx(BAR(B_W), B_BYTES);
                for (int t = 0; t < 9; t++) bulk_g2s(smem_u32(smem + OFF_B + t * B_TAP), a.wprep + (size_t)t * 4096, B_TAP, BAR(B_W));
            }
        }
        __syncwarp();
        mbar_wait(BAR(B_W), 0);
        if (PAIR) {
            if (leader) mbar_wait_cluster(BAR(B_WPEER), 0);                       // the peer's half has landed too
            else if (el) mbar_arrive_cluster(mapa_u32(BAR(B_WPEER), 0));
            __syncwarp();
        }
        constexpr uint32_t IDESC = PAIR ? idesc_h16(256, 64, 0, 0) : idesc_h16(128, 64, 0, 0);
        const uint64_t bd0 = smem_desc(smem_u32(smem + OFF_B), 1024, 128);
        const uint32_t b_lo = (uint32_t)bd0, b_hi = (uint32_t)(bd0 >> 32);
        const uint64_t ad0 = smem_desc_sw128(0, HW * 128);
        const uint32_t a_lo0 = (uint32_t)ad0, a_hi = (uint32_t)(ad0 >> 32);
        const bool tr = a.trace != nullptr && el;
        long long tr_te = 0, tr_full = 0, tr_issue = 0;
        // SM cycles against wall time over the tile loop of CTA 0: the clock the kernel actually ran at
        // (back-to-back tensor kernels run under the power cap, an isolated launch does not)
        const long long tr_c0 = tr ? clock64() : 0;
        unsigned long long tr_g0 = 0;
        if (tr) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr_g0));
        int it = 0;
        for (; it < (leader ? nit : 0); it++) {                 // the peer's MMA warp only hands over its weights
            const int s = it % NA, b = it & 1;
            const long long t0 = tr ? clock64() : 0;
            if (PAIR) mbar_wait_cluster(BAR(B_TEMPTY + b), ((it >> 1) & 1) ^ 1);
            else mbar_wait(BAR(B_TEMPTY + b), ((it >> 1) & 1) ^ 1);   // accumulator drained by the epilogue
            const long long t1 = tr ? clock64() : 0;
            if (PAIR) mbar_wait_cluster(BAR(B_FULL_A + s), (it / NA) & 1);
            else mbar_wait(BAR(B_FULL_A + s), (it / NA) & 1);         // operand tile landed
            const long long t2 = tr ? clock64() : 0;
            if (tr && b == 0) *tr_issue_t0 = t2;
            fence_after_sync();
            const uint32_t a_lo = a_lo0 | ((smem_u32(smem + OFF_A + s * A_STAGE) & 0x3FFFFu) >> 4);
            const uint32_t d0 = tmem_base + b * 64;
#pragma unroll
            for (int tap = 0; tap < 9; tap++) {
                const int dy = tap / 3, dx = tap % 3;            // halo coordinates of the tap
#pragma unroll
                for (int j = 0; j < 4; j++) {                    // K = 64 input channels = 4 x 16
                    const uint64_t bd = ((uint64_t)b_hi << 32) | (b_lo + (tap * B_TAP + j * 2 * 1024) / 16);
                    const uint64_t ad = ((uint64_t)a_hi << 32) | (a_lo + (dy * HW + dx) * 8 + j * 2);
                    if (PAIR) mma_ss_el_pair(d0, ad, bd, IDESC, (tap | j) != 0, el);
                    else mma_ss_el(d0, ad, bd, IDESC, (tap | j) != 0, el);
                }
            }
            if (el) {
                if (PAIR) {
                    mma_commit_pair(BAR(B_EMPTY_A + s));
                    mma_commit_pair(BAR(B_TFULL + b));
                } else {
                    mma_commit(BAR(B_EMPTY_A + s));
                    mma_commit(BAR(B_TFULL + b));
                }
            }
            __syncwarp();
            if (tr) { tr_te += t1 - t0; tr_full += t2 - t1; tr_issue += clock64() - t2; }
        }
        if (tr) {
            atomicAdd(a.trace + 2, (unsigned long long)tr_te); atomicAdd(a.trace + 3, (unsigned long long)tr_full);
            atomicAdd(a.trace + 4, (unsigned long long)tr_issue); atomicAdd(a.trace + 8, (unsigned long long)it);
            if (blockIdx.x == 0) {
                unsigned long long g1;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
                atomicAdd(a.trace + 10, (unsigned long long)(clock64() - tr_c0));
                atomicAdd(a.trace + 11, g1 - tr_g0);
            }
        }
    } else {
        // ===== epilogue: warps 0-3 take even tiles (accumulator 0), warps 4-7 odd tiles (accumulator 1) =====
        const int set = warp >> 2, wq = warp & 3;
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
        const bool tr = a.trace != nullptr && threadIdx.x == 0;
        long long tr_wait = 0, tr_work = 0;
        const long long tr_start = tr ? clock64() : 0;
        int ntr = 0;
        for (int it = set; it < nit; it += 2) {
            const int tile = tile_of(blockIdx.x + it * gridDim.x);
            const int sz = it % NZ;
            const int ty = (tile / a.tiles_x) * TH, tx = (tile % a.tiles_x) * TW;
            const long long t0 = tr ? clock64() : 0;
            mbar_wait(BAR(B_TFULL + set), (it >> 1) & 1);
            const long long t1 = tr ? clock64() : 0;
            if (tr && leader) {      // first MMA of the tile issued -> its completion seen here (same SM, same clock)
                atomicAdd(a.trace + 12, (unsigned long long)(t1 - *tr_issue_t0)); atomicAdd(a.trace + 13, 1ull);
            }
            fence_after_sync();
            epilogue_tile<EPI>(a, tmem_base + ((uint32_t)(wq * 32) << 16) + set * 64, stg, smem + OFF_Z + sz * Z_STAGE,
                               BAR(B_FULL_Z + sz), (it / NZ) & 1, lane, wq, ty, tx, acc0, acc1, cf);
            if (DGRAD && lane == 0) mbar_arrive(BAR(B_EMPTY_Z + sz));
            fence_before_sync();
            if (PAIR) {
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(mapa_u32(BAR(B_TEMPTY + set), 0));
            } else {
                mbar_arrive(BAR(B_TEMPTY + set));
            }
            if (tr) { tr_wait += t1 - t0; tr_work += clock64() - t1; ntr++; }
        }
        if (tr) {
            atomicAdd(a.trace + 5, (unsigned long long)tr_wait); atomicAdd(a.trace + 6, (unsigned long long)tr_work);
            atomicAdd(a.trace + 7, (unsigned long long)(clock64() - tr_start)); atomicAdd(a.trace + 9, (unsigned long long)ntr);
        }
        if ((EPI == EPI_STATS || EPI == EPI_DGRAD_BN) && a.stats) {
#pragma unroll
            for (int j = 0; j < 8; j++) {
                acc0[j] += __shfl_xor_sync(0xffffffffu, acc0[j], 8);
                acc0[j] += __shfl_xor_sync(0xffffffffu, acc0[j], 16);
                acc1[j] += __shfl_xor_sync(0xffffffffu, acc1[j], 8);
                acc1[j] += __shfl_xor_sync(0xffffffffu, acc1[j], 16);
            }
            // The eight epilogue warps' sums are combined in shared memory first and the CTA issues ONE atomic per
            // value: FP64 atomics on the same 128 addresses serialise in L2 (measured: 8 x 148 atomics per address
            // cost 25-35 us at the end of every launch, profiles/r01_conv64_role_trace_clock.log).
            float *red = reinterpret_cast<float *>(stg);          // this warp's staging area, free after its last tile
            if (lane < 8) {
#pragma unroll
                for (int j = 0; j < 8; j++) { red[lane * 8 + j] = acc0[j]; red[64 + lane * 8 + j] = acc1[j]; }
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");        // the 8 epilogue warps only
            if (threadIdx.x < 128) {
                double t = 0.0;
#pragma unroll
                for (int w = 0; w < NEPI; w++)
                    t += (double)reinterpret_cast<const float *>(smem + OFF_EPI + w * EPI_WARP)[threadIdx.x];
                atomicAdd(a.stats + threadIdx.x, t);
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (PAIR) cluster_sync_all();          // neither CTA leaves while the other may still arrive on its barriers
    if (warp == MMA_WARP) {
        if (PAIR) tmem_dealloc_pair(tmem_base, TMEM_COLS);
        else tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// NHWC BF16 image [H][W][64] as a 3-D tensor (channels, x, y), 128-byte swizzle, box = boxw x boxh pixels
int make_tensor_map(CUtensorMap *m, const void *base, int H, int W, int boxw, int boxh)
{
    const cuuint64_t dims[3] = {64, (cuuint64_t)W, (cuuint64_t)H};
    const cuuint64_t strides[2] = {128, (cuuint64_t)W * 128};
    const cuuint32_t box[3] = {64, (cuuint32_t)boxw, (cuuint32_t)boxh};
    const cuuint32_t estr[3] = {1, 1, 1};
    // the driver entry point is resolved through the runtime, so the library carries no link-time
    // dependency on libcuda.so.1 (it must load, and export its symbols, on a machine without a driver)
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn encode = nullptr;
    if (!encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) {
            set_error("cuTensorMapEncodeTiled is not available from the installed driver");
            return F2F_ECUDA;
        }
        encode = (encode_fn)fn;
    }
    const CUresult r = encode(m, (tc05::H16_FORMAT ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16), 3, const_cast<void *>(base), dims, strides, box,
                              estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d) for a %dx%d image at %p", (int)r, W, H, base);
        return F2F_ECUDA;
    }
    return F2F_OK;
}

template <int PRO, int EPI>
static int launch_conv64t(ConvArgs a, cudaStream_t st)
{
    F2F_CUDA_CHECK((ensure_dyn_smem<k_conv64t<PRO, EPI, false>>(SMEM_BYTES)));
    F2F_CUDA_CHECK((ensure_dyn_smem<k_conv64t<PRO, EPI, true>>(SMEM_BYTES)));
    a.tiles_x = cdiv(a.W, TW);
    a.tiles_y = cdiv(a.H, TH);
    CUtensorMap tm_in, tm_z;
    int rc = make_tensor_map(&tm_in, a.in, a.H, a.W, HW, HH);
    if (rc) return rc;
    constexpr bool DGRAD = EPI == EPI_DGRAD_BN || EPI == EPI_DGRAD_RELU;
    rc = make_tensor_map(&tm_z, DGRAD ? (const void *)a.zprev : (const void *)a.in, a.H, a.W, TW, TH);
    if (rc) return rc;
    const int sms = sm_count();
    const int ntiles = a.tiles_x * a.tiles_y;
    const int grid = ntiles < sms ? ntiles : sms;
    const dim3 block(PRO == dn::PRO_BN_RELU ? NTHREADS_BN : NTHREADS_RAW);
    prof_begin(0, st);
    a.pdl_trigger = (g_pdl >> 2) & 1;
    a.rev = g_tile_rev;
    if (g_conv_pair)      // an even number of CTAs: the last pair's second CTA may only have all-zero tiles
        F2F_CUDA_CHECK(launch_pdl_pairs(1, k_conv64t<PRO, EPI, true>, dim3((grid + 1) & ~1), block, SMEM_BYTES, st, a, tm_in, tm_z));
    else
        F2F_CUDA_CHECK(launch_pdl(1, k_conv64t<PRO, EPI, false>, dim3(grid), block, SMEM_BYTES, st, a, tm_in, tm_z));
    prof_end(0, st);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}
}  // namespace dnt

// 3x3 64->64 convolution of a 16-bit image with 16-bit weights whose FP32 accumulators are stored (or added, scaled) to
// an FP32 NHWC image: one term of the split-operand products of dncnn_net.cu's tensor-core precise mode
int conv64_f32(const void *d_in, const void *d_wprep, float *d_out32, float scale, int accumulate, int H, int W, cudaStream_t st)
{
    dn::ConvArgs a = {};
    a.in = (const tc05::h16 *)d_in; a.wprep = (const tc05::h16 *)d_wprep; a.out32 = d_out32; a.oscale = scale; a.oacc = accumulate;
    a.H = H; a.W = W;
    return dnt::launch_conv64t<dn::PRO_RAW, dn::EPI_F32>(a, st);
}

namespace dn {

// Weight layout transform: PyTorch conv weight [oc][ic][3][3] fp32 -> the smem image of B for
//   forward (transpose = 0): B_tap[n = oc][k = ic] = w[oc][ic][tap]
//   dgrad   (transpose = 1): B_tap[n = ic][k = oc] = w[oc][ic][8 - tap]   (flipped filter, transposed channels)
// image order: [tap][k chunk of 8][n][8 k] bf16.
__global__ void k_prep_w64(const float *__restrict__ w, h16 *__restrict__ out, int transpose)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 9 * 64 * 64) return;
    const int kk = idx & 7, n = (idx >> 3) & 63, kc = (idx >> 9) & 7, tap = idx >> 12;
    const int k = kc * 8 + kk;
    float v;
    if (!transpose) v = w[((size_t)n * 64 + k) * 9 + tap];
    else v = w[((size_t)k * 64 + n) * 9 + (8 - tap)];
    out[idx] = tc05::float_to_h16(v);
}

}  // namespace dn
}  // namespace f2f

using namespace f2f;
using namespace f2f::dn;
using f2f::dnt::TW;
using f2f::dnt::TH;

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

void f2f_dncnn_set_trace(unsigned long long *d_counters) { g_conv_trace = d_counters; }
void f2f_set_pdl(int on) { f2f::g_pdl = on; }
void f2f_set_serpentine(int on) { f2f::g_serpentine = on; f2f::g_tile_rev = 0; }
void f2f_set_conv_pair(int on) { f2f::g_conv_pair = on; }
int f2f_dncnn_storage_dtype(void) { return tc05::H16_FORMAT ? 0 : 1; }

int f2f_dncnn_prep_w64(const float *d_w, void *d_wprep, int transpose, void *stream)
{
    if (!d_w || !d_wprep) { set_error("f2f_dncnn_prep_w64: bad arguments"); return F2F_EINVAL; }
    k_prep_w64<<<cdiv(9 * 64 * 64, 256), 256, 0, (cudaStream_t)stream>>>(d_w, (h16 *)d_wprep, transpose);
    g_dncnn_launches++;
    F2F_LAUNCH_CHECK();
    return F2F_OK;
}

int f2f_dncnn_conv64(const void *d_in, void *d_out, const void *d_wprep, const float *d_scale, const float *d_shift,
                     double *d_stats, int H, int W, void *stream)
{
    if (!d_in || !d_out || !d_wprep || H < 1 || W < 1) { set_error("f2f_dncnn_conv64: bad arguments"); return F2F_EINVAL; }
    ConvArgs a = {};
    a.in = (const h16 *)d_in; a.out = (h16 *)d_out; a.wprep = (const h16 *)d_wprep;
    a.scale = d_scale; a.shift = d_shift; a.stats = d_stats;
    a.H = H; a.W = W; a.tiles_x = cdiv(W, TW); a.tiles_y = cdiv(H, TH);
    a.trace = g_conv_trace;
    cudaStream_t st = (cudaStream_t)stream;
    const bool bn = d_scale != nullptr;
    if (bn && d_stats) return f2f::dnt::launch_conv64t<PRO_BN_RELU, EPI_STATS>(a, st);
    if (bn) return f2f::dnt::launch_conv64t<PRO_BN_RELU, EPI_STORE>(a, st);
    if (d_stats) return f2f::dnt::launch_conv64t<PRO_RAW, EPI_STATS>(a, st);
    return f2f::dnt::launch_conv64t<PRO_RAW, EPI_STORE>(a, st);
}

int f2f_dncnn_dgrad64(const void *d_dz, void *d_gout, const void *d_wprep_t, const void *d_zprev, const float *d_pscale,
                      const float *d_pshift, const float *d_pmean, const float *d_pinvstd, double *d_stats, int H, int W,
                      void *stream)
{
    if (!d_dz || !d_gout || !d_wprep_t || !d_zprev || H < 1 || W < 1) { set_error("f2f_dncnn_dgrad64: bad arguments"); return F2F_EINVAL; }
    ConvArgs a = {};
    a.in = (const h16 *)d_dz; a.out = (h16 *)d_gout; a.wprep = (const h16 *)d_wprep_t;
    a.zprev = (const h16 *)d_zprev; a.p_scale = d_pscale; a.p_shift = d_pshift; a.p_mean = d_pmean; a.p_invstd = d_pinvstd;
    a.stats = d_stats;
    a.H = H; a.W = W; a.tiles_x = cdiv(W, TW); a.tiles_y = cdiv(H, TH);
    a.trace = g_conv_trace;
    cudaStream_t st = (cudaStream_t)stream;
    if (d_pscale) return f2f::dnt::launch_conv64t<PRO_RAW, EPI_DGRAD_BN>(a, st);
    return f2f::dnt::launch_conv64t<PRO_RAW, EPI_DGRAD_RELU>(a, st);
}

}  // extern "C"
