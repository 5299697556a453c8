/*
 * f2f_b200.h -- C ABI of libf2f_b200.so, the B200 (sm_100a) implementation of
 * the frame2frame hot path.  Plain pointers and sizes only; no torch types.
 *
 * Conventions
 *  - "d_" pointers are device pointers owned by the caller (e.g. torch
 *    tensors' data_ptr()); "h_" pointers are host pointers.
 *  - `stream` is a cudaStream_t passed as void* (0 = default stream).
 *  - every function returns F2F_OK (0) or a negative F2F_E* code; the text of
 *    the last error of the calling thread is available from f2f_last_error().
 *    Where the reference exits/aborts (xmalloc.c:16-22, mask.c:238-241,
 *    main.c:189-193) this library returns an error code instead.
 *  - images are row-major float32, `nx` = width, `ny` = height, exactly as in
 *    the reference's C code; flows are planar (u1 plane then u2 plane), the
 *    layout main.c:173-174 allocates and `read_flow(...).permute(2,0,1)`
 *    (blind_denoising.py:206-209) produces.
 *  - threading: calls act on the CURRENT CUDA device (cudaSetDevice) and are
 *    serialised where they touch library-owned state: the TV-L1 entry points
 *    hold one process-wide lock for the duration of a call (their cached staging
 *    buffers, streams and convergence flags are shared and are rebuilt when the
 *    current device changes), so concurrent callers are safe but do not overlap.
 *    The DnCNN / F2F-step entry points keep no per-call state besides what the
 *    caller passes in; the measurement and development switches
 *    (f2f_profile_*, f2f_dncnn_set_trace / _set_debug_dump, f2f_set_*) are
 *    process-global and must not be flipped while another thread is inside a call.
 *    One host thread per GPU -- how torch.distributed runs it -- needs nothing else.
 *
 * Citations are into the reference tree (gauenk/frame2frame).
 */
#ifndef F2F_B200_H
#define F2F_B200_H

#include <stddef.h>
#include <stdint.h>
#include <stdbool.h>

#ifdef __cplusplus
extern "C" {
#endif

#define F2F_OK 0
#define F2F_EINVAL (-1)   /* bad argument                                   */
#define F2F_ESIGMA (-2)   /* "GaussianSmooth: sigma too large" mask.c:238   */
#define F2F_ENOMEM (-3)   /* workspace too small / allocation failed        */
#define F2F_ECUDA (-4)    /* CUDA runtime error (see f2f_last_error)        */

const char *f2f_last_error(void);
/* Library/ABI version and the SM architecture the kernels were built for. */
int f2f_version(void);
int f2f_built_sm(void);

/* ------------------------------------------------------------------ TV-L1 */

/* Parameters of tvl1flow (main.c:86-99 positional arguments 5..13). */
typedef struct f2f_tvl1_params {
    float tau;      /* time step                  default 0.25 (main.c:27) */
    float lambda;   /* data term weight           default 0.15 (main.c:28) */
    float theta;    /* tightness                  default 0.3  (main.c:29) */
    int nscales;    /* pyramid levels             default 100  (main.c:30) */
    int fscale;     /* finest scale computed      default 0    (main.c:31) */
    float zfactor;  /* pyramid zoom factor        default 0.5  (main.c:32) */
    int nwarps;     /* warps per scale            default 5    (main.c:33) */
    float epsilon;  /* stopping threshold         default 0.01 (main.c:34) */
    int verbose;    /* print per-warp iteration counts to stderr (main.c:35) */
} f2f_tvl1_params;

/* main.c:25-35 */
void f2f_tvl1_default_params(f2f_tvl1_params *p);
/* main.c:102-141 (out-of-range values reset to defaults) followed by the
 * image-size dependent clamp of nscales/fscale, main.c:159-163.  Must be
 * applied before f2f_tvl1_flow_batch, like main() does before calling
 * Dual_TVL1_optic_flow_multiscale. */
int f2f_tvl1_validate_params(f2f_tvl1_params *p, int nx, int ny);
/* zoom.c:24-36 */
void f2f_tvl1_zoom_size(int nx, int ny, int *nxx, int *nyy, float factor);

/* Bytes of device workspace needed to process `nslots` frame pairs
 * concurrently (params already validated).  Returns 0 on bad arguments. */
size_t f2f_tvl1_workspace_bytes(int nx, int ny, const f2f_tvl1_params *p, int nslots);

/* Replaces Dual_TVL1_optic_flow_multiscale (tvl1flow_lib.c:356-372) for a
 * batch of independent frame pairs resident in HBM.
 *   d_I0, d_I1 : [npairs][ny][nx] float32 (any value range; normalised inside
 *                as tvl1flow_lib.c:314-348 does)
 *   d_u        : [npairs][2][ny][nx] float32 out (u1 plane, u2 plane)
 *   h_iters    : optional host int32 [npairs][nscales][nwarps]: inner
 *                iterations run per (scale, warp) -- what `verbose` prints at
 *                tvl1flow_lib.c:259-262; rows of scales < fscale are 0.
 *   d_workspace: >= f2f_tvl1_workspace_bytes(nx, ny, p, nslots) for some
 *                nslots >= 1; pairs are processed nslots at a time.
 * Synchronises `stream` before returning iff h_iters != NULL or verbose. */
int f2f_tvl1_flow_batch(const float *d_I0, const float *d_I1, float *d_u,
                        int npairs, int nx, int ny, const f2f_tvl1_params *p,
                        int32_t *h_iters, void *d_workspace, size_t ws_bytes,
                        void *stream);

/* Both directions of every pair in one pass -- what tvl1flow.sh:11-18 runs as two processes per frame (fflow: I0 = frame
 * i, I1 = frame i+1; bflow: the reverse).  The joint normalisation, the pre-smoothing and the two pyramids
 * (tvl1flow_lib.c:393-414) do not depend on the direction and are built once per pair.
 *   d_A, d_B : [npairs][ny][nx]
 *   d_u      : [npairs][2][2][ny][nx]: direction 0 = flow from A to B (I0 = A, I1 = B), direction 1 = from B to A
 *   h_iters  : optional [npairs][2][nscales][nwarps]
 * Results are bit-identical to two f2f_tvl1_flow_batch calls. */
int f2f_tvl1_flow_bidir(const float *d_A, const float *d_B, float *d_u, int npairs, int nx, int ny,
                        const f2f_tvl1_params *p, int32_t *h_iters, void *d_workspace, size_t ws_bytes,
                        void *stream);

/* Same, for HOST buffers: uploads the frames, runs the batch and downloads the
 * flows (device memory is allocated inside and cached between calls; release
 * with f2f_release()).  This is the call a host program that used to run one
 * `tvl1flow` process per pair (tvl1flow.sh:9-19) makes instead.
 * max_slots <= 0 picks a default. */
int f2f_tvl1_flow_batch_host(const float *h_I0, const float *h_I1, float *h_u,
                             int npairs, int nx, int ny, const f2f_tvl1_params *p,
                             int32_t *h_iters, int max_slots);

/* Drop-in for the reference symbol, same signature and host-pointer semantics
 * (tvl1flow_lib.c:356-372): a main.c linked against this library instead of
 * #including tvl1flow_lib.c runs unchanged.  Aborts on error like the
 * reference does. */
void Dual_TVL1_optic_flow_multiscale(float *I0, float *I1, float *u1, float *u2,
                                     const int nxx, const int nyy, const float tau,
                                     const float lambda, const float theta,
                                     const int nscales, const int fscale,
                                     const float zfactor, const int warps,
                                     const float epsilon, const bool verbose);

/* Stage-level entry points (device pointers), exported so that each kernel can
 * be parity-tested against the reference function it replaces. */
int f2f_tvl1_gaussian(float *d_I, int nx, int ny, double sigma, float *d_tmp,
                      void *stream);                       /* mask.c:223-339  */
/* Allocation-free variant of f2f_tvl1_gaussian.  NO counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid allocation).  The intermediate image of the
 * row pass lives in a file-scope `static __device__` array of F2F_NOALLOC_MAX_PIXELS floats
 * (8.3 MB, held by every GPU that loads the library until exit); the call allocates nothing and
 * calls no allocating library routine.  nx * ny > F2F_NOALLOC_MAX_PIXELS returns F2F_EINVAL.
 * Every call on a device shares the array: calls that overlap on the GPU (two streams, two host
 * threads) RACE on it -- serialise them on one stream.  Nothing in the array must start at zero
 * (the row pass writes every element the column pass reads), so it is not cleared between calls.
 * Result: bit-identical to f2f_tvl1_gaussian. */
#define F2F_NOALLOC_MAX_PIXELS (1920 * 1080)
int f2f_tvl1_gaussian_noalloc(float *d_I, int nx, int ny, double sigma, void *stream);
int f2f_tvl1_zoom_out(const float *d_I, float *d_out, int nx, int ny, float factor,
                      float *d_tmp2, void *stream);        /* zoom.c:43-81 (d_tmp2: 2*nx*ny) */
int f2f_tvl1_zoom_in_scaled(const float *d_I, float *d_out, int nx, int ny, int nxx,
                            int nyy, float zfactor, void *stream); /* zoom.c:89-115 + tvl1flow_lib.c:442-446 */
int f2f_tvl1_centered_gradient(const float *d_I, float *d_dx, float *d_dy, int nx,
                               int ny, void *stream);      /* mask.c:156-215  */
int f2f_tvl1_normalize(const float *d_I0, const float *d_I1, float *d_I0n,
                       float *d_I1n, int nx, int ny, void *d_scratch8,
                       void *stream);                      /* tvl1flow_lib.c:314-348 */

/* ------------------------------------------------- frame-to-frame step */

/* Replaces WarpedLoss.warp + WarpedLoss.occlusion_mask (blind_denoising.py:49-113,
 * lib/frame2frame/warped_loss.py:145-211) with one kernel.
 *   d_x      [B,C,H,W] frame to warp (the previous frame)
 *   d_flow   [B,2,H,W] flow, channel 0 = horizontal; the occlusion test uses
 *            batch item 0's flow for every item and only mask[0,0] is dilated,
 *            exactly as the reference code does (blind_denoising.py:92-103)
 *   d_in_mask optional [B,C,H,W] multiplied into the mask (warped_loss.py:222-223)
 *   d_warped [B,C,H,W] out: bilinear sample, zeros padding
 *   d_valid  optional out: the mask warp() returns (1 where the sample of a
 *            ones image is >= 0.9999, blind_denoising.py:73-76)
 *   d_mask   optional out: valid * (1 - occluded), blind_denoising.py:112
 *   align_corners: 0 = as run by the reference under torch >= 1.3
 *            (grid_sample default), 1 = the convention its grid scaling assumes
 *   occ_f64: 1 = float64 sum/compare of blind_denoising.py:88-94,
 *            0 = float32 of warped_loss.py:184-190 */
int f2f_warp_mask(const float *d_x, const float *d_flow, const float *d_in_mask,
                  float *d_warped, float *d_valid, float *d_mask, int B, int C, int H,
                  int W, int align_corners, int occ_f64, void *stream);
/* WarpedLoss.occlusion_mask alone (blind_denoising.py:81-113). */
int f2f_occlusion_mask(const float *d_flow, const float *d_old_mask, float *d_mask,
                       int B, int C, int H, int W, int occ_f64, void *stream);
/* Masked loss and its gradient wrt d_in in one pass; *d_loss (device double)
 * is overwritten.  kind 0: sum|m*in-m*w| (blind_denoising.py:121);
 * 1: mean (.)^2; 2: mean sqrt((.)^2+1e-6) (warped_loss.py:239-248).
 * d_grad may be NULL. */
int f2f_masked_loss(const float *d_in, const float *d_warped, const float *d_mask,
                    float *d_grad, double *d_loss, size_t n, int kind, void *stream);

/* ------------------------------------------------------------------ DnCNN */
/* Layer-level entry points of the 17-layer DnCNN (lib/frame2frame/dncnn/net.py:6-39).
 * Activations are NHWC BF16 ([H][W][64], one pixel = 128 B); B = 1 as in the
 * reference's frame loop (blind_denoising.py:187-230). */

/* conv weight [64][64][3][3] FP32 (PyTorch layout, net.py:16) -> BF16 operand image
 * [9 taps][8][64][8] (147456 B) for the forward (transpose=0) or the input-gradient
 * (transpose=1: flipped taps, transposed channels) tensor-core convolution. */
int f2f_dncnn_prep_w64(const float *d_w, void *d_wprep, int transpose, void *stream);

/* out = conv3x3(act(in)), 64->64, zero padding 1, no bias (net.py:16).
 *   act = identity if d_scale == NULL, else relu(in*scale[c] + shift[c]): the
 *   BatchNorm+ReLU of the layer below (net.py:17-18) folded into operand staging.
 *   d_stats (optional, 128 doubles, accumulated into): per-channel sum and sum of
 *   squares of the stored BF16 outputs = train-mode batch statistics. */
int f2f_dncnn_conv64(const void *d_in, void *d_out, const void *d_wprep, const float *d_scale,
                     const float *d_shift, double *d_stats, int H, int W, void *stream);

/* Input gradient of a 64->64 conv fused with the backward of the ReLU (and BatchNorm
 * sums) below it: g = conv3x3(dz, W^T flipped) masked by the layer below's ReLU.
 *   d_zprev: with BN (d_pscale != NULL) the pre-BN output z of the layer below, mask is
 *   z*pscale+pshift > 0 and d_stats accumulates sum g and sum g*xhat (xhat = (z-pmean)*pinvstd);
 *   without BN (layer 0) its post-ReLU activation, mask is > 0. */
int f2f_dncnn_dgrad64(const void *d_dz, void *d_gout, const void *d_wprep_t, const void *d_zprev,
                      const float *d_pscale, const float *d_pshift, const float *d_pmean,
                      const float *d_pinvstd, double *d_stats, int H, int W, void *stream);

/* Weight gradient of a 64->64 conv (tcgen05, pixels as the contraction) fused with the
 * BatchNorm backward of the layer: dz = A*g + B*z + C per channel is formed while the operand
 * is staged and written back over d_g (so f2f_dncnn_dgrad64 can consume it next).
 *   d_act    : the layer's input, z of the layer below (d_coef_in = its 448-float coefficient
 *              block: scale[64], shift[64], mean[64], invstd[64], A[64], B[64], C[64]; BN+ReLU
 *              applied on load) or a post-ReLU activation (d_coef_in = NULL)
 *   d_g, d_z : masked output gradient (in) / dz (out), and the layer's pre-BN output
 *   d_coef   : the layer's own coefficient block (A, B, C used)
 *   d_partial: scratch, (number of SMs) * 49152 floats; d_dw: [64][64][3][3] FP32 (PyTorch layout). */
int f2f_dncnn_wgrad64(const void *d_act, const float *d_coef_in, void *d_g, const void *d_z,
                      const float *d_coef, float *d_partial, float *d_dw, int H, int W,
                      int accumulate, void *stream);

/* Whole network.  d_params / d_grads: flat FP32 vector in DnCNN.state_dict() order (conv0.w,
 * then conv_k.w, bn_k.weight, bn_k.bias for k = 1..15, conv16.w; f2f_dncnn_param_count() =
 * 556032 floats, channels = 1).  d_bn_running: [15][2][64] running mean / var; d_nbt: [15]
 * num_batches_tracked (int64) or NULL.  d_x, d_y: [B][H][W] FP32.
 *   train = 1: model.train() forward (blind_denoising.py:212-218) -- batch statistics over
 *              B*H*W, running stats updated (momentum 0.1, unbiased variance), activations kept
 *              in the workspace for f2f_dncnn_backward;
 *   train = 0: model.eval() forward (blind_denoising.py:226-230);
 *   residual : 0 = y is the noise estimate (the paper script's dncnn.models.DnCNN),
 *              1 = y = x - noise (lib/frame2frame/dncnn/net.py:31). */
size_t f2f_dncnn_param_count(void);
size_t f2f_dncnn_workspace_bytes(int B, int H, int W, int train);
/* Byte offsets inside the workspace, for layer-by-layer parity tests: out[0] saved input x
 * (FP32), out[1] a_0 (NHWC BF16), out[2] z_1 (z_k at out[2] + (k-1)*out[3] in train mode, at
 * out[2] + (k&1)*out[3] in eval mode), out[3] bytes of one activation, out[4], out[5] the two
 * gradient buffers, out[6] coefficient blocks (15 x 448 floats), out[7] forward batch sums
 * (15 x 128 doubles). */
void f2f_dncnn_workspace_layout(int B, int H, int W, int train, size_t out[8]);
/* Test hook: when d_buf != NULL, f2f_dncnn_backward copies its intermediate gradients into it
 * (31 slots of out[3] bytes: slot 2(k-1) = masked gradient g_k entering layer k's BatchNorm
 * backward, slot 2(k-1)+1 = dz_k, slot 30 = g_0), so each backward kernel can be checked against
 * the reference formula on its own inputs.  NULL switches it off. */
void f2f_dncnn_set_debug_dump(void *d_buf);
/* Development aid: when d_counters != NULL (16 device uint64, zeroed by the caller), the 64->64
 * convolution / input-gradient kernels add per-role SM-cycle totals over all CTAs: [0],[1] operand
 * producer waiting for a free slot / staging a tile (one thread of group 0); [2],[3],[4] MMA
 * issuer waiting for a drained accumulator / a staged tile / issuing; [5],[6] epilogue waiting for
 * accumulators / draining a tile; [7] epilogue loop total; [8] tiles.  NULL switches it off. */
void f2f_dncnn_set_trace(unsigned long long *d_counters);
/* Precision mode of the whole-network entry points (f2f_dncnn_forward / _backward / f2f_step_train and
 * f2f_dncnn_workspace_bytes / _layout, which must be asked again after a switch):
 *   0 (default): activations, gradients and tensor-core operands in 16 bits, FP32 accumulation (tcgen05);
 *   1 ("precise"): everything in FP32 on the CUDA cores (csrc/dncnn_precise.cuh) -- the arithmetic of the reference's
 *      FP32 torch path (blind_denoising.py:211-230) up to summation order, ~25x slower; the setting in which the
 *      online loop is held to north_star's 0.05 dB PSNR tolerance independently of 16-bit rounding.
 *   2 ("split operand"): every FP32 operand as an FP16 hi + lo/2048 pair, three tcgen05 products per GEMM, FP32
 *      activations between layers (csrc/dncnn_precise2.cuh): FP32-class arithmetic on the tensor cores, ~5x slower
 *      than mode 0.  FP16 builds only; a BF16 build falls back to mode 1.
 * F2F_PRECISE in the environment sets the initial value.  Process-global, like the other switches. */
void f2f_dncnn_set_precise(int mode);
int f2f_dncnn_get_precise(void);
/* Programmatic dependent launch of the per-layer DnCNN kernel chain: bit mask, default 3 (launch attribute on the big
 * and the small kernels, no early trigger); 0 launches the same kernels fully serialised (F2F_PDL in the environment
 * sets the initial value).  Results are identical, only the gaps between kernels differ (csrc/f2f_common.cuh). */
void f2f_set_pdl(int on);
/* Serpentine tile order of the 64->64 layer chain (default 1, F2F_SERPENTINE in the environment): consecutive launches
 * walk the image in opposite directions (forward: odd layers bottom-up; backward: weight gradient bottom-up, input
 * gradient top-down), so every launch starts on the rows its predecessor touched last, which the 126 MB L2 still
 * holds -- a 1080p activation is 265 MB.  0 = every launch top-down.  Same results up to FP32 summation order of
 * the weight-gradient partials and BatchNorm sums. */
void f2f_set_serpentine(int on);
/* Experimental: on != 0 runs the 64->64 convolutions (forward and input gradient) as CTA pairs -- clusters of two
 * CTAs issuing one tcgen05.mma.cta_group::2 M = 256 instruction stream, each CTA supplying half of the weights
 * (F2F_CONV_PAIR in the environment sets the initial value, default 0).  Same results bit for bit. */
void f2f_set_conv_pair(int on);
/* 16-bit storage type the library was built with for DnCNN activations, gradients and tensor-core operands:
 * 1 = FP16 (default), 0 = BF16 (-DF2F_STORE_BF16, frame2frame_b200/build.py F2F_STORE_BF16=1).  Callers that
 * pre-pack operand buffers (f2f_dncnn_conv64 / dgrad64 / wgrad64) must use this type. */
int f2f_dncnn_storage_dtype(void);
int f2f_dncnn_forward(const float *d_params, float *d_bn_running, long long *d_nbt,
                      const float *d_x, float *d_y, void *d_ws, size_t ws_bytes, int B, int H,
                      int W, int train, int residual, void *stream);
/* Gradient of a scalar loss wrt every parameter given d_dy = dL/dy of the last train-mode
 * forward on the same workspace (loss.backward(), blind_denoising.py:221).  d_grads is
 * overwritten (accumulate = 0) or added to (accumulate = 1). */
int f2f_dncnn_backward(const float *d_params, float *d_grads, const float *d_dy, void *d_ws,
                       size_t ws_bytes, int B, int H, int W, int residual, int accumulate,
                       void *stream);
/* The same for a network with `channels` image channels (1 or 3; the reference's loader defaults to 3,
 * lib/frame2frame/dncnn/io.py:64-67): d_x, d_y, d_dy are [B][channels][H][W] FP32 (NCHW like the reference's tensors),
 * the flat parameter / gradient vectors have f2f_dncnn_param_count_c(channels) = 556032 + 1152 * (channels - 1) floats
 * (conv0.weight [64][channels][3][3] first, conv16.weight [channels][64][3][3] last).  channels = 1 is exactly the
 * functions above; other channel counts run the first / last layer on generic CUDA-core kernels. */
size_t f2f_dncnn_param_count_c(int channels);
size_t f2f_dncnn_workspace_bytes_c(int B, int H, int W, int train, int channels);
int f2f_dncnn_forward_c(const float *d_params, float *d_bn_running, long long *d_nbt,
                        const float *d_x, float *d_y, void *d_ws, size_t ws_bytes, int B, int H,
                        int W, int train, int residual, int channels, void *stream);
int f2f_dncnn_backward_c(const float *d_params, float *d_grads, const float *d_dy, void *d_ws,
                         size_t ws_bytes, int B, int H, int W, int residual, int accumulate,
                         int channels, void *stream);
/* torch.optim.Adam step (amsgrad = False, weight decay added to the gradient) on one flat
 * tensor; `step` counts from 1 (blind_denoising.py:163-165, 222). */
int f2f_adam_step(float *d_p, const float *d_g, float *d_m, float *d_v, size_t n,
                  long long step, float lr, float beta1, float beta2, float eps,
                  float weight_decay, void *stream);

/* One inner iteration of the reference's fine-tuning loop (blind_denoising.py:216-222) in one
 * call: out = curr - net(curr) in train mode, loss = masked loss of out against the pre-warped
 * previous frame (f2f_warp_mask output; kinds as f2f_masked_loss), backward, Adam step `step`.
 * d_out, d_dout: [H][W] scratch (d_out holds the train-mode output afterwards); *d_loss: device
 * double. */
int f2f_step_train(float *d_params, float *d_grads, float *d_m, float *d_v, float *d_bn_running,
                   long long *d_nbt, const float *d_curr, const float *d_warped,
                   const float *d_mask, float *d_out, float *d_dout, double *d_loss, void *d_ws,
                   size_t ws_bytes, int H, int W, int loss_kind, long long step, float lr,
                   float beta1, float beta2, float eps, float weight_decay, void *stream);

/* Measurement hooks for bench.py.  While enabled, the TV-L1 batch brackets the
 * finest-scale inner-iteration launches with CUDA events on the caller's
 * stream (and synchronises at the end of each group to read them).
 * f2f_profile_read: out[0] = ms spent in those launches, out[1] = sum over
 * frame pairs of inner iterations run at scale 0, out[2] = pixels of scale 0,
 * out[3] = number of those launches, out[4] = kernels launched by the library
 * since load (all entry points). */
void f2f_profile_enable(int on);
void f2f_profile_read(double out[5]);

/* Same for the DnCNN tensor-core kernels: while on, every 64->64 convolution / input-gradient
 * launch (class 0) and weight-gradient launch (class 1) is bracketed with CUDA events on its
 * stream.  f2f_profile_dncnn(1) also resets the counters.  f2f_profile_dncnn_read synchronises
 * those events: out[0] = ms in class-0 launches, out[1] = their number, out[2], out[3] = class 1. */
void f2f_profile_dncnn(int on);
void f2f_profile_dncnn_read(double out[4]);

/* Compares the iteration kernel's in-line correctly-rounded division with IEEE
 * division on n pseudo-random operand pairs in the ranges the solver produces;
 * *h_mismatches receives the number of differing results (expected 0). */
int f2f_selftest_division(unsigned long long n, unsigned long long *h_mismatches);
/* The same for the dual update's (float)hypot(a, b) (tvl1flow_lib.c:247-248): the in-line correctly rounded double
 * square root (no slow-path call) against the library sqrt(), float and double results, n pseudo-random gradient
 * pairs incl. zeros; *h_mismatches receives the number of differing results (expected 0). */
int f2f_selftest_hypot(unsigned long long n, unsigned long long *h_mismatches);


/* Release every cached device allocation of the calling process. */
void f2f_release(void);

#ifdef __cplusplus
}
#endif
#endif /* F2F_B200_H */
