/* C ABI of libdcgan_b200.so -- the hand-written sm_100a kernels of the B200-native DCGAN framework.
 *
 * Every entry point takes raw device pointers plus a CUDA stream, launches asynchronously on that
 * stream and returns 0 on success or a negative DCGAN_ERR_* code.  None of them synchronises.  The
 * Python package binds these with ctypes (distributed_tensorflow_for_dcgan_b200/ops/_lib.py).
 *
 * Tensor conventions: activations are NHWC bf16; master weights / gradients / optimizer state are
 * f32 in TensorFlow layouts (conv HWIO, deconv [kh,kw,Cout,Cin], linear [in,out]) so that the
 * checkpoint maps one-to-one on the reference's variable names.
 * Geometry is TensorFlow's 5x5 / stride 2 / SAME: one pixel of padding before, two after.
 */
#ifndef DCGAN_B200_H_
#define DCGAN_B200_H_

#include <cuda_runtime_api.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DCGAN_ERR_SHAPE (-1)
#define DCGAN_ERR_DRIVER (-2)
#define DCGAN_ERR_LAUNCH (-3)
#define DCGAN_ERR_LIMIT (-4)

#define DCGAN_ACT_NONE 0
#define DCGAN_ACT_RELU 1
#define DCGAN_ACT_LRELU 2
#define DCGAN_ACT_TANH 3
#define DCGAN_ACT_SIGMOID 4

/* Reads and clears the sticky device-side error flag (0 = none, 1 = mbarrier wait timed out,
 * 2 = peer signal wait timed out).  Synchronises the device. */
int dcgan_last_device_error(void);

/* Stream-ordered read of the same flag into pinned host memory (4 bytes; neither synchronises nor clears): a
 * training loop fetches it with the losses of every step and stops when it is non-zero, so that a timed-out
 * peer wait -- which makes a rank skip its shard's update -- cannot go unnoticed. */
int dcgan_copy_error_flag(int* host_dst, cudaStream_t stream);
/* Stream-ordered zero fill of `bytes` bytes (cudaMemsetAsync: a memset node inside a captured graph). */
int dcgan_zero_async(void* ptr, size_t bytes, cudaStream_t stream);

/* Bring-up switch for the MN-major UMMA descriptor convention (0 = CuTe convention). */
void dcgan_set_mn_desc_variant(int v);

/* Programmatic dependent launch of the hot-path kernels (each kernel's set-up overlaps the tail of its
 * predecessor).  Off by default -- on this step it measured 3 % slower (the early CTAs take SM slots from the
 * side-stream filter-gradient kernels); DCGAN_PDL=1 in the environment or this call turns it on. */
void dcgan_set_pdl(int v);

/* Profiling switch for the implicit-GEMM kernels.  Bit 7 (128) selects the first-generation kernels (cp.async
 * gather, csrc/igemm_sm100.cu) instead of the TMA-fed persistent ones (csrc/igemm2_sm100.cu); the other bits
 * only act on the first generation: bit 0 = two CTAs per SM (default on). */
void dcgan_set_igemm_tuning(int v);
/* Second-generation kernels: cap on the persistent grid (0 = every CTA slot of the device), ring configuration
 * (0 = default: two CTAs per SM for tiles of 64 / 128 columns; 2 = one CTA per SM with the deepest ring; 3 = 64-column
 * tiles with two k-blocks per ring slot). */
void dcgan_set_igemm_max_ctas(int v);
void dcgan_set_igemm_stage_mode(int v);
/* Tile width (64 / 128 / 256 output channels) used wherever it divides the layer; 0 = the library's choice. */
void dcgan_set_igemm_tile(int v);
/* Records the measured-best (tile width, split-K factor) of one layer shape; later calls of that shape which leave
 * both to the library (block_n = 0, splits = 0) use it instead of the cost model.  kind: 0 strided conv (fprop*),
 * 1 transposed conv (tconv*), 2 filter gradient (there H, W = the small grid, Cin = Cg, Cout = Cd).  block_n <= 0
 * forgets the entry.  variant: 0 = default kernel of that width, 1 = two k-blocks per ring slot (64 columns only; what
 * dcgan_set_igemm_stage_mode(3) forces globally).  ops/autotune.py fills the table by timing the candidates once per shape. */
int dcgan_get_igemm_ksplit(void); /* the factor forced by dcgan_set_igemm_ksplit, 0 = none */
int dcgan_igemm_plan_set(int kind, int NB, int H, int W, int Cin, int Cout, int block_n, int ksplit, int variant);
/* Bring-up only: per-CTA cycle counters [CTA][16] written by a library built with -DIGEMM2_TIMING (else ignored). */
void dcgan_set_igemm_debug(long long* counters);
/* Split-K scratch of the second-generation kernels for the calling host thread: `bytes` of device memory whose
 * first 64 KiB are zero (arrival counters; the kernels leave them zero), at least 1 MiB + 64 KiB.  Kernels that
 * may run concurrently (different streams) must not share a workspace.  NULL returns to the library's default
 * buffer (48 MiB per device, allocated on first need -- not possible inside a stream capture). */
int dcgan_set_workspace(void* ptr, size_t bytes);
/* Split-K factor of the strided / transposed convs (a thread-block cluster shares one output tile and reduces
 * its partial accumulators through distributed shared memory): 0 = chosen per layer (default), 1/2/4 = forced.
 * Tests and profiling only. */
void dcgan_set_igemm_ksplit(int v);
/* CTA count up to which the filter-gradient kernel keeps doubling its split-K factor (default 148). */
void dcgan_set_wgrad_cta_target(int v);
/* CTA caps of the 3-channel tensor-core kernels (0 = default): forward / data-gradient kernel, filter-gradient
 * kernel.  Profiling only. */
void dcgan_set_c3_tuning(int fprop_blocks, int wgrad_blocks);
/* Cap on the block count of dcgan_adam (0 = one block per 1024 parameters): a capped launch walks its range
 * with a grid-stride loop and leaves SMs to kernels running beside it. */
void dcgan_set_adam_blocks(int v);
/* Host-only: the launch plan (tile width, split-K factor, grid, partial tiles fit the stage ring) that
 * dcgan_conv_fprop* (transposed = 0) / dcgan_conv_tconv* (transposed = 1) would use.  Makes no CUDA call. */
int dcgan_debug_plan_conv(int transposed, int NB, int H, int W, int Cin, int Cout, int* block_n, int* ksplit,
                          int* grid_xyz, int* partials_fit);

/* ---- implicit-GEMM convolutions on tcgen05 / TMEM / TMA (csrc/igemm_sm100.cu) ------------------ */

/* Strided forward conv 5x5 s2 SAME: out[NB,H/2,W/2,Cout] = act(conv(x[NB,H,W,Cin]) + bias).
 * w_nk is bf16 [25][Cout][Cin] (tap-major, reduction channel contiguous).
 * Same math as tf.nn.conv2d at reference distriubted_model.py:183 and as the data gradient of the
 * transposed conv at :200.  block_n <= 0 picks a tile width. */
int dcgan_conv_fprop(const void* x, const void* w_nk, const float* bias, void* out, int NB, int H, int W,
                     int Cin, int Cout, int act, int block_n, cudaStream_t stream);

/* Same, with the filter in its TensorFlow layout [25][Cin][Cout] (HWIO; also the deconv layout
 * [kh,kw,Cout_d,Cin_d] read as a data-gradient filter): fed to the MMA as an MN-major operand, so no
 * transposed weight copy is needed. */
int dcgan_conv_fprop_kn(const void* x, const void* w_kn, const float* bias, void* out, int NB, int H, int W,
                        int Cin, int Cout, int act, int block_n, cudaStream_t stream);

/* Transposed conv 5x5 s2 (TF crop 1 before / 2 after): out[NB,2H,2W,Cout] = act(deconv(x) + bias).
 * w_nk is bf16 [25][Cout][Cin].  Same math as tf.nn.conv2d_transpose at reference
 * distriubted_model.py:200 and as the data gradient of conv2d at :183. */
int dcgan_conv_tconv(const void* x, const void* w_nk, const float* bias, void* out, int NB, int H, int W,
                     int Cin, int Cout, int act, int block_n, cudaStream_t stream);

/* Variants that also accumulate the batch-norm moments of their (pre-activation) output in the epilogue:
 * stats[stat_groups][2][Cout] (f32, zeroed by the caller) += per-channel (sum, sum of squares); rows of each
 * group must be a multiple of 128.  Replaces a separate dcgan_bn_stats pass (distriubted_model.py:37). */
int dcgan_conv_fprop_kn_stats(const void* x, const void* w_kn, const float* bias, void* out, float* stats,
                              int stat_groups, int NB, int H, int W, int Cin, int Cout, int block_n,
                              cudaStream_t stream);
int dcgan_conv_tconv_stats(const void* x, const void* w_nk, const float* bias, void* out, float* stats,
                           int stat_groups, int NB, int H, int W, int Cin, int Cout, int block_n,
                           cudaStream_t stream);

/* Data-gradient convs (no bias, no activation) that also accumulate, in their epilogue, the batch-norm backward
 * reductions of the layer whose activation gradient they produce: with pre = that layer's saved pre-activations
 * [stat_groups * rows][Cout] and saved = its (mean, rstd) [stat_groups][2][Cout],
 *   dy = out * act'(gamma * xhat + beta),  red[slab][0][c] += sum dy,  red[slab][1][c] += sum dy * xhat
 * -- the result of dcgan_bn_bwd_reduce without its extra pass over out and pre (autodiff of the reference's
 * batch_norm, distriubted_model.py:15-52).  The NB images are `slabs` equal groups; slab g reads the statistics
 * of group min(g, stat_groups - 1); (NB / slabs) images must fill whole 128-row tiles. */
int dcgan_conv_tconv_bnred(const void* x, const void* w_nk, void* out, const void* pre, const float* saved,
                           const float* gamma, const float* beta, float* red, int slabs, int stat_groups, int bn_act,
                           int NB, int H, int W, int Cin, int Cout, int block_n, cudaStream_t stream);
int dcgan_conv_fprop_kn_bnred(const void* x, const void* w_kn, void* out, const void* pre, const float* saved,
                              const float* gamma, const float* beta, float* red, int slabs, int stat_groups,
                              int bn_act, int NB, int H, int W, int Cin, int Cout, int block_n,
                              cudaStream_t stream);

/* dcgan_conv_*_bnred plus the rest of that layer's batch-norm backward pass in the same launch: once every CTA has
 * stored its tiles and added its reductions (a grid-wide barrier inside the kernel), the grid rewrites out (= da) as
 * dpre = gamma * rstd * (dy - mean(dy) - xhat * mean(dy * xhat)) (bf16, shape of out) and CTA 0 writes dgamma / dbeta
 * (optional) = the reductions summed over the first param_slabs slabs -- dcgan_bn_bwd_apply without its launch.
 * Uses the calling thread's workspace (dcgan_set_workspace); two kernels of this family (these and dcgan_conv_*_bn,
 * dcgan_bn_forward_fused, dcgan_bn_backward_fused) must not run concurrently on one device: launch them on one stream. */
int dcgan_conv_tconv_bnbwd(const void* x, const void* w_nk, void* out, const void* pre, const float* saved,
                           const float* gamma, const float* beta, float* red, void* dpre, float* dgamma, float* dbeta,
                           int slabs, int stat_groups, int param_slabs, int bn_act, int NB, int H, int W, int Cin,
                           int Cout, int block_n, cudaStream_t stream);
int dcgan_conv_fprop_kn_bnbwd(const void* x, const void* w_kn, void* out, const void* pre, const float* saved,
                              const float* gamma, const float* beta, float* red, void* dpre, float* dgamma,
                              float* dbeta, int slabs, int stat_groups, int param_slabs, int bn_act, int NB, int H,
                              int W, int Cin, int Cout, int block_n, cudaStream_t stream);
/* Data gradient of a strided conv (= transposed conv of the upstream gradient) into a layer that is bias +
 * activation without batch norm (d_h0_conv, distriubted_model.py:118): out = tconv(x) * act'(act_saved) is the gradient
 * w.r.t. that layer's pre-activation, dbias[Cout] += its column sums over the first bias_slabs of `slabs` equal image
 * groups (the slabs that belong to d_loss); slab g reads the saved activations of group min(g, stat_groups - 1).
 * dcgan_conv_tconv followed by dcgan_act_bwd, in one launch. */
int dcgan_conv_tconv_actbwd(const void* x, const void* w_nk, void* out, const void* act_saved, float* dbias, int slabs,
                            int stat_groups, int bias_slabs, int act, int NB, int H, int W, int Cin, int Cout,
                            int block_n, cudaStream_t stream);
/* Forward conv + training-mode batch norm + activation in one launch (reference: conv2d / deconv2d followed by
 * batch_norm(train=True) and relu / lrelu, distriubted_model.py:91-121): pre = conv(x) + bias (bf16, kept for the
 * backward pass), sums[stat_groups][2][Cout] += its moments (zero it first), then behind a grid-wide barrier
 * act_out = act(gamma * (pre - mean) * rstd + beta), saved[stat_groups][2][Cout] = (mean, rstd), and the EMA shadows
 * (optional, [stat_groups][Cout]) advance with `decay`.  Same restrictions as dcgan_conv_*_bnbwd. */
int dcgan_conv_fprop_kn_bn(const void* x, const void* w_kn, const float* bias, void* pre, void* act_out, float* sums,
                           const float* gamma, const float* beta, float* saved, float* ema_mean, float* ema_var,
                           int stat_groups, int NB, int H, int W, int Cin, int Cout, float eps, float decay, int bn_act,
                           int block_n, cudaStream_t stream);
int dcgan_conv_tconv_bn(const void* x, const void* w_nk, const float* bias, void* pre, void* act_out, float* sums,
                        const float* gamma, const float* beta, float* saved, float* ema_mean, float* ema_var,
                        int stat_groups, int NB, int H, int W, int Cin, int Cout, float eps, float decay, int bn_act,
                        int block_n, cudaStream_t stream);

/* Weight gradient of either layer kind: dw[25][Cg][Cd] (f32, += -- zero it first; the second-generation kernels add each element once, split-K partials are
 * summed in a fixed order, so the result is reproducible bit for bit) += sum over pixels of big[NB,2Hs,2Ws,Cg] (gathered at 2q+k-1) * small[NB,Hs,Ws,Cd].
 * conv2d:  big = layer input, small = output gradient  -> dw is HWIO.
 * deconv:  big = output gradient, small = layer input  -> dw is [kh,kw,Cout,Cin].
 * (autodiff of reference image_train.py:109-112).  splits <= 0 picks a split-K factor. */
int dcgan_conv_wgrad(const void* big, const void* small, float* dw, int NB, int Hs, int Ws, int Cg, int Cd,
                     int splits, int block_n, cudaStream_t stream);

/* ---- 3-channel layers on the CUDA cores (csrc/conv_c3.cu) ------------------------------------------
 * Filter layout for all three is f32 [25][3][C] (= d_h0_conv/w HWIO and g_h4/w [kh,kw,3,C]). */

/* out[N,H/2,W/2,C] = act(conv5x5s2(x[N,H,W,3]) + bias): d_h0 forward (distriubted_model.py:118) and
 * the g_h4 data gradient. */
int dcgan_c3_fprop(const void* x, const float* w, const float* bias, void* out, int N, int H, int W, int C,
                   int act, cudaStream_t stream);
/* out[N,2H,2W,3] from x[N,H,W,C]: g_h4 forward (distriubted_model.py:109; mode 0: act(v+bias)) and the
 * d_h0 data gradient (mode 1: v * (1 - aux^2), the tanh backward against aux = generator output).
 * colsum (optional, f32[3], zeroed by the caller) += per-channel sum of the results = g_h4's bias gradient. */
int dcgan_c3_tconv(const void* x, const float* w, const float* bias, const void* aux, void* out, float* colsum,
                   int N, int H, int W, int C, int act, int mode, cudaStream_t stream);
/* dw[25][3][C] += sum x3[N,2Hs,2Ws,3] (gathered) * y[N,Hs,Ws,C]   (f32 atomics; zero first). */
int dcgan_c3_wgrad(const void* x3, const void* y, float* dw, int N, int Hs, int Ws, int C, cudaStream_t stream);

/* Tensor-core versions (csrc/conv_c3_tc.cu): im2col built in shared memory in the UMMA layout (K padded to
 * 5x16), tcgen05.mma with TMEM accumulators.  Output width (W/2, resp. Ws) must be 32, 64 or 128. */
int dcgan_c3_fprop_tc(const void* x, const float* w, const float* bias, void* out, int N, int H, int W, int C,
                      int act, cudaStream_t stream);
int dcgan_c3_wgrad_tc(const void* x3, const void* y, float* dw, int N, int Hs, int Ws, int C, cudaStream_t stream);
/* C=64 -> 3 transposed conv as a plain GEMM Y[q][tap*3+co] on tcgen05 followed by a col2im in shared memory;
 * same contract as dcgan_c3_tconv but the filter comes as the bf16 shadow [75][64]. */
int dcgan_c3_tconv_tc(const void* x, const void* w_bf16, const float* bias, const void* aux, void* out, float* colsum,
                      int N, int H, int W, int C, int act, int mode, cudaStream_t stream);

/* ---- batch norm fused with its activation (csrc/bn_act.cu; reference distriubted_model.py:15-52) --
 * x/y/da/dpre are bf16 [groups*rows_per_group][C]; sums/saved/red are f32 [groups][2][C]. */
int dcgan_bn_stats(const void* x, float* sums, int groups, int rows_per_group, int C, cudaStream_t stream);
int dcgan_bn_apply(const void* x, void* y, const float* sums, const float* gamma, const float* beta, float* saved,
                   float* ema_mean, float* ema_var, int groups, int rows_per_group, int C, float eps, float decay,
                   int act, int use_ema, cudaStream_t stream);
int dcgan_bn_bwd_reduce(const void* da, const void* pre, const float* saved, const float* gamma, const float* beta,
                        float* red, int groups, int stat_groups, int rows_per_group, int C, int act,
                        cudaStream_t stream);
int dcgan_bn_bwd_apply(const void* da, const void* pre, void* dpre, const float* saved, const float* gamma,
                       const float* beta, const float* red, float* dgamma, float* dbeta, int groups,
                       int stat_groups, int param_groups, int rows_per_group, int C, int act, cudaStream_t stream);
/* dcgan_bn_stats + dcgan_bn_apply (training mode) in one launch, and dcgan_bn_bwd_reduce + dcgan_bn_bwd_apply in
 * one launch: the statistics pass, a grid-wide barrier inside the kernel, then the elementwise pass.  For the layers
 * whose producer is not one of the convs above (which carry the batch norm in their own launch: dcgan_conv_*_bn,
 * dcgan_conv_*_bnbwd).  `barrier` = two ints of device memory that are zero (the kernel leaves them zero); sums / red
 * must be zero on entry.  Not to be run concurrently with another kernel of the grid-barrier family on one device. */
int dcgan_bn_forward_fused(const void* x, void* y, float* sums, const float* gamma, const float* beta, float* saved,
                           float* ema_mean, float* ema_var, int* barrier, int groups, int rows_per_group, int C,
                           float eps, float decay, int act, cudaStream_t st);
int dcgan_bn_backward_fused(const void* da, const void* pre, void* dpre, const float* saved, const float* gamma,
                            const float* beta, float* red, float* dgamma, float* dbeta, int* barrier, int groups,
                            int stat_groups, int param_groups, int rows_per_group, int C, int act, cudaStream_t st);
/* Batch-norm forward (statistics + normalise + activation) for callers that cannot allocate memory.
 * No counterpart in the reference.  Same result as dcgan_bn_stats followed by dcgan_bn_apply, but the
 * per-channel sums live in a 64 KiB array built into the library (groups*2*C <= 16384 floats, otherwise
 * DCGAN_ERR_LIMIT), so no buffer is asked of the caller and nothing is allocated.  Meant for hosts that
 * only hold buffers lent by another runtime's pool, or that must not call cudaMalloc because it stalls
 * the device.  The first call on each device looks the array's address up -- make it outside stream
 * capture.  All calls share the array: it is cleared per call, and calls that overlap on the GPU race. */
int dcgan_bn_forward_static(const void* x, void* y, const float* gamma, const float* beta, float* saved,
                            float* ema_mean, float* ema_var, int groups, int rows_per_group, int C, float eps,
                            float decay, int act, cudaStream_t stream);
/* dpre = da * act'(a) for a layer without batch norm (d_h0), plus its bias gradient. */
int dcgan_act_bwd(const void* da, const void* a, void* dpre, float* dbias, int rows, int a_rows, int C, int act,
                  int bias_rows, cudaStream_t stream);

/* ---- skinny layers, losses, optimizer, helpers (csrc/heads_optim.cu) ------------------------------ */
int dcgan_lin_fwd(const float* z, const void* W_bf16, const float* bias, void* out, int NB, int K, int F,
                  cudaStream_t stream);
int dcgan_lin_wgrad(const float* z, const void* dy, float* dW, float* dbias, int NB, int K, int F,
                    cudaStream_t stream);
int dcgan_head_fwd(const void* h, const float* w, const float* b, float* logits, int rows, int F,
                   cudaStream_t stream);
/* losses[3] = {d_loss_real, d_loss_fake, g_loss} (image_train.py:91-96); dlogit[3B]; advances
 * counters[0] (Adam t) and counters[1] (global_step) when advance != 0. */
int dcgan_gan_loss(const float* logits, float* losses, float* dlogit, float* probs, long long* counters, int B,
                   float grad_div, int advance, cudaStream_t stream);
int dcgan_head_bwd(const void* h, const float* w, const float* dlogit, void* dh, float* dW, float* db, int B,
                   int F, cudaStream_t stream);
/* gan_loss + head_bwd in one launch (the simultaneous-update step): the logit gradients are recomputed from
 * `logits` where they are needed and block (0,0) writes losses / dlogit (optional) / probs (optional) / counters
 * exactly as dcgan_gan_loss would (image_train.py:91-96). */
int dcgan_head_loss_bwd(const void* h, const float* w, const float* logits, float* losses, float* dlogit,
                        float* probs, long long* counters, void* dh, float* dW, float* db, int B, int F,
                        float grad_div, int advance, cudaStream_t st);
/* TensorFlow-form Adam (image_train.py:109-112) over a flat range; t is read from counters[0]. */
int dcgan_adam(float* p, const float* g, float* m, float* v, void* p_bf16, const long long* counters, size_t n,
               float lr, float b1, float b2, float eps, float gscale, cudaStream_t stream);
int dcgan_cast_bf16(const float* src, void* dst, size_t n, cudaStream_t stream);
int dcgan_transpose_taps(const void* src, void* dst, int taps, int R, int C, cudaStream_t stream);
int dcgan_colsum(const void* x, float* out, size_t rows, int C, cudaStream_t stream);
/* float64 records -> bf16 (image_input.py:48,118 decode_raw + cast), with an optional affine map. */
int dcgan_decode_f64(const double* src, void* dst, size_t n, float scale, float shift, cudaStream_t stream);
/* The same with float32 output (the input pipeline's device-side decode: raw float64 records are copied to the GPU
 * as stored and cast here, data/loader.py DevicePrefetcher(gpu_decode=True)). */
int dcgan_decode_f64_f32(const double* src, float* dst, size_t n, float scale, float shift, cudaStream_t stream);
int dcgan_f32_to_bf16(const float* src, void* dst, size_t n, float scale, float shift, cudaStream_t stream);
/* inverse_transform + merge of the sample grid (image_train.py:205-219) into f32 [rows*H][cols*W][C]. */
int dcgan_grid_merge(const void* img, float* out, int N, int H, int W, int C, int cols, cudaStream_t stream);
int dcgan_count_zeros(const void* x, unsigned long long* count, size_t n, cudaStream_t stream);

/* ---- in-kernel collectives over NVLink peer memory (csrc/comm_sm100.cu) --------------------------------
 * Replace the reference's gRPC parameter push/pull.  *_ptrs are host arrays of `world`
 * device pointers (one per rank, peer-mapped symmetric buffers); `state` is 2 local uint32 per slot. */

/* Polls (about 0.7 us each) before a wait for the peers' signals gives up and raises the error flag; default 2^24
 * (~12 s).  A training driver whose chief pauses between steps (checkpoint, sample queue) raises it.  Applies to
 * the current device. */
int dcgan_comm_set_peer_spin_limit(unsigned polls);

/* Sync replicas: reduce-scatter of f32 gradients + average + TF-form Adam on this rank's shard of
 * [lo,hi) + all-gather of the updated bf16 parameters, in one kernel.  mc_* = NVLS multicast addresses
 * (gradients / bf16 shadow / f32 master) or NULL for plain P2P loads and stores.  Elements >= f32_from are
 * read in f32 by their consumers, so their master value is gathered as well (master_ptrs); elements in
 * [pull_lo,pull_hi) are not pushed because their first consumer pulls them (dcgan_comm_lin_fwd_gather). */
int dcgan_comm_rs_adam_ag(const void* const* grad_ptrs, const void* const* bf16_ptrs, const void* const* sig_ptrs,
                          const void* const* master_ptrs, float* master, float* m, float* v,
                          const long long* counters, unsigned* state, int rank, int world, size_t lo, size_t hi,
                          float lr, float b1, float b2, float eps, int slot, const float* mc_grad, void* mc_bf16,
                          float* mc_master, size_t f32_from, size_t pull_lo, size_t pull_hi, int blocks,
                          cudaStream_t stream);
/* g_h0_lin (distriubted_model.py:88) with the parameter all-gather fused into the consumer: every slice of
 * Matrix is loaded over NVLink from the bf16 shadow of the rank that owns it. */
int dcgan_comm_lin_fwd_gather(const void* const* bf16_ptrs, const float* z, const float* bias, void* out, int NB,
                              int K, int F, size_t w_off, size_t lo, size_t hi, int rank, int world,
                              cudaStream_t stream);
/* Mean all-reduce of a small f32 vector (cross-GPU batch-norm statistics). */
int dcgan_comm_allreduce_small(const void* const* staging_ptrs, const void* const* sig_ptrs, float* data, int n,
                               int staging_stride, unsigned* state, int rank, int world, int slot,
                               cudaStream_t stream);
/* Async parameter-server mode: apply this worker's gradients to the owners' shards through peer memory
 * (unlocked, stale -- reference image_train.py:109-112 with use_locking=False) and refresh the local copy.
 * [lo, hi) = the elements of the n-element arena this launch applies (the step may be split, e.g. D's variables
 * while G's backward pass still runs); advance != 0 on the one launch per step that advances the owners' Adam step
 * counts and global_step. */
int dcgan_comm_async_apply(const void* const* master_ptrs, const void* const* bf16_ptrs, const void* const* m_ptrs,
                           const void* const* v_ptrs, const float* grad, void* local_bf16, float* local_master,
                           long long* const* peer_counters_dev, int rank, int world, size_t n, size_t lo, size_t hi,
                           int advance, float lr, float b1, float b2, float eps, int blocks, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DCGAN_B200_H_ */
