/*
 * libgabril_b200 — C ABI of the B200-native GABRIL training-step kernels.
 *
 * Drop-in boundary for the hot path of Gonglitian/GABRIL-CARLA `vlm_gaze/train/train_bc.py`
 * (BCTrainer.compute_loss + backward + optimizer step).  The reference has no native code on this
 * path: each entry point below replaces a run of PyTorch ATen/cuDNN calls, cited
 * by reference file:line.  The Python mirror of the reference API that binds these symbols
 * through ctypes lives in gabril-carla_b200/ (see INTEGRATION.md).
 *
 * Conventions
 *   - every function returns int status: 0 = GABRIL_OK, < 0 = error enum below;
 *     gabril_last_error() gives a thread-local message.  Nothing throws, nothing syncs the
 *     device, nothing allocates memory the caller can see.
 *   - all pointers are DEVICE pointers owned by the caller (PyTorch on the host side);
 *     work is ordered only by the cudaStream_t argument.
 *   - activations: NHWC bf16.  conv / linear weights: [Cout][R*S][Cin] bf16 (linear: R*S = 1).
 *     weight gradients: fp32, same element order.  Heat maps, masks, losses: fp32.
 *   - sm_100a only; gabril_init() returns GABRIL_ERR_UNSUPPORTED_ARCH elsewhere.
 */
#ifndef GABRIL_B200_H_
#define GABRIL_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st* cudaStream_t;
#endif

#define GABRIL_OK 0
#define GABRIL_ERR_BAD_SHAPE (-1)
#define GABRIL_ERR_UNSUPPORTED_ARCH (-2)
#define GABRIL_ERR_CUDA (-3)
#define GABRIL_ERR_DRIVER (-4)
#define GABRIL_ERR_NULL (-5)
#define GABRIL_ERR_KERNEL_TIMEOUT (-6)

/* ---- lifecycle ------------------------------------------------------------------------- */
int gabril_version(void);
const char* gabril_last_error(void);
/* Checks the current device is sm_100 and resolves the TMA descriptor encoders. */
int gabril_init(void);
int gabril_tc_init(void);
/* Reads and clears the device-side protocol-error flag (mbarrier wait timeouts). 0 = clean. */
int gabril_poll_error(void);

/* ---- dense contractions: tcgen05 / TMEM / TMA implicit GEMM ------------------------------ */
/* Launch attribute of the NEXT conv / linear call below made by the calling host thread (consumed
 * by it; the call must produce a bf16 output whose channel count is a multiple of 32).
 * Packed ReLU sign bits: one uint32 word per 32 consecutive channels of one pixel, word index =
 * (element offset into the output tensor) / 32; bit j (j < 16) = (element 2j of the chunk, as stored
 * in bf16, > 0), bit 16 + j = (element 2j + 1 > 0) — i.e. the low / high halves of the chunk's sixteen
 * packed bf16x2 words — the predicate of nn.ReLU's backward (linear_models.py:78-102,191-207).
 *   sign_bits_out: the call also writes the sign words of everything it stores (forward);
 *   gate_bits:     the call reads these words instead of its `gate` tensor (backward): 4 bytes of
 *                  HBM traffic per 32 channels instead of 64.
 * Either may be NULL.  The buffers must be zero wherever the output tensor is never stored
 * (zero borders). */
int gabril_conv_epilogue_bits(const void* gate_bits, void* sign_bits_out);
/* Bias gradient fused into the NEXT weight-gradient launch of the calling host thread (gabril_conv_wgrad,
 * gabril_conv_wgrad_flat, gabril_conv1_s4d_wgrad; consumed by it): dbias[c % mod] += sum over every output pixel
 * of dy[pixel][c] (mod = 0: dbias[c]) - the bias gradient of nn.Conv2d / nn.Linear backward (linear_models.py:141-207,
 * train_bc.py:79-82), taken from the dY tiles the launch stages in shared memory anyway instead of a separate
 * gabril_colsum_bf16 pass over dy.  The caller zeroes dbias (it lives in the flat gradient arena).  Cout must be
 * 32, 64 or a multiple of 128; conv_1's space-to-depth form passes mod = C1 (its 4*C1 columns are four quadrants
 * of the same channels). */
int gabril_conv_wgrad_dbias_next(float* dbias, int mod);
/* One Residual block forward - linear_models.py:78-102 (`x + _block(x)` with the block's in-place ReLU already
 * applied to x) - in ONE launch: h = relu(conv3x3(x, w3; pad 1, no bias)) [N,H,W,Ch] and
 * out = relu(x + conv1x1(h, w1; no bias)) [N,H,W,C], both stored as bf16 together with their packed sign words
 * (layout of gabril_conv_epilogue_bits; bits_h: one word per pixel, bits_out: C / 32 per pixel).  The
 * intermediate goes from TMEM through registers into shared memory as the second MMA's operand instead of
 * through HBM.  Bit-identical to gabril_conv_fwd(3x3, relu) followed by gabril_conv_fwd(1x1, resid = x, relu).
 * Built for C = 128, Ch = 32 (num_hiddens / num_residual_hiddens of the reference configs); other widths return
 * GABRIL_ERR_BAD_SHAPE and the caller uses the two convolutions. */
int gabril_res_block_fwd(const void* x, const void* w3, const void* w1, void* h, void* bits_h, void* out,
                         void* bits_out, int N, int H, int W, int C, int Ch, cudaStream_t stream);
/* nn.Conv2d / nn.Linear forward — linear_models.py:141-207, train_bc.py:79-82.
 * y = relu?( conv(x, w) + bias + resid ) * rowmask ; out_f32_atomic = 1: y is fp32 and the raw
 * products are atomically added (split-K linear; the caller zeroes y); = 2: y is fp32
 * [planes][rows][Cout], every split stores its partial product into its own plane with plain
 * stores (deterministic; planes = gabril_conv_fwd_splits(Cin, R, S, splits), summed in order by
 * gabril_bias_act_cast_planes).  Epilogue terms are ignored in both fp32 modes. */
int gabril_conv_fwd_splits(int Cin, int R, int S, int splits);
/* The same nn.Conv2d forward (linear_models.py:141-207) for NARROW stride-1 convolutions (Cout <= 64, W even, conv_6
 * of the encoder) with two adjacent output pixels stacked along the GEMM's N dimension: tile row = every second pixel
 * p, columns = [y(p) | y(p+1)]; y(p+1) = sum_t' x(p + t') w(t' - 1), so one activation tile per tap t' in 0..S and the
 * pair-packed weight w_pair[2*Cout][R*(S+1)][Cin] (gabril_conv_pack_wpairs: [w(r,t') | w(r,t'-1)], zero where out of
 * range) give both; an M = 128 tensor-core instruction is not proportionally cheaper below N = 128.  Bit-identical to
 * gabril_conv_fwd (the extra taps add exact zeros).  y = relu?(conv(x, w) + bias), bf16 NHWC. */
int gabril_conv_pack_wpairs(const void* w, void* w_pair, int Cout, int R, int S, int Cin, cudaStream_t stream);
int gabril_conv_fwd_pairs(const void* x, const void* w_pair, const float* bias, void* y, int N, int H, int W,
                          int Cin, int Cout, int R, int S, int pad, int relu, cudaStream_t stream);
/* ... with the per-pixel multiplier of the in-encoder gaze dropout (linear_models.py:194-199), rowmask fp32 [N*H*W]. */
int gabril_conv_fwd_pairs_masked(const void* x, const void* w_pair, const float* bias, const float* rowmask, void* y,
                                 int N, int H, int W, int Cin, int Cout, int R, int S, int pad, int relu,
                                 cudaStream_t stream);
int gabril_conv_fwd(const void* x, const void* w, const float* bias, const void* resid,
                    const float* rowmask, void* y, int N, int H, int W, int Cin, int Cout, int R,
                    int S, int stride, int pad, int relu, int out_f32_atomic, int splits,
                    cudaStream_t stream);
/* autograd of the above w.r.t. x (cuDNN dgrad in the reference).
 * dx = ( conv_transpose(dy, w) + addend ) gated by (gate > 0), times rowmask. */
int gabril_conv_dgrad(const void* dy, const void* w, const void* addend, const void* gate,
                      const float* rowmask, void* dx, int N, int H, int W, int Cin, int Cout,
                      int R, int S, int stride, int pad, cudaStream_t stream);
/* autograd w.r.t. w (cuDNN wgrad): dw (fp32, [Cout][R*S][Cin]) += dy^T * im2col(x). */
int gabril_conv_wgrad(const void* dy, const void* x, float* dw, int N, int H, int W, int Cin,
                      int Cout, int R, int S, int stride, int pad, int direct_store,
                      cudaStream_t stream);
/* ---- frame formatting: data_loader_robomimic.py:144-157 (_gather_last_s_frames) + :179-202
 * (_format_obs_image).  frames: uint8 [B, L, H, W, 3]; the S frames ending at `center` are taken
 * (indices clamped into [0, L-1], bit-exact integer gather). ---------------------------------- */
/* -> fp32 [B, S*C', H, W] (C' = 1 gray / 3 RGB), u8/255 then the reference's left-to-right gray
 * sum, every operation rounded separately (bit-exact vs the reference's fp32 result). W % 4 == 0. */
int gabril_obs_format(const uint8_t* frames, float* out, int B, int L, int S, int center, int H,
                      int W, int grayscale, cudaStream_t stream);
/* -> bf16 patch matrix [B*OH*OW, Kp] of the k4/s2/p1 conv_1 (linear_models.py:141-146), column
 * (r*4+s)*S*C' + frame*C' + c, zero padded to Kp: formatting and im2col fused in one pass. */
int gabril_obs_patches(const uint8_t* frames, void* patches_bf16, int B, int L, int S, int center,
                       int H, int W, int grayscale, int Kp, cudaStream_t stream);
/* Same patches, rows in space-to-depth order over the one-pixel-padded conv_1 output: row
 * ((b*SH + sh)*SW + sw)*4 + q holds output pixel (2*sh + q/2 - 1, 2*sw + q%2 - 1); rows of padding
 * pixels are left untouched (allocate the matrix zeroed).  Feeds gabril_conv1_s2d_fwd. */
int gabril_obs_patches_s2d(const uint8_t* frames, void* patches_bf16, int B, int L, int S, int center,
                           int H, int W, int grayscale, int Kp, int SH, int SW, cudaStream_t stream);

/* ---- gaze heat maps: GazePreprocessor.forward (data_loader_robomimic.py:85-139) and
 * build_stack_heatmaps (:204-278) in one kernel.  gaze: fp32 [B, L, P*2] ((x, y) in [0,1], negative
 * = padding).  ktab: [S][ktab_stride] taps of the 1-D Gaussian per stack step (Lk[s] taps, odd);
 * agg_coeff: [S][S] lower-triangular causal coefficients (NULL = ones); step_coeff: [S] per-step
 * multipliers of the multiscale mode (NULL = none); normalize_base: per-frame min-max before the
 * aggregation (alpha_decay mode); aggregate = 0 returns the per-frame maps (gaze.temporal_flag =
 * false).  out: fp32 [B, S, H, W].  Pixel indices are bit-exact; the row-band quirk of the
 * reference's padding is reproduced. */
int gabril_gaze_splat(const float* gaze, float* out, float* ext_ws /* [B][32] fp32 workspace */,
                      const float* ktab, const int* Lk,
                      int ktab_stride, const float* agg_coeff, const float* step_coeff,
                      int normalize_base, int aggregate, int B, int L, int S, int center, int P,
                      int H, int W, cudaStream_t stream);
/* Allocation-free variant of gabril_gaze_splat.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid it).  Same arguments and results as
 * gabril_gaze_splat minus ext_ws: the [B][32] fp32 extrema workspace is a file-scope
 * `static __device__ __align__(256)` array inside the library (csrc/gaze_splat.cu), 512 KB, held by every
 * GPU that loads the library until the process exits.  Limit: B <= gabril_gaze_splat_static_max_batch()
 * = 4096 samples per call; a larger B returns GABRIL_ERR_BAD_SHAPE with a message in gabril_last_error().
 * The call performs no cudaMalloc / cudaFree / allocating library routine; its set-up (once per device,
 * on the first call with that device current) is one cudaGetSymbolAddress.  Every call shares the one
 * array, which keeps the previous call's contents (nothing in it has to start at zero: each call
 * rewrites every slot it reads): calls that OVERLAP on the GPU - different streams, or different host
 * threads - race on it; serialise them on one stream or use gabril_gaze_splat with separate workspaces. */
int gabril_gaze_splat_static(const float* gaze, float* out, const float* ktab, const int* Lk,
                             int ktab_stride, const float* agg_coeff, const float* step_coeff,
                             int normalize_base, int aggregate, int B, int L, int S, int center, int P,
                             int H, int W, cudaStream_t stream);
int gabril_gaze_splat_static_max_batch(void);

/* ---- gaze-mask dropout: apply_gmd_dropout (gaze_utils.py:33-52).  g: fp32 [B, S, H, W];
 * A: the uniform draw fp32 [B, h*w]; K_ws: fp32 [B, h*w] workspace; mnmx_ws: 2 x uint32 workspace.
 * rowmask[b, j] = (A < 0.7 Khat + 0.3) (train) or the keep value itself (test_mode); keep_out
 * (optional) receives 0.7 Khat + 0.3.  min / max are global over the local batch. */
int gabril_gmd_mask(const float* g, const float* A, float* K_ws, unsigned int* mnmx_ws,
                    float* rowmask, float* keep_out, int B, int S, int H, int W, int h, int w,
                    float drop_p, int test_mode, cudaStream_t stream);
/* The same with the per-sample gaze selection of train_bc.py:258 (gg_for_cat = ivg * gg): samples
 * with sel[b] <= 0 contribute an all-zero heat map (sel may be NULL = all selected). */
int gabril_gmd_mask_sel(const float* g, const float* sel, const float* A, float* K_ws,
                        unsigned int* mnmx_ws, float* rowmask, float* keep_out, int B, int S, int H,
                        int W, int h, int w, float drop_p, int test_mode, cudaStream_t stream);
/* The same, writing the mask in the storage order of the activation it multiplies - the in-encoder
 * dropout IGMD of linear_models.py:194-199 (after conv_1 + ReLU at 90x160, after conv_2 + ReLU at
 * 45x80): PLAIN [B, h*w]; S2D [B, SH, SW, 4] (pixel (y, x) at ((y+1)>>1, (x+1)>>1, ((y+1)&1)*2 +
 * ((x+1)&1)), the patch-row order of gabril_obs_patches_s2d / gabril_conv1_s2d_fwd); BORDERED
 * [B, SH, SW] (pixel (y, x) at (y+1, x+1)).  Cells that hold no pixel are not written. */
enum { GABRIL_MASK_PLAIN = 0, GABRIL_MASK_S2D = 1, GABRIL_MASK_BORDERED = 2 };
int gabril_gmd_mask_layout(const float* g, const float* sel, const float* A, float* K_ws,
                           unsigned int* mnmx_ws, float* rowmask, float* keep_out, int B, int S,
                           int H, int W, int h, int w, float drop_p, int test_mode, int layout,
                           int SH, int SW, cudaStream_t stream);
/* The same with GMD's uniform draw made inside the kernel when A is NULL (train mode): Philox4x32-10 keyed
 * by rng_state[0] (seed) with the counter (rng_state[1] = step, draw_id, element index), 24-bit uniforms in
 * [0, 1) like torch.rand (gaze_utils.py:36).  The state lives in device memory and gabril_rng_advance
 * bumps the step once per training step, so the whole step (draw included) is CUDA-graph capturable. */
int gabril_gmd_mask_rng(const float* g, const float* sel, const float* A, const unsigned long long* rng_state,
                        int draw_id, float* K_ws, unsigned int* mnmx_ws, float* rowmask, float* keep_out, int B,
                        int S, int H, int W, int h, int w, float drop_p, int test_mode, int layout, int SH, int SW,
                        cudaStream_t stream);
int gabril_rng_advance(unsigned long long* rng_state, cudaStream_t stream);
/* out[b, c, j] = z[b, c, j] * rowmask[b, j] on the reference's fp32 NCHW layout. */
int gabril_rowmask_apply_nchw(const float* z, const float* rowmask, float* out, int B, int C, int n,
                              cudaStream_t stream);

/* ---- get_gaze_mask (gaze_utils.py:7-30): |z| channel sum -> softmax(./beta) -> bicubic up-sample
 * (fp32 index math, A = -0.75) -> per-sample min-max without eps.  z is bf16 [B, h*w, C]
 * (z_nhwc_bf16 = 1) or fp32 [B, C, h*w].  stats (optional): [B][4] = min, max, argmin, argmax. */
int gabril_gaze_mask_fwd(const void* z, int z_nhwc_bf16, float* mask_out, float* stats, int B, int C,
                         int h, int w, int H, int W, float beta, cudaStream_t stream);
/* autograd of the above: grad_mask fp32 [B, H*W] -> dz in the layout of z. */
int gabril_gaze_mask_bwd(const void* z, int z_nhwc_bf16, const float* grad_mask, void* dz, int B,
                         int C, int h, int w, int H, int W, float beta, cudaStream_t stream);
/* Reg/Teacher MSE regulariser (train_bc.py:145-150,168-169) forward + backward in one pass:
 * loss_part[b] = sum_i (g1 - get_gaze_mask(z))^2, dz = grad_scale * d(sum)/dz.  g1 plane of sample b
 * at g1 + b * g1_bstride floats; sel (optional) [B]: samples with sel <= 0 are skipped. */
int gabril_reg_mse(const void* z, int z_nhwc_bf16, const float* g1, long long g1_bstride,
                   const float* sel, void* dz, float* loss_part, float* stats, int B, int C, int h,
                   int w, int H, int W, float beta, float grad_scale, cudaStream_t stream);

/* The regulariser with the configured distance (gaze.prob_dist_type, train_bc.py:152-169):
 * GABRIL_DIST_MSE as above; KL / TV / JS first divide g1 and the mask by their detached sums + 1e-8,
 * loss_part[b] is then the per-sample sum of the divergence (the caller's mean over the selected
 * samples = sum_b loss_part / n_selected).  scale_dev (optional): device scalar multiplied into
 * grad_scale, e.g. sel_stats + 1 of gabril_ivg_select when the number of selected samples is only
 * known on the device. */
enum { GABRIL_DIST_MSE = 0, GABRIL_DIST_KL = 1, GABRIL_DIST_TV = 2, GABRIL_DIST_JS = 3 };
int gabril_reg_dist(const void* z, int z_nhwc_bf16, const float* g1, long long g1_bstride,
                    const float* sel, const float* scale_dev, int dist, void* dz, float* loss_part,
                    float* stats, int B, int C, int h, int w, int H, int W, float beta,
                    float grad_scale, cudaStream_t stream);
/* gaze.ratio < 1 (train_bc.py:230-240): sel[b] = 1 iff frac(|sum(obs_b) * 123456.789|) < ratio, with
 * obs the formatted frame stack of gabril_obs_format (uint8 frames [B, L, H, W, 3], window of S
 * frames ending at `center`); ratio >= 1 selects all, <= 0 none.  sel_stats[0] = number selected,
 * sel_stats[1] = B / number selected. */
int gabril_ivg_select(const uint8_t* frames, float* sel, float* sel_stats, int B, int L, int S,
                      int center, int H, int W, int grayscale, float ratio, cudaStream_t stream);

/* ---- device-resident demonstration store: replaces SequenceDataset.get_item
 * (vlm_gaze/robomimic/utils/dataset.py:503-633, frame_stack = S, seq_length = 1, pad_frame_stack)
 * and the DataLoader's host->device copy of the batch dict (train_bc.py:204-207).  The store holds
 * every step of every demo back to back: frames uint8 [N, H, W, 3], gaze fp32 [N, P*2], actions
 * fp32 [N, A], demo_first int64 [N] = first step of the demo step i belongs to. ------------------ */
/* widx[b][s] = max(step_idx[b] - S + 1 + s, demo_first[step_idx[b]])  (dataset.py:609-628: the
 * first frame of a demo repeated as front padding); step_idx clamped into [0, n_steps-1]. */
int gabril_store_windows(const long long* step_idx, const long long* demo_first, long long* widx,
                         int B, int S, long long n_steps, cudaStream_t stream);
/* dst[r][:] = src[widx[r]][:] for r < rows (gaze and action rows of the batch dict). */
int gabril_store_gather_rows(const float* src, const long long* widx, float* dst, long long rows,
                             int row_len, cudaStream_t stream);
/* gabril_obs_patches_s2d reading frame widx[b][s] of the store instead of frame s of batch row b:
 * the uint8 batch tensor [B, S, H, W, 3] is never materialised. */
int gabril_obs_patches_s2d_store(const uint8_t* store, const long long* widx, void* patches_bf16,
                                 int B, int S, int H, int W, int grayscale, int Kp, int SH, int SW,
                                 cudaStream_t stream);
/* gabril_ivg_select over the same window table. */
int gabril_ivg_select_store(const uint8_t* store, const long long* widx, float* sel, float* sel_stats,
                            int B, int S, int H, int W, int grayscale, float ratio,
                            cudaStream_t stream);

/* ---- heads, losses, optimiser ------------------------------------------------------------- */
/* db[c] += sum_rows dy[row][c]  (bias gradient of nn.Conv2d / nn.Linear); dy bf16 [M, C]. */
int gabril_colsum_bf16(const void* dy, float* db, long long M, int C, cudaStream_t stream);
/* y = act(x + bias) cast to bf16 (and optionally kept in fp32); x fp32 [M, C]. */
int gabril_bias_act_cast(const float* x, const float* bias, void* y_bf16, float* y_f32, long long M,
                         int C, int relu, cudaStream_t stream);
/* The same over `planes` split-K partial planes x[s][M][C] (gabril_conv_fwd with out_f32_atomic = 2),
 * summed in plane order: a deterministic replacement of the fp32 atomic accumulation. */
int gabril_bias_act_cast_planes(const float* x, int planes, const float* bias, void* y_bf16, float* y_f32,
                                long long M, int C, int relu, cudaStream_t stream);
/* nn.MSELoss on the action head (train_bc.py:292): logits = raw[:, :A] + bias;
 * part[b] = sum_k (logits - actions)^2; dlogits (bf16 [B, ld_d], zero padded) = 2 (l - a) gscale;
 * dbias[k] += sum_b dlogits. */
int gabril_mse_head(const float* raw, int ld_raw, const float* bias, const float* actions,
                    long long act_bstride, int A, float* logits_out, void* dlogits_bf16, int ld_d,
                    float* dbias, float* part, int B, float gscale, cudaStream_t stream);
/* losses[0] = actor, [1] = reg, [2] = lambda * reg + actor (train_bc.py:297); deterministic. */
/* gaze.method = GRIL (train_bc.py:84-86,182-192,293-294): loss of the gaze-coordinate head on z_flat.
 * pred = raw[:, :P2] + bias over the samples with sel > 0 (sel NULL = all); reg = clamp(mse(pred, target) + 1e-8,
 * 0, 100) -> reg_part[0] (rest 0); dpred (bf16 [B, ld_d]) = gscale * d reg / d pred; dbias += column sums. */
int gabril_gril_head(const float* raw, int ld_raw, const float* bias, const float* target, long long tgt_bstride,
                     const float* sel, int P2, void* dpred_bf16, int ld_d, float* dbias, float* reg_part, int B,
                     float gscale, cudaStream_t stream);
/* gaze.method = AGIL (train_bc.py:268-272): z = where(sel > 0, 0.5 (z + z_agil), z) [* rowmask per pixel], bf16
 * [B, n, C] in place; and its backward: dz_agil = [sel > 0] 0.5 dz, dz = (sel > 0 ? 0.5 : 1) dz in place. */
int gabril_agil_combine(void* z, const void* z_agil, const float* sel, const float* rowmask, int B, int n, int C,
                        cudaStream_t stream);
int gabril_agil_split(void* dz, void* dz_agil, const float* sel, int B, int n, int C, cudaStream_t stream);
int gabril_finalize_losses(const float* actor_part, const float* reg_part, int B, float actor_scale,
                           float reg_scale, float lambda, float* losses, cudaStream_t stream);
/* reg additionally multiplied by the device scalar reg_scale_dev[0] (NULL = 1). */
int gabril_finalize_losses_sel(const float* actor_part, const float* reg_part, int B,
                               float actor_scale, float reg_scale, const float* reg_scale_dev,
                               float lambda, float* losses, cudaStream_t stream);
/* torch.optim.Adam (decoupled = 0) / AdamW (1) step `step` (1-based) over a flat fp32 arena
 * (optim.py:11-28); grads are multiplied by grad_scale first; the bf16 shadow is refreshed. */
int gabril_adam(float* p, const float* g, float* m, float* v, void* shadow_bf16, long long n,
                float lr, float beta1, float beta2, float eps, float weight_decay, int decoupled,
                int step, float grad_scale, cudaStream_t stream);
int gabril_cast_f32_bf16(const float* x, void* y_bf16, long long n, cudaStream_t stream);
/* t[row][c] *= scale[row] in place; t bf16 [rows, C] (GMD backward, gaze_utils.py:51: dz = dz' * M). */
int gabril_rowscale_bf16(void* t, const float* scale, long long rows, int C, cudaStream_t stream);

/* ---- flat-slab convolutions (stride 1) over zero-bordered NHWC tensors: one TMA slab per filter
 * row, the S taps of the row are descriptor views shifted by whole pixel rows.  Replaces the same
 * nn.Conv2d forward / backward (linear_models.py:141-207) as gabril_conv_fwd/_dgrad/_wgrad; the
 * k4/s2/p1 conv_2 runs through it as a 2x2 stride-1 convolution over the space-to-depth conv_1
 * output.  See csrc/tc_plan.cu for the exact index conventions. ------------------------------- */
int gabril_conv_flat(const void* x, const void* w, const float* bias, const void* resid,
                     const void* gate, void* y, int N, int Hg, int Wg, int Cx, int Cy, int R, int S,
                     int valid_h, int valid_w, int o_rows, int o_pitch, int o_off_h, int o_off_w,
                     int relu, int transpose_w, cudaStream_t stream);
int gabril_conv_wgrad_flat(const void* dy, const void* x, float* dw, int N, int Hg, int Wg, int Cx,
                           int Cy, int R, int S, int dy_row_off, cudaStream_t stream);
int gabril_conv1_s2d_fwd(const void* patches, const void* w, const float* bias, void* y, int N,
                         int SH, int SW, int img_h, int img_w, int Kp, int Cout, int relu,
                         cudaStream_t stream);
/* The two forward layers with a per-pixel multiplier applied after bias / ReLU (IGMD,
 * linear_models.py:194-199).  rowmask is indexed by the pixel index of the OUTPUT storage: patch-row
 * order for conv_1 (GABRIL_MASK_S2D), bordered [N, o_rows, o_pitch] for the flat convolution
 * (GABRIL_MASK_BORDERED); NULL = no mask. */
int gabril_conv_flat_masked(const void* x, const void* w, const float* bias, const void* resid,
                            const void* gate, const float* rowmask, void* y, int N, int Hg, int Wg,
                            int Cx, int Cy, int R, int S, int valid_h, int valid_w, int o_rows,
                            int o_pitch, int o_off_h, int o_off_w, int relu, int transpose_w,
                            cudaStream_t stream);
int gabril_conv1_s2d_fwd_masked(const void* patches, const void* w, const float* bias,
                                const float* rowmask, void* y, int N, int SH, int SW, int img_h,
                                int img_w, int Kp, int Cout, int relu, cudaStream_t stream);

/* ---- conv_1 without a patch matrix (csrc/conv1_s4d.cu) ------------------------------------------
 * Replaces _gather_last_s_frames + _format_obs_image (data_loader_robomimic.py:144-157,179-202)
 * feeding Encoder._conv_1 + ReLU (linear_models.py:141-146,191-193).  The uint8 frames are written
 * once as a 4x4 space-to-depth bf16 image on conv_1's own space-to-depth grid,
 *   x4[n, u, v, (a*4 + b)*C0 + ci] = obs[n, ci, 4(u-1) + a, 4(v-1) + b]   (0 outside), SH = H/4 + 2,
 *   SW = W/4 + 2, Cp >= 16*C0 channels per cell; the buffer holds N*SH*SW + 1 cells (the last one zero);
 * conv_1 is then a 2x2 stride-1 flat convolution with 4*C1 output channels (quadrant, co) whose weights
 * w4 [4*C1][2][2*Cp] (bf16) gabril_conv1_s4d_expand derives from the fp32 master w1 [C1][K1p]
 * (column (kh*4 + kw)*C0 + ci); gabril_conv1_s4d_wgrad accumulates dw4 (fp32, same shape) and
 * gabril_conv1_s4d_fold adds its four quadrants into dw1 [C1][K1p].  rowmask: one multiplier per
 * conv_1 pixel in space-to-depth order [N, SH, SW, 4] (in-encoder gaze dropout) or NULL. */
int gabril_obs_s4d(const uint8_t* frames, void* x4, int B, int L, int S, int center, int H, int W,
                   int grayscale, int Cp, int SH, int SW, cudaStream_t stream);
int gabril_obs_s4d_store(const uint8_t* store, const long long* widx, void* x4, int B, int S, int H,
                         int W, int grayscale, int Cp, int SH, int SW, cudaStream_t stream);
/* The same with the gaze heat maps folded in (gray stacks; heat fp32 [B, S, H, W], sel = ivg fp32 [B] or NULL;
 * widx = NULL: batch layout, else the frame store through the window table):
 *   mode 1  gaze.method = Mask (train_bc.py:260-261) and AGIL's second encoder (:270):
 *           obs * (ivg * heat + (1 - ivg));
 *   mode 2  gaze.method = ViSaRL (train_bc.py:262-264): channels (obs, ivg * heat), C0 = 2 S. */
int gabril_obs_s4d_gaze(const uint8_t* frames, const long long* widx, const float* heat, const float* sel, int mode,
                        void* x4, int B, int L, int S, int center, int H, int W, int grayscale, int Cp, int SH,
                        int SW, cudaStream_t stream);
int gabril_conv1_s4d_expand(const float* w1, void* w4, int C1, int C0, int K1p, int Cp,
                            cudaStream_t stream);
int gabril_conv1_s4d_fwd(const void* x4, const void* w4, const float* bias, const float* rowmask,
                         void* y, int N, int SH, int SW, int Cp, int C1, int img_h, int img_w, int relu,
                         cudaStream_t stream);
int gabril_conv1_s4d_wgrad(const void* dy, const void* x4, float* dw4, int N, int SH, int SW, int Cp,
                           int C1, cudaStream_t stream);
int gabril_conv1_s4d_fold(const float* dw4, float* dw1, int C1, int C0, int K1p, int Cp,
                          cudaStream_t stream);

/* ---- gaze-predictor pre-training (train_gaze_predictor.py:83-101): AutoEncoder = Encoder + Decoder
 * (linear_models.py:210-282, 356-367), MSE against the last aggregated heat map.
 * nn.ConvTranspose2d with weight Wt[Cin_t][Cout_t][R][S] runs on the convolution kernels through the
 * internal weight w[Cin_t][R*S][Cout_t]: forward = data gradient of the conv Cout_t -> Cin_t (+ bias,
 * ReLU), backward-data = that conv's forward (+ ReLU gate), weight gradient = gabril_conv_wgrad with the
 * roles of x and dy exchanged.  x [N,OH,OW,Cin_t] -> y [N,H,W,Cout_t]; k4/s2/p0|p1 and stride 1. */
int gabril_convt_fwd(const void* x, const void* w, const float* bias, void* y, int N, int H, int W,
                     int Cout_t, int Cin_t, int R, int S, int stride, int pad, int relu,
                     cudaStream_t stream);
int gabril_convt_dgrad(const void* dy, const void* w, const void* gate, void* dx, int N, int H, int W,
                       int Cout_t, int Cin_t, int R, int S, int stride, int pad, cudaStream_t stream);
/* _conv_trans_4 (C -> 1 channel, k4/s2/p1) + nn.MSELoss: x bf16 [B,h,w,C], wp bf16 [C][Kp] (tap ky*4+kx),
 * target plane b at target + b*target_bstride ([2h][2w] fp32).  Writes out (optional) and
 * dout = 2*(out - target)*gscale as fp32 [B,2h,2w]; ADDS the per-sample squared error into
 * loss_part[B] and sum(dout) into dbias[1] (both zeroed by the caller). */
int gabril_gp_head_fwd(const void* x, const void* wp, int Kp, const float* bias, const float* target,
                       long long target_bstride, float* out, float* dout, float* loss_part,
                       float* dbias, int B, int h, int w, int C, float gscale, cudaStream_t stream);
/* dout fp32 [B,2h,2w] -> bf16 patch matrix [B*h*w, Kp] of the k4/s2/p1 window of every input pixel
 * (columns 16.. zero): the backward of _conv_trans_4 then is two patch GEMMs. */
int gabril_gp_dout_patches(const float* dout, void* patches_bf16, int Kp, int B, int h, int w,
                           cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* GABRIL_B200_H_ */
