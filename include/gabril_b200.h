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
/* nn.Conv2d / nn.Linear forward — linear_models.py:141-207, train_bc.py:79-82.
 * y = relu?( conv(x, w) + bias + resid ) * rowmask ; out_f32_atomic: y is fp32 and the raw
 * products are atomically added (split-K linear), epilogue terms are then ignored. */
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
/* CUDA-core cross-check versions of the three contracts (bring-up only). */
int gabril_conv_fwd_naive(const void* x, const void* w, const float* bias, const void* resid,
                          const float* rowmask, void* y, int N, int H, int W, int Cin, int Cout,
                          int R, int S, int stride, int pad, int relu, int out_f32_atomic,
                          int splits, cudaStream_t stream);
int gabril_conv_dgrad_naive(const void* dy, const void* w, const void* addend, const void* gate,
                            const float* rowmask, void* dx, int N, int H, int W, int Cin, int Cout,
                            int R, int S, int stride, int pad, cudaStream_t stream);
int gabril_conv_wgrad_naive(const void* dy, const void* x, float* dw, int N, int H, int W, int Cin,
                            int Cout, int R, int S, int stride, int pad, int direct_store,
                            cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* GABRIL_B200_H_ */
