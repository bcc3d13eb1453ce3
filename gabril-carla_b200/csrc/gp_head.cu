// Gaze-predictor head (BASELINE config 4): the last Decoder layer
//   _conv_trans_4 = nn.ConvTranspose2d(num_hiddens/4, 1, kernel_size=4, stride=2, padding=1)
// (vlm_gaze/models/linear_models.py:257-263) fused with the nn.MSELoss of
// GazePredictorTrainer.compute_loss (vlm_gaze/train/train_gaze_predictor.py:83-101).
// One output channel is far below a UMMA tile, and the layer is HBM-bound (921 600 B of bf16
// input per sample for 57 600 outputs), so the forward is a gather-form CUDA-core kernel:
//   out(oy, ox) = bias + sum_c sum_{(ky,kx) of the pixel's parity class} x(iy, ix, c) * Wt[c][ky][kx]
// with oy = 2*iy - 1 + ky.  The same pass emits the squared-error partial sums, dL/dout and its sum
// (the bias gradient).  gp_dout_patches then lays dL/dout out as the 4x4/stride-2 patch matrix of
// the input pixels, which turns the two backward products into the tcgen05 patch GEMMs conv_1
// already uses: dx = gate(patches . W^T) (gabril_convt_dgrad, R = S = 1) and dW = x^T . patches
// (gabril_conv_wgrad).
#include "common.cuh"

namespace {

constexpr int kHeadThreads = 256;

// x [B, h, w, C] bf16 (C <= 64, C % 8 == 0), wp [C][Kp] bf16 (tap ky*4+kx in columns 0..15)
template <int C>
__global__ void __launch_bounds__(kHeadThreads)
gp_head_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ wp, int Kp,
                   const float* __restrict__ bias, const float* __restrict__ target,
                   long long t_bstride, float* __restrict__ out, float* __restrict__ dout,
                   float* __restrict__ loss_part, float* __restrict__ dbias, int B, int h, int w,
                   float gscale) {
  __shared__ float ws[16 * C];  // [tap][c]
  __shared__ float red[kHeadThreads / 32];
  for (int i = threadIdx.x; i < 16 * C; i += blockDim.x) {
    const int tap = i / C, c = i - tap * C;
    ws[i] = __bfloat162float(wp[(long long)c * Kp + tap]);
  }
  __syncthreads();
  const int H = 2 * h, W = 2 * w;
  const int b = blockIdx.y;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  float sq = 0.f, dsum = 0.f;
  if (idx < H * W) {
    const int oy = idx / W, ox = idx - oy * W;
    const int py = (oy + 1) & 1, px = (ox + 1) & 1;
    const int iy0 = (oy + 1 - py) >> 1, ix0 = (ox + 1 - px) >> 1;
    float acc = bias[0];
#pragma unroll
    for (int ry = 0; ry < 2; ++ry) {
      const int iy = iy0 - ry, ky = py + 2 * ry;
      if (iy < 0 || iy >= h) continue;
#pragma unroll
      for (int rx = 0; rx < 2; ++rx) {
        const int ix = ix0 - rx, kx = px + 2 * rx;
        if (ix < 0 || ix >= w) continue;
        const uint4* xp = reinterpret_cast<const uint4*>(x + (((long long)b * h + iy) * w + ix) * C);
        const float* wt = ws + (ky * 4 + kx) * C;
#pragma unroll
        for (int v = 0; v < C / 8; ++v) {
          const uint4 q = __ldg(xp + v);
          acc = fmaf(bf16lo(q.x), wt[8 * v + 0], acc); acc = fmaf(bf16hi(q.x), wt[8 * v + 1], acc);
          acc = fmaf(bf16lo(q.y), wt[8 * v + 2], acc); acc = fmaf(bf16hi(q.y), wt[8 * v + 3], acc);
          acc = fmaf(bf16lo(q.z), wt[8 * v + 4], acc); acc = fmaf(bf16hi(q.z), wt[8 * v + 5], acc);
          acc = fmaf(bf16lo(q.w), wt[8 * v + 6], acc); acc = fmaf(bf16hi(q.w), wt[8 * v + 7], acc);
        }
      }
    }
    const long long o = (long long)b * H * W + idx;
    if (out) out[o] = acc;
    const float d = acc - __ldg(target + (long long)b * t_bstride + idx);
    sq = d * d;
    const float g = 2.f * d * gscale;
    dout[o] = g;
    dsum = g;
  }
  // block sums: squared error of this sample, bias gradient
  sq = warp_sum(sq);
  dsum = warp_sum(dsum);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) red[wid] = sq;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < kHeadThreads / 32; ++i) s += red[i];
    atomicAdd(loss_part + b, s);
  }
  __syncthreads();
  if (lane == 0) red[wid] = dsum;
  __syncthreads();
  if (threadIdx.x == 0 && dbias) {
    float s = 0.f;
    for (int i = 0; i < kHeadThreads / 32; ++i) s += red[i];
    atomicAdd(dbias, s);
  }
}

// dout fp32 [B, 2h, 2w] -> bf16 patch matrix [B*h*w, Kp]: column ky*4+kx of row (b, iy, ix) is
// dout(b, 2*iy - 1 + ky, 2*ix - 1 + kx) (0 outside the map); columns 16.. are zero.
__global__ void __launch_bounds__(256)
gp_dout_patches_kernel(const float* __restrict__ dout, bf16* __restrict__ patches, int Kp, int B,
                       int h, int w) {
  const int H = 2 * h, W = 2 * w;
  const long long total = (long long)B * h * w * 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int ky = (int)(i & 3);
    const long long pix = i >> 2;
    const int ix = (int)(pix % w);
    const long long t = pix / w;
    const int iy = (int)(t % h);
    const int b = (int)(t / h);
    const int oy = 2 * iy - 1 + ky;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (oy >= 0 && oy < H) {
      const float* row = dout + ((long long)b * H + oy) * W;
#pragma unroll
      for (int kx = 0; kx < 4; ++kx) {
        const int ox = 2 * ix - 1 + kx;
        if (ox >= 0 && ox < W) v[kx] = __ldg(row + ox);
      }
    }
    bf16* prow = patches + pix * Kp;
    *reinterpret_cast<uint2*>(prow + 4 * ky) = make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]));
    for (int k = 16 + 4 * ky; k < Kp; k += 16) *reinterpret_cast<uint2*>(prow + k) = make_uint2(0u, 0u);
  }
}

}  // namespace

extern "C" {

int gabril_gp_head_fwd(const void* x, const void* wp, int Kp, const float* bias, const float* target,
                       long long target_bstride, float* out, float* dout, float* loss_part,
                       float* dbias, int B, int h, int w, int C, float gscale, cudaStream_t stream) {
  if (!x || !wp || !bias || !target || !dout || !loss_part) return GABRIL_ERR_NULL;
  if (B <= 0 || Kp < 16 || (C != 32 && C != 64)) return GABRIL_ERR_BAD_SHAPE;
  const int HW = 4 * h * w;
  const dim3 grid((HW + kHeadThreads - 1) / kHeadThreads, B);
  if (C == 32)
    gp_head_fwd_kernel<32><<<grid, kHeadThreads, 0, stream>>>(
        (const bf16*)x, (const bf16*)wp, Kp, bias, target, target_bstride, out, dout, loss_part, dbias, B, h, w, gscale);
  else
    gp_head_fwd_kernel<64><<<grid, kHeadThreads, 0, stream>>>(
        (const bf16*)x, (const bf16*)wp, Kp, bias, target, target_bstride, out, dout, loss_part, dbias, B, h, w, gscale);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_gp_dout_patches(const float* dout, void* patches, int Kp, int B, int h, int w,
                           cudaStream_t stream) {
  if (!dout || !patches) return GABRIL_ERR_NULL;
  if (B <= 0 || Kp < 16 || Kp % 16) return GABRIL_ERR_BAD_SHAPE;
  const long long total = (long long)B * h * w * 4;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  gp_dout_patches_kernel<<<(int)blocks, 256, 0, stream>>>(dout, (bf16*)patches, Kp, B, h, w);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

}  // extern "C"
