// Bring-up / cross-check kernels: the same contracts as gabril_conv_fwd / _dgrad / _wgrad written
// as direct convolutions on CUDA cores (one thread per output element).  They exist so that the
// tcgen05 path can be validated on-device layer by layer (tests/test_conv_gpu.py) and are selected
// only when the host sets GABRIL_NAIVE_CONV=1; they are never a silent fallback.
#include "common.cuh"

namespace {

__global__ void conv_fwd_naive_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                      const float* __restrict__ bias,
                                      const bf16* __restrict__ resid,
                                      const float* __restrict__ rowmask, void* __restrict__ y,
                                      int N, int H, int W, int Cin, int Cout, int R, int S,
                                      int stride, int pad, int OH, int OW, int relu, int out_f32) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)N * OH * OW * Cout;
  if (idx >= total) return;
  const int co = (int)(idx % Cout);
  const long long pix = idx / Cout;
  const int ow = (int)(pix % OW);
  const int oh = (int)((pix / OW) % OH);
  const int n = (int)(pix / ((long long)OW * OH));
  float acc = 0.f;
  for (int r = 0; r < R; ++r) {
    const int ih = oh * stride - pad + r;
    if (ih < 0 || ih >= H) continue;
    for (int s = 0; s < S; ++s) {
      const int iw = ow * stride - pad + s;
      if (iw < 0 || iw >= W) continue;
      const bf16* xp = x + (((long long)n * H + ih) * W + iw) * Cin;
      const bf16* wp = w + ((long long)co * R * S + r * S + s) * Cin;
      for (int ci = 0; ci < Cin; ++ci) acc += __bfloat162float(xp[ci]) * __bfloat162float(wp[ci]);
    }
  }
  if (out_f32) {
    atomicAdd(reinterpret_cast<float*>(y) + idx, acc);
    return;
  }
  if (bias) acc += bias[co];
  if (resid) acc += __bfloat162float(resid[idx]);
  if (relu) acc = fmaxf(acc, 0.f);
  if (rowmask) acc *= rowmask[pix];
  reinterpret_cast<bf16*>(y)[idx] = __float2bfloat16(acc);
}

__global__ void conv_dgrad_naive_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ w,
                                        const bf16* __restrict__ addend,
                                        const bf16* __restrict__ gate,
                                        const float* __restrict__ rowmask, bf16* __restrict__ dx,
                                        int N, int H, int W, int Cin, int Cout, int R, int S,
                                        int stride, int pad, int OH, int OW) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)N * H * W * Cin;
  if (idx >= total) return;
  const int ci = (int)(idx % Cin);
  const long long pix = idx / Cin;
  const int iw = (int)(pix % W);
  const int ih = (int)((pix / W) % H);
  const int n = (int)(pix / ((long long)W * H));
  float acc = 0.f;
  for (int r = 0; r < R; ++r) {
    const int t = ih + pad - r;
    if (t < 0 || t % stride) continue;
    const int oh = t / stride;
    if (oh >= OH) continue;
    for (int s = 0; s < S; ++s) {
      const int u = iw + pad - s;
      if (u < 0 || u % stride) continue;
      const int ow = u / stride;
      if (ow >= OW) continue;
      const bf16* dp = dy + (((long long)n * OH + oh) * OW + ow) * Cout;
      const bf16* wp = w + ((long long)r * S + s) * Cin + ci;
      for (int co = 0; co < Cout; ++co)
        acc += __bfloat162float(dp[co]) * __bfloat162float(wp[(long long)co * R * S * Cin]);
    }
  }
  if (addend) acc += __bfloat162float(addend[idx]);
  if (gate) acc = __bfloat162float(gate[idx]) > 0.f ? acc : 0.f;
  if (rowmask) acc *= rowmask[pix];
  dx[idx] = __float2bfloat16(acc);
}

// one block per (co, tap, ci) output element, block-wide reduction over pixels
__global__ void conv_wgrad_naive_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x,
                                        float* __restrict__ dw, int N, int H, int W, int Cin,
                                        int Cout, int R, int S, int stride, int pad, int OH,
                                        int OW) {
  const long long e = blockIdx.x;
  const int ci = (int)(e % Cin);
  const int tap = (int)((e / Cin) % (R * S));
  const int co = (int)(e / ((long long)Cin * R * S));
  const int r = tap / S, s = tap % S;
  float acc = 0.f;
  const long long P = (long long)N * OH * OW;
  for (long long pidx = threadIdx.x; pidx < P; pidx += blockDim.x) {
    const int ow = (int)(pidx % OW);
    const int oh = (int)((pidx / OW) % OH);
    const int n = (int)(pidx / ((long long)OW * OH));
    const int ih = oh * stride - pad + r, iw = ow * stride - pad + s;
    if (ih < 0 || ih >= H || iw < 0 || iw >= W) continue;
    acc += __bfloat162float(dy[pidx * Cout + co]) *
           __bfloat162float(x[(((long long)n * H + ih) * W + iw) * Cin + ci]);
  }
  __shared__ float red[32];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) dw[e] += v;
  }
}

}  // namespace

extern "C" {

int gabril_conv_fwd_naive(const void* x, const void* w, const float* bias, const void* resid,
                          const float* rowmask, void* y, int N, int H, int W, int Cin, int Cout,
                          int R, int S, int stride, int pad, int relu, int out_f32_atomic,
                          int splits, cudaStream_t stream) {
  (void)splits;
  const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
  const long long total = (long long)N * OH * OW * Cout;
  conv_fwd_naive_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(
      (const bf16*)x, (const bf16*)w, bias, (const bf16*)resid, rowmask, y, N, H, W, Cin, Cout, R,
      S, stride, pad, OH, OW, relu, out_f32_atomic);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_conv_dgrad_naive(const void* dy, const void* w, const void* addend, const void* gate,
                            const float* rowmask, void* dx, int N, int H, int W, int Cin, int Cout,
                            int R, int S, int stride, int pad, cudaStream_t stream) {
  const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
  const long long total = (long long)N * H * W * Cin;
  conv_dgrad_naive_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(
      (const bf16*)dy, (const bf16*)w, (const bf16*)addend, (const bf16*)gate, rowmask, (bf16*)dx,
      N, H, W, Cin, Cout, R, S, stride, pad, OH, OW);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_conv_wgrad_naive(const void* dy, const void* x, float* dw, int N, int H, int W, int Cin,
                            int Cout, int R, int S, int stride, int pad, int direct_store,
                            cudaStream_t stream) {
  const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
  const long long total = (long long)Cout * R * S * Cin;
  if (direct_store) cudaMemsetAsync(dw, 0, total * sizeof(float), stream);
  conv_wgrad_naive_kernel<<<(unsigned)total, 256, 0, stream>>>(
      (const bf16*)dy, (const bf16*)x, dw, N, H, W, Cin, Cout, R, S, stride, pad, OH, OW);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

}  // extern "C"
