// Shared status codes and small device helpers for libgabril_b200.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>

#include "../../include/gabril_b200.h"

#define GABRIL_CHECK_LAUNCH()                                   \
  do {                                                          \
    cudaError_t e__ = cudaGetLastError();                       \
    if (e__ != cudaSuccess) return gabril_set_cuda_error(e__);  \
  } while (0)

int gabril_set_cuda_error(cudaError_t e);
void gabril_set_error(const char* msg);

typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Order-preserving float <-> uint encoding so atomicMin/atomicMax work on floats.
__device__ __forceinline__ unsigned int f32_to_ordered(float f) {
  unsigned int u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_f32(unsigned int u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

// u8 / 255.0f with the bits of the IEEE division (data_loader_robomimic.py:182): one Newton correction of the
// reciprocal product.  Bit-exact for all 256 inputs (tests/test_ops_gpu.py::test_obs_format_bit_exact compares fp32 outputs);
// a 256-entry shared-memory table of the same values made gabril_obs_s4d bound by the L1 data pipe (24
// bank-conflicted lookups per 16 output bytes).
__device__ __forceinline__ float u8_over_255(uint32_t v) {
  const float x = (float)v, rcp = 1.0f / 255.0f;
  const float q0 = __fmul_rn(x, rcp);
  const float rem = __fmaf_rn(-q0, 255.0f, x);
  return __fmaf_rn(rem, rcp, q0);
}

// ---- PyTorch upsample_bicubic2d index math (fp32, A = -0.75, align_corners = False) -------------
__device__ __forceinline__ float cubic1(float x) {
  const float A = -0.75f;
  return ((A + 2.f) * x - (A + 3.f)) * x * x + 1.f;
}
__device__ __forceinline__ float cubic2(float x) {
  const float A = -0.75f;
  return ((A * x - 5.f * A) * x + 8.f * A) * x - 4.f * A;
}
// taps i0-1 .. i0+2 clamped into [0, in-1]; weights in tap order.
__device__ __forceinline__ void bicubic_taps(int dst, float scale, int in_size, int idx[4],
                                             float w[4]) {
  const float src = scale * ((float)dst + 0.5f) - 0.5f;
  const float fl = floorf(src);
  const int i0 = (int)fl;
  const float t = src - fl;
  w[0] = cubic2(t + 1.f);
  w[1] = cubic1(t);
  const float t2 = 1.f - t;
  w[2] = cubic1(t2);
  w[3] = cubic2(t2 + 1.f);
#pragma unroll
  for (int k = 0; k < 4; ++k) idx[k] = min(max(i0 - 1 + k, 0), in_size - 1);
}
