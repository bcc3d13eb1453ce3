// conv_1 (k4 / s2 / p1 over the raw frame stack) without a patch matrix.
//
// Reference ops replaced: _gather_last_s_frames + _format_obs_image
// (vlm_gaze/data_utils/data_loader_robomimic.py:144-157,179-202) feeding Encoder._conv_1 + ReLU
// (vlm_gaze/models/linear_models.py:141-146,191-193).
//
// The frames are written ONCE as a 4x4 space-to-depth bf16 image X4 on the SAME [SH, SW] grid as
// conv_1's space-to-depth output a1s (SH = H/4 + 2, SW = W/4 + 2 for 180x320: 47 x 82):
//     X4[n, u, v, (a*4 + b)*C0 + ci] = obs[n, ci, 4(u-1) + a, 4(v-1) + b]        (0 outside the image)
// Output super-pixel (sh, sw) = the four conv_1 pixels (2sh+qh-1, 2sw+qw-1) reads input rows
// 4sh-3 .. 4sh+2, i.e. cells (sh + r', sw + s'), r', s' in {0, 1}: conv_1 + space-to-depth is a 2x2
// stride-1 convolution over X4 with 4*C1 output channels (qh, qw, co) whose weights
//     W4[(qh,qw,co)][r'][(s', a, b, ci)] = W1[co, ci, 4r' + a - 2qh - 1, 4s' + b - 2qw - 1]   (0 if out of 0..3)
// are a 4x zero-inflated copy of conv_1's (tensor work is 0.6 % of the step; HBM bytes are what this
// layer costs).  Cells of row u = 0 / column v = 0 are all zero, so the flat (wrapping) index q + SW + 1
// needs no extra border, and two horizontally adjacent cells are ONE row of an overlapping-row tensor
// map (row pitch = one cell), which makes (s', a, b, ci) a single 64-channel (gray) contraction chunk.
// Traffic per sample (gray, stack 2): 345.6 KB u8 read + 247 KB written, instead of the 864 KB patch
// matrix written once and read by conv_1 forward and conv_1 wgrad.
#include <algorithm>

#include "common.cuh"
#include "ptx.cuh"

namespace {

__device__ __forceinline__ int frame_index(int s, int S, int center, int L) {
  return min(max(center - S + 1 + s, 0), L - 1);   // data_loader_robomimic.py:151-152
}
__device__ __forceinline__ float to_gray(float r, float g, float b) {
  // ((0.299 R) + (0.587 G)) + (0.114 B), every op rounded separately (data_loader_robomimic.py:192-193)
  return __fadd_rn(__fadd_rn(__fmul_rn(0.299f, r), __fmul_rn(0.587f, g)), __fmul_rn(0.114f, b));
}

// One thread = one cell row (u, v, a): 4 pixels x S frames -> 4*C0 consecutive bf16 of the cell.
// MODE (gray stacks only; train_bc.py:252-264, heat = the gaze heat maps [B, S, H, W], sel = ivg or NULL = all):
//   1  gaze.method = Mask / AGIL's second encoder: obs * (ivg * heat + (1 - ivg))
//   2  gaze.method = ViSaRL: channels (obs_0 .. obs_S-1, ivg*heat_0 .. ivg*heat_S-1), C0 = 2 S
template <int S_, bool GRAY, int MODE>
__global__ void __launch_bounds__(384)
obs_s4d_kernel(const uint8_t* __restrict__ frames, const long long* __restrict__ widx,
               const float* __restrict__ heat, const float* __restrict__ sel,
               bf16* __restrict__ out, long long cells, int L, int center, int H, int W, int SH, int SW,
               int Cp) {
  constexpr int CP = GRAY ? 1 : 3;
  constexpr int C0 = S_ * CP * (MODE == 2 ? 2 : 1);
  // One CTA pass = one row of cells (b, u): the (b, u) split is one 32-bit division per pass instead of two
  // 64-bit divisions per work item (they were most of the kernel's instructions: 3.0 TB/s, ALU-bound).
  const int rows = (int)(cells / SW), items = SW * 4;
  for (int row = blockIdx.x; row < rows; row += gridDim.x)
  for (int it = threadIdx.x; it < items; it += blockDim.x) {
    const int a = it & 3, v = it >> 2;
    const int u = row % SH;
    const long long b = row / SH;
    const long long cell = (long long)row * SW + v;
    const int y = 4 * (u - 1) + a, x0 = 4 * (v - 1);
    const bool ok = u >= 1 && v >= 1 && y < H && x0 < W;
    float px[4][C0];
    if constexpr (GRAY && MODE == 0) {
      // Plain gray stacks (the headline path): the kernel is bound by instruction issue (ncu: 330 instructions per
      // 16 output bytes, issue slots 77 % busy, DRAM 35 %), so the /255 and gray arithmetic runs on packed pairs
      // of pixels (mul / fma / add .f32x2: same operations, same roundings) and padding cells skip it all.
      if (!ok) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
          for (int c = 0; c < C0; ++c) px[k][c] = 0.f;
      } else {
        const uint64_t rcp2 = ptx::pack2(__float_as_uint(1.0f / 255.0f), __float_as_uint(1.0f / 255.0f));
        const uint64_t m255 = ptx::pack2(__float_as_uint(-255.0f), __float_as_uint(-255.0f));
        const uint64_t cr = ptx::pack2(__float_as_uint(0.299f), __float_as_uint(0.299f));
        const uint64_t cg = ptx::pack2(__float_as_uint(0.587f), __float_as_uint(0.587f));
        const uint64_t cb = ptx::pack2(__float_as_uint(0.114f), __float_as_uint(0.114f));
#pragma unroll
        for (int fs = 0; fs < S_; ++fs) {
          const long long fr = widx ? __ldg(widx + b * S_ + fs) : b * L + frame_index(fs, S_, center, L);
          const uint32_t* src = reinterpret_cast<const uint32_t*>(frames + ((fr * H + y) * (long long)W + x0) * 3);
          const uint32_t w0 = __ldg(src), w1 = __ldg(src + 1), w2 = __ldg(src + 2);
          // bytes: p0 = (w0.0 w0.1 w0.2) p1 = (w0.3 w1.0 w1.1) p2 = (w1.2 w1.3 w2.0) p3 = (w2.1 w2.2 w2.3)
          auto unit2 = [&](uint32_t ba, uint32_t bb) {   // (ba / 255, bb / 255), see u8_over_255
            const uint64_t x = ptx::pack2(__float_as_uint((float)ba), __float_as_uint((float)bb));
            const uint64_t q0 = ptx::mul2(x, rcp2);
            const uint64_t rem = ptx::fma2(q0, m255, x);
            return ptx::fma2(rem, rcp2, q0);
          };
          auto gray2 = [&](uint64_t r, uint64_t g, uint64_t bl) {   // ((0.299 R) + (0.587 G)) + (0.114 B)
            return ptx::add2(ptx::add2(ptx::mul2(cr, r), ptx::mul2(cg, g)), ptx::mul2(cb, bl));
          };
          const uint64_t g01 = gray2(unit2(w0 & 0xffu, w0 >> 24), unit2((w0 >> 8) & 0xffu, w1 & 0xffu),
                                     unit2((w0 >> 16) & 0xffu, (w1 >> 8) & 0xffu));
          const uint64_t g23 = gray2(unit2((w1 >> 16) & 0xffu, (w2 >> 8) & 0xffu), unit2(w1 >> 24, (w2 >> 16) & 0xffu),
                                     unit2(w2 & 0xffu, w2 >> 24));
          px[0][fs] = __uint_as_float((uint32_t)g01); px[1][fs] = __uint_as_float((uint32_t)(g01 >> 32));
          px[2][fs] = __uint_as_float((uint32_t)g23); px[3][fs] = __uint_as_float((uint32_t)(g23 >> 32));
        }
      }
    } else {
#pragma unroll
    for (int fs = 0; fs < S_; ++fs) {
      uint32_t wv[3] = {0u, 0u, 0u};
      if (ok) {
        // batch layout [B, L, H, W, 3], or (widx != NULL) a frame store [N, H, W, 3] addressed through
        // the window table widx[b][fs] of gabril_store_windows
        const long long fr = widx ? __ldg(widx + b * S_ + fs) : b * L + frame_index(fs, S_, center, L);
        const uint32_t* src = reinterpret_cast<const uint32_t*>(frames + ((fr * H + y) * (long long)W + x0) * 3);
        wv[0] = __ldg(src); wv[1] = __ldg(src + 1); wv[2] = __ldg(src + 2);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float ch[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const int byte = 3 * k + c;
          ch[c] = u8_over_255((wv[byte >> 2] >> (8 * (byte & 3))) & 0xffu);
        }
        if (GRAY) {
          px[k][fs] = ok ? to_gray(ch[0], ch[1], ch[2]) : 0.f;
          if (MODE != 0) {
            float gv = 0.f, iv = 1.f;
            if (ok) {
              gv = __ldg(heat + ((b * S_ + fs) * H + y) * (long long)W + x0 + k);
              if (sel) iv = __ldg(sel + b) > 0.f ? 1.f : 0.f;
            }
            if (MODE == 1) {   // x * (ivg * g + (1 - ivg) * 1), every op rounded like the reference's torch ops
              const float m = __fadd_rn(__fmul_rn(iv, gv), __fsub_rn(1.f, iv));
              px[k][fs] = __fmul_rn(px[k][fs], m);
            } else {
              px[k][S_ + fs] = ok ? __fmul_rn(iv, gv) : 0.f;
            }
          }
        } else {
          px[k][fs * 3 + 0] = ok ? ch[0] : 0.f;
          px[k][fs * 3 + 1] = ok ? ch[1] : 0.f;
          px[k][fs * 3 + 2] = ok ? ch[2] : 0.f;
        }
      }
    }
    }   // general path
    uint32_t* d32 = reinterpret_cast<uint32_t*>(out + cell * Cp + a * (4 * C0));
    const float* pf = &px[0][0];
    if ((C0 & 1) == 0) {   // 4*C0 bf16 = C0/2 16-byte stores
      uint4* d4 = reinterpret_cast<uint4*>(d32);
#pragma unroll
      for (int k = 0; k < C0 / 2; ++k)
        d4[k] = make_uint4(pack_bf16x2(pf[8 * k], pf[8 * k + 1]), pack_bf16x2(pf[8 * k + 2], pf[8 * k + 3]),
                           pack_bf16x2(pf[8 * k + 4], pf[8 * k + 5]), pack_bf16x2(pf[8 * k + 6], pf[8 * k + 7]));
    } else {
#pragma unroll
      for (int k = 0; k < 2 * C0; ++k) d32[k] = pack_bf16x2(pf[2 * k], pf[2 * k + 1]);
    }
    if (a == 3)   // channel padding of the cell (RGB: 96 -> 128)
      for (int k = 16 * C0; k < Cp; k += 2) *reinterpret_cast<uint32_t*>(out + cell * Cp + k) = 0u;
  }
}

// W4[(q, co)][r'][(s', a, b, ci)] (bf16) from the fp32 master W1[co][(kh*4 + kw)*C0 + ci] (row pitch K1p).
__global__ void conv1_expand_kernel(const float* __restrict__ w1, bf16* __restrict__ w4, int C1, int C0,
                                    int K1p, int Cp) {
  const int total = 4 * C1 * 2 * 2 * Cp;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int kk = i % (2 * Cp);
    const int rp = (i / (2 * Cp)) & 1;
    const int m = i / (4 * Cp);
    const int q = m / C1, co = m - q * C1;
    const int qh = q >> 1, qw = q & 1;
    const int sp = kk / Cp, k = kk - sp * Cp;
    float val = 0.f;
    if (k < 16 * C0) {
      const int ab = k / C0, ci = k - ab * C0;
      const int kh = 4 * rp + (ab >> 2) - 2 * qh - 1, kw = 4 * sp + (ab & 3) - 2 * qw - 1;
      if (kh >= 0 && kh < 4 && kw >= 0 && kw < 4) val = w1[(long long)co * K1p + (kh * 4 + kw) * C0 + ci];
    }
    w4[i] = __float2bfloat16(val);
  }
}

// dW1[co][(kh*4 + kw)*C0 + ci] += sum over the four quadrants of dW4 (fixed order: deterministic)
__global__ void conv1_fold_kernel(const float* __restrict__ dw4, float* __restrict__ dw1, int C1, int C0,
                                  int K1p, int Cp) {
  const int total = C1 * 16 * C0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int ci = i % C0;
    const int tap = (i / C0) & 15;
    const int co = i / (16 * C0);
    const int kh = tap >> 2, kw = tap & 3;
    float acc = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int eh = kh + 2 * (q >> 1) + 1, ew = kw + 2 * (q & 1) + 1;
      const int rp = eh >> 2, a = eh & 3, sp = ew >> 2, b = ew & 3;
      acc += dw4[((long long)(q * C1 + co) * 2 + rp) * (2 * Cp) + sp * Cp + (a * 4 + b) * C0 + ci];
    }
    dw1[(long long)co * K1p + tap * C0 + ci] += acc;
  }
}

int obs_s4d_impl(const uint8_t* frames, const long long* widx, const float* heat, const float* sel, int mode,
                 void* x4, int B, int L, int S, int center, int H, int W, int grayscale, int Cp, int SH, int SW,
                 cudaStream_t stream) {
  const int C0 = S * (grayscale ? 1 : 3) * (mode == 2 ? 2 : 1);
  if (!frames || !x4 || B <= 0 || S < 1 || S > 4 || (W & 3) || (H & 3) || SH != H / 4 + 2 || SW != W / 4 + 2 ||
      Cp < 16 * C0 || (Cp & 7) || mode < 0 || mode > 2)
    return GABRIL_ERR_BAD_SHAPE;
  // the reference multiplies / concatenates a [B, S, H, W] heat-map stack with the [B, S*C', H, W] frames: the
  // shapes only agree for grayscale stacks (train_bc.py:261,264 raise for RGB)
  if (mode != 0 && (!grayscale || !heat)) return GABRIL_ERR_BAD_SHAPE;
  const long long cells = (long long)B * SH * SW;
  const int grid = (int)std::min<long long>((long long)B * SH, 148 * 16);
  const int threads = std::min(384, (SW * 4 + 31) / 32 * 32);   // one row of cells per pass (W = 320: 328 items, 352 threads)
#define LAUNCH_S4D(SS, GG, MM)                                                                            \
  obs_s4d_kernel<SS, GG, MM><<<grid, threads, 0, stream>>>(frames, widx, heat, sel, (bf16*)x4, cells, L, center, H, \
                                                       W, SH, SW, Cp)
#define LAUNCH_S4D_S(GG, MM)                                                                       \
  switch (S) { case 1: LAUNCH_S4D(1, GG, MM); break; case 2: LAUNCH_S4D(2, GG, MM); break;         \
               case 3: LAUNCH_S4D(3, GG, MM); break; default: LAUNCH_S4D(4, GG, MM); break; }
  if (!grayscale) { LAUNCH_S4D_S(false, 0) }
  else if (mode == 0) { LAUNCH_S4D_S(true, 0) }
  else if (mode == 1) { LAUNCH_S4D_S(true, 1) }
  else { LAUNCH_S4D_S(true, 2) }
#undef LAUNCH_S4D_S
#undef LAUNCH_S4D
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

}  // namespace

extern "C" {

int gabril_obs_s4d(const uint8_t* frames, void* x4, int B, int L, int S, int center, int H, int W,
                   int grayscale, int Cp, int SH, int SW, cudaStream_t stream) {
  if (L < 1 || center < 0) return GABRIL_ERR_BAD_SHAPE;
  return obs_s4d_impl(frames, nullptr, nullptr, nullptr, 0, x4, B, L, S, center, H, W, grayscale, Cp, SH, SW, stream);
}

int gabril_obs_s4d_gaze(const uint8_t* frames, const long long* widx, const float* heat, const float* sel, int mode,
                        void* x4, int B, int L, int S, int center, int H, int W, int grayscale, int Cp, int SH,
                        int SW, cudaStream_t stream) {
  if (!widx && (L < 1 || center < 0)) return GABRIL_ERR_BAD_SHAPE;
  return obs_s4d_impl(frames, widx, heat, sel, mode, x4, B, widx ? S : L, S, widx ? S - 1 : center, H, W, grayscale,
                      Cp, SH, SW, stream);
}

int gabril_obs_s4d_store(const uint8_t* store, const long long* widx, void* x4, int B, int S, int H, int W,
                         int grayscale, int Cp, int SH, int SW, cudaStream_t stream) {
  if (!widx) return GABRIL_ERR_BAD_SHAPE;
  return obs_s4d_impl(store, widx, nullptr, nullptr, 0, x4, B, S, S, S - 1, H, W, grayscale, Cp, SH, SW, stream);
}

int gabril_conv1_s4d_expand(const float* w1, void* w4, int C1, int C0, int K1p, int Cp, cudaStream_t stream) {
  if (!w1 || !w4 || C1 <= 0 || C0 <= 0 || K1p < 16 * C0 || Cp < 16 * C0) return GABRIL_ERR_BAD_SHAPE;
  const int total = 4 * C1 * 4 * Cp;
  conv1_expand_kernel<<<(total + 255) / 256, 256, 0, stream>>>(w1, (bf16*)w4, C1, C0, K1p, Cp);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_conv1_s4d_fold(const float* dw4, float* dw1, int C1, int C0, int K1p, int Cp, cudaStream_t stream) {
  if (!dw4 || !dw1 || C1 <= 0 || C0 <= 0 || K1p < 16 * C0 || Cp < 16 * C0) return GABRIL_ERR_BAD_SHAPE;
  const int total = C1 * 16 * C0;
  conv1_fold_kernel<<<(total + 255) / 256, 256, 0, stream>>>(dw4, dw1, C1, C0, K1p, Cp);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

}  // extern "C"
