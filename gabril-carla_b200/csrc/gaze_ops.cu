// Memory-bound kernels of the GABRIL training step (sm_100a): frame formatting, gaze heat-map
// synthesis, gaze-mask dropout and the activation-map regulariser (forward + backward).
// Reference ops replaced (Gonglitian/GABRIL-CARLA, vlm_gaze/):
//   data_utils/data_loader_robomimic.py:144-157  _gather_last_s_frames      (integer, bit-exact)
//   data_utils/data_loader_robomimic.py:179-202  _format_obs_image          gabril_obs_format / _patches
//   data_utils/data_loader_robomimic.py:85-139   GazePreprocessor.forward   gabril_gaze_splat
//   data_utils/data_loader_robomimic.py:204-278  build_stack_heatmaps       gabril_gaze_splat
//   data_utils/gaze_utils.py:7-30                get_gaze_mask              gabril_gaze_mask_fwd/_bwd
//   data_utils/gaze_utils.py:33-52               apply_gmd_dropout          gabril_gmd_keep / _mask
//   train/train_bc.py:145-171                    Reg/Teacher MSE loss       gabril_reg_mse
// All of them are HBM-bound: one pass over the inputs, 16-byte accesses, per-sample tables in
// shared memory, warp-shuffle reductions.
#include <cfloat>

#include "common.cuh"

namespace {

constexpr int kMaxS = 8;     // frame-stack length supported by the splat kernel
constexpr int kMaxPts = 16;  // gaze points per frame

__device__ __forceinline__ int frame_index(int s, int S, int center, int L) {
  // data_loader_robomimic.py:151-152: window [center-S+1, center], each index clamped to [0, L-1]
  return min(max(center - S + 1 + s, 0), L - 1);
}

// ---- block reductions (all threads must call; result broadcast) --------------------------------
__device__ __forceinline__ float block_sum(float v, float* scratch) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  float r = 0.f;
  for (int i = 0; i < nw; ++i) r += scratch[i];  // fixed order: deterministic
  return r;
}
__device__ __forceinline__ float block_max(float v, float* scratch) {
  v = warp_max(v);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  float r = scratch[0];
  for (int i = 1; i < nw; ++i) r = fmaxf(r, scratch[i]);
  return r;
}
__device__ __forceinline__ float block_min(float v, float* scratch) {
  v = warp_min(v);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  float r = scratch[0];
  for (int i = 1; i < nw; ++i) r = fminf(r, scratch[i]);
  return r;
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

// =================================================================================================
// K1a: uint8 NHWC frame stack -> fp32 NCHW (the reference-facing layout of _format_obs_image)
// =================================================================================================
__device__ __forceinline__ float u8_to_unit(uint32_t v) { return __fdiv_rn((float)v, 255.0f); }
__device__ __forceinline__ float to_gray(float r, float g, float b) {
  // ((0.299 R) + (0.587 G)) + (0.114 B), every op rounded separately like the reference's
  // three torch multiplies and two adds (data_loader_robomimic.py:192-193)
  return __fadd_rn(__fadd_rn(__fmul_rn(0.299f, r), __fmul_rn(0.587f, g)), __fmul_rn(0.114f, b));
}

template <bool GRAY>
__global__ void __launch_bounds__(256)
obs_format_kernel(const uint8_t* __restrict__ frames, float* __restrict__ out, int B, int L, int S,
                  int center, int H, int W) {
  const int W4 = W >> 2;
  const long long total = (long long)B * S * H * W4;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int x4 = (int)(idx % W4);
    long long t = idx / W4;
    const int y = (int)(t % H);
    t /= H;
    const int s = (int)(t % S);
    const int b = (int)(t / S);
    const int f = frame_index(s, S, center, L);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(
        frames + ((((long long)b * L + f) * H + y) * W + x4 * 4) * 3);
    const uint32_t w0 = __ldg(src), w1 = __ldg(src + 1), w2 = __ldg(src + 2);
    const uint32_t by[12] = {w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u, w0 >> 24,
                             w1 & 255u, (w1 >> 8) & 255u, (w1 >> 16) & 255u, w1 >> 24,
                             w2 & 255u, (w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24};
    const long long plane = (long long)H * W;
    if (GRAY) {
      float4 o;
      o.x = to_gray(u8_to_unit(by[0]), u8_to_unit(by[1]), u8_to_unit(by[2]));
      o.y = to_gray(u8_to_unit(by[3]), u8_to_unit(by[4]), u8_to_unit(by[5]));
      o.z = to_gray(u8_to_unit(by[6]), u8_to_unit(by[7]), u8_to_unit(by[8]));
      o.w = to_gray(u8_to_unit(by[9]), u8_to_unit(by[10]), u8_to_unit(by[11]));
      float* dst = out + ((long long)b * S + s) * plane + (long long)y * W + x4 * 4;
      *reinterpret_cast<float4*>(dst) = o;
    } else {
      float* dst = out + (((long long)b * S + s) * 3) * plane + (long long)y * W + x4 * 4;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float4 o = make_float4(u8_to_unit(by[c]), u8_to_unit(by[3 + c]), u8_to_unit(by[6 + c]),
                               u8_to_unit(by[9 + c]));
        *reinterpret_cast<float4*>(dst + c * plane) = o;
      }
    }
  }
}

// =================================================================================================
// K1b: uint8 NHWC frame stack -> bf16 patch matrix of conv_1 (k4 s2 p1), fused format + im2col.
// Row = output pixel (b, oh, ow); column = (r*4 + s)*C0 + (frame*C' + c), zero padded to Kp.
// One thread = one (pixel, filter row r): 4 taps x C0 values.
// =================================================================================================
template <int S_, bool GRAY>
__global__ void __launch_bounds__(256)
obs_patches_kernel(const uint8_t* __restrict__ frames, bf16* __restrict__ out, int B, int L,
                   int center, int H, int W, int OH, int OW, int Kp) {
  constexpr int CP = GRAY ? 1 : 3;
  constexpr int C0 = S_ * CP;
  const long long total = (long long)B * OH * OW * 4;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(idx & 3);
    long long pix = idx >> 2;
    const int ow = (int)(pix % OW);
    long long t = pix / OW;
    const int oh = (int)(t % OH);
    const int b = (int)(t / OH);
    const int ih = 2 * oh - 1 + r;
    float v[4][C0];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const int iw = 2 * ow - 1 + s;
      const bool ok = (ih >= 0) && (ih < H) && (iw >= 0) && (iw < W);
#pragma unroll
      for (int fs = 0; fs < S_; ++fs) {
        float rr = 0.f, gg = 0.f, bb = 0.f;
        if (ok) {
          const int f = frame_index(fs, S_, center, L);
          const uint8_t* src = frames + ((((long long)b * L + f) * H + ih) * W + iw) * 3;
          rr = u8_to_unit(__ldg(src));
          gg = u8_to_unit(__ldg(src + 1));
          bb = u8_to_unit(__ldg(src + 2));
        }
        if (GRAY) {
          v[s][fs] = ok ? to_gray(rr, gg, bb) : 0.f;
        } else {
          v[s][fs * 3 + 0] = rr;
          v[s][fs * 3 + 1] = gg;
          v[s][fs * 3 + 2] = bb;
        }
      }
    }
    bf16* dst = out + pix * Kp + r * (4 * C0);
    const float* vf = &v[0][0];
    if ((4 * C0) % 8 == 0) {
#pragma unroll
      for (int i = 0; i < (4 * C0) / 8; ++i) {
        uint4 o;
        o.x = pack_bf16x2(vf[8 * i + 0], vf[8 * i + 1]);
        o.y = pack_bf16x2(vf[8 * i + 2], vf[8 * i + 3]);
        o.z = pack_bf16x2(vf[8 * i + 4], vf[8 * i + 5]);
        o.w = pack_bf16x2(vf[8 * i + 6], vf[8 * i + 7]);
        *reinterpret_cast<uint4*>(dst + 8 * i) = o;
      }
    } else {
#pragma unroll
      for (int i = 0; i < (4 * C0) / 4; ++i) {
        uint2 o;
        o.x = pack_bf16x2(vf[4 * i + 0], vf[4 * i + 1]);
        o.y = pack_bf16x2(vf[4 * i + 2], vf[4 * i + 3]);
        *reinterpret_cast<uint2*>(dst + 4 * i) = o;
      }
    }
    if (r == 3) {
      for (int k = 16 * C0; k < Kp; k += 4)
        *reinterpret_cast<uint2*>(out + pix * Kp + k) = make_uint2(0u, 0u);
    }
  }
}

// Tiled variant (even C0, W*3 a multiple of 16): one CTA = one sample x kPatchRows output rows.
// Phase 1 copies the 2*kPatchRows+2 input rows of every stacked frame into shared memory with
// 16-byte loads, phase 2 converts each input pixel ONCE (u8 -> /255 -> gray -> bf16) into a
// zero-padded [row][x+1][C0] image, phase 3 emits the patch rows: the 4*C0 values of one
// (pixel, filter row) are contiguous in that image, so they leave as 16-byte stores, four
// consecutive threads filling one 64/128-byte patch row.
constexpr int kPatchRows = 6;

template <int S_, bool GRAY>
__global__ void __launch_bounds__(256)
obs_patches_tiled_kernel(const uint8_t* __restrict__ frames, const long long* __restrict__ widx,
                         bf16* __restrict__ out, int L, int center, int H, int W, int OH, int OW,
                         int Kp, int SH, int SW) {
  constexpr int CP = GRAY ? 1 : 3;
  constexpr int C0 = S_ * CP;
  constexpr int NR = 2 * kPatchRows + 2;
  extern __shared__ __align__(16) uint8_t psm[];
  const int row_bytes = W * 3;
  const int img_row = (W + 2) * C0;  // bf16 elements
  uint8_t* raw = psm;                                        // [NR][S_][row_bytes]
  bf16* img = reinterpret_cast<bf16*>(psm + ((NR * S_ * row_bytes + 15) & ~15));  // [NR][img_row]
  const int tiles = (OH + kPatchRows - 1) / kPatchRows;
  const int b = blockIdx.x / tiles;
  const int oh0 = (blockIdx.x - b * tiles) * kPatchRows;
  const int ih0 = 2 * oh0 - 1;
  __shared__ float lut[256];
  lut[threadIdx.x & 255] = u8_to_unit(threadIdx.x & 255);
  // ---- phase 1 ----
  const int vec_per_row = row_bytes / 16;
  for (int i = threadIdx.x; i < NR * S_ * vec_per_row; i += blockDim.x) {
    const int v = i % vec_per_row;
    const int rs = i / vec_per_row;
    const int fs = rs % S_, r = rs / S_;
    const int ih = ih0 + r;
    if (ih >= 0 && ih < H) {
      // batch layout [B, L, H, W, 3], or (widx != NULL) a frame store [N, H, W, 3] addressed through
      // the window table widx[b][fs] of gabril_store_windows: no batch tensor is materialised
      const long long fr = widx ? __ldg(widx + (long long)b * S_ + fs)
                                : (long long)b * L + frame_index(fs, S_, center, L);
      const uint4* src = reinterpret_cast<const uint4*>(
          frames + (fr * H + ih) * (long long)row_bytes);
      reinterpret_cast<uint4*>(raw + (size_t)rs * row_bytes)[v] = __ldg(src + v);
    }
  }
  __syncthreads();
  // ---- phase 2: four pixels per thread (12 raw bytes = three 32-bit words per frame); u8/255 comes
  //      from a 256-entry table of the correctly rounded quotients (same bits as the division) ----
  const int quads = W >> 2;
  for (int i = threadIdx.x; i < NR * (quads + 1); i += blockDim.x) {
    const int qd = i % (quads + 1);
    const int r = i / (quads + 1);
    bf16* drow = img + r * img_row;
    if (qd == quads) {  // the two zero-padding columns x = -1 and x = W
#pragma unroll
      for (int c = 0; c < C0; ++c) { drow[c] = __float2bfloat16(0.f); drow[(W + 1) * C0 + c] = __float2bfloat16(0.f); }
      continue;
    }
    const int ih = ih0 + r;
    const bool ok = (ih >= 0) && (ih < H);
    float px[4][C0];
#pragma unroll
    for (int fs = 0; fs < S_; ++fs) {
      uint32_t wv[3] = {0u, 0u, 0u};
      if (ok) {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(raw + (size_t)(r * S_ + fs) * row_bytes) + 3 * qd;
        wv[0] = src[0]; wv[1] = src[1]; wv[2] = src[2];
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float ch[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const int byte = 3 * k + c;
          ch[c] = lut[(wv[byte >> 2] >> (8 * (byte & 3))) & 0xffu];
        }
        if (GRAY) {
          px[k][fs] = ok ? to_gray(ch[0], ch[1], ch[2]) : 0.f;
        } else {
          px[k][fs * 3 + 0] = ok ? ch[0] : 0.f;
          px[k][fs * 3 + 1] = ok ? ch[1] : 0.f;
          px[k][fs * 3 + 2] = ok ? ch[2] : 0.f;
        }
      }
    }
    // 4 pixels x C0 bf16 = 2*C0 words, contiguous from column (4*qd + 1)
    uint32_t* d32 = reinterpret_cast<uint32_t*>(drow + (4 * qd + 1) * C0);
    const float* pf = &px[0][0];
#pragma unroll
    for (int k = 0; k < 2 * C0; ++k) d32[k] = pack_bf16x2(pf[2 * k], pf[2 * k + 1]);
  }
  __syncthreads();
  // ---- phase 3 ----
  const int rows_here = min(kPatchRows, OH - oh0);
  for (int i = threadIdx.x; i < rows_here * OW * 4; i += blockDim.x) {
    const int r4 = i & 3;
    const int pl = i >> 2;
    const int ow = pl % OW;
    const int ohl = pl / OW;
    long long pix = ((long long)b * OH + oh0 + ohl) * OW + ow;
    if (SH > 0) {
      // space-to-depth row order over the one-pixel-padded output: pixel (oh, ow) is quadrant
      // ((oh+1)&1, (ow+1)&1) of super-pixel ((oh+1)>>1, (ow+1)>>1) of an SH x SW grid
      const int ph = oh0 + ohl + 1, pw = ow + 1;
      pix = ((((long long)b * SH + (ph >> 1)) * SW + (pw >> 1)) << 2) + ((ph & 1) << 1) + (pw & 1);
    }
    const uint2* src = reinterpret_cast<const uint2*>(img + (2 * ohl + r4) * img_row + 2 * ow * C0);
    uint4* dst = reinterpret_cast<uint4*>(out + pix * Kp + r4 * (4 * C0));
#pragma unroll
    for (int k = 0; k < C0 / 2; ++k) {
      const uint2 lo = src[2 * k], hi = src[2 * k + 1];
      dst[k] = make_uint4(lo.x, lo.y, hi.x, hi.y);
    }
    if (r4 == 3) {
      uint4* z = reinterpret_cast<uint4*>(out + pix * Kp + 16 * C0);
      for (int k = 0; k < (Kp - 16 * C0) / 8; ++k) z[k] = make_uint4(0u, 0u, 0u, 0u);
    }
  }
}

// =================================================================================================
// K6: gaze-mask dropout.  gmd_keep: K[b, j] = mean_s bicubic_down(g[b, s])[j] + batch-global
// min / max (ordered-uint atomics).  gmd_mask: rowmask[b, j] = (A < 0.7 Khat + 0.3) or the keep
// probability itself in test mode.
// =================================================================================================
__global__ void __launch_bounds__(256)
gmd_keep_kernel(const float* __restrict__ g, const float* __restrict__ sel, float* __restrict__ K,
                unsigned int* __restrict__ mnmx, int B, int S, int H, int W, int h, int w) {
  __shared__ float scratch[32];
  const int n = h * w;
  const long long total = (long long)B * n;
  const float sy = (float)H / (float)h, sx = (float)W / (float)w;
  float lo = FLT_MAX, hi = -FLT_MAX;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(idx % n);
    const int b = (int)(idx / n);
    const int oy = j / w, ox = j - oy * w;
    int iy[4], ix[4];
    float wy[4], wx[4];
    bicubic_taps(oy, sy, H, iy, wy);
    bicubic_taps(ox, sx, W, ix, wx);
    float acc = 0.f;
    // samples that do not use gaze (ivg = 0) see an all-zero heat map (train_bc.py:258)
    const int Suse = (sel && __ldg(sel + b) <= 0.f) ? 0 : S;
    for (int s = 0; s < Suse; ++s) {
      const float* gp = g + ((long long)b * S + s) * H * W;
      float rows[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float* rp = gp + (long long)iy[r] * W;
        rows[r] = __ldg(rp + ix[0]) * wx[0] + __ldg(rp + ix[1]) * wx[1] + __ldg(rp + ix[2]) * wx[2] +
                  __ldg(rp + ix[3]) * wx[3];
      }
      const float v = rows[0] * wy[0] + rows[1] * wy[1] + rows[2] * wy[2] + rows[3] * wy[3];
      acc = (s == 0) ? v : __fadd_rn(acc, v);
    }
    const float k = (S == 1) ? acc : __fdiv_rn(acc, (float)S);
    K[idx] = k;
    lo = fminf(lo, k);
    hi = fmaxf(hi, k);
  }
  lo = block_min(lo, scratch);
  hi = block_max(hi, scratch);
  if (threadIdx.x == 0) {
    atomicMin(&mnmx[0], f32_to_ordered(lo));
    atomicMax(&mnmx[1], f32_to_ordered(hi));
  }
}

__global__ void gmd_init_kernel(unsigned int* mnmx) {
  mnmx[0] = 0xFFFFFFFFu;
  mnmx[1] = 0u;
}

// Philox4x32-10 (Salmon et al., SC'11): counter (c0..c3), key (k0, k1) -> four uniform 32-bit words.
// Used for GMD's uniform draw A ~ U[0,1) (gaze_utils.py:36 draws it with torch.rand) when the caller
// passes no A: the draw then lives inside the kernel, so a GMD / IGMD step is CUDA-graph capturable.
// The stream is (seed, step counter read from device memory, draw id, element / 4).
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

__global__ void gmd_rng_advance_kernel(unsigned long long* rng) { rng[1] += 1ull; }

__global__ void __launch_bounds__(256)
gmd_mask_kernel(const float* __restrict__ K, const float* __restrict__ A,
                const unsigned long long* __restrict__ rng, int draw_id,
                const unsigned int* __restrict__ mnmx, float* __restrict__ rowmask,
                float* __restrict__ keep_out, long long total, float drop_p, int test_mode,
                int layout, int h, int w, int SH, int SW) {
  const float mn = ordered_to_f32(mnmx[0]), mx = ordered_to_f32(mnmx[1]);
  const float den = __fadd_rn(__fsub_rn(mx, mn), 1e-8f);
  const float one_minus = (float)(1.0 - (double)drop_p);
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const float kh = __fdiv_rn(__fsub_rn(K[idx], mn), den);
    const float kt = __fadd_rn(__fmul_rn(drop_p, kh), one_minus);
    if (keep_out) keep_out[idx] = kt;
    long long o = idx;
    if (layout != GABRIL_MASK_PLAIN) {
      // the storage order of the activation the mask multiplies (pixel (oy, ox) sits at (oy+1, ox+1)
      // of a zero-bordered grid; conv_1's output additionally space-to-depth by 2)
      const int n = h * w;
      const long long b = idx / n;
      const int j = (int)(idx - b * n), oy = j / w + 1, ox = j % w + 1;
      o = (layout == GABRIL_MASK_S2D)
              ? ((b * SH + (oy >> 1)) * SW + (ox >> 1)) * 4 + ((oy & 1) * 2 + (ox & 1))
              : (b * SH + oy) * SW + ox;
    }
    float a_draw = 0.f;
    if (!test_mode) {
      if (A) {
        a_draw = A[idx];
      } else {
        const unsigned long long seed = rng[0], step = rng[1];
        const uint4 r = philox4x32_10(make_uint4((uint32_t)(idx >> 2), (uint32_t)(idx >> 34), (uint32_t)step,
                                                 (uint32_t)(step >> 32) ^ ((uint32_t)draw_id << 24)),
                                      make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
        const int lane4 = (int)(idx & 3);
        const uint32_t bits = lane4 == 0 ? r.x : lane4 == 1 ? r.y : lane4 == 2 ? r.z : r.w;
        a_draw = (float)(bits >> 8) * (1.0f / 16777216.0f);   // 24 random mantissa bits: [0, 1)
      }
    }
    rowmask[o] = test_mode ? kt : ((a_draw < kt) ? 1.f : 0.f);
  }
}

// z[B, C, n] fp32 (reference layout) times rowmask[B, n]
__global__ void __launch_bounds__(256)
rowmask_apply_nchw_kernel(const float* __restrict__ z, const float* __restrict__ rowmask,
                          float* __restrict__ out, int B, int C, int n) {
  const long long total = (long long)B * C * n;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(idx % n);
    const int b = (int)(idx / ((long long)C * n));
    out[idx] = z[idx] * rowmask[(long long)b * n + j];
  }
}

// =================================================================================================
// Device-resident demonstration store.  robomimic's SequenceDataset with
// frame_stack = S, seq_length = 1 returns, for the sampled step i of a demo that starts at step d,
// the steps max(i - S + 1 + s, d), s = 0..S-1 (the demo's first frame repeated as front padding:
// vlm_gaze/robomimic/utils/dataset.py:609-628, pad_sequence(pad_same=True)).  The store keeps all
// steps of all demos back to back in HBM, demo_first[i] = d; a batch is a table of frame indices
// and every consumer (patch kernel, content hash, gaze / action row gather) reads through it.
// Integer work: bit-exact.
// =================================================================================================
__global__ void __launch_bounds__(256)
store_windows_kernel(const long long* __restrict__ step_idx, const long long* __restrict__ demo_first,
                     long long* __restrict__ widx, int B, int S, long long n_steps) {
  const long long total = (long long)B * S;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int s = (int)(i % S);
    long long c = step_idx[i / S];
    c = c < 0 ? 0 : (c > n_steps - 1 ? n_steps - 1 : c);
    const long long t = c - (S - 1) + s, d = demo_first[c];
    widx[i] = t > d ? t : d;
  }
}

__global__ void __launch_bounds__(256)
store_gather_rows_kernel(const float* __restrict__ src, const long long* __restrict__ widx,
                         float* __restrict__ dst, long long rows, int row_len) {
  const long long total = rows * row_len;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / row_len;
    dst[i] = __ldg(src + widx[r] * row_len + (i - r * row_len));
  }
}

// =================================================================================================
// gaze.ratio < 1 (train_bc.py:230-240): per-sample content hash r = frac(|sum(obs) * 123456.789|)
// in fp32, sample uses gaze iff r < ratio.  sum(obs) is over the formatted frame stack (the exact
// fp32 values of obs_format).  One CTA per sample, fixed summation order.  For any frame whose
// normalised pixel sum exceeds 2^23 / 123456.789 = 68 the fp32 product has no fractional bits, so
// r = 0 whatever the order of the additions; below that the reference's own
// result depends on the reduction order of the device it runs on.
// sel_stats[0] = number selected, sel_stats[1] = B / number selected (inf when none: the
// reference's mean over an empty selection is NaN as well).
// =================================================================================================
__global__ void __launch_bounds__(256)
ivg_key_kernel(const uint8_t* __restrict__ frames, const long long* __restrict__ widx,
               float* __restrict__ sel, int B, int L, int S, int center, int H, int W, int gray,
               float ratio) {
  __shared__ float scratch[32];
  const int b = blockIdx.x;
  const long long HW = (long long)H * W;
  float acc = 0.f;
  for (int s = 0; s < S; ++s) {
    int t = center - S + 1 + s;
    t = t < 0 ? 0 : (t > L - 1 ? L - 1 : t);
    const uint8_t* fp = frames + (widx ? widx[(long long)b * S + s] : (long long)b * L + t) * HW * 3;
    for (long long i = threadIdx.x; i < HW; i += blockDim.x) {
      const float r = __fdiv_rn((float)fp[3 * i + 0], 255.f), g = __fdiv_rn((float)fp[3 * i + 1], 255.f),
                  bl = __fdiv_rn((float)fp[3 * i + 2], 255.f);
      if (gray) acc += __fadd_rn(__fadd_rn(__fmul_rn(0.299f, r), __fmul_rn(0.587f, g)), __fmul_rn(0.114f, bl));
      else acc += r + g + bl;
    }
  }
  const float key = block_sum(acc, scratch);
  if (threadIdx.x == 0) {
    const float v = fabsf(__fmul_rn(key, 123456.789f));
    const float r = v - truncf(v);
    sel[b] = (ratio >= 1.f) ? 1.f : ((ratio <= 0.f) ? 0.f : (r < ratio ? 1.f : 0.f));
  }
}

__global__ void __launch_bounds__(256) ivg_count_kernel(const float* __restrict__ sel,
                                                        float* __restrict__ stats, int B) {
  __shared__ float scratch[32];
  float c = 0.f;
  for (int i = threadIdx.x; i < B; i += blockDim.x) c += sel[i] > 0.f ? 1.f : 0.f;
  c = block_sum(c, scratch);
  if (threadIdx.x == 0) { stats[0] = c; stats[1] = __fdiv_rn((float)B, c); }
}

// =================================================================================================
// K5: get_gaze_mask forward / backward and the fused Reg-MSE loss + gradient.
// One CTA = one sample; thread x owns output column x (blockDim = W rounded to a warp multiple).
// Shared: p[n] softmax, t[h][W] row-interpolated map, sacc[h][W] y-adjoint accumulators,
// bicubic tap tables for both axes, dp[n].
// =================================================================================================
// MODE_KL / MODE_TV / MODE_JS: train_bc.py:152-167 - both maps divided by their (detached) sums
// + 1e-8, then KL(g1 || g2), sum |g1 - g2| or the Jensen-Shannon form with eps = 1e-6 inside the logs;
// loss_part[b] is the per-sample sum, the caller divides by the number of selected samples.
enum { MODE_MSE = 0, MODE_MASK_FWD = 1, MODE_MASK_BWD = 2, MODE_KL = 3, MODE_TV = 4, MODE_JS = 5 };

struct RegArgs {
  const void* z;         // bf16 [B, n, C] (NHWC) or fp32 [B, C, n] (NCHW)
  const float* g1;       // MODE_MSE: target plane of sample b at g1 + b * g1_bstride
  long long g1_bstride;
  const float* sel;      // optional [B]: 0 => sample skipped (gaze.ratio < 1)
  float* mask_out;       // MODE_MASK_FWD: [B, H*W]
  const float* grad_in;  // MODE_MASK_BWD: [B, H*W]
  void* dz;              // gradient w.r.t. z, same layout as z (MODE_MSE / MODE_MASK_BWD)
  float* loss_part;      // MODE_MSE: [B] per-sample sum of squared differences
  float* stats;          // optional [B][4]: min, max, argmin, argmax of the up-sampled map
  const float* scale_dev;  // optional device scalar multiplied into grad_scale (B / n_selected)
  float beta, grad_scale;
  int B, C, h, w, H, W;
};

// blockDim = Wr * YS (Wr = W rounded up to a warp multiple): thread (x, yq) owns output column x
// and the row slab yq of YS.  Walking down a column the four source rows of the bicubic y-taps change
// only every H/h rows, so their tmap values (and, in the backward pass, the four y-adjoint
// accumulators) live in registers and shared memory is touched only when the tap window moves.
// RC: rows handled per inner chunk = the length of a tap-window run (H / h = 9 for 180 -> 20): a chunk of 8
// walked a 9-row run in two rounds of eight predicated bodies, i.e. 16 issue slots for 9 pixels.
template <bool NHWC_BF16, int MODE, int MAXT, int RC>
__global__ void __launch_bounds__(MAXT, MAXT <= 640 ? 2 : 1) reg_kernel(const __grid_constant__ RegArgs a) {
  extern __shared__ __align__(16) float sm[];
  constexpr bool LOSS = MODE == MODE_MSE || MODE >= MODE_KL;  // fused loss + gradient modes
  constexpr bool DISTN = MODE >= MODE_KL;                     // sum-normalised divergences
  const int h = a.h, w = a.w, H = a.H, W = a.W, C = a.C, n = h * w, n4 = (n + 3) & ~3;
  float* p = sm;                                        // [n4]
  float* tmap = p + n4;                                 // [h][W]
  float* sacc = tmap + h * W;                           // [2][h][W] (one copy per row slab)
  float* wyt = sacc + 2 * h * W;                        // [H][4]
  int* iyt = reinterpret_cast<int*>(wyt + 4 * H);       // [H][4]
  float* wxt = reinterpret_cast<float*>(iyt + 4 * H);   // [W][4]
  int* ixt = reinterpret_cast<int*>(wxt + 4 * W);       // [W][4]
  float* dp = reinterpret_cast<float*>(ixt + 4 * W);    // [n4]
  int* xrange = reinterpret_cast<int*>(dp + n4);        // [w][2]
  int* ystart = xrange + 2 * w;                         // [runs + 1] first row of every tap-window run
  __shared__ float scratch[32];
  __shared__ int nruns_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nwarp = blockDim.x >> 5;
  const int b = blockIdx.x;
  const bool skip = (a.sel != nullptr) && (a.sel[b] <= 0.f);
  const int Wr = (W + 31) & ~31, YS = blockDim.x / Wr;
  const int x = tid % Wr, yq = tid / Wr;
  const int Hq = (H + YS - 1) / YS;
  const int ya = min(yq * Hq, H), yb = min(ya + Hq, H);

  const bf16* zb16 = reinterpret_cast<const bf16*>(a.z) + (long long)b * n * C;
  const float* zf32 = reinterpret_cast<const float*>(a.z) + (long long)b * n * C;

  if (LOSS && skip) {
    // sample not selected: zero gradient, zero loss
    if (NHWC_BF16) {
      uint4* d = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(a.dz) + (long long)b * n * C);
      for (int i = tid; i < n * C / 8; i += blockDim.x) d[i] = make_uint4(0, 0, 0, 0);
    } else {
      float* d = reinterpret_cast<float*>(a.dz) + (long long)b * n * C;
      for (int i = tid; i < n * C; i += blockDim.x) d[i] = 0.f;
    }
    if (tid == 0) a.loss_part[b] = 0.f;
    return;
  }

  // ---- tap tables ----
  const float sy = (float)h / (float)H, sx = (float)w / (float)W;
  for (int y = tid; y < H; y += blockDim.x) {
    int id[4]; float wt[4];
    bicubic_taps(y, sy, h, id, wt);
#pragma unroll
    for (int k = 0; k < 4; ++k) { iyt[4 * y + k] = id[k]; wyt[4 * y + k] = wt[k]; }
  }
  for (int xx = tid; xx < W; xx += blockDim.x) {
    int id[4]; float wt[4];
    bicubic_taps(xx, sx, w, id, wt);
#pragma unroll
    for (int k = 0; k < 4; ++k) { ixt[4 * xx + k] = id[k]; wxt[4 * xx + k] = wt[k]; }
  }
  if (tid < w) { xrange[2 * tid] = W; xrange[2 * tid + 1] = -1; }

  // ---- s_j = sum_c |z| (fp32) ----
  if (NHWC_BF16) {
    // C/8 lanes per position, 16-byte loads, shuffle reduce inside the lane group
    const int lpp = C >> 3;              // lanes per position (power of two <= 32)
    const int ppw = 32 / lpp;            // positions per warp step
    const int sub = lane % lpp, posl = lane / lpp;
    for (int j0 = wid * ppw; j0 < n; j0 += nwarp * ppw) {
      const int j = j0 + posl;
      float acc = 0.f;
      if (j < n) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(zb16 + (long long)j * C) + sub);
        acc = fabsf(bf16lo(v.x)) + fabsf(bf16hi(v.x)) + fabsf(bf16lo(v.y)) + fabsf(bf16hi(v.y)) +
              fabsf(bf16lo(v.z)) + fabsf(bf16hi(v.z)) + fabsf(bf16lo(v.w)) + fabsf(bf16hi(v.w));
      }
      for (int o = lpp >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (j < n && sub == 0) p[j] = acc;
    }
  } else {
    for (int j = tid; j < n; j += blockDim.x) {
      float acc = 0.f;
      for (int c = 0; c < C; ++c) acc += fabsf(__ldg(zf32 + (long long)c * n + j));
      p[j] = acc;
    }
  }
  __syncthreads();

  // source columns x' <- the contiguous range of output columns whose taps touch it
  if (MODE != MODE_MASK_FWD && tid < W) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      atomicMin(&xrange[2 * ixt[4 * tid + k]], tid);
      atomicMax(&xrange[2 * ixt[4 * tid + k] + 1], tid);
    }
  }

  // runs of consecutive output rows that share one y-tap window (H / h rows each)
  if (wid == 0) {
    int base = 0;
    for (int c = 0; c < H; c += 32) {
      const int y = c + lane;
      bool f = false;
      if (y < H) f = (y == 0) || iyt[4 * y] != iyt[4 * y - 4] || iyt[4 * y + 3] != iyt[4 * y - 1];
      const unsigned m = __ballot_sync(0xffffffffu, f);
      if (f) ystart[base + __popc(m & ((2u << lane) - 1u)) - 1] = y;
      base += __popc(m);
    }
    if (lane == 0) { ystart[base] = H; nruns_s = base; }
  }

  // ---- softmax(s / beta) over the n positions ----
  float lmax = -FLT_MAX;
  for (int j = tid; j < n; j += blockDim.x) { const float v = __fdiv_rn(p[j], a.beta); p[j] = v; lmax = fmaxf(lmax, v); }
  const float smax = block_max(lmax, scratch);
  float lsum = 0.f;
  for (int j = tid; j < n; j += blockDim.x) { const float e = expf(p[j] - smax); p[j] = e; lsum += e; }
  const float ssum = block_sum(lsum, scratch);
  for (int j = tid; j < n; j += blockDim.x) p[j] = __fdiv_rn(p[j], ssum);
  __syncthreads();

  // ---- x interpolation of every source row: tmap[y'][x] (rows dealt over the YS slabs) ----
  const bool col_ok = x < W && yq < YS;
  if (col_ok) {
    const int4 ii = *reinterpret_cast<const int4*>(ixt + 4 * x);
    const float4 ww = *reinterpret_cast<const float4*>(wxt + 4 * x);
    for (int yy = yq; yy < h; yy += YS) {
      const float* pr = p + yy * w;
      tmap[yy * W + x] = pr[ii.x] * ww.x + pr[ii.y] * ww.y + pr[ii.z] * ww.z + pr[ii.w] * ww.w;
      if (MODE != MODE_MASK_FWD) { sacc[yy * W + x] = 0.f; sacc[(h + yy) * W + x] = 0.f; }
    }
  }
  __syncthreads();

  const int4* iy4 = reinterpret_cast<const int4*>(iyt);
  const float4* wy4 = reinterpret_cast<const float4*>(wyt);
  const int nruns = nruns_s;
  // register window over the four source rows of the current run
  int4 cur = make_int4(0, 0, 0, 0);
  float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f;
  auto load_window = [&](int y) {
    cur = iy4[y];
    t0 = tmap[cur.x * W + x]; t1 = tmap[cur.y * W + x]; t2 = tmap[cur.z * W + x]; t3 = tmap[cur.w * W + x];
  };
  auto u_of = [&](const float4& wv) { return t0 * wv.x + t1 * wv.y + t2 * wv.z + t3 * wv.w; };

  // ---- pass 1: extrema (first occurrence wins ties, like torch's CPU max/min) ----
  float vmin = FLT_MAX, vmax = -FLT_MAX;
  int imin = 0x7fffffff, imax = 0x7fffffff;
  float usum = 0.f, g1sum = 0.f;  // DISTN: column sums of the raw map and of the target
  if (col_ok) {
    const float* g1c = DISTN ? a.g1 + (long long)b * a.g1_bstride + x : nullptr;
    for (int r = 0; r < nruns; ++r) {
      const int y0 = max(ya, ystart[r]), y1 = min(yb, ystart[r + 1]);
      if (y0 >= y1) continue;
      load_window(y0);
      for (int yc = y0; yc < y1; yc += RC) {
#pragma unroll
        for (int k = 0; k < RC; ++k) {
          const int y = yc + k;
          if (y < y1) {
            const float u = u_of(wy4[y]);
            if (u < vmin) { vmin = u; imin = y; }
            if (u > vmax) { vmax = u; imax = y; }
            if (DISTN) { usum += u; g1sum += __ldg(g1c + y * W); }
          }
        }
      }
    }
    if (imin != 0x7fffffff) imin = imin * W + x;
    if (imax != 0x7fffffff) imax = imax * W + x;
  }
  // block arg-reduce
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, vmin, o);
    const int oi = __shfl_xor_sync(0xffffffffu, imin, o);
    if (ov < vmin || (ov == vmin && oi < imin)) { vmin = ov; imin = oi; }
    const float ov2 = __shfl_xor_sync(0xffffffffu, vmax, o);
    const int oi2 = __shfl_xor_sync(0xffffffffu, imax, o);
    if (ov2 > vmax || (ov2 == vmax && oi2 < imax)) { vmax = ov2; imax = oi2; }
  }
  __shared__ float r_min[32], r_max[32];
  __shared__ int r_imin[32], r_imax[32];
  if (lane == 0) { r_min[wid] = vmin; r_imin[wid] = imin; r_max[wid] = vmax; r_imax[wid] = imax; }
  __syncthreads();
  vmin = r_min[0]; imin = r_imin[0]; vmax = r_max[0]; imax = r_imax[0];
  for (int i = 1; i < nwarp; ++i) {
    if (r_min[i] < vmin || (r_min[i] == vmin && r_imin[i] < imin)) { vmin = r_min[i]; imin = r_imin[i]; }
    if (r_max[i] > vmax || (r_max[i] == vmax && r_imax[i] < imax)) { vmax = r_max[i]; imax = r_imax[i]; }
  }
  if (imin >= H * W) imin = 0;  // all-NaN map: keep the indices addressable, values stay NaN
  if (imax >= H * W) imax = 0;
  const float mn = vmin, mx = vmax;
  const float rng = __fsub_rn(mx, mn);
  if (a.stats && tid == 0) {
    a.stats[4 * b + 0] = mn; a.stats[4 * b + 1] = mx;
    a.stats[4 * b + 2] = (float)imin; a.stats[4 * b + 3] = (float)imax;
  }

  if (MODE == MODE_MASK_FWD) {
    if (col_ok) {
      float* mo = a.mask_out + (long long)b * H * W;
      for (int r = 0; r < nruns; ++r) {
        const int y0 = max(ya, ystart[r]), y1 = min(yb, ystart[r + 1]);
        if (y0 >= y1) continue;
        load_window(y0);
        for (int y = y0; y < y1; ++y)  // reference-facing op: exact division
          mo[y * W + x] = __fdiv_rn(__fsub_rn(u_of(wy4[y]), mn), rng);
      }
    }
    return;
  }

  // ---- pass 2: g2, loss, d(loss)/du accumulated through the y-adjoint into sacc ----
  // (training path: the two divisions by rng are multiplications by 1/rng; 0/0 still gives NaN)
  const float inv_rng = __fdiv_rn(1.f, rng);
  const float gs = a.scale_dev ? a.grad_scale * __ldg(a.scale_dev) : a.grad_scale;
  float iS1 = 0.f, iS2 = 0.f;
  if (DISTN) {
    // S1 = sum g1 + 1e-8; S2 = sum (u - mn) / rng + 1e-8 (block sums in double: the second one
    // is a difference of two sums of 57 600 terms)
    __shared__ double dred[2][32];
    double du = usum, dg = g1sum;
    for (int o = 16; o > 0; o >>= 1) {
      du += __shfl_xor_sync(0xffffffffu, du, o);
      dg += __shfl_xor_sync(0xffffffffu, dg, o);
    }
    if (lane == 0) { dred[0][wid] = du; dred[1][wid] = dg; }
    __syncthreads();
    du = 0.0; dg = 0.0;
    for (int i = 0; i < nwarp; ++i) { du += dred[0][i]; dg += dred[1][i]; }
    const double S2 = (du - (double)(H * W) * (double)mn) / (double)rng + 1e-8;
    const double S1 = dg + 1e-8;
    iS1 = (float)(1.0 / S1);
    iS2 = (float)(1.0 / S2);
  }
  float sq = 0.f, E = 0.f, F = 0.f;
  if (col_ok) {
    const float* gsrc = LOSS ? a.g1 + (long long)b * a.g1_bstride
                                           : a.grad_in + (long long)b * H * W;
    gsrc += x;
    for (int r = 0; r < nruns; ++r) {
      const int y0 = max(ya, ystart[r]), y1 = min(yb, ystart[r + 1]);
      if (y0 >= y1) continue;
      load_window(y0);
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
      for (int yc = y0; yc < y1; yc += RC) {
        float gv[RC];
#pragma unroll
        for (int k = 0; k < RC; ++k) gv[k] = (yc + k < y1) ? __ldg(gsrc + (yc + k) * W) : 0.f;
#pragma unroll
        for (int k = 0; k < RC; ++k) {
          if (yc + k < y1) {
            const float4 wv = wy4[yc + k];
            const float g2 = __fsub_rn(u_of(wv), mn) * inv_rng;
            float e;
            if (MODE == MODE_MSE) {
              const float d = g2 - gv[k];
              sq = fmaf(d, d, sq);
              e = 2.f * d * gs;
            } else if (MODE == MODE_TV) {
              const float d = gv[k] * iS1 - g2 * iS2;
              sq += fabsf(d);
              e = (d > 0.f ? -iS2 : (d < 0.f ? iS2 : 0.f)) * gs;
            } else if (MODE == MODE_KL) {
              const float pa = gv[k] * iS1, pb = g2 * iS2;
              const float ib = __fdiv_rn(1.f, pb + 1e-6f);
              sq = fmaf(pa, logf((pa + 1e-6f) * ib), sq);
              e = -pa * ib * iS2 * gs;
            } else if (MODE == MODE_JS) {
              const float pa = gv[k] * iS1, pb = g2 * iS2;
              const float im = __fdiv_rn(1.f, 0.5f * (pa + pb) + 1e-6f);
              const float la = logf((pa + 1e-6f) * im), lb = logf((pb + 1e-6f) * im);
              sq += 0.5f * (pa * la + pb * lb);
              // d/dpb of 0.5 [pa log((pa+e)/(m+e)) + pb log((pb+e)/(m+e))], m = (pa + pb) / 2
              e = 0.5f * (lb + __fdiv_rn(pb, pb + 1e-6f) - 0.5f * (pa + pb) * im) * iS2 * gs;
            } else {
              e = gv[k];
            }
            E += e;
            F = fmaf(e, g2, F);
            const float du = e * inv_rng;
            a0 = fmaf(wv.x, du, a0); a1 = fmaf(wv.y, du, a1); a2 = fmaf(wv.z, du, a2); a3 = fmaf(wv.w, du, a3);
          }
        }
      }
      // slab-private copy of the accumulators: plain read-modify-write, no atomics
      float* sa = sacc + (yq & 1) * h * W + x;
      sa[cur.x * W] += a0; sa[cur.y * W] += a1; sa[cur.z * W] += a2; sa[cur.w * W] += a3;
    }
  }
  const float Esum = block_sum(E, scratch);
  const float Fsum = block_sum(F, scratch);
  if (LOSS) {
    const float sqsum = block_sum(sq, scratch);
    if (tid == 0) a.loss_part[b] = sqsum;
  }
  __syncthreads();
  if (tid == 0) {
    // max/min backward: the whole -sum term goes to the single arg index
    const int ymin = imin / W, xmin = imin - ymin * W;
    const int ymax = imax / W, xmax = imax - ymax * W;
    const float dmin = -(Esum - Fsum) * inv_rng, dmax = -Fsum * inv_rng;
    for (int k = 0; k < 4; ++k) sacc[iyt[4 * ymin + k] * W + xmin] += wyt[4 * ymin + k] * dmin;
    for (int k = 0; k < 4; ++k) sacc[iyt[4 * ymax + k] * W + xmax] += wyt[4 * ymax + k] * dmax;
  }
  __syncthreads();
  // ---- x-adjoint: dp[y'][x'] = sum_x sum_k [ix[x][k] == x'] wx[x][k] sacc[y'][x] ----
  for (int j = tid; j < n; j += blockDim.x) {
    const int yy = j / w, xs = j - yy * w;
    float acc = 0.f;
    for (int xx = xrange[2 * xs]; xx <= xrange[2 * xs + 1]; ++xx) {
      const float sv = sacc[yy * W + xx] + sacc[(h + yy) * W + xx];
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (ixt[4 * xx + k] == xs) acc = fmaf(wxt[4 * xx + k], sv, acc);
    }
    dp[j] = acc;
  }
  __syncthreads();
  // ---- softmax backward: ds_j = p_j (dp_j - sum_k p_k dp_k) / beta ----
  float ldot = 0.f;
  for (int j = tid; j < n; j += blockDim.x) ldot = fmaf(p[j], dp[j], ldot);
  const float dot = block_sum(ldot, scratch);
  for (int j = tid; j < n; j += blockDim.x) dp[j] = p[j] * (dp[j] - dot) / a.beta;
  __syncthreads();
  // ---- dz = sign(z) ds ----
  if (NHWC_BF16) {
    bf16* dzb = reinterpret_cast<bf16*>(a.dz) + (long long)b * n * C;
    const int vec_per_pos = C >> 3;
    for (int i = tid; i < n * vec_per_pos; i += blockDim.x) {
      const int j = i / vec_per_pos;
      const float ds = dp[j];
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(zb16) + i);
      auto sg = [&](float zv) { return zv > 0.f ? ds : (zv < 0.f ? -ds : 0.f); };
      uint4 o;
      o.x = pack_bf16x2(sg(bf16lo(v.x)), sg(bf16hi(v.x)));
      o.y = pack_bf16x2(sg(bf16lo(v.y)), sg(bf16hi(v.y)));
      o.z = pack_bf16x2(sg(bf16lo(v.z)), sg(bf16hi(v.z)));
      o.w = pack_bf16x2(sg(bf16lo(v.w)), sg(bf16hi(v.w)));
      reinterpret_cast<uint4*>(dzb)[i] = o;
    }
  } else {
    float* dzf = reinterpret_cast<float*>(a.dz) + (long long)b * n * C;
    for (int i = tid; i < n * C; i += blockDim.x) {
      const int j = i % n;
      const float zv = __ldg(zf32 + i), ds = dp[j];
      dzf[i] = zv > 0.f ? ds : (zv < 0.f ? -ds : 0.f);
    }
  }
}

size_t reg_smem_bytes(int h, int w, int H, int W) {
  const size_t n4 = ((size_t)h * w + 3) & ~(size_t)3;
  return sizeof(float) * (n4 + 3 * (size_t)h * W + 8 * (size_t)H + 8 * (size_t)W + n4 + 2 * (size_t)w + (size_t)H + 1);
}

template <bool NHWC, int MODE, int MAXT, int RC>
int launch_reg_rc(const RegArgs& a, int threads, size_t smem, cudaStream_t stream) {
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(reg_kernel<NHWC, MODE, MAXT, RC>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return gabril_set_cuda_error(e);
    attr_done = true;
  }
  reg_kernel<NHWC, MODE, MAXT, RC><<<a.B, threads, smem, stream>>>(a);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

template <bool NHWC, int MODE, int MAXT>
int launch_reg_t(const RegArgs& a, int threads, size_t smem, cudaStream_t stream) {
  // tap-window runs are exactly H / h rows when h divides H (180 / 20 = 9): one chunk per run
  if (a.H % a.h == 0 && a.H / a.h == 9) return launch_reg_rc<NHWC, MODE, MAXT, 9>(a, threads, smem, stream);
  return launch_reg_rc<NHWC, MODE, MAXT, 8>(a, threads, smem, stream);
}

template <bool NHWC, int MODE>
int launch_reg(const RegArgs& a, cudaStream_t stream) {
  if (a.B <= 0 || a.W > 1024 || a.W < 1 || a.h < 1 || a.w < 1 || (a.W & 3)) return GABRIL_ERR_BAD_SHAPE;
  if (NHWC) {
    const int lpp = a.C / 8;
    if (a.C % 8 || lpp > 32 || (lpp & (lpp - 1))) return GABRIL_ERR_BAD_SHAPE;
  }
  const size_t smem = reg_smem_bytes(a.h, a.w, a.H, a.W);
  if (smem > 200 * 1024) return GABRIL_ERR_BAD_SHAPE;
  const int Wr = ((a.W + 31) / 32) * 32;
  // two row slabs per column when the block still fits twice on an SM
  if (2 * Wr <= 640) return launch_reg_t<NHWC, MODE, 640>(a, 2 * Wr, smem, stream);
  if (Wr <= 640) return launch_reg_t<NHWC, MODE, 640>(a, Wr, smem, stream);
  return launch_reg_t<NHWC, MODE, 1024>(a, Wr, smem, stream);
}

int grid_for(long long total, int threads, int max_blocks = 148 * 16) {
  long long blocks = (total + threads - 1) / threads;
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace

extern "C" {

int gabril_obs_format(const uint8_t* frames, float* out, int B, int L, int S, int center, int H,
                      int W, int grayscale, cudaStream_t stream) {
  if (!frames || !out) return GABRIL_ERR_NULL;
  if (B <= 0 || S <= 0 || L <= 0 || W % 4) return GABRIL_ERR_BAD_SHAPE;
  const long long total = (long long)B * S * H * (W / 4);
  const int grid = grid_for(total, 256, 148 * 32);
  if (grayscale)
    obs_format_kernel<true><<<grid, 256, 0, stream>>>(frames, out, B, L, S, center, H, W);
  else
    obs_format_kernel<false><<<grid, 256, 0, stream>>>(frames, out, B, L, S, center, H, W);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

static int obs_patches_impl(const uint8_t* frames, void* patches, int B, int L, int S, int center,
                            int H, int W, int grayscale, int Kp, int SH, int SW, cudaStream_t stream,
                            const long long* widx = nullptr) {
  if (!frames || !patches) return GABRIL_ERR_NULL;
  const int C0 = S * (grayscale ? 1 : 3);
  if (B <= 0 || S < 1 || S > 4 || Kp < 16 * C0 || Kp % 4) return GABRIL_ERR_BAD_SHAPE;
  const int OH = (H + 2 - 4) / 2 + 1, OW = (W + 2 - 4) / 2 + 1;
  const long long total = (long long)B * OH * OW * 4;
  const int grid = grid_for(total, 256, 148 * 32);
  bf16* out = reinterpret_cast<bf16*>(patches);
  if (C0 % 2 == 0 && (W * 3) % 16 == 0 && W % 4 == 0 && (Kp - 16 * C0) % 8 == 0) {
    constexpr int NR = 2 * kPatchRows + 2;
    const size_t smem = (size_t)((NR * S * W * 3 + 15) & ~15) + (size_t)NR * (W + 2) * C0 * 2;
    const int tgrid = B * ((OH + kPatchRows - 1) / kPatchRows);
    if (smem <= 160 * 1024) {
#define GABRIL_TILED_CASE(SS, GG)                                                                  \
  do {                                                                                             \
    static bool attr_done = false;                                                                 \
    if (!attr_done) {                                                                              \
      cudaFuncSetAttribute(obs_patches_tiled_kernel<SS, GG>,                                       \
                           cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);               \
      attr_done = true;                                                                            \
    }                                                                                              \
    obs_patches_tiled_kernel<SS, GG><<<tgrid, 256, smem, stream>>>(frames, widx, out, L, center,   \
                                                                   H, W, OH, OW, Kp, SH, SW);      \
  } while (0)
      if (grayscale && S == 2) GABRIL_TILED_CASE(2, true);
      else if (grayscale && S == 4) GABRIL_TILED_CASE(4, true);
      else if (!grayscale && S == 2) GABRIL_TILED_CASE(2, false);
      else GABRIL_TILED_CASE(4, false);
#undef GABRIL_TILED_CASE
      GABRIL_CHECK_LAUNCH();
      return GABRIL_OK;
    }
  }
  if (SH > 0 || widx) return GABRIL_ERR_BAD_SHAPE;  // s2d row order / store addressing: tiled kernel only
#define GABRIL_PATCH_CASE(SS, GG)                                                              \
  obs_patches_kernel<SS, GG><<<grid, 256, 0, stream>>>(frames, out, B, L, center, H, W, OH, OW, Kp)
  if (grayscale) {
    switch (S) {
      case 1: GABRIL_PATCH_CASE(1, true); break;
      case 2: GABRIL_PATCH_CASE(2, true); break;
      case 3: GABRIL_PATCH_CASE(3, true); break;
      default: GABRIL_PATCH_CASE(4, true); break;
    }
  } else {
    switch (S) {
      case 1: GABRIL_PATCH_CASE(1, false); break;
      case 2: GABRIL_PATCH_CASE(2, false); break;
      case 3: GABRIL_PATCH_CASE(3, false); break;
      default: GABRIL_PATCH_CASE(4, false); break;
    }
  }
#undef GABRIL_PATCH_CASE
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_obs_patches(const uint8_t* frames, void* patches, int B, int L, int S, int center, int H,
                       int W, int grayscale, int Kp, cudaStream_t stream) {
  return obs_patches_impl(frames, patches, B, L, S, center, H, W, grayscale, Kp, 0, 0, stream);
}

int gabril_obs_patches_s2d(const uint8_t* frames, void* patches, int B, int L, int S, int center,
                           int H, int W, int grayscale, int Kp, int SH, int SW, cudaStream_t stream) {
  const int OH = (H + 2 - 4) / 2 + 1, OW = (W + 2 - 4) / 2 + 1;
  if (SH < (OH + 2) / 2 || SW < (OW + 2) / 2) return GABRIL_ERR_BAD_SHAPE;
  return obs_patches_impl(frames, patches, B, L, S, center, H, W, grayscale, Kp, SH, SW, stream);
}

int gabril_gmd_mask(const float* g, const float* A, float* K_ws, unsigned int* mnmx_ws,
                    float* rowmask, float* keep_out, int B, int S, int H, int W, int h, int w,
                    float drop_p, int test_mode, cudaStream_t stream) {
  return gabril_gmd_mask_sel(g, nullptr, A, K_ws, mnmx_ws, rowmask, keep_out, B, S, H, W, h, w, drop_p,
                             test_mode, stream);
}

int gabril_gmd_mask_sel(const float* g, const float* sel, const float* A, float* K_ws,
                        unsigned int* mnmx_ws, float* rowmask, float* keep_out, int B, int S, int H,
                        int W, int h, int w, float drop_p, int test_mode, cudaStream_t stream) {
  return gabril_gmd_mask_layout(g, sel, A, K_ws, mnmx_ws, rowmask, keep_out, B, S, H, W, h, w, drop_p,
                                test_mode, GABRIL_MASK_PLAIN, 0, 0, stream);
}

int gabril_gmd_mask_layout(const float* g, const float* sel, const float* A, float* K_ws,
                           unsigned int* mnmx_ws, float* rowmask, float* keep_out, int B, int S,
                           int H, int W, int h, int w, float drop_p, int test_mode, int layout,
                           int SH, int SW, cudaStream_t stream) {
  return gabril_gmd_mask_rng(g, sel, A, nullptr, 0, K_ws, mnmx_ws, rowmask, keep_out, B, S, H, W, h, w, drop_p,
                             test_mode, layout, SH, SW, stream);
}

int gabril_rng_advance(unsigned long long* rng_state, cudaStream_t stream) {
  if (!rng_state) return GABRIL_ERR_NULL;
  gmd_rng_advance_kernel<<<1, 1, 0, stream>>>(rng_state);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_gmd_mask_rng(const float* g, const float* sel, const float* A, const unsigned long long* rng_state,
                        int draw_id, float* K_ws, unsigned int* mnmx_ws, float* rowmask, float* keep_out, int B,
                        int S, int H, int W, int h, int w, float drop_p, int test_mode, int layout, int SH, int SW,
                        cudaStream_t stream) {
  if (layout == GABRIL_MASK_S2D && (2 * SH < h + 2 || 2 * SW < w + 2)) return GABRIL_ERR_BAD_SHAPE;
  if (layout == GABRIL_MASK_BORDERED && (SH < h + 2 || SW < w + 2)) return GABRIL_ERR_BAD_SHAPE;
  if (layout < GABRIL_MASK_PLAIN || layout > GABRIL_MASK_BORDERED) return GABRIL_ERR_BAD_SHAPE;
  if (!g || !K_ws || !mnmx_ws || !rowmask || (!A && !rng_state && !test_mode)) return GABRIL_ERR_NULL;
  if (B <= 0 || S < 1) return GABRIL_ERR_BAD_SHAPE;
  const long long total = (long long)B * h * w;
  gmd_init_kernel<<<1, 1, 0, stream>>>(mnmx_ws);
  gmd_keep_kernel<<<grid_for(total, 256, 148 * 8), 256, 0, stream>>>(g, sel, K_ws, mnmx_ws, B, S, H, W, h, w);
  gmd_mask_kernel<<<grid_for(total, 256, 148 * 8), 256, 0, stream>>>(K_ws, A, rng_state, draw_id, mnmx_ws, rowmask,
                                                                     keep_out, total, drop_p, test_mode, layout,
                                                                     h, w, SH, SW);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_rowmask_apply_nchw(const float* z, const float* rowmask, float* out, int B, int C, int n,
                              cudaStream_t stream) {
  if (!z || !rowmask || !out) return GABRIL_ERR_NULL;
  const long long total = (long long)B * C * n;
  rowmask_apply_nchw_kernel<<<grid_for(total, 256, 148 * 16), 256, 0, stream>>>(z, rowmask, out, B, C, n);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_gaze_mask_fwd(const void* z, int z_nhwc_bf16, float* mask_out, float* stats, int B, int C,
                         int h, int w, int H, int W, float beta, cudaStream_t stream) {
  if (!z || !mask_out) return GABRIL_ERR_NULL;
  RegArgs a = {};
  a.z = z; a.mask_out = mask_out; a.stats = stats; a.beta = beta;
  a.B = B; a.C = C; a.h = h; a.w = w; a.H = H; a.W = W;
  return z_nhwc_bf16 ? launch_reg<true, MODE_MASK_FWD>(a, stream)
                     : launch_reg<false, MODE_MASK_FWD>(a, stream);
}

int gabril_gaze_mask_bwd(const void* z, int z_nhwc_bf16, const float* grad_mask, void* dz, int B,
                         int C, int h, int w, int H, int W, float beta, cudaStream_t stream) {
  if (!z || !grad_mask || !dz) return GABRIL_ERR_NULL;
  RegArgs a = {};
  a.z = z; a.grad_in = grad_mask; a.dz = dz; a.beta = beta;
  a.B = B; a.C = C; a.h = h; a.w = w; a.H = H; a.W = W;
  return z_nhwc_bf16 ? launch_reg<true, MODE_MASK_BWD>(a, stream)
                     : launch_reg<false, MODE_MASK_BWD>(a, stream);
}

int gabril_reg_mse(const void* z, int z_nhwc_bf16, const float* g1, long long g1_bstride,
                   const float* sel, void* dz, float* loss_part, float* stats, int B, int C, int h,
                   int w, int H, int W, float beta, float grad_scale, cudaStream_t stream) {
  return gabril_reg_dist(z, z_nhwc_bf16, g1, g1_bstride, sel, nullptr, GABRIL_DIST_MSE, dz, loss_part,
                         stats, B, C, h, w, H, W, beta, grad_scale, stream);
}

int gabril_reg_dist(const void* z, int z_nhwc_bf16, const float* g1, long long g1_bstride,
                    const float* sel, const float* scale_dev, int dist, void* dz, float* loss_part,
                    float* stats, int B, int C, int h, int w, int H, int W, float beta,
                    float grad_scale, cudaStream_t stream) {
  if (!z || !g1 || !dz || !loss_part) return GABRIL_ERR_NULL;
  RegArgs a = {};
  a.z = z; a.g1 = g1; a.g1_bstride = g1_bstride; a.sel = sel; a.dz = dz; a.loss_part = loss_part;
  a.stats = stats; a.beta = beta; a.grad_scale = grad_scale; a.scale_dev = scale_dev;
  a.B = B; a.C = C; a.h = h; a.w = w; a.H = H; a.W = W;
  switch (dist) {
    case GABRIL_DIST_MSE:
      return z_nhwc_bf16 ? launch_reg<true, MODE_MSE>(a, stream) : launch_reg<false, MODE_MSE>(a, stream);
    case GABRIL_DIST_KL:
      return z_nhwc_bf16 ? launch_reg<true, MODE_KL>(a, stream) : launch_reg<false, MODE_KL>(a, stream);
    case GABRIL_DIST_TV:
      return z_nhwc_bf16 ? launch_reg<true, MODE_TV>(a, stream) : launch_reg<false, MODE_TV>(a, stream);
    case GABRIL_DIST_JS:
      return z_nhwc_bf16 ? launch_reg<true, MODE_JS>(a, stream) : launch_reg<false, MODE_JS>(a, stream);
    default:
      return GABRIL_ERR_BAD_SHAPE;
  }
}

int gabril_ivg_select(const uint8_t* frames, float* sel, float* sel_stats, int B, int L, int S,
                      int center, int H, int W, int grayscale, float ratio, cudaStream_t stream) {
  if (!frames || !sel || !sel_stats) return GABRIL_ERR_NULL;
  if (B <= 0 || B > 65535 * 16 || S < 1 || L < 1 || H < 1 || W < 1) return GABRIL_ERR_BAD_SHAPE;
  ivg_key_kernel<<<B, 256, 0, stream>>>(frames, nullptr, sel, B, L, S, center, H, W, grayscale, ratio);
  ivg_count_kernel<<<1, 256, 0, stream>>>(sel, sel_stats, B);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

// ---- device-resident demonstration store -----------------------------------------
int gabril_store_windows(const long long* step_idx, const long long* demo_first, long long* widx,
                         int B, int S, long long n_steps, cudaStream_t stream) {
  if (!step_idx || !demo_first || !widx) return GABRIL_ERR_NULL;
  if (B <= 0 || S < 1 || S > kMaxS || n_steps < 1) return GABRIL_ERR_BAD_SHAPE;
  const long long total = (long long)B * S;
  store_windows_kernel<<<grid_for(total, 256, 148 * 8), 256, 0, stream>>>(step_idx, demo_first, widx,
                                                                         B, S, n_steps);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_store_gather_rows(const float* src, const long long* widx, float* dst, long long rows,
                             int row_len, cudaStream_t stream) {
  if (!src || !widx || !dst) return GABRIL_ERR_NULL;
  if (rows <= 0 || row_len < 1) return GABRIL_ERR_BAD_SHAPE;
  const long long total = rows * row_len;
  store_gather_rows_kernel<<<grid_for(total, 256, 148 * 8), 256, 0, stream>>>(src, widx, dst, rows,
                                                                             row_len);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_obs_patches_s2d_store(const uint8_t* store, const long long* widx, void* patches, int B,
                                 int S, int H, int W, int grayscale, int Kp, int SH, int SW,
                                 cudaStream_t stream) {
  if (!widx) return GABRIL_ERR_NULL;
  const int OH = (H + 2 - 4) / 2 + 1, OW = (W + 2 - 4) / 2 + 1;
  if (SH < (OH + 2) / 2 || SW < (OW + 2) / 2) return GABRIL_ERR_BAD_SHAPE;
  return obs_patches_impl(store, patches, B, S, S, S - 1, H, W, grayscale, Kp, SH, SW, stream, widx);
}

int gabril_ivg_select_store(const uint8_t* store, const long long* widx, float* sel, float* sel_stats,
                            int B, int S, int H, int W, int grayscale, float ratio,
                            cudaStream_t stream) {
  if (!store || !widx || !sel || !sel_stats) return GABRIL_ERR_NULL;
  if (B <= 0 || B > 65535 * 16 || S < 1 || H < 1 || W < 1) return GABRIL_ERR_BAD_SHAPE;
  ivg_key_kernel<<<B, 256, 0, stream>>>(store, widx, sel, B, S, S, S - 1, H, W, grayscale, ratio);
  ivg_count_kernel<<<1, 256, 0, stream>>>(sel, sel_stats, B);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

}  // extern "C"
