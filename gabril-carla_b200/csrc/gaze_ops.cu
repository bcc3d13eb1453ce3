// Memory-bound kernels of the GABRIL training step (sm_100a): frame formatting, gaze heat-map
// synthesis and gaze-mask dropout (the activation-map regulariser lives in reg_dist.cu).
// Reference ops replaced (Gonglitian/GABRIL-CARLA, vlm_gaze/):
//   data_utils/data_loader_robomimic.py:144-157  _gather_last_s_frames      (integer, bit-exact)
//   data_utils/data_loader_robomimic.py:179-202  _format_obs_image          gabril_obs_format / _patches
//   data_utils/data_loader_robomimic.py:85-139   GazePreprocessor.forward   gabril_gaze_splat
//   data_utils/data_loader_robomimic.py:204-278  build_stack_heatmaps       gabril_gaze_splat
//   data_utils/gaze_utils.py:33-52               apply_gmd_dropout          gabril_gmd_keep / _mask
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

// =================================================================================================
// K1a: uint8 NHWC frame stack -> fp32 NCHW (the reference-facing layout of _format_obs_image)
// =================================================================================================
__device__ __forceinline__ float u8_to_unit(uint32_t v) { return u8_over_255(v); }
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
