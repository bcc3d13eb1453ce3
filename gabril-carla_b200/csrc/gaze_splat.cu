// K2: gaze heat maps (sm_100a, HBM-write-bound).  Closed form of the reference's
//   scatter_add_ delta map -> two 1-D Gaussian conv2d passes (with the row-band quirk of their
//   padding) -> per-frame min-max (+1e-8) -> causal aggregation over the frame stack
//   -> second min-max (+1e-8)
// (vlm_gaze/data_utils/data_loader_robomimic.py:85-139 GazePreprocessor.forward, :204-278
// build_stack_heatmaps), straight from the gaze coordinates:
//   raw_s(y,x) = coeff_s * sum_p band_s(y) k_s[y_p - y + R_s] * k_s[x_p - x + R_s]
// Two launches:
//   splat_extrema  one CTA per sample: pixel indices (bit-exact integer math), the per-frame
//                  extrema of raw_s and of the aggregated maps -> 4*S floats per sample
//   splat_write    (sample, row-slab) CTAs: re-evaluate the closed form and stream the fp32 maps
//                  out, one coalesced 128-byte line per warp store.
// Thread x owns image column x: its k_s[x_p - x + R] factors stay in registers, the row factors
// are broadcast reads from shared memory.
#include <algorithm>
#include <cfloat>

#include "common.cuh"

namespace {

constexpr int kMaxS = 8;
constexpr int kMaxPts = 16;
constexpr int kThreads = 512;  // max; the launch uses W rounded up to a warp multiple

struct SplatArgs {
  const float* gaze;  // [B, L, P*2]
  float* out;         // [B, S, H, W]
  float* ext;         // [B, 4, kMaxS]: mn, mx (raw), amn, amx (aggregated)
  const float* ktab;  // [S][ktab_stride]
  int Lk[kMaxS];
  int ktab_stride;
  float agg[kMaxS][kMaxS];
  float step_coeff[kMaxS];
  int use_step_coeff, normalize_base, aggregate;
  int B, L, S, center, P, H, W, nsplit;
};

__device__ __forceinline__ int frame_index(int s, int S, int center, int L) {
  return min(max(center - S + 1 + s, 0), L - 1);  // data_loader_robomimic.py:151-152
}

// Shared per-CTA tables: valid flags, integer pixel coordinates, row factors ky[s][p][y].
template <int S>
struct Tables {
  int xi[S * kMaxPts], yi[S * kMaxPts], ok[S * kMaxPts];
};

template <int S>
__device__ __forceinline__ void load_points(const SplatArgs& a, int b, Tables<S>& t, float* ky) {
  const int P = a.P, H = a.H, W = a.W;
  if (threadIdx.x < S * P) {
    const int s = threadIdx.x / P, p = threadIdx.x - s * P;
    const int f = frame_index(s, S, a.center, a.L);
    const float* gp = a.gaze + (((long long)b * a.L + f) * P + p) * 2;
    const float gx = gp[0], gy = gp[1];
    // (:107-111) valid = x>=0 & y>=0; index = trunc(clamp(v,0,1) * (size-1)) clamped; the
    // product is ONE fp32 multiply, truncation toward zero: bit-exact integer indices
    t.ok[threadIdx.x] = (gx >= 0.f) && (gy >= 0.f);
    const float cx = fminf(fmaxf(gx, 0.f), 1.f), cy = fminf(fmaxf(gy, 0.f), 1.f);
    t.xi[threadIdx.x] = min(max((int)__fmul_rn(cx, (float)(W - 1)), 0), W - 1);
    t.yi[threadIdx.x] = min(max((int)__fmul_rn(cy, (float)(H - 1)), 0), H - 1);
  }
  __syncthreads();
  const int Hs = (H + 3) & ~3;  // row stride: four consecutive rows are one aligned float4
  for (int i = threadIdx.x; i < S * P * Hs; i += blockDim.x) {
    const int sp = i / Hs, y = i - sp * Hs;
    const int s = sp / P;
    const int Lk = a.Lk[s], R = Lk >> 1;
    const int d = t.yi[sp] - y + R;
    const bool in = t.ok[sp] && d >= 0 && d < Lk && y >= R && y < H - R;  // row band [R, H-R)
    ky[i] = in ? __ldg(a.ktab + s * a.ktab_stride + d) : 0.f;
  }
  __syncthreads();
}

// column factors of thread column x: kx[s][p]
template <int S, int PMAX>
__device__ __forceinline__ void column_factors(const SplatArgs& a, const Tables<S>& t, int x,
                                               float (&kx)[S][PMAX]) {
#pragma unroll
  for (int s = 0; s < S; ++s) {
    const int Lk = a.Lk[s], R = Lk >> 1;
#pragma unroll
    for (int p = 0; p < PMAX; ++p) {
      float v = 0.f;
      if (p < a.P) {
        const int d = t.xi[s * a.P + p] - x + R;
        if (t.ok[s * a.P + p] && d >= 0 && d < Lk) v = __ldg(a.ktab + s * a.ktab_stride + d);
      }
      kx[s][p] = v;
    }
  }
}

template <int S, int PMAX>
__device__ __forceinline__ void raw_at(const SplatArgs& a, const float* ky, const float (&kx)[S][PMAX],
                                       int y, float (&raw)[S]) {
  const int P = a.P, Hs = (a.H + 3) & ~3;
#pragma unroll
  for (int s = 0; s < S; ++s) {
    float acc = 0.f;
#pragma unroll
    for (int p = 0; p < PMAX; ++p)
      if (p < P) acc = fmaf(ky[(s * P + p) * Hs + y], kx[s][p], acc);
    raw[s] = a.use_step_coeff ? __fmul_rn(acc, a.step_coeff[s]) : acc;
  }
}

// four consecutive rows y4 .. y4+3 (y4 % 4 == 0): one 16-byte broadcast read per (s, p)
template <int S, int PMAX>
__device__ __forceinline__ void raw_at4(const SplatArgs& a, const float* ky, const float (&kx)[S][PMAX],
                                        int y4, float (&raw)[S][4]) {
  const int P = a.P, Hs = (a.H + 3) & ~3;
#pragma unroll
  for (int s = 0; s < S; ++s) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int p = 0; p < PMAX; ++p)
      if (p < P) {
        const float4 k4 = *reinterpret_cast<const float4*>(ky + (s * P + p) * Hs + y4);
        a0 = fmaf(k4.x, kx[s][p], a0); a1 = fmaf(k4.y, kx[s][p], a1);
        a2 = fmaf(k4.z, kx[s][p], a2); a3 = fmaf(k4.w, kx[s][p], a3);
      }
    if (a.use_step_coeff) {
      const float c = a.step_coeff[s];
      a0 = __fmul_rn(a0, c); a1 = __fmul_rn(a1, c); a2 = __fmul_rn(a2, c); a3 = __fmul_rn(a3, c);
    }
    raw[s][0] = a0; raw[s][1] = a1; raw[s][2] = a2; raw[s][3] = a3;
  }
}

// base_s = (raw - mn) / ((mx - mn) + 1e-8) (:134-136); agg_s = sum_{j<=s} c[s][j] base_j with every
// product and sum rounded separately, j ascending (:265-274)
// The two min-max divisions are multiplications by the per-(sample, frame) reciprocal
// inv = 1 / ((mx - mn) + 1e-8): <= 1 ulp away from the reference's division (heat-map tolerance 1e-6).
template <int S>
__device__ __forceinline__ void agg_at(const SplatArgs& a, const float (&raw)[S], const float (&mn)[S],
                                       const float (&inv)[S], float (&agg)[S]) {
  float base[S];
#pragma unroll
  for (int s = 0; s < S; ++s)
    base[s] = a.normalize_base ? __fmul_rn(__fsub_rn(raw[s], mn[s]), inv[s]) : raw[s];
#pragma unroll
  for (int s = 0; s < S; ++s) {
    if (!a.aggregate || s == 0) {
      agg[s] = base[s];
    } else {
      float acc = __fmul_rn(a.agg[s][0], base[0]);
#pragma unroll
      for (int j = 1; j <= s; ++j) acc = __fadd_rn(acc, __fmul_rn(a.agg[s][j], base[j]));
      agg[s] = acc;
    }
  }
}

template <int S>
__device__ __forceinline__ void block_minmax(float (&lo)[S], float (&hi)[S], float (*red)[2][kMaxS]) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
#pragma unroll
  for (int s = 0; s < S; ++s) {
    const float l = warp_min(lo[s]), h = warp_max(hi[s]);
    if (lane == 0) { red[wid][0][s] = l; red[wid][1][s] = h; }
  }
  __syncthreads();
#pragma unroll
  for (int s = 0; s < S; ++s) {
    float l = red[0][0][s], h = red[0][1][s];
    for (int w = 1; w < nw; ++w) { l = fminf(l, red[w][0][s]); h = fmaxf(h, red[w][1][s]); }
    lo[s] = l;
    hi[s] = h;
  }
}

template <int S>
__device__ __forceinline__ void band_of(const SplatArgs& a, int& ylo, int& yhi) {
  ylo = a.H;
  yhi = 0;
#pragma unroll
  for (int s = 0; s < S; ++s) {
    const int R = a.Lk[s] >> 1;
    ylo = min(ylo, R);
    yhi = max(yhi, a.H - R);
  }
}

template <int S>
__device__ __forceinline__ void make_inv(const SplatArgs& a, const float (&mn)[S], const float (&mx)[S],
                                         float (&inv)[S]) {
#pragma unroll
  for (int s = 0; s < S; ++s)
    inv[s] = a.normalize_base ? __fdiv_rn(1.f, __fadd_rn(__fsub_rn(mx[s], mn[s]), 1e-8f)) : 1.f;
}

template <int S, int PMAX>
__global__ void __launch_bounds__(kThreads) splat_extrema_kernel(const __grid_constant__ SplatArgs a) {
  extern __shared__ __align__(16) float ky[];  // [S*P][Hs]
  __shared__ Tables<S> t;
  __shared__ float red[kThreads / 32][2][kMaxS];
  const int b = blockIdx.x, H = a.H, W = a.W;
  load_points<S>(a, b, t, ky);
  int ylo, yhi;
  band_of<S>(a, ylo, yhi);
  const bool has_outside = (ylo > 0) || (yhi < H);  // rows where every raw_s is exactly 0
  // fast path: the band is a whole number of 4-row groups (sigma 30 / 20 at H = 180: [60,120), [40,140))
  const bool quad = has_outside && ((ylo | yhi) & 3) == 0;

  float mn[S], mx[S];
#pragma unroll
  for (int s = 0; s < S; ++s) { mn[s] = has_outside ? 0.f : FLT_MAX; mx[s] = has_outside ? 0.f : -FLT_MAX; }
  if (a.normalize_base) {
    for (int x = threadIdx.x; x < W; x += blockDim.x) {
      float kx[S][PMAX];
      column_factors<S, PMAX>(a, t, x, kx);
      if (quad) {
        for (int y = ylo; y < yhi; y += 4) {
          float raw[S][4];
          raw_at4<S, PMAX>(a, ky, kx, y, raw);
#pragma unroll
          for (int s = 0; s < S; ++s) {
            mn[s] = fminf(fminf(mn[s], fminf(raw[s][0], raw[s][1])), fminf(raw[s][2], raw[s][3]));
            mx[s] = fmaxf(fmaxf(mx[s], fmaxf(raw[s][0], raw[s][1])), fmaxf(raw[s][2], raw[s][3]));
          }
        }
      } else {
        for (int y = ylo; y < yhi; ++y) {
          float raw[S];
          raw_at<S, PMAX>(a, ky, kx, y, raw);
#pragma unroll
          for (int s = 0; s < S; ++s) { mn[s] = fminf(mn[s], raw[s]); mx[s] = fmaxf(mx[s], raw[s]); }
        }
      }
    }
    block_minmax<S>(mn, mx, red);
  } else {
#pragma unroll
    for (int s = 0; s < S; ++s) { mn[s] = 0.f; mx[s] = 1.f; }
  }
  float inv[S];
  make_inv<S>(a, mn, mx, inv);
  float amn[S], amx[S];
  float zero_raw[S], agg0[S];
#pragma unroll
  for (int s = 0; s < S; ++s) zero_raw[s] = 0.f;
  agg_at<S>(a, zero_raw, mn, inv, agg0);
#pragma unroll
  for (int s = 0; s < S; ++s) { amn[s] = has_outside ? agg0[s] : FLT_MAX; amx[s] = has_outside ? agg0[s] : -FLT_MAX; }
  if (a.aggregate) {
    for (int x = threadIdx.x; x < W; x += blockDim.x) {
      float kx[S][PMAX];
      column_factors<S, PMAX>(a, t, x, kx);
      if (quad) {
        for (int y = ylo; y < yhi; y += 4) {
          float raw4[S][4];
          raw_at4<S, PMAX>(a, ky, kx, y, raw4);
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            float raw[S], ag[S];
#pragma unroll
            for (int s = 0; s < S; ++s) raw[s] = raw4[s][r];
            agg_at<S>(a, raw, mn, inv, ag);
#pragma unroll
            for (int s = 0; s < S; ++s) { amn[s] = fminf(amn[s], ag[s]); amx[s] = fmaxf(amx[s], ag[s]); }
          }
        }
      } else {
        for (int y = ylo; y < yhi; ++y) {
          float raw[S], ag[S];
          raw_at<S, PMAX>(a, ky, kx, y, raw);
          agg_at<S>(a, raw, mn, inv, ag);
#pragma unroll
          for (int s = 0; s < S; ++s) { amn[s] = fminf(amn[s], ag[s]); amx[s] = fmaxf(amx[s], ag[s]); }
        }
      }
    }
    block_minmax<S>(amn, amx, red);
  }
  if (threadIdx.x == 0) {
    float* e = a.ext + (long long)b * 4 * kMaxS;
#pragma unroll
    for (int s = 0; s < S; ++s) {
      e[s] = mn[s]; e[kMaxS + s] = mx[s]; e[2 * kMaxS + s] = amn[s]; e[3 * kMaxS + s] = amx[s];
    }
  }
}

// grid (B, nsplit): CTA (b, split) streams the 4-row groups [g0, g1) of sample b
template <int S, int PMAX>
__global__ void __launch_bounds__(kThreads) splat_write_kernel(const __grid_constant__ SplatArgs a) {
  extern __shared__ __align__(16) float ky[];
  __shared__ Tables<S> t;
  const int b = blockIdx.x, split = blockIdx.y, H = a.H, W = a.W;
  load_points<S>(a, b, t, ky);
  int ylo, yhi;
  band_of<S>(a, ylo, yhi);
  float mn[S], mx[S], amn[S], ainv[S], inv[S];
  const float* e = a.ext + (long long)b * 4 * kMaxS;
#pragma unroll
  for (int s = 0; s < S; ++s) {
    mn[s] = e[s]; mx[s] = e[kMaxS + s]; amn[s] = e[2 * kMaxS + s];
    ainv[s] = __fdiv_rn(1.f, __fadd_rn(__fsub_rn(e[3 * kMaxS + s], amn[s]), 1e-8f));
  }
  make_inv<S>(a, mn, mx, inv);
  auto finish = [&](float v, int s) {
    return a.aggregate ? __fmul_rn(__fsub_rn(v, amn[s]), ainv[s]) : v;
  };
  float zero_raw[S], agg0[S], out0[S];
#pragma unroll
  for (int s = 0; s < S; ++s) zero_raw[s] = 0.f;
  agg_at<S>(a, zero_raw, mn, inv, agg0);
#pragma unroll
  for (int s = 0; s < S; ++s) out0[s] = finish(agg0[s], s);

  const int groups = (H + 3) >> 2;
  const int per = (groups + a.nsplit - 1) / a.nsplit;
  const int y0 = split * per * 4, y1 = min(y0 + per * 4, H);
  float* outb = a.out + (long long)b * S * H * W;
  for (int x = threadIdx.x; x < W; x += blockDim.x) {
    float kx[S][PMAX];
    column_factors<S, PMAX>(a, t, x, kx);
    for (int y = y0; y < y1; y += 4) {
      const int nr = min(4, y1 - y);
      if (y + 4 <= ylo || y >= yhi) {  // whole group outside the band: constant rows
        for (int r = 0; r < nr; ++r)
#pragma unroll
          for (int s = 0; s < S; ++s) outb[((long long)s * H + y + r) * W + x] = out0[s];
      } else {
        // ky is zero outside the band, so the closed form covers partially covered groups too
        float raw4[S][4];
        raw_at4<S, PMAX>(a, ky, kx, y, raw4);
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          if (r < nr) {
            float raw[S], ag[S];
#pragma unroll
            for (int s = 0; s < S; ++s) raw[s] = raw4[s][r];
            agg_at<S>(a, raw, mn, inv, ag);
#pragma unroll
            for (int s = 0; s < S; ++s) outb[((long long)s * H + y + r) * W + x] = finish(ag[s], s);
          }
        }
      }
    }
  }
}

template <int S, int PMAX>
int launch_splat(const SplatArgs& a, cudaStream_t stream) {
  const size_t smem = sizeof(float) * (size_t)S * a.P * ((a.H + 3) & ~3);
  if (smem > 96 * 1024) return GABRIL_ERR_BAD_SHAPE;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e1 = cudaFuncSetAttribute(splat_extrema_kernel<S, PMAX>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    cudaError_t e2 = cudaFuncSetAttribute(splat_write_kernel<S, PMAX>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    if (e1 != cudaSuccess) return gabril_set_cuda_error(e1);
    if (e2 != cudaSuccess) return gabril_set_cuda_error(e2);
    attr_done = true;
  }
  const int threads = std::min(kThreads, (a.W + 31) / 32 * 32);
  splat_extrema_kernel<S, PMAX><<<a.B, threads, smem, stream>>>(a);
  splat_write_kernel<S, PMAX><<<dim3(a.B, a.nsplit), threads, smem, stream>>>(a);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

}  // namespace

extern "C" int gabril_gaze_splat(const float* gaze, float* out, float* ext_ws, const float* ktab,
                                 const int* Lk, int ktab_stride, const float* agg_coeff,
                                 const float* step_coeff, int normalize_base, int aggregate, int B,
                                 int L, int S, int center, int P, int H, int W,
                                 cudaStream_t stream) {
  if (!gaze || !out || !ktab || !Lk || !ext_ws) return GABRIL_ERR_NULL;
  if (B <= 0 || S < 1 || S > kMaxS || P < 1 || P > kMaxPts || S * P > 512)
    return GABRIL_ERR_BAD_SHAPE;
  SplatArgs a;
  a.gaze = gaze; a.out = out; a.ext = ext_ws; a.ktab = ktab; a.ktab_stride = ktab_stride;
  for (int s = 0; s < kMaxS; ++s) {
    a.Lk[s] = s < S ? Lk[s] : 1;
    a.step_coeff[s] = (step_coeff && s < S) ? step_coeff[s] : 1.f;
    for (int j = 0; j < kMaxS; ++j)
      a.agg[s][j] = (agg_coeff && s < S && j <= s) ? agg_coeff[s * S + j] : (j <= s ? 1.f : 0.f);
    // the reference's conv2d needs H >= kernel size (valid vertical pass); kernels are odd
    if (s < S && (Lk[s] > H || Lk[s] < 1 || Lk[s] > ktab_stride || (Lk[s] & 1) == 0))
      return GABRIL_ERR_BAD_SHAPE;
  }
  a.use_step_coeff = step_coeff != nullptr;
  a.normalize_base = normalize_base; a.aggregate = aggregate;
  a.B = B; a.L = L; a.S = S; a.center = center; a.P = P; a.H = H; a.W = W;
  int nsplit = 1;  // row slabs per sample: enough CTAs to keep every SM streaming stores
  while ((long long)B * nsplit < 8 * 148 && nsplit < 16 && nsplit * 8 <= H) nsplit *= 2;
  a.nsplit = nsplit;
#define GABRIL_SPLAT_CASE(SS)                                                     \
  case SS: return P <= 8 ? launch_splat<SS, 8>(a, stream) : launch_splat<SS, 16>(a, stream)
  switch (S) {
    GABRIL_SPLAT_CASE(1); GABRIL_SPLAT_CASE(2); GABRIL_SPLAT_CASE(3); GABRIL_SPLAT_CASE(4);
    GABRIL_SPLAT_CASE(5); GABRIL_SPLAT_CASE(6); GABRIL_SPLAT_CASE(7); GABRIL_SPLAT_CASE(8);
  }
#undef GABRIL_SPLAT_CASE
  return GABRIL_ERR_BAD_SHAPE;
}

// ---- allocation-free variant ------------------------------------------------------------------
// gabril_gaze_splat with the extrema workspace in a file-scope device array instead of a caller
// buffer, for callers that cannot allocate (CUDA-graph capture, serving loops).  No counterpart in
// the reference.  The array is 512 KB per GPU that
// loads the library, held until exit.  Every call shares it: calls that overlap on the GPU (two
// streams, two threads) race on it.  Nothing in it has to start at zero - the extrema launch writes
// every slot [b][0..3][0..S) that the write launch reads - so it is never cleared.
namespace {

constexpr int kStaticMaxB = 4096;  // samples per call
constexpr int kStaticMaxDev = 64;

static __device__ __align__(256) float g_splat_ext_static[kStaticMaxB * 4 * kMaxS];

// In host code g_splat_ext_static is only the symbol's host-side stand-in: its device address is
// looked up once per device, with that device current, and that pointer is what the kernels get.
float* splat_static_ws() {
  static float* ws[kStaticMaxDev] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { gabril_set_cuda_error(e); return nullptr; }
  if (dev < 0 || dev >= kStaticMaxDev) {
    gabril_set_error("gabril_gaze_splat_static: device ordinal beyond 63");
    return nullptr;
  }
  float* p = __atomic_load_n(&ws[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    void* sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_splat_ext_static);
    if (e != cudaSuccess || !sym) {
      gabril_set_cuda_error(e != cudaSuccess ? e : cudaErrorInvalidSymbol);
      return nullptr;
    }
    p = static_cast<float*>(sym);
    __atomic_store_n(&ws[dev], p, __ATOMIC_RELEASE);  // racing threads store the same address
  }
  return p;
}

}  // namespace

extern "C" int gabril_gaze_splat_static_max_batch(void) { return kStaticMaxB; }

extern "C" int gabril_gaze_splat_static(const float* gaze, float* out, const float* ktab, const int* Lk,
                                        int ktab_stride, const float* agg_coeff, const float* step_coeff,
                                        int normalize_base, int aggregate, int B, int L, int S,
                                        int center, int P, int H, int W, cudaStream_t stream) {
  if (B > kStaticMaxB) {
    gabril_set_error("gabril_gaze_splat_static: B exceeds the static workspace (4096 samples per call); "
                     "split the batch or use gabril_gaze_splat with a caller workspace");
    return GABRIL_ERR_BAD_SHAPE;
  }
  float* ws = splat_static_ws();
  if (!ws) return GABRIL_ERR_CUDA;
  return gabril_gaze_splat(gaze, out, ws, ktab, Lk, ktab_stride, agg_coeff, step_coeff, normalize_base,
                           aggregate, B, L, S, center, P, H, W, stream);
}
