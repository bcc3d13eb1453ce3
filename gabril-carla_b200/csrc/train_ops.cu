// Small fused kernels around the dense contractions of the GABRIL training step (sm_100a):
// bias gradients, bias + activation + cast, the action head's MSE loss and its gradient, the
// loss finaliser and the fused Adam update.  Reference ops replaced (vlm_gaze/):
//   train/train_bc.py:79-82,290-292   pre_actor / actor bias adds, ReLU, nn.MSELoss
//   train/train_bc.py:297             total = lambda * reg + actor
//   train/common/optim.py:11-28       torch.optim.Adam / AdamW (foreach implementation)
//   autograd of nn.Conv2d / nn.Linear bias (column sums of the output gradient)
#include "common.cuh"

namespace {

// db[c] += sum_rows dy[row][c];  dy bf16 [M, C], C % 8 == 0, C <= 2048.
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const bf16* __restrict__ dy, float* __restrict__ db, long long M, int C) {
  extern __shared__ float sm[];  // [rows_per_block][C]
  const int vec = C >> 3;                       // uint4 per row
  const int rpb = blockDim.x / vec;             // rows handled per block step
  const int cg = threadIdx.x % vec, rl = threadIdx.x / vec;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (rl < rpb) {
    // four independent 16-byte loads in flight per thread (ncu: one load per iteration left the
    // kernel latency-bound at 4.2-4.5 TB/s with 21 % issue activity); same per-thread summation order
    const long long stride = (long long)gridDim.x * rpb;
    long long row = (long long)blockIdx.x * rpb + rl;
#define GABRIL_COLSUM_ACC(v)                                                                     \
  acc[0] += bf16lo(v.x); acc[1] += bf16hi(v.x); acc[2] += bf16lo(v.y); acc[3] += bf16hi(v.y);    \
  acc[4] += bf16lo(v.z); acc[5] += bf16hi(v.z); acc[6] += bf16lo(v.w); acc[7] += bf16hi(v.w)
    for (; row + 3 * stride < M; row += 4 * stride) {
      const uint4 v0 = __ldg(reinterpret_cast<const uint4*>(dy + row * C) + cg);
      const uint4 v1 = __ldg(reinterpret_cast<const uint4*>(dy + (row + stride) * C) + cg);
      const uint4 v2 = __ldg(reinterpret_cast<const uint4*>(dy + (row + 2 * stride) * C) + cg);
      const uint4 v3 = __ldg(reinterpret_cast<const uint4*>(dy + (row + 3 * stride) * C) + cg);
      GABRIL_COLSUM_ACC(v0); GABRIL_COLSUM_ACC(v1); GABRIL_COLSUM_ACC(v2); GABRIL_COLSUM_ACC(v3);
    }
    for (; row < M; row += stride) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(dy + row * C) + cg);
      GABRIL_COLSUM_ACC(v);
    }
#undef GABRIL_COLSUM_ACC
#pragma unroll
    for (int i = 0; i < 8; ++i) sm[rl * C + cg * 8 + i] = acc[i];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int r = 0; r < rpb; ++r) s += sm[r * C + c];
    atomicAdd(db + c, s);
  }
}

// y = act(sum_s x[s] + bias) as bf16;  x fp32 [planes][M, C] (C % 4 == 0).
// Eight lanes share one float4 of the output: lane l adds the split-K planes l, l + 8, ... in order (four loads
// in flight), then a fixed xor-shuffle tree folds the eight partial sums: the summation order depends only on
// `planes`, so the result is bit-reproducible, and a small batch (few outputs, up to 148 planes) still has
// enough loads in flight to hide the memory latency.
__global__ void __launch_bounds__(256)
bias_act_cast_kernel(const float* __restrict__ x, const float* __restrict__ bias,
                     bf16* __restrict__ y, float* __restrict__ y_f32, long long M, int C, int relu,
                     int planes) {
  const long long total4 = M * C / 4;
  const int sub = threadIdx.x & 7;
  const long long n_iter = (total4 + 31) / 32 * 32;   // whole warps stay in the loop: the shuffles need every lane
  for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3; i < n_iter;
       i += ((long long)gridDim.x * blockDim.x) >> 3) {
    const bool live = i < total4;
    const float4* src = reinterpret_cast<const float4*>(x) + (live ? i : 0);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s0 = sub; s0 < planes; s0 += 32) {
      float4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        t[u] = (live && s0 + 8 * u < planes) ? __ldg(src + (long long)(s0 + 8 * u) * total4) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int u = 0; u < 4; ++u) { v.x += t[u].x; v.y += t[u].y; v.z += t[u].z; v.w += t[u].w; }
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      v.x += __shfl_xor_sync(0xffffffffu, v.x, o); v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
      v.z += __shfl_xor_sync(0xffffffffu, v.z, o); v.w += __shfl_xor_sync(0xffffffffu, v.w, o);
    }
    if (!live || sub != 0) continue;
    if (bias) {
      const float4 bb = __ldg(reinterpret_cast<const float4*>(bias + (int)((i * 4) % C)));
      v.x += bb.x; v.y += bb.y; v.z += bb.z; v.w += bb.w;
    }
    if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
    if (y_f32) reinterpret_cast<float4*>(y_f32)[i] = v;
    reinterpret_cast<uint2*>(y)[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
}

// logits = raw[:, :A] + bias; part[b] = sum_k (logits - a)^2; dlogits = 2 (logits - a) gscale
__global__ void __launch_bounds__(128)
mse_head_kernel(const float* __restrict__ raw, int ld_raw, const float* __restrict__ bias,
                const float* __restrict__ actions, long long act_bstride, int A,
                float* __restrict__ logits_out, bf16* __restrict__ dlogits, int ld_d,
                float* __restrict__ dbias, float* __restrict__ part, int B, float gscale) {
  __shared__ float sdb[64];
  if (threadIdx.x < 64) sdb[threadIdx.x] = 0.f;
  __syncthreads();
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = b < B;
  float sq = 0.f;
  for (int k = 0; k < A; ++k) {   // A is block-uniform: every lane reaches the shuffles
    float g = 0.f;
    if (live) {
      const float l = raw[(long long)b * ld_raw + k] + bias[k];
      const float d = l - actions[(long long)b * act_bstride + k];
      sq = fmaf(d, d, sq);
      g = 2.f * d * gscale;
      if (logits_out) logits_out[(long long)b * A + k] = l;
      if (dlogits) dlogits[(long long)b * ld_d + k] = __float2bfloat16(g);
    }
    if (dbias) {   // one shared atomic per warp instead of one per sample
      g = warp_sum(g);
      if ((threadIdx.x & 31) == 0) atomicAdd(&sdb[k], g);
    }
  }
  if (live) {
    if (dlogits)
      for (int k = A; k < ld_d; ++k) dlogits[(long long)b * ld_d + k] = __float2bfloat16(0.f);
    part[b] = sq;
  }
  __syncthreads();
  if (dbias && threadIdx.x < A) atomicAdd(dbias + threadIdx.x, sdb[threadIdx.x]);
}

// GRIL (train_bc.py:182-192): pred = raw[:, :P2] + bias over the samples that use gaze (sel > 0 or sel NULL);
// reg = clamp(mean((pred - target)^2) + 1e-8, 0, 100) -> reg_part[0] (the other entries 0, so that
// finalize_losses reads it with reg_scale 1); dpred (bf16 [B, ld_d], zero padded) = d(gscale * reg)/dpred,
// zero when the clamp saturates or no sample uses gaze; dbias[k] += sum_b dpred.  One block: fixed order.
__global__ void __launch_bounds__(1024)
gril_head_kernel(const float* __restrict__ raw, int ld_raw, const float* __restrict__ bias,
                 const float* __restrict__ target, long long tgt_bstride, const float* __restrict__ sel, int P2,
                 bf16* __restrict__ dpred, int ld_d, float* __restrict__ dbias, float* __restrict__ reg_part, int B,
                 float gscale) {
  __shared__ float s_sq[1024], s_n[1024], s_db[32];
  float sq = 0.f, n = 0.f;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    if (sel && sel[b] <= 0.f) continue;
    n += 1.f;
    for (int k = 0; k < P2; ++k) {
      const float d = raw[(long long)b * ld_raw + k] + bias[k] - target[(long long)b * tgt_bstride + k];
      sq = fmaf(d, d, sq);
    }
  }
  s_sq[threadIdx.x] = sq; s_n[threadIdx.x] = n;
  if (threadIdx.x < 32) s_db[threadIdx.x] = 0.f;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) { s_sq[threadIdx.x] += s_sq[threadIdx.x + o]; s_n[threadIdx.x] += s_n[threadIdx.x + o]; }
    __syncthreads();
  }
  const float n_sel = s_n[0];
  const float mse = n_sel > 0.f ? s_sq[0] / (n_sel * (float)P2) + 1e-8f : 0.f;
  const bool live = n_sel > 0.f && mse >= 0.f && mse <= 100.f;      // inside the clamp: gradient flows
  const float gs = live ? 2.f * gscale / (n_sel * (float)P2) : 0.f;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    reg_part[b] = (b == 0) ? (n_sel > 0.f ? fminf(fmaxf(mse, 0.f), 100.f) : 0.f) : 0.f;
    const bool use = !(sel && sel[b] <= 0.f);
    for (int k = 0; k < ld_d; ++k) {
      float g = 0.f;
      if (use && k < P2) {
        g = gs * (raw[(long long)b * ld_raw + k] + bias[k] - target[(long long)b * tgt_bstride + k]);
        if (dbias) atomicAdd(&s_db[k], g);
      }
      dpred[(long long)b * ld_d + k] = __float2bfloat16(g);
    }
  }
  __syncthreads();
  if (dbias && threadIdx.x < P2) dbias[threadIdx.x] += s_db[threadIdx.x];
}

// gaze.method = AGIL (train_bc.py:268-272): z = where(ivg > 0, 0.5 (z + z_agil), z) [* rowmask], bf16 NHWC in place
__global__ void __launch_bounds__(256)
agil_combine_kernel(bf16* __restrict__ z, const bf16* __restrict__ z2, const float* __restrict__ sel,
                    const float* __restrict__ rowmask, long long total2, int per_sample2, int C2) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total2;
       i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / per_sample2;
    const uint32_t a = reinterpret_cast<const uint32_t*>(z)[i], c = reinterpret_cast<const uint32_t*>(z2)[i];
    float lo = bf16lo(a), hi = bf16hi(a);
    if (!sel || sel[b] > 0.f) { lo = 0.5f * (lo + bf16lo(c)); hi = 0.5f * (hi + bf16hi(c)); }
    if (rowmask) { const float m = rowmask[i / C2]; lo *= m; hi *= m; }
    reinterpret_cast<uint32_t*>(z)[i] = pack_bf16x2(lo, hi);
  }
}

// its backward: dz_agil = [ivg > 0] 0.5 dz; dz = (ivg > 0 ? 0.5 : 1) dz (in place)
__global__ void __launch_bounds__(256)
agil_split_kernel(bf16* __restrict__ dz, bf16* __restrict__ dz2, const float* __restrict__ sel, long long total2,
                  int per_sample2) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total2;
       i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / per_sample2;
    const uint32_t a = reinterpret_cast<const uint32_t*>(dz)[i];
    const bool use = !sel || sel[b] > 0.f;
    const float lo = bf16lo(a), hi = bf16hi(a);
    reinterpret_cast<uint32_t*>(dz2)[i] = use ? pack_bf16x2(0.5f * lo, 0.5f * hi) : 0u;
    if (use) reinterpret_cast<uint32_t*>(dz)[i] = pack_bf16x2(0.5f * lo, 0.5f * hi);
  }
}

// losses[0] = actor = sum(actor_part) / (B*A); losses[1] = reg = sum(reg_part) * reg_scale;
// losses[2] = lambda * reg + actor.  One block, fixed summation order (deterministic).
__global__ void __launch_bounds__(256)
finalize_losses_kernel(const float* __restrict__ actor_part, const float* __restrict__ reg_part,
                       int B, float actor_scale, float reg_scale, const float* __restrict__ reg_scale_dev,
                       float lambda, float* __restrict__ losses) {
  __shared__ float sa[256], sr[256];
  float a = 0.f, r = 0.f;
  for (int i = threadIdx.x; i < B; i += 256) {
    a += actor_part[i];
    if (reg_part) r += reg_part[i];
  }
  sa[threadIdx.x] = a;
  sr[threadIdx.x] = r;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { sa[threadIdx.x] += sa[threadIdx.x + o]; sr[threadIdx.x] += sr[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float actor = sa[0] * actor_scale;
    const float reg = reg_scale_dev ? sr[0] * reg_scale * reg_scale_dev[0] : sr[0] * reg_scale;
    losses[0] = actor;
    losses[1] = reg;
    losses[2] = lambda * reg + actor;
  }
}

// torch.optim.Adam / AdamW single step over a flat fp32 arena; also refreshes the bf16 shadow
// the tensor-core kernels read.  g is scaled by gscale first (1 / world_size for data parallel).
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
            float* __restrict__ v, bf16* __restrict__ shadow, long long n4, float lr, float b1,
            float b2, float eps, float wd, int decoupled, float bc1, float bc2_sqrt, float gscale) {
  const float step_size = lr / bc1;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg4 = __ldg(reinterpret_cast<const float4*>(g) + i);
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float pa[4] = {pp.x, pp.y, pp.z, pp.w}, ga[4] = {gg4.x, gg4.y, gg4.z, gg4.w};
    float ma[4] = {mm.x, mm.y, mm.z, mm.w}, va[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float gk = ga[k] * gscale;
      if (wd != 0.f) {
        if (decoupled) pa[k] *= (1.f - lr * wd);
        else gk = fmaf(wd, pa[k], gk);
      }
      ma[k] = fmaf(1.f - b1, gk - ma[k], ma[k]);          // m.lerp_(g, 1 - b1)
      va[k] = fmaf(1.f - b2, gk * gk, b2 * va[k]);        // v.mul_(b2).addcmul_(g, g, 1 - b2)
      const float denom = sqrtf(va[k]) / bc2_sqrt + eps;
      pa[k] -= step_size * (ma[k] / denom);
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
    reinterpret_cast<float4*>(m)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
    reinterpret_cast<float4*>(v)[i] = make_float4(va[0], va[1], va[2], va[3]);
    if (shadow)
      reinterpret_cast<uint2*>(shadow)[i] =
          make_uint2(pack_bf16x2(pa[0], pa[1]), pack_bf16x2(pa[2], pa[3]));
  }
}

__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float* __restrict__ x, bf16* __restrict__ y, long long n4) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
    reinterpret_cast<uint2*>(y)[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
}

// t[row][c] *= scale[row]  (GMD backward: dz = dz' * M);  t bf16 [rows, C], C % 8 == 0
__global__ void __launch_bounds__(256)
rowscale_bf16_kernel(bf16* __restrict__ t, const float* __restrict__ scale, long long rows, int C) {
  const int vec = C >> 3;
  const long long total = rows * vec;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const float m = __ldg(scale + i / vec);
    uint4 v = reinterpret_cast<uint4*>(t)[i];
    v.x = pack_bf16x2(bf16lo(v.x) * m, bf16hi(v.x) * m);
    v.y = pack_bf16x2(bf16lo(v.y) * m, bf16hi(v.y) * m);
    v.z = pack_bf16x2(bf16lo(v.z) * m, bf16hi(v.z) * m);
    v.w = pack_bf16x2(bf16lo(v.w) * m, bf16hi(v.w) * m);
    reinterpret_cast<uint4*>(t)[i] = v;
  }
}

int blocks_for(long long work, int per_block, int cap) {
  long long b = (work + per_block - 1) / per_block;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace

extern "C" {

int gabril_colsum_bf16(const void* dy, float* db, long long M, int C, cudaStream_t stream) {
  if (!dy || !db) return GABRIL_ERR_NULL;
  if (C % 8 || C > 2048 || C < 8 || M <= 0) return GABRIL_ERR_BAD_SHAPE;
  const int vec = C / 8, rpb = 256 / vec;
  if (rpb < 1) return GABRIL_ERR_BAD_SHAPE;
  const size_t smem = sizeof(float) * (size_t)rpb * C;
  const int grid = blocks_for(M, rpb * 8, 148 * 4);
  colsum_bf16_kernel<<<grid, 256, smem, stream>>>(reinterpret_cast<const bf16*>(dy), db, M, C);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_bias_act_cast(const float* x, const float* bias, void* y_bf16, float* y_f32, long long M,
                         int C, int relu, cudaStream_t stream) {
  return gabril_bias_act_cast_planes(x, 1, bias, y_bf16, y_f32, M, C, relu, stream);
}

int gabril_bias_act_cast_planes(const float* x, int planes, const float* bias, void* y_bf16, float* y_f32,
                                long long M, int C, int relu, cudaStream_t stream) {
  if (!x || !y_bf16) return GABRIL_ERR_NULL;
  if (C % 4 || M <= 0 || planes < 1) return GABRIL_ERR_BAD_SHAPE;
  const int grid = blocks_for(M * C / 4 * 8, 256, 148 * 8);
  bias_act_cast_kernel<<<grid, 256, 0, stream>>>(x, bias, reinterpret_cast<bf16*>(y_bf16), y_f32, M,
                                                 C, relu, planes);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_mse_head(const float* raw, int ld_raw, const float* bias, const float* actions,
                    long long act_bstride, int A, float* logits_out, void* dlogits_bf16, int ld_d,
                    float* dbias, float* part, int B, float gscale, cudaStream_t stream) {
  if (!raw || !bias || !actions || !part) return GABRIL_ERR_NULL;
  if (A < 1 || A > 64 || B <= 0 || (dlogits_bf16 && ld_d < A)) return GABRIL_ERR_BAD_SHAPE;
  mse_head_kernel<<<(B + 127) / 128, 128, 0, stream>>>(raw, ld_raw, bias, actions, act_bstride, A,
                                                       logits_out,
                                                       reinterpret_cast<bf16*>(dlogits_bf16), ld_d,
                                                       dbias, part, B, gscale);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_gril_head(const float* raw, int ld_raw, const float* bias, const float* target, long long tgt_bstride,
                     const float* sel, int P2, void* dpred_bf16, int ld_d, float* dbias, float* reg_part, int B,
                     float gscale, cudaStream_t stream) {
  if (!raw || !bias || !target || !dpred_bf16 || !reg_part) return GABRIL_ERR_NULL;
  if (P2 < 1 || P2 > 32 || ld_d < P2 || B <= 0) return GABRIL_ERR_BAD_SHAPE;
  gril_head_kernel<<<1, 1024, 0, stream>>>(raw, ld_raw, bias, target, tgt_bstride, sel, P2,
                                           reinterpret_cast<bf16*>(dpred_bf16), ld_d, dbias, reg_part, B, gscale);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_agil_combine(void* z, const void* z_agil, const float* sel, const float* rowmask, int B, int n, int C,
                        cudaStream_t stream) {
  if (!z || !z_agil) return GABRIL_ERR_NULL;
  if (B <= 0 || (C & 1)) return GABRIL_ERR_BAD_SHAPE;
  const long long total2 = (long long)B * n * C / 2;
  agil_combine_kernel<<<blocks_for(total2, 256, 148 * 8), 256, 0, stream>>>(
      reinterpret_cast<bf16*>(z), reinterpret_cast<const bf16*>(z_agil), sel, rowmask, total2, n * C / 2, C / 2);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_agil_split(void* dz, void* dz_agil, const float* sel, int B, int n, int C, cudaStream_t stream) {
  if (!dz || !dz_agil) return GABRIL_ERR_NULL;
  if (B <= 0 || (C & 1)) return GABRIL_ERR_BAD_SHAPE;
  const long long total2 = (long long)B * n * C / 2;
  agil_split_kernel<<<blocks_for(total2, 256, 148 * 8), 256, 0, stream>>>(
      reinterpret_cast<bf16*>(dz), reinterpret_cast<bf16*>(dz_agil), sel, total2, n * C / 2);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_finalize_losses(const float* actor_part, const float* reg_part, int B, float actor_scale,
                           float reg_scale, float lambda, float* losses, cudaStream_t stream) {
  return gabril_finalize_losses_sel(actor_part, reg_part, B, actor_scale, reg_scale, nullptr, lambda,
                                    losses, stream);
}

int gabril_finalize_losses_sel(const float* actor_part, const float* reg_part, int B,
                               float actor_scale, float reg_scale, const float* reg_scale_dev,
                               float lambda, float* losses, cudaStream_t stream) {
  if (!actor_part || !losses) return GABRIL_ERR_NULL;
  finalize_losses_kernel<<<1, 256, 0, stream>>>(actor_part, reg_part, B, actor_scale, reg_scale,
                                                reg_scale_dev, lambda, losses);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_adam(float* p, const float* g, float* m, float* v, void* shadow_bf16, long long n,
                float lr, float beta1, float beta2, float eps, float weight_decay, int decoupled,
                int step, float grad_scale, cudaStream_t stream) {
  if (!p || !g || !m || !v) return GABRIL_ERR_NULL;
  if (n <= 0 || n % 4 || step < 1) return GABRIL_ERR_BAD_SHAPE;
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  const int grid = blocks_for(n / 4, 256, 148 * 8);
  adam_kernel<<<grid, 256, 0, stream>>>(p, g, m, v, reinterpret_cast<bf16*>(shadow_bf16), n / 4, lr,
                                        beta1, beta2, eps, weight_decay, decoupled, (float)bc1,
                                        (float)sqrt(bc2), grad_scale);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_rowscale_bf16(void* t, const float* scale, long long rows, int C, cudaStream_t stream) {
  if (!t || !scale) return GABRIL_ERR_NULL;
  if (C % 8 || rows <= 0) return GABRIL_ERR_BAD_SHAPE;
  rowscale_bf16_kernel<<<blocks_for(rows * (C / 8), 256, 148 * 8), 256, 0, stream>>>(
      reinterpret_cast<bf16*>(t), scale, rows, C);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

int gabril_cast_f32_bf16(const float* x, void* y, long long n, cudaStream_t stream) {
  if (!x || !y) return GABRIL_ERR_NULL;
  if (n <= 0 || n % 4) return GABRIL_ERR_BAD_SHAPE;
  cast_f32_bf16_kernel<<<blocks_for(n / 4, 256, 148 * 8), 256, 0, stream>>>(
      x, reinterpret_cast<bf16*>(y), n / 4);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

}  // extern "C"
