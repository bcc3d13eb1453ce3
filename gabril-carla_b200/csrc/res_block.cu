// One Residual block forward in ONE launch (reference: vlm_gaze/models/linear_models.py:78-102):
//   h   = relu(conv3x3(x; 128 -> 32, pad 1, no bias))          stored bf16 + sign words (backward needs them)
//   out = relu(x + conv1x1(h; 32 -> 128, no bias))             stored bf16 + sign words
// with x the already rectified input (the block's in-place ReLU).  As two tc_gemm launches the
// 1x1 convolution was a pure HBM round trip: h written and re-read, x re-read for the skip, at 2-6 % tensor
// pipe.  Here the 32-channel intermediate goes TMEM -> registers -> shared memory (as the A operand of the
// second MMA) and never waits for HBM.
//
// Warp roles (384 threads, one persistent CTA per SM):
//   warp 0        TMEM owner, single-thread tcgen05.mma issuer (3x3: 9 k-blocks of 8 MMAs 128x32x16 per tile;
//                 1x1: 2 MMAs 128x128x16 from the h tile in shared memory)
//   warps 1-3     TMA producers: im2col tiles of x into a 3-stage ring (one filter tap per stage); both weight
//                 tensors once per CTA
//   warps 4-7     epilogue group 0: acc1 -> ReLU -> bf16 -> h tile in shared memory (64-byte swizzle, K-major)
//                 + h / sign words to HBM; then column chunk 0 of the PREVIOUS tile's output
//   warps 8-11    epilogue group 1: column chunks 1-3 of the output (skip add, ReLU, sign words)
// Both accumulators are double buffered in TMEM (2 x 32 + 2 x 128 columns), so the 3x3 MMAs of tile i+1
// overlap both epilogues of tile i.
#include "res_block.cuh"

#include <algorithm>

#include "common.cuh"
#include "ptx.cuh"
#include "tc_epi.cuh"

namespace rb {

namespace {

constexpr int kThreads = 384;
constexpr int kProducers = 3;
constexpr uint32_t kABox = 16384;              // 128 pixels x 64 channels bf16
constexpr uint32_t kAStage = 2 * kABox;        // one filter tap: both channel chunks
constexpr uint32_t kW3Tile = 4096;             // 32 couts x 64 cin
constexpr uint32_t kW3Bytes = 18 * kW3Tile;    // 9 taps x 2 channel chunks
constexpr uint32_t kW1Bytes = 8192;            // 128 couts x 32 cin
constexpr uint32_t kTBytes = 8192;             // 128 pixels x 32 channels
constexpr uint32_t kOffW3 = kStages * kAStage;
constexpr uint32_t kOffW1 = kOffW3 + kW3Bytes;
constexpr uint32_t kOffT = kOffW1 + kW1Bytes;
constexpr uint32_t kAcc2Col = 64;              // TMEM: acc1 buffers at columns 0 / 32, acc2 at 64 / 192

__device__ __forceinline__ int fdiv(int n, const tc::FastDiv& d) {
  return (int)((__umulhi((uint32_t)n, d.mul) + (uint32_t)n) >> d.shr);
}
__device__ __forceinline__ void fail(const Params& p, int code) {
  if (p.err_flag) atomicExch(p.err_flag, code);
}

__global__ void __launch_bounds__(kThreads, 1) res_block_fwd_kernel(const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  // full[6] empty[6] wfull a1f[2] a1e[2] tr[2] tf[2] a2f[2] a2e[2]
  __shared__ __align__(8) uint64_t bars[2 * kStages + 1 + 12];
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t full0 = ptx::smem_u32(&bars[0]), empty0 = ptx::smem_u32(&bars[kStages]);
  const uint32_t wfull = ptx::smem_u32(&bars[2 * kStages]);
  const uint32_t a1f = wfull + 8, a1e = a1f + 16, tr = a1e + 16, tf = tr + 16, a2f = tf + 16, a2e = a2f + 16;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { ptx::mbar_init(full0 + 8 * s, 1); ptx::mbar_init(empty0 + 8 * s, 1); }
    ptx::mbar_init(wfull, 1);
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(a1f + 8 * b, 1);
      ptx::mbar_init(a1e + 8 * b, 4);   // the four warps of group 0
      ptx::mbar_init(tr + 8 * b, 4);
      ptx::mbar_init(tf + 8 * b, 1);
      ptx::mbar_init(a2f + 8 * b, 1);
      ptx::mbar_init(a2e + 8 * b, 8);   // both groups
    }
    ptx::fence_mbar_init();
  }
  ptx::pdl_launch_dependents();
  if (warp == 0) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_base_s), 512);
    ptx::tmem_relinquish();
  }
  ptx::pdl_wait();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  const int n_mine = p.tiles > (int)blockIdx.x ? (p.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp >= 1 && warp <= 3) {
    // ===================== TMA producers =====================
    if (ptx::elect_one()) {
      const int me = warp - 1;
      ptx::tma_prefetch_desc(&p.mapX);
      if (me == 0) {
        ptx::mbar_arrive_expect_tx(wfull, kW3Bytes + kW1Bytes);
        for (int tap = 0; tap < 9; ++tap)
          for (int ch = 0; ch < 2; ++ch)
            ptx::tma_load_3d(smem_base + kOffW3 + (uint32_t)(tap * 2 + ch) * kW3Tile, &p.mapW3, wfull, ch * 64, tap, 0);
        ptx::tma_load_3d(smem_base + kOffW1, &p.mapW1, wfull, 0, 0, 0);
      }
      int st = 0;
      uint32_t st_ph = 0;
      bool ok = true;
      for (int j = 0; j < n_mine && ok; ++j) {
        const int m0 = ((int)blockIdx.x + j * (int)gridDim.x) * 128;
        const int n_img = fdiv(m0, p.fd_hw);
        const int rem = m0 - n_img * p.HW;
        const int oh = fdiv(rem, p.fd_w);
        const int ow = rem - oh * p.W;
        for (int tap = 0; tap < 9; ++tap) {
          const int s = st;
          const uint32_t ph = st_ph;
          if (++st == kStages) { st = 0; st_ph ^= 1; }
          if (s % kProducers != me) continue;   // a stage always belongs to the same producer (parity waits)
          if (!ptx::mbar_wait(empty0 + 8 * s, ph ^ 1)) { fail(p, 201); ok = false; break; }
          const int rp = tap / 3, sp = tap - rp * 3;
          if (p.debug & 1) { ptx::mbar_arrive(full0 + 8 * s); continue; }
          ptx::mbar_arrive_expect_tx(full0 + 8 * s, kAStage);
#pragma unroll
          for (int ch = 0; ch < 2; ++ch)
            ptx::tma_load_im2col_4d(smem_base + (uint32_t)s * kAStage + (uint32_t)ch * kABox, &p.mapX, full0 + 8 * s,
                                    ch * 64, ow - 1, oh - 1, n_img, (uint16_t)sp, (uint16_t)rp);
        }
      }
    }
  } else if (warp == 0) {
    // ===================== MMA issuer =====================
    if (ptx::elect_one()) {
      const uint32_t a1_hi = (uint32_t)(p.desc_a1_hi >> 32), b1_hi = (uint32_t)(p.desc_b1_hi >> 32);
      const uint32_t a2_hi = (uint32_t)(p.desc_a2_hi >> 32), b2_hi = (uint32_t)(p.desc_b2_hi >> 32);
      const uint32_t a1_lo0 = (uint32_t)p.desc_a1_hi + (smem_base >> 4);
      const uint32_t b1_lo0 = (uint32_t)p.desc_b1_hi + ((smem_base + kOffW3) >> 4);
      const uint32_t a2_lo0 = (uint32_t)p.desc_a2_hi + ((smem_base + kOffT) >> 4);
      const uint32_t b2_lo0 = (uint32_t)p.desc_b2_hi + ((smem_base + kOffW1) >> 4);
      bool ok = ptx::mbar_wait(wfull, 0);
      if (!ok) fail(p, 202);
      ptx::tc_fence_after();
      // the 1x1 convolution of this CTA's j-th tile: A = h tile (shared memory), B = W1
      auto mma2 = [&](int j) -> bool {
        const int b = j & 1, u = j >> 1;
        if (!ptx::mbar_wait(a2e + 8 * b, (u & 1) ^ 1)) { fail(p, 203); return false; }
        if (!ptx::mbar_wait(tr + 8 * b, u & 1)) { fail(p, 204); return false; }
        ptx::tc_fence_after();
        const uint32_t d = tmem_base + kAcc2Col + (uint32_t)b * 128u;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const uint64_t ad = ((uint64_t)a2_hi << 32) | (uint64_t)(a2_lo0 + (uint32_t)b * (kTBytes >> 4) + k * 2u);
          const uint64_t bd = ((uint64_t)b2_hi << 32) | (uint64_t)(b2_lo0 + k * 2u);
          ptx::umma_bf16(d, ad, bd, p.idesc2, k > 0 ? 1u : 0u);
        }
        ptx::umma_commit(tf + 8 * b);    // the h tile may be overwritten
        ptx::umma_commit(a2f + 8 * b);   // the output accumulator is complete
        return true;
      };
      int s = 0;
      uint32_t ph = 0;
      for (int j = 0; j < n_mine && ok; ++j) {
        const int b = j & 1, u = j >> 1;
        if (!ptx::mbar_wait(a1e + 8 * b, (u & 1) ^ 1)) { fail(p, 205); ok = false; break; }
        ptx::tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)b * 32u;
        for (int tap = 0; tap < 9; ++tap) {
          if (!ptx::mbar_wait(full0 + 8 * s, ph)) { fail(p, 206); ok = false; break; }
          ptx::tc_fence_after();
          const uint32_t a_lo = a1_lo0 + (uint32_t)s * (kAStage >> 4);
          const uint32_t b_lo = b1_lo0 + (uint32_t)(tap * 2) * (kW3Tile >> 4);
#pragma unroll
          for (int k = 0; k < 8; ++k) {   // two 64-channel chunks x four K = 16 steps
            const uint32_t ch = (uint32_t)(k >> 2), kk = (uint32_t)(k & 3);
            const uint64_t ad = ((uint64_t)a1_hi << 32) | (uint64_t)(a_lo + ch * (kABox >> 4) + kk * 2u);
            const uint64_t bd = ((uint64_t)b1_hi << 32) | (uint64_t)(b_lo + ch * (kW3Tile >> 4) + kk * 2u);
            if (!(p.debug & 2)) ptx::umma_bf16(d, ad, bd, p.idesc1, (tap > 0 || k > 0) ? 1u : 0u);
          }
          ptx::umma_commit(empty0 + 8 * s);
          if (++s == kStages) { s = 0; ph ^= 1; }
        }
        if (!ok) break;
        ptx::umma_commit(a1f + 8 * b);
        if (j > 0) ok = mma2(j - 1);
      }
      if (ok && n_mine > 0) mma2(n_mine - 1);
    }
  } else {
    // ===================== epilogue warps =====================
    const int q = warp & 3, grp = (warp - 4) >> 2;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    constexpr uint32_t kEF2 = tc::EF_RESID | tc::EF_RELU | tc::EF_BITS;
    // output column chunks [c0, c1) of this CTA's j-th tile
    auto epi2 = [&](int j, int c0, int c1) -> bool {
      const int b = j & 1, u = j >> 1;
      const long long row = (long long)((int)blockIdx.x + j * (int)gridDim.x) * 128 + q * 32 + lane;
      const bool live = row < p.M;
      bool ok = ptx::mbar_wait(a2f + 8 * b, u & 1);
      if (!ok) fail(p, 207);
      if (!__all_sync(0xffffffffu, ok)) return false;
      ptx::tc_fence_after();
#pragma unroll 1
      for (int c = c0; c < c1 && !(p.debug & 4); ++c) {
        const long long off = row * 128 + c * 32;
        tc::epi_item_bf16<kEF2>(kEF2, lane_addr + kAcc2Col + (uint32_t)b * 128u + (uint32_t)c * 32u, live, nullptr,
                                p.x + off, nullptr, 0u, 1.0f, 0xffffffffu, p.out + off, p.bits_out + (off >> 5));
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(a2e + 8 * b);
      return true;
    };
    if (grp == 0) {
      for (int j = 0; j < n_mine; ++j) {
        const int b = j & 1, u = j >> 1;
        const long long row = (long long)((int)blockIdx.x + j * (int)gridDim.x) * 128 + q * 32 + lane;
        const bool live = row < p.M;
        bool ok = ptx::mbar_wait(a1f + 8 * b, u & 1);
        if (!ok) fail(p, 208);
        if (!__all_sync(0xffffffffu, ok)) break;
        ptx::tc_fence_after();
        uint32_t v[32];
        ptx::tmem_ld_32x32(lane_addr + (uint32_t)b * 32u, v);
        ptx::tmem_ld_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(a1e + 8 * b);   // acc1 is in registers: the next 3x3 tile may start
        uint32_t w[16];
        uint32_t s0 = 0u, s1 = 0u;
#pragma unroll
        for (int i = 0; i < 16; ++i) w[i] = ptx::relu_bf16x2(ptx::cvt_bf16x2(ptx::pack2(v[2 * i], v[2 * i + 1])));
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
          s0 |= ptx::gt0_mask_bf16x2(w[i]) & (0x00010001u << i);
          s1 |= ptx::gt0_mask_bf16x2(w[i + 1]) & (0x00010001u << (i + 1));
        }
        // the h tile buffer must have been consumed by the 1x1 MMAs of tile j - 2
        ok = ptx::mbar_wait(tf + 8 * b, (u & 1) ^ 1);
        if (!ok) fail(p, 209);
        if (!__all_sync(0xffffffffu, ok)) break;
        // K-major A operand, 64-byte rows in the 64-byte swizzle: 16-byte chunk c of row r at c ^ ((r >> 1) & 3)
        const int r = q * 32 + lane;
        const uint32_t trow = smem_base + kOffT + (uint32_t)b * kTBytes + (uint32_t)r * 64u;
        const uint32_t sw = (uint32_t)((r >> 1) & 3) << 4;
        ptx::sts128(trow + (0u ^ sw), w[0], w[1], w[2], w[3]);
        ptx::sts128(trow + (16u ^ sw), w[4], w[5], w[6], w[7]);
        ptx::sts128(trow + (32u ^ sw), w[8], w[9], w[10], w[11]);
        ptx::sts128(trow + (48u ^ sw), w[12], w[13], w[14], w[15]);
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(tr + 8 * b);
        if (live) {
          ptx::stg256(p.h + row * 32, w);
          ptx::stg256(p.h + row * 32 + 16, w + 8);
          p.bits_h[row] = s0 | s1;
        }
        if (j > 0 && !epi2(j - 1, 0, 1)) break;
      }
      if (n_mine > 0) epi2(n_mine - 1, 0, 1);
    } else {
      for (int j = 0; j < n_mine; ++j)
        if (!epi2(j, 1, 4)) break;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tmem_base, 512);
}

}  // namespace

size_t smem_bytes() { return (size_t)kOffT + 2 * kTBytes + 1024; }

int launch(const Params& p, int grid, cudaStream_t stream) {
  static cudaError_t attr_err = cudaFuncSetAttribute(reinterpret_cast<const void*>(res_block_fwd_kernel),
                                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes());
  if (attr_err != cudaSuccess) return gabril_set_cuda_error(attr_err);
  if (p.tiles <= 0) return GABRIL_OK;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)std::min(grid, p.tiles));
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem_bytes();
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  Params q = p;
  void* args[1] = {(void*)&q};
  cudaError_t le = cudaLaunchKernelExC(&cfg, reinterpret_cast<const void*>(res_block_fwd_kernel), args);
  if (le != cudaSuccess) return gabril_set_cuda_error(le);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

}  // namespace rb
