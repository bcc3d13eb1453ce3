// tcgen05 / TMEM / TMA implicit-GEMM kernel and its host-side planners.
// See tc_gemm.cuh for the operand regimes.  Reference ops replaced:
//   nn.Conv2d forward/backward in Encoder (vlm_gaze/models/linear_models.py:141-207),
//   Residual 3x3 / 1x1 convs (linear_models.py:78-102), pre_actor Linear (train_bc.py:79).
#include "tc_gemm.cuh"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "common.cuh"
#include "ptx.cuh"

namespace tc {

// ---------------------------------------------------------------------------------------------
// device
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void report_timeout(const Params& p, int code) {
  if (p.err_flag) atomicExch(p.err_flag, code);
}

__global__ void __launch_bounds__(kThreads) tc_gemm_kernel(const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * kMaxStages + 1];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cls = p.n_classes > 1 ? (int)blockIdx.z : 0;
  const int split = p.n_classes > 1 ? 0 : (int)blockIdx.z;
  const int M_total = p.M_total[cls];
  const int m0 = blockIdx.x * 128;
  if (m0 >= M_total) return;
  const int kb_begin = split * p.kb_per_split;
  const int kb_end = min(kb_begin + p.kb_per_split, p.kb_total);
  if (kb_begin >= kb_end) return;

  int n_tile, tg;
  if (p.kind == KIND_WGRAD) {
    n_tile = blockIdx.y / p.tap_groups;
    tg = blockIdx.y - n_tile * p.tap_groups;
  } else {
    n_tile = blockIdx.y;
    tg = 0;
  }
  const int n0 = n_tile * p.block_n;
  const int tap0 = tg * p.n_acc;
  const int ntaps = (p.kind == KIND_WGRAD) ? min(p.n_acc, p.taps - tap0) : 1;

  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t full0 = ptx::smem_u32(&bars[0]);
  const uint32_t empty0 = ptx::smem_u32(&bars[kMaxStages]);
  const uint32_t tfull = ptx::smem_u32(&bars[2 * kMaxStages]);
  const uint32_t stage_bytes = p.a_stage_bytes + p.b_stage_bytes;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      ptx::mbar_init(full0 + 8 * s, 1);
      ptx::mbar_init(empty0 + 8 * s, 1);
    }
    ptx::mbar_init(tfull, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_base_s), p.tmem_cols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    // ===================== TMA producer (one thread) =====================
    if (lane == 0) {
      ptx::tma_prefetch_desc(&p.mapA[cls]);
      ptx::tma_prefetch_desc(&p.mapB);
      const int ohw = p.OHW[cls], oww = p.OW[cls];
      int it = 0;
      if (p.kind == KIND_FWD) {
        int w_base = 0, h_base = 0, n_img = 0;
        if (p.a_im2col) {
          n_img = m0 / ohw;
          const int rem = m0 - n_img * ohw;
          const int oh = rem / oww;
          const int ow = rem - oh * oww;
          w_base = ow * p.trav_stride + p.lower_w[cls];
          h_base = oh * p.trav_stride + p.lower_h[cls];
        }
        const uint32_t tx = p.a_boxes * p.a_box_bytes + p.b_boxes * p.b_box_bytes;
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int s = it % p.stages;
          const uint32_t ph = (it / p.stages) & 1;
          if (!ptx::mbar_wait(empty0 + 8 * s, ph ^ 1)) { report_timeout(p, 101); break; }
          const uint32_t a_dst = smem_base + s * stage_bytes;
          const uint32_t b_dst = a_dst + p.a_stage_bytes;
          const uint32_t full = full0 + 8 * s;
          ptx::mbar_arrive_expect_tx(full, tx);
          const int tap = kb / p.chunks_per_tap;
          const int chunk = kb - tap * p.chunks_per_tap;
          const int rp = tap / p.S, sp = tap - rp * p.S;
          if (p.a_im2col)
            ptx::tma_load_im2col_4d(a_dst, &p.mapA[cls], full, chunk * p.kc, w_base, h_base, n_img,
                                    (uint16_t)sp, (uint16_t)rp);
          else
            ptx::tma_load_2d(a_dst, &p.mapA[0], full, kb * p.kc, m0);
          const int wtap =
              (p.r_base[cls] + p.r_step * rp) * p.S_w + p.s_base[cls] + p.s_step * sp;
          if (!p.b_mn) {
            ptx::tma_load_3d(b_dst, &p.mapB, full, chunk * p.kc, wtap, n0);
          } else {
            for (int i = 0; i < p.b_boxes; ++i)
              ptx::tma_load_3d(b_dst + i * p.b_box_bytes, &p.mapB, full, n0 + i * p.b_box_w, wtap,
                               chunk * p.kc);
          }
        }
      } else {
        const uint32_t tx = p.a_boxes * p.a_box_bytes + ntaps * p.b_boxes * p.b_box_bytes;
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int s = it % p.stages;
          const uint32_t ph = (it / p.stages) & 1;
          if (!ptx::mbar_wait(empty0 + 8 * s, ph ^ 1)) { report_timeout(p, 102); break; }
          const uint32_t a_dst = smem_base + s * stage_bytes;
          const uint32_t b_dst = a_dst + p.a_stage_bytes;
          const uint32_t full = full0 + 8 * s;
          ptx::mbar_arrive_expect_tx(full, tx);
          const int p0 = kb * p.kc;
          for (int i = 0; i < p.a_boxes; ++i)
            ptx::tma_load_2d(a_dst + i * p.a_box_bytes, &p.mapA[0], full, m0 + i * p.a_box_w, p0);
          if (p.a_im2col) {
            const int n_img = p0 / ohw;
            const int rem = p0 - n_img * ohw;
            const int oh = rem / oww;
            const int ow = rem - oh * oww;
            const int w_base = ow * p.trav_stride + p.lower_w[0];
            const int h_base = oh * p.trav_stride + p.lower_h[0];
            for (int t = 0; t < ntaps; ++t) {
              const int tap = tap0 + t;
              const int rp = tap / p.S, sp = tap - rp * p.S;
              for (int i = 0; i < p.b_boxes; ++i)
                ptx::tma_load_im2col_4d(b_dst + (t * p.b_boxes + i) * p.b_box_bytes, &p.mapB, full,
                                        n0 + i * p.b_box_w, w_base, h_base, n_img, (uint16_t)sp,
                                        (uint16_t)rp);
            }
          } else {
            for (int i = 0; i < p.b_boxes; ++i)
              ptx::tma_load_2d(b_dst + i * p.b_box_bytes, &p.mapB, full, n0 + i * p.b_box_w, p0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      int it = 0;
      bool ok = true;
      for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
        const int s = it % p.stages;
        const uint32_t ph = (it / p.stages) & 1;
        if (!ptx::mbar_wait(full0 + 8 * s, ph)) { report_timeout(p, 103); ok = false; break; }
        ptx::tc_fence_after();
        const uint32_t a_base = (smem_base + s * stage_bytes) >> 4;
        const uint32_t b_base = a_base + (p.a_stage_bytes >> 4);
        for (int j = 0; j < p.n_mma; ++j) {
          if (p.mma_acc[j] >= ntaps) continue;
          const uint64_t a_desc = p.desc_a_hi | (uint64_t)((a_base + p.mma_a_off[j]) & 0x3FFFu);
          const uint64_t b_desc = p.desc_b_hi | (uint64_t)((b_base + p.mma_b_off[j]) & 0x3FFFu);
          const uint32_t accumulate = (it > 0 || !p.mma_first[j]) ? 1u : 0u;
          ptx::umma_bf16(tmem_base + p.mma_dcol[j], a_desc, b_desc, p.idesc, accumulate);
        }
        ptx::umma_commit(empty0 + 8 * s);
      }
      if (ok) ptx::umma_commit(tfull);
    }
  } else {
    // ===================== epilogue (4 warps, one TMEM lane quarter each) =====================
    const int q = warp & 3;
    const int row = m0 + q * 32 + lane;
    const bool row_ok = row < M_total;
    bool ok = ptx::mbar_wait(tfull, 0);
    if (!ok) report_timeout(p, 104);
    ok = __all_sync(0xffffffffu, ok);
    if (ok) {
      ptx::tc_fence_after();
      long long orow = row;
      if (p.remap && row_ok) {
        const int hw = p.Hs[cls] * p.Ws[cls];
        const int n = row / hw;
        const int rem = row - n * hw;
        const int jh = rem / p.Ws[cls];
        const int jw = rem - jh * p.Ws[cls];
        orow = ((long long)n * p.OHf + (2 * jh + p.ah[cls])) * p.OWf + (2 * jw + p.aw[cls]);
      }
      const float rmask = (p.rowmask && row_ok) ? p.rowmask[orow] : 1.0f;
      for (int a = 0; a < ntaps; ++a) {
        for (int c = 0; c < p.block_n; c += 32) {
          uint32_t v[32];
          ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + a * p.block_n + c, v);
          ptx::tmem_ld_wait();
          const int col0 = n0 + c;
          if (!row_ok || col0 >= p.N_total) continue;
          float f[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
          if (p.out_mode == OUT_BF16) {
            const long long off = orow * p.ldo + col0;
            if (p.bias) {
              const float4* b4 = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const float4 b = __ldg(b4 + i);
                f[4 * i + 0] += b.x; f[4 * i + 1] += b.y; f[4 * i + 2] += b.z; f[4 * i + 3] += b.w;
              }
            }
            if (p.resid) {
              const uint4* r4 = reinterpret_cast<const uint4*>(p.resid + off);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 r = __ldg(r4 + i);
                f[8 * i + 0] += bf16lo(r.x); f[8 * i + 1] += bf16hi(r.x);
                f[8 * i + 2] += bf16lo(r.y); f[8 * i + 3] += bf16hi(r.y);
                f[8 * i + 4] += bf16lo(r.z); f[8 * i + 5] += bf16hi(r.z);
                f[8 * i + 6] += bf16lo(r.w); f[8 * i + 7] += bf16hi(r.w);
              }
            }
            if (p.relu) {
#pragma unroll
              for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.0f);
            }
            if (p.gate) {
              const uint4* g4 = reinterpret_cast<const uint4*>(p.gate + off);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 g = __ldg(g4 + i);
                f[8 * i + 0] = bf16lo(g.x) > 0.f ? f[8 * i + 0] : 0.f;
                f[8 * i + 1] = bf16hi(g.x) > 0.f ? f[8 * i + 1] : 0.f;
                f[8 * i + 2] = bf16lo(g.y) > 0.f ? f[8 * i + 2] : 0.f;
                f[8 * i + 3] = bf16hi(g.y) > 0.f ? f[8 * i + 3] : 0.f;
                f[8 * i + 4] = bf16lo(g.z) > 0.f ? f[8 * i + 4] : 0.f;
                f[8 * i + 5] = bf16hi(g.z) > 0.f ? f[8 * i + 5] : 0.f;
                f[8 * i + 6] = bf16lo(g.w) > 0.f ? f[8 * i + 6] : 0.f;
                f[8 * i + 7] = bf16hi(g.w) > 0.f ? f[8 * i + 7] : 0.f;
              }
            }
            if (p.rowmask) {
#pragma unroll
              for (int i = 0; i < 32; ++i) f[i] *= rmask;
            }
            uint4* o4 = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(p.out) + off);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 o;
              o.x = pack_bf16x2(f[8 * i + 0], f[8 * i + 1]);
              o.y = pack_bf16x2(f[8 * i + 2], f[8 * i + 3]);
              o.z = pack_bf16x2(f[8 * i + 4], f[8 * i + 5]);
              o.w = pack_bf16x2(f[8 * i + 6], f[8 * i + 7]);
              o4[i] = o;
            }
          } else {
            float* o = reinterpret_cast<float*>(p.out) + orow * p.ldo +
                       (long long)(tap0 + a) * p.tap_stride + col0;
            if (p.out_mode == OUT_F32_STORE) {
              float4* o4 = reinterpret_cast<float4*>(o);
#pragma unroll
              for (int i = 0; i < 8; ++i)
                o4[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i)
                ptx::red_add_v4_f32(o + 4 * i, f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
            }
          }
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, p.tmem_cols);
}

// ---------------------------------------------------------------------------------------------
// host: launch
// ---------------------------------------------------------------------------------------------
static std::once_flag g_attr_once;

int launch(const Params& p, dim3 grid, size_t smem_bytes, cudaStream_t stream) {
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(g_attr_once, [] {
    attr_err = cudaFuncSetAttribute(tc_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    227 * 1024 - 256);
  });
  if (attr_err != cudaSuccess) return gabril_set_cuda_error(attr_err);
  tc_gemm_kernel<<<grid, kThreads, smem_bytes, stream>>>(p);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

}  // namespace tc
