// tcgen05 / TMEM / TMA implicit-GEMM kernel and its host-side planners.
// See tc_gemm.cuh for the operand regimes.  Reference ops replaced:
//   nn.Conv2d forward/backward in Encoder (vlm_gaze/models/linear_models.py:141-207),
//   Residual 3x3 / 1x1 convs (linear_models.py:78-102), pre_actor Linear (train_bc.py:79).
#include "tc_gemm.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "common.cuh"
#include "ptx.cuh"
#include "tc_epi.cuh"

namespace tc {

// ---------------------------------------------------------------------------------------------
// device
// ---------------------------------------------------------------------------------------------
__device__ unsigned long long g_prof[16];

__device__ __forceinline__ void report_timeout(const Params& p, int code) {
  if (p.err_flag) atomicExch(p.err_flag, code);
}

struct Tile {
  int cls, m0, n0, tap0, ntaps, kb_begin, kb_end, split;
  int mi, y;   // m tile index over the concatenated classes; n tile [x tap group] index
};

__device__ __forceinline__ int fdiv(int n, const FastDiv& d) {
  return (int)((__umulhi((uint32_t)n, d.mul) + (uint32_t)n) >> d.shr);
}

// tile id -> (m tile over the concatenated classes, n tile [x tap group], split)
__device__ __forceinline__ Tile decode_tile(const Params& p, int t) {
  Tile r;
  if (p.simple_tiles) {
    // one class, one column tile, no split (most forward / data-gradient launches): tile = m tile.  The
    // general decode below is ~80 instructions, paid per tile by the producers, the issuer and every epilogue
    // warp - all single warps whose instruction count bounds short-K layers.
    r.mi = t; r.y = 0; r.cls = 0; r.m0 = t * 128 * p.m_sub; r.n0 = 0; r.tap0 = 0; r.ntaps = 1; r.split = 0;
    r.kb_begin = 0; r.kb_end = p.kb_total;
    return r;
  }
  const int rest = fdiv(t, p.fd_tiles_m);
  int mi = t - rest * p.tiles_m_total;
  const int split = fdiv(rest, p.fd_n_tiles);
  const int y = rest - split * p.n_tiles;
  r.mi = mi;
  r.y = y;
  int cls = 0;
  if (p.cls_min_tiles > 0) {
    // parity classes interleaved (tile 4i + c = tile i of class c), so that the CTAs working at any moment
    // read the same region of dY through all four class maps: class-major order re-read dY from DRAM once
    // per class whenever it is larger than L2 (conv_3 dgrad: 997 MB read for a 231 MB tensor).  The
    // classes' unequal tails follow class by class.
    const int body = p.cls_min_tiles * p.n_classes;
    if (mi < body) {
      const int i = fdiv(mi, p.fd_n_classes);
      cls = mi - i * p.n_classes;
      mi = i;
    } else {
      mi -= body;
      while (cls + 1 < p.n_classes && mi >= p.tiles_m[cls] - p.cls_min_tiles) { mi -= p.tiles_m[cls] - p.cls_min_tiles; ++cls; }
      mi += p.cls_min_tiles;
    }
  } else {
    while (cls + 1 < p.n_classes && mi >= p.tiles_m[cls]) { mi -= p.tiles_m[cls]; ++cls; }
  }
  r.cls = cls;
  r.m0 = mi * 128 * p.m_sub;
  int n_tile = y, tg = 0;
  if (p.kind == KIND_WGRAD) {
    n_tile = fdiv(y, p.fd_tap_groups);
    tg = y - n_tile * p.tap_groups;
  }
  r.n0 = n_tile * p.block_n;
  r.tap0 = tg * p.n_acc;
  r.ntaps = (p.kind == KIND_WGRAD) ? min(p.n_acc, p.taps - r.tap0) : 1;
  r.split = split;
  r.kb_begin = split * p.kb_per_split;
  r.kb_end = min(r.kb_begin + p.kb_per_split, p.kb_total);
  return r;
}

// Persistent CTA: tiles blockIdx.x, blockIdx.x + gridDim.x, ...  The smem stage ring runs across
// tile boundaries and the accumulator lives in a ring of acc_bufs TMEM buffers, so the producer
// and the MMA issuer work on tile i+1 while the epilogue warps drain tile i.
// kEpiGroups = 1: 256 threads, two CTAs per SM (each other's epilogue hides behind the other's MMAs);
// kEpiGroups = 2: 384 threads for the single-occupancy launches (flat-slab layers with resident weights,
// 256-row tiles, 512-column wgrad accumulators), where nothing else covers the epilogue's latencies.
// kEF: the epilogue this instantiation is compiled for: one EPI_VARIANTS mask (bf16 output, straight-line
// code), EF_F32 (fp32 stores / atomics only: wgrad, split-K), or EF_RUNTIME (everything, runtime flags).
template <int kEpiGroups, uint32_t kEF>
__global__ void __launch_bounds__(32 * (kFirstEpilogueWarp + 4 * kEpiGroups), kEpiGroups == 1 ? 2 : 1)
tc_gemm_kernel(const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * kMaxStages + 2 * kMaxAcc + 1];
  __shared__ uint32_t tmem_base_s;
  // bias vector of the layer: read by every epilogue item; from global memory each read was an exposed
  // L1 / L2 round trip on a warp that has nothing else to issue (ncu: a quarter of the epilogue's samples)
  __shared__ __align__(16) float bias_s[kBiasSmem];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t full0 = ptx::smem_u32(&bars[0]);
  const uint32_t empty0 = ptx::smem_u32(&bars[kMaxStages]);
  const uint32_t tfull0 = ptx::smem_u32(&bars[2 * kMaxStages]);
  const uint32_t tempty0 = ptx::smem_u32(&bars[2 * kMaxStages + kMaxAcc]);
  const uint32_t wfull = ptx::smem_u32(&bars[2 * kMaxStages + 2 * kMaxAcc]);  // resident weights landed
  const uint32_t stage_bytes = p.a_stage_bytes + p.b_stage_bytes;

  // fused bias gradient (WGRAD, one tile per CTA): this CTA's epilogue warps also read the dY tile of every stage
  bool dbias_cta = p.dbias != nullptr;
  uint32_t dy_center_off = 0;
  if (dbias_cta && p.shift_dy) {
    // shifted-dY form: only the CTAs whose tap group holds the centre tap stage dY itself
    const Tile t0 = decode_tile(p, blockIdx.x);
    const int ci = p.center_tap - t0.tap0;
    dbias_cta = ci >= 0 && ci < t0.ntaps;
    dy_center_off = (uint32_t)ci * p.b_box_bytes;
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      ptx::mbar_init(full0 + 8 * s, 1);
      // a stage is free once its MMAs have completed (tcgen05.commit) and, with the fused bias gradient, once
      // every epilogue warp has read (or passed on) the dY tile
      ptx::mbar_init(empty0 + 8 * s, dbias_cta ? 1 + 4 * kEpiGroups : 1);
    }
    for (int b = 0; b < p.acc_bufs; ++b) {
      ptx::mbar_init(tfull0 + 8 * b, 1);
      // one arrival per epilogue warp that drains the buffer: every group, or only the tile's owner
      ptx::mbar_init(tempty0 + 8 * b, p.epi_tile_split ? 4 : 4 * kEpiGroups);
    }
    ptx::mbar_init(wfull, 1);
    ptx::fence_mbar_init();
  }
  // Programmatic dependent launch: let the next kernel of the stream be scheduled while this one runs (its
  // CTAs become resident as ours retire and wait below), and wait for the previous kernel's results before
  // the first global-memory access.  Launch latency and this prologue overlap the predecessor's tail.
  ptx::pdl_launch_dependents();
  if (warp == kMmaWarp) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_base_s), p.tmem_cols);
    ptx::tmem_relinquish();
  }
  ptx::pdl_wait();
  const bool bias_sm = p.bias != nullptr && p.N_total <= kBiasSmem;
  if (warp >= kFirstEpilogueWarp && bias_sm)
    for (int i = threadIdx.x - 32 * kFirstEpilogueWarp; i < p.N_total; i += 128 * kEpiGroups)
      bias_s[i] = p.bias[p.bias_mod > 0 ? i % p.bias_mod : i];
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp >= kFirstProducerWarp && warp < kFirstEpilogueWarp) {
    // ===================== TMA producers (one thread per warp) =====================
    // Issuing a tensor-map copy costs the issuing thread hundreds of cycles, so the smem stages
    // are dealt round-robin to kProducers threads in different warps.  A stage is always served
    // by the same thread (in iteration order), which keeps the parity waits unambiguous.
    if (ptx::elect_one()) {
      const int me = warp - kFirstProducerWarp;
      ptx::tma_prefetch_desc(&p.mapA[0]);
      ptx::tma_prefetch_desc(&p.mapB);
      int st = 0;
      uint32_t st_ph = 0;
      bool ok = true;
      if (p.a_flat && p.kind == KIND_FWD && p.b_resident && me == 0) {
        // every weight tile of the layer, once per CTA, in traversal order (r, s, chunk)
        const int chunks = p.chunks_per_tap;
        ptx::mbar_arrive_expect_tx(wfull, (uint32_t)(p.flat_R * p.flat_S * chunks) * p.b_tap_bytes);
        for (int r = 0; r < p.flat_R; ++r)
          for (int sx = 0; sx < p.flat_S; ++sx) {
            const int wtap = (p.r_base[0] + p.r_step * r) * p.S_w + p.s_base[0] + p.s_step * sx;
            for (int ch = 0; ch < chunks; ++ch) {
              const uint32_t dst =
                  smem_base + p.resident_off + (uint32_t)((r * p.flat_S + sx) * chunks + ch) * p.b_tap_bytes;
              if (!p.b_mn) {
                ptx::tma_load_3d(dst, &p.mapB, wfull, ch * p.kc, wtap, 0);
              } else {
                for (int i = 0; i < p.b_boxes; ++i)
                  ptx::tma_load_3d(dst + i * p.b_box_bytes, &p.mapB, wfull, i * p.b_box_w, wtap, ch * p.kc);
              }
            }
          }
      }
      for (int t = blockIdx.x; t < p.total_tiles && ok; t += gridDim.x) {
        const Tile tl = decode_tile(p, t);
        const int cls = tl.cls, m0 = tl.m0, n0 = tl.n0;
        const int ohw = p.OHW[cls], oww = p.OW[cls];
        if (p.a_flat && p.kind == KIND_FWD) {
          const int chunks = p.chunks_per_tap;
          const uint32_t tx = p.slab_tx_bytes + (p.b_resident ? 0u : (uint32_t)p.flat_S * p.b_tap_bytes);
          int r = tl.kb_begin / chunks, chunk = tl.kb_begin - (tl.kb_begin / chunks) * chunks;
          for (int kb = tl.kb_begin; kb < tl.kb_end; ++kb) {
            const int s = st;
            const uint32_t ph = st_ph;
            const int chunk_c = chunk, r_c = r;
            if (++st == p.stages) { st = 0; st_ph ^= 1; }
            if (++chunk == chunks) { chunk = 0; ++r; }
            if (s % kProducers != me) continue;
            if (!ptx::mbar_wait(empty0 + 8 * s, ph ^ 1)) { report_timeout(p, 106); ok = false; break; }
            const uint32_t a_dst = smem_base + s * stage_bytes;
            const uint32_t b_dst = a_dst + p.a_stage_bytes;
            const uint32_t full = full0 + 8 * s;
            if (p.debug & 1) { ptx::mbar_arrive(full); continue; }
            ptx::mbar_arrive_expect_tx(full, tx);
            ptx::tma_load_2d(a_dst, &p.mapA[0], full, chunk_c * p.kc, m0 + r_c * p.flat_pitch);
            if (!p.b_resident) {
              for (int sx = 0; sx < p.flat_S; ++sx) {
                const int wtap = (p.r_base[0] + p.r_step * r_c) * p.S_w + p.s_base[0] + p.s_step * sx;
                const uint32_t dst = b_dst + sx * p.b_tap_bytes;
                if (!p.b_mn) {
                  ptx::tma_load_3d(dst, &p.mapB, full, chunk_c * p.kc, wtap, n0);
                } else {
                  for (int i = 0; i < p.b_boxes; ++i)
                    ptx::tma_load_3d(dst + i * p.b_box_bytes, &p.mapB, full, n0 + i * p.b_box_w, wtap,
                                     chunk_c * p.kc);
                }
              }
            }
          }
        } else if (p.a_flat) {
          // WGRAD flat: A = dY rows of the k-block, B = one pixel slab per filter row of the group
          const int r0 = tl.tap0 / p.flat_S, nrows = tl.ntaps / p.flat_S;
          const uint32_t tx = p.a_boxes * p.a_box_bytes + (uint32_t)(nrows * p.b_boxes) * p.slab_tx_bytes;
          for (int kb = tl.kb_begin; kb < tl.kb_end; ++kb) {
            const int s = st;
            const uint32_t ph = st_ph;
            if (++st == p.stages) { st = 0; st_ph ^= 1; }
            if (s % kProducers != me) continue;
            if (!ptx::mbar_wait(empty0 + 8 * s, ph ^ 1)) { report_timeout(p, 107); ok = false; break; }
            const uint32_t a_dst = smem_base + s * stage_bytes;
            const uint32_t b_dst = a_dst + p.a_stage_bytes;
            const uint32_t full = full0 + 8 * s;
            ptx::mbar_arrive_expect_tx(full, tx);
            const int p0 = kb * p.kc;
            for (int i = 0; i < p.a_boxes; ++i)
              ptx::tma_load_2d(a_dst + i * p.a_box_bytes, &p.mapA[0], full, m0 + i * p.a_box_w,
                               p0 + p.dy_row_off);
            for (int rr = 0; rr < nrows; ++rr)
              for (int i = 0; i < p.b_boxes; ++i)
                ptx::tma_load_2d(b_dst + (uint32_t)(rr * p.b_boxes + i) * p.b_tap_bytes, &p.mapB, full,
                                 n0 + i * p.b_box_w, p0 + (r0 + rr) * p.flat_pitch);
          }
        } else if (p.kind == KIND_FWD) {
          int w_base = 0, h_base = 0, n_img = 0;
          if (p.a_im2col) {
            n_img = fdiv(m0, p.fd_ohw[cls]);
            const int rem = m0 - n_img * ohw;
            const int oh = fdiv(rem, p.fd_ow[cls]);
            const int ow = rem - oh * oww;
            w_base = ow * p.trav_stride_w + p.lower_w[cls];
            h_base = oh * p.trav_stride_h + p.lower_h[cls];
          }
          const uint32_t tx = p.a_boxes * p.a_box_bytes + p.b_boxes * p.b_box_bytes;
          int tap = tl.kb_begin / p.chunks_per_tap;
          int chunk = tl.kb_begin - tap * p.chunks_per_tap;
          int rp = tap / p.S, sp = tap - rp * p.S;
          // loop bookkeeping in opaque registers (two of three iterations only skip: keep them short)
          int n_st = p.stages, cpt = p.chunks_per_tap, fS = p.S;
          asm volatile("" : "+r"(n_st), "+r"(cpt), "+r"(fS));
          const int kb_end = tl.kb_end;
          for (int kb = tl.kb_begin; kb < kb_end; ++kb) {
            const int s = st;
            const uint32_t ph = st_ph;
            const int chunk_c = chunk, rp_c = rp, sp_c = sp;
            if (++st == n_st) { st = 0; st_ph ^= 1; }
            if (++chunk == cpt) { chunk = 0; if (++sp == fS) { sp = 0; ++rp; } }
            if (s % kProducers != me) continue;  // a stage always belongs to the same producer
            if (!ptx::mbar_wait(empty0 + 8 * s, ph ^ 1)) { report_timeout(p, 101); ok = false; break; }
            const uint32_t a_dst = smem_base + s * stage_bytes;
            const uint32_t b_dst = a_dst + p.a_stage_bytes;
            const uint32_t full = full0 + 8 * s;
            if (p.debug & 1) { ptx::mbar_arrive(full); continue; }
            // feed experiments: bit 6 = no weight tile, bit 7 = no activation tile (the MMAs then read stale data)
            ptx::mbar_arrive_expect_tx(full, tx - ((p.debug & 64) ? p.b_boxes * p.b_box_bytes : 0u) -
                                                 ((p.debug & 128) ? p.a_boxes * p.a_box_bytes : 0u));
            if (p.debug & 128) {
            } else if (p.a_im2col)
              ptx::tma_load_im2col_4d(a_dst, &p.mapA[cls], full, chunk_c * p.kc, w_base, h_base,
                                      n_img, (uint16_t)sp_c, (uint16_t)rp_c);
            else
              ptx::tma_load_2d(a_dst, &p.mapA[0], full, kb * p.kc, m0);
            const int wtap =
                (p.r_base[cls] + p.r_step * rp_c) * p.S_w + p.s_base[cls] + p.s_step * sp_c;
            if (p.debug & 64) {
            } else if (p.cls_n) {
              // stacked parity classes: box i = the weight tap of class i, all Cin columns
              for (int i = 0; i < p.b_boxes; ++i)
                ptx::tma_load_3d(b_dst + i * p.b_box_bytes, &p.mapB, full, 0,
                                 (p.r_base[i] + p.r_step * rp_c) * p.S_w + p.s_base[i] + p.s_step * sp_c,
                                 chunk_c * p.kc);
            } else if (!p.b_mn) {
              ptx::tma_load_3d(b_dst, &p.mapB, full, chunk_c * p.kc, wtap, n0);
            } else {
              for (int i = 0; i < p.b_boxes; ++i)
                ptx::tma_load_3d(b_dst + i * p.b_box_bytes, &p.mapB, full, n0 + i * p.b_box_w,
                                 wtap, chunk_c * p.kc);
            }
          }
        } else if (p.shift_dy) {
          // WGRAD, operands exchanged, shift on dY: A = the activation rows of the k-block (shared by every
          // tap), B = one im2col tile of dY per tap, walked like a data gradient (tap (r, s) reads dY at
          // pixel + (pad - r, pad - s) = traversal offset (R-1-r, S-1-s) from the lower corner)
          const uint32_t tx = p.a_boxes * p.a_box_bytes + (uint32_t)tl.ntaps * p.b_box_bytes;
          const int Rf = p.taps / p.S;
          for (int kb = tl.kb_begin; kb < tl.kb_end; ++kb) {
            const int s = st;
            const uint32_t ph = st_ph;
            if (++st == p.stages) { st = 0; st_ph ^= 1; }
            if (s % kProducers != me) continue;
            if (!ptx::mbar_wait(empty0 + 8 * s, ph ^ 1)) { report_timeout(p, 111); ok = false; break; }
            const uint32_t a_dst = smem_base + s * stage_bytes;
            const uint32_t b_dst = a_dst + p.a_stage_bytes;
            const uint32_t full = full0 + 8 * s;
            ptx::mbar_arrive_expect_tx(full, tx);
            const int p0 = kb * p.kc;
            for (int i = 0; i < p.a_boxes; ++i)
              ptx::tma_load_2d(a_dst + i * p.a_box_bytes, &p.mapA[0], full, m0 + i * p.a_box_w, p0);
            const int n_img = fdiv(p0, p.fd_ohw[0]);
            const int rem = p0 - n_img * ohw;
            const int ih = fdiv(rem, p.fd_ow[0]);
            const int iw = rem - ih * oww;
            const int w_base = iw + p.lower_w[0], h_base = ih + p.lower_h[0];
            for (int tt = 0; tt < tl.ntaps; ++tt) {
              const int tap = tl.tap0 + tt;
              const int r = tap / p.S, sx = tap - r * p.S;
              ptx::tma_load_im2col_4d(b_dst + (uint32_t)tt * p.b_box_bytes, &p.mapB, full, n0, w_base, h_base,
                                      n_img, (uint16_t)(p.S - 1 - sx), (uint16_t)(Rf - 1 - r));
            }
          }
        } else if (p.swap_ab) {
          // WGRAD, operands exchanged: A = one im2col activation tile per tap, B = the dY rows
          const uint32_t tx = tl.ntaps * p.a_boxes * p.a_box_bytes + p.b_boxes * p.b_box_bytes;
          for (int kb = tl.kb_begin; kb < tl.kb_end; ++kb) {
            const int s = st;
            const uint32_t ph = st_ph;
            if (++st == p.stages) { st = 0; st_ph ^= 1; }
            if (s % kProducers != me) continue;
            if (!ptx::mbar_wait(empty0 + 8 * s, ph ^ 1)) { report_timeout(p, 109); ok = false; break; }
            const uint32_t a_dst = smem_base + s * stage_bytes;
            const uint32_t b_dst = a_dst + p.a_stage_bytes;
            const uint32_t full = full0 + 8 * s;
            ptx::mbar_arrive_expect_tx(full, tx);
            const int p0 = kb * p.kc;
            for (int i = 0; i < p.b_boxes; ++i)
              ptx::tma_load_2d(b_dst + i * p.b_box_bytes, &p.mapB, full, n0 + i * p.b_box_w, p0);
            if (p.a_im2col) {
              const int n_img = fdiv(p0, p.fd_ohw[0]);
              const int rem = p0 - n_img * ohw;
              const int oh = fdiv(rem, p.fd_ow[0]);
              const int ow = rem - oh * oww;
              const int w_base = ow * p.trav_stride_w + p.lower_w[0];
              const int h_base = oh * p.trav_stride_h + p.lower_h[0];
              for (int tt = 0; tt < tl.ntaps; ++tt) {
                const int tap = tl.tap0 + tt;
                const int rp = tap / p.S, sp = tap - rp * p.S;
                for (int i = 0; i < p.a_boxes; ++i)
                  ptx::tma_load_im2col_4d(a_dst + (tt * p.a_boxes + i) * p.a_box_bytes, &p.mapA[0],
                                          full, m0 + i * p.a_box_w, w_base, h_base, n_img,
                                          (uint16_t)sp, (uint16_t)rp);
              }
            } else {
              for (int i = 0; i < p.a_boxes; ++i)
                ptx::tma_load_2d(a_dst + i * p.a_box_bytes, &p.mapA[0], full, m0 + i * p.a_box_w, p0);
            }
          }
        } else {
          const uint32_t tx = p.a_boxes * p.a_box_bytes + tl.ntaps * p.b_boxes * p.b_box_bytes;
          for (int kb = tl.kb_begin; kb < tl.kb_end; ++kb) {
            const int s = st;
            const uint32_t ph = st_ph;
            if (++st == p.stages) { st = 0; st_ph ^= 1; }
            if (s % kProducers != me) continue;  // a stage always belongs to the same producer
            if (!ptx::mbar_wait(empty0 + 8 * s, ph ^ 1)) { report_timeout(p, 102); ok = false; break; }
            const uint32_t a_dst = smem_base + s * stage_bytes;
            const uint32_t b_dst = a_dst + p.a_stage_bytes;
            const uint32_t full = full0 + 8 * s;
            ptx::mbar_arrive_expect_tx(full, tx);
            const int p0 = kb * p.kc;
            for (int i = 0; i < p.a_boxes; ++i)
              ptx::tma_load_2d(a_dst + i * p.a_box_bytes, &p.mapA[0], full, m0 + i * p.a_box_w, p0);
            if (p.a_im2col) {
              const int n_img = fdiv(p0, p.fd_ohw[0]);
              const int rem = p0 - n_img * ohw;
              const int oh = fdiv(rem, p.fd_ow[0]);
              const int ow = rem - oh * oww;
              const int w_base = ow * p.trav_stride_w + p.lower_w[0];
              const int h_base = oh * p.trav_stride_h + p.lower_h[0];
              for (int tt = 0; tt < tl.ntaps; ++tt) {
                const int tap = tl.tap0 + tt;
                const int rp = tap / p.S, sp = tap - rp * p.S;
                for (int i = 0; i < p.b_boxes; ++i)
                  ptx::tma_load_im2col_4d(b_dst + (tt * p.b_boxes + i) * p.b_box_bytes, &p.mapB,
                                          full, n0 + i * p.b_box_w, w_base, h_base, n_img,
                                          (uint16_t)sp, (uint16_t)rp);
              }
            } else {
              for (int i = 0; i < p.b_boxes; ++i)
                ptx::tma_load_2d(b_dst + i * p.b_box_bytes, &p.mapB, full, n0 + i * p.b_box_w, p0);
            }
          }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ===================== MMA issuer (one thread) =====================
    if (ptx::elect_one()) {
      int it = 0, tcount = 0;
      bool ok = true;
      int mma_k = p.mma_k, mma_groups = p.mma_groups, g_taps = p.g_taps;
      uint32_t a_kstep = p.a_kstep, b_kstep = p.b_kstep, b_gstep = p.b_gstep, d_gstep = p.d_gstep;
      uint32_t a_gstep = p.a_gstep;  // swapped wgrad: per tap; FWD with m_sub = 2: per 128-row sub-tile
      uint32_t a_hi32 = (uint32_t)(p.desc_a_hi >> 32), b_hi32 = (uint32_t)(p.desc_b_hi >> 32);
      uint32_t stage16 = stage_bytes >> 4;
      uint32_t a_lo_const = (uint32_t)p.desc_a_hi + (smem_base >> 4);
      uint32_t b_lo_const = (uint32_t)p.desc_b_hi + ((smem_base + p.a_stage_bytes) >> 4);
      uint32_t idesc = p.idesc;
      int stages = p.stages;
      // opaque register copies: the compiler otherwise re-loads these from the constant bank in every k-block
      asm volatile("" : "+r"(mma_k), "+r"(mma_groups), "+r"(g_taps), "+r"(a_kstep), "+r"(b_kstep), "+r"(b_gstep),
                        "+r"(d_gstep), "+r"(a_gstep), "+r"(a_hi32), "+r"(b_hi32), "+r"(stage16), "+r"(a_lo_const),
                        "+r"(b_lo_const), "+r"(idesc), "+r"(stages));
      const bool dbg_skip = (p.debug & 2) != 0;
      int s = 0;
      uint32_t ph = 0;
      long long pw = 0, pi = 0, pc = 0, pt = 0, pn = 0;
      const bool prof = (p.debug & 4) != 0;
      // lean loop state (see `fast` below): descriptor low words / barrier address of the current stage
      const bool fast = !p.a_flat && (mma_k == 4 || mma_k == 2) && (p.debug & ~(3 | 64 | 128)) == 0;
      const bool k4 = mma_k == 4;   // kc = 64: four K-steps per k-block; kc = 32: two
      uint32_t a_cur = a_lo_const, b_cur = b_lo_const, full_cur = full0;
      uint32_t a_k2 = 2 * a_kstep, a_k3 = 3 * a_kstep, b_k2 = 2 * b_kstep, b_k3 = 3 * b_kstep;
      // opaque copies: the compiler otherwise re-materialises these from the constant bank in every iteration
      asm volatile("" : "+r"(a_k2), "+r"(a_k3), "+r"(b_k2), "+r"(b_k3));
      // flat-slab forward regime: lean loop state
      const bool fast_flat = p.a_flat && p.kind == KIND_FWD && mma_k == 4 && (p.debug & ~3) == 0 &&
                             p.kb_per_split == p.kb_total;
      const bool fast_flatw = p.a_flat && p.kind == KIND_WGRAD && mma_k == 4 && p.debug == 0;
      int ff_kb = p.kb_total, ff_S = p.flat_S, ff_chunks = p.chunks_per_tap, ff_res = p.b_resident;
      uint32_t ff_tap16 = p.b_tap_bytes >> 4;
      uint32_t ff_sxb = ff_res ? (uint32_t)ff_chunks * ff_tap16 : ff_tap16;
      uint32_t ff_rowskip = (uint32_t)((ff_S - 1) * ff_chunks) * ff_tap16;
      asm volatile("" : "+r"(ff_kb), "+r"(ff_S), "+r"(ff_chunks), "+r"(ff_res), "+r"(ff_tap16), "+r"(ff_sxb),
                        "+r"(ff_rowskip));
      for (int t = blockIdx.x; t < p.total_tiles && ok; t += gridDim.x, ++tcount) {
        const long long c0 = prof ? clock64() : 0;
        const Tile tl = decode_tile(p, t);
        const int buf = tcount & (p.acc_bufs - 1);
        const int use = tcount >> p.acc_shift;
        // the epilogue must have drained this accumulator buffer
        if (!ptx::mbar_wait(tempty0 + 8 * buf, (use & 1) ^ 1)) { report_timeout(p, 105); break; }
        ptx::tc_fence_after();
        if (prof) pt += clock64() - c0;
        const uint32_t d_base = tmem_base + buf * p.acc_stride;
        if (p.a_flat) {
          const int chunks = p.chunks_per_tap, FS = p.flat_S;
          const uint32_t tap16 = p.b_tap_bytes >> 4;
          if (p.kind == KIND_FWD) {
            if (p.b_resident && tcount == 0) {
              if (!ptx::mbar_wait(wfull, 0)) { report_timeout(p, 108); break; }
              ptx::tc_fence_after();
            }
            const uint32_t w_lo = (uint32_t)p.desc_b_hi + ((smem_base + p.resident_off) >> 4);
            if (fast_flat) {
              // Lean issue loop of the flat-slab forward regime (a tile always walks all k-blocks from 0): the
              // stage's descriptor words, its barrier and the resident weight tile advance by addition.
              // conv_2 forward with TMA copies and MMAs both stubbed out took 0.395 of its 0.498 ms alone.
              uint32_t wk = w_lo;
              int chunk = 0;
              for (int kb = 0; kb < ff_kb; ++kb) {
                if (!ptx::mbar_wait(full_cur, ph)) { report_timeout(p, 103); ok = false; break; }
                ptx::tc_fence_after();
                uint32_t as = a_cur, bb = ff_res ? wk : b_cur;
                for (int sx = 0; sx < ff_S; ++sx, as += 8u, bb += ff_sxb) {
                  if (dbg_skip) continue;
                  ptx::umma_bf16(d_base, ((uint64_t)a_hi32 << 32) | as, ((uint64_t)b_hi32 << 32) | bb, idesc, (kb > 0 || sx > 0) ? 1u : 0u);
                  ptx::umma_bf16(d_base, ((uint64_t)a_hi32 << 32) | (as + a_kstep), ((uint64_t)b_hi32 << 32) | (bb + b_kstep), idesc, 1u);
                  ptx::umma_bf16(d_base, ((uint64_t)a_hi32 << 32) | (as + a_k2), ((uint64_t)b_hi32 << 32) | (bb + b_k2), idesc, 1u);
                  ptx::umma_bf16(d_base, ((uint64_t)a_hi32 << 32) | (as + a_k3), ((uint64_t)b_hi32 << 32) | (bb + b_k3), idesc, 1u);
                }
                ptx::umma_commit(full_cur + 8 * kMaxStages);
                a_cur += stage16; b_cur += stage16; full_cur += 8;
                if (++s == stages) { s = 0; ph ^= 1; a_cur = a_lo_const; b_cur = b_lo_const; full_cur = full0; }
                // resident weights are stored (filter row, tap column, chunk): next chunk, or next filter row
                wk += ff_tap16;
                if (++chunk == ff_chunks) { chunk = 0; wk += ff_rowskip; }
              }
              if (ok) ptx::umma_commit(tfull0 + 8 * buf);
              continue;
            }
            int r = tl.kb_begin / chunks, chunk = tl.kb_begin - (tl.kb_begin / chunks) * chunks;
            for (int kb = tl.kb_begin; kb < tl.kb_end; ++kb) {
              if (!ptx::mbar_wait(full0 + 8 * s, ph)) { report_timeout(p, 103); ok = false; break; }
              ptx::tc_fence_after();
              const uint32_t a_lo = a_lo_const + (uint32_t)s * stage16;
              const uint32_t b_lo = b_lo_const + (uint32_t)s * stage16;
              for (int sx = 0; sx < FS; ++sx) {
                const uint32_t as = a_lo + sx * 8u;  // view of the slab from pixel row sx on
                const uint32_t bb = p.b_resident ? w_lo + (uint32_t)((r * FS + sx) * chunks + chunk) * tap16
                                                 : b_lo + sx * tap16;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const uint64_t a_desc = ((uint64_t)a_hi32 << 32) | (uint64_t)(as + j * a_kstep);
                  const uint64_t b_desc = ((uint64_t)b_hi32 << 32) | (uint64_t)(bb + j * b_kstep);
                  if (!dbg_skip) ptx::umma_bf16(d_base, a_desc, b_desc, idesc, (kb > tl.kb_begin || sx > 0 || j > 0) ? 1u : 0u);
                }
              }
              ptx::umma_commit(empty0 + 8 * s);
              if (++s == stages) { s = 0; ph ^= 1; }
              if (++chunk == chunks) { chunk = 0; ++r; }
            }
          } else if (fast_flatw) {
            // lean issue loop of the flat weight gradient (one tile per CTA, long k loop)
            const int nrows = tl.ntaps / FS;
            uint32_t row16 = (uint32_t)p.b_boxes * tap16, bn = (uint32_t)p.block_n;
            int swp = p.swap_ab;
            asm volatile("" : "+r"(row16), "+r"(bn), "+r"(swp));
            const int kb_end = tl.kb_end;
            for (int kb = tl.kb_begin; kb < kb_end; ++kb) {
              if (!ptx::mbar_wait(full_cur, ph)) { report_timeout(p, 103); ok = false; break; }
              ptx::tc_fence_after();
              const uint32_t acc0 = kb > tl.kb_begin ? 1u : 0u;
              uint32_t brow = b_cur, dcol = d_base;
              for (int rr = 0; rr < nrows; ++rr, brow += row16) {
                uint32_t bb = brow;
                for (int sx = 0; sx < ff_S; ++sx, bb += 8u, dcol += bn) {
                  const uint64_t a0 = ((uint64_t)a_hi32 << 32) | a_cur, a1 = ((uint64_t)a_hi32 << 32) | (a_cur + a_kstep);
                  const uint64_t a2 = ((uint64_t)a_hi32 << 32) | (a_cur + a_k2), a3 = ((uint64_t)a_hi32 << 32) | (a_cur + a_k3);
                  const uint64_t b0 = ((uint64_t)b_hi32 << 32) | bb, b1 = ((uint64_t)b_hi32 << 32) | (bb + b_kstep);
                  const uint64_t b2 = ((uint64_t)b_hi32 << 32) | (bb + b_k2), b3 = ((uint64_t)b_hi32 << 32) | (bb + b_k3);
                  if (swp) {   // exchanged form (narrow dY): the slab view is the M = 128 operand, dY the N = Cy one
                    ptx::umma_bf16(dcol, b0, a0, idesc, acc0); ptx::umma_bf16(dcol, b1, a1, idesc, 1u);
                    ptx::umma_bf16(dcol, b2, a2, idesc, 1u); ptx::umma_bf16(dcol, b3, a3, idesc, 1u);
                  } else {
                    ptx::umma_bf16(dcol, a0, b0, idesc, acc0); ptx::umma_bf16(dcol, a1, b1, idesc, 1u);
                    ptx::umma_bf16(dcol, a2, b2, idesc, 1u); ptx::umma_bf16(dcol, a3, b3, idesc, 1u);
                  }
                }
              }
              ptx::umma_commit(full_cur + 8 * kMaxStages);
              a_cur += stage16; b_cur += stage16; full_cur += 8;
              if (++s == stages) { s = 0; ph ^= 1; a_cur = a_lo_const; b_cur = b_lo_const; full_cur = full0; }
            }
          } else {
            const int nrows = tl.ntaps / FS;
            for (int kb = tl.kb_begin; kb < tl.kb_end; ++kb) {
              if (!ptx::mbar_wait(full0 + 8 * s, ph)) { report_timeout(p, 103); ok = false; break; }
              ptx::tc_fence_after();
              const uint32_t a_lo = a_lo_const + (uint32_t)s * stage16;
              const uint32_t b_lo = b_lo_const + (uint32_t)s * stage16;
              for (int rr = 0; rr < nrows; ++rr)
                for (int sx = 0; sx < FS; ++sx) {
                  const uint32_t bb = b_lo + (uint32_t)(rr * p.b_boxes) * tap16 + sx * 8u;
                  const uint32_t dcol = d_base + (uint32_t)((rr * FS + sx) * p.block_n);
#pragma unroll
                  for (int j = 0; j < 4; ++j) {
                    const uint64_t a_desc = ((uint64_t)a_hi32 << 32) | (uint64_t)(a_lo + j * a_kstep);
                    const uint64_t b_desc = ((uint64_t)b_hi32 << 32) | (uint64_t)(bb + j * b_kstep);
                    // exchanged form (narrow dY): the slab view is the M = 128 operand, dY the N = Cy one
                    if (p.swap_ab) ptx::umma_bf16(dcol, b_desc, a_desc, idesc, (kb > tl.kb_begin || j > 0) ? 1u : 0u);
                    else ptx::umma_bf16(dcol, a_desc, b_desc, idesc, (kb > tl.kb_begin || j > 0) ? 1u : 0u);
                  }
                }
              ptx::umma_commit(empty0 + 8 * s);
              if (++s == stages) { s = 0; ph ^= 1; }
            }
          }
          if (ok) ptx::umma_commit(tfull0 + 8 * buf);
          continue;
        }
        if (fast) {
          // Lean issue loop (kc = 64: four K-steps per k-block; no debug / profiling branches).  This thread's
          // INSTRUCTION COUNT per k-block bounds every narrow (N <= 64) or short-K layer: with the TMA copies
          // and the MMAs both stubbed out conv_6 forward still took 0.32 of its 0.38 ms (~580 clk per k-block
          // for ~140 instructions, most of them parameter re-loads from the constant bank).  Here the stage's
          // descriptor words advance by addition and every parameter lives in a register.
          int ng = mma_groups;
          uint32_t idesc_last = idesc;
          if (g_taps > 0) {
            ng = min(mma_groups, (tl.ntaps + g_taps - 1) / g_taps);
            if (p.shift_dy) {
              const int nt = tl.ntaps - (ng - 1) * g_taps;   // taps of the last group: N = nt x Cout
              idesc_last = (idesc & ~(0x3Fu << 17)) | ((uint32_t)((nt * p.block_n) >> 3) << 17);
            }
          }
          const int kb_end = tl.kb_end;
          for (int kb = tl.kb_begin; kb < kb_end; ++kb) {
            if (!ptx::mbar_wait(full_cur, ph)) { report_timeout(p, 103); ok = false; break; }
            ptx::tc_fence_after();
            const uint32_t acc0 = kb > tl.kb_begin ? 1u : 0u;
            if (dbg_skip) {
            } else if (ng == 1) {
              ptx::umma_bf16(d_base, ((uint64_t)a_hi32 << 32) | a_cur, ((uint64_t)b_hi32 << 32) | b_cur, idesc_last, acc0);
              ptx::umma_bf16(d_base, ((uint64_t)a_hi32 << 32) | (a_cur + a_kstep), ((uint64_t)b_hi32 << 32) | (b_cur + b_kstep), idesc_last, 1u);
              if (k4) {
                ptx::umma_bf16(d_base, ((uint64_t)a_hi32 << 32) | (a_cur + a_k2), ((uint64_t)b_hi32 << 32) | (b_cur + b_k2), idesc_last, 1u);
                ptx::umma_bf16(d_base, ((uint64_t)a_hi32 << 32) | (a_cur + a_k3), ((uint64_t)b_hi32 << 32) | (b_cur + b_k3), idesc_last, 1u);
              }
            } else {
              uint32_t ag = a_cur, bg = b_cur, dg = d_base;
              for (int g = 0; g < ng; ++g, ag += a_gstep, bg += b_gstep, dg += d_gstep) {
                const uint32_t id = g == ng - 1 ? idesc_last : idesc;
                ptx::umma_bf16(dg, ((uint64_t)a_hi32 << 32) | ag, ((uint64_t)b_hi32 << 32) | bg, id, acc0);
                ptx::umma_bf16(dg, ((uint64_t)a_hi32 << 32) | (ag + a_kstep), ((uint64_t)b_hi32 << 32) | (bg + b_kstep), id, 1u);
                if (k4) {
                  ptx::umma_bf16(dg, ((uint64_t)a_hi32 << 32) | (ag + a_k2), ((uint64_t)b_hi32 << 32) | (bg + b_k2), id, 1u);
                  ptx::umma_bf16(dg, ((uint64_t)a_hi32 << 32) | (ag + a_k3), ((uint64_t)b_hi32 << 32) | (bg + b_k3), id, 1u);
                }
              }
            }
            ptx::umma_commit(full_cur + 8 * kMaxStages);   // the stage's `empty` barrier
            a_cur += stage16; b_cur += stage16; full_cur += 8;
            if (++s == stages) { s = 0; ph ^= 1; a_cur = a_lo_const; b_cur = b_lo_const; full_cur = full0; }
          }
          if (ok) ptx::umma_commit(tfull0 + 8 * buf);
          continue;
        }
        for (int kb = tl.kb_begin; kb < tl.kb_end; ++kb, ++it) {
          const long long c1 = prof ? clock64() : 0;
          if (!ptx::mbar_wait(full0 + 8 * s, ph)) { report_timeout(p, 103); ok = false; break; }
          ptx::tc_fence_after();
          const long long c2 = prof ? clock64() : 0;
          // descriptors differ only in their low word (14-bit address field; smem addresses >> 4
          // stay below 2^14, so plain adds never carry into the LBO field)
          const uint32_t a_lo = a_lo_const + (uint32_t)s * stage16;
          const uint32_t b_lo = b_lo_const + (uint32_t)s * stage16;
          for (int g = 0; g < mma_groups; ++g) {
            if (g * g_taps >= tl.ntaps) break;
            const uint32_t bg = b_lo + g * b_gstep, dg = d_base + g * d_gstep, ag = a_lo + g * a_gstep;
            uint32_t idesc_g = idesc;
            if (p.shift_dy) {
              // N of this group = its taps x Cout (bits 17..22 of the instruction descriptor hold N >> 3)
              const int nt = min(g_taps, tl.ntaps - g * g_taps);
              idesc_g = (idesc & ~(0x3Fu << 17)) | ((uint32_t)((nt * p.block_n) >> 3) << 17);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (j < mma_k) {
                const uint64_t a_desc = ((uint64_t)a_hi32 << 32) | (uint64_t)(ag + j * a_kstep);
                const uint64_t b_desc = ((uint64_t)b_hi32 << 32) | (uint64_t)(bg + j * b_kstep);
                const uint32_t accumulate = (kb > tl.kb_begin || j > 0) ? 1u : 0u;
                if (!dbg_skip) ptx::umma_bf16(dg, a_desc, b_desc, idesc_g, accumulate);
              }
            }
          }
          const long long c3 = prof ? clock64() : 0;
          ptx::umma_commit(empty0 + 8 * s);
          if (prof) { const long long c4 = clock64(); pw += c2 - c1; pi += c3 - c2; pc += c4 - c3; ++pn; }
          if (++s == stages) { s = 0; ph ^= 1; }
        }
        if (ok) ptx::umma_commit(tfull0 + 8 * buf);
      }
      if (prof) {
        atomicAdd(&g_prof[0], (unsigned long long)pw); atomicAdd(&g_prof[1], (unsigned long long)pi);
        atomicAdd(&g_prof[2], (unsigned long long)pc); atomicAdd(&g_prof[3], (unsigned long long)pt);
        atomicAdd(&g_prof[4], (unsigned long long)pn);
      }
    }
  } else {
    // ===================== epilogue (kEpiGroups x 4 warps, one TMEM lane quarter each) =====================
    // Work item = one 32-row x 32-column chunk of an accumulator.  The warp is alone (or nearly) on its
    // scheduler, so its cost is INSTRUCTIONS (measured: ~2.5 clk per instruction): the bf16 path works
    // on packed pairs (f32x2 adds, bf16x2 max / compare), gates through byte-permute sign replication
    // and moves 32 bytes per access.  With two groups the items of a tile alternate between them (or
    // whole tiles do, when a tile is a single item).
    const int q = warp & 3;
    const int grp = (warp - kFirstEpilogueWarp) >> 2;
    if (dbias_cta) {
      // ---- bias gradient: column sums of the dY tiles of this CTA's split ----
      // Every CTA that shares (m tile, split) - the tap groups and column tiles of the same dY rows - stages
      // the same dY tiles: they deal the k-blocks round-robin, so nobody becomes the straggler of the launch
      // (one reader CTA per group lengthened conv_5 / conv_6 by 20 %).
      // Thread = (16-byte chunk column, pixel phase): it owns 8 consecutive channels and every pstep-th pixel.
      // The eight lanes of a 16-byte-per-lane shared load phase read one pixel's 128 contiguous bytes
      // (permuted by the swizzle XOR): conflict-free.
      const Tile tl = decode_tile(p, blockIdx.x);
      const int etid = threadIdx.x - 32 * kFirstEpilogueWarp, nthr = 128 * kEpiGroups;
      const int nc = p.dy_cols >> 3;                 // chunk columns: 4, 8 or 16
      const int cpb = p.dy_row_bytes >> 4;           // chunks per box row: 4 or 8
      const int col = etid & (nc - 1), prow = etid / nc, pstep = nthr / nc;
      const uint32_t cj = (uint32_t)(col & (cpb - 1));
      const uint32_t tile_off = p.dy_off + dy_center_off + (uint32_t)(col / cpb) * p.dy_box_bytes;
      const bool wide = p.dy_row_bytes == 128;
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (etid < kBiasSmem) bias_s[etid] = 0.f;
      ptx::bar_sync(1, nthr);
      int s = 0;
      uint32_t ph = 0;
      const int share_n = p.shift_dy ? 1 : p.swap_ab ? p.tiles_m_total * p.n_tiles : p.n_tiles;
      int turn = p.shift_dy ? 0 : p.swap_ab ? tl.mi * p.n_tiles + tl.y : tl.y;   // k-blocks until this CTA's next turn
      for (int kb = tl.kb_begin; kb < tl.kb_end; ++kb) {
        if (!ptx::mbar_wait(full0 + 8 * s, ph)) { report_timeout(p, 110); break; }
        const uint32_t st = smem_base + (uint32_t)s * stage_bytes + tile_off;
        const bool mine = turn == 0;
        turn = mine ? share_n - 1 : turn - 1;
#pragma unroll 4
        for (int px = prow; mine && px < p.kc; px += pstep) {
          const uint32_t f = wide ? (uint32_t)(px & 7) : (uint32_t)((px >> 1) & 3);
          const uint4 v = ptx::lds128(st + (uint32_t)px * (uint32_t)p.dy_row_bytes + ((cj ^ f) << 4));
          acc[0] += bf16lo(v.x); acc[1] += bf16hi(v.x); acc[2] += bf16lo(v.y); acc[3] += bf16hi(v.y);
          acc[4] += bf16lo(v.z); acc[5] += bf16hi(v.z); acc[6] += bf16lo(v.w); acc[7] += bf16hi(v.w);
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(empty0 + 8 * s);
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
      for (int off = nc; off < 32; off <<= 1)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], off);
      if (lane < nc)
#pragma unroll
        for (int i = 0; i < 8; ++i) atomicAdd(&bias_s[col * 8 + i], acc[i]);
      ptx::bar_sync(1, nthr);
      if (etid < p.dy_cols && bias_s[etid] != 0.f) {
        const int c = (p.swap_ab ? 0 : tl.m0) + etid;
        ptx::red_add_f32(p.dbias + (p.dbias_mod > 0 ? c % p.dbias_mod : c), bias_s[etid]);
      }
    }
    int tcount = 0;
    const bool eprof = (p.debug & 4) != 0 && warp == kFirstEpilogueWarp && lane == 0;
    long long e_wait = 0, e_tile = 0, e_n = 0, e_pre = 0;
    const int n_chunks = (p.block_n + 31) >> 5;
    const bool out_bf16 = p.out_mode == OUT_BF16;
    const uint32_t epi_flags = p.epi_flags;
    const float* bias_base = bias_sm ? bias_s : p.bias;
    const bf16* resid_base = p.resid;
    const bf16* gate_base = p.gate;
    uint32_t* bits_base = p.bits_out;
    // One UNIT = one 128-row sub-tile of a tile.  The row geometry and the global loads of a unit (sign words of
    // the gate, row mask) are issued one unit AHEAD of its items: issued next to their first use, the sign-word
    // read was an exposed DRAM round trip per tile - 48 % of this warp's samples in res3x3 dgrad (ncu source page).
    struct RowPre {
      bool row_ok;
      long long orow, off0;
      uint32_t keep_row;
      float rm4[4];      // rm4[0]: the row's multiplier; quadrant mode: one per quadrant
      uint32_t gw[8];    // sign words of every 32-column chunk of this row: 4 bytes per chunk instead of 64
      int q_sh, q_sw;    // quadrant mode: super-pixel coordinates of this row
      uint32_t vmask;    // stacked parity classes: bit c = pixel (2jh + ah_c, 2jw + aw_c) exists
    };
    auto prep = [&](const Tile& tl, int sub) {
      RowPre r;
      const int cls = tl.cls, n0 = tl.n0;
      const int row = tl.m0 + sub * 128 + q * 32 + lane;
      bool row_ok = row < p.M_total[cls];
      long long orow = row;
      r.vmask = 0u;
      if (p.cls_n) {
        // row = base pixel (n, jh, jw); orow = pixel (2jh, 2jw) of the output, class c adds ah_c * OWf + aw_c
        const int hw = p.Hs[0] * p.Ws[0];
        const int n = fdiv(row, p.fd_hw[0]);
        const int rem = row - n * hw;
        const int jh = fdiv(rem, p.fd_ws[0]);
        const int jw = rem - jh * p.Ws[0];
        orow = ((long long)n * p.OHf + 2 * jh) * p.OWf + 2 * jw;
        const bool h1 = 2 * jh + 1 < p.OHf, w1 = 2 * jw + 1 < p.OWf;
        r.vmask = row_ok ? (1u | (w1 ? 2u : 0u) | (h1 ? 4u : 0u) | (h1 && w1 ? 8u : 0u)) : 0u;
      } else if (p.remap && row_ok) {
        const int hw = p.Hs[cls] * p.Ws[cls];
        const int n = fdiv(row, p.fd_hw[cls]);
        const int rem = row - n * hw;
        const int jh = fdiv(rem, p.fd_ws[cls]);
        const int jw = rem - jh * p.Ws[cls];
        orow = ((long long)n * p.OHf + (jh * p.remap_sh + p.ah[cls])) * p.OWf +
               (jw * p.remap_sw + p.aw[cls]);
        if (p.valid_h > 0 && (jh >= p.valid_h || jw >= p.valid_w)) row_ok = false;
      }
      bool zero_row = false;
      r.q_sh = 0; r.q_sw = 0;
      if (p.quad_C > 0) {
        const int spr = fdiv(row, p.fd_border_w);
        r.q_sw = row - spr * p.border_w;
        r.q_sh = spr - fdiv(spr, p.fd_border_h) * p.border_h;
      } else if (p.s2d_H > 0 && row_ok) {
        const int qd = row & 3, sp = row >> 2;
        const int spr = fdiv(sp, p.fd_border_w);
        const int sw = sp - spr * p.border_w, sh = spr - fdiv(spr, p.fd_border_h) * p.border_h;
        const int ih = 2 * sh + (qd >> 1) - 1, iw = 2 * sw + (qd & 1) - 1;
        zero_row = ih < 0 || ih >= p.s2d_H || iw < 0 || iw >= p.s2d_W;
      } else if (p.border_h > 0 && row_ok) {
        const int hw = p.border_h * p.border_w;
        const int rem = row - fdiv(row, p.fd_border_hw) * hw;
        const int y = fdiv(rem, p.fd_border_w);
        const int x = rem - y * p.border_w;
        zero_row = (y == 0) || (y == p.border_h - 1) || (x == 0) || (x == p.border_w - 1);
      }
      r.rm4[0] = 1.0f; r.rm4[1] = 1.0f; r.rm4[2] = 1.0f; r.rm4[3] = 1.0f;
      if (p.rowmask && row_ok) {
        if (p.quad_C > 0) {
          const float4 r4 = __ldg(reinterpret_cast<const float4*>(p.rowmask) + orow);
          r.rm4[0] = r4.x; r.rm4[1] = r4.y; r.rm4[2] = r4.z; r.rm4[3] = r4.w;
        } else if (p.pair_C > 0) {   // a row holds two adjacent pixels: one multiplier per half
          const float2 r2 = __ldg(reinterpret_cast<const float2*>(p.rowmask) + orow);
          r.rm4[0] = r2.x; r.rm4[1] = r2.y;
        } else {
          r.rm4[0] = p.rowmask[orow];
        }
      }
      r.keep_row = zero_row ? 0u : 0xffffffffu;
      r.off0 = orow * p.ldo + n0;
#pragma unroll
      for (int ci = 0; ci < 8; ++ci) r.gw[ci] = 0u;
      if (p.out_mode == OUT_BF16 && row_ok && p.gate_bits) {
        if (p.cls_n) {
          const int cpc = p.ldo >> 5;   // 32-column chunks per class
#pragma unroll
          for (int ci = 0; ci < 8; ++ci) {
            const int c = ci / cpc;     // cpc = 2 (Cin = 64): classes 0 .. 3
            if ((r.vmask >> c) & 1u)
              r.gw[ci] = __ldg(p.gate_bits + ((r.off0 + ((c >> 1) * p.OWf + (c & 1)) * p.ldo + (ci - c * cpc) * 32) >> 5));
          }
        } else {
          const uint32_t* gb = p.gate_bits + (r.off0 >> 5);
#pragma unroll
          for (int ci = 0; ci < 4; ++ci)
            if (ci * 32 < p.block_n && n0 + ci * 32 < p.N_total) r.gw[ci] = __ldg(gb + ci);
        }
      }
      r.row_ok = row_ok;
      r.orow = orow;
      return r;
    };
    const int tstep = (int)gridDim.x * (p.epi_tile_split ? kEpiGroups : 1);   // with the tile split the groups alternate tiles
    const int tcstep = p.epi_tile_split ? kEpiGroups : 1;
    int t = blockIdx.x + (p.epi_tile_split ? grp * (int)gridDim.x : 0);
    tcount = p.epi_tile_split ? grp : 0;
    int sub = 0;
    bool have = t < p.total_tiles;
    Tile tl;
    RowPre cur;
    if (have) { tl = decode_tile(p, t); cur = prep(tl, 0); }
#pragma unroll 1
    while (have) {
      const long long ec0 = eprof ? clock64() : 0;
      // the unit after this one: its loads fly while this one's items run
      int nt = t, nsub = sub + 1, ntc = tcount;
      if (nsub == p.m_sub) { nsub = 0; nt = t + tstep; ntc = tcount + tcstep; }
      const bool nhave = nt < p.total_tiles;
      Tile ntl = tl;
      RowPre nxt = cur;
      if (nhave) {
        if (nsub == 0) ntl = decode_tile(p, nt);
        nxt = prep(ntl, nsub);
      }
      const int n0 = tl.n0;
      const int buf = tcount & (p.acc_bufs - 1);
      const int use = tcount >> p.acc_shift;
      if (sub == 0) {
        const long long ec1 = eprof ? clock64() : 0;
        if (eprof) e_pre += ec1 - ec0;
        bool ok = ptx::mbar_wait(tfull0 + 8 * buf, use & 1);
        if (eprof) e_wait += clock64() - ec1;
        if (!ok) report_timeout(p, 104);
        ok = __all_sync(0xffffffffu, ok);
        if (!ok) break;
        ptx::tc_fence_after();
      }
      {
        const bool row_ok = cur.row_ok;
        const long long orow = cur.orow, off0 = cur.off0;
        const uint32_t keep_row = cur.keep_row;
        const int q_sh = cur.q_sh, q_sw = cur.q_sw;
        const float* rm4 = cur.rm4;
        const uint32_t* gw = cur.gw;
        int item = sub * tl.ntaps * n_chunks;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16) + buf * p.acc_stride + sub * p.block_n;
        // The item loop is deliberately NOT unrolled: a rolled body of ~150 instructions stays in the
        // instruction cache across chunks and tiles.
        for (int a = 0; a < tl.ntaps; ++a) {
#pragma unroll 1
          for (int ci = 0; ci < n_chunks; ++ci, ++item) {
            if (!p.epi_tile_split && (item & (kEpiGroups - 1)) != grp) continue;   // the other group's chunk
            const int c = ci * 32;
            const int col0 = n0 + c;
            bool live = row_ok && col0 < p.N_total;
            long long coff = c;   // offset of this chunk from the row's first element
            if (p.cls_n) {
              const int cpc = p.ldo >> 5, k = ci / cpc;
              live = ((cur.vmask >> k) & 1u) != 0u;
              coff = (long long)((k >> 1) * p.OWf + (k & 1)) * p.ldo + (ci - k * cpc) * 32;
            }
            uint32_t keep = keep_row;
            float rmask = rm4[0];
            if (p.pair_C > 0 && col0 >= p.pair_C) rmask = rm4[1];
            if (p.quad_C > 0) {
              const int qd = fdiv(col0, p.fd_quad);
              const int ih = 2 * q_sh + (qd >> 1) - 1, iw = 2 * q_sw + (qd & 1) - 1;
              keep = (ih < 0 || ih >= p.s2d_H || iw < 0 || iw >= p.s2d_W) ? 0u : 0xffffffffu;
              rmask = qd == 0 ? rm4[0] : qd == 1 ? rm4[1] : qd == 2 ? rm4[2] : rm4[3];
            }
            if constexpr ((kEF & EF_F32) == 0u) {
              if ((kEF & EF_RUNTIME) == 0u || out_bf16) {
                const long long off = off0 + coff;
                const uint32_t gword = ci == 0 ? gw[0] : ci == 1 ? gw[1] : ci == 2 ? gw[2] : ci == 3 ? gw[3] :
                                       ci == 4 ? gw[4] : ci == 5 ? gw[5] : ci == 6 ? gw[6] : gw[7];
                epi_item_bf16<kEF>(epi_flags, t_lane + a * p.block_n + c, live, bias_base + col0, resid_base + off,
                                   gate_base + off, gword, rmask, keep, reinterpret_cast<bf16*>(p.out) + off,
                                   bits_base + (off >> 5));
                continue;
              }
            }
            if constexpr ((kEF & (EF_F32 | EF_RUNTIME)) != 0u) {
            uint32_t v[32];
            __syncwarp();
            ptx::tmem_ld_32x32(t_lane + a * p.block_n + c, v);
            ptx::tmem_ld_wait();
            if (!live) continue;
            {
              float f[32];
#pragma unroll
              for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
              if (p.swap_ab) {
                // lane = cin (row), column = cout: 32 lanes write 32 consecutive cin of one cout
                float* o = reinterpret_cast<float*>(p.out) + (long long)col0 * p.ldo +
                           (long long)(tl.tap0 + a) * p.tap_stride + orow;
                if (p.out_mode == OUT_F32_STORE) {
#pragma unroll
                  for (int i = 0; i < 32; ++i) o[(long long)i * p.ldo] = f[i];
                } else {
#pragma unroll
                  for (int i = 0; i < 32; ++i) atomicAdd(o + (long long)i * p.ldo, f[i]);
                }
              } else {
                float* o = reinterpret_cast<float*>(p.out) + orow * p.ldo +
                           (long long)(tl.tap0 + a) * p.tap_stride + col0 + tl.split * p.split_stride;
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
            }   // fp32 output path
          }
        }
      }
      if (eprof) { e_tile += clock64() - ec0; ++e_n; }
      if (sub == p.m_sub - 1) {
        // accumulator buffer drained: hand it back to the MMA issuer
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(tempty0 + 8 * buf);
      }
      t = nt; sub = nsub; tcount = ntc; tl = ntl; cur = nxt; have = nhave;
    }
    if (eprof) {
      atomicAdd(&g_prof[8], (unsigned long long)e_pre); atomicAdd(&g_prof[9], (unsigned long long)e_wait);
      atomicAdd(&g_prof[11], (unsigned long long)e_tile); atomicAdd(&g_prof[12], (unsigned long long)e_n);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) ptx::tmem_dealloc(tmem_base, p.tmem_cols);
}

// ---------------------------------------------------------------------------------------------
// host: launch
// ---------------------------------------------------------------------------------------------
static std::once_flag g_attr_once;

static thread_local const void* t_next_gate_bits = nullptr;
static thread_local void* t_next_bits_out = nullptr;

static thread_local float* t_next_dbias = nullptr;
static thread_local int t_next_dbias_mod = 0;

void set_next_dbias(float* dbias, int mod) {
  t_next_dbias = dbias;
  t_next_dbias_mod = mod;
}

void set_next_bits(const void* gate_bits, void* bits_out) {
  t_next_gate_bits = gate_bits;
  t_next_bits_out = bits_out;
}

int launch(const Params& p_in, int num_sms, size_t smem_bytes, cudaStream_t stream) {
  Params p = p_in;
  if (t_next_gate_bits || t_next_bits_out) {
    const void* gb = t_next_gate_bits;
    void* bo = t_next_bits_out;
    t_next_gate_bits = nullptr;
    t_next_bits_out = nullptr;
    // sign words cover 32 consecutive channels of one row of a bf16 output
    if (p.out_mode != OUT_BF16 || (p.ldo & 31) || (p.block_n & 31)) {
      gabril_set_error("gabril_conv_epilogue_bits: the next launch has no bf16 output with 32-aligned rows");
      return GABRIL_ERR_BAD_SHAPE;
    }
    if (gb) { p.gate_bits = (const uint32_t*)gb; p.gate = nullptr; }
    p.bits_out = (uint32_t*)bo;
  }
  if (t_next_dbias) {
    float* db = t_next_dbias;
    const int mod = t_next_dbias_mod;
    t_next_dbias = nullptr;
    t_next_dbias_mod = 0;
    const bool dy_is_b = p.swap_ab && !p.a_flat;   // the flat exchanged form keeps dY in the stage's first region
    const int cols = p.swap_ab ? p.N_total : std::min(p.M_total[0], 128);
    const int row_bytes = 2 * (dy_is_b ? p.b_box_w : p.a_box_w);
    if (p.kind != KIND_WGRAD || p.total_tiles > num_sms || p.kc != 64 || (row_bytes != 64 && row_bytes != 128) ||
        (cols != 32 && cols != 64 && cols != 128) || (!p.swap_ab && p.M_total[0] % cols) || cols > kBiasSmem) {
      gabril_set_error("gabril_conv_wgrad_dbias_next: the next launch is not a weight gradient with 32 / 64 / 128-wide dY tiles");
      return GABRIL_ERR_BAD_SHAPE;
    }
    p.dbias = db;
    p.dbias_mod = mod;
    p.dy_cols = cols;
    p.dy_row_bytes = row_bytes;
    p.dy_off = dy_is_b ? p.a_stage_bytes : 0u;
    p.dy_box_bytes = dy_is_b ? p.b_box_bytes : p.a_box_bytes;
  }
  auto magic = [](int d) {
    FastDiv f{0u, 0u};
    if (d > 1) {
      uint32_t l = 0;
      while ((1ull << l) < (unsigned long long)d) ++l;
      f.mul = (uint32_t)((((1ull << l) - (unsigned long long)d) << 32) / (unsigned long long)d + 1ull);
      f.shr = l;
    }
    return f;
  };
  if (p.m_sub < 1) p.m_sub = 1;
  if (p.acc_bufs < 1) p.acc_bufs = 1;
  p.simple_tiles = (p.kind == KIND_FWD && p.n_classes == 1 && p.n_tiles == 1 && p.n_splits == 1) ? 1 : 0;
  p.acc_shift = 0;
  while ((1 << p.acc_shift) < p.acc_bufs) ++p.acc_shift;
  if ((1 << p.acc_shift) != p.acc_bufs || p.acc_bufs > kMaxAcc || p.acc_bufs * p.acc_stride > 512) {
    gabril_set_error("tc::launch: accumulator ring must be a power of two <= 8 buffers within 512 TMEM columns");
    return GABRIL_ERR_BAD_SHAPE;
  }
  // a tile that is a single 32-column item cannot be shared by two epilogue groups: they alternate tiles
  p.epi_flags = (p.bias ? EF_BIAS : 0u) | (p.resid ? EF_RESID : 0u) | (p.relu ? EF_RELU : 0u) |
                (p.gate_bits ? EF_GBITS : 0u) | (p.rowmask ? EF_RMASK : 0u) | (p.border_h > 0 ? EF_KEEP : 0u) |
                (p.bits_out ? EF_BITS : 0u) | (p.gate ? EF_GATE : 0u);
  p.epi_mode = 0;
#define EPI_PICK(idx, F) if (p.epi_flags == (F)) p.epi_mode = idx;
  EPI_VARIANTS(EPI_PICK)
#undef EPI_PICK
  if (p.epi_groups != 2) p.epi_groups = 1;
  p.epi_tile_split = (p.epi_groups > 1 && p.kind == KIND_FWD && p.m_sub == 1 && p.block_n <= 32 && p.acc_bufs >= 2) ? 1 : 0;
  if (p.out_mode == OUT_BF16 &&
      ((reinterpret_cast<uintptr_t>(p.out) | reinterpret_cast<uintptr_t>(p.resid) | reinterpret_cast<uintptr_t>(p.gate)) & 31)) {
    gabril_set_error("tc::launch: bf16 output / residual / gate tensors must be 32-byte aligned");
    return GABRIL_ERR_BAD_SHAPE;
  }
  p.fd_tiles_m = magic(p.tiles_m_total);
  p.fd_n_classes = magic(p.n_classes);
  p.fd_n_tiles = magic(p.n_tiles);
  p.fd_tap_groups = magic(p.tap_groups);
  p.fd_border_w = magic(p.border_w);
  p.fd_border_h = magic(p.border_h);
  p.fd_border_hw = magic(p.border_h * p.border_w);
  p.fd_quad = magic(p.quad_C);
  for (int c = 0; c < 4; ++c) {
    p.fd_hw[c] = magic(p.Hs[c] * p.Ws[c]);
    p.fd_ws[c] = magic(p.Ws[c]);
    p.fd_ohw[c] = magic(p.OHW[c]);
    p.fd_ow[c] = magic(p.OW[c]);
  }
  // kernel table: [group count - 1][variant]; variant 0 = general form, EPI_VARIANTS indices, last = fp32 only
  typedef void (*KernelFn)(const Params);
  constexpr int kF32Slot = 12;
#define EPI_FN(idx, F) tc_gemm_kernel<kG, (F)>,
  static const KernelFn table[2][kF32Slot + 1] = {
#define kG 1
      {tc_gemm_kernel<1, EF_RUNTIME>, EPI_VARIANTS(EPI_FN) tc_gemm_kernel<1, EF_F32>},
#undef kG
#define kG 2
      {tc_gemm_kernel<2, EF_RUNTIME>, EPI_VARIANTS(EPI_FN) tc_gemm_kernel<2, EF_F32>},
#undef kG
  };
#undef EPI_FN
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(g_attr_once, [] {
    // static smem: barriers, TMEM base, bias vector (< 2 KB)
    for (int g = 0; g < 2 && attr_err == cudaSuccess; ++g)
      for (int v = 0; v <= kF32Slot && attr_err == cudaSuccess; ++v)
        attr_err = cudaFuncSetAttribute(reinterpret_cast<const void*>(table[g][v]),
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048);
  });
  if (attr_err != cudaSuccess) return gabril_set_cuda_error(attr_err);
  if (p.total_tiles <= 0) return GABRIL_OK;
  // one persistent CTA per SM (smem / TMEM are sized for single occupancy)
  const unsigned grid = (unsigned)std::min(p.total_tiles, num_sms);
  const int slot = (p.debug & 32) ? 0 : (p.out_mode != OUT_BF16 ? kF32Slot : p.epi_mode);
  const KernelFn fn = table[p.epi_groups - 1][slot];
  static int use_pdl = -1;
  if (use_pdl < 0) { const char* e = getenv("GABRIL_TC_PDL"); use_pdl = (e && e[0] == '0') ? 0 : 1; }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(32 * (kFirstEpilogueWarp + 4 * p.epi_groups));
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = use_pdl ? 1 : 0;
  void* args[1] = {(void*)&p};
  cudaError_t le = cudaLaunchKernelExC(&cfg, reinterpret_cast<const void*>(fn), args);
  if (le != cudaSuccess) return gabril_set_cuda_error(le);

  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

}  // namespace tc

extern "C" int gabril_tc_prof(unsigned long long* out16, int reset) {
  if (cudaMemcpyFromSymbol(out16, tc::g_prof, 16 * sizeof(unsigned long long)) != cudaSuccess) return -1;
  if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(tc::g_prof, z, sizeof(z)); }
  return 0;
}
