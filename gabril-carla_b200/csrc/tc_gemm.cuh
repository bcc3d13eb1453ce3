// tcgen05 / TMEM / TMA implicit-GEMM engine shared by every dense contraction of the
// GABRIL training step (reference: vlm_gaze/models/linear_models.py:124-207 Encoder convs,
// vlm_gaze/train/train_bc.py:79-82 pre_actor Linear) — forward, dgrad and wgrad.
//
// One kernel, two operand regimes:
//   KIND_FWD   D[pixels, Cout] = A[pixels, K] * B[Cout, K]^T      A K-major (TMA im2col or tiled),
//              (forward conv, dgrad as a conv over dY, linear)     B K-major or MN-major weight tile
//   KIND_WGRAD D[Cout, Cin]    = dY[pixels, Cout]^T * X[pixels, Cin]   both MN-major, contraction
//              (one accumulator per filter tap, several per CTA)       over pixels, split over CTAs
// Warp roles: warp0 = TMEM owner + single-thread MMA issuer, warps1..3 = TMA producers, warps 4.. =
// epilogue (tcgen05.ld -> fused bias/residual/ReLU/gate/mask -> global), one or two groups of four.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace tc {

constexpr int kMaxStages = 8;
constexpr int kMaxAcc = 8;             // TMEM accumulator ring depth (power of two)
constexpr int kBiasSmem = 256;         // bias vectors up to this many columns are staged in shared memory
constexpr int kProducers = 3;          // TMA-issuing warps
constexpr int kMmaWarp = 0;            // TMEM owner + single-thread MMA issuer
constexpr int kFirstProducerWarp = 1;  // warps 1..3
constexpr int kFirstEpilogueWarp = 4;  // warps 4..7 (group 0), 8..11 (group 1); TMEM lane quarter = warp % 4

enum Kind { KIND_FWD = 0, KIND_WGRAD = 1 };

// Terms of the bf16 epilogue (see epi_item_bf16 in tc_gemm.cu) and the combinations compiled as
// straight-line variants (index, mask): every conv / linear launch of the training step hits one of them.
enum : uint32_t {
  EF_BIAS = 1u, EF_RESID = 2u, EF_RELU = 4u, EF_GBITS = 8u, EF_RMASK = 16u, EF_KEEP = 32u, EF_BITS = 64u,
  EF_GATE = 128u, EF_F32 = 1u << 30, EF_RUNTIME = 1u << 31
};
#define EPI_VARIANTS(X)                                                                     \
  X(1, EF_BIAS | EF_RELU | EF_BITS)             /* conv_2 .. conv_5 forward */              \
  X(2, EF_BIAS | EF_RELU | EF_BITS | EF_KEEP)   /* conv_1 forward (zero border quadrants) */ \
  X(3, EF_RELU | EF_BITS)                       /* residual 3x3 forward */                  \
  X(4, EF_RESID | EF_RELU | EF_BITS)            /* residual 1x1 forward + skip */           \
  X(5, EF_GBITS)                                /* data gradients through a ReLU */         \
  X(6, EF_RESID | EF_GBITS)                     /* residual 3x3 dgrad + skip gradient */    \
  X(7, EF_BIAS)                                 /* conv_6 forward */                        \
  X(8, EF_BIAS | EF_RMASK)                      /* conv_6 forward with the GMD mask */      \
  X(9, EF_RESID)                                /* pre_actor dgrad + regulariser gradient */ \
  X(10, 0u)                                     /* plain dgrad */                           \
  X(11, EF_BIAS | EF_RELU)                      /* actor.0 forward */
enum OutMode { OUT_BF16 = 0, OUT_F32_STORE = 1, OUT_F32_ATOMIC = 2 };

// n / d for 0 <= n < 2^31 as (umulhi(n, mul) + n) >> shr (round-up multiplier, Granlund-Montgomery):
// the epilogue warps decode a tile and a pixel per tile, and a runtime 32-bit division is a serial
// chain of ~30 instructions on a warp that runs alone on its scheduler.
struct FastDiv {
  uint32_t mul, shr;
};

struct Params {
  CUtensorMap mapA[4];  // FWD: activation (one per strided-dgrad parity class); WGRAD: dY (tiled)
  CUtensorMap mapB;     // FWD: weights [Cout][tap][Cin];  WGRAD: activation (im2col or tiled)
  int kind, n_classes;
  int a_im2col;  // FWD: A through im2col (else tiled 2D). WGRAD: B through im2col (else tiled 2D)
  int b_mn;      // FWD: weight tile is MN-major (dgrad: contraction over Cout)
  int M_total[4];
  int N_total;
  int block_n;
  int n_acc;       // accumulators per CTA (WGRAD: taps per CTA; FWD: 1)
  int tap_groups;  // WGRAD: ceil(taps / n_acc)
  int taps;        // filter taps walked by the traversal
  int kc;          // contraction elements per k-block
  int kb_total, kb_per_split;
  int stages;
  uint32_t a_stage_bytes, b_stage_bytes;
  uint32_t a_box_bytes, b_box_bytes;
  int a_boxes, b_boxes;  // TMA boxes per stage (B: per tap for WGRAD)
  int a_box_w, b_box_w;  // elements one box covers along the MN dimension
  // im2col traversal geometry (per parity class)
  int OW[4], OHW[4], lower_w[4], lower_h[4];
  int trav_stride_w, trav_stride_h;
  int S;  // traversal filter width: tap -> (r', s') = (tap / S, tap % S)
  int chunks_per_tap;
  // weight tap mapping r = r_base + r_step*r', s = s_base + s_step*s', wtap = r*S_w + s
  int r_base[4], s_base[4], r_step, s_step, S_w;
  // MMA schedule of one k-block
  // MMA schedule of one k-block: mma_groups accumulator groups x mma_k K-steps of 16.
  //   A start = stage A base + j*a_kstep;  B start = stage B base + g*b_gstep + j*b_kstep
  //   (16-byte units);  accumulator column = g*d_gstep;  group g covers taps [g*g_taps, ...)
  int mma_k, mma_groups, g_taps;
  uint32_t a_kstep, b_kstep, b_gstep, d_gstep;
  // WGRAD with exchanged operands (narrow Cout): A = activation (im2col, one tile per tap, M = Cin),
  // B = dY (one tile for all taps, N = Cout).  Group g = tap g: A start += g*a_gstep (b_gstep = 0);
  // TMEM lane = cin, column = cout, so the epilogue writes element (cout, tap, cin) transposed.
  int swap_ab;
  uint32_t a_gstep;
  // WGRAD with exchanged operands and the SHIFT ON dY (stride-1 "same" convolutions with narrow Cout):
  //   dW[co][r,s][ci] = sum_q x[q][ci] * dY[q - (r - pad, s - pad)][co]
  // A = the activation rows of the k-block, ONE tile for every tap (plain 2-D map, M = Cin = 128); B = one
  // im2col tile of dY per tap (the data-gradient traversal), consecutive in the stage, so one MMA covers g_taps
  // taps with N = g_taps * Cout (the last group of a CTA is narrower: its N is patched into the instruction
  // descriptor).  Against the form above this loads x once per k-block instead of once per tap.
  // center_tap = (pad, pad): its tile is dY itself (bias gradient).
  int shift_dy, center_tap;
  // FWD regime, data gradient of a k4 / s2 / pad-0 convolution with the FOUR INPUT-PARITY CLASSES STACKED ALONG N:
  // without padding every class (ah, aw) reads the same 2x2 neighbourhood of dY, dY[jh-1+r', jw-1+s'], and differs
  // only in its weight taps (kh, kw) = (ah + 2 - 2r', aw + 2 - 2s').  One tile = 128 base pixels (jh, jw) x
  // (4 classes x Cin) columns: the dY tile is staged once instead of once per class, one N = 4*Cin MMA replaces
  // four N = Cin ones (narrow MMAs cost the tensor pipe clearly more than N / 128 of a wide one), and there are a quarter
  // of the tiles.  B box i of a stage = class i's weight tap (r_base[i], s_base[i]); the epilogue scatters column
  // chunk c to pixel (2jh + ah, 2jw + aw) of class c / (Cin / 32).
  int cls_n;
  uint64_t desc_a_hi, desc_b_hi;  // smem descriptors with the start-address field zero
  uint32_t idesc;
  uint32_t tmem_cols;
  // epilogue
  int out_mode;
  void* out;
  long long ldo;
  const float* bias;            // per output column, or null
  const __nv_bfloat16* resid;   // added before ReLU / gate (same indexing as out), or null
  const __nv_bfloat16* gate;    // out = gate > 0 ? out : 0 (ReLU backward), or null
  const float* rowmask;         // per output row multiplier (GMD mask), or null
  // packed sign bits (1 bit per element, word (off >> 5) covers elements off .. off + 31 of the
  // tensor indexed like out): gate_bits replaces the `gate` tensor read (64 B -> 4 B per 32-column
  // chunk); bits_out receives (stored bf16 value > 0) for every chunk this launch stores
  const uint32_t* gate_bits;
  uint32_t* bits_out;
  int relu;
  // remap: tile row (n, jh, jw) over the Hs x Ws grid of its class -> row
  //        (n*OHf + jh*remap_sh + ah)*OWf + jw*remap_sw + aw of the output tensor
  //        (strided-dgrad parity classes; writes into the interior of a pre-padded tensor)
  int remap, remap_sh, remap_sw;
  int Hs[4], Ws[4], ah[4], aw[4], OHf, OWf;
  // border_h/w > 0: rows are pixels of a border_h x border_w padded grid; border pixels store 0.
  // s2d_H/W > 0 instead: rows are (super-pixel, quadrant) of a border_h x border_w space-to-depth
  // grid over an s2d_H x s2d_W image padded by one pixel; rows of padding pixels store 0.
  int border_h, border_w, s2d_H, s2d_W;
  // quad_C > 0 (with border_h/w = the grid and s2d_H/W = the image): rows are super-pixels (n, sh, sw) and
  // the columns four quadrants (qh, qw) of quad_C channels = pixels (2sh+qh-1, 2sw+qw-1) of the image;
  // quadrants outside it store 0, rowmask is indexed [row*4 + quadrant], bias by column % bias_mod.
  int quad_C, bias_mod;
  // pair_C > 0 (gabril_conv_fwd_pairs): a row holds two adjacent pixels of pair_C channels; rowmask has one entry per
  // pixel, i.e. [row*2 + half]
  int pair_C;
  // remap validity: rows with jh >= valid_h or jw >= valid_w are not stored (0 = no check)
  int valid_h, valid_w;
  // flat-slab mode: the A rows (FWD) / B rows (WGRAD) are the flat pixels of a zero-bordered
  // tensor.  Filter row r is ONE 2-D TMA slab of (tile rows + flat_S - 1) pixels starting at
  // row0 + r * flat_pitch; tap (r, s) is that slab viewed from row s on (descriptor start
  // address + s * 128 bytes: the 128-byte swizzle is a function of the absolute shared-memory
  // address, so whole-row shifts of the start address are legal).
  int a_flat, flat_pitch, flat_R, flat_S;
  int dy_row_off;           // WGRAD flat: dY row of contraction index q is q + dy_row_off
  int rows_per_group;       // WGRAD flat: filter rows per CTA (taps per CTA = rows_per_group * flat_S)
  int b_resident;           // FWD flat: all weight tiles are loaded once per CTA
  uint32_t b_tap_bytes;     // bytes of one weight tap tile (FWD flat) / one slab (WGRAD flat)
  uint32_t slab_tx_bytes;   // bytes one slab load transfers
  uint32_t resident_off;    // offset of the resident weight region from the aligned smem base
  // persistent tile schedule: tile -> (m tile [classes concatenated], n tile, split)
  int tiles_m[4], tiles_m_total, n_tiles, n_splits, total_tiles;
  int simple_tiles;          // FWD, one class, one column tile, no split: tile id = m tile (filled by launch())
  int cls_min_tiles;         // > 0: the classes' m tiles are interleaved (decode_tile); = min over classes of tiles_m
  int acc_bufs, acc_stride;  // TMEM accumulator ring (1, 2, 4 or 8 buffers of acc_stride columns)
  int acc_shift;             // log2(acc_bufs), filled by launch()
  // FWD regime: 128-row sub-tiles per CTA tile (1 or 2).  With 2, one A box of 256 rows and ONE weight
  // tile feed two accumulators (columns [0, block_n) and [block_n, 2*block_n) of the buffer), which cuts
  // the shared-memory fill per MAC by a quarter for 128-wide layers (the L2 -> SM path bounds them).
  int m_sub;
  // epilogue groups split the work by whole tiles (a tile is a single 32-column item) instead of by items
  int epi_tile_split;
  int epi_groups;   // 1 (256 threads, two CTAs per SM) or 2 (384 threads, single-occupancy launches)
  uint32_t epi_flags;   // EF_* terms of this launch's bf16 epilogue, filled by launch()
  int epi_mode;         // index into EPI_VARIANTS, 0 = general form
  long long tap_stride;  // WGRAD: element (m, tap, n) at out + m*ldo + tap*tap_stride + n
  long long split_stride;  // FWD split-K with plain stores: split s writes its partial at out + s*split_stride
  // magic numbers of the divisions above, filled by launch() from the fields they belong to
  FastDiv fd_n_classes;
  FastDiv fd_tiles_m, fd_n_tiles, fd_tap_groups, fd_border_w, fd_border_h, fd_border_hw, fd_quad;
  FastDiv fd_hw[4], fd_ws[4];    // remap: Hs*Ws, Ws per class
  FastDiv fd_ohw[4], fd_ow[4];   // im2col traversal: OHW, OW per class
  // WGRAD: bias gradient fused into the weight-gradient launch (gabril_conv_wgrad_dbias_next).  The dY tile
  // of every k-block is in shared memory anyway (A operand, or B in the exchanged form); the CTAs of tap
  // group 0 / column tile 0 let their epilogue warps - idle until the k loop ends - add up its columns:
  //   dbias[(col0 + c) % dbias_mod] += sum_pixels dY[pixel][c]
  // which replaces a separate pass that re-read every dY from HBM (2.2 GB per step at batch 1024).
  float* dbias;
  int dbias_mod;
  uint32_t dy_off, dy_box_bytes;   // the dY tile inside a stage: boxes of [kc pixels][dy_row_bytes / 2 channels]
  int dy_row_bytes, dy_cols;       // 64 or 128 (= the TMA swizzle span); channels summed per tile (<= 128)
  int* err_flag;
  int debug;  // perf experiments only: bit0 = producer skips the TMA copies, bit1 = issuer skips the MMAs
};

// Fills the launch geometry and launches; returns a negative gabril status on failure.
int launch(const Params& p, int num_sms, size_t smem_bytes, cudaStream_t stream);

// Launch attribute of the NEXT launch() on the calling host thread (consumed by it): see
// gabril_conv_epilogue_bits in include/gabril_b200.h.
void set_next_bits(const void* gate_bits, void* bits_out);
// Same mechanism: the next WGRAD launch() of this thread also accumulates the bias gradient into dbias.
void set_next_dbias(float* dbias, int mod);

}  // namespace tc
