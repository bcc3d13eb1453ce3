// Host-side planners for the tcgen05 implicit-GEMM kernel: they translate a convolution /
// linear layer (forward, dgrad, wgrad) into TMA tensor maps, UMMA descriptors and a launch grid.
// C-ABI entry points: gabril_conv_fwd / gabril_conv_dgrad / gabril_conv_wgrad (declared in
// include/gabril_b200.h).  Activations are NHWC bf16, weights are [Cout][R*S][Cin] bf16,
// weight gradients are fp32 in the same [Cout][R*S][Cin] order.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "res_block.cuh"
#include "tc_gemm.cuh"

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const int*, const int*,
                                   cuuint32_t, cuuint32_t, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn g_encode_tiled = nullptr;
EncodeIm2colFn g_encode_im2col = nullptr;
int g_driver_version = 0;
int g_num_sms = 148;
int* g_err_flag = nullptr;  // device int, polled by gabril_poll_error

int resolve_driver() {
  // cuTensorMapEncode* are driver entry points and need a context current on the CALLING thread.  A
  // host thread that never issued a runtime call (autograd's backward thread on device 0: torch skips
  // cudaSetDevice when the device already is the thread's default) has none -> error 201; cudaFree(0)
  // binds the primary context of the thread's current device.  Once per thread.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    cudaFree(0);
    ctx_bound = true;
  }
  if (g_encode_tiled && g_encode_im2col) return GABRIL_OK;
  cudaDriverEntryPointQueryResult qres;
  void* fn = nullptr;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) return GABRIL_ERR_DRIVER;
  g_encode_tiled = reinterpret_cast<EncodeTiledFn>(fn);
  fn = nullptr;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) return GABRIL_ERR_DRIVER;
  g_encode_im2col = reinterpret_cast<EncodeIm2colFn>(fn);
  cudaDriverGetVersion(&g_driver_version);
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
  if (!g_err_flag) {
    if (cudaMalloc(&g_err_flag, sizeof(int)) != cudaSuccess) return GABRIL_ERR_CUDA;
    cudaMemset(g_err_flag, 0, sizeof(int));
  }
  return GABRIL_OK;
}

CUtensorMapSwizzle swizzle_for_bytes(int inner_bytes) {
  if (inner_bytes >= 128) return CU_TENSOR_MAP_SWIZZLE_128B;
  if (inner_bytes >= 64) return CU_TENSOR_MAP_SWIZZLE_64B;
  if (inner_bytes >= 32) return CU_TENSOR_MAP_SWIZZLE_32B;
  return CU_TENSOR_MAP_SWIZZLE_NONE;
}
// UMMA smem-descriptor layout_type for a swizzle span in bytes.
int layout_type_for_bytes(int inner_bytes) {
  if (inner_bytes >= 128) return 2;
  if (inner_bytes >= 64) return 4;
  if (inner_bytes >= 32) return 6;
  return 0;
}

int tiled_map(CUtensorMap* m, const void* ptr, int rank, const uint64_t* dims,
              const uint64_t* strides_bytes /* rank-1 */, const uint32_t* box, int inner_bytes) {
  cuuint64_t gd[5];
  cuuint64_t gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  CUresult r = g_encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank,
                              const_cast<void*>(ptr), gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              swizzle_for_bytes(inner_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "[gabril] cuTensorMapEncodeTiled failed: %d (rank %d, box0 %u)\n", (int)r, rank,
            box[0]);
    return GABRIL_ERR_DRIVER;
  }
  return GABRIL_OK;
}

// NHWC activation seen as (C, W, H, N); the traversal walks `box_pix` base pixels with `stride`.
int im2col_map(CUtensorMap* m, const void* ptr, int N, int H, int W, int C, int lower_w,
               int lower_h, int upper_w, int upper_h, int box_c, int box_pix, int stride, int stride_h = 0) {
  if (stride_h <= 0) stride_h = stride;
  cuuint64_t gd[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t gs[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  int lo[2] = {lower_w, lower_h};
  int up[2] = {upper_w, upper_h};
  cuuint32_t es[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride_h, 1};
  CUresult r = g_encode_im2col(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), gd,
                               gs, lo, up, (cuuint32_t)box_c, (cuuint32_t)box_pix, es,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for_bytes(box_c * 2),
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr,
            "[gabril] cuTensorMapEncodeIm2col failed: %d (N%d H%d W%d C%d lo(%d,%d) up(%d,%d) "
            "box_c %d box_pix %d stride %d)\n",
            (int)r, N, H, W, C, lower_w, lower_h, upper_w, upper_h, box_c, box_pix, stride);
    return GABRIL_ERR_DRIVER;
  }
  // Same driver quirk CUTLASS works around for im2col maps over tensors below 128 KiB.
  if (g_driver_version <= 13010 && (size_t)N * H * W * C * 2 < 131072)
    reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
  return GABRIL_OK;
}

uint64_t desc_hi(uint32_t lbo_bytes, uint32_t sbo_bytes, int layout_type) {
  uint64_t d = 0;
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  d |= (uint64_t)layout_type << 61;
  return d;
}

uint32_t make_idesc(int a_mn, int b_mn, int M, int N) {
  uint32_t d = 0;
  d |= 1u << 4;   // D format fp32
  d |= 1u << 7;   // A bf16
  d |= 1u << 10;  // B bf16
  d |= (uint32_t)(a_mn ? 1 : 0) << 15;
  d |= (uint32_t)(b_mn ? 1 : 0) << 16;
  d |= (uint32_t)(N >> 3) << 17;
  d |= (uint32_t)(M >> 4) << 24;
  return d;
}

uint32_t round_up(uint32_t v, uint32_t a) { return (v + a - 1) / a * a; }
uint32_t pow2_cols(uint32_t c) {
  uint32_t r = 32;
  while (r < c) r <<= 1;
  return r;
}

// N-side channel counts the MN-major box split supports: 32, 64, or a multiple of 128.
bool mn_ok(int c) { return c == 32 || c == 64 || (c >= 128 && c % 128 == 0); }
// Output-column counts of the K-major forward: a multiple of 32 up to 128, or a multiple of 128.
bool fwd_cout_ok(int c) { return c % 32 == 0 && (c <= 128 || c % 128 == 0); }

void zero_params(tc::Params& p) {
  memset(&p, 0, sizeof(p));
  static int dbg = -1;
  if (dbg < 0) { const char* e = getenv("GABRIL_TC_DEBUG"); dbg = e ? atoi(e) : 0; }
  p.debug = dbg;
}

// Persistent single-occupancy CTA: the stage ring takes what shared memory allows (ring runs
// across tiles, so even one-k-block tiles benefit from depth).
int g_occ = 0;  // CTAs per SM the FWD-regime launches are sized for (GABRIL_TC_OCC, default 2)
int occ_setting() {
  if (!g_occ) {
    const char* e = getenv("GABRIL_TC_OCC");
    g_occ = (e && atoi(e) == 1) ? 1 : 2;
  }
  return g_occ;
}

void pick_stages(tc::Params& p, size_t* smem_bytes, int occ) {
  const uint32_t stage = p.a_stage_bytes + p.b_stage_bytes;
  int st = (int)(((occ == 2 ? 104u : 208u) * 1024u) / stage);
  st = std::max(2, std::min(st, tc::kMaxStages));
  p.stages = st;
  *smem_bytes = (size_t)st * stage + 1024;
}

// Tile schedule + TMEM accumulator ring.  m_tiles[c] = 128-row tiles of class c; n_tiles_y = n
// tiles (x tap groups for WGRAD); splits = split-K factor.
void finalize_schedule(tc::Params& p, const int* m_tiles, int n_tiles_y, int splits, int acc_cols,
                       int occ) {
  p.tiles_m_total = 0;
  for (int c = 0; c < p.n_classes; ++c) { p.tiles_m[c] = m_tiles[c]; p.tiles_m_total += m_tiles[c]; }
  p.cls_min_tiles = 0;
  if (p.n_classes > 1) {
    static int inter = -1;
    if (inter < 0) { const char* e = getenv("GABRIL_TC_CLS_INTERLEAVE"); inter = (e && e[0] == '0') ? 0 : 1; }
    int mn = m_tiles[0];
    for (int c = 1; c < p.n_classes; ++c) mn = std::min(mn, m_tiles[c]);
    if (inter) p.cls_min_tiles = mn;
  }
  p.n_tiles = n_tiles_y;
  p.n_splits = splits;
  p.total_tiles = p.tiles_m_total * n_tiles_y * splits;
  p.acc_stride = (int)pow2_cols((uint32_t)acc_cols);
  // as many accumulator buffers as this CTA's share of TMEM holds (GABRIL_TC_ACC caps it): a tile's
  // buffer is busy from its first MMA until the epilogue has drained it, and for short-K tiles that
  // round trip (MMA latency + tcgen05.ld + stores) is several times the MMA issue time of the tile
  static int acc_max = 0;
  if (!acc_max) { const char* e = getenv("GABRIL_TC_ACC"); acc_max = e ? std::max(1, std::min(atoi(e), tc::kMaxAcc)) : tc::kMaxAcc; }
  p.acc_bufs = 1;
  while (p.acc_bufs * 2 <= acc_max && p.acc_stride * p.acc_bufs * 2 <= 512 / occ) p.acc_bufs *= 2;
  p.tmem_cols = (uint32_t)(p.acc_stride * p.acc_bufs);
  // single-occupancy launches get a second group of epilogue warps (GABRIL_TC_EPI=1 switches it off)
  static int epi_max = 0;
  if (!epi_max) { const char* e = getenv("GABRIL_TC_EPI"); epi_max = (e && atoi(e) == 1) ? 1 : 2; }
  p.epi_groups = (occ == 1 && epi_max == 2) ? 2 : 1;
}

struct ActDims { int N, H, W, C; };

// Common K-major / MN-major schedule helpers ---------------------------------------------------
// FWD regime: A K-major with kc elements per k-block; B K-major or MN-major.
// 128-row sub-tiles per CTA tile for a FWD-regime launch over M rows (GABRIL_TC_MSUB: 1, 2; default 2):
// two only when every SM still gets several tiles and the accumulators fit a single-occupancy CTA.
// Measured at batch 1024 (tools/exp_tiles.sh): 256-row tiles pay off only where the k loop is long and
// the tile wide (conv_5 fwd / dgrad: K = 3200, N = 128: 0.54 -> 0.48 ms); shorter or narrower layers lose
// more to the single-occupancy CTA (no second CTA to hide the epilogue) than the smaller feed saves.
// With the lean issue loop (tc_gemm.cu) eight MMAs per k-block no longer saturate the issuing thread, and the
// threshold moved from 40 to 16 k-blocks: conv_4 dgrad 0.243 -> 0.226 ms, conv_6 dgrad 0.277 -> 0.258 ms,
// conv_3 / conv_4 forward 2 %; N = 64 tiles (conv_6 forward, conv_3 dgrad) still lose.
int pick_m_sub(long long M, int block_n, int kb_total) {
  static int want = 0;   // 1: never, 2 (default): where it pays, 3: always (tests)
  if (!want) { const char* e = getenv("GABRIL_TC_MSUB"); want = e ? std::max(1, std::min(atoi(e), 3)) : 2; }
  if (want == 1 || block_n > 128) return 1;
  if (want == 3) return 2;
  return (block_n == 128 && kb_total >= 16 && M >= 256ll * 4 * g_num_sms) ? 2 : 1;
}

void fwd_schedule(tc::Params& p, int kc, int block_n, int b_mn, int m_sub = 1) {
  const int a_sw = kc * 2;  // 128 or 64 byte swizzle span
  p.kc = kc;
  p.block_n = block_n;
  p.n_acc = 1;
  p.m_sub = m_sub;
  p.a_boxes = 1;
  p.a_box_bytes = 128 * m_sub * a_sw;
  p.a_stage_bytes = round_up(p.a_box_bytes, 1024);
  p.desc_a_hi = desc_hi(16, 8 * a_sw, layout_type_for_bytes(a_sw));
  if (!b_mn) {
    p.b_boxes = 1;
    p.b_box_w = block_n;
    p.b_box_bytes = block_n * a_sw;
    p.desc_b_hi = desc_hi(16, 8 * a_sw, layout_type_for_bytes(a_sw));
  } else {
    const int bw = std::min(64, block_n);
    const int b_sw = bw * 2;
    p.b_boxes = block_n / bw;
    p.b_box_w = bw;
    p.b_box_bytes = kc * b_sw;
    p.desc_b_hi = desc_hi(p.b_box_bytes, 8 * b_sw, layout_type_for_bytes(b_sw));
  }
  p.b_stage_bytes = round_up(p.b_boxes * p.b_box_bytes, 1024);
  p.mma_k = kc / 16;
  // one MMA group per sub-tile: A start advances by one 128-row tile, the accumulator by block_n columns
  p.mma_groups = m_sub;
  p.g_taps = m_sub > 1 ? 0 : 1;
  p.a_gstep = m_sub > 1 ? (uint32_t)((128 * a_sw) >> 4) : 0u;
  p.a_kstep = 32 >> 4;
  p.b_kstep = (uint32_t)((b_mn ? 16 * (std::min(64, block_n) * 2) : 32) >> 4);
  p.b_gstep = 0;
  p.d_gstep = m_sub > 1 ? (uint32_t)block_n : 0u;
  p.idesc = make_idesc(0, b_mn, 128, block_n);
}

int weight_map(CUtensorMap* m, const void* w, int Cout, int T, int Cin, int b_mn, int kc,
               int block_n) {
  uint64_t dims[3] = {(uint64_t)Cin, (uint64_t)T, (uint64_t)Cout};
  uint64_t strides[2] = {(uint64_t)Cin * 2, (uint64_t)T * Cin * 2};
  if (!b_mn) {
    uint32_t box[3] = {(uint32_t)kc, 1, (uint32_t)block_n};
    return tiled_map(m, w, 3, dims, strides, box, kc * 2);
  }
  const int bw = std::min(64, block_n);
  uint32_t box[3] = {(uint32_t)bw, 1, (uint32_t)kc};
  return tiled_map(m, w, 3, dims, strides, box, bw * 2);
}

// w[Cout][R*S][Cin] -> w_pair[2*Cout][R*(S+1)][Cin]: rows [0, Cout) = w(r, t') for t' < S, rows [Cout, 2*Cout) =
// w(r, t' - 1) for t' >= 1, zeros elsewhere (gabril_conv_fwd_pairs)
__global__ void pack_wpairs_kernel(const uint4* __restrict__ w, uint4* __restrict__ wp, int Cout, int R, int S, int Cin8) {
  // one 16-byte group of 8 input channels per thread
  const int total = 2 * Cout * R * (S + 1) * Cin8;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c8 = i % Cin8;
  int t = i / Cin8;
  const int sp = t % (S + 1); t /= (S + 1);
  const int r = t % R;
  const int co2 = t / R;
  const int half = co2 / Cout, co = co2 - half * Cout, s0 = sp - half;
  wp[i] = (s0 >= 0 && s0 < S) ? __ldg(w + ((long long)co * R * S + r * S + s0) * Cin8 + c8) : make_uint4(0u, 0u, 0u, 0u);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
extern "C" {

int gabril_tc_init(void) { return resolve_driver(); }

int gabril_conv_epilogue_bits(const void* gate_bits, void* sign_bits_out) {
  tc::set_next_bits(gate_bits, sign_bits_out);
  return GABRIL_OK;
}

// One Residual block forward (linear_models.py:78-102) in one launch: h = relu(conv3x3(x)), out = relu(x +
// conv1x1(h)); x is the rectified block input (NHWC bf16, 128 channels), w3 [32][9][128], w1 [128][1][32].
int gabril_res_block_fwd(const void* x, const void* w3, const void* w1, void* h, void* bits_h, void* out,
                         void* bits_out, int N, int H, int W, int C, int Ch, cudaStream_t stream) {
  int rc = resolve_driver();
  if (rc) return rc;
  if (!x || !w3 || !w1 || !h || !bits_h || !out || !bits_out) return GABRIL_ERR_NULL;
  const long long M = (long long)N * H * W;
  if (C != 128 || Ch != 32 || N <= 0 || H < 1 || W < 1 || M >= (1ll << 31)) {
    gabril_set_error("gabril_res_block_fwd: built for 128 -> 32 -> 128 channels");
    return GABRIL_ERR_BAD_SHAPE;
  }
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(h) | reinterpret_cast<uintptr_t>(out)) & 31)
    return GABRIL_ERR_BAD_SHAPE;
  rb::Params p;
  memset(&p, 0, sizeof(p));
  rc = im2col_map(&p.mapX, x, N, H, W, C, -1, -1, -1, -1, 64, 128, 1);
  if (rc) return rc;
  rc = weight_map(&p.mapW3, w3, Ch, 9, C, 0, 64, Ch);
  if (rc) return rc;
  rc = weight_map(&p.mapW1, w1, C, 1, Ch, 0, 32, C);
  if (rc) return rc;
  p.x = (const bf16*)x; p.h = (bf16*)h; p.bits_h = (uint32_t*)bits_h; p.out = (bf16*)out; p.bits_out = (uint32_t*)bits_out;
  p.M = (int)M; p.HW = H * W; p.W = W; p.tiles = (int)((M + 127) / 128);
  auto magic = [](int d) {
    tc::FastDiv f{0u, 0u};
    if (d > 1) {
      uint32_t l = 0;
      while ((1ull << l) < (unsigned long long)d) ++l;
      f.mul = (uint32_t)((((1ull << l) - (unsigned long long)d) << 32) / (unsigned long long)d + 1ull);
      f.shr = l;
    }
    return f;
  };
  p.fd_hw = magic(p.HW); p.fd_w = magic(p.W);
  p.desc_a1_hi = desc_hi(16, 1024, 2);   // K-major, 128-byte rows
  p.desc_b1_hi = desc_hi(16, 1024, 2);
  p.desc_a2_hi = desc_hi(16, 512, 4);    // K-major, 64-byte rows
  p.desc_b2_hi = desc_hi(16, 512, 4);
  p.idesc1 = make_idesc(0, 0, 128, Ch);
  p.idesc2 = make_idesc(0, 0, 128, C);
  p.err_flag = g_err_flag;
  static int dbg = -1;
  if (dbg < 0) { const char* e = getenv("GABRIL_RB_DEBUG"); dbg = e ? atoi(e) : 0; }
  p.debug = dbg;
  return rb::launch(p, g_num_sms, stream);
}

int gabril_conv_wgrad_dbias_next(float* dbias, int mod) {
  tc::set_next_dbias(dbias, mod);
  return GABRIL_OK;
}

int gabril_poll_error(void) {
  if (!g_err_flag) return 0;
  int v = 0;
  if (cudaMemcpy(&v, g_err_flag, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess)
    return GABRIL_ERR_CUDA;
  if (v) cudaMemset(g_err_flag, 0, sizeof(int));
  return v;
}

// y[N,OH,OW,Cout] = act( conv(x[N,H,W,Cin], w[Cout][R*S][Cin]) + bias + resid ) * rowmask
// Cin == 0 mod 32, Cout in {32, 64, 128, 256...} (multiple of 32).  R = S = 1 with H = W = 1 and
// N = rows gives a plain linear layer (x[rows, Cin]).
static int conv_fwd_impl(const void* x, const void* w, const float* bias, const void* resid,
                         const void* gate, const float* rowmask, void* y, int N, int H, int W, int Cin,
                         int Cout, int R, int S, int stride, int pad, int relu, int out_f32_atomic,
                         int splits, cudaStream_t stream) {
  int rc = resolve_driver();
  if (rc) return rc;
  if (Cin % 32 || !fwd_cout_ok(Cout) || N <= 0) return GABRIL_ERR_BAD_SHAPE;
  const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
  const int kc = (Cin % 64 == 0) ? 64 : 32;
  const int block_n = std::min(Cout, 128);
  tc::Params p;
  zero_params(p);
  p.kind = tc::KIND_FWD;
  p.n_classes = 1;
  p.b_mn = 0;
  const long long M = (long long)N * OH * OW;
  const int m_sub = (out_f32_atomic || splits > 1) ? 1 : pick_m_sub(M, block_n, R * S * (Cin / kc));
  fwd_schedule(p, kc, block_n, 0, m_sub);
  const bool plain = (R == 1 && S == 1 && stride == 1 && pad == 0);
  p.a_im2col = plain ? 0 : 1;
  p.M_total[0] = (int)M;
  p.N_total = Cout;
  p.taps = R * S;
  p.S = S;
  p.chunks_per_tap = Cin / kc;
  p.kb_total = p.taps * p.chunks_per_tap;
  if (splits < 1) splits = 1;
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  splits = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.OW[0] = OW; p.OHW[0] = OH * OW; p.lower_w[0] = -pad; p.lower_h[0] = -pad;
  p.trav_stride_w = stride; p.trav_stride_h = stride;
  p.r_base[0] = 0; p.s_base[0] = 0; p.r_step = 1; p.s_step = 1; p.S_w = S;
  if (plain) {
    uint64_t dims[2] = {(uint64_t)Cin, (uint64_t)M};
    uint64_t strides[1] = {(uint64_t)Cin * 2};
    uint32_t box[2] = {(uint32_t)kc, (uint32_t)(128 * m_sub)};
    rc = tiled_map(&p.mapA[0], x, 2, dims, strides, box, kc * 2);
  } else {
    rc = im2col_map(&p.mapA[0], x, N, H, W, Cin, -pad, -pad, pad - (S - 1), pad - (R - 1), kc, 128 * m_sub,
                    stride);
  }
  if (rc) return rc;
  rc = weight_map(&p.mapB, w, Cout, R * S, Cin, 0, kc, block_n);
  if (rc) return rc;
  // out_f32_atomic: 0 = bf16 output with the fused epilogue; 1 = fp32 atomic accumulation into y (the
  // caller zeroes it); 2 = fp32 plain stores, split s into its own plane y[s][M][Cout] (deterministic:
  // gabril_bias_act_cast_planes sums the planes in order; y holds gabril_conv_fwd_splits(...) planes)
  p.out_mode = out_f32_atomic == 1 ? tc::OUT_F32_ATOMIC : out_f32_atomic == 2 ? tc::OUT_F32_STORE : tc::OUT_BF16;
  if (out_f32_atomic == 2) p.split_stride = M * Cout;
  p.out = y; p.ldo = Cout;
  p.bias = bias; p.resid = (const bf16*)resid; p.gate = (const bf16*)gate; p.rowmask = rowmask;
  p.relu = relu;
  p.err_flag = g_err_flag;
  size_t smem;
  const int occ = m_sub > 1 ? 1 : occ_setting();
  pick_stages(p, &smem, occ);
  const int mt[1] = {(int)((M + 128 * m_sub - 1) / (128 * m_sub))};
  finalize_schedule(p, mt, Cout / block_n, splits, m_sub * block_n, occ);
  return tc::launch(p, g_num_sms * occ, smem, stream);
}

// Number of split-K planes gabril_conv_fwd uses for a requested split count (planes of the
// out_f32_atomic = 2 mode).
int gabril_conv_fwd_splits(int Cin, int R, int S, int splits) {
  const int kc = (Cin % 64 == 0) ? 64 : 32;
  const int kb_total = R * S * (Cin / kc);
  if (splits < 1) splits = 1;
  const int per = (kb_total + splits - 1) / splits;
  return (kb_total + per - 1) / per;
}

int gabril_conv_fwd(const void* x, const void* w, const float* bias, const void* resid,
                    const float* rowmask, void* y, int N, int H, int W, int Cin, int Cout, int R,
                    int S, int stride, int pad, int relu, int out_f32_atomic, int splits,
                    cudaStream_t stream) {
  return conv_fwd_impl(x, w, bias, resid, nullptr, rowmask, y, N, H, W, Cin, Cout, R, S, stride, pad, relu,
                       out_f32_atomic, splits, stream);
}

static bool dgrad_stack_disabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("GABRIL_DGRAD_NO_STACK"); v = (e && e[0] == '1') ? 1 : 0; }
  return v == 1;
}

int gabril_conv_pack_wpairs(const void* w, void* w_pair, int Cout, int R, int S, int Cin, cudaStream_t stream) {
  if (!w || !w_pair) return GABRIL_ERR_NULL;
  if (Cout <= 0 || R <= 0 || S <= 0 || Cin <= 0) return GABRIL_ERR_BAD_SHAPE;
  if (Cin % 8 || ((reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(w_pair)) & 15)) return GABRIL_ERR_BAD_SHAPE;
  const long long total = 2ll * Cout * R * (S + 1) * (Cin / 8);
  if (total >= (1ll << 31)) return GABRIL_ERR_BAD_SHAPE;
  pack_wpairs_kernel<<<(unsigned)((total + 127) / 128), 128, 0, stream>>>((const uint4*)w, (uint4*)w_pair, Cout, R, S, Cin / 8);
  GABRIL_CHECK_LAUNCH();
  return GABRIL_OK;
}

// Two adjacent output pixels along N (see include/gabril_b200.h): a stride-1 convolution Cin -> Cout <= 64 runs as
// Cin -> 2*Cout over every second pixel with an R x (S+1) filter; output row (n, oh, ow') of 2*Cout columns IS the
// NHWC pixel pair (2ow', 2ow'+1), so the epilogue stores as usual.
static int conv_fwd_pairs_impl(const void* x, const void* w_pair, const float* bias, const float* rowmask, void* y,
                               int N, int H, int W, int Cin, int Cout, int R, int S, int pad, int relu,
                               cudaStream_t stream) {
  int rc = resolve_driver();
  if (rc) return rc;
  if (!x || !w_pair || !y) return GABRIL_ERR_NULL;
  if (Cin % 64 || (Cout != 32 && Cout != 64) || N <= 0 || (W & 1) || R != 2 * pad + 1 || S != 2 * pad + 1) {
    gabril_set_error("gabril_conv_fwd_pairs: stride-1 same convolution, Cin % 64 == 0, Cout 32 or 64, even W");
    return GABRIL_ERR_BAD_SHAPE;
  }
  const int kc = 64, block_n = 2 * Cout, S2 = S + 1, W2 = W / 2;
  const long long M = (long long)N * H * W2;
  if (M >= (1ll << 31)) return GABRIL_ERR_BAD_SHAPE;
  tc::Params p;
  zero_params(p);
  p.kind = tc::KIND_FWD;
  p.n_classes = 1;
  p.b_mn = 0;
  fwd_schedule(p, kc, block_n, 0, 1);
  p.a_im2col = 1;
  p.M_total[0] = (int)M;
  p.N_total = block_n;
  p.taps = R * S2;
  p.S = S2;
  p.chunks_per_tap = Cin / kc;
  p.kb_total = p.taps * p.chunks_per_tap;
  p.kb_per_split = p.kb_total;
  p.OW[0] = W2; p.OHW[0] = H * W2; p.lower_w[0] = -pad; p.lower_h[0] = -pad;
  p.trav_stride_w = 2; p.trav_stride_h = 1;
  p.r_base[0] = 0; p.s_base[0] = 0; p.r_step = 1; p.s_step = 1; p.S_w = S2;
  // the filter covers columns p - pad .. p + pad + 1: right padding pad + 1, width S + 1
  rc = im2col_map(&p.mapA[0], x, N, H, W, Cin, -pad, -pad, (pad + 1) - (S2 - 1), pad - (R - 1), kc, 128, 2, 1);
  if (rc) return rc;
  rc = weight_map(&p.mapB, w_pair, block_n, R * S2, Cin, 0, kc, block_n);
  if (rc) return rc;
  p.out_mode = tc::OUT_BF16;
  p.out = y; p.ldo = block_n;
  p.bias = bias; p.bias_mod = bias ? Cout : 0;
  p.rowmask = rowmask; p.pair_C = Cout;
  p.relu = relu;
  p.err_flag = g_err_flag;
  size_t smem;
  const int occ = occ_setting();
  pick_stages(p, &smem, occ);
  const int mt[1] = {(int)((M + 127) / 128)};
  finalize_schedule(p, mt, 1, 1, block_n, occ);
  return tc::launch(p, g_num_sms * occ, smem, stream);
}

int gabril_conv_fwd_pairs(const void* x, const void* w_pair, const float* bias, void* y, int N, int H, int W,
                          int Cin, int Cout, int R, int S, int pad, int relu, cudaStream_t stream) {
  return conv_fwd_pairs_impl(x, w_pair, bias, nullptr, y, N, H, W, Cin, Cout, R, S, pad, relu, stream);
}
// rowmask: one multiplier per output pixel (fp32 [N*H*W], 8-byte aligned), applied after bias / ReLU - the
// in-encoder gaze dropout of linear_models.py:194-199
int gabril_conv_fwd_pairs_masked(const void* x, const void* w_pair, const float* bias, const float* rowmask, void* y,
                                 int N, int H, int W, int Cin, int Cout, int R, int S, int pad, int relu,
                                 cudaStream_t stream) {
  if (rowmask && (reinterpret_cast<uintptr_t>(rowmask) & 7)) return GABRIL_ERR_BAD_SHAPE;
  return conv_fwd_pairs_impl(x, w_pair, bias, rowmask, y, N, H, W, Cin, Cout, R, S, pad, relu, stream);
}

// dx[N,H,W,Cin] = gate( conv_transpose(dy[N,OH,OW,Cout], w[Cout][R*S][Cin]) + addend )
//   gate: dx = gate_tensor > 0 ? dx : 0 (ReLU backward of the layer below), rowmask: per-pixel
//   multiplier (GMD).  stride 1 (any R,S,pad) and the k4/s2/p1 encoder convs are supported.
static int conv_dgrad_impl(const void* dy, const void* w, const float* bias, int relu, const void* addend,
                           const void* gate, const float* rowmask, void* dx, int N, int H, int W, int Cin,
                           int Cout, int R, int S, int stride, int pad, cudaStream_t stream) {
  int rc = resolve_driver();
  if (rc) return rc;
  if (!mn_ok(Cin) || Cout % 32 || N <= 0) return GABRIL_ERR_BAD_SHAPE;
  const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
  const int kc = (Cout % 64 == 0) ? 64 : 32;  // contraction over Cout
  const int block_n = std::min(Cin, 128);
  tc::Params p;
  zero_params(p);
  p.kind = tc::KIND_FWD;
  p.b_mn = 1;
  const int m_sub = pick_m_sub((long long)N * H * W / (stride * stride), block_n,
                               (stride == 1 ? R * S : 4) * (Cout / kc));
  const int tile_rows = 128 * m_sub;
  fwd_schedule(p, kc, block_n, 1, m_sub);
  p.N_total = Cin;
  p.chunks_per_tap = Cout / kc;
  p.trav_stride_w = 1; p.trav_stride_h = 1;
  p.S_w = S;
  const bool plain = (R == 1 && S == 1 && stride == 1 && pad == 0);
  if (stride == 1) {
    p.n_classes = 1;
    p.a_im2col = plain ? 0 : 1;
    const long long M = (long long)N * H * W;
    p.M_total[0] = (int)M;
    p.taps = R * S;
    p.S = S;
    p.OW[0] = W; p.OHW[0] = H * W;
    p.lower_w[0] = pad - (S - 1); p.lower_h[0] = pad - (R - 1);
    p.r_base[0] = R - 1; p.s_base[0] = S - 1; p.r_step = -1; p.s_step = -1;
    if (plain) {
      uint64_t dims[2] = {(uint64_t)Cout, (uint64_t)M};
      uint64_t strides[1] = {(uint64_t)Cout * 2};
      uint32_t box[2] = {(uint32_t)kc, (uint32_t)tile_rows};
      rc = tiled_map(&p.mapA[0], dy, 2, dims, strides, box, kc * 2);
    } else {
      rc = im2col_map(&p.mapA[0], dy, N, OH, OW, Cout, p.lower_w[0], p.lower_h[0], -pad, -pad, kc,
                      tile_rows, 1);
    }
    if (rc) return rc;
  } else if (stride == 2 && R == 4 && S == 4 && pad == 0 && Cin == 64 && kc == 64 && m_sub == 1 && !bias && !rowmask &&
             !gate && !dgrad_stack_disabled()) {   // (per-column / per-row epilogue terms are not class-aware)
    // the four parity classes stacked along N (tc::Params::cls_n): one launch geometry, N = 4 * Cin = 256
    p.cls_n = 4;
    p.n_classes = 1;
    p.a_im2col = 1;
    p.remap = 1; p.remap_sh = 2; p.remap_sw = 2;
    p.OHf = H; p.OWf = W;
    p.taps = 4;
    p.S = 2;
    p.r_step = -2; p.s_step = -2;
    fwd_schedule(p, kc, 4 * Cin, 1, 1);   // four 64-column weight boxes per stage, one per class
    p.N_total = 4 * Cin;
    for (int c = 0; c < 4; ++c) {
      const int ah = c >> 1, aw = c & 1;
      p.ah[c] = ah; p.aw[c] = aw;
      p.r_base[c] = (ah & 1) + 2; p.s_base[c] = (aw & 1) + 2;
    }
    const int Hs = (H + 1) / 2, Ws = (W + 1) / 2;   // base grid = class (0, 0), the largest
    p.Hs[0] = Hs; p.Ws[0] = Ws;
    p.M_total[0] = N * Hs * Ws;
    p.OW[0] = Ws; p.OHW[0] = Hs * Ws;
    p.lower_w[0] = -1; p.lower_h[0] = -1;
    rc = im2col_map(&p.mapA[0], dy, N, OH, OW, Cout, -1, -1, -1 + Ws - OW, -1 + Hs - OH, kc, 128, 1);
    if (rc) return rc;
    p.kb_total = p.taps * p.chunks_per_tap;
    p.kb_per_split = p.kb_total;
    rc = weight_map(&p.mapB, w, Cout, R * S, Cin, 1, kc, Cin);
    if (rc) return rc;
    p.out_mode = tc::OUT_BF16;
    p.out = dx; p.ldo = Cin;
    p.bias = bias; p.resid = (const bf16*)addend; p.gate = (const bf16*)gate; p.rowmask = rowmask;
    p.relu = relu;
    p.err_flag = g_err_flag;
    size_t smem1;
    pick_stages(p, &smem1, 1);
    const int mt1[1] = {(p.M_total[0] + 127) / 128};
    finalize_schedule(p, mt1, 1, 1, 4 * Cin, 1);
    return tc::launch(p, g_num_sms, smem1, stream);
  } else if (stride == 2 && R == 4 && S == 4 && (pad == 1 || pad == 0)) {
    // Input pixel ih = 2*jh + a_h receives exactly two filter rows: kh = b_h + 2 - 2r' with
    // b_h = (a_h + pad) & 1, read from dy row jh + c_h - 1 + r', c_h = (a_h + pad - b_h) / 2
    // (same along w).  Four stride-1 sub-convolutions, one per input parity class.
    p.n_classes = 4;
    p.a_im2col = 1;
    p.remap = 1; p.remap_sh = 2; p.remap_sw = 2;
    p.OHf = H; p.OWf = W;
    p.taps = 4;
    p.S = 2;
    p.r_step = -2; p.s_step = -2;
    for (int c = 0; c < 4; ++c) {
      const int ah = c >> 1, aw = c & 1;
      const int bh = (ah + pad) & 1, bw_ = (aw + pad) & 1;
      const int ch = (ah + pad - bh) / 2, cw = (aw + pad - bw_) / 2;
      const int Hs = (H - ah + 1) / 2, Ws = (W - aw + 1) / 2;
      p.ah[c] = ah; p.aw[c] = aw; p.Hs[c] = Hs; p.Ws[c] = Ws;
      p.M_total[c] = N * Hs * Ws;
      p.OW[c] = Ws; p.OHW[c] = Hs * Ws;
      p.lower_w[c] = cw - 1; p.lower_h[c] = ch - 1;
      p.r_base[c] = bh + 2; p.s_base[c] = bw_ + 2;
      const int upper_w = p.lower_w[c] + Ws - OW, upper_h = p.lower_h[c] + Hs - OH;
      rc = im2col_map(&p.mapA[c], dy, N, OH, OW, Cout, p.lower_w[c], p.lower_h[c], upper_w,
                      upper_h, kc, tile_rows, 1);
      if (rc) return rc;
    }
  } else {
    return GABRIL_ERR_BAD_SHAPE;
  }
  p.kb_total = p.taps * p.chunks_per_tap;
  p.kb_per_split = p.kb_total;
  rc = weight_map(&p.mapB, w, Cout, R * S, Cin, 1, kc, block_n);
  if (rc) return rc;
  p.out_mode = tc::OUT_BF16;
  p.out = dx; p.ldo = Cin;
  p.bias = bias; p.resid = (const bf16*)addend; p.gate = (const bf16*)gate; p.rowmask = rowmask;
  p.relu = relu;
  p.err_flag = g_err_flag;
  size_t smem;
  const int occ = m_sub > 1 ? 1 : occ_setting();
  pick_stages(p, &smem, occ);
  int mt[4] = {0, 0, 0, 0};
  for (int c = 0; c < p.n_classes; ++c) mt[c] = (p.M_total[c] + tile_rows - 1) / tile_rows;
  finalize_schedule(p, mt, Cin / block_n, 1, m_sub * block_n, occ);
  return tc::launch(p, g_num_sms * occ, smem, stream);
}

int gabril_conv_dgrad(const void* dy, const void* w, const void* addend, const void* gate,
                      const float* rowmask, void* dx, int N, int H, int W, int Cin, int Cout,
                      int R, int S, int stride, int pad, cudaStream_t stream) {
  return conv_dgrad_impl(dy, w, nullptr, 0, addend, gate, rowmask, dx, N, H, W, Cin, Cout, R, S, stride, pad,
                         stream);
}

// nn.ConvTranspose2d (Decoder, linear_models.py:236-263) on the same kernels.  A transposed
// convolution with weight Wt[Cin_t][Cout_t][R][S] is the data gradient of the convolution
// Cout_t -> Cin_t whose internal weight is w[Cin_t][R*S][Cout_t] (= the "conv" layout of Wt):
//   forward   y[N,H,W,Cout_t]  = act(dgrad(x[N,OH,OW,Cin_t]) + bias)   (H, W: output size, so
//             output_padding needs no special case: it only changes which H, W the caller asks for)
//   backward  dx[N,OH,OW,Cin_t] = gate(conv_fwd(dy[N,H,W,Cout_t]))
//   weights   dw = gabril_conv_wgrad(dy := x, x := dy)
int gabril_convt_fwd(const void* x, const void* w, const float* bias, void* y, int N, int H, int W,
                     int Cout_t, int Cin_t, int R, int S, int stride, int pad, int relu,
                     cudaStream_t stream) {
  return conv_dgrad_impl(x, w, bias, relu, nullptr, nullptr, nullptr, y, N, H, W, Cout_t, Cin_t, R, S, stride,
                         pad, stream);
}
int gabril_convt_dgrad(const void* dy, const void* w, const void* gate, void* dx, int N, int H, int W,
                       int Cout_t, int Cin_t, int R, int S, int stride, int pad, cudaStream_t stream) {
  return conv_fwd_impl(dy, w, nullptr, nullptr, gate, nullptr, dx, N, H, W, Cout_t, Cin_t, R, S, stride, pad, 0,
                       0, 1, stream);
}

// Narrow layers (Cout <= 64): with A = dY^T the UMMA M dimension would be 25-50 % filled, so the
// operands are exchanged: A = im2col(x) of one tap (M = 128 input channels), B = dY (N = Cout), one
// accumulator of Cout columns per tap (up to 512 / Cout taps per CTA share one dY tile).
static bool wgrad_swap_disabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("GABRIL_WGRAD_NO_SWAP"); v = (e && e[0] == '1') ? 1 : 0; }
  return v == 1;
}

static int conv_wgrad_swapped(const void* dy, const void* x, float* dw, int N, int H, int W, int Cin,
                              int Cout, int R, int S, int stride, int pad, int direct_store,
                              cudaStream_t stream) {
  const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
  const long long P = (long long)N * OH * OW;
  const int T = R * S, KP = 64;
  tc::Params p;
  zero_params(p);
  p.kind = tc::KIND_WGRAD;
  p.n_classes = 1;
  p.swap_ab = 1;
  const bool plain = (R == 1 && S == 1 && stride == 1 && pad == 0);
  p.a_im2col = plain ? 0 : 1;
  p.kc = KP;
  // A = x^T of one tap: M = 128 input channels = two 64-channel boxes of KP pixels
  const int aw = 64, a_sw = aw * 2;
  p.a_box_w = aw;
  p.a_boxes = 2;
  p.a_box_bytes = KP * a_sw;
  const uint32_t a_tap_bytes = p.a_boxes * p.a_box_bytes;
  p.desc_a_hi = desc_hi(p.a_box_bytes, 8 * a_sw, layout_type_for_bytes(a_sw));
  // B = dY: N = Cout columns in one box
  const int block_n = Cout, bw = Cout, b_sw = bw * 2;
  p.block_n = block_n;
  p.b_box_w = bw;
  p.b_boxes = 1;
  p.b_box_bytes = KP * b_sw;
  p.b_stage_bytes = round_up(p.b_box_bytes, 1024);
  p.desc_b_hi = desc_hi(p.b_box_bytes, 8 * b_sw, layout_type_for_bytes(b_sw));
  // One tap per CTA: sharing the dY tile between several taps (TMEM would hold 512 / Cout of them)
  // makes the smem stages so large that only two fit, and the shallow pipeline costs more than the
  // re-read of the small dY tile saves (measured at batch 1024: conv_6 0.48 ms with 5 taps per CTA,
  // 0.42 ms with 1; res3x3 0.20 -> 0.15 ms).
  int n_acc = 1;
  const int groups = (T + n_acc - 1) / n_acc;
  p.n_acc = n_acc;
  p.taps = T;
  p.tap_groups = groups;
  p.S = S;
  p.a_stage_bytes = round_up(n_acc * a_tap_bytes, 1024);
  p.mma_k = KP / 16;
  p.mma_groups = n_acc;
  p.g_taps = 1;
  p.a_kstep = (uint32_t)((16 * a_sw) >> 4);
  p.b_kstep = (uint32_t)((16 * b_sw) >> 4);
  p.a_gstep = a_tap_bytes >> 4;
  p.b_gstep = 0;
  p.d_gstep = (uint32_t)block_n;
  p.idesc = make_idesc(1, 1, 128, block_n);
  p.M_total[0] = Cin;
  p.N_total = Cout;
  p.kb_total = (int)((P + KP - 1) / KP);
  p.OW[0] = OW; p.OHW[0] = OH * OW; p.lower_w[0] = -pad; p.lower_h[0] = -pad;
  p.trav_stride_w = stride; p.trav_stride_h = stride;
  int rc;
  {
    uint64_t dims[2] = {(uint64_t)Cout, (uint64_t)P};
    uint64_t strides[1] = {(uint64_t)Cout * 2};
    uint32_t box[2] = {(uint32_t)bw, (uint32_t)KP};
    rc = tiled_map(&p.mapB, dy, 2, dims, strides, box, b_sw);
    if (rc) return rc;
  }
  if (plain) {
    uint64_t dims[2] = {(uint64_t)Cin, (uint64_t)P};
    uint64_t strides[1] = {(uint64_t)Cin * 2};
    uint32_t box[2] = {(uint32_t)aw, (uint32_t)KP};
    rc = tiled_map(&p.mapA[0], x, 2, dims, strides, box, a_sw);
  } else {
    rc = im2col_map(&p.mapA[0], x, N, H, W, Cin, -pad, -pad, pad - (S - 1), pad - (R - 1), aw, KP,
                    stride);
  }
  if (rc) return rc;
  const int m_tiles = Cin / 128, n_tiles = 1;
  const int occ = (n_acc * block_n <= 256 && 2 * (p.a_stage_bytes + p.b_stage_bytes) <= 100 * 1024) ? 2 : 1;
  int splits = 1;
  if (!direct_store) {
    const int base = m_tiles * n_tiles * p.tap_groups;
    splits = std::max(1, std::min(p.kb_total, (occ * g_num_sms) / base));
  }
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  splits = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.out_mode = direct_store ? tc::OUT_F32_STORE : tc::OUT_F32_ATOMIC;
  p.out = dw;
  p.ldo = (long long)T * Cin;
  p.tap_stride = Cin;
  p.err_flag = g_err_flag;
  size_t smem;
  pick_stages(p, &smem, occ);
  const int mt[1] = {m_tiles};
  finalize_schedule(p, mt, n_tiles * p.tap_groups, splits, n_acc * block_n, occ);
  p.acc_bufs = 1;
  p.tmem_cols = (uint32_t)p.acc_stride;
  return tc::launch(p, 1 << 30, smem, stream);
}

// Narrow stride-1 "same" convolutions (conv_6: 128 -> 64, 5x5; Residual 3x3: 128 -> 32) with the shift on dY
// (tc::Params::shift_dy): the activation tile of a k-block is loaded once for all taps and each tap's
// operand is an im2col tile of the narrow dY, so a CTA holds several taps (N = taps x Cout per MMA) without
// the 16 KB-per-tap activation tiles that made multi-tap CTAs lose in conv_wgrad_swapped.  Shared-memory fill
// per k-block of 64 pixels: conv_6 16 + 5 x 8 KB for 5 taps (was 5 x 24 KB), res3x3 16 + 9 x 4 KB for all 9
// taps (was 9 x 20 KB) - these launches were bound by the L2 -> SM feed.
static bool wgrad_shift_disabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("GABRIL_WGRAD_NO_SHIFT"); v = (e && e[0] == '1') ? 1 : 0; }
  return v == 1;
}

static int conv_wgrad_shifted(const void* dy, const void* x, float* dw, int N, int H, int W, int Cin,
                              int Cout, int R, int S, int pad, int direct_store, cudaStream_t stream) {
  const long long P = (long long)N * H * W;   // contraction over the INPUT pixels (= output pixels here)
  const int T = R * S, KP = 64;
  tc::Params p;
  zero_params(p);
  p.kind = tc::KIND_WGRAD;
  p.n_classes = 1;
  p.swap_ab = 1;     // TMEM lane = cin, column = (tap, cout): the exchanged form's epilogue
  p.shift_dy = 1;
  p.center_tap = pad * S + pad;
  p.kc = KP;
  // A = x rows of the k-block: M = 128 input channels = two 64-channel boxes of KP pixels
  const int aw = 64, a_sw = aw * 2;
  p.a_box_w = aw;
  p.a_boxes = 2;
  p.a_box_bytes = KP * a_sw;
  p.a_stage_bytes = round_up(p.a_boxes * p.a_box_bytes, 1024);
  p.desc_a_hi = desc_hi(p.a_box_bytes, 8 * a_sw, layout_type_for_bytes(a_sw));
  // B = dY of one tap: KP pixels x Cout channels, one box; taps are b_box_bytes apart (whole swizzle atoms),
  // which is the MN-major leading-dimension stride, so g_taps taps are the N dimension of one instruction
  const int block_n = Cout, b_sw = Cout * 2;
  p.block_n = block_n;
  p.b_box_w = Cout;
  p.b_boxes = 1;
  p.b_box_bytes = KP * b_sw;
  p.desc_b_hi = desc_hi(p.b_box_bytes, 8 * b_sw, layout_type_for_bytes(b_sw));
  // taps per CTA: TMEM columns, a stage of at most ~56 KB (three or four of them in flight), balanced groups
  int n_max = std::min(T, std::min(512 / Cout, (int)((56u * 1024u - p.a_stage_bytes) / p.b_box_bytes)));
  n_max = std::max(1, n_max);
  const int groups = (T + n_max - 1) / n_max;
  const int n_acc = (T + groups - 1) / groups;
  int G = std::max(1, std::min(n_acc, 256 / Cout));   // taps per MMA (N <= 256), balanced
  const int ng = (n_acc + G - 1) / G;
  G = (n_acc + ng - 1) / ng;
  p.n_acc = n_acc;
  p.taps = T;
  p.tap_groups = groups;
  p.S = S;
  p.b_stage_bytes = round_up((uint32_t)n_acc * p.b_box_bytes, 1024);
  p.mma_k = KP / 16;
  p.mma_groups = ng;
  p.g_taps = G;
  p.a_kstep = (uint32_t)((16 * a_sw) >> 4);
  p.b_kstep = (uint32_t)((16 * b_sw) >> 4);
  p.a_gstep = 0;
  p.b_gstep = (uint32_t)((G * p.b_box_bytes) >> 4);
  p.d_gstep = (uint32_t)(G * block_n);
  p.idesc = make_idesc(1, 1, 128, G * block_n);
  p.M_total[0] = Cin;
  p.N_total = Cout;
  p.kb_total = (int)((P + KP - 1) / KP);
  // k-block start pixel -> (n, ih, iw) of x; dY is read at (ih, iw) + lower corner + traversal offset
  p.OW[0] = W; p.OHW[0] = H * W; p.lower_w[0] = pad - (S - 1); p.lower_h[0] = pad - (R - 1);
  p.trav_stride_w = 1; p.trav_stride_h = 1;
  int rc;
  {
    uint64_t dims[2] = {(uint64_t)Cin, (uint64_t)P};
    uint64_t strides[1] = {(uint64_t)Cin * 2};
    uint32_t box[2] = {(uint32_t)aw, (uint32_t)KP};
    rc = tiled_map(&p.mapA[0], x, 2, dims, strides, box, a_sw);
    if (rc) return rc;
  }
  // one base position per x pixel: the traversal box has the tensor's own extent (upper corner = lower corner)
  rc = im2col_map(&p.mapB, dy, N, H, W, Cout, p.lower_w[0], p.lower_h[0], p.lower_w[0], p.lower_h[0], Cout, KP, 1);
  if (rc) return rc;
  const int m_tiles = Cin / 128;
  const int occ = 1;   // 288 / 320 accumulator columns round up to all 512
  int splits = 1;
  if (!direct_store) splits = std::max(1, std::min(p.kb_total, g_num_sms / (m_tiles * groups)));
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  splits = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.out_mode = direct_store ? tc::OUT_F32_STORE : tc::OUT_F32_ATOMIC;
  p.out = dw;
  p.ldo = (long long)T * Cin;
  p.tap_stride = Cin;
  p.err_flag = g_err_flag;
  size_t smem;
  pick_stages(p, &smem, occ);
  const int mt[1] = {m_tiles};
  finalize_schedule(p, mt, groups, splits, n_acc * block_n, occ);
  p.acc_bufs = 1;
  p.tmem_cols = (uint32_t)p.acc_stride;
  return tc::launch(p, 1 << 30, smem, stream);
}

// dw[Cout][R*S][Cin] (fp32) += sum_pixels dy[p, Cout]^T * im2col(x)[p, (tap, Cin)].
// The caller zeroes dw (it lives in the flat gradient arena).  Also serves linear layers
// (R = S = 1, H = W = 1, N = rows): dw[Cout][Cin] += dy[rows,Cout]^T x[rows,Cin].
int gabril_conv_wgrad(const void* dy, const void* x, float* dw, int N, int H, int W, int Cin,
                      int Cout, int R, int S, int stride, int pad, int direct_store,
                      cudaStream_t stream) {
  int rc = resolve_driver();
  if (rc) return rc;
  if (!mn_ok(Cin) || Cout % 32 || N <= 0) return GABRIL_ERR_BAD_SHAPE;
  const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
  const long long P = (long long)N * OH * OW;  // contraction length (output pixels)
  const int T = R * S;
  const int KP = 64;
  if (Cout <= 64 && Cin == 128 && stride == 1 && T > 1 && R == 2 * pad + 1 && S == 2 * pad + 1 && P < (1ll << 31) &&
      !wgrad_shift_disabled())
    return conv_wgrad_shifted(dy, x, dw, N, H, W, Cin, Cout, R, S, pad, direct_store, stream);
  if (Cout <= 64 && Cin % 128 == 0 && !wgrad_swap_disabled())
    return conv_wgrad_swapped(dy, x, dw, N, H, W, Cin, Cout, R, S, stride, pad, direct_store, stream);
  tc::Params p;
  zero_params(p);
  p.kind = tc::KIND_WGRAD;
  p.n_classes = 1;
  const bool plain = (R == 1 && S == 1 && stride == 1 && pad == 0);
  p.a_im2col = plain ? 0 : 1;
  p.kc = KP;
  // A = dy^T: M = Cout rows padded to 128 with out-of-range (zero-filled) boxes.
  const int aw = std::min(64, Cout), a_sw = aw * 2;
  p.a_box_w = aw;
  p.a_boxes = 128 / aw;
  p.a_box_bytes = KP * a_sw;
  p.a_stage_bytes = round_up(p.a_boxes * p.a_box_bytes, 1024);
  p.desc_a_hi = desc_hi(p.a_box_bytes, 8 * a_sw, layout_type_for_bytes(a_sw));
  // B = x (per tap): N = block_n channels.
  const int block_n = std::min(Cin, 128);
  const int bw = std::min(64, block_n), b_sw = bw * 2;
  p.block_n = block_n;
  p.b_box_w = bw;
  p.b_boxes = block_n / bw;
  p.b_box_bytes = KP * b_sw;
  p.desc_b_hi = desc_hi(p.b_box_bytes, 8 * b_sw, layout_type_for_bytes(b_sw));
  const uint32_t b_tap_bytes = p.b_boxes * p.b_box_bytes;
  // accumulators: the taps of ONE N = 256 instruction per CTA.  More taps per CTA (TMEM has room
  // for 512 columns) re-use the dY tile further but leave only two 80 KB smem stages; 48 KB stages
  // four deep are faster (batch 1024: conv_5 0.63 -> 0.50 ms, conv_4 0.31 -> 0.22 ms).
  int n_acc = std::min(T, 256 / block_n);
  while (n_acc > 1 && 2 * (p.a_stage_bytes + n_acc * b_tap_bytes) > 200 * 1024) --n_acc;
  // One UMMA covers G consecutive taps: their MN-major B tiles sit b_tap_bytes apart, which is a
  // whole number of swizzle atoms, so N = G * block_n (<= 256) is a single instruction.
  const int G = std::max(1, std::min(n_acc, 256 / block_n));
  if (n_acc > G) n_acc = n_acc / G * G;
  p.n_acc = n_acc;
  p.taps = T;
  p.tap_groups = (T + n_acc - 1) / n_acc;
  p.S = S;
  p.b_stage_bytes = round_up(n_acc * b_tap_bytes, 1024);
  p.mma_k = KP / 16;
  p.mma_groups = (n_acc + G - 1) / G;
  p.g_taps = G;
  p.a_kstep = (uint32_t)((16 * a_sw) >> 4);
  p.b_kstep = (uint32_t)((16 * b_sw) >> 4);
  p.b_gstep = (uint32_t)((G * b_tap_bytes) >> 4);
  p.d_gstep = (uint32_t)(G * block_n);
  p.idesc = make_idesc(1, 1, 128, G * block_n);
  p.M_total[0] = Cout;
  p.N_total = Cin;
  p.kb_total = (int)((P + KP - 1) / KP);
  p.OW[0] = OW; p.OHW[0] = OH * OW; p.lower_w[0] = -pad; p.lower_h[0] = -pad;
  p.trav_stride_w = stride; p.trav_stride_h = stride;
  {
    uint64_t dims[2] = {(uint64_t)Cout, (uint64_t)P};
    uint64_t strides[1] = {(uint64_t)Cout * 2};
    uint32_t box[2] = {(uint32_t)aw, (uint32_t)KP};
    rc = tiled_map(&p.mapA[0], dy, 2, dims, strides, box, a_sw);
    if (rc) return rc;
  }
  if (plain) {
    uint64_t dims[2] = {(uint64_t)Cin, (uint64_t)P};
    uint64_t strides[1] = {(uint64_t)Cin * 2};
    uint32_t box[2] = {(uint32_t)bw, (uint32_t)KP};
    rc = tiled_map(&p.mapB, x, 2, dims, strides, box, b_sw);
  } else {
    rc = im2col_map(&p.mapB, x, N, H, W, Cin, -pad, -pad, pad - (S - 1), pad - (R - 1), bw, KP,
                    stride);
  }
  if (rc) return rc;
  const int m_tiles = (Cout + 127) / 128, n_tiles = Cin / block_n;
  // every tile is its own CTA here (long split-K loops, up to 512 TMEM columns)
  const int occ = (n_acc * block_n <= 256 && 2 * (p.a_stage_bytes + p.b_stage_bytes) <= 100 * 1024) ? 2 : 1;
  int splits = 1;
  if (!direct_store) {
    // whole waves only: the CTA count never exceeds what is resident at once
    const int base = m_tiles * n_tiles * p.tap_groups;
    splits = std::max(1, std::min(p.kb_total, (occ * g_num_sms) / base));
  }
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  splits = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.out_mode = direct_store ? tc::OUT_F32_STORE : tc::OUT_F32_ATOMIC;
  p.out = dw;
  p.ldo = (long long)T * Cin;
  p.tap_stride = Cin;
  p.err_flag = g_err_flag;
  size_t smem;
  pick_stages(p, &smem, occ);
  const int mt[1] = {m_tiles};
  finalize_schedule(p, mt, n_tiles * p.tap_groups, splits, n_acc * block_n, occ);
  p.acc_bufs = 1;
  p.tmem_cols = (uint32_t)p.acc_stride;
  return tc::launch(p, 1 << 30, smem, stream);
}

// ---------------------------------------------------------------------------------------------
// Flat-slab convolutions over zero-bordered tensors (see Params::a_flat).
// ---------------------------------------------------------------------------------------------
// out(q) = sum_{r,s} x_flat[q + r*Wg + s] . W(r, s) for every flat pixel q of x[N, Hg, Wg, Cx]
// (stride 1; the zero border of x supplies the padding).  Pixel q = (n, i, j) is stored at
// y[(n*o_rows + i + o_off_h)*o_pitch + j + o_off_w][:] when i < valid_h and j < valid_w
// (valid_h == 0: every pixel is stored at its own index).
//   transpose_w == 0 (forward):  w[Cy][R*S][Cx], tap (r, s)           -> Cy output channels
//   transpose_w == 1 (dgrad):    w[Cx][R*S][Cy], tap (R-1-r, S-1-s)   -> Cy output channels,
//                                contraction over w's leading dimension
// epilogue: + bias[Cy], + resid (indexed like y), ReLU, gate (y = gate > 0 ? y : 0, indexed like y).
int gabril_conv_flat(const void* x, const void* w, const float* bias, const void* resid,
                     const void* gate, void* y, int N, int Hg, int Wg, int Cx, int Cy, int R, int S,
                     int valid_h, int valid_w, int o_rows, int o_pitch, int o_off_h, int o_off_w,
                     int relu, int transpose_w, cudaStream_t stream) {
  return gabril_conv_flat_masked(x, w, bias, resid, gate, nullptr, y, N, Hg, Wg, Cx, Cy, R, S, valid_h,
                                 valid_w, o_rows, o_pitch, o_off_h, o_off_w, relu, transpose_w, stream);
}

// rowmask: per OUTPUT-STORAGE pixel multiplier (index = pixel index of y after the remap), applied
// after bias / ReLU - the in-encoder gaze dropout of linear_models.py:194-199.
// x_pitch_c: channel pitch between consecutive flat rows of x (0 = Cx).  A pitch below Cx makes rows
// overlap: row q = cells q and q + 1 of a tensor with x_pitch_c channels per cell (conv1_s4d.cu).
// quad_C > 0: the output columns are four quadrants (qh, qw) of quad_C channels, i.e. row (n, sh, sw) holds
// the pixels (2sh+qh-1, 2sw+qw-1) of a quad_H x quad_W image; pixels outside it store zeros, the bias
// vector has quad_C entries, and rowmask has one entry per (row, quadrant).
static int conv_flat_impl(const void* x, const void* w, const float* bias, const void* resid,
                          const void* gate, const float* rowmask, void* y, int N, int Hg, int Wg,
                          int Cx, int Cy, int R, int S, int valid_h, int valid_w, int o_rows,
                          int o_pitch, int o_off_h, int o_off_w, int relu, int transpose_w,
                          int x_pitch_c, int quad_C, int quad_H, int quad_W, cudaStream_t stream) {
  int rc = resolve_driver();
  if (rc) return rc;
  if (Cx % 64 || N <= 0 || R < 1 || S < 1 || S > 8) return GABRIL_ERR_BAD_SHAPE;
  if (transpose_w ? !mn_ok(Cy) : !fwd_cout_ok(Cy)) return GABRIL_ERR_BAD_SHAPE;
  const int kc = 64;
  const int block_n = std::min(Cy, 128);
  tc::Params p;
  zero_params(p);
  p.kind = tc::KIND_FWD;
  p.n_classes = 1;
  p.b_mn = transpose_w ? 1 : 0;
  fwd_schedule(p, kc, block_n, p.b_mn);
  p.a_flat = 1;
  p.flat_pitch = Wg; p.flat_R = R; p.flat_S = S;
  const long long M = (long long)N * Hg * Wg;
  if (M >= (1ll << 31)) return GABRIL_ERR_BAD_SHAPE;
  p.M_total[0] = (int)M;
  p.N_total = Cy;
  p.taps = R * S; p.S = S; p.S_w = S;
  p.chunks_per_tap = Cx / kc;
  p.kb_total = R * p.chunks_per_tap;  // one k-block = one (filter row, channel chunk) slab
  p.kb_per_split = p.kb_total;
  if (transpose_w) { p.r_base[0] = R - 1; p.s_base[0] = S - 1; p.r_step = -1; p.s_step = -1; }
  else { p.r_base[0] = 0; p.s_base[0] = 0; p.r_step = 1; p.s_step = 1; }
  // A: the activation as [rows][Cx]; a slab = 128 + S - 1 rows of one 64-channel chunk
  const int slab_rows = 128 + S - 1;
  {
    uint64_t dims[2] = {(uint64_t)Cx, (uint64_t)M};
    uint64_t strides[1] = {(uint64_t)(x_pitch_c > 0 ? x_pitch_c : Cx) * 2};
    uint32_t box[2] = {(uint32_t)kc, (uint32_t)slab_rows};
    rc = tiled_map(&p.mapA[0], x, 2, dims, strides, box, kc * 2);
    if (rc) return rc;
  }
  p.slab_tx_bytes = (uint32_t)slab_rows * 128u;
  p.a_box_bytes = p.slab_tx_bytes;
  p.a_stage_bytes = round_up(p.slab_tx_bytes, 1024);
  // B: weight tap tiles
  if (!transpose_w) rc = weight_map(&p.mapB, w, Cy, R * S, Cx, 0, kc, block_n);
  else rc = weight_map(&p.mapB, w, Cx, R * S, Cy, 1, kc, block_n);
  if (rc) return rc;
  p.b_tap_bytes = (uint32_t)p.b_boxes * p.b_box_bytes;
  const int n_tiles = Cy / block_n;
  const uint32_t resident_bytes = (uint32_t)(R * S * p.chunks_per_tap) * p.b_tap_bytes;
  p.b_resident = (n_tiles == 1 && resident_bytes <= 96u * 1024u) ? 1 : 0;
  p.b_stage_bytes = p.b_resident ? 0u : (uint32_t)S * p.b_tap_bytes;
  const uint32_t stage = p.a_stage_bytes + p.b_stage_bytes;
  const uint32_t budget = 224u * 1024u - 2048u - (p.b_resident ? resident_bytes : 0u);
  int st = (int)(budget / stage);
  st = std::min(st, tc::kMaxStages);
  if (st < 2) return GABRIL_ERR_BAD_SHAPE;
  p.stages = st;
  p.resident_off = (uint32_t)st * stage;
  const size_t smem = (size_t)st * stage + (p.b_resident ? resident_bytes : 0u) + 1024;
  // epilogue
  p.out_mode = tc::OUT_BF16;
  p.out = y; p.ldo = Cy;
  p.bias = bias; p.resid = (const bf16*)resid; p.gate = (const bf16*)gate; p.rowmask = rowmask;
  p.relu = relu;
  if (valid_h > 0) {
    p.remap = 1; p.remap_sh = 1; p.remap_sw = 1;
    p.Hs[0] = Hg; p.Ws[0] = Wg; p.OHf = o_rows; p.OWf = o_pitch; p.ah[0] = o_off_h; p.aw[0] = o_off_w;
    p.valid_h = valid_h; p.valid_w = valid_w;
  }
  if (quad_C > 0) {
    if (quad_C % 32 || 4 * quad_C != Cy || valid_h > 0) return GABRIL_ERR_BAD_SHAPE;
    p.quad_C = quad_C;
    p.border_h = Hg; p.border_w = Wg; p.s2d_H = quad_H; p.s2d_W = quad_W;
    p.bias_mod = quad_C;
  }
  p.err_flag = g_err_flag;
  const int mt[1] = {(int)((M + 127) / 128)};
  finalize_schedule(p, mt, n_tiles, 1, block_n, 1);
  return tc::launch(p, g_num_sms, smem, stream);
}

int gabril_conv_flat_masked(const void* x, const void* w, const float* bias, const void* resid,
                            const void* gate, const float* rowmask, void* y, int N, int Hg, int Wg,
                            int Cx, int Cy, int R, int S, int valid_h, int valid_w, int o_rows,
                            int o_pitch, int o_off_h, int o_off_w, int relu, int transpose_w,
                            cudaStream_t stream) {
  return conv_flat_impl(x, w, bias, resid, gate, rowmask, y, N, Hg, Wg, Cx, Cy, R, S, valid_h, valid_w, o_rows,
                        o_pitch, o_off_h, o_off_w, relu, transpose_w, 0, 0, 0, 0, stream);
}

// conv_1 + ReLU over the 4x4 space-to-depth frame image (conv1_s4d.cu): x4 [N*SH*SW (+1), Cp] bf16,
// w4 [4*C1][2][2*Cp] bf16 (gabril_conv1_s4d_expand), bias fp32 [C1], rowmask fp32 [N*SH*SW*4] or NULL
// (in-encoder gaze dropout, one entry per conv_1 pixel in space-to-depth order) -> y [N, SH, SW, 4*C1].
int gabril_conv1_s4d_fwd(const void* x4, const void* w4, const float* bias, const float* rowmask, void* y,
                         int N, int SH, int SW, int Cp, int C1, int img_h, int img_w, int relu,
                         cudaStream_t stream) {
  return conv_flat_impl(x4, w4, bias, nullptr, nullptr, rowmask, y, N, SH, SW, 2 * Cp, 4 * C1, 2, 1, 0, 0, SH,
                        SW, 0, 0, relu, 0, Cp, C1, img_h, img_w, stream);
}

// dw[Cy][R*S][Cx] (fp32) += sum_q dy_flat[q + dy_row_off, Cy]^T . x_flat[q + r*Wg + s, Cx]
// over every flat pixel q of the [N, Hg, Wg] grid; dy and x are zero-bordered tensors with that
// same grid (rows past the end read as zero).  The caller zeroes dw.
static int conv_wgrad_flat_impl(const void* dy, const void* x, float* dw, int N, int Hg, int Wg, int Cx,
                                int Cy, int R, int S, int dy_row_off, int x_pitch_c, cudaStream_t stream) {
  int rc = resolve_driver();
  if (rc) return rc;
  if (!mn_ok(Cx) || Cx < 64 || Cy % 32 || N <= 0 || S > 8) return GABRIL_ERR_BAD_SHAPE;
  const long long P = (long long)N * Hg * Wg;
  if (P >= (1ll << 31)) return GABRIL_ERR_BAD_SHAPE;
  // conv_2 on the space-to-depth grid (2x2, dY one pixel down-right of its flat row, zero-bordered dY): the same
  // sum as the weight gradient with the shift on dY of a 2x2 convolution with "pad" 1 - one N = 4 * Cy = 256 MMA
  // per K step instead of four N = 64 ones, x staged once (the flat form wraps where im2col zero-fills; on a
  // zero-bordered dY both read zeros)
  if (R == 2 && S == 2 && dy_row_off == Wg + 1 && Cx == 128 && Cy == 64 && x_pitch_c == 0 && !wgrad_shift_disabled())
    return conv_wgrad_shifted(dy, x, dw, N, Hg, Wg, Cx, Cy, 2, 2, 1, 0, stream);
  const int KP = 64, T = R * S;
  tc::Params p;
  zero_params(p);
  p.kind = tc::KIND_WGRAD;
  p.n_classes = 1;
  p.a_flat = 1;
  p.flat_pitch = Wg; p.flat_R = R; p.flat_S = S;
  p.dy_row_off = dy_row_off;
  p.kc = KP;
  // Narrow dY (Cy <= 64) over 128 input channels: with A = dy^T half (or three quarters) of every MMA's M rows
  // would be padding, so the operands are exchanged like in conv_wgrad_swapped: the slab views are the M = 128
  // operand, dY the N = Cy one, and tap t accumulates [Cx lanes][Cy columns] (conv_2: 16 MMAs of 128 x 64 x 16
  // per k-block instead of 16 of 128 x 128 x 16).  The shared-memory layout and the producer stay as they are.
  static int swap_on = -1;
  if (swap_on < 0) { const char* e = getenv("GABRIL_WGRAD_FLAT_NO_SWAP"); swap_on = (e && e[0] == '1') ? 0 : 1; }
  const bool swap = swap_on && Cy <= 64 && Cx == 128;
  p.swap_ab = swap ? 1 : 0;
  // A = dy^T: M = Cy rows padded to 128 with out-of-range (zero-filled) boxes (exchanged form: one box, N = Cy)
  const int aw = std::min(64, Cy), a_sw = aw * 2;
  p.a_box_w = aw;
  p.a_boxes = swap ? 1 : 128 / aw;
  p.a_box_bytes = KP * a_sw;
  p.a_stage_bytes = round_up(p.a_boxes * p.a_box_bytes, 1024);
  p.desc_a_hi = desc_hi(p.a_box_bytes, 8 * a_sw, layout_type_for_bytes(a_sw));
  // B = x slabs: 64-channel boxes of KP + S - 1 pixel rows
  int block_n = swap ? Cy : std::min(Cx, 128);   // accumulator columns per tap
  while (S * block_n > 512) block_n >>= 1;
  if (block_n < (swap ? 32 : 64) || (swap && block_n != Cy)) return GABRIL_ERR_BAD_SHAPE;
  const int slab_rows = KP + S - 1;
  p.block_n = block_n;
  p.b_box_w = 64;
  p.b_boxes = swap ? Cx / 64 : block_n / 64;
  p.slab_tx_bytes = (uint32_t)slab_rows * 128u;
  p.b_tap_bytes = round_up(p.slab_tx_bytes, 1024);  // one slab box
  p.b_box_bytes = p.b_tap_bytes;
  p.desc_b_hi = desc_hi(p.b_tap_bytes, 1024, 2);
  int rpg = std::max(1, std::min(R, 512 / (S * block_n)));  // filter rows per CTA
  while (rpg > 1 && 2 * (p.a_stage_bytes + (uint32_t)(rpg * p.b_boxes) * p.b_tap_bytes) > 200u * 1024u) --rpg;
  p.rows_per_group = rpg;
  p.n_acc = rpg * S;
  p.taps = T;
  p.tap_groups = (R + rpg - 1) / rpg;
  p.S = S;
  p.b_stage_bytes = (uint32_t)(rpg * p.b_boxes) * p.b_tap_bytes;
  p.mma_k = KP / 16; p.mma_groups = 1; p.g_taps = 1;
  p.a_kstep = (uint32_t)((16 * a_sw) >> 4);
  p.b_kstep = (uint32_t)((16 * 128) >> 4);
  p.idesc = make_idesc(1, 1, 128, block_n);
  p.M_total[0] = swap ? Cx : Cy;
  p.N_total = swap ? Cy : Cx;
  p.kb_total = (int)((P + KP - 1) / KP);
  {
    uint64_t dims[2] = {(uint64_t)Cy, (uint64_t)P};
    uint64_t strides[1] = {(uint64_t)Cy * 2};
    uint32_t box[2] = {(uint32_t)aw, (uint32_t)KP};
    rc = tiled_map(&p.mapA[0], dy, 2, dims, strides, box, a_sw);
    if (rc) return rc;
  }
  {
    uint64_t dims[2] = {(uint64_t)Cx, (uint64_t)P};
    uint64_t strides[1] = {(uint64_t)(x_pitch_c > 0 ? x_pitch_c : Cx) * 2};
    uint32_t box[2] = {64u, (uint32_t)slab_rows};
    rc = tiled_map(&p.mapB, x, 2, dims, strides, box, 128);
    if (rc) return rc;
  }
  const int m_tiles = swap ? 1 : (Cy + 127) / 128, n_tiles = swap ? 1 : Cx / block_n;
  const int base = m_tiles * n_tiles * p.tap_groups;
  int splits = std::max(1, std::min(p.kb_total, g_num_sms / base));  // one whole wave
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  splits = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.out_mode = tc::OUT_F32_ATOMIC;
  p.out = dw;
  p.ldo = (long long)T * Cx;   // exchanged form: element (cy, tap, cx) from lane cx, column cy (kernel: swap_ab)
  p.tap_stride = Cx;
  p.err_flag = g_err_flag;
  size_t smem;
  pick_stages(p, &smem, 1);
  const int mt[1] = {m_tiles};
  finalize_schedule(p, mt, n_tiles * p.tap_groups, splits, p.n_acc * block_n, 1);
  p.acc_bufs = 1;
  p.tmem_cols = (uint32_t)p.acc_stride;
  return tc::launch(p, 1 << 30, smem, stream);
}

int gabril_conv_wgrad_flat(const void* dy, const void* x, float* dw, int N, int Hg, int Wg, int Cx,
                           int Cy, int R, int S, int dy_row_off, cudaStream_t stream) {
  return conv_wgrad_flat_impl(dy, x, dw, N, Hg, Wg, Cx, Cy, R, S, dy_row_off, 0, stream);
}

// dw4 [4*C1][2][2*Cp] fp32 += da1s [N, SH, SW, 4*C1]^T . (overlapping rows of x4): the weight gradient of
// gabril_conv1_s4d_fwd; gabril_conv1_s4d_fold sums its four quadrants into conv_1's own gradient.
int gabril_conv1_s4d_wgrad(const void* dy, const void* x4, float* dw4, int N, int SH, int SW, int Cp, int C1,
                           cudaStream_t stream) {
  return conv_wgrad_flat_impl(dy, x4, dw4, N, SH, SW, 2 * Cp, 4 * C1, 2, 1, 0, Cp, stream);
}

// conv_1 as a patch GEMM whose rows are in space-to-depth order: row = ((n*SH + sh)*SW + sw)*4 + q
// is pixel (2*sh + q/2 - 1, 2*sw + q%2 - 1) of the img_h x img_w output; rows of padding pixels
// store exact zeros (they are the zero border the next layer's flat-slab convolution reads).
int gabril_conv1_s2d_fwd(const void* patches, const void* w, const float* bias, void* y, int N,
                         int SH, int SW, int img_h, int img_w, int Kp, int Cout, int relu,
                         cudaStream_t stream) {
  return gabril_conv1_s2d_fwd_masked(patches, w, bias, nullptr, y, N, SH, SW, img_h, img_w, Kp, Cout,
                                     relu, stream);
}

// rowmask: one multiplier per patch row (space-to-depth order [N, SH, SW, 4]), after bias / ReLU.
int gabril_conv1_s2d_fwd_masked(const void* patches, const void* w, const float* bias,
                                const float* rowmask, void* y, int N, int SH, int SW, int img_h,
                                int img_w, int Kp, int Cout, int relu, cudaStream_t stream) {
  int rc = resolve_driver();
  if (rc) return rc;
  if (Kp % 32 || !fwd_cout_ok(Cout) || N <= 0) return GABRIL_ERR_BAD_SHAPE;
  const int kc = (Kp % 64 == 0) ? 64 : 32;
  const int block_n = std::min(Cout, 128);
  tc::Params p;
  zero_params(p);
  p.kind = tc::KIND_FWD;
  p.n_classes = 1;
  const long long M = (long long)N * SH * SW * 4;
  if (M >= (1ll << 31)) return GABRIL_ERR_BAD_SHAPE;
  const int m_sub = pick_m_sub(M, block_n, Kp / kc);
  fwd_schedule(p, kc, block_n, 0, m_sub);
  p.M_total[0] = (int)M;
  p.N_total = Cout;
  p.taps = 1; p.S = 1; p.S_w = 1; p.r_step = 1; p.s_step = 1;
  p.chunks_per_tap = Kp / kc;
  p.kb_total = p.chunks_per_tap;
  p.kb_per_split = p.kb_total;
  p.OW[0] = 1; p.OHW[0] = 1;
  p.trav_stride_w = 1; p.trav_stride_h = 1;
  uint64_t dims[2] = {(uint64_t)Kp, (uint64_t)M};
  uint64_t strides[1] = {(uint64_t)Kp * 2};
  uint32_t box[2] = {(uint32_t)kc, (uint32_t)(128 * m_sub)};
  rc = tiled_map(&p.mapA[0], patches, 2, dims, strides, box, kc * 2);
  if (rc) return rc;
  rc = weight_map(&p.mapB, w, Cout, 1, Kp, 0, kc, block_n);
  if (rc) return rc;
  p.out_mode = tc::OUT_BF16;
  p.out = y; p.ldo = Cout;
  p.bias = bias; p.relu = relu; p.rowmask = rowmask;
  p.border_h = SH; p.border_w = SW; p.s2d_H = img_h; p.s2d_W = img_w;
  p.err_flag = g_err_flag;
  size_t smem;
  const int occ = m_sub > 1 ? 1 : occ_setting();
  pick_stages(p, &smem, occ);
  const int mt[1] = {(int)((M + 128 * m_sub - 1) / (128 * m_sub))};
  finalize_schedule(p, mt, Cout / block_n, 1, m_sub * block_n, occ);
  return tc::launch(p, g_num_sms * occ, smem, stream);
}

}  // extern "C"
