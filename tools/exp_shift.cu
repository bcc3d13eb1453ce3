// Experiment (not product code): do UMMA shared-memory descriptors accept a start address that is
// shifted by whole 128-byte rows inside a 128B-swizzled TMA tile?  If yes, one TMA-loaded slab of
// pixels serves every filter tap of a row as a shifted view (no im2col re-load per tap).
//   mode 0: A K-major  [pixels x 64ch], start shifted by `shift` rows (M shift)
//   mode 1: B MN-major [pixels(K) x 64ch(N)], start shifted by `shift` rows (K shift)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I../gabril-carla_b200/csrc exp_shift.cu -o exp_shift -lcuda
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>

#include "ptx.cuh"

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static uint64_t desc_hi(uint32_t lbo, uint32_t sbo, int layout) {
  uint64_t d = 0;
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}
static uint32_t make_idesc(int a_mn, int b_mn, int M, int N) {
  uint32_t d = 0;
  d |= 1u << 4; d |= 1u << 7; d |= 1u << 10;
  d |= (uint32_t)(a_mn ? 1 : 0) << 15;
  d |= (uint32_t)(b_mn ? 1 : 0) << 16;
  d |= (uint32_t)(N >> 3) << 17;
  d |= (uint32_t)(M >> 4) << 24;
  return d;
}

struct P {
  CUtensorMap mapX, mapW;
  int mode, shift, use_base_offset;
  uint64_t a_hi, b_hi;
  uint32_t idesc;
  float* out;  // [128][64]
};

__global__ void __launch_bounds__(128) k(const __grid_constant__ P p) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar[2];
  __shared__ uint32_t tb;
  const uint32_t base = (ptx::smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t xs = base;             // X slab: 192 rows x 128 B = 24 KB
  const uint32_t ws = base + 32 * 1024; // W / dY tile
  const uint32_t b0 = ptx::smem_u32(&bar[0]), b1 = ptx::smem_u32(&bar[1]);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { ptx::mbar_init(b0, 1); ptx::mbar_init(b1, 1); ptx::fence_mbar_init(); }
  if (warp == 0) { ptx::tmem_alloc(ptx::smem_u32(&tb), 64); ptx::tmem_relinquish(); }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tm = tb;
  if (threadIdx.x == 0) {
    if (p.mode == 0) {
      ptx::mbar_arrive_expect_tx(b0, 192 * 128 + 64 * 128);
      ptx::tma_load_2d(xs, &p.mapX, b0, 0, 0);   // X rows 0..191
      ptx::tma_load_2d(ws, &p.mapW, b0, 0, 0);   // W [64 n][64 k] K-major
    } else {
      ptx::mbar_arrive_expect_tx(b0, 192 * 128 + 2 * 64 * 128);
      ptx::tma_load_2d(xs, &p.mapX, b0, 0, 0);            // X [pix][64 ch]
      ptx::tma_load_2d(ws, &p.mapW, b0, 0, 0);            // dY [64 pix][cout 0..63]
      ptx::tma_load_2d(ws + 64 * 128, &p.mapW, b0, 64, 0);  // dY [64 pix][cout 64..127]
    }
    ptx::mbar_wait(b0, 0);
    ptx::tc_fence_after();
    const uint32_t sh_addr = xs + p.shift * 128;
    const uint64_t bo = p.use_base_offset ? ((uint64_t)((sh_addr >> 7) & 7) << 49) : 0;
    if (p.mode == 0) {
      for (int j = 0; j < 4; ++j) {
        const uint64_t ad = p.a_hi | bo | (uint64_t)(((sh_addr + j * 32) >> 4) & 0x3FFF);
        const uint64_t bd = p.b_hi | (uint64_t)(((ws + j * 32) >> 4) & 0x3FFF);
        ptx::umma_bf16(tm, ad, bd, p.idesc, j > 0);
      }
    } else {
      for (int j = 0; j < 4; ++j) {  // K = 64 pixels, 16 per MMA
        const uint64_t ad = p.a_hi | (uint64_t)(((ws + j * 16 * 128) >> 4) & 0x3FFF);
        const uint64_t bd = p.b_hi | bo | (uint64_t)(((sh_addr + j * 16 * 128) >> 4) & 0x3FFF);
        ptx::umma_bf16(tm, ad, bd, p.idesc, j > 0);
      }
    }
    ptx::umma_commit(b1);
  }
  __syncwarp();
  ptx::mbar_wait(b1, 0);
  ptx::tc_fence_after();
  for (int c = 0; c < 64; c += 32) {
    uint32_t v[32];
    ptx::tmem_ld_32x32(tm + ((uint32_t)(warp * 32) << 16) + c, v);
    ptx::tmem_ld_wait();
    for (int i = 0; i < 32; ++i) p.out[(warp * 32 + lane) * 64 + c + i] = __uint_as_float(v[i]);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tm, 64);
}

int main() {
  cudaFree(0);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  EncodeTiledFn enc = (EncodeTiledFn)fn;
  const int R = 256;
  std::vector<__nv_bfloat16> hX(R * 64), hW(64 * 64), hY(64 * 128);
  std::vector<float> fX(R * 64), fW(64 * 64), fY(64 * 128);
  srand(1);
  for (int i = 0; i < R * 64; ++i) { fX[i] = (float)(rand() % 9 - 4); hX[i] = __float2bfloat16(fX[i]); }
  for (int i = 0; i < 64 * 64; ++i) { fW[i] = (float)(rand() % 5 - 2); hW[i] = __float2bfloat16(fW[i]); }
  for (int i = 0; i < 64 * 128; ++i) { fY[i] = (float)(rand() % 5 - 2); hY[i] = __float2bfloat16(fY[i]); }
  __nv_bfloat16 *dX, *dW, *dY;
  float* dO;
  cudaMalloc(&dX, hX.size() * 2); cudaMalloc(&dW, hW.size() * 2); cudaMalloc(&dY, hY.size() * 2);
  cudaMalloc(&dO, 128 * 64 * 4);
  cudaMemcpy(dX, hX.data(), hX.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dW, hW.data(), hW.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dY, hY.data(), hY.size() * 2, cudaMemcpyHostToDevice);
  auto mk = [&](CUtensorMap* m, void* ptr, uint64_t inner, uint64_t rows, uint32_t bi, uint32_t br) {
    cuuint64_t gd[2] = {inner, rows};
    cuuint64_t gs[1] = {inner * 2};
    cuuint32_t bx[2] = {bi, br}, es[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, gd, gs, bx, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  };
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  std::vector<float> hO(128 * 64);
  for (int mode = 0; mode < 2; ++mode) {
    for (int ubo = 0; ubo < 2; ++ubo) {
      for (int shift : {0, 1, 2, 3, 5, 7, 8, 9, 13, 42}) {
        P p;
        memset(&p, 0, sizeof(p));
        p.mode = mode; p.shift = shift; p.use_base_offset = ubo; p.out = dO;
        mk(&p.mapX, dX, 64, R, 64, 192);
        if (mode == 0) {
          mk(&p.mapW, dW, 64, 64, 64, 64);
          p.a_hi = desc_hi(16, 1024, 2);
          p.b_hi = desc_hi(16, 1024, 2);
          p.idesc = make_idesc(0, 0, 128, 64);
        } else {
          mk(&p.mapW, dY, 128, 64, 64, 64);
          p.a_hi = desc_hi(64 * 128, 1024, 2);   // A = dY^T MN-major, two 64-wide boxes
          p.b_hi = desc_hi(64 * 128, 1024, 2);   // B = X MN-major, one 64-wide box
          p.idesc = make_idesc(1, 1, 128, 64);
        }
        cudaMemset(dO, 0, 128 * 64 * 4);
        k<<<1, 128, 60 * 1024>>>(p);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mode %d shift %d: CUDA error %s\n", mode, shift, cudaGetErrorString(e)); return 1; }
        cudaMemcpy(hO.data(), dO, hO.size() * 4, cudaMemcpyDeviceToHost);
        double maxerr = 0;
        int bad = 0;
        for (int m = 0; m < 128; ++m)
          for (int n = 0; n < 64; ++n) {
            double ref = 0;
            if (mode == 0) for (int kk = 0; kk < 64; ++kk) ref += fX[(m + shift) * 64 + kk] * fW[n * 64 + kk];
            else for (int pp = 0; pp < 64; ++pp) ref += fY[pp * 128 + m] * fX[(pp + shift) * 64 + n];
            const double er = fabs(ref - hO[m * 64 + n]);
            if (er > maxerr) maxerr = er;
            if (er > 1e-3) ++bad;
          }
        printf("mode %d base_offset %d shift %2d: maxerr %.1f bad %d/8192\n", mode, ubo, shift, maxerr, bad);
      }
    }
  }
  return 0;
}
