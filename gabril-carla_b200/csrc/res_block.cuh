// Fused Residual block forward (res_block.cu): launch parameters filled by gabril_res_block_fwd (tc_plan.cu).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>

#include "tc_gemm.cuh"

namespace rb {

constexpr int kStages = 4;   // k-blocks in flight: one filter tap = two 64-channel im2col tiles (32 KB)

struct Params {
  CUtensorMap mapX;    // x [N, H, W, 128] in im2col mode: 64 channels x 128 pixels, pad 1
  CUtensorMap mapW3;   // w3 [32][9][128]: boxes of 32 couts x 64 cin of one tap
  CUtensorMap mapW1;   // w1 [128][1][32]: one box
  const __nv_bfloat16* x;   // also the skip operand
  __nv_bfloat16* h;         // [M, 32]
  uint32_t* bits_h;         // [M]
  __nv_bfloat16* out;       // [M, 128]
  uint32_t* bits_out;       // [M * 4]
  int M, HW, W, tiles;
  tc::FastDiv fd_hw, fd_w;
  uint64_t desc_a1_hi, desc_b1_hi, desc_a2_hi, desc_b2_hi;   // smem descriptors, start address zero
  uint32_t idesc1, idesc2;
  int* err_flag;
  int debug;   // perf experiments (GABRIL_RB_DEBUG): 1 = no TMA loads, 2 = no 3x3 MMAs, 4 = no output epilogue work
};

size_t smem_bytes();
int launch(const Params& p, int grid, cudaStream_t stream);

}  // namespace rb
