// The bf16 epilogue item shared by the tcgen05 kernels (tc_gemm.cu, res_block.cu).
#pragma once
#include "common.cuh"
#include "ptx.cuh"
#include "tc_gemm.cuh"

namespace tc {

// ---------------------------------------------------------------------------------------------
// bf16 epilogue of one 32-row x 32-column accumulator chunk (one warp).
// The epilogue warp is alone (or nearly) on its scheduler: every instruction's fixed latency, every
// branch on a runtime flag and every parameter load from the constant bank is exposed (ncu: stall_wait /
// branch_resolving / short_scoreboard dominate, not memory).  So the chunk is compiled once per
// COMBINATION of epilogue terms that the network uses (straight-line code, in-place packed math) and the
// kernel picks the variant with one switch per chunk; EF_RUNTIME keeps the general form for the rest.
// ---------------------------------------------------------------------------------------------
template <uint32_t F>
__device__ __forceinline__ void epi_item_bf16(uint32_t rt, uint32_t taddr, bool live, const float* bias,
                                              const bf16* resid, const bf16* gate, uint32_t gword, float rmask,
                                              uint32_t keep, bf16* out, uint32_t* bits) {
  constexpr bool RT = (F & EF_RUNTIME) != 0;
#define EPI_HAS(flag) (RT ? ((rt & (flag)) != 0u) : ((F & (flag)) != 0u))
  uint32_t v[32];
  // residual / gate operands: issued before the accumulator read so that their latency overlaps it
  uint32_t rp[16], gp[16];
  if (EPI_HAS(EF_RESID) && live) { ptx::ldg256(resid, rp); ptx::ldg256(resid + 16, rp + 8); }
  if (EPI_HAS(EF_GATE) && live) { ptx::ldg256(gate, gp); ptx::ldg256(gate + 16, gp + 8); }
  __syncwarp();
  ptx::tmem_ld_32x32(taddr, v);
  ptx::tmem_ld_wait();
  if (!live) return;
  uint64_t acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = ptx::pack2(v[2 * i], v[2 * i + 1]);
  if (EPI_HAS(EF_BIAS)) {
    const float4* b4 = reinterpret_cast<const float4*>(bias);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 b = b4[i];
      acc[2 * i] = ptx::add2(acc[2 * i], ptx::pack2(__float_as_uint(b.x), __float_as_uint(b.y)));
      acc[2 * i + 1] = ptx::add2(acc[2 * i + 1], ptx::pack2(__float_as_uint(b.z), __float_as_uint(b.w)));
    }
  }
  if (EPI_HAS(EF_RESID)) {
#pragma unroll
    for (int i = 0; i < 16; ++i)   // bf16 -> fp32 is a shift / mask of the packed word
      acc[i] = ptx::add2(acc[i], ptx::pack2(rp[i] << 16, rp[i] & 0xffff0000u));
  }
  if (EPI_HAS(EF_RMASK)) {   // masks are >= 0, so scaling before the ReLU equals scaling after it
    const uint64_t rmask2 = ptx::pack2(__float_as_uint(rmask), __float_as_uint(rmask));
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = ptx::mul2(acc[i], rmask2);
  }
  uint32_t w[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) w[i] = ptx::cvt_bf16x2(acc[i]);
  if (EPI_HAS(EF_RELU)) {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = ptx::relu_bf16x2(w[i]);
  }
  if (EPI_HAS(EF_GATE)) {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] &= ptx::gt0_mask_bf16x2(gp[i]);
  }
  if (EPI_HAS(EF_GBITS)) {
    // sign word layout: bit k = element 2k, bit 16 + k = element 2k + 1 (the two halves of packed word
    // k).  Shift the pair of bits to the top of bytes 0/2 (k < 8) or 1/3, then one byte permute with sign
    // replication expands them to a 0xffff / 0 mask per half.
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] &= ptx::prmt(gword << (7 - i), 0u, 0xaa88u);
#pragma unroll
    for (int i = 8; i < 16; ++i) w[i] &= ptx::prmt(gword << (15 - i), 0u, 0xbb99u);
  }
  if (EPI_HAS(EF_KEEP)) {   // border / padding pixels of a zero-bordered output store exact zeros
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] &= keep;
  }
  ptx::stg256(out, w);
  ptx::stg256(out + 16, w + 8);
  if (EPI_HAS(EF_BITS)) {
    // ReLU's backward predicate on the stored bf16 value; two partial words keep the OR chains short
    uint32_t s0 = 0u, s1 = 0u;
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      s0 |= ptx::gt0_mask_bf16x2(w[i]) & (0x00010001u << i);
      s1 |= ptx::gt0_mask_bf16x2(w[i + 1]) & (0x00010001u << (i + 1));
    }
    *bits = s0 | s1;
  }
#undef EPI_HAS
}

}  // namespace tc
