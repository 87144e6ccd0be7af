// Shared pieces of the attention kernels (attention.cu: backward, decode, first-generation forward;
// attn_fwd2.cu: ping-pong forward).
#pragma once
#include "tc_common.cuh"

namespace dla {
namespace attn {

constexpr int BQ = 128;
constexpr int BKV = 128;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

struct Params {
  int B, T, Hq, Hk, D;
  float scale;
  int causal;
  const int32_t* seqlens;
  // packed sequences (several documents per row): seg_start[b, t] = index of the first token of t's document, so the
  // keys a query may see are [seg_start[q], q].  Non-decreasing along t; nullptr = one document per row.
  const int32_t* seg_start;
  int dq_tiled;      // dq_acc holds [B, Hq, ceil(T/128), HD/32] blocks of [128 rows x 32 fp32] in the swizzled staging image
  int debug_flags;   // timing experiments only (dla_attn_set_bwd_variant >= 16): bit 0 = skip the dQ reduce-add
  // forward
  bf16* o;
  int64_t o_sb, o_st, o_sh;
  float* lse;        // [B, Hq, T]
  // backward
  const float* delta;  // [B, Hq, T]
  float* dq_acc;       // [B, T, Hq, D] fp32
  bf16* dk;
  bf16* dv;
  int64_t dkv_sb, dkv_st, dkv_sh;
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 2^x on the FMA / ALU pipes (no MUFU): x is split into the nearest integer n and a remainder f in [-0.5, 0.5] with the
// 1.5 * 2^23 rounding trick, 2^f comes from its degree-4 Taylor polynomial (relative error < 5e-5, far below a bf16 ulp) and
// n is added to the exponent field.  At head_dim 128 the forward needs as many MUFU cycles for its exponentials as tensor
// cycles for its MMAs; evaluating half of them here halves the MUFU time (ncu: XU pipe 47 % busy, FMA 14 %).
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -125.f);
  const float t = x + 12582912.f;
  const float f = x - (t - 12582912.f);
  float p = fmaf(f, 0.009618129f, 0.05550411f);
  p = fmaf(p, f, 0.2402265f);
  p = fmaf(p, f, 0.6931472f);
  p = fmaf(p, f, 1.0f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  bf162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// 4-D map over a [B, T, H, D] view with arbitrary element strides; box = 64 (d) x 1 (head) x 128 (tokens) x 1 (batch)
inline int make_qkv_map(CUtensorMap* m, const void* base, int B, int T, int H, int D, int64_t sb, int64_t st, int64_t sh) {
  const uint64_t dims[4] = {(uint64_t)D, (uint64_t)H, (uint64_t)T, (uint64_t)B};
  const uint64_t strides[3] = {(uint64_t)sh, (uint64_t)st, (uint64_t)sb};
  const uint32_t box[4] = {64, 1, 128, 1};
  return make_tmap_4d_bf16(m, base, dims, strides, box);
}

// attn_fwd2.cu
int attn_fwd2_dispatch(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV, const Params& p, cudaStream_t s);
// attn_bwd2.cu: pipelined backward (transposed score tiles, P^T / dS^T consumed from TMEM)
int attn_bwd2_dispatch(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV, const CUtensorMap& tmdO,
                       const CUtensorMap& tmdQ, const Params& p, cudaStream_t s);
// attn_fwd3.cu
int attn_fwd3_dispatch(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV, const Params& p, cudaStream_t s);

}  // namespace attn
}  // namespace dla
