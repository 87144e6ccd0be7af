// One-token decode attention fused with RoPE and the KV-cache append.
//
// For every sequence b the rollout engine has a fresh qkv row (one token).  The eager formulation needs seven
// launches per layer: RoPE, two scatter writes into the cache (plus two index casts), a copy that makes q
// contiguous, and the attention kernel.  Here one block per (kv head, sequence) does all of it:
//   1. rotate the new K row and the G = Hq/Hk query rows that share it (pairs (j, j + rot/2), cos/sin table lookup),
//      write K and V into the cache at position lens[b]-1;
//   2. scores: every lane owns whole cache positions (a K row is 2*HD contiguous bytes), dots against the G
//      rotated queries held in shared memory — no cross-lane reduction;
//   3. softmax per query head (one warp per head);
//   4. P.V with 16-byte loads of V, each thread owning an 8-wide slice of HD for a stripe of positions and
//      accumulating all G heads, then a cross-stripe reduction through shared memory.
// The K/V rows of a kv head are read once for its G query heads (the unfused kernel read them G times).
//
// Reference path replaced: HF generate()'s per-token DynamicCache.update + apply_rotary_pos_emb + SDPA
// (R/src/training/train_rlhf.py:123-124, R/src/training/generate_teacher_data.py:73-79; SURVEY K16, worklist #11).
#include <math.h>

#include "common.cuh"

namespace dla {
namespace decode {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

template <int HD, int G>
__global__ void __launch_bounds__(kThreads)
decode_attn_kernel(bf16* __restrict__ qkv, int64_t qkv_stride, bf16* __restrict__ kc, bf16* __restrict__ vc,
                   bf16* __restrict__ o, const int32_t* __restrict__ lens, const float* __restrict__ cos_t,
                   const float* __restrict__ sin_t, int S_max, int Hq, int Hk, int rot_dim, float scale) {
  extern __shared__ float sm[];
  float* q_s = sm;                        // [G][HD] rotated queries, pre-scaled
  float* sc = q_s + G * HD;               // [G][S_max] scores / probabilities
  float* part = sc + (size_t)G * S_max;   // [kThreads / (HD/8)][G][HD] partial outputs
  __shared__ float inv_sum[G];

  const int hk = blockIdx.x, b = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int len = min(lens[b], S_max);
  if (len <= 0) {
    for (int i = tid; i < G * HD; i += kThreads) o[((int64_t)b * Hq + hk * G + i / HD) * HD + i % HD] = __float2bfloat16(0.f);
    return;
  }
  const int pos = len - 1;
  bf16* row = qkv + (int64_t)b * qkv_stride;
  const int half = rot_dim >> 1;

  // ---- 1. rotate q (G heads) and k, append k / v ---------------------------------------------------------
  // work items: (head slot, pair index) with head slot 0..G-1 = queries, G = key; HD/2 pairs per head cover
  // (j, j + half) for j < half and the untouched tail (rot_dim + 2j', rot_dim + 2j' + 1) beyond the rotary part
  for (int it = tid; it < (G + 1) * (HD / 2); it += kThreads) {
    const int slot = it / (HD / 2), j = it % (HD / 2);
    const bf16* src = slot < G ? row + (int64_t)(hk * G + slot) * HD : row + (int64_t)(Hq + hk) * HD;
    int i0, i1;
    float v0, v1;
    if (j < half) {
      i0 = j;
      i1 = j + half;
      const float a = __bfloat162float(src[i0]), c = __bfloat162float(src[i1]);
      const float cs = cos_t[(int64_t)pos * half + j], sn = sin_t[(int64_t)pos * half + j];
      v0 = a * cs - c * sn;
      v1 = c * cs + a * sn;
    } else {
      i0 = rot_dim + 2 * (j - half);
      i1 = i0 + 1;
      v0 = __bfloat162float(src[i0]);
      v1 = __bfloat162float(src[i1]);
    }
    // keep the bf16 rounding of the unfused path (q and k are bf16 tensors after RoPE there)
    const bf16 r0 = __float2bfloat16(v0), r1 = __float2bfloat16(v1);
    if (slot < G) {
      q_s[slot * HD + i0] = __bfloat162float(r0) * scale;
      q_s[slot * HD + i1] = __bfloat162float(r1) * scale;
    } else {
      bf16* dst = kc + (((int64_t)b * S_max + pos) * Hk + hk) * HD;
      dst[i0] = r0;
      dst[i1] = r1;
    }
  }
  for (int i = tid; i < HD; i += kThreads)
    vc[(((int64_t)b * S_max + pos) * Hk + hk) * HD + i] = row[(int64_t)(Hq + Hk + hk) * HD + i];
  __syncthreads();      // cache row and q_s visible to the whole block

  // ---- 2. scores: one cache position per thread.  All 2*HD bytes of the K row are requested before the first FMA
  // (one L2 round trip per position instead of one per 16-byte chunk).
  for (int s = tid; s < len; s += kThreads) {
    const bf16* kp = kc + (((int64_t)b * S_max + s) * Hk + hk) * HD;
    bf16x8 kraw[HD / 8];
#pragma unroll
    for (int c = 0; c < HD / 8; ++c) kraw[c] = ld8(kp + c * 8);
    float d[G];
#pragma unroll
    for (int g = 0; g < G; ++g) d[g] = 0.f;
#pragma unroll
    for (int c = 0; c < HD / 8; ++c) {
      float kv[8];
      unpack8(kraw[c], kv);
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const float4 qa = *reinterpret_cast<const float4*>(q_s + g * HD + c * 8);
        const float4 qb = *reinterpret_cast<const float4*>(q_s + g * HD + c * 8 + 4);
        d[g] += kv[0] * qa.x + kv[1] * qa.y + kv[2] * qa.z + kv[3] * qa.w + kv[4] * qb.x + kv[5] * qb.y + kv[6] * qb.z +
                kv[7] * qb.w;
      }
    }
#pragma unroll
    for (int g = 0; g < G; ++g) sc[(size_t)g * S_max + s] = d[g];
  }
  __syncthreads();

  // ---- 3. softmax, one warp per query head ------------------------------------------------------------------
  for (int g = warp; g < G; g += kWarps) {
    float* p = sc + (size_t)g * S_max;
    float mx = -INFINITY;
    for (int s = lane; s < len; s += 32) mx = fmaxf(mx, p[s]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int s = lane; s < len; s += 32) {
      const float e = __expf(p[s] - mx);
      p[s] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    if (lane == 0) inv_sum[g] = 1.f / sum;
  }
  __syncthreads();

  // ---- 4. P.V: thread = (8-wide slice of HD, stripe of positions) --------------------------------------------
  constexpr int kChunks = HD / 8, kStripes = kThreads / kChunks;
  const int ch = tid % kChunks, stripe = tid / kChunks;
  float acc[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[g][e] = 0.f;
  // positions of this stripe in batches of kBatch independent 16-byte loads
  constexpr int kBatch = 8;
  for (int s0 = stripe; s0 < len; s0 += kStripes * kBatch) {
    bf16x8 raw[kBatch];
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int s = s0 + u * kStripes;
      if (s < len) raw[u] = ld8(vc + (((int64_t)b * S_max + s) * Hk + hk) * HD + ch * 8);
    }
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int s = s0 + u * kStripes;
      if (s < len) {
        float vv[8];
        unpack8(raw[u], vv);
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const float pr = sc[(size_t)g * S_max + s];
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[g][e] += pr * vv[e];
        }
      }
    }
  }
#pragma unroll
  for (int g = 0; g < G; ++g)
#pragma unroll
    for (int e = 0; e < 8; ++e) part[((size_t)stripe * G + g) * HD + ch * 8 + e] = acc[g][e];
  __syncthreads();
  for (int i = tid; i < G * HD; i += kThreads) {
    const int g = i / HD, dd = i % HD;
    float t = 0.f;
#pragma unroll 4
    for (int st = 0; st < kStripes; ++st) t += part[((size_t)st * G + g) * HD + dd];
    o[((int64_t)b * Hq + hk * G + g) * HD + dd] = __float2bfloat16(t * inv_sum[g]);
  }
}

template <int HD, int G>
int launch(bf16* qkv, int64_t qkv_stride, bf16* kc, bf16* vc, bf16* o, const int32_t* lens, const float* cos_t,
           const float* sin_t, int B, int S_max, int Hq, int Hk, int rot_dim, float scale, cudaStream_t st) {
  constexpr int kStripes = kThreads / (HD / 8);
  const size_t smem = ((size_t)G * HD + (size_t)G * S_max + (size_t)kStripes * G * HD) * sizeof(float);
  DLA_REQUIRE(smem <= 220 * 1024, DLA_E_TOOLARGE,
              "decode_attn_fused: S_max=%d with %d query heads per kv head needs %zu B of shared memory (limit 220 KiB)", S_max, G,
              smem);
  DLA_CUDA(cudaFuncSetAttribute(decode_attn_kernel<HD, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  decode_attn_kernel<HD, G><<<dim3(Hk, B), kThreads, smem, st>>>(qkv, qkv_stride, kc, vc, o, lens, cos_t, sin_t, S_max, Hq, Hk,
                                                               rot_dim, scale);
  return check_launch("decode_attn_fused");
}

}  // namespace decode
}  // namespace dla

using namespace dla;

extern "C" {

int dla_decode_attn_fused(void* qkv, int64_t qkv_stride, void* kcache, void* vcache, void* o, const int32_t* lens,
                          const float* cos_t, const float* sin_t, int B, int S_max, int Hq, int Hk, int D, int rot_dim,
                          float scale, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(B > 0 && S_max > 0 && Hk > 0 && Hq % Hk == 0, DLA_E_BADARG, "decode_attn_fused: bad shape B=%d S_max=%d Hq=%d Hk=%d",
              B, S_max, Hq, Hk);
  DLA_REQUIRE(D == 64 || D == 128, DLA_E_UNSUPP, "decode_attn_fused: head_dim %d not in {64, 128}", D);
  DLA_REQUIRE(rot_dim >= 0 && rot_dim <= D && rot_dim % 2 == 0 && (rot_dim == 0 || (cos_t != nullptr && sin_t != nullptr)),
              DLA_E_BADARG, "decode_attn_fused: rot_dim=%d must be even, <= head_dim, with tables when non-zero", rot_dim);
  DLA_REQUIRE(qkv_stride % 8 == 0 && reinterpret_cast<uintptr_t>(kcache) % 16 == 0 && reinterpret_cast<uintptr_t>(vcache) % 16 == 0,
              DLA_E_BADARG, "decode_attn_fused: cache bases must be 16-byte aligned and the qkv pitch a multiple of 8");
  const int G = Hq / Hk;
  cudaStream_t st = (cudaStream_t)stream;
  bf16 *q = (bf16*)qkv, *kc = (bf16*)kcache, *vc = (bf16*)vcache, *op = (bf16*)o;
#define DLA_DEC(HDv, Gv) \
  return decode::launch<HDv, Gv>(q, qkv_stride, kc, vc, op, lens, cos_t, sin_t, B, S_max, Hq, Hk, rot_dim, scale, st)
  if (D == 128) {
    switch (G) {
      case 1: DLA_DEC(128, 1);
      case 2: DLA_DEC(128, 2);
      case 4: DLA_DEC(128, 4);
      case 8: DLA_DEC(128, 8);
    }
  } else {
    switch (G) {
      case 1: DLA_DEC(64, 1);
      case 2: DLA_DEC(64, 2);
      case 4: DLA_DEC(64, 4);
      case 8: DLA_DEC(64, 8);
    }
  }
#undef DLA_DEC
  DLA_REQUIRE(false, DLA_E_UNSUPP, "decode_attn_fused: %d query heads per kv head not in {1, 2, 4, 8}", G);
  return 0;
}

}  // extern "C"
