// Causal GQA flash attention for sm_100a: tcgen05.mma with TMEM accumulators, TMA-fed, forward and
// backward, plus the single-token KV-cache decode kernel used by the rollout engine.
//
// Forward (one CTA = one 128-row query tile of one head):
//   warp 0   TMA producer: Q once, then K/V tiles through a 2-stage ring
//   warp 1   MMA issuer  : S = Q K^T (128x128, fp32 in TMEM, double-buffered), O_j = P V
//   warps 2-5 softmax    : one thread owns one query row (TMEM lane): online max/sum, P -> smem (bf16,
//                          128B-swizzled K-major), running O in registers (rescaled by alpha each tile)
// Backward (one CTA = one 128-row key/value tile of one KV head, loops over the group's query heads
// and the query tiles at or below the diagonal): S and dP from two MMAs, P/dS written to smem by the
// row-owner threads, then dV += P^T dO, dK += dS^T Q (TMEM-resident across the whole loop) and
// dQ = dS K (red.global.add.f32 into an fp32 scratch).
//
// Reference path replaced: PyTorch SDPA under HF Llama (R/src/training/train_dpo.py:32) and
// HF generate's per-token SDPA (R/src/training/train_rlhf.py:124).
#include <math.h>

#include "attn_common.cuh"

namespace dla {
namespace attn {

constexpr int kFwdThreads = 320;     // TMA warp, MMA warp, 2 x 4 softmax / row-owner warps (forward and backward)

// byte offset of element (row, col) in a [128 rows x 128 cols] bf16 tile stored as two 128B-swizzled
// K-major blocks of 64 columns (the layout TMA SWIZZLE_128B produces and UMMA consumes)
__device__ __forceinline__ uint32_t sw128_offset(int row, int col) {
  const int blk = col >> 6;
  const int chunk = (col & 63) >> 3;
  return (uint32_t)(blk * 16384 + row * 128 + ((chunk ^ (row & 7)) << 4) + ((col & 7) << 1));
}

// ============================================================================== forward
template <int HD>
__global__ void __launch_bounds__(kFwdThreads, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const Params p) {
  constexpr int NB = HD / 64;                 // 64-column blocks per head-dim
  constexpr int Q_BYTES = BQ * HD * 2;
  constexpr int KV_BYTES = BKV * HD * 2;      // one of K or V
  constexpr int STAGE_BYTES = 2 * KV_BYTES;
  constexpr int P_BYTES = BQ * BKV * 2;
  constexpr int kStages = 2;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + Q_BYTES;
  uint8_t* sP = sKV + kStages * STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + P_BYTES);
  uint64_t* q_full = bars;           // 1
  uint64_t* kv_full = bars + 1;      // [2]
  uint64_t* kv_empty = bars + 3;     // [2]
  uint64_t* s_full = bars + 5;       // [2]
  uint64_t* s_free = bars + 7;       // [2]
  uint64_t* p_full = bars + 9;       // 1
  uint64_t* o_full = bars + 10;      // 1
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 11);
  float* mx_ex = reinterpret_cast<float*>(bars + 16);   // [2 tiles][2 warpgroups][128 rows] row-max exchange
  float* l_ex = mx_ex + 512;                            // [2 warpgroups][128 rows] row-sum exchange

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qi = (int)gridDim.x - 1 - (int)blockIdx.x;   // heaviest (last) query tiles first
  const int h = blockIdx.y, b = blockIdx.z;
  const int hk = h / (p.Hq / p.Hk);
  const int q0 = qi * BQ;
  int len = p.T;
  if (p.seqlens != nullptr) len = min(len, p.seqlens[b]);
  int n_tiles = (len + BKV - 1) / BKV;
  if (p.causal) n_tiles = min(n_tiles, qi + 1);
  if (q0 >= len) n_tiles = 0;

  if (threadIdx.x == 0) {
    tc::prefetch_tmap(&tmQ);
    tc::prefetch_tmap(&tmK);
    tc::prefetch_tmap(&tmV);
    tc::mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&kv_full[s], 1);
      tc::mbar_init(&kv_empty[s], 1);
      tc::mbar_init(&s_full[s], 1);
      tc::mbar_init(&s_free[s], 256);
    }
    tc::mbar_init(p_full, 256);
    tc::mbar_init(o_full, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc<1>(tmem_holder, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const uint32_t tmem_S = tmem_base;           // 2 x 128 columns
  const uint32_t tmem_O = tmem_base + 256;     // HD columns

  if (warp == 0) {
    if (lane == 0 && n_tiles > 0) {
      tc::mbar_expect_tx(q_full, Q_BYTES);
#pragma unroll
      for (int c = 0; c < NB; ++c) tc::tma_load_4d(sQ + c * 16384, &tmQ, q_full, c * 64, h, q0, b);
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j & 1;
        tc::mbar_wait(&kv_empty[st], ((j >> 1) & 1) ^ 1);
        uint8_t* sK = sKV + st * STAGE_BYTES;
        uint8_t* sV = sK + KV_BYTES;
        tc::mbar_expect_tx(&kv_full[st], STAGE_BYTES);
#pragma unroll
        for (int c = 0; c < NB; ++c) tc::tma_load_4d(sK + c * 16384, &tmK, &kv_full[st], c * 64, hk, j * BKV, b);
#pragma unroll
        for (int c = 0; c < NB; ++c) tc::tma_load_4d(sV + c * 16384, &tmV, &kv_full[st], c * 64, hk, j * BKV, b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && n_tiles > 0) {
      constexpr uint32_t idesc_s = tc::make_idesc_bf16(128, BKV, 0, 0);
      constexpr uint32_t idesc_pv = tc::make_idesc_bf16(128, HD, 0, 1);
      const uint32_t aQ = tc::smem_u32(sQ), aP = tc::smem_u32(sP);
      auto issue_s = [&](int j) {
        const int st = j & 1, buf = j & 1;
        tc::mbar_wait(&kv_full[st], (j >> 1) & 1);
        tc::mbar_wait(&s_free[buf], ((j >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        const uint32_t aK = tc::smem_u32(sKV + st * STAGE_BYTES);
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) {
          const uint32_t off = (ks >> 2) * 16384 + (ks & 3) * 32;
          tc::umma_bf16<1>(tmem_S + buf * 128, tc::make_smem_desc(aQ + off, 16, 1024),
                           tc::make_smem_desc(aK + off, 16, 1024), idesc_s, ks > 0 ? 1u : 0u);
        }
        tc::umma_commit(&s_full[buf]);
      };
      tc::mbar_wait(q_full, 0);
      issue_s(0);
      for (int j = 0; j < n_tiles; ++j) {
        if (j + 1 < n_tiles) issue_s(j + 1);
        const int st = j & 1;
        tc::mbar_wait(p_full, j & 1);
        tc::tc_fence_after();
        const uint32_t aV = tc::smem_u32(sKV + st * STAGE_BYTES + KV_BYTES);
#pragma unroll
        for (int ks = 0; ks < BKV / 16; ++ks) {
          const uint32_t offP = (ks >> 2) * 16384 + (ks & 3) * 32;
          tc::umma_bf16<1>(tmem_O, tc::make_smem_desc(aP + offP, 16, 1024),
                           tc::make_smem_desc(aV + ks * 2048, 16384, 1024), idesc_pv, ks > 0 ? 1u : 0u);
        }
        tc::umma_commit(o_full);
        tc::umma_commit(&kv_empty[st]);
      }
    }
  } else {
    // ------------------------------ softmax / accumulate / epilogue
    // Two warpgroups share every query row: thread (wg, r) owns key columns [64 wg, 64 wg + 64) of each
    // S tile and output columns [wg HD/2, (wg+1) HD/2).  Two warps per scheduler hide each other's FFMA /
    // MUFU latency; the row max is exchanged through shared memory once per tile, the row sum only at
    // the end (both halves rescale with the same running max).
    const int wg = (warp - 2) >> 2;
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;
    const int q_idx = q0 + r;
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const float scale2 = p.scale * kLog2e;
    constexpr int OC = HD / 2;
    float acc[OC];
#pragma unroll
    for (int i = 0; i < OC; ++i) acc[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;

    for (int j = 0; j < n_tiles; ++j) {
      const int kv0 = j * BKV;
      const int buf = j & 1;
      tc::mbar_wait(&s_full[buf], (j >> 1) & 1);
      tc::tc_fence_after();
      uint32_t v[64];
      const uint32_t tS = tmem_S + lane_off + buf * 128 + wg * 64;
      tc::tmem_ld_32x32(tS, v);
      tc::tmem_ld_32x32(tS + 32, v + 32);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&s_free[buf]);          // S lives in registers now: the MMA warp may refill this buffer
      const bool need_mask = (p.causal && (kv0 + BKV - 1 > q0)) || (kv0 + BKV > len);
      int limit = 64;                          // number of valid columns among my 64
      if (need_mask) {
        int lim = len;
        if (p.causal) lim = min(lim, q_idx + 1);
        limit = lim - (kv0 + wg * 64);
      }
      float mx = -INFINITY;
      if (!need_mask) {
#pragma unroll
        for (int jj = 0; jj < 64; ++jj) mx = fmaxf(mx, __uint_as_float(v[jj]));
      } else {
#pragma unroll
        for (int jj = 0; jj < 64; ++jj) mx = fmaxf(mx, jj < limit ? __uint_as_float(v[jj]) : -INFINITY);
      }
      float* ex = mx_ex + (j & 1) * 256;
      ex[wg * 128 + r] = mx;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      mx = fmaxf(mx, ex[(wg ^ 1) * 128 + r]);
      const float m_new = fmaxf(m_run, mx * scale2);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = ex2_approx(m_run - m_use);
      const float neg_m = -m_use;
      float sum = 0.f;
      uint32_t pk[32];
      if (!need_mask) {
#pragma unroll
        for (int jj = 0; jj < 64; jj += 2) {
          const float p0 = ex2_approx(fmaf(__uint_as_float(v[jj]), scale2, neg_m));
          const float p1 = ex2_approx(fmaf(__uint_as_float(v[jj + 1]), scale2, neg_m));
          sum += p0 + p1;
          pk[jj >> 1] = pack_bf16x2(p0, p1);
        }
      } else {
#pragma unroll
        for (int jj = 0; jj < 64; jj += 2) {
          const float p0 = jj < limit ? ex2_approx(fmaf(__uint_as_float(v[jj]), scale2, neg_m)) : 0.f;
          const float p1 = jj + 1 < limit ? ex2_approx(fmaf(__uint_as_float(v[jj + 1]), scale2, neg_m)) : 0.f;
          sum += p0 + p1;
          pk[jj >> 1] = pack_bf16x2(p0, p1);
        }
      }
      uint8_t* prow = sP + wg * 16384 + r * 128;
#pragma unroll
      for (int c = 0; c < 8; ++c)
        *reinterpret_cast<uint4*>(prow + ((c ^ (r & 7)) << 4)) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
      l_run = l_run * alpha + sum;
      m_run = m_new;
      tc::fence_proxy_async();
      tc::mbar_arrive(p_full);
      // fold this tile's P V into the running output
      tc::mbar_wait(o_full, j & 1);
      tc::tc_fence_after();
#pragma unroll
      for (int ch = 0; ch < OC / 32; ++ch) {
        uint32_t o[32];
        tc::tmem_ld_32x32(tmem_O + lane_off + wg * OC + ch * 32, o);
        tc::tmem_ld_wait();
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) acc[ch * 32 + jj] = fmaf(acc[ch * 32 + jj], alpha, __uint_as_float(o[jj]));
      }
      tc::tc_fence_before();
    }
    // epilogue: combine the two halves' row sums, normalise, store
    l_ex[wg * 128 + r] = l_run;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    const float l_tot = l_ex[r] + l_ex[128 + r];
    if (q_idx < p.T) {
      const bool valid = (q_idx < len) && (l_tot > 0.f);
      const float inv = valid ? 1.f / l_tot : 0.f;
      bf16* dst = p.o + (int64_t)b * p.o_sb + (int64_t)q_idx * p.o_st + (int64_t)h * p.o_sh + wg * OC;
#pragma unroll
      for (int c = 0; c < OC; c += 8) {
        float f[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) f[t] = acc[c + t] * inv;
        st8(dst + c, pack8(f));
      }
      if (p.lse != nullptr && wg == 0)
        p.lse[((int64_t)b * p.Hq + h) * p.T + q_idx] = valid ? (m_run + log2f(l_tot)) * kLn2 : 0.f;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc<1>(tmem_base, 512);
  }
}

// ============================================================================== backward helpers
// delta[b,h,t] = sum_d O * dO ; also zeroes the fp32 dQ accumulator row.
__global__ void attn_bwd_prep_kernel(const bf16* __restrict__ o, const bf16* __restrict__ dout, float* __restrict__ delta,
                                     float* __restrict__ dq_acc, int B, int T, int Hq, int D, int64_t o_sb, int64_t o_st,
                                     int64_t o_sh) {
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);   // over B*T*Hq
  if (row >= (int64_t)B * T * Hq) return;
  const int lane = threadIdx.x & 31;
  const int h = (int)(row % Hq);
  const int64_t bt = row / Hq;
  const int t = (int)(bt % T), b = (int)(bt / T);
  const bf16* po = o + b * o_sb + t * o_st + h * o_sh;
  const bf16* pd = dout + b * o_sb + t * o_st + h * o_sh;
  float s = 0.f;
  for (int c = lane * 4; c < D; c += 128) {
    const uint2 a = *reinterpret_cast<const uint2*>(po + c);
    const uint2 g = *reinterpret_cast<const uint2*>(pd + c);
    const float2 a0 = __bfloat1622float2(*reinterpret_cast<const bf162*>(&a.x));
    const float2 a1 = __bfloat1622float2(*reinterpret_cast<const bf162*>(&a.y));
    const float2 g0 = __bfloat1622float2(*reinterpret_cast<const bf162*>(&g.x));
    const float2 g1 = __bfloat1622float2(*reinterpret_cast<const bf162*>(&g.y));
    s += a0.x * g0.x + a0.y * g0.y + a1.x * g1.x + a1.y * g1.y;
    *reinterpret_cast<float4*>(dq_acc + row * D + c) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  s = warp_sum(s);
  if (lane == 0) delta[((int64_t)b * Hq + h) * T + t] = s;
}

__global__ void attn_bwd_dq_cast_kernel(const float* __restrict__ dq_acc, bf16* __restrict__ dq, int64_t rows, int D,
                                        int Hq, int T, int64_t sb, int64_t st, int64_t sh, int tiled) {
  const int per_row = D >> 3;
  const int64_t total = rows * per_row;
  const int nT = (T + 127) >> 7;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = idx / per_row;
    const int c = (int)(idx % per_row) * 8;
    const int h = (int)(row % Hq);
    const int64_t bt = row / Hq;
    const int t = (int)(bt % T);
    const int64_t b = bt / T;
    float f[8];
    if (!tiled) {
      *reinterpret_cast<float4*>(f) = *reinterpret_cast<const float4*>(dq_acc + row * D + c);
      *reinterpret_cast<float4*>(f + 4) = *reinterpret_cast<const float4*>(dq_acc + row * D + c + 4);
    } else {
      // [b, h, t / 128, c / 32] blocks of [128 rows x 32 fp32]; inside a block the eight 16-byte chunks of a row are
      // XOR-swizzled with the row index (the shared-memory staging image of attn_bwd2_kernel)
      const int r = t & 127, blk = c >> 5, ch = (c & 31) >> 2;
      const float* base = dq_acc + ((((b * Hq + h) * nT + (t >> 7)) * (D >> 5) + blk) * 128 + r) * 32;
      *reinterpret_cast<float4*>(f) = *reinterpret_cast<const float4*>(base + ((ch ^ (r & 7)) << 2));
      *reinterpret_cast<float4*>(f + 4) = *reinterpret_cast<const float4*>(base + (((ch + 1) ^ (r & 7)) << 2));
    }
    st8(dq + b * sb + t * st + h * sh + c, pack8(f));
  }
}

// ============================================================================== backward main
template <int HD>
__global__ void __launch_bounds__(kFwdThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                const __grid_constant__ CUtensorMap tmdQ, const Params p) {
  constexpr int NB = HD / 64;
  constexpr int T_BYTES = 128 * HD * 2;       // one [128 x HD] operand tile
  constexpr int P_BYTES = 128 * 128 * 2;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sK = smem;
  uint8_t* sV = sK + T_BYTES;
  uint8_t* sQ = sV + T_BYTES;
  uint8_t* sdO = sQ + T_BYTES;
  uint8_t* sP = sdO + T_BYTES;
  uint8_t* sdS = sP + P_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sdS + P_BYTES);
  uint64_t* kv_full = bars;        // 1
  uint64_t* qdo_full = bars + 1;   // 1
  uint64_t* qdo_empty = bars + 2;  // 1
  uint64_t* sdp_full = bars + 3;   // 1
  uint64_t* sdp_free = bars + 4;   // 128
  uint64_t* pds_full = bars + 5;   // 128
  uint64_t* dq_full = bars + 6;    // 1
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 7);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D grid, longest-processing-time first: kv tile 0 (most query tiles at or below it) of every (kv head, batch)
  // runs before any kv tile 1, ...; the launch ends with the shortest CTAs
  const int n_hb = p.Hk * p.B;
  const int j = (int)blockIdx.x / n_hb, hb = (int)blockIdx.x % n_hb;
  const int hk = hb % p.Hk, b = hb / p.Hk;
  const int G = p.Hq / p.Hk;
  const int kv0 = j * BKV;
  int len = p.T;
  if (p.seqlens != nullptr) len = min(len, p.seqlens[b]);
  int nq = (len + BQ - 1) / BQ;
  const int i_begin = p.causal ? j : 0;
  if (p.seg_start != nullptr) {
    // packed rows: query tiles whose FIRST row's document starts after this kv tile see none of its keys, and
    // neither does any later tile (seg_start is non-decreasing)
    int i_end = i_begin;
    while (i_end < nq && p.seg_start[(int64_t)b * p.T + i_end * BQ] <= kv0 + BKV - 1) ++i_end;
    nq = i_end;
  }
  const int iters_per_head = (kv0 < len && nq > i_begin) ? (nq - i_begin) : 0;
  const int n_iters = iters_per_head * G;

  if (threadIdx.x == 0) {
    tc::prefetch_tmap(&tmQ);
    tc::prefetch_tmap(&tmK);
    tc::prefetch_tmap(&tmV);
    tc::prefetch_tmap(&tmdO);
    tc::mbar_init(kv_full, 1);
    tc::mbar_init(qdo_full, 1);
    tc::mbar_init(qdo_empty, 1);
    tc::mbar_init(sdp_full, 1);
    tc::mbar_init(sdp_free, 256);
    tc::mbar_init(pds_full, 256);
    tc::mbar_init(dq_full, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc<1>(tmem_holder, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const uint32_t tmem_S = tmem_base;              // 128 cols; reused for dQ (HD cols)
  const uint32_t tmem_dP = tmem_base + 128;       // 128 cols
  const uint32_t tmem_dV = tmem_base + 256;       // HD cols
  const uint32_t tmem_dK = tmem_base + 256 + HD;  // HD cols

  if (warp == 0) {
    if (lane == 0 && n_iters > 0) {
      tc::mbar_expect_tx(kv_full, 2 * T_BYTES);
#pragma unroll
      for (int c = 0; c < NB; ++c) tc::tma_load_4d(sK + c * 16384, &tmK, kv_full, c * 64, hk, kv0, b);
#pragma unroll
      for (int c = 0; c < NB; ++c) tc::tma_load_4d(sV + c * 16384, &tmV, kv_full, c * 64, hk, kv0, b);
      for (int it = 0; it < n_iters; ++it) {
        const int g = it / iters_per_head, i = i_begin + it % iters_per_head;
        const int h = hk * G + g;
        tc::mbar_wait(qdo_empty, (it & 1) ^ 1);
        tc::mbar_expect_tx(qdo_full, 2 * T_BYTES);
#pragma unroll
        for (int c = 0; c < NB; ++c) tc::tma_load_4d(sQ + c * 16384, &tmQ, qdo_full, c * 64, h, i * BQ, b);
#pragma unroll
        for (int c = 0; c < NB; ++c) tc::tma_load_4d(sdO + c * 16384, &tmdO, qdo_full, c * 64, h, i * BQ, b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && n_iters > 0) {
      constexpr uint32_t idesc_s = tc::make_idesc_bf16(128, 128, 0, 0);     // S, dP : K-major x K-major
      constexpr uint32_t idesc_t = tc::make_idesc_bf16(128, HD, 1, 1);      // dV, dK: MN-major x MN-major
      constexpr uint32_t idesc_q = tc::make_idesc_bf16(128, HD, 0, 1);      // dQ    : K-major x MN-major
      const uint32_t aK = tc::smem_u32(sK), aV = tc::smem_u32(sV), aQ = tc::smem_u32(sQ), adO = tc::smem_u32(sdO);
      const uint32_t aP = tc::smem_u32(sP), adS = tc::smem_u32(sdS);
      tc::mbar_wait(kv_full, 0);
      for (int it = 0; it < n_iters; ++it) {
        tc::mbar_wait(qdo_full, it & 1);
        tc::mbar_wait(sdp_free, (it & 1) ^ 1);
        tc::tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) {
          const uint32_t off = (ks >> 2) * 16384 + (ks & 3) * 32;
          tc::umma_bf16<1>(tmem_S, tc::make_smem_desc(aQ + off, 16, 1024), tc::make_smem_desc(aK + off, 16, 1024),
                           idesc_s, ks > 0 ? 1u : 0u);
        }
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) {
          const uint32_t off = (ks >> 2) * 16384 + (ks & 3) * 32;
          tc::umma_bf16<1>(tmem_dP, tc::make_smem_desc(adO + off, 16, 1024), tc::make_smem_desc(aV + off, 16, 1024),
                           idesc_s, ks > 0 ? 1u : 0u);
        }
        tc::umma_commit(sdp_full);
        tc::mbar_wait(pds_full, it & 1);
        tc::tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {   // contraction over the 128 query rows
          tc::umma_bf16<1>(tmem_dV, tc::make_smem_desc(aP + ks * 2048, 16384, 1024),
                           tc::make_smem_desc(adO + ks * 2048, 16384, 1024), idesc_t, (it > 0 || ks > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          tc::umma_bf16<1>(tmem_dK, tc::make_smem_desc(adS + ks * 2048, 16384, 1024),
                           tc::make_smem_desc(aQ + ks * 2048, 16384, 1024), idesc_t, (it > 0 || ks > 0) ? 1u : 0u);
        }
        // Q and dO are dead once dV / dK retire: release them now so the TMA load of the next query tile overlaps the
        // dQ MMA and its epilogue (there is no shared memory left for a second Q/dO stage)
        tc::umma_commit(qdo_empty);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {   // dQ = dS K : contraction over the 128 kv rows
          const uint32_t off = (ks >> 2) * 16384 + (ks & 3) * 32;
          tc::umma_bf16<1>(tmem_S, tc::make_smem_desc(adS + off, 16, 1024), tc::make_smem_desc(aK + ks * 2048, 16384, 1024),
                           idesc_q, ks > 0 ? 1u : 0u);
        }
        tc::umma_commit(dq_full);
      }
    }
  } else {
    // Row-owner threads, two warpgroups per row: thread (wg, r) handles key columns [64 wg, 64 wg + 64) of
    // S / dP (-> P, dS in shared memory) and columns [wg HD/2, (wg+1) HD/2) of dQ, dK, dV.
    const int wg = (warp - 2) >> 2;
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const float scale2 = p.scale * kLog2e;
    constexpr int OC = HD / 2;
    for (int it = 0; it < n_iters; ++it) {
      const int g = it / iters_per_head, i = i_begin + it % iters_per_head;
      const int h = hk * G + g;
      const int q0 = i * BQ;
      const int q_idx = q0 + r;
      const bool row_valid = q_idx < len;
      float neg_lse2 = 0.f, dlt = 0.f;
      if (row_valid) {
        const int64_t o = ((int64_t)b * p.Hq + h) * p.T + q_idx;
        neg_lse2 = -p.lse[o] * kLog2e;
        dlt = p.delta[o];
      }
      int tile_lo = 0, row_lo = 0;
      if (p.seg_start != nullptr) {
        tile_lo = p.seg_start[(int64_t)b * p.T + min(q0 + BQ - 1, max(len - 1, 0))];     // largest start in this query tile
        if (row_valid) row_lo = p.seg_start[(int64_t)b * p.T + q_idx];
      }
      const bool need_mask = (p.causal && (kv0 + BKV - 1 > q0)) || (kv0 + BKV > len) || (q0 + BQ > len) || (kv0 < tile_lo);
      int limit = 64, lo = 0;                      // valid columns of this thread's 64: [lo, limit)
      if (need_mask) {
        int lim = len;
        if (p.causal) lim = min(lim, q_idx + 1);
        limit = row_valid ? lim - (kv0 + wg * 64) : 0;
        lo = row_lo - (kv0 + wg * 64);
      }
      tc::mbar_wait(sdp_full, it & 1);
      tc::tc_fence_after();
      if (it > 0) {     // the previous iteration's dQ reduce must have finished READING the staging area
        if (threadIdx.x == 64) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("bar.sync 2, 256;" ::: "memory");
      }
      uint8_t* prow = sP + wg * 16384 + r * 128;
      uint8_t* drow = sdS + wg * 16384 + r * 128;
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t s[32], dp[32];
        tc::tmem_ld_32x32(tmem_S + lane_off + wg * 64 + ch * 32, s);
        tc::tmem_ld_32x32(tmem_dP + lane_off + wg * 64 + ch * 32, dp);
        tc::tmem_ld_wait();
        uint32_t pk[16], dk_[16];
        if (!need_mask) {
#pragma unroll
          for (int jj = 0; jj < 32; jj += 2) {
            const float p0 = ex2_approx(fmaf(__uint_as_float(s[jj]), scale2, neg_lse2));
            const float p1 = ex2_approx(fmaf(__uint_as_float(s[jj + 1]), scale2, neg_lse2));
            const float d0 = p0 * (__uint_as_float(dp[jj]) - dlt) * p.scale;
            const float d1 = p1 * (__uint_as_float(dp[jj + 1]) - dlt) * p.scale;
            pk[jj >> 1] = pack_bf16x2(p0, p1);
            dk_[jj >> 1] = pack_bf16x2(d0, d1);
          }
        } else {
#pragma unroll
          for (int jj = 0; jj < 32; jj += 2) {
            const int e = ch * 32 + jj;
            const float p0 = (e < limit && e >= lo) ? ex2_approx(fmaf(__uint_as_float(s[jj]), scale2, neg_lse2)) : 0.f;
            const float p1 = (e + 1 < limit && e + 1 >= lo) ? ex2_approx(fmaf(__uint_as_float(s[jj + 1]), scale2, neg_lse2)) : 0.f;
            const float d0 = p0 * (__uint_as_float(dp[jj]) - dlt) * p.scale;
            const float d1 = p1 * (__uint_as_float(dp[jj + 1]) - dlt) * p.scale;
            pk[jj >> 1] = pack_bf16x2(p0, p1);
            dk_[jj >> 1] = pack_bf16x2(d0, d1);
          }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int off = (((ch * 4 + c) ^ (r & 7)) << 4);
          *reinterpret_cast<uint4*>(prow + off) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
          *reinterpret_cast<uint4*>(drow + off) = make_uint4(dk_[4 * c], dk_[4 * c + 1], dk_[4 * c + 2], dk_[4 * c + 3]);
        }
      }
      tc::tc_fence_before();
      tc::fence_proxy_async();
      tc::mbar_arrive(pds_full);
      // dQ tile -> staged in shared memory (the P / dS buffers are idle once this iteration's MMAs retired)
      // -> ONE bulk reduce-add per 32-column block into the fp32 accumulator.  Per-lane red.global.add was
      // ~1 lane/clk/SM: 16 K atomics per iteration dominated the kernel.
      tc::mbar_wait(dq_full, it & 1);
      tc::tc_fence_after();
#pragma unroll
      for (int ch = 0; ch < OC / 32; ++ch) {
        uint32_t v[32];
        tc::tmem_ld_32x32(tmem_S + lane_off + wg * OC + ch * 32, v);
        tc::tmem_ld_wait();
        uint8_t* blk = sP + (wg * (OC / 32) + ch) * 16384 + r * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c)
          *reinterpret_cast<uint4*>(blk + ((c ^ (r & 7)) << 4)) = make_uint4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
      }
      tc::tc_fence_before();
      tc::mbar_arrive(sdp_free);               // TMEM S / dP may be overwritten by the next iteration's MMAs
      tc::fence_proxy_async();
      asm volatile("bar.sync 2, 256;" ::: "memory");
      if (threadIdx.x == 64) {
#pragma unroll
        for (int blk = 0; blk < HD / 32; ++blk) tc::tma_reduce_add_4d(&tmdQ, sP + blk * 16384, blk * 32, h, q0, b);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    if (threadIdx.x == 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    // dK / dV of this tile
    const int kv_idx = kv0 + r;
    bf16* pdk = p.dk + (int64_t)b * p.dkv_sb + (int64_t)kv_idx * p.dkv_st + (int64_t)hk * p.dkv_sh + wg * OC;
    bf16* pdv = p.dv + (int64_t)b * p.dkv_sb + (int64_t)kv_idx * p.dkv_st + (int64_t)hk * p.dkv_sh + wg * OC;
    if (n_iters == 0) {
      if (kv_idx < p.T) {
        const float zf[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        const bf16x8 z = pack8(zf);
#pragma unroll
        for (int c = 0; c < OC; c += 8) {
          st8(pdk + c, z);
          st8(pdv + c, z);
        }
      }
    } else {
      // the last dq_full wait above already ordered every MMA (commits complete in issue order)
#pragma unroll
      for (int ch = 0; ch < OC / 32; ++ch) {
        uint32_t v[32], u[32];
        tc::tmem_ld_32x32(tmem_dV + lane_off + wg * OC + ch * 32, v);
        tc::tmem_ld_32x32(tmem_dK + lane_off + wg * OC + ch * 32, u);
        tc::tmem_ld_wait();
        if (kv_idx < p.T) {
#pragma unroll
          for (int c = 0; c < 32; c += 8) {
            float f[8], e[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) {
              f[t] = __uint_as_float(v[c + t]);
              e[t] = __uint_as_float(u[c + t]);
            }
            st8(pdv + ch * 32 + c, pack8(f));
            st8(pdk + ch * 32 + c, pack8(e));
          }
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc<1>(tmem_base, 512);
  }
}

// ============================================================================== decode
// One block per (batch, query head): scores over the cached keys, softmax, weighted sum of values.
template <int HD>
__global__ void __launch_bounds__(128)
attn_decode_kernel(const bf16* __restrict__ q, const bf16* __restrict__ kc, const bf16* __restrict__ vc,
                   bf16* __restrict__ o, const int32_t* __restrict__ lens, int S_max, int Hq, int Hk, float scale) {
  extern __shared__ float sc[];     // [S_max] scores, then [4][HD] partial outputs
  __shared__ float red[33];
  const int h = blockIdx.x, b = blockIdx.y;
  const int hk = h / (Hq / Hk);
  const int len = min(lens[b], S_max);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bf16* qp = q + ((int64_t)b * Hq + h) * HD;
  // each lane keeps HD/32 query elements
  float qv[HD / 32];
#pragma unroll
  for (int i = 0; i < HD / 32; ++i) qv[i] = __bfloat162float(qp[lane * (HD / 32) + i]);
  float mx = -INFINITY;
  for (int s = warp; s < len; s += 4) {
    const bf16* kp = kc + (((int64_t)b * S_max + s) * Hk + hk) * HD + lane * (HD / 32);
    float d = 0.f;
#pragma unroll
    for (int i = 0; i < HD / 32; ++i) d += qv[i] * __bfloat162float(kp[i]);
    d = warp_sum(d) * scale;
    if (lane == 0) sc[s] = d;
    mx = fmaxf(mx, d);
  }
  __syncthreads();
  // block max / sum
  __shared__ float smx[4];
  if (lane == 0) smx[warp] = mx;
  __syncthreads();
  mx = fmaxf(fmaxf(smx[0], smx[1]), fmaxf(smx[2], smx[3]));
  float sum = 0.f;
  for (int s = threadIdx.x; s < len; s += 128) {
    const float e = __expf(sc[s] - mx);
    sc[s] = e;
    sum += e;
  }
  sum = block_sum(sum, red);
  __syncthreads();
  // output: thread t owns dims [t*HD/128 ...) — use all 128 threads over HD dims x 128/HD kv slices
  constexpr int SL = 128 / HD > 0 ? 128 / HD : 1;   // kv slices per dim
  const int d = threadIdx.x % HD, sl = threadIdx.x / HD;
  float acc = 0.f;
  if (sl < SL) {
    for (int s = sl; s < len; s += SL) acc += sc[s] * __bfloat162float(vc[(((int64_t)b * S_max + s) * Hk + hk) * HD + d]);
  }
  float* part = sc + S_max;
  if (sl < SL) part[sl * HD + d] = acc;
  __syncthreads();
  if (threadIdx.x < HD) {
    float t = 0.f;
    for (int x = 0; x < SL; ++x) t += part[x * HD + threadIdx.x];
    o[((int64_t)b * Hq + h) * HD + threadIdx.x] = __float2bfloat16(len > 0 ? t / sum : 0.f);
  }
}

}  // namespace attn
}  // namespace dla

using namespace dla;
using namespace dla::attn;

static int g_fwd_variant = 0;     // 0 = auto (ping-pong kernel), 1 = first-generation kernel, 2 = ping-pong kernel
// 0 = auto: the pipelined kernel (attn_bwd2.cu) from T = 2048 on, the first-generation kernel below (measured on B200,
// 32/8 heads, D = 128: 307 vs 335 us at B=8 T=1024, 1006 vs 913 us at B=2 T=4096); 1 = first generation; 2 = pipelined
static int g_bwd_variant = 0;

extern "C" {

int dla_attn_set_fwd_variant(int v) {
  g_fwd_variant = v;
  return 0;
}

int dla_attn_set_bwd_variant(int v) {
  g_bwd_variant = v;
  return 0;
}

int dla_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int T, int Hq, int Hk, int D,
                 int64_t q_sb, int64_t q_st, int64_t q_sh, int64_t kv_sb, int64_t kv_st, int64_t kv_sh, int64_t o_sb,
                 int64_t o_st, int64_t o_sh, const int32_t* seqlens, float scale, int causal, void* stream) {
  return dla_attn_fwd_seg(q, k, v, o, lse, B, T, Hq, Hk, D, q_sb, q_st, q_sh, kv_sb, kv_st, kv_sh, o_sb, o_st, o_sh, seqlens,
                          nullptr, scale, causal, stream);
}

int dla_attn_fwd_seg(const void* q, const void* k, const void* v, void* o, float* lse, int B, int T, int Hq, int Hk, int D,
                     int64_t q_sb, int64_t q_st, int64_t q_sh, int64_t kv_sb, int64_t kv_st, int64_t kv_sh, int64_t o_sb,
                     int64_t o_st, int64_t o_sh, const int32_t* seqlens, const int32_t* seg_start, float scale, int causal,
                     void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(seg_start == nullptr || (causal && g_fwd_variant != 1 && g_fwd_variant != 3), DLA_E_UNSUPP,
              "attn_fwd: packed rows (seg_start) need causal attention and the ping-pong forward kernel");
  DLA_REQUIRE(D == 64 || D == 128, DLA_E_UNSUPP, "attn_fwd: head_dim %d not supported (64 or 128)", D);
  DLA_REQUIRE(Hq % Hk == 0 && B > 0 && T > 0, DLA_E_BADARG, "attn_fwd: bad shape B=%d T=%d Hq=%d Hk=%d", B, T, Hq, Hk);
  DLA_REQUIRE(o_st % 8 == 0 && o_sh % 8 == 0 && o_sb % 8 == 0, DLA_E_BADARG, "attn_fwd: output strides must be multiples of 8");
  CUtensorMap tmQ, tmK, tmV;
  int rc;
  if ((rc = make_qkv_map(&tmQ, q, B, T, Hq, D, q_sb, q_st, q_sh))) return rc;
  if ((rc = make_qkv_map(&tmK, k, B, T, Hk, D, kv_sb, kv_st, kv_sh))) return rc;
  if ((rc = make_qkv_map(&tmV, v, B, T, Hk, D, kv_sb, kv_st, kv_sh))) return rc;
  Params p{};
  p.B = B; p.T = T; p.Hq = Hq; p.Hk = Hk; p.D = D; p.scale = scale; p.causal = causal; p.seqlens = seqlens;
  p.seg_start = seg_start;
  p.o = (bf16*)o; p.o_sb = o_sb; p.o_st = o_st; p.o_sh = o_sh; p.lse = lse;
  p.debug_flags = g_fwd_variant >= 16 ? (g_fwd_variant & 15) : 0;     // A/B switches of the ping-pong kernel (bit 2: two-pass softmax only)
  dim3 grid((T + BQ - 1) / BQ, Hq, B);
  cudaStream_t s = (cudaStream_t)stream;
  if (g_fwd_variant == 3) return attn_fwd3_dispatch(tmQ, tmK, tmV, p, s);     // early-S ping-pong kernel (attn_fwd3.cu)
  if (g_fwd_variant != 1) return attn_fwd2_dispatch(tmQ, tmK, tmV, p, s);     // ping-pong kernel (attn_fwd2.cu)
  if (D == 128) {
    constexpr int smem = 128 * 128 * 2 + 2 * 2 * 128 * 128 * 2 + 128 * 128 * 2 + 1024 + 4096;
    static bool set = false;
    if (!set) { DLA_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); set = true; }
    attn_fwd_kernel<128><<<grid, kFwdThreads, smem, s>>>(tmQ, tmK, tmV, p);
  } else {
    constexpr int smem = 128 * 64 * 2 + 2 * 2 * 128 * 64 * 2 + 128 * 128 * 2 + 1024 + 4096;
    static bool set = false;
    if (!set) { DLA_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); set = true; }
    attn_fwd_kernel<64><<<grid, kFwdThreads, smem, s>>>(tmQ, tmK, tmV, p);
  }
  return check_launch("attn_fwd");
}

int dla_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout, const float* lse, void* dq,
                 void* dk, void* dv, float* dq_acc, float* delta, int B, int T, int Hq, int Hk, int D, int64_t q_sb,
                 int64_t q_st, int64_t q_sh, int64_t kv_sb, int64_t kv_st, int64_t kv_sh, int64_t o_sb, int64_t o_st,
                 int64_t o_sh, const int32_t* seqlens, float scale, int causal, void* stream) {
  return dla_attn_bwd_seg(q, k, v, o, dout, lse, dq, dk, dv, dq_acc, delta, B, T, Hq, Hk, D, q_sb, q_st, q_sh, kv_sb, kv_st, kv_sh,
                          o_sb, o_st, o_sh, seqlens, nullptr, scale, causal, stream);
}

int dla_attn_bwd_seg(const void* q, const void* k, const void* v, const void* o, const void* dout, const float* lse, void* dq,
                     void* dk, void* dv, float* dq_acc, float* delta, int B, int T, int Hq, int Hk, int D, int64_t q_sb,
                     int64_t q_st, int64_t q_sh, int64_t kv_sb, int64_t kv_st, int64_t kv_sh, int64_t o_sb, int64_t o_st,
                     int64_t o_sh, const int32_t* seqlens, const int32_t* seg_start, float scale, int causal, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(seg_start == nullptr || causal, DLA_E_UNSUPP, "attn_bwd: packed rows (seg_start) need causal attention");
  DLA_REQUIRE(D == 64 || D == 128, DLA_E_UNSUPP, "attn_bwd: head_dim %d not supported (64 or 128)", D);
  DLA_REQUIRE(Hq % Hk == 0 && B > 0 && T > 0, DLA_E_BADARG, "attn_bwd: bad shape");
  CUtensorMap tmQ, tmK, tmV, tmdO;
  int rc;
  if ((rc = make_qkv_map(&tmQ, q, B, T, Hq, D, q_sb, q_st, q_sh))) return rc;
  if ((rc = make_qkv_map(&tmK, k, B, T, Hk, D, kv_sb, kv_st, kv_sh))) return rc;
  if ((rc = make_qkv_map(&tmV, v, B, T, Hk, D, kv_sb, kv_st, kv_sh))) return rc;
  if ((rc = make_qkv_map(&tmdO, dout, B, T, Hq, D, o_sb, o_st, o_sh))) return rc;
  CUtensorMap tmdQ;
  {   // fp32 accumulator [B, T, Hq, D], contiguous
    const uint64_t dims[4] = {(uint64_t)D, (uint64_t)Hq, (uint64_t)T, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)D, (uint64_t)Hq * D, (uint64_t)T * Hq * D};
    const uint32_t box[4] = {32, 1, 128, 1};
    if ((rc = make_tmap_4d(&tmdQ, dq_acc, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dims, strides, box))) return rc;
  }
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t rows = (int64_t)B * T * Hq;
  attn_bwd_prep_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, s>>>((const bf16*)o, (const bf16*)dout, delta, dq_acc, B, T, Hq, D,
                                                                  o_sb, o_st, o_sh);
  if ((rc = check_launch("attn_bwd.prep"))) return rc;
  Params p{};
  p.B = B; p.T = T; p.Hq = Hq; p.Hk = Hk; p.D = D; p.scale = scale; p.causal = causal; p.seqlens = seqlens;
  p.seg_start = seg_start;
  p.lse = const_cast<float*>(lse); p.delta = delta; p.dq_acc = dq_acc;
  // contract: dq has q's strides, dk/dv have k/v's strides
  p.dk = (bf16*)dk; p.dv = (bf16*)dv; p.dkv_sb = kv_sb; p.dkv_st = kv_st; p.dkv_sh = kv_sh;
  dim3 grid((unsigned)(((T + BKV - 1) / BKV) * Hk * B));
  p.debug_flags = g_bwd_variant >= 16 ? (g_bwd_variant & 15) : 0;
  const bool gen2 = (g_bwd_variant >= 2 || (g_bwd_variant == 0 && T >= 2048)) && T < 65536;
  p.dq_tiled = gen2 ? 1 : 0;
  if (gen2 && T % BQ != 0) {
    // the padded rows of each (batch, head)'s last query tile are part of the tiled accumulator: the pre-pass zeroed
    // only the first B*T*Hq*D floats, clear the rest
    const int64_t have = rows * D, need = (int64_t)B * Hq * ((T + BQ - 1) / BQ) * BQ * D;
    DLA_CUDA(cudaMemsetAsync(dq_acc + have, 0, (size_t)(need - have) * sizeof(float), s));
  }
  if (gen2) {
    if ((rc = attn_bwd2_dispatch(tmQ, tmK, tmV, tmdO, tmdQ, p, s))) return rc;
  } else if (D == 128) {
    constexpr int smem = 4 * 128 * 128 * 2 + 2 * 128 * 128 * 2 + 1024 + 256;
    static bool set = false;
    if (!set) { DLA_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); set = true; }
    attn_bwd_kernel<128><<<grid, kFwdThreads, smem, s>>>(tmQ, tmK, tmV, tmdO, tmdQ, p);
  } else {
    constexpr int smem = 4 * 128 * 64 * 2 + 2 * 128 * 128 * 2 + 1024 + 256;
    static bool set = false;
    if (!set) { DLA_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); set = true; }
    attn_bwd_kernel<64><<<grid, kFwdThreads, smem, s>>>(tmQ, tmK, tmV, tmdO, tmdQ, p);
  }
  if ((rc = check_launch("attn_bwd"))) return rc;
  int64_t blocks = (rows * (D / 8) + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 16) blocks = (int64_t)kNumSMs * 16;
  attn_bwd_dq_cast_kernel<<<(unsigned)blocks, 256, 0, s>>>(dq_acc, (bf16*)dq, rows, D, Hq, T, q_sb, q_st, q_sh, p.dq_tiled);
  return check_launch("attn_bwd.dq_cast");
}

int dla_attn_decode(const void* q, const void* kcache, const void* vcache, void* o, const int32_t* lens, int B, int S_max,
                    int Hq, int Hk, int D, float scale, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(D == 64 || D == 128, DLA_E_UNSUPP, "attn_decode: head_dim %d not supported (64 or 128)", D);
  DLA_REQUIRE(Hq % Hk == 0, DLA_E_BADARG, "attn_decode: Hq %% Hk != 0");
  const size_t smem = (size_t)(S_max + 4 * 128) * sizeof(float);
  DLA_REQUIRE(smem <= 200 * 1024, DLA_E_TOOLARGE, "attn_decode: S_max=%d exceeds the shared-memory score buffer", S_max);
  dim3 grid(Hq, B);
  cudaStream_t s = (cudaStream_t)stream;
  if (D == 128) {
    DLA_CUDA(cudaFuncSetAttribute(attn_decode_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attn_decode_kernel<128><<<grid, 128, smem, s>>>((const bf16*)q, (const bf16*)kcache, (const bf16*)vcache, (bf16*)o, lens, S_max, Hq, Hk, scale);
  } else {
    DLA_CUDA(cudaFuncSetAttribute(attn_decode_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attn_decode_kernel<64><<<grid, 128, smem, s>>>((const bf16*)q, (const bf16*)kcache, (const bf16*)vcache, (bf16*)o, lens, S_max, Hq, Hk, scale);
  }
  return check_launch("attn_decode");
}

}  // extern "C"
