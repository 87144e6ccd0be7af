// Flash attention backward, second generation: transposed score tiles, P^T / dS^T consumed from TMEM, and a
// software pipeline in which the softmax threads and the tensor core always have work from different stages.
//
// The first kernel (attention.cu) ran every query tile as one serial chain — S, dP -> P, dS through shared
// memory -> dV, dK, dQ -> dQ staging — with single-buffered Q / dO, so the tensor pipe idled during the exponentials
// and the other way round (ncu: 16 % tensor activity, 231 TFLOP/s at T = 1024).  Here, per CTA = one 128-row key/value
// tile of one KV head (looping over the group's query heads and the query tiles that can see it):
//   * the score tiles are computed TRANSPOSED: S^T = K Q^T and dP^T = V dO^T, [kv lanes x q columns] in TMEM.  One
//     thread owns one key row; the per-query statistics (log-sum-exp, delta, document start) are broadcast reads from
//     shared memory.
//   * P^T and dS^T are written back over their own fp32 columns as packed bf16 (tcgen05.st) and feed
//     dV += P^T dO and dK += dS^T Q as the A operand FROM TMEM — no shared-memory round trip, no P buffer at all.
//     Only dS goes to shared memory (once, in the layout it is produced in: it is the MN-major A operand of dQ = dS K).
//   * the shared memory this frees holds a second Q / dO stage, so the next query tile's loads and its S^T MMA run
//     under the current tile's work.
//   * MMA order per query tile i:  dV(i), S^T(i+1) | dK(i), dQ(i) | dP^T(i+1).  The softmax threads compute
//     P(i+1) = exp2(S^T(i+1) - lse) — the MUFU-bound half — while dK(i) / dQ(i) run, drain dQ(i) (TMEM -> shared
//     memory -> one bulk reduce-add per 32 columns into the fp32 accumulator) while dP^T(i+1) runs, and finish with the
//     cheap half dS = P (dP - delta).
//   * the dQ reduce-add is asynchronous: the engine reads a staging buffer only as fast as L2 absorbs the atomics
//     (~2500 cycles for 64 KB when measured synchronously — as long as all five MMAs of a tile), so nobody waits for
//     it at the point of issue.  Half of dQ is staged in a buffer of its own, the other half in the dS buffer; the
//     "buffer has been read" waits sit immediately before each buffer's next writer (one tile later for the first,
//     the dS store of the next tile for the second).
// TMEM (512 columns): S^T / P^T 128 | dP^T / dS^T / dQ 128 | dV HD | dK HD.
// Shared memory at HD = 128: K, V 64 KB + 2 x Q 64 KB + dO 32 KB + dS 32 KB + dQ staging 32 KB.  Q is double-buffered
// because S^T(i+1) is issued a whole tile ahead; dO(i+1) is only needed by dP^T(i+1), two MMAs after dV(i) released it.
//
// Reference path replaced: the SDPA backward under HF Llama (R/src/training/train_dpo.py:118).
#include <math.h>

#include "attn_common.cuh"

namespace dla {
namespace attn {

constexpr int kBwd2Threads = 320;     // warp 0 loader, warp 1 MMA issuer, warps 2-9 two softmax warpgroups

template <int HD>
__global__ void __launch_bounds__(kBwd2Threads, 1)
attn_bwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                 const __grid_constant__ CUtensorMap tmdQ, const Params p) {
  constexpr int NB = HD / 64;
  constexpr int T_BYTES = 128 * HD * 2;       // one [128 x HD] bf16 operand tile
  constexpr int DS_BYTES = 128 * 128 * 2;     // dS^T tile, also 2 x [128 x 32] fp32 dQ staging slots
  constexpr int OC = HD / 2;                  // dQ / dK / dV columns per warpgroup
  constexpr int NPASS = HD / 64;              // dQ drain passes (32 columns per warpgroup per pass)

  // The HD = 128 layout fills the 227 KB an SM offers to within 400 bytes, so there is no room for the usual
  // "round the base up to 1024" slack: the kernel has no static shared memory, the dynamic window therefore starts on
  // a 1024-byte boundary of the shared address space (checked below; the swizzled TMA / UMMA layouts depend on it).
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((tc::smem_u32(smem) & 1023u) != 0u) __trap();
  uint8_t* sK = smem;
  uint8_t* sV = sK + T_BYTES;
  uint8_t* sQ = sV + T_BYTES;                 // [2][T_BYTES]
  uint8_t* sdO = sQ + 2 * T_BYTES;            // [T_BYTES]
  uint8_t* sdS = sdO + T_BYTES;               // [DS_BYTES]  dS^T; second dQ staging half when HD = 128
  uint8_t* sX = sdS + DS_BYTES;               // [DS_BYTES]  first dQ staging half: 2 x [128 x 32] fp32
  float* s_nlse2 = reinterpret_cast<float*>(sX + DS_BYTES);    // [2][128]  -lse * log2(e) of the stage's query rows
  float* s_delta = s_nlse2 + 256;                              // [2][128]
  uint16_t* s_seg = reinterpret_cast<uint16_t*>(s_delta + 256);   // [2][128]  document start of each query row (T < 65536)
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_seg + 256);
  uint64_t* kv_full = bars;          // 1
  uint64_t* q_full = bars + 1;       // [2] Q of a stage landed (tx bytes)
  uint64_t* q_empty = bars + 3;      // [2] Q stage may be refilled (MMA commit after dK)
  uint64_t* aux_full = bars + 5;     // [2] lse / delta / seg of a stage written (32 arrivals)
  uint64_t* s_full = bars + 7;       // S^T(i) complete
  uint64_t* dp_full = bars + 8;      // dP^T(i) complete
  uint64_t* p_ready = bars + 9;      // P^T(i) in TMEM (256 arrivals)
  uint64_t* ds_ready = bars + 10;    // dS^T(i) in TMEM: dK may be issued (256 arrivals)
  uint64_t* dq_full = bars + 11;     // dQ(i) complete
  uint64_t* dq_drained = bars + 12;  // dQ(i) read out of TMEM (256 arrivals)
  uint64_t* do_full = bars + 13;     // dO(i) landed (tx bytes)
  uint64_t* do_empty = bars + 14;    // dO buffer may be refilled (MMA commit after dV)
  uint64_t* dss_ready = bars + 15;   // dS^T(i) in shared memory: dQ may be issued (256 arrivals)
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D grid, longest-processing-time first (kv tile 0 sees the most query tiles)
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
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&q_full[s], 1);
      tc::mbar_init(&q_empty[s], 1);
      tc::mbar_init(&aux_full[s], 32);
    }
    tc::mbar_init(s_full, 1);
    tc::mbar_init(dp_full, 1);
    tc::mbar_init(p_ready, 256);
    tc::mbar_init(ds_ready, 256);
    tc::mbar_init(dq_full, 1);
    tc::mbar_init(dq_drained, 256);
    tc::mbar_init(do_full, 1);
    tc::mbar_init(do_empty, 1);
    tc::mbar_init(dss_ready, 256);
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc<1>(tmem_holder, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const uint32_t tmem_S = tmem_base;              // S^T fp32 [0,128); P^T bf16 pairs in [0,32) and [64,96)
  const uint32_t tmem_dP = tmem_base + 128;       // dP^T fp32; dS^T bf16 pairs in [+0,+32) and [+64,+96); dQ fp32 [+0,+HD)
  const uint32_t tmem_dV = tmem_base + 256;
  const uint32_t tmem_dK = tmem_base + 256 + HD;

  if (warp == 0) {
    // ------------------------------------------------------------------ loader: TMA for K, V, Q, dO; per-query
    // statistics of the stage through plain loads (whole warp, 4 query rows per lane)
    if (n_iters > 0) {
      if (lane == 0) {
        tc::mbar_expect_tx(kv_full, 2 * T_BYTES);
#pragma unroll
        for (int c = 0; c < NB; ++c) tc::tma_load_4d(sK + c * 16384, &tmK, kv_full, c * 64, hk, kv0, b);
#pragma unroll
        for (int c = 0; c < NB; ++c) tc::tma_load_4d(sV + c * 16384, &tmV, kv_full, c * 64, hk, kv0, b);
      }
      for (int it = 0; it < n_iters; ++it) {
        const int g = it / iters_per_head, i = i_begin + it % iters_per_head;
        const int h = hk * G + g;
        const int st = it & 1;
        tc::mbar_wait(&q_empty[st], ((it >> 1) & 1) ^ 1);
        if (lane == 0) {
          tc::mbar_expect_tx(&q_full[st], T_BYTES);
#pragma unroll
          for (int c = 0; c < NB; ++c) tc::tma_load_4d(sQ + st * T_BYTES + c * 16384, &tmQ, &q_full[st], c * 64, h, i * BQ, b);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int qr = lane * 4 + k;
          const int q_idx = i * BQ + qr;
          float nl = 0.f, dl = 0.f;
          if (q_idx < len) {
            const int64_t o = ((int64_t)b * p.Hq + h) * p.T + q_idx;
            nl = -p.lse[o] * kLog2e;
            dl = p.delta[o];
          }
          // rows past the end repeat the last valid row's start, so entry 127 is always the tile's largest start
          int sg = 0;
          if (p.seg_start != nullptr && len > 0) sg = p.seg_start[(int64_t)b * p.T + min(q_idx, len - 1)];
          s_nlse2[st * 128 + qr] = nl;
          s_delta[st * 128 + qr] = dl;
          s_seg[st * 128 + qr] = (uint16_t)min(sg, 65535);
        }
        tc::mbar_arrive(&aux_full[st]);
        // dO(it): its single buffer is released by dV(it-1); the waits of this loop follow the MMA order
        // dV(0), dK(0), dV(1), dK(1), ... so none of them holds back a load that could already start
        tc::mbar_wait(do_empty, (it & 1) ^ 1);
        if (lane == 0) {
          tc::mbar_expect_tx(do_full, T_BYTES);
#pragma unroll
          for (int c = 0; c < NB; ++c) tc::tma_load_4d(sdO + c * 16384, &tmdO, do_full, c * 64, h, i * BQ, b);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0 && n_iters > 0) {
      constexpr uint32_t idesc_s = tc::make_idesc_bf16(128, 128, 0, 0);     // S^T, dP^T: K-major x K-major
      constexpr uint32_t idesc_ts = tc::make_idesc_bf16(128, HD, 0, 1);     // dV, dK   : A from TMEM x MN-major B
      constexpr uint32_t idesc_q = tc::make_idesc_bf16(128, HD, 1, 1);      // dQ       : MN-major A (dS^T) x MN-major B (K)
      const uint32_t aK = tc::smem_u32(sK), aV = tc::smem_u32(sV), adS = tc::smem_u32(sdS);
      auto issue_st = [&](uint32_t tmem_d, uint32_t aA, uint32_t aB) {       // D[kv, q] = A[kv, d] B[q, d]^T
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) {
          const uint32_t off = (ks >> 2) * 16384 + (ks & 3) * 32;
          tc::umma_bf16<1>(tmem_d, tc::make_smem_desc(aA + off, 16, 1024), tc::make_smem_desc(aB + off, 16, 1024), idesc_s,
                           ks > 0 ? 1u : 0u);
        }
      };
      auto issue_acc = [&](uint32_t tmem_d, uint32_t tmem_a, uint32_t aB, bool first) {   // D[kv, d] += A^T(TMEM)[kv, q] B[q, d]
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t ta = tmem_a + (ks < 4 ? ks * 8 : 64 + (ks - 4) * 8);
          tc::umma_bf16_ts(tmem_d, ta, tc::make_smem_desc(aB + ks * 2048, 16384, 1024), idesc_ts, (!first || ks > 0) ? 1u : 0u);
        }
      };
      const uint32_t adO = tc::smem_u32(sdO);
      tc::mbar_wait(kv_full, 0);
      tc::mbar_wait(&q_full[0], 0);
      tc::tc_fence_after();
      issue_st(tmem_S, aK, tc::smem_u32(sQ));
      tc::umma_commit(s_full);
      tc::mbar_wait(do_full, 0);
      tc::tc_fence_after();
      issue_st(tmem_dP, aV, adO);
      tc::umma_commit(dp_full);
      for (int it = 0; it < n_iters; ++it) {
        const int st = it & 1;
        const uint32_t aQ = tc::smem_u32(sQ + st * T_BYTES);
        const bool more = it + 1 < n_iters;
        tc::mbar_wait(p_ready, it & 1);
        tc::tc_fence_after();
        issue_acc(tmem_dV, tmem_S, adO, it == 0);                       // dV += P^T dO
        tc::umma_commit(do_empty);                                      // dO(i) is dead: dP^T(i) and dV(i) have read it
        if (more) {
          tc::mbar_wait(&q_full[st ^ 1], ((it + 1) >> 1) & 1);
          tc::tc_fence_after();
          issue_st(tmem_S, aK, tc::smem_u32(sQ + (st ^ 1) * T_BYTES));  // S^T(i+1): the pipe is in order, P^T(i) was consumed above
          tc::umma_commit(s_full);
        }
        tc::mbar_wait(ds_ready, it & 1);
        tc::tc_fence_after();
        issue_acc(tmem_dK, tmem_dP, aQ, it == 0);                       // dK += dS^T Q
        tc::umma_commit(&q_empty[st]);                                  // Q(i) is dead once S^T(i) and dK(i) have read it
        tc::mbar_wait(dss_ready, it & 1);                               // dS in shared memory (written after the TMEM copy)
        tc::tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)                                  // dQ = dS K (contraction over the 128 kv rows)
          tc::umma_bf16<1>(tmem_dP, tc::make_smem_desc(adS + ks * 2048, 16384, 1024), tc::make_smem_desc(aK + ks * 2048, 16384, 1024),
                           idesc_q, ks > 0 ? 1u : 0u);
        tc::umma_commit(dq_full);
        if (more) {
          tc::mbar_wait(dq_drained, it & 1);
          tc::mbar_wait(do_full, (it + 1) & 1);
          tc::tc_fence_after();
          issue_st(tmem_dP, aV, adO);                                    // dP^T(i+1)
          tc::umma_commit(dp_full);
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax / epilogue threads
    // Phases A and B: thread (wg, r) owns key row r and the query columns [64 wg, 64 wg + 64) of S^T / dP^T.
    // Drain and final epilogue: thread (wg, r) owns TMEM lane r and the head-dim columns [wg OC, (wg + 1) OC).
    const int wg = (warp - 2) >> 2;
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const float scale2 = p.scale * kLog2e;
    const int kv_idx = kv0 + r;
    float Pv[64];

    auto phase_a = [&](int it) {
      const int i = i_begin + it % iters_per_head;
      const int st = it & 1;
      const int qc0 = i * BQ + wg * 64;              // first query index of this thread's columns
      tc::mbar_wait(&aux_full[st], (it >> 1) & 1);
      const bool seg_mask = p.seg_start != nullptr && (int)s_seg[st * 128 + 127] > kv0;
      const bool need_mask = (p.causal && i == j) || (kv0 + BKV > len) || (i * BQ + BQ > len) || seg_mask;
      int c_lo = 0, c_hi = 64;
      if (need_mask) {
        if (p.causal) c_lo = max(0, kv_idx - qc0);
        c_hi = kv_idx < len ? min(64, len - qc0) : 0;
      }
      tc::mbar_wait(s_full, it & 1);
      tc::tc_fence_after();
      uint32_t pk[32];
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t s[32];
        tc::tmem_ld_32x32(tmem_S + lane_off + wg * 64 + ch * 32, s);
        tc::tmem_ld_wait();
        const float* nl = s_nlse2 + st * 128 + wg * 64 + ch * 32;
        if (!need_mask) {
#pragma unroll
          for (int jj = 0; jj < 32; jj += 4) {
            const float4 n4 = *reinterpret_cast<const float4*>(nl + jj);
            Pv[ch * 32 + jj] = ex2_approx(fmaf(__uint_as_float(s[jj]), scale2, n4.x));
            Pv[ch * 32 + jj + 1] = ex2_approx(fmaf(__uint_as_float(s[jj + 1]), scale2, n4.y));
            Pv[ch * 32 + jj + 2] = ex2_approx(fmaf(__uint_as_float(s[jj + 2]), scale2, n4.z));
            Pv[ch * 32 + jj + 3] = ex2_approx(fmaf(__uint_as_float(s[jj + 3]), scale2, n4.w));
          }
        } else {
          const uint16_t* sg = s_seg + st * 128 + wg * 64 + ch * 32;
#pragma unroll
          for (int jj = 0; jj < 32; ++jj) {
            const int c = ch * 32 + jj;
            const bool ok = c >= c_lo && c < c_hi && (!seg_mask || kv_idx >= (int)sg[jj]);
            Pv[c] = ok ? ex2_approx(fmaf(__uint_as_float(s[jj]), scale2, nl[jj])) : 0.f;
          }
        }
#pragma unroll
        for (int jj = 0; jj < 32; jj += 2) pk[ch * 16 + (jj >> 1)] = pack_bf16x2(Pv[ch * 32 + jj], Pv[ch * 32 + jj + 1]);
      }
      // both of this warpgroup's S^T chunks are in registers: P^T overwrites the FIRST 32 columns of its own 64
      tc::tmem_st_32x32(tmem_S + lane_off + wg * 64, pk);
      tc::tmem_st_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(p_ready);
    };

    auto phase_b = [&](int it) {
      const int st = it & 1;
      tc::mbar_wait(dp_full, it & 1);
      tc::tc_fence_after();
      uint32_t dk_[32];
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t dp[32];
        tc::tmem_ld_32x32(tmem_dP + lane_off + wg * 64 + ch * 32, dp);
        tc::tmem_ld_wait();
        const float* dl = s_delta + st * 128 + wg * 64 + ch * 32;
#pragma unroll
        for (int jj = 0; jj < 32; jj += 4) {
          const float4 d4 = *reinterpret_cast<const float4*>(dl + jj);
          const float e0 = Pv[ch * 32 + jj] * (__uint_as_float(dp[jj]) - d4.x) * p.scale;
          const float e1 = Pv[ch * 32 + jj + 1] * (__uint_as_float(dp[jj + 1]) - d4.y) * p.scale;
          const float e2 = Pv[ch * 32 + jj + 2] * (__uint_as_float(dp[jj + 2]) - d4.z) * p.scale;
          const float e3 = Pv[ch * 32 + jj + 3] * (__uint_as_float(dp[jj + 3]) - d4.w) * p.scale;
          dk_[ch * 16 + (jj >> 1)] = pack_bf16x2(e0, e1);
          dk_[ch * 16 + (jj >> 1) + 1] = pack_bf16x2(e2, e3);
        }
      }
      tc::tmem_st_32x32(tmem_dP + lane_off + wg * 64, dk_);
      tc::tmem_st_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(ds_ready);                // dK(i) can start from the TMEM copy while the shared-memory copy waits below
      if (NPASS == 2 && it > 0) {
        // the dS buffer staged the second half of dQ(it-1): its reduce-add must have READ it before it is overwritten
        if (threadIdx.x == 64) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("bar.sync 2, 256;" ::: "memory");
      }
      // dS^T row r, query columns [64 wg, 64 wg + 64): one 128-byte swizzled row of chunk `wg`
      uint8_t* drow = sdS + wg * 16384 + r * 128;
#pragma unroll
      for (int c = 0; c < 8; ++c)
        *reinterpret_cast<uint4*>(drow + ((c ^ (r & 7)) << 4)) = make_uint4(dk_[4 * c], dk_[4 * c + 1], dk_[4 * c + 2], dk_[4 * c + 3]);
      tc::fence_proxy_async();
      tc::mbar_arrive(dss_ready);
    };

    auto drain = [&](int it) {
      const int g = it / iters_per_head, i = i_begin + it % iters_per_head;
      const int h = hk * G + g;
      tc::mbar_wait(dq_full, it & 1);
      tc::tc_fence_after();
      uint32_t v[NPASS][32];
#pragma unroll
      for (int pass = 0; pass < NPASS; ++pass) tc::tmem_ld_32x32(tmem_dP + lane_off + wg * OC + pass * 32, v[pass]);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(dq_drained);             // the dP^T / dQ columns may be overwritten by dP^T(i+1)
      if (NPASS == 1 && it > 0) {
        // HD = 64: the only staging buffer is sX; the previous tile's reduce-add must have read it
        if (threadIdx.x == 64) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("bar.sync 2, 256;" ::: "memory");
      }
      // (HD = 128: phase B of this tile already waited for every earlier reduce-add, sX included)
#pragma unroll
      for (int pass = 0; pass < NPASS; ++pass) {
        uint8_t* blk = (pass == 0 ? sX : sdS) + wg * 16384 + r * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c)
          *reinterpret_cast<uint4*>(blk + ((c ^ (r & 7)) << 4)) =
              make_uint4(v[pass][4 * c], v[pass][4 * c + 1], v[pass][4 * c + 2], v[pass][4 * c + 3]);
      }
      tc::fence_proxy_async();
      asm volatile("bar.sync 2, 256;" ::: "memory");
      if (threadIdx.x == 64 && !(p.debug_flags & 1)) {
        // the fp32 accumulator is laid out as the staging image itself (one contiguous 16 KB block per 32 columns of a
        // query tile), so each reduce-add is a single linear 16 KB transfer instead of 128 strided 128-byte rows
        const int64_t tile = ((int64_t)b * p.Hq + h) * ((p.T + BQ - 1) / BQ) + i;
#pragma unroll
        for (int pass = 0; pass < NPASS; ++pass)
#pragma unroll
          for (int w = 0; w < 2; ++w)
            tc::bulk_reduce_add_f32(p.dq_acc + (tile * (HD / 32) + (w * NPASS + pass)) * 4096, (pass == 0 ? sX : sdS) + w * 16384, 16384);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    };

    if (n_iters > 0) phase_a(0);
    for (int it = 0; it < n_iters; ++it) {
      phase_b(it);
      if (it + 1 < n_iters) phase_a(it + 1);
      drain(it);
    }
    if (threadIdx.x == 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");

    // ---- dK / dV of this tile: lane r = key row, columns [wg OC, (wg + 1) OC)
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
      // the last dq_full wait ordered every MMA issued before it (commits complete in issue order)
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

template <int HD>
static int launch_bwd2(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV, const CUtensorMap& tmdO,
                       const CUtensorMap& tmdQ, const Params& p, cudaStream_t s) {
  constexpr int smem = 5 * 128 * HD * 2 + 2 * 128 * 128 * 2 + 2 * 256 * 4 + 256 * 2 + 256;   // tiles, dS, dQ staging, lse / delta / seg, barriers
  static bool set = false;
  if (!set) {
    DLA_CUDA(cudaFuncSetAttribute(attn_bwd2_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    set = true;
  }
  dim3 grid((unsigned)(((p.T + BKV - 1) / BKV) * p.Hk * p.B));
  attn_bwd2_kernel<HD><<<grid, kBwd2Threads, smem, s>>>(tmQ, tmK, tmV, tmdO, tmdQ, p);
  return check_launch("attn_bwd2");
}

int attn_bwd2_dispatch(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV, const CUtensorMap& tmdO,
                       const CUtensorMap& tmdQ, const Params& p, cudaStream_t s) {
  return p.D == 128 ? launch_bwd2<128>(tmQ, tmK, tmV, tmdO, tmdQ, p, s) : launch_bwd2<64>(tmQ, tmK, tmV, tmdO, tmdQ, p, s);
}

}  // namespace attn
}  // namespace dla
