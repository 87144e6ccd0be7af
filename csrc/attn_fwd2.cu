// Flash attention forward, second generation: two 128-row query tiles per CTA in ping-pong.
//
// The first kernel (attention.cu) handled one query tile per CTA, so every kv step was a serial chain
// S = QK^T -> softmax -> P (through shared memory) -> PV -> O read back into registers: the tensor pipe sat
// idle during the softmax and the other way round (ncu: 12-16% tensor activity).  Here
//   * a CTA owns query tiles A and B (256 rows of one head).  Warp 0 = TMA producer, warp 1 = MMA issuer,
//     warps 2-5 = softmax of tile A, warps 6-9 = softmax of tile B; one thread per query row, so row max and row
//     sum need no cross-thread traffic at all.
//   * the MMA warp alternates S_A, S_B, PV_A, S_A', PV_B, S_B', ...: while one tile is in its softmax, the
//     tensor pipe works on the other tile.
//   * TMEM holds everything the tensor core produces or consumes besides K/V: S_A, S_B (128 fp32 columns each),
//     O_A, O_B (HD columns each) = 512 columns.  P (bf16) is written by the softmax threads with tcgen05.st over
//     the first 64 columns of its own S and consumed as the A operand FROM TMEM — no shared-memory round trip.
//   * O stays in TMEM across the kv loop.  The running max only moves when a row's new maximum exceeds the
//     reference by more than 2^8 (then O and the running sum are rescaled through tcgen05.ld/st); otherwise the
//     stale reference is kept and P <= 256, exact after the final division by the row sum.
//   * S is read from TMEM twice (max pass, exp pass) instead of being held in 128 registers.
// MUFU (16 ex2/clk/SM) and the tensor pipe are balanced at ~50% tensor utilisation for HD = 128; that is the
// ceiling of this formulation without polynomial exp2 emulation.
//
// Reference path replaced: PyTorch SDPA under HF Llama (R/src/training/train_dpo.py:32).
#include <math.h>

#include "attn_common.cuh"

namespace dla {
namespace attn {

constexpr int kFwd2Threads = 320;
constexpr float kRescaleThreshold = 8.f;     // log2 units
// Half of the exponentials of unmasked tiles on the FMA pipe (ex2_poly) instead of the MUFU.  Measured on B200 and left OFF:
// 390 vs 327 us at B=2 T=4096, 672 vs 537 us at T=8192 — with two softmax warps per scheduler the loop is bound by
// instruction issue and dependency latency (issue active 36 %, stalls: long scoreboard, fixed-latency wait), not by MUFU
// throughput, so nine extra FMA/ALU instructions per element cost more than the MUFU slot they free.
constexpr bool kPolyExp = false;

template <int HD>
__global__ void __launch_bounds__(kFwd2Threads, 1)
attn_fwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const Params p) {
  constexpr int NB = HD / 64;
  constexpr int TILE_BYTES = 128 * HD * 2;      // one Q, K or V tile
  constexpr int kStages = 2;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;                                   // [2][TILE_BYTES]
  uint8_t* sK = sQ + 2 * TILE_BYTES;                    // [kStages][TILE_BYTES]
  uint8_t* sV = sK + kStages * TILE_BYTES;              // [kStages][TILE_BYTES]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + kStages * TILE_BYTES);
  uint64_t* q_full = bars;            // 1
  uint64_t* k_full = bars + 1;        // [2]
  uint64_t* k_empty = bars + 3;       // [2]
  uint64_t* v_full = bars + 5;        // [2]
  uint64_t* v_empty = bars + 7;       // [2]
  uint64_t* s_full = bars + 9;        // [2] per query tile
  uint64_t* p_full = bars + 11;       // [2] per query tile, 128 arrivals
  uint64_t* o_done = bars + 13;       // [2] per query tile
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D grid, longest-processing-time first: ALL (head, batch) instances of the heaviest query-tile pair are
  // scheduled before any lighter pair, so the tail of the launch consists of the shortest CTAs
  const int n_pairs = (p.T + 255) >> 8, n_hb = p.Hq * p.B;
  const int pi = n_pairs - 1 - (int)blockIdx.x / n_hb;
  const int hb = (int)blockIdx.x % n_hb;
  const int h = hb % p.Hq, b = hb / p.Hq;
  const int hk = h / (p.Hq / p.Hk);
  int len = p.T;
  if (p.seqlens != nullptr) len = min(len, p.seqlens[b]);
  const int n_all = (len + 127) >> 7;
  // packed rows: kv tiles that end before the document of the pair's FIRST query row starts are skipped for both
  // tiles (seg_start is non-decreasing, so that row has the smallest start); all loop counters below are relative
  // to that first tile `jb`
  int jb = 0;
  if (p.seg_start != nullptr && pi * 256 < len) jb = p.seg_start[(int64_t)b * p.T + pi * 256] >> 7;
  int nX[2];
#pragma unroll
  for (int x = 0; x < 2; ++x) {
    const int t = 2 * pi + x;
    int n = p.causal ? min(n_all, t + 1) : n_all;
    if (t * 128 >= len) n = 0;
    nX[x] = max(n - jb, 0);
  }
  const int n = max(nX[0], nX[1]);

  if (threadIdx.x == 0) {
    tc::prefetch_tmap(&tmQ);
    tc::prefetch_tmap(&tmK);
    tc::prefetch_tmap(&tmV);
    tc::mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&k_full[s], 1);
      tc::mbar_init(&k_empty[s], 1);
      tc::mbar_init(&v_full[s], 1);
      tc::mbar_init(&v_empty[s], 1);
      tc::mbar_init(&s_full[s], 1);
      tc::mbar_init(&p_full[s], 128);
      tc::mbar_init(&o_done[s], 1);
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc<1>(tmem_holder, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && n > 0) {
      tc::mbar_expect_tx(q_full, 2 * TILE_BYTES);
#pragma unroll
      for (int x = 0; x < 2; ++x)
#pragma unroll
        for (int c = 0; c < NB; ++c)
          tc::tma_load_4d(sQ + x * TILE_BYTES + c * 16384, &tmQ, q_full, c * 64, h, (2 * pi + x) * 128, b);
      for (int j = 0; j < n; ++j) {
        const int st = j & 1;
        const uint32_t ph = ((j >> 1) & 1) ^ 1;
        tc::mbar_wait(&k_empty[st], ph);
        tc::mbar_expect_tx(&k_full[st], TILE_BYTES);
#pragma unroll
        for (int c = 0; c < NB; ++c) tc::tma_load_4d(sK + st * TILE_BYTES + c * 16384, &tmK, &k_full[st], c * 64, hk, (jb + j) * 128, b);
        tc::mbar_wait(&v_empty[st], ph);
        tc::mbar_expect_tx(&v_full[st], TILE_BYTES);
#pragma unroll
        for (int c = 0; c < NB; ++c) tc::tma_load_4d(sV + st * TILE_BYTES + c * 16384, &tmV, &v_full[st], c * 64, hk, (jb + j) * 128, b);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0 && n > 0) {
      constexpr uint32_t idesc_s = tc::make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_pv = tc::make_idesc_bf16(128, HD, 0, 1);
      auto issue_s = [&](int x, int j) {           // S_x = Q_x K_j^T
        const uint32_t aQ = tc::smem_u32(sQ + x * TILE_BYTES), aK = tc::smem_u32(sK + (j & 1) * TILE_BYTES);
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) {
          const uint32_t off = (ks >> 2) * 16384 + (ks & 3) * 32;
          tc::umma_bf16<1>(tmem_base + x * 128, tc::make_smem_desc(aQ + off, 16, 1024), tc::make_smem_desc(aK + off, 16, 1024),
                           idesc_s, ks > 0 ? 1u : 0u);
        }
        tc::umma_commit(&s_full[x]);
      };
      auto issue_pv = [&](int x, int j) {          // O_x (+)= P_x V_j, P read from TMEM (bf16 pairs, 8 columns per k-step)
        const uint32_t aV = tc::smem_u32(sV + (j & 1) * TILE_BYTES);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          tc::umma_bf16_ts(tmem_base + 256 + x * 128, tmem_base + x * 128 + ks * 8, tc::make_smem_desc(aV + ks * 2048, 16384, 1024),
                           idesc_pv, (j > 0 || ks > 0) ? 1u : 0u);
        tc::umma_commit(&o_done[x]);
      };
      tc::mbar_wait(q_full, 0);
      tc::mbar_wait(&k_full[0], 0);
      tc::tc_fence_after();
      if (nX[0] > 0) issue_s(0, 0);
      if (nX[1] > 0) issue_s(1, 0);
      tc::umma_commit(&k_empty[0]);
      for (int j = 0; j < n; ++j) {
        const int st = j & 1;
        tc::mbar_wait(&v_full[st], (j >> 1) & 1);
        const bool more = j + 1 < n;
        if (more) tc::mbar_wait(&k_full[st ^ 1], ((j + 1) >> 1) & 1);
#pragma unroll
        for (int x = 0; x < 2; ++x) {
          if (j < nX[x]) {
            tc::mbar_wait(&p_full[x], j & 1);
            tc::tc_fence_after();
            issue_pv(x, j);
            if (j + 1 < nX[x]) issue_s(x, j + 1);
          }
        }
        tc::umma_commit(&v_empty[st]);
        if (more) tc::umma_commit(&k_empty[st ^ 1]);
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax of tile x, one thread per row
    const int x = (warp - 2) >> 2;
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
    const int r = quarter * 32 + lane;
    const int q0 = (2 * pi + x) * 128;
    const int q_idx = q0 + r;
    const int nt = nX[x];
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const uint32_t tS = tmem_base + lane_off + x * 128;
    const uint32_t tO = tmem_base + lane_off + 256 + x * 128;
    const float scale2 = p.scale * kLog2e;
    float m_ref = -INFINITY, l_run = 0.f;
    // first key this row may see (start of its document when rows are packed)
    const int row_lo = (p.seg_start != nullptr && q_idx < len) ? p.seg_start[(int64_t)b * p.T + q_idx] : 0;
    // the last row of the tile has the largest start: below it, some row of the tile needs the lower mask
    const int tile_lo = (p.seg_start != nullptr) ? p.seg_start[(int64_t)b * p.T + min(q0 + 127, max(len - 1, 0))] : 0;

    for (int j = 0; j < nt; ++j) {
      const int kv0 = (jb + j) * 128;
      tc::mbar_wait(&s_full[x], j & 1);
      tc::tc_fence_after();
      const bool need_mask = (p.causal && kv0 + 127 > q0) || (kv0 + 128 > len) || (kv0 < tile_lo);
      int limit = 128, lo = 0;                    // valid columns of this row in this tile: [lo, limit)
      if (need_mask) {
        int lim = len;
        if (p.causal) lim = min(lim, q_idx + 1);
        limit = lim - kv0;
        lo = row_lo - kv0;
      }
      float s4[4] = {0.f, 0.f, 0.f, 0.f};
      bool done = false;
      // ---- optimistic single pass (every kv tile after the first): exponentiate against the CURRENT reference while the
      // row maximum is tracked on the side; the next 32-column chunk is requested from TMEM before the current one is
      // processed, so the TMEM latency hides under the MUFU work.  P stays in registers until the maximum is known: if no
      // row of the warp moved by more than the rescale threshold (the usual case — the reference only has to be within 2^8)
      // it is committed as is; otherwise nothing was stored, S is still intact, and the two-pass path below redoes the tile.
      if (j > 0 && !__any_sync(0xffffffffu, m_ref == -INFINITY) && !(p.debug_flags & 4)) {
        const float neg_m = -m_ref;
        uint32_t pk[64];
        uint32_t v[2][32];
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        tc::tmem_ld_32x32(tS, v[0]);
        tc::tmem_ld_wait_for(v[0]);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          if (ch < 3) tc::tmem_ld_32x32(tS + (ch + 1) * 32, v[(ch + 1) & 1]);
          const uint32_t* cur = v[ch & 1];
          if (!need_mask) {
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const float a0 = __uint_as_float(cur[i]), a1 = __uint_as_float(cur[i + 1]);
              mx4[(i >> 1) & 3] = fmaxf(mx4[(i >> 1) & 3], fmaxf(a0, a1));
              // every other pair is exponentiated on the FMA pipe instead of the MUFU (see ex2_poly)
              const bool poly = kPolyExp && ((i >> 1) & 1);
              const float p0 = poly ? ex2_poly(fmaf(a0, scale2, neg_m)) : ex2_approx(fmaf(a0, scale2, neg_m));
              const float p1 = poly ? ex2_poly(fmaf(a1, scale2, neg_m)) : ex2_approx(fmaf(a1, scale2, neg_m));
              s4[(i >> 1) & 3] += p0 + p1;
              pk[ch * 16 + (i >> 1)] = pack_bf16x2(p0, p1);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const int c = ch * 32 + i;
              const bool ok0 = c < limit && c >= lo, ok1 = c + 1 < limit && c + 1 >= lo;
              const float a0 = ok0 ? __uint_as_float(cur[i]) : -INFINITY, a1 = ok1 ? __uint_as_float(cur[i + 1]) : -INFINITY;
              mx4[(i >> 1) & 3] = fmaxf(mx4[(i >> 1) & 3], fmaxf(a0, a1));
              const float p0 = ok0 ? ex2_approx(fmaf(a0, scale2, neg_m)) : 0.f;
              const float p1 = ok1 ? ex2_approx(fmaf(a1, scale2, neg_m)) : 0.f;
              s4[(i >> 1) & 3] += p0 + p1;
              pk[ch * 16 + (i >> 1)] = pack_bf16x2(p0, p1);
            }
          }
          if (ch < 3) tc::tmem_ld_wait_for(v[(ch + 1) & 1]);
        }
        const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * scale2;
        if (!__any_sync(0xffffffffu, mx > m_ref + kRescaleThreshold)) {
          tc::tmem_st_32x32(tS, pk);
          tc::tmem_st_32x32(tS + 32, pk + 32);
          done = true;
        } else {
          s4[0] = s4[1] = s4[2] = s4[3] = 0.f;
        }
      }
      if (!done) {
      // ---- pass 1: row maximum.  The four 32-column loads are in flight together (one wait), and four
      // independent max chains keep the FMNMX latency off the critical path of this lone warp.
      float mx;
      {
        uint32_t v[128];
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) tc::tmem_ld_32x32(tS + ch * 32, v + ch * 32);
        tc::tmem_ld_wait();
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        if (!need_mask) {
#pragma unroll
          for (int i = 0; i < 128; ++i) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(v[i]));
        } else {
#pragma unroll
          for (int i = 0; i < 128; ++i) m4[i & 3] = fmaxf(m4[i & 3], (i < limit && i >= lo) ? __uint_as_float(v[i]) : -INFINITY);
        }
        mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      }
      mx *= scale2;
      // ---- move the reference only when it is far off; O and l follow
      const bool need = mx > m_ref + kRescaleThreshold;
      if (j > 0 && __any_sync(0xffffffffu, need)) {
        // S_x(j) is complete, hence so is PV_x(j-1) (same in-order pipe, committed together): O_x is quiescent
        const float alpha = need ? ex2_approx(m_ref - mx) : 1.f;
#pragma unroll
        for (int ch = 0; ch < HD / 32; ++ch) {
          uint32_t o[32];
          tc::tmem_ld_32x32(tO + ch * 32, o);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tc::tmem_st_32x32(tO + ch * 32, o);
        }
        l_run *= alpha;
      }
      if (need) m_ref = mx;
      const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
      // ---- pass 2: P = exp2(S * scale2 - m_ref) -> bf16 pairs -> TMEM columns [0, 64) of S_x.
      // Both 64-column halves are requested up front; P of half 0 overwrites S[0,32) (already in registers),
      // P of half 1 overwrites S[32,64).
      {
        uint32_t v[128];
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) tc::tmem_ld_32x32(tS + ch * 32, v + ch * 32);
        tc::tmem_ld_wait();
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t pk[32];
          if (!need_mask) {
#pragma unroll
            for (int i = 0; i < 64; i += 2) {
              const float p0 = ex2_approx(fmaf(__uint_as_float(v[half * 64 + i]), scale2, neg_m));
              const float p1 = ex2_approx(fmaf(__uint_as_float(v[half * 64 + i + 1]), scale2, neg_m));
              s4[(i >> 1) & 3] += p0 + p1;
              pk[i >> 1] = pack_bf16x2(p0, p1);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 64; i += 2) {
              const int c = half * 64 + i;
              const float p0 = (c < limit && c >= lo) ? ex2_approx(fmaf(__uint_as_float(v[c]), scale2, neg_m)) : 0.f;
              const float p1 = (c + 1 < limit && c + 1 >= lo) ? ex2_approx(fmaf(__uint_as_float(v[c + 1]), scale2, neg_m)) : 0.f;
              s4[(i >> 1) & 3] += p0 + p1;
              pk[i >> 1] = pack_bf16x2(p0, p1);
            }
          }
          tc::tmem_st_32x32(tS + half * 32, pk);
        }
      }
      }
      const float sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
      tc::tmem_st_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&p_full[x]);
      l_run += sum;
    }

    // ---- epilogue: O / l -> bf16 rows, log-sum-exp
    if (nt > 0) {
      tc::mbar_wait(&o_done[x], (nt - 1) & 1);
      tc::tc_fence_after();
    }
    // TMEM loads are warp-wide (.sync.aligned): every lane takes part, only the global stores are predicated
    const bool in_range = q_idx < p.T;
    const bool valid = nt > 0 && q_idx < len && l_run > 0.f;
    const float inv = valid ? 1.f / l_run : 0.f;
    bf16* dst = p.o + (int64_t)b * p.o_sb + (int64_t)q_idx * p.o_st + (int64_t)h * p.o_sh;
#pragma unroll
    for (int ch = 0; ch < HD / 32; ++ch) {
      uint32_t o[32];
      if (nt > 0) {
        tc::tmem_ld_32x32(tO + ch * 32, o);
        tc::tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = 0u;
      }
      if (in_range) {
#pragma unroll
        for (int c = 0; c < 32; c += 8) {
          const uint32_t w0 = pack_bf16x2(__uint_as_float(o[c]) * inv, __uint_as_float(o[c + 1]) * inv);
          const uint32_t w1 = pack_bf16x2(__uint_as_float(o[c + 2]) * inv, __uint_as_float(o[c + 3]) * inv);
          const uint32_t w2 = pack_bf16x2(__uint_as_float(o[c + 4]) * inv, __uint_as_float(o[c + 5]) * inv);
          const uint32_t w3 = pack_bf16x2(__uint_as_float(o[c + 6]) * inv, __uint_as_float(o[c + 7]) * inv);
          // one 16-byte store (the compiler split the struct store into four 4-byte ones)
          asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(dst + ch * 32 + c), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
        }
      }
    }
    if (in_range && p.lse != nullptr) p.lse[((int64_t)b * p.Hq + h) * p.T + q_idx] = valid ? (m_ref + log2f(l_run)) * kLn2 : 0.f;
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc<1>(tmem_base, 512);
  }
}

template <int HD>
int launch_fwd2(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV, const Params& p, cudaStream_t s) {
  constexpr int smem = 6 * 128 * HD * 2 + 1024 + 256;
  static bool set = false;
  if (!set) {
    DLA_CUDA(cudaFuncSetAttribute(attn_fwd2_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    set = true;
  }
  dim3 grid((unsigned)(((p.T + 255) / 256) * p.Hq * p.B));
  attn_fwd2_kernel<HD><<<grid, kFwd2Threads, smem, s>>>(tmQ, tmK, tmV, p);
  return check_launch("attn_fwd2");
}

int attn_fwd2_dispatch(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV, const Params& p, cudaStream_t s) {
  return p.D == 128 ? launch_fwd2<128>(tmQ, tmK, tmV, p, s) : launch_fwd2<64>(tmQ, tmK, tmV, p, s);
}

}  // namespace attn
}  // namespace dla
