// Flash attention forward, third generation: the ping-pong kernel of attn_fwd2.cu with S released early.
//
// In attn_fwd2 P aliases S in TMEM, so for one query tile the tensor pipe must run PV_x(j) before S_x(j+1) may be
// written: after handing over P a softmax warp group waits for two MMAs (plus whatever the other tile queued), 24%
// of all warp samples in ncu.  Here
//   * a softmax thread reads its S row ONCE (128 registers) and frees the TMEM buffer immediately (s_free): the MMA
//     warp issues S_x(j+1) = Q_x K_{j+1}^T while the exponentials of step j are still being computed;
//   * P goes through shared memory again (bf16, 128B-swizzled K-major, 32 KB per query tile).  To pay for it V keeps a
//     single stage (its next tile is only needed one softmax later); K stays double buffered because S runs ahead:
//     Q 64 KB + K 64 KB + V 32 KB + P 64 KB = 224 KB;
//   * the MMA warp is an event loop over (s_free, k_full) -> issue S and (p_full, v_full) -> issue PV for both tiles,
//     instead of a fixed A, B, A, B order;
//   * O stays in TMEM with the lazy (2^8) rescale; because S no longer orders PV, a softmax thread waits for o_done of
//     step j-1 before it touches O or overwrites P.
//
// MEASURED (B200, profiles/README.md): correct, but slower than attn_fwd2 — 0.147 vs 0.129 ms at T=1024 and 0.358 vs
// 0.295 ms at T=4096 (B*T = 8192, 32/8 heads, D=128).  The shared-memory round trip of P, the single V stage and the
// 128-register S row cost more than the early S issue gains, so dla_attn_fwd keeps attn_fwd2 as its default and this
// kernel is reachable only through dla_attn_set_fwd_variant(3) (kept as the record of the experiment, with its test).
//
// Reference path replaced: PyTorch SDPA under HF Llama (R/src/training/train_dpo.py:32).
#include <math.h>

#include "attn_common.cuh"

namespace dla {
namespace attn {

constexpr int kFwd3Threads = 320;     // TMA warp, MMA warp, 4 softmax warps per query tile
constexpr float kRescale3 = 8.f;     // log2 units
constexpr int P_TILE_BYTES = 128 * 128 * 2;

template <int HD>
__global__ void __launch_bounds__(kFwd3Threads, 1)
attn_fwd3_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const Params p) {
  constexpr int NB = HD / 64;
  constexpr int TILE_BYTES = 128 * HD * 2;      // one Q, K or V tile

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;                                   // [2][TILE_BYTES]
  uint8_t* sK = sQ + 2 * TILE_BYTES;                    // [2][TILE_BYTES]
  uint8_t* sV = sK + 2 * TILE_BYTES;                    // [1][TILE_BYTES]
  uint8_t* sP = sV + TILE_BYTES;                        // [2][P_TILE_BYTES]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * P_TILE_BYTES);
  uint64_t* q_full = bars;            // 1
  uint64_t* k_full = bars + 1;        // [2]
  uint64_t* k_empty = bars + 3;       // [2]
  uint64_t* v_full = bars + 5;        // 1
  uint64_t* v_empty = bars + 6;       // 1
  uint64_t* s_full = bars + 7;        // [2] per query tile
  uint64_t* s_free = bars + 9;        // [2] per query tile, 128 arrivals
  uint64_t* p_full = bars + 11;       // [2] per query tile, 128 arrivals
  uint64_t* o_done = bars + 13;       // [2] per query tile
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1-D grid, longest-processing-time first (see attn_fwd2.cu)
  const int n_pairs = (p.T + 255) >> 8, n_hb = p.Hq * p.B;
  const int pi = n_pairs - 1 - (int)blockIdx.x / n_hb;
  const int hb = (int)blockIdx.x % n_hb;
  const int h = hb % p.Hq, b = hb / p.Hq;
  const int hk = h / (p.Hq / p.Hk);
  int len = p.T;
  if (p.seqlens != nullptr) len = min(len, p.seqlens[b]);
  const int n_all = (len + 127) >> 7;
  int n0_ = p.causal ? min(n_all, 2 * pi + 1) : n_all;
  int n1_ = p.causal ? min(n_all, 2 * pi + 2) : n_all;
  if (2 * pi * 128 >= len) n0_ = 0;
  if ((2 * pi + 1) * 128 >= len) n1_ = 0;
  const int nA = n0_, nB = n1_;
  const int n = max(nA, nB);

  if (threadIdx.x == 0) {
    tc::prefetch_tmap(&tmQ);
    tc::prefetch_tmap(&tmK);
    tc::prefetch_tmap(&tmV);
    tc::mbar_init(q_full, 1);
    tc::mbar_init(v_full, 1);
    tc::mbar_init(v_empty, 1);
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&k_full[s], 1);
      tc::mbar_init(&k_empty[s], 1);
      tc::mbar_init(&s_full[s], 1);
      tc::mbar_init(&s_free[s], 128);
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
    // ------------------------------------------------------------------ TMA producer: Q, K(0), then K(j+1) ahead of V(j)
    if (lane == 0 && n > 0) {
      tc::mbar_expect_tx(q_full, 2 * TILE_BYTES);
#pragma unroll
      for (int x = 0; x < 2; ++x)
#pragma unroll
        for (int c = 0; c < NB; ++c)
          tc::tma_load_4d(sQ + x * TILE_BYTES + c * 16384, &tmQ, q_full, c * 64, h, (2 * pi + x) * 128, b);
      auto load_k = [&](int j) {
        const int st = j & 1;
        tc::mbar_wait(&k_empty[st], ((j >> 1) & 1) ^ 1);
        tc::mbar_expect_tx(&k_full[st], TILE_BYTES);
#pragma unroll
        for (int c = 0; c < NB; ++c) tc::tma_load_4d(sK + st * TILE_BYTES + c * 16384, &tmK, &k_full[st], c * 64, hk, j * 128, b);
      };
      load_k(0);
      for (int j = 0; j < n; ++j) {
        if (j + 1 < n) load_k(j + 1);
        tc::mbar_wait(v_empty, (j & 1) ^ 1);
        tc::mbar_expect_tx(v_full, TILE_BYTES);
#pragma unroll
        for (int c = 0; c < NB; ++c) tc::tma_load_4d(sV + c * 16384, &tmV, v_full, c * 64, hk, j * 128, b);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer: event loop over both query tiles
    if (lane == 0 && n > 0) {
      constexpr uint32_t idesc_s = tc::make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_pv = tc::make_idesc_bf16(128, HD, 0, 1);
      const uint32_t aV = tc::smem_u32(sV);
      int ns[2] = {0, 0}, npv[2] = {0, 0};       // next kv step whose S / PV has not been issued yet, per query tile
      const int nx[2] = {nA, nB};
      int k_rel = 0, v_rel = 0;                  // next K / V tile to hand back to the producer
      tc::mbar_wait(q_full, 0);
      uint32_t idle = 0;                         // polls without progress: a protocol bug must abort the grid, not hang the GPU
      while (npv[0] < nA || npv[1] < nB) {
        const int before = ns[0] + ns[1] + npv[0] + npv[1];
#pragma unroll
        for (int x = 0; x < 2; ++x) {
          // ---- S_x(j): needs K(j) in shared memory and S_x read out by the softmax of step j-1
          const int j = ns[x];
          if (j < nx[x] && (j == 0 || tc::mbar_try_wait(&s_free[x], (j - 1) & 1)) &&
              tc::mbar_try_wait(&k_full[j & 1], (j >> 1) & 1)) {
            tc::tc_fence_after();
            const uint32_t aQ = tc::smem_u32(sQ + x * TILE_BYTES), aK = tc::smem_u32(sK + (j & 1) * TILE_BYTES);
#pragma unroll
            for (int ks = 0; ks < HD / 16; ++ks) {
              const uint32_t off = (ks >> 2) * 16384 + (ks & 3) * 32;
              tc::umma_bf16<1>(tmem_base + x * 128, tc::make_smem_desc(aQ + off, 16, 1024), tc::make_smem_desc(aK + off, 16, 1024),
                               idesc_s, ks > 0 ? 1u : 0u);
            }
            tc::umma_commit(&s_full[x]);
            ns[x] = j + 1;
          }
          // ---- PV_x(i): needs P_x(i) in shared memory and V(i)
          const int i = npv[x];
          if (i < nx[x] && i < ns[x] && tc::mbar_try_wait(&p_full[x], i & 1) && tc::mbar_try_wait(v_full, i & 1)) {
            tc::tc_fence_after();
            const uint32_t aP = tc::smem_u32(sP + x * P_TILE_BYTES);
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
              const uint32_t offP = (ks >> 2) * 16384 + (ks & 3) * 32;
              tc::umma_bf16<1>(tmem_base + 256 + x * 128, tc::make_smem_desc(aP + offP, 16, 1024),
                               tc::make_smem_desc(aV + ks * 2048, 16384, 1024), idesc_pv, (i > 0 || ks > 0) ? 1u : 0u);
            }
            tc::umma_commit(&o_done[x]);
            npv[x] = i + 1;
          }
        }
        // K(k_rel) is free once every tile that uses it has issued its S; likewise V(v_rel) and PV
        while (k_rel < n && (k_rel >= nA || ns[0] > k_rel) && (k_rel >= nB || ns[1] > k_rel)) {
          tc::umma_commit(&k_empty[k_rel & 1]);
          ++k_rel;
        }
        while (v_rel < n && (v_rel >= nA || npv[0] > v_rel) && (v_rel >= nB || npv[1] > v_rel)) {
          tc::umma_commit(v_empty);
          ++v_rel;
        }
        if (ns[0] + ns[1] + npv[0] + npv[1] != before) idle = 0;
        else if (++idle > (1u << 27)) __trap();
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax of tile x, one thread per row
    const int x = (warp - 2) >> 2;
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
    const int r = quarter * 32 + lane;
    const int q0 = (2 * pi + x) * 128;
    const int q_idx = q0 + r;
    const int nt = x == 0 ? nA : nB;
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const uint32_t tS = tmem_base + lane_off + x * 128;
    const uint32_t tO = tmem_base + lane_off + 256 + x * 128;
    uint8_t* prow = sP + x * P_TILE_BYTES + r * 128;      // + blk * 16384 + ((chunk ^ (r & 7)) << 4)
    const float scale2 = p.scale * kLog2e;
    float m_ref = -INFINITY, l_run = 0.f;

    for (int j = 0; j < nt; ++j) {
      const int kv0 = j * 128;
      tc::mbar_wait(&s_full[x], j & 1);
      tc::tc_fence_after();
      // the whole S row comes into registers once; the TMEM buffer goes straight back to the MMA warp
      uint32_t v[128];
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) tc::tmem_ld_32x32(tS + ch * 32, v + ch * 32);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&s_free[x]);
      const bool need_mask = (p.causal && kv0 + 127 > q0) || (kv0 + 128 > len);
      if (need_mask) {
        int lim = len;
        if (p.causal) lim = min(lim, q_idx + 1);
        const int limit = lim - kv0;              // valid columns of this row in this tile
#pragma unroll
        for (int i = 0; i < 128; ++i)
          if (i >= limit) v[i] = 0xff800000u;     // -inf: exp2 gives exactly 0
      }
      float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < 128; ++i) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(v[i]));
      const float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * scale2;
      // PV_x(j-1) must have retired before O is rescaled or P(j-1) is overwritten (S no longer orders it)
      if (j > 0) {
        tc::mbar_wait(&o_done[x], (j - 1) & 1);
        tc::tc_fence_after();
      }
      const bool need = mx > m_ref + kRescale3;
      if (j > 0 && __any_sync(0xffffffffu, need)) {
        const float alpha = need ? ex2_approx(m_ref - mx) : 1.f;
        // 8 columns at a time: the S row (128 registers) stays live across this rare path
#pragma unroll 1
        for (int ch = 0; ch < HD / 8; ++ch) {
          uint32_t o[8];
          tc::tmem_ld_32x32_x8(tO + ch * 8, o);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tc::tmem_st_32x32_x8(tO + ch * 8, o);
        }
        tc::tmem_st_wait();
        l_run *= alpha;
      }
      if (need) m_ref = mx;
      const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
      float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 16; ++c) {              // 16 chunks of 8 columns = one 16-byte store each
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float p0 = ex2_approx(fmaf(__uint_as_float(v[c * 8 + 2 * e]), scale2, neg_m));
          const float p1 = ex2_approx(fmaf(__uint_as_float(v[c * 8 + 2 * e + 1]), scale2, neg_m));
          s4[e] += p0 + p1;
          pk[e] = pack_bf16x2(p0, p1);
        }
        *reinterpret_cast<uint4*>(prow + (c >> 3) * 16384 + (((c & 7) ^ (r & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
      tc::tc_fence_before();
      tc::fence_proxy_async();
      tc::mbar_arrive(&p_full[x]);
      l_run += (s4[0] + s4[1]) + (s4[2] + s4[3]);
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
int launch_fwd3(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV, const Params& p, cudaStream_t s) {
  constexpr int smem = 5 * 128 * HD * 2 + 2 * P_TILE_BYTES + 1024 + 256;
  static bool set = false;
  if (!set) {
    DLA_CUDA(cudaFuncSetAttribute(attn_fwd3_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    set = true;
  }
  dim3 grid((unsigned)(((p.T + 255) / 256) * p.Hq * p.B));
  attn_fwd3_kernel<HD><<<grid, kFwd3Threads, smem, s>>>(tmQ, tmK, tmV, p);
  return check_launch("attn_fwd3");
}

int attn_fwd3_dispatch(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV, const Params& p, cudaStream_t s) {
  return p.D == 128 ? launch_fwd3<128>(tmQ, tmK, tmV, p, s) : launch_fwd3<64>(tmQ, tmK, tmV, p, s);
}

}  // namespace attn
}  // namespace dla
