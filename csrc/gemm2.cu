// 2-CTA (cta_group::2) variant of the bf16 GEMM: a cluster of two CTAs on one TPC computes a 256x256 tile
// with ONE tcgen05.mma.cta_group::2 stream issued by the leader CTA.  Each CTA stages only its 128 rows of A
// and its 128-row half of B per k-block (32 KiB instead of 48 KiB), so the tensor core's shared-memory
// operand traffic per SM drops by a third and the ring deepens to 6 stages; accumulators stay split across
// the two CTAs' TMEM (128 lanes x 256 columns each, double buffered).
//
//   warp 0  TMA producer (both CTAs; transaction bytes are signalled on the LEADER's full barrier)
//   warp 1  MMA issuer (leader only); tcgen05.commit multicasts "slot free" / "accumulator ready" to both CTAs
//   warps 2-5 epilogue (each CTA drains its own TMEM half through staged smem + TMA store / reduce-add)
//
// kSwiGLU (gate/up projection of a Llama MLP, B = [2F, K] with the gate rows first): the cluster tile is made of the
// SAME 128 feature columns of gate (CTA 0's half of B) and up (CTA 1's half), so each CTA ends up with gate in TMEM
// columns [0,128) and up in [128,256) for its 128 rows and the epilogue emits act = silu(gate) * up directly — the
// separate SwiGLU pass (read 2F, write F per token) disappears; gate/up themselves are stored only when a backward
// will need them.  Replaces K5 of SURVEY §2.6.2 as a stand-alone kernel on the forward path.
//
// kFuse == 3 (QKV projection of a Llama block, head_dim 128, full rotary): the 256-column tile is two heads; the epilogue
// rotates (c, c + 64) column pairs of the q and k heads by the row's position on the fp32 accumulators and stores bf16
// once, so RoPE is neither a separate pass over the packed projection nor a second rounding.
//
// kFuse == 2 (dgrad of the down projection): the epilogue turns its d(act) tile into d(gate) and d(up) on the spot — it
// reads the saved gate / up values of the same rows and columns and stores both halves of d[gate | up]; d(act) itself
// never reaches memory and the separate SwiGLU-backward pass disappears.
//
// Same operand-layout support as gemm.cu (K-major / MN-major A and B).  Selected by dla_gemm_bf16 when the
// problem has at least one full wave of 256x256 tiles (see gemm.cu) or forced with dla_gemm_set_variant(2).
#include <math.h>

#include <atomic>

#include "tc_common.cuh"

namespace dla {
namespace gemm2 {

constexpr int BM = 128;            // A rows per CTA (cluster tile M = 256)
constexpr int BNH = 128;           // B rows per CTA (cluster tile N = 256)
constexpr int BN = 256;
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int kStages = 6;
constexpr int A_BYTES = BM * BK * 2;     // 16 KiB
constexpr int B_BYTES = BNH * BK * 2;    // 16 KiB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int kThreads = 192;
constexpr int kEpiWarps = 4;
constexpr int EPI_BYTES = kEpiWarps * 2 * 4096;
constexpr int SMEM_BYTES = kStages * STAGE_BYTES + EPI_BYTES + 1024 + 256;
constexpr int GROUP_M = 8;

struct Params {
  int M, N, K;
  int c_fp32, accumulate;
  int num_m_tiles, num_n_tiles;      // in 256 x 256 cluster tiles
  // wave-quantisation fix: the last (partial) wave of `tail_tiles` tiles is computed as tail_tiles x tail_splits work
  // items over disjoint K slices, one per cluster, combined by TMA reduce-add (those tiles are zero-filled by
  // zero_tail_kernel first unless the caller accumulates).  tail_splits <= 1: off.
  int tail_tiles, tail_splits;
  int F;            // kFuse 1: features per half (N = 2F), num_n_tiles counts 128-wide feature blocks; kFuse 2: N = F
  int store_gu;     // kFuse 1: also store gate / up (bf16) through tmC for the backward
  const bf16* gu;   // kFuse 2: saved [M, 2F] gate | up
  int64_t ldgu;
  const bf16* bias; // kFuse 0: [N] added to every output row (forward linears with a bias; never with accumulate / tail split)
  // kFuse 3: rotary embedding of the first n_rot_heads 128-wide heads of each output row (packed QKV projection)
  const float* cos_t;          // [max_pos, 64]
  const float* sin_t;
  const int32_t* positions;    // [M]
  int n_rot_heads;
};

// work item w -> (tile, K-block range).  Items below `full` are whole tiles; the rest are K slices of the tail tiles,
// laid out slice-major so that the slices of one tile sit on different clusters of the same (last) wave.
struct Work {
  int tile, kb0, kb1;
  bool split;
};
__device__ __forceinline__ Work work_item(int w, const Params& p, int num_tiles, int num_kb) {
  Work x;
  const int full = num_tiles - (p.tail_splits > 1 ? p.tail_tiles : 0);
  if (w < full) {
    x.tile = w;
    x.kb0 = 0;
    x.kb1 = num_kb;
    x.split = false;
  } else {
    const int i = w - full;
    const int slice = i / p.tail_tiles;
    const int per = (num_kb + p.tail_splits - 1) / p.tail_splits;
    x.tile = full + i % p.tail_tiles;
    x.kb0 = slice * per;
    x.kb1 = min(num_kb, x.kb0 + per);
    x.split = true;
  }
  return x;
}

__device__ __forceinline__ uint32_t pack2(uint32_t a, uint32_t b) {
  bf162 v = __floats2bfloat162_rn(__uint_as_float(a), __uint_as_float(b));
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ void tile_coords(int t, int num_m, int num_n, int& m_blk, int& n_blk) {
  const int per_group = GROUP_M * num_n;
  const int g = t / per_group;
  const int first_m = g * GROUP_M;
  const int gsz = min(num_m - first_m, GROUP_M);
  const int r = t - g * per_group;
  m_blk = first_m + r % gsz;
  n_blk = r / gsz;
}

__device__ __forceinline__ uint32_t mapa_u32(uint32_t local_smem_addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(cta));
  return r;
}

// 2-CTA TMA load: data into THIS CTA's smem, completion bytes onto the barrier at cluster address `bar_cluster`
__device__ __forceinline__ void tma_load_2d_to(void* smem_dst, const CUtensorMap* m, uint32_t bar_cluster, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          tc::smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}

template <bool kAMN, bool kBMN, int kFuse = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
             const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmD, const Params p) {
  constexpr bool kSwiGLU = kFuse == 1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* epi_smem = smem + kStages * STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_smem + EPI_BYTES);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tfull_bar = empty_bar + kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = tc::cluster_ctarank();
  const bool leader = cta == 0;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int num_kb = (p.K + BK - 1) / BK;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;
  const int num_items = num_tiles + (p.tail_splits > 1 ? p.tail_tiles * (p.tail_splits - 1) : 0);

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmA);
    tc::prefetch_tmap(&tmB);
    tc::prefetch_tmap(&tmC);
    for (int s = 0; s < kStages; ++s) {
      tc::mbar_init(&full_bar[s], 1);        // leader's producer arrives once (+ tx bytes of both CTAs)
      tc::mbar_init(&empty_bar[s], 1);       // one multicast commit per use
    }
    for (int a = 0; a < 2; ++a) {
      tc::mbar_init(&tfull_bar[a], 1);
      tc::mbar_init(&tempty_bar[a], 2 * kEpiWarps);   // epilogue warps of BOTH CTAs (used in the leader only)
    }
    tc::fence_barrier_init();
  }
  tc::cluster_sync();                          // peer barriers are initialised before anyone signals them
  if (warp == 1) tc::tmem_alloc<2>(tmem_holder, 512);
  tc::tc_fence_before();
  tc::cluster_sync();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = cluster_id; w < num_items; w += n_clusters) {
        const Work wk = work_item(w, p, num_tiles, num_kb);
        int m_blk, n_blk;
        tile_coords(wk.tile, p.num_m_tiles, p.num_n_tiles, m_blk, n_blk);
        const int m0 = m_blk * 256 + (int)cta * BM;
        // kSwiGLU: CTA 0 stages the gate rows of feature block n_blk, CTA 1 the up rows of the same block
        const int n0 = kSwiGLU ? (int)cta * p.F + n_blk * BNH : n_blk * BN + (int)cta * BNH;
        for (int kb = wk.kb0; kb < wk.kb1; ++kb) {
          tc::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          const uint32_t bar = mapa_u32(tc::smem_u32(&full_bar[stage]), 0);     // the leader's barrier
          if (leader) tc::mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
          const int k0 = kb * BK;
          if constexpr (!kAMN) {
            tma_load_2d_to(sa, &tmA, bar, k0, m0);
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c) tma_load_2d_to(sa + c * (64 * BK * 2), &tmA, bar, m0 + c * 64, k0);
          }
          if constexpr (!kBMN) {
            tma_load_2d_to(sb, &tmB, bar, k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BNH / 64; ++c) tma_load_2d_to(sb + c * (64 * BK * 2), &tmB, bar, n0 + c * 64, k0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = tc::make_idesc_bf16(256, BN, kAMN ? 1 : 0, kBMN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int w = cluster_id; w < num_items; w += n_clusters, ++it) {
        const Work wk = work_item(w, p, num_tiles, num_kb);
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        tc::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc::tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = wk.kb0; kb < wk.kb1; ++kb) {
          tc::mbar_wait(&full_bar[stage], phase);
          tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t da = kAMN ? tc::make_smem_desc(sa + k * 2048, 64 * BK * 2, 1024) : tc::make_smem_desc(sa + k * 32, 16, 1024);
            const uint64_t db = kBMN ? tc::make_smem_desc(sb + k * 2048, 64 * BK * 2, 1024) : tc::make_smem_desc(sb + k * 32, 16, 1024);
            tc::umma_bf16<2>(tmem_d, da, db, idesc, (kb > wk.kb0 || k > 0) ? 1u : 0u);
          }
          tc::umma_commit_2sm(&empty_bar[stage], 0b11);                      // both CTAs' smem slots are reusable
          if (kb == wk.kb1 - 1) tc::umma_commit_2sm(&tfull_bar[acc], 0b11);  // both CTAs' accumulator halves are ready
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5 of both CTAs) =====================
    const int quarter = warp & 3;
    const uint32_t tempty_leader0 = mapa_u32(tc::smem_u32(&tempty_bar[0]), 0);
    const uint32_t tempty_leader1 = mapa_u32(tc::smem_u32(&tempty_bar[1]), 0);
    int it = 0;
    for (int w = cluster_id; w < num_items; w += n_clusters, ++it) {
      const Work wk = work_item(w, p, num_tiles, num_kb);
      const bool add = p.accumulate != 0 || wk.split;       // K slices of a tail tile meet in C by reduce-add
      int m_blk, n_blk;
      tile_coords(wk.tile, p.num_m_tiles, p.num_n_tiles, m_blk, n_blk);
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      tc::mbar_wait(&tfull_bar[acc], acc_phase);
      tc::tc_fence_after();
      const int col_base = n_blk * BN;
      const int r0 = m_blk * 256 + (int)cta * BM + quarter * 32;
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BN;
      const int gc = p.c_fp32 ? 32 : 64;
      const int n_groups = BN / gc;
      uint8_t* stage_base = epi_smem + (warp - 2) * 2 * 4096;
      if constexpr (kSwiGLU) {
        // act[:, n_blk*128 + 64 gi .. +64) = silu(gate) * up from TMEM columns [64 gi, +64) and [128 + 64 gi, +64)
        int nbuf = 0;
        auto emit = [&](const CUtensorMap* tm, const uint32_t* r, int c0) {      // 64 fp32 values of this lane's row -> bf16 tile store
          uint8_t* buf = stage_base + (nbuf & 1) * 4096;
          ++nbuf;
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          uint8_t* myrow = buf + lane * 128;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint4 v;
            v.x = pack2(r[8 * c], r[8 * c + 1]);
            v.y = pack2(r[8 * c + 2], r[8 * c + 3]);
            v.z = pack2(r[8 * c + 4], r[8 * c + 5]);
            v.w = pack2(r[8 * c + 6], r[8 * c + 7]);
            *reinterpret_cast<uint4*>(myrow + ((c ^ (lane & 7)) << 4)) = v;
          }
          tc::fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            if (r0 < p.M) tc::tma_store_2d(tm, buf, c0, r0);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        };
#pragma unroll 1
        for (int gi = 0; gi < 2; ++gi) {
          uint32_t g[64], u[64];
          tc::tmem_ld_32x32(taddr + gi * 64, g);
          tc::tmem_ld_32x32(taddr + gi * 64 + 32, g + 32);
          tc::tmem_ld_32x32(taddr + 128 + gi * 64, u);
          tc::tmem_ld_32x32(taddr + 128 + gi * 64 + 32, u + 32);
          tc::tmem_ld_wait();
          const int fc0 = n_blk * BNH + gi * 64;
          if (p.store_gu) {
            emit(&tmC, g, fc0);
            emit(&tmC, u, p.F + fc0);
          }
#pragma unroll
          for (int c = 0; c < 64; ++c) {
            const float gv = __uint_as_float(g[c]);
            g[c] = __float_as_uint(gv * __fdividef(1.f, 1.f + __expf(-gv)) * __uint_as_float(u[c]));   // fast reciprocal: the result is rounded to bf16
          }
          emit(&tmD, g, fc0);
        }
      } else if constexpr (kFuse == 2) {
        // d(gate), d(up) from this CTA's d(act) tile [128 rows x 256 feature columns] and the saved gate / up
        const int row = r0 + lane;
        int nbuf = 0;
        auto emit = [&](const uint32_t* r, int c0) {
          uint8_t* buf = stage_base + (nbuf & 1) * 4096;
          ++nbuf;
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          uint8_t* myrow = buf + lane * 128;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint4 v;
            v.x = pack2(r[8 * c], r[8 * c + 1]);
            v.y = pack2(r[8 * c + 2], r[8 * c + 3]);
            v.z = pack2(r[8 * c + 4], r[8 * c + 5]);
            v.w = pack2(r[8 * c + 6], r[8 * c + 7]);
            *reinterpret_cast<uint4*>(myrow + ((c ^ (lane & 7)) << 4)) = v;
          }
          tc::fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            if (r0 < p.M) tc::tma_store_2d(&tmD, buf, c0, r0);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        };
#pragma unroll 1
        for (int gi = 0; gi < 4; ++gi) {
          const int c0 = col_base + gi * 64;
          if (c0 >= p.F) break;                     // (F % 64 == 0: a group is all in or all out)
          // request the saved gate / up values of this lane's row first: they arrive while the TMEM load is in flight
          uint4 gq[8], uq[8];
          const bool rv = row < p.M;
          const bf16* gp = p.gu + (int64_t)(rv ? row : 0) * p.ldgu + c0;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            gq[c] = *reinterpret_cast<const uint4*>(gp + c * 8);
            uq[c] = *reinterpret_cast<const uint4*>(gp + p.F + c * 8);
          }
          uint32_t d[64], o[64];
          tc::tmem_ld_32x32(taddr + gi * 64, d);
          tc::tmem_ld_32x32(taddr + gi * 64 + 32, d + 32);
          tc::tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            float g8[8], u8[8];
            unpack8(*reinterpret_cast<const bf16x8*>(&gq[c]), g8);
            unpack8(*reinterpret_cast<const bf16x8*>(&uq[c]), u8);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const float da = __uint_as_float(d[c * 8 + e]);
              const float sg = __fdividef(1.f, 1.f + __expf(-g8[e]));
              const float silu = g8[e] * sg;
              o[c * 8 + e] = __float_as_uint(da * silu);                                       // d(up)
              d[c * 8 + e] = __float_as_uint(da * u8[e] * sg * (1.f + g8[e] * (1.f - sg)));   // d(gate)
            }
          }
          emit(d, c0);
          emit(o, p.F + c0);
        }
      } else if constexpr (kFuse == 3) {
        const int row = r0 + lane;
        const int pos = row < p.M ? p.positions[row] : 0;
        const float4* ct = reinterpret_cast<const float4*>(p.cos_t + (int64_t)pos * 64);
        const float4* st = reinterpret_cast<const float4*>(p.sin_t + (int64_t)pos * 64);
        uint8_t* lo_row = stage_base + lane * 128;              // columns [0, 64) of the head, this lane's row
        uint8_t* hi_row = stage_base + 4096 + lane * 128;       // columns [64, 128)
#pragma unroll 1
        for (int hh = 0; hh < 2; ++hh) {
          const int c0 = col_base + hh * 128;
          if (c0 >= p.N) break;                     // (N % 128 == 0: a head is all in or all out)
          const bool rot = (c0 >> 7) < p.n_rot_heads;
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the previous head's stores have read both buffers
          __syncwarp();
#pragma unroll 1
          for (int q = 0; q < 2; ++q) {             // 32 rotation pairs at a time: (c, c + 64) for c in [32q, 32q + 32)
            uint32_t lo[32], hi[32];
            tc::tmem_ld_32x32(taddr + hh * 128 + q * 32, lo);
            tc::tmem_ld_32x32(taddr + hh * 128 + 64 + q * 32, hi);
            tc::tmem_ld_wait();
            if (rot) {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 c4 = __ldg(ct + q * 8 + j), s4 = __ldg(st + q * 8 + j);
                const float cc[4] = {c4.x, c4.y, c4.z, c4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float a = __uint_as_float(lo[4 * j + e]), b = __uint_as_float(hi[4 * j + e]);
                  lo[4 * j + e] = __float_as_uint(a * cc[e] - b * ss[e]);
                  hi[4 * j + e] = __float_as_uint(b * cc[e] + a * ss[e]);
                }
              }
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {           // 32 bf16 = four 16-byte chunks of the 128-byte staging row
              const int chunk = ((q * 4 + c) ^ (lane & 7)) << 4;
              *reinterpret_cast<uint4*>(lo_row + chunk) =
                  make_uint4(pack2(lo[8 * c], lo[8 * c + 1]), pack2(lo[8 * c + 2], lo[8 * c + 3]), pack2(lo[8 * c + 4], lo[8 * c + 5]),
                             pack2(lo[8 * c + 6], lo[8 * c + 7]));
              *reinterpret_cast<uint4*>(hi_row + chunk) =
                  make_uint4(pack2(hi[8 * c], hi[8 * c + 1]), pack2(hi[8 * c + 2], hi[8 * c + 3]), pack2(hi[8 * c + 4], hi[8 * c + 5]),
                             pack2(hi[8 * c + 6], hi[8 * c + 7]));
            }
          }
          tc::fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            if (r0 < p.M) {
              tc::tma_store_2d(&tmC, stage_base, c0, r0);
              tc::tma_store_2d(&tmC, stage_base + 4096, c0 + 64, r0);
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
      } else
#pragma unroll 1
      for (int gi = 0; gi < n_groups; ++gi) {
        uint8_t* buf = stage_base + (gi & 1) * 4096;
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        __syncwarp();
        uint8_t* myrow = buf + lane * 128;
        if (p.c_fp32) {
          uint32_t r[32];
          tc::tmem_ld_32x32(taddr + gi * 32, r);
          tc::tmem_ld_wait();
          if (p.bias != nullptr) tc::add_bias_row<32>(r, p.bias, col_base + gi * 32, p.N);
#pragma unroll
          for (int c = 0; c < 8; ++c)
            *reinterpret_cast<uint4*>(myrow + ((c ^ (lane & 7)) << 4)) = make_uint4(r[4 * c], r[4 * c + 1], r[4 * c + 2], r[4 * c + 3]);
        } else {
          uint32_t r[64];
          tc::tmem_ld_32x32(taddr + gi * 64, r);
          tc::tmem_ld_32x32(taddr + gi * 64 + 32, r + 32);
          tc::tmem_ld_wait();
          if (p.bias != nullptr) tc::add_bias_row<64>(r, p.bias, col_base + gi * 64, p.N);
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint4 v;
            v.x = pack2(r[8 * c], r[8 * c + 1]);
            v.y = pack2(r[8 * c + 2], r[8 * c + 3]);
            v.z = pack2(r[8 * c + 4], r[8 * c + 5]);
            v.w = pack2(r[8 * c + 6], r[8 * c + 7]);
            *reinterpret_cast<uint4*>(myrow + ((c ^ (lane & 7)) << 4)) = v;
          }
        }
        tc::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          const int c0 = col_base + gi * gc;
          if (c0 < p.N && r0 < p.M) {
            if (add) tc::tma_reduce_add_2d(&tmC, buf, c0, r0);
            else tc::tma_store_2d(&tmC, buf, c0, r0);
          }
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) {      // free this accumulator buffer in the LEADER (its MMA warp owns the schedule)
        asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(acc ? tempty_leader1 : tempty_leader0) : "memory");
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc::tc_fence_before();
  tc::cluster_sync();                          // neither CTA may exit (or free TMEM) while the peer still uses it
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc<2>(tmem_base, 512);
  }
}

// zero-fills the tail tiles (block = one 128-row half of one tile) ahead of their reduce-add slices
__global__ void __launch_bounds__(256)
zero_tail_kernel(uint8_t* __restrict__ C, int64_t ld_bytes, int esz, int M, int N, int num_m, int num_n, int first_tile) {
  int m_blk, n_blk;
  tile_coords(first_tile + (int)(blockIdx.x >> 1), num_m, num_n, m_blk, n_blk);
  const int r0 = m_blk * 256 + (int)(blockIdx.x & 1) * 128, c0 = n_blk * BN;
  const int rows = min(128, M - r0), cols = min(BN, N - c0);
  if (rows <= 0 || cols <= 0) return;
  const int vec_per_row = cols * esz / 16;                 // row pitch and tile origin are 16-byte aligned (TMA contract)
  const int tail_bytes = cols * esz - vec_per_row * 16;
  for (int i = threadIdx.x; i < rows * vec_per_row; i += blockDim.x) {
    const int r = i / vec_per_row, v = i - r * vec_per_row;
    *reinterpret_cast<uint4*>(C + (int64_t)(r0 + r) * ld_bytes + (int64_t)c0 * esz + v * 16) = make_uint4(0, 0, 0, 0);
  }
  if (tail_bytes > 0)
    for (int i = threadIdx.x; i < rows * tail_bytes; i += blockDim.x) {
      const int r = i / tail_bytes, b = i - r * tail_bytes;
      C[(int64_t)(r0 + r) * ld_bytes + (int64_t)c0 * esz + vec_per_row * 16 + b] = 0;
    }
}

template <bool kAMN, bool kBMN, int kFuse = 0>
static int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const CUtensorMap& tmD, const Params& p,
                  cudaStream_t stream) {
  auto kern = gemm2_kernel<kAMN, kBMN, kFuse>;
  static std::atomic<unsigned long long> attr_devices{0};      // per-device function attribute, as in gemm.cu
  int dev = 0;
  DLA_CUDA(cudaGetDevice(&dev));
  const unsigned long long bit = 1ULL << (dev & 63);
  if (!(attr_devices.load(std::memory_order_relaxed) & bit)) {
    DLA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    attr_devices.fetch_or(bit, std::memory_order_relaxed);
  }
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;
  const int clusters = num_tiles < kNumSMs / 2 ? num_tiles : kNumSMs / 2;
  kern<<<2 * clusters, kThreads, SMEM_BYTES, stream>>>(tmA, tmB, tmC, tmD, p);
  return check_launch(kFuse == 1 ? "gemm2_swiglu" : kFuse == 2 ? "gemm2_swiglu_bwd" : kFuse == 3 ? "gemm2_rope" : "gemm2_bf16");
}

}  // namespace gemm2

static int g_tail_split = 1;
void gemm2_set_tail_split(int on) { g_tail_split = on; }

// called from dla_gemm_bf16 (gemm.cu)
int gemm2_dispatch(const void* A, const void* B, void* C, const void* bias, int M, int N, int K, int64_t lda, int64_t ldb,
                   int64_t ldc, int a_mn, int b_mn, int c_fp32, int accumulate, cudaStream_t stream) {
  using namespace gemm2;
  CUtensorMap tmA, tmB, tmC;
  int rc;
  if (!a_mn) rc = make_tmap_2d_bf16(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM);
  else       rc = make_tmap_2d_bf16(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, BK);
  if (rc) return rc;
  if (!b_mn) rc = make_tmap_2d_bf16(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, BNH);
  else       rc = make_tmap_2d_bf16(&tmB, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, BK);
  if (rc) return rc;
  const int esz = c_fp32 ? 4 : 2;
  rc = make_tmap_2d(&tmC, C, c_fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, esz, (uint64_t)N,
                    (uint64_t)M, (uint64_t)ldc, c_fp32 ? 32 : 64, 32);
  if (rc) return rc;
  Params p{};
  p.M = M; p.N = N; p.K = K; p.c_fp32 = c_fp32; p.accumulate = accumulate;
  p.bias = reinterpret_cast<const bf16*>(bias);
  p.num_m_tiles = (M + 255) / 256;
  p.num_n_tiles = (N + BN - 1) / BN;
  // last partial wave: if it fills at most half of the clusters, split its tiles along K so the wave is full
  const int T = p.num_m_tiles * p.num_n_tiles, Cn = kNumSMs / 2, num_kb = (K + BK - 1) / BK;
  const int R = T > Cn ? T % Cn : 0;
  // Only where the slices add into storage that accumulates anyway (gradient buckets, fp32 outputs): a bf16 activation
  // tile keeps its single fp32-accumulated rounding, identical to the unsplit kernel.
  if (g_tail_split && bias == nullptr && (accumulate || c_fp32) && R > 0 && 2 * R <= Cn && num_kb >= 16) {
    int sp = Cn / R;
    if (sp > 8) sp = 8;
    if (sp > num_kb / 8) sp = num_kb / 8;
    while (sp > 1 && (sp - 1) * ((num_kb + sp - 1) / sp) >= num_kb) --sp;       // no empty K slice
    if (sp > 1) {
      p.tail_tiles = R;
      p.tail_splits = sp;
      if (!accumulate) {
        zero_tail_kernel<<<2 * R, 256, 0, stream>>>(reinterpret_cast<uint8_t*>(C), ldc * esz, esz, M, N, p.num_m_tiles, p.num_n_tiles, T - R);
        if (int zrc = check_launch("gemm2.zero_tail")) return zrc;
      }
    }
  }
  if (!a_mn && !b_mn) return launch<false, false>(tmA, tmB, tmC, tmC, p, stream);
  if (!a_mn && b_mn) return launch<false, true>(tmA, tmB, tmC, tmC, p, stream);
  if (a_mn && !b_mn) return launch<true, false>(tmA, tmB, tmC, tmC, p, stream);
  return launch<true, true>(tmA, tmB, tmC, tmC, p, stream);
}

// C[M, N] = rope(A[M,K] B[N,K]^T): the first n_rot_heads 128-wide heads of every row rotated by positions[row]
int gemm2_rope_dispatch(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc,
                        const float* cos_t, const float* sin_t, const int32_t* positions, int n_rot_heads, cudaStream_t stream) {
  using namespace gemm2;
  CUtensorMap tmA, tmB, tmC;
  int rc;
  if ((rc = make_tmap_2d_bf16(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM))) return rc;
  if ((rc = make_tmap_2d_bf16(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, BNH))) return rc;
  if ((rc = make_tmap_2d(&tmC, C, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (uint64_t)N, (uint64_t)M, (uint64_t)ldc, 64, 32))) return rc;
  Params p{};
  p.M = M; p.N = N; p.K = K; p.c_fp32 = 0; p.accumulate = 0;
  p.num_m_tiles = (M + 255) / 256;
  p.num_n_tiles = (N + BN - 1) / BN;
  p.cos_t = cos_t; p.sin_t = sin_t; p.positions = positions; p.n_rot_heads = n_rot_heads;
  return launch<false, false, 3>(tmA, tmB, tmC, tmC, p, stream);
}

// act[M, F] = silu(A W_gate^T) * (A W_up^T) with W = [gate rows | up rows] = B[2F, K]; gu (nullable) receives gate | up
int gemm2_swiglu_dispatch(const void* A, const void* B, void* GU, void* ACT, int M, int F, int K, int64_t lda, int64_t ldb,
                          int64_t ldgu, int64_t ldact, cudaStream_t stream) {
  using namespace gemm2;
  CUtensorMap tmA, tmB, tmC, tmD;
  int rc;
  if ((rc = make_tmap_2d_bf16(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM))) return rc;
  if ((rc = make_tmap_2d_bf16(&tmB, B, (uint64_t)K, (uint64_t)(2 * F), (uint64_t)ldb, BK, BNH))) return rc;
  if ((rc = make_tmap_2d(&tmD, ACT, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (uint64_t)F, (uint64_t)M, (uint64_t)ldact, 64, 32))) return rc;
  if (GU != nullptr) {
    if ((rc = make_tmap_2d(&tmC, GU, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (uint64_t)(2 * F), (uint64_t)M, (uint64_t)ldgu, 64, 32))) return rc;
  } else {
    tmC = tmD;
  }
  Params p{};
  p.M = M; p.N = 2 * F; p.K = K; p.c_fp32 = 0; p.accumulate = 0;
  p.num_m_tiles = (M + 255) / 256;
  p.num_n_tiles = F / BNH;
  p.F = F;
  p.store_gu = GU != nullptr;
  return launch<false, false, 1>(tmA, tmB, tmC, tmD, p, stream);
}

// d[gate | up][M, 2F] = swiglu_bwd(dY[M, H] @ Wdown[H, F], GU[M, 2F]) — the dgrad of the down projection with the SwiGLU
// backward as its epilogue
int gemm2_swiglu_bwd_dispatch(const void* dY, const void* Wdown, const void* GU, void* dGU, int M, int F, int H, int64_t lddy,
                              int64_t ldw, int64_t ldgu, int64_t lddgu, cudaStream_t stream) {
  using namespace gemm2;
  CUtensorMap tmA, tmB, tmD;
  int rc;
  if ((rc = make_tmap_2d_bf16(&tmA, dY, (uint64_t)H, (uint64_t)M, (uint64_t)lddy, BK, BM))) return rc;
  if ((rc = make_tmap_2d_bf16(&tmB, Wdown, (uint64_t)F, (uint64_t)H, (uint64_t)ldw, 64, BK))) return rc;      // MN-major B
  if ((rc = make_tmap_2d(&tmD, dGU, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (uint64_t)(2 * F), (uint64_t)M, (uint64_t)lddgu, 64, 32))) return rc;
  Params p{};
  p.M = M; p.N = F; p.K = H; p.c_fp32 = 0; p.accumulate = 0;
  p.num_m_tiles = (M + 255) / 256;
  p.num_n_tiles = (F + BN - 1) / BN;
  p.F = F;
  p.gu = reinterpret_cast<const bf16*>(GU);
  p.ldgu = ldgu;
  return launch<false, true, 2>(tmA, tmB, tmD, tmD, p, stream);
}

}  // namespace dla
