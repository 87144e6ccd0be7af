// bf16 GEMM on 5th-gen tensor cores: C[M,N] (+)= A[M,K] * B[N,K]^T, fp32 accumulation in TMEM.
//
// Persistent, warp-specialised, one CTA per SM:
//   warp 0      TMA producer  (cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier tx)
//   warp 1      MMA issuer    (one elected lane: tcgen05.mma 128x256x16, tcgen05.commit)
//   warps 2..5  epilogue      (tcgen05.ld from TMEM -> convert -> global), overlapped with the next
//                              tile's MMAs through two 256-column TMEM accumulators.
// Operand layouts: K-major or MN-major for either operand (forward / dgrad / wgrad of a Linear run
// on the same kernel without materialised transposes).  The epilogue can instead reduce each tile to
// log-softmax partials (fused lm_head log-prob).
//
// Reference path replaced: nn.Linear -> cuBLASLt (R/src/training/train_dpo.py:32) and
// log_softmax + gather over the vocab (R/src/training/train_dpo.py:33-38).
#include <math.h>
#include <stdlib.h>

#include <atomic>
#include <mutex>
#include <unordered_map>

#include "tc_common.cuh"

namespace dla {

// ------------------------------------------------------------------ host: tensor-map encode
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  });
  return fn;
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1, uint64_t ld_elems,
                      uint32_t box0, uint32_t box1) {
  PFN_encodeTiled fn = get_encode_fn();
  DLA_REQUIRE(fn != nullptr, DLA_E_DRIVER, "cuTensorMapEncodeTiled entry point not available");
  DLA_REQUIRE(((uintptr_t)base & 15) == 0 && (ld_elems * 2) % 16 == 0, DLA_E_BADARG,
              "tensor map: base must be 16-byte aligned and the row pitch a multiple of 8 elements (ld=%llu)",
              (unsigned long long)ld_elems);
  cuuint64_t dims[2] = {dim0, dim1};
  cuuint64_t strides[1] = {ld_elems * 2};
  cuuint32_t box[2] = {box0, box1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DLA_REQUIRE(r == CUDA_SUCCESS, DLA_E_DRIVER, "cuTensorMapEncodeTiled failed with %d (dims %llu x %llu, ld %llu, box %u x %u)",
              (int)r, (unsigned long long)dim0, (unsigned long long)dim1, (unsigned long long)ld_elems, box0, box1);
  return 0;
}

int make_tmap_2d(CUtensorMap* out, const void* base, CUtensorMapDataType dtype, int esize, uint64_t dim0, uint64_t dim1,
                 uint64_t ld_elems, uint32_t box0, uint32_t box1) {
  PFN_encodeTiled fn = get_encode_fn();
  DLA_REQUIRE(fn != nullptr, DLA_E_DRIVER, "cuTensorMapEncodeTiled entry point not available");
  DLA_REQUIRE(((uintptr_t)base & 15) == 0 && (ld_elems * esize) % 16 == 0, DLA_E_BADARG, "tensor map: unaligned base or pitch");
  cuuint64_t dims[2] = {dim0, dim1};
  cuuint64_t strides[1] = {ld_elems * (uint64_t)esize};
  cuuint32_t box[2] = {box0, box1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, dtype, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DLA_REQUIRE(r == CUDA_SUCCESS, DLA_E_DRIVER, "cuTensorMapEncodeTiled failed with %d (dims %llu x %llu, ld %llu, box %u x %u)",
              (int)r, (unsigned long long)dim0, (unsigned long long)dim1, (unsigned long long)ld_elems, box0, box1);
  return 0;
}

int make_tmap_4d_bf16(CUtensorMap* out, const void* base, const uint64_t dims[4], const uint64_t strides_elems[3],
                      const uint32_t box[4]) {
  return make_tmap_4d(out, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dims, strides_elems, box);
}

int make_tmap_4d(CUtensorMap* out, const void* base, CUtensorMapDataType dtype, int esize, const uint64_t dims[4],
                 const uint64_t strides_elems[3], const uint32_t box[4]) {
  PFN_encodeTiled fn = get_encode_fn();
  DLA_REQUIRE(fn != nullptr, DLA_E_DRIVER, "cuTensorMapEncodeTiled entry point not available");
  DLA_REQUIRE(((uintptr_t)base & 15) == 0, DLA_E_BADARG, "tensor map: base must be 16-byte aligned");
  cuuint64_t d[4], st[3];
  cuuint32_t bx[4], es[4] = {1, 1, 1, 1};
  for (int i = 0; i < 4; ++i) { d[i] = dims[i]; bx[i] = box[i]; }
  for (int i = 0; i < 3; ++i) {
    st[i] = strides_elems[i] * (uint64_t)esize;
    DLA_REQUIRE(st[i] % 16 == 0, DLA_E_BADARG, "tensor map: stride %d (%llu elements) must be a multiple of 8 elements", i,
                (unsigned long long)strides_elems[i]);
  }
  CUresult r = fn(out, dtype, 4, const_cast<void*>(base), d, st, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DLA_REQUIRE(r == CUDA_SUCCESS, DLA_E_DRIVER, "cuTensorMapEncodeTiled(4d) failed with %d", (int)r);
  return 0;
}

// ------------------------------------------------------------------ kernel
namespace gemm {

constexpr int BM = 128;          // rows of A per CTA tile (= TMEM lanes)
constexpr int BN = 256;          // columns per tile (= UMMA N, one TMEM accumulator)
constexpr int BK = 64;           // 64 bf16 = 128 B = one swizzle span
constexpr int UMMA_K = 16;
constexpr int kStages = 4;
constexpr int A_BYTES = BM * BK * 2;   // 16 KiB
constexpr int B_BYTES = BN * BK * 2;   // 32 KiB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int kThreads = 192;
constexpr int kEpiWarps = 4;
constexpr int EPI_BYTES = kEpiWarps * 2 * 4096;
constexpr int SMEM_BYTES = kStages * STAGE_BYTES + EPI_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr int GROUP_M = 16;

enum Epi { kStore = 0, kLogprob = 1 };

struct Params {
  int M, N, K;
  void* C;
  int64_t ldc;
  int c_fp32, accumulate;
  int store_mode;     // 0 = TMA store / reduce-add through staged smem, 1 = direct st.global from the TMEM layout
  int num_m_tiles, num_n_tiles;
  int k_splits;       // > 1: every output tile is computed by k_splits CTAs over disjoint K ranges and reduce-added (C pre-zeroed)
  const bf16* bias;   // [N] added to every output row in the epilogue (by K slice 0 only when split); nullptr: none
  // log-prob epilogue
  const int64_t* labels;
  float2* partials;   // [num_n_tiles, M] (max, sumexp)
  float* picked;      // [M] logit at the label
  // all-gather fused mode (path (b)): B's rows live sharded across ranks inside a flat unit buffer; the
  // first ag_ncopy CTAs pull them from the owners' memory over NVLink into `ag_full` tile by tile while the
  // other CTAs run the GEMM, each B tile being consumed as soon as its flag is published
  const bf16* const* ag_peers;   // [world] every rank's shard of the unit (ag_S elements each)
  bf16* ag_full;                 // local gathered unit buffer (B = ag_full + ag_b_off)
  int64_t ag_S, ag_b_off;
  uint32_t* ag_flags;            // [num_n_tiles] epoch-stamped "tile is local" flags
  uint32_t ag_epoch;
  int ag_ncopy, ag_world, ag_rank;
};

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

// AG mode: all m-tiles of one n-tile are consecutive, n-tiles start at this rank's own rows (already local)
__device__ __forceinline__ void tile_coords_ag(int t, int num_m, int num_n, int first_n, int& m_blk, int& n_blk) {
  m_blk = t % num_m;
  n_blk = (t / num_m + first_n) % num_n;
}

__device__ __forceinline__ int ag_first_tile(const Params& p) {
  // first n-tile whose rows start inside this rank's shard
  const int64_t tile_elems = (int64_t)BN * p.K;
  const int64_t my_lo = p.ag_S * p.ag_rank;
  int64_t t = (my_lo - p.ag_b_off + tile_elems - 1) / tile_elems;
  if (t < 0) t = 0;
  return (int)(t % p.num_n_tiles);
}

// Copy role of the all-gather fused GEMM: one thread drives TMA bulk copies peer HBM -> smem -> local HBM
// through a ring of 32 KiB buffers (several MB in flight chip-wide hides the ~2 us NVLink latency), then
// publishes the tile's flag with release semantics.
__device__ void ag_copy_role(const Params& p, uint8_t* smem, uint64_t* bars) {
  constexpr int kBuf = 6, kPiece = 32768, kAhead = kBuf - 2;
  if (threadIdx.x != 0) return;
  for (int b = 0; b < kBuf; ++b) tc::mbar_init(&bars[b], 1);
  tc::fence_barrier_init();
  const int first_n = ag_first_tile(p);
  const int64_t tile_elems = (int64_t)BN * p.K;
  const int64_t b_total = (int64_t)p.N * p.K;
  uint32_t phase_bits = 0;
  long long issued = 0, stored = 0;
  struct Piece { const bf16* src; bf16* dst; uint32_t bytes; };
  Piece ring[kBuf];
  for (int seq = blockIdx.x; seq < p.num_n_tiles; seq += p.ag_ncopy) {
    const int n_blk = (seq + first_n) % p.num_n_tiles;
    int64_t e = (int64_t)n_blk * tile_elems;                       // element offset inside B
    const int64_t e_end = (e + tile_elems < b_total) ? e + tile_elems : b_total;
    while (e < e_end || stored < issued) {
      if (e < e_end && issued - stored < kAhead) {
        const int64_t flat = p.ag_b_off + e;                       // offset inside the unit
        const int owner = (int)(flat / p.ag_S);
        const int64_t in_shard = flat - (int64_t)owner * p.ag_S;
        int64_t n = e_end - e;
        if (n > kPiece / 2) n = kPiece / 2;
        if (n > p.ag_S - in_shard) n = p.ag_S - in_shard;                   // do not straddle an owner boundary
        const int b = (int)(issued % kBuf);
        if (issued >= kBuf) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // store of piece issued-kBuf drained
        ring[b] = Piece{p.ag_peers[owner] + in_shard, p.ag_full + flat, (uint32_t)(n * 2)};
        tc::mbar_expect_tx(&bars[b], ring[b].bytes);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         tc::smem_u32(smem + b * kPiece)),
                     "l"(ring[b].src), "r"(ring[b].bytes), "r"(tc::smem_u32(&bars[b]))
                     : "memory");
        ++issued;
        e += n;
      } else {
        const int b = (int)(stored % kBuf);
        tc::mbar_wait(&bars[b], (phase_bits >> b) & 1u);
        phase_bits ^= 1u << b;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(ring[b].dst),
                     "r"(tc::smem_u32(smem + b * kPiece)), "r"(ring[b].bytes)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        ++stored;
      }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // the tile's bytes are in local HBM
    __threadfence();
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p.ag_flags + n_blk), "r"(p.ag_epoch) : "memory");
  }
}

template <bool kAMN, bool kBMN, int kEpi>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* epi_smem = smem + kStages * STAGE_BYTES;              // 4 epilogue warps x 2 x 4 KiB staging
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_smem + EPI_BYTES);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tfull_bar = empty_bar + kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int n_copy = p.ag_ncopy;
  if (n_copy > 0 && (int)blockIdx.x < n_copy) {      // lowest block ids are scheduled first: copies never starve
    ag_copy_role(p, smem, full_bar);
    return;
  }
  const int cta = (int)blockIdx.x - n_copy, n_cta = (int)gridDim.x - n_copy;
  const int ag_first = n_copy > 0 ? ag_first_tile(p) : 0;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = (p.K + BK - 1) / BK;
  const int out_tiles = p.num_m_tiles * p.num_n_tiles;
  // split-K (small-M problems: too few output tiles to fill the machine): work item w = (K slice w / out_tiles,
  // output tile w % out_tiles); slices of one tile land in C through the reduce-add epilogue
  const int k_splits = p.k_splits > 1 ? p.k_splits : 1;
  const int kb_per = (num_kb + k_splits - 1) / k_splits;
  const int num_tiles = out_tiles * k_splits;

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmA);
    tc::prefetch_tmap(&tmB);
    for (int s = 0; s < kStages; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      tc::mbar_init(&tfull_bar[a], 1);
      tc::mbar_init(&tempty_bar[a], kEpiWarps);
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc<1>(tmem_holder, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = cta; t < num_tiles; t += n_cta) {
        int m_blk, n_blk;
        if (n_copy > 0) {
          tile_coords_ag(t, p.num_m_tiles, p.num_n_tiles, ag_first, m_blk, n_blk);
          // wait until the copy CTAs have landed this B tile in local memory, then order the generic-proxy
          // flag read before the async-proxy (TMA) reads of the tile
          // Liveness: the copy CTAs are the lowest block ids of a grid of at most one CTA per SM, so they are resident
          // whenever a compute CTA is (the hardware dispatches blocks in id order) — unless another kernel holds SMs
          // for the whole time or a peer is dead.  The wait is therefore bounded by wall time (%globaltimer): after
          // 30 s the kernel traps, which surfaces on the host as a launch failure instead of a GPU that never returns.
          uint32_t seen;
          unsigned long long t0 = 0;
          do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.ag_flags + n_blk) : "memory");
            if (seen != p.ag_epoch) {
              unsigned long long now;
              asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
              if (t0 == 0) t0 = now;
              if (now - t0 > 30000000000ULL) __trap();
              __nanosleep(100);
            }
          } while (seen != p.ag_epoch);
          asm volatile("fence.proxy.async;" ::: "memory");
        } else {
          tile_coords(t % out_tiles, p.num_m_tiles, p.num_n_tiles, m_blk, n_blk);
        }
        const int m0 = m_blk * BM, n0 = n_blk * BN;
        const int kb0 = (t / out_tiles) * kb_per, kb1 = min(num_kb, kb0 + kb_per);
        for (int kb = kb0; kb < kb1; ++kb) {
          tc::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          tc::mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
          const int k0 = kb * BK;
          if constexpr (!kAMN) {
            tc::tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c) tc::tma_load_2d(sa + c * (64 * BK * 2), &tmA, &full_bar[stage], m0 + c * 64, k0);
          }
          if constexpr (!kBMN) {
            tc::tma_load_2d(sb, &tmB, &full_bar[stage], k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c) tc::tma_load_2d(sb + c * (64 * BK * 2), &tmB, &full_bar[stage], n0 + c * 64, k0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::make_idesc_bf16(BM, BN, kAMN ? 1 : 0, kBMN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = cta; t < num_tiles; t += n_cta, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        tc::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc::tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        const int kb0 = (t / out_tiles) * kb_per, kb1 = min(num_kb, kb0 + kb_per);
        for (int kb = kb0; kb < kb1; ++kb) {
          tc::mbar_wait(&full_bar[stage], phase);
          tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // K-major: step 16 elements = 32 B inside the 128 B swizzle span.
            // MN-major: step 16 K-rows of 128 B = 2048 B; LBO = 8 KiB between 64-wide MN chunks.
            const uint64_t da = kAMN ? tc::make_smem_desc(sa + k * 2048, 64 * BK * 2, 1024)
                                     : tc::make_smem_desc(sa + k * 32, 16, 1024);
            const uint64_t db = kBMN ? tc::make_smem_desc(sb + k * 2048, 64 * BK * 2, 1024)
                                     : tc::make_smem_desc(sb + k * 32, 16, 1024);
            tc::umma_bf16<1>(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          tc::umma_commit(&empty_bar[stage]);       // smem slot reusable once these MMAs retire
          if (kb == kb1 - 1) tc::umma_commit(&tfull_bar[acc]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
    int it = 0;
    for (int t = cta; t < num_tiles; t += n_cta, ++it) {
      int m_blk, n_blk;
      if (n_copy > 0) tile_coords_ag(t, p.num_m_tiles, p.num_n_tiles, ag_first, m_blk, n_blk);
      else tile_coords(t % out_tiles, p.num_m_tiles, p.num_n_tiles, m_blk, n_blk);
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      tc::mbar_wait(&tfull_bar[acc], acc_phase);
      tc::tc_fence_after();
      const int row = m_blk * BM + quarter * 32 + lane;
      const int col_base = n_blk * BN;
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BN;

      if constexpr (kEpi == kStore) {
       const bf16* bias = (t / out_tiles == 0) ? p.bias : nullptr;      // split-K: slice 0 carries the bias
       if (p.store_mode == 0) {
        // TMEM -> registers -> 128B-swizzled staging rows -> TMA store (or TMA reduce-add when accumulating).
        // The output never touches the LSU path: uncoalesced per-row st.global from the TMEM layout used
        // to eat ~45% of the L1TEX data pipe, which the tensor core shares for its operand reads.
        const int gc = p.c_fp32 ? 32 : 64;                 // output columns per 128-byte staging row
        const int n_groups = BN / gc;
        uint8_t* stage_base = epi_smem + (warp - 2) * 2 * 4096;
        const int r0 = m_blk * BM + quarter * 32;
#pragma unroll 1
        for (int gi = 0; gi < n_groups; ++gi) {
          uint8_t* buf = stage_base + (gi & 1) * 4096;
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // store from 2 groups ago read `buf`
          __syncwarp();
          uint8_t* myrow = buf + lane * 128;
          if (p.c_fp32) {
            uint32_t r[32];
            tc::tmem_ld_32x32(taddr + gi * 32, r);
            tc::tmem_ld_wait();
            if (bias != nullptr) tc::add_bias_row<32>(r, bias, col_base + gi * 32, p.N);
#pragma unroll
            for (int c = 0; c < 8; ++c)
              *reinterpret_cast<uint4*>(myrow + ((c ^ (lane & 7)) << 4)) = make_uint4(r[4 * c], r[4 * c + 1], r[4 * c + 2], r[4 * c + 3]);
          } else {
            uint32_t r[64];
            tc::tmem_ld_32x32(taddr + gi * 64, r);
            tc::tmem_ld_32x32(taddr + gi * 64 + 32, r + 32);
            tc::tmem_ld_wait();
            if (bias != nullptr) tc::add_bias_row<64>(r, bias, col_base + gi * 64, p.N);
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
            if (c0 < p.N && r0 < p.M) {       // TMA clips the box at the tensor bounds
              if (p.accumulate) tc::tma_reduce_add_2d(&tmC, buf, c0, r0);
              else tc::tma_store_2d(&tmC, buf, c0, r0);
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
       } else {
#pragma unroll 1
        for (int ch = 0; ch < BN / 32; ++ch) {
          uint32_t r[32];
          tc::tmem_ld_32x32(taddr + ch * 32, r);
          tc::tmem_ld_wait();
          const int col0 = col_base + ch * 32;
          if (bias != nullptr && col0 < p.N) tc::add_bias_row<32>(r, bias, col0, p.N);
          if (row < p.M && col0 < p.N) {
            const bool full = (col0 + 32 <= p.N);
            if (p.c_fp32) {
              float* dst = reinterpret_cast<float*>(p.C) + (int64_t)row * p.ldc + col0;
              if (full && (p.ldc & 3) == 0) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                  float4 v = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
                  if (p.accumulate) {
                    const float4 o = *reinterpret_cast<const float4*>(dst + j);
                    v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
                  }
                  *reinterpret_cast<float4*>(dst + j) = v;
                }
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (col0 + j < p.N) dst[j] = __uint_as_float(r[j]) + (p.accumulate ? dst[j] : 0.f);
              }
            } else {
              bf16* dst = reinterpret_cast<bf16*>(p.C) + (int64_t)row * p.ldc + col0;
              if (full && (p.ldc & 7) == 0) {
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                  float f[8];
#pragma unroll
                  for (int q = 0; q < 8; ++q) f[q] = __uint_as_float(r[j + q]);
                  if (p.accumulate) {
                    float o[8];
                    unpack8(ld8(dst + j), o);
#pragma unroll
                    for (int q = 0; q < 8; ++q) f[q] += o[q];
                  }
                  st8(dst + j, pack8(f));
                }
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (col0 + j < p.N)
                    dst[j] = __float2bfloat16(__uint_as_float(r[j]) + (p.accumulate ? __bfloat162float(dst[j]) : 0.f));
              }
            }
          }
        }
       }
      } else {
        // ---- fused log-softmax partials: (max, sum exp(x - max)) of this row over the tile's columns
        constexpr float kLog2e = 1.4426950408889634f;
        float mx = -INFINITY, sum = 0.f, picked = 0.f;
        bool has_pick = false;
        int rel = -1;
        if (row < p.M) {
          const int64_t lab = p.labels[row];
          if (lab >= col_base && lab < col_base + BN) rel = (int)(lab - col_base);
        }
#pragma unroll 1
        for (int ch = 0; ch < BN / 32; ++ch) {
          uint32_t r[32];
          tc::tmem_ld_32x32(taddr + ch * 32, r);
          tc::tmem_ld_wait();
          const int col0 = col_base + ch * 32;
          if (col0 >= p.N) continue;
          const int nvalid = min(32, p.N - col0);
          float cm = -INFINITY;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < nvalid) cm = fmaxf(cm, __uint_as_float(r[j]));
          const float nm = fmaxf(mx, cm);
          float cs = 0.f;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < nvalid) cs += exp2f((__uint_as_float(r[j]) - nm) * kLog2e);
          sum = sum * exp2f((mx - nm) * kLog2e) + cs;
          mx = nm;
          if (p.C != nullptr && row < p.M) {
            // keep the logits (bf16) for the backward: it then skips the recomputation GEMM.  64 contiguous bytes per
            // lane and chunk; a tile's stores are spread over its 24 us of MMA time, so the LSU cost is noise.
            bf16* dst = reinterpret_cast<bf16*>(p.C) + (int64_t)row * p.ldc + col0;
            if (nvalid == 32) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                float f[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) f[q] = __uint_as_float(r[j + q]);
                st8(dst + j, pack8(f));
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < nvalid) dst[j] = __float2bfloat16(__uint_as_float(r[j]));
            }
          }
          if ((rel >> 5) == ch && rel >= 0) {
            // r[rel & 31] by a 5-level select tree on the index bits: static register indices only.  (The obvious compare
            // chain was compiled into a dynamically indexed read of a local-memory copy of r, i.e. 128 bytes of local
            // stores per lane and chunk for every row, taken or not.)
            const int jj = rel & 31;
            uint32_t s16[16], s8[8], s4[4], s2[2];
#pragma unroll
            for (int k = 0; k < 16; ++k) s16[k] = (jj & 16) ? r[k + 16] : r[k];
#pragma unroll
            for (int k = 0; k < 8; ++k) s8[k] = (jj & 8) ? s16[k + 8] : s16[k];
#pragma unroll
            for (int k = 0; k < 4; ++k) s4[k] = (jj & 4) ? s8[k + 4] : s8[k];
#pragma unroll
            for (int k = 0; k < 2; ++k) s2[k] = (jj & 2) ? s4[k + 2] : s4[k];
            picked = __uint_as_float((jj & 1) ? s2[1] : s2[0]);
            has_pick = true;
          }
        }
        if (row < p.M) {
          p.partials[(int64_t)n_blk * p.M + row] = make_float2(mx, sum);
          if (has_pick) p.picked[row] = picked;
        }
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&tempty_bar[acc]);
    }
  }

  if (kEpi == kStore && warp >= 2 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc<1>(tmem_base, 512);
  }
}

// Combine per-tile (max, sumexp) partials into lse and the label log-prob.
__global__ void logprob_combine_kernel(const float2* __restrict__ partials, const float* __restrict__ picked,
                                       const int64_t* __restrict__ labels, int M, int num_n_tiles,
                                       float* __restrict__ logp, float* __restrict__ lse) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  float mx = -INFINITY, sum = 0.f;
  for (int t = 0; t < num_n_tiles; ++t) {
    const float2 v = partials[(int64_t)t * M + m];
    const float nm = fmaxf(mx, v.x);
    sum = sum * __expf(mx - nm) + v.y * __expf(v.x - nm);
    mx = nm;
  }
  const float l = mx + logf(sum);
  if (lse != nullptr) lse[m] = l;
  logp[m] = labels[m] >= 0 ? picked[m] - l : 0.f;
}

// In place on a bf16 logits chunk: g * (onehot - softmax)
__global__ void __launch_bounds__(256)
logprob_grad_kernel(bf16* __restrict__ logits, const float* __restrict__ lse, const int64_t* __restrict__ labels,
                    const float* __restrict__ grow, int Mc, int V, int64_t ld) {
  const int m = blockIdx.y;
  if (m >= Mc) return;
  const int64_t lab = labels[m];
  const float g = lab >= 0 ? grow[m] : 0.f;   // ignored positions (label < 0) carry no gradient
  const float l = lse[m];
  bf16* row = logits + (int64_t)m * ld;
  const int nvec = V >> 3;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += gridDim.x * blockDim.x) {
    float f[8];
    unpack8(ld8(row + i * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float pr = __expf(f[j] - l);
      f[j] = g * (((int64_t)(i * 8 + j) == lab ? 1.f : 0.f) - pr);
    }
    st8(row + i * 8, pack8(f));
  }
  if (blockIdx.x == 0 && threadIdx.x < (V & 7)) {
    const int c = (nvec << 3) + threadIdx.x;
    const float pr = __expf(__bfloat162float(row[c]) - l);
    row[c] = __float2bfloat16(g * ((c == lab ? 1.f : 0.f) - pr));
  }
}

static int g_variant = 0;
static int g_split_k = 1;        // small-M split-K (dla_gemm_set_variant(3) turns it off for A/B runs)
static int g_auto_2cta = 1;      // measured on B200 (profiles/kernels_r1_2cta.jsonl, tools/gpu/ab_k.sh)

template <bool kAMN, bool kBMN, int kEpi>
static int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const Params& p, cudaStream_t stream) {
  auto kern = gemm_kernel<kAMN, kBMN, kEpi>;
  // the opt-in to > 48 KB of dynamic shared memory is a per-device function attribute: set it once on every device
  // this process launches on (the allocation-free entry point promises to work on each device it is called on)
  static std::atomic<unsigned long long> attr_devices{0};
  int dev = 0;
  DLA_CUDA(cudaGetDevice(&dev));
  const unsigned long long bit = 1ULL << (dev & 63);
  if (!(attr_devices.load(std::memory_order_relaxed) & bit)) {
    DLA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    attr_devices.fetch_or(bit, std::memory_order_relaxed);
  }
  const int num_tiles = p.num_m_tiles * p.num_n_tiles * (p.k_splits > 1 ? p.k_splits : 1);
  int grid = num_tiles < kNumSMs ? num_tiles : kNumSMs;
  if (p.ag_ncopy > 0) grid = kNumSMs;     // copy + compute CTAs must all be co-resident (1 CTA / SM)
  kern<<<grid, kThreads, SMEM_BYTES, stream>>>(tmA, tmB, tmC, p);
  return check_launch("gemm_bf16");
}

static int build_maps(CUtensorMap* tmA, CUtensorMap* tmB, const void* A, const void* B, int M, int N, int K,
                      int64_t lda, int64_t ldb, int a_mn, int b_mn) {
  int rc;
  if (!a_mn) rc = make_tmap_2d_bf16(tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM);
  else       rc = make_tmap_2d_bf16(tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, BK);
  if (rc) return rc;
  if (!b_mn) rc = make_tmap_2d_bf16(tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, BN);
  else       rc = make_tmap_2d_bf16(tmB, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, BK);
  return rc;
}

}  // namespace gemm
}  // namespace dla

namespace dla {
int gemm2_dispatch(const void* A, const void* B, void* C, const void* bias, int M, int N, int K, int64_t lda, int64_t ldb,
                   int64_t ldc, int a_mn, int b_mn, int c_fp32, int accumulate, cudaStream_t stream);
int gemm2_swiglu_dispatch(const void* A, const void* B, void* GU, void* ACT, int M, int F, int K, int64_t lda, int64_t ldb,
                          int64_t ldgu, int64_t ldact, cudaStream_t stream);
int gemm2_swiglu_bwd_dispatch(const void* dY, const void* Wdown, const void* GU, void* dGU, int M, int F, int H, int64_t lddy,
                              int64_t ldw, int64_t ldgu, int64_t lddgu, cudaStream_t stream);
int gemm2_rope_dispatch(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc,
                        const float* cos_t, const float* sin_t, const int32_t* positions, int n_rot_heads, cudaStream_t stream);
void gemm2_set_tail_split(int on);
}

using namespace dla;
using namespace dla::gemm;

static int gemm_entry(const void* A, const void* B, void* C, const void* bias, int M, int N, int K, int64_t lda, int64_t ldb,
                      int64_t ldc, int a_mn_major, int b_mn_major, int c_fp32, int accumulate, void* stream);

extern "C" {

int dla_gemm_set_variant(int variant) {
  if (variant == 3) { g_split_k = 0; return 0; }     // 3 / 4: small-M split-K off / on, the kernel choice is untouched
  if (variant == 4) { g_split_k = 1; return 0; }
  if (variant == 5) { gemm2_set_tail_split(0); return 0; }   // 5 / 6: last-wave K split of the 2-CTA kernel off / on
  if (variant == 6) { gemm2_set_tail_split(1); return 0; }
  g_variant = variant;
  return 0;
}

int dla_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc,
                  int a_mn_major, int b_mn_major, int c_fp32, int accumulate, void* stream) {
  return gemm_entry(A, B, C, nullptr, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, c_fp32, accumulate, stream);
}

int dla_gemm_bias_bf16(const void* A, const void* B, void* C, const void* bias, int M, int N, int K, int64_t lda, int64_t ldb,
                       int64_t ldc, int c_fp32, void* stream) {
  DLA_REQUIRE(bias != nullptr, DLA_E_BADARG, "gemm_bias: bias is required (use dla_gemm_bf16 without one)");
  return gemm_entry(A, B, C, bias, M, N, K, lda, ldb, ldc, 0, 0, c_fp32, 0, stream);
}

}  // extern "C"

static int gemm_entry(const void* A, const void* B, void* C, const void* bias, int M, int N, int K, int64_t lda, int64_t ldb,
                      int64_t ldc, int a_mn_major, int b_mn_major, int c_fp32, int accumulate, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(M > 0 && N > 0 && K > 0, DLA_E_BADARG, "gemm: empty problem %dx%dx%d", M, N, K);
  DLA_REQUIRE(K % 8 == 0, DLA_E_BADARG, "gemm: K=%d must be a multiple of 8", K);
  DLA_REQUIRE(!a_mn_major || M % 8 == 0, DLA_E_BADARG, "gemm: MN-major A needs M %% 8 == 0 (M=%d)", M);
  DLA_REQUIRE(!b_mn_major || N % 8 == 0, DLA_E_BADARG, "gemm: MN-major B needs N %% 8 == 0 (N=%d)", N);
  CUtensorMap tmA, tmB;
  int rc = build_maps(&tmA, &tmB, A, B, M, N, K, lda, ldb, a_mn_major, b_mn_major);
  if (rc) return rc;
  Params p{};
  p.M = M; p.N = N; p.K = K; p.C = C; p.ldc = ldc; p.c_fp32 = c_fp32; p.accumulate = accumulate;
  p.bias = reinterpret_cast<const bf16*>(bias);
  p.num_m_tiles = (M + BM - 1) / BM;
  p.num_n_tiles = (N + BN - 1) / BN;
  const int64_t esz = c_fp32 ? 4 : 2;
  const bool tma_ok = (((uintptr_t)C & 15) == 0) && ((ldc * esz) % 16 == 0) && g_variant != 1;
  p.store_mode = tma_ok ? 0 : 1;
  // 2-CTA (cta_group::2) kernel: 256x256 cluster tiles; worth it once there is about a wave of them
  const int64_t tiles2 = (int64_t)((M + 255) / 256) * ((N + 255) / 256);
  static const int max_k_2cta = [] {
    const char* e = getenv("DLA_GEMM_2CTA_MAX_K");       // A/B knob.  In isolation the 1-CTA kernel wins above K = 16K, inside the
                                                         // power-capped step the 2-CTA kernel (less shared-memory traffic per flop) does
    return e != nullptr ? atoi(e) : 32768;
  }();
  if (tma_ok && (g_variant == 2 || (g_variant == 0 && g_auto_2cta && tiles2 >= kNumSMs / 2 && M >= 256 && K <= max_k_2cta)))
    return gemm2_dispatch(A, B, C, bias, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, c_fp32, accumulate, (cudaStream_t)stream);
  CUtensorMap tmC{};
  if (tma_ok) {
    rc = make_tmap_2d(&tmC, C, c_fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (int)esz, (uint64_t)N,
                      (uint64_t)M, (uint64_t)ldc, c_fp32 ? 32 : 64, 32);
    if (rc) return rc;
  }
  cudaStream_t s = (cudaStream_t)stream;
  // split-K when the output has far fewer tiles than SMs and K is long enough to share (prefill / small-batch
  // forward: M = 256 gives 32 tiles for a 4096-wide projection).  Slices meet in C through TMA reduce-add, so C is
  // zero-filled first unless the caller accumulates anyway.
  const int out_tiles = p.num_m_tiles * p.num_n_tiles, num_kb = (K + BK - 1) / BK;
  if (g_split_k && tma_ok && out_tiles * 2 <= kNumSMs && num_kb >= 16) {
    int splits = kNumSMs / out_tiles;
    if (splits > 8) splits = 8;
    if (splits > num_kb / 8) splits = num_kb / 8;
    while (splits > 1 && (splits - 1) * ((num_kb + splits - 1) / splits) >= num_kb) --splits;   // no empty K slice (it would never signal its epilogue)
    if (splits > 1) {
      p.k_splits = splits;
      if (!accumulate) {
        DLA_CUDA(cudaMemset2DAsync(C, (size_t)ldc * esz, 0, (size_t)N * esz, (size_t)M, s));
        p.accumulate = 1;
      }
    }
  }
  if (!a_mn_major && !b_mn_major) return launch<false, false, kStore>(tmA, tmB, tmC, p, s);
  if (!a_mn_major && b_mn_major) return launch<false, true, kStore>(tmA, tmB, tmC, p, s);
  if (a_mn_major && !b_mn_major) return launch<true, false, kStore>(tmA, tmB, tmC, p, s);
  return launch<true, true, kStore>(tmA, tmB, tmC, p, s);
}

extern "C" {

int dla_gemm_swiglu_bf16(const void* A, const void* B, void* GU, void* ACT, int M, int F, int K, int64_t lda, int64_t ldb,
                         int64_t ldgu, int64_t ldact, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(M > 0 && F > 0 && K > 0 && K % 8 == 0, DLA_E_BADARG, "gemm_swiglu: bad shape M=%d F=%d K=%d", M, F, K);
  DLA_REQUIRE(F % 128 == 0 && M >= 256, DLA_E_UNSUPP, "gemm_swiglu: needs F %% 128 == 0 and M >= 256 (F=%d, M=%d)", F, M);
  DLA_REQUIRE(((uintptr_t)ACT & 15) == 0 && (ldact * 2) % 16 == 0 && (GU == nullptr || ((((uintptr_t)GU & 15) == 0) && (ldgu * 2) % 16 == 0)),
              DLA_E_BADARG, "gemm_swiglu: outputs must be 16-byte aligned with 16-byte row pitch");
  return gemm2_swiglu_dispatch(A, B, GU, ACT, M, F, K, lda, ldb, ldgu, ldact, (cudaStream_t)stream);
}

int dla_gemm_rope_bf16(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc,
                       const float* cos_t, const float* sin_t, const int32_t* positions, int n_rot_heads, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(M > 0 && N > 0 && K > 0 && K % 8 == 0, DLA_E_BADARG, "gemm_rope: bad shape %dx%dx%d", M, N, K);
  DLA_REQUIRE(N % 128 == 0 && n_rot_heads >= 0 && n_rot_heads * 128 <= N, DLA_E_UNSUPP,
              "gemm_rope: needs 128-wide heads (N=%d, n_rot_heads=%d)", N, n_rot_heads);
  DLA_REQUIRE(cos_t != nullptr && sin_t != nullptr && positions != nullptr, DLA_E_BADARG, "gemm_rope: tables and positions are required");
  DLA_REQUIRE(((uintptr_t)C & 15) == 0 && (ldc * 2) % 16 == 0 && (((uintptr_t)cos_t | (uintptr_t)sin_t) & 15) == 0, DLA_E_BADARG,
              "gemm_rope: C and the tables must be 16-byte aligned, C with a 16-byte row pitch");
  return gemm2_rope_dispatch(A, B, C, M, N, K, lda, ldb, ldc, cos_t, sin_t, positions, n_rot_heads, (cudaStream_t)stream);
}

int dla_gemm_swiglu_bwd_bf16(const void* dY, const void* Wdown, const void* GU, void* dGU, int M, int F, int H, int64_t lddy,
                             int64_t ldw, int64_t ldgu, int64_t lddgu, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(M >= 256 && F > 0 && H > 0 && H % 8 == 0 && F % 64 == 0, DLA_E_UNSUPP,
              "gemm_swiglu_bwd: needs M >= 256, H %% 8 == 0, F %% 64 == 0 (M=%d F=%d H=%d)", M, F, H);
  DLA_REQUIRE((((uintptr_t)GU | (uintptr_t)dGU) & 15) == 0 && ldgu % 8 == 0 && lddgu % 8 == 0, DLA_E_BADARG,
              "gemm_swiglu_bwd: gate|up buffers must be 16-byte aligned with row pitches that are multiples of 8");
  return gemm2_swiglu_bwd_dispatch(dY, Wdown, GU, dGU, M, F, H, lddy, ldw, ldgu, lddgu, (cudaStream_t)stream);
}

int dla_ag_gemm_bf16(const void* A, const void* const* peer_shards, void* unit_full, int64_t shard_elems, int64_t b_offset,
                     void* C, int M, int N, int K, int64_t lda, int64_t ldc, int world, int rank, uint32_t* flags,
                     uint32_t epoch, int n_copy_ctas, int c_fp32, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(M > 0 && N > 0 && K > 0 && K % 8 == 0, DLA_E_BADARG, "ag_gemm: bad shape %dx%dx%d", M, N, K);
  DLA_REQUIRE(world >= 1 && rank >= 0 && rank < world && shard_elems % 8 == 0 && b_offset % 8 == 0, DLA_E_BADARG,
              "ag_gemm: bad world/rank/shard (%d/%d/%lld)", world, rank, (long long)shard_elems);
  DLA_REQUIRE(b_offset + (int64_t)N * K <= shard_elems * world, DLA_E_BADARG, "ag_gemm: B exceeds the unit");
  DLA_REQUIRE(n_copy_ctas >= 1 && n_copy_ctas <= 64 && flags != nullptr, DLA_E_BADARG, "ag_gemm: n_copy_ctas in [1,64] and flags required");
  bf16* B = reinterpret_cast<bf16*>(unit_full) + b_offset;
  CUtensorMap tmA, tmB;
  int rc = build_maps(&tmA, &tmB, A, B, M, N, K, lda, K, 0, 0);
  if (rc) return rc;
  Params p{};
  p.M = M; p.N = N; p.K = K; p.C = C; p.ldc = ldc; p.c_fp32 = c_fp32; p.accumulate = 0;
  p.num_m_tiles = (M + BM - 1) / BM;
  p.num_n_tiles = (N + BN - 1) / BN;
  const int64_t esz = c_fp32 ? 4 : 2;
  DLA_REQUIRE((((uintptr_t)C & 15) == 0) && ((ldc * esz) % 16 == 0), DLA_E_BADARG, "ag_gemm: C must be 16-byte aligned with a 16-byte pitch");
  p.store_mode = 0;
  CUtensorMap tmC{};
  rc = make_tmap_2d(&tmC, C, c_fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (int)esz, (uint64_t)N,
                    (uint64_t)M, (uint64_t)ldc, c_fp32 ? 32 : 64, 32);
  if (rc) return rc;
  p.ag_peers = reinterpret_cast<const bf16* const*>(peer_shards);
  p.ag_full = reinterpret_cast<bf16*>(unit_full);
  p.ag_S = shard_elems; p.ag_b_off = b_offset; p.ag_flags = flags; p.ag_epoch = epoch;
  p.ag_ncopy = n_copy_ctas; p.ag_world = world; p.ag_rank = rank;
  return launch<false, false, kStore>(tmA, tmB, tmC, p, (cudaStream_t)stream);
}

size_t dla_lmhead_scratch_bytes(int M, int V) {
  const size_t nt = (size_t)(V + BN - 1) / BN;
  return nt * (size_t)M * sizeof(float2) + (size_t)M * sizeof(float) + 256;
}

int dla_lmhead_logprob(const void* X, const void* W, const int64_t* labels, int M, int V, int K, int64_t ldx,
                       int64_t ldw, void* scratch, float* logp, float* lse, void* stream) {
  return dla_lmhead_logprob_keep(X, W, labels, M, V, K, ldx, ldw, scratch, logp, lse, nullptr, 0, stream);
}

int dla_lmhead_logprob_keep(const void* X, const void* W, const int64_t* labels, int M, int V, int K, int64_t ldx,
                            int64_t ldw, void* scratch, float* logp, float* lse, void* logits_out, int64_t ld_logits,
                            void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(logits_out == nullptr || (ld_logits % 8 == 0 && ld_logits >= V && ((uintptr_t)logits_out & 15) == 0), DLA_E_BADARG,
              "lmhead_logprob: kept logits need a 16-byte aligned buffer with a row pitch that is a multiple of 8 and >= V");
  DLA_REQUIRE(M > 0 && V > 0 && K > 0 && K % 8 == 0, DLA_E_BADARG, "lmhead_logprob: bad shape M=%d V=%d K=%d", M, V, K);
  DLA_REQUIRE(scratch != nullptr && logp != nullptr, DLA_E_BADARG, "lmhead_logprob: scratch and logp are required");
  CUtensorMap tmA, tmB;
  int rc = build_maps(&tmA, &tmB, X, W, M, V, K, ldx, ldw, 0, 0);
  if (rc) return rc;
  Params p{};
  p.M = M; p.N = V; p.K = K;
  p.num_m_tiles = (M + BM - 1) / BM;
  p.num_n_tiles = (V + BN - 1) / BN;
  p.labels = labels;
  p.partials = reinterpret_cast<float2*>(scratch);
  p.picked = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(scratch) + (size_t)p.num_n_tiles * M * sizeof(float2));
  p.C = logits_out;
  p.ldc = ld_logits;
  cudaStream_t s = (cudaStream_t)stream;
  rc = launch<false, false, kLogprob>(tmA, tmB, CUtensorMap{}, p, s);
  if (rc) return rc;
  logprob_combine_kernel<<<(M + 255) / 256, 256, 0, s>>>(p.partials, p.picked, labels, M, p.num_n_tiles, logp, lse);
  return check_launch("lmhead_logprob.combine");
}

int dla_logprob_grad_inplace(void* logits, const float* lse, const int64_t* labels, const float* grow, int Mc, int V,
                             int64_t ld, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(ld % 8 == 0, DLA_E_BADARG, "logprob_grad: ld=%lld must be a multiple of 8", (long long)ld);
  if (Mc == 0) return 0;
  int bx = ((V >> 3) + 255) / 256;
  if (bx > 64) bx = 64;
  if (bx < 1) bx = 1;
  dim3 grid(bx, Mc);
  logprob_grad_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((bf16*)logits, lse, labels, grow, Mc, V, ld);
  return check_launch("logprob_grad");
}

}  // extern "C"

// ====================================================================================================
// Allocation-free variant of the fused lm_head log-prob.
//
// Who needs it: code that scores tokens from inside a process that cannot obtain device memory on
// demand — typically an inference / scoring server that handed (almost) all of HBM to its own KV-cache
// or tensor pool at start-up, so a cudaMalloc for our tile-partials work area would fail or fragment that
// pool, and callers that run under a fixed memory budget audited ahead of time.  For them the work area
// lives inside the library image: a file-scope __device__ array of a fixed, modest size.  Every GPU that
// loads the library holds these bytes until the process exits, whether or not the entry point is called.
//
// Limits (stated, checked): M <= DLA_NOALLOC_MAX_ROWS rows per call and ceil(V / 256) <= DLA_NOALLOC_MAX_TILES
// vocabulary tiles, i.e. V <= 131072 (Llama-3's 128256 fits).  Work area: 512 * 1024 * 8 B + 1024 * 4 B ~ 4 MiB.
// ====================================================================================================
#define DLA_NOALLOC_MAX_ROWS 1024
#define DLA_NOALLOC_MAX_TILES 512

struct __align__(256) NoAllocLogprobArea {
  float2 partials[DLA_NOALLOC_MAX_TILES * DLA_NOALLOC_MAX_ROWS];   // [tile][row] (max, sum-exp)
  float picked[DLA_NOALLOC_MAX_ROWS];                              // logit at the label
};
static __device__ __align__(256) NoAllocLogprobArea g_noalloc_logprob_area[1];

namespace {
constexpr int kMaxDevices = 64;
// Device addresses of the array above, one per GPU.  In host code `g_noalloc_logprob_area` is only a
// handle for the runtime, never something a kernel may dereference: the real address differs per device
// and has to be asked for with that device current.
NoAllocLogprobArea* g_noalloc_dev_ptr[kMaxDevices] = {};
std::mutex g_noalloc_mu;

int noalloc_area_for_current_device(NoAllocLogprobArea** out) {
  int dev = -1;
  DLA_CUDA(cudaGetDevice(&dev));
  DLA_REQUIRE(dev >= 0 && dev < kMaxDevices, DLA_E_UNSUPP, "no-alloc log-prob: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_dev_ptr[dev] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_logprob_area);
    if (e != cudaSuccess || p == nullptr) {
      dla::set_error("no-alloc log-prob: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : DLA_E_DRIVER;
    }
    g_noalloc_dev_ptr[dev] = reinterpret_cast<NoAllocLogprobArea*>(p);
  }
  *out = g_noalloc_dev_ptr[dev];
  return 0;
}
}  // namespace

extern "C" {

// Same result as dla_lmhead_logprob, but the tile partials go to the library's built-in work area:
// no cudaMalloc / cudaFree / allocating library call happens here, on any call.
//   * The FIRST call on each device looks up that device's address of the work area (and sets the
//     kernel's shared-memory attribute).  Make that first call during start-up, before any stream capture.
//   * All calls on one GPU share the one work area, which keeps whatever the previous call left in it.
//     The part that must start from zero (the per-row "picked logit" slots, written only for rows whose
//     label falls in a tile) is cleared on `stream` at the start of every call; the partials are fully
//     rewritten for the rows and tiles a call reads back.
//   * Because the area is shared, two calls whose GPU work overlaps in time (different streams, or
//     different threads) race on it.  Issue them on one stream, or otherwise order them.
int dla_lmhead_logprob_noalloc(const void* X, const void* W, const int64_t* labels, int M, int V, int K, int64_t ldx,
                               int64_t ldw, float* logp, float* lse, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(M > 0 && V > 0 && K > 0 && K % 8 == 0, DLA_E_BADARG, "lmhead_logprob_noalloc: bad shape M=%d V=%d K=%d", M, V, K);
  DLA_REQUIRE(logp != nullptr, DLA_E_BADARG, "lmhead_logprob_noalloc: logp is required");
  const int n_tiles = (V + BN - 1) / BN;
  DLA_REQUIRE(M <= DLA_NOALLOC_MAX_ROWS, DLA_E_TOOLARGE,
              "lmhead_logprob_noalloc: M=%d rows exceed the built-in work area (max %d rows per call); split the call or "
              "use dla_lmhead_logprob with caller scratch", M, DLA_NOALLOC_MAX_ROWS);
  DLA_REQUIRE(n_tiles <= DLA_NOALLOC_MAX_TILES, DLA_E_TOOLARGE,
              "lmhead_logprob_noalloc: V=%d needs %d vocabulary tiles, the built-in work area holds %d (V <= %d)", V, n_tiles,
              DLA_NOALLOC_MAX_TILES, DLA_NOALLOC_MAX_TILES * BN);
  NoAllocLogprobArea* area = nullptr;
  int rc = noalloc_area_for_current_device(&area);
  if (rc) return rc;
  CUtensorMap tmA, tmB;
  rc = build_maps(&tmA, &tmB, X, W, M, V, K, ldx, ldw, 0, 0);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  DLA_CUDA(cudaMemsetAsync(area->picked, 0, sizeof(float) * (size_t)M, s));
  Params p{};
  p.M = M; p.N = V; p.K = K;
  p.num_m_tiles = (M + BM - 1) / BM;
  p.num_n_tiles = n_tiles;
  p.labels = labels;
  p.partials = area->partials;     // [n_tiles, M] packed with pitch M, exactly like the caller-scratch path
  p.picked = area->picked;
  rc = launch<false, false, kLogprob>(tmA, tmB, CUtensorMap{}, p, s);
  if (rc) return rc;
  logprob_combine_kernel<<<(M + 255) / 256, 256, 0, s>>>(p.partials, p.picked, labels, M, p.num_n_tiles, logp, lse);
  return check_launch("lmhead_logprob_noalloc.combine");
}

int dla_lmhead_logprob_noalloc_max_rows(void) { return DLA_NOALLOC_MAX_ROWS; }
int dla_lmhead_logprob_noalloc_max_vocab(void) { return DLA_NOALLOC_MAX_TILES * BN; }

}  // extern "C"
