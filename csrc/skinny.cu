// Weight-streaming GEMM for decode: y[M,N] = x[M,K] @ W[N,K]^T (+ bias) with M <= 32 rows.
//
// One decode step of the rollout engine multiplies a handful of rows (one per sequence) with every weight
// matrix of the model, so the op is a pure HBM stream of W (16 GB per token at Llama-3-8B): the persistent
// 128x256 tcgen05 tiles of gemm.cu give 16-112 CTAs for these shapes and leave most of the memory system idle.
// Here the grid is over N (16 or 32 weight rows per CTA), the eight warps of a CTA interleave over K in steps of
// 32 (so the CTA reads 512 contiguous bytes of each weight row per iteration), and the partial sums are combined
// through shared memory — deterministic, no atomics.  The math runs on mma.sync.m16n8k16 with the weight rows on
// the M side and the batch on the N side; the operands come straight from 16-byte global loads because the K
// index inside one MMA may be any permutation as long as A and B use the same one: lane (g,t) feeds actual
// k = 8t..8t+3 to the first MMA of a step and 8t+4..8t+7 to the second.  Tensor throughput is irrelevant here
// (2*M flops per weight byte pair); what is measured is GB/s against the HBM copy peak.
//
// Reference path replaced: the per-token nn.Linear calls inside transformers' generate()
// (R/src/training/train_rlhf.py:123-124, R/src/training/generate_teacher_data.py:73-79).
#include "common.cuh"

namespace dla {
namespace skinny {

constexpr int kWarps = 8;

__device__ __forceinline__ uint4 ld_stream(const bf16* p) {
  uint4 v;
  asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// NS: 16-row weight stripes per CTA (1 or 2); MB: groups of 8 batch rows (1, 2 or 4).
template <int NS, int MB>
__global__ void __launch_bounds__(kWarps * 32)
skinny_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, void* __restrict__ y, const bf16* __restrict__ bias,
              const bf16* __restrict__ residual, int M, int N, int K, int64_t ldx, int64_t ldw, int64_t ldy, int64_t ldr,
              int out_fp32) {
  constexpr int kRows = NS * 16, kCols = MB * 8;
  __shared__ float red[kWarps][kRows][kCols + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int n0 = blockIdx.x * kRows;

  const bf16* wrow[NS * 2];
#pragma unroll
  for (int i = 0; i < NS * 2; ++i) wrow[i] = w + (int64_t)min(n0 + i * 8 + g, N - 1) * ldw + t * 8;
  const bf16* xrow[MB];
  bool xok[MB];
#pragma unroll
  for (int j = 0; j < MB; ++j) {
    xok[j] = j * 8 + g < M;
    xrow[j] = x + (int64_t)min(j * 8 + g, M - 1) * ldx + t * 8;
  }

  float acc[NS][MB][4];
#pragma unroll
  for (int s = 0; s < NS; ++s)
#pragma unroll
    for (int j = 0; j < MB; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[s][j][e] = 0.f;

  const int ksteps = K >> 5;
  auto load_w = [&](int ks, uint4 (&a)[NS * 2]) {
#pragma unroll
    for (int i = 0; i < NS * 2; ++i) a[i] = ld_stream(wrow[i] + (ks << 5));
  };
  auto load_x = [&](int ks, uint4 (&b)[MB]) {
#pragma unroll
    for (int j = 0; j < MB; ++j) {
      b[j] = *reinterpret_cast<const uint4*>(xrow[j] + (ks << 5));
      if (!xok[j]) b[j] = make_uint4(0, 0, 0, 0);
    }
  };
  auto mma_step = [&](const uint4 (&a)[NS * 2], const uint4 (&b)[MB]) {
#pragma unroll
    for (int s = 0; s < NS; ++s)
#pragma unroll
      for (int j = 0; j < MB; ++j) {
        mma16816(acc[s][j], a[2 * s].x, a[2 * s + 1].x, a[2 * s].y, a[2 * s + 1].y, b[j].x, b[j].y);
        mma16816(acc[s][j], a[2 * s].z, a[2 * s + 1].z, a[2 * s].w, a[2 * s + 1].w, b[j].z, b[j].w);
      }
  };
  // U steps of loads are issued back to back before their MMAs: 16 B x (2 NS + MB) x U per lane in flight
  constexpr int U = MB == 4 ? 2 : 4;
  int ks = warp;
  {
    // Programmatic dependent launch: this grid may start while its predecessor (the kernel that produces x)
    // is still draining.  The weights do not depend on it, so the first batch of the weight stream is
    // requested before griddepcontrol.wait; x, the residual and y are only touched after it.
    pdl_launch_dependents();
    uint4 a[U][NS * 2], b[U][MB];
    const bool full = ks + (U - 1) * kWarps < ksteps;
    if (full) {
#pragma unroll
      for (int u = 0; u < U; ++u) load_w(ks + u * kWarps, a[u]);
    }
    pdl_wait();
    if (full) {
#pragma unroll
      for (int u = 0; u < U; ++u) load_x(ks + u * kWarps, b[u]);
#pragma unroll
      for (int u = 0; u < U; ++u) mma_step(a[u], b[u]);
      ks += U * kWarps;
    }
  }
  for (; ks + (U - 1) * kWarps < ksteps; ks += U * kWarps) {
    uint4 a[U][NS * 2], b[U][MB];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      load_w(ks + u * kWarps, a[u]);
      load_x(ks + u * kWarps, b[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) mma_step(a[u], b[u]);
  }
  for (; ks < ksteps; ks += kWarps) {
    uint4 a[NS * 2], b[MB];
    load_w(ks, a);
    load_x(ks, b);
    mma_step(a, b);
  }

  // accumulator (row g / g+8 of the stripe, batch 2t / 2t+1 of the group) -> red[warp][row][batch]
#pragma unroll
  for (int s = 0; s < NS; ++s)
#pragma unroll
    for (int j = 0; j < MB; ++j) {
      red[warp][s * 16 + g][j * 8 + 2 * t] = acc[s][j][0];
      red[warp][s * 16 + g][j * 8 + 2 * t + 1] = acc[s][j][1];
      red[warp][s * 16 + g + 8][j * 8 + 2 * t] = acc[s][j][2];
      red[warp][s * 16 + g + 8][j * 8 + 2 * t + 1] = acc[s][j][3];
    }
  __syncthreads();
  for (int i = threadIdx.x; i < kRows * kCols; i += kWarps * 32) {
    const int r = i % kRows, m = i / kRows;        // weight row fastest: stores of one batch row are contiguous
    const int n = n0 + r;
    if (m >= M || n >= N) continue;
    float v = 0.f;
#pragma unroll
    for (int wp = 0; wp < kWarps; ++wp) v += red[wp][r][m];
    if (bias != nullptr) v += __bfloat162float(bias[n]);
    if (residual != nullptr) v += __bfloat162float(residual[(int64_t)m * ldr + n]);
    if (out_fp32) reinterpret_cast<float*>(y)[(int64_t)m * ldy + n] = v;
    else reinterpret_cast<bf16*>(y)[(int64_t)m * ldy + n] = __float2bfloat16(v);
  }
}

template <int NS, int MB>
cudaError_t launch(const bf16* x, const bf16* w, void* y, const bf16* bias, const bf16* residual, int M, int N, int K,
                   int64_t ldx, int64_t ldw, int64_t ldy, int64_t ldr, int out_fp32, int early_start, cudaStream_t st) {
  const int rows = NS * 16;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((N + rows - 1) / rows);
  cfg.blockDim = dim3(kWarps * 32);
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = early_start ? 1 : 0;      // without the attribute the grid starts after its predecessor, as usual
  return cudaLaunchKernelEx(&cfg, skinny_kernel<NS, MB>, x, w, y, bias, residual, M, N, K, ldx, ldw, ldy, ldr, out_fp32);
}

}  // namespace skinny
}  // namespace dla

using namespace dla;

extern "C" {

int dla_gemm_skinny_max_rows(void) { return 32; }

int dla_gemm_skinny_bf16(const void* x, const void* w, void* y, const void* bias, const void* residual, int M, int N, int K,
                         int64_t ldx, int64_t ldw, int64_t ldy, int64_t ldr, int out_fp32, int weights_static,
                         void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(M > 0 && M <= 32 && N > 0 && K > 0, DLA_E_BADARG, "gemm_skinny: M=%d must be in 1..32 (N=%d K=%d)", M, N, K);
  DLA_REQUIRE(K % 32 == 0 && ldx % 8 == 0 && ldw % 8 == 0, DLA_E_BADARG,
              "gemm_skinny: K=%d must be a multiple of 32 and the row pitches (%lld, %lld) multiples of 8", K, (long long)ldx,
              (long long)ldw);
  DLA_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w)) % 16 == 0, DLA_E_BADARG,
              "gemm_skinny: operands must be 16-byte aligned");
  const bf16* xp = static_cast<const bf16*>(x);
  const bf16* wp = static_cast<const bf16*>(w);
  const bf16* bp = static_cast<const bf16*>(bias);
  const bf16* rp = static_cast<const bf16*>(residual);
  cudaError_t err = cudaSuccess;
  cudaStream_t st = (cudaStream_t)stream;
  const bool wide = (N + 31) / 32 >= 2 * 148;      // enough 32-row CTAs for two per SM; else use 16-row stripes
#define DLA_SKINNY(MBv)                                                                                     \
  do {                                                                                                      \
    if (wide) err = skinny::launch<2, MBv>(xp, wp, y, bp, rp, M, N, K, ldx, ldw, ldy, ldr, out_fp32, weights_static, st); \
    else err = skinny::launch<1, MBv>(xp, wp, y, bp, rp, M, N, K, ldx, ldw, ldy, ldr, out_fp32, weights_static, st);      \
  } while (0)
  if (M <= 8) DLA_SKINNY(1);
  else if (M <= 16) DLA_SKINNY(2);
  else DLA_SKINNY(4);
#undef DLA_SKINNY
  DLA_CUDA(err);
  return check_launch("gemm_skinny");
}

}  // extern "C"
