// Scalar heads on the final hidden states: the reward head (pool -> dropout -> Linear(H, 1)) and the per-token
// value head.  mode 0 = last non-pad token, 1 = masked mean over tokens, 2 = one scalar per token.
//
// Forward is one block per sequence (modes 0/1) or one warp per token (mode 2); no pooled [B, H] tensor, no
// fp32 copy of the hidden states.  Backward writes the dense dhidden the backbone's backward expects and
// accumulates dw / db in fp32.  By linearity mean-pool-then-dot equals the masked mean of per-token dots, so
// mode 1 never materialises the pooled vector either.
//
// Reference: R/src/models/reward_model.py:44 (Sequential(Dropout, Linear)), :54-64 (forward + pooling);
// SURVEY K10 / worklist #15; value head = worklist #20.
#include "common.cuh"

namespace dla {
namespace heads {

// w_eff[h] = w[h] * keep[b, h] (dropout mask already scaled by 1/(1-p); keep may be null)
__device__ __forceinline__ float dot_row(const bf16* __restrict__ x, const bf16* __restrict__ w, const float* __restrict__ keep, int H,
                                         int tid, int nthreads) {
  float acc = 0.f;
  for (int c = tid * 8; c < H; c += nthreads * 8) {
    float xv[8], wv[8];
    unpack8(ld8(x + c), xv);
    unpack8(ld8(w + c), wv);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc += xv[e] * wv[e] * (keep != nullptr ? keep[c + e] : 1.f);
  }
  return acc;
}

__global__ void __launch_bounds__(256)
pooled_fwd_kernel(const bf16* __restrict__ hidden, const int64_t* __restrict__ mask, const bf16* __restrict__ w,
                  const bf16* __restrict__ bias, const float* __restrict__ keep, float* __restrict__ out, int T, int H, int mode) {
  __shared__ float red[33];
  __shared__ int s_cnt;
  const int b = blockIdx.x;
  int cnt = 0;
  for (int t = threadIdx.x; t < T; t += blockDim.x) cnt += mask[(int64_t)b * T + t] != 0;
  const float total = block_sum((float)cnt, red);
  if (threadIdx.x == 0) s_cnt = (int)(total + 0.5f);
  __syncthreads();
  const int n = s_cnt;
  const float* kp = keep != nullptr ? keep + (int64_t)b * H : nullptr;
  float acc = 0.f;
  if (mode == 0) {
    const int idx = max(n - 1, 0);
    acc = dot_row(hidden + ((int64_t)b * T + idx) * H, w, kp, H, threadIdx.x, blockDim.x);
  } else {
    for (int t = 0; t < T; ++t) {
      if (mask[(int64_t)b * T + t] == 0) continue;
      acc += dot_row(hidden + ((int64_t)b * T + t) * H, w, kp, H, threadIdx.x, blockDim.x);
    }
    acc /= (float)max(n, 1);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) out[b] = acc + (bias != nullptr ? __bfloat162float(bias[0]) : 0.f);
}

// mode 2: one warp per token row
__global__ void __launch_bounds__(256)
token_fwd_kernel(const bf16* __restrict__ hidden, const bf16* __restrict__ w, const bf16* __restrict__ bias, float* __restrict__ out,
                 int64_t M, int H) {
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= M) return;
  float acc = dot_row(hidden + row * H, w, nullptr, H, threadIdx.x & 31, 32);
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) out[row] = acc + (bias != nullptr ? __bfloat162float(bias[0]) : 0.f);
}

// dhidden[b, t, :] = coef(b, t) * w_eff;  dw[h] += sum_{b,t} coef(b, t) * keep * hidden[b, t, h];  db += sum dout.
// coef = dout[b] * (t == last)            (mode 0)
//      = dout[b] * mask[b, t] / n_b       (mode 1)
//      = dout[b, t]                       (mode 2)
// grid (T-chunks, B): each block owns kRowsPerBlock token rows of one sequence; dw partials go out with one
// fp32 atomicAdd per (block, h).
constexpr int kRowsPerBlock = 32;

__global__ void __launch_bounds__(256)
head_bwd_kernel(const bf16* __restrict__ hidden, const int64_t* __restrict__ mask, const bf16* __restrict__ w,
                const float* __restrict__ keep, const float* __restrict__ dout, bf16* __restrict__ dhidden, float* __restrict__ dw,
                float* __restrict__ db, int T, int H, int mode) {
  __shared__ float s_coef[kRowsPerBlock];
  __shared__ float red[33];
  const int b = blockIdx.y, t0 = blockIdx.x * kRowsPerBlock;
  const int rows = min(kRowsPerBlock, T - t0);
  if (mode != 2) {
    int cnt = 0;
    for (int t = threadIdx.x; t < T; t += blockDim.x) cnt += mask[(int64_t)b * T + t] != 0;
    const int n = (int)(block_sum((float)cnt, red) + 0.5f);
    if (threadIdx.x < rows) {
      const int t = t0 + threadIdx.x;
      float c;
      if (mode == 0) c = (t == max(n - 1, 0)) ? dout[b] : 0.f;
      else c = mask[(int64_t)b * T + t] != 0 ? dout[b] / (float)max(n, 1) : 0.f;
      s_coef[threadIdx.x] = c;
    }
  } else if (threadIdx.x < rows) {
    s_coef[threadIdx.x] = dout[(int64_t)b * T + t0 + threadIdx.x];
  }
  __syncthreads();
  const float* kp = keep != nullptr ? keep + (int64_t)b * H : nullptr;
  for (int c = threadIdx.x * 8; c < H; c += blockDim.x * 8) {
    float wv[8], kv[8], g[8];
    unpack8(ld8(w + c), wv);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      kv[e] = kp != nullptr ? kp[c + e] : 1.f;
      wv[e] *= kv[e];
      g[e] = 0.f;
    }
    for (int r = 0; r < rows; ++r) {
      const float cf = s_coef[r];
      const int64_t off = ((int64_t)b * T + t0 + r) * H + c;
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = cf * wv[e];
      if (dhidden != nullptr) st8(dhidden + off, pack8(o));
      if (cf != 0.f && dw != nullptr) {
        float xv[8];
        unpack8(ld8(hidden + off), xv);
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] += cf * kv[e] * xv[e];
      }
    }
    if (dw != nullptr) {
#pragma unroll
      for (int e = 0; e < 8; ++e)
        if (g[e] != 0.f) atomicAdd(dw + c + e, g[e]);
    }
  }
  if (db != nullptr) {
    // every dout element is counted once: mode 2 by the block that owns the row, modes 0/1 by the first T-chunk
    float s = 0.f;
    if (mode == 2) {
      if (threadIdx.x < rows) s = s_coef[threadIdx.x];
    } else if (blockIdx.x == 0 && threadIdx.x == 0) {
      s = dout[b];
    }
    s = block_sum(s, red);
    if (threadIdx.x == 0 && s != 0.f) atomicAdd(db, s);
  }
}

}  // namespace heads
}  // namespace dla

using namespace dla;

extern "C" {

int dla_scalar_head_fwd(const void* hidden, const int64_t* mask, const void* w, const void* bias, const float* keep, float* out, int B,
                        int T, int H, int mode, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(B > 0 && T > 0 && H > 0 && H % 8 == 0, DLA_E_BADARG, "scalar_head_fwd: bad shape B=%d T=%d H=%d (H %% 8)", B, T, H);
  DLA_REQUIRE(mode >= 0 && mode <= 2 && (mode == 2 || mask != nullptr), DLA_E_BADARG, "scalar_head_fwd: mode %d / mask", mode);
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 2) {
    const int64_t M = (int64_t)B * T;
    heads::token_fwd_kernel<<<(unsigned)((M + 7) / 8), 256, 0, st>>>((const bf16*)hidden, (const bf16*)w, (const bf16*)bias, out, M, H);
  } else {
    heads::pooled_fwd_kernel<<<B, 256, 0, st>>>((const bf16*)hidden, mask, (const bf16*)w, (const bf16*)bias, keep, out, T, H, mode);
  }
  return check_launch("scalar_head_fwd");
}

int dla_scalar_head_bwd(const void* hidden, const int64_t* mask, const void* w, const float* keep, const float* dout, void* dhidden,
                        float* dw, float* db, int B, int T, int H, int mode, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(B > 0 && T > 0 && H > 0 && H % 8 == 0, DLA_E_BADARG, "scalar_head_bwd: bad shape B=%d T=%d H=%d (H %% 8)", B, T, H);
  DLA_REQUIRE(mode >= 0 && mode <= 2 && (mode == 2 || mask != nullptr), DLA_E_BADARG, "scalar_head_bwd: mode %d / mask", mode);
  dim3 grid((T + heads::kRowsPerBlock - 1) / heads::kRowsPerBlock, B);
  heads::head_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)hidden, mask, (const bf16*)w, keep, dout, (bf16*)dhidden,
                                                                 dw, db, T, H, mode);
  return check_launch("scalar_head_bwd");
}

}  // extern "C"
