// Fused AdamW over flat shards, global grad-norm pieces, dtype casts.
// Replaces torch.optim.AdamW foreach kernels + clip_grad_norm_.
#include "common.cuh"

namespace dla {

constexpr int kSumsqBlocks = kNumSMs * 8;
constexpr int kSumsqThreads = 256;

// bf16 Adam moments (70B sizing: 4 instead of 8 bytes of moments per parameter).  Round-to-nearest would freeze
// v once (1-beta2) g^2 drops under half a bf16 ulp of v, so the store rounds stochastically: the low 16 bits of
// the fp32 value plus a 16-bit hash of (element, step) carry into the kept half.
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}
__device__ __forceinline__ bf16 to_bf16_stochastic(float x, uint32_t rnd16) {
  const uint32_t bits = __float_as_uint(x) + (rnd16 & 0xFFFFu);
  __nv_bfloat16_raw r;
  r.x = (unsigned short)(bits >> 16);
  return bf16(r);
}
template <bool kBF16>
__device__ __forceinline__ void load_moment8(const void* base, int64_t o, float* out) {
  if (kBF16) {
    unpack8(ld8(reinterpret_cast<const bf16*>(base) + o), out);
  } else {
    const float* f = reinterpret_cast<const float*>(base) + o;
    *reinterpret_cast<float4*>(out) = *reinterpret_cast<const float4*>(f);
    *reinterpret_cast<float4*>(out + 4) = *reinterpret_cast<const float4*>(f + 4);
  }
}
template <bool kBF16>
__device__ __forceinline__ void store_moment8(void* base, int64_t o, const float* in, uint32_t seed) {
  if (kBF16) {
    bf16x8 pk;
    bf16* e = reinterpret_cast<bf16*>(&pk);
    const uint32_t h = mix32(seed ^ (uint32_t)o ^ (uint32_t)(o >> 32) * 0x9E3779B9u);
#pragma unroll
    for (int j = 0; j < 8; ++j) e[j] = to_bf16_stochastic(in[j], mix32(h + j));
    st8(reinterpret_cast<bf16*>(base) + o, pk);
  } else {
    float* f = reinterpret_cast<float*>(base) + o;
    *reinterpret_cast<float4*>(f) = *reinterpret_cast<const float4*>(in);
    *reinterpret_cast<float4*>(f + 4) = *reinterpret_cast<const float4*>(in + 4);
  }
}

// 16 bytes stored through the NVSwitch multicast address: lands in every rank's copy of the buffer
__device__ __forceinline__ void multimem_st8(bf16* mc, const bf16x8& v) {
  const uint4 u = *reinterpret_cast<const uint4*>(&v);
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(__uint_as_float(u.x)),
               "f"(__uint_as_float(u.y)), "f"(__uint_as_float(u.z)), "f"(__uint_as_float(u.w)) : "memory");
}

// One thread = 8 consecutive elements: 16 B of grad, 3 x 32 B of fp32 state in, 3 x 32 B + 16 B out
// (2 x 16 B each way for the moments when they are bf16).  kMC: `param` is a MULTICAST address — the updated bf16
// shard is written into every rank's resident parameters by the same store (the post-step all-gather disappears).
// A clip coefficient of exactly 0 means "non-finite gradient norm": the whole update, weight decay included, is skipped
// (g * 0 would still poison the moments with NaN).
template <bool kGradF32, bool kMomBF16, bool kMC = false>
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ master, void* __restrict__ m, void* __restrict__ v,
             const void* __restrict__ grad, bf16* __restrict__ param, int64_t n, float lr, float beta1,
             float beta2, float eps, float wd, float inv_bc1, float inv_sqrt_bc2,
             const float* __restrict__ grad_scale_dev, float prescale) {
  const float coef = grad_scale_dev != nullptr ? *grad_scale_dev : 1.f;
  if (coef == 0.f) return;
  const float gs = prescale * coef;
  const uint32_t step_seed = __float_as_uint(inv_sqrt_bc2) * 0x9E3779B9u;   // changes every optimizer step
  const int64_t nvec = n >> 3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t o = i * 8;
    float g[8], p[8], mm[8], vv[8];
    if (kGradF32) {
      *reinterpret_cast<float4*>(g) = *reinterpret_cast<const float4*>((const float*)grad + o);
      *reinterpret_cast<float4*>(g + 4) = *reinterpret_cast<const float4*>((const float*)grad + o + 4);
    } else {
      unpack8(ld8_stream((const bf16*)grad + o), g);
    }
    *reinterpret_cast<float4*>(p) = *reinterpret_cast<const float4*>(master + o);
    *reinterpret_cast<float4*>(p + 4) = *reinterpret_cast<const float4*>(master + o + 4);
    load_moment8<kMomBF16>(m, o, mm);
    load_moment8<kMomBF16>(v, o, vv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = g[j] * gs;
      mm[j] = beta1 * mm[j] + (1.f - beta1) * gj;
      vv[j] = beta2 * vv[j] + (1.f - beta2) * gj * gj;
      const float denom = sqrtf(vv[j]) * inv_sqrt_bc2 + eps;
      p[j] = p[j] * (1.f - lr * wd) - lr * inv_bc1 * (mm[j] / denom);
    }
    *reinterpret_cast<float4*>(master + o) = *reinterpret_cast<float4*>(p);
    *reinterpret_cast<float4*>(master + o + 4) = *reinterpret_cast<float4*>(p + 4);
    store_moment8<kMomBF16>(m, o, mm, step_seed);
    store_moment8<kMomBF16>(v, o, vv, step_seed ^ 0xA5A5A5A5u);
    if (kMC) multimem_st8(param + o, pack8(p));
    else if (param != nullptr) st8(param + o, pack8(p));
  }
  // tail (n % 8) handled by the first threads of block 0 (never taken with kMC: shards are multiples of 8)
  if (!kMC && blockIdx.x == 0 && threadIdx.x < (n & 7)) {
    const int64_t o = (nvec << 3) + threadIdx.x;
    float gj = (kGradF32 ? ((const float*)grad)[o] : __bfloat162float(((const bf16*)grad)[o])) * gs;
    const float m0 = kMomBF16 ? __bfloat162float(((const bf16*)m)[o]) : ((const float*)m)[o];
    const float v0 = kMomBF16 ? __bfloat162float(((const bf16*)v)[o]) : ((const float*)v)[o];
    float mj = beta1 * m0 + (1.f - beta1) * gj;
    float vj = beta2 * v0 + (1.f - beta2) * gj * gj;
    float pj = master[o] * (1.f - lr * wd) - lr * inv_bc1 * (mj / (sqrtf(vj) * inv_sqrt_bc2 + eps));
    if (kMomBF16) {
      ((bf16*)m)[o] = to_bf16_stochastic(mj, mix32(step_seed + (uint32_t)o));
      ((bf16*)v)[o] = to_bf16_stochastic(vj, mix32(step_seed + (uint32_t)o + 0x51ED27u));
    } else {
      ((float*)m)[o] = mj;
      ((float*)v)[o] = vj;
    }
    master[o] = pj;
    if (param != nullptr) param[o] = __float2bfloat16(pj);
  }
}

template <bool kF32>
__global__ void __launch_bounds__(kSumsqThreads)
sumsq_partial_kernel(const void* __restrict__ x, int64_t n, float* __restrict__ partials) {
  __shared__ float red[33];
  float acc = 0.f;
  if (kF32) {
    const int64_t nvec = n >> 2;
    const float4* p = reinterpret_cast<const float4*>(x);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
      const float4 t = p[i];
      acc += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
      const float t = ((const float*)x)[(nvec << 2) + threadIdx.x];
      acc += t * t;
    }
  } else {
    const int64_t nvec = n >> 3;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
      float f[8];
      unpack8(ld8_stream((const bf16*)x + i * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc += f[j] * f[j];
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 7)) {
      const float t = __bfloat162float(((const bf16*)x)[(nvec << 3) + threadIdx.x]);
      acc += t * t;
    }
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
}

// deterministic final reduce of the per-block partials (single block)
__global__ void __launch_bounds__(1024)
sumsq_final_kernel(const float* __restrict__ partials, int nblocks, float* __restrict__ out, int accumulate) {
  __shared__ float red[33];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nblocks; i += blockDim.x) acc += partials[i];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) out[0] = accumulate ? out[0] + acc : acc;
}

__global__ void clip_coef_kernel(const float* __restrict__ sumsq, float max_norm, float* __restrict__ coef,
                                 float* __restrict__ norm_out) {
  const float norm = sqrtf(sumsq[0]);
  if (norm_out != nullptr) norm_out[0] = norm;
  float c = 1.f;
  if (max_norm > 0.f) c = fminf(1.f, max_norm / (norm + 1e-6f));
  if (!(norm == norm) || isinf(norm)) c = 0.f;  // non-finite gradients: skip the update
  coef[0] = c;
}

template <typename S, typename D>
__global__ void cast_kernel(const S* __restrict__ src, D* __restrict__ dst, int64_t n, float scale) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v;
    if constexpr (sizeof(S) == 4) v = (float)src[i]; else v = __bfloat162float(src[i]);
    v *= scale;
    if constexpr (sizeof(D) == 4) dst[i] = v; else dst[i] = __float2bfloat16(v);
  }
}

}  // namespace dla

using namespace dla;

extern "C" {

int dla_sumsq_blocks(void) { return kSumsqBlocks; }

int dla_adamw(float* master, void* m, void* v, const void* grad, void* param, int64_t n, float lr, float beta1,
              float beta2, float eps, float weight_decay, float bias_c1, float bias_c2,
              const float* grad_scale_dev, float grad_prescale, int flags, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(bias_c1 > 0.f && bias_c2 > 0.f, DLA_E_BADARG, "adamw: bias corrections must be > 0");
  if (n == 0) return 0;
  int64_t blocks = ((n >> 3) + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > (int64_t)kNumSMs * 16) blocks = (int64_t)kNumSMs * 16;
  const float inv_bc1 = 1.f / bias_c1, inv_sqrt_bc2 = 1.f / sqrtf(bias_c2);
  const bool g32 = (flags & DLA_ADAMW_GRAD_FP32) != 0, mom16 = (flags & DLA_ADAMW_MOMENTS_BF16) != 0;
  cudaStream_t st = (cudaStream_t)stream;
#define DLA_ADAMW_LAUNCH(G, M)                                                                                              \
  adamw_kernel<G, M><<<(int)blocks, 256, 0, st>>>(master, m, v, grad, (bf16*)param, n, lr, beta1, beta2, eps, weight_decay, \
                                                  inv_bc1, inv_sqrt_bc2, grad_scale_dev, grad_prescale)
  if (g32 && mom16) DLA_ADAMW_LAUNCH(true, true);
  else if (g32) DLA_ADAMW_LAUNCH(true, false);
  else if (mom16) DLA_ADAMW_LAUNCH(false, true);
  else DLA_ADAMW_LAUNCH(false, false);
#undef DLA_ADAMW_LAUNCH
  return check_launch("adamw");
}

int dla_adamw_mc(float* master, void* m, void* v, const void* grad, void* mc_param, int64_t n, float lr, float beta1,
                 float beta2, float eps, float weight_decay, float bias_c1, float bias_c2, const float* grad_scale_dev,
                 float grad_prescale, int flags, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(bias_c1 > 0.f && bias_c2 > 0.f, DLA_E_BADARG, "adamw_mc: bias corrections must be > 0");
  DLA_REQUIRE(mc_param != nullptr && n % 8 == 0 && (reinterpret_cast<uintptr_t>(mc_param) & 15) == 0, DLA_E_BADARG,
              "adamw_mc: needs a 16-byte aligned multicast address and n %% 8 == 0");
  if (n == 0) return 0;
  int64_t blocks = ((n >> 3) + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 16) blocks = (int64_t)kNumSMs * 16;
  const float inv_bc1 = 1.f / bias_c1, inv_sqrt_bc2 = 1.f / sqrtf(bias_c2);
  const bool g32 = (flags & DLA_ADAMW_GRAD_FP32) != 0, mom16 = (flags & DLA_ADAMW_MOMENTS_BF16) != 0;
  cudaStream_t st = (cudaStream_t)stream;
#define DLA_ADAMW_MC_LAUNCH(G, M)                                                                                     \
  adamw_kernel<G, M, true><<<(int)blocks, 256, 0, st>>>(master, m, v, grad, (bf16*)mc_param, n, lr, beta1, beta2, eps, \
                                                        weight_decay, inv_bc1, inv_sqrt_bc2, grad_scale_dev, grad_prescale)
  if (g32 && mom16) DLA_ADAMW_MC_LAUNCH(true, true);
  else if (g32) DLA_ADAMW_MC_LAUNCH(true, false);
  else if (mom16) DLA_ADAMW_MC_LAUNCH(false, true);
  else DLA_ADAMW_MC_LAUNCH(false, false);
#undef DLA_ADAMW_MC_LAUNCH
  return check_launch("adamw_mc");
}

int dla_sumsq(const void* x, int64_t n, int is_fp32, float* scratch, float* out, int accumulate, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(scratch != nullptr && out != nullptr, DLA_E_BADARG, "sumsq: scratch and out are required");
  const int per = is_fp32 ? 4 : 8;
  int64_t blocks = ((n / per) + kSumsqThreads - 1) / kSumsqThreads;
  if (blocks < 1) blocks = 1;
  if (blocks > kSumsqBlocks) blocks = kSumsqBlocks;
  if (is_fp32)
    sumsq_partial_kernel<true><<<(int)blocks, kSumsqThreads, 0, (cudaStream_t)stream>>>(x, n, scratch);
  else
    sumsq_partial_kernel<false><<<(int)blocks, kSumsqThreads, 0, (cudaStream_t)stream>>>(x, n, scratch);
  int rc = check_launch("sumsq.partial");
  if (rc) return rc;
  sumsq_final_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(scratch, (int)blocks, out, accumulate);
  return check_launch("sumsq.final");
}

int dla_clip_coef(const float* sumsq, float max_norm, float* coef_out, float* norm_out, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  clip_coef_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(sumsq, max_norm, coef_out, norm_out);
  return check_launch("clip_coef");
}

int dla_cast(const void* src, void* dst, int64_t n, int src_is_fp32, int dst_is_fp32, float scale, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  if (n == 0) return 0;
  int64_t blocks = (n + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 16) blocks = (int64_t)kNumSMs * 16;
  cudaStream_t s = (cudaStream_t)stream;
  if (src_is_fp32 && dst_is_fp32) cast_kernel<float, float><<<(int)blocks, 256, 0, s>>>((const float*)src, (float*)dst, n, scale);
  else if (src_is_fp32) cast_kernel<float, bf16><<<(int)blocks, 256, 0, s>>>((const float*)src, (bf16*)dst, n, scale);
  else if (dst_is_fp32) cast_kernel<bf16, float><<<(int)blocks, 256, 0, s>>>((const bf16*)src, (float*)dst, n, scale);
  else cast_kernel<bf16, bf16><<<(int)blocks, 256, 0, s>>>((const bf16*)src, (bf16*)dst, n, scale);
  return check_launch("cast");
}

}  // extern "C"
