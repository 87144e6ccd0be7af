// Distillation KL on logits rows: KL(mean_k softmax(t_k / T) || softmax(s / T)) summed over the vocabulary, and its
// gradient with respect to the student logits, without materialising any [rows, V] probability tensor.
//
// One block per row, two passes over the (1 + K) logit rows:
//   pass 1  online max / sum-exp of the student row and of every teacher row (one (m, z) pair each);
//   pass 2  p_t = mean_k exp(t_k/T - m_k) / z_k,  log p_s = s/T - m_s - log z_s,
//           kl += p_t * (log p_t - log p_s)                      (0 where p_t == 0, like torch.xlogy)
//           dl  = (exp(log p_s) - p_t) * coef[row] / T           -> bf16, feeds the lm_head dgrad / wgrad GEMMs
// The PyTorch formulation (log_softmax, K x softmax, stack, mean, kl_div, sum, mask) makes ~6 + 3K passes over
// [rows, V] fp32 tensors and keeps K + 2 of them alive.
//
// Reference: R/src/training/train_distill.py:132-144 (worklist #17).
#include <math.h>

#include "common.cuh"

namespace dla {
namespace distill {

constexpr int kThreads = 1024;
constexpr int kMaxTeachers = 8;

struct Teachers {
  const float* p[kMaxTeachers];
  int64_t ld[kMaxTeachers];
};

struct MZ {
  float m, z;
};

__device__ __forceinline__ void online(MZ& a, float x) {
  if (x > a.m) {
    a.z = a.z * __expf(a.m - x) + 1.f;
    a.m = x;
  } else {
    a.z += __expf(x - a.m);
  }
}
__device__ __forceinline__ MZ merge(MZ a, MZ b) {
  const float m = fmaxf(a.m, b.m);
  MZ r;
  r.m = m;
  r.z = (a.m == -INFINITY ? 0.f : a.z * __expf(a.m - m)) + (b.m == -INFINITY ? 0.f : b.z * __expf(b.m - m));
  return r;
}

__device__ __forceinline__ MZ block_merge(MZ v, MZ* sh) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MZ other;
    other.m = __shfl_xor_sync(0xffffffffu, v.m, o);
    other.z = __shfl_xor_sync(0xffffffffu, v.z, o);
    v = merge(v, other);
  }
  __syncthreads();
  if (lane == 0) sh[wid] = v;
  __syncthreads();
  MZ t = lane < (kThreads >> 5) ? sh[lane] : MZ{-INFINITY, 0.f};
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MZ other;
    other.m = __shfl_xor_sync(0xffffffffu, t.m, o);
    other.z = __shfl_xor_sync(0xffffffffu, t.z, o);
    t = merge(t, other);
  }
  return t;
}

template <int K>
__global__ void __launch_bounds__(kThreads)
distill_kl_kernel(const float* __restrict__ student, int64_t ld_s, Teachers teachers, const float* __restrict__ coef, int V,
                  float inv_temp, float* __restrict__ kl_out, bf16* __restrict__ dl, int64_t ld_d) {
  __shared__ MZ sh[32];
  __shared__ float red[33];
  const int row = blockIdx.x;
  const float* s = student + (int64_t)row * ld_s;
  const float* t[K];
#pragma unroll
  for (int k = 0; k < K; ++k) t[k] = teachers.p[k] + (int64_t)row * teachers.ld[k];

  MZ ms = {-INFINITY, 0.f}, mt[K];
#pragma unroll
  for (int k = 0; k < K; ++k) mt[k] = MZ{-INFINITY, 0.f};
  for (int i = threadIdx.x; i < V; i += kThreads) {
    online(ms, s[i] * inv_temp);
#pragma unroll
    for (int k = 0; k < K; ++k) online(mt[k], t[k][i] * inv_temp);
  }
  ms = block_merge(ms, sh);
  float tscale[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    mt[k] = block_merge(mt[k], sh);
    tscale[k] = 1.f / (mt[k].z * (float)K);
  }
  const float log_zs = __logf(ms.z);
  const float cf = coef != nullptr ? coef[row] : 1.f;

  float kl = 0.f;
  for (int i = threadIdx.x; i < V; i += kThreads) {
    const float lps = s[i] * inv_temp - ms.m - log_zs;
    float pt = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) pt += __expf(t[k][i] * inv_temp - mt[k].m) * tscale[k];
    if (kl_out != nullptr && pt > 0.f) kl += pt * (__logf(pt) - lps);
    if (dl != nullptr) dl[(int64_t)row * ld_d + i] = __float2bfloat16((__expf(lps) - pt) * cf * inv_temp);
  }
  if (kl_out != nullptr) {
    kl = block_sum(kl, red);
    if (threadIdx.x == 0) kl_out[row] = kl;
  }
}

}  // namespace distill
}  // namespace dla

using namespace dla;

extern "C" {

int dla_distill_max_teachers(void) { return distill::kMaxTeachers; }

int dla_distill_kl(const float* student, int64_t ld_s, const float* const* teacher_ptrs, const int64_t* teacher_ld, int n_teachers,
                   const float* coef, int rows, int V, float temperature, float* kl_out, void* dl, int64_t ld_d, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(rows > 0 && V > 0 && temperature > 0.f, DLA_E_BADARG, "distill_kl: bad shape rows=%d V=%d T=%g", rows, V, temperature);
  DLA_REQUIRE(n_teachers >= 1 && n_teachers <= distill::kMaxTeachers, DLA_E_UNSUPP, "distill_kl: %d teachers (1..%d supported)",
              n_teachers, distill::kMaxTeachers);
  DLA_REQUIRE(kl_out != nullptr || dl != nullptr, DLA_E_BADARG, "distill_kl: nothing to compute");
  distill::Teachers ts = {};
  for (int k = 0; k < n_teachers; ++k) {
    ts.p[k] = teacher_ptrs[k];          // host array of device pointers
    ts.ld[k] = teacher_ld[k];
  }
  cudaStream_t st = (cudaStream_t)stream;
  const float it = 1.f / temperature;
#define DLA_KL(Kv)                                                                                                                 \
  case Kv:                                                                                                                         \
    distill::distill_kl_kernel<Kv><<<rows, distill::kThreads, 0, st>>>(student, ld_s, ts, coef, V, it, kl_out, (bf16*)dl, ld_d); \
    break
  switch (n_teachers) {
    DLA_KL(1);
    DLA_KL(2);
    DLA_KL(3);
    DLA_KL(4);
    DLA_KL(5);
    DLA_KL(6);
    DLA_KL(7);
    DLA_KL(8);
  }
#undef DLA_KL
  return check_launch("distill_kl");
}

}  // extern "C"
