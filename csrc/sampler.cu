// On-device next-token sampling for the rollout engine: temperature, nucleus (top-p) filter and the
// multinomial draw in ONE kernel per batch, no sort and no [B, V] scratch.
//
// One block per row.  p_i = exp((l_i - max) / T).  The nucleus is {i : p_i >= tau*} where tau* is the largest
// threshold whose kept mass still reaches top_p * Z — found by bisection over tau (each step is one pass over
// the row, which stays in L2), instead of HF's sort + cumsum + scatter over V = 128256.  The draw walks the
// kept set in index order with a block-wide prefix over per-thread partial sums.
//
// Reference path replaced: transformers' TemperatureLogitsWarper + TopPLogitsWarper + torch.multinomial inside
// generate() (R/src/training/generate_teacher_data.py:73-79, R/src/eval/eval_alignment.py:71-77).
#include <math.h>

#include "common.cuh"

namespace dla {
namespace sampler {

constexpr int kThreads = 1024;

__device__ __forceinline__ float load_logit(const void* base, int fp32, int64_t i) {
  return fp32 ? reinterpret_cast<const float*>(base)[i] : __bfloat162float(reinterpret_cast<const bf16*>(base)[i]);
}

__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float t = lane < (kThreads >> 5) ? red[lane] : -INFINITY;
  return warp_max(t);
}

// splitmix64 -> uniform in [0, 1)
__device__ __forceinline__ float uniform01(uint64_t seed, uint64_t offset, uint64_t row) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * (offset * 0x100000001B3ull + row + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(z >> 40) * (1.0f / 16777216.0f);
}

__global__ void __launch_bounds__(kThreads)
sample_kernel(const void* __restrict__ logits, int fp32, int V, int64_t ld, float inv_temp, float top_p, uint64_t seed,
              uint64_t offset, const int64_t* __restrict__ offset_dev, int64_t* __restrict__ out) {
  if (offset_dev != nullptr) offset += (uint64_t)*offset_dev;       // replayed graphs advance the counter on the device
  __shared__ float red[33];
  __shared__ float s_part[kThreads];
  __shared__ int s_pick;
  const int row = blockIdx.x, tid = threadIdx.x;
  const int64_t base = (int64_t)row * ld;
  // contiguous slice per thread so the final walk is in index order
  const int per = (V + kThreads - 1) / kThreads;
  const int lo = min(tid * per, V), hi = min(lo + per, V);

  float mx = -INFINITY;
  int arg = lo;
  for (int i = lo; i < hi; ++i) {
    const float l = load_logit(logits, fp32, base + i);
    if (l > mx) { mx = l; arg = i; }
  }
  const float row_max = block_max(mx, red);
  if (inv_temp <= 0.f) {                                   // greedy: lowest index among the maxima
    __shared__ int s_arg;
    if (tid == 0) s_arg = V;
    __syncthreads();
    if (mx == row_max && hi > lo) atomicMin(&s_arg, arg);
    __syncthreads();
    if (tid == 0) out[row] = s_arg;
    return;
  }
  float z = 0.f;
  for (int i = lo; i < hi; ++i) z += __expf((load_logit(logits, fp32, base + i) - row_max) * inv_temp);
  const float Z = block_sum(z, red);
  // bisection on the probability threshold: keep(tau) = sum of p_i >= tau, monotonically non-increasing in tau
  float t_lo = 0.f, t_hi = 1.0f;                           // p_max = 1 after the max shift
  const float need = top_p * Z;
  if (top_p < 1.f) {
    for (int it = 0; it < 24; ++it) {
      const float tau = 0.5f * (t_lo + t_hi);
      float kept = 0.f;
      for (int i = lo; i < hi; ++i) {
        const float p = __expf((load_logit(logits, fp32, base + i) - row_max) * inv_temp);
        kept += p >= tau ? p : 0.f;
      }
      kept = block_sum(kept, red);
      if (kept >= need) t_lo = tau; else t_hi = tau;       // uniform branch: every thread sees the same sum
    }
  }
  const float tau = t_lo;
  float mine = 0.f;
  for (int i = lo; i < hi; ++i) {
    const float p = __expf((load_logit(logits, fp32, base + i) - row_max) * inv_temp);
    mine += p >= tau ? p : 0.f;
  }
  s_part[tid] = mine;
  const float kept_total = block_sum(mine, red);
  const float u = uniform01(seed, offset, (uint64_t)row) * kept_total;
  if (tid == 0) s_pick = -1;
  __syncthreads();
  if (tid == 0) {                                           // locate the slice that contains u
    float run = 0.f;
    int t = 0;
    for (; t < kThreads - 1; ++t) {
      if (run + s_part[t] > u) break;
      run += s_part[t];
    }
    s_pick = t;
    red[0] = run;
  }
  __syncthreads();
  if (tid == s_pick) {
    float run = red[0];
    int choice = -1, last_kept = -1;
    for (int i = lo; i < hi; ++i) {
      const float p = __expf((load_logit(logits, fp32, base + i) - row_max) * inv_temp);
      if (p >= tau) {
        last_kept = i;
        run += p;
        if (run > u) { choice = i; break; }
      }
    }
    if (choice < 0) choice = last_kept >= 0 ? last_kept : arg;   // rounding at the upper end
    out[row] = choice;
  }
}

}  // namespace sampler
}  // namespace dla

using namespace dla;

extern "C" {

size_t dla_sampler_scratch_bytes(int B, int V) {
  (void)B;
  (void)V;
  return 0;      // the bisection re-reads the logits row (L2 resident); no work area is needed
}

int dla_sample_top_p(const void* logits, int logits_fp32, int B, int V, int64_t ld, float temperature, float top_p,
                     uint64_t seed, uint64_t offset, const int64_t* offset_dev, int64_t* out, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(B > 0 && V > 0 && out != nullptr, DLA_E_BADARG, "sample_top_p: bad shape B=%d V=%d", B, V);
  DLA_REQUIRE(top_p > 0.f, DLA_E_BADARG, "sample_top_p: top_p must be > 0");
  const float inv_temp = temperature > 0.f ? 1.f / temperature : 0.f;
  sampler::sample_kernel<<<B, sampler::kThreads, 0, (cudaStream_t)stream>>>(logits, logits_fp32, V, ld, inv_temp,
                                                                            top_p > 1.f ? 1.f : top_p, seed, offset, offset_dev, out);
  return check_launch("sample_top_p");
}

}  // extern "C"
