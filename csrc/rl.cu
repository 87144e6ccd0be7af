// RLHF objectives as single launches: sequence-level REINFORCE (the reference's "PPO"), token-level clipped PPO
// with a clipped value loss, and GAE with advantage whitening.  The tensors are [B] or [B, T] fp32 (at most a few
// 10^4 elements), so each op is one block: value, logged statistics and gradients come out of the same pass
// instead of the ~25 ATen launches the PyTorch formulation costs.
//
// Reference: R/src/training/train_rlhf.py:149-153 (kl / rewards / advantages / loss); clipped PPO + GAE are the
// `ppo.*` keys the reference only declares (R/config/rlhf_config.yaml:14,17; SURVEY worklist #20).
#include <math.h>

#include "common.cuh"

namespace dla {
namespace rl {

constexpr int kThreads = 1024;

// out[0] = loss = -mean((r - base) * pi), out[1] = mean(kl), out[2] = mean(r); r = score - c * (pi - ref).
// base = *baseline when given (global cross-rank baseline) else the batch mean of r.  dpol = -(r - base) / B.
__global__ void __launch_bounds__(kThreads)
reinforce_kernel(const float* __restrict__ pol, const float* __restrict__ ref, const float* __restrict__ score, int B,
                 float kl_coef, const float* __restrict__ baseline, float* __restrict__ out, float* __restrict__ kl_out,
                 float* __restrict__ rew_out, float* __restrict__ dpol) {
  __shared__ float red[33];
  float sr = 0.f, skl = 0.f;
  for (int i = threadIdx.x; i < B; i += kThreads) {
    const float kl = pol[i] - ref[i];
    const float r = score[i] - kl_coef * kl;
    if (kl_out != nullptr) kl_out[i] = kl;
    if (rew_out != nullptr) rew_out[i] = r;
    sr += r;
    skl += kl;
  }
  sr = block_sum(sr, red);
  skl = block_sum(skl, red);
  const float inv = 1.f / (float)B;
  const float base = baseline != nullptr ? *baseline : sr * inv;
  float loss = 0.f;
  for (int i = threadIdx.x; i < B; i += kThreads) {
    const float a = (score[i] - kl_coef * (pol[i] - ref[i])) - base;
    loss += -a * pol[i];
    if (dpol != nullptr) dpol[i] = -a * inv;
  }
  loss = block_sum(loss, red);
  if (threadIdx.x == 0) {
    out[0] = loss * inv;
    out[1] = skl * inv;
    out[2] = sr * inv;
  }
}

// Token-level clipped PPO.  All inputs [N] fp32, mask 0/1.  values/old_values/returns may be null (policy term only).
// out = {loss, pg_loss, vf_loss, approx_kl, clipfrac}; dlogp / dvalues are d loss / d input.
__global__ void __launch_bounds__(kThreads)
ppo_kernel(const float* __restrict__ new_lp, const float* __restrict__ old_lp, const float* __restrict__ adv,
           const float* __restrict__ values, const float* __restrict__ old_values, const float* __restrict__ returns,
           const float* __restrict__ mask, int N, float clip, float vclip, float vf_coef, float* __restrict__ out,
           float* __restrict__ dlogp, float* __restrict__ dvalues) {
  __shared__ float red[33];
  float cnt = 0.f;
  for (int i = threadIdx.x; i < N; i += kThreads) cnt += mask[i];
  cnt = block_sum(cnt, red);
  const float inv = 1.f / fmaxf(cnt, 1.f);
  float pg = 0.f, vf = 0.f, akl = 0.f, cf = 0.f;
  for (int i = threadIdx.x; i < N; i += kThreads) {
    const float m = mask[i];
    const float d = new_lp[i] - old_lp[i];
    const float ratio = __expf(d);
    const float rc = fminf(fmaxf(ratio, 1.f - clip), 1.f + clip);
    const float a = adv[i];
    const float l1 = -a * ratio, l2 = -a * rc;
    pg += fmaxf(l1, l2) * m;
    akl += 0.5f * d * d * m;
    const bool outside = ratio < 1.f - clip || ratio > 1.f + clip;
    cf += (fabsf(ratio - 1.f) > clip ? 1.f : 0.f) * m;
    // inside the clip range both branches coincide; outside, the clipped branch is constant in logp
    const float g = (!outside || l1 > l2) ? -a * ratio : 0.f;
    if (dlogp != nullptr) dlogp[i] = g * m * inv;
    if (values != nullptr) {
      const float v = values[i], ov = old_values[i], R = returns[i];
      const float dv = v - ov;
      const float vc = ov + fminf(fmaxf(dv, -vclip), vclip);
      const float e1 = (v - R) * (v - R), e2 = (vc - R) * (vc - R);
      vf += 0.5f * fmaxf(e1, e2) * m;
      const bool vin = fabsf(dv) <= vclip;
      const float gv = (vin || e1 > e2) ? (v - R) : 0.f;
      if (dvalues != nullptr) dvalues[i] = vf_coef * gv * m * inv;
    } else if (dvalues != nullptr) {
      dvalues[i] = 0.f;
    }
  }
  pg = block_sum(pg, red);
  vf = block_sum(vf, red);
  akl = block_sum(akl, red);
  cf = block_sum(cf, red);
  if (threadIdx.x == 0) {
    out[0] = (pg + vf_coef * vf) * inv;
    out[1] = pg * inv;
    out[2] = vf * inv;
    out[3] = akl * inv;
    out[4] = cf * inv;
  }
}

// GAE over [B, T]: delta_t = r_t + gamma * V_{t+1} * m_{t+1} - V_t; A_t = delta_t + gamma * lam * m_{t+1} * A_{t+1};
// A *= m; returns = A + V.  whiten: A <- (A - mean_m(A)) / max(std_m(A), 1e-6) * m (returns keep the raw A).
__global__ void __launch_bounds__(kThreads)
gae_kernel(const float* __restrict__ rewards, const float* __restrict__ values, const float* __restrict__ mask,
           float* __restrict__ adv, float* __restrict__ ret, int B, int T, float gamma, float lam, int whiten) {
  __shared__ float red[33];
  for (int b = threadIdx.x; b < B; b += kThreads) {
    float last = 0.f, nv = 0.f, nm = 0.f;
    for (int t = T - 1; t >= 0; --t) {
      const int64_t i = (int64_t)b * T + t;
      const float v = values[i], m = mask[i];
      const float delta = rewards[i] + gamma * nv * nm - v;
      last = delta + gamma * lam * nm * last;
      adv[i] = last * m;
      ret[i] = last * m + v;
      nv = v;
      nm = m;
    }
  }
  if (!whiten) return;
  __syncthreads();
  const int N = B * T;
  float s = 0.f, c = 0.f;
  for (int i = threadIdx.x; i < N; i += kThreads) {
    s += adv[i] * mask[i];
    c += mask[i];
  }
  s = block_sum(s, red);
  c = block_sum(c, red);
  const float inv = 1.f / fmaxf(c, 1.f), mean = s * inv;
  float var = 0.f;
  for (int i = threadIdx.x; i < N; i += kThreads) {
    const float d = adv[i] - mean;
    var += d * d * mask[i];
  }
  var = block_sum(var, red);
  const float rstd = 1.f / fmaxf(sqrtf(var * inv), 1e-6f);
  for (int i = threadIdx.x; i < N; i += kThreads) adv[i] = (adv[i] - mean) * rstd * mask[i];
}

}  // namespace rl
}  // namespace dla

using namespace dla;

extern "C" {

int dla_reinforce_loss(const float* pol, const float* ref, const float* score, int B, float kl_coef, const float* baseline,
                       float* out, float* kl_out, float* rew_out, float* dpol, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(B > 0 && out != nullptr, DLA_E_BADARG, "reinforce_loss: bad batch %d", B);
  rl::reinforce_kernel<<<1, rl::kThreads, 0, (cudaStream_t)stream>>>(pol, ref, score, B, kl_coef, baseline, out, kl_out, rew_out, dpol);
  return check_launch("reinforce_loss");
}

int dla_ppo_loss(const float* new_logp, const float* old_logp, const float* adv, const float* values, const float* old_values,
                 const float* returns, const float* mask, int N, float clip, float vclip, float vf_coef, float* out, float* dlogp,
                 float* dvalues, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(N > 0 && out != nullptr, DLA_E_BADARG, "ppo_loss: bad size %d", N);
  DLA_REQUIRE(values == nullptr || (old_values != nullptr && returns != nullptr), DLA_E_BADARG,
              "ppo_loss: values need old_values and returns");
  rl::ppo_kernel<<<1, rl::kThreads, 0, (cudaStream_t)stream>>>(new_logp, old_logp, adv, values, old_values, returns, mask, N, clip,
                                                              vclip, vf_coef, out, dlogp, dvalues);
  return check_launch("ppo_loss");
}

int dla_gae(const float* rewards, const float* values, const float* mask, float* adv, float* ret, int B, int T, float gamma,
            float lam, int whiten, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(B > 0 && T > 0, DLA_E_BADARG, "gae: bad shape B=%d T=%d", B, T);
  rl::gae_kernel<<<1, rl::kThreads, 0, (cudaStream_t)stream>>>(rewards, values, mask, adv, ret, B, T, gamma, lam, whiten);
  return check_launch("gae");
}

}  // extern "C"
