// Sequence-level losses: DPO, Bradley-Terry pairwise, masked per-sequence means.
// One launch each, forward value + gradients + logged statistics together
// (reference: 5 tiny ATen kernels per loss; R/src/training/train_dpo.py:42-44).
#include "common.cuh"

namespace dla {

__device__ __forceinline__ float log_sigmoid(float z) {
  // log(sigmoid(z)) = min(z,0) - log1p(exp(-|z|))
  return fminf(z, 0.f) - log1pf(__expf(-fabsf(z)));
}
__device__ __forceinline__ float sigmoidf_(float z) { return 1.f / (1.f + __expf(-z)); }

__global__ void __launch_bounds__(256)
dpo_loss_kernel(const float* __restrict__ pp, const float* __restrict__ pn, const float* __restrict__ rp,
                const float* __restrict__ rn, int B, float beta, float ls, float scale, float* __restrict__ out,
                float* __restrict__ dpp, float* __restrict__ dpn) {
  __shared__ float red[33];
  float loss = 0.f, pref = 0.f, margin = 0.f;
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    const float z = beta * ((pp[i] - pn[i]) - (rp[i] - rn[i]));
    loss += -(1.f - ls) * log_sigmoid(z) - ls * log_sigmoid(-z);
    // d/dz = -(1-ls)*sigmoid(-z) + ls*sigmoid(z)
    const float dz = (-(1.f - ls) * sigmoidf_(-z) + ls * sigmoidf_(z)) * beta * scale;
    if (dpp != nullptr) dpp[i] = dz;
    if (dpn != nullptr) dpn[i] = -dz;
    pref += (pp[i] > pn[i]) ? 1.f : 0.f;
    margin += z;
  }
  loss = block_sum(loss, red);
  pref = block_sum(pref, red);
  margin = block_sum(margin, red);
  if (threadIdx.x == 0) {
    out[0] = loss * scale;
    out[1] = pref * scale;
    out[2] = margin * scale;
  }
}

__global__ void __launch_bounds__(256)
pairwise_loss_kernel(const float* __restrict__ c, const float* __restrict__ r, int B, float scale,
                     float* __restrict__ out, float* __restrict__ dc, float* __restrict__ dr) {
  __shared__ float red[33];
  float loss = 0.f, acc = 0.f;
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    const float z = c[i] - r[i];
    loss += -log_sigmoid(z);
    const float dz = -sigmoidf_(-z) * scale;
    if (dc != nullptr) dc[i] = dz;
    if (dr != nullptr) dr[i] = -dz;
    acc += (c[i] > r[i]) ? 1.f : 0.f;
  }
  loss = block_sum(loss, red);
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) {
    out[0] = loss * scale;
    out[1] = acc * scale;
  }
}

// one warp per sequence
__global__ void seq_mean_fwd_kernel(const float* __restrict__ tok, const int64_t* __restrict__ mask,
                                    float* __restrict__ seq, int B, int T, int sum_not_mean) {
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;
  const int lane = threadIdx.x & 31;
  float s = 0.f, cnt = 0.f;
  for (int t = lane; t < T; t += 32) {
    const float mk = mask[(int64_t)b * T + t] != 0 ? 1.f : 0.f;
    s += tok[(int64_t)b * T + t] * mk;
    cnt += mk;
  }
  s = warp_sum(s);
  cnt = warp_sum(cnt);
  if (lane == 0) seq[b] = sum_not_mean ? s : s / fmaxf(cnt, 1.f);
}

__global__ void seq_mean_bwd_kernel(const float* __restrict__ dseq, const int64_t* __restrict__ mask,
                                    float* __restrict__ dtok, int B, int T, int sum_not_mean) {
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;
  const int lane = threadIdx.x & 31;
  float cnt = 0.f;
  for (int t = lane; t < T; t += 32) cnt += mask[(int64_t)b * T + t] != 0 ? 1.f : 0.f;
  cnt = warp_sum(cnt);
  const float g = sum_not_mean ? dseq[b] : dseq[b] / fmaxf(cnt, 1.f);
  for (int t = lane; t < T; t += 32) dtok[(int64_t)b * T + t] = mask[(int64_t)b * T + t] != 0 ? g : 0.f;
}

}  // namespace dla

using namespace dla;

extern "C" {

int dla_dpo_loss(const float* pol_pos, const float* pol_neg, const float* ref_pos, const float* ref_neg, int B,
                 float beta, float label_smoothing, float scale, float* out, float* dpol_pos, float* dpol_neg,
                 void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(B > 0 && out != nullptr, DLA_E_BADARG, "dpo_loss: B must be > 0");
  dpo_loss_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(pol_pos, pol_neg, ref_pos, ref_neg, B, beta, label_smoothing,
                                                      scale, out, dpol_pos, dpol_neg);
  return check_launch("dpo_loss");
}

int dla_pairwise_loss(const float* chosen, const float* rejected, int B, float scale, float* out, float* dchosen,
                      float* drejected, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(B > 0 && out != nullptr, DLA_E_BADARG, "pairwise_loss: B must be > 0");
  pairwise_loss_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(chosen, rejected, B, scale, out, dchosen, drejected);
  return check_launch("pairwise_loss");
}

int dla_seq_mean_fwd(const float* tok, const int64_t* mask, float* seq, int B, int T, int sum_not_mean, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  if (B == 0) return 0;
  seq_mean_fwd_kernel<<<(B + 7) / 8, 256, 0, (cudaStream_t)stream>>>(tok, mask, seq, B, T, sum_not_mean);
  return check_launch("seq_mean_fwd");
}

int dla_seq_mean_bwd(const float* dseq, const int64_t* mask, float* dtok, int B, int T, int sum_not_mean, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  if (B == 0) return 0;
  seq_mean_bwd_kernel<<<(B + 7) / 8, 256, 0, (cudaStream_t)stream>>>(dseq, mask, dtok, B, T, sum_not_mean);
  return check_launch("seq_mean_bwd");
}

}  // extern "C"
