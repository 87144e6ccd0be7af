/*
 * dla_b200 native library — C ABI of the sm_100a kernels.
 *
 * Conventions
 *   - every entry point returns 0 on success, a cudaError_t (>0) or a negative DLA_E_* code on
 *     failure; dla_last_error() returns a human-readable message for the calling thread.
 *   - pointers are device pointers unless a comment says "host"; "bf16" buffers hold
 *     __nv_bfloat16; strides/leading dimensions are in ELEMENTS.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Nothing here
 *     synchronises the host; nothing allocates unless stated.
 *   - "scratch" arguments are caller-provided global-memory work areas; their required sizes
 *     come from the matching *_scratch_bytes() query.
 *
 * The reference framework (nikhil-lalgudi/distributed-llm-alignment, "R/") ships no native code;
 * each declaration cites the Python line whose library kernels it replaces.
 */
#ifndef DLA_B200_H_
#define DLA_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLA_E_BADARG   (-1)
#define DLA_E_UNSUPP   (-2)
#define DLA_E_DRIVER   (-3)
#define DLA_E_TOOLARGE (-4)

const char* dla_last_error(void);
int dla_version(void);
/* One process drives one GPU: tell the library which (call once after torch.cuda.set_device).  Every
 * entry point binds that device to its calling thread, so autograd worker threads launch on it too. */
int dla_set_device(int device);
/* Number of kernel launches issued through this library since load (all threads). */
long long dla_launch_count(void);

/* ---------------------------------------------------------------- normalisation (K3) */
/* y = rmsnorm(x [+ delta]) * w ; optionally writes resid_out = x + delta.  R: HF LlamaRMSNorm
 * reached from R/src/training/train_dpo.py:32.  x,delta,resid_out,y: bf16 [M,H]; w: bf16 [H];
 * rstd: fp32 [M].  delta/resid_out may be NULL. H % 8 == 0, H <= 16384. */
int dla_rmsnorm_fwd(const void* x, const void* delta, const void* w, void* resid_out, void* y,
                    float* rstd, int64_t M, int H, float eps, void* stream);
/* dx = d(rmsnorm)/dx * dy [+ dresid]; dw += / = sum_rows(dy * xhat).  x is the normalised
 * input (resid_out of fwd).  scratch: fp32 [dla_norm_bwd_blocks() * H] (2H per block for layernorm). */
int dla_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd,
                    const void* dresid, void* dx, void* dw, float* scratch, int accumulate_dw,
                    int64_t M, int H, void* stream);
int dla_norm_bwd_blocks(void);
/* LayerNorm (GPT-2 / phi-2 family): y = (x-mean)*rstd*w + b. stats: fp32 [2*M] (mean, rstd). */
int dla_layernorm_fwd(const void* x, const void* w, const void* b, void* y, float* stats,
                      int64_t M, int H, float eps, void* stream);
int dla_layernorm_bwd(const void* dy, const void* x, const void* w, const float* stats,
                      void* dx, void* dw, void* db, float* scratch, int accumulate,
                      int64_t M, int H, void* stream);

/* ---------------------------------------------------------------- RoPE (K4) */
/* In-place rotate-half RoPE on the first n_heads heads of each row of a packed [M, row_stride]
 * bf16 buffer (q heads then k heads of a fused QKV projection).  cos/sin: fp32 [max_pos, rot_dim/2];
 * positions: int32 [M].  sign=+1 forward, -1 backward (inverse rotation of the gradient). */
int dla_rope_inplace(void* x, const float* cos_t, const float* sin_t, const int32_t* positions,
                     int64_t M, int n_heads, int head_dim, int rot_dim, int64_t row_stride,
                     float sign, void* stream);
/* Out-of-place form (full rotary, rot_dim == head_dim): dst receives the rotated first n_heads heads of src and a verbatim
 * copy of the n_copy_heads heads after them (the v part).  The backward uses it on its incoming gradient, which it may
 * not modify: one read + one write instead of a clone followed by the in-place pass. */
int dla_rope_copy(const void* src, void* dst, const float* cos_t, const float* sin_t, const int32_t* positions,
                  int64_t M, int n_heads, int n_copy_heads, int head_dim, int64_t src_stride,
                  int64_t dst_stride, float sign, void* stream);

/* ---------------------------------------------------------------- activations (K5) */
/* h = silu(gu[:, :F]) * gu[:, F:] ; gu bf16 [M, 2F], h bf16 [M, F]. */
int dla_swiglu_fwd(const void* gu, void* h, int64_t M, int F, void* stream);
int dla_swiglu_bwd(const void* dh, const void* gu, void* dgu, int64_t M, int F, void* stream);
/* tanh-GELU with optional bias add fused: y = gelu(x + b). */
int dla_gelu_fwd(const void* x, const void* bias, void* y, int64_t M, int F, void* stream);
int dla_gelu_bwd(const void* dy, const void* x, const void* bias, void* dx, int64_t M, int F, void* stream);
/* y = a + b (+ bias broadcast over rows) in bf16; used for residual / bias adds. */
int dla_add_bias(const void* a, const void* b, const void* bias, void* y, int64_t M, int H, void* stream);
/* db (bf16 [H]) (+)= column sums of dy [M,H]; scratch fp32 [dla_norm_bwd_blocks()*H]. */
int dla_bias_grad(const void* dy, void* db, float* scratch, int accumulate, int64_t M, int H, void* stream);

/* ---------------------------------------------------------------- embedding (K17) */
int dla_embedding_fwd(const int64_t* ids, const void* table, const void* pos_table, void* out,
                      int64_t M, int H, int V, int T, void* stream);
/* dtable[ids[m]] += dout[m] (bf16x2 atomics); dpos_table likewise for learned positions. */
int dla_embedding_bwd(const int64_t* ids, const void* dout, void* dtable, void* dpos_table,
                      int64_t M, int H, int V, int T, void* stream);

/* ---------------------------------------------------------------- optimizer (K13, K14) */
/* Fused AdamW on a flat shard: reads bf16 grad, updates fp32 master/m/v, writes the bf16 param.
 * grad_scale_dev (nullable): device fp32 scalar multiplied into the gradient (clip factor), so no
 * host sync sits between the norm and the update.  R: torch.optim.AdamW at
 * R/src/training/train_dpo.py:73-77,120.
 * flags: DLA_ADAMW_GRAD_FP32 = grad is fp32 (else bf16); DLA_ADAMW_MOMENTS_BF16 = m and v are bf16 buffers, stored
 * with stochastic rounding (halves the moment memory: the 70B-on-8-GPU sizing). */
#define DLA_ADAMW_GRAD_FP32     1
#define DLA_ADAMW_MOMENTS_BF16  2
int dla_adamw(float* master, void* m, void* v, const void* grad, void* param, int64_t n,
              float lr, float beta1, float beta2, float eps, float weight_decay,
              float bias_c1, float bias_c2, const float* grad_scale_dev, float grad_prescale,
              int flags, void* stream);
/* As dla_adamw, but `mc_param` is the NVSwitch MULTICAST address of the rank's shard inside the symmetric parameter
 * arena: each 16-byte result is written with one multimem.st and lands in every rank's resident copy, so no all-gather
 * follows the step.  n % 8 == 0.  R: the post-step parameter all-gather of FSDP/ZeRO-3. */
int dla_adamw_mc(float* master, void* m, void* v, const void* grad, void* mc_param, int64_t n,
                 float lr, float beta1, float beta2, float eps, float weight_decay,
                 float bias_c1, float bias_c2, const float* grad_scale_dev, float grad_prescale,
                 int flags, void* stream);
/* out[0] (+)= sum(x^2). scratch: fp32 [dla_sumsq_blocks()].  R: clip_grad_norm_ at
 * R/src/training/utils.py:121-123. */
int dla_sumsq(const void* x, int64_t n, int is_fp32, float* scratch, float* out, int accumulate,
              void* stream);
int dla_sumsq_blocks(void);
/* coef = min(1, max_norm / (sqrt(sumsq) + 1e-6)); norm_out = sqrt(sumsq). */
int dla_clip_coef(const float* sumsq, float max_norm, float* coef_out, float* norm_out, void* stream);
int dla_cast(const void* src, void* dst, int64_t n, int src_is_fp32, int dst_is_fp32, float scale, void* stream);

/* ---------------------------------------------------------------- losses (K8, K9, K11) */
/* DPO sigmoid loss over B pairs with gradients wrt the policy log-probs and the preference
 * rate, one launch.  out: fp32 [3] = {loss, preference_rate, mean_margin}; scale = 1/(global B).
 * R: R/src/training/train_dpo.py:42-44,130-132. */
int dla_dpo_loss(const float* pol_pos, const float* pol_neg, const float* ref_pos,
                 const float* ref_neg, int B, float beta, float label_smoothing, float scale,
                 float* out, float* dpol_pos, float* dpol_neg, void* stream);
/* Bradley-Terry pairwise loss + grads + accuracy.  out fp32[2] = {loss, accuracy}.
 * R: R/src/models/reward_model.py:67-68. */
int dla_pairwise_loss(const float* chosen, const float* rejected, int B, float scale, float* out,
                      float* dchosen, float* drejected, void* stream);
/* per-sequence masked mean of token log-probs: seq[b] = sum_t(tok*mask)/max(sum mask,1);
 * bwd: dtok = dseq[b]*mask/len.  R: R/src/training/train_dpo.py:38. */
int dla_seq_mean_fwd(const float* tok, const int64_t* mask, float* seq, int B, int T, int sum_not_mean, void* stream);
int dla_seq_mean_bwd(const float* dseq, const int64_t* mask, float* dtok, int B, int T, int sum_not_mean, void* stream);

/* ---------------------------------------------------------------- GEMM (K1) */
/* C[M,N] (+)= A[M,K] * B[N,K]^T with bf16 operands and fp32 accumulation in TMEM
 * (tcgen05.mma, TMA-fed, persistent).  a_mn_major: A is stored [K, M] (M contiguous) instead of
 * [M, K]; b_mn_major: B is stored [K, N] (N contiguous) instead of [N, K].  c_fp32 selects the
 * output type; accumulate adds into C.  K % 8 == 0; lda/ldb % 8 == 0; pointers 16-byte aligned.
 * R: nn.Linear -> cuBLASLt under R/src/training/train_dpo.py:32. */
int dla_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K,
                  int64_t lda, int64_t ldb, int64_t ldc, int a_mn_major, int b_mn_major,
                  int c_fp32, int accumulate, void* stream);
/* C[M,N] = A[M,K] * B[N,K]^T + bias[N] (bf16 bias, K-major operands): the bias add rides in the epilogue of either GEMM
 * kernel on the fp32 accumulators, so a biased linear (GPT-2's c_attn / c_proj) is one launch and one rounding.
 * R: nn.Linear(bias=True) / HF Conv1D under R/src/training/train_sft.py:66. */
int dla_gemm_bias_bf16(const void* A, const void* B, void* C, const void* bias, int M, int N, int K,
                       int64_t lda, int64_t ldb, int64_t ldc, int c_fp32, void* stream);
/* QKV projection with RoPE in the epilogue (2-CTA kernel, head_dim 128, full rotary): C[M,N] = A[M,K] B[N,K]^T with the
 * first n_rot_heads 128-wide heads of every row rotated (rotate-half pairs (c, c+64)) by positions[row] on the fp32
 * accumulators; cos/sin fp32 [max_pos, 64], positions int32 [M].  N % 128 == 0 (DLA_E_UNSUPP otherwise: run
 * dla_gemm_bf16 + dla_rope_inplace).  R: HF LlamaAttention q_proj/k_proj + apply_rotary_pos_emb under
 * R/src/training/train_dpo.py:32 (K1 + K4). */
int dla_gemm_rope_bf16(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc,
                       const float* cos_t, const float* sin_t, const int32_t* positions, int n_rot_heads, void* stream);
/* Gate/up projection of a gated MLP with the activation in the GEMM epilogue (2-CTA kernel): B = [2F, K] holds the gate
 * rows then the up rows; ACT[M, F] = silu(A B_gate^T) * (A B_up^T), computed from the fp32 accumulators.  GU (nullable,
 * [M, 2F] bf16) additionally receives gate | up for a backward pass.  Needs F % 128 == 0, M >= 256 (DLA_E_UNSUPP
 * otherwise: run dla_gemm_bf16 + dla_swiglu_fwd).  R: the HF Llama MLP under R/src/training/train_dpo.py:32 (K1 + K5). */
int dla_gemm_swiglu_bf16(const void* A, const void* B, void* GU, void* ACT, int M, int F, int K, int64_t lda,
                         int64_t ldb, int64_t ldgu, int64_t ldact, void* stream);
/* Backward counterpart: dGU[M, 2F] = d[gate | up] from dY[M, H] (gradient of the down projection's output), the down
 * weight Wdown[H, F] and the saved GU[M, 2F]: the dgrad GEMM dY Wdown whose epilogue applies the SwiGLU backward, so
 * d(act) never reaches memory.  Needs M >= 256, F % 64 == 0 (DLA_E_UNSUPP otherwise: dla_gemm_bf16 + dla_swiglu_bwd). */
int dla_gemm_swiglu_bwd_bf16(const void* dY, const void* Wdown, const void* GU, void* dGU, int M, int F, int H,
                             int64_t lddy, int64_t ldw, int64_t ldgu, int64_t lddgu, void* stream);
/* 0 = auto, 1 = 1-CTA 128x256, 2 = 2-CTA pair 256x256 (debug / benchmarking switch). */
int dla_gemm_set_variant(int variant);

/* ---------------------------------------------------------------- fused lm_head log-prob (K6, K7) */
/* logp[m] = log_softmax(X[m,:] . W^T)[labels[m]], lse[m] = logsumexp — the [M,V] logits never
 * reach memory: the GEMM epilogue reduces each 128x256 tile to (max, sumexp, picked) partials.
 * scratch: dla_lmhead_scratch_bytes(M, V).  labels < 0 => logp = 0.
 * R: R/src/training/train_dpo.py:31-39 (log_softmax + gather over the vocab). */
int dla_lmhead_logprob(const void* X, const void* W, const int64_t* labels, int M, int V, int K,
                       int64_t ldx, int64_t ldw, void* scratch, float* logp, float* lse,
                       void* stream);
/* Same, additionally storing the bf16 logits (row pitch ld_logits >= V, multiple of 8) when logits_out != NULL: a
 * memory-for-compute trade for the backward, which then skips its recomputation GEMM (2 bytes x M x V). */
int dla_lmhead_logprob_keep(const void* X, const void* W, const int64_t* labels, int M, int V, int K,
                            int64_t ldx, int64_t ldw, void* scratch, float* logp, float* lse,
                            void* logits_out, int64_t ld_logits, void* stream);
size_t dla_lmhead_scratch_bytes(int M, int V);
/* In place on a bf16 logits chunk [Mc, V]: logits <- grow[m] * (onehot(label) - softmax(logits)),
 * using the saved lse (backward of the op above, chunked). */
int dla_logprob_grad_inplace(void* logits, const float* lse, const int64_t* labels,
                             const float* grow, int Mc, int V, int64_t ld, void* stream);
/* dla_lmhead_logprob for callers that cannot allocate device memory (no counterpart in the reference,
 * which has no native code and lets PyTorch allocate).  Think of a scoring / inference server that gave
 * nearly all of HBM to its own cache pool at start-up: a cudaMalloc for our work area would fail or eat
 * into that pool, so here the tile partials live in a fixed ~4 MiB __device__ array inside the library
 * image (held on every GPU that loads the library, until the process exits).  This entry point never
 * calls cudaMalloc/cudaFree or an allocating library routine; the caller still owns inputs and outputs.
 *   limits : M <= 1024 rows per call and V <= 131072 (dla_lmhead_logprob_noalloc_max_rows/_max_vocab);
 *            larger sizes are rejected with DLA_E_TOOLARGE and a message, nothing is launched.
 *   set-up : the first call on each device resolves that device's address of the array — make it once
 *            at start-up, before any stream capture.
 *   sharing: all calls on a GPU use the same array and inherit its previous contents; the slots that must
 *            start at zero are cleared on `stream` at the top of every call.  Calls whose GPU work overlaps
 *            (other streams / threads) race on it: keep them on one stream or order them yourself. */
int dla_lmhead_logprob_noalloc(const void* X, const void* W, const int64_t* labels, int M, int V,
                               int K, int64_t ldx, int64_t ldw, float* logp, float* lse, void* stream);
int dla_lmhead_logprob_noalloc_max_rows(void);
int dla_lmhead_logprob_noalloc_max_vocab(void);

/* ---------------------------------------------------------------- attention (K2, K16) */
/* Causal GQA flash attention, tcgen05: q [B,T,Hq,D], k/v [B,T,Hk,D] bf16 with explicit element
 * strides (batch, token, head); D == 128 or 64.  seqlens (nullable int32 [B]) masks right padding.
 * lse: fp32 [B,Hq,T]. */
/* Forward kernel choice: 0 = auto (two-query-tile ping-pong kernel, csrc/attn_fwd2.cu), 1 = first-generation
 * one-tile kernel (csrc/attention.cu), 2 = ping-pong, 3 = ping-pong with early S release / P through shared memory
 * (csrc/attn_fwd3.cu; measured slower, experimental).  Process-wide; for benchmarks and A/B tests. */
int dla_attn_set_fwd_variant(int v);
/* Backward kernel choice: 0 = auto (pipelined kernel with transposed score tiles, csrc/attn_bwd2.cu, from T = 2048 on;
 * first-generation serial kernel, csrc/attention.cu, below), 1 = first generation, 2 = pipelined. */
int dla_attn_set_bwd_variant(int v);
int dla_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse,
                 int B, int T, int Hq, int Hk, int D,
                 int64_t q_sb, int64_t q_st, int64_t q_sh, int64_t kv_sb, int64_t kv_st, int64_t kv_sh,
                 int64_t o_sb, int64_t o_st, int64_t o_sh,
                 const int32_t* seqlens, float scale, int causal, void* stream);
/* Backward. dq_acc: fp32 scratch of B * Hq * ceil(T/128)*128 * D elements (zeroed inside); delta: fp32 scratch [B,Hq,T].
 * dq must have q's strides, dk/dv must have k/v's strides; dout must have o's strides. */
int dla_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout,
                 const float* lse, void* dq, void* dk, void* dv, float* dq_acc, float* delta,
                 int B, int T, int Hq, int Hk, int D,
                 int64_t q_sb, int64_t q_st, int64_t q_sh, int64_t kv_sb, int64_t kv_st, int64_t kv_sh,
                 int64_t o_sb, int64_t o_st, int64_t o_sh,
                 const int32_t* seqlens, float scale, int causal, void* stream);
/* Packed rows (several documents per row, `data.packing: true` of R/config/sft_config.yaml:16): as dla_attn_fwd /
 * dla_attn_bwd with seg_start (nullable int32 [B,T], contiguous): seg_start[b,t] = index of the first token of t's
 * document, non-decreasing along t.  A query sees the keys [seg_start[q], q]; kv tiles (forward) and query tiles
 * (backward) that lie entirely outside that range are skipped, not just masked.  Causal only. */
int dla_attn_fwd_seg(const void* q, const void* k, const void* v, void* o, float* lse,
                     int B, int T, int Hq, int Hk, int D,
                     int64_t q_sb, int64_t q_st, int64_t q_sh, int64_t kv_sb, int64_t kv_st, int64_t kv_sh,
                     int64_t o_sb, int64_t o_st, int64_t o_sh,
                     const int32_t* seqlens, const int32_t* seg_start, float scale, int causal, void* stream);
int dla_attn_bwd_seg(const void* q, const void* k, const void* v, const void* o, const void* dout,
                     const float* lse, void* dq, void* dk, void* dv, float* dq_acc, float* delta,
                     int B, int T, int Hq, int Hk, int D,
                     int64_t q_sb, int64_t q_st, int64_t q_sh, int64_t kv_sb, int64_t kv_st, int64_t kv_sh,
                     int64_t o_sb, int64_t o_st, int64_t o_sh,
                     const int32_t* seqlens, const int32_t* seg_start, float scale, int causal, void* stream);
/* Single-token decode attention over a contiguous KV cache [B, S_max, Hk, D] with per-row lengths. */
int dla_attn_decode(const void* q, const void* kcache, const void* vcache, void* o,
                    const int32_t* lens, int B, int S_max, int Hq, int Hk, int D, float scale,
                    void* stream);

/* ---------------------------------------------------------------- decode (weight-streaming) GEMM */
/* y[M,N] = x[M,K] @ w[N,K]^T (+ bias[N]) (+ residual[M,N]) for M <= dla_gemm_skinny_max_rows() (32): grid over N,
 * K split across the warps of a CTA, mma.sync on 16-byte global loads; bound by the HBM stream of w.  K % 32 == 0,
 * row pitches multiples of 8 elements, 16-byte aligned bases.  weights_static != 0 launches with programmatic
 * stream serialization: the grid may start before its predecessor ends and prefetches w ahead of
 * griddepcontrol.wait — only legal when no earlier kernel in the stream writes w (decode, not training).
 * R: per-token nn.Linear inside generate(), R/src/training/train_rlhf.py:123-124. */
int dla_gemm_skinny_bf16(const void* x, const void* w, void* y, const void* bias, const void* residual, int M,
                         int N, int K, int64_t ldx, int64_t ldw, int64_t ldy, int64_t ldr, int out_fp32,
                         int weights_static, void* stream);
int dla_gemm_skinny_max_rows(void);

/* One-token decode attention fused with RoPE and the KV-cache append (csrc/decode.cu): qkv is the fresh
 * [B, (Hq+2Hk)*D] projection of one token per sequence (row pitch qkv_stride elements); lens[b] counts the valid
 * positions INCLUDING this token.  The kernel rotates q and k (rot_dim leading dims of each head, table
 * cos_t/sin_t [max_pos, rot_dim/2]; rot_dim 0 = no RoPE), writes k/v at position lens[b]-1 of the
 * [B, S_max, Hk, D] caches and attends over lens[b] positions; o is [B, Hq, D].  D in {64,128}, Hq/Hk in {1,2,4,8}.
 * R: DynamicCache.update + apply_rotary_pos_emb + SDPA per generated token, R/src/training/train_rlhf.py:123-124. */
int dla_decode_attn_fused(void* qkv, int64_t qkv_stride, void* kcache, void* vcache, void* o, const int32_t* lens,
                          const float* cos_t, const float* sin_t, int B, int S_max, int Hq, int Hk, int D,
                          int rot_dim, float scale, void* stream);

/* ---------------------------------------------------------------- scalar heads, RL objectives, distillation */
/* Reward / value head (csrc/heads.cu).  hidden bf16 [B,T,H], mask int64 [B,T], w bf16 [H], bias bf16 [1] (nullable),
 * keep fp32 [B,H] dropout mask already scaled by 1/(1-p) (nullable).  mode 0: out[b] = w.(keep*hidden[b, n_b-1]);
 * mode 1: masked mean pooling; mode 2: out[b,t] per token (mask/keep unused).  bwd writes dense dhidden (nullable)
 * and ACCUMULATES fp32 dw[H], db[1] (nullable; caller zeroes).  R: R/src/models/reward_model.py:44,54-64. */
int dla_scalar_head_fwd(const void* hidden, const int64_t* mask, const void* w, const void* bias, const float* keep,
                        float* out, int B, int T, int H, int mode, void* stream);
int dla_scalar_head_bwd(const void* hidden, const int64_t* mask, const void* w, const float* keep, const float* dout,
                        void* dhidden, float* dw, float* db, int B, int T, int H, int mode, void* stream);

/* Sequence-level REINFORCE: out = {loss, mean kl, mean reward}; kl_out/rew_out [B] and dpol [B] nullable; baseline
 * (device scalar, nullable) overrides the batch-mean baseline.  R: R/src/training/train_rlhf.py:149-153. */
int dla_reinforce_loss(const float* pol, const float* ref, const float* score, int B, float kl_coef,
                       const float* baseline, float* out, float* kl_out, float* rew_out, float* dpol, void* stream);
/* Token-level clipped PPO + clipped value loss over N masked elements: out = {loss, pg, vf, approx_kl, clipfrac};
 * values/old_values/returns nullable together.  GAE (+ optional advantage whitening) over [B,T].  Beyond the
 * reference: the `ppo.*` keys of R/config/rlhf_config.yaml:14,17 (SURVEY worklist #20). */
int dla_ppo_loss(const float* new_logp, const float* old_logp, const float* adv, const float* values,
                 const float* old_values, const float* returns, const float* mask, int N, float clip, float vclip,
                 float vf_coef, float* out, float* dlogp, float* dvalues, void* stream);
int dla_gae(const float* rewards, const float* values, const float* mask, float* adv, float* ret, int B, int T,
            float gamma, float lam, int whiten, void* stream);

/* Distillation KL per logits row (csrc/distill.cu): kl_out[r] = KL(mean_k softmax(t_k/T) || softmax(s/T)) and/or
 * dl[r,:] = (softmax(s/T) - mean_k softmax(t_k/T)) * coef[r] / T in bf16.  student / teachers fp32 [rows, V] with
 * pitches in elements; teacher_ptrs / teacher_ld are HOST arrays of n_teachers (<= dla_distill_max_teachers())
 * device pointers / pitches.  R: R/src/training/train_distill.py:132-144. */
int dla_distill_kl(const float* student, int64_t ld_s, const float* const* teacher_ptrs, const int64_t* teacher_ld,
                   int n_teachers, const float* coef, int rows, int V, float temperature, float* kl_out, void* dl,
                   int64_t ld_d, void* stream);
int dla_distill_max_teachers(void);

/* ---------------------------------------------------------------- sampling (K16) */
/* Per-row temperature + top-p + multinomial on fp32/bf16 logits [B,V]; writes int64 token ids.
 * seed/offset: counter RNG; offset_dev (nullable) is a device int64 added to offset inside the kernel, so a
 * captured decode graph draws fresh numbers on every replay.  No work area is needed. */
int dla_sample_top_p(const void* logits, int logits_fp32, int B, int V, int64_t ld, float temperature,
                     float top_p, uint64_t seed, uint64_t offset, const int64_t* offset_dev, int64_t* out,
                     void* stream);
size_t dla_sampler_scratch_bytes(int B, int V);

/* ---------------------------------------------------------------- peer-memory fused collectives */
/* (a) reduce-scatter + cast/scale + grad-norm partial + sharded AdamW in one kernel: every rank
 * reads the W peer copies of ITS shard of the bf16 gradient bucket over NVLink (peer_grads: device
 * array of W device pointers to the full-size buckets), reduces in fp32, optionally accumulates
 * sum(g^2) and applies AdamW to its fp32 shard, writing the updated bf16 shard into every peer's
 * parameter buffer (peer_params, nullable => local only).  mode: 0 = reduce only into grad_out,
 * 1 = reduce + sumsq, 2 = reduce + adamw.  R: NCCL reduce_scatter + clip + AdamW,
 * R/src/training/train_dpo.py:118-120. */
int dla_fused_rs_adamw(const void* const* peer_grads, void* const* peer_params, int world, int rank,
                       int64_t shard_elems, int64_t shard_offset, float grad_scale,
                       float* grad_out_fp32, float* sumsq_out, float* master, float* m, float* v,
                       float lr, float beta1, float beta2, float eps, float weight_decay,
                       float bias_c1, float bias_c2, const float* clip_coef_dev, int mode,
                       void* stream);
/* All-gather by peer stores: src[0..n) (bf16) -> element offset dst_offset of every rank's buffer
 * (peers: device array of `world` base pointers).  n and dst_offset multiples of 8. */
int dla_peer_push(const void* src, void* const* peers, int world, int rank, int64_t dst_offset,
                  int64_t n, void* stream);
/* All-reduce (sum * scale) of the bf16 range [offset, offset+n) of a symmetric buffer, in place on every rank:
 * each rank reduces 1/world of the range from all peers over NVLink and stores the result into all of them.
 * Bracket with dla_peer_barrier on both sides.  n % (8*world) == 0.  R: DDP's bucketed NCCL all_reduce under
 * accelerator.backward (R/src/training/train_dpo.py:118; SURVEY C1, worklist #14). */
int dla_peer_allreduce(void* const* peers, int world, int rank, int64_t offset, int64_t n, float scale, void* stream);
/* Signal-pad barrier across ranks (system-scope release/acquire flags in peer memory).
 * pads: device array of W pointers to each rank's uint32 pad [W * channels]. */
int dla_peer_barrier(uint32_t* const* pads, int world, int rank, int channel, uint32_t epoch,
                     int timeout_ms, void* stream);
/* 1 if any peer-flag wait of this process timed out since load (synchronises the device). */
int dla_peer_status(int* timed_out);
/* (b) all-gather fused GEMM: C[M,N] = A[M,K] * B[N,K]^T where B = unit_full + b_offset and the unit's
 * elements are sharded contiguously across ranks (rank r owns [r*shard_elems, (r+1)*shard_elems); peer_shards:
 * device array of `world` pointers to every rank's shard in peer-mapped memory).  ONE kernel: the first
 * n_copy_ctas CTAs TMA-bulk-copy B's rows from their owners over NVLink into unit_full, tile by tile, publishing
 * epoch-stamped flags; the other CTAs run the tcgen05 GEMM and consume each B tile as soon as its flag shows
 * `epoch` (tiles owned by this rank first).  flags: uint32 [ceil(N/256)], reused across calls with a fresh epoch.
 * R: NCCL all_gather_into_tensor + cuBLAS issued by FSDP/ZeRO-3 under R/src/training/train_dpo.py:32. */
int dla_ag_gemm_bf16(const void* A, const void* const* peer_shards, void* unit_full, int64_t shard_elems,
                     int64_t b_offset, void* C, int M, int N, int K, int64_t lda, int64_t ldc, int world,
                     int rank, uint32_t* flags, uint32_t epoch, int n_copy_ctas, int c_fp32, void* stream);
/* Plain peer gather of a flat range [offset, offset+n) of a sharded unit into unit_full (no GEMM): used for the
 * small non-GEMM parameters of a streamed unit. */
int dla_peer_gather(const void* const* peer_shards, void* unit_full, int64_t shard_elems, int64_t offset, int64_t n,
                    int world, void* stream);
/* (a) v2 — one unit's gradient reduce-scatter, meant to be launched on a side stream DURING backward as soon as this
 * rank's bucket for the unit is final.  The unit owns `world` flag words at pads[*] + flag_word: block 0 publishes
 * `epoch` to every peer, all blocks wait for all ranks' flags (timeout_ms of wall time, 0 = forever), then the rank's
 * shard [shard_offset, +shard_elems) is summed into grad_out_fp32 (x scale) and sum(out^2) is added to sumsq_out
 * (nullable).  mc_grads != NULL: the sum is done inside NVSwitch by multimem.ld_reduce on the multicast address of
 * the gradient arena (bf16 result, fp32 accumulation); else W peer loads from peer_grads.  max_blocks <= 0: 2 per SM.
 * R: the bucketed reduce-scatter DeepSpeed/FSDP overlap with backward (R/config/deepspeed_zero3.json:13-14,
 * R/src/training/train_dpo.py:118). */
int dla_unit_reduce(uint32_t* const* pads, int flag_word, const void* mc_grads, const void* const* peer_grads,
                    int world, int rank, int64_t shard_elems, int64_t shard_offset, float scale,
                    float* grad_out_fp32, float* sumsq_out, uint32_t epoch, int timeout_ms, int max_blocks,
                    void* stream);
/* dla_unit_reduce with accumulate != 0: grad_out_fp32 += scale * sum (streamed ZeRO-3 units fold every micro-step's
 * reduce-scatter into the fp32 gradient shard).  R: R/config/deepspeed_zero3.json:5-15 (stage 3, overlap_comm). */
int dla_unit_reduce_acc(uint32_t* const* pads, int flag_word, const void* mc_grads, const void* const* peer_grads,
                        int world, int rank, int64_t shard_elems, int64_t shard_offset, float scale,
                        float* grad_out_fp32, float* sumsq_out, uint32_t epoch, int timeout_ms, int max_blocks,
                        int accumulate, void* stream);
/* One half of a flag exchange on a row of `world` flag words at word offset flag_word of every pad: mode 1 publishes
 * `epoch` to every peer, mode 2 waits until all ranks have published it to this rank.  Brackets the reuse of a
 * gradient-bucket ring slot that peers read directly. */
int dla_peer_flag_row(uint32_t* const* pads, int world, int rank, int flag_word, uint32_t epoch, int mode,
                      int timeout_ms, void* stream);
/* Global grad-norm + clip coefficient without an NCCL launch: exchanges every rank's partial sum(g^2) over peer memory
 * (bufs: per-rank fp32 [2][world]), sums in rank order, writes norm = sqrt(total) * prescale and
 * coef = min(1, max_norm / (norm + 1e-6)) (0 when non-finite), and resets *sumsq_local to 0.
 * R: clip_grad_norm_'s all-reduce (R/src/training/utils.py:121-123). */
int dla_peer_norm_clip(float* const* bufs, uint32_t* const* pads, int world, int rank, float* sumsq_local,
                       float prescale, float max_norm, uint32_t epoch, int channel, float* coef, float* norm_out,
                       void* stream);
/* All-gather by multicast stores: src[0..n) (bf16) -> element offset dst_offset of EVERY rank's copy (mc_dst is the
 * multicast base address of the symmetric buffer). */
int dla_peer_push_mc(const void* src, void* mc_dst, int64_t dst_offset, int64_t n, void* stream);
/* dla_peer_allreduce through the switch: multimem.ld_reduce of this rank's slice + multimem.st of the result. */
int dla_peer_allreduce_mc(void* mc, int world, int rank, int64_t offset, int64_t n, float scale, void* stream);
/* Mirror the "a peer-flag wait timed out" bit into a word of pinned (device-mapped) host memory so that the training
 * loop can poll it every step without synchronising the device.  NULL switches the mirror off. */
int dla_peer_watch(void* host_mapped_word);
/* (c) cross-rank all-gather of per-sequence log-probs fused with the DPO loss: each rank writes its
 * [4, B] block into every peer's gather buffer (fp32 [2][world][4][1024], half = epoch parity), exchanges flags, then
 * reduces loss / preference rate / margin over all W*B pairs and produces its local gradients. */
int dla_fused_gather_dpo(float* const* peer_bufs, uint32_t* const* pads, int world, int rank,
                         const float* pol_pos, const float* pol_neg, const float* ref_pos,
                         const float* ref_neg, int B, float beta, float label_smoothing,
                         uint32_t epoch, float* out, float* dpol_pos, float* dpol_neg, void* stream);

/* (c) for the RLHF phase, sequence-level REINFORCE over the GLOBAL rollout batch: gathers every rank's
 * (policy log-prob, reference log-prob, score) [3, B] block over peer memory (peer_bufs: fp32 [2][world][3][1024]) and
 * computes r = score - kl_coef * (pol - ref), the global baseline mean(r), out = {loss, mean kl, mean r, mean score,
 * baseline} over all W*B sequences and dpol[B] = -(r - baseline) / B for the local ones.
 * R: R/src/training/train_rlhf.py:149-153 (rank-local there). */
int dla_fused_gather_reinforce(float* const* peer_bufs, uint32_t* const* pads, int world, int rank, const float* pol,
                               const float* ref, const float* score, int B, float kl_coef, uint32_t epoch,
                               float* out, float* dpol, void* stream);
/* (c) for clipped PPO: advantage whitening with GLOBAL mean / variance.  Each rank contributes the masked sums of its
 * adv[n] (and up to 13 extra per-rank sums in `stats`, returned summed over ranks in stats_out), then normalises adv
 * in place.  bufs: fp32 [2][world][16].  R: the `ppo.*` keys of R/config/rlhf_config.yaml:12-22. */
int dla_fused_gather_whiten(float* const* bufs, uint32_t* const* pads, int world, int rank, float* adv,
                            const float* mask, int n, const float* stats, int nstats, uint32_t epoch,
                            float* stats_out, void* stream);

/* Tiny all-gather over peer memory (eval losses, per-rank statistics): out[r*n + i] = rank r's src[i].  bufs: device
 * array of per-rank fp32 buffers [2][world][nmax] (the half is picked by the epoch parity); pads/channel/epoch as
 * for dla_peer_barrier.  R: accelerator.gather at R/src/training/train_sft.py:39, train_reward.py:47. */
int dla_peer_allgather_small(float* const* bufs, uint32_t* const* pads, int world, int rank, const float* src, int n,
                             int nmax, uint32_t epoch, int channel, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DLA_B200_H_ */
