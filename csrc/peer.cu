// Collectives fused with their neighbouring compute, issued from inside the kernels over NVLink 5 /
// NVSwitch peer memory (symmetric buffers: every rank holds device pointers to every peer's copy).
//
//   dla_peer_barrier      signal-pad barrier: st.release.sys flags into every peer, ld.acquire.sys spin
//                         with a clock-based timeout (a dead peer must not hang the GPU forever)
//   dla_fused_rs_adamw    (a) gradient reduce-scatter fused with cast/scale, the grad-norm partial and
//                         (optionally) the sharded AdamW update + push of the updated bf16 shard into
//                         every peer's resident parameter buffer (the post-step all-gather disappears)
//   dla_unit_reduce       (a) v2: one unit's reduce-scatter launched DURING backward: per-unit epoch flags replace the
//                         global barrier, the sum happens inside NVSwitch (multimem.ld_reduce on the multicast
//                         address of the gradient arena) or, without multicast, by W peer loads
//   dla_peer_norm_clip    the 4-byte grad-norm exchange + clip coefficient in one single-block kernel (no NCCL launch)
//   dla_peer_push         all-gather by P2P stores of an updated shard into all peers' buffers (multimem.st when a
//                         multicast address is given: one store, the switch replicates it)
//   dla_fused_gather_dpo  (c) all-gather of per-sequence log-probs fused with the DPO loss and the
//                         global preference-rate / margin reduction
//
// Reference equivalents: NCCL reduce_scatter / all_gather issued by FSDP/DeepSpeed under
// accelerator.backward + clip_grad_norm_ + optimizer.step (R/src/training/train_dpo.py:118-121) and
// rank-local ATen loss math (R/src/training/train_dpo.py:42-44, 130-132).
#include "common.cuh"

namespace dla {
namespace peer {

constexpr int kMaxWorld = 16;
// set to 1 by any kernel whose wait on a peer flag timed out (dead / wedged peer); read with dla_peer_status()
__device__ uint32_t g_peer_timeout_flag = 0;
// optional mirror in pinned host memory (dla_peer_watch): the training loop polls it without touching the device
__device__ uint32_t* g_peer_status_host = nullptr;

// Debug build (DLA_BUILD_DEFINES="-DDLA_PEER_DEBUG" python -m dla_b200.ops.build --force): sequence-number assertions on
// the signal pad.  Epochs on a flag row advance by exactly one per use and no rank can be more than one use ahead of a
// peer (every use ends in a wait for all ranks), so a flag that jumps, repeats or runs ahead by two means a protocol bug
// (a missed launch on some rank, two owners of one row, a row reused before its readers left) — trap instead of
// computing on torn data.  SURVEY §5.2: the stand-in for a race detector on the one protocol compute-sanitizer
// cannot see (it is cross-GPU).
#ifdef DLA_PEER_DEBUG
#define DLA_PEER_ASSERT(cond, what, a, b)                                                             \
  do {                                                                                                \
    if (!(cond)) {                                                                                    \
      printf("dla peer protocol violation: %s (%u vs %u) block %d thread %d\n", what, (unsigned)(a), \
             (unsigned)(b), (int)blockIdx.x, (int)threadIdx.x);                                       \
      __trap();                                                                                       \
    }                                                                                                 \
  } while (0)
#else
#define DLA_PEER_ASSERT(cond, what, a, b) ((void)0)
#endif

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void note_timeout(uint32_t* status) {
  atomicExch(status != nullptr ? status : &g_peer_timeout_flag, 1u);
  uint32_t* h = g_peer_status_host;
  if (h != nullptr) {
    *reinterpret_cast<volatile uint32_t*>(h) = 1u;
    __threadfence_system();
  }
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// 16-byte load that bypasses L1 (peer data is written by another GPU; never reuse a stale line)
__device__ __forceinline__ uint4 ld_peer16(const void* p) {
  uint4 r;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_peer16(void* p, const uint4& v) {
  asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- multimem: one instruction addresses the same offset of every rank's copy through the NVSwitch multicast object
// 8 bf16 summed over all ranks inside the switch (fp32 accumulation, result rounded to bf16)
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
  return r;
}
// 16 bytes stored into every rank's copy
__device__ __forceinline__ void multimem_st16(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w)) : "memory");
}

// spin until *flag >= epoch (wrap-safe); false on timeout.  The limit is wall time (%globaltimer, ns), so it does not
// stretch when the SM clock drops under a power cap.
__device__ __forceinline__ bool wait_flag(const uint32_t* flag, uint32_t epoch, unsigned long long timeout_ns) {
  const unsigned long long t0 = globaltimer_ns();
  uint32_t seen;
  while ((int32_t)((seen = ld_acquire_sys(flag)) - epoch) < 0) {
    if (timeout_ns > 0 && globaltimer_ns() - t0 > timeout_ns) return false;
    __nanosleep(64);
  }
  DLA_PEER_ASSERT((int32_t)(seen - epoch) <= 1, "a peer is two or more uses ahead on this flag", seen, epoch);
  return true;
}

// pads[r] -> rank r's pad: uint32 [channels][world].  Rank `me` publishes `epoch` into slot [channel][me]
// of every peer, then waits until its own pad shows >= epoch from everyone.  status[0] is set to 1 on timeout.
__device__ __forceinline__ bool barrier_body(uint32_t* const* pads, int world, int me, int channel, uint32_t epoch,
                                             unsigned long long timeout_ns) {
  const int t = threadIdx.x;
  bool ok = true;
  if (t < world) {
    // my own slot of my own pad holds what I published last time: the sequence must advance by exactly one
    DLA_PEER_ASSERT(t != me || ld_acquire_sys(pads[me] + channel * world + me) + 1u == epoch,
                    "barrier epoch does not follow the previous one", ld_acquire_sys(pads[me] + channel * world + me), epoch);
    __threadfence_system();
    st_release_sys(pads[t] + channel * world + me, epoch);
    ok = wait_flag(pads[me] + channel * world + t, epoch, timeout_ns);
  }
  return ok;
}

// One-sided halves of the flag exchange on an arbitrary row of `world` flag words (word offset flag_word in every pad):
// mode 1 = publish `epoch` into slot `me` of every peer's row, mode 2 = wait until every rank's slot of MY row shows it.
// Used around the gradient-bucket ring of streamed units: "I have finished reading your bucket" / "everybody has
// finished reading mine, it may be refilled".
__global__ void flag_row_kernel(uint32_t* const* pads, int world, int me, int flag_word, uint32_t epoch, int mode,
                                unsigned long long timeout_ns) {
  const int t = threadIdx.x;
  if (t >= world) return;
  if (mode == 1) {
    __threadfence_system();
    st_release_sys(pads[t] + flag_word + me, epoch);
  } else if (!wait_flag(pads[me] + flag_word + t, epoch, timeout_ns)) {
    note_timeout(nullptr);
  }
}

__global__ void barrier_kernel(uint32_t* const* pads, int world, int me, int channel, uint32_t epoch,
                               unsigned long long timeout_ns, uint32_t* status) {
  const bool ok = barrier_body(pads, world, me, channel, epoch, timeout_ns);
  if (!ok) note_timeout(status);
}

// ------------------------------------------------------------------------------- (a) RS + AdamW
struct RsParams {
  const bf16* const* peer_grads;   // [world] base pointers of every rank's full-size gradient bucket
  bf16* const* peer_params;        // [world] base pointers of every rank's resident bf16 parameters (nullable)
  int world, rank;
  int64_t shard_elems, shard_offset;   // this rank's shard: elements [shard_offset, shard_offset + shard_elems)
  float grad_scale;
  float* grad_out;                 // fp32 [shard_elems] reduced + scaled gradient (modes 0/1)
  float* sumsq_out;                // += sum(g^2) of the reduced, scaled gradient (mode 1)
  float *master, *m, *v;           // fp32 [shard_elems] (mode 2)
  float lr, beta1, beta2, eps, wd, inv_bc1, inv_sqrt_bc2;
  const float* clip_coef;          // device scalar (nullable)
  int mode;
};

template <int kMode>
__global__ void __launch_bounds__(256)
fused_rs_kernel(const RsParams p) {
  __shared__ float red[33];
  const int64_t nvec = p.shard_elems >> 3;
  const float clip = (kMode == 2 && p.clip_coef != nullptr) ? *p.clip_coef : 1.f;
  float ss = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = p.shard_offset + i * 8;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    // issue all peer loads first (NVLink latency ~2 us: keep world x 16 B in flight per thread)
    uint4 raw[kMaxWorld];
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r)
      if (r < p.world) raw[r] = ld_peer16(p.peer_grads[(r + p.rank) % p.world] + e);
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r) {
      if (r < p.world) {
        float f[8];
        unpack8(*reinterpret_cast<const bf16x8*>(&raw[r]), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= p.grad_scale;
    if (kMode <= 1) {
      float* o = p.grad_out + i * 8;
      *reinterpret_cast<float4*>(o) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      *reinterpret_cast<float4*>(o + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
      if (kMode == 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += acc[j] * acc[j];
      }
    } else {
      float w[8], mm[8], vv[8];
      const int64_t o = i * 8;
      *reinterpret_cast<float4*>(w) = *reinterpret_cast<const float4*>(p.master + o);
      *reinterpret_cast<float4*>(w + 4) = *reinterpret_cast<const float4*>(p.master + o + 4);
      *reinterpret_cast<float4*>(mm) = *reinterpret_cast<const float4*>(p.m + o);
      *reinterpret_cast<float4*>(mm + 4) = *reinterpret_cast<const float4*>(p.m + o + 4);
      *reinterpret_cast<float4*>(vv) = *reinterpret_cast<const float4*>(p.v + o);
      *reinterpret_cast<float4*>(vv + 4) = *reinterpret_cast<const float4*>(p.v + o + 4);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float g = acc[j] * clip;
        mm[j] = p.beta1 * mm[j] + (1.f - p.beta1) * g;
        vv[j] = p.beta2 * vv[j] + (1.f - p.beta2) * g * g;
        w[j] = w[j] * (1.f - p.lr * p.wd) - p.lr * p.inv_bc1 * (mm[j] / (sqrtf(vv[j]) * p.inv_sqrt_bc2 + p.eps));
      }
      *reinterpret_cast<float4*>(p.master + o) = *reinterpret_cast<float4*>(w);
      *reinterpret_cast<float4*>(p.master + o + 4) = *reinterpret_cast<float4*>(w + 4);
      *reinterpret_cast<float4*>(p.m + o) = *reinterpret_cast<float4*>(mm);
      *reinterpret_cast<float4*>(p.m + o + 4) = *reinterpret_cast<float4*>(mm + 4);
      *reinterpret_cast<float4*>(p.v + o) = *reinterpret_cast<float4*>(vv);
      *reinterpret_cast<float4*>(p.v + o + 4) = *reinterpret_cast<float4*>(vv + 4);
      if (p.peer_params != nullptr) {
        const bf16x8 packed = pack8(w);
        const uint4 u = *reinterpret_cast<const uint4*>(&packed);
#pragma unroll
        for (int r = 0; r < kMaxWorld; ++r)
          if (r < p.world) st_peer16(p.peer_params[(r + p.rank) % p.world] + e, u);
      }
    }
  }
  if (kMode == 1) {
    ss = block_sum(ss, red);
    if (threadIdx.x == 0) atomicAdd(p.sumsq_out, ss);
  }
}

// all-gather by stores: src[0..n) -> every peer's buffer at element offset dst_offset
__global__ void __launch_bounds__(256)
peer_push_kernel(const bf16* __restrict__ src, bf16* const* peers, int world, int rank, int64_t dst_offset, int64_t n) {
  const int64_t nvec = n >> 3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const uint4 u = *reinterpret_cast<const uint4*>(src + i * 8);
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r)
      if (r < world) st_peer16(peers[(r + rank) % world] + dst_offset + i * 8, u);
  }
}

// all-reduce by peer loads + peer stores (DDP gradients): this rank owns 1/world of [offset, offset+n):
// it sums that slice over all ranks in fp32, scales, and writes the bf16 result into EVERY rank's buffer.
__global__ void __launch_bounds__(256)
peer_allreduce_kernel(bf16* const* peers, int world, int rank, int64_t offset, int64_t n, float scale) {
  const int64_t slice = n / world;
  const int64_t base = offset + slice * rank;
  const int64_t nvec = slice >> 3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = base + i * 8;
    uint4 raw[kMaxWorld];
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r)
      if (r < world) raw[r] = ld_peer16(peers[(r + rank) % world] + e);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r) {
      if (r < world) {
        float f[8];
        unpack8(*reinterpret_cast<const bf16x8*>(&raw[r]), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= scale;
    const bf16x8 packed = pack8(acc);
    const uint4 u = *reinterpret_cast<const uint4*>(&packed);
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r)
      if (r < world) st_peer16(peers[(r + rank) % world] + e, u);
  }
}

// gather a flat range of a sharded unit: element e comes from rank e / S at offset e % S
__global__ void __launch_bounds__(256)
peer_gather_kernel(const bf16* const* peers, bf16* __restrict__ full, int64_t S, int64_t offset, int64_t n) {
  const int64_t nvec = n >> 3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = offset + i * 8;
    const int owner = (int)(e / S);
    const uint4 u = ld_peer16(peers[owner] + (e - (int64_t)owner * S));
    *reinterpret_cast<uint4*>(full + e) = u;
  }
}

// ------------------------------------------------------------------------------- (c) gather + DPO loss
__device__ __forceinline__ float log_sigmoid(float z) { return fminf(z, 0.f) - log1pf(__expf(-fabsf(z))); }
__device__ __forceinline__ float sigmoid_(float z) { return 1.f / (1.f + __expf(-z)); }

// peer_bufs[r]: fp32 [2][world][4][Bmax] on rank r; the half is chosen by the epoch parity so a rank that has already
// entered call k+1 writes the other half while a slower peer still reads call k (nobody can be two calls ahead: the
// flag exchange of call k+1 needs every rank to have finished call k).  Single block.
__global__ void __launch_bounds__(256)
fused_gather_dpo_kernel(float* const* peer_bufs, uint32_t* const* pads, int world, int me, const float* pp, const float* pn,
                        const float* rp, const float* rn, int B, int Bmax, float beta, float ls, uint32_t epoch, int channel,
                        unsigned long long timeout_ns, float* out, float* dpp, float* dpn, uint32_t* status) {
  __shared__ float red[33];
  __shared__ int s_ok;
  // 1. publish my [4, B] block into every peer's gather buffer (slot `me` of this epoch's half)
  const int64_t half = (int64_t)(epoch & 1u) * world * 4 * Bmax;
  for (int idx = threadIdx.x; idx < 4 * B; idx += blockDim.x) {
    const int which = idx / B, i = idx % B;
    const float v = which == 0 ? pp[i] : which == 1 ? pn[i] : which == 2 ? rp[i] : rn[i];
    for (int r = 0; r < world; ++r) {
      float* dst = peer_bufs[r] + half + ((int64_t)me * 4 + which) * Bmax + i;
      asm volatile("st.global.relaxed.sys.f32 [%0], %1;" ::"l"(dst), "f"(v) : "memory");
    }
  }
  __syncthreads();
  // 2. flag exchange: everyone's block has landed in my buffer once all flags show this epoch
  const bool ok = barrier_body(pads, world, me, channel, epoch, timeout_ns);
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (!ok) s_ok = 0;
  __syncthreads();
  if (!s_ok) {
    if (threadIdx.x == 0) note_timeout(status);
    return;
  }
  // 3. global statistics over world * B pairs, local gradients for my B pairs
  const float* mine = peer_bufs[me] + half;
  float loss = 0.f, pref = 0.f, margin = 0.f;
  const int total = world * B;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int r = idx / B, i = idx % B;
    const float* blk = mine + (int64_t)r * 4 * Bmax;
    float a, b, c, d;
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(a) : "l"(blk + i) : "memory");
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(b) : "l"(blk + Bmax + i) : "memory");
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(c) : "l"(blk + 2 * Bmax + i) : "memory");
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(d) : "l"(blk + 3 * Bmax + i) : "memory");
    const float z = beta * ((a - b) - (c - d));
    loss += -(1.f - ls) * log_sigmoid(z) - ls * log_sigmoid(-z);
    pref += a > b ? 1.f : 0.f;
    margin += z;
    if (r == me) {
      // gradient of the LOCAL-batch mean (the engine divides by world when it averages over ranks)
      const float dz = (-(1.f - ls) * sigmoid_(-z) + ls * sigmoid_(z)) * beta / (float)B;
      dpp[i] = dz;
      dpn[i] = -dz;
    }
  }
  loss = block_sum(loss, red);
  pref = block_sum(pref, red);
  margin = block_sum(margin, red);
  if (threadIdx.x == 0) {
    const float inv = 1.f / (float)total;
    out[0] = loss * inv;      // global mean loss
    out[1] = pref * inv;      // global preference rate
    out[2] = margin * inv;    // global mean margin
  }
}

// Tiny all-gather (eval losses, per-rank statistics): every rank stores its n floats into slot `me`
// of every peer's buffer, flags are exchanged, then the world x n block is copied out.  bufs[r]: fp32
// [2][world][nmax] on rank r — the half is chosen by the epoch parity, so a fast rank that already entered the
// next call writes the other half while a slow one still reads (nobody can be two calls ahead of a peer: the flag
// exchange of call k+1 needs every rank to have finished call k).  Single block.
__global__ void __launch_bounds__(256)
peer_allgather_small_kernel(float* const* bufs, uint32_t* const* pads, int world, int me, const float* src, int n, int nmax,
                            uint32_t epoch, int channel, unsigned long long timeout_ns, float* out) {
  __shared__ int s_ok;
  const int64_t half = (int64_t)(epoch & 1u) * world * nmax;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = src[i];
    for (int r = 0; r < world; ++r) {
      float* dst = bufs[r] + half + (int64_t)me * nmax + i;
      asm volatile("st.global.relaxed.sys.f32 [%0], %1;" ::"l"(dst), "f"(v) : "memory");
    }
  }
  __syncthreads();
  const bool ok = barrier_body(pads, world, me, channel, epoch, timeout_ns);
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (!ok) s_ok = 0;
  __syncthreads();
  if (!s_ok) {
    if (threadIdx.x == 0) note_timeout(nullptr);
    return;
  }
  const float* mine = bufs[me] + half;
  for (int idx = threadIdx.x; idx < world * n; idx += blockDim.x) {
    const int r = idx / n, i = idx - r * n;
    float v;
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(v) : "l"(mine + (int64_t)r * nmax + i) : "memory");
    out[idx] = v;
  }
}

// ------------------------------------------------------------------------------- (a) v2: per-unit RS in backward
// One unit's reduce-scatter, launched on a side stream as soon as THIS rank's bucket for the unit is final.  Instead of
// a global barrier the unit has its own row of flags in the signal pad: block 0 publishes "my bucket of unit u, step
// `epoch`, is complete" to every peer, every block waits until all W ranks have said so, then the rank's shard is
// summed — inside NVSwitch with one multimem.ld_reduce per 16 bytes when the arena has a multicast address (the
// NVLink receive side carries S elements instead of (W-1) x S), or by W peer loads otherwise.  Blocks wait only on
// flags written by OTHER GPUs, never on each other, so the grid need not be co-resident.
struct UnitRsParams {
  uint32_t* const* pads;          // [world] signal pads
  int flag_word;                  // word offset of this unit's [world] flags inside a pad
  const bf16* mc_grads;           // multicast address of the gradient arena (nullptr: P2P loads)
  const bf16* const* peer_grads;  // [world] unicast base pointers
  int world, rank;
  int64_t shard_elems, shard_offset;
  float scale;
  float* grad_out;                // fp32 [shard_elems]
  float* sumsq_out;               // += sum(out^2) (nullable)
  uint32_t epoch;
  unsigned long long timeout_ns;
  int accumulate;                 // grad_out += result (streamed units fold every micro-step into the fp32 shard)
};

template <bool kMC>
__global__ void __launch_bounds__(256)
unit_rs_kernel(const UnitRsParams p) {
  __shared__ float red[33];
  __shared__ int s_ok;
  const int t = threadIdx.x;
  if (blockIdx.x == 0 && t < p.world) {
    DLA_PEER_ASSERT(t != p.rank || ld_acquire_sys(p.pads[p.rank] + p.flag_word + p.rank) + 1u == p.epoch,
                    "unit flag epoch does not follow the previous one", ld_acquire_sys(p.pads[p.rank] + p.flag_word + p.rank), p.epoch);
    __threadfence_system();
    st_release_sys(p.pads[t] + p.flag_word + p.rank, p.epoch);
  }
  if (t == 0) s_ok = 1;
  __syncthreads();
  if (t < p.world && !wait_flag(p.pads[p.rank] + p.flag_word + t, p.epoch, p.timeout_ns)) s_ok = 0;
  __syncthreads();
  if (!s_ok) {
    if (t == 0) note_timeout(nullptr);
    return;
  }
  const int64_t nvec = p.shard_elems >> 3;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float ss = 0.f;
  if (kMC) {
    constexpr int kU = 4;       // 4 x 16 B switch reductions in flight per thread
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + t; i0 < nvec; i0 += stride * kU) {
      uint4 raw[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int64_t i = i0 + u * stride;
        if (i < nvec) raw[u] = multimem_ld_reduce_bf16x8(p.mc_grads + p.shard_offset + i * 8);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int64_t i = i0 + u * stride;
        if (i < nvec) {
          float f[8];
          unpack8(*reinterpret_cast<const bf16x8*>(&raw[u]), f);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            f[j] *= p.scale;
            ss += f[j] * f[j];
          }
          float* o = p.grad_out + i * 8;
          if (p.accumulate) {
            const float4 a0 = *reinterpret_cast<const float4*>(o), a1 = *reinterpret_cast<const float4*>(o + 4);
            f[0] += a0.x; f[1] += a0.y; f[2] += a0.z; f[3] += a0.w;
            f[4] += a1.x; f[5] += a1.y; f[6] += a1.z; f[7] += a1.w;
          }
          *reinterpret_cast<float4*>(o) = make_float4(f[0], f[1], f[2], f[3]);
          *reinterpret_cast<float4*>(o + 4) = make_float4(f[4], f[5], f[6], f[7]);
        }
      }
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + t; i < nvec; i += stride) {
      const int64_t e = p.shard_offset + i * 8;
      uint4 raw[kMaxWorld];
#pragma unroll
      for (int r = 0; r < kMaxWorld; ++r)
        if (r < p.world) raw[r] = ld_peer16(p.peer_grads[(r + p.rank) % p.world] + e);
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
      for (int r = 0; r < kMaxWorld; ++r) {
        if (r < p.world) {
          float f[8];
          unpack8(*reinterpret_cast<const bf16x8*>(&raw[r]), f);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        acc[j] *= p.scale;
        ss += acc[j] * acc[j];
      }
      float* o = p.grad_out + i * 8;
      if (p.accumulate) {
        const float4 a0 = *reinterpret_cast<const float4*>(o), a1 = *reinterpret_cast<const float4*>(o + 4);
        acc[0] += a0.x; acc[1] += a0.y; acc[2] += a0.z; acc[3] += a0.w;
        acc[4] += a1.x; acc[5] += a1.y; acc[6] += a1.z; acc[7] += a1.w;
      }
      *reinterpret_cast<float4*>(o) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      *reinterpret_cast<float4*>(o + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
  }
  if (p.sumsq_out != nullptr) {
    ss = block_sum(ss, red);
    if (t == 0) atomicAdd(p.sumsq_out, ss);
  }
}

// Grad-norm exchange + clip coefficient without an NCCL launch: every rank stores its partial
// sum(g^2) into slot `me` of every peer's [2][world] buffer (half = epoch parity, see peer_allgather_small_kernel),
// flags are exchanged, then each rank adds the W partials IN RANK ORDER (bitwise the same total everywhere) and derives
// norm and coefficient.  The local partial is reset to 0 for the next step's accumulation.  One warp.
__global__ void __launch_bounds__(32)
peer_norm_clip_kernel(float* const* bufs, uint32_t* const* pads, int world, int me, float* sumsq_local, float prescale_sq,
                      float max_norm, uint32_t epoch, int channel, unsigned long long timeout_ns, float* coef, float* norm_out) {
  const int t = threadIdx.x;
  const int half = (int)(epoch & 1u) * world;
  if (t < world) {
    const float v = *sumsq_local;
    asm volatile("st.global.relaxed.sys.f32 [%0], %1;" ::"l"(bufs[t] + half + me), "f"(v) : "memory");
  }
  __syncwarp();
  const bool ok = barrier_body(pads, world, me, channel, epoch, timeout_ns);
  const unsigned bad = __ballot_sync(0xffffffffu, !ok);
  if (bad != 0u) {
    if (t == 0) {
      note_timeout(nullptr);
      coef[0] = 0.f;          // a step whose norm is unknown must not be applied
    }
    return;
  }
  if (t == 0) {
    float tot = 0.f;
    for (int r = 0; r < world; ++r) {
      float v;
      asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(v) : "l"(bufs[me] + half + r) : "memory");
      tot += v;
    }
    const float norm = sqrtf(tot * prescale_sq);
    if (norm_out != nullptr) norm_out[0] = norm;
    float c = 1.f;
    if (max_norm > 0.f) c = fminf(1.f, max_norm / (norm + 1e-6f));
    if (!(norm == norm) || isinf(norm)) c = 0.f;
    coef[0] = c;
    *sumsq_local = 0.f;
  }
}

// all-gather by ONE multicast store per 16 bytes: src[0..n) -> element offset dst_offset of every rank's copy
__global__ void __launch_bounds__(256)
peer_push_mc_kernel(const bf16* __restrict__ src, bf16* mc_dst, int64_t dst_offset, int64_t n) {
  const int64_t nvec = n >> 3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    const uint4 u = *reinterpret_cast<const uint4*>(src + i * 8);
    multimem_st16(mc_dst + dst_offset + i * 8, u);
  }
}

// DDP all-reduce through the switch: this rank owns 1/world of the range, reads the sum of all copies with
// multimem.ld_reduce and writes it back into all copies with multimem.st
__global__ void __launch_bounds__(256)
peer_allreduce_mc_kernel(bf16* mc, int world, int rank, int64_t offset, int64_t n, float scale) {
  const int64_t slice = n / world;
  const int64_t base = offset + slice * rank;
  const int64_t nvec = slice >> 3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    uint4 raw = multimem_ld_reduce_bf16x8(mc + base + i * 8);
    if (scale != 1.f) {
      float f[8];
      unpack8(*reinterpret_cast<const bf16x8*>(&raw), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] *= scale;
      const bf16x8 pk = pack8(f);
      raw = *reinterpret_cast<const uint4*>(&pk);
    }
    multimem_st16(mc + base + i * 8, raw);
  }
}

// ------------------------------------------------------------------------------- (c) gather + REINFORCE / PPO
// Sequence-level REINFORCE (the reference's "PPO" update, R/src/training/train_rlhf.py:149-153) over the GLOBAL rollout
// batch: every rank publishes its [3, B] block (policy log-prob, reference log-prob, reward-model score), flags are
// exchanged, then each rank computes r_i = score_i - c * (pol_i - ref_i) for all world x B sequences, the global
// baseline mean(r), the logged global means of kl / r / score, and the gradient for its own B sequences.
// peer_bufs[r]: fp32 [2][world][3][Bmax] (half = epoch parity).  out = {loss(global mean), mean kl, mean r, mean score,
// baseline}.  Single block.
__global__ void __launch_bounds__(256)
fused_gather_reinforce_kernel(float* const* peer_bufs, uint32_t* const* pads, int world, int me, const float* pol,
                              const float* ref, const float* score, int B, int Bmax, float kl_coef, uint32_t epoch, int channel,
                              unsigned long long timeout_ns, float* out, float* dpol) {
  __shared__ float red[33];
  __shared__ int s_ok;
  const int64_t half = (int64_t)(epoch & 1u) * world * 3 * Bmax;
  for (int idx = threadIdx.x; idx < 3 * B; idx += blockDim.x) {
    const int which = idx / B, i = idx % B;
    const float v = which == 0 ? pol[i] : which == 1 ? ref[i] : score[i];
    for (int r = 0; r < world; ++r) {
      float* dst = peer_bufs[r] + half + ((int64_t)me * 3 + which) * Bmax + i;
      asm volatile("st.global.relaxed.sys.f32 [%0], %1;" ::"l"(dst), "f"(v) : "memory");
    }
  }
  __syncthreads();
  const bool ok = barrier_body(pads, world, me, channel, epoch, timeout_ns);
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (!ok) s_ok = 0;
  __syncthreads();
  if (!s_ok) {
    if (threadIdx.x == 0) note_timeout(nullptr);
    return;
  }
  const float* mine = peer_bufs[me] + half;
  const int total = world * B;
  float sr = 0.f, skl = 0.f, ssc = 0.f;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int r = idx / B, i = idx % B;
    const float* blk = mine + (int64_t)r * 3 * Bmax;
    float a, b, c;
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(a) : "l"(blk + i) : "memory");
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(b) : "l"(blk + Bmax + i) : "memory");
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(c) : "l"(blk + 2 * Bmax + i) : "memory");
    const float kl = a - b;
    sr += c - kl_coef * kl;
    skl += kl;
    ssc += c;
  }
  sr = block_sum(sr, red);
  skl = block_sum(skl, red);
  ssc = block_sum(ssc, red);
  const float inv = 1.f / (float)total;
  const float base = sr * inv;
  float loss = 0.f;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int r = idx / B, i = idx % B;
    const float* blk = mine + (int64_t)r * 3 * Bmax;
    float a, b, c;
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(a) : "l"(blk + i) : "memory");
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(b) : "l"(blk + Bmax + i) : "memory");
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(c) : "l"(blk + 2 * Bmax + i) : "memory");
    const float adv = (c - kl_coef * (a - b)) - base;
    loss += -adv * a;
    // gradient of the LOCAL-batch mean (the engine's divide-by-world makes it the global-mean gradient)
    if (r == me) dpol[i] = -adv / (float)B;
  }
  loss = block_sum(loss, red);
  if (threadIdx.x == 0) {
    out[0] = loss * inv;
    out[1] = skl * inv;
    out[2] = sr * inv;
    out[3] = ssc * inv;
    out[4] = base;
  }
}

// Advantage whitening + logged statistics over the GLOBAL rollout batch for clipped PPO: each rank publishes
// {sum(A m), sum(A^2 m), sum(m), stats[0..nstats)} (stats: per-rank SUMS the caller wants added up, e.g. kl, reward,
// score, and their count), flags are exchanged, every rank forms the global mean / variance and normalises its local
// advantages in place: A <- (A - mean) / max(std, 1e-6) * m.  stats_out[k] = sum over ranks of stats[k].
// bufs[r]: fp32 [2][world][kWhitenSlots].  Single block.
constexpr int kWhitenSlots = 16;
__global__ void __launch_bounds__(1024)
fused_gather_whiten_kernel(float* const* bufs, uint32_t* const* pads, int world, int me, float* adv, const float* mask, int n,
                           const float* stats, int nstats, uint32_t epoch, int channel, unsigned long long timeout_ns,
                           float* stats_out) {
  __shared__ float red[33];
  __shared__ float tot[kWhitenSlots];
  __shared__ int s_ok;
  float s1 = 0.f, s2 = 0.f, c = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float m = mask[i], a = adv[i];
    s1 += a * m;
    s2 += a * a * m;
    c += m;
  }
  s1 = block_sum(s1, red);
  s2 = block_sum(s2, red);
  c = block_sum(c, red);
  const int64_t half = (int64_t)(epoch & 1u) * world * kWhitenSlots;
  if (threadIdx.x < 3 + nstats) {
    const int k = threadIdx.x;
    const float v = k == 0 ? s1 : k == 1 ? s2 : k == 2 ? c : stats[k - 3];
    for (int r = 0; r < world; ++r) {
      float* dst = bufs[r] + half + (int64_t)me * kWhitenSlots + k;
      asm volatile("st.global.relaxed.sys.f32 [%0], %1;" ::"l"(dst), "f"(v) : "memory");
    }
  }
  __syncthreads();
  const bool ok = barrier_body(pads, world, me, channel, epoch, timeout_ns);
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (!ok) s_ok = 0;
  __syncthreads();
  if (!s_ok) {
    if (threadIdx.x == 0) note_timeout(nullptr);
    return;
  }
  if (threadIdx.x < 3 + nstats) {
    float acc = 0.f;
    for (int r = 0; r < world; ++r) {       // rank order: the same value on every rank
      float v;
      asm volatile("ld.global.relaxed.sys.f32 %0, [%1];"
                   : "=f"(v) : "l"(bufs[me] + half + (int64_t)r * kWhitenSlots + threadIdx.x) : "memory");
      acc += v;
    }
    tot[threadIdx.x] = acc;
    if (threadIdx.x >= 3) stats_out[threadIdx.x - 3] = acc;
  }
  __syncthreads();
  const float inv = 1.f / fmaxf(tot[2], 1.f);
  const float mean = tot[0] * inv;
  const float var = fmaxf(tot[1] * inv - mean * mean, 0.f);
  const float rstd = 1.f / fmaxf(sqrtf(var), 1e-6f);
  for (int i = threadIdx.x; i < n; i += blockDim.x) adv[i] = (adv[i] - mean) * rstd * mask[i];
}

static unsigned long long timeout_ns_of(int timeout_ms) {
  return timeout_ms > 0 ? (unsigned long long)timeout_ms * 1000000ULL : 0ULL;
}

}  // namespace peer
}  // namespace dla

using namespace dla;
using namespace dla::peer;

extern "C" {

int dla_peer_barrier(uint32_t* const* pads, int world, int rank, int channel, uint32_t epoch, int timeout_ms, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, DLA_E_BADARG, "peer_barrier: bad world/rank %d/%d", world, rank);
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pads, world, rank, channel, epoch, timeout_ns_of(timeout_ms),
                                                    nullptr);
  return check_launch("peer_barrier");
}

int dla_peer_status(int* timed_out) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  uint32_t v = 0;
  DLA_CUDA(cudaMemcpyFromSymbol(&v, g_peer_timeout_flag, sizeof(v)));
  *timed_out = (int)v;
  return 0;
}

int dla_fused_rs_adamw(const void* const* peer_grads, void* const* peer_params, int world, int rank, int64_t shard_elems,
                       int64_t shard_offset, float grad_scale, float* grad_out_fp32, float* sumsq_out, float* master, float* m,
                       float* v, float lr, float beta1, float beta2, float eps, float weight_decay, float bias_c1, float bias_c2,
                       const float* clip_coef_dev, int mode, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, DLA_E_BADARG, "fused_rs: bad world/rank %d/%d", world, rank);
  DLA_REQUIRE(shard_elems % 8 == 0 && shard_offset % 8 == 0, DLA_E_BADARG, "fused_rs: shard size/offset must be multiples of 8 elements");
  DLA_REQUIRE(mode >= 0 && mode <= 2, DLA_E_BADARG, "fused_rs: mode %d", mode);
  DLA_REQUIRE(mode == 2 ? (master && m && v && bias_c1 > 0 && bias_c2 > 0) : (grad_out_fp32 != nullptr), DLA_E_BADARG,
              "fused_rs: missing buffers for mode %d", mode);
  DLA_REQUIRE(mode != 1 || sumsq_out != nullptr, DLA_E_BADARG, "fused_rs: sumsq_out required in mode 1");
  if (shard_elems == 0) return 0;
  RsParams p{};
  p.peer_grads = reinterpret_cast<const bf16* const*>(peer_grads);
  p.peer_params = reinterpret_cast<bf16* const*>(peer_params);
  p.world = world; p.rank = rank; p.shard_elems = shard_elems; p.shard_offset = shard_offset; p.grad_scale = grad_scale;
  p.grad_out = grad_out_fp32; p.sumsq_out = sumsq_out; p.master = master; p.m = m; p.v = v;
  p.lr = lr; p.beta1 = beta1; p.beta2 = beta2; p.eps = eps; p.wd = weight_decay;
  if (mode == 2) { p.inv_bc1 = 1.f / bias_c1; p.inv_sqrt_bc2 = 1.f / sqrtf(bias_c2); }
  p.clip_coef = clip_coef_dev; p.mode = mode;
  int64_t blocks = ((shard_elems >> 3) + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 8) blocks = (int64_t)kNumSMs * 8;
  cudaStream_t s = (cudaStream_t)stream;
  if (mode == 0) fused_rs_kernel<0><<<(int)blocks, 256, 0, s>>>(p);
  else if (mode == 1) fused_rs_kernel<1><<<(int)blocks, 256, 0, s>>>(p);
  else fused_rs_kernel<2><<<(int)blocks, 256, 0, s>>>(p);
  return check_launch("fused_rs_adamw");
}

int dla_peer_push(const void* src, void* const* peers, int world, int rank, int64_t dst_offset, int64_t n, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && n % 8 == 0 && dst_offset % 8 == 0, DLA_E_BADARG, "peer_push: bad arguments");
  if (n == 0) return 0;
  int64_t blocks = ((n >> 3) + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 8) blocks = (int64_t)kNumSMs * 8;
  peer_push_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((const bf16*)src, reinterpret_cast<bf16* const*>(peers), world, rank, dst_offset, n);
  return check_launch("peer_push");
}

int dla_peer_allreduce(void* const* peers, int world, int rank, int64_t offset, int64_t n, float scale, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, DLA_E_BADARG, "peer_allreduce: bad world/rank");
  DLA_REQUIRE(n % (8 * world) == 0 && offset % 8 == 0, DLA_E_BADARG, "peer_allreduce: n must be a multiple of 8*world elements");
  if (n == 0) return 0;
  int64_t blocks = ((n / world >> 3) + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 8) blocks = (int64_t)kNumSMs * 8;
  if (blocks < 1) blocks = 1;
  peer_allreduce_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<bf16* const*>(peers), world, rank, offset, n, scale);
  return check_launch("peer_allreduce");
}

int dla_peer_gather(const void* const* peer_shards, void* unit_full, int64_t shard_elems, int64_t offset, int64_t n, int world,
                    void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && shard_elems % 8 == 0 && offset % 8 == 0 && n % 8 == 0, DLA_E_BADARG,
              "peer_gather: sizes and offsets must be multiples of 8 elements");
  if (n == 0) return 0;
  int64_t blocks = ((n >> 3) + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 8) blocks = (int64_t)kNumSMs * 8;
  peer_gather_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const bf16* const*>(peer_shards),
                                                                   reinterpret_cast<bf16*>(unit_full), shard_elems, offset, n);
  return check_launch("peer_gather");
}

int dla_fused_gather_dpo(float* const* peer_bufs, uint32_t* const* pads, int world, int rank, const float* pol_pos,
                         const float* pol_neg, const float* ref_pos, const float* ref_neg, int B, float beta,
                         float label_smoothing, uint32_t epoch, float* out, float* dpol_pos, float* dpol_neg, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world && B > 0 && B <= 1024, DLA_E_BADARG,
              "fused_gather_dpo: bad world/rank/B %d/%d/%d", world, rank, B);
  fused_gather_dpo_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(peer_bufs, pads, world, rank, pol_pos, pol_neg, ref_pos, ref_neg, B,
                                                              1024, beta, label_smoothing, epoch, /*channel=*/1,
                                                              timeout_ns_of(20000), out, dpol_pos, dpol_neg, nullptr);
  return check_launch("fused_gather_dpo");
}

int dla_peer_allgather_small(float* const* bufs, uint32_t* const* pads, int world, int rank, const float* src, int n, int nmax,
                             uint32_t epoch, int channel, float* out, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world && n > 0 && n <= nmax, DLA_E_BADARG,
              "peer_allgather_small: bad world/rank/n %d/%d/%d (nmax %d)", world, rank, n, nmax);
  peer_allgather_small_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(bufs, pads, world, rank, src, n, nmax, epoch, channel,
                                                                  timeout_ns_of(20000), out);
  return check_launch("peer_allgather_small");
}

int dla_peer_watch(void* host_mapped_word) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  uint32_t* hp = reinterpret_cast<uint32_t*>(host_mapped_word);
  DLA_CUDA(cudaMemcpyToSymbol(g_peer_status_host, &hp, sizeof(hp)));
  return 0;
}

int dla_peer_flag_row(uint32_t* const* pads, int world, int rank, int flag_word, uint32_t epoch, int mode, int timeout_ms,
                      void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world && flag_word >= 0 && (mode == 1 || mode == 2),
              DLA_E_BADARG, "peer_flag_row: bad arguments (world %d rank %d mode %d)", world, rank, mode);
  flag_row_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pads, world, rank, flag_word, epoch, mode, timeout_ns_of(timeout_ms));
  return check_launch("peer_flag_row");
}

int dla_unit_reduce(uint32_t* const* pads, int flag_word, const void* mc_grads, const void* const* peer_grads, int world, int rank,
                    int64_t shard_elems, int64_t shard_offset, float scale, float* grad_out_fp32, float* sumsq_out, uint32_t epoch,
                    int timeout_ms, int max_blocks, void* stream) {
  return dla_unit_reduce_acc(pads, flag_word, mc_grads, peer_grads, world, rank, shard_elems, shard_offset, scale, grad_out_fp32,
                             sumsq_out, epoch, timeout_ms, max_blocks, 0, stream);
}

int dla_unit_reduce_acc(uint32_t* const* pads, int flag_word, const void* mc_grads, const void* const* peer_grads, int world,
                        int rank, int64_t shard_elems, int64_t shard_offset, float scale, float* grad_out_fp32, float* sumsq_out,
                        uint32_t epoch, int timeout_ms, int max_blocks, int accumulate, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, DLA_E_BADARG, "unit_reduce: bad world/rank %d/%d", world, rank);
  DLA_REQUIRE(shard_elems % 8 == 0 && shard_offset % 8 == 0 && flag_word >= 0, DLA_E_BADARG,
              "unit_reduce: shard size/offset must be multiples of 8 elements");
  DLA_REQUIRE(grad_out_fp32 != nullptr && (mc_grads != nullptr || peer_grads != nullptr), DLA_E_BADARG, "unit_reduce: missing buffers");
  UnitRsParams p{};
  p.pads = pads; p.flag_word = flag_word; p.mc_grads = reinterpret_cast<const bf16*>(mc_grads);
  p.peer_grads = reinterpret_cast<const bf16* const*>(peer_grads);
  p.world = world; p.rank = rank; p.shard_elems = shard_elems; p.shard_offset = shard_offset; p.scale = scale;
  p.grad_out = grad_out_fp32; p.sumsq_out = sumsq_out; p.epoch = epoch; p.timeout_ns = timeout_ns_of(timeout_ms);
  p.accumulate = accumulate;
  const int per_thread = mc_grads != nullptr ? 4 : 1;
  int64_t blocks = ((shard_elems >> 3) + 256 * per_thread - 1) / (256 * per_thread);
  if (max_blocks <= 0) max_blocks = kNumSMs * 2;
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks < 1) blocks = 1;        // an empty shard still takes part in the flag exchange
  if (mc_grads != nullptr) unit_rs_kernel<true><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p);
  else unit_rs_kernel<false><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p);
  return check_launch("unit_reduce");
}

int dla_peer_norm_clip(float* const* bufs, uint32_t* const* pads, int world, int rank, float* sumsq_local, float prescale,
                       float max_norm, uint32_t epoch, int channel, float* coef, float* norm_out, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world && sumsq_local && coef, DLA_E_BADARG,
              "peer_norm_clip: bad arguments");
  peer_norm_clip_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(bufs, pads, world, rank, sumsq_local, prescale * prescale, max_norm, epoch,
                                                           channel, timeout_ns_of(30000), coef, norm_out);
  return check_launch("peer_norm_clip");
}

int dla_peer_push_mc(const void* src, void* mc_dst, int64_t dst_offset, int64_t n, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(mc_dst != nullptr && n % 8 == 0 && dst_offset % 8 == 0, DLA_E_BADARG, "peer_push_mc: bad arguments");
  if (n == 0) return 0;
  int64_t blocks = ((n >> 3) + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 4) blocks = (int64_t)kNumSMs * 4;
  peer_push_mc_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((const bf16*)src, (bf16*)mc_dst, dst_offset, n);
  return check_launch("peer_push_mc");
}

int dla_peer_allreduce_mc(void* mc, int world, int rank, int64_t offset, int64_t n, float scale, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(mc != nullptr && world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, DLA_E_BADARG, "peer_allreduce_mc: bad world/rank");
  DLA_REQUIRE(n % (8 * world) == 0 && offset % 8 == 0, DLA_E_BADARG, "peer_allreduce_mc: n must be a multiple of 8*world elements");
  if (n == 0) return 0;
  int64_t blocks = ((n / world >> 3) + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 4) blocks = (int64_t)kNumSMs * 4;
  if (blocks < 1) blocks = 1;
  peer_allreduce_mc_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((bf16*)mc, world, rank, offset, n, scale);
  return check_launch("peer_allreduce_mc");
}

int dla_fused_gather_reinforce(float* const* peer_bufs, uint32_t* const* pads, int world, int rank, const float* pol,
                               const float* ref, const float* score, int B, float kl_coef, uint32_t epoch, float* out, float* dpol,
                               void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world && B > 0 && B <= 1024, DLA_E_BADARG,
              "fused_gather_reinforce: bad world/rank/B %d/%d/%d", world, rank, B);
  fused_gather_reinforce_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(peer_bufs, pads, world, rank, pol, ref, score, B, 1024, kl_coef,
                                                                    epoch, /*channel=*/5, timeout_ns_of(20000), out, dpol);
  return check_launch("fused_gather_reinforce");
}

int dla_fused_gather_whiten(float* const* bufs, uint32_t* const* pads, int world, int rank, float* adv, const float* mask, int n,
                            const float* stats, int nstats, uint32_t epoch, float* stats_out, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world && n > 0, DLA_E_BADARG,
              "fused_gather_whiten: bad world/rank/n %d/%d/%d", world, rank, n);
  DLA_REQUIRE(nstats >= 0 && nstats + 3 <= kWhitenSlots && (nstats == 0 || (stats && stats_out)), DLA_E_BADARG,
              "fused_gather_whiten: at most %d statistics", kWhitenSlots - 3);
  fused_gather_whiten_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(bufs, pads, world, rank, adv, mask, n, stats, nstats, epoch,
                                                                  /*channel=*/6, timeout_ns_of(20000), stats_out);
  return check_launch("fused_gather_whiten");
}

}  // extern "C"
