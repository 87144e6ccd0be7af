// Memory-bound fused ops: RMSNorm/LayerNorm fwd+bwd (with residual add), RoPE, SwiGLU, GELU,
// bias helpers, embedding gather / scatter-add.  All bf16 I/O, fp32 math, 16-byte vector access.
// Replaces the eager ATen chains the reference runs through transformers.
#include "common.cuh"

namespace dla {

constexpr int kNormThreads = 256;
constexpr int kMaxVec = 8;               // 8 vectors x 8 elems x 256 threads = H <= 16384
constexpr int kNormBwdBlocks = kNumSMs * 4;

// ------------------------------------------------------------------------------- RMSNorm fwd
template <int VPT>
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ delta, const bf16* __restrict__ w,
                   bf16* __restrict__ resid_out, bf16* __restrict__ y, float* __restrict__ rstd,
                   int64_t M, int H, float eps) {
  // One block per row at a time, rows strided over the grid.  The raw 16-byte vectors of the NEXT row are requested
  // before the current row's reduction (two block-wide syncs) so the loads of one row overlap the reduce + store
  // phase of the previous one; without that the memory system idled during every reduction.
  __shared__ float red[33];
  const int nvec = H >> 3;
  bf16x8 nx[VPT], nd[VPT];
  auto fetch = [&](int64_t row) {
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        nx[i] = ld8_stream(x + row * H + c * 8);
        if (delta != nullptr) nd[i] = ld8_stream(delta + row * H + c * 8);
      }
    }
  };
  int64_t row = blockIdx.x;
  if (row < M) fetch(row);
  for (; row < M; row += gridDim.x) {
    float vals[VPT][8];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        unpack8(nx[i], vals[i]);
        if (delta != nullptr) {
          float d[8];
          unpack8(nd[i], d);
#pragma unroll
          for (int j = 0; j < 8; ++j) vals[i][j] += d[j];
          if (resid_out != nullptr) {
            bf16x8 r = pack8(vals[i]);
            st8(resid_out + row * H + c * 8, r);
            unpack8(r, vals[i]);  // normalise exactly what later kernels will re-read
          }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += vals[i][j] * vals[i][j];
      }
    }
    if (row + gridDim.x < M) fetch(row + gridDim.x);      // in flight during the reduction and the stores below
    ss = block_sum(ss, red);
    const float r = rsqrtf(ss / (float)H + eps);
    if (threadIdx.x == 0) rstd[row] = r;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        float wv[8], o[8];
        unpack8(ld8(w + c * 8), wv);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = vals[i][j] * r * wv[j];
        st8(y + row * H + c * 8, pack8(o));
      }
    }
  }
}

// ------------------------------------------------------------------------------- RMSNorm bwd
// dx = rstd * (dy*w - xhat * mean(dy*w*xhat)) [+ dresid];  partial dw per block -> scratch.
template <int VPT>
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, const bf16* __restrict__ w,
                   const float* __restrict__ rstd, const bf16* __restrict__ dresid,
                   bf16* __restrict__ dx, float* __restrict__ scratch, int64_t M, int H) {
  __shared__ float red[33];
  const int nvec = H >> 3;
  float dwacc[VPT][8];
  float wv[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int c = i * kNormThreads + threadIdx.x;
#pragma unroll
    for (int j = 0; j < 8; ++j) dwacc[i][j] = 0.f;
    if (c < nvec) unpack8(ld8(w + c * 8), wv[i]);
  }
  // raw vectors of the next row are requested before this row's reduction (same scheme as the forward)
  bf16x8 nx[VPT], ng[VPT], nr[VPT];
  auto fetch = [&](int64_t row) {
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        nx[i] = ld8_stream(x + row * H + c * 8);
        ng[i] = ld8_stream(dy + row * H + c * 8);
        if (dresid != nullptr) nr[i] = ld8_stream(dresid + row * H + c * 8);
      }
    }
  };
  int64_t row = blockIdx.x;
  if (row < M) fetch(row);
  for (; row < M; row += gridDim.x) {
    const float r = rstd[row];
    float xh[VPT][8], g[VPT][8], dres[VPT][8];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        unpack8(nx[i], xh[i]);
        unpack8(ng[i], g[i]);
        if (dresid != nullptr) unpack8(nr[i], dres[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] *= r;
          dwacc[i][j] += g[i][j] * xh[i][j];
          g[i][j] *= wv[i][j];
          dot += g[i][j] * xh[i][j];
        }
      }
    }
    if (row + gridDim.x < M) fetch(row + gridDim.x);
    dot = block_sum(dot, red) / (float)H;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = r * (g[i][j] - xh[i][j] * dot);
        if (dresid != nullptr) {
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += dres[i][j];
        }
        st8(dx + row * H + c * 8, pack8(o));
      }
    }
  }
  float* out = scratch + (int64_t)blockIdx.x * H;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int c = i * kNormThreads + threadIdx.x;
    if (c < nvec) {
      *reinterpret_cast<float4*>(out + c * 8) = make_float4(dwacc[i][0], dwacc[i][1], dwacc[i][2], dwacc[i][3]);
      *reinterpret_cast<float4*>(out + c * 8 + 4) = make_float4(dwacc[i][4], dwacc[i][5], dwacc[i][6], dwacc[i][7]);
    }
  }
}

// column sums of a [nrows, H] fp32 partial matrix -> bf16 vector (optionally accumulated)
__global__ void colsum_partials_kernel(const float* __restrict__ partials, int nrows, int H, int64_t pitch,
                                       bf16* __restrict__ out, int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= H) return;
  float s = 0.f;
  for (int r = 0; r < nrows; ++r) s += partials[(int64_t)r * pitch + c];
  if (accumulate) s += __bfloat162float(out[c]);
  out[c] = __float2bfloat16(s);
}

// ------------------------------------------------------------------------------- LayerNorm
template <int VPT>
__global__ void __launch_bounds__(kNormThreads)
layernorm_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, const bf16* __restrict__ b,
                     bf16* __restrict__ y, float* __restrict__ stats, int64_t M, int H, float eps) {
  __shared__ float red[33];
  const int nvec = H >> 3;
  for (int64_t row = blockIdx.x; row < M; row += gridDim.x) {
    float vals[VPT][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        unpack8(ld8_stream(x + row * H + c * 8), vals[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += vals[i][j];
      }
    }
    const float mean = block_sum(s, red) / (float)H;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          vals[i][j] -= mean;
          ss += vals[i][j] * vals[i][j];
        }
      }
    }
    const float r = rsqrtf(block_sum(ss, red) / (float)H + eps);
    if (threadIdx.x == 0) {
      stats[2 * row] = mean;
      stats[2 * row + 1] = r;
    }
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        float wv[8], bv[8], o[8];
        unpack8(ld8(w + c * 8), wv);
        if (b != nullptr) unpack8(ld8(b + c * 8), bv);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = vals[i][j] * r * wv[j] + (b != nullptr ? bv[j] : 0.f);
        st8(y + row * H + c * 8, pack8(o));
      }
    }
  }
}

template <int VPT>
__global__ void __launch_bounds__(kNormThreads)
layernorm_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, const bf16* __restrict__ w,
                     const float* __restrict__ stats, bf16* __restrict__ dx, float* __restrict__ scratch,
                     int64_t M, int H) {
  __shared__ float red[33];
  const int nvec = H >> 3;
  float dwacc[VPT][8], dbacc[VPT][8], wv[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int c = i * kNormThreads + threadIdx.x;
#pragma unroll
    for (int j = 0; j < 8; ++j) dwacc[i][j] = dbacc[i][j] = 0.f;
    if (c < nvec) unpack8(ld8(w + c * 8), wv[i]);
  }
  for (int64_t row = blockIdx.x; row < M; row += gridDim.x) {
    const float mean = stats[2 * row], r = stats[2 * row + 1];
    float xh[VPT][8], g[VPT][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        unpack8(ld8_stream(x + row * H + c * 8), xh[i]);
        unpack8(ld8_stream(dy + row * H + c * 8), g[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xh[i][j] - mean) * r;
          dwacc[i][j] += g[i][j] * xh[i][j];
          dbacc[i][j] += g[i][j];
          g[i][j] *= wv[i][j];
          s1 += g[i][j];
          s2 += g[i][j] * xh[i][j];
        }
      }
    }
    s1 = block_sum(s1, red) / (float)H;
    s2 = block_sum(s2, red) / (float)H;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = r * (g[i][j] - s1 - xh[i][j] * s2);
        st8(dx + row * H + c * 8, pack8(o));
      }
    }
  }
  float* out = scratch + (int64_t)blockIdx.x * 2 * H;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int c = i * kNormThreads + threadIdx.x;
    if (c < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        out[c * 8 + j] = dwacc[i][j];
        out[H + c * 8 + j] = dbacc[i][j];
      }
    }
  }
}

// ------------------------------------------------------------------------------- RoPE
// grid: x over (head, 8-pair vector) of one token row, y over token rows (grid-stride): 32-bit index math only
__global__ void __launch_bounds__(256)
rope_kernel(const bf16* src, bf16* x, const float* __restrict__ cos_t, const float* __restrict__ sin_t,
            const int32_t* __restrict__ positions, int64_t M, int n_heads, int n_copy_heads, int head_dim, int rot_dim,
            int64_t src_stride, int64_t row_stride, float sign) {
  // src == x: in place.  Otherwise x receives the rotated q/k heads and a verbatim copy of the n_copy_heads heads that
  // follow them (the v part of a packed projection), so the backward needs no clone of its incoming gradient.
  const int half = rot_dim >> 1;
  const int vec_per_head = half >> 3;  // 8 pairs per work item
  const int item = blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= (n_heads + n_copy_heads) * vec_per_head) return;
  const int h = item / vec_per_head, vi = item - h * vec_per_head;
  for (int64_t m = blockIdx.y; m < M; m += gridDim.y) {
    const bf16* in = src + m * src_stride + (int64_t)h * head_dim + vi * 8;
    bf16* base = x + m * row_stride + (int64_t)h * head_dim + vi * 8;
    if (h >= n_heads) {
      st8(base, ld8(in));
      st8(base + half, ld8(in + half));
      continue;
    }
    const int pos = positions[m];
    float a[8], b[8], c[8], s[8];
    unpack8(ld8(in), a);
    unpack8(ld8(in + half), b);
    const float4* cp = reinterpret_cast<const float4*>(cos_t + (int64_t)pos * half + vi * 8);
    const float4* sp = reinterpret_cast<const float4*>(sin_t + (int64_t)pos * half + vi * 8);
    *reinterpret_cast<float4*>(c) = cp[0];
    *reinterpret_cast<float4*>(c + 4) = cp[1];
    *reinterpret_cast<float4*>(s) = sp[0];
    *reinterpret_cast<float4*>(s + 4) = sp[1];
    float o1[8], o2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sj = s[j] * sign;
      o1[j] = a[j] * c[j] - b[j] * sj;
      o2[j] = b[j] * c[j] + a[j] * sj;
    }
    st8(base, pack8(o1));
    st8(base + half, pack8(o2));
  }
}

// ------------------------------------------------------------------------------- SwiGLU / GELU
// grid: x over 8-element vectors of a row, y over rows (grid-stride) — no 64-bit divisions per element; the
// sigmoid uses the fast reciprocal (the result is rounded to bf16 anyway)
__device__ __forceinline__ float fast_sigmoid(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }

__global__ void __launch_bounds__(256)
swiglu_fwd_kernel(const bf16* __restrict__ gu, bf16* __restrict__ h, int64_t M, int F) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (c >= F) return;
  for (int64_t m = blockIdx.y; m < M; m += gridDim.y) {
    float g[8], u[8], o[8];
    unpack8(ld8_stream(gu + m * 2 * F + c), g);
    unpack8(ld8_stream(gu + m * 2 * F + F + c), u);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = g[j] * fast_sigmoid(g[j]) * u[j];
    st8(h + m * F + c, pack8(o));
  }
}

__global__ void __launch_bounds__(256)
swiglu_bwd_kernel(const bf16* __restrict__ dh, const bf16* __restrict__ gu, bf16* __restrict__ dgu, int64_t M, int F) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (c >= F) return;
  for (int64_t m = blockIdx.y; m < M; m += gridDim.y) {
    float g[8], u[8], d[8], dg[8], du[8];
    unpack8(ld8_stream(gu + m * 2 * F + c), g);
    unpack8(ld8_stream(gu + m * 2 * F + F + c), u);
    unpack8(ld8_stream(dh + m * F + c), d);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sg = fast_sigmoid(g[j]);
      du[j] = d[j] * g[j] * sg;
      dg[j] = d[j] * u[j] * sg * (1.f + g[j] * (1.f - sg));
    }
    st8(dgu + m * 2 * F + c, pack8(dg));
    st8(dgu + m * 2 * F + F + c, pack8(du));
  }
}

__device__ __forceinline__ float gelu_tanh_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  return 0.5f * x * (1.f + tanhf(k0 * (x + k1 * x * x * x)));
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  const float t = tanhf(k0 * (x + k1 * x * x * x));
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * k0 * (1.f + 3.f * k1 * x * x);
}

__global__ void gelu_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ bias, bf16* __restrict__ y,
                                int64_t M, int F) {
  const int fv = F >> 3;
  const int64_t total = M * fv;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % fv) * 8;
    float v[8], b[8], o[8];
    unpack8(ld8_stream(x + idx * 8), v);
    if (bias != nullptr) unpack8(ld8(bias + c), b);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = gelu_tanh_f(v[j] + (bias != nullptr ? b[j] : 0.f));
    st8(y + idx * 8, pack8(o));
  }
}

__global__ void gelu_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, const bf16* __restrict__ bias,
                                bf16* __restrict__ dx, int64_t M, int F) {
  const int fv = F >> 3;
  const int64_t total = M * fv;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % fv) * 8;
    float v[8], b[8], g[8], o[8];
    unpack8(ld8_stream(x + idx * 8), v);
    unpack8(ld8_stream(dy + idx * 8), g);
    if (bias != nullptr) unpack8(ld8(bias + c), b);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = g[j] * gelu_tanh_grad(v[j] + (bias != nullptr ? b[j] : 0.f));
    st8(dx + idx * 8, pack8(o));
  }
}

__global__ void add_bias_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b, const bf16* __restrict__ bias,
                                bf16* __restrict__ y, int64_t M, int H) {
  const int hv = H >> 3;
  const int64_t total = M * hv;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % hv) * 8;
    float v[8], w[8], o[8];
    unpack8(ld8_stream(a + idx * 8), v);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = v[j];
    if (b != nullptr) {
      unpack8(ld8_stream(b + idx * 8), w);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += w[j];
    }
    if (bias != nullptr) {
      unpack8(ld8(bias + c), w);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += w[j];
    }
    st8(y + idx * 8, pack8(o));
  }
}

// per-block column partial sums of dy [M,H] -> scratch[block][H]
template <int VPT>
__global__ void __launch_bounds__(kNormThreads)
colsum_rows_kernel(const bf16* __restrict__ dy, float* __restrict__ scratch, int64_t M, int H) {
  const int nvec = H >> 3;
  float acc[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  for (int64_t row = blockIdx.x; row < M; row += gridDim.x) {
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = i * kNormThreads + threadIdx.x;
      if (c < nvec) {
        float g[8];
        unpack8(ld8_stream(dy + row * H + c * 8), g);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] += g[j];
      }
    }
  }
  float* out = scratch + (int64_t)blockIdx.x * H;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int c = i * kNormThreads + threadIdx.x;
    if (c < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) out[c * 8 + j] = acc[i][j];
    }
  }
}

// ------------------------------------------------------------------------------- embedding
__global__ void embedding_fwd_kernel(const int64_t* __restrict__ ids, const bf16* __restrict__ table,
                                     const bf16* __restrict__ pos_table, bf16* __restrict__ out,
                                     int64_t M, int H, int V, int T) {
  const int hv = H >> 3;
  const int64_t total = M * hv;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = idx / hv;
    const int c = (int)(idx % hv) * 8;
    int64_t id = ids[m];
    id = id < 0 ? 0 : (id >= V ? V - 1 : id);
    bf16x8 v = ld8(table + id * H + c);
    if (pos_table != nullptr) {
      float a[8], p[8];
      unpack8(v, a);
      unpack8(ld8(pos_table + (m % T) * H + c), p);
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] += p[j];
      v = pack8(a);
    }
    st8(out + m * H + c, v);
  }
}

__global__ void embedding_bwd_kernel(const int64_t* __restrict__ ids, const bf16* __restrict__ dout,
                                     bf16* __restrict__ dtable, bf16* __restrict__ dpos, int64_t M, int H,
                                     int V, int T) {
  const int hv = H >> 3;
  const int64_t total = M * hv;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = idx / hv;
    const int c = (int)(idx % hv) * 8;
    int64_t id = ids[m];
    if (id < 0 || id >= V) continue;
    const bf16x8 g = ld8_stream(dout + m * H + c);
    bf162* dst = reinterpret_cast<bf162*>(dtable + id * H + c);
#pragma unroll
    for (int j = 0; j < 4; ++j) atomicAdd(dst + j, g.v[j]);
    if (dpos != nullptr) {
      bf162* dp = reinterpret_cast<bf162*>(dpos + (m % T) * H + c);
#pragma unroll
      for (int j = 0; j < 4; ++j) atomicAdd(dp + j, g.v[j]);
    }
  }
}

static inline int ew_grid(int64_t work_items, int threads) {
  int64_t b = (work_items + threads - 1) / threads;
  const int64_t cap = (int64_t)kNumSMs * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace dla

using namespace dla;

#define DLA_VPT_DISPATCH(H, KERNEL, GRID, STREAM, ...)                                         \
  do {                                                                                         \
    const int nv_ = ((H) / 8 + kNormThreads - 1) / kNormThreads;                               \
    if (nv_ <= 1) KERNEL<1><<<GRID, kNormThreads, 0, STREAM>>>(__VA_ARGS__);                   \
    else if (nv_ <= 2) KERNEL<2><<<GRID, kNormThreads, 0, STREAM>>>(__VA_ARGS__);              \
    else if (nv_ <= 4) KERNEL<4><<<GRID, kNormThreads, 0, STREAM>>>(__VA_ARGS__);              \
    else KERNEL<8><<<GRID, kNormThreads, 0, STREAM>>>(__VA_ARGS__);                            \
  } while (0)

extern "C" {

int dla_norm_bwd_blocks(void) { return kNormBwdBlocks; }

int dla_rmsnorm_fwd(const void* x, const void* delta, const void* w, void* resid_out, void* y, float* rstd,
                    int64_t M, int H, float eps, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(H % 8 == 0 && H <= kMaxVec * 8 * kNormThreads, DLA_E_BADARG, "rmsnorm_fwd: H=%d must be a multiple of 8 and <= 16384", H);
  if (M == 0) return 0;
  const int grid = (int)(M < (int64_t)kNumSMs * 8 ? M : (int64_t)kNumSMs * 8);
  DLA_VPT_DISPATCH(H, rmsnorm_fwd_kernel, grid, (cudaStream_t)stream, (const bf16*)x, (const bf16*)delta, (const bf16*)w, (bf16*)resid_out, (bf16*)y, rstd, M, H, eps);
  return check_launch("rmsnorm_fwd");
}

int dla_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd, const void* dresid,
                    void* dx, void* dw, float* scratch, int accumulate_dw, int64_t M, int H, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(H % 8 == 0 && H <= kMaxVec * 8 * kNormThreads, DLA_E_BADARG, "rmsnorm_bwd: bad H=%d", H);
  if (M == 0) return 0;
  const int grid = (int)(M < kNormBwdBlocks ? M : kNormBwdBlocks);
  DLA_VPT_DISPATCH(H, rmsnorm_bwd_kernel, grid, (cudaStream_t)stream, (const bf16*)dy, (const bf16*)x, (const bf16*)w, rstd, (const bf16*)dresid, (bf16*)dx, scratch, M, H);
  int rc = check_launch("rmsnorm_bwd");
  if (rc) return rc;
  if (dw != nullptr) {
    colsum_partials_kernel<<<(H + 255) / 256, 256, 0, (cudaStream_t)stream>>>(scratch, grid, H, H, (bf16*)dw, accumulate_dw);
    rc = check_launch("rmsnorm_bwd.colsum");
  }
  return rc;
}

int dla_layernorm_fwd(const void* x, const void* w, const void* b, void* y, float* stats, int64_t M, int H,
                      float eps, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(H % 8 == 0 && H <= kMaxVec * 8 * kNormThreads, DLA_E_BADARG, "layernorm_fwd: bad H=%d", H);
  if (M == 0) return 0;
  const int grid = (int)(M < (int64_t)kNumSMs * 8 ? M : (int64_t)kNumSMs * 8);
  DLA_VPT_DISPATCH(H, layernorm_fwd_kernel, grid, (cudaStream_t)stream, (const bf16*)x, (const bf16*)w, (const bf16*)b, (bf16*)y, stats, M, H, eps);
  return check_launch("layernorm_fwd");
}

int dla_layernorm_bwd(const void* dy, const void* x, const void* w, const float* stats, void* dx, void* dw,
                      void* db, float* scratch, int accumulate, int64_t M, int H, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(H % 8 == 0 && H <= kMaxVec * 8 * kNormThreads, DLA_E_BADARG, "layernorm_bwd: bad H=%d", H);
  if (M == 0) return 0;
  const int grid = (int)(M < kNormBwdBlocks ? M : kNormBwdBlocks);
  DLA_VPT_DISPATCH(H, layernorm_bwd_kernel, grid, (cudaStream_t)stream, (const bf16*)dy, (const bf16*)x, (const bf16*)w, stats, (bf16*)dx, scratch, M, H);
  int rc = check_launch("layernorm_bwd");
  if (rc) return rc;
  if (dw != nullptr) {
    colsum_partials_kernel<<<(H + 255) / 256, 256, 0, (cudaStream_t)stream>>>(scratch, grid, H, 2 * (int64_t)H, (bf16*)dw, accumulate);
    if ((rc = check_launch("layernorm_bwd.dw"))) return rc;
  }
  if (db != nullptr) {
    colsum_partials_kernel<<<(H + 255) / 256, 256, 0, (cudaStream_t)stream>>>(scratch + H, grid, H, 2 * (int64_t)H, (bf16*)db, accumulate);
    rc = check_launch("layernorm_bwd.db");
  }
  return rc;
}

int dla_rope_inplace(void* x, const float* cos_t, const float* sin_t, const int32_t* positions, int64_t M,
                     int n_heads, int head_dim, int rot_dim, int64_t row_stride, float sign, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(rot_dim % 16 == 0 && rot_dim <= head_dim && head_dim % 8 == 0 && row_stride % 8 == 0, DLA_E_BADARG,
              "rope: rot_dim=%d must be a multiple of 16 and <= head_dim=%d", rot_dim, head_dim);
  const int64_t work = M * n_heads * (rot_dim / 16);
  if (work == 0) return 0;
  const int items = n_heads * (rot_dim / 16);
  const int threads = items >= 256 ? 256 : ((items + 31) / 32) * 32;
  const dim3 grid((items + threads - 1) / threads, (unsigned)(M < 32768 ? M : 32768));
  rope_kernel<<<grid, threads, 0, (cudaStream_t)stream>>>((const bf16*)x, (bf16*)x, cos_t, sin_t, positions, M, n_heads, 0, head_dim,
                                                          rot_dim, row_stride, row_stride, sign);
  return check_launch("rope");
}

int dla_rope_copy(const void* src, void* dst, const float* cos_t, const float* sin_t, const int32_t* positions, int64_t M,
                  int n_heads, int n_copy_heads, int head_dim, int64_t src_stride, int64_t dst_stride, float sign, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(head_dim % 16 == 0 && src_stride % 8 == 0 && dst_stride % 8 == 0 && src != dst, DLA_E_BADARG,
              "rope_copy: head_dim=%d must be a multiple of 16, pitches multiples of 8, src != dst", head_dim);
  if (M == 0 || n_heads + n_copy_heads == 0) return 0;
  const int items = (n_heads + n_copy_heads) * (head_dim / 16);
  const int threads = items >= 256 ? 256 : ((items + 31) / 32) * 32;
  const dim3 grid((items + threads - 1) / threads, (unsigned)(M < 32768 ? M : 32768));
  rope_kernel<<<grid, threads, 0, (cudaStream_t)stream>>>((const bf16*)src, (bf16*)dst, cos_t, sin_t, positions, M, n_heads,
                                                          n_copy_heads, head_dim, head_dim, src_stride, dst_stride, sign);
  return check_launch("rope_copy");
}

int dla_swiglu_fwd(const void* gu, void* h, int64_t M, int F, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(F % 8 == 0, DLA_E_BADARG, "swiglu: F=%d must be a multiple of 8", F);
  if (M == 0) return 0;
  const dim3 grid((F / 8 + 255) / 256, (unsigned)(M < 32768 ? M : 32768));
  swiglu_fwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)gu, (bf16*)h, M, F);
  return check_launch("swiglu_fwd");
}

int dla_swiglu_bwd(const void* dh, const void* gu, void* dgu, int64_t M, int F, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(F % 8 == 0, DLA_E_BADARG, "swiglu: F=%d must be a multiple of 8", F);
  if (M == 0) return 0;
  const dim3 grid((F / 8 + 255) / 256, (unsigned)(M < 32768 ? M : 32768));
  swiglu_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)dh, (const bf16*)gu, (bf16*)dgu, M, F);
  return check_launch("swiglu_bwd");
}

int dla_gelu_fwd(const void* x, const void* bias, void* y, int64_t M, int F, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(F % 8 == 0, DLA_E_BADARG, "gelu: F=%d must be a multiple of 8", F);
  if (M == 0) return 0;
  gelu_fwd_kernel<<<ew_grid(M * (F / 8), 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)x, (const bf16*)bias, (bf16*)y, M, F);
  return check_launch("gelu_fwd");
}

int dla_gelu_bwd(const void* dy, const void* x, const void* bias, void* dx, int64_t M, int F, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(F % 8 == 0, DLA_E_BADARG, "gelu: F=%d must be a multiple of 8", F);
  if (M == 0) return 0;
  gelu_bwd_kernel<<<ew_grid(M * (F / 8), 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)dy, (const bf16*)x, (const bf16*)bias, (bf16*)dx, M, F);
  return check_launch("gelu_bwd");
}

int dla_add_bias(const void* a, const void* b, const void* bias, void* y, int64_t M, int H, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(H % 8 == 0, DLA_E_BADARG, "add_bias: H=%d must be a multiple of 8", H);
  if (M == 0) return 0;
  add_bias_kernel<<<ew_grid(M * (H / 8), 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)a, (const bf16*)b, (const bf16*)bias, (bf16*)y, M, H);
  return check_launch("add_bias");
}

int dla_bias_grad(const void* dy, void* db, float* scratch, int accumulate, int64_t M, int H, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(H % 8 == 0 && H <= kMaxVec * 8 * kNormThreads, DLA_E_BADARG, "bias_grad: bad H=%d", H);
  if (M == 0) return 0;
  const int grid = (int)(M < kNormBwdBlocks ? M : kNormBwdBlocks);
  DLA_VPT_DISPATCH(H, colsum_rows_kernel, grid, (cudaStream_t)stream, (const bf16*)dy, scratch, M, H);
  int rc = check_launch("bias_grad");
  if (rc) return rc;
  colsum_partials_kernel<<<(H + 255) / 256, 256, 0, (cudaStream_t)stream>>>(scratch, grid, H, H, (bf16*)db, accumulate);
  return check_launch("bias_grad.colsum");
}

int dla_embedding_fwd(const int64_t* ids, const void* table, const void* pos_table, void* out, int64_t M, int H,
                      int V, int T, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(H % 8 == 0, DLA_E_BADARG, "embedding: H=%d must be a multiple of 8", H);
  if (M == 0) return 0;
  embedding_fwd_kernel<<<ew_grid(M * (H / 8), 256), 256, 0, (cudaStream_t)stream>>>(ids, (const bf16*)table, (const bf16*)pos_table, (bf16*)out, M, H, V, T < 1 ? 1 : T);
  return check_launch("embedding_fwd");
}

int dla_embedding_bwd(const int64_t* ids, const void* dout, void* dtable, void* dpos_table, int64_t M, int H,
                      int V, int T, void* stream) {
  if (int rc_ = ::dla::ensure_device()) return rc_;
  DLA_REQUIRE(H % 8 == 0, DLA_E_BADARG, "embedding: H=%d must be a multiple of 8", H);
  if (M == 0) return 0;
  embedding_bwd_kernel<<<ew_grid(M * (H / 8), 256), 256, 0, (cudaStream_t)stream>>>(ids, (const bf16*)dout, (bf16*)dtable, (bf16*)dpos_table, M, H, V, T < 1 ? 1 : T);
  return check_launch("embedding_bwd");
}

}  // extern "C"
