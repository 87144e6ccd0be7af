// Shared device/host helpers for the dla_b200 sm_100a kernels.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "../include/dla_b200.h"

namespace dla {

void set_error(const char* fmt, ...);
extern std::atomic<long long> g_launches;
int ensure_device();

inline int check_launch(const char* what) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: launch failed: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

#define DLA_REQUIRE(cond, code, ...)   \
  do {                                 \
    if (!(cond)) {                     \
      ::dla::set_error(__VA_ARGS__);   \
      return (code);                   \
    }                                  \
  } while (0)

#define DLA_CUDA(call)                                                         \
  do {                                                                         \
    cudaError_t e_ = (call);                                                   \
    if (e_ != cudaSuccess) {                                                   \
      ::dla::set_error("%s failed: %s", #call, cudaGetErrorString(e_));        \
      return (int)e_;                                                          \
    }                                                                          \
  } while (0)

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

struct __align__(16) bf16x8 {
  bf162 v[4];
};

// 16-byte accesses go through uint4: a struct copy of four bf162 members is split by nvcc into four 4-byte
// LDG/STG (cuobjdump showed 676 STG.E vs 8 STG.E.128 in elementwise.o), which quadruples LSU wavefronts and
// L2 write requests.
__device__ __forceinline__ bf16x8 ld8(const bf16* p) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  bf16x8 r;
  *reinterpret_cast<uint4*>(&r) = u;
  return r;
}
__device__ __forceinline__ void st8(bf16* p, const bf16x8& v) {
  *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(&v);
}

// streaming (read-once) 16-byte load that does not pollute L1
__device__ __forceinline__ bf16x8 ld8_stream(const bf16* p) {
  bf16x8 r;
  uint32_t* u = reinterpret_cast<uint32_t*>(&r);
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3])
               : "l"(p));
  return r;
}

__device__ __forceinline__ void unpack8(const bf16x8& in, float* f) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(in.v[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ bf16x8 pack8(const float* f) {
  bf16x8 o;
#pragma unroll
  for (int i = 0; i < 4; ++i) o.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return o;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum; `red` is a shared float[33]. Every thread gets the result.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from a previous use
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float t = (lane < nw) ? red[lane] : 0.f;
  t = warp_sum(t);
  return t;
}

// ---- programmatic dependent launch (PDL) -----------------------------------------------------------------------
// A kernel launched with launch_pdl may become resident while its stream predecessor is still running; it must
// execute pdl_wait() before touching anything the predecessor produces (the wait returns once the predecessor grid
// has completed and flushed).  pdl_launch_dependents() lets the successor start its own pre-wait prologue early.
// Both are no-ops for kernels launched the ordinary way.  Used along the decode chain (one token = ~260 short
// dependent kernels), where launch latency and ramp-up would otherwise be paid between every pair.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace dla
