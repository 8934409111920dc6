// Shared device/host helpers for the sm_100a kernels of the PC-sampling hot path.
#pragma once
#include <cstdlib>
#include <utility>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#define B200_OK 0
#define B200_ERR_ARG (-1)
#define B200_ERR_UNSUPPORTED (-2)
#define B200_ERR_DRIVER (-3)
#define B200_ERR_TOO_LARGE (-4)   /* size beyond the capacity of a static-scratch entry point */

// Every launcher returns B200_OK or -(1000 + cudaError) so the host side can raise.
#define B200_RETURN_LAST()                                        \
  do {                                                            \
    cudaError_t e__ = cudaGetLastError();                         \
    return e__ == cudaSuccess ? B200_OK : -(1000 + (int)e__);     \
  } while (0)

#define B200_DT_F32 0
#define B200_DT_BF16 1

#define B200_NUM_SMS 148

namespace b200 {

typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// strict fp32 path: IEEE division and the accurate expf (the error of __expf = ex2.approx(x log2 e) grows with |x| and
// is visible against the fp64 arbiter of tests/test_score_gpu.py)
__device__ __forceinline__ float silu(float v) { return v / (1.0f + expf(-v)); }
// bf16 paths: MUFU.EX2 + MUFU.RCP (2 ulp) instead of the IEEE division -- the result is rounded to 8 bits anyway
#ifdef B200_SILU_EXP
__device__ __forceinline__ float silu_fast(float v) { return __fdividef(v, 1.0f + __expf(-v)); }
#else
// x sigmoid(x) = h + h tanh(h), h = x / 2: ONE MUFU op (tanh.approx.f32, relative error 2^-11) instead of EX2 + RCP; the
// GroupNorm apply pass issues 8 of these per 16 bytes and was at 55-65 % of the MUFU rate of the throttled clock
__device__ __forceinline__ float silu_fast(float v) {
  const float h = 0.5f * v;
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
  return fmaf(h, t, h);
}
#endif
template <typename T> __device__ __forceinline__ float silu_t(float v);
template <> __device__ __forceinline__ float silu_t<float>(float v) { return silu(v); }
template <> __device__ __forceinline__ float silu_t<__nv_bfloat16>(float v) { return silu_fast(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum; every thread gets the result.  `red` is >= 32 floats of shared memory.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.0f;
  r = warp_sum(r);
  return r;
}

// 8-wide vector of T (16 B for bf16, 32 B for fp32) as floats.
template <typename T> struct Vec8;
template <> struct Vec8<float> {
  struct raw_t { float4 a, b; };
  static __device__ __forceinline__ raw_t load_raw(const float* p) {
    raw_t r; r.a = *reinterpret_cast<const float4*>(p); r.b = *reinterpret_cast<const float4*>(p + 4); return r;
  }
  static __device__ __forceinline__ void unpack(const raw_t& r, float* v) {
    v[0] = r.a.x; v[1] = r.a.y; v[2] = r.a.z; v[3] = r.a.w; v[4] = r.b.x; v[5] = r.b.y; v[6] = r.b.z; v[7] = r.b.w;
  }
  static __device__ __forceinline__ void load(const float* p, float* v) {
    float4 a = *reinterpret_cast<const float4*>(p);
    float4 b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void store(float* p, const float* v) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
};
template <> struct Vec8<bf16> {
  typedef uint4 raw_t;                       // as loaded: conversion can wait until the value is needed (registers)
  static __device__ __forceinline__ raw_t load_raw(const bf16* p) { return *reinterpret_cast<const uint4*>(p); }
  static __device__ __forceinline__ void unpack(const raw_t& r, float* v) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
  static __device__ __forceinline__ void load(const bf16* p, float* v) {
    uint4 r = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
  static __device__ __forceinline__ void store(bf16* p, const float* v) {
    uint4 r;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = r;
  }
};

// ---------------- Philox4x32-10 (counter-based RNG for the fused noise path) ----------------
struct Philox {
  static __device__ __forceinline__ uint4 round10(uint4 ctr, uint2 key) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
      uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
      ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
      key.x += W0; key.y += W1;
    }
    return ctr;
  }
  // Four N(0,1) variates for (seed, stream offset, index of the 4-element group).
  static __device__ __forceinline__ float4 normal4(uint64_t seed, uint64_t offset, uint64_t group) {
    uint4 c = make_uint4((uint32_t)group, (uint32_t)(group >> 32), (uint32_t)offset, (uint32_t)(offset >> 32));
    uint2 k = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
    uint4 r = round10(c, k);
    // Box-Muller on (0,1] uniforms.
    const float S = 2.3283064365386963e-10f;  // 2^-32
    float u0 = ((float)r.x + 1.0f) * S, u1 = (float)r.y * S;
    float u2 = ((float)r.z + 1.0f) * S, u3 = (float)r.w * S;
    u0 = fminf(u0, 1.0f); u2 = fminf(u2, 1.0f);
    float ra = sqrtf(-2.0f * __logf(u0)), rb = sqrtf(-2.0f * __logf(u2));
    float sa, ca, sb, cb;
    __sincosf(6.283185307179586f * u1, &sa, &ca);
    __sincosf(6.283185307179586f * u3, &sb, &cb);
    return make_float4(ra * ca, ra * sa, rb * cb, rb * sb);
  }
};

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a PER-DEVICE property of a kernel: set it once per (kernel
// instantiation, device ordinal), not once per process -- a second engine on another GPU of the same process would
// otherwise fail its large-shared-memory launches.  `state` is the caller's static array (zero-initialised);
// entries: 0 = not set, 1 = set, < 0 = -(1000 + cudaError).  Racing threads both set the attribute (idempotent).
#define B200_MAX_DEVICES 64
template <typename K>
inline int b200_set_smem_once(int* state, K kern, int bytes) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= B200_MAX_DEVICES) return B200_ERR_ARG;
  if (state[dev] == 0) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    state[dev] = e == cudaSuccess ? 1 : -(1000 + (int)e);
  }
  return state[dev] == 1 ? B200_OK : state[dev];
}


// ---- programmatic dependent launch ------------------------------------------------------------------------------------
// A kernel launched through b200_launch_pdl may be scheduled while its predecessor in the stream is still draining
// (its CTAs take SMs as the predecessor's CTAs exit) and must call b200_pdl_wait() before it touches anything the
// predecessor wrote; everything before the wait (barrier init, TMEM allocation, descriptor prefetch, index math)
// overlaps the predecessor's tail and the launch latency.  The wait is a no-op under a plain launch.
// Off by default: measured on the C3 bench it changes nothing (64.0 vs 65.0 ms per PC step; the GPU sits at its power
// cap, so filling the 2 us launch gaps is paid back in clocks -- DESIGN.md "experiments").  B200_PDL=1 turns it on.
__device__ __forceinline__ void b200_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// Lets the NEXT kernel's CTAs be scheduled as soon as every CTA of this grid has got here (they still wait for this
// grid's completion at their own b200_pdl_wait): call it once the CTA holds all the resources it will ever take.
__device__ __forceinline__ void b200_pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
inline bool b200_pdl_enabled() {
  static const bool on = [] { const char* e = getenv("B200_PDL"); return e && e[0] == '1'; }();
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t b200_launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                   Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = b200_pdl_enabled() ? 1 : 0;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

}  // namespace b200
