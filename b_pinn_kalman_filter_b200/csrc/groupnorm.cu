// GroupNorm (+ optional SiLU) on channels-last activations for sm_100a, split into
//   gn_partial : per-(sample, split, group) sum / sum-of-squares   (reads x once: 1 pass, warp/shared reductions)
//   gn_apply   : y = act((x - mean) * rstd * gamma + beta)          (reads x, writes y; folds the partials itself)
// Replaces torch's nn.GroupNorm(num_groups=min(C//4,32), eps=1e-6) + SiLU pairs of the score networks
// (models/layerspp.py:219,231,243,264; models/layers.py:618-642; models/ncsnpp.py:194-228).
// Both kernels accept TWO sources concatenated along C, so the skip-connection `torch.cat([h, hs.pop()], 1)`
// (models/ncsnpp.py:318) is never materialised before normalisation.
// HBM-bound: stats 1 read, apply 1 read + 1 write of the activation (2 B/elem bf16, 4 B/elem fp32).
#include "common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

struct GnArgs {
  const void* x1; const void* x2; int C1, C2;   // sources [B, HW, C1] and [B, HW, C2] (x2 may be null, C2 = 0)
  int B, HW, G, nsplit;
  float* partial;                               // [B, nsplit, G, 2]
  const float* gamma; const float* beta; float eps; int act;  // act: 0 none, 1 silu
  void* out;                                    // [B, HW, C1+C2]
  // alternative to `partial`: per-channel statistics written by the producing convolutions' epilogues
  // (conv_tc.cu `stats`): st_i = [B][T_i][C_i][2]
  const float* st1; const float* st2; int T1, T2;
};

template <typename T> struct AccT { using type = float; };
template <> struct AccT<float> { using type = double; };
__device__ __forceinline__ float gn_rsqrt(double v) { return (float)(1.0 / sqrt(v)); }
__device__ __forceinline__ float gn_rsqrt(float v) { return rsqrtf(v); }

// Thread layout shared by both kernels: `nthr` threads sweep pixels; nthr is a multiple of CV = C/8 so that each
// thread always owns the same 8 channels (cv = tid % CV) and steps over pixels.
template <typename T>
__global__ void __launch_bounds__(256) gn_partial_kernel(GnArgs a) {
  extern __shared__ float sh[];  // [ppt][2C] per-thread partials (fixed-order fold: bit-reproducible, no atomics)
  const int C = a.C1 + a.C2, CV = C >> 3;
  const int b = blockIdx.x / a.nsplit, sp = blockIdx.x % a.nsplit;
  const int ppt = blockDim.x / CV;              // pixels per sweep
  const int cv = threadIdx.x % CV, pl = threadIdx.x / CV;
  const int c0 = cv << 3;
  const int per = (a.HW + a.nsplit - 1) / a.nsplit;
  const int p_begin = sp * per, p_end = min(a.HW, p_begin + per);
  const T* src; int cs, cc;
  if (c0 < a.C1) { src = (const T*)a.x1; cs = a.C1; cc = c0; } else { src = (const T*)a.x2; cs = a.C2; cc = c0 - a.C1; }
  src += (size_t)b * a.HW * cs + cc;
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s[j] = 0.f; q[j] = 0.f; }
  int p = p_begin + pl;
  for (; p + 3 * ppt < p_end; p += 4 * ppt) {     // 4 independent 16/32-byte loads in flight per thread
    float v[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u) Vec8<T>::load(src + (size_t)(p + u * ppt) * cs, v[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int j = 0; j < 8; ++j) { s[j] += v[u][j]; q[j] = fmaf(v[u][j], v[u][j], q[j]); }
  }
  for (; p < p_end; p += ppt) {
    float v[8];
    Vec8<T>::load(src + (size_t)p * cs, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] += v[j]; q[j] = fmaf(v[j], v[j], q[j]); }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) { sh[pl * 2 * C + c0 + j] = s[j]; sh[pl * 2 * C + C + c0 + j] = q[j]; }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) {
    float t = 0.f;
    for (int r = 0; r < ppt; ++r) t += sh[r * 2 * C + i];
    sh[i] = t;   // row 0 is only read by this thread's own column here
  }
  __syncthreads();
  const int cpg = C / a.G;
  for (int g = threadIdx.x; g < a.G; g += blockDim.x) {
    float ts = 0.f, tq = 0.f;
    for (int j = 0; j < cpg; ++j) { ts += sh[g * cpg + j]; tq += sh[C + g * cpg + j]; }
    float* o = a.partial + (((size_t)b * a.nsplit + sp) * a.G + g) * 2;
    o[0] = ts; o[1] = tq;
  }
}

// MINB: resident blocks per SM the compiler must allow.  The kernel is a stream, and what it lacked was warps (ncu: 79
// registers = 3 blocks = 31 % warps active, 5.2 TB/s): the four loads in flight stay as loaded and are converted one at a
// time, which brings the bf16 kernel to 64 registers without a spill = 4 blocks per SM (32x32x128: 101.8 -> 96.3 us at
// 1024 images, 12.6 -> 11.3 us = 6.0 TB/s at 128).  A fifth block (48 registers) measured slightly worse again (98.2 us).
template <typename T, int MINB>
__global__ void __launch_bounds__(256, MINB) gn_apply_kernel(GnArgs a) {
  extern __shared__ float sh[];  // [G] mean, [G] rstd
  const int C = a.C1 + a.C2, CV = C >> 3;
  const int b = blockIdx.x / a.nsplit, sp = blockIdx.x % a.nsplit;
  const int cpg = C / a.G;
  b200_pdl_wait();
  b200_pdl_trigger();
  // Statistics are folded in fp64 on the strict fp32 path and in fp32 on the bf16 path: B200's fp64 rate is 1/64 of
  // fp32, and 2 (T + 1) C double additions + a double division and square root per group cost a block about as many
  // cycles as streaming its 64 KB share of a 16x16 activation (the sums are over <= 16 K values of O(1): fp32 leaves
  // ~1e-6 relative error on rstd, two orders below the bf16 rounding of the output).
  using acc_t = typename AccT<T>::type;
  acc_t* chs = reinterpret_cast<acc_t*>(sh + 2 * a.G);      // [C] sums, [C] sums of squares (stats path only)
  if (a.st1) {
    // fold the slots per channel with all threads (coalesced over channels), then the channels per group
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const bool first = c < a.C1;
      const float* base = first ? a.st1 + ((size_t)b * a.T1 * a.C1 + c) * 2
                                : a.st2 + ((size_t)b * a.T2 * a.C2 + (c - a.C1)) * 2;
      const int T = first ? a.T1 : a.T2, cs = first ? a.C1 : a.C2;
      acc_t ts = 0, tq = 0;
      for (int k = 0; k < T; ++k) {
        const float2 v = __ldg(reinterpret_cast<const float2*>(base + (size_t)k * cs * 2));
        ts += (acc_t)v.x; tq += (acc_t)v.y;
      }
      chs[c] = ts; chs[C + c] = tq;
    }
    __syncthreads();
  }
  for (int g = threadIdx.x; g < a.G; g += blockDim.x) {
    acc_t ts = 0, tq = 0;
    if (a.st1) {
      for (int c = g * cpg; c < (g + 1) * cpg; ++c) { ts += chs[c]; tq += chs[C + c]; }   // may straddle the sources
    } else
    for (int k = 0; k < a.nsplit; ++k) {
      const float* o = a.partial + (((size_t)b * a.nsplit + k) * a.G + g) * 2;
      ts += (acc_t)o[0]; tq += (acc_t)o[1];
    }
    const acc_t rn = (acc_t)1 / ((acc_t)a.HW * (acc_t)cpg);
    const acc_t mean = ts * rn;
    acc_t var = tq * rn - mean * mean;
    if (var < (acc_t)0) var = 0;
    sh[g] = (float)mean;
    sh[a.G + g] = gn_rsqrt(var + (acc_t)a.eps);
  }
  __syncthreads();
  const int ppt = blockDim.x / CV;
  const int cv = threadIdx.x % CV, pl = threadIdx.x / CV;
  const int c0 = cv << 3;
  float sc[8], sf[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c0 + j, g = c / cpg;
    const float ga = a.gamma ? a.gamma[c] : 1.0f, be = a.beta ? a.beta[c] : 0.0f;
    sc[j] = sh[a.G + g] * ga;
    sf[j] = be - sh[g] * sc[j];
  }
  const int per = (a.HW + a.nsplit - 1) / a.nsplit;
  const int p_begin = sp * per, p_end = min(a.HW, p_begin + per);
  const T* src; int cs, cc;
  if (c0 < a.C1) { src = (const T*)a.x1; cs = a.C1; cc = c0; } else { src = (const T*)a.x2; cs = a.C2; cc = c0 - a.C1; }
  src += (size_t)b * a.HW * cs + cc;
  T* dst = (T*)a.out + (size_t)b * a.HW * C + c0;
  int p = p_begin + pl;
  for (; p + 3 * ppt < p_end; p += 4 * ppt) {
    // four loads in flight, kept as loaded; each is converted only when its turn comes (32 fewer live registers)
    typename Vec8<T>::raw_t raw[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) raw[u] = Vec8<T>::load_raw(src + (size_t)(p + u * ppt) * cs);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float v[8];
      Vec8<T>::unpack(raw[u], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float y = fmaf(v[j], sc[j], sf[j]);
        v[j] = a.act ? silu_t<T>(y) : y;
      }
      Vec8<T>::store(dst + (size_t)(p + u * ppt) * C, v);
    }
  }
  for (; p < p_end; p += ppt) {
    float v[8];
    Vec8<T>::load(src + (size_t)p * cs, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float y = fmaf(v[j], sc[j], sf[j]);
      v[j] = a.act ? silu_t<T>(y) : y;
    }
    Vec8<T>::store(dst + (size_t)p * C, v);
  }
}

// Small images (HW <= 64: the 8x8 and 4x4 levels, 47 of the 75 applies of a C3 forward).  gn_apply_kernel gives each
// image a block that folds the statistics through shared memory behind two barriers and then streams 8-32 KB: four
// dependent round trips for a few hundred loads, 18-23 us per launch at B = 1024 where the bytes need 3-13.  Here a
// thread owns 8 channels x 8 pixels and derives the mean / rstd of the group(s) its channels belong to by itself -- the
// same sums in the same order as the block-wide fold, so the output is bit-identical -- from max(8, C / G) channel
// entries per slot, fetched as independent 16-byte loads: no shared memory, no barrier, two round trips,
// B * C / 8 * HW / 8 independent threads.  C / G in {4, 8, 16} (every shipped network); other shapes keep the block
// kernel.
// (scale, shift) of the 8 channels c0 .. c0 + 7 of image b from the producers' statistics, by one thread: the SPAN =
// max(8, C / G) channels around c0 (whole groups; SPAN | C1, so one source) arrive as SPAN / 2 independent 16-byte
// loads per slot; slots are summed in order per channel, channels in order per group -- the order of the block kernels.
template <typename T, int CPG>
__device__ __forceinline__ void gn_thread_coef(const GnArgs& a, int b, int c0, float* sc, float* sf) {
  using acc_t = typename AccT<T>::type;
  constexpr int SPAN = CPG > 8 ? CPG : 8;
  const int s0 = c0 / SPAN * SPAN;
  const bool first = s0 < a.C1;
  const float* base = first ? a.st1 + ((size_t)b * a.T1 * a.C1 + s0) * 2 : a.st2 + ((size_t)b * a.T2 * a.C2 + (s0 - a.C1)) * 2;
  const int nslot = first ? a.T1 : a.T2, scs = first ? a.C1 : a.C2;
  acc_t ts[SPAN], tq[SPAN];
#pragma unroll
  for (int i = 0; i < SPAN; ++i) { ts[i] = 0; tq[i] = 0; }
  for (int k = 0; k < nslot; ++k) {
    float4 v[SPAN / 2];
#pragma unroll
    for (int i = 0; i < SPAN / 2; ++i) v[i] = __ldg(reinterpret_cast<const float4*>(base + (size_t)k * scs * 2) + i);
#pragma unroll
    for (int i = 0; i < SPAN / 2; ++i) {
      ts[2 * i] += (acc_t)v[i].x; tq[2 * i] += (acc_t)v[i].y;
      ts[2 * i + 1] += (acc_t)v[i].z; tq[2 * i + 1] += (acc_t)v[i].w;
    }
  }
  float gam[8], bet[8];
  if (a.gamma) {
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(a.gamma + c0)), g1 = __ldg(reinterpret_cast<const float4*>(a.gamma + c0) + 1);
    gam[0] = g0.x; gam[1] = g0.y; gam[2] = g0.z; gam[3] = g0.w; gam[4] = g1.x; gam[5] = g1.y; gam[6] = g1.z; gam[7] = g1.w;
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) gam[j] = 1.0f;
  }
  if (a.beta) {
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(a.beta + c0)), b1 = __ldg(reinterpret_cast<const float4*>(a.beta + c0) + 1);
    bet[0] = b0.x; bet[1] = b0.y; bet[2] = b0.z; bet[3] = b0.w; bet[4] = b1.x; bet[5] = b1.y; bet[6] = b1.z; bet[7] = b1.w;
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) bet[j] = 0.0f;
  }
  const acc_t rn = (acc_t)1 / ((acc_t)a.HW * (acc_t)CPG);
#pragma unroll
  for (int gi = 0; gi < SPAN / CPG; ++gi) {
    acc_t gs = 0, gq = 0;
#pragma unroll
    for (int i = 0; i < CPG; ++i) { gs += ts[gi * CPG + i]; gq += tq[gi * CPG + i]; }
    const acc_t m = gs * rn;
    acc_t var = gq * rn - m * m;
    if (var < (acc_t)0) var = 0;
    const float mean = (float)m, rstd = gn_rsqrt(var + (acc_t)a.eps);
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if ((c0 - s0 + j) / CPG == gi) {
        sc[j] = rstd * gam[j];
        sf[j] = bet[j] - mean * sc[j];
      }
  }
}

template <typename T, int CPG>
__global__ void __launch_bounds__(256) gn_apply_small_kernel(GnArgs a, int total) {
  const int idx = blockIdx.x * 256 + threadIdx.x;
  b200_pdl_wait();
  b200_pdl_trigger();
  if (idx >= total) return;
  const int C = a.C1 + a.C2, CV = C >> 3;
  const int PS = (a.HW + 7) >> 3;                           // pixel strips of 8 per image
  const int cv = idx % CV, rest = idx / CV, ps = rest % PS, b = rest / PS;
  const int c0 = cv << 3;
  float sc[8], sf[8];
  gn_thread_coef<T, CPG>(a, b, c0, sc, sf);
  const T* src; int cs, cc;
  if (c0 < a.C1) { src = (const T*)a.x1; cs = a.C1; cc = c0; } else { src = (const T*)a.x2; cs = a.C2; cc = c0 - a.C1; }
  src += (size_t)b * a.HW * cs + cc;
  T* dst = (T*)a.out + (size_t)b * a.HW * C + c0;
  const int p0 = ps * 8, np = min(8, a.HW - p0);
  if (np == 8) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float v[4][8];
#pragma unroll
      for (int u = 0; u < 4; ++u) Vec8<T>::load(src + (size_t)(p0 + h * 4 + u) * cs, v[u]);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float y = fmaf(v[u][j], sc[j], sf[j]);
          v[u][j] = a.act ? silu_t<T>(y) : y;
        }
        Vec8<T>::store(dst + (size_t)(p0 + h * 4 + u) * C, v[u]);
      }
    }
  } else {
    for (int p = p0; p < p0 + np; ++p) {
      float v[8];
      Vec8<T>::load(src + (size_t)p * cs, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float y = fmaf(v[j], sc[j], sf[j]);
        v[j] = a.act ? silu_t<T>(y) : y;
      }
      Vec8<T>::store(dst + (size_t)p * C, v);
    }
  }
}

// GroupNorm coefficients for the GroupNorm-fused convolution (b200_conv2d_tc_gn): the same statistics fold as
// gn_apply_kernel<bf16> (fp32, slots then channels in a fixed order: identical mean / rstd), written per (image,
// channel) as (scale / 2, shift / 2) so that the consumer evaluates silu(x scale + shift) = h + h tanh(h) with
// h = fma(x, scale / 2, shift / 2) -- halving is exact, so the fused path reproduces gn_apply's output bit for bit.
__global__ void __launch_bounds__(256) gn_coef_kernel(GnArgs a, float* __restrict__ coef) {
  extern __shared__ float sh[];  // [G] mean, [G] rstd, [C] sums, [C] sums of squares
  const int C = a.C1 + a.C2, cpg = C / a.G;
  const int b = blockIdx.x;
  float* chs = sh + 2 * a.G;
  b200_pdl_wait();
  b200_pdl_trigger();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const bool first = c < a.C1;
    const float* base = first ? a.st1 + ((size_t)b * a.T1 * a.C1 + c) * 2
                              : a.st2 + ((size_t)b * a.T2 * a.C2 + (c - a.C1)) * 2;
    const int nslot = first ? a.T1 : a.T2, cs = first ? a.C1 : a.C2;
    float ts = 0, tq = 0;
    for (int k = 0; k < nslot; ++k) {
      const float2 v = __ldg(reinterpret_cast<const float2*>(base + (size_t)k * cs * 2));
      ts += v.x; tq += v.y;
    }
    chs[c] = ts; chs[C + c] = tq;
  }
  __syncthreads();
  for (int g = threadIdx.x; g < a.G; g += blockDim.x) {
    float ts = 0, tq = 0;
    for (int c = g * cpg; c < (g + 1) * cpg; ++c) { ts += chs[c]; tq += chs[C + c]; }
    const float rn = 1.0f / ((float)a.HW * (float)cpg);
    const float mean = ts * rn;
    float var = tq * rn - mean * mean;
    if (var < 0.0f) var = 0.0f;
    sh[g] = mean;
    sh[a.G + g] = gn_rsqrt(var + a.eps);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int g = c / cpg;
    const float ga = a.gamma ? a.gamma[c] : 1.0f, be = a.beta ? a.beta[c] : 0.0f;
    const float sc = sh[a.G + g] * ga;
    const float sf = be - sh[g] * sc;
    reinterpret_cast<float2*>(coef)[(size_t)b * C + c] = make_float2(0.5f * sc, 0.5f * sf);
  }
}

// gn_coef without the block: one thread per (image, 8 channels), for up to 16 statistics slots (images up to 32x32).
// Same sums in the same order -> the same bits; 11.6 -> 4 us per launch at B = 1024.
template <int CPG>
__global__ void __launch_bounds__(256) gn_coef_small_kernel(GnArgs a, float* __restrict__ coef, int total) {
  const int idx = blockIdx.x * 256 + threadIdx.x;
  b200_pdl_wait();
  b200_pdl_trigger();
  if (idx >= total) return;
  const int C = a.C1 + a.C2, CV = C >> 3;
  const int cv = idx % CV, b = idx / CV, c0 = cv << 3;
  float sc[8], sf[8];
  gn_thread_coef<bf16, CPG>(a, b, c0, sc, sf);
  float4* o = reinterpret_cast<float4*>(coef + ((size_t)b * C + c0) * 2);
#pragma unroll
  for (int i = 0; i < 4; ++i) o[i] = make_float4(0.5f * sc[2 * i], 0.5f * sf[2 * i], 0.5f * sc[2 * i + 1], 0.5f * sf[2 * i + 1]);
}

// out[b][c] = sum over the T slots of st[b][t][c] (sum, sum of squares): at 128x128 / 256x256 a producing convolution
// leaves 256 / 1024 slots per image (two per 8 x 16 pixel tile), and every block of gn_apply / gn_coef used to fold all
// of them for itself -- 1 MB of statistics per block and a 1024-step dependent sum in front of 0.5 MB of payload
// (CelebA-HQ 256, B = 8: 101 us per GroupNorm where the tensor streams in 45).  Folded ONCE here, in a fixed order:
// thread (lane l of 32, channel c) adds slots l, l + 32, ..., then the 32 partials are added in order.
__global__ void __launch_bounds__(512) gn_fold_kernel(const float* __restrict__ st, float* __restrict__ out, int T, int C) {
  __shared__ float2 part[32][16];
  const int cblocks = (C + 15) / 16;
  const int cl = threadIdx.x & 15, l = threadIdx.x >> 4;              // 16 channels (128 B per slot) x 32 slot lanes
  const int b = blockIdx.x / cblocks, c = (blockIdx.x % cblocks) * 16 + cl;
  float s = 0.f, q = 0.f;
  if (c < C) {
    const float2* base = reinterpret_cast<const float2*>(st) + (size_t)b * T * C + c;
    for (int t = l; t < T; t += 32) {
      const float2 v = __ldg(base + (size_t)t * C);
      s += v.x; q += v.y;
    }
  }
  part[l][cl] = make_float2(s, q);
  __syncthreads();
  if (l == 0 && c < C) {
    float2 a = part[0][cl];
#pragma unroll
    for (int i = 1; i < 32; ++i) { a.x += part[i][cl].x; a.y += part[i][cl].y; }
    reinterpret_cast<float2*>(out)[(size_t)b * C + c] = a;
  }
}

static int gn_threads(int C) {
  const int CV = C / 8;
  if (CV <= 0 || CV > 256) return 0;
  return (256 / CV) * CV;
}

}  // namespace b200

using namespace b200;

// b200_gn_apply_stats takes the thread-per-group kernel for HW <= 64 unless switched off (A/B runs, bit-identity tests)
static int g_gn_small = 1;
extern "C" int b200_gn_small_enable(int on) { const int was = g_gn_small; g_gn_small = on ? 1 : 0; return was; }

extern "C" int b200_gn_nsplit(int B, int HW) {
  // At most ONE wave of blocks (gn_apply_kernel<bf16>: 256 threads, up to 4 blocks per SM), never fewer than
  // 128 pixels per block: every block folds the image's statistics before it streams (up to 16 slots x C x 8 B), so
  // more, smaller blocks are slower, not faster -- measured inside a CUDA graph (profiles/prof_gn.py, NS_LIST): 128
  // images of 32x32x128: 23.9 us with 10 splits (the old "8 blocks per SM, several waves" rule), 11.4 us = 5.9 TB/s
  // with 4; 1024 images: 102.9 us with 2 splits, 98.0 us with 1.
  int want = (B200_NUM_SMS * 4) / (B > 0 ? B : 1);
  int maxs = HW / 128 > 0 ? HW / 128 : 1;
  int s = want < maxs ? want : maxs;
  if (s < 1) s = 1;
  if (s > 128) s = 128;       // (8 images of 256 x 256: 74 blocks per image fill the wave; 32 left half the SMs idle)
  return s;
}

static int gn_fill(GnArgs& a, const void* x1, int C1, const void* x2, int C2, int B, int HW, int G, int nsplit,
                   float* partial) {
  const int C = C1 + C2;
  if (!x1 || !partial || B < 0 || HW <= 0 || G <= 0 || nsplit <= 0 || C1 <= 0 || C2 < 0 || (C2 > 0 && !x2)) return B200_ERR_ARG;
  if (C1 % 8 || C2 % 8 || C % G || gn_threads(C) == 0) return B200_ERR_UNSUPPORTED;
  a.x1 = x1; a.x2 = x2; a.C1 = C1; a.C2 = C2; a.B = B; a.HW = HW; a.G = G; a.nsplit = nsplit; a.partial = partial;
  a.gamma = nullptr; a.beta = nullptr; a.eps = 0.f; a.act = 0; a.out = nullptr;
  a.st1 = nullptr; a.st2 = nullptr; a.T1 = 0; a.T2 = 0;
  return B200_OK;
}

extern "C" int b200_gn_partial(const void* x1, int C1, const void* x2, int C2, int dtype, int B, int HW, int G,
                               int nsplit, float* partial, cudaStream_t st) {
  GnArgs a;
  int rc = gn_fill(a, x1, C1, x2, C2, B, HW, G, nsplit, partial);
  if (rc || B == 0) return rc;
  const int C = C1 + C2, thr = gn_threads(C);
  const size_t smem = (size_t)(thr / (C / 8)) * 2 * C * sizeof(float);   // <= 256 * 16 * 4 = 16 KB
  if (dtype == B200_DT_F32) gn_partial_kernel<float><<<B * nsplit, thr, smem, st>>>(a);
  else if (dtype == B200_DT_BF16) gn_partial_kernel<bf16><<<B * nsplit, thr, smem, st>>>(a);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

extern "C" int b200_gn_apply(const void* x1, int C1, const void* x2, int C2, int dtype, int B, int HW, int G,
                             int nsplit, const float* partial, const float* gamma, const float* beta, float eps,
                             int act, void* out, cudaStream_t st) {
  GnArgs a;
  int rc = gn_fill(a, x1, C1, x2, C2, B, HW, G, nsplit, const_cast<float*>(partial));
  if (rc) return rc;
  if (!out) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  a.gamma = gamma; a.beta = beta; a.eps = eps; a.act = act; a.out = out;
  const int C = C1 + C2, thr = gn_threads(C);
  if (G % 2) return B200_ERR_UNSUPPORTED;              // keeps the double scratch 8-byte aligned
  const size_t smem = 2 * G * sizeof(float) + 2 * C * sizeof(double);
  cudaError_t le;
  if (dtype == B200_DT_F32) le = b200_launch_pdl(gn_apply_kernel<float, 3>, dim3(B * nsplit), dim3(thr), smem, st, a);
  else if (dtype == B200_DT_BF16) le = b200_launch_pdl(gn_apply_kernel<bf16, 4>, dim3(B * nsplit), dim3(thr), smem, st, a);
  else return B200_ERR_ARG;
  if (le != cudaSuccess) return -(1000 + (int)le);
  B200_RETURN_LAST();
}

// GroupNorm(+SiLU) apply from the per-channel statistics the producing tensor-core convolutions leave behind
// (b200_conv2d_tc `stats`): no statistics pass over the activation at all.
extern "C" int b200_gn_apply_stats(const void* x1, int C1, const float* st1, int T1, const void* x2, int C2,
                                   const float* st2, int T2, int dtype, int B, int HW, int G, int nsplit,
                                   const float* gamma, const float* beta, float eps, int act, void* out,
                                   cudaStream_t st) {
  GnArgs a;
  float dummy;
  int rc = gn_fill(a, x1, C1, x2, C2, B, HW, G, nsplit, &dummy);
  if (rc) return rc;
  if (!out || !st1 || T1 <= 0 || (C2 > 0 && (!st2 || T2 <= 0))) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  a.partial = nullptr; a.st1 = st1; a.st2 = st2; a.T1 = T1; a.T2 = T2;
  a.gamma = gamma; a.beta = beta; a.eps = eps; a.act = act; a.out = out;
  const int C = C1 + C2, thr = gn_threads(C);
  if (G % 2) return B200_ERR_UNSUPPORTED;              // keeps the double scratch 8-byte aligned
  cudaError_t le;
  const long long small_total = (long long)B * (C / 8) * ((HW + 7) / 8);
  const int cpg = C % G == 0 ? C / G : 0, span = cpg > 8 ? cpg : 8;
  if (g_gn_small && HW <= 64 && (cpg == 4 || cpg == 8 || cpg == 16) && C1 % span == 0 && C2 % span == 0 && T1 <= 4 &&
      (C2 == 0 || T2 <= 4) && small_total < (1ll << 31) && (dtype == B200_DT_F32 || dtype == B200_DT_BF16)) {
    const dim3 grid((unsigned)((small_total + 255) / 256)), blk(256);
    const int tot = (int)small_total;
#define B200_GN_SMALL(T)                                                                                   \
  (cpg == 4 ? b200_launch_pdl(gn_apply_small_kernel<T, 4>, grid, blk, 0, st, a, tot)                       \
            : cpg == 8 ? b200_launch_pdl(gn_apply_small_kernel<T, 8>, grid, blk, 0, st, a, tot)            \
                       : b200_launch_pdl(gn_apply_small_kernel<T, 16>, grid, blk, 0, st, a, tot))
    le = dtype == B200_DT_F32 ? B200_GN_SMALL(float) : B200_GN_SMALL(bf16);
#undef B200_GN_SMALL
    if (le != cudaSuccess) return -(1000 + (int)le);
    B200_RETURN_LAST();
  }
  const size_t smem = 2 * G * sizeof(float) + 2 * C * sizeof(double);
  if (dtype == B200_DT_F32) le = b200_launch_pdl(gn_apply_kernel<float, 3>, dim3(B * nsplit), dim3(thr), smem, st, a);
  else if (dtype == B200_DT_BF16) le = b200_launch_pdl(gn_apply_kernel<bf16, 4>, dim3(B * nsplit), dim3(thr), smem, st, a);
  else return B200_ERR_ARG;
  if (le != cudaSuccess) return -(1000 + (int)le);
  B200_RETURN_LAST();
}

// (scale / 2, shift / 2) per (image, channel) of GroupNorm over cat(x1, x2) from the producers' statistics by-products:
// coef = [B][C1 + C2][2] fp32, consumed by b200_conv2d_tc_gn (which applies GroupNorm + SiLU inside its operand
// pipeline).  bf16 activations only (the statistics are those of bf16 outputs).
extern "C" int b200_gn_coef(const float* st1, int T1, int C1, const float* st2, int T2, int C2, int B, int HW, int G,
                            const float* gamma, const float* beta, float eps, float* coef, cudaStream_t st) {
  const int C = C1 + C2;
  if (!st1 || !coef || T1 <= 0 || C1 <= 0 || C2 < 0 || (C2 > 0 && (!st2 || T2 <= 0)) || B < 0 || HW <= 0 || G <= 0)
    return B200_ERR_ARG;
  if (C1 % 8 || C2 % 8 || C % G) return B200_ERR_UNSUPPORTED;
  if (B == 0) return B200_OK;
  GnArgs a;
  a.x1 = nullptr; a.x2 = nullptr; a.C1 = C1; a.C2 = C2; a.B = B; a.HW = HW; a.G = G; a.nsplit = 1; a.partial = nullptr;
  a.gamma = gamma; a.beta = beta; a.eps = eps; a.act = 1; a.out = nullptr;
  a.st1 = st1; a.st2 = st2; a.T1 = T1; a.T2 = T2;
  cudaError_t le;
  const int cpg = C / G, span = cpg > 8 ? cpg : 8;
  const long long tot = (long long)B * (C / 8);
  if (g_gn_small && (cpg == 4 || cpg == 8 || cpg == 16) && C1 % span == 0 && C2 % span == 0 && T1 <= 16 &&
      (C2 == 0 || T2 <= 16) && tot < (1ll << 31)) {
    const dim3 grid((unsigned)((tot + 255) / 256)), blk(256);
    le = cpg == 4 ? b200_launch_pdl(gn_coef_small_kernel<4>, grid, blk, 0, st, a, coef, (int)tot)
                  : cpg == 8 ? b200_launch_pdl(gn_coef_small_kernel<8>, grid, blk, 0, st, a, coef, (int)tot)
                             : b200_launch_pdl(gn_coef_small_kernel<16>, grid, blk, 0, st, a, coef, (int)tot);
  } else {
    const size_t smem = (size_t)(2 * G + 2 * C) * sizeof(float);
    le = b200_launch_pdl(gn_coef_kernel, dim3(B), dim3(256), smem, st, a, coef);
  }
  if (le != cudaSuccess) return -(1000 + (int)le);
  B200_RETURN_LAST();
}

// Folds the per-tile statistics slots a producing convolution left behind, [B][T][C][2] -> [B][1][C][2] (fixed order), so
// that the GroupNorm consumers of large images (b200_gn_apply_stats / b200_gn_coef with T = 1) do not each re-fold them.
extern "C" int b200_gn_fold_stats(const float* stats, float* out, int B, int T, int C, cudaStream_t st) {
  if (!stats || !out || B < 0 || T <= 0 || C <= 0) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  gn_fold_kernel<<<B * ((C + 15) / 16), 512, 0, st>>>(stats, out, T, C);
  B200_RETURN_LAST();
}
