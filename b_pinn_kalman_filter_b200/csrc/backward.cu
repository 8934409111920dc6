// Input-gradient (vector-Jacobian product w.r.t. the image) pieces of the score networks for sm_100a: what DPS needs
// (inverse/conditional_sampling.py:116-136 differentiates logp(y | x0_hat(x_t)) through the whole U-Net w.r.t. x_t).
// Convolution backward-data reuses the forward convolution kernels with flipped / transposed weights (engine side);
// this file holds the rest:
//   gn_bwd_partial / gn_bwd_apply  GroupNorm(+SiLU) backward over one or two concatenated sources
//   axpby                          out = sa * a + sb * b          (residual branches, gradient accumulation)
//   nearest_up2_bwd                adjoint of nearest x2 upsample (models/layers.py:585)
//   zero_stuff2                    dy -> odd positions of a zero tensor: turns the stride-2 conv's backward-data into
//                                  a stride-1 "same" convolution (models/layers.py:596-608)
//   attention backward             dq, dk, dv of softmax(q k^T s) v     (models/layers.py:566-570)
// No weight gradients anywhere: DPS differentiates w.r.t. the input only.
#include "common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

struct GnBwdArgs {
  const void* x1; const void* x2; int C1, C2;
  int B, HW, G, nsplit;
  const float* part_fwd;        // [B, nsplit, G, 2]  sum, sum of squares of x     (b200_gn_partial)
  float* part_bwd;              // [B, nsplit, G, 2]  sum t, sum t * xhat          t = dz * gamma
  const float* gamma; const float* beta; float eps; int act;
  const void* dy;               // [B, HW, C1 + C2]
  void* dx1; void* dx2;         // gradients of the sources (dx2 null when C2 == 0)
  int acc1, acc2;               // 1: add to what dx already holds
  float* chan_sums;             // optional [B, nsplit, C, 2]: (sum dz xhat, sum dz) per channel = the pieces of d gamma, d beta
};

__device__ __forceinline__ void gn_bwd_stats(const GnBwdArgs& a, int b, float* sh) {
  const int C = a.C1 + a.C2, cpg = C / a.G;
  for (int g = threadIdx.x; g < a.G; g += blockDim.x) {
    double ts = 0.0, tq = 0.0;
    for (int k = 0; k < a.nsplit; ++k) {
      const float* o = a.part_fwd + (((size_t)b * a.nsplit + k) * a.G + g) * 2;
      ts += (double)o[0]; tq += (double)o[1];
    }
    const double n = (double)a.HW * cpg;
    const double mean = ts / n;
    double var = tq / n - mean * mean;
    if (var < 0.0) var = 0.0;
    sh[g] = (float)mean;
    sh[a.G + g] = (float)(1.0 / sqrt(var + (double)a.eps));
  }
}

// dz = d(loss)/d(z) = dy * act'(z), with z = gamma * xhat + beta
__device__ __forceinline__ float gn_bwd_dz(float dy, float xhat, float ga, float be, int act) {
  if (act) {
    const float z = fmaf(ga, xhat, be);
    const float sg = 1.0f / (1.0f + __expf(-z));
    dy *= sg * (1.0f + z * (1.0f - sg));
  }
  return dy;
}
// t = d(loss)/d(xhat) = dz * gamma
__device__ __forceinline__ float gn_bwd_t(float dy, float xhat, float ga, float be, int act) {
  return gn_bwd_dz(dy, xhat, ga, be, act) * ga;
}

template <typename T>
__global__ void __launch_bounds__(256) gn_bwd_partial_kernel(GnBwdArgs a) {
  extern __shared__ float sh[];   // [2G] mean, rstd | [ppt][2C] per-thread partials
  const int C = a.C1 + a.C2, CV = C >> 3, cpg = C / a.G;
  const int b = blockIdx.x / a.nsplit, sp = blockIdx.x % a.nsplit;
  gn_bwd_stats(a, b, sh);
  __syncthreads();
  float* acc = sh + 2 * a.G;
  const int ppt = blockDim.x / CV, cv = threadIdx.x % CV, pl = threadIdx.x / CV, c0 = cv << 3;
  const int per = (a.HW + a.nsplit - 1) / a.nsplit;
  const int p_begin = sp * per, p_end = min(a.HW, p_begin + per);
  const T* src; int cs, cc;
  if (c0 < a.C1) { src = (const T*)a.x1; cs = a.C1; cc = c0; } else { src = (const T*)a.x2; cs = a.C2; cc = c0 - a.C1; }
  src += (size_t)b * a.HW * cs + cc;
  const T* dy = (const T*)a.dy + (size_t)b * a.HW * C + c0;
  float mu[8], rs[8], ga[8], be[8], s1[8], s2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c0 + j, g = c / cpg;
    mu[j] = sh[g]; rs[j] = sh[a.G + g];
    ga[j] = a.gamma ? a.gamma[c] : 1.0f; be[j] = a.beta ? a.beta[c] : 0.0f;
    s1[j] = 0.f; s2[j] = 0.f;
  }
  // two pixels per trip, all four loads requested before any is used: at 92 registers only two blocks fit an SM, and with
  // one pixel per trip (two loads in flight per thread) the sweep ran at half the HBM rate of the forward apply
  int p = p_begin + pl;
  for (; p + ppt < p_end; p += 2 * ppt) {
    float v[2][8], d[2][8];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      Vec8<T>::load(src + (size_t)(p + u * ppt) * cs, v[u]);
      Vec8<T>::load(dy + (size_t)(p + u * ppt) * C, d[u]);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xh = (v[u][j] - mu[j]) * rs[j];
        const float dz = gn_bwd_dz(d[u][j], xh, ga[j], be[j], a.act);   // per-channel sums of dz: gamma comes in at the fold
        s1[j] += dz; s2[j] = fmaf(dz, xh, s2[j]);
      }
  }
  for (; p < p_end; p += ppt) {
    float v[8], d[8];
    Vec8<T>::load(src + (size_t)p * cs, v);
    Vec8<T>::load(dy + (size_t)p * C, d);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float xh = (v[j] - mu[j]) * rs[j];
      const float dz = gn_bwd_dz(d[j], xh, ga[j], be[j], a.act);
      s1[j] += dz; s2[j] = fmaf(dz, xh, s2[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) { acc[pl * 2 * C + c0 + j] = s1[j]; acc[pl * 2 * C + C + c0 + j] = s2[j]; }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) {
    float t = 0.f;
    for (int r = 0; r < ppt; ++r) t += acc[r * 2 * C + i];
    acc[i] = t;
  }
  __syncthreads();
  if (a.chan_sums) {
    // the same sums are the block's share of d beta (sum dz) and d gamma (sum dz xhat): training plans take them from
    // here instead of a second sweep over x and dy (b200_gn_param_grad)
    float* cs_out = a.chan_sums + ((size_t)b * a.nsplit + sp) * C * 2;
    for (int c = threadIdx.x; c < C; c += blockDim.x) { cs_out[2 * c] = acc[C + c]; cs_out[2 * c + 1] = acc[c]; }
  }
  for (int g = threadIdx.x; g < a.G; g += blockDim.x) {
    float t1 = 0.f, t2 = 0.f;
    for (int j = 0; j < cpg; ++j) {
      const int c = g * cpg + j;
      const float gc = a.gamma ? a.gamma[c] : 1.0f;
      t1 = fmaf(gc, acc[c], t1); t2 = fmaf(gc, acc[C + c], t2);
    }
    float* o = a.part_bwd + (((size_t)b * a.nsplit + sp) * a.G + g) * 2;
    o[0] = t1; o[1] = t2;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) gn_bwd_apply_kernel(GnBwdArgs a) {
  extern __shared__ float sh[];   // [2G] mean, rstd | [2G] m1, m2
  const int C = a.C1 + a.C2, CV = C >> 3, cpg = C / a.G;
  const int b = blockIdx.x / a.nsplit, sp = blockIdx.x % a.nsplit;
  gn_bwd_stats(a, b, sh);
  for (int g = threadIdx.x; g < a.G; g += blockDim.x) {
    double t1 = 0.0, t2 = 0.0;
    for (int k = 0; k < a.nsplit; ++k) {
      const float* o = a.part_bwd + (((size_t)b * a.nsplit + k) * a.G + g) * 2;
      t1 += (double)o[0]; t2 += (double)o[1];
    }
    const double n = (double)a.HW * cpg;
    sh[2 * a.G + g] = (float)(t1 / n);
    sh[3 * a.G + g] = (float)(t2 / n);
  }
  __syncthreads();
  const int ppt = blockDim.x / CV, cv = threadIdx.x % CV, pl = threadIdx.x / CV, c0 = cv << 3;
  const int per = (a.HW + a.nsplit - 1) / a.nsplit;
  const int p_begin = sp * per, p_end = min(a.HW, p_begin + per);
  const T* src; T* dst; int cs, cc, accum;
  if (c0 < a.C1) { src = (const T*)a.x1; dst = (T*)a.dx1; cs = a.C1; cc = c0; accum = a.acc1; }
  else { src = (const T*)a.x2; dst = (T*)a.dx2; cs = a.C2; cc = c0 - a.C1; accum = a.acc2; }
  if (!dst) return;
  src += (size_t)b * a.HW * cs + cc;
  dst += (size_t)b * a.HW * cs + cc;
  const T* dy = (const T*)a.dy + (size_t)b * a.HW * C + c0;
  float mu[8], rs[8], ga[8], be[8], m1[8], m2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c0 + j, g = c / cpg;
    mu[j] = sh[g]; rs[j] = sh[a.G + g]; m1[j] = sh[2 * a.G + g]; m2[j] = sh[3 * a.G + g];
    ga[j] = a.gamma ? a.gamma[c] : 1.0f; be[j] = a.beta ? a.beta[c] : 0.0f;
  }
  int p = p_begin + pl;
  for (; p + ppt < p_end; p += 2 * ppt) {          // two pixels per trip: every load is in flight before the first use
    float v[2][8], d[2][8], o[2][8];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      Vec8<T>::load(src + (size_t)(p + u * ppt) * cs, v[u]);
      Vec8<T>::load(dy + (size_t)(p + u * ppt) * C, d[u]);
      if (accum) Vec8<T>::load(dst + (size_t)(p + u * ppt) * cs, o[u]);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xh = (v[u][j] - mu[j]) * rs[j];
        const float t = gn_bwd_t(d[u][j], xh, ga[j], be[j], a.act);
        const float g = rs[j] * (t - m1[j] - xh * m2[j]);
        o[u][j] = accum ? o[u][j] + g : g;
      }
      Vec8<T>::store(dst + (size_t)(p + u * ppt) * cs, o[u]);
    }
  }
  for (; p < p_end; p += ppt) {
    float v[8], d[8], o[8];
    Vec8<T>::load(src + (size_t)p * cs, v);
    Vec8<T>::load(dy + (size_t)p * C, d);
    if (accum) Vec8<T>::load(dst + (size_t)p * cs, o);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float xh = (v[j] - mu[j]) * rs[j];
      const float t = gn_bwd_t(d[j], xh, ga[j], be[j], a.act);
      const float g = rs[j] * (t - m1[j] - xh * m2[j]);
      o[j] = accum ? o[j] + g : g;
    }
    Vec8<T>::store(dst + (size_t)p * cs, o);
  }
}

static int gn_bwd_threads(int C) {
  const int CV = C / 8;
  if (CV <= 0 || CV > 256) return 0;
  return (256 / CV) * CV;
}

template <typename T>
__global__ void axpby_kernel(const T* __restrict__ a, float sa, const T* __restrict__ b, float sb, T* __restrict__ out,
                             long long nvec) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float x[8], y[8];
    Vec8<T>::load(a + i * 8, x);
    if (b) {
      Vec8<T>::load(b + i * 8, y);
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] = fmaf(sa, x[j], sb * y[j]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] *= sa;
    }
    Vec8<T>::store(out + i * 8, x);
  }
}

// dx[b, h, w, :] (+)= sum of the 2x2 block of dy
template <typename T>
__global__ void nearest_up2_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int B, int H, int W, int C8, int acc) {
  const long long total = (long long)B * H * W * C8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % C8);
    long long r = i / C8;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H);
    const long long b = r / H;
    float s[8], v[8];
    if (acc) Vec8<T>::load(dx + i * 8, s);
    else {
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] = 0.f;
    }
#pragma unroll
    for (int dh = 0; dh < 2; ++dh)
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        Vec8<T>::load(dy + ((((b * 2 * H) + 2 * h + dh) * (2 * W) + 2 * w + dw) * C8 + c8) * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) s[j] += v[j];
      }
    Vec8<T>::store(dx + i * 8, s);
  }
}

// out [B, Ho, Wo, C] = 0 except out[b, 2i+1, 2j+1, :] = dy[b, i, j, :]
template <typename T>
__global__ void zero_stuff2_kernel(const T* __restrict__ dy, T* __restrict__ out, int B, int Hi, int Wi, int Ho, int Wo,
                                   int C8) {
  const long long total = (long long)B * Ho * Wo * C8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % C8);
    long long r = i / C8;
    const int w = (int)(r % Wo); r /= Wo;
    const int h = (int)(r % Ho);
    const long long b = r / Ho;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
    if ((h & 1) && (w & 1) && (h >> 1) < Hi && (w >> 1) < Wi)
      Vec8<T>::load(dy + (((b * Hi + (h >> 1)) * Wi + (w >> 1)) * C8 + c8) * 8, v);
    Vec8<T>::store(out + i * 8, v);
  }
}

// ---- attention backward (one warp per row; fp32 math; P and dS kept in a global scratch between the two passes) ----
constexpr int AB_ROWS = 8;

// pass 1, per query row i: p = softmax(s q_i K^T), dP_j = dO_i . v_j, dS = p (dP - sum p dP), dq_i = s dS K
template <typename T>
__global__ void __launch_bounds__(AB_ROWS * 32) attn_bwd_q_kernel(const T* __restrict__ qkv, const T* __restrict__ dout,
                                                                  T* __restrict__ dqkv, float* __restrict__ P,
                                                                  float* __restrict__ dS, int L, int C, float scale) {
  extern __shared__ float sm[];   // [rows][C] q | [rows][C] dO | [rows][L] p | [rows][L] dS
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles = (L + AB_ROWS - 1) / AB_ROWS;
  const int b = blockIdx.x / tiles, i = (blockIdx.x % tiles) * AB_ROWS + warp;
  if (i >= L) return;
  float* sq = sm + warp * C;
  float* sd = sm + AB_ROWS * C + warp * C;
  float* sp = sm + 2 * AB_ROWS * C + warp * L;
  float* ss = sm + 2 * AB_ROWS * C + AB_ROWS * L + warp * L;
  const int ld = 3 * C;
  const T* base = qkv + (size_t)b * L * ld;
  for (int c = lane * 8; c < C; c += 256) {
    float t[8];
    Vec8<T>::load(base + (size_t)i * ld + c, t);
#pragma unroll
    for (int j = 0; j < 8; ++j) sq[c + j] = t[j];
    Vec8<T>::load(dout + ((size_t)b * L + i) * C + c, t);
#pragma unroll
    for (int j = 0; j < 8; ++j) sd[c + j] = t[j];
  }
  __syncwarp();
  float mx = -INFINITY;
  for (int j = lane; j < L; j += 32) {
    const T* kr = base + (size_t)j * ld + C;
    const T* vr = base + (size_t)j * ld + 2 * C;
    float s = 0.f, dp = 0.f;
    for (int c = 0; c < C; c += 8) {
      float t[8];
      Vec8<T>::load(kr + c, t);
#pragma unroll
      for (int u = 0; u < 8; ++u) s = fmaf(sq[c + u], t[u], s);
      Vec8<T>::load(vr + c, t);
#pragma unroll
      for (int u = 0; u < 8; ++u) dp = fmaf(sd[c + u], t[u], dp);
    }
    s *= scale;
    sp[j] = s; ss[j] = dp;
    mx = fmaxf(mx, s);
  }
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float den = 0.f;
  for (int j = lane; j < L; j += 32) { const float e = __expf(sp[j] - mx); sp[j] = e; den += e; }
  for (int o = 16; o; o >>= 1) den += __shfl_xor_sync(0xffffffffu, den, o);
  const float inv = 1.0f / den;
  float dot = 0.f;
  for (int j = lane; j < L; j += 32) { const float p = sp[j] * inv; sp[j] = p; dot = fmaf(p, ss[j], dot); }
  for (int o = 16; o; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
  float* Pr = P + ((size_t)b * L + i) * L;
  float* Sr = dS + ((size_t)b * L + i) * L;
  for (int j = lane; j < L; j += 32) {
    const float d = sp[j] * (ss[j] - dot);
    ss[j] = d;
    Pr[j] = sp[j]; Sr[j] = d;
  }
  __syncwarp();
  // dq_i = scale * sum_j dS_ij k_j : lanes own channel octets
  for (int c = lane * 8; c < C; c += 256) {
    float acc[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) acc[u] = 0.f;
    for (int j = 0; j < L; ++j) {
      float t[8];
      Vec8<T>::load(base + (size_t)j * ld + C + c, t);
      const float d = ss[j];
#pragma unroll
      for (int u = 0; u < 8; ++u) acc[u] = fmaf(d, t[u], acc[u]);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) acc[u] *= scale;
    Vec8<T>::store(dqkv + ((size_t)b * L + i) * ld + c, acc);
  }
}

// pass 2, per key row j: dv_j = sum_i P_ij dO_i, dk_j = s sum_i dS_ij q_i
template <typename T>
__global__ void __launch_bounds__(AB_ROWS * 32) attn_bwd_kv_kernel(const T* __restrict__ qkv, const T* __restrict__ dout,
                                                                   T* __restrict__ dqkv, const float* __restrict__ P,
                                                                   const float* __restrict__ dS, int L, int C,
                                                                   float scale) {
  extern __shared__ float sm[];   // [rows][L] P column | [rows][L] dS column
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles = (L + AB_ROWS - 1) / AB_ROWS;
  const int b = blockIdx.x / tiles, j = (blockIdx.x % tiles) * AB_ROWS + warp;
  if (j >= L) return;
  float* sp = sm + warp * L;
  float* ss = sm + AB_ROWS * L + warp * L;
  for (int i = lane; i < L; i += 32) {
    sp[i] = P[((size_t)b * L + i) * L + j];
    ss[i] = dS[((size_t)b * L + i) * L + j];
  }
  __syncwarp();
  const int ld = 3 * C;
  const T* base = qkv + (size_t)b * L * ld;
  for (int c = lane * 8; c < C; c += 256) {
    float av[8], ak[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { av[u] = 0.f; ak[u] = 0.f; }
    for (int i = 0; i < L; ++i) {
      float t[8];
      Vec8<T>::load(dout + ((size_t)b * L + i) * C + c, t);
      const float p = sp[i];
#pragma unroll
      for (int u = 0; u < 8; ++u) av[u] = fmaf(p, t[u], av[u]);
      Vec8<T>::load(base + (size_t)i * ld + c, t);
      const float d = ss[i];
#pragma unroll
      for (int u = 0; u < 8; ++u) ak[u] = fmaf(d, t[u], ak[u]);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) ak[u] *= scale;
    Vec8<T>::store(dqkv + ((size_t)b * L + j) * ld + C + c, ak);
    Vec8<T>::store(dqkv + ((size_t)b * L + j) * ld + 2 * C + c, av);
  }
}

static inline int bw_grid(long long n) {
  long long blocks = (n + 255) / 256, cap = (long long)B200_NUM_SMS * 32;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace b200

using namespace b200;

static int gn_bwd_fill(GnBwdArgs& a, const void* x1, int C1, const void* x2, int C2, int B, int HW, int G, int nsplit,
                       const float* part_fwd, float* part_bwd, const float* gamma, const float* beta, float eps, int act,
                       const void* dy) {
  const int C = C1 + C2;
  if (!x1 || !part_fwd || !part_bwd || !dy || B < 0 || HW <= 0 || G <= 0 || nsplit <= 0 || C1 <= 0 || C2 < 0 ||
      (C2 > 0 && !x2)) return B200_ERR_ARG;
  if (C1 % 8 || C2 % 8 || C % G || gn_bwd_threads(C) == 0) return B200_ERR_UNSUPPORTED;
  a.x1 = x1; a.x2 = x2; a.C1 = C1; a.C2 = C2; a.B = B; a.HW = HW; a.G = G; a.nsplit = nsplit;
  a.part_fwd = part_fwd; a.part_bwd = part_bwd; a.gamma = gamma; a.beta = beta; a.eps = eps; a.act = act; a.dy = dy;
  a.dx1 = nullptr; a.dx2 = nullptr; a.acc1 = 0; a.acc2 = 0; a.chan_sums = nullptr;
  return B200_OK;
}

// Backward of GroupNorm(+SiLU) (models/layers.py:618-642, layerspp.py:243,264): per-(sample, split, group) sums of
// t = dy act'(z) gamma and t * xhat.  part_fwd comes from b200_gn_partial on the same sources.
// chan_sums (may be NULL): [B][nsplit][C1 + C2][2] = (sum dz xhat, sum dz) per channel and block: summed over B and nsplit
// they are d gamma / d beta (the training plans' GroupNorm parameter gradients, at no extra pass)
extern "C" int b200_gn_bwd_partial2(const void* x1, int C1, const void* x2, int C2, int dtype, int B, int HW, int G,
                                    int nsplit, const float* part_fwd, const float* gamma, const float* beta, float eps,
                                    int act, const void* dy, float* part_bwd, float* chan_sums, cudaStream_t st) {
  GnBwdArgs a;
  int rc = gn_bwd_fill(a, x1, C1, x2, C2, B, HW, G, nsplit, part_fwd, part_bwd, gamma, beta, eps, act, dy);
  if (rc || B == 0) return rc;
  a.chan_sums = chan_sums;
  const int C = C1 + C2, thr = gn_bwd_threads(C);
  const size_t smem = (size_t)2 * G * sizeof(float) + (size_t)(thr / (C / 8)) * 2 * C * sizeof(float);
  if (dtype == B200_DT_F32) gn_bwd_partial_kernel<float><<<B * nsplit, thr, smem, st>>>(a);
  else if (dtype == B200_DT_BF16) gn_bwd_partial_kernel<bf16><<<B * nsplit, thr, smem, st>>>(a);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

extern "C" int b200_gn_bwd_partial(const void* x1, int C1, const void* x2, int C2, int dtype, int B, int HW, int G,
                                   int nsplit, const float* part_fwd, const float* gamma, const float* beta, float eps,
                                   int act, const void* dy, float* part_bwd, cudaStream_t st) {
  return b200_gn_bwd_partial2(x1, C1, x2, C2, dtype, B, HW, G, nsplit, part_fwd, gamma, beta, eps, act, dy, part_bwd, nullptr, st);
}

// dx = rstd (t - mean_g(t) - xhat mean_g(t xhat)), written (acc = 0) or added (acc = 1) to dx1 / dx2 (either may be NULL).
extern "C" int b200_gn_bwd_apply(const void* x1, int C1, const void* x2, int C2, int dtype, int B, int HW, int G,
                                 int nsplit, const float* part_fwd, const float* part_bwd, const float* gamma,
                                 const float* beta, float eps, int act, const void* dy, void* dx1, int acc1, void* dx2,
                                 int acc2, cudaStream_t st) {
  GnBwdArgs a;
  int rc = gn_bwd_fill(a, x1, C1, x2, C2, B, HW, G, nsplit, part_fwd, const_cast<float*>(part_bwd), gamma, beta, eps, act, dy);
  if (rc || B == 0) return rc;
  a.dx1 = dx1; a.dx2 = dx2; a.acc1 = acc1; a.acc2 = acc2;
  const int C = C1 + C2, thr = gn_bwd_threads(C);
  const size_t smem = (size_t)4 * G * sizeof(float);
  if (dtype == B200_DT_F32) gn_bwd_apply_kernel<float><<<B * nsplit, thr, smem, st>>>(a);
  else if (dtype == B200_DT_BF16) gn_bwd_apply_kernel<bf16><<<B * nsplit, thr, smem, st>>>(a);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

extern "C" int b200_axpby(const void* a, float sa, const void* b, float sb, void* out, int dtype, long long n,
                          cudaStream_t st) {
  if (!a || !out || n < 0 || (n % 8)) return B200_ERR_ARG;
  if (n == 0) return B200_OK;
  if (dtype == B200_DT_F32) axpby_kernel<float><<<bw_grid(n / 8), 256, 0, st>>>((const float*)a, sa, (const float*)b, sb, (float*)out, n / 8);
  else if (dtype == B200_DT_BF16) axpby_kernel<bf16><<<bw_grid(n / 8), 256, 0, st>>>((const bf16*)a, sa, (const bf16*)b, sb, (bf16*)out, n / 8);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

extern "C" int b200_nearest_up2_bwd_nhwc(const void* dy, void* dx, int dtype, int B, int H, int W, int C, int acc,
                                         cudaStream_t st) {
  if (!dy || !dx || B < 0 || C % 8) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  const long long total = (long long)B * H * W * (C / 8);
  if (dtype == B200_DT_F32) nearest_up2_bwd_kernel<float><<<bw_grid(total), 256, 0, st>>>((const float*)dy, (float*)dx, B, H, W, C / 8, acc);
  else if (dtype == B200_DT_BF16) nearest_up2_bwd_kernel<bf16><<<bw_grid(total), 256, 0, st>>>((const bf16*)dy, (bf16*)dx, B, H, W, C / 8, acc);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

extern "C" int b200_zero_stuff2_nhwc(const void* dy, void* out, int dtype, int B, int Hi, int Wi, int Ho, int Wo, int C,
                                     cudaStream_t st) {
  if (!dy || !out || B < 0 || C % 8 || Ho < 2 * Hi || Wo < 2 * Wi) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  const long long total = (long long)B * Ho * Wo * (C / 8);
  if (dtype == B200_DT_F32) zero_stuff2_kernel<float><<<bw_grid(total), 256, 0, st>>>((const float*)dy, (float*)out, B, Hi, Wi, Ho, Wo, C / 8);
  else if (dtype == B200_DT_BF16) zero_stuff2_kernel<bf16><<<bw_grid(total), 256, 0, st>>>((const bf16*)dy, (bf16*)out, B, Hi, Wi, Ho, Wo, C / 8);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

// dqkv [B, L, 3C] from qkv [B, L, 3C] and dout [B, L, C]; scratch: 2 * B * L * L floats.
extern "C" int b200_attention_bwd(const void* qkv, const void* dout, void* dqkv, float* scratch, int dtype, int B, int L,
                                  int C, float scale, cudaStream_t st) {
  if (!qkv || !dout || !dqkv || !scratch || B < 0 || L <= 0 || C <= 0 || C % 8) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  // bf16, L = C = 256 (every shipped network's attention): tensor-core path (attn_bwd_mma.cu); B200_ATTN_BWD_MMA=0: A/B
  static const bool mma_on = [] { const char* e = getenv("B200_ATTN_BWD_MMA"); return !(e && e[0] == '0'); }();
  if (mma_on && dtype == B200_DT_BF16 && b200_attention_bwd_mma_supported(L, C))
    return b200_attention_bwd_mma(qkv, dout, dqkv, scratch, B, L, C, scale, st);
  float* P = scratch;
  float* dS = scratch + (size_t)B * L * L;
  const int tiles = (L + AB_ROWS - 1) / AB_ROWS;
  const size_t sm1 = (size_t)(2 * AB_ROWS * C + 2 * AB_ROWS * L) * sizeof(float);
  const size_t sm2 = (size_t)(2 * AB_ROWS * L) * sizeof(float);
  if (sm1 > 200 * 1024) return B200_ERR_UNSUPPORTED;
  cudaError_t e;
  if (dtype == B200_DT_F32) {
    e = cudaFuncSetAttribute(attn_bwd_q_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1);
    if (e != cudaSuccess) return -(1000 + (int)e);
    attn_bwd_q_kernel<float><<<B * tiles, AB_ROWS * 32, sm1, st>>>((const float*)qkv, (const float*)dout, (float*)dqkv, P, dS, L, C, scale);
    attn_bwd_kv_kernel<float><<<B * tiles, AB_ROWS * 32, sm2, st>>>((const float*)qkv, (const float*)dout, (float*)dqkv, P, dS, L, C, scale);
  } else if (dtype == B200_DT_BF16) {
    e = cudaFuncSetAttribute(attn_bwd_q_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1);
    if (e != cudaSuccess) return -(1000 + (int)e);
    attn_bwd_q_kernel<bf16><<<B * tiles, AB_ROWS * 32, sm1, st>>>((const bf16*)qkv, (const bf16*)dout, (bf16*)dqkv, P, dS, L, C, scale);
    attn_bwd_kv_kernel<bf16><<<B * tiles, AB_ROWS * 32, sm2, st>>>((const bf16*)qkv, (const bf16*)dout, (bf16*)dqkv, P, dS, L, C, scale);
  } else return B200_ERR_ARG;
  B200_RETURN_LAST();
}
