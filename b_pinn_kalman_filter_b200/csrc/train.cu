// Training-step kernels for sm_100a: what `loss.backward(); optimize_fn(...); ema.update(...)` of the reference's
// step_fn (losses.py:165-222, 48-66; models/ema.py:36-56) needs beyond the forward / input-gradient kernels.
//
// Weight gradients reuse the tensor-core convolution: dW[co][(tap, ci)] = sum_p dY[p][co] X[p + tap][ci] is a 1x1
// "convolution" over K = pixels whose fake image rows are the Cout rows of dY^T and whose "weights" are the transposed
// im2col XT[(tap, ci)][p].  transpose2d and im2col_t below produce those two K-contiguous operands (memory-bound, one
// read + one / taps writes); b200_conv2d_tc (or the SIMT conv for odd shapes and the strict fp32 mode) contracts them
// and its output IS the packed weight layout [Cout][kh][kw][Cin].
#include "common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

// ---- out[c][r] = in[r][c]  (R x C -> C x R), 32 x 32 tiles through shared memory ----------------------------------
template <typename T>
__global__ void __launch_bounds__(256) transpose2d_kernel(const T* __restrict__ in, T* __restrict__ out, long long R, int C) {
  __shared__ T tile[32][33];
  const long long r0 = (long long)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const long long r = r0 + ty + j;
    const int c = c0 + tx;
    if (r < R && c < C) tile[ty + j][tx] = in[r * C + c];
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int c = c0 + ty + j;
    const long long r = r0 + tx;
    if (r < R && c < C) out[(long long)c * R + r] = tile[tx][ty + j];
  }
}

// ---- transposed im2col: out[(tap * Ct + c)][p] = x[b, ho * stride + kh - pad, wo * stride + kw - pad, c] ------------
// p = (b * Ho + ho) * Wo + wo; zero outside the image; two channels-last sources concatenated along c.
template <typename T>
__global__ void __launch_bounds__(256) im2col_t_kernel(const T* __restrict__ x1, int C1, const T* __restrict__ x2, int C2,
                                                       T* __restrict__ out, int B, int H, int W, int Ho, int Wo, int ks,
                                                       int stride, int pad) {
  __shared__ T tile[32][33];
  const int Ct = C1 + C2;
  const long long P = (long long)B * Ho * Wo;
  const long long p0 = (long long)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32, tap = blockIdx.z;
  const int kh = tap / ks, kw = tap - kh * ks;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const long long p = p0 + ty + j;
    const int c = c0 + tx;
    T v = from_f<T>(0.0f);
    if (p < P && c < Ct) {
      const int wo = (int)(p % Wo);
      const long long t = p / Wo;
      const int ho = (int)(t % Ho);
      const long long b = t / Ho;
      const int h = ho * stride + kh - pad, w = wo * stride + kw - pad;
      if (h >= 0 && h < H && w >= 0 && w < W) {
        const long long px = (b * H + h) * W + w;
        v = c < C1 ? x1[px * C1 + c] : x2[px * C2 + (c - C1)];
      }
    }
    tile[ty + j][tx] = v;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int c = c0 + ty + j;
    const long long p = p0 + tx;
    if (p < P && c < Ct) out[((long long)tap * Ct + c) * P + p] = tile[tx][ty + j];
  }
}

// bf16, 16-byte accesses on both sides: a block moves 64 pixels x 64 channels of one tap.  Loads: one uint4 = 8 channels
// of a pixel (coalesced over channels); stores: one uint4 = 8 consecutive pixels of a channel row (coalesced over
// pixels).  The 2-byte scatter version above ran at ~0.6 TB/s; the im2col is 9x the activation, so it matters.
// Needs C1 % 8 == C2 % 8 == 0 and P % 8 == 0.
__global__ void __launch_bounds__(256) im2col_t_bf16v_kernel(const bf16* __restrict__ x1, int C1, const bf16* __restrict__ x2,
                                                             int C2, bf16* __restrict__ out, int B, int H, int W, int Ho,
                                                             int Wo, int ks, int stride, int pad) {
  __shared__ __align__(16) bf16 tile[64][72];          // [pixel][channel], 144-byte pitch: conflict-free uint4 rows
  const int Ct = C1 + C2;
  const long long P = (long long)B * Ho * Wo;
  const long long p0 = (long long)blockIdx.x * 64;
  const int c0 = blockIdx.y * 64, tap = blockIdx.z;
  const int kh = tap / ks, kw = tap - kh * ks;
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int idx = threadIdx.x + it * 256;              // 64 pixels x 8 vectors
    const int pl = idx >> 3, v = idx & 7;
    const long long p = p0 + pl;
    const int c = c0 + v * 8;
    uint4 val = make_uint4(0u, 0u, 0u, 0u);
    if (p < P && c < Ct) {
      const int wo = (int)(p % Wo);
      const long long t = p / Wo;
      const int ho = (int)(t % Ho);
      const long long b = t / Ho;
      const int h = ho * stride + kh - pad, w = wo * stride + kw - pad;
      if (h >= 0 && h < H && w >= 0 && w < W) {
        const long long px = (b * H + h) * W + w;
        val = c < C1 ? __ldg(reinterpret_cast<const uint4*>(x1 + px * C1 + c))
                     : __ldg(reinterpret_cast<const uint4*>(x2 + px * C2 + (c - C1)));
      }
    }
    *reinterpret_cast<uint4*>(&tile[pl][v * 8]) = val;
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int idx = threadIdx.x + it * 256;              // 64 channels x 8 pixel-vectors
    const int cl = idx >> 3, v = idx & 7;
    const int c = c0 + cl;
    const long long p = p0 + v * 8;
    if (c < Ct && p < P) {
      uint4 o;
      bf16* ob = reinterpret_cast<bf16*>(&o);
#pragma unroll
      for (int j = 0; j < 8; ++j) ob[j] = tile[v * 8 + j][cl];
      *reinterpret_cast<uint4*>(out + ((long long)tap * Ct + c) * P + p) = o;
    }
  }
}

// ---- out[b][c] (+)= scale * sum_r in[b][r][c]   (bias / time-embedding gradients: column sums of a cotangent) --------
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ in, float* __restrict__ out, long long R, int C,
                                                     long long out_stride, float scale, int accumulate) {
  __shared__ float red[8][33];
  const long long b = blockIdx.y;
  const int c = blockIdx.x * 32 + (threadIdx.x & 31), ty = threadIdx.x >> 5;
  float s = 0.0f;
  if (c < C) {
    const T* p = in + b * R * C + c;
    for (long long r = ty; r < R; r += 8) s += to_f(p[r * C]);
  }
  red[ty][threadIdx.x & 31] = s;
  __syncthreads();
  if (ty == 0 && c < C) {
    float t = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += red[j][threadIdx.x & 31];
    float* o = out + b * out_stride + c;
    *o = accumulate ? *o + t * scale : t * scale;
  }
}

// ---- GroupNorm affine gradients: dgamma[c] = sum dz xhat, dbeta[c] = sum dz, dz = dy * act'(xhat gamma + beta) ------
// partial[b][c][2] over the pixels of one sample; stats = b200_gn_partial output [B][nsplit][G][2] of the SAME x.
template <typename T>
__global__ void __launch_bounds__(256) gn_param_grad_kernel(const T* __restrict__ x1, int C1, const T* __restrict__ x2, int C2,
                                                            const T* __restrict__ dy, const float* __restrict__ stats,
                                                            int nsplit, int G, int HW, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float eps, int act,
                                                            float* __restrict__ partial) {
  __shared__ float red[8][33][2];
  const int C = C1 + C2, cpg = C / G;
  const long long b = blockIdx.y;
  const int c = blockIdx.x * 32 + (threadIdx.x & 31), ty = threadIdx.x >> 5;
  float sg = 0.0f, sb = 0.0f;
  if (c < C) {
    const int g = c / cpg;
    double ts = 0.0, tq = 0.0;
    for (int k = 0; k < nsplit; ++k) {
      const float* o = stats + ((b * nsplit + k) * G + g) * 2;
      ts += o[0]; tq += o[1];
    }
    const double rn = 1.0 / ((double)HW * cpg);
    const double mean = ts * rn;
    double var = tq * rn - mean * mean;
    if (var < 0.0) var = 0.0;
    const float mu = (float)mean, rstd = (float)(1.0 / sqrt(var + (double)eps));
    const float ga = gamma[c], be = beta[c];
    const T* xp = c < C1 ? x1 + b * HW * C1 + c : x2 + b * HW * C2 + (c - C1);
    const int xs = c < C1 ? C1 : C2;
    const T* dp = dy + b * HW * C + c;
    for (int p = ty; p < HW; p += 8) {
      const float xhat = (to_f(xp[(long long)p * xs]) - mu) * rstd;
      float dz = to_f(dp[(long long)p * C]);
      if (act) {
        const float z = fmaf(xhat, ga, be);
        const float sgm = 1.0f / (1.0f + expf(-z));
        dz *= sgm * (1.0f + z * (1.0f - sgm));
      }
      sg = fmaf(dz, xhat, sg); sb += dz;
    }
  }
  red[ty][threadIdx.x & 31][0] = sg; red[ty][threadIdx.x & 31][1] = sb;
  __syncthreads();
  if (ty == 0 && c < C) {
    float a = 0.0f, d = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { a += red[j][threadIdx.x & 31][0]; d += red[j][threadIdx.x & 31][1]; }
    partial[(b * C + c) * 2] = a; partial[(b * C + c) * 2 + 1] = d;
  }
}

// ---- small fp32 GEMM for the time-embedding MLP's backward: C = alpha op(A) op(B) + beta C -------------------------
// op(A): M x K (A stored K x M when ta), op(B): K x N (B stored N x K when tb); 16 x 16 tiles.
__global__ void __launch_bounds__(256) gemm_f32_kernel(float* __restrict__ Cm, const float* __restrict__ A,
                                                       const float* __restrict__ Bm, int M, int N, int K, int ta, int tb,
                                                       float alpha, float beta) {
  __shared__ float sa[16][17], sb[16][17];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m = blockIdx.y * 16 + ty, n = blockIdx.x * 16 + tx;
  float acc = 0.0f;
  for (int k0 = 0; k0 < K; k0 += 16) {
    const int ka = k0 + tx, kb = k0 + ty;
    sa[ty][tx] = (m < M && ka < K) ? (ta ? A[(long long)ka * M + m] : A[(long long)m * K + ka]) : 0.0f;
    sb[ty][tx] = (kb < K && n < N) ? (tb ? Bm[(long long)n * K + kb] : Bm[(long long)kb * N + n]) : 0.0f;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) acc = fmaf(sa[ty][k], sb[k][tx], acc);
    __syncthreads();
  }
  if (m < M && n < N) {
    float* o = Cm + (long long)m * N + n;
    *o = beta == 0.0f ? alpha * acc : alpha * acc + beta * *o;
  }
}

__global__ void silu_f32_kernel(float* __restrict__ y, const float* __restrict__ x, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float z = x[i];
    y[i] = z / (1.0f + expf(-z));
  }
}

// dx = dy * silu'(x)
__global__ void silu_bwd_kernel(float* __restrict__ dx, const float* __restrict__ dy, const float* __restrict__ x, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float z = x[i], s = 1.0f / (1.0f + expf(-z));
    dx[i] = dy[i] * s * (1.0f + z * (1.0f - s));
  }
}

// ---- score-matching loss and its cotangent (losses.py:69-163) -------------------------------------------------------
// residual = alpha_b * out + beta_b * z;  losses_b = r * sum_e residual^2 * w_b  (r = 1/n_per for reduce_mean, 0.5 for
// the sum form);  loss = mean_b losses_b;  d loss / d out = (1/B) w_b r 2 residual alpha_b.  Covers the continuous SDE
// loss with / without likelihood weighting, the SMLD and the DDPM losses (the host picks alpha, beta, w).
// One block per sample: writes losses[b] and the cotangent `dout` (already multiplied by mul[b], the per-sample factor
// the network head applies to its output: h / sigma, models/ncsnpp.py:377-379).
__global__ void __launch_bounds__(256) score_loss_kernel(float* __restrict__ losses, float* __restrict__ dout,
                                                         const float* __restrict__ out, const float* __restrict__ z,
                                                         const float* __restrict__ alpha, const float* __restrict__ beta,
                                                         const float* __restrict__ wgt, const float* __restrict__ mul,
                                                         float r, float inv_batch, long long n_per) {
  __shared__ float red[32];
  const long long b = blockIdx.x;
  const float al = alpha[b], be = beta[b], w = wgt[b], m = mul ? mul[b] : 1.0f;
  const float g = inv_batch * w * r * 2.0f * al * m;
  float acc = 0.0f;
  for (long long i = threadIdx.x; i < n_per; i += blockDim.x) {
    const float res = fmaf(al, out[b * n_per + i], be * z[b * n_per + i]);
    acc = fmaf(res, res, acc);
    dout[b * n_per + i] = g * res;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) losses[b] = acc * r * w;
}

// mean of n floats (single block) -> out[0]
__global__ void __launch_bounds__(256) mean_kernel(float* __restrict__ out, const float* __restrict__ in, int n) {
  __shared__ float red[32];
  float a = 0.0f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) a += in[i];
  a = block_sum(a, red);
  if (threadIdx.x == 0) out[0] = a / (float)n;
}

// ---- dropout (nn.Dropout in train mode, models/layerspp.py:265): y = x * keep / (1 - p), keep ~ Bernoulli(1 - p) ----
// The mask is a pure function of (seed, offset + *step * 16, element), so the backward pass regenerates it.
template <typename T>
__global__ void __launch_bounds__(256) dropout_kernel(T* __restrict__ y, const T* __restrict__ x, long long n, float p,
                                                      uint64_t seed, uint64_t offset, const int* __restrict__ step) {
  const uint64_t off = offset + (step ? (uint64_t)(*step) * 16 : 0);
  const float inv = 1.0f / (1.0f - p);
  const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g * 4 < n; g += (long long)gridDim.x * blockDim.x) {
    const uint4 r = Philox::round10(make_uint4((uint32_t)g, (uint32_t)((uint64_t)g >> 32), (uint32_t)off, (uint32_t)(off >> 32)), key);
    const uint32_t u[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long e = g * 4 + j;
      if (e < n) {
        const bool keep = (float)u[j] * 2.3283064365386963e-10f >= p;
        y[e] = from_f<T>(keep ? to_f(x[e]) * inv : 0.0f);
      }
    }
  }
}

// ---- optimizer: clip_grad_norm_ + Adam + EMA in one pass over the flat parameter buffer -----------------------------
// (losses.py:48-66 -> torch.optim.Adam; models/ema.py:36-56).  hyper (device, refreshed by the host each step):
//   [0] lr  [1] beta1  [2] beta2  [3] eps  [4] weight_decay  [5] 1 - beta1^t  [6] 1 - beta2^t  [7] max_norm (< 0: off)
//   [8] 1 - ema_decay_t      sumsq[0] = sum of squares of the (already all-reduced) gradient.
// Block partials land in a per-device scratch array and the LAST block to finish adds them in index order: the same bits
// every run (a float atomicAdd per block made the clip factor, and with it every parameter, depend on block timing).
// One array per device: calls on different streams of one device must not overlap (one training stream per device).
constexpr int SUMSQ_MAX_BLOCKS = 148 * 16;
__device__ float g_sumsq_part[SUMSQ_MAX_BLOCKS];
__device__ unsigned int g_sumsq_done;
__global__ void __launch_bounds__(256) sumsq_f32_kernel(float* __restrict__ out, const float* __restrict__ g, long long n) {
  __shared__ float red[32];
  __shared__ bool last;
  float a = 0.0f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) a = fmaf(g[i], g[i], a);
  a = block_sum(a, red);
  if (threadIdx.x == 0) {
    g_sumsq_part[blockIdx.x] = a;
    __threadfence();
    last = atomicAdd(&g_sumsq_done, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  float t = 0.0f;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) t += __ldcg(&g_sumsq_part[i]);   // fixed assignment and order
  t = block_sum(t, red);
  if (threadIdx.x == 0) { out[0] = t; g_sumsq_done = 0u; }
}

__global__ void __launch_bounds__(256) adam_ema_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                                                       float* __restrict__ ema, const float* __restrict__ g,
                                                       const float* __restrict__ hyper, const float* __restrict__ sumsq,
                                                       float grad_scale, long long n) {
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4], bc1 = hyper[5], bc2 = hyper[6];
  const float max_norm = hyper[7], omd = hyper[8];
  float clip = 1.0f;
  if (max_norm >= 0.0f) {
    const float total = sqrtf(sumsq[0]) * grad_scale;          // norm of the scaled gradient
    clip = fminf(max_norm / (total + 1e-6f), 1.0f);              // torch.nn.utils.clip_grad_norm_
  }
  const float gs = grad_scale * clip, step = lr / bc1, rs2 = rsqrtf(bc2);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float pi = p[i];
    float gi = g[i] * gs;
    if (wd != 0.0f) gi = fmaf(wd, pi, gi);
    const float mi = fmaf(b1, m[i], (1.0f - b1) * gi);
    const float vi = fmaf(b2, v[i], (1.0f - b2) * gi * gi);
    m[i] = mi; v[i] = vi;
    pi -= step * mi / (sqrtf(vi) * rs2 + eps);
    p[i] = pi;
    if (ema) { const float s = ema[i]; ema[i] = s - omd * (s - pi); }
  }
}

}  // namespace b200

using namespace b200;

static inline int tr_grid(long long n) {
  const long long want = (n + 255) / 256;
  return (int)(want < (long long)B200_NUM_SMS * 16 ? (want > 0 ? want : 1) : (long long)B200_NUM_SMS * 16);
}

// out[C][R] = in[R][C]^T  (dtype: B200_DT_F32 / B200_DT_BF16)
extern "C" int b200_transpose2d(const void* in, void* out, int dtype, long long R, int C, cudaStream_t st) {
  if (!in || !out || R <= 0 || C <= 0 || (R + 31) / 32 > 0x7fffffffLL || (C + 31) / 32 > 65535) return B200_ERR_ARG;
  dim3 grid((unsigned)((R + 31) / 32), (unsigned)((C + 31) / 32));
  if (dtype == B200_DT_F32) transpose2d_kernel<float><<<grid, 256, 0, st>>>((const float*)in, (float*)out, R, C);
  else if (dtype == B200_DT_BF16) transpose2d_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)in, (bf16*)out, R, C);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

// out[(tap, c)][p] = x[b, ho*stride + kh - pad, wo*stride + kw - pad, c] over p = (b, ho, wo): the K-contiguous "weight"
// operand of the weight-gradient contraction.  x1 / x2: channels-last [B, H, W, C1 / C2] (x2 may be NULL, C2 = 0).
extern "C" int b200_im2col_t(const void* x1, int C1, const void* x2, int C2, void* out, int dtype, int B, int H, int W,
                             int Ho, int Wo, int ksize, int stride, int pad, cudaStream_t st) {
  if (!x1 || !out || C1 <= 0 || C2 < 0 || (C2 > 0 && !x2) || B <= 0 || H <= 0 || W <= 0 || Ho <= 0 || Wo <= 0 ||
      ksize < 1 || ksize > 7 || stride < 1) return B200_ERR_ARG;
  const long long P = (long long)B * Ho * Wo;
  if ((P + 31) / 32 > 0x7fffffffLL || (C1 + C2 + 31) / 32 > 65535) return B200_ERR_UNSUPPORTED;
  dim3 grid((unsigned)((P + 31) / 32), (unsigned)((C1 + C2 + 31) / 32), (unsigned)(ksize * ksize));
  if (dtype == B200_DT_F32)
    im2col_t_kernel<float><<<grid, 256, 0, st>>>((const float*)x1, C1, (const float*)x2, C2, (float*)out, B, H, W, Ho, Wo, ksize, stride, pad);
  else if (dtype == B200_DT_BF16) {
    const bool al = (((uintptr_t)x1 | (uintptr_t)x2 | (uintptr_t)out) & 15) == 0;
    if (al && C1 % 8 == 0 && C2 % 8 == 0 && P % 8 == 0) {
      dim3 g64((unsigned)((P + 63) / 64), (unsigned)((C1 + C2 + 63) / 64), (unsigned)(ksize * ksize));
      im2col_t_bf16v_kernel<<<g64, 256, 0, st>>>((const bf16*)x1, C1, (const bf16*)x2, C2, (bf16*)out, B, H, W, Ho, Wo, ksize, stride, pad);
    } else {
      im2col_t_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)x1, C1, (const bf16*)x2, C2, (bf16*)out, B, H, W, Ho, Wo, ksize, stride, pad);
    }
  }
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

// out[b * out_stride + c] (+)= scale * sum_r in[b][r][c]; in: [Bn][R][C] of `dtype`, out: fp32 rows `out_stride` apart
// (out_stride >= C: lets a per-sample column sum land in a column slice of a wider [Bn][total] table)
extern "C" int b200_colsum(const void* in, float* out, int dtype, int Bn, long long R, int C, long long out_stride,
                           float scale, int accumulate, cudaStream_t st) {
  if (!in || !out || Bn <= 0 || R <= 0 || C <= 0 || Bn > 65535 || out_stride < C) return B200_ERR_ARG;
  dim3 grid((unsigned)((C + 31) / 32), (unsigned)Bn);
  if (dtype == B200_DT_F32) colsum_kernel<float><<<grid, 256, 0, st>>>((const float*)in, out, R, C, out_stride, scale, accumulate);
  else if (dtype == B200_DT_BF16) colsum_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)in, out, R, C, out_stride, scale, accumulate);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

// partial[B][C1 + C2][2] = per-sample (dgamma, dbeta) of GroupNorm(+SiLU) over cat(x1, x2); stats from b200_gn_partial
extern "C" int b200_gn_param_grad(const void* x1, int C1, const void* x2, int C2, const void* dy, int dtype, int B, int HW,
                                  int G, int nsplit, const float* stats, const float* gamma, const float* beta, float eps,
                                  int act, float* partial, cudaStream_t st) {
  const int C = C1 + C2;
  if (!x1 || !dy || !stats || !gamma || !beta || !partial || B <= 0 || B > 65535 || HW <= 0 || G <= 0 || C % G || (C2 > 0 && !x2))
    return B200_ERR_ARG;
  dim3 grid((unsigned)((C + 31) / 32), (unsigned)B);
  if (dtype == B200_DT_F32)
    gn_param_grad_kernel<float><<<grid, 256, 0, st>>>((const float*)x1, C1, (const float*)x2, C2, (const float*)dy, stats, nsplit, G, HW, gamma, beta, eps, act, partial);
  else if (dtype == B200_DT_BF16)
    gn_param_grad_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)x1, C1, (const bf16*)x2, C2, (const bf16*)dy, stats, nsplit, G, HW, gamma, beta, eps, act, partial);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

// C[M][N] = alpha * op(A) op(B) + beta * C  (fp32; ta: A stored [K][M]; tb: B stored [N][K])
extern "C" int b200_gemm_f32(float* C, const float* A, const float* Bm, int M, int N, int K, int ta, int tb, float alpha,
                             float beta, cudaStream_t st) {
  if (!C || !A || !Bm || M <= 0 || N <= 0 || K <= 0 || (M + 15) / 16 > 65535) return B200_ERR_ARG;
  dim3 grid((unsigned)((N + 15) / 16), (unsigned)((M + 15) / 16));
  gemm_f32_kernel<<<grid, 256, 0, st>>>(C, A, Bm, M, N, K, ta, tb, alpha, beta);
  B200_RETURN_LAST();
}

extern "C" int b200_silu_f32(float* y, const float* x, long long n, cudaStream_t st) {
  if (!y || !x || n <= 0) return B200_ERR_ARG;
  silu_f32_kernel<<<tr_grid(n), 256, 0, st>>>(y, x, n);
  B200_RETURN_LAST();
}

extern "C" int b200_silu_bwd_f32(float* dx, const float* dy, const float* x, long long n, cudaStream_t st) {
  if (!dx || !dy || !x || n <= 0) return B200_ERR_ARG;
  silu_bwd_kernel<<<tr_grid(n), 256, 0, st>>>(dx, dy, x, n);
  B200_RETURN_LAST();
}

// losses[B], loss[1] (mean of losses) and the cotangent dout[B][n_per] of the score-matching losses (losses.py:69-163)
extern "C" int b200_score_loss(float* losses, float* loss, float* dout, const float* out, const float* z,
                               const float* alpha, const float* beta, const float* weight, const float* mul, float r,
                               int B, long long n_per, cudaStream_t st) {
  if (!losses || !loss || !dout || !out || !z || !alpha || !beta || !weight || B <= 0 || n_per <= 0) return B200_ERR_ARG;
  score_loss_kernel<<<B, 256, 0, st>>>(losses, dout, out, z, alpha, beta, weight, mul, r, 1.0f / (float)B, n_per);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return -(1000 + (int)e);
  mean_kernel<<<1, 256, 0, st>>>(loss, losses, B);
  B200_RETURN_LAST();
}

// y = dropout(x; p) with the mask of (seed, offset + *step * 16); the same call on a cotangent is the backward
extern "C" int b200_dropout(void* y, const void* x, int dtype, long long n, float p, uint64_t seed, uint64_t offset,
                            const int* step, cudaStream_t st) {
  if (!y || !x || n <= 0 || p < 0.0f || p >= 1.0f) return B200_ERR_ARG;
  if (dtype == B200_DT_F32) dropout_kernel<float><<<tr_grid((n + 3) / 4), 256, 0, st>>>((float*)y, (const float*)x, n, p, seed, offset, step);
  else if (dtype == B200_DT_BF16) dropout_kernel<bf16><<<tr_grid((n + 3) / 4), 256, 0, st>>>((bf16*)y, (const bf16*)x, n, p, seed, offset, step);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

// sumsq[0] = sum g^2, bit-reproducible
extern "C" int b200_grad_sumsq(float* sumsq, const float* g, long long n, cudaStream_t st) {
  if (!sumsq || !g || n <= 0) return B200_ERR_ARG;
  sumsq_f32_kernel<<<tr_grid(n), 256, 0, st>>>(sumsq, g, n);          // tr_grid <= SUMSQ_MAX_BLOCKS
  B200_RETURN_LAST();
}

// clip_grad_norm_ + Adam step + EMA update over flat fp32 buffers (ema may be NULL); hyper: 9 floats on the device
// (see adam_ema_kernel); grad_scale multiplies g first (1 / world after a summing all-reduce)
extern "C" int b200_adam_ema(float* p, float* m, float* v, float* ema, const float* g, const float* hyper,
                             const float* sumsq, float grad_scale, long long n, cudaStream_t st) {
  if (!p || !m || !v || !g || !hyper || !sumsq || n <= 0) return B200_ERR_ARG;
  adam_ema_kernel<<<tr_grid(n), 256, 0, st>>>(p, m, v, ema, g, hyper, sumsq, grad_scale, n);
  B200_RETURN_LAST();
}

// ---- out[i] = sum_s in[s][i]  (the S partial products of b200_gemm_tc_splitk; 8 elements per thread, fixed order) -------
namespace b200 {
template <typename T>
__global__ void __launch_bounds__(256) sum_partials_kernel(const T* __restrict__ in, float* __restrict__ out, int S, long long n) {
  for (long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) * 8; i < n; i += (long long)gridDim.x * blockDim.x * 8) {
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.0f;
    for (int s = 0; s < S; ++s) {
      float v[8];
      Vec8<T>::load(in + (size_t)s * n + i, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
    *reinterpret_cast<float4*>(out + i) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    *reinterpret_cast<float4*>(out + i + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
}
}  // namespace b200

extern "C" int b200_sum_partials(const void* in, float* out, int dtype, int S, long long n, cudaStream_t st) {
  if (!in || !out || S <= 0 || n <= 0 || n % 8) return B200_ERR_ARG;
  if (dtype == B200_DT_F32) sum_partials_kernel<float><<<tr_grid(n / 8), 256, 0, st>>>((const float*)in, out, S, n);
  else if (dtype == B200_DT_BF16) sum_partials_kernel<bf16><<<tr_grid(n / 8), 256, 0, st>>>((const bf16*)in, out, S, n);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

// ---- batched strided copy --------------------------------------------------------------------------------------------
// After every optimizer step the plans' derived weight copies have to follow the flat fp32 parameters: forward weights
// re-packed to [Cout][kh][kw][Cin] bf16 (NIN transposed, q/k/v and the 1x1 shortcuts stacked), backward-data weights
// flipped and transposed, and before it the per-layer weight gradients scattered from their GEMM layouts into the flat
// gradient buffer.  As ~1500 torch copy / permute / cast calls that was 20 of the 76 ms of a C3 training step, all of it
// host time.  Each of those is a <= 4-D strided copy with a scale and a type change: the host writes them ONCE into a
// descriptor table (models/copylist.py) and one launch executes the whole table.
namespace b200 {
struct CopyDesc {
  const void* src; const void* src2; void* dst;        // dst[i] = scale * (src[j(i)] + src2[j(i)])   (src2 may be null)
  long long ss[4], ds[4];                              // element strides of src (src2 alike) and dst, dimension 3 fastest
  int n[4];                                            // extents
  int flip;                                            // bit k: source index along dimension k runs backwards
  float scale;
  int sdt, ddt;                                        // B200_DT_F32 / B200_DT_BF16
};
constexpr int COPY_CHUNK = 2048;                       // elements per block

__global__ void __launch_bounds__(256) copy_batch_kernel(const CopyDesc* __restrict__ descs, const int2* __restrict__ blocks) {
  const int2 bk = blocks[blockIdx.x];                  // (descriptor, chunk)
  const CopyDesc d = descs[bk.x];
  const long long total = (long long)d.n[0] * d.n[1] * d.n[2] * d.n[3];
  const long long e0 = (long long)bk.y * COPY_CHUNK;
  const long long e1 = e0 + COPY_CHUNK < total ? e0 + COPY_CHUNK : total;
  for (long long e = e0 + threadIdx.x; e < e1; e += blockDim.x) {
    long long r = e;
    const int i3 = (int)(r % d.n[3]); r /= d.n[3];
    const int i2 = (int)(r % d.n[2]); r /= d.n[2];
    const int i1 = (int)(r % d.n[1]);
    const int i0 = (int)(r / d.n[1]);
    const int j0 = (d.flip & 1) ? d.n[0] - 1 - i0 : i0, j1 = (d.flip & 2) ? d.n[1] - 1 - i1 : i1;
    const int j2 = (d.flip & 4) ? d.n[2] - 1 - i2 : i2, j3 = (d.flip & 8) ? d.n[3] - 1 - i3 : i3;
    const long long so = j0 * d.ss[0] + j1 * d.ss[1] + j2 * d.ss[2] + j3 * d.ss[3];
    const long long dof = i0 * d.ds[0] + i1 * d.ds[1] + i2 * d.ds[2] + i3 * d.ds[3];
    float v = d.sdt == B200_DT_F32 ? ((const float*)d.src)[so] : __bfloat162float(((const bf16*)d.src)[so]);
    if (d.src2) v += d.sdt == B200_DT_F32 ? ((const float*)d.src2)[so] : __bfloat162float(((const bf16*)d.src2)[so]);
    v *= d.scale;
    if (d.ddt == B200_DT_F32) ((float*)d.dst)[dof] = v; else ((bf16*)d.dst)[dof] = __float2bfloat16_rn(v);
  }
}
}  // namespace b200

extern "C" int b200_copy_desc_bytes(void) { return (int)sizeof(b200::CopyDesc); }
extern "C" int b200_copy_chunk(void) { return b200::COPY_CHUNK; }
// descs: device array of CopyDesc (layout above; b200_copy_desc_bytes each); blocks: device array of nblocks (descriptor,
// chunk) int pairs covering every descriptor's ceil(elements / b200_copy_chunk()) chunks
extern "C" int b200_copy_batch(const void* descs, const void* blocks, int nblocks, cudaStream_t st) {
  if (!descs || !blocks || nblocks < 0) return B200_ERR_ARG;
  if (nblocks == 0) return B200_OK;
  b200::copy_batch_kernel<<<nblocks, 256, 0, st>>>((const b200::CopyDesc*)descs, (const int2*)blocks);
  B200_RETURN_LAST();
}
