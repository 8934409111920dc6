// Small memory-bound kernels around the score network: layout changes at the API boundary, the time embedding
// (models/layerspp.py:32-41 Gaussian Fourier, models/layers.py:500-514 sinusoidal), the tiny temb MLP / Dense_0
// projections (models/ncsnpp.py:86-91, 251-255; models/layerspp.py:227,263), residual combine and nearest x2.
#include "common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

// x: [B, C, H, W] fp32  ->  out: [B, H, W, Cp] T with out = x*scale + shift, channels c >= C zero-filled.
// One thread = one 8-channel group of one pixel (a single 16 / 32-byte store; the padded groups are pure zero stores).
template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, T* __restrict__ out, int B, int C, int HW, int Cp,
                                    float scale, float shift) {
  const int G = Cp >> 3;
  const long long total = (long long)B * HW * G;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    long long r = i / G;
    const int p = (int)(r % HW);
    const long long b = r / HW;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = g * 8 + j;
      v[j] = c < C ? fmaf(__ldg(x + (b * C + c) * HW + p), scale, shift) : 0.0f;
    }
    Vec8<T>::store(out + i * 8, v);
  }
}

// x: [B, HW, Cs] T (first C channels used) -> out [B, C, HW] fp32, out = x * mul[b] (mul may be null)
template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ x, float* __restrict__ out, int B, int C, int HW, int Cs,
                                    const float* __restrict__ mul) {
  const long long total = (long long)B * C * HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int p = (int)(i % HW);
    long long r = i / HW;
    int c = (int)(r % C);
    long long b = r / C;
    float v = to_f(x[(b * HW + p) * Cs + c]);
    out[i] = mul ? v * mul[b] : v;
  }
}

// Time / noise-level embedding.  labels: [B] (or a per-step table read at *step).  kind 0: Gaussian Fourier
// [sin(2 pi log(sigma) W), cos(...)] (dim 2*nf); kind 1: sinusoidal positional (dim nf).
// Also emits out_mul[b] = 1/used_sigma (kind 0: 1/label; kind 1: 1/sigmas[(long)label]) or 1 when !scale_by_sigma
// -- the `h / used_sigmas` of models/ncsnpp.py:377-379.
__global__ void temb_embed_kernel(const float* __restrict__ labels, const int* __restrict__ step, int kind,
                                  const float* __restrict__ Wf, int nf, const float* __restrict__ sigmas, int n_sigmas,
                                  int scale_by_sigma, float* __restrict__ emb, float* __restrict__ out_mul, int B) {
  const int b = blockIdx.x;
  const float lab = step ? labels[*step] : labels[b];
  if (kind == 0) {
    const float lg = logf(lab);
    for (int i = threadIdx.x; i < nf; i += blockDim.x) {
      // x[:, None] * W[None, :] * 2 * np.pi   (left-to-right float products, layerspp.py:40)
      const float proj = lg * Wf[i] * 2.0f * 3.14159265358979323846f;
      emb[(size_t)b * 2 * nf + i] = sinf(proj);
      emb[(size_t)b * 2 * nf + nf + i] = cosf(proj);
    }
    if (threadIdx.x == 0 && out_mul) out_mul[b] = scale_by_sigma ? 1.0f / lab : 1.0f;
  } else {
    const int half = nf / 2;
    const float f = logf(10000.0f) / (float)(half - 1);
    for (int i = threadIdx.x; i < half; i += blockDim.x) {
      const float e = lab * expf((float)i * -f);
      emb[(size_t)b * nf + i] = sinf(e);
      emb[(size_t)b * nf + half + i] = cosf(e);
    }
    if ((nf & 1) && threadIdx.x == 0) emb[(size_t)b * nf + nf - 1] = 0.0f;
    if (threadIdx.x == 0 && out_mul) {
      float m = 1.0f;
      if (scale_by_sigma && sigmas) {
        long long idx = (long long)lab;
        idx = idx < 0 ? 0 : (idx >= n_sigmas ? n_sigmas - 1 : idx);
        m = 1.0f / sigmas[idx];
      }
      out_mul[b] = m;
    }
  }
}

// out[b, n] = sum_k act(x[b, k]) * W[n, k] + bias[n]   (fp32; one warp per output, K-contiguous weights)
__global__ void __launch_bounds__(256) linear_kernel(const float* __restrict__ x, const float* __restrict__ W,
                                                     const float* __restrict__ bias, float* __restrict__ out, int B,
                                                     int K, int N, int act_in, int act_out) {
  extern __shared__ float sx[];  // act(x[b, :])
  const int b = blockIdx.y;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    float v = x[(size_t)b * K + k];
    sx[k] = act_in ? silu(v) : v;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int n = blockIdx.x * nw + warp; n < N; n += gridDim.x * nw) {
    const float* wr = W + (size_t)n * K;
    float acc = 0.f;
    for (int k = lane; k < K; k += 32) acc = fmaf(sx[k], __ldg(wr + k), acc);
    acc = warp_sum(acc);
    if (lane == 0) {
      acc += bias ? bias[n] : 0.0f;
      out[(size_t)b * N + n] = act_out ? silu(acc) : acc;
    }
  }
}

// Tiled fp32 GEMM for the same op when the batch is large: 64 samples x 64 outputs per block, K in chunks of 16,
// 4x4 outputs per thread.  (The per-sample kernel above re-reads W once per sample.)
__global__ void __launch_bounds__(256) linear_tiled_kernel(const float* __restrict__ x, const float* __restrict__ W,
                                                           const float* __restrict__ bias, float* __restrict__ out,
                                                           int B, int K, int N, int act_in, int act_out) {
  __shared__ __align__(16) float As[16][64 + 4];
  __shared__ __align__(16) float Bs[16][64 + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int b0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int lr = tid >> 2, lk = (tid & 3) * 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += 16) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), w = make_float4(0.f, 0.f, 0.f, 0.f);
    if (b0 + lr < B) a = __ldg(reinterpret_cast<const float4*>(x + (size_t)(b0 + lr) * K + k0 + lk));
    if (n0 + lr < N) w = __ldg(reinterpret_cast<const float4*>(W + (size_t)(n0 + lr) * K + k0 + lk));
    if (act_in) { a.x = silu(a.x); a.y = silu(a.y); a.z = silu(a.z); a.w = silu(a.w); }
    __syncthreads();
    As[lk][lr] = a.x; As[lk + 1][lr] = a.y; As[lk + 2][lr] = a.z; As[lk + 3][lr] = a.w;
    Bs[lk][lr] = w.x; Bs[lk + 1][lr] = w.y; Bs[lk + 2][lr] = w.z; Bs[lk + 3][lr] = w.w;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const float4 av = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 bv = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float aa[4] = {av.x, av.y, av.z, av.w}, bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(aa[i], bb[j], acc[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int b = b0 + ty * 4 + i;
    if (b >= B) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j] + (bias ? bias[n] : 0.0f);
      out[(size_t)b * N + n] = act_out ? silu(v) : v;
    }
  }
}

// out = (a + b) * s    (8-wide)
template <typename T>
__global__ void add_scale_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out, long long nvec,
                                 float s) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float x[8], y[8];
    Vec8<T>::load(a + i * 8, x);
    Vec8<T>::load(b + i * 8, y);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = (x[j] + y[j]) * s;
    Vec8<T>::store(out + i * 8, x);
  }
}

// nearest-neighbour x2 upsample, channels-last (F.interpolate(..., 'nearest'), models/layers.py:585)
template <typename T>
__global__ void nearest_up2_kernel(const T* __restrict__ x, T* __restrict__ out, int B, int H, int W, int C8) {
  const long long total = (long long)B * (2 * H) * (2 * W) * C8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int c8 = (int)(i % C8);
    long long r = i / C8;
    int ox = (int)(r % (2 * W)); r /= (2 * W);
    int oy = (int)(r % (2 * H));
    long long b = r / (2 * H);
    const uint4* src4;
    if (sizeof(T) == 2) {
      src4 = reinterpret_cast<const uint4*>(x + (((b * H + (oy >> 1)) * W + (ox >> 1)) * C8 + c8) * 8);
      *reinterpret_cast<uint4*>(out + i * 8) = *src4;
    } else {
      float v[8];
      Vec8<T>::load(x + (((b * H + (oy >> 1)) * W + (ox >> 1)) * C8 + c8) * 8, v);
      Vec8<T>::store(out + i * 8, v);
    }
  }
}

static inline int grid_for(long long work) {
  long long want = (work + 255) / 256;
  long long cap = (long long)B200_NUM_SMS * 16;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_nchw_to_nhwc(const float* x, void* out, int dtype, int B, int C, int H, int W, int Cp, float scale,
                                 float shift, cudaStream_t st) {
  if (!x || !out || B < 0 || C <= 0 || Cp < C) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  if (Cp % 8) return B200_ERR_ARG;
  const long long total = (long long)B * H * W * (Cp / 8);
  if (dtype == B200_DT_F32) nchw_to_nhwc_kernel<float><<<grid_for(total), 256, 0, st>>>(x, (float*)out, B, C, H * W, Cp, scale, shift);
  else if (dtype == B200_DT_BF16) nchw_to_nhwc_kernel<bf16><<<grid_for(total), 256, 0, st>>>(x, (bf16*)out, B, C, H * W, Cp, scale, shift);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

extern "C" int b200_nhwc_to_nchw(const void* x, float* out, int dtype, int B, int C, int H, int W, int Cs,
                                 const float* mul, cudaStream_t st) {
  if (!x || !out || B < 0 || C <= 0 || Cs < C) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  const long long total = (long long)B * H * W * C;
  if (dtype == B200_DT_F32) nhwc_to_nchw_kernel<float><<<grid_for(total), 256, 0, st>>>((const float*)x, out, B, C, H * W, Cs, mul);
  else if (dtype == B200_DT_BF16) nhwc_to_nchw_kernel<bf16><<<grid_for(total), 256, 0, st>>>((const bf16*)x, out, B, C, H * W, Cs, mul);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

extern "C" int b200_temb_embed(const float* labels, const int* step, int kind, const float* fourier_w, int nf,
                               const float* sigmas, int n_sigmas, int scale_by_sigma, float* emb, float* out_mul, int B,
                               cudaStream_t st) {
  if (!labels || !emb || B < 0 || nf <= 0 || (kind == 0 && !fourier_w) || (kind != 0 && kind != 1)) return B200_ERR_ARG;
  if (kind == 1 && nf < 4) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  temb_embed_kernel<<<B, 128, 0, st>>>(labels, step, kind, fourier_w, nf, sigmas, n_sigmas, scale_by_sigma, emb, out_mul, B);
  B200_RETURN_LAST();
}

extern "C" int b200_linear_f32(const float* x, const float* W, const float* bias, float* out, int B, int K, int N,
                               int act_in, int act_out, cudaStream_t st) {
  if (!x || !W || !out || B < 0 || K <= 0 || N <= 0 || K > 8192) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  if (B >= 16 && K % 16 == 0) {
    dim3 grid((N + 63) / 64, (B + 63) / 64);
    linear_tiled_kernel<<<grid, 256, 0, st>>>(x, W, bias, out, B, K, N, act_in, act_out);
    B200_RETURN_LAST();
  }
  int gx = (N + 7) / 8;
  if (gx > 64) gx = 64;
  dim3 grid(gx, B);
  linear_kernel<<<grid, 256, K * sizeof(float), st>>>(x, W, bias, out, B, K, N, act_in, act_out);
  B200_RETURN_LAST();
}

extern "C" int b200_add_scale(const void* a, const void* b, void* out, int dtype, long long n, float s,
                              cudaStream_t st) {
  if (!a || !b || !out || n < 0 || (n % 8)) return B200_ERR_ARG;
  if (n == 0) return B200_OK;
  if (dtype == B200_DT_F32) add_scale_kernel<float><<<grid_for(n / 8), 256, 0, st>>>((const float*)a, (const float*)b, (float*)out, n / 8, s);
  else if (dtype == B200_DT_BF16) add_scale_kernel<bf16><<<grid_for(n / 8), 256, 0, st>>>((const bf16*)a, (const bf16*)b, (bf16*)out, n / 8, s);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}

extern "C" int b200_nearest_up2_nhwc(const void* x, void* out, int dtype, int B, int H, int W, int C, cudaStream_t st) {
  if (!x || !out || B < 0 || C % 8) return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  const long long total = (long long)B * 4 * H * W * (C / 8);
  if (dtype == B200_DT_F32) nearest_up2_kernel<float><<<grid_for(total), 256, 0, st>>>((const float*)x, (float*)out, B, H, W, C / 8);
  else if (dtype == B200_DT_BF16) nearest_up2_kernel<bf16><<<grid_for(total), 256, 0, st>>>((const bf16*)x, (bf16*)out, B, H, W, C / 8);
  else return B200_ERR_ARG;
  B200_RETURN_LAST();
}
