// Single-head spatial self-attention core on CUDA cores (fp32 math): softmax(q k^T * C^-1/2) v per image.
// Replaces the two einsums + softmax of AttnBlockpp / AttnBlock (models/layerspp.py:82-86, models/layers.py:566-570).
// Role: strict-fp32 path and the small-L cases (4x4 / 8x8 bottleneck attention, L = 16 / 64) that do not fill a
// tcgen05 tile; the L = 256 bf16 case runs on tensor cores (conv_tc.cu batched GEMM + fused row softmax).
// One warp per query row; K/V rows stream through L1/L2 (per image they are only L*C*sizeof(T) <= 256 KB).
#include "common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

constexpr int AT_ROWS = 8;  // query rows (= warps) per block

template <typename T>
__global__ void __launch_bounds__(AT_ROWS * 32) attn_simt_kernel(const T* __restrict__ q, const T* __restrict__ k,
                                                                 const T* __restrict__ v, T* __restrict__ out, int L,
                                                                 int C, int ldq, int ldk, int ldv, int ldo,
                                                                 float scale) {
  extern __shared__ float sm[];  // [AT_ROWS][C] q rows, then [AT_ROWS][L] scores
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles = (L + AT_ROWS - 1) / AT_ROWS;
  const int b = blockIdx.x / tiles, i = (blockIdx.x % tiles) * AT_ROWS + warp;
  float* sq = sm + warp * C;
  float* ss = sm + AT_ROWS * C + warp * L;
  if (i >= L) return;  // whole warp exits together; no block-level sync below
  const T* qr = q + ((size_t)b * L + i) * ldq;
  for (int c = lane * 8; c < C; c += 256) {
    float t[8];
    Vec8<T>::load(qr + c, t);
#pragma unroll
    for (int j = 0; j < 8; ++j) sq[c + j] = t[j];
  }
  __syncwarp();
  float mx = -INFINITY;
  for (int j = lane; j < L; j += 32) {
    const T* kr = k + ((size_t)b * L + j) * ldk;
    float acc = 0.f;
    for (int c = 0; c < C; c += 8) {
      float t[8];
      Vec8<T>::load(kr + c, t);
#pragma unroll
      for (int u = 0; u < 8; ++u) acc = fmaf(sq[c + u], t[u], acc);
    }
    acc *= scale;
    ss[j] = acc;
    mx = fmaxf(mx, acc);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
  for (int j = lane; j < L; j += 32) {
    float e = expf(ss[j] - mx);
    ss[j] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  const float inv = 1.0f / sum;
  __syncwarp();
  for (int c = lane * 8; c < C; c += 256) {
    float acc[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) acc[u] = 0.f;
    for (int j = 0; j < L; ++j) {
      float t[8];
      Vec8<T>::load(v + ((size_t)b * L + j) * ldv + c, t);
      const float p = ss[j];
#pragma unroll
      for (int u = 0; u < 8; ++u) acc[u] = fmaf(p, t[u], acc[u]);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) acc[u] *= inv;
    Vec8<T>::store(out + ((size_t)b * L + i) * ldo + c, acc);
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_attention_simt(const void* q, const void* k, const void* v, void* out, int dtype, int B, int L,
                                   int C, int ldq, int ldk, int ldv, int ldo, float scale, cudaStream_t st) {
  if (!q || !k || !v || !out || B < 0 || L <= 0 || C <= 0 || (C % 8) || (ldq % 8) || (ldk % 8) || (ldv % 8) || (ldo % 8))
    return B200_ERR_ARG;
  if (B == 0) return B200_OK;
  const size_t smem = (size_t)AT_ROWS * (C + L) * sizeof(float);
  if (smem > 200 * 1024) return B200_ERR_UNSUPPORTED;
  const int tiles = (L + AT_ROWS - 1) / AT_ROWS;
  if (dtype == B200_DT_F32) {
    if (smem > 48 * 1024) cudaFuncSetAttribute(attn_simt_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    attn_simt_kernel<float><<<B * tiles, AT_ROWS * 32, smem, st>>>((const float*)q, (const float*)k, (const float*)v,
                                                                  (float*)out, L, C, ldq, ldk, ldv, ldo, scale);
  } else if (dtype == B200_DT_BF16) {
    if (smem > 48 * 1024) cudaFuncSetAttribute(attn_simt_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    attn_simt_kernel<bf16><<<B * tiles, AT_ROWS * 32, smem, st>>>((const bf16*)q, (const bf16*)k, (const bf16*)v,
                                                                 (bf16*)out, L, C, ldq, ldk, ldv, ldo, scale);
  } else {
    return B200_ERR_ARG;
  }
  B200_RETURN_LAST();
}
