// Attention backward on tensor cores (bf16 operands, fp32 accumulate) for the input-gradient / training passes:
//   P = softmax(s Q K^T), O = P V  (models/layerspp.py:82-86)  ->  dV = P^T dO, dP = dO V^T,
//   dS = P o (dP - rowsum(P o dP)),  dQ = s dS K,  dK = s dS^T Q.
// The CUDA-core version (backward.cu: one warp per row) took 3.4 ms per launch at B = 128 -- 20 of the 94 ms of a C3
// training step and 2.9 of the 12.3 ms of a DPS right-hand side.  These are five 256^3 GEMMs per image: warp-level MMA
// (nvcuda::wmma, m16n16k16) fed straight from global memory (K, V, Q, dO of one image are 0.5 MB, L2 resident) is
// plenty; two passes as before, P and dS handed from the query-block pass to the key-block pass in bf16.
// Shapes: L % 64 == 0, L <= 256, C % 16 == 0, C <= 256 (every shipped network: L = C = 256).
#include <mma.h>

#include "common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {
using namespace nvcuda;

constexpr int AM_ROWS = 64;        // query (pass 1) / key (pass 2) rows per CTA
constexpr int AM_THREADS = 256;    // 8 warps: warp w -> row tile w % 4, column tiles (w / 4) * (N / 32) ...

typedef wmma::fragment<wmma::matrix_a, 16, 16, 16, bf16, wmma::row_major> FragA;
typedef wmma::fragment<wmma::matrix_a, 16, 16, 16, bf16, wmma::col_major> FragAT;
typedef wmma::fragment<wmma::matrix_b, 16, 16, 16, bf16, wmma::row_major> FragB;
typedef wmma::fragment<wmma::matrix_b, 16, 16, 16, bf16, wmma::col_major> FragBT;
typedef wmma::fragment<wmma::accumulator, 16, 16, 16, float> FragC;

// out[64][N] (fp32, shared, pitch N) = A[64][K] * B, this warp's share: row tile rt, column tiles ct0 .. ct0 + NT - 1.
// A: row-major (lda) or, TRANS_A, stored [K][64] (element (m, k) at a[k * lda + m]).
// B: B_COL: element (k, n) at b[n * ldb + k] (i.e. the transpose of a row-major [N][K] matrix); else row-major [K][N].
template <int NT, bool TRANS_A, bool B_COL>
__device__ __forceinline__ void warp_gemm(float* out, int N, const bf16* a, int lda, const bf16* b, int ldb, int K,
                                          int rt, int ct0) {
  FragC acc[NT];
#pragma unroll
  for (int t = 0; t < NT; ++t) wmma::fill_fragment(acc[t], 0.0f);
  for (int k = 0; k < K; k += 16) {
    FragA fa; FragAT fat;
    if (TRANS_A) wmma::load_matrix_sync(fat, a + (size_t)k * lda + rt * 16, lda);
    else wmma::load_matrix_sync(fa, a + (size_t)(rt * 16) * lda + k, lda);
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int n0 = (ct0 + t) * 16;
      if (B_COL) {
        FragBT fb;
        wmma::load_matrix_sync(fb, b + (size_t)n0 * ldb + k, ldb);
        if (TRANS_A) wmma::mma_sync(acc[t], fat, fb, acc[t]); else wmma::mma_sync(acc[t], fa, fb, acc[t]);
      } else {
        FragB fb;
        wmma::load_matrix_sync(fb, b + (size_t)k * ldb + n0, ldb);
        if (TRANS_A) wmma::mma_sync(acc[t], fat, fb, acc[t]); else wmma::mma_sync(acc[t], fa, fb, acc[t]);
      }
    }
  }
#pragma unroll
  for (int t = 0; t < NT; ++t) wmma::store_matrix_sync(out + (size_t)(rt * 16) * N + (ct0 + t) * 16, acc[t], N, wmma::mem_row_major);
}

// fp32 [64][N] staging -> bf16 rows of a strided global tensor, scaled
__device__ __forceinline__ void store_rows(bf16* dst, long long ld, const float* src, int N, float scale) {
  for (int i = threadIdx.x * 8; i < AM_ROWS * N; i += AM_THREADS * 8) {
    const int r = i / N, c = i - r * N;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = src[i + j] * scale;
    Vec8<bf16>::store(dst + (size_t)r * ld + c, v);
  }
}

// pass 1: one CTA = 64 query rows of one image.  LT = L / 16, CT = C / 16 column tiles.
template <int L, int C>
__global__ void __launch_bounds__(AM_THREADS) attn_bwd_q_mma_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ dout,
                                                                    bf16* __restrict__ dqkv, bf16* __restrict__ Pg,
                                                                    bf16* __restrict__ dSg, float scale) {
  extern __shared__ __align__(32) uint8_t smraw[];
  float* S = reinterpret_cast<float*>(smraw);                 // [64][L] scores, later the dQ staging ([64][C], C <= L)
  float* dP = S + AM_ROWS * L;                                // [64][L]
  bf16* dSb = reinterpret_cast<bf16*>(dP + AM_ROWS * L);      // [64][L]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int blocks = L / AM_ROWS;
  const int b = blockIdx.x / blocks, i0 = (blockIdx.x % blocks) * AM_ROWS;
  constexpr int ld = 3 * C;
  const bf16* base = qkv + (size_t)b * L * ld;
  const bf16* q = base + (size_t)i0 * ld;
  const bf16* dO = dout + ((size_t)b * L + i0) * C;
  const int rt = warp & 3, half = warp >> 2;
  warp_gemm<L / 32, false, true>(S, L, q, ld, base + C, ld, C, rt, half * (L / 32));          // Q K^T
  warp_gemm<L / 32, false, true>(dP, L, dO, C, base + 2 * C, ld, C, rt, half * (L / 32));     // dO V^T
  __syncthreads();
  for (int r = warp; r < AM_ROWS; r += AM_THREADS / 32) {
    float* sr = S + (size_t)r * L;
    const float* dr = dP + (size_t)r * L;
    float mx = -INFINITY;
    for (int j = lane; j < L; j += 32) mx = fmaxf(mx, sr[j] * scale);
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float den = 0.f;
    for (int j = lane; j < L; j += 32) { const float e = __expf(sr[j] * scale - mx); sr[j] = e; den += e; }
    for (int o = 16; o; o >>= 1) den += __shfl_xor_sync(0xffffffffu, den, o);
    const float inv = 1.0f / den;
    float dot = 0.f;
    for (int j = lane; j < L; j += 32) { const float p = sr[j] * inv; sr[j] = p; dot = fmaf(p, dr[j], dot); }
    for (int o = 16; o; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    bf16* pg = Pg + ((size_t)b * L + i0 + r) * L;
    bf16* sg = dSg + ((size_t)b * L + i0 + r) * L;
    for (int j = lane; j < L; j += 32) {
      const bf16 pb = __float2bfloat16_rn(sr[j]);
      const bf16 db = __float2bfloat16_rn(sr[j] * (dr[j] - dot));
      pg[j] = pb; sg[j] = db; dSb[(size_t)r * L + j] = db;
    }
  }
  __syncthreads();
  warp_gemm<C / 32, false, false>(S, C, dSb, L, base + C, ld, L, rt, half * (C / 32));        // dS K  -> [64][C]
  __syncthreads();
  store_rows(dqkv + ((size_t)b * L + i0) * ld, ld, S, C, scale);
}

// pass 2: one CTA = 64 key rows of one image: dV = P^T dO, dK = s dS^T Q
template <int L, int C>
__global__ void __launch_bounds__(AM_THREADS) attn_bwd_kv_mma_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ dout,
                                                                     bf16* __restrict__ dqkv, const bf16* __restrict__ Pg,
                                                                     const bf16* __restrict__ dSg, float scale) {
  extern __shared__ __align__(32) uint8_t smraw[];
  float* stage = reinterpret_cast<float*>(smraw);             // [64][C]
  const int warp = threadIdx.x >> 5;
  const int blocks = L / AM_ROWS;
  const int b = blockIdx.x / blocks, j0 = (blockIdx.x % blocks) * AM_ROWS;
  constexpr int ld = 3 * C;
  const bf16* base = qkv + (size_t)b * L * ld;
  const bf16* Pb = Pg + (size_t)b * L * L + j0;               // element (m = key, k = query) at Pb[k * L + m]
  const bf16* Sb = dSg + (size_t)b * L * L + j0;
  const int rt = warp & 3, half = warp >> 2;
  warp_gemm<C / 32, true, false>(stage, C, Pb, L, dout + (size_t)b * L * C, C, L, rt, half * (C / 32));
  __syncthreads();
  store_rows(dqkv + ((size_t)b * L + j0) * ld + 2 * C, ld, stage, C, 1.0f);
  __syncthreads();
  warp_gemm<C / 32, true, false>(stage, C, Sb, L, base, ld, L, rt, half * (C / 32));
  __syncthreads();
  store_rows(dqkv + ((size_t)b * L + j0) * ld + C, ld, stage, C, scale);
}

}  // namespace b200

using namespace b200;

// 1 when b200_attention_bwd takes the tensor-core path for this shape (bf16 only)
extern "C" int b200_attention_bwd_mma_supported(int L, int C) { return (L == 256 && C == 256) ? 1 : 0; }

// scratch: at least 2 * B * L * L bf16 (the caller's fp32 scratch of b200_attention_bwd is four times that)
extern "C" int b200_attention_bwd_mma(const void* qkv, const void* dout, void* dqkv, void* scratch, int B, int L, int C,
                                      float scale, cudaStream_t st) {
  if (!qkv || !dout || !dqkv || !scratch || B < 0) return B200_ERR_ARG;
  if (!b200_attention_bwd_mma_supported(L, C)) return B200_ERR_UNSUPPORTED;
  if (B == 0) return B200_OK;
  bf16* P = static_cast<bf16*>(scratch);
  bf16* dS = P + (size_t)B * L * L;
  constexpr int LL = 256, CC = 256;
  const int sm1 = 2 * AM_ROWS * LL * 4 + AM_ROWS * LL * 2, sm2 = AM_ROWS * CC * 4;
  static int st1[B200_MAX_DEVICES], st2[B200_MAX_DEVICES];
  if (int rc = b200_set_smem_once(st1, attn_bwd_q_mma_kernel<LL, CC>, sm1)) return rc;
  if (int rc = b200_set_smem_once(st2, attn_bwd_kv_mma_kernel<LL, CC>, sm2)) return rc;
  const int grid = B * (L / AM_ROWS);
  attn_bwd_q_mma_kernel<LL, CC><<<grid, AM_THREADS, sm1, st>>>((const bf16*)qkv, (const bf16*)dout, (bf16*)dqkv, P, dS, scale);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return -(1000 + (int)e);
  attn_bwd_kv_mma_kernel<LL, CC><<<grid, AM_THREADS, sm2, st>>>((const bf16*)qkv, (const bf16*)dout, (bf16*)dqkv, P, dS, scale);
  B200_RETURN_LAST();
}
