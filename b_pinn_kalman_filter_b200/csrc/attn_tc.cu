// Single-head spatial self-attention on the tcgen05 tensor cores for the one shape every shipped network uses:
// L = H*W = 256 tokens (16x16 feature map), C = 256 channels, bf16.  softmax(q k^T C^-1/2) v per image
// (AttnBlockpp.forward models/layerspp.py:82-86, AttnBlock.forward models/layers.py:566-570).
//
// One CTA per (image, 128-query block).  q, k, v are column slices of the fused projection [B, L, 3C]:
//   TMA   : Q block [128 x 256] and K as two [128 keys x 256] halves -> shared memory (SW128, 4 channel chunks each)
//   MMA 1 : S_j = Q K_j^T  (M = 128 queries, N = 128 keys, K = 256 channels), j = 0, 1 -> TMEM columns [0, 256)
//   TMA   : as soon as S_j is committed, V_j ([128 keys x 256]) overwrites K_j's buffer (overlaps the softmax)
//   SIMT  : 4 warps, one query row per thread: row max, exp, row sum straight out of TMEM; P (bf16) is written over
//           the Q buffer in the K-major SW128 layout the second MMA reads
//   MMA 2 : O = P V  (M = 128, N = 256 channels, K = 256 keys); V is consumed in its natural [key][channel] layout
//           as an MN-major B operand (no transpose anywhere) -> TMEM columns [256, 512)
//   SIMT  : O / rowsum -> bf16 -> global [B, L, C]
// Shared memory: 3 x 64 KB.  Everything between the two GEMMs stays on chip; the [L, L] score matrix never exists
// in HBM (the reference materialises it twice: einsum output and softmax output).
#include "tc_common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

constexpr int AT_L = 256, AT_C = 256, AT_BM = 128;
constexpr int AT_CHUNK = 128 * 64 * 2;              // one [128 rows x 64 elements] SW128 chunk: 16 KB
constexpr int AT_BUF = 4 * AT_CHUNK;                // [128 x 256] bf16: 64 KB
constexpr int AT_SMEM = 3 * AT_BUF + 256 + 1024;

__global__ void __launch_bounds__(192, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap map, bf16* __restrict__ out, float scale_log2e) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base, sK0 = base + AT_BUF, sK1 = base + 2 * AT_BUF;     // P aliases Q; V_j aliases K_j
  const uint32_t bar0 = base + 3 * AT_BUF;
  enum { B_QK0 = 0, B_K1, B_S0, B_S1, B_V0, B_V1, B_P, B_O, NBAR };
  auto bar = [&](int i) { return bar0 + 8u * i; };
  const uint32_t tmem_slot = bar0 + 8u * NBAR;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));
  uint8_t* sP = smem_raw + (sQ - smem_u32(smem_raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x >> 1, qb = blockIdx.x & 1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NBAR; ++i) mbar_init(bar(i), i == B_P ? 4 : 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_ptr;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(bar(B_QK0), 2 * AT_BUF);
      for (int kc = 0; kc < 4; ++kc) tma_load_3d(sQ + kc * AT_CHUNK, &map, bar(B_QK0), kc * 64, qb * AT_BM, b);
      for (int kc = 0; kc < 4; ++kc) tma_load_3d(sK0 + kc * AT_CHUNK, &map, bar(B_QK0), AT_C + kc * 64, 0, b);
      mbar_expect_tx(bar(B_K1), AT_BUF);
      for (int kc = 0; kc < 4; ++kc) tma_load_3d(sK1 + kc * AT_CHUNK, &map, bar(B_K1), AT_C + kc * 64, 128, b);
      // V_j replaces K_j once S_j no longer needs it
      mbar_wait(bar(B_S0), 0);
      mbar_expect_tx(bar(B_V0), AT_BUF);
      for (int cc = 0; cc < 4; ++cc) tma_load_3d(sK0 + cc * AT_CHUNK, &map, bar(B_V0), 2 * AT_C + cc * 64, 0, b);
      mbar_wait(bar(B_S1), 0);
      mbar_expect_tx(bar(B_V1), AT_BUF);
      for (int cc = 0; cc < 4; ++cc) tma_load_3d(sK1 + cc * AT_CHUNK, &map, bar(B_V1), 2 * AT_C + cc * 64, 128, b);
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // S_j = Q K_j^T : A, B K-major, N = 128
      const uint32_t idesc_s = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      for (int j = 0; j < 2; ++j) {
        mbar_wait(bar(j == 0 ? B_QK0 : B_K1), 0);
        tc_fence_after();
        const uint32_t sK = j == 0 ? sK0 : sK1;
        for (int kc = 0; kc < 4; ++kc) {
          const uint64_t da = make_sw128_desc(sQ + kc * AT_CHUNK), db = make_sw128_desc(sK + kc * AT_CHUNK);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma_bf16(tmem + j * 128, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc_s, (kc | k) != 0 ? 1u : 0u);
        }
        tc_commit(bar(j == 0 ? B_S0 : B_S1));
      }
      // O = P V : A = P K-major (keys contiguous), B = V MN-major (channels contiguous), N = 256
      const uint32_t idesc_o = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((uint32_t)(256 >> 3) << 17) |
                               ((uint32_t)(128 >> 4) << 24);
      mbar_wait(bar(B_P), 0);
      tc_fence_after();
      for (int j = 0; j < 2; ++j) {
        mbar_wait(bar(j == 0 ? B_V0 : B_V1), 0);
        tc_fence_after();
        const uint32_t sV = j == 0 ? sK0 : sK1;
#pragma unroll
        for (int t = 0; t < 8; ++t) {               // 16 keys per MMA
          const int key = j * 128 + t * 16;
          // A: P chunk (key / 64), 32 B per 16 keys inside the 128 B row
          const uint64_t da = make_sw128_desc(sQ + (key >> 6) * AT_CHUNK) + (uint64_t)(2 * ((key & 63) >> 4));
          // B: V_j rows [t*16, t*16+16): 2 KB per 16 keys; 64-channel groups 16 KB apart (LBO), 8-key atoms 1 KB (SBO)
          const uint64_t db = make_sw128_desc_ex(sV + t * 2048, AT_CHUNK, 1024);
          tc_mma_bf16(tmem + 256, da, db, idesc_o, (j | t) != 0 ? 1u : 0u);
        }
      }
      tc_commit(bar(B_O));
    }
  } else {
    const int q = warp & 3;
    const int r = q * 32 + lane;                               // query row inside the block == TMEM lane
    const uint32_t t_row = tmem + ((uint32_t)(q * 32) << 16);
    mbar_wait(bar(B_S0), 0);
    mbar_wait(bar(B_S1), 0);
    tc_fence_after();
    float mx = -INFINITY;
#pragma unroll 1
    for (int c0 = 0; c0 < AT_L; c0 += 32) {
      uint32_t v[32];
      tc_ld32(t_row + c0, v);
      tc_wait_ld();
#pragma unroll
      for (int j = 0; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(v[j]));
    }
    const float mb = mx * scale_log2e;
    float sum = 0.f;
#pragma unroll 1
    for (int c0 = 0; c0 < AT_L; c0 += 32) {
      uint32_t v[32];
      tc_ld32(t_row + c0, v);
      tc_wait_ld();
      uint8_t* chunk = sP + (c0 >> 6) * AT_CHUNK + r * 128;
#pragma unroll
      for (int u = 0; u < 4; ++u) {                            // 8 keys = one 16-byte unit
        uint4 pk;
        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float e0 = exp2f(fmaf(__uint_as_float(v[u * 8 + 2 * i]), scale_log2e, -mb));
          const float e1 = exp2f(fmaf(__uint_as_float(v[u * 8 + 2 * i + 1]), scale_log2e, -mb));
          hh[i] = __floats2bfloat162_rn(e0, e1);
          // the row sum uses the values the MMA will see (bf16-rounded), so rows of P / sum add up to 1
          const float2 f = __bfloat1622float2(hh[i]);
          sum += f.x + f.y;
        }
        const int unit = ((c0 & 63) >> 3) + u;                 // 16-byte unit index inside the 128 B row
        *reinterpret_cast<uint4*>(chunk + ((unit ^ (r & 7)) << 4)) = pk;
      }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(bar(B_P));
    const float inv = 1.0f / sum;
    mbar_wait(bar(B_O), 0);
    tc_fence_after();
    bf16* orow = out + ((size_t)b * AT_L + qb * AT_BM + r) * AT_C;
#pragma unroll 1
    for (int c0 = 0; c0 < AT_C; c0 += 32) {
      uint32_t v[32];
      tc_ld32(t_row + 256 + c0, v);
      tc_wait_ld();
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        uint4 pk;
        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          hh[i] = __floats2bfloat162_rn(__uint_as_float(v[u * 8 + 2 * i]) * inv, __uint_as_float(v[u * 8 + 2 * i + 1]) * inv);
        *reinterpret_cast<uint4*>(orow + c0 + u * 8) = pk;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512));
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_attention_tc_supported(int L, int C) { return (L == AT_L && C == AT_C) ? 1 : 0; }

// qkv: [B, L, 3C] bf16 (q | k | v along the channel axis); out: [B, L, C] bf16.
extern "C" int b200_attention_tc(const void* qkv, void* out, int B, int L, int C, float scale, cudaStream_t st) {
  if (!qkv || !out || B < 0) return B200_ERR_ARG;
  if (!b200_attention_tc_supported(L, C)) return B200_ERR_UNSUPPORTED;
  if (B == 0) return B200_OK;
  static int attr_state[B200_MAX_DEVICES];
  if (int rc = b200_set_smem_once(attr_state, attn_tc_kernel, AT_SMEM)) return rc;
  CUtensorMap map;
  const uint64_t dims[3] = {(uint64_t)3 * C, (uint64_t)L, (uint64_t)B};
  const uint64_t strides[2] = {(uint64_t)3 * C * 2, (uint64_t)L * 3 * C * 2};
  const uint32_t box[3] = {64, 128, 1};
  int rc = tc_tensor_map(&map, qkv, 3, dims, strides, box, 0);
  if (rc) return rc;
  attn_tc_kernel<<<B * 2, 192, AT_SMEM, st>>>(map, (bf16*)out, scale * 1.4426950408889634f);
  B200_RETURN_LAST();
}
