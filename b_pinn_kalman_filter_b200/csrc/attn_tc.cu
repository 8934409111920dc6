// Single-head spatial self-attention on the tcgen05 tensor cores for the one shape every shipped network uses:
// L = H*W = 256 tokens (16x16 feature map), C = 256 channels, bf16.  softmax(q k^T C^-1/2) v per image
// (AttnBlockpp.forward models/layerspp.py:82-86, AttnBlock.forward models/layers.py:566-570).
//
// A tile is (image, 128-query block); q, k, v are column slices of the fused projection [B, L, 3C].  Persistent CTAs
// (one per SM) take tiles round-robin; per tile
//   MMA 1 : S = Q K^T  (M = 128 queries, N = 2 x 128 keys, K = 256 channels)            -> a 256-column TMEM accumulator
//   SIMT  : one query row per thread: row max, exp2, row sum straight out of TMEM; P (bf16) goes to shared memory in
//           the K-major SW128 layout the second MMA reads
//   MMA 2 : O = P V  (M = 128, N = 2 x 128 channels, K = 256 keys); V is consumed in its natural [key][channel] layout
//           as an MN-major B operand (no transpose anywhere)                              -> the same accumulator
//   SIMT  : O / rowsum -> bf16 -> global [B, L, C]
// Everything between the two GEMMs stays on chip; the [L, L] score matrix never exists in HBM (the reference
// materialises it twice: einsum output and softmax output).
//
// Operands stream; nothing but P is resident:
//   * a tile's operands arrive as 16 KB slots of an 8-slot TMA ring: per 64-channel slice [Q 128x64], [K 0..127 x64],
//     [K 128..255 x64] for MMA 1, then per 64 keys [V 64 x ch 0..127], [V 64 x ch 128..255] for MMA 2; one producer
//     thread runs ahead across tiles in the order the issuer consumes (QK(0), QK(1), PV(0), QK(2), PV(1), ...);
//   * TMEM is two 256-column accumulators; S_n and, once the softmax has read it, O_n share accumulator n & 1, so
//     QK(n+1) runs on the tensor core while the softmax of tile n runs on the CUDA cores;
//   * two groups of four SIMT warps alternate tiles (softmax + drain of tile n on group n & 1).  P is ONE 64 KB
//     buffer: group n & 1 writes P_n after PV(n-1) has retired;
//   * output rows leave through a 4 KB per-warp staging area as whole 128-byte lines.
// History (profiles/r02n_attention.txt): the first kernel kept Q, K, V of one tile resident in three aliasing 64 KB
// buffers, one CTA per tile, every phase back to back: 151 us per launch at B = 1024.  This one: 104 us, against an HBM
// floor of 82 us (403 MB of qkv in, 134 MB out: at 128 flop/B the op is HBM-bound once qkv no longer fits L2).
#include "tc_common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

constexpr int AT_L = 256, AT_C = 256, AT_BM = 128;
constexpr int AT_CHUNK = 128 * 64 * 2;              // one [128 rows x 64 elements] SW128 chunk: 16 KB
constexpr int AT_BUF = 4 * AT_CHUNK;                // [128 x 256] bf16: 64 KB

// -DB200_ATTN_TRACE: CTA 0 stamps clock64() at its phase boundaries into the buffer given to b200_attention_tc_trace
// (profiles/prof_attn_phases.py prints the timeline); compiled out otherwise.
#ifdef B200_ATTN_TRACE
__device__ long long* g_attn_trace = nullptr;
#define ATTN_TRACE(slot) do { if (trace) trace[(slot)] = clock64(); } while (0)
#else
#define ATTN_TRACE(slot) do { } while (0)
#endif
constexpr int AS_SLOT = AT_CHUNK;                   // 16 KB
constexpr int AS_NSLOT = 8;
constexpr int AS_THREADS = 320;                     // warp 0 TMA, warp 1 MMA, warps 2-5 / 6-9 the two SIMT groups
constexpr int AS_STAGING = 8 * 4096;
constexpr int AS_SMEM = AT_BUF + AS_NSLOT * AS_SLOT + AS_STAGING + 256 + 1024;

__global__ void __launch_bounds__(AS_THREADS, 1)
attn_tc_stream_kernel(const __grid_constant__ CUtensorMap map, const __grid_constant__ CUtensorMap map_v,
                      bf16* __restrict__ out, float scale_log2e, int ntiles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t sPa = base;                                   // P [128 queries x 256 keys] bf16, 4 SW128 chunks
  const uint32_t ring = base + AT_BUF;
  const uint32_t bar0 = ring + AS_NSLOT * AS_SLOT + AS_STAGING;
  enum { B_FULL = 0, B_EMPTY = AS_NSLOT, B_S = 2 * AS_NSLOT, B_O = B_S + 2, B_FREE = B_O + 2, B_P = B_FREE + 2, NBAR = B_P + 1 };
  auto bar = [&](int i) { return bar0 + 8u * i; };
  const uint32_t tmem_slot = bar0 + 8u * NBAR;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nt = (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // tiles of this CTA
#ifdef B200_ATTN_TRACE
  long long* trace = (blockIdx.x == 0 && (threadIdx.x & 31) == 0 && (threadIdx.x < 96 || threadIdx.x == 192)) ? g_attn_trace : nullptr;
#endif

  if (threadIdx.x == 0) {
    for (int i = 0; i < NBAR; ++i) mbar_init(bar(i), i >= B_FREE ? 4 : 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
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
      uint32_t it = 0;
      auto acquire = [&]() {                                   // next slot, free again, armed for 16 KB
        const uint32_t s = it & (AS_NSLOT - 1);
        if (it >= AS_NSLOT) mbar_wait(bar(B_EMPTY + s), ((it / AS_NSLOT) - 1u) & 1u);
        mbar_expect_tx(bar(B_FULL + s), AS_SLOT);
        ++it;
        return s;
      };
      auto load_qk = [&](int n) {
        const int tile = (int)blockIdx.x + n * (int)gridDim.x, b = tile >> 1, qb = tile & 1;
        for (int kc = 0; kc < 4; ++kc) {
          uint32_t s = acquire();
          tma_load_3d(ring + s * AS_SLOT, &map, bar(B_FULL + s), kc * 64, qb * AT_BM, b);
          s = acquire();
          tma_load_3d(ring + s * AS_SLOT, &map, bar(B_FULL + s), AT_C + kc * 64, 0, b);
          s = acquire();
          tma_load_3d(ring + s * AS_SLOT, &map, bar(B_FULL + s), AT_C + kc * 64, 128, b);
        }
      };
      auto load_v = [&](int n) {
        const int tile = (int)blockIdx.x + n * (int)gridDim.x, b = tile >> 1;
        for (int vc = 0; vc < 4; ++vc)
          for (int h = 0; h < 2; ++h) {
            const uint32_t s = acquire();
            tma_load_3d(ring + s * AS_SLOT, &map_v, bar(B_FULL + s), 2 * AT_C + h * 128, vc * 64, b);
            tma_load_3d(ring + s * AS_SLOT + AS_SLOT / 2, &map_v, bar(B_FULL + s), 2 * AT_C + h * 128 + 64, vc * 64, b);
          }
      };
      ATTN_TRACE(16);
      if (nt > 0) load_qk(0);
      for (int n = 0; n < nt; ++n) {
        if (n + 1 < nt) load_qk(n + 1);
        ATTN_TRACE(32 * n + 17);
        load_v(n);
        ATTN_TRACE(32 * n + 18);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint32_t idesc_o = idesc_s | (1u << 16);                     // B = V, MN-major (channels contiguous)
      uint32_t it = 0;
      auto take = [&]() {                                                // next slot, filled
        const uint32_t s = it & (AS_NSLOT - 1);
        mbar_wait(bar(B_FULL + s), (it / AS_NSLOT) & 1u);
        ++it;
        return s;
      };
      for (int n = 0; n <= nt; ++n) {
        if (n < nt) {                                                    // S_n = Q K^T into accumulator n & 1
          const int g = n & 1, u = n >> 1;
          ATTN_TRACE(32 * n + 0);
          if (u >= 1) { mbar_wait(bar(B_FREE + g), (uint32_t)(u - 1) & 1u); tc_fence_after(); }
          ATTN_TRACE(32 * n + 1);
          for (int kc = 0; kc < 4; ++kc) {
            const uint32_t s0 = take(), s1 = take(), s2 = take();
            tc_fence_after();
            const uint64_t da = make_sw128_desc(ring + s0 * AS_SLOT), db0 = make_sw128_desc(ring + s1 * AS_SLOT),
                           db1 = make_sw128_desc(ring + s2 * AS_SLOT);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint32_t acc = (kc | k) != 0 ? 1u : 0u;
              tc_mma_bf16(tmem + g * 256, da + (uint64_t)(2 * k), db0 + (uint64_t)(2 * k), idesc_s, acc);
              tc_mma_bf16(tmem + g * 256 + 128, da + (uint64_t)(2 * k), db1 + (uint64_t)(2 * k), idesc_s, acc);
            }
            tc_commit(bar(B_EMPTY + s0));
            tc_commit(bar(B_EMPTY + s1));
            tc_commit(bar(B_EMPTY + s2));
          }
          tc_commit(bar(B_S + g));
          ATTN_TRACE(32 * n + 2);
        }
        if (n >= 1) {                                                    // O_m = P_m V over S_m's columns
          const int m = n - 1, g = m & 1;
          ATTN_TRACE(32 * m + 3);
          mbar_wait(bar(B_P), (uint32_t)m & 1u);
          tc_fence_after();
          ATTN_TRACE(32 * m + 4);
          for (int vc = 0; vc < 4; ++vc) {
            const uint32_t s0 = take(), s1 = take();
            tc_fence_after();
#pragma unroll
            for (int t = 0; t < 4; ++t) {                                // 16 keys per MMA
              const uint64_t da = make_sw128_desc(sPa + vc * AT_CHUNK) + (uint64_t)(2 * t);
              // V slot: 2 boxes [64 keys x 64 channels] 8 KB apart (LBO); 8-key atoms 1 KB apart (SBO)
              const uint32_t acc = (vc | t) != 0 ? 1u : 0u;
              tc_mma_bf16(tmem + g * 256, da, make_sw128_desc_ex(ring + s0 * AS_SLOT + t * 2048, AS_SLOT / 2, 1024), idesc_o, acc);
              tc_mma_bf16(tmem + g * 256 + 128, da, make_sw128_desc_ex(ring + s1 * AS_SLOT + t * 2048, AS_SLOT / 2, 1024), idesc_o, acc);
            }
            tc_commit(bar(B_EMPTY + s0));
            tc_commit(bar(B_EMPTY + s1));
          }
          tc_commit(bar(B_O + g));
          ATTN_TRACE(32 * m + 5);
        }
      }
    }
  } else {
    const int g = (warp - 2) >> 2;                             // SIMT group == accumulator
    const int q = warp & 3;                                    // TMEM lane quarter this warp may read
    const int r = q * 32 + lane;                               // query row inside the tile == TMEM lane
    const uint32_t t_row = tmem + ((uint32_t)(q * 32) << 16) + g * 256;
    uint8_t* sP = base_ptr;
    uint8_t* stage = base_ptr + AT_BUF + AS_NSLOT * AS_SLOT + (warp - 2) * 4096;
    for (int n = g; n < nt; n += 2) {
      const uint32_t ph = (uint32_t)(n >> 1) & 1u;
      const int tile = (int)blockIdx.x + n * (int)gridDim.x;
      ATTN_TRACE(32 * n + 8);
      mbar_wait(bar(B_S + g), ph);
      tc_fence_after();
      ATTN_TRACE(32 * n + 9);
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
      // P is free once PV(n-1) -- the other group's tile -- has retired.  PV(n+1) cannot be issued before this group
      // has written P_n, so the phase waited for here cannot be lapped.
      ATTN_TRACE(32 * n + 10);
      if (n >= 1) mbar_wait(bar(B_O + (g ^ 1)), (uint32_t)((n - 1) >> 1) & 1u);
      ATTN_TRACE(32 * n + 11);
#pragma unroll 1
      for (int c0 = 0; c0 < AT_L; c0 += 32) {
        uint32_t v[32];
        tc_ld32(t_row + c0, v);
        tc_wait_ld();
        uint8_t* chunk = sP + (c0 >> 6) * AT_CHUNK + r * 128;
#pragma unroll
        for (int uu = 0; uu < 4; ++uu) {                       // 8 keys = one 16-byte unit
          uint4 pk;
          __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float e0 = exp2f(fmaf(__uint_as_float(v[uu * 8 + 2 * i]), scale_log2e, -mb));
            const float e1 = exp2f(fmaf(__uint_as_float(v[uu * 8 + 2 * i + 1]), scale_log2e, -mb));
            hh[i] = __floats2bfloat162_rn(e0, e1);
            // the row sum uses the values the MMA will see (bf16-rounded), so rows of P / sum add up to 1
            const float2 f = __bfloat1622float2(hh[i]);
            sum += f.x + f.y;
          }
          const int unit = ((c0 & 63) >> 3) + uu;              // 16-byte unit index inside the 128 B row
          *reinterpret_cast<uint4*>(chunk + ((unit ^ (r & 7)) << 4)) = pk;
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(B_P));
      ATTN_TRACE(32 * n + 12);
      const float inv = 1.0f / sum;
      mbar_wait(bar(B_O + g), ph);
      tc_fence_after();
      ATTN_TRACE(32 * n + 13);
      // O rows, 64 channels at a time: row-per-thread into the warp's staging area (16-byte units XOR-ed with the
      // row, conflict free both ways), then out as whole 128-byte lines, four rows per instruction
      bf16* obase = out + ((size_t)tile * AT_BM + q * 32) * AT_C;
#pragma unroll 1
      for (int c0 = 0; c0 < AT_C; c0 += 64) {
#pragma unroll
        for (int hc = 0; hc < 2; ++hc) {
          uint32_t v[32];
          tc_ld32(t_row + c0 + hc * 32, v);
          tc_wait_ld();
#pragma unroll
          for (int uu = 0; uu < 4; ++uu) {
            uint4 pk;
            __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              hh[i] = __floats2bfloat162_rn(__uint_as_float(v[uu * 8 + 2 * i]) * inv, __uint_as_float(v[uu * 8 + 2 * i + 1]) * inv);
            const int unit = hc * 4 + uu;
            *reinterpret_cast<uint4*>(stage + lane * 128 + ((unit ^ (lane & 7)) << 4)) = pk;
          }
        }
        if (c0 == AT_C - 64) {                                 // last TMEM read of the tile: accumulator g may take S_{n+2}
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(B_FREE + g));
        } else {
          __syncwarp();
        }
#pragma unroll
        for (int rr = 0; rr < 32; rr += 4) {
          const int row = rr + (lane >> 3), unit = lane & 7;
          const uint4 pk = *reinterpret_cast<const uint4*>(stage + row * 128 + ((unit ^ (row & 7)) << 4));
          *reinterpret_cast<uint4*>(obase + (size_t)row * AT_C + c0 + unit * 8) = pk;
        }
        __syncwarp();
      }
      ATTN_TRACE(32 * n + 14);
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

#ifdef B200_ATTN_TRACE
extern "C" int b200_attention_tc_trace(long long* buf) {
  return cudaMemcpyToSymbol(g_attn_trace, &buf, sizeof(buf)) == cudaSuccess ? B200_OK : B200_ERR_DRIVER;
}
#endif

extern "C" int b200_attention_tc_supported(int L, int C) { return (L == AT_L && C == AT_C) ? 1 : 0; }

// qkv: [B, L, 3C] bf16 (q | k | v along the channel axis); out: [B, L, C] bf16.
extern "C" int b200_attention_tc(const void* qkv, void* out, int B, int L, int C, float scale, cudaStream_t st) {
  if (!qkv || !out || B < 0) return B200_ERR_ARG;
  if (!b200_attention_tc_supported(L, C)) return B200_ERR_UNSUPPORTED;
  if (B == 0) return B200_OK;
  CUtensorMap map, map_v;
  const uint64_t dims[3] = {(uint64_t)3 * C, (uint64_t)L, (uint64_t)B};
  const uint64_t strides[2] = {(uint64_t)3 * C * 2, (uint64_t)L * 3 * C * 2};
  const uint32_t box[3] = {64, 128, 1}, box_v[3] = {64, 64, 1};
  int rc = tc_tensor_map(&map, qkv, 3, dims, strides, box, 0);
  if (rc) return rc;
  rc = tc_tensor_map(&map_v, qkv, 3, dims, strides, box_v, 0);
  if (rc) return rc;
  static int attr_state_s[B200_MAX_DEVICES];
  if (int rc2 = b200_set_smem_once(attr_state_s, attn_tc_stream_kernel, AS_SMEM)) return rc2;
  const int ntiles = B * 2, grid = ntiles < B200_NUM_SMS ? ntiles : B200_NUM_SMS;
  attn_tc_stream_kernel<<<grid, AS_THREADS, AS_SMEM, st>>>(map, map_v, (bf16*)out, scale * 1.4426950408889634f, ntiles);
  B200_RETURN_LAST();
}
