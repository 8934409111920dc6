// Strict-mode (fp32 activations and weights) convolution on the tensor cores: tcgen05.mma kind::tf32 with the
// three-term split  x w ~= xh wh + xl wh + xh wl  (xh = tf32(x), xl = x - xh, likewise w), fp32 accumulation in TMEM.
// Replaces cuDNN's fp32 path behind nn.Conv2d 3x3 / 1x1 and NIN (models/layers.py:96-118, 537-546) for
// `precision="fp32"`: until now those ran on conv_simt.cu (26-29 TFLOP/s).  Every product of two tf32 numbers is exact
// in fp32, the dropped xl wl term and the tf32 rounding of xl / wl are each <= 2^-22 of |x w|, so the result is as
// close to the exact convolution as an fp32 FMA loop is (measured: tests/test_conv_tc32_gpu.py); `terms = 1` keeps only
// xh wh = plain TF32 (what `torch.backends.cudnn.allow_tf32` gives the reference).
//
// Same implicit GEMM as conv_tc.cu: M = 128 pixels (one 4-D TMA box {32 c, TW, TH, TB} per filter tap, zero padding =
// TMA out-of-bounds fill), N = 128 output channels, one K block = 32 fp32 channels = one 128-byte swizzle row.
// Per stage (64 KB, 3 stages): the raw activation tile, its low part, and the hi / lo weight tiles (split once on the
// host side, b200_split_tf32).  Roles: warp 0 TMA producer, warp 1 MMA issuer (12 MMAs of K = 8 per stage), warps 2-9
// epilogue (TMEM -> register sums -> + bias + temb + residual, scale -> fp32 NHWC), warps 10-13 transform: rewrite the landed activation
// tile as xh in place (cvt.rna.tf32: low 13 bits zero, so whatever the tensor core does with them is the identity) and
// xl beside it.
//
// Accumulation.  The tensor core rounds its fp32 accumulator once per MMA instruction, and a 3x3 x 128-channel tile is
// 432 instructions deep: left in ONE accumulator the result is ~6e-6 of the output scale off the exact convolution,
// six times what cuDNN's fp32 kernels leave (measured; the split itself is good to 2^-22).  So the K loop is cut into
// groups of `flush` K blocks: the issuer alternates between two TMEM accumulators, starting each group from zero, and
// the eight epilogue warps drain every finished group into fp32 REGISTER sums (64 per thread) while the next group is
// being multiplied.  Each partial sum sees 12 x flush rounding steps at a fraction of the final magnitude, and the
// partials meet in round-to-nearest FADDs.
#include <cuda.h>

#include "tc_common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

constexpr int T32_BM = 128, T32_BN = 128, T32_BK = 32;
constexpr int T32_STAGES = 3;
constexpr int T32_TILE = T32_BM * 128;                    // 16 KB: 128 rows of 128 B
constexpr int T32_STAGE = 4 * T32_TILE;                   // A (-> hi), A lo, W hi, W lo
constexpr int T32_BAR_OFF = T32_STAGES * T32_STAGE;
constexpr int T32_SMEM = T32_BAR_OFF + 256 + 1024;        // barriers + 1024-byte alignment slack
constexpr int T32_THREADS = 448;       // producer, issuer, 8 epilogue warps, 4 transform warps

struct Conv32Params {
  int B, H, W, C1, C2, Cout;
  int taps_w, pad;
  int TW, TH, TB, tiles_w, tiles_h, num_m, num_n;
  int terms;                                              // 3: split, 1: plain TF32
  int flush;                                              // K blocks per TMEM accumulation group
  int w_batched;                                          // B operand per image: 3-D weight maps {K, N, B} (batched GEMM)
  const float* bias; const float* temb; int temb_stride; const float* res; float out_scale;
  float* out;
};

__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc) : "memory");
}

__global__ void __launch_bounds__(T32_THREADS, 1)
conv_tc32_kernel(const __grid_constant__ CUtensorMap mapA1, const __grid_constant__ CUtensorMap mapA2,
                 const __grid_constant__ CUtensorMap mapWh, const __grid_constant__ CUtensorMap mapWl, const Conv32Params p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* const sgen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar0 = base + T32_BAR_OFF;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto xf_bar = [&](int s) { return bar0 + 8u * (T32_STAGES + s); };
  auto empty_bar = [&](int s) { return bar0 + 8u * (2 * T32_STAGES + s); };
  auto tfull_bar = [&](int a) { return bar0 + 8u * (3 * T32_STAGES + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (3 * T32_STAGES + 2 + a); };
  const uint32_t tmem_slot = bar0 + 8u * (3 * T32_STAGES + 4);
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(sgen + (tmem_slot - base));

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int Ctot = p.C1 + p.C2;
  const int kb_per_tap = Ctot / T32_BK;
  const int taps = p.taps_w * p.taps_w;
  const int num_kb = taps * kb_per_tap;
  const int num_tiles = p.num_m * p.num_n;
  const bool split = p.terms == 3;

  if (threadIdx.x == 0) {
    for (int s = 0; s < T32_STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(xf_bar(s), 4); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA1) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapWh) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(2 * T32_BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  auto m_origin = [&](int mt, int& b0, int& h0, int& w0) {
    if (p.TB > 1) { b0 = mt * p.TB; h0 = 0; w0 = 0; return; }
    const int per_img = p.tiles_w * p.tiles_h;
    b0 = mt / per_img;
    const int r = mt - b0 * per_img;
    h0 = (r / p.tiles_w) * p.TH;
    w0 = (r % p.tiles_w) * p.TW;
  };

  if (warp == 0) {
    // ================= TMA producer =================
    const bool elected = tc_elect_one();
    int stage = 0; uint32_t phase = 0;
    const uint32_t bytes = (split ? 3u : 2u) * T32_TILE;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int nt = tile % p.num_n, mt = tile / p.num_n;
      int b0, h0, w0;
      m_origin(mt, b0, h0, w0);
      for (int tap = 0; tap < taps; ++tap) {
        const int r = tap / p.taps_w, s = tap - r * p.taps_w;
        for (int kb = 0; kb < kb_per_tap; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          if (elected) {
            const uint32_t sa = base + stage * T32_STAGE;
            const int c = kb * T32_BK;
            mbar_expect_tx(full_bar(stage), bytes);
            if (c < p.C1) tma_load_4d(sa, &mapA1, full_bar(stage), c, w0 + s - p.pad, h0 + r - p.pad, b0);
            else tma_load_4d(sa, &mapA2, full_bar(stage), c - p.C1, w0 + s - p.pad, h0 + r - p.pad, b0);
            if (p.w_batched) {
              tma_load_3d(sa + 2 * T32_TILE, &mapWh, full_bar(stage), c, nt * T32_BN, b0);
              if (split) tma_load_3d(sa + 3 * T32_TILE, &mapWl, full_bar(stage), c, nt * T32_BN, b0);
            } else {
              tma_load_2d(sa + 2 * T32_TILE, &mapWh, full_bar(stage), tap * Ctot + c, nt * T32_BN);
              if (split) tma_load_2d(sa + 3 * T32_TILE, &mapWl, full_bar(stage), tap * Ctot + c, nt * T32_BN);
            }
          }
          if (++stage == T32_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (whole warp waits, one elected lane issues) =================
    const bool elected = tc_elect_one();
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    // instruction descriptor: D = f32 (1 << 4), A = B = tf32 (2 << 7, 2 << 10), both K-major, N = 128, M = 128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(T32_BN >> 3) << 17) | ((uint32_t)(T32_BM >> 4) << 24);
    int stage = 0; uint32_t phase = 0;
    int acc = 0; uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      for (int kb0 = 0; kb0 < num_kb; kb0 += p.flush) {          // one accumulation group per TMEM accumulator
        mbar_wait(tempty_bar(acc), acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_u + (uint32_t)(acc * T32_BN);
        const int kb1 = min(num_kb, kb0 + p.flush);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(xf_bar(stage), phase);
          tc_fence_after();
          if (elected) {
            const uint32_t sa = base + stage * T32_STAGE;
            const uint64_t dah = make_sw128_desc(sa), dal = make_sw128_desc(sa + T32_TILE);
            const uint64_t dbh = make_sw128_desc(sa + 2 * T32_TILE), dbl = make_sw128_desc(sa + 3 * T32_TILE);
#pragma unroll
            for (int k = 0; k < T32_BK / 8; ++k) {             // K = 8 tf32 = 32 B = 2 descriptor units per MMA
              const uint64_t o = (uint64_t)(2 * k);
              const uint32_t first = (kb != kb0 || k != 0) ? 1u : 0u;
              if (split) {                                     // small terms first
                tc_mma_tf32(d_tmem, dal + o, dbh + o, idesc, first);
                tc_mma_tf32(d_tmem, dah + o, dbl + o, idesc, 1u);
                tc_mma_tf32(d_tmem, dah + o, dbh + o, idesc, 1u);
              } else {
                tc_mma_tf32(d_tmem, dah + o, dbh + o, idesc, first);
              }
            }
            tc_commit(empty_bar(stage));
          }
          if (++stage == T32_STAGES) { stage = 0; phase ^= 1; }
        }
        if (elected) tc_commit(tfull_bar(acc));
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp < 10) {
    // ================= epilogue warps (2..9): TMEM lane quadrant warp & 3, column half (warp - 2) >> 2 =================
    const int q = warp & 3, half = (warp - 2) >> 2;
    const int m_local = q * 32 + lane;
    constexpr int CH = T32_BN / 2;                            // 64 columns per thread
    int acc = 0; uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int nt = tile % p.num_n, mt = tile / p.num_n;
      int b0, h0, w0;
      m_origin(mt, b0, h0, w0);
      const int lw = m_local % p.TW, lh = (m_local / p.TW) % p.TH, lb = m_local / (p.TW * p.TH);
      const int b = b0 + lb, h = h0 + lh, w = w0 + lw;
      const bool valid = b < p.B;
      const int col0 = nt * T32_BN + half * CH;
      const size_t px = valid ? (((size_t)b * p.H + h) * p.W + w) * p.Cout + (size_t)col0 : 0;
      float sum[CH];
#pragma unroll
      for (int j = 0; j < CH; ++j) sum[j] = 0.0f;
      for (int kb0 = 0; kb0 < num_kb; kb0 += p.flush) {
        mbar_wait(tfull_bar(acc), acc_phase);
        tc_fence_after();
        const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * T32_BN + half * CH);
        uint32_t r0[32], r1[32];
        tc_ld32(t_addr, r0);
        tc_ld32(t_addr + 32, r1);
        tc_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(acc));           // the accumulator is free for the group after next
#pragma unroll
        for (int j = 0; j < 32; ++j) { sum[j] += __uint_as_float(r0[j]); sum[32 + j] += __uint_as_float(r1[j]); }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      if (valid) {
        const float* bs = p.bias ? p.bias + col0 : nullptr;
        const float* tb = p.temb ? p.temb + (size_t)b * p.temb_stride + col0 : nullptr;
#pragma unroll
        for (int j = 0; j < CH; j += 4) {
          float4 v = make_float4(sum[j], sum[j + 1], sum[j + 2], sum[j + 3]);
          if (bs) { const float4 t = __ldg(reinterpret_cast<const float4*>(bs + j)); v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w; }
          if (tb) { const float4 t = __ldg(reinterpret_cast<const float4*>(tb + j)); v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w; }
          if (p.res) { const float4 t = __ldg(reinterpret_cast<const float4*>(p.res + px + j)); v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w; }
          v.x *= p.out_scale; v.y *= p.out_scale; v.z *= p.out_scale; v.w *= p.out_scale;
          *reinterpret_cast<float4*>(p.out + px + j) = v;
        }
      }
    }
  } else {
    // ================= transform warps (10..13): x -> (tf32(x), x - tf32(x)); thread = one pixel row of the tile ======
    const int row = (int)threadIdx.x - 10 * 32;              // 0 .. 127
    int stage = 0; uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(full_bar(stage), phase);
        if (split) {
          uint8_t* const rh = sgen + stage * T32_STAGE + row * 128;
          uint8_t* const rl = rh + T32_TILE;
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int o = (u ^ (row & 7)) << 4;               // (any order: hi and lo keep the tile's own swizzle)
            const float4 x = *reinterpret_cast<const float4*>(rh + o);
            uint4 hi;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi.x) : "f"(x.x));
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi.y) : "f"(x.y));
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi.z) : "f"(x.z));
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi.w) : "f"(x.w));
            hi.x &= 0xffffe000u; hi.y &= 0xffffe000u; hi.z &= 0xffffe000u; hi.w &= 0xffffe000u;
            const float4 lo = make_float4(x.x - __uint_as_float(hi.x), x.y - __uint_as_float(hi.y),
                                          x.z - __uint_as_float(hi.z), x.w - __uint_as_float(hi.w));
            *reinterpret_cast<uint4*>(rh + o) = hi;
            *reinterpret_cast<float4*>(rl + o) = lo;
          }
          fence_proxy_async_smem();                           // generic-proxy writes -> visible to the MMA's reads
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(xf_bar(stage));
        if (++stage == T32_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(2 * T32_BN));
  }
}

// rows of n fp32 values -> softmax in place; one warp per row (the scores of the strict-mode attention)
__global__ void __launch_bounds__(256) softmax_rows_kernel(float* __restrict__ x, long long rows, int n) {
  const long long r = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  float* row = x + r * n;
  float mx = -INFINITY;
  for (int j = lane; j < n; j += 32) mx = fmaxf(mx, row[j]);
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float den = 0.f;
  for (int j = lane; j < n; j += 32) { const float e = expf(row[j] - mx); row[j] = e; den += e; }
  for (int o = 16; o; o >>= 1) den += __shfl_xor_sync(0xffffffffu, den, o);
  const float inv = 1.0f / den;
  for (int j = lane; j < n; j += 32) row[j] *= inv;
}

// Operands of the strict-mode attention from the fused projection qkv [B][L][3C]: k split into (hi, lo) as dense [B][L][C],
// v transposed to [B][C][L] (the K-major B operand of p v) and split.  One 32 x 32 tile per block (256 threads).
__global__ void __launch_bounds__(256) attn32_prep_kernel(const float* __restrict__ qkv, float* __restrict__ k_hi,
                                                          float* __restrict__ k_lo, float* __restrict__ vt_hi,
                                                          float* __restrict__ vt_lo, int L, int C) {
  __shared__ float tile[32][33];
  const int l0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const size_t b = blockIdx.z;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8
  const float* src = qkv + b * (size_t)L * 3 * C;
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int l = l0 + ty + j, c = c0 + tx;
    const float kv = src[(size_t)l * 3 * C + C + c];
    uint32_t h;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(kv));
    h &= 0xffffe000u;
    const size_t o = (b * L + l) * C + c;
    k_hi[o] = __uint_as_float(h);
    k_lo[o] = kv - __uint_as_float(h);
    tile[ty + j][tx] = src[(size_t)l * 3 * C + 2 * C + c];
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int c = c0 + ty + j, l = l0 + tx;
    const float v = tile[tx][ty + j];
    uint32_t h;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(v));
    h &= 0xffffe000u;
    const size_t o = (b * C + c) * L + l;
    vt_hi[o] = __uint_as_float(h);
    vt_lo[o] = v - __uint_as_float(h);
  }
}

// w -> (tf32(w), w - tf32(w)), elementwise (weights are split once per load_state_dict, not per launch)
__global__ void split_tf32_kernel(const float* __restrict__ w, float* __restrict__ hi, float* __restrict__ lo, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float x = w[i];
    uint32_t h;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
    h &= 0xffffe000u;
    hi[i] = __uint_as_float(h);
    lo[i] = x - __uint_as_float(h);
  }
}

// 128-pixel M tile as a TMA box {TW, TH, TB}: part of a row, whole rows of one image, or several whole images
// (the same tiling as conv_tc.cu)
static bool tile_box(int B, int H, int W, int& TW, int& TH, int& TB, int& tiles_w, int& tiles_h) {
  if (W >= T32_BM) {
    if (W % T32_BM) return false;
    TW = T32_BM; TH = 1; TB = 1; tiles_w = W / T32_BM; tiles_h = H;
    return true;
  }
  if (W <= 0 || T32_BM % W) return false;
  const int rows = T32_BM / W;
  TW = W; tiles_w = 1;
  if (rows <= H) {
    if (H % rows) return false;
    TH = rows; TB = 1; tiles_h = H / rows;
    return true;
  }
  if (rows % H) return false;
  TH = H; TB = rows / H; tiles_h = 1;
  return TB <= 256;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_split_tf32(const float* w, float* hi, float* lo, long long n, cudaStream_t st) {
  if (!w || !hi || !lo || n < 0) return B200_ERR_ARG;
  if (n == 0) return B200_OK;
  long long want = (n + 255) / 256;
  const int blocks = (int)(want < (long long)B200_NUM_SMS * 8 ? want : (long long)B200_NUM_SMS * 8);
  split_tf32_kernel<<<blocks, 256, 0, st>>>(w, hi, lo, n);
  B200_RETURN_LAST();
}

extern "C" int b200_conv2d_tc32_supported(int C1, int C2, int B, int H, int W, int Cout, int ksize) {
  if (C1 <= 0 || C1 % T32_BK || C2 < 0 || C2 % T32_BK || Cout <= 0 || Cout % T32_BN || B <= 0) return 0;
  if (ksize != 1 && ksize != 3) return 0;
  int TW, TH, TB, tw, th;
  return tile_box(B, H, W, TW, TH, TB, tw, th) ? 1 : 0;
}

extern "C" int b200_conv2d_tc32(const float* x1, int C1, const float* x2, int C2, const float* w_hi, const float* w_lo,
                                const float* bias, const float* temb, int temb_stride, const float* res, float out_scale,
                                float* out, int B, int H, int W, int Cout, int ksize, int terms, cudaStream_t st) {
  if (!x1 || !w_hi || !out || (C2 > 0 && !x2) || (terms != 1 && terms != 3) || (terms == 3 && !w_lo)) return B200_ERR_ARG;
  // K blocks (32 channels) per accumulation group: B200_TC32_FLUSH overrides (accuracy / speed measurements)
  static const int flush_env = [] { const char* e = getenv("B200_TC32_FLUSH"); return e ? atoi(e) : 0; }();
  if (!b200_conv2d_tc32_supported(C1, C2, B, H, W, Cout, ksize)) return B200_ERR_UNSUPPORTED;
  Conv32Params p;
  p.B = B; p.H = H; p.W = W; p.C1 = C1; p.C2 = C2; p.Cout = Cout; p.taps_w = ksize; p.pad = ksize / 2;
  tile_box(B, H, W, p.TW, p.TH, p.TB, p.tiles_w, p.tiles_h);
  p.num_m = p.TB > 1 ? (B + p.TB - 1) / p.TB : B * p.tiles_w * p.tiles_h;
  p.num_n = Cout / T32_BN;
  p.terms = terms;
  p.w_batched = 0;
  // terms = 3: every K block is drained (12 MMAs per partial sum): the full networks then sit as close to the fp64 run of
  // the reference as the reference's own fp32 run does (profiles/r02h_strict_accuracy.txt: DDPM 64x64 2.0e-6 with 1,
  // 3.4e-6 with 2, 6.6e-6 with 4 K blocks per group; reference fp32 2.1e-6) for 7 % of the speed.  terms = 1: the tf32
  // rounding of the operands (1e-3) is all that matters.
  p.flush = flush_env > 0 ? flush_env : (terms == 3 ? 1 : 8);
  p.bias = bias; p.temb = temb; p.temb_stride = temb_stride; p.res = res; p.out_scale = out_scale; p.out = out;
  auto act_map32 = [&](CUtensorMap* m, const float* ptr, int C) {
    const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)C * 4, (uint64_t)W * C * 4, (uint64_t)H * W * C * 4};
    const uint32_t box[4] = {(uint32_t)T32_BK, (uint32_t)p.TW, (uint32_t)p.TH, (uint32_t)p.TB};
    return tc_tensor_map(m, ptr, 4, dims, strides, box, 0, nullptr, 1);
  };
  const int Ktot = ksize * ksize * (C1 + C2);
  auto w_map32 = [&](CUtensorMap* m, const float* ptr) {
    const uint64_t dims[2] = {(uint64_t)Ktot, (uint64_t)Cout};
    const uint64_t strides[1] = {(uint64_t)Ktot * 4};
    const uint32_t box[2] = {(uint32_t)T32_BK, (uint32_t)T32_BN};
    return tc_tensor_map(m, ptr, 2, dims, strides, box, 1, nullptr, 1);
  };
  CUtensorMap a1, a2, wh, wl;
  int rc = act_map32(&a1, x1, C1);
  if (rc) return rc;
  if (C2 > 0) { rc = act_map32(&a2, x2, C2); if (rc) return rc; } else a2 = a1;
  rc = w_map32(&wh, w_hi);
  if (rc) return rc;
  if (terms == 3) { rc = w_map32(&wl, w_lo); if (rc) return rc; } else wl = wh;
  static int attr_state[B200_MAX_DEVICES];
  if (int r2 = b200_set_smem_once(attr_state, conv_tc32_kernel, T32_SMEM)) return r2;
  const int tiles = p.num_m * p.num_n;
  const int grid = tiles < B200_NUM_SMS ? tiles : B200_NUM_SMS;
  conv_tc32_kernel<<<grid, T32_THREADS, T32_SMEM, st>>>(a1, a2, wh, wl, p);
  B200_RETURN_LAST();
}

// Batched GEMM on the same kernel: out[b][m][n] = scale * sum_k a[b][m][k] * bm[b][n][k]  (fp32; bm pre-split into hi / lo
// by b200_split_tf32; a's rows `lda` floats apart).  The two products of the strict-mode attention (q k^T and p v with
// v transposed, models/layerspp.py:82-86).  M % 128 == N % 128 == K % 32 == 0.
extern "C" int b200_gemm_tc32_batched(const float* a, long long lda, const float* b_hi, const float* b_lo, float* out, int B,
                                      int M, int N, int K, float scale, int terms, cudaStream_t st) {
  if (!a || !b_hi || !out || (terms != 1 && terms != 3) || (terms == 3 && !b_lo) || B < 0 || lda < K) return B200_ERR_ARG;
  if (M <= 0 || M % T32_BM || N <= 0 || N % T32_BN || K <= 0 || K % T32_BK || (lda * 4) % 16) return B200_ERR_UNSUPPORTED;
  if (B == 0) return B200_OK;
  static const int flush_env = [] { const char* e = getenv("B200_TC32_FLUSH"); return e ? atoi(e) : 0; }();
  Conv32Params p;
  p.B = B; p.H = M / T32_BM; p.W = T32_BM; p.C1 = K; p.C2 = 0; p.Cout = N; p.taps_w = 1; p.pad = 0;
  p.TW = T32_BM; p.TH = 1; p.TB = 1; p.tiles_w = 1; p.tiles_h = p.H;
  p.num_m = B * p.H; p.num_n = N / T32_BN;
  p.terms = terms; p.flush = flush_env > 0 ? flush_env : (terms == 3 ? 1 : 8); p.w_batched = 1;
  p.bias = nullptr; p.temb = nullptr; p.temb_stride = 0; p.res = nullptr; p.out_scale = scale; p.out = out;
  CUtensorMap am, wh, wl;
  {
    const uint64_t dims[4] = {(uint64_t)K, (uint64_t)T32_BM, (uint64_t)p.H, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)lda * 4, (uint64_t)T32_BM * lda * 4, (uint64_t)M * lda * 4};
    const uint32_t box[4] = {(uint32_t)T32_BK, (uint32_t)T32_BM, 1u, 1u};
    if (int rc = tc_tensor_map(&am, a, 4, dims, strides, box, 0, nullptr, 1)) return rc;
  }
  auto bmap = [&](CUtensorMap* m, const float* ptr) {
    const uint64_t dims[3] = {(uint64_t)K, (uint64_t)N, (uint64_t)B};
    const uint64_t strides[2] = {(uint64_t)K * 4, (uint64_t)N * K * 4};
    const uint32_t box[3] = {(uint32_t)T32_BK, (uint32_t)T32_BN, 1u};
    return tc_tensor_map(m, ptr, 3, dims, strides, box, 1, nullptr, 1);
  };
  if (int rc = bmap(&wh, b_hi)) return rc;
  if (terms == 3) { if (int rc = bmap(&wl, b_lo)) return rc; } else wl = wh;
  static int attr_state[B200_MAX_DEVICES];
  if (int r2 = b200_set_smem_once(attr_state, conv_tc32_kernel, T32_SMEM)) return r2;
  const int tiles = p.num_m * p.num_n;
  conv_tc32_kernel<<<tiles < B200_NUM_SMS ? tiles : B200_NUM_SMS, T32_THREADS, T32_SMEM, st>>>(am, am, wh, wl, p);
  B200_RETURN_LAST();
}

extern "C" int b200_softmax_rows_f32(float* x, long long rows, int n, cudaStream_t st) {
  if (!x || rows < 0 || n <= 0) return B200_ERR_ARG;
  if (rows == 0) return B200_OK;
  softmax_rows_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(x, rows, n);
  B200_RETURN_LAST();
}

extern "C" int b200_attn32_prep(const float* qkv, float* k_hi, float* k_lo, float* vt_hi, float* vt_lo, int B, int L, int C,
                                cudaStream_t st) {
  if (!qkv || !k_hi || !k_lo || !vt_hi || !vt_lo || B < 0 || B > 65535) return B200_ERR_ARG;
  if (L <= 0 || L % 32 || C <= 0 || C % 32) return B200_ERR_UNSUPPORTED;
  if (B == 0) return B200_OK;
  attn32_prep_kernel<<<dim3(L / 32, C / 32, B), 256, 0, st>>>(qkv, k_hi, k_lo, vt_hi, vt_lo, L, C);
  B200_RETURN_LAST();
}
