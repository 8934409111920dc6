// upfirdn2d for sm_100a: zero-insert upsample x`up`, pad, 2-D FIR with the flipped filter, decimate x`down`.
// Replaces op/upfirdn2d_kernel.cu:49-207 (+ host dispatch :209-368) of the reference.
//
// out[p, oy, ox] = sum_{ky,kx} U[oy*down + ky, ox*down + kx] * k[kh-1-ky, kw-1-kx]
// U = in upsampled by zero insertion and shifted by (pad_y0, pad_x0); zero outside.
// Negative pads crop, exactly like the index arithmetic of the reference kernel.
//
// Two data layouts:
//   * planes  : [n_planes, H, W] fp32 (the NCHW view `input.reshape(-1, H, W, 1)` of op/upfirdn2d.py:99) -- op API.
//   * nhwc    : [B, H, W, C] fp32/bf16 channels-last, 8 channels per thread -- used inside the score engine.
// Both are HBM-bound: algorithmic bytes = 4*(in_elems + out_elems) (planes).
#include "common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

__host__ __device__ __forceinline__ int floor_div(int a, int b) {
  int q = a / b, r = a % b;
  return (r != 0 && ((r < 0) != (b < 0))) ? q - 1 : q;
}
__host__ __device__ __forceinline__ int pmod(int a, int b) {
  int r = a % b;
  return r < 0 ? r + b : r;
}

struct UfdTile {
  int planes, in_h, in_w, out_h, out_w, pad_x0, pad_y0;
  int toh, tow, ppb, tiles_x, tiles_y, tih, tiw;  // tih/tiw: smem tile extents (upper bounds)
};

// Shared-memory tiled kernel.  One block = `ppb` consecutive planes x (toh x tow) outputs.
template <int UP, int DOWN, int KH, int KW>
__global__ void __launch_bounds__(256) upfirdn2d_planes_kernel(const float* __restrict__ in, const float* __restrict__ k,
                                                               float* __restrict__ out, UfdTile t) {
  extern __shared__ float sx[];
  __shared__ float4 swt[4];  // UP==2: taps regrouped by output parity class (ky0, kx0)
  // flipped taps in registers: w[ky][kx] = k[KH-1-ky][KW-1-kx]
  float w[KH][KW];
#pragma unroll
  for (int a = 0; a < KH; ++a)
#pragma unroll
    for (int b = 0; b < KW; ++b) w[a][b] = __ldg(k + (KH - 1 - a) * KW + (KW - 1 - b));
  if (UP == 2 && threadIdx.x < 4) {
    const int ky0 = threadIdx.x >> 1, kx0 = threadIdx.x & 1;
    auto tap = [&](int ky, int kx) { return __ldg(k + (KH - 1 - ky) * KW + (KW - 1 - kx)); };
    swt[threadIdx.x] = make_float4(tap(ky0, kx0), tap(ky0, kx0 + 2), tap(ky0 + 2, kx0), tap(ky0 + 2, kx0 + 2));
  }

  int bid = blockIdx.x;
  const int tx = bid % t.tiles_x; bid /= t.tiles_x;
  const int ty = bid % t.tiles_y; bid /= t.tiles_y;
  const int p0 = bid * t.ppb;
  const int np = min(t.ppb, t.planes - p0);
  const int ox0 = tx * t.tow, oy0 = ty * t.toh;
  const int tow = min(t.tow, t.out_w - ox0), toh = min(t.toh, t.out_h - oy0);

  // input window needed by this tile
  const int ix0 = floor_div(ox0 * DOWN - t.pad_x0 + UP - 1, UP);
  const int iy0 = floor_div(oy0 * DOWN - t.pad_y0 + UP - 1, UP);
  // UP==1, DOWN==2: ix0 + pad_x0 = 2*ox0 is even, so (ox*2 - pad_x0 - ix0) is even -> 8-byte smem reads.
  const int tiw = t.tiw, tih = t.tih;

  // stage: coalesced along x, zero fill outside the image
  const int per_plane = tih * tiw;
  for (int i = threadIdx.x; i < np * per_plane; i += blockDim.x) {
    int pl = i / per_plane, r = i - pl * per_plane;
    int yy = r / tiw, xx = r - yy * tiw;
    int iy = iy0 + yy, ix = ix0 + xx;
    float v = 0.0f;
    if (iy >= 0 && iy < t.in_h && ix >= 0 && ix < t.in_w)
      v = __ldg(in + ((size_t)(p0 + pl) * t.in_h + iy) * t.in_w + ix);
    sx[i] = v;
  }
  __syncthreads();

  const int per_out = toh * tow;
  for (int i = threadIdx.x; i < np * per_out; i += blockDim.x) {
    int pl = i / per_out, r = i - pl * per_out;
    int oyl = r / tow, oxl = r - oyl * tow;
    int oy = oy0 + oyl, ox = ox0 + oxl;
    const float* sp = sx + pl * per_plane;
    float acc = 0.0f;
    if (UP == 1 && DOWN == 2 && KW == 4) {
      const int cx = ox * 2 - t.pad_x0 - ix0;  // even by construction
      const int cy = oy * 2 - t.pad_y0 - iy0;
#pragma unroll
      for (int ky = 0; ky < KH; ++ky) {
        const float2* row = reinterpret_cast<const float2*>(sp + (cy + ky) * tiw + cx);
        float2 a = row[0], b = row[1];
        acc = fmaf(a.x, w[ky][0], acc); acc = fmaf(a.y, w[ky][1], acc);
        acc = fmaf(b.x, w[ky][2], acc); acc = fmaf(b.y, w[ky][3], acc);
      }
    } else if (UP == 2) {
      // two taps per axis; which two depends on the parity of (pad - o)
      const int ky0 = (t.pad_y0 - oy) & 1, kx0 = (t.pad_x0 - ox) & 1;
      const float4 wq = swt[ky0 * 2 + kx0];
      const int yy = ((oy + ky0 - t.pad_y0) >> 1) - iy0, xx = ((ox + kx0 - t.pad_x0) >> 1) - ix0;
      const float* r0 = sp + yy * tiw + xx;
      acc = r0[0] * wq.x;
      acc = fmaf(r0[1], wq.y, acc);
      acc = fmaf(r0[tiw], wq.z, acc);
      acc = fmaf(r0[tiw + 1], wq.w, acc);
    } else {
      const int cx = ox * DOWN - t.pad_x0 - ix0, cy = oy * DOWN - t.pad_y0 - iy0;
#pragma unroll
      for (int ky = 0; ky < KH; ++ky)
#pragma unroll
        for (int kx = 0; kx < KW; ++kx) acc = fmaf(sp[(cy + ky) * tiw + cx + kx], w[ky][kx], acc);
    }
    out[((size_t)(p0 + pl) * t.out_h + oy) * t.out_w + ox] = acc;
  }
}

// Generic fallback: any up/down/filter size, one output per thread, taps straight from global/L1.
__global__ void upfirdn2d_planes_generic_kernel(const float* __restrict__ in, const float* __restrict__ k,
                                                float* __restrict__ out, long long total, int in_h, int in_w,
                                                int out_h, int out_w, int kh, int kw, int up_x, int up_y,
                                                int down_x, int down_y, int pad_x0, int pad_y0) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int ox = (int)(i % out_w);
    long long r = i / out_w;
    int oy = (int)(r % out_h);
    long long p = r / out_h;
    const float* ip = in + p * (long long)in_h * in_w;
    float acc = 0.0f;
    for (int ky = pmod(pad_y0 - oy * down_y, up_y); ky < kh; ky += up_y) {
      int iy = (oy * down_y + ky - pad_y0) / up_y;
      if (iy < 0 || iy >= in_h) continue;
      for (int kx = pmod(pad_x0 - ox * down_x, up_x); kx < kw; kx += up_x) {
        int ix = (ox * down_x + kx - pad_x0) / up_x;
        if (ix < 0 || ix >= in_w) continue;
        acc = fmaf(__ldg(ip + (long long)iy * in_w + ix), __ldg(k + (kh - 1 - ky) * kw + (kw - 1 - kx)), acc);
      }
    }
    out[i] = acc;
  }
}

// Channels-last variant: one thread = 8 channels of one output pixel.
template <typename T>
__global__ void __launch_bounds__(256) upfirdn2d_nhwc_kernel(const T* __restrict__ in, const float* __restrict__ k,
                                                             T* __restrict__ out, long long total_vec, int C8, int in_h,
                                                             int in_w, int out_h, int out_w, int kh, int kw, int up,
                                                             int down, int pad_x0, int pad_y0) {
  __shared__ float sk[64];
  if (threadIdx.x < kh * kw) {
    int ky = threadIdx.x / kw, kx = threadIdx.x % kw;
    sk[threadIdx.x] = k[(kh - 1 - ky) * kw + (kw - 1 - kx)];
  }
  __syncthreads();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total_vec; i += (long long)gridDim.x * blockDim.x) {
    int c8 = (int)(i % C8);
    long long r = i / C8;
    int ox = (int)(r % out_w); r /= out_w;
    int oy = (int)(r % out_h);
    long long b = r / out_h;
    const T* ip = in + b * (long long)in_h * in_w * C8 * 8 + c8 * 8;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.0f;
    for (int ky = pmod(pad_y0 - oy * down, up); ky < kh; ky += up) {
      int iy = (oy * down + ky - pad_y0) / up;
      if (iy < 0 || iy >= in_h) continue;
      for (int kx = pmod(pad_x0 - ox * down, up); kx < kw; kx += up) {
        int ix = (ox * down + kx - pad_x0) / up;
        if (ix < 0 || ix >= in_w) continue;
        float v[8];
        Vec8<T>::load(ip + ((long long)iy * in_w + ix) * C8 * 8, v);
        float wk = sk[ky * kw + kx];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(v[j], wk, acc[j]);
      }
    }
    Vec8<T>::store(out + i * 8, acc);
  }
}

// Channels-last 4x4-filter variant, register blocked: one thread = 8 channels of a 2x2 block of output pixels.  Each
// input vector of the (WIN x WIN) window is loaded once and scattered into the outputs it feeds, which cuts the loads
// per output from 16 to 9 (down 2), 2.25 (up 2) and 6.25 (pad-only).
// The kernel is bound by instruction issue and by loads in flight, not by bytes, so those are what was engineered
// (ncu source page of the first version, up-2 shape: 639 instructions per thread of which 128 FFMA, 171 IMAD, 54 IADD3,
// 50 SHF, 46 ISETP -- profiles/r02o_fir.txt):
//   * the padding is a template parameter: with even block origins the COLUMN tap of window column wx for output
//     column c is K0 + wx * UP - c * DOWN, a compile-time number; rows are a real loop for UP == 1 (one window row of
//     16/32-byte loads in flight, bounded registers) with one uniform test + one 32-byte tap-row fetch per (row,
//     output row);
//   * index arithmetic is 32-bit: the thread index is decomposed with host-computed multiply-shift divisors (a shift
//     when the channel count is one of the templated 128 / 256 / 512), and every load address is ONE per-thread
//     64-bit window origin + a row offset + a column offset that is an immediate for the templated channel counts.
//     The loads are inline PTX (`@p ld.global.nc.v4` after zeroing the destination) because the compiler otherwise
//     sinks the whole address computation under each load's predicate: 9-10 instructions per load;
//   * the accumulation is packed: fma.rn.f32x2 (FFMA2) on channel pairs -- the same IEEE fma per lane, half the issue
//     slots; a bf16 pair unpacks into exactly such a register pair with two ALU ops;
//   * MINB = 3 (<= 80 registers, 24 warps per SM) for the bf16 shapes: the UP == 1 loops have one window row in
//     flight per warp and need the warps (down-2 3.7 -> 4.4 TB/s; a two-row software pipeline at 124 registers gave
//     the same and nothing on top).
// B = 1024, bf16: up-2 16x16x256 216 -> 127 us (5.3 TB/s, 0.81 of HBM), down-2 32x32x128 97 -> 70 us (4.8 TB/s, 0.73),
// pad-only 16x16x256 161 -> 111 us (2.6 TB/s: 16 taps per output at stride 1, still issue-bound).
struct FastDiv {
  uint32_t d, mul, shr;   // n / d == umulhi(n, mul) >> shr for n < 2^31 (d >= 2); d == 1 is passed through
};
static inline FastDiv make_fastdiv(uint32_t d) {
  FastDiv f{d, 0u, 0u};
  if (d <= 1) return f;
  uint32_t s = 0;
  while ((1ull << s) < d) ++s;                                  // ceil(log2 d)
  f.mul = (uint32_t)(((1ull << (31 + s)) + d - 1) / d);         // ceil(2^(31+s) / d) < 2^32
  f.shr = s - 1;
  return f;
}
__device__ __forceinline__ uint32_t fd_div(uint32_t n, const FastDiv& f) {
  return f.d == 1 ? n : (__umulhi(n, f.mul) >> f.shr);
}
__device__ __forceinline__ void ffma2(float2& acc, const float2 v, const float2 w) {
  unsigned long long a = *reinterpret_cast<unsigned long long*>(&acc);
  const unsigned long long x = *reinterpret_cast<const unsigned long long*>(&v),
                           y = *reinterpret_cast<const unsigned long long*>(&w);
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(a) : "l"(x), "l"(y));
  acc = *reinterpret_cast<float2*>(&a);
}
// eight channels as loaded (no conversion), and channel pair p of them as a float2
template <typename T> struct Raw8;
template <> struct Raw8<float> {
  float4 lo, hi;
  static __device__ __forceinline__ Raw8 load(const float* p) {
    Raw8 r; r.lo = *reinterpret_cast<const float4*>(p); r.hi = *reinterpret_cast<const float4*>(p + 4); return r;
  }
  static __device__ __forceinline__ Raw8 zero() { Raw8 r; r.lo = r.hi = make_float4(0.f, 0.f, 0.f, 0.f); return r; }
  // eight channels at byte address p + off when ok, zeros otherwise; the address is formed unconditionally
  static __device__ __forceinline__ Raw8 load_if(const char* p, int off, int ok) {
    Raw8 r;
    asm("{\n .reg .pred q;\n setp.ne.s32 q, %9, 0;\n"
        " mov.b32 %0, 0; mov.b32 %1, 0; mov.b32 %2, 0; mov.b32 %3, 0; mov.b32 %4, 0; mov.b32 %5, 0; mov.b32 %6, 0; mov.b32 %7, 0;\n"
        " @q ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%8];\n @q ld.global.nc.v4.f32 {%4, %5, %6, %7}, [%8+16];\n}"
        : "=f"(r.lo.x), "=f"(r.lo.y), "=f"(r.lo.z), "=f"(r.lo.w), "=f"(r.hi.x), "=f"(r.hi.y), "=f"(r.hi.z), "=f"(r.hi.w)
        : "l"(p + off), "r"(ok));
    return r;
  }
  __device__ __forceinline__ float2 pair(int p) const {
    return p == 0 ? make_float2(lo.x, lo.y) : p == 1 ? make_float2(lo.z, lo.w) : p == 2 ? make_float2(hi.x, hi.y)
                                                                                       : make_float2(hi.z, hi.w);
  }
  static __device__ __forceinline__ void store(float* p, const float2* v) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0].x, v[0].y, v[1].x, v[1].y);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[2].x, v[2].y, v[3].x, v[3].y);
  }
};
template <> struct Raw8<bf16> {
  uint4 w;
  static __device__ __forceinline__ Raw8 load(const bf16* p) { Raw8 r; r.w = *reinterpret_cast<const uint4*>(p); return r; }
  static __device__ __forceinline__ Raw8 zero() { Raw8 r; r.w = make_uint4(0u, 0u, 0u, 0u); return r; }
  static __device__ __forceinline__ Raw8 load_if(const char* p, int off, int ok) {
    Raw8 r;
    asm("{\n .reg .pred q;\n setp.ne.s32 q, %5, 0;\n mov.b32 %0, 0; mov.b32 %1, 0; mov.b32 %2, 0; mov.b32 %3, 0;\n"
        " @q ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];\n}"
        : "=r"(r.w.x), "=r"(r.w.y), "=r"(r.w.z), "=r"(r.w.w)
        : "l"(p + off), "r"(ok));
    return r;
  }
  __device__ __forceinline__ float2 pair(int p) const {
    const uint32_t x = p == 0 ? w.x : p == 1 ? w.y : p == 2 ? w.z : w.w;
    return make_float2(__uint_as_float(x << 16), __uint_as_float(x & 0xffff0000u));
  }
  static __device__ __forceinline__ void store(bf16* p, const float2* v) {
    uint4 r;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[i].x, v[i].y);
    *reinterpret_cast<uint4*>(p) = r;
  }
};

template <typename T, int UP, int DOWN, int PAD0, int MINB, int C8T>
__global__ void __launch_bounds__(256, MINB) upfirdn2d_nhwc4_kernel(const T* __restrict__ in, const float* __restrict__ k,
                                                              T* __restrict__ out, uint32_t total, FastDiv dC8,
                                                              FastDiv dbw, FastDiv dbh, int in_h, int in_w, int out_h,
                                                              int out_w) {
  static_assert(UP == 1 || (UP == 2 && DOWN == 1), "block origins are even: UP in {1, 2}");
  constexpr int K0 = PAD0 % UP;
  constexpr int WIN = (DOWN + 3) / UP + 1;      // input rows / cols that can touch a 2-wide output block
  __shared__ float2 skp[16];                    // flipped taps, each duplicated into an FFMA2 operand pair
  if (threadIdx.x < 16) {
    const float w = k[(3 - (threadIdx.x >> 2)) * 4 + (3 - (threadIdx.x & 3))];
    skp[threadIdx.x] = make_float2(w, w);
  }
  __syncthreads();
  // C8T != 0: the channel count is a compile-time number (a power of two): column offsets become load immediates and
  // the first division a shift
  const uint32_t C8 = C8T ? (uint32_t)C8T : dC8.d;
  const int C = (int)C8 * 8;
  // everything below is int arithmetic on element offsets inside ONE image (the host guarantees < 2^31); the only
  // 64-bit operations are the per-image base and one widening add per load / store
  const int rowstride = in_w * C, in_img = in_h * rowstride;
  const int orow = out_w * C, out_img = out_h * orow;
  for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
    const uint32_t r1 = C8T ? i / (uint32_t)(C8T ? C8T : 1) : fd_div(i, dC8), c8 = i - r1 * C8;
    const uint32_t r2 = fd_div(r1, dbw);
    const int ox0 = (int)(r1 - r2 * dbw.d) * 2;
    const uint32_t b = fd_div(r2, dbh);
    const int oy0 = (int)(r2 - b * dbh.d) * 2;
    const int uy0 = oy0 * DOWN - PAD0, ux0 = ox0 * DOWN - PAD0;       // upsampled-grid coordinate of tap (0, 0)
    const int iy0 = UP == 1 ? uy0 : ((uy0 + 1) >> 1), ix0 = UP == 1 ? ux0 : ((ux0 + 1) >> 1);
    // byte address of the window origin (may lie outside the image: never dereferenced there)
    const char* p00 = reinterpret_cast<const char*>(in + (size_t)b * (uint32_t)in_img + c8 * 8) +
                      (long long)(iy0 * rowstride + ix0 * C) * (int)sizeof(T);
    int colok[WIN];
#pragma unroll
    for (int wx = 0; wx < WIN; ++wx) colok[wx] = (unsigned)(ix0 + wx) < (unsigned)in_w;
    float2 acc[2][2][4];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int p = 0; p < 4; ++p) acc[a][c][p] = make_float2(0.f, 0.f);
    auto load_row = [&](int wy, Raw8<T>* v) {                       // the whole window row is requested at once
      const int row_ok = (unsigned)(iy0 + wy) < (unsigned)in_h;
      const char* rp = p00 + (long long)(wy * rowstride) * (int)sizeof(T);
#pragma unroll
      for (int wx = 0; wx < WIN; ++wx) v[wx] = Raw8<T>::load_if(rp, wx * C * (int)sizeof(T), row_ok & colok[wx]);
    };
    auto use_row = [&](int wy, const Raw8<T>* v) {
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        const int ky = K0 + wy * UP - a * DOWN;                      // uniform over the block; static when unrolled
        if ((unsigned)ky >= 4u) continue;
        float2 wrow[4];
        *reinterpret_cast<float4*>(&wrow[0]) = *reinterpret_cast<const float4*>(&skp[ky * 4]);
        *reinterpret_cast<float4*>(&wrow[2]) = *reinterpret_cast<const float4*>(&skp[ky * 4 + 2]);
#pragma unroll
        for (int wx = 0; wx < WIN; ++wx)
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const int kx = K0 + wx * UP - c * DOWN;                  // compile-time
            if (kx < 0 || kx >= 4) continue;
#pragma unroll
            for (int p = 0; p < 4; ++p) ffma2(acc[a][c][p], v[wx].pair(p), wrow[kx]);
          }
      }
    };
    // UP == 1: rows stay a real loop (unrolled, the compiler hoists all 25-36 loads and needs 140+ registers).
    // UP == 2 (3 x 3 window) is small enough to unroll.
#pragma unroll(UP == 1 ? 1 : WIN)
    for (int wy = 0; wy < WIN; ++wy) {
      Raw8<T> v[WIN];
      load_row(wy, v);
      use_row(wy, v);
    }
    T* op = out + (size_t)b * (uint32_t)out_img + c8 * 8;
    const int o00 = oy0 * orow + ox0 * C;
    const bool ok_y1 = oy0 + 1 < out_h, ok_x1 = ox0 + 1 < out_w;      // (oy0, ox0) itself always exists
    Raw8<T>::store(op + o00, acc[0][0]);
    if (ok_x1) Raw8<T>::store(op + (o00 + C), acc[0][1]);
    if (ok_y1) Raw8<T>::store(op + (o00 + orow), acc[1][0]);
    if (ok_y1 && ok_x1) Raw8<T>::store(op + (o00 + orow + C), acc[1][1]);
  }
}

// ---- fast planes kernels for the two resampling shapes of the score networks (4x4 filter, factor 2) ------------
// No shared memory: every thread owns a small output block, issues all of its (vector) loads up front and keeps the
// window in registers; halo overlap between neighbouring threads is served by L1.  HBM traffic = algorithmic.
// down 2, pad0 = 1 (downsample_2d, models/up_or_down_sampling.py:227-257): 2 x 4 outputs from a 6 x 10 window.
__global__ void __launch_bounds__(256) upfirdn2d_down2_kernel(const float* __restrict__ in, const float* __restrict__ k,
                                                              float* __restrict__ out, long long total, int in_h,
                                                              int in_w, int out_h, int out_w) {
  float w[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) w[a][b] = __ldg(k + (3 - a) * 4 + (3 - b));
  const int bw = out_w >> 2, bh = (out_h + 1) >> 1;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int tx = (int)(i % bw);
    long long r = i / bw;
    const int ty = (int)(r % bh);
    const long long pl = r / bh;
    const int ox0 = tx * 4, oy0 = ty * 2;
    const float* ip = in + pl * (long long)in_h * in_w;
    const int ix0 = ox0 * 2, iy0 = oy0 * 2 - 1;
    float win[6][10];
#pragma unroll
    for (int y = 0; y < 6; ++y) {
      const int iy = iy0 + y;
      if (iy >= 0 && iy < in_h) {
        const float* rp = ip + (long long)iy * in_w + ix0;
        const float4 a = __ldg(reinterpret_cast<const float4*>(rp));
        const float4 b = __ldg(reinterpret_cast<const float4*>(rp + 4));
        win[y][0] = ix0 > 0 ? __ldg(rp - 1) : 0.0f;
        win[y][1] = a.x; win[y][2] = a.y; win[y][3] = a.z; win[y][4] = a.w;
        win[y][5] = b.x; win[y][6] = b.y; win[y][7] = b.z; win[y][8] = b.w;
        win[y][9] = ix0 + 8 < in_w ? __ldg(rp + 8) : 0.0f;
      } else {
#pragma unroll
        for (int x = 0; x < 10; ++x) win[y][x] = 0.0f;
      }
    }
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      if (oy0 + a >= out_h) break;
      float o[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float acc = 0.0f;
#pragma unroll
        for (int ky = 0; ky < 4; ++ky)
#pragma unroll
          for (int kx = 0; kx < 4; ++kx) acc = fmaf(win[2 * a + ky][2 * c + kx], w[ky][kx], acc);
        o[c] = acc;
      }
      *reinterpret_cast<float4*>(out + (pl * out_h + oy0 + a) * (long long)out_w + ox0) = make_float4(o[0], o[1], o[2], o[3]);
    }
  }
}

// up 2, pad0 = 2 (upsample_2d, models/up_or_down_sampling.py:195-224): 2 x 8 outputs from a 3 x 6 window.
__global__ void __launch_bounds__(256) upfirdn2d_up2_kernel(const float* __restrict__ in, const float* __restrict__ k,
                                                            float* __restrict__ out, long long total, int in_h, int in_w,
                                                            int out_h, int out_w) {
  float w[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) w[a][b] = __ldg(k + (3 - a) * 4 + (3 - b));
  const int bw = in_w >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int tx = (int)(i % bw);
    long long r = i / bw;
    const int j = (int)(r % in_h);                    // input row: produces output rows 2j, 2j+1
    const long long pl = r / in_h;
    const int i0 = tx * 4;
    const float* ip = in + pl * (long long)in_h * in_w;
    float win[3][6];
#pragma unroll
    for (int y = 0; y < 3; ++y) {
      const int iy = j - 1 + y;
      if (iy >= 0 && iy < in_h) {
        const float* rp = ip + (long long)iy * in_w + i0;
        const float4 a = __ldg(reinterpret_cast<const float4*>(rp));
        win[y][0] = i0 > 0 ? __ldg(rp - 1) : 0.0f;
        win[y][1] = a.x; win[y][2] = a.y; win[y][3] = a.z; win[y][4] = a.w;
        win[y][5] = i0 + 4 < in_w ? __ldg(rp + 4) : 0.0f;
      } else {
#pragma unroll
        for (int x = 0; x < 6; ++x) win[y][x] = 0.0f;
      }
    }
    // out[2j+a][2i+c] = sum_{dy,dx in {0,1}} in[j-1+a+dy][i-1+c+dx] * w[2dy+a][2dx+c]
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      float o[8];
#pragma unroll
      for (int x = 0; x < 4; ++x)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          float acc = 0.0f;
#pragma unroll
          for (int dy = 0; dy < 2; ++dy)
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) acc = fmaf(win[a + dy][x + c + dx], w[2 * dy + a][2 * dx + c], acc);
          o[2 * x + c] = acc;
        }
      float* op = out + (pl * out_h + 2 * j + a) * (long long)out_w + 2 * i0;
      // one 256-bit streaming store per thread and row (sm_100): a whole 32-byte sector per request instead of two
      // half-sector requests 16 bytes apart; written once, never re-read
      asm volatile("st.global.cs.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(op), "f"(o[0]), "f"(o[1]), "f"(o[2]),
                   "f"(o[3]), "f"(o[4]), "f"(o[5]), "f"(o[6]), "f"(o[7]) : "memory");
    }
  }
}

template <int UP, int DOWN>
static int launch_tiled(const float* in, const float* k, float* out, int planes, int in_h, int in_w, int out_h,
                        int out_w, int pad_x0, int pad_y0, cudaStream_t st) {
  UfdTile t;
  t.planes = planes; t.in_h = in_h; t.in_w = in_w; t.out_h = out_h; t.out_w = out_w;
  t.pad_x0 = pad_x0; t.pad_y0 = pad_y0;
  t.tow = out_w < 64 ? out_w : 64;
  t.toh = out_h < 32 ? out_h : 32;
  if (UP == 2) { t.tow = out_w < 128 ? out_w : 128; }
  t.tiles_x = ceil_div(out_w, t.tow);
  t.tiles_y = ceil_div(out_h, t.toh);
  int per = t.toh * t.tow;
  t.ppb = per >= 1024 ? 1 : (1024 / per);
  if (t.ppb > planes) t.ppb = planes;
  t.tiw = ((t.tow - 1) * DOWN + 4 - 1) / UP + 2 + ((UP == 1 && DOWN == 2) ? 1 : 0);
  if (UP == 1 && DOWN == 2) t.tiw += (t.tiw & 1);
  t.tih = ((t.toh - 1) * DOWN + 4 - 1) / UP + 2;
  size_t smem = (size_t)t.ppb * t.tih * t.tiw * sizeof(float);
  if (smem > 200 * 1024) return B200_ERR_UNSUPPORTED;
  auto kern = upfirdn2d_planes_kernel<UP, DOWN, 4, 4>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  long long nblk = (long long)ceil_div(planes, t.ppb) * t.tiles_x * t.tiles_y;
  if (nblk > 0x7fffffffLL) return B200_ERR_UNSUPPORTED;
  kern<<<(unsigned)nblk, 256, smem, st>>>(in, k, out, t);
  B200_RETURN_LAST();
}

}  // namespace b200

using namespace b200;

extern "C" int b200_upfirdn2d_out_size(int in_size, int k, int up, int down, int pad0, int pad1) {
  return (in_size * up + pad0 + pad1 - k) / down + 1;
}

extern "C" int b200_upfirdn2d_f32(const float* in, const float* k, float* out, int n_planes, int in_h, int in_w, int kh,
                                  int kw, int up_x, int up_y, int down_x, int down_y, int pad_x0, int pad_x1,
                                  int pad_y0, int pad_y1, cudaStream_t st) {
  if (!in || !k || !out || n_planes < 0 || in_h <= 0 || in_w <= 0 || kh <= 0 || kw <= 0 || up_x <= 0 || up_y <= 0 ||
      down_x <= 0 || down_y <= 0)
    return B200_ERR_ARG;
  int out_h = (in_h * up_y + pad_y0 + pad_y1 - kh) / down_y + 1;
  int out_w = (in_w * up_x + pad_x0 + pad_x1 - kw) / down_x + 1;
  if (out_h <= 0 || out_w <= 0) return B200_ERR_ARG;
  if (n_planes == 0) return B200_OK;
  const bool al16 = ((((uintptr_t)in | (uintptr_t)out) & 15) == 0);
  if (kh == 4 && kw == 4 && al16 && up_x == 1 && up_y == 1 && down_x == 2 && down_y == 2 && pad_x0 == 1 && pad_y0 == 1 &&
      in_w % 4 == 0 && out_w % 4 == 0 && out_w * 2 == in_w) {
    const long long total = (long long)n_planes * ((out_h + 1) / 2) * (out_w / 4);
    long long want = (total + 255) / 256;
    const int blocks = (int)(want < (long long)B200_NUM_SMS * 64 ? want : (long long)B200_NUM_SMS * 64);
    upfirdn2d_down2_kernel<<<blocks, 256, 0, st>>>(in, k, out, total, in_h, in_w, out_h, out_w);
    B200_RETURN_LAST();
  }
  if (kh == 4 && kw == 4 && al16 && up_x == 2 && up_y == 2 && down_x == 1 && down_y == 1 && pad_x0 == 2 && pad_y0 == 2 &&
      in_w % 4 == 0 && out_w == 2 * in_w && out_h == 2 * in_h) {
    const long long total = (long long)n_planes * in_h * (in_w / 4);
    long long want = (total + 255) / 256;
    const int blocks = (int)(want < (long long)B200_NUM_SMS * 64 ? want : (long long)B200_NUM_SMS * 64);
    upfirdn2d_up2_kernel<<<blocks, 256, 0, st>>>(in, k, out, total, in_h, in_w, out_h, out_w);
    B200_RETURN_LAST();
  }
  if (kh == 4 && kw == 4 && up_x == up_y && down_x == down_y) {
    int rc = B200_ERR_UNSUPPORTED;
    if (up_x == 1 && down_x == 1) rc = launch_tiled<1, 1>(in, k, out, n_planes, in_h, in_w, out_h, out_w, pad_x0, pad_y0, st);
    else if (up_x == 2 && down_x == 1) rc = launch_tiled<2, 1>(in, k, out, n_planes, in_h, in_w, out_h, out_w, pad_x0, pad_y0, st);
    else if (up_x == 1 && down_x == 2) rc = launch_tiled<1, 2>(in, k, out, n_planes, in_h, in_w, out_h, out_w, pad_x0, pad_y0, st);
    if (rc != B200_ERR_UNSUPPORTED) return rc;
  }
  long long total = (long long)n_planes * out_h * out_w;
  int blocks = (int)((total + 255) / 256 < (long long)B200_NUM_SMS * 16 ? (total + 255) / 256 : (long long)B200_NUM_SMS * 16);
  upfirdn2d_planes_generic_kernel<<<blocks, 256, 0, st>>>(in, k, out, total, in_h, in_w, out_h, out_w, kh, kw, up_x,
                                                          up_y, down_x, down_y, pad_x0, pad_y0);
  B200_RETURN_LAST();
}

extern "C" int b200_upfirdn2d_nhwc(const void* in, const float* k, void* out, int dtype, int batch, int in_h, int in_w,
                                   int channels, int kh, int kw, int up, int down, int pad0, int pad1, cudaStream_t st) {
  if (!in || !k || !out || channels % 8 != 0 || kh * kw > 64 || up <= 0 || down <= 0) return B200_ERR_ARG;
  int out_h = (in_h * up + pad0 + pad1 - kh) / down + 1;
  int out_w = (in_w * up + pad0 + pad1 - kw) / down + 1;
  if (out_h <= 0 || out_w <= 0) return B200_ERR_ARG;
  long long total = (long long)batch * out_h * out_w * (channels / 8);
  if (total == 0) return B200_OK;
  // register-blocked kernel for the three shapes the score networks use (up_or_down_sampling.py:195-257 and
  // conv_downsample_2d :144-178): up 2 / pad (2,1), down 2 / pad (1,1), pad-only / pad (2,2)
  const int mode = (kh == 4 && kw == 4) ? ((up == 2 && down == 1 && pad0 == 2) ? 1 : (up == 1 && down == 2 && pad0 == 1) ? 2
                                                        : (up == 1 && down == 1 && pad0 == 2) ? 3 : 0) : 0;
  const long long tot4 = (long long)batch * ((out_h + 1) / 2) * ((out_w + 1) / 2) * (channels / 8);
  if (mode && tot4 < (1ll << 31) && (long long)in_h * in_w * channels < (1ll << 31) &&
      (long long)out_h * out_w * channels < (1ll << 31)) {
    long long w4 = (tot4 + 255) / 256;
    const int blk = (int)(w4 < (long long)B200_NUM_SMS * 32 ? w4 : (long long)B200_NUM_SMS * 32);
    const FastDiv dC8 = make_fastdiv((uint32_t)(channels / 8)), dbw = make_fastdiv((uint32_t)((out_w + 1) / 2)),
                  dbh = make_fastdiv((uint32_t)((out_h + 1) / 2));
#define B200_UFD4(T, U, D, P0, MB, CT) upfirdn2d_nhwc4_kernel<T, U, D, P0, MB, CT><<<blk, 256, 0, st>>>((const T*)in, k, (T*)out, (uint32_t)tot4, dC8, dbw, dbh, in_h, in_w, out_h, out_w)
#define B200_UFD4_C(T, U, D, P0, MB)                                                       \
  do {                                                                                     \
    if (channels == 128) B200_UFD4(T, U, D, P0, MB, 16);                                    \
    else if (channels == 256) B200_UFD4(T, U, D, P0, MB, 32);                               \
    else if (channels == 512) B200_UFD4(T, U, D, P0, MB, 64);                               \
    else B200_UFD4(T, U, D, P0, MB, 0);                                                     \
  } while (0)
    if (dtype == B200_DT_F32) {
      if (mode == 1) B200_UFD4(float, 2, 1, 2, 2, 0); else if (mode == 2) B200_UFD4(float, 1, 2, 1, 2, 0); else B200_UFD4(float, 1, 1, 2, 2, 0);
    } else if (dtype == B200_DT_BF16) {
      if (mode == 1) B200_UFD4_C(bf16, 2, 1, 2, 3); else if (mode == 2) B200_UFD4_C(bf16, 1, 2, 1, 3); else B200_UFD4_C(bf16, 1, 1, 2, 3);
    } else {
      return B200_ERR_ARG;
    }
#undef B200_UFD4_C
#undef B200_UFD4
    B200_RETURN_LAST();
  }
  long long want = (total + 255) / 256;
  int blocks = (int)(want < (long long)B200_NUM_SMS * 32 ? want : (long long)B200_NUM_SMS * 32);
  if (dtype == B200_DT_F32)
    upfirdn2d_nhwc_kernel<float><<<blocks, 256, 0, st>>>((const float*)in, k, (float*)out, total, channels / 8, in_h,
                                                         in_w, out_h, out_w, kh, kw, up, down, pad0, pad0);
  else if (dtype == B200_DT_BF16)
    upfirdn2d_nhwc_kernel<bf16><<<blocks, 256, 0, st>>>((const bf16*)in, k, (bf16*)out, total, channels / 8, in_h, in_w,
                                                        out_h, out_w, kh, kw, up, down, pad0, pad0);
  else
    return B200_ERR_ARG;
  B200_RETURN_LAST();
}
