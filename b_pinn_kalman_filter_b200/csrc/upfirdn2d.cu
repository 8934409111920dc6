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
// per output from 16 to 9 (down 2), 2.25 (up 2) and 6.25 (pad-only) and keeps the kernel on the HBM roofline instead
// of the L1 one.
// PAD0 is a template parameter: with even block origins the tap index of window row wy is PAD0 % UP + wy * UP, a
// compile-time number, so every "does this tap exist" test folds away (the runtime-pad version spent 66 branches and
// 2.25x the necessary FMAs per thread on them and ran at 25 % of HBM for the up-2 shape).
// NA = output rows per thread (x 2 columns): 2 for the up-2 shape (3 x 3 window), 1 for down-2 / pad-only, whose
// 6 x 6 / 5 x 5 windows for two rows cost > 140 registers once the taps are static.
template <typename T, int UP, int DOWN, int PAD0, int NA>
__global__ void __launch_bounds__(256) upfirdn2d_nhwc4_kernel(const T* __restrict__ in, const float* __restrict__ k,
                                                              T* __restrict__ out, long long total, int C8, int in_h,
                                                              int in_w, int out_h, int out_w, int pad0_rt) {
  static_assert(UP == 1 || NA == 2, "odd block origins need UP == 1");
  // PAD0 < 0: padding (and with it the tap pattern) is a runtime value -- measured faster for the 6 x 6 / 5 x 5
  // windows of down-2 / pad-only (3.4 / 1.9 TB/s vs 2.0 / 1.1 with static taps, which cost 140+ registers)
  const int pad0 = PAD0 >= 0 ? PAD0 : pad0_rt;
  constexpr int K0 = PAD0 >= 0 ? PAD0 % UP : 0;
  constexpr int WINY = ((NA - 1) * DOWN + 3) / UP + 1;
  constexpr int WIN = (DOWN + 3) / UP + 1;      // input rows/cols that can touch a 2-wide output block
  __shared__ float sk[16];
  if (threadIdx.x < 16) sk[threadIdx.x] = k[(3 - (threadIdx.x >> 2)) * 4 + (3 - (threadIdx.x & 3))];   // flipped taps
  __syncthreads();
  const int bw = (out_w + 1) >> 1, bh = (out_h + NA - 1) / NA;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % C8);
    long long r = i / C8;
    const int ox0 = (int)(r % bw) * 2; r /= bw;
    const int oy0 = (int)(r % bh) * NA;
    const long long b = r / bh;
    const T* ip = in + b * (long long)in_h * in_w * C8 * 8 + c8 * 8;
    const int uy0 = oy0 * DOWN - pad0, ux0 = ox0 * DOWN - pad0;       // upsampled-grid coordinate of tap (0, 0)
    const int iy0 = floor_div(uy0 + UP - 1, UP), ix0 = floor_div(ux0 + UP - 1, UP);
    float acc[NA][2][8];
#pragma unroll
    for (int a = 0; a < NA; ++a)
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[a][c][j] = 0.0f;
    // NA == 1: rows stay a real loop (6 loads in flight, ~60 registers); fully unrolled the compiler hoists all 24
    // loads and needs 146 registers.  NA == 2 (3 x 3 window) is small enough to unroll.
#pragma unroll(NA == 1 ? 1 : WINY)
    for (int wy = 0; wy < WINY; ++wy) {
      const int iy = iy0 + wy;
      const bool row_ok = iy >= 0 && iy < in_h;
      const int ky_a = PAD0 >= 0 ? K0 + wy * UP : iy * UP - uy0;       // tap row for output row oy0 (-DOWN per row)
      // the whole window row is requested before any of it is used (no control flow between the loads): WIN
      // independent 16/32-byte loads in flight per thread instead of one
      float v[WIN][8];
#pragma unroll
      for (int wx = 0; wx < WIN; ++wx) {
        const int ix = ix0 + wx;
        if (row_ok && ix >= 0 && ix < in_w) Vec8<T>::load(ip + ((long long)iy * in_w + ix) * C8 * 8, v[wx]);
        else {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[wx][j] = 0.0f;
        }
      }
#pragma unroll
      for (int wx = 0; wx < WIN; ++wx) {
        const int kx_a = PAD0 >= 0 ? K0 + wx * UP : (ix0 + wx) * UP - ux0;
#pragma unroll
        for (int a = 0; a < NA; ++a) {
          const int ky = ky_a - a * DOWN;
          if ((unsigned)ky >= 4u) continue;
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const int kx = kx_a - c * DOWN;
            if ((unsigned)kx >= 4u) continue;
            const float wk = sk[ky * 4 + kx];
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[a][c][j] = fmaf(v[wx][j], wk, acc[a][c][j]);
          }
        }
      }
    }
#pragma unroll
    for (int a = 0; a < NA; ++a)
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int oy = oy0 + a, ox = ox0 + c;
        if (oy < out_h && ox < out_w)
          Vec8<T>::store(out + ((b * out_h + oy) * (long long)out_w + ox) * C8 * 8 + c8 * 8, acc[a][c]);
      }
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
  const int mode = (kh == 4 && kw == 4) ? ((up == 2 && down == 1 && pad0 == 2) ? 1 : (up == 1 && down == 2) ? 2
                                                        : (up == 1 && down == 1) ? 3 : 0) : 0;
  if (mode) {
    const int na = 2;
    const long long tot4 = (long long)batch * ((out_h + na - 1) / na) * ((out_w + 1) / 2) * (channels / 8);
    long long w4 = (tot4 + 255) / 256;
    const int blk = (int)(w4 < (long long)B200_NUM_SMS * 32 ? w4 : (long long)B200_NUM_SMS * 32);
#define B200_UFD4(T, U, D, P0, NA) upfirdn2d_nhwc4_kernel<T, U, D, P0, NA><<<blk, 256, 0, st>>>((const T*)in, k, (T*)out, tot4, channels / 8, in_h, in_w, out_h, out_w, pad0)
    if (dtype == B200_DT_F32) {
      if (mode == 1) B200_UFD4(float, 2, 1, 2, 2); else if (mode == 2) B200_UFD4(float, 1, 2, -1, 2); else B200_UFD4(float, 1, 1, -1, 2);
    } else if (dtype == B200_DT_BF16) {
      if (mode == 1) B200_UFD4(bf16, 2, 1, 2, 2); else if (mode == 2) B200_UFD4(bf16, 1, 2, -1, 2); else B200_UFD4(bf16, 1, 1, -1, 2);
    } else {
      return B200_ERR_ARG;
    }
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
