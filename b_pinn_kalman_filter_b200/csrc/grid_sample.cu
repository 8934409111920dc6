// Bilinear / trilinear grid sampling with first and second derivatives for sm_100a.
// Replaces, for op.grid_sample (op/grid_sample.py:15-22 <- models/flownet.py:8-25):
//   forward   F.grid_sample(mode='bilinear')                     (op/grid_sample.py:38-39, 107-108)
//   backward  aten::grid_sampler_{2d,3d}_backward                (op/grid_sample.py:54-60, 124-130)
//   grad2     gridsample_grad2.grad2_{2d,3d}                     (op/grid_sample_kernel.cu:27-533, grid_sample.cpp:26-56)
// One templated body for DIM = 2, 3.  A sample point has 2^DIM corners k; bit d of k says whether the corner is the
// upper one along axis d (axis 0 = x = fastest input index).  With t_d the fractional offset,
//   w_k = prod_d (bit_d ? t_d : 1 - t_d),   dw_k/dt_d = s_d prod_{e != d} (...),   s_d = bit_d ? +1 : -1.
// forward   out(c)   = sum_k w_k v_k(c)                                        v_k = input at corner k (0 outside)
// backward  gI[k]   += w_k gO(c);     gG_d = m_d sum_c gO(c) sum_k dw_k/dt_d v_k(c)          m_d = d(index)/d(grid)
// grad2     with a_k = g2_gI at corner k and G_d = g2_gG_d m_d:
//           ggO(c)   = sum_k w_k a_k(c) + sum_k T_k v_k(c),           T_k = sum_d G_d dw_k/dt_d
//           gInp[k] += T_k gO(c)
//           gGrid_d  = m_d sum_c gO(c) [ sum_k dw_k/dt_d a_k(c) + sum_k v_k(c) sum_{e != d} G_e d2w_k/dt_d dt_e ]
// Layouts: input [N, C, (D,) H, W] fp32, grid [N, (D',) H', W', DIM] fp32 (x, y[, z]), output [N, C, (D',) H', W'].
// One thread per sample point, looping over channels: reads of gO / writes of out are coalesced across points,
// corner gathers hit L1/L2; the scatter sides use atomicAdd (fire-and-forget RED) into caller-zeroed buffers.
#include "common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

template <int DIM>
struct GsGeom {
  long long idx[1 << DIM];   // element offset of corner k inside one (n, c) volume, -1 when outside
  float w[1 << DIM];
  float t[DIM];
  float m[DIM];              // d(index)/d(grid coordinate), 0 where 'border' clipped
};

// source index of a normalised coordinate and its derivative (ATen grid_sampler conventions: align_corners maps
// [-1, 1] to pixel centres 0 .. size-1, otherwise to pixel edges; 'border' clamps and zeroes the derivative).
__device__ __forceinline__ float gs_index(float x, int size, int border, int align, float& mult) {
  float v;
  if (align) { mult = 0.5f * (size - 1); v = (x + 1.0f) * 0.5f * (size - 1); }
  else       { mult = 0.5f * size;       v = ((x + 1.0f) * size - 1.0f) * 0.5f; }
  if (border) {
    if (v <= 0.0f) { v = 0.0f; mult = 0.0f; }
    else if (v >= (float)(size - 1)) { v = (float)(size - 1); mult = 0.0f; }
  }
  return v;
}

template <int DIM>
__device__ __forceinline__ void gs_geom(GsGeom<DIM>& g, const float* __restrict__ coord, const int* size, int border,
                                        int align) {
  int lo[DIM];
#pragma unroll
  for (int d = 0; d < DIM; ++d) {
    const float v = gs_index(coord[d], size[d], border, align, g.m[d]);
    const float f = floorf(v);
    lo[d] = (int)f;
    g.t[d] = v - f;
  }
#pragma unroll
  for (int k = 0; k < (1 << DIM); ++k) {
    float w = 1.0f;
    long long off = 0, stride = 1;
    bool in = true;
#pragma unroll
    for (int d = 0; d < DIM; ++d) {
      const int bit = (k >> d) & 1;
      const int i = lo[d] + bit;
      w *= bit ? g.t[d] : 1.0f - g.t[d];
      in = in && i >= 0 && i < size[d];
      off += (long long)i * stride;
      stride *= size[d];
    }
    g.w[k] = w;
    g.idx[k] = in ? off : -1;
  }
}

// dw_k/dt_d
template <int DIM>
__device__ __forceinline__ float gs_dw(const GsGeom<DIM>& g, int k, int d) {
  float r = ((k >> d) & 1) ? 1.0f : -1.0f;
#pragma unroll
  for (int e = 0; e < DIM; ++e)
    if (e != d) r *= ((k >> e) & 1) ? g.t[e] : 1.0f - g.t[e];
  return r;
}
// d2w_k/dt_d dt_e  (d != e)
template <int DIM>
__device__ __forceinline__ float gs_d2w(const GsGeom<DIM>& g, int k, int d, int e) {
  float r = (((k >> d) & 1) ? 1.0f : -1.0f) * (((k >> e) & 1) ? 1.0f : -1.0f);
#pragma unroll
  for (int f = 0; f < DIM; ++f)
    if (f != d && f != e) r *= ((k >> f) & 1) ? g.t[f] : 1.0f - g.t[f];
  return r;
}

struct GsShape {
  int N, C;
  int in_size[3];     // W, H, D (x, y, z order); unused entries 1
  long long in_vol;   // elements of one (n, c) input volume
  long long out_pts;  // sample points per batch item
  int border, align;
};

template <int DIM>
__global__ void __launch_bounds__(256) gs_fwd_kernel(float* __restrict__ out, const float* __restrict__ in,
                                                     const float* __restrict__ grid, GsShape s) {
  const long long total = (long long)s.N * s.out_pts;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long n = i / s.out_pts, p = i - n * s.out_pts;
    float co[DIM];
#pragma unroll
    for (int d = 0; d < DIM; ++d) co[d] = __ldg(grid + i * DIM + d);
    GsGeom<DIM> g;
    gs_geom<DIM>(g, co, s.in_size, s.border, s.align);
    const float* ip = in + n * s.C * s.in_vol;
    float* op = out + n * s.C * s.out_pts + p;
    for (int c = 0; c < s.C; ++c, ip += s.in_vol, op += s.out_pts) {
      float acc = 0.0f;
#pragma unroll
      for (int k = 0; k < (1 << DIM); ++k)
        if (g.idx[k] >= 0) acc = fmaf(g.w[k], __ldg(ip + g.idx[k]), acc);
      *op = acc;
    }
  }
}

template <int DIM>
__global__ void __launch_bounds__(256) gs_bwd_kernel(float* __restrict__ g_in, float* __restrict__ g_grid,
                                                     const float* __restrict__ g_out, const float* __restrict__ in,
                                                     const float* __restrict__ grid, GsShape s) {
  const long long total = (long long)s.N * s.out_pts;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long n = i / s.out_pts, p = i - n * s.out_pts;
    float co[DIM];
#pragma unroll
    for (int d = 0; d < DIM; ++d) co[d] = __ldg(grid + i * DIM + d);
    GsGeom<DIM> g;
    gs_geom<DIM>(g, co, s.in_size, s.border, s.align);
    float dw[1 << DIM][DIM];
#pragma unroll
    for (int k = 0; k < (1 << DIM); ++k)
#pragma unroll
      for (int d = 0; d < DIM; ++d) dw[k][d] = gs_dw<DIM>(g, k, d);
    float gg[DIM];
#pragma unroll
    for (int d = 0; d < DIM; ++d) gg[d] = 0.0f;
    const float* ip = in + n * s.C * s.in_vol;
    float* gi = g_in ? g_in + n * s.C * s.in_vol : nullptr;
    const float* go = g_out + n * s.C * s.out_pts + p;
    for (int c = 0; c < s.C; ++c, ip += s.in_vol, go += s.out_pts) {
      const float o = __ldg(go);
#pragma unroll
      for (int k = 0; k < (1 << DIM); ++k)
        if (g.idx[k] >= 0) {
          if (gi) atomicAdd(gi + g.idx[k], g.w[k] * o);
          if (g_grid) {
            const float v = __ldg(ip + g.idx[k]) * o;
#pragma unroll
            for (int d = 0; d < DIM; ++d) gg[d] = fmaf(dw[k][d], v, gg[d]);
          }
        }
      if (gi) gi += s.in_vol;
    }
    if (g_grid) {
#pragma unroll
      for (int d = 0; d < DIM; ++d) g_grid[i * DIM + d] = g.m[d] * gg[d];
    }
  }
}

template <int DIM>
__global__ void __launch_bounds__(256) gs_grad2_kernel(float* __restrict__ gg_out, float* __restrict__ g_in,
                                                       float* __restrict__ g_grid, const float* __restrict__ g2_gin,
                                                       const float* __restrict__ g2_ggrid, const float* __restrict__ g_out,
                                                       const float* __restrict__ in, const float* __restrict__ grid,
                                                       GsShape s) {
  const long long total = (long long)s.N * s.out_pts;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long n = i / s.out_pts, p = i - n * s.out_pts;
    float co[DIM], G[DIM];
#pragma unroll
    for (int d = 0; d < DIM; ++d) co[d] = __ldg(grid + i * DIM + d);
    GsGeom<DIM> g;
    gs_geom<DIM>(g, co, s.in_size, s.border, s.align);
#pragma unroll
    for (int d = 0; d < DIM; ++d) G[d] = (g2_ggrid ? __ldg(g2_ggrid + i * DIM + d) : 0.0f) * g.m[d];
    float dw[1 << DIM][DIM], T[1 << DIM], X[1 << DIM][DIM];
#pragma unroll
    for (int k = 0; k < (1 << DIM); ++k) {
      T[k] = 0.0f;
#pragma unroll
      for (int d = 0; d < DIM; ++d) {
        dw[k][d] = gs_dw<DIM>(g, k, d);
        T[k] = fmaf(G[d], dw[k][d], T[k]);
        float x = 0.0f;
#pragma unroll
        for (int e = 0; e < DIM; ++e)
          if (e != d) x = fmaf(G[e], gs_d2w<DIM>(g, k, d, e), x);
        X[k][d] = x;
      }
    }
    float gg[DIM];
#pragma unroll
    for (int d = 0; d < DIM; ++d) gg[d] = 0.0f;
    const long long nc0 = n * s.C;
    for (int c = 0; c < s.C; ++c) {
      const float* ip = in + (nc0 + c) * s.in_vol;
      const float* ap = g2_gin ? g2_gin + (nc0 + c) * s.in_vol : nullptr;
      const float o = __ldg(g_out + (nc0 + c) * s.out_pts + p);
      float acc = 0.0f;
#pragma unroll
      for (int k = 0; k < (1 << DIM); ++k)
        if (g.idx[k] >= 0) {
          const float v = __ldg(ip + g.idx[k]);
          const float a = ap ? __ldg(ap + g.idx[k]) : 0.0f;
          acc = fmaf(g.w[k], a, fmaf(T[k], v, acc));
          if (g_in) atomicAdd(g_in + (nc0 + c) * s.in_vol + g.idx[k], T[k] * o);
#pragma unroll
          for (int d = 0; d < DIM; ++d) gg[d] = fmaf(o, fmaf(dw[k][d], a, X[k][d] * v), gg[d]);
        }
      if (gg_out) gg_out[(nc0 + c) * s.out_pts + p] = acc;
    }
    if (g_grid) {
#pragma unroll
      for (int d = 0; d < DIM; ++d) g_grid[i * DIM + d] = g.m[d] * gg[d];
    }
  }
}

static inline int gs_shape(GsShape& s, int dim, int N, int C, int D, int H, int W, int Do, int Ho, int Wo, int padding_mode,
                           int align_corners) {
  if (N < 0 || C < 0 || H <= 0 || W <= 0 || Ho < 0 || Wo < 0 || (dim == 3 && (D <= 0 || Do < 0))) return B200_ERR_ARG;
  if (padding_mode != 0 && padding_mode != 1) return B200_ERR_UNSUPPORTED;     // 0 zeros, 1 border (op/grid_sample.py:16)
  s.N = N; s.C = C;
  s.in_size[0] = W; s.in_size[1] = H; s.in_size[2] = dim == 3 ? D : 1;
  s.in_vol = (long long)W * H * (dim == 3 ? D : 1);
  s.out_pts = (long long)Wo * Ho * (dim == 3 ? Do : 1);
  s.border = padding_mode; s.align = align_corners ? 1 : 0;
  return B200_OK;
}

static inline int gs_blocks(const GsShape& s) {
  long long want = ((long long)s.N * s.out_pts + 255) / 256;
  long long cap = (long long)B200_NUM_SMS * 32;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_grid_sample_fwd(int dim, float* out, const float* in, const float* grid, int N, int C, int D, int H,
                                    int W, int Do, int Ho, int Wo, int padding_mode, int align_corners, cudaStream_t st) {
  GsShape s;
  int rc = gs_shape(s, dim, N, C, D, H, W, Do, Ho, Wo, padding_mode, align_corners);
  if (rc) return rc;
  if (!out || !in || !grid || (dim != 2 && dim != 3)) return B200_ERR_ARG;
  if ((long long)s.N * s.out_pts == 0 || C == 0) return B200_OK;
  if (dim == 2) gs_fwd_kernel<2><<<gs_blocks(s), 256, 0, st>>>(out, in, grid, s);
  else gs_fwd_kernel<3><<<gs_blocks(s), 256, 0, st>>>(out, in, grid, s);
  B200_RETURN_LAST();
}

// g_in (may be NULL) must be zero-filled by the caller: the kernel accumulates into it.  g_grid may be NULL.
extern "C" int b200_grid_sample_bwd(int dim, float* g_in, float* g_grid, const float* g_out, const float* in,
                                    const float* grid, int N, int C, int D, int H, int W, int Do, int Ho, int Wo,
                                    int padding_mode, int align_corners, cudaStream_t st) {
  GsShape s;
  int rc = gs_shape(s, dim, N, C, D, H, W, Do, Ho, Wo, padding_mode, align_corners);
  if (rc) return rc;
  if (!g_out || !in || !grid || (dim != 2 && dim != 3)) return B200_ERR_ARG;
  if ((long long)s.N * s.out_pts == 0) return B200_OK;
  if (dim == 2) gs_bwd_kernel<2><<<gs_blocks(s), 256, 0, st>>>(g_in, g_grid, g_out, in, grid, s);
  else gs_bwd_kernel<3><<<gs_blocks(s), 256, 0, st>>>(g_in, g_grid, g_out, in, grid, s);
  B200_RETURN_LAST();
}

// Second derivative: cotangents g2_gin (shape of input, may be NULL = zeros) and g2_ggrid (shape of grid, may be NULL)
// of the backward's two outputs -> gg_out (shape of g_out), g_in (shape of input, zero-filled by the caller,
// accumulated), g_grid (shape of grid).  Any output may be NULL.
extern "C" int b200_grid_sample_grad2(int dim, float* gg_out, float* g_in, float* g_grid, const float* g2_gin,
                                      const float* g2_ggrid, const float* g_out, const float* in, const float* grid,
                                      int N, int C, int D, int H, int W, int Do, int Ho, int Wo, int padding_mode,
                                      int align_corners, cudaStream_t st) {
  GsShape s;
  int rc = gs_shape(s, dim, N, C, D, H, W, Do, Ho, Wo, padding_mode, align_corners);
  if (rc) return rc;
  if (!g_out || !in || !grid || (dim != 2 && dim != 3)) return B200_ERR_ARG;
  if ((long long)s.N * s.out_pts == 0) return B200_OK;
  if (dim == 2) gs_grad2_kernel<2><<<gs_blocks(s), 256, 0, st>>>(gg_out, g_in, g_grid, g2_gin, g2_ggrid, g_out, in, grid, s);
  else gs_grad2_kernel<3><<<gs_blocks(s), 256, 0, st>>>(gg_out, g_in, g_grid, g2_gin, g2_ggrid, g_out, in, grid, s);
  B200_RETURN_LAST();
}
