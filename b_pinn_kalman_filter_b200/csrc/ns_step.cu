// Navier-Stokes simulator step (pinn_kalman/simulator.py:54-57) as shared-memory stencil kernels for sm_100a.
// Replaces op/ns_step_kernel.cu:96-234 (update_gradient / cip_advect / velocity_update / pressure_update kernels,
// one 1-cell block each with batch as blockDim) and the 9-launch host composites in op/ns_step.cpp:45-102.
//
// Fields are [B, C, H, W] fp32; x is the fast (W) index, y the slow (H) index (identical to the reference's
// `y * gridDim.x + x` indexing whenever H == W, the only case the reference indexes correctly).
// Arithmetic: fp32 with per-launch reciprocals (ns_math.cuh explains why and what that costs against the reference's
// mixed float/double form: ~1e-7 relative); exact-zero advecting velocity yields NaN like the reference.
//
// Kernels (one block = one 64x32 tile of one sample, halo staged in shared memory):
//   ns_cip_kernel       out = CIP-advect(field, grad(field), vel)        field: 16 B/cell   (update_density)
//   ns_velocity_kernel  v* = v - dt grad(p);  out = CIP(v*_c, v*)        20 B/cell          (update_velocity)
//   ns_vstar_kernel     v* only (building block of the reference-bug-compatible path)
//   ns_pressure_kernel  Jacobi-style pressure relaxation                 16 B/cell          (update_pressure)
#include "common.cuh"
#include "ns_math.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

constexpr int NS_TX = 64, NS_TY = 32, NS_THREADS = 256;

struct NsTile {
  int x0, y0;   // first interior cell of the tile
  int nx, ny;   // field extent
};

// Stage a (W x H) window whose top-left is (wx0, wy0); cells outside the field are left untouched (never read).
// Loads are issued in batches of NS_MLP per thread before any of them is stored: with one load in flight per thread the
// kernels were latency-bound (profiles/r01_aux.txt), not bandwidth- or issue-bound.
constexpr int NS_MLP = 10;
template <int W, int H>
__device__ __forceinline__ void ns_stage(float* s, const float* __restrict__ g, int wx0, int wy0, int nx, int ny) {
  constexpr int N = W * H;
  for (int base = threadIdx.x; base < N; base += NS_THREADS * NS_MLP) {
    float r[NS_MLP];
    bool ok[NS_MLP];
#pragma unroll
    for (int u = 0; u < NS_MLP; ++u) {
      const int i = base + u * NS_THREADS;
      const int yy = i / W, xx = i - yy * W;
      const int gx = wx0 + xx, gy = wy0 + yy;
      ok[u] = i < N && gx >= 0 && gx < nx && gy >= 0 && gy < ny;
      r[u] = ok[u] ? __ldg(g + (size_t)gy * nx + gx) : 0.0f;
    }
#pragma unroll
    for (int u = 0; u < NS_MLP; ++u)
      if (ok[u]) s[base + u * NS_THREADS] = r[u];
  }
}

__device__ __forceinline__ void ns_tile_coords(int nx, int ny, int& b, int& x0, int& y0) {
  const int tiles_x = (nx + NS_TX - 1) / NS_TX, tiles_y = (ny + NS_TY - 1) / NS_TY;
  int bid = blockIdx.x;
  x0 = (bid % tiles_x) * NS_TX; bid /= tiles_x;
  y0 = (bid % tiles_y) * NS_TY; bid /= tiles_y;
  b = bid;
}

// out[b] = CIP(field[b], vel[b]).  field_bstride / out_bstride in elements (lets callers address single planes).
__global__ void __launch_bounds__(NS_THREADS) ns_cip_kernel(float* __restrict__ out, const float* __restrict__ field,
                                                            const float* __restrict__ vel, long long field_bstride,
                                                            long long out_bstride, int nx, int ny, float dt, float dx) {
  constexpr int HALO = 2, W = NS_TX + 2 * HALO, H = NS_TY + 2 * HALO;
  __shared__ float sf[W * H];
  int b, x0, y0;
  ns_tile_coords(nx, ny, b, x0, y0);
  const float* fp = field + (size_t)b * field_bstride;
  const float* up = vel + (size_t)b * 2 * nx * ny;
  const float* vp = up + (size_t)nx * ny;
  // the advecting velocity of this thread's 8 cells: issued before the staging so both are in flight together
  constexpr int PER = NS_TX * NS_TY / NS_THREADS;
  float uu[PER], vv[PER];
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = threadIdx.x + j * NS_THREADS, x = x0 + (i % NS_TX), y = y0 + (i / NS_TX);
    const bool in = x < nx && y < ny;
    uu[j] = in ? __ldg(up + (size_t)y * nx + x) : 1.0f;
    vv[j] = in ? __ldg(vp + (size_t)y * nx + x) : 1.0f;
  }
  ns_stage<W, H>(sf, fp, x0 - HALO, y0 - HALO, nx, ny);
  __syncthreads();
  const NsConst k = ns_const(dt, dx);
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = threadIdx.x + j * NS_THREADS;
    const int lx = i % NS_TX, ly = i / NS_TX, x = x0 + lx, y = y0 + ly;
    if (x < nx && y < ny) {
      const NsCell g = ns_cell(x, y, (ly + HALO) * W + lx + HALO, W, nx, ny, uu[j], vv[j], k);
      out[(size_t)b * out_bstride + (size_t)y * nx + x] = ns_cip(sf, g, k);
    }
  }
}

// v* = v - dt * grad(p)   (op/ns_step_kernel.cu:180-202 fed by :96-112)
__global__ void __launch_bounds__(NS_THREADS) ns_vstar_kernel(float* __restrict__ vstar, const float* __restrict__ vel,
                                                              const float* __restrict__ pres, int nx, int ny, float dt,
                                                              float dx) {
  constexpr int HALO = 1, W = NS_TX + 2 * HALO, H = NS_TY + 2 * HALO;
  __shared__ float sp[W * H];
  int b, x0, y0;
  ns_tile_coords(nx, ny, b, x0, y0);
  const size_t plane = (size_t)nx * ny;
  ns_stage<W, H>(sp, pres + b * plane, x0 - HALO, y0 - HALO, nx, ny);
  __syncthreads();
  const float rd = 1.0f / dx;
  for (int i = threadIdx.x; i < NS_TX * NS_TY; i += NS_THREADS) {
    const int lx = i % NS_TX, ly = i / NS_TX, x = x0 + lx, y = y0 + ly;
    if (x < nx && y < ny) {
      const int c = (ly + HALO) * W + lx + HALO;
      size_t o = (size_t)b * 2 * plane + (size_t)y * nx + x;
      vstar[o] = __ldg(vel + o) - ns_ddx(sp, c, x, nx, rd) * dt;
      vstar[o + plane] = __ldg(vel + o + plane) - ns_ddy(sp, c, y, ny, W, rd) * dt;
    }
  }
}

// Fused update_velocity with per-sample semantics (== reference for B = 1): op/ns_step.cpp:59-92 in one launch.
__global__ void __launch_bounds__(NS_THREADS) ns_velocity_kernel(float* __restrict__ out, const float* __restrict__ vel,
                                                                 const float* __restrict__ pres, int nx, int ny,
                                                                 float dt, float dx) {
  constexpr int HV = 2, WV = NS_TX + 2 * HV, HHV = NS_TY + 2 * HV;   // v* window (CIP needs +-2)
  constexpr int HP = 3, WP = NS_TX + 2 * HP, HHP = NS_TY + 2 * HP;   // pressure window (+1 for its gradient)
  __shared__ float sp[WP * HHP];
  __shared__ float su[WV * HHV];
  __shared__ float sv[WV * HHV];
  int b, x0, y0;
  ns_tile_coords(nx, ny, b, x0, y0);
  const size_t plane = (size_t)nx * ny;
  const float* up = vel + (size_t)b * 2 * plane;
  const float* vp = up + plane;
  ns_stage<WP, HHP>(sp, pres + b * plane, x0 - HP, y0 - HP, nx, ny);
  ns_stage<WV, HHV>(su, up, x0 - HV, y0 - HV, nx, ny);
  ns_stage<WV, HHV>(sv, vp, x0 - HV, y0 - HV, nx, ny);
  __syncthreads();
  const NsConst k = ns_const(dt, dx);
  // v* in place over the whole window (every window cell inside the field)
  for (int i = threadIdx.x; i < WV * HHV; i += NS_THREADS) {
    const int lx = i % WV, ly = i / WV, x = x0 - HV + lx, y = y0 - HV + ly;
    if (x >= 0 && x < nx && y >= 0 && y < ny) {
      const int c = (ly + HP - HV) * WP + lx + HP - HV;
      su[i] = su[i] - ns_ddx(sp, c, x, nx, k.rd) * dt;
      sv[i] = sv[i] - ns_ddy(sp, c, y, ny, WP, k.rd) * dt;
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NS_TX * NS_TY; i += NS_THREADS) {
    const int lx = i % NS_TX, ly = i / NS_TX, x = x0 + lx, y = y0 + ly;
    if (x < nx && y < ny) {
      const int c = (ly + HV) * WV + lx + HV;
      const NsCell g = ns_cell(x, y, c, WV, nx, ny, su[c], sv[c], k);
      size_t o = (size_t)b * 2 * plane + (size_t)y * nx + x;
      out[o] = ns_cip(su, g, k);
      out[o + plane] = ns_cip(sv, g, k);
    }
  }
}

// op/ns_step_kernel.cu:204-234
__global__ void __launch_bounds__(NS_THREADS) ns_pressure_kernel(float* __restrict__ out, const float* __restrict__ pres,
                                                                 const float* __restrict__ vel, int nx, int ny, float dt,
                                                                 float dx) {
  constexpr int HALO = 1, W = NS_TX + 2 * HALO, H = NS_TY + 2 * HALO;
  __shared__ float sp[W * H];
  __shared__ float su[W * H];
  __shared__ float sv[W * H];
  int b, x0, y0;
  ns_tile_coords(nx, ny, b, x0, y0);
  const size_t plane = (size_t)nx * ny;
  ns_stage<W, H>(sp, pres + b * plane, x0 - HALO, y0 - HALO, nx, ny);
  ns_stage<W, H>(su, vel + (size_t)b * 2 * plane, x0 - HALO, y0 - HALO, nx, ny);
  ns_stage<W, H>(sv, vel + (size_t)b * 2 * plane + plane, x0 - HALO, y0 - HALO, nx, ny);
  __syncthreads();
  const NsConst k = ns_const(dt, dx);
  for (int i = threadIdx.x; i < NS_TX * NS_TY; i += NS_THREADS) {
    const int lx = i % NS_TX, ly = i / NS_TX, x = x0 + lx, y = y0 + ly;
    if (x < nx && y < ny)
      out[b * plane + (size_t)y * nx + x] = ns_pressure(sp, su, sv, (ly + HALO) * W + lx + HALO, x, y, W, nx, ny, k);
  }
}

// ---- vectorised strip kernels (W % 4 == 0, 16-byte aligned planes) --------------------------------------------------
// One thread = 4 consecutive cells of a row, neighbourhood in registers (ns_math.cuh "strip form").  The shared-memory
// kernels above ran at 80-85 % issue-slot utilisation with 160-475 instructions per cell (profiles/r01_ns_ncu.txt):
// per-element staging, per-cell wall cases and window index arithmetic, not the stencil, were the cost.  Here loads are
// float4 / float2, walls are ghost cells fixed once per strip, and the upwind choice is a handful of selects.
constexpr int NS4_BX = 16, NS4_BY = 16;

__device__ __forceinline__ void ns_load_win(float (&w)[5][8], const float* __restrict__ f, int x0, int y, int nx, int ny) {
#pragma unroll
  for (int r = 0; r < 5; ++r) {
    const int gy = y - 2 + r;
    float2 a = make_float2(0.f, 0.f), c = a;
    float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
    if (gy >= 0 && gy < ny) {
      const float* rp = f + (size_t)gy * nx + x0;
      if (x0 > 0) a = __ldg(reinterpret_cast<const float2*>(rp - 2));
      b = __ldg(reinterpret_cast<const float4*>(rp));
      if (x0 + 4 < nx) c = __ldg(reinterpret_cast<const float2*>(rp + 4));
    }
    w[r][0] = a.x; w[r][1] = a.y; w[r][2] = b.x; w[r][3] = b.y; w[r][4] = b.z; w[r][5] = b.w; w[r][6] = c.x; w[r][7] = c.y;
  }
}

__global__ void __launch_bounds__(NS4_BX * NS4_BY) ns_cip4_kernel(float* __restrict__ out, const float* __restrict__ field,
                                                                  const float* __restrict__ vel, long long field_bstride,
                                                                  long long out_bstride, int nx, int ny, float dt, float dx) {
  const int x0 = (blockIdx.x * NS4_BX + threadIdx.x) * 4, y = blockIdx.y * NS4_BY + threadIdx.y, b = blockIdx.z;
  if (x0 >= nx || y >= ny) return;
  const size_t plane = (size_t)nx * ny, o = (size_t)y * nx + x0;
  const float4 u4 = __ldg(reinterpret_cast<const float4*>(vel + (size_t)b * 2 * plane + o));
  const float4 v4 = __ldg(reinterpret_cast<const float4*>(vel + (size_t)b * 2 * plane + plane + o));
  float w[5][8];
  ns_load_win(w, field + (size_t)b * field_bstride, x0, y, nx, ny);
  ns_fix_ghosts(w, x0, y, nx, ny);
  const float uu[4] = {u4.x, u4.y, u4.z, u4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w};
  float r[4];
  ns_cip_strip(w, uu, vv, x0, y, nx, ny, ns_const(dt, dx), r);
  *reinterpret_cast<float4*>(out + (size_t)b * out_bstride + o) = make_float4(r[0], r[1], r[2], r[3]);
}

// row y at x0-1 .. x0+4 and rows y-1 / y+1 at x0 .. x0+3, REFLECTED at the walls (pressure stencil)
__device__ __forceinline__ void ns_load_cross_reflect(const float* __restrict__ f, int x0, int y, int nx, int ny,
                                                      float (&c)[6], float (&d)[4], float (&u)[4]) {
  const float* rp = f + (size_t)y * nx + x0;
  const float4 b = __ldg(reinterpret_cast<const float4*>(rp));
  const float4 lo = __ldg(reinterpret_cast<const float4*>(f + (size_t)ns_reflect(y - 1, ny) * nx + x0));
  const float4 hi = __ldg(reinterpret_cast<const float4*>(f + (size_t)ns_reflect(y + 1, ny) * nx + x0));
  c[0] = x0 > 0 ? __ldg(rp - 1) : b.y;
  c[5] = x0 + 4 < nx ? __ldg(rp + 4) : b.z;
  c[1] = b.x; c[2] = b.y; c[3] = b.z; c[4] = b.w;
  d[0] = lo.x; d[1] = lo.y; d[2] = lo.z; d[3] = lo.w;
  u[0] = hi.x; u[1] = hi.y; u[2] = hi.z; u[3] = hi.w;
}

__global__ void __launch_bounds__(NS4_BX * NS4_BY) ns_pressure4_kernel(float* __restrict__ out, const float* __restrict__ pres,
                                                                       const float* __restrict__ vel, int nx, int ny,
                                                                       float dt, float dx) {
  const int x0 = (blockIdx.x * NS4_BX + threadIdx.x) * 4, y = blockIdx.y * NS4_BY + threadIdx.y, b = blockIdx.z;
  if (x0 >= nx || y >= ny) return;
  const size_t plane = (size_t)nx * ny;
  float pc[6], pd[4], pu[4], uc[6], ud[4], uu[4], vc[6], vd[4], vu[4], r[4];
  ns_load_cross_reflect(pres + b * plane, x0, y, nx, ny, pc, pd, pu);
  ns_load_cross_reflect(vel + (size_t)b * 2 * plane, x0, y, nx, ny, uc, ud, uu);
  ns_load_cross_reflect(vel + (size_t)b * 2 * plane + plane, x0, y, nx, ny, vc, vd, vu);
  ns_pressure_strip(pc, pd, pu, uc, ud, uu, vc, vd, vu, ns_const(dt, dx), r);
  *reinterpret_cast<float4*>(out + b * plane + (size_t)y * nx + x0) = make_float4(r[0], r[1], r[2], r[3]);
}

// update_velocity in one launch: phase 1 writes v* = v - dt grad(p) of the tile + halo (2 rows, 4 columns: keeps the
// strips 16-byte aligned) into shared memory straight from vector loads; phase 2 advects u* and v* by (u*, v*) with the
// same strip code as the density kernel, its windows read from shared memory.
constexpr int NSV_W = NS_TX + 8, NSV_H = NS_TY + 4, NSV_STRIPS = NSV_W / 4;

__device__ __forceinline__ void ns_load_win_smem(float (&w)[5][8], const float* s, int lx0, int ly) {
#pragma unroll
  for (int r = 0; r < 5; ++r) {
    const float* rp = s + (ly + r) * NSV_W + lx0 + 2;      // window column 0 = tile column lx0 - 2 = smem column lx0 + 2
    const float2 a = *reinterpret_cast<const float2*>(rp);
    const float4 b = *reinterpret_cast<const float4*>(rp + 2);
    const float2 c = *reinterpret_cast<const float2*>(rp + 6);
    w[r][0] = a.x; w[r][1] = a.y; w[r][2] = b.x; w[r][3] = b.y; w[r][4] = b.z; w[r][5] = b.w; w[r][6] = c.x; w[r][7] = c.y;
  }
}

__global__ void __launch_bounds__(NS_THREADS, 3) ns_velocity4_kernel(float* __restrict__ out, const float* __restrict__ vel,
                                                                  const float* __restrict__ pres, int nx, int ny,
                                                                  float dt, float dx) {
  __shared__ __align__(16) float su[NSV_W * NSV_H];
  __shared__ __align__(16) float sv[NSV_W * NSV_H];
  const int tx0 = blockIdx.x * NS_TX, ty0 = blockIdx.y * NS_TY, b = blockIdx.z;
  const size_t plane = (size_t)nx * ny;
  const float* pp = pres + b * plane;
  const float* up = vel + (size_t)b * 2 * plane;
  const float* vp = up + plane;
  const NsConst k = ns_const(dt, dx);
#pragma unroll
  for (int it = 0; it < (NSV_STRIPS * NSV_H + NS_THREADS - 1) / NS_THREADS; ++it) {
    const int s = threadIdx.x + it * NS_THREADS;
    const int sy = s / NSV_STRIPS, sx = s - sy * NSV_STRIPS;
    const int gy = ty0 - 2 + sy, gx0 = tx0 - 4 + sx * 4;
    if (s < NSV_STRIPS * NSV_H && gy >= 0 && gy < ny && gx0 >= 0 && gx0 < nx) {
      const size_t o = (size_t)gy * nx + gx0;
      const float4 pb = __ldg(reinterpret_cast<const float4*>(pp + o));
      const float4 u4 = __ldg(reinterpret_cast<const float4*>(up + o));
      const float4 v4 = __ldg(reinterpret_cast<const float4*>(vp + o));
      float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), hi = lo;
      float pl = 0.f, pr = 0.f;
      if (gy > 0) lo = __ldg(reinterpret_cast<const float4*>(pp + o - nx));
      if (gy < ny - 1) hi = __ldg(reinterpret_cast<const float4*>(pp + o + nx));
      if (gx0 > 0) pl = __ldg(pp + o - 1);
      if (gx0 + 4 < nx) pr = __ldg(pp + o + 4);
      float pc[6] = {pl, pb.x, pb.y, pb.z, pb.w, pr};
      float pd[4] = {lo.x, lo.y, lo.z, lo.w}, pu[4] = {hi.x, hi.y, hi.z, hi.w};
      if (gx0 == 0) pc[0] = 2.0f * pc[1] - pc[2];
      if (gx0 + 4 == nx) pc[5] = 2.0f * pc[4] - pc[3];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (gy == 0) pd[c] = 2.0f * pc[c + 1] - pu[c];
        if (gy == ny - 1) pu[c] = 2.0f * pc[c + 1] - pd[c];
      }
      const float u[4] = {u4.x, u4.y, u4.z, u4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
      float us[4], vs[4];
      ns_vstar_strip(pc, pd, pu, u, v, k, us, vs);
      *reinterpret_cast<float4*>(su + sy * NSV_W + sx * 4) = make_float4(us[0], us[1], us[2], us[3]);
      *reinterpret_cast<float4*>(sv + sy * NSV_W + sx * 4) = make_float4(vs[0], vs[1], vs[2], vs[3]);
    }
  }
  __syncthreads();
  const int lx0 = (threadIdx.x % NS4_BX) * 4, x0 = tx0 + lx0;
#pragma unroll
  for (int pass = 0; pass < NS_TY / NS4_BY; ++pass) {
    const int ly = threadIdx.x / NS4_BX + pass * NS4_BY, y = ty0 + ly;
    if (x0 < nx && y < ny) {
      float w[5][8], uu[4], vv[4], r[4];
      const size_t o = (size_t)b * 2 * plane + (size_t)y * nx + x0;
      ns_load_win_smem(w, su, lx0, ly);
#pragma unroll
      for (int c = 0; c < 4; ++c) uu[c] = w[2][c + 2];
      const float4 vc = *reinterpret_cast<const float4*>(sv + (ly + 2) * NSV_W + lx0 + 4);
      vv[0] = vc.x; vv[1] = vc.y; vv[2] = vc.z; vv[3] = vc.w;
      ns_fix_ghosts(w, x0, y, nx, ny);
      ns_cip_strip(w, uu, vv, x0, y, nx, ny, k, r);
      *reinterpret_cast<float4*>(out + o) = make_float4(r[0], r[1], r[2], r[3]);
      ns_load_win_smem(w, sv, lx0, ly);
      ns_fix_ghosts(w, x0, y, nx, ny);
      ns_cip_strip(w, uu, vv, x0, y, nx, ny, k, r);
      *reinterpret_cast<float4*>(out + o + plane) = make_float4(r[0], r[1], r[2], r[3]);
    }
  }
}

static inline bool ns_al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static inline bool ns_grid_ok(int B, int H) { return B <= 65535 && (H + NS_TY - 1) / NS_TY <= 65535 && (H + NS4_BY - 1) / NS4_BY <= 65535; }
static inline dim3 ns4_grid(int B, int H, int W) { return dim3((W / 4 + NS4_BX - 1) / NS4_BX, (H + NS4_BY - 1) / NS4_BY, B); }

static inline unsigned ns_grid(int B, int H, int W) {
  return (unsigned)B * ((W + NS_TX - 1) / NS_TX) * ((H + NS_TY - 1) / NS_TY);
}

}  // namespace b200

using namespace b200;

static inline int ns_check(const void* a, const void* b, const void* c, int B, int H, int W) {
  if (!a || !b || !c || B < 0 || H < 2 || W < 2) return B200_ERR_ARG;
  if ((long long)B * ((W + NS_TX - 1) / NS_TX) * ((H + NS_TY - 1) / NS_TY) > 0x7fffffffLL) return B200_ERR_UNSUPPORTED;
  return B200_OK;
}

extern "C" int b200_ns_cip_advect(float* out, const float* field, const float* vel, long long field_bstride,
                                  long long out_bstride, int B, int H, int W, float dt, float dx, cudaStream_t st) {
  int rc = ns_check(out, field, vel, B, H, W);
  if (rc || B == 0) return rc;
  if (W % 4 == 0 && ns_al16(out) && ns_al16(field) && ns_al16(vel) && field_bstride % 4 == 0 && out_bstride % 4 == 0 &&
      ns_grid_ok(B, H)) {
    ns_cip4_kernel<<<ns4_grid(B, H, W), dim3(NS4_BX, NS4_BY), 0, st>>>(out, field, vel, field_bstride, out_bstride, W, H,
                                                                      dt, dx);
    B200_RETURN_LAST();
  }
  ns_cip_kernel<<<ns_grid(B, H, W), NS_THREADS, 0, st>>>(out, field, vel, field_bstride, out_bstride, W, H, dt, dx);
  B200_RETURN_LAST();
}

extern "C" int b200_ns_update_density(float* dens_out, const float* dens, const float* vel, int B, int H, int W,
                                      float dt, float dx, cudaStream_t st) {
  return b200_ns_cip_advect(dens_out, dens, vel, (long long)H * W, (long long)H * W, B, H, W, dt, dx, st);
}

extern "C" int b200_ns_velocity_nonadvect(float* vstar, const float* vel, const float* pres, int B, int H, int W,
                                          float dt, float dx, cudaStream_t st) {
  int rc = ns_check(vstar, vel, pres, B, H, W);
  if (rc || B == 0) return rc;
  ns_vstar_kernel<<<ns_grid(B, H, W), NS_THREADS, 0, st>>>(vstar, vel, pres, W, H, dt, dx);
  B200_RETURN_LAST();
}

// bug_compat = 0: per-sample semantics in ONE launch (scratch unused, may be NULL).
// bug_compat = 1: reproduces the reference's behaviour for B > 1 (op/ns_step.cpp:70-72 hands non-contiguous u/v
//   views to kernels that assume a dense batch stride: sample b advects flat plane b (as "u") and b+1 (as "v")
//   of v*).  Needs `scratch` of B*2*H*W floats.
extern "C" int b200_ns_update_velocity(float* vel_out, const float* vel, const float* pres, float* scratch, int B,
                                       int H, int W, float dt, float dx, int bug_compat, cudaStream_t st) {
  int rc = ns_check(vel_out, vel, pres, B, H, W);
  if (rc || B == 0) return rc;
  if (!bug_compat && W % 4 == 0 && ns_al16(vel_out) && ns_al16(vel) && ns_al16(pres) && ns_grid_ok(B, H)) {
    ns_velocity4_kernel<<<dim3((W + NS_TX - 1) / NS_TX, (H + NS_TY - 1) / NS_TY, B), NS_THREADS, 0, st>>>(vel_out, vel, pres,
                                                                                                       W, H, dt, dx);
    B200_RETURN_LAST();
  }
  if (!bug_compat) {
    ns_velocity_kernel<<<ns_grid(B, H, W), NS_THREADS, 0, st>>>(vel_out, vel, pres, W, H, dt, dx);
    B200_RETURN_LAST();
  }
  if (!scratch) return B200_ERR_ARG;
  const long long plane = (long long)H * W;
  ns_vstar_kernel<<<ns_grid(B, H, W), NS_THREADS, 0, st>>>(scratch, vel, pres, W, H, dt, dx);
  rc = b200_ns_cip_advect(vel_out, scratch, scratch, plane, 2 * plane, B, H, W, dt, dx, st);
  if (rc) return rc;
  return b200_ns_cip_advect(vel_out + plane, scratch + plane, scratch, plane, 2 * plane, B, H, W, dt, dx, st);
}

extern "C" int b200_ns_update_pressure(float* pres_out, const float* pres, const float* vel, int B, int H, int W,
                                       float dt, float dx, cudaStream_t st) {
  int rc = ns_check(pres_out, pres, vel, B, H, W);
  if (rc || B == 0) return rc;
  if (W % 4 == 0 && ns_al16(pres_out) && ns_al16(pres) && ns_al16(vel) && ns_grid_ok(B, H)) {
    ns_pressure4_kernel<<<ns4_grid(B, H, W), dim3(NS4_BX, NS4_BY), 0, st>>>(pres_out, pres, vel, W, H, dt, dx);
    B200_RETURN_LAST();
  }
  ns_pressure_kernel<<<ns_grid(B, H, W), NS_THREADS, 0, st>>>(pres_out, pres, vel, W, H, dt, dx);
  B200_RETURN_LAST();
}

extern "C" int b200_ns_step_fused(float* dens_out, float* vel_out, float* pres_out, const float* dens, const float* vel,
                                  const float* pres, int B, int H, int W, float dt, float dx, cudaStream_t st) {
  if (!dens_out || !vel_out || !pres_out || !dens) return B200_ERR_ARG;
  if (dens_out == dens || vel_out == vel || pres_out == pres) return B200_ERR_ARG;
  int rc = b200_ns_update_velocity(vel_out, vel, pres, nullptr, B, H, W, dt, dx, 0, st);
  if (rc) return rc;
  rc = b200_ns_update_pressure(pres_out, pres, vel_out, B, H, W, dt, dx, st);
  if (rc) return rc;
  return b200_ns_update_density(dens_out, dens, vel_out, B, H, W, dt, dx, st);
}
