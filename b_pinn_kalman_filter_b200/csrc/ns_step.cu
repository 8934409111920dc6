// Navier-Stokes simulator step (pinn_kalman/simulator.py:54-57) as shared-memory stencil kernels for sm_100a.
// Replaces op/ns_step_kernel.cu:96-234 (update_gradient / cip_advect / velocity_update / pressure_update kernels,
// one 1-cell block each with batch as blockDim) and the 9-launch host composites in op/ns_step.cpp:45-102.
//
// Fields are [B, C, H, W] fp32; x is the fast (W) index, y the slow (H) index (identical to the reference's
// `y * gridDim.x + x` indexing whenever H == W, the only case the reference indexes correctly).
// Arithmetic mirrors the reference's mixed float/double expressions (its double literals promote sub-expressions),
// so results agree to ~1e-6 relative; exact-zero advecting velocity yields inf/NaN exactly like the reference.
//
// Kernels (one block = one 64x32 tile of one sample, halo staged in shared memory):
//   ns_cip_kernel       out = CIP-advect(field, grad(field), vel)        field: 16 B/cell   (update_density)
//   ns_velocity_kernel  v* = v - dt grad(p);  out = CIP(v*_c, v*)        20 B/cell          (update_velocity)
//   ns_vstar_kernel     v* only (building block of the reference-bug-compatible path)
//   ns_pressure_kernel  Jacobi-style pressure relaxation                 16 B/cell          (update_pressure)
#include "common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

constexpr int NS_TX = 64, NS_TY = 32, NS_THREADS = 256;

struct NsTile {
  int x0, y0;   // first interior cell of the tile
  int nx, ny;   // field extent
};

// Stage a (w x h) window whose top-left is (wx0, wy0); cells outside the field are left untouched (never read).
__device__ __forceinline__ void ns_stage(float* s, const float* __restrict__ g, int wx0, int wy0, int w, int h, int nx,
                                         int ny) {
  for (int i = threadIdx.x; i < w * h; i += NS_THREADS) {
    int yy = i / w, xx = i - yy * w;
    int gx = wx0 + xx, gy = wy0 + yy;
    if (gx >= 0 && gx < nx && gy >= 0 && gy < ny) s[i] = __ldg(g + (size_t)gy * nx + gx);
  }
}

// Window accessor: s holds cells [wx0, wx0+w) x [wy0, ...).
struct NsWin {
  const float* s; int wx0, wy0, w, nx, ny;
  __device__ __forceinline__ float at(int x, int y) const { return s[(y - wy0) * w + (x - wx0)]; }
  // op/ns_step_kernel.cu:51-75 (central inside, one-sided at the walls; float arithmetic)
  __device__ __forceinline__ float ddx(int x, int y, float dx) const {
    if (x == 0) return (at(x + 1, y) - at(x, y)) / dx;
    if (x == nx - 1) return (at(x, y) - at(x - 1, y)) / dx;
    return (at(x + 1, y) - at(x - 1, y)) / dx / 2;
  }
  __device__ __forceinline__ float ddy(int x, int y, float dx) const {
    if (y == 0) return (at(x, y + 1) - at(x, y)) / dx;
    if (y == ny - 1) return (at(x, y) - at(x, y - 1)) / dx;
    return (at(x, y + 1) - at(x, y - 1)) / dx / 2;
  }
};

__device__ __forceinline__ int ns_reflect(int i, int n) { return i < 0 ? -i : (i > n - 1 ? 2 * n - 2 - i : i); }
__device__ __forceinline__ int ns_sign(float v) { return v < 0.0f ? -1 : (v > 0.0f ? 1 : 0); }

// CIP (cubic-interpolated pseudo-particle) advection of the windowed field by (u, v) at cell (x, y).
// Follows op/ns_step_kernel.cu:114-158 term by term, including which sub-expressions are evaluated in double.
__device__ __forceinline__ float ns_cip(const NsWin& f, int x, int y, float u, float v, float dt, float dx) {
  const int xs = ns_sign(u), ys = ns_sign(v);
  const int xm = ns_reflect(x - xs, f.nx), ym = ns_reflect(y - ys, f.ny);
  const float c00 = f.at(x, y), c10 = f.at(xm, y), c01 = f.at(x, ym), c11 = f.at(xm, ym);
  const float gx00 = f.ddx(x, y, dx), gx10 = f.ddx(xm, y, dx), gx01 = f.ddx(x, ym, dx);
  const float gy00 = f.ddy(x, y, dx), gy01 = f.ddy(x, ym, dx), gy10 = f.ddy(xm, y, dx);
  const float tmp1 = c00 - c01 - c10 + c11;
  const float tmp2 = c10 - c00;
  const float tmp3 = c01 - c00;
  const float xsd = xs * dx * dx * dx;
  const float ysd = ys * dx * dx * dx;
  const float a = (float)(((double)(xs * (gx10 + gx00) * dx) - 2.0 * (double)(-tmp2)) / (double)xsd);
  const float b = (float)(((double)(ys * (gy01 + gy00) * dx) - 2.0 * (double)(-tmp3)) / (double)ysd);
  const float c = (-tmp1 - xs * (gx01 - gx00) * dx) / ysd;
  const float d = (-tmp1 - ys * (gy10 - gy00) * dx) / xsd;
  const float e = (float)((3.0 * (double)tmp2 + (double)xs * ((double)gx10 + 2.0 * (double)gx00) * (double)dx) / (double)dx / (double)dx);
  const float ff = (float)((3.0 * (double)tmp3 + (double)ys * ((double)gy01 + 2.0 * (double)gy00) * (double)dx) / (double)dx / (double)dx);
  const float g = (-(gy10 - gy00) + c * dx * dx) / (xs * dx);
  const float X = -u * dt, Y = -v * dt;
  return ((a * X + c * Y + e) * X + g * Y + gx00) * X + ((b * Y + d * X + ff) * Y + gy00) * Y + c00;
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
  ns_stage(sf, fp, x0 - HALO, y0 - HALO, W, H, nx, ny);
  __syncthreads();
  NsWin f{sf, x0 - HALO, y0 - HALO, W, nx, ny};
  for (int i = threadIdx.x; i < NS_TX * NS_TY; i += NS_THREADS) {
    int x = x0 + (i % NS_TX), y = y0 + (i / NS_TX);
    if (x < nx && y < ny) {
      size_t o = (size_t)y * nx + x;
      out[(size_t)b * out_bstride + o] = ns_cip(f, x, y, __ldg(up + o), __ldg(vp + o), dt, dx);
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
  ns_stage(sp, pres + b * plane, x0 - HALO, y0 - HALO, W, H, nx, ny);
  __syncthreads();
  NsWin p{sp, x0 - HALO, y0 - HALO, W, nx, ny};
  for (int i = threadIdx.x; i < NS_TX * NS_TY; i += NS_THREADS) {
    int x = x0 + (i % NS_TX), y = y0 + (i / NS_TX);
    if (x < nx && y < ny) {
      size_t o = (size_t)b * 2 * plane + (size_t)y * nx + x;
      vstar[o] = __ldg(vel + o) - p.ddx(x, y, dx) * dt;
      vstar[o + plane] = __ldg(vel + o + plane) - p.ddy(x, y, dx) * dt;
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
  ns_stage(sp, pres + b * plane, x0 - HP, y0 - HP, WP, HHP, nx, ny);
  ns_stage(su, up, x0 - HV, y0 - HV, WV, HHV, nx, ny);
  ns_stage(sv, vp, x0 - HV, y0 - HV, WV, HHV, nx, ny);
  __syncthreads();
  NsWin p{sp, x0 - HP, y0 - HP, WP, nx, ny};
  // v* in place over the whole window (every window cell inside the field)
  for (int i = threadIdx.x; i < WV * HHV; i += NS_THREADS) {
    int x = x0 - HV + (i % WV), y = y0 - HV + (i / WV);
    if (x >= 0 && x < nx && y >= 0 && y < ny) {
      su[i] = su[i] - p.ddx(x, y, dx) * dt;
      sv[i] = sv[i] - p.ddy(x, y, dx) * dt;
    }
  }
  __syncthreads();
  NsWin fu{su, x0 - HV, y0 - HV, WV, nx, ny}, fv{sv, x0 - HV, y0 - HV, WV, nx, ny};
  for (int i = threadIdx.x; i < NS_TX * NS_TY; i += NS_THREADS) {
    int x = x0 + (i % NS_TX), y = y0 + (i / NS_TX);
    if (x < nx && y < ny) {
      const float u = fu.at(x, y), v = fv.at(x, y);
      size_t o = (size_t)b * 2 * plane + (size_t)y * nx + x;
      out[o] = ns_cip(fu, x, y, u, v, dt, dx);
      out[o + plane] = ns_cip(fv, x, y, u, v, dt, dx);
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
  ns_stage(sp, pres + b * plane, x0 - HALO, y0 - HALO, W, H, nx, ny);
  ns_stage(su, vel + (size_t)b * 2 * plane, x0 - HALO, y0 - HALO, W, H, nx, ny);
  ns_stage(sv, vel + (size_t)b * 2 * plane + plane, x0 - HALO, y0 - HALO, W, H, nx, ny);
  __syncthreads();
  NsWin p{sp, x0 - HALO, y0 - HALO, W, nx, ny}, fu{su, x0 - HALO, y0 - HALO, W, nx, ny},
      fv{sv, x0 - HALO, y0 - HALO, W, nx, ny};
  for (int i = threadIdx.x; i < NS_TX * NS_TY; i += NS_THREADS) {
    int x = x0 + (i % NS_TX), y = y0 + (i / NS_TX);
    if (x < nx && y < ny) {
      const int xu = ns_reflect(x + 1, nx), xd = ns_reflect(x - 1, nx);
      const int yu = ns_reflect(y + 1, ny), yd = ns_reflect(y - 1, ny);
      const float sxx = fu.at(xu, y) - fu.at(xd, y), sxy = fv.at(xu, y) - fv.at(xd, y);   // sub_x.{x,y}
      const float syx = fu.at(x, yu) - fu.at(x, yd), syy = fv.at(x, yu) - fv.at(x, yd);   // sub_y.{x,y}
      const float aver = (float)(0.25 * (double)(p.at(xd, y) + p.at(xu, y) + p.at(x, yd) + p.at(x, yu)));
      const double pred = (double)aver + (double)(sxx * sxx + syy * syy + (syx * sxy)) / 8.0 -
                          (double)(dx * (sxx + syy) / (8 * dt));
      out[b * plane + (size_t)y * nx + x] = (float)pred;
    }
  }
}

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
  if (!bug_compat) {
    ns_velocity_kernel<<<ns_grid(B, H, W), NS_THREADS, 0, st>>>(vel_out, vel, pres, W, H, dt, dx);
    B200_RETURN_LAST();
  }
  if (!scratch) return B200_ERR_ARG;
  const long long plane = (long long)H * W;
  ns_vstar_kernel<<<ns_grid(B, H, W), NS_THREADS, 0, st>>>(scratch, vel, pres, W, H, dt, dx);
  ns_cip_kernel<<<ns_grid(B, H, W), NS_THREADS, 0, st>>>(vel_out, scratch, scratch, plane, 2 * plane, W, H, dt, dx);
  ns_cip_kernel<<<ns_grid(B, H, W), NS_THREADS, 0, st>>>(vel_out + plane, scratch + plane, scratch, plane, 2 * plane, W,
                                                         H, dt, dx);
  B200_RETURN_LAST();
}

extern "C" int b200_ns_update_pressure(float* pres_out, const float* pres, const float* vel, int B, int H, int W,
                                       float dt, float dx, cudaStream_t st) {
  int rc = ns_check(pres_out, pres, vel, B, H, W);
  if (rc || B == 0) return rc;
  ns_pressure_kernel<<<ns_grid(B, H, W), NS_THREADS, 0, st>>>(pres_out, pres, vel, W, H, dt, dx);
  B200_RETURN_LAST();
}
