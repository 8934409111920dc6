// Per-cell arithmetic of the Navier-Stokes step, shared by the CUDA kernels (ns_step.cu) and by the CPU test that
// pins it to the oracle without a GPU (tests/test_ns_math_cpu.py compiles this header with g++).
//
// The reference (op/ns_step_kernel.cu:51-75, 114-158, 204-234) writes the stencils with IEEE divisions and with double
// literals that promote sub-expressions to fp64.  That form is issue-bound (about 15 clk/cell/SM on B200: six fp64
// divisions and ~15 fp32 divisions per CIP evaluation).  Here every division is a multiplication by a reciprocal
// computed once per launch and everything stays in fp32; the upwind signs are +-1, so 1/(xs*dx^3) = xs/dx^3 exactly.
// Against the oracle (which keeps the reference's promotions) the fields agree to ~1e-7 relative -- the cancelling sums
// are scaled by (u dt)^k / dx^k <= 0.125 before they reach the output -- well inside north_star's 1e-5.
// An exactly zero advecting velocity gives NaN, as in the reference (0/0 in its `a` / `b` coefficients).
#pragma once
#include <math.h>
#ifdef __CUDACC__
#define NS_HD __host__ __device__ __forceinline__
#else
#define NS_HD inline
#endif

namespace b200 {

NS_HD int ns_reflect(int i, int n) { return i < 0 ? -i : (i > n - 1 ? 2 * n - 2 - i : i); }
NS_HD int ns_sign(float v) { return v < 0.0f ? -1 : (v > 0.0f ? 1 : 0); }

struct NsConst {
  float dt, dx, rd, dtrd, k8;
};
NS_HD NsConst ns_const(float dt, float dx) {
  NsConst k;
  k.dt = dt; k.dx = dx; k.rd = 1.0f / dx; k.dtrd = dt * k.rd;
  k.k8 = dx / (8 * dt);
  return k;
}

// Difference taps along an axis of extent n at coordinate i (pitch p between neighbours): central inside, one-sided at
// the walls (op/ns_step_kernel.cu:51-75).
NS_HD void ns_taps(int i, int n, int p, float rd, int& lo, int& hi, float& s) {
  const bool l = i > 0, h = i < n - 1;
  lo = l ? -p : 0;
  hi = h ? p : 0;
  s = (l && h) ? 0.5f * rd : rd;
}

// Geometry of one cell's CIP stencil inside a window of pitch w; shared by every field advected by the same velocity.
struct NsCell {
  int c, ox, oy;             // centre index, offsets to the upwind column / row (reflected at the walls)
  int xl, xr, xlm, xrm;      // d/dx taps at column x and at column xm
  int yl, yr, ylm, yrm;      // d/dy taps at row y and at row ym
  float sx, sxm, sy, sym;    // 1 (one-sided) or 0.5 (central): differences scaled to gradient * dx
  float fxs, fys, p, q;      // upwind signs; p = -|u| dt/dx, q = -|v| dt/dx
};

NS_HD NsCell ns_cell(int x, int y, int c, int w, int nx, int ny, float u, float v, const NsConst& k) {
  NsCell g;
  const int xs = ns_sign(u), ys = ns_sign(v);
  const int xm = ns_reflect(x - xs, nx), ym = ns_reflect(y - ys, ny);
  g.c = c; g.ox = xm - x; g.oy = (ym - y) * w;
  ns_taps(x, nx, 1, 1.0f, g.xl, g.xr, g.sx);
  ns_taps(xm, nx, 1, 1.0f, g.xlm, g.xrm, g.sxm);
  ns_taps(y, ny, w, 1.0f, g.yl, g.yr, g.sy);
  ns_taps(ym, ny, w, 1.0f, g.ylm, g.yrm, g.sym);
  g.fxs = (float)xs; g.fys = (float)ys;
  g.p = -fabsf(u) * k.dtrd; g.q = -fabsf(v) * k.dtrd;
  return g;
}

// The CIP polynomial (op/ns_step_kernel.cu:130-158) in upwind-normalised form.  With X = -u dt, Y = -v dt the reference
// evaluates  a X^3 + c X^2 Y + e X^2 + g X Y + gx00 X + b Y^3 + d X Y^2 + f Y^2 + gy00 Y + c00  with coefficients that
// carry 1/(xs dx^3), 1/dx^2, 1/(xs dx).  Substituting p = xs X/dx = -|u| dt/dx, q = ys Y/dx and the oriented, dx-scaled
// gradients h = xs * gx * dx (xs^2 = 1) removes every power of dx:
//   x part   ((h10 + h00 + 2 t2) p + (3 t2 + h10 + 2 h00)) p^2 + h00 p        t2 = c10 - c00
//   cross    (C p + D q + C - hy10 + hy00) p q      C = hx00 - hx01 - t1,  D = hy00 - hy10 - t1,  t1 = c00-c01-c10+c11
// Same polynomial, ~50 instead of ~80 flops per cell; agreement with the reference form ~1e-7 (tests/test_ns_math_cpu).
// d* are differences scaled so that kx * d = xs * gradient * dx (kx = xs for pre-scaled, 0.5 xs for raw central ones).
NS_HD float ns_cip_eval(float c00, float c10, float c01, float c11, float dx00, float dx10, float dx01, float dy00,
                        float dy01, float dy10, float kx, float ky, float p, float q) {
  const float hx00 = kx * dx00, hx10 = kx * dx10, hx01 = kx * dx01;
  const float hy00 = ky * dy00, hy01 = ky * dy01, hy10 = ky * dy10;
  const float t1 = c00 - c01 - c10 + c11, t2 = c10 - c00, t3 = c01 - c00;
  const float C = hx00 - hx01 - t1, D = hy00 - hy10 - t1, G = C + (hy00 - hy10);
  const float Ax = fmaf(2.0f, t2, hx10 + hx00), Ex = fmaf(3.0f, t2, fmaf(2.0f, hx00, hx10));
  const float Ay = fmaf(2.0f, t3, hy01 + hy00), Ey = fmaf(3.0f, t3, fmaf(2.0f, hy00, hy01));
  const float xp = fmaf(fmaf(Ax, p, Ex), p, hx00) * p;
  const float yp = fmaf(fmaf(Ay, q, Ey), q, hy00) * q;
  const float cr = fmaf(C, p, fmaf(D, q, G)) * (p * q);
  return (kx == 0.0f || ky == 0.0f) ? NAN : (xp + yp) + cr + c00;
}

// CIP (cubic-interpolated pseudo-particle) advection of the field in window s: op/ns_step_kernel.cu:114-158.
NS_HD float ns_cip(const float* s, const NsCell& g, const NsConst& k) {
  const float* p = s + g.c;
  const float* px = p + g.ox;
  const float* py = p + g.oy;
  const float c00 = p[0], c10 = px[0], c01 = py[0], c11 = px[g.oy];
  const float gx00 = (p[g.xr] - p[g.xl]) * g.sx, gx10 = (px[g.xrm] - px[g.xlm]) * g.sxm;
  const float gx01 = (py[g.xr] - py[g.xl]) * g.sx;
  const float gy00 = (p[g.yr] - p[g.yl]) * g.sy, gy01 = (py[g.yrm] - py[g.ylm]) * g.sym;
  const float gy10 = (px[g.yr] - px[g.yl]) * g.sy;
  return ns_cip_eval(c00, c10, c01, c11, gx00, gx10, gx01, gy00, gy01, gy10, g.fxs, g.fys, g.p, g.q);
}

// d/dx, d/dy of the window field at (x, y) whose window index is c.
NS_HD float ns_ddx(const float* s, int c, int x, int nx, float rd) {
  int lo, hi; float sc;
  ns_taps(x, nx, 1, rd, lo, hi, sc);
  return (s[c + hi] - s[c + lo]) * sc;
}
NS_HD float ns_ddy(const float* s, int c, int y, int ny, int w, float rd) {
  int lo, hi; float sc;
  ns_taps(y, ny, w, rd, lo, hi, sc);
  return (s[c + hi] - s[c + lo]) * sc;
}

// Pressure relaxation, op/ns_step_kernel.cu:204-234.  p, u, v are windows of the same pitch, c the cell's index.
NS_HD float ns_pressure(const float* p, const float* u, const float* v, int c, int x, int y, int w, int nx, int ny,
                        const NsConst& k) {
  const int xu = ns_reflect(x + 1, nx) - x, xd = ns_reflect(x - 1, nx) - x;
  const int yu = (ns_reflect(y + 1, ny) - y) * w, yd = (ns_reflect(y - 1, ny) - y) * w;
  const float sxx = u[c + xu] - u[c + xd], sxy = v[c + xu] - v[c + xd];
  const float syx = u[c + yu] - u[c + yd], syy = v[c + yu] - v[c + yd];
  const float aver = 0.25f * (p[c + xd] + p[c + xu] + p[c + yd] + p[c + yu]);
  return aver + (sxx * sxx + syy * syy + syx * sxy) * 0.125f - (sxx + syy) * k.k8;
}

// ---- strip form: one thread owns 4 consecutive cells of a row and keeps the 5 x 8 neighbourhood in registers ------
// w[r][j] = field(y - 2 + r, x0 - 2 + j).  Cells outside the field hold anything; ns_fix_ghosts then writes the one
// ghost column / row next to a wall as the linear extrapolation 2 f0 - f1, which makes the CENTRAL difference across
// the wall equal the reference's one-sided difference there ((f1 - f0)/dx, op/ns_step_kernel.cu:51-75), so the strip
// code needs no wall cases at all.
NS_HD void ns_fix_ghosts(float (&w)[5][8], int x0, int y, int nx, int ny) {
  if (x0 == 0) {
#pragma unroll
    for (int r = 0; r < 5; ++r) w[r][1] = 2.0f * w[r][2] - w[r][3];
  }
  if (x0 + 4 == nx) {
#pragma unroll
    for (int r = 0; r < 5; ++r) w[r][6] = 2.0f * w[r][5] - w[r][4];
  }
  if (y < 2 || y >= ny - 2) {      // a real branch: 4 rows of the field, not 32 predicated FMAs for every strip
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (y == 0) w[1][j] = 2.0f * w[2][j] - w[3][j];
      if (y == 1) w[0][j] = 2.0f * w[1][j] - w[2][j];
      if (y == ny - 1) w[3][j] = 2.0f * w[2][j] - w[1][j];
      if (y == ny - 2) w[4][j] = 2.0f * w[3][j] - w[2][j];
    }
  }
}

// CIP of the strip's 4 cells (after ns_fix_ghosts); the upwind neighbour is picked with selects on the velocity signs.
NS_HD void ns_cip_strip(const float (&w)[5][8], const float (&uu)[4], const float (&vv)[4], int x0, int y, int nx,
                        int ny, const NsConst& k, float (&o)[4]) {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const int j = c + 2, x = x0 + c;
    const int xs = ns_sign(uu[c]), ys = ns_sign(vv[c]);
    // upwind column / row lies at -1 (true) or +1; reflected at the walls (op/ns_step_kernel.cu:77-83)
    const bool xn = (xs > 0 && x > 0) || (xs < 0 && x == nx - 1);
    const bool yn = (ys > 0 && y > 0) || (ys < 0 && y == ny - 1);
    const float c00 = w[2][j];
    const float c10 = xn ? w[2][j - 1] : w[2][j + 1];
    const float c01 = yn ? w[1][j] : w[3][j];
    const float c11 = yn ? (xn ? w[1][j - 1] : w[1][j + 1]) : (xn ? w[3][j - 1] : w[3][j + 1]);
    const float dx00 = w[2][j + 1] - w[2][j - 1];
    const float dx10 = (xn ? c00 : w[2][j + 2]) - (xn ? w[2][j - 2] : c00);
    const float dx01 = (yn ? w[1][j + 1] : w[3][j + 1]) - (yn ? w[1][j - 1] : w[3][j - 1]);
    const float dy00 = w[3][j] - w[1][j];
    const float dy01 = (yn ? c00 : w[4][j]) - (yn ? w[0][j] : c00);
    const float dy10 = (xn ? w[3][j - 1] : w[3][j + 1]) - (xn ? w[1][j - 1] : w[1][j + 1]);
    o[c] = ns_cip_eval(c00, c10, c01, c11, dx00, dx10, dx01, dy00, dy01, dy10, 0.5f * (float)xs, 0.5f * (float)ys,
                       -fabsf(uu[c]) * k.dtrd, -fabsf(vv[c]) * k.dtrd);
  }
}

// v* = v - dt grad(p) for a strip.  pc = p(y, x0-1 .. x0+4), pd / pu = p(y-1 / y+1, x0 .. x0+3), walls already
// extrapolated by the loader (2 p0 - p1).
NS_HD void ns_vstar_strip(const float (&pc)[6], const float (&pd)[4], const float (&pu)[4], const float (&u)[4],
                          const float (&v)[4], const NsConst& k, float (&us)[4], float (&vs)[4]) {
  const float h = 0.5f * k.rd;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    us[c] = u[c] - ((pc[c + 2] - pc[c]) * h) * k.dt;
    vs[c] = v[c] - ((pu[c] - pd[c]) * h) * k.dt;
  }
}

// Pressure relaxation of a strip.  *c = row y at x0-1 .. x0+4, *d / *u = rows y-1 / y+1 at x0 .. x0+3, all REFLECTED at
// the walls by the loader (op/ns_step_kernel.cu:204-234).
NS_HD void ns_pressure_strip(const float (&pc)[6], const float (&pd)[4], const float (&pu)[4], const float (&uc)[6],
                             const float (&ud)[4], const float (&uu)[4], const float (&vc)[6], const float (&vd)[4],
                             const float (&vu)[4], const NsConst& k, float (&o)[4]) {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float sxx = uc[c + 2] - uc[c], sxy = vc[c + 2] - vc[c];
    const float syx = uu[c] - ud[c], syy = vu[c] - vd[c];
    const float aver = 0.25f * (pc[c] + pc[c + 2] + pd[c] + pu[c]);
    o[c] = aver + (sxx * sxx + syy * syy + syx * sxy) * 0.125f - (sxx + syy) * k.k8;
  }
}

}  // namespace b200
