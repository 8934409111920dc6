/*
 * ns_step_ref.c -- CPU restatement (plain C) of the reference's Navier-Stokes step.  TEST INFRASTRUCTURE ONLY:
 * imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg, never by the product package.
 *
 * The reference has NO CPU implementation of this op (op/ns_step.cpp:46-47,60-61,95-96 are CHECK_CUDA), so this file
 * restates its CUDA kernels with the launch geometry turned into loops:
 *   blockIdx.x -> x in [0, gridDim.x), blockIdx.y -> y in [0, gridDim.y), threadIdx.x -> batch b
 *   gridDim.x = size(2) (H), gridDim.y = size(3) (W); memory index = batch + y*gridDim.x + x
 *                                                             (op/ns_step_kernel.cu:30-47, 248-254)
 * The reference only indexes correctly when H == W; this oracle requires H == W.
 * Expression types (float vs double literal promotion) follow the reference statement by statement.
 * PARITY: pinned on the GPU box against the reference's own CUDA extension, built unmodified from
 * /root/reference/op/ns_step{.cpp,_kernel.cu} into oracle/_ref/ns_step_forward by oracle/build_ref_ext.py:
 * tests/test_ref_ext_gpu.py compares this file (and the product kernels) with update_density / update_velocity /
 * update_pressure at B = 1 for four grid sizes, with the B = 3 bug-compatible composition, and over a rollout.
 * Those tests FAIL (not skip) when CUDA is present and oracle/_ref is missing.
 *
 * Build: make -C oracle   (gcc -O2 -fopenmp -ffp-contract=off -shared)
 */
#include <stddef.h>

typedef struct { int nx, ny; } dims_t; /* nx = gridDim.x, ny = gridDim.y */

static inline float getf(const float* f, dims_t d, size_t batch, int x, int y) { return f[batch + (size_t)y * d.nx + x]; }

/* op/ns_step_kernel.cu:49-75 */
static inline float diff_x(const float* f, dims_t d, size_t batch, int x, int y, float dx) {
  if (x == 0) return (getf(f, d, batch, x + 1, y) - getf(f, d, batch, x, y)) / dx;
  else if (x == d.nx - 1) return (getf(f, d, batch, x, y) - getf(f, d, batch, x - 1, y)) / dx;
  else return (getf(f, d, batch, x + 1, y) - getf(f, d, batch, x - 1, y)) / dx / 2;
}
static inline float diff_y(const float* f, dims_t d, size_t batch, int x, int y, float dx) {
  if (y == 0) return (getf(f, d, batch, x, y + 1) - getf(f, d, batch, x, y)) / dx;
  else if (y == d.ny - 1) return (getf(f, d, batch, x, y) - getf(f, d, batch, x, y - 1)) / dx;
  else return (getf(f, d, batch, x, y + 1) - getf(f, d, batch, x, y - 1)) / dx / 2;
}
/* :77-83 */
static inline int clamp_x(int x, dims_t d) { return x < 0 ? -x : (x > d.nx - 1 ? 2 * d.nx - 2 - x : x); }
static inline int clamp_y(int y, dims_t d) { return y < 0 ? -y : (y > d.ny - 1 ? 2 * d.ny - 2 - y : y); }
/* :85-94 */
static inline int sgn(float v) { return v < 0.0 ? -1 : (v > 0.0 ? 1 : 0); }

/* update_gradient_kernel :96-112 -- dense [B,1,H,W] addressing from `field` */
void ref_update_gradient(float* df_dx, float* df_dy, const float* field, int B, int H, int W, float dx) {
  dims_t d = {H, W};
#pragma omp parallel for collapse(2)
  for (int b = 0; b < B; ++b)
    for (int y = 0; y < d.ny; ++y)
      for (int x = 0; x < d.nx; ++x) {
        size_t batch = (size_t)b * d.nx * d.ny;
        size_t loc = batch + (size_t)y * d.nx + x;
        df_dx[loc] = diff_x(field, d, batch, x, y, dx);
        df_dy[loc] = diff_y(field, d, batch, x, y, dx);
      }
}

/* cip_advect_kernel :114-158.  vel is [B,2,H,W] (get_v :38-47). */
void ref_cip_advect(float* dens_n, const float* dens_c, const float* dens_dx, const float* dens_dy, const float* vel_c,
                    int B, int H, int W, float dt, float dx) {
  dims_t d = {H, W};
#pragma omp parallel for collapse(2)
  for (int b = 0; b < B; ++b)
    for (int y = 0; y < d.ny; ++y)
      for (int x = 0; x < d.nx; ++x) {
        size_t plane = (size_t)d.nx * d.ny;
        size_t batch = (size_t)b * plane;
        float vx = vel_c[(size_t)b * plane * 2 + (size_t)y * d.nx + x];
        float vy = vel_c[(size_t)b * plane * 2 + plane + (size_t)y * d.nx + x];
        int x_s = sgn(vx), y_s = sgn(vy);
        int x_m = clamp_x(x - x_s, d), y_m = clamp_y(y - y_s, d);
#define G(f, xx, yy) getf(f, d, batch, xx, yy)
        float tmp1 = G(dens_c, x, y) - G(dens_c, x, y_m) - G(dens_c, x_m, y) + G(dens_c, x_m, y_m);
        float tmp2 = G(dens_c, x_m, y) - G(dens_c, x, y);
        float tmp3 = G(dens_c, x, y_m) - G(dens_c, x, y);
        float x_s_denom = x_s * dx * dx * dx;
        float y_s_denom = y_s * dx * dx * dx;
        float a = (x_s * (G(dens_dx, x_m, y) + G(dens_dx, x, y)) * dx - 2.0 * (-tmp2)) / x_s_denom;
        float bq = (y_s * (G(dens_dy, x, y_m) + G(dens_dy, x, y)) * dx - 2.0 * (-tmp3)) / y_s_denom;
        float c = (-tmp1 - x_s * (G(dens_dx, x, y_m) - G(dens_dx, x, y)) * dx) / y_s_denom;
        float dd = (-tmp1 - y_s * (G(dens_dy, x_m, y) - G(dens_dy, x, y)) * dx) / x_s_denom;
        float e = (3.0 * tmp2 + x_s * (G(dens_dx, x_m, y) + 2.0 * G(dens_dx, x, y)) * dx) / dx / dx;
        float f = (3.0 * tmp3 + y_s * (G(dens_dy, x, y_m) + 2.0 * G(dens_dy, x, y)) * dx) / dx / dx;
        float g = (-(G(dens_dy, x_m, y) - G(dens_dy, x, y)) + c * dx * dx) / (x_s * dx);
        float X = -vx * dt;
        float Y = -vy * dt;
        dens_n[batch + (size_t)y * d.nx + x] =
            (((a * X + c * Y + e) * X + g * Y + G(dens_dx, x, y)) * X + ((bq * Y + dd * X + f) * Y + G(dens_dy, x, y)) * Y +
             G(dens_c, x, y));
#undef G
      }
}

/* velocity_update_kernel :180-202 */
void ref_velocity_nonadvect(float* vel_n, const float* vel_c, const float* pres_dx, const float* pres_dy, int B, int H,
                            int W, float dt) {
  dims_t d = {H, W};
#pragma omp parallel for collapse(2)
  for (int b = 0; b < B; ++b)
    for (int y = 0; y < d.ny; ++y)
      for (int x = 0; x < d.nx; ++x) {
        size_t plane = (size_t)d.nx * d.ny;
        size_t batch0 = (size_t)b * plane * 2, batch1 = batch0 + plane, pb = (size_t)b * plane;
        size_t o = (size_t)y * d.nx + x;
        vel_n[batch0 + o] = vel_c[batch0 + o] - pres_dx[pb + o] * dt;
        vel_n[batch1 + o] = vel_c[batch1 + o] - pres_dy[pb + o] * dt;
      }
}

/* pressure_update_kernel :204-234 */
void ref_update_pressure(float* pres_n, const float* pres_c, const float* vel_c, int B, int H, int W, float dt, float dx) {
  dims_t d = {H, W};
#pragma omp parallel for collapse(2)
  for (int b = 0; b < B; ++b)
    for (int y = 0; y < d.ny; ++y)
      for (int x = 0; x < d.nx; ++x) {
        size_t plane = (size_t)d.nx * d.ny;
        size_t batch = (size_t)b * plane;
        const float* u = vel_c + (size_t)b * plane * 2;
        const float* v = u + plane;
        int x_u = clamp_x(x + 1, d), x_d = clamp_x(x - 1, d), y_u = clamp_y(y + 1, d), y_d = clamp_y(y - 1, d);
#define V(f, xx, yy) f[(size_t)(yy)*d.nx + (xx)]
        float sub_x_x = V(u, x_u, y) - V(u, x_d, y), sub_x_y = V(v, x_u, y) - V(v, x_d, y);
        float sub_y_x = V(u, x, y_u) - V(u, x, y_d), sub_y_y = V(v, x, y_u) - V(v, x, y_d);
#undef V
        float aver_p = 0.25 * (getf(pres_c, d, batch, x_d, y) + getf(pres_c, d, batch, x_u, y) +
                               getf(pres_c, d, batch, x, y_d) + getf(pres_c, d, batch, x, y_u));
        float pred_p = (aver_p + (sub_x_x * sub_x_x + sub_y_y * sub_y_y + (sub_y_x * sub_x_y)) / 8.0 -
                        dx * (sub_x_x + sub_y_y) / (8 * dt));
        pres_n[batch + (size_t)y * d.nx + x] = pred_p;
      }
}

/* Host composites, op/ns_step.cpp:45-102.  tmp buffers are caller-provided (each B*H*W floats unless noted). */
void ref_update_density(float* dens_n, const float* dens_c, const float* vel_c, float* tmp_dx, float* tmp_dy, int B,
                        int H, int W, float dt, float dx) {
  ref_update_gradient(tmp_dx, tmp_dy, dens_c, B, H, W, dx);
  ref_cip_advect(dens_n, dens_c, tmp_dx, tmp_dy, vel_c, B, H, W, dt, dx);
}

/* bug_compat = 1 reproduces what the reference does for any B: `u`/`v` are views into vel_n (data pointers
 * vel_n and vel_n + H*W) that the kernels address with a DENSE batch stride H*W (op/ns_step.cpp:70-72).
 * bug_compat = 0 is the per-sample reading (identical at B = 1): u/v are gathered into dense planes first.
 * vel_n: B*2*H*W scratch; u_n, v_n, tmp_dx, tmp_dy, uc, vc: B*H*W each; out: B*2*H*W. */
void ref_update_velocity(float* out, const float* vel_c, const float* pres_c, float* vel_n, float* tmp_dx, float* tmp_dy,
                         float* u_n, float* v_n, float* uc, float* vc, int B, int H, int W, float dt, float dx,
                         int bug_compat) {
  size_t plane = (size_t)H * W;
  ref_update_gradient(tmp_dx, tmp_dy, pres_c, B, H, W, dx);
  ref_velocity_nonadvect(vel_n, vel_c, tmp_dx, tmp_dy, B, H, W, dt);
  const float *u = vel_n, *v = vel_n + plane;
  if (!bug_compat) {
    for (int b = 0; b < B; ++b)
      for (size_t i = 0; i < plane; ++i) {
        uc[b * plane + i] = vel_n[(size_t)b * 2 * plane + i];
        vc[b * plane + i] = vel_n[(size_t)b * 2 * plane + plane + i];
      }
    u = uc; v = vc;
  }
  ref_update_gradient(tmp_dx, tmp_dy, u, B, H, W, dx);
  ref_cip_advect(u_n, u, tmp_dx, tmp_dy, vel_n, B, H, W, dt, dx);
  ref_update_gradient(tmp_dx, tmp_dy, v, B, H, W, dx);
  ref_cip_advect(v_n, v, tmp_dx, tmp_dy, vel_n, B, H, W, dt, dx);
  for (int b = 0; b < B; ++b) /* torch::cat({u_n, v_n}, 1) */
    for (size_t i = 0; i < plane; ++i) {
      out[(size_t)b * 2 * plane + i] = u_n[b * plane + i];
      out[(size_t)b * 2 * plane + plane + i] = v_n[b * plane + i];
    }
}
