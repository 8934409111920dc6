// Host harness around csrc/ns_math.cuh: the per-cell arithmetic of the CUDA kernels run over whole fields on the CPU
// (window = the field itself, pitch = nx), so tests/test_ns_math_cpu.py can pin it to the oracle without a GPU.
#include "../b_pinn_kalman_filter_b200/csrc/ns_math.cuh"
#include <vector>
using namespace b200;

extern "C" void ns_host_density(float* out, const float* f, const float* vel, int B, int ny, int nx, float dt, float dx) {
  const NsConst k = ns_const(dt, dx);
  const size_t pl = (size_t)nx * ny;
  for (int b = 0; b < B; ++b)
    for (int y = 0; y < ny; ++y)
      for (int x = 0; x < nx; ++x) {
        const int c = y * nx + x;
        const NsCell g = ns_cell(x, y, c, nx, nx, ny, vel[b * 2 * pl + c], vel[b * 2 * pl + pl + c], k);
        out[b * pl + c] = ns_cip(f + b * pl, g, k);
      }
}

extern "C" void ns_host_velocity(float* out, const float* vel, const float* p, int B, int ny, int nx, float dt, float dx) {
  const NsConst k = ns_const(dt, dx);
  const size_t pl = (size_t)nx * ny;
  std::vector<float> us(pl), vs(pl);
  for (int b = 0; b < B; ++b) {
    for (int y = 0; y < ny; ++y)
      for (int x = 0; x < nx; ++x) {
        const int c = y * nx + x;
        us[c] = vel[b * 2 * pl + c] - ns_ddx(p + b * pl, c, x, nx, k.rd) * dt;
        vs[c] = vel[b * 2 * pl + pl + c] - ns_ddy(p + b * pl, c, y, ny, nx, k.rd) * dt;
      }
    for (int y = 0; y < ny; ++y)
      for (int x = 0; x < nx; ++x) {
        const int c = y * nx + x;
        const NsCell g = ns_cell(x, y, c, nx, nx, ny, us[c], vs[c], k);
        out[b * 2 * pl + c] = ns_cip(us.data(), g, k);
        out[b * 2 * pl + pl + c] = ns_cip(vs.data(), g, k);
      }
  }
}

extern "C" void ns_host_pressure(float* out, const float* p, const float* vel, int B, int ny, int nx, float dt, float dx) {
  const NsConst k = ns_const(dt, dx);
  const size_t pl = (size_t)nx * ny;
  for (int b = 0; b < B; ++b)
    for (int y = 0; y < ny; ++y)
      for (int x = 0; x < nx; ++x) {
        const int c = y * nx + x;
        out[b * pl + c] = ns_pressure(p + b * pl, vel + b * 2 * pl, vel + b * 2 * pl + pl, c, x, y, nx, nx, ny, k);
      }
}

// ---- strip form (what the vectorised kernels run): windows filled with NaN outside the field, so any read of a cell
// the ghost construction does not define would poison the result.
static void load_win(float (&w)[5][8], const float* f, int x0, int y, int nx, int ny) {
  for (int r = 0; r < 5; ++r)
    for (int j = 0; j < 8; ++j) {
      const int gy = y - 2 + r, gx = x0 - 2 + j;
      w[r][j] = (gy >= 0 && gy < ny && gx >= 0 && gx < nx) ? f[(size_t)gy * nx + gx] : NAN;
    }
}

extern "C" void ns_host_density_strip(float* out, const float* f, const float* vel, int B, int ny, int nx, float dt, float dx) {
  const NsConst k = ns_const(dt, dx);
  const size_t pl = (size_t)nx * ny;
  for (int b = 0; b < B; ++b)
    for (int y = 0; y < ny; ++y)
      for (int x0 = 0; x0 < nx; x0 += 4) {
        float w[5][8], uu[4], vv[4], o[4];
        load_win(w, f + b * pl, x0, y, nx, ny);
        ns_fix_ghosts(w, x0, y, nx, ny);
        for (int c = 0; c < 4; ++c) { uu[c] = vel[b * 2 * pl + y * nx + x0 + c]; vv[c] = vel[b * 2 * pl + pl + y * nx + x0 + c]; }
        ns_cip_strip(w, uu, vv, x0, y, nx, ny, k, o);
        for (int c = 0; c < 4; ++c) out[b * pl + y * nx + x0 + c] = o[c];
      }
}

// p(y, x) with the wall extrapolation (ghost = 2 p0 - p1) the v* loader applies
static float p_ext(const float* p, int x, int y, int nx, int ny) {
  if (x < 0) return 2.0f * p[(size_t)y * nx] - p[(size_t)y * nx + 1];
  if (x >= nx) return 2.0f * p[(size_t)y * nx + nx - 1] - p[(size_t)y * nx + nx - 2];
  if (y < 0) return 2.0f * p[x] - p[nx + x];
  if (y >= ny) return 2.0f * p[(size_t)(ny - 1) * nx + x] - p[(size_t)(ny - 2) * nx + x];
  return p[(size_t)y * nx + x];
}

extern "C" void ns_host_velocity_strip(float* out, const float* vel, const float* p, int B, int ny, int nx, float dt, float dx) {
  const NsConst k = ns_const(dt, dx);
  const size_t pl = (size_t)nx * ny;
  std::vector<float> us(pl), vs(pl);
  for (int b = 0; b < B; ++b) {
    const float* pp = p + b * pl;
    for (int y = 0; y < ny; ++y)
      for (int x0 = 0; x0 < nx; x0 += 4) {
        float pc[6], pd[4], pu[4], u[4], v[4], a[4], c[4];
        for (int j = 0; j < 6; ++j) pc[j] = p_ext(pp, x0 - 1 + j, y, nx, ny);
        for (int j = 0; j < 4; ++j) {
          pd[j] = p_ext(pp, x0 + j, y - 1, nx, ny); pu[j] = p_ext(pp, x0 + j, y + 1, nx, ny);
          u[j] = vel[b * 2 * pl + y * nx + x0 + j]; v[j] = vel[b * 2 * pl + pl + y * nx + x0 + j];
        }
        ns_vstar_strip(pc, pd, pu, u, v, k, a, c);
        for (int j = 0; j < 4; ++j) { us[y * nx + x0 + j] = a[j]; vs[y * nx + x0 + j] = c[j]; }
      }
    for (int y = 0; y < ny; ++y)
      for (int x0 = 0; x0 < nx; x0 += 4) {
        float w[5][8], uu[4], vv[4], o[4];
        for (int c = 0; c < 4; ++c) { uu[c] = us[y * nx + x0 + c]; vv[c] = vs[y * nx + x0 + c]; }
        load_win(w, us.data(), x0, y, nx, ny);
        ns_fix_ghosts(w, x0, y, nx, ny);
        ns_cip_strip(w, uu, vv, x0, y, nx, ny, k, o);
        for (int c = 0; c < 4; ++c) out[b * 2 * pl + y * nx + x0 + c] = o[c];
        load_win(w, vs.data(), x0, y, nx, ny);
        ns_fix_ghosts(w, x0, y, nx, ny);
        ns_cip_strip(w, uu, vv, x0, y, nx, ny, k, o);
        for (int c = 0; c < 4; ++c) out[b * 2 * pl + pl + y * nx + x0 + c] = o[c];
      }
  }
}

extern "C" void ns_host_pressure_strip(float* out, const float* p, const float* vel, int B, int ny, int nx, float dt, float dx) {
  const NsConst k = ns_const(dt, dx);
  const size_t pl = (size_t)nx * ny;
  for (int b = 0; b < B; ++b) {
    const float* f[3] = {p + b * pl, vel + b * 2 * pl, vel + b * 2 * pl + pl};
    for (int y = 0; y < ny; ++y)
      for (int x0 = 0; x0 < nx; x0 += 4) {
        float c[3][6], d[3][4], u[3][4], o[4];
        const int yd = ns_reflect(y - 1, ny), yu = ns_reflect(y + 1, ny);
        for (int q = 0; q < 3; ++q) {
          for (int j = 0; j < 6; ++j) c[q][j] = f[q][(size_t)y * nx + ns_reflect(x0 - 1 + j, nx)];
          for (int j = 0; j < 4; ++j) { d[q][j] = f[q][(size_t)yd * nx + x0 + j]; u[q][j] = f[q][(size_t)yu * nx + x0 + j]; }
        }
        ns_pressure_strip(c[0], d[0], u[0], c[1], d[1], u[1], c[2], d[2], u[2], k, o);
        for (int j = 0; j < 4; ++j) out[b * pl + y * nx + x0 + j] = o[j];
      }
  }
}
