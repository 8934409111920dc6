// Device side of the adaptive Runge-Kutta integrator behind get_ode_sampler / likelihood / DPS (the reference hands
// these to scipy.integrate.solve_ivp with a D2H + H2D of the whole state per evaluation: sampling.py:459-473,
// likelihood.py:86-99, inverse/conditional_sampling.py:10-29).  State y lives in HBM as fp64 (scipy integrates in
// fp64), stage derivatives K[j] as the fp32 the score network produced; combinations are evaluated in fp64.
//   stage : x32 = fp32(y + h * sum_j a[j] K[j])      (and optionally y_new in fp64)
//   error : partial[b] = sum over the block's elements of (h * sum_j e[j] K[j] / (atol + rtol*max(|y|,|y_new|)))^2
// Both are one pass over the operands (HBM-bound): 8 + 4*ns (+8) + 4 bytes per element.
#include "common.cuh"
#include "../../include/b200_sde.h"

namespace b200 {

struct RkCoef { double c[B200_RK_MAX_STAGES]; };

constexpr int RK_THREADS = 256;

template <int NS>
__global__ void __launch_bounds__(RK_THREADS) rk_stage_kernel(float* __restrict__ x32, double* __restrict__ y_out,
                                                              const double* __restrict__ y, const float* __restrict__ K,
                                                              long long kstride, RkCoef a, double h, long long n) {
  const long long stride = (long long)gridDim.x * RK_THREADS;
  for (long long i = (long long)blockIdx.x * RK_THREADS + threadIdx.x; i < n; i += stride) {
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < NS; ++j) acc = fma(a.c[j], (double)__ldg(K + j * kstride + i), acc);
    const double v = fma(h, acc, y[i]);
    if (x32) x32[i] = (float)v;
    if (y_out) y_out[i] = v;
  }
}

template <int NS>
__global__ void __launch_bounds__(RK_THREADS) rk_error_kernel(double* __restrict__ partial, const double* __restrict__ y,
                                                              const double* __restrict__ y_new,
                                                              const float* __restrict__ K, long long kstride, RkCoef e,
                                                              double h, double atol, double rtol, long long n) {
  __shared__ double red[RK_THREADS / 32];
  const long long stride = (long long)gridDim.x * RK_THREADS;
  double s = 0.0;
  for (long long i = (long long)blockIdx.x * RK_THREADS + threadIdx.x; i < n; i += stride) {
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < NS; ++j) acc = fma(e.c[j], (double)__ldg(K + j * kstride + i), acc);
    const double scale = atol + fmax(fabs(y[i]), fabs(y_new[i])) * rtol;
    const double r = acc * h / scale;
    s = fma(r, r, s);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < RK_THREADS / 32; ++w) t += red[w];      // fixed order: the sum is reproducible
    partial[blockIdx.x] = t;
  }
}

static inline unsigned rk_grid(long long n) {
  long long g = (n + RK_THREADS - 1) / RK_THREADS;
  const long long cap = B200_RK_PARTIALS;
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace b200

using namespace b200;

template <int NS> static void rk_stage_launch(float* x32, double* y_out, const double* y, const float* K, long long ks,
                                              const RkCoef& a, double h, long long n, cudaStream_t st) {
  rk_stage_kernel<NS><<<rk_grid(n), RK_THREADS, 0, st>>>(x32, y_out, y, K, ks, a, h, n);
}

extern "C" int b200_rk_stage(float* x32_out, double* y_out, const double* y, const float* K, long long kstride, int ns,
                             const double* a_host, double h, long long n, cudaStream_t st) {
  if ((!x32_out && !y_out) || !y || !K || !a_host || ns < 1 || ns > B200_RK_MAX_STAGES || n < 0 || kstride < n)
    return B200_ERR_ARG;
  if (n == 0) return B200_OK;
  RkCoef a{};
  for (int j = 0; j < ns; ++j) a.c[j] = a_host[j];
  switch (ns) {
    case 1: rk_stage_launch<1>(x32_out, y_out, y, K, kstride, a, h, n, st); break;
    case 2: rk_stage_launch<2>(x32_out, y_out, y, K, kstride, a, h, n, st); break;
    case 3: rk_stage_launch<3>(x32_out, y_out, y, K, kstride, a, h, n, st); break;
    case 4: rk_stage_launch<4>(x32_out, y_out, y, K, kstride, a, h, n, st); break;
    case 5: rk_stage_launch<5>(x32_out, y_out, y, K, kstride, a, h, n, st); break;
    case 6: rk_stage_launch<6>(x32_out, y_out, y, K, kstride, a, h, n, st); break;
    default: rk_stage_launch<7>(x32_out, y_out, y, K, kstride, a, h, n, st); break;
  }
  B200_RETURN_LAST();
}

extern "C" int b200_rk_error(double* partial, const double* y, const double* y_new, const float* K, long long kstride,
                             int ns, const double* e_host, double h, double atol, double rtol, long long n,
                             cudaStream_t st) {
  if (!partial || !y || !y_new || !K || !e_host || ns < 1 || ns > B200_RK_MAX_STAGES || n < 0 || kstride < n)
    return B200_ERR_ARG;
  // all B200_RK_PARTIALS slots are defined after the call: unused ones are zeroed
  cudaError_t e = cudaMemsetAsync(partial, 0, sizeof(double) * B200_RK_PARTIALS, st);
  if (e != cudaSuccess) return -(1000 + (int)e);
  if (n == 0) return B200_OK;
  RkCoef c{};
  for (int j = 0; j < ns; ++j) c.c[j] = e_host[j];
  // zero coefficients cost a plane of reads each; the caller passes all stages, so keep one instantiation
  if (ns == 7) rk_error_kernel<7><<<rk_grid(n), RK_THREADS, 0, st>>>(partial, y, y_new, K, kstride, c, h, atol, rtol, n);
  else if (ns == 4) rk_error_kernel<4><<<rk_grid(n), RK_THREADS, 0, st>>>(partial, y, y_new, K, kstride, c, h, atol, rtol, n);
  else return B200_ERR_UNSUPPORTED;
  B200_RETURN_LAST();
}
