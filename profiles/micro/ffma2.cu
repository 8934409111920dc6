// FFMA vs FFMA2 (fma.rn.f32x2) issue throughput on sm_100a.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -o ffma2 ffma2.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b) {
    unsigned long long r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b) {
    unsigned long long r; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
template <int MODE> __global__ void k(float* out, int iters, float s) {
    float a[16]; unsigned long long p[8];
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
    for (int i = 0; i < 8; ++i) { float2 v = make_float2(a[2*i], a[2*i+1]); p[i] = *reinterpret_cast<unsigned long long*>(&v); }
    float2 sv = make_float2(s, s); unsigned long long sp = *reinterpret_cast<unsigned long long*>(&sv);
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(a[i]) : "f"(s));
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], sp, sp);
        } else if (MODE == 2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = add2(p[i], sp);
        } else if (MODE == 3) {
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = mul2(p[i], sp);
        } else {   // mixed: 8 FFMA2 + 8 integer ops
#pragma unroll
            for (int i = 0; i < 8; ++i) { p[i] = fma2(p[i], sp, sp); asm volatile("lop3.b32 %0, %0, %1, %1, 0x96;" : "+r"(*reinterpret_cast<unsigned*>(&a[i])) : "r"(it)); }
        }
    }
    float acc = 0; for (int i = 0; i < 16; ++i) acc += a[i];
    for (int i = 0; i < 8; ++i) { float2 v = *reinterpret_cast<float2*>(&p[i]); acc += v.x + v.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int MODE> void run(const char* name, float* out) {
    int iters = 4096; cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * 4, 512>>>(out, iters, 1.0001f); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<MODE><<<148 * 4, 512>>>(out, iters, 1.0001f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double lanes = 148.0 * 4 * 512 * (double)iters * 16;   // fp32 lane-operations (16 per iter per thread in every mode)
    printf("%-28s %.3f ms  %.2f T lane-ops/s\n", name, ms, lanes / ms * 1e-9);
}
int main() { float* out; cudaMalloc(&out, 148 * 4 * 512 * 4);
    run<0>("16 x FFMA", out); run<1>("8 x FFMA2", out); run<2>("8 x FADD2", out); run<3>("8 x FMUL2", out); run<4>("8 x FFMA2 + 8 x LOP3", out);
    return 0; }
