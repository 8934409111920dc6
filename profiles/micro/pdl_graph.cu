// Per-node cost of a chain of small kernels replayed as a CUDA graph, with and without programmatic dependent launch.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o pdl_graph pdl_graph.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void small(float* x, int n, int spin) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { float v = x[i]; for (int k = 0; k < spin; ++k) v = v * 1.0001f + 0.5f; x[i] = v; }
}
static float run(bool pdl, int nodes, int blocks, int spin, float* x, int n) {
    cudaStream_t s; cudaStreamCreate(&s);
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < nodes; ++i) {
        cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(256); cfg.stream = s;
        cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
        cudaLaunchKernelEx(&cfg, small, x, n, spin);
    }
    cudaStreamEndCapture(s, &g);
    if (cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) { printf("instantiate failed\n"); return -1; }
    for (int i = 0; i < 3; ++i) cudaGraphLaunch(ge, s);
    cudaStreamSynchronize(s);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, s); for (int i = 0; i < 10; ++i) cudaGraphLaunch(ge, s); cudaEventRecord(e1, s); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    return ms / 10 / nodes * 1e3f;
}
int main() {
    int n = 148 * 8 * 256; float* x; cudaMalloc(&x, n * 4); cudaMemset(x, 0, n * 4);
    for (int spin : {0, 2000}) for (int blocks : {148, 148 * 8}) {
        float a = run(false, 400, blocks, spin, x, n), b = run(true, 400, blocks, spin, x, n);
        printf("blocks %5d spin %5d: %.2f us per node plain, %.2f us with programmatic dependent launch (%s)\n", blocks, spin, a, b, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
