// DFMA / fp64 special-function throughput micro-benchmark (roofline denominator for the fp64 pipe).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts_dev/fp64_peak scripts_dev/fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP>
__global__ void k_dfma(double* out, double a, double b, int iters) {
    double x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) x[i] = fma(x[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_div(double* out, double a, int iters) {
    double x = 1.0 + threadIdx.x * 1e-3, y = 2.0 + threadIdx.x * 1e-3;
    for (int it = 0; it < iters; ++it) { x = a / x + 1.0; y = a / y + 1.0; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x + y;
}
__global__ void k_sincos(double* out, int iters) {
    double x = 0.1 + threadIdx.x * 1e-3, acc = 0;
    for (int it = 0; it < iters; ++it) { double s, c; sincos(x, &s, &c); acc += s * c; x += 1e-3; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
int main() {
    int sm; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
    double* out; cudaMalloc(&out, sizeof(double) * sm * 8 * 1024);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms;
    for (int tpb : {128, 256, 512, 1024}) {
        const int iters = 1 << 14, grid = sm * (2048 / tpb);
        k_dfma<8><<<grid, tpb>>>(out, 1.0000001, 1e-9, iters);
        cudaEventRecord(e0);
        for (int r = 0; r < 5; ++r) k_dfma<8><<<grid, tpb>>>(out, 1.0000001, 1e-9, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
        double fl = 5.0 * grid * tpb * (double)iters * 8 * 2;
        printf("dfma tpb=%d grid=%d: %.2f TFLOP/s\n", tpb, grid, fl / (ms * 1e-3) / 1e12);
    }
    {
        const int iters = 1 << 12, tpb = 256, grid = sm * 8;
        k_div<<<grid, tpb>>>(out, 3.0, iters);
        cudaEventRecord(e0); k_div<<<grid, tpb>>>(out, 3.0, iters); cudaEventRecord(e1);
        cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
        printf("ddiv: %.2f Gdiv/s\n", 2.0 * grid * tpb * (double)iters / (ms * 1e-3) / 1e9);
        k_sincos<<<grid, tpb>>>(out, iters);
        cudaEventRecord(e0); k_sincos<<<grid, tpb>>>(out, iters); cudaEventRecord(e1);
        cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
        printf("sincos: %.2f G/s\n", 1.0 * grid * tpb * (double)iters / (ms * 1e-3) / 1e9);
    }
    printf("sm=%d err=%s\n", sm, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
