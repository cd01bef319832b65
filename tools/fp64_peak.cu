// FP64 peak microbenchmark: DFMA issue rate and DMMA (mma.sync ... f64) rate on
// this GPU, so the FP64 side of the roofline has a measured denominator.  tcgen05 has no FP64 kind; the FP64
// tensor path on sm_100a is the warp-level mma.sync.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak fp64_peak.cu && ./fp64_peak
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;

__global__ void __launch_bounds__(256) dfma_kernel(double *out, double a, double b)
{
    double acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        acc[i] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < ITERS; ++it)
    {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            acc[i] = fma(acc[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) dmma884_kernel(double *out, double a, double b)
{
    double c[4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        c[i][0] = c[i][1] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < ITERS; ++it)
    {
#pragma unroll
        for (int i = 0; i < 4; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1])
                         : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
        s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) dmma1688_kernel(double *out, double a, double b)
{
    double c[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            c[i][j] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < ITERS; ++it)
    {
#pragma unroll
        for (int i = 0; i < 4; ++i)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                         : "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            s += c[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class K>
double time_kernel(K kernel, int grid, double *out)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int w = 0; w < 3; ++w)
        kernel<<<grid, 256>>>(out, 1.0000001, 1e-9);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 10; ++r)
    {
        cudaEventRecord(e0);
        kernel<<<grid, 256>>>(out, 1.0000001, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        best = ms < best ? ms : best;
    }
    return best * 1e-3;
}

int main()
{
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int grid = prop.multiProcessorCount * 8;
    double *out;
    cudaMalloc(&out, sizeof(double) * grid * 256);
    const double threads = (double)grid * 256;
    const double t_fma = time_kernel(dfma_kernel, grid, out);
    const double t_884 = time_kernel(dmma884_kernel, grid, out);
    const double t_1688 = time_kernel(dmma1688_kernel, grid, out);
    const double f_fma = threads * ITERS * 8 * 2 / t_fma;
    const double f_884 = (threads / 32) * ITERS * 4 * (2.0 * 8 * 8 * 4) / t_884;
    const double f_1688 = (threads / 32) * ITERS * 4 * (2.0 * 16 * 8 * 8) / t_1688;
    printf("{\"gpu\": \"%s\", \"sm_count\": %d, \"fp64_fma_tflops\": %.3f, \"fp64_dmma_m8n8k4_tflops\": %.3f, "
           "\"fp64_dmma_m16n8k8_tflops\": %.3f}\n",
           prop.name, prop.multiProcessorCount, f_fma / 1e12, f_884 / 1e12, f_1688 / 1e12);
    if (cudaGetLastError() != cudaSuccess)
        return 1;
    return 0;
}
